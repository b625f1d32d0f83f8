"""B200-native drop-in for the `llama_cpp` package that /root/reference/api.py:7 imports
(`from llama_cpp import Llama`). Put this directory's parent on PYTHONPATH and the reference's
api.py runs unchanged on the new engine."""
from .llama import Llama

__version__ = "0.2.77+b200"
__all__ = ["Llama", "__version__"]
