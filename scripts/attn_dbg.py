import sys, os
sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", "llama-fastapi-k8s-gpu_b200"))
import torch
from llama_cpp import _cabi
lib = _cabi.load_library(os.environ.get("B200_GGUF_LIB"))
n_head, n_kv, hd, bt = 32, 8, 128, 64
dev = torch.device("cuda:0")
ctx, ns = 150, 8
nblk = (ctx + bt - 1) // bt + 1
k = torch.randn(nblk, n_kv, bt, hd, device=dev).half(); v = torch.randn(nblk, n_kv, bt, hd, device=dev).half()
q = torch.randn(n_head, hd, device=dev); out = torch.empty(n_head, hd, device=dev)
table = torch.arange(nblk, dtype=torch.int32, device=dev); pos = torch.tensor([ctx - 1], dtype=torch.int32, device=dev)
flush = torch.empty(512 << 20, dtype=torch.uint8, device=dev)
for i in range(6):
    if i >= 3: flush.zero_(); torch.cuda.synchronize()
    rc = lib.b2_attention_decode_static(0, q.data_ptr(), k.data_ptr(), v.data_ptr(), table.data_ptr(), nblk, pos.data_ptr(),
                                        n_head, n_kv, hd, bt, ns, out.data_ptr(), torch.cuda.current_stream().cuda_stream)
    assert rc == 0
