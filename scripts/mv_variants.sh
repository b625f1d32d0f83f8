S="qkv_q4k o_q4k gateup_q4k down_q4k down_q6k head_q6k gateup_q8"
export MV_PLANES=2
TAG=base python scripts/mv_micro.py 100 1
TAG=planes1 MV_PLANES=1 python scripts/mv_micro.py 100 1 $S
TAG=nopdl python scripts/mv_micro.py 100 0 $S
TAG=unroll1 B200_GGUF_LIB=llama-fastapi-k8s-gpu_b200/lib_u1/libb200gguf.so python scripts/mv_micro.py 100 1 $S
TAG=skip B200_GGUF_LIB=llama-fastapi-k8s-gpu_b200/lib_skip/libb200gguf.so python scripts/mv_micro.py 100 1 $S
