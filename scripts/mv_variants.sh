S="qkv_q4k o_q4k gateup_q4k down_q4k down_q6k head_q6k"
export MV_PLANES=2
TAG=base python scripts/mv_micro.py 100 1 $S
TAG=8x2 B200_GGUF_LIB=llama-fastapi-k8s-gpu_b200/lib_8x2/libb200gguf.so python scripts/mv_micro.py 100 1 $S
