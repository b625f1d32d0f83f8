S="qkv_q4k o_q4k gateup_q4k down_q4k down_q6k head_q6k gateup_q8"
export MV_PLANES=2
TAG=rows4 python scripts/mv_micro.py 100 1 $S
TAG=rows2 B2_MV_NO_ROWS4=1 python scripts/mv_micro.py 100 1 $S
TAG=rows4_u4 B200_GGUF_LIB=llama-fastapi-k8s-gpu_b200/lib_u4/libb200gguf.so python scripts/mv_micro.py 100 1 $S
