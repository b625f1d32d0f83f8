S="qkv_q4k o_q4k gateup_q4k down_q4k down_q6k head_q6k gateup_q8"
export MV_PLANES=2
TAG=base python scripts/mv_micro.py 100 1
TAG=solo B2_MV_SOLO=1 python scripts/mv_micro.py 100 1 $S
TAG=2cta MV_CTAS=296 python scripts/mv_micro.py 100 1 $S
TAG=nopdl python scripts/mv_micro.py 100 0 $S
TAG=planes1 MV_PLANES=1 python scripts/mv_micro.py 100 1 $S
for t in 8192 10240; do TAG=tile$t B2_MV_TILE=$t python scripts/mv_micro.py 100 1 gateup_q4k down_q4k down_q6k head_q6k; done
export B200_GGUF_LIB=llama-fastapi-k8s-gpu_b200/lib_skip/libb200gguf.so
TAG=skip python scripts/mv_micro.py 100 1 $S
TAG=skip_solo B2_MV_SOLO=1 python scripts/mv_micro.py 100 1 $S
