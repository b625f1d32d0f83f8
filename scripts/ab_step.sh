# A/B of whole-step time between library builds (B200_GGUF_LIB picks the .so) and knobs
P=llama-fastapi-k8s-gpu_b200
run() { timeout 200 python bench.py --steps 30 --warmup 5 --no-cpu-baseline --batch-extra 0 --no-parity-check 2>&1 | grep -o '"ms_per_step": [0-9.]*'; }
echo "== lib"; run
echo "== unroll2"; B200_GGUF_LIB=$P/lib_v1/libb200gguf.so run
echo "== stages 24"; B2_MV_STAGES=24 run
echo "== stages 16"; B2_MV_STAGES=16 run
echo "== lib"; run
