P=llama-fastapi-k8s-gpu_b200
for st in 1 2 3 4 5; do
B200_GGUF_LIB=$P/lib_as$st/libb200gguf.so B2_TIMELINE=1 timeout 120 python bench.py --steps 10 --warmup 5 --no-cpu-baseline --batch-extra 0 --no-parity-check 2>&1 | grep -E "timeline\] (attn)" | head -1 | cut -c1-150
done
