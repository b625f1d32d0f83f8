# A/B of whole-step time between library builds (B200_GGUF_LIB picks the .so)
P=llama-fastapi-k8s-gpu_b200
for v in lib lib_head lib lib_head; do
  echo "== $v"
  B200_GGUF_LIB=$P/$v/libb200gguf.so timeout 200 python bench.py --steps 30 --warmup 5 --no-cpu-baseline --batch-extra 0 --no-parity-check 2>&1 | grep -o '"ms_per_step": [0-9.]*'
done
B2_TIMELINE=1 timeout 200 python bench.py --steps 10 --warmup 5 --no-cpu-baseline --batch-extra 0 --no-parity-check 2>&1 | grep timeline | head -8
