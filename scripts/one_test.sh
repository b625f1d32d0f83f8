P=llama-fastapi-k8s-gpu_b200
timeout 900 python -m pytest tests -m gpu -q -x 2>&1 | tail -3
for v in lib lib_v1 lib lib_v1; do
  echo "== $v"
  B200_GGUF_LIB=$P/$v/libb200gguf.so timeout 200 python bench.py --steps 30 --warmup 5 --no-cpu-baseline --batch-extra 0 --no-parity-check 2>&1 | grep -o '"ms_per_step": [0-9.]*'
done
