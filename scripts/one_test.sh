P=llama-fastapi-k8s-gpu_b200
timeout 900 python -m pytest tests -m gpu -q -x 2>&1 | tail -3
for v in lib lib_head; do
  for w in llama3-8b-q4_k_m-bs1 llama3-8b-q8_0-bs1 tinyllama-1.1b-q4_0-bs1; do
  echo "== $v $w"
  B200_GGUF_LIB=$P/$v/libb200gguf.so timeout 200 python bench.py --workload $w --steps 30 --warmup 5 --no-cpu-baseline --batch-extra 0 --no-parity-check 2>&1 | grep -o '"ms_per_step": [0-9.]*'
  done
done
