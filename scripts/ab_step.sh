P=llama-fastapi-k8s-gpu_b200
for v in lib_v1 lib_head lib_v1; do
  echo "== $v"
  B200_GGUF_LIB=$P/$v/libb200gguf.so timeout 120 python bench.py --steps 30 --warmup 5 --no-cpu-baseline --batch-extra 0 --no-parity-check 2>&1 | grep -o '"ms_per_step": [0-9.]*'
done
echo "== lib, no tma2d"
B2_MV_NO_TMA2D=1 timeout 120 python bench.py --steps 30 --warmup 5 --no-cpu-baseline --batch-extra 0 --no-parity-check 2>&1 | grep -o '"ms_per_step": [0-9.]*'
