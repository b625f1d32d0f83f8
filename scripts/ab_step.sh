# A/B of whole-step time between library builds (B200_GGUF_LIB picks the .so)
P=llama-fastapi-k8s-gpu_b200
for v in lib_head lib_cur lib; do
  for rep in 1 2; do
    echo "== $v rep $rep"
    B200_GGUF_LIB=$P/$v/libb200gguf.so python bench.py --steps 30 --warmup 5 --no-cpu-baseline --batch-extra 0 --no-parity-check 2>&1 | grep -o '"ms_per_step": [0-9.]*'
  done
done
echo "== timeline lib"
B2_TIMELINE=1 python bench.py --steps 10 --warmup 5 --no-cpu-baseline --batch-extra 0 --no-parity-check 2>&1 | grep -iE "timeline|qkv|attn|gate|down|lm_head|ms_per_step" | head -20
