for m in default 0; do
  if [ $m = 0 ]; then export B2_ATTN_MMA=0; fi
  for sp in 1 2; do B2_BATCH_ATTN_SPLITS=$sp timeout 300 python bench.py --workload llama3-8b-q8_0-bs32 --no-cpu-baseline --no-parity-check --steps 32 --warmup 4 2>/dev/null | python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('mma=$m splits=$sp', d['ms_per_step'], d['value'])"; done
done
