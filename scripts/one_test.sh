timeout 1200 python -m pytest tests -m gpu -q -x 2>&1 | tail -2
timeout 600 python bench.py --workload llama3-8b-q8_0-bs32 --no-cpu-baseline --steps 32 --warmup 4 2>/dev/null | python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('bs32', d['ms_per_step'], d['value'], 'prefill p50', d.get('prefill_p50_ms') or d.get('batched'))"
