timeout 1200 python -m pytest tests -m gpu -q 2>&1 | tail -6
timeout 200 python bench.py --steps 30 --warmup 5 --no-cpu-baseline --batch-extra 0 --no-parity-check 2>&1 | grep -o '"ms_per_step": [0-9.]*\|"prefill_ms": [0-9.]*'
