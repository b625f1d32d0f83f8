for rep in 1 2; do timeout 200 python bench.py --steps 30 --warmup 5 --no-cpu-baseline --batch-extra 0 --no-parity-check 2>&1 | grep -o '"ms_per_step": [0-9.]*'; done
B2_TIMELINE=1 timeout 200 python bench.py --steps 10 --warmup 5 --no-cpu-baseline --batch-extra 0 --no-parity-check 2>&1 | grep "timeline" | head -9
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
