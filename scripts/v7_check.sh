timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
for rep in 1 2; do timeout 200 python bench.py --steps 30 --warmup 5 --no-cpu-baseline --batch-extra 0 --no-parity-check 2>&1 | grep -o '"ms_per_step": [0-9.]*'; done
timeout 900 python bench.py --workload mistral-7b-q4_k_m-8k --no-cpu-baseline --batch-extra 0 2>&1 | grep -o '"ms_per_step": [0-9.]*\|frac_of_logit_std": [0-9.]*\|Error.*'
timeout 600 python bench.py --workload llama3-8b-q8_0-bs32 --no-cpu-baseline 2>&1 | grep -o '"ms_per_step": [0-9.]*'
