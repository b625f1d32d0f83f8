B2_ATTN_MMA=1 B2_TIMELINE=1 timeout 120 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --batch-extra 0 --no-parity-check 2>&1 | grep -E "timeline\] |ms_per_step" | head -9 | cut -c1-200
