B2_ATTN_MMA=0 timeout 900 python bench.py --workload mistral-7b-q4_k_m-8k --no-cpu-baseline --batch-extra 0 > gpurun_out/mis_scalar.json 2> gpurun_out/mis_scalar.err
timeout 900 python bench.py --workload mistral-7b-q4_k_m-8k --no-cpu-baseline --batch-extra 0 --no-parity-check > gpurun_out/mis_mma.json 2> gpurun_out/mis_mma.err
for ns in 16 32; do B2_ATTN_SPLITS=$ns B2_ATTN_MMA=0 timeout 900 python bench.py --workload mistral-7b-q4_k_m-8k --no-cpu-baseline --batch-extra 0 --no-parity-check 2>/dev/null | grep -o '"ms_per_step": [0-9.]*'; done
