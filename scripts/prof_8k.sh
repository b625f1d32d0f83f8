B="python bench.py --workload mistral-7b-q4_k_m-8k --steps 2 --warmup 1 --ncu-mode --no-cpu-baseline --batch-extra 0 --no-parity-check"
timeout 600 $B > gpurun_out/r02_8k_plain.log 2>&1 || exit 1
timeout 900 ncu --set full --clock-control none --import-source on --kernel-name 'regex:attn_decode_kernel' --launch-skip 40 --launch-count 1 -f -o gpurun_out/r02_attn8k $B > gpurun_out/r02_8k_ncu.log 2>&1
ls -la gpurun_out/r02_attn8k.ncu-rep
