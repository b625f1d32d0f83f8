# r02 profiling pass: plain run first (must exit 0), then the launch list, then one full capture of a decode layer
set -x
B="python bench.py --steps 2 --warmup 1 --ncu-mode --no-cpu-baseline --batch-extra 0 --no-parity-check"
timeout 300 $B > gpurun_out/r02_plain.log 2>&1 || exit 1
timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 1400 --csv \
  --log-file gpurun_out/r02_launches.csv $B > gpurun_out/r02_ncu_list.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on --kernel-name 'regex:matvec_kernel|attn_decode_kernel' \
  --launch-skip 200 --launch-count 6 -f -o gpurun_out/r02_decode $B > gpurun_out/r02_ncu_full.log 2>&1
ls -la gpurun_out/r02_*
