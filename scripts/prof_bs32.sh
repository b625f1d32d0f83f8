B="python bench.py --workload llama3-8b-q8_0-bs32 --steps 2 --warmup 1 --ncu-mode --no-cpu-baseline"
timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none --launch-skip 9700 -c 1100 --csv \
  --log-file gpurun_out/r02_bs32_launches.csv $B > gpurun_out/r02_bs32_ncu_list.log 2>&1
