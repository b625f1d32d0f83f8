timeout 1500 python bench.py --workload llama3-70b-q4_k_m-tp --steps 64 --warmup 8 > gpurun_out/r02_bench_70b_tp1.json 2> gpurun_out/tp1n1.err
tail -2 gpurun_out/tp1n1.err
