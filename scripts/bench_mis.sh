timeout 1200 python bench.py --workload mistral-7b-q4_k_m-8k > gpurun_out/r02_bench_mistral_8k.json 2> gpurun_out/r02_bench_mistral.err
tail -2 gpurun_out/r02_bench_mistral.err
