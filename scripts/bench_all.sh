timeout 600 python bench.py > gpurun_out/r02_bench_default.json 2> gpurun_out/r02_bench_default.err
timeout 600 python bench.py --workload llama3-8b-q8_0-bs32 > gpurun_out/r02_bench_q8_0_bs32.json 2> gpurun_out/r02_bench_q8_0_bs32.err
timeout 600 python bench.py --workload tinyllama-1.1b-q4_0-bs1 > gpurun_out/r02_bench_tinyllama_q4_0.json 2> gpurun_out/r02_bench_tinyllama.err
timeout 900 python bench.py --workload mistral-7b-q4_k_m-8k > gpurun_out/r02_bench_mistral_8k.json 2> gpurun_out/r02_bench_mistral.err
timeout 1500 python bench.py --workload llama3-70b-q4_k_m-tp --steps 64 --warmup 8 > gpurun_out/r02_bench_70b_tp1.json 2> gpurun_out/tp1n1.err
