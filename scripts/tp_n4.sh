timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29551 bench.py --gpus 4 --workload llama3-70b-q4_k_m-tp --steps 64 --warmup 8 --no-cpu-baseline > gpurun_out/r02_bench_70b_tp4.json 2> gpurun_out/tp4.err
tail -2 gpurun_out/tp4.err
