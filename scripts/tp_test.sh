timeout 600 python -m pytest tests/test_gpu_tp.py -m gpu -x -q 2>&1 | tail -2
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29542 bench.py --gpus 2 --workload llama3-70b-q4_k_m-tp --steps 64 --warmup 8 --no-cpu-baseline > gpurun_out/r02_bench_70b_tp2.json 2> gpurun_out/tp2.err
tail -2 gpurun_out/tp2.err
