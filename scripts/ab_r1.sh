cd _r1 && for rep in 1 2; do python bench.py --steps 30 --warmup 5 --no-cpu-baseline --batch-extra 0 2>&1 | grep -o '"ms_per_step": [0-9.]*'; done
python bench.py --steps 10 --warmup 3 --no-cpu-baseline --batch-extra 0 --profile-kernels 2>&1 | grep -v '^{' | tail -15
cd .. && for rep in 1 2; do B200_GGUF_LIB=llama-fastapi-k8s-gpu_b200/lib_head/libb200gguf.so python bench.py --steps 30 --warmup 5 --no-cpu-baseline --batch-extra 0 --no-parity-check 2>&1 | grep -o '"ms_per_step": [0-9.]*'; done
cd _r1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 700 --csv --log-file ../gpurun_out/r1_launches.csv python bench.py --steps 1 --warmup 1 --ncu-mode --no-cpu-baseline --batch-extra 0 > /dev/null 2>&1
