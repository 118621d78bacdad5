#!/bin/bash
# 8 GPUs: config 5b (Kepler jet transport, 4 variables of order 8) sharded strong over the ranks, 2048 trajectories per GPU
nvidia-smi -L | wc -l
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29513 bench.py --config 5b --ntraj 16384 --gpus 8 --steps 1 --warmup 1 --no-cpu-baseline --scaling strong > gpurun_out/r02_bench_c5b_n8_strong.json 2> gpurun_out/r02_bench_c5b_n8_strong.err; echo rc=$?; tail -1 gpurun_out/r02_bench_c5b_n8_strong.json | cut -c1-330; tail -3 gpurun_out/r02_bench_c5b_n8_strong.err
