#!/bin/bash
# 8 GPUs: strong scaling of ONE 2^20 ensemble with the reversed queue (torchrun/NCCL) + tb_integrate_multi
nvidia-smi -L | wc -l
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 8 --steps 2 --warmup 3 --no-fast --no-cpu-baseline --scaling strong > gpurun_out/r02_bench_n8_strong_v2.json 2> gpurun_out/r02_bench_n8_strong_v2.err; echo rc=$?; cut -c1-330 gpurun_out/r02_bench_n8_strong_v2.json; tail -3 gpurun_out/r02_bench_n8_strong_v2.err
python tools/multi_abi_bench.py 8 > gpurun_out/r02_multi_abi_n8_v2.txt 2>&1; cat gpurun_out/r02_multi_abi_n8_v2.txt
