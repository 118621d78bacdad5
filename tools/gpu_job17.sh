#!/bin/bash
# 8 GPUs: strong scaling of ONE 2^20-trajectory Kepler ensemble (torchrun/NCCL) and through tb_integrate_multi (one process)
nvidia-smi -L | wc -l
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29521 bench.py --gpus 8 --steps 2 --warmup 3 --no-fast --scaling strong > gpurun_out/r02_bench_n8_strong.json 2> gpurun_out/r02_bench_n8_strong.err; echo rc=$?; cut -c1-330 gpurun_out/r02_bench_n8_strong.json; tail -2 gpurun_out/r02_bench_n8_strong.err
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29522 bench.py --gpus 8 --steps 2 --warmup 3 --no-fast --no-cpu-baseline > gpurun_out/r02_bench_n8_weak.json 2> gpurun_out/r02_bench_n8_weak.err; echo rc=$?; cut -c1-330 gpurun_out/r02_bench_n8_weak.json
python tools/multi_abi_bench.py 8 2>&1 | grep devices > gpurun_out/r02_multi_abi_n8.txt; cat gpurun_out/r02_multi_abi_n8.txt
