#!/bin/bash
# 2 GPUs: tb_integrate_multi test, sharded test, bench weak + strong (torchrun, NCCL)
nvidia-smi -L
timeout 600 python -m pytest tests/test_gpu_parity2.py::test_integrate_multi_matches_single_device tests/test_sharding.py -m gpu -x -q 2>&1 | tail -4
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 2 --warmup 3 --no-fast > gpurun_out/r02_bench_n2_weak.json 2> gpurun_out/r02_bench_n2_weak.err; echo rc=$?; cut -c1-330 gpurun_out/r02_bench_n2_weak.json
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --steps 2 --warmup 3 --no-fast --scaling strong > gpurun_out/r02_bench_n2_strong.json 2> gpurun_out/r02_bench_n2_strong.err; echo rc=$?; cut -c1-330 gpurun_out/r02_bench_n2_strong.json; tail -3 gpurun_out/r02_bench_n2_strong.err
python tools/multi_abi_bench.py 2 > gpurun_out/r02_multi_abi_n2.txt 2>&1; cat gpurun_out/r02_multi_abi_n2.txt
