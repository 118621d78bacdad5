#!/bin/bash
timeout 900 python -m pytest tests/test_tn.py -m gpu -x -q 2>&1 | tail -5
python bench.py --config 5a --ntraj 8192 --steps 1 --warmup 1 --cpu-seconds 8 > gpurun_out/r02_bench_c5a.json 2> gpurun_out/r02_bench_c5a.err; echo rc=$?; cut -c1-260 gpurun_out/r02_bench_c5a.json; tail -2 gpurun_out/r02_bench_c5a.err
python bench.py --config 5b --ntraj 256 --steps 1 --warmup 1 --cpu-seconds 8 > gpurun_out/r02_bench_c5b.json 2> gpurun_out/r02_bench_c5b.err; echo rc=$?; cut -c1-260 gpurun_out/r02_bench_c5b.json; tail -2 gpurun_out/r02_bench_c5b.err
python bench.py --config 4 > gpurun_out/r02_bench_c4.json 2> gpurun_out/r02_bench_c4.err; echo rc=$?; cut -c1-260 gpurun_out/r02_bench_c4.json; tail -2 gpurun_out/r02_bench_c4.err
python tools/tmem_sweep.py 1048576 tools/sweep_cfg_e.json 0 2>&1 | tail -4
