#!/bin/bash
# verify skip-one-pow stepsize, pow reciprocal finalize and the reversed queue: parity, sweep, bench, full GPU tests
nvidia-smi -L
timeout 900 python tools/tmem_sweep.py 1048576 tools/sweep_cfg_f.json 0 2>&1 | tail -5
timeout 900 python bench.py --no-fast > gpurun_out/r02_bench_c1_v3.json 2> gpurun_out/r02_bench_c1_v3.err; echo rc=$?; cut -c1-600 gpurun_out/r02_bench_c1_v3.json
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -5
