#!/bin/bash
# TaylorN kernel: paired entries (x^2 | y^2, x/r^3 | y/r^3) and the fused sincos! staging
nvidia-smi -L
timeout 600 python -m pytest tests/test_tn.py tests/test_zz_no_alloc.py -m gpu -x -q 2>&1 | tail -2
echo "== 5b N=592 paired"; timeout 300 python bench.py --config 5b --ntraj 592 --no-cpu-baseline --no-e2e --steps 2 --warmup 3 2>/dev/null | cut -c1-200
timeout 600 python bench.py --config 5a > gpurun_out/r02_bench_c5a_v2.json 2> gpurun_out/r02_bench_c5a_v2.err; echo rc=$?; cut -c1-200 gpurun_out/r02_bench_c5a_v2.json
