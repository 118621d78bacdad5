#!/bin/bash
# cluster kernel: out-of-line division fallback in solcoeff!, step size from the largest coefficients (reltol = 0)
nvidia-smi -L
timeout 600 python -m pytest tests/test_wide.py -m gpu -x -q 2>&1 | tail -2
TB_WIDE_INTERP=1 timeout 600 python -m pytest tests/test_wide.py -m gpu -x -q 2>&1 | tail -2
timeout 600 python bench.py --config 4 > gpurun_out/r02_bench_c4_v3.json 2> gpurun_out/r02_bench_c4_v3.err; echo rc=$?; cut -c1-200 gpurun_out/r02_bench_c4_v3.json
