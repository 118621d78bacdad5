#!/bin/bash
# ncu of the shared-memory cluster kernel (many-spin 100, 1024 members) + GPU tests of the new primitives
timeout 600 ncu --set full --clock-control none --import-source on -k regex:wide2 -c 1 -f -o gpurun_out/r02_prof_wide2 \
    python tools/bench_configs.py manyspin --scale=0.0625 > gpurun_out/r02_ncu_wide2.log 2>&1
tail -2 gpurun_out/r02_ncu_wide2.log
timeout 600 python -m pytest tests/test_primitives.py -m gpu -x -q 2>&1 | tail -4
