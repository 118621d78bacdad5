#!/bin/bash
timeout 600 ncu --set full --clock-control none --import-source on -k regex:wide2 -c 1 -f -o gpurun_out/r02_prof_wide2b \
    python tools/bench_configs.py manyspin --scale=0.0625 > gpurun_out/r02_ncu_wide2.log 2>&1
tail -2 gpurun_out/r02_ncu_wide2.log
