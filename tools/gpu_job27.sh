#!/bin/bash
nvidia-smi -L
timeout 400 ncu --set full --clock-control none --import-source on -k regex:tn_kernel -c 1 -f -o gpurun_out/r02_prof_tn2_smem \
    python bench.py --config 5a --ntraj 444 --no-cpu-baseline --no-e2e --steps 1 --warmup 0 > gpurun_out/r02_ncu_tn2_smem.log 2>&1; echo rc=$?
tail -3 gpurun_out/r02_ncu_tn2_smem.log | cut -c1-300
