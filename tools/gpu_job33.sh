#!/bin/bash
# full ncu capture of the shipped TaylorN kernel on config 5a at full occupancy (two waves of 16 CTAs per SM)
nvidia-smi -L
timeout 500 ncu --set full --clock-control none --import-source on -k regex:tn_kernel -c 1 -f -o gpurun_out/r02_prof_tn2_shipped \
    python bench.py --config 5a --ntraj 4736 --no-cpu-baseline --no-e2e --steps 1 --warmup 0 > gpurun_out/r02_ncu_tn2_shipped.log 2>&1; echo rc=$?
