#!/bin/bash
# round-2 first GPU job: TMEM microbenchmark, GPU tests, ncu re-capture of the shipped parity kernel
set -x
./tools/tmem_bench > gpurun_out/r02_tmem_bench.txt 2>&1
tail -50 gpurun_out/r02_tmem_bench.txt
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r02_gputests_start.log 2>&1; tail -3 gpurun_out/r02_gputests_start.log
timeout 600 ncu --set full --clock-control none --import-source on -k regex:tb_ensemble -s 3 -c 1 -f -o gpurun_out/r02_prof_parity_shipped \
    python bench.py --no-cpu-baseline --no-e2e --no-fast --steps 1 --warmup 3 --ntraj 131072 > gpurun_out/r02_ncu_full.log 2>&1
tail -2 gpurun_out/r02_ncu_full.log
timeout 300 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none -k regex:tb_ensemble -s 3 -c 1 --csv --log-file gpurun_out/r02_traffic_parity.csv \
    python bench.py --no-cpu-baseline --no-e2e --no-fast --steps 1 --warmup 3 > gpurun_out/r02_ncu_traffic.log 2>&1
tail -3 gpurun_out/r02_traffic_parity.csv
