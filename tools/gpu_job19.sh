#!/bin/bash
# evidence for the shipped kernels: launch list of the default bench command, full capture + DRAM traffic of the
# Kepler parity kernel, full captures of the cluster (many-spin) and TaylorN (n=4, d=8) kernels, bench lines 4/5a/5b
nvidia-smi -L
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02_launches_bench_default.csv \
    python bench.py > gpurun_out/r02_ncu_launches.log 2>&1; echo rc=$?; tail -4 gpurun_out/r02_launches_bench_default.csv | cut -c1-200
timeout 600 ncu --set full --clock-control none --import-source on -k regex:tb_ensemble -s 3 -c 1 -f -o gpurun_out/r02_prof_parity_final \
    python bench.py --no-cpu-baseline --no-e2e --no-fast --steps 1 --warmup 3 --ntraj 131072 > gpurun_out/r02_ncu_full_final.log 2>&1; echo rc=$?
timeout 300 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none -k regex:tb_ensemble -s 3 -c 1 --csv --log-file gpurun_out/r02_traffic_parity_final.csv \
    python bench.py --no-cpu-baseline --no-e2e --no-fast --steps 1 --warmup 3 > gpurun_out/r02_ncu_traffic_final.log 2>&1; echo rc=$?
tail -3 gpurun_out/r02_traffic_parity_final.csv
timeout 600 ncu --set full --clock-control none --import-source on -k regex:wide2 -c 1 -f -o gpurun_out/r02_prof_wide2_final \
    python bench.py --config 4 --ntraj 1024 --no-cpu-baseline --steps 1 --warmup 1 > gpurun_out/r02_ncu_wide2_final.log 2>&1; echo rc=$?
timeout 600 ncu --set full --clock-control none --import-source on -k regex:tn_kernel -c 1 -f -o gpurun_out/r02_prof_tn4_final \
    python bench.py --config 5b --ntraj 512 --no-cpu-baseline --steps 1 --warmup 1 > gpurun_out/r02_ncu_tn4_final.log 2>&1; echo rc=$?
for c in 4 5a 5b; do
timeout 900 python bench.py --config $c > gpurun_out/r02_bench_c${c}.json 2> gpurun_out/r02_bench_c${c}.err; echo rc=$?; cut -c1-250 gpurun_out/r02_bench_c${c}.json
done
