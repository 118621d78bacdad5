#!/bin/bash
# Round-end measurement recipe (run under gpurun): bench line, reference arm, ncu launch list of the SAME command,
# one full capture of the dominant kernel on the bench workload.
set -x
python bench.py > gpurun_out/bench_final.json 2> gpurun_out/bench_final.err || exit 1
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_reference.json 2> gpurun_out/bench_reference.err
ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/launches_final.csv \
    python bench.py --no-cpu-baseline > gpurun_out/ncu_launch_final.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:tb_ensemble -s 3 -c 1 -f -o gpurun_out/prof_final \
    python bench.py --no-cpu-baseline --no-e2e --no-fast --steps 1 --warmup 3 > gpurun_out/ncu_full_final.log 2>&1
tail -2 gpurun_out/ncu_full_final.log
