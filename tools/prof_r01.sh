#!/bin/bash
# ncu recipe of round 1 (run under gpurun): launch list of the default bench + one full capture of the fused kernel
set -x
ARGS="--ntraj 56832 --periods 20 --steps 1 --warmup 1 --no-cpu-baseline --no-e2e"
MODE=${1:-parity}
python bench.py $ARGS --mode $MODE > gpurun_out/plain_$MODE.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 50 --csv --log-file gpurun_out/launches_$MODE.csv python bench.py $ARGS --mode $MODE > gpurun_out/ncu_launch_$MODE.log 2>&1
python bench.py $ARGS --mode $MODE > gpurun_out/plain2_$MODE.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:tb_ensemble -c 1 -f -o gpurun_out/prof_$MODE python bench.py $ARGS --mode $MODE > gpurun_out/ncu_full_$MODE.log 2>&1
tail -3 gpurun_out/plain_$MODE.log gpurun_out/ncu_full_$MODE.log
