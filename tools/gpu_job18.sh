#!/bin/bash
# after the skip-one-pow guard: full GPU tests, config 1 (parity + fast), configs 2/3 with both queue orders
nvidia-smi -L
timeout 1800 python -m pytest tests -m gpu -x -q 2>&1 | tail -5
timeout 900 python bench.py > gpurun_out/r02_bench_c1_v4.json 2> gpurun_out/r02_bench_c1_v4.err; echo rc=$?; cut -c1-300 gpurun_out/r02_bench_c1_v4.json
for c in 2 3; do for q in forward reverse; do
timeout 900 python bench.py --config $c --queue $q --no-cpu-baseline > gpurun_out/r02_bench_c${c}_${q}.json 2> gpurun_out/r02_bench_c${c}_${q}.err; echo rc=$?; cut -c1-200 gpurun_out/r02_bench_c${c}_${q}.json
done; done
