#!/bin/bash
# TaylorN kernel with the jets in shared memory (n = 2): parity tests for both placements, bench 5a
nvidia-smi -L
timeout 900 python -m pytest tests/test_tn.py tests/test_zz_no_alloc.py -m gpu -x -q 2>&1 | tail -3
TB_TN_GLOBAL_JETS=1 timeout 900 python -m pytest tests/test_tn.py tests/test_zz_no_alloc.py -m gpu -x -q 2>&1 | tail -3
run() { echo "== $*"; env "$@" timeout 900 python bench.py --config 5a --ntraj 16384 --no-cpu-baseline --steps 2 --warmup 3 2> gpurun_out/tn.err | python -c "
import sys, json
d = json.loads(sys.stdin.read().strip().splitlines()[-1])
print(d['value'], d['roofline']['frac'], d['ms_per_step'])"; tail -2 gpurun_out/tn.err; }
run TB_TN_GLOBAL_JETS=1
run X=1
timeout 900 python bench.py --config 5a > gpurun_out/r02_bench_c5a_smem.json 2> gpurun_out/r02_bench_c5a_smem.err; echo rc=$?; cut -c1-250 gpurun_out/r02_bench_c5a_smem.json
timeout 600 ncu --set full --clock-control none --import-source on -k regex:tn_kernel -c 1 -f -o gpurun_out/r02_prof_tn2_smem \
    python bench.py --config 5a --ntraj 2048 --no-cpu-baseline --steps 1 --warmup 1 > gpurun_out/r02_ncu_tn2_smem.log 2>&1; echo rc=$?
