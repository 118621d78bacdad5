#!/bin/bash
# cluster kernel specialised by NVRTC: parity tests for the three builds, variants, bench line, ncu capture
nvidia-smi -L
timeout 900 python -m pytest tests/test_wide.py -m gpu -x -q 2>&1 | tail -2
TB_WIDE_LC_SMEM=1 timeout 900 python -m pytest tests/test_wide.py -m gpu -x -q 2>&1 | tail -2
TB_WIDE_INTERP=1 timeout 900 python -m pytest tests/test_wide.py -m gpu -x -q 2>&1 | tail -2
run() { echo "== $*"; env "$@" timeout 600 python bench.py --config 4 --no-cpu-baseline --steps 2 --warmup 3 2> gpurun_out/w2.err | python -c "
import sys, json
d = json.loads(sys.stdin.read().strip().splitlines()[-1])
k = d['config'].get('kernel')
print(d['value'], d['roofline']['frac'], {x: k[x] for x in ('smem_bytes', 'ctas_per_sm', 'generated', 'regs', 'local_bytes', 'coupling_in_smem')})"; tail -2 gpurun_out/w2.err; }
run TB_WIDE_INTERP=1
run X=1
run TB_WIDE_LCREG=0
run TB_WIDE_LCREG=44
run TB_WIDE_LC_SMEM=1
timeout 900 python bench.py --config 4 > gpurun_out/r02_bench_c4_gen.json 2> gpurun_out/r02_bench_c4_gen.err; echo rc=$?; cut -c1-250 gpurun_out/r02_bench_c4_gen.json
timeout 600 ncu --set full --clock-control none --import-source on -k regex:tb_w2 -c 1 -f -o gpurun_out/r02_prof_w2gen \
    python bench.py --config 4 --ntraj 1184 --no-cpu-baseline --steps 1 --warmup 1 > gpurun_out/r02_ncu_w2gen.log 2>&1; echo rc=$?
