#!/bin/bash
# balanced TaylorN kernel (tb_tn2.cuh) against the thread-per-slot kernel: parity, then throughput at small sizes
nvidia-smi -L
timeout 600 python -m pytest tests/test_tn.py tests/test_zz_no_alloc.py tests/test_lyap_userjac.py -m gpu -x -q 2>&1 | tail -3
TB_TN_V1=1 timeout 600 python -m pytest tests/test_tn.py -m gpu -x -q 2>&1 | tail -2
run() { cfg=$1; n=$2; shift 2; echo "== $cfg N=$n $*"; env "$@" timeout 300 python bench.py --config $cfg --ntraj $n --no-cpu-baseline --no-e2e --steps 2 --warmup 3 2> gpurun_out/tn.err | python -c "
import sys, json
d = json.loads(sys.stdin.read().strip().splitlines()[-1])
print(d['value'], d['roofline']['frac'], d['ms_per_step'])"; tail -2 gpurun_out/tn.err; }
run 5a 8192 TB_TN_V1=1
run 5a 8192 X=1
run 5b 592 TB_TN_V1=1
run 5b 592 X=1
