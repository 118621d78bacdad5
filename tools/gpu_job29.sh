#!/bin/bash
nvidia-smi -L
run() { cfg=$1; n=$2; shift 2; echo "== $cfg N=$n $*"; env "$@" timeout 300 python bench.py --config $cfg --ntraj $n --no-cpu-baseline --no-e2e --steps 2 --warmup 3 2> gpurun_out/tn.err | python -c "
import sys, json
d = json.loads(sys.stdin.read().strip().splitlines()[-1])
print(d['value'], d['roofline']['frac'], d['ms_per_step'], d.get('parity_sample'))"; tail -2 gpurun_out/tn.err; }
run 5a 8192 X=1
run 5b 592 X=1
