#!/bin/bash
nvidia-smi -L
run() { echo "== $*"; env "$@" timeout 300 python bench.py --config 4 --no-cpu-baseline --no-e2e --steps 2 --warmup 3 2>/dev/null | python -c "
import sys, json
d = json.loads(sys.stdin.read().strip().splitlines()[-1])
print(d['value'], d['roofline']['frac'], d['config']['kernel']['regs'])"; }
run TB_WIDE_LCREG=66
run TB_WIDE_LCB=8
run TB_WIDE_LCB=24
