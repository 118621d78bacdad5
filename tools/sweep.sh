#!/bin/bash
# usage: tools/sweep.sh "<mode> <kflags> <reg_budget> <block_threads>" ...   (run under gpurun)
for cfg in "$@"; do
  set -- $cfg
  timeout 300 python bench.py --mode $1 --kflags $2 --reg-budget $3 --block-threads $4 --steps 2 --warmup 3 --no-cpu-baseline --no-e2e 2>&1 | tail -1 | python -c "
import sys,json
try:
    d=json.loads(sys.stdin.read()); print('$cfg', '->', '%.4g steps/s' % d['value'], 'frac %.3f' % d['roofline']['frac'], d['config']['kernel'], flush=True)
except Exception as e: print('$cfg', 'FAILED', e, flush=True)"
done
