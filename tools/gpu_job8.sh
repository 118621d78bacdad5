#!/bin/bash
# software-pipelining flags on the other generated kernels; ptxas -O levels on the Kepler kernel; new GPU tests
echo '[{"name": "256 thr pf8 ma4 (ptxas default)", "kw": {"flags": 1208221696}}]' > /tmp/c.json
for O in -O2 -O1; do TB_PTXAS_EXTRA=$O python tools/tmem_sweep.py 1048576 /tmp/c.json 0 2>&1 | sed "s/^/ptxas $O: /"; done
python tools/bench_configs.py pendulum lorenz 2>&1 | cut -c1-330
python tools/bench_configs.py pendulum lorenz --kflags=1207959552 2>&1 | cut -c1-330
python tools/bench_configs.py pendulum lorenz --kflags=603979776 2>&1 | cut -c1-330
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -15
