#!/usr/bin/env python3
"""Precompile (on the CPU box) the kernels tools/tmem_sweep.py times, so the GPU box finds the cubins in kcache/."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import taylorintegration_b200 as tb
from taylorintegration_b200 import _lib as L, problems as P
from concurrent.futures import ThreadPoolExecutor
tape = P.kepler_tape()
cfgs = [dict(flags=262144 | 524288, reg_budget=4 * 29), dict(flags=262144 | 524288, reg_budget=3 * 29),
        dict(flags=262144 | 524288, reg_budget=2 * 29), dict(flags=262144 | 524288, reg_budget=3 * 29, tmem_budget=2 * 29),
        dict(flags=262144 | 524288, reg_budget=3 * 29, tmem_budget=29)]
def one(a):
    mode, kw = a
    k = tb.Kernel(tape, 28, L.KIND_PLAIN, mode, **kw)
    return mode, kw, k.info()
with ThreadPoolExecutor(6) as ex:
    for r in ex.map(one, [(m, kw) for m in (L.MODE_PARITY, L.MODE_FAST) for kw in cfgs]):
        print(r, flush=True)
