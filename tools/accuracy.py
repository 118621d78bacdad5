#!/usr/bin/env python3
"""Fast mode vs parity mode (== oracle bitwise) on the Kepler ensemble: relative final-state difference
and relative energy error after `periods` periods.  Run on a GPU box."""
import sys, os, json
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import taylorintegration_b200 as tb
from taylorintegration_b200 import _lib as L, problems as P
N = int(sys.argv[1]) if len(sys.argv) > 1 else 2048
out = {}
for periods in (100.0, 1000.0):
    tape = P.kepler_tape()
    q0, mu, e = P.kepler_ensemble(N)
    tmax = 2 * np.pi * periods
    sols = {}
    for mode in ("parity", "fast"):
        sols[mode] = tb.taylorinteg(tape, q0, 0.0, tmax, 28, 1e-20, mu, maxsteps=100000, mode=mode)
    def energy(x):
        return 0.5 * (x[2] ** 2 + x[3] ** 2) - 1.0 / np.sqrt(x[0] ** 2 + x[1] ** 2)
    E0 = energy(q0)
    a, b = sols["parity"].xfinal, sols["fast"].xfinal
    scale = np.sqrt((a ** 2).sum(0))
    rel = np.abs(a - b).max(0) / np.maximum(np.abs(a).max(0), 1e-300)
    relq0 = np.abs(a - q0).max(0) / np.abs(q0).max(0)
    res = {"periods": periods, "N": N}
    for lo, hi in ((0, 0.3), (0.3, 0.6), (0.6, 0.8), (0.8, 0.9001)):
        m = (e >= lo) & (e < hi)
        res[f"e[{lo},{hi})"] = {"max_rel_fast_vs_parity": float(rel[m].max()), "median": float(np.median(rel[m])),
                               "max_rel_return_parity": float(relq0[m].max()),
                               "max_dE_parity": float(np.abs((energy(a) - E0) / E0)[m].max()),
                               "max_dE_fast": float(np.abs((energy(b) - E0) / E0)[m].max())}
    res["steps_equal_frac"] = float((sols["parity"].nsteps == sols["fast"].nsteps).mean())
    print(json.dumps(res), flush=True)
