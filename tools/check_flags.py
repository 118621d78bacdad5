#!/usr/bin/env python3
"""GPU check: a kernel built with extra TB_CF_* flags returns the bits of the default build (parity) / stays within 1e-9 (fast).
usage: python tools/check_flags.py FLAGS [REG_BUDGET]"""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import taylorintegration_b200 as tb
from taylorintegration_b200 import problems as P
flags = int(sys.argv[1]); rb = int(sys.argv[2]) if len(sys.argv) > 2 else 0
q0, mu, _ = P.kepler_ensemble(3000)
tm = 2 * np.pi * 3
for mode in ("parity", "fast"):
    a = tb.taylorinteg(P.kepler_tape(), q0, 0.0, tm, 28, 1e-20, mu, maxsteps=3000, mode=mode, kernel_opts={"full_outputs": True})
    b = tb.taylorinteg(P.kepler_tape(), q0, 0.0, tm, 28, 1e-20, mu, maxsteps=3000, mode=mode, kernel_opts={"flags": flags, "reg_budget": rb})
    same = np.array_equal(a.xfinal, b.xfinal) and np.array_equal(a.nsteps, b.nsteps)
    rel = float(np.max(np.abs(a.xfinal - b.xfinal) / np.abs(a.xfinal).max(0)))
    print(f"flags {flags} budget {rb} {mode}: identical={same} max rel diff {rel:.3e} steps equal={np.array_equal(a.nsteps, b.nsteps)}")
    assert same if mode == "parity" else rel < 1e-9
