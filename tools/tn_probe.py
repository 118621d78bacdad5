#!/usr/bin/env python3
"""Short jet-transport run for profiling: kepler1! in 4 variables of order 8 (config 5b shape), N nominal orbits, tmax."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import taylorintegration_b200 as tb
from taylorintegration_b200 import problems as P
N = int(sys.argv[1]) if len(sys.argv) > 1 else 148
tmax = float(sys.argv[2]) if len(sys.argv) > 2 else 0.5
nv, d = (int(sys.argv[3]), int(sys.argv[4])) if len(sys.argv) > 4 else (4, 8)
tape = P.kepler_tape() if nv == 4 else P.pendulum_tape(events=False)
jt = tb.JetTransport(tape, 28, nv, d)
if nv == 4:
    q, mu, _ = P.kepler_ensemble(N)
else:
    q, mu = np.stack([np.linspace(0.1, 3.0, N), np.zeros(N)]), None
q0 = jt.initial(q)
for _ in range(2):
    t0 = time.perf_counter()
    sol = jt.integrate(q0, 0.0, tmax, 1e-20, mu, maxsteps=5000)
    wall = time.perf_counter() - t0
    ms = tb._lib.lib().tb_last_kernel_ms()
print(f"N {N} steps {int(sol.nsteps.sum())} kernel {ms:.1f} ms  {sol.nsteps.sum() / ms * 1e3:.4g} steps/s  ok {(sol.status == 0).all()}")
