#!/usr/bin/env python3
"""Lorenz Lyapunov (config 3) kernel time for several TB_CF_* flag sets.  usage: python tools/lorenz_flags.py 0 1024 ..."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import taylorintegration_b200 as tb
from taylorintegration_b200 import problems as P
N = 1 << 16
rho = 28.0 + 64.0 * np.arange(N) / (N - 1)
par = np.stack([np.full(N, 16.0), rho, np.full(N, 4.0)])
q0 = np.tile(np.array([[19.0], [20.0], [50.0]]), (1, N))
ref = None
for fl in [int(x) for x in sys.argv[1:]] or [0]:
    for _ in range(2):
        sol = tb.lyap_taylorinteg(P.lorenz_tape(), q0, 0.0, 100.0, 28, 1e-20, par, maxsteps=20000, output="final", kernel_opts={"flags": fl})
    ms = tb._lib.lib().tb_last_kernel_ms()
    ref = ref or sol
    print(f"flags {fl}: {ms:.1f} ms, {sol.nsteps.sum() / ms * 1e3:.4g} steps/s, identical: {np.array_equal(ref.lam, sol.lam)}", flush=True)
