#!/usr/bin/env python3
"""tb_integrate_multi (single process, one host thread per GPU) on the Kepler headline ensemble: wall time of the
host-buffer call for 1..G devices (strong scaling of ONE 2^20-trajectory ensemble).  usage: multi_abi_bench.py G"""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import taylorintegration_b200 as tb
from taylorintegration_b200 import problems as P
G = int(sys.argv[1]) if len(sys.argv) > 1 else 2
N = 1 << 20
q0, mu, _ = P.kepler_ensemble(N)
tmax = 2 * np.pi * 100
base = None
for g in sorted({1, 2, G}):
    devs = list(range(g))
    for it in range(2):
        t0 = time.perf_counter()
        s = tb.taylorinteg(P.kepler_tape(), q0, 0.0, tmax, 28, 1e-20, mu, maxsteps=10_000, devices=devs, kernel_opts={"sort_by_cost": 1})
        wall = time.perf_counter() - t0
    steps = int(s.nsteps.sum())
    if base is None:
        base = (s.xfinal.copy(), wall)
    same = bool(np.array_equal(s.xfinal, base[0]))
    print(f"devices {devs}: wall {wall:.3f} s, {steps / wall:.4g} steps/s (pageable host buffers), speed-up {base[1] / wall:.2f}, "
          f"slowest kernel {tb._lib.lib().tb_last_kernel_ms():.1f} ms, same bits as 1 device: {same}", flush=True)
