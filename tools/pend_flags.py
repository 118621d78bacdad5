#!/usr/bin/env python3
"""Pendulum + events (config 2) kernel time for several TB_CF_* flag sets.  usage: python tools/pend_flags.py 0 1024 ..."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import taylorintegration_b200 as tb
from taylorintegration_b200 import problems as P
N = 1 << 20
tape = P.pendulum_tape(events=True)
q0 = P.pendulum_ensemble(N)
q0 = q0[0] if isinstance(q0, tuple) else q0
ref = None
for fl in [int(x) for x in sys.argv[1:]] or [0]:
    for _ in range(2):
        sol = tb.taylorinteg_events(tape, q0, 0.0, 100.0, 25, 1e-20, maxsteps=2000, event_capacity=64, output="final",
                                    kernel_opts={"flags": fl})
    ms = tb._lib.lib().tb_last_kernel_ms()
    if ref is None:
        ref = sol
    same = np.array_equal(ref.xfinal, sol.xfinal) and np.array_equal(ref.tevents, sol.tevents)
    print(f"flags {fl}: {ms:.1f} ms, {sol.nsteps.sum() / ms * 1e3:.4g} steps/s, identical to flags[0]: {same}", flush=True)
