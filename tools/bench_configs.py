#!/usr/bin/env python3
"""Throughput of BASELINE.json configs 2-5 through the public host-buffer API (one JSON line each).
Config 1 (the headline) is bench.py.  Timing: device time of the single fused launch (CUDA events inside the library, tb_last_kernel_ms) after one
warm-up call; the wall clock of the whole host-buffer call (H2D, launch, D2H from pageable memory) is reported beside it.  Flop models: SURVEY.md section 8(d).

    python tools/bench_configs.py [pendulum] [lorenz] [manyspin] [jet2] [jet4] [--scale S]
"""
import json
import math
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import taylorintegration_b200 as tb  # noqa: E402
from taylorintegration_b200 import problems as P  # noqa: E402

args = [a for a in sys.argv[1:] if not a.startswith("--")]
scale = 1.0
for a in sys.argv[1:]:
    if a.startswith("--scale"):
        scale = float(a.split("=")[1])
kflags = 0
for a in sys.argv[1:]:
    if a.startswith("--kflags"):
        kflags = int(a.split("=")[1])
KOPTS = {"flags": kflags} if kflags else None
which = args or ["pendulum", "lorenz", "manyspin", "jet2", "jet4"]
mode = "fast" if "--fast" in sys.argv else "parity"
peak_tf, _ = tb.fp64_peak(0.3)


WALL = []


def timed(fn):
    fn()  # warm-up (NVRTC compile, module load, allocations)
    t0 = time.perf_counter()
    r = fn()
    wall = time.perf_counter() - t0
    ms = tb._lib.lib().tb_last_kernel_ms()
    WALL.append(wall)
    return r, (ms * 1e-3 if ms > 0 else wall)


def report(name, workload, steps, secs, flops_per_step, ok, extra=None):
    ach = steps * flops_per_step / secs / 1e12
    line = {"config": name, "workload": workload, "mode": mode, "value": steps / secs, "unit": "trajectory-steps/s",
            "kernel_seconds": secs, "api_wall_seconds": WALL[-1], "trajectory_steps": int(steps), "flops_per_step": flops_per_step,
            "roofline": {"bound": "fp64", "achieved": ach, "peak": peak_tf, "unit": "TFLOP/s", "frac": ach / peak_tf},
            "all_ok": bool(ok)}
    if extra:
        line.update(extra)
    print(json.dumps(line), flush=True)


if "pendulum" in which:  # config 2
    N = int((1 << 20) * scale)
    tape = P.pendulum_tape(events=True)
    q0 = P.pendulum_ensemble(N)
    q0 = q0[0] if isinstance(q0, tuple) else q0
    sol, secs = timed(lambda: tb.taylorinteg_events(tape, q0, 0.0, 100.0, 25, 1e-20, maxsteps=2000, event_capacity=64,
                                                    mode=mode, output="final", kernel_opts=KOPTS))
    nev = int(sol.nevents.sum())
    report("2 pendulum+events", f"N={N}, order 25, abstol 1e-20, tmax 100, g = x[2], event capacity 64", sol.nsteps.sum(), secs,
           1685, (sol.status == 0).all(), {"events": nev})

if "lorenz" in which:  # config 3
    N = int((1 << 16) * scale)
    tape = P.lorenz_tape()
    rho = 28.0 + 64.0 * np.arange(N) / max(N - 1, 1)
    par = np.stack([np.full(N, 16.0), rho, np.full(N, 4.0)])
    q0 = np.tile(np.array([[19.0], [20.0], [50.0]]), (1, N))
    sol, secs = timed(lambda: tb.lyap_taylorinteg(tape, q0, 0.0, 100.0, 28, 1e-20, par, maxsteps=20000, mode=mode,
                                                  output="final", kernel_opts=KOPTS))
    report("3 lorenz lyapunov", f"N={N} rho in [28,92], sigma 16, beta 4, order 28, abstol 1e-20, tmax 100", sol.nsteps.sum(), secs,
           24700, (sol.status == 0).all(), {"sum_lambda_mean": float(sol.lam.sum(0).mean())})

if "manyspin" in which:  # config 4
    N = int((1 << 14) * scale)
    nn = 100
    tape = P.manyspin_tape(nn)
    q0 = P.manyspin_ensemble(nn, N)
    sol, secs = timed(lambda: tb.taylorinteg(tape, q0, 0.0, 10.0, 25, 1e-20, maxsteps=5000, mode=mode))
    norm_err = float(np.abs((sol.xfinal.reshape(nn, 3, -1) ** 2).sum(1) - 1).max())
    report("4 many-spin", f"N={N} members x {nn} spins (300 components), order 25, abstol 1e-20, tmax 10", sol.nsteps.sum(), secs,
           6.5e5, (sol.status == 0).all(), {"max_spin_norm_error": norm_err})

if "jet2" in which:  # config 5a
    N = int((1 << 16) * scale)
    tape = P.pendulum_tape(events=False)
    jt = tb.JetTransport(tape, 28, 2, 8, tb._lib.MODE_FAST if mode == "fast" else tb._lib.MODE_PARITY)
    th = np.random.Generator(np.random.PCG64(20261104)).uniform(0.1, 3.0, N)
    th[0] = 1.3
    q0 = jt.initial(np.stack([th, np.zeros(N)]))
    T = 7.019250311844546
    tr = np.arange(0, 49) * (T / 8)
    sol, secs = timed(lambda: jt.integrate(q0, abstol=1e-20, trange=tr, maxsteps=2000))
    pairs = math.comb(2 * 2 + 8, 8)
    fl = 2 * pairs * 2 * sum(range(1, 28))  # sincos!: 2 polynomial products per (k, i) pair
    report("5a jet transport pendulum n=2 d=8", f"N={N}, 45 coefficients, order 28, trange 0:T/8:6T", sol.nsteps.sum(), secs, fl,
           (sol.status == 0).all())

if "jet4" in which:  # config 5b
    N = int(2048 * scale)
    tape = P.kepler_tape()
    jt = tb.JetTransport(tape, 28, 4, 8, tb._lib.MODE_FAST if mode == "fast" else tb._lib.MODE_PARITY)
    q, mu, _ = P.kepler_ensemble(N)
    q0 = jt.initial(q)
    sol, secs = timed(lambda: jt.integrate(q0, 0.0, 2 * np.pi, 1e-20, mu, maxsteps=5000))
    pairs = math.comb(2 * 4 + 8, 8)
    prods = sum((k // 2 + 1) * 2 + k + 2 * k for k in range(28))  # 2 sqr!, pow!, 2 div! polynomial products per order
    report("5b jet transport kepler n=4 d=8", f"N={N}, 495 coefficients, order 28, one period", sol.nsteps.sum(), secs, 2 * pairs * prods,
           (sol.status == 0).all())
