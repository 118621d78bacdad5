#!/usr/bin/env python3
"""Generates the committed fixtures of tests/golden/ (run in the build container, where /root/reference exists).

1. reference_kats.json -- the known answers the reference's OWN tests hold for the hot path.  The reference is pure
   Julia and cannot run here, so nothing is executed: every literal is looked up (regex) on the cited line of the
   reference test file and the script FAILS if a literal is no longer there.  Consumers: tests/test_golden.py
   (oracle pinned against them, CPU) and the -m gpu tests (CUDA path against them).
2. oracle_vectors.json -- outputs of the pinned CPU oracle (oracle/tb_oracle.cpp, TB math) on small seeded cases:
   first-step jets, step sizes, final states, step counts, event times, Lyapunov spectra.  The GPU box has no
   /root/reference; the -m gpu tests compare the CUDA path bit-for-bit with these vectors (hex floats), and the CPU
   suite re-derives them to catch oracle drift.

    python tests/golden/make_golden.py          # rewrites both files
"""
import json
import os
import re
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)
REF = "/root/reference"

# (key, file, line, regex on that line, value extractor)
KATS = [
    ("xsq_last_coeff_is_3_pow_order_plus_1", "test/one_ode.jl", 203, r"x0T\.coeffs\[end\] == 3\.0\^\(_order \+ 1\)", None),
    ("xsq_many_last_coeff", "test/many_ode.jl", 35, r"q0T\[1\]\.coeffs\[end\] == 3\.0\^\(_order \+ 1\)", None),
    ("xsq_psol_powers_of_three", "test/many_ode.jl", 164, r"psol\[1, 1\] == Taylor1\(\[3\.0\^i for i = 1:_order\+1\]\)", None),
    ("xsq_len_tv_tmax_0.3", "test/one_ode.jl", 219, r"length\(tv\) == (\d+)", int),
    ("xsq_len_tv_tmax_0.33", "test/one_ode.jl", 239, r"length\(tv\) == (\d+)", int),
    ("xsq_many_len_tv_tmax_0.3", "test/many_ode.jl", 193, r"length\(tv\) == (\d+)", int),
    ("xsq_many_len_tv_tmax_0.33", "test/many_ode.jl", 217, r"length\(tv\) == (\d+)", int),
    ("pendulum_period_theta_1.3", "test/rootfinding.jl", 30, r"Tend = ([0-9.]+)", float),
    ("lorenz_lambda1_t20", "test/lyapunov.jl", 208, r"λv\[end, 1\], ([-0-9.]+)", float),
    ("lorenz_lambda2_t20", "test/lyapunov.jl", 209, r"λv\[end, 2\], ([-0-9.]+)", float),
    ("lorenz_lambda3_t20", "test/lyapunov.jl", 210, r"λv\[end, 3\], ([-0-9.]+)", float),
]


def reference_kats():
    out = {"_comment": "literals held by the reference's own tests; extracted, not executed (no julia in the image)",
           "_reference": "PerezHz/TaylorIntegration.jl v0.18.14 (Project.toml), TaylorSeries compat 0.22.1"}
    for key, rel, line, rx, conv in KATS:
        with open(os.path.join(REF, rel), encoding="utf-8") as f:
            text = f.readlines()[line - 1]
        m = re.search(rx, text)
        if not m:
            raise SystemExit(f"{rel}:{line} no longer matches {rx!r}: {text.strip()!r}")
        out[key] = {"cite": f"{rel}:{line}", "source": text.strip(), "value": conv(m.group(1)) if conv else True}
    return out


def hexarr(a):
    return [float(x).hex() for x in np.asarray(a, dtype=np.float64).reshape(-1)]


def oracle_vectors():
    import oracle
    from taylorintegration_b200 import _lib as L, problems as P
    v = {"_comment": "outputs of oracle/tb_oracle.cpp (TB math, -O2 -ffp-contract=off); hex floats, row-major"}
    # Kepler order 28: first-step jets + step sizes for 5 eccentricities, and final states after 3 periods
    tape = P.kepler_tape()
    q0, mu, e = P.kepler_ensemble(4096)
    idx = np.array([0, 1000, 2048, 3640, 4095])
    q = np.ascontiguousarray(q0[:, idx])
    jets, dts = [], []
    for j in range(idx.size):
        xj, _, dt = oracle.jet(tape, 28, q[:, j], mu, abstol=1e-20)
        jets.append(xj)
        dts.append(dt)
    s = oracle.integrate(tape, 28, q, 0.0, 2 * np.pi * 3, mu, maxsteps=2000)
    v["kepler28"] = {"ensemble": "problems.kepler_ensemble(4096)", "indices": idx.tolist(), "q0": hexarr(q),
                     "jets_shape": [idx.size, 4, 29], "jets": hexarr(np.stack(jets)), "dt": hexarr(dts),
                     "tmax_periods": 3, "xfinal_shape": [4, idx.size], "xfinal": hexarr(s.xfinal),
                     "nsteps": s.nsteps.tolist()}
    # pendulum events, theta = 1.3, order 28 to 3T (test/rootfinding.jl:106-122)
    T = 7.019250311844546
    pt = P.pendulum_tape()
    s = oracle.integrate(pt, 28, [1.3, 0.0], 0.0, 3 * T, kind=L.KIND_EVENTS, abstol=1e-20, maxsteps=1000)
    ne = int(s.nevents[0])
    v["pendulum_events28"] = {"q0": [1.3, 0.0], "tmax": (3 * T).hex(), "nevents": ne, "nsteps": int(s.nsteps[0]),
                              "tevents": hexarr(s.tevents[:ne, 0]), "xevents": hexarr(s.xevents[:, :ne, 0]),
                              "gresids": hexarr(s.gresids[:ne, 0]), "xfinal": hexarr(s.xfinal[:, 0])}
    # Lorenz Lyapunov, rho = 45.92, t = 20 (test/lyapunov.jl:191-215)
    lt = P.lorenz_tape()
    s = oracle.integrate(lt, 28, [19.0, 20.0, 50.0], 0.0, 20.0, [16.0, 45.92, 4.0], kind=L.KIND_LYAP, maxsteps=2000)
    v["lorenz_lyap28"] = {"q0": [19.0, 20.0, 50.0], "params": [16.0, 45.92, 4.0], "tmax": 20.0,
                          "nsteps": int(s.nsteps[0]), "lam": hexarr(s.lam[:, 0]), "xfinal": hexarr(s.xfinal[:, 0])}
    # xdot = x^2 from 3.0, order 28: dense polynomial of the first step and trange output
    import taylorintegration_b200 as tb
    xsq = tb.taylorize(lambda dx, x, p, t: dx.__setitem__(0, x[0] ** 2), 1, 0)
    xj, _, dt = oracle.jet(xsq, 28, [3.0], abstol=1e-20)
    s = oracle.integrate(xsq, 28, [3.0], 0.0, 0.33, abstol=1e-20, steps=True)
    v["xsq28"] = {"jet": hexarr(xj[0]), "dt": float(dt).hex(), "nsteps_tmax_0.33": int(s.nsteps[0]),
                  "tv": hexarr(s.t[: int(s.nsteps[0]) + 1, 0]), "xv": hexarr(s.x[0, : int(s.nsteps[0]) + 1, 0])}
    return v


def main():
    kats = reference_kats()
    with open(os.path.join(HERE, "reference_kats.json"), "w", encoding="utf-8") as f:
        json.dump(kats, f, indent=1, ensure_ascii=False)
        f.write("\n")
    vec = oracle_vectors()
    with open(os.path.join(HERE, "oracle_vectors.json"), "w") as f:
        json.dump(vec, f, indent=0)
        f.write("\n")
    print("wrote reference_kats.json (%d literals) and oracle_vectors.json" % (len(kats) - 2))


if __name__ == "__main__":
    main()
