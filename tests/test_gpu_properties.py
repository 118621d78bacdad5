"""Size-independent properties at BASELINE.json's full sizes, and the edge cases of the ensemble interface (GPU).

The oracle cannot integrate 2^20 trajectories in seconds, so the full-size run is checked through invariants of the
problem (energy, angular momentum, closed orbits), through consistency across launches of different size (a trajectory's
result must not depend on its neighbours), and by spot-checking a strided sample against the oracle bit-for-bit.
"""
import numpy as np
import pytest

import oracle
import taylorintegration_b200 as tb
from taylorintegration_b200 import _lib as L, problems as P

pytestmark = pytest.mark.gpu


def _energy(q):
    return 0.5 * (q[2] ** 2 + q[3] ** 2) - 1.0 / np.hypot(q[0], q[1])


def _angmom(q):
    return q[0] * q[3] - q[1] * q[2]


def test_kepler_full_size_invariants_and_sample_parity():
    N = 1 << 20  # the north-star ensemble (BASELINE.json metric), 10 periods to keep the oracle sample short
    q0, mu, e = P.kepler_ensemble(N)
    tmax = 2 * np.pi * 10
    s = tb.taylorinteg(P.kepler_tape(), q0, 0.0, tmax, 28, 1e-20, mu, maxsteps=10_000)
    assert not s.status.any() and np.all(s.tfinal == tmax)
    # test/taylorize.jl:2430-2433: energy conserved to 1e-12 (relative to |E| = 1/2 here); same for angular momentum
    assert np.max(np.abs(_energy(s.xfinal) - _energy(q0))) < 1e-12
    assert np.max(np.abs(_angmom(s.xfinal) - _angmom(q0)) / np.abs(_angmom(q0))) < 1e-12
    assert np.max(np.abs(s.xfinal - q0)) < 1e-8          # ten closed orbits
    # steps per period grow with the eccentricity (docs/src/kepler.md): monotone up to the +-1 of the last partial step
    ns = s.nsteps.astype(np.int64)
    assert ns[0] < ns[N // 2] < ns[-1] and np.all(np.diff(ns[:: 4096]) >= -2)
    # a strided sample against the oracle, bit for bit
    idx = np.linspace(0, N - 1, 64).round().astype(np.int64)
    o = oracle.integrate(P.kepler_tape(), 28, np.ascontiguousarray(q0[:, idx]), 0.0, tmax, mu, maxsteps=10_000, nthreads=8)
    assert np.array_equal(s.xfinal[:, idx], o.xfinal) and np.array_equal(s.nsteps[idx], o.nsteps)
    # neighbours do not matter: the same trajectories in a small launch give the same bits
    small = tb.taylorinteg(P.kepler_tape(), np.ascontiguousarray(q0[:, idx]), 0.0, tmax, 28, 1e-20, mu, maxsteps=10_000)
    assert np.array_equal(small.xfinal, s.xfinal[:, idx]) and np.array_equal(small.nsteps, s.nsteps[idx])


def test_lorenz_full_size_trace_identity():
    # the exponents of a Lorenz flow sum to the trace of the Jacobian, -(1 + sigma + beta), whatever rho is
    N = 1 << 16
    rho = 28.0 + 64.0 * np.arange(N) / (N - 1)
    par = np.stack([np.full(N, 16.0), rho, np.full(N, 4.0)])
    q0 = np.tile(np.array([[19.0], [20.0], [50.0]]), (1, N))
    s = tb.lyap_taylorinteg(P.lorenz_tape(), q0, 0.0, 5.0, 28, 1e-20, par, maxsteps=20_000)
    assert not s.status.any() and np.all(s.tfinal == 5.0)
    assert np.max(np.abs(s.lam.sum(axis=0) + 21.0)) < 1e-9
    assert np.all(s.lam[1] > s.lam[2])
    idx = np.array([0, 17, 4097, N - 1])
    o = oracle.integrate(P.lorenz_tape(), 28, np.ascontiguousarray(q0[:, idx]), 0.0, 5.0, np.ascontiguousarray(par[:, idx]),
                         kind=L.KIND_LYAP, maxsteps=20_000)
    assert np.array_equal(s.lam[:, idx], o.lam) and np.array_equal(s.xfinal[:, idx], o.xfinal)


@pytest.mark.parametrize("n", [1, 31, 33, 257])
def test_ragged_ensembles(n):
    q0, mu, _ = P.kepler_ensemble(max(n, 2))
    q0 = np.ascontiguousarray(q0[:, :n])
    s = tb.taylorinteg(P.kepler_tape(), q0, 0.0, 1.0, 28, 1e-20, mu, maxsteps=500)
    o = oracle.integrate(P.kepler_tape(), 28, q0, 0.0, 1.0, mu, maxsteps=500)
    assert np.array_equal(s.xfinal, o.xfinal) and np.array_equal(s.nsteps, o.nsteps) and not s.status.any()
    par = np.array([16.0, 45.92, 4.0])
    lq = np.tile(np.array([[19.0], [20.0], [50.0]]), (1, n)) + 1e-3 * np.arange(n)[None, :]
    g = tb.lyap_taylorinteg(P.lorenz_tape(), lq, 0.0, 0.3, 28, 1e-20, par, maxsteps=500)  # teams of 3 lanes: 10 per warp
    r = oracle.integrate(P.lorenz_tape(), 28, lq, 0.0, 0.3, par, kind=L.KIND_LYAP, maxsteps=500)
    assert np.array_equal(g.lam, r.lam) and np.array_equal(g.xfinal, r.xfinal) and np.array_equal(g.nsteps, r.nsteps)


def test_empty_ensemble_and_degenerate_interval():
    q0 = np.zeros((4, 0))
    s = tb.taylorinteg(P.kepler_tape(), q0, 0.0, 1.0, 28, 1e-20, np.array([-1.0]))
    assert s.xfinal.shape == (4, 0) and s.nsteps.shape == (0,)
    q, mu, _ = P.kepler_ensemble(5)
    s = tb.taylorinteg(P.kepler_tape(), q, 2.0, 2.0, 28, 1e-20, mu)     # t0 == tmax: no step (taylorinteg.jl:356)
    assert np.array_equal(s.xfinal, q) and not s.nsteps.any() and np.all(s.tfinal == 2.0)


def test_status_flags_follow_the_reference_warnings():
    q, mu, _ = P.kepler_ensemble(40)
    s = tb.taylorinteg(P.kepler_tape(), q, 0.0, 100.0, 28, 1e-20, mu, maxsteps=7)   # taylorinteg.jl:372-377
    assert np.all(s.status & L.ST_MAXSTEPS) and np.all(s.nsteps == 7) and np.all(s.tfinal < 100.0)
    o = oracle.integrate(P.kepler_tape(), 28, q, 0.0, 100.0, mu, maxsteps=7)
    assert np.array_equal(s.xfinal, o.xfinal) and np.array_equal(s.tfinal, o.tfinal) and np.array_equal(s.status, o.status)
    bad = q.copy()
    bad[0, 3] = np.nan
    s = tb.taylorinteg(P.kepler_tape(), bad, 0.0, 1.0, 28, 1e-20, mu, maxsteps=50)
    o = oracle.integrate(P.kepler_tape(), 28, bad, 0.0, 1.0, mu, maxsteps=50)
    assert np.array_equal(s.status, o.status) and s.status[3] != 0 and not s.status[:3].any()
    assert np.array_equal(s.nsteps, o.nsteps)
