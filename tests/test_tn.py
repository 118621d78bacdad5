"""TaylorN jet transport: oracle known-answer tests (CPU) and GPU == oracle parity.

Goldens follow the reference's own tests: test/jettransport.jl:19-66 (x' = x^2 with a TaylorN initial
condition against the closed form), 705-805 (pendulum: jet at xi = 0 equals the scalar run, jet evaluated at a
displacement equals the displaced integration), test/taylorize.jl:2352-2433 (Kepler in 4 variables, energy).
"""
import math

import numpy as np
import pytest

import oracle
import taylorintegration_b200 as tb
from taylorintegration_b200 import problems as P
from taylorintegration_b200.jettransport import evaluate
from taylorintegration_b200.tape import taylorize

T_PEND = 7.019250311844546  # test/rootfinding.jl:30


def x2(dx, x, p, t):
    dx[0] = x[0] ** 2


def test_monomial_tables():
    for n, d in ((1, 4), (2, 3), (2, 8), (3, 4), (4, 8)):
        mon = oracle.tn_monomials(n, d)
        assert len(mon) == math.comb(n + d, d)
        deg = mon.sum(1)
        assert (np.diff(deg) >= 0).all()  # ascending total degree
        for g in range(d + 1):  # inside a degree: lexicographically descending exponents
            blk = [tuple(r) for r in mon[deg == g]]
            assert blk == sorted(blk, reverse=True)
            assert len(set(blk)) == len(blk) == math.comb(n + g - 1, g)
    assert oracle.tn_monomials(2, 2).tolist() == [[0, 0], [1, 0], [0, 1], [2, 0], [1, 1], [0, 2]]


def test_library_tables_match_oracle():
    jt = tb.JetTransport(P.kepler_tape(), 28, 4, 8)  # no GPU needed to build tables
    assert jt.ncoef == 495
    assert np.array_equal(jt.monomials, oracle.tn_monomials(4, 8))


def _x2_exact(mon, t):
    a = 1 - 3 * t
    out = np.zeros(len(mon))
    for m, e in enumerate(mon):
        j = int(e[0])
        if e[1:].sum() == 0:
            out[m] = 3 * t ** j / a ** (j + 1) + (t ** (j - 1) / a ** j if j >= 1 else 0.0)
    return out


def test_oracle_x2_closed_form():
    """x' = x^2, x0 = 3 + xi_1, 2 variables of order 5: x(t) = x0/(1 - x0 t) (test/jettransport.jl:19-66)."""
    tape = taylorize(x2, 1, 0, name="x2")
    mon = oracle.tn_monomials(2, 5)
    q0 = np.zeros((1, len(mon), 1))
    q0[0, 0, 0], q0[0, 1, 0] = 3.0, 1.0
    r = oracle.integrate_tn(tape, 28, 2, 5, q0, 0.0, 0.3, maxsteps=500, abstol=1e-20)
    assert r["tfinal"][0] == 0.3 and r["status"][0] == 0
    ex = _x2_exact(mon, 0.3)
    assert np.allclose(r["xfinal"][0, :, 0], ex, rtol=1e-10, atol=1e-10)


def test_oracle_pendulum_jet_vs_displaced():
    """Pendulum, 2 variables: constant term == scalar run; jet(xi) == run from q0 + xi (jettransport.jl:705-805)."""
    tape = P.pendulum_tape(events=False)
    mon = oracle.tn_monomials(2, 4)
    q0 = np.zeros((2, len(mon), 1))
    q0[0, 0, 0], q0[0, 1, 0], q0[1, 2, 0] = 1.3, 1.0, 1.0
    r = oracle.integrate_tn(tape, 28, 2, 4, q0, 0.0, T_PEND, maxsteps=500, abstol=1e-20)
    s = oracle.integrate(tape, 28, np.array([1.3, 0.0]), 0.0, T_PEND, maxsteps=500)
    assert np.allclose(r["xfinal"][:, 0, 0], s.xfinal[:, 0], rtol=0, atol=1e-13)
    assert abs(r["xfinal"][0, 0, 0] - 1.3) < 1e-13  # closes after one period
    xi = np.array([1e-3, -0.5e-3])
    s2 = oracle.integrate(tape, 28, np.array([1.3, 0.0]) + xi, 0.0, T_PEND, maxsteps=500)
    assert np.allclose(evaluate(r["xfinal"], mon, xi)[:, 0], s2.xfinal[:, 0], rtol=0, atol=1e-11)  # O(|xi|^5)


def test_oracle_kepler_jet_energy():
    """Kepler, 4 variables of order 2: energy of the jet's constant term is conserved (taylorize.jl:2430-2433)."""
    tape = P.kepler_tape()
    mon = oracle.tn_monomials(4, 2)
    q0 = np.zeros((4, len(mon), 1))
    q0[:, 0, 0] = [0.2, 0.0, 0.0, 3.0]
    for c in range(4):
        q0[c, 1 + c, 0] = 1.0
    r = oracle.integrate_tn(tape, 28, 4, 2, q0, 0.0, 2 * np.pi * 2, params=[-1.0], maxsteps=2000, abstol=1e-20)
    x = r["xfinal"][:, 0, 0]
    e0 = 0.5 * 9.0 - 1 / 0.2
    e1 = 0.5 * (x[2] ** 2 + x[3] ** 2) - 1 / np.hypot(x[0], x[1])
    assert abs((e1 - e0) / e0) < 1e-12
    s = oracle.integrate(tape, 28, np.array([0.2, 0.0, 0.0, 3.0]), 0.0, 2 * np.pi * 2, params=[-1.0], maxsteps=2000)
    assert np.allclose(x, s.xfinal[:, 0], rtol=0, atol=1e-9)


# ---- GPU --------------------------------------------------------------------------------------------------
@pytest.mark.gpu
def test_gpu_x2_bitwise_and_closed_form():
    tape = taylorize(x2, 1, 0, name="x2")
    jt = tb.JetTransport(tape, 28, 2, 5)
    N = 37
    q0 = np.zeros((1, jt.ncoef, N))
    q0[0, 0, :] = 3.0 - 0.01 * np.arange(N)
    q0[0, 1, :] = 1.0
    sol = jt.integrate(q0, 0.0, 0.3, 1e-20, maxsteps=500)
    ref = oracle.integrate_tn(tape, 28, 2, 5, q0, 0.0, 0.3, maxsteps=500, abstol=1e-20)
    assert np.array_equal(sol.nsteps, ref["nsteps"]) and np.array_equal(sol.tfinal, ref["tfinal"])
    assert np.array_equal(sol.xfinal, ref["xfinal"])  # parity mode is bit-exact
    assert np.allclose(sol.xfinal[0, :, 0], _x2_exact(jt.monomials, 0.3), rtol=1e-10, atol=1e-10)


@pytest.mark.gpu
def test_gpu_pendulum_order8_trange_bitwise():
    """Config 5a: pendulum, 2 variables of order 8, trange = 0:T/8:T (docs/src/pendulum.md:92-106)."""
    tape = P.pendulum_tape(events=False)
    jt = tb.JetTransport(tape, 28, 2, 8)
    th = np.array([1.3, 0.4, 2.9, 1.0])
    q0 = jt.initial(np.stack([th, np.zeros_like(th)]))
    tr = np.linspace(0.0, T_PEND, 9)
    sol = jt.integrate(q0, abstol=1e-20, trange=tr, maxsteps=500)
    ref = oracle.integrate_tn(tape, 28, 2, 8, q0, trange=tr, maxsteps=500, abstol=1e-20)
    assert np.array_equal(sol.nsteps, ref["nsteps"])
    assert np.array_equal(sol.x_trange, ref["x_trange"], equal_nan=True)
    assert np.array_equal(sol.xfinal, ref["xfinal"])
    xi = np.array([2e-3, -1e-3])
    disp = tb.taylorinteg(tape, np.stack([th + xi[0], np.zeros_like(th) + xi[1]]), 0.0, T_PEND, 28, 1e-20, maxsteps=500)
    assert np.allclose(sol.evaluate(xi), disp.xfinal, rtol=0, atol=1e-9)


@pytest.mark.gpu
def test_gpu_kepler_4vars_bitwise():
    """Config 5b (small varorder): kepler1! in 4 variables (test/taylorize.jl:2352-2367), per-trajectory check."""
    tape = P.kepler_tape()
    jt = tb.JetTransport(tape, 28, 4, 3)
    q, mu, _ = P.kepler_ensemble(6)
    q0 = jt.initial(q)
    sol = jt.integrate(q0, 0.0, 2 * np.pi, 1e-20, mu, maxsteps=2000)
    ref = oracle.integrate_tn(tape, 28, 4, 3, q0, 0.0, 2 * np.pi, params=mu, maxsteps=2000, abstol=1e-20)
    assert np.array_equal(sol.nsteps, ref["nsteps"]) and (sol.status == 0).all()
    assert np.array_equal(sol.xfinal, ref["xfinal"])


@pytest.mark.gpu
def test_gpu_fast_mode_close_to_parity():
    tape = P.pendulum_tape(events=False)
    a = tb.JetTransport(tape, 25, 2, 6)
    b = tb.JetTransport(tape, 25, 2, 6, mode=tb._lib.MODE_FAST)
    q0 = a.initial(np.array([[1.3, 0.7], [0.0, 0.2]]))
    sa, sb = a.integrate(q0, 0.0, 10.0, 1e-20, maxsteps=500), b.integrate(q0, 0.0, 10.0, 1e-20, maxsteps=500)
    assert np.array_equal(sa.nsteps, sb.nsteps)
    scale = np.abs(sa.xfinal).max()
    assert np.abs(sa.xfinal - sb.xfinal).max() <= 1e-11 * scale


@pytest.mark.gpu
def test_gpu_kepler_4vars_order8_config5b_bitwise():
    """BASELINE config 5b at its stated shape: kepler1! in 4 variables of order 8 (495 coefficients per component),
    time order 28; three nominal orbits over a quarter period, bit for bit against the oracle.  The constant term of
    the jet is the scalar orbit, and the jet evaluated at a displacement matches the displaced integration."""
    tape = P.kepler_tape()
    jt = tb.JetTransport(tape, 28, 4, 8)
    assert jt.ncoef == 495
    q, mu, _ = P.kepler_ensemble(5)
    q = np.ascontiguousarray(q[:, [0, 2, 4]])          # e = 0, 0.45, 0.9
    q0 = jt.initial(q)
    tmax = np.pi / 2
    sol = jt.integrate(q0, 0.0, tmax, 1e-20, mu, maxsteps=2000)
    ref = oracle.integrate_tn(tape, 28, 4, 8, q0, 0.0, tmax, params=mu, maxsteps=2000, abstol=1e-20)
    assert np.array_equal(sol.nsteps, ref["nsteps"]) and (sol.status == 0).all() and np.all(sol.tfinal == tmax)
    assert np.array_equal(sol.xfinal, ref["xfinal"])
    xi = np.array([1e-3, -2e-3, 1.5e-3, -0.5e-3])
    disp = tb.taylorinteg(tape, q + xi[:, None], 0.0, tmax, 28, 1e-20, mu, maxsteps=2000)
    assert np.allclose(sol.evaluate(xi)[:, :2], disp.xfinal[:, :2], rtol=0, atol=1e-10)  # e = 0.9 needs |xi| smaller


def test_schedule_pairs_independent_entries():
    """tn_schedule: kepler1! = [x^2 | y^2], +, ^1.5, identity x2, mu*x, mu*y moved ahead, [x/r^3 | y/r^3]; the
    pendulum's sincos! stays alone (it fuses its two sums internally)."""
    import ctypes as C
    from taylorintegration_b200 import _lib as L

    def sched(jt):
        buf = (C.c_uint32 * 64)()
        n = L.lib().tb_tn_schedule(jt._t, buf, 64)
        return [(e & 0xffff, None if (e >> 16) == 0xffff else e >> 16) for e in list(buf)[:n]]

    k = sched(tb.JetTransport(P.kepler_tape(), 28, 2, 2))
    assert k == [(0, 1), (2, None), (3, None), (4, None), (5, None), (6, None), (8, None), (7, 9)]
    p = sched(tb.JetTransport(P.pendulum_tape(events=False), 25, 2, 2))
    assert all(q is None for _, q in p) and [j for j, _ in p] == list(range(len(p)))
