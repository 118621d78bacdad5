"""GPU parity: the fused sm_100a kernels (parity mode, through the C ABI) against the CPU oracle on the same
seeded inputs.  Floating-point bar (BASELINE.json north_star): first-step coefficients and step sizes within
2 ulp, final states within 1e-12 relative, Lyapunov exponents within 1e-6.  Parity mode is built to be
bit-identical to the oracle (same operations in the same order, shared elementary functions), so these tests
assert exact equality and would catch any drift long before the stated tolerances."""
import mpmath as mp
import numpy as np
import pytest

import oracle
import taylorintegration_b200 as tb
from taylorintegration_b200 import _lib as L, problems as P
from taylorintegration_b200 import cos, exp, log, sin, sqrt

pytestmark = pytest.mark.gpu
T_PEND = 7.019250311844546
mp.mp.prec = 200


def _ulps(a, b):
    return np.max(np.abs(a - b) / np.maximum(np.spacing(np.abs(b)), 1e-300))


def test_first_step_jets_and_stepsize_kepler():
    tape = P.kepler_tape()
    q0, mu, _ = P.kepler_ensemble(257)
    k = tb.Kernel(tape, 28)
    xj, dt = k.jet(q0, mu)
    for i in range(0, 257, 16):
        xo, _, dto = oracle.jet(tape, 28, q0[:, i], mu)
        assert np.array_equal(xj[:, :, i], xo)           # bar: <= 2 ulp
        assert dt[i] == dto
        # step size from the GPU's own coefficients in 200-bit arithmetic: (eps/|x_k|)^(1/k), stepsize.jl:107-111
        want = min(mp.power(mp.mpf(1e-20) / abs(mp.mpf(float(xj[c, kk, i]))), mp.mpf(1) / kk)
                   for c in range(4) for kk in (27, 28) if xj[c, kk, i] != 0.0)
        assert abs(mp.mpf(float(dt[i])) - want) <= 2 * mp.mpf(float(np.spacing(dt[i])))


def test_xsquared_goldens_on_gpu():
    # test/one_ode.jl:203-205, 219, 239
    t = tb.taylorize(lambda dx, x, p, t: dx.__setitem__(0, x[0] ** 2), 1, 0)
    xj, dt = tb.Kernel(t, 28).jet(np.array([3.0]))
    assert np.array_equal(xj[0, :, 0], 3.0 ** np.arange(1, 30))
    s = tb.taylorinteg(t, np.array([3.0]), 0.0, 0.3, 28, 1e-20)
    assert len(s.t) == 14 and s.t[-1] == 0.3 and np.array_equal(s.p[0, 0], 3.0 ** np.arange(1, 30))
    s = tb.taylorinteg(t, np.array([3.0]), 0.0, 0.33, 28, 1e-20)
    assert len(s.t) == 28 and s.t[-1] == 0.33
    with pytest.warns(UserWarning):
        s = tb.taylorinteg(t, np.array([1.0]), 0.0, 1.0, 28, 1e-20)  # pole: zero step / maxsteps
    assert s.t[-1] < 1.0


def test_kepler_ensemble_final_states():
    tape = P.kepler_tape()
    q0, mu, _ = P.kepler_ensemble(300)
    tmax = 2 * np.pi * 5
    g = tb.taylorinteg(tape, q0, 0.0, tmax, 28, 1e-20, mu, maxsteps=2000)
    o = oracle.integrate(tape, 28, q0, 0.0, tmax, mu, maxsteps=2000, nthreads=8)
    assert np.array_equal(g.nsteps, o.nsteps) and np.array_equal(g.status, o.status)
    assert np.array_equal(g.tfinal, o.tfinal) and np.all(g.tfinal == tmax)
    assert np.array_equal(g.xfinal, o.xfinal)  # bar: 1e-12 relative


def test_kepler_per_step_and_dense_output():
    tape = P.kepler_tape()
    q0, mu, _ = P.kepler_ensemble(40)
    tmax = 2 * np.pi
    g = tb.taylorinteg(tape, q0, 0.0, tmax, 28, 1e-20, mu, maxsteps=100, output="steps")
    o = oracle.integrate(tape, 28, q0, 0.0, tmax, mu, maxsteps=100, steps=True, dense=True)
    assert np.array_equal(g.nsteps, o.nsteps)
    for i in range(40):
        n = int(g.nsteps[i])
        assert np.array_equal(g.t[: n + 1, i], o.t[: n + 1, i])
        assert np.array_equal(g.x[:, : n + 1, i], o.x[:, : n + 1, i])
        assert np.array_equal(g.p[:, :, :n, i], o.p[:, :, :n, i])
    # dense evaluation sol(t) on the GPU vs Horner on the host
    s = g.trajectory(7)
    tq = np.linspace(0.0, tmax, 23)
    xq = s(tq)
    for j, t in enumerate(tq):
        idx = min(np.searchsorted(s.t, t, side="right") - 1, len(s.t) - 2)
        ref = [np.polynomial.polynomial.polyval(t - s.t[idx], s.p[idx, c]) for c in range(4)]
        assert np.allclose(xq[j], ref, rtol=1e-13, atol=1e-14)


def test_maxsteps_and_backward():
    tape = P.kepler_tape()
    q0, mu, _ = P.kepler_ensemble(33)
    with pytest.warns(UserWarning):
        g = tb.taylorinteg(tape, q0, 0.0, 100.0, 28, 1e-20, mu, maxsteps=7)
    o = oracle.integrate(tape, 28, q0, 0.0, 100.0, mu, maxsteps=7)
    assert np.all(g.status & L.ST_MAXSTEPS) and np.all(g.nsteps == 7)
    assert np.array_equal(g.xfinal, o.xfinal) and np.array_equal(g.tfinal, o.tfinal)
    g = tb.taylorinteg(tape, q0, 0.0, -3.0, 28, 1e-20, mu)
    o = oracle.integrate(tape, 28, q0, 0.0, -3.0, mu)
    assert np.array_equal(g.xfinal, o.xfinal) and np.all(g.tfinal == -3.0)


def test_trange_variant():
    tape = P.kepler_tape()
    q0, mu, _ = P.kepler_ensemble(50)
    tr = np.linspace(0.0, 7.0, 41)
    g = tb.taylorinteg(tape, q0, tr, 28, 1e-20, mu)
    o = oracle.integrate(tape, 28, q0, params=mu, trange=tr)
    assert np.array_equal(g.x_trange, o.x_trange) and not np.any(np.isnan(g.x_trange))
    # NaN past a blow-up (test/one_ode.jl:101)
    t = tb.taylorize(lambda dx, x, p, t: dx.__setitem__(0, x[0] ** 2), 1, 0)
    tr = np.linspace(0.0, 0.34, 18)
    with pytest.warns(UserWarning):
        g = tb.taylorinteg(t, np.array([[3.0, 2.0]]), tr, 28, 1e-20, maxsteps=100)
    o = oracle.integrate(t, 28, np.array([[3.0, 2.0]]), trange=tr, maxsteps=100)
    assert np.array_equal(g.x_trange, o.x_trange, equal_nan=True) and np.isnan(g.x_trange[0, -1, 0])


def test_pendulum_events():
    # test/rootfinding.jl:106-122 on the GPU + ensemble parity
    tape = P.pendulum_tape()
    s = tb.taylorinteg_events(tape, np.array([1.3, 0.0]), 0.0, 3 * T_PEND, 28, 1e-20, maxsteps=1000)
    assert len(s.tevents) == 5
    assert np.max(np.abs(s.tevents - np.arange(1, 6) * T_PEND / 2)) < 1e-13
    assert np.max(np.abs(s.gresids)) < np.finfo(float).eps
    q0 = P.pendulum_ensemble(200)
    g = tb.taylorinteg_events(tape, q0, 0.0, 30.0, 25, 1e-20, maxsteps=1000, event_capacity=64, output="final")
    o = oracle.integrate(tape, 25, q0, 0.0, 30.0, kind=L.KIND_EVENTS, maxsteps=1000, event_capacity=64, nthreads=8)
    assert np.array_equal(g.nevents, o.nevents) and np.array_equal(g.nsteps, o.nsteps)
    assert np.array_equal(g.xfinal, o.xfinal)
    for i in range(200):
        n = g.nevents[i]
        assert np.array_equal(g.tevents[:n, i], o.tevents[:n, i])
        assert np.array_equal(g.xevents[:, :n, i], o.xevents[:, :n, i])
        assert np.array_equal(g.gresids[:n, i], o.gresids[:n, i])


def test_events_with_trange_variant():
    # taylorinteg(f!, g, q0, trange, ...): src/rootfinding.jl:440-584
    tape = P.pendulum_tape()
    tr = np.linspace(0.0, 3 * T_PEND, 25)
    q0 = np.stack([np.linspace(0.4, 2.9, 70), np.zeros(70)])
    g = tb.taylorinteg_events(tape, q0, None, None, 28, 1e-20, trange=tr, maxsteps=1000, event_capacity=16)
    o = oracle.integrate(tape, 28, q0, kind=L.KIND_EVENTS, trange=tr, abstol=1e-20, maxsteps=1000, event_capacity=16)
    assert np.array_equal(g.nevents, o.nevents) and g.nevents.max() > 3
    for i in range(70):
        ne = g.nevents[i]
        assert np.array_equal(g.tevents[:ne, i], o.tevents[:ne, i]) and np.array_equal(g.xevents[:, :ne, i], o.xevents[:, :ne, i])
        assert np.array_equal(g.gresids[:ne, i], o.gresids[:ne, i])
    assert np.array_equal(g.x_trange, o.x_trange) and np.array_equal(g.nsteps, o.nsteps) and np.array_equal(g.xfinal, o.xfinal)
    one = tb.taylorinteg(P.pendulum, lambda dx, x, p, t: (True, x[1]), np.array([1.3, 0.0]), tr, 28, 1e-20, maxsteps=1000)
    assert one.x.shape == (25, 2) and len(one.tevents) == 5
    assert np.max(np.abs(one.tevents - np.arange(1, 6) * T_PEND / 2)) < 1e-13


def test_python_function_front_end_with_event_and_gate():
    # the reference's call shape taylorinteg(f, g, q0, t0, tmax, order, abstol) (rootfinding.jl:301-345)
    g_gate = lambda dx, x, p, t: (x[0] > 0, x[1])  # noqa: E731  test/taylorize.jl:2549-2603
    s = tb.taylorinteg(P.pendulum, g_gate, np.array([1.3, 0.0]), 0.0, 3 * T_PEND, 28, 1e-20, maxsteps=1000)
    assert len(s.tevents) == 2 and np.max(np.abs(s.tevents - np.array([T_PEND, 2 * T_PEND]))) < 1e-13


def test_lorenz_lyapunov():
    tape = P.lorenz_tape()
    par = np.array([16.0, 45.92, 4.0])
    s = tb.lyap_taylorinteg(tape, np.array([19.0, 20.0, 50.0]), 0.0, 20.0, 28, 1e-20, par, maxsteps=2000)
    lam = s.lam[-1]
    assert abs(lam[0] - 1.47167) < 1e-4 and abs(lam[1] + 0.00830) < 1e-4 and abs(lam[2] + 22.46336) < 1e-4  # test/lyapunov.jl:205-210
    q0, pars = P.lorenz_ensemble(65)
    g = tb.lyap_taylorinteg(tape, q0, 0.0, 5.0, 28, 1e-20, pars, maxsteps=2000)
    o = oracle.integrate(tape, 28, q0, 0.0, 5.0, pars, kind=L.KIND_LYAP, maxsteps=2000, nthreads=8)
    assert np.array_equal(g.nsteps, o.nsteps)
    assert np.array_equal(g.xfinal, o.xfinal)
    assert np.array_equal(g.lam, o.lam)  # bar: 1e-6
    # the call above ran the column-parallel kernel (teams of 3 lanes, TB_KIND_LYAPCOL); the single-thread kernel
    # (TB_KIND_LYAP) must give the same bits, and per-step output must agree between the two as well
    g1 = tb.lyap_taylorinteg(tape, q0, 0.0, 5.0, 28, 1e-20, pars, maxsteps=2000, kernel_opts={"columns": False})
    assert np.array_equal(g1.nsteps, o.nsteps) and np.array_equal(g1.xfinal, o.xfinal) and np.array_equal(g1.lam, o.lam)
    s1 = tb.lyap_taylorinteg(tape, np.array([19.0, 20.0, 50.0]), 0.0, 20.0, 28, 1e-20, par, maxsteps=2000,
                             kernel_opts={"columns": False})
    assert np.array_equal(s.t, s1.t) and np.array_equal(s.x, s1.x) and np.array_equal(s.lam, s1.lam)
    # fast mode through the column kernel: exponents within the reference's own tolerance
    f = tb.lyap_taylorinteg(tape, np.array([19.0, 20.0, 50.0]), 0.0, 20.0, 28, 1e-20, par, maxsteps=2000, mode="fast")
    assert np.max(np.abs(f.lam[-1] - lam)) < 1e-4


def test_lyapunov_trange_variant():
    # lyap_taylorinteg(f!, q0, trange, ...): src/lyapunovspectrum.jl:431-596
    tape = P.lorenz_tape()
    tr = np.linspace(0.0, 2.0, 41)
    q0, pars = P.lorenz_ensemble(23)
    g = tb.lyap_taylorinteg(tape, q0, tr, 28, 1e-20, pars, maxsteps=2000)
    o = oracle.integrate(tape, 28, q0, params=pars, kind=L.KIND_LYAP, trange=tr, maxsteps=2000, nthreads=8)
    assert np.array_equal(g.x_trange, o.x_trange) and np.array_equal(g.lam_trange, o.lam_trange)
    assert np.array_equal(g.lam, o.lam) and np.array_equal(g.nsteps, o.nsteps) and np.array_equal(g.xfinal, o.xfinal)
    one = tb.lyap_taylorinteg(tape, np.array([19.0, 20.0, 50.0]), tr, 28, 1e-20, np.array([16.0, 45.92, 4.0]), maxsteps=2000)
    assert one.x.shape == (41, 3) and one.lam.shape == (41, 3) and not np.isnan(one.lam).any()
    assert abs(one.lam[-1].sum() + 21.0) < 1e-9


def test_lyapunov_column_tape_structure():
    from taylorintegration_b200.tape import lyap_column_tape
    col = lyap_column_tape(P.lorenz_tape())
    assert col is not None and col.dof == 6 and col.npar == 3 and col.jac is None
    assert lyap_column_tape(P.kepler_tape()) is None  # no Jacobian requested


def test_every_primitive_matches_oracle():
    def rhs(dx, x, p, t):
        a = sqrt(x[0] * x[0] + 1.0)
        b = exp(-0.1 * x[1]) + log(a + 2.0)
        c = sin(x[2]) * cos(t) - x[0] / a
        dx[0] = x[1] / (1.0 + b ** 2)
        dx[1] = c - p[0] * x[0] ** 3.0
        dx[2] = (b + 1.0) ** 0.5 - 0.3 * x[2] + p[1] / (2.0 + cos(x[0]))
    tape = tb.taylorize(rhs, 3, 2)
    rng = np.random.default_rng(5)
    q0 = rng.uniform(-1, 1, (3, 70))
    par = rng.uniform(0.5, 1.5, (2, 70))
    for order in (5, 20):
        k = tb.Kernel(tape, order)
        xj, dt = k.jet(q0, par, t0=0.3)
        for i in range(0, 70, 7):
            xo, _, dto = oracle.jet(tape, order, q0[:, i], par[:, i], t0=0.3, params_per_traj=0)
            assert np.all(np.isfinite(xo)) and np.array_equal(xj[:, :, i], xo) and dt[i] == dto
    g = tb.taylorinteg(tape, q0, 0.3, 4.0, 20, 1e-18, par, maxsteps=3000)
    o = oracle.integrate(tape, 20, q0, 0.3, 4.0, par, abstol=1e-18, maxsteps=3000)
    assert np.array_equal(g.xfinal, o.xfinal) and np.array_equal(g.nsteps, o.nsteps)


def test_fast_mode_is_close_to_parity_mode():
    tape = P.kepler_tape()
    q0, mu, _ = P.kepler_ensemble(128)
    tmax = 2 * np.pi * 10
    a = tb.taylorinteg(tape, q0, 0.0, tmax, 28, 1e-20, mu, maxsteps=5000)
    b = tb.taylorinteg(tape, q0, 0.0, tmax, 28, 1e-20, mu, maxsteps=5000, mode="fast")
    rel = np.max(np.abs(a.xfinal - b.xfinal)) / np.max(np.abs(a.xfinal))
    assert rel < 1e-9, rel  # not bit-faithful: two roundings of the same recurrences drift apart (hard part 2)
    assert np.max(np.abs(a.nsteps - b.nsteps)) <= 1


def test_final_only_build_matches_full_outputs_build():
    """output='final' selects the TB_CF_FINAL_ONLY build of the plain kernel; it must return the bits of the build that
    also carries the per-step / dense / trange code, and the C-ABI must refuse per-step outputs on it."""
    q0, mu, _ = P.kepler_ensemble(300)
    tm = 2 * np.pi * 2
    for mode in ("parity", "fast"):
        a = tb.taylorinteg(P.kepler_tape(), q0, 0.0, tm, 28, 1e-20, mu, maxsteps=2000, mode=mode)
        b = tb.taylorinteg(P.kepler_tape(), q0, 0.0, tm, 28, 1e-20, mu, maxsteps=2000, mode=mode,
                           kernel_opts={"full_outputs": True})
        c = tb.taylorinteg(P.kepler_tape(), q0, 0.0, tm, 28, 1e-20, mu, maxsteps=2000, mode=mode, output="steps")
        assert np.array_equal(a.xfinal, b.xfinal) and np.array_equal(a.nsteps, b.nsteps)
        assert np.array_equal(a.xfinal, c.xfinal) and np.array_equal(a.tfinal, c.tfinal)
    with pytest.raises(L.TaylorB200Error):
        tb.taylorinteg(P.kepler_tape(), q0, 0.0, tm, 28, 1e-20, mu, maxsteps=50, output="steps",
                       kernel_opts={"flags": 262144})
