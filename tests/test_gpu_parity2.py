"""GPU parity, round 2: the step-size branches, event orders and edge cases that only ran on the oracle before.

Every case is `array_equal` against oracle/tb_oracle.cpp through the C ABI (parity mode):
  * reltol > 0 (stepsize.jl:69-73; test/lyapunov.jl:217-245), minstepsize / maxstepsize clamp (taylorstep.jl:79)
  * `_second_stepsize` (stepsize.jl:91-97, 119-132): constant RHS (test/taylorize.jl:303-424), falling ball
    (test/many_ode.jl:337-362)
  * zero step at a pole (taylorinteg.jl:360-363), NaN propagation through Julia's min (stepsize.jl:74, 86)
  * two components whose last coefficients are within 2^-40 (the tie window of the parity fast path)
  * eventorder = 2, newtoniter = 2 forward and backward (test/rootfinding.jl:301-339)
"""
import numpy as np
import pytest

import oracle
import taylorintegration_b200 as tb
from taylorintegration_b200 import _lib as L, problems as P

pytestmark = pytest.mark.gpu
T_PEND = 7.019250311844546


def _same(g, o, fields=("xfinal", "tfinal", "nsteps", "status")):
    for f in fields:
        assert np.array_equal(getattr(g, f), getattr(o, f), equal_nan=True), f


@pytest.mark.parametrize("reltol", [1e-20, 1e-15, 1e-12])
def test_reltol_branch_kepler_and_lorenz_lyapunov(reltol):
    q0, mu, _ = P.kepler_ensemble(96)
    g = tb.taylorinteg(P.kepler_tape(), q0, 0.0, 12.0, 28, 1e-20, mu, maxsteps=3000, reltol=reltol)
    o = oracle.integrate(P.kepler_tape(), 28, q0, 0.0, 12.0, mu, maxsteps=3000, reltol=reltol, nthreads=8)
    _same(g, o)
    lq, par = P.lorenz_ensemble(33)
    g = tb.lyap_taylorinteg(P.lorenz_tape(), lq, 0.0, 2.0, 28, 1e-20, par, maxsteps=2000, reltol=reltol)
    o = oracle.integrate(P.lorenz_tape(), 28, lq, 0.0, 2.0, par, kind=L.KIND_LYAP, maxsteps=2000, reltol=reltol, nthreads=8)
    _same(g, o)
    assert np.array_equal(g.lam, o.lam)
    pq = P.pendulum_ensemble(64)
    g = tb.taylorinteg_events(P.pendulum_tape(), pq, 0.0, 20.0, 25, 1e-20, maxsteps=1000, event_capacity=32, reltol=reltol,
                              output="final")
    o = oracle.integrate(P.pendulum_tape(), 25, pq, 0.0, 20.0, kind=L.KIND_EVENTS, maxsteps=1000, event_capacity=32,
                         reltol=reltol, nthreads=8)
    _same(g, o)
    assert np.array_equal(g.nevents, o.nevents)


def test_stepsize_clamp_min_and_max():
    q0, mu, _ = P.kepler_ensemble(64)
    for kw in (dict(maxstepsize=0.05), dict(minstepsize=0.5), dict(minstepsize=0.01, maxstepsize=0.2)):
        g = tb.taylorinteg(P.kepler_tape(), q0, 0.0, 9.0, 28, 1e-20, mu, maxsteps=5000, **kw)
        o = oracle.integrate(P.kepler_tape(), 28, q0, 0.0, 9.0, mu, maxsteps=5000, nthreads=8, **kw)
        _same(g, o)
    assert np.all(g.tfinal == 9.0)


def test_second_stepsize_constant_rhs_and_falling_ball():
    # x' = -10 from x(1) = 10 to t = 20, order 20 (test/taylorize.jl:303-424): all of x[P-1], x[P] vanish, so
    # h == Inf and the second step-size control decides (stepsize.jl:91-97)
    t = tb.taylorize(lambda dx, x, p, t: dx.__setitem__(0, -10.0 + 0.0 * x[0]), 1, 0)
    q = np.array([[10.0, 3.0, -7.5]])
    g = tb.taylorinteg(t, q, 1.0, 20.0, 20, 1e-20, maxsteps=500, output="steps")
    o = oracle.integrate(t, 20, q, 1.0, 20.0, maxsteps=500, steps=True)
    _same(g, o)
    assert np.all(g.tfinal == 20.0) and abs(g.xfinal[0, 0] - (10.0 - 10.0 * 19.0)) < 10 * np.finfo(float).eps * 190
    for i in range(3):
        n = int(g.nsteps[i])
        assert np.array_equal(g.t[: n + 1, i], o.t[: n + 1, i]) and np.array_equal(g.x[:, : n + 1, i], o.x[:, : n + 1, i])

    # falling ball (test/many_ode.jl:337-362): x' = v, v' = -g, order 10: exact parabola, zero high coefficients
    def ball(dx, x, p, t):
        dx[0] = x[1]
        dx[1] = -9.81 + 0.0 * x[0]
    tb_ = tb.taylorize(ball, 2, 0)
    q = np.array([[10.0, 2.0], [0.0, 1.5]])
    g = tb.taylorinteg(tb_, q, 0.0, 1.0, 10, 1e-20, maxsteps=500)
    o = oracle.integrate(tb_, 10, q, 0.0, 1.0, maxsteps=500)
    _same(g, o)
    assert abs(g.xfinal[0, 0] - (10.0 - 0.5 * 9.81)) < 1e-13 and abs(g.xfinal[1, 1] - (1.5 - 9.81)) < 1e-13


def test_zero_step_and_nan_propagation():
    # x' = x^2 from x0 = 1 runs into the pole at t = 1: the step size underflows to zero (taylorinteg.jl:360-363)
    t = tb.taylorize(lambda dx, x, p, t: dx.__setitem__(0, x[0] ** 2), 1, 0)
    with pytest.warns(UserWarning):
        g = tb.taylorinteg(t, np.array([[1.0, 3.0, 0.5]]), 0.0, 2.5, 28, 1e-20, maxsteps=5000)
    o = oracle.integrate(t, 28, np.array([[1.0, 3.0, 0.5]]), 0.0, 2.5, maxsteps=5000)
    _same(g, o)
    assert np.all(g.status & (L.ST_ZERO_STEP | L.ST_MAXSTEPS | L.ST_NONFINITE)) and np.all(g.tfinal < 2.5)
    # x' = x/x from x0 = 0: NaN coefficients; Julia's min propagates them, the run ends after ONE step with NaN time
    t2 = tb.taylorize(lambda dx, x, p, t: dx.__setitem__(0, x[0] / x[0]), 1, 0)
    g = tb.taylorinteg(t2, np.array([[0.0, 2.0]]), 0.0, 1.0, 20, 1e-20, maxsteps=50)
    o = oracle.integrate(t2, 20, np.array([[0.0, 2.0]]), 0.0, 1.0, maxsteps=50)
    _same(g, o)
    assert g.nsteps[0] == 1 and np.isnan(g.tfinal[0]) and g.tfinal[1] == 1.0 and g.xfinal[0, 1] == 3.0
    # the rolled generator (variant 1) and the event / Lyapunov drivers take the same path
    g1 = tb.taylorinteg(t2, np.array([[0.0, 2.0]]), 0.0, 1.0, 20, 1e-20, maxsteps=50, kernel_opts={"flags": 4})
    _same(g1, o)


def test_tie_window_of_the_parity_stepsize():
    # two uncoupled copies of x' = x^2 whose initial values differ by a few ulp: |x_c[P]| within 2^-40 of each other,
    # so the parity fast path must evaluate BOTH candidates like the reference's loop over components
    def f(dx, x, p, t):
        dx[0] = x[0] ** 2
        dx[1] = x[1] ** 2
    t = tb.taylorize(f, 2, 0)
    base = np.array([3.0, 0.7, 1.9, 2.5])
    q = np.stack([base, base * (1 + np.array([1, -3, 2, 5]) * 2.0 ** -50)])
    g = tb.taylorinteg(t, q, 0.0, 0.2, 28, 1e-20, maxsteps=500)
    o = oracle.integrate(t, 28, q, 0.0, 0.2, maxsteps=500)
    _same(g, o)
    k = tb.Kernel(t, 28)
    xj, dt = k.jet(q)
    for i in range(4):
        _, _, dto = oracle.jet(t, 28, q[:, i])
        assert dt[i] == dto


def test_eventorder2_newtoniter2_forward_and_backward_on_gpu():
    tape = P.pendulum_tape()
    kw = dict(maxsteps=1000, eventorder=2, newtoniter=2)
    with pytest.warns(UserWarning, match="Newton-Raphson"):
        s = tb.taylorinteg_events(tape, np.array([1.3, 0.0]), 0.0, 3 * T_PEND, 28, 1e-20, **kw)
    tvr = np.arange(1, 6) * T_PEND / 2
    assert s.tevents.shape == (5,) and np.max(np.abs(s.tevents - tvr)) < 1e-13 and np.linalg.norm(s.gresids) < 1e-15
    with pytest.warns(UserWarning, match="Newton-Raphson"):
        b = tb.taylorinteg_events(tape, s.x[-1], 3 * T_PEND, 0.0, 28, 1e-20, **kw)
    assert b.t[0] > b.t[-1] and b.tevents[0] > b.tevents[-1] and np.linalg.norm(b.gresids) < 1e-14
    assert np.max(np.abs(b.tevents[1:] - tvr[::-1])) < 1e-13 and np.max(np.abs(b.xevents[1:] - s.xevents[::-1])) < 1e-13
    # ensemble, both directions and event orders 0..2, bit for bit against the oracle
    q0 = P.pendulum_ensemble(96)
    for eo, (t0, t1) in ((1, (0.0, 25.0)), (2, (0.0, 25.0)), (2, (25.0, 0.0)), (0, (25.0, 0.0))):
        import warnings
        with warnings.catch_warnings():
            warnings.simplefilter("ignore")
            g = tb.taylorinteg_events(tape, q0, t0, t1, 25, 1e-20, maxsteps=1000, eventorder=eo, newtoniter=3,
                                      event_capacity=40, output="final")
        o = oracle.integrate(tape, 25, q0, t0, t1, kind=L.KIND_EVENTS, maxsteps=1000, eventorder=eo, newtoniter=3,
                             event_capacity=40, nthreads=8)
        _same(g, o)
        assert np.array_equal(g.nevents, o.nevents) and g.nevents.max() > 2
        for i in range(96):
            n = g.nevents[i]
            assert np.array_equal(g.tevents[:n, i], o.tevents[:n, i]) and np.array_equal(g.gresids[:n, i], o.gresids[:n, i])
            assert np.array_equal(g.xevents[:, :n, i], o.xevents[:, :n, i])


def test_pendulum_events_full_size_sample_parity():
    # BASELINE config 2 at its full size: 2^20 initial conditions, order 25, events on x[2] = 0, capacity 64
    N = 1 << 20
    q0 = P.pendulum_ensemble(N)
    kw = dict(maxsteps=5000, event_capacity=64)
    g = tb.taylorinteg_events(P.pendulum_tape(), q0, 0.0, 100.0, 25, 1e-20, output="final", **kw)
    assert not g.status.any() and np.all(g.tfinal == 100.0)
    # energy is conserved and every event sits on the section x[2] = 0
    en = lambda q: 0.5 * q[1] ** 2 - np.cos(q[0])  # noqa: E731
    assert np.max(np.abs(en(g.xfinal) - en(q0))) < 1e-12
    mask = np.arange(64)[:, None] < g.nevents[None, :]
    assert np.max(np.abs(g.gresids[mask])) < 4 * np.finfo(float).eps and np.max(np.abs(g.xevents[1][mask])) < 1e-13
    assert np.all(np.diff(g.tevents, axis=0)[mask[1:]] > 0)
    assert g.tevents[0, 0] == pytest.approx(T_PEND / 2, abs=1e-13)  # trajectory 0 is the golden theta = 1.3
    idx = np.linspace(0, N - 1, 96).round().astype(np.int64)
    o = oracle.integrate(P.pendulum_tape(), 25, np.ascontiguousarray(q0[:, idx]), 0.0, 100.0, kind=L.KIND_EVENTS,
                         nthreads=8, **kw)
    assert np.array_equal(g.nevents[idx], o.nevents) and np.array_equal(g.nsteps[idx], o.nsteps)
    assert np.array_equal(g.xfinal[:, idx], o.xfinal)
    for j, i in enumerate(idx):
        n = o.nevents[j]
        assert np.array_equal(g.tevents[:n, i], o.tevents[:n, j]) and np.array_equal(g.xevents[:, :n, i], o.xevents[:, :n, j])
        assert np.array_equal(g.gresids[:n, i], o.gresids[:n, j])


def test_device_entry_point_rejects_what_the_host_entry_point_rejects():
    """tb_integrate_device shares the argument / build-flag checks of tb_integrate."""
    import ctypes as C
    import torch
    lib = L.lib()
    dev = torch.device("cuda", 0)
    q0, mu, _ = P.kepler_ensemble(64)
    d_q0, d_mu = torch.from_numpy(q0).to(dev), torch.from_numpy(mu).to(dev)
    d_x = torch.empty((4, 64), dtype=torch.float64, device=dev)
    d_tr = torch.linspace(0, 1, 5, dtype=torch.float64, device=dev)
    work = torch.zeros(256, dtype=torch.uint8, device=dev)
    o = L.TbOut()
    o.xfinal = C.cast(d_x.data_ptr(), L._DP)
    opts = tb.make_opts(maxsteps=50)
    kf = tb.Kernel(P.kepler_tape(), 28, flags=262144)

    def call(k, trange=None, nt=0, out=o, params=d_mu.data_ptr(), q=d_q0.data_ptr(), op=opts):
        return lib.tb_integrate_device(k.handle, 64, q, params, 0.0, 1.0, trange, nt, C.byref(op), C.byref(out),
                                       work.data_ptr(), None)
    assert call(kf) == 0
    assert call(kf, d_tr.data_ptr(), 5) == L.TB_ERR_ARG          # TB_CF_FINAL_ONLY kernel given a trange
    o2 = L.TbOut()
    o2.xfinal, o2.xv = o.xfinal, C.cast(d_x.data_ptr(), L._DP)
    assert call(kf, out=o2) == L.TB_ERR_ARG                      # ... or per-step output
    k = tb.Kernel(P.kepler_tape(), 28)
    assert call(k, d_tr.data_ptr(), 1) == L.TB_ERR_ARG           # nt < 2
    assert call(k, params=None) == L.TB_ERR_ARG and call(k, q=None) == L.TB_ERR_ARG
    ke = tb.Kernel(P.pendulum_tape(), 25, L.KIND_EVENTS)
    oe = tb.make_opts(maxsteps=50, eventorder=26)
    assert lib.tb_integrate_device(ke.handle, 64, d_q0.data_ptr(), None, 0.0, 1.0, None, 0, C.byref(oe), C.byref(o),
                                   work.data_ptr(), None) == L.TB_ERR_ARG
    torch.cuda.synchronize()


def test_integrate_multi_matches_single_device():
    """tb_integrate_multi (one ensemble over the GPUs of this process, round-robin blocks): same bits as the 1-GPU
    call whatever the device list.  On a 1-GPU box the list [0, 0] still runs two host threads over two shards."""
    ndev = L.lib().tb_device_count()
    devs = list(range(ndev)) if ndev >= 2 else [0, 0]
    q0, mu, _ = P.kepler_ensemble(10_000)
    tmax = 2 * np.pi * 2
    one = tb.taylorinteg(P.kepler_tape(), q0, 0.0, tmax, 28, 1e-20, mu, maxsteps=2000)
    for block in (4096, 1000, 7):
        m = tb.taylorinteg(P.kepler_tape(), q0, 0.0, tmax, 28, 1e-20, mu, maxsteps=2000, devices=devs, block=block)
        _same(m, one)
    pq = P.pendulum_ensemble(3000)
    one = tb.taylorinteg_events(P.pendulum_tape(), pq, 0.0, 20.0, 25, 1e-20, maxsteps=1000, event_capacity=16, output="final")
    m = tb.taylorinteg_events(P.pendulum_tape(), pq, 0.0, 20.0, 25, 1e-20, maxsteps=1000, event_capacity=16, output="final",
                              devices=devs, block=512)
    _same(m, one)
    assert np.array_equal(m.nevents, one.nevents)
    mask = np.arange(16)[:, None] < one.nevents[None, :]
    assert np.array_equal(m.tevents[mask], one.tevents[mask]) and np.array_equal(m.xevents[:, mask], one.xevents[:, mask])
    lq, par = P.lorenz_ensemble(700)
    one = tb.lyap_taylorinteg(P.lorenz_tape(), lq, 0.0, 1.0, 28, 1e-20, par, maxsteps=2000)
    m = tb.lyap_taylorinteg(P.lorenz_tape(), lq, 0.0, 1.0, 28, 1e-20, par, maxsteps=2000, devices=devs, block=64)
    _same(m, one)
    assert np.array_equal(m.lam, one.lam)
    tr = np.linspace(0.0, 3.0, 13)
    one = tb.taylorinteg(P.kepler_tape(), q0[:, :900], tr, 28, 1e-20, mu)
    m = tb.taylorinteg(P.kepler_tape(), q0[:, :900], tr, 28, 1e-20, mu, devices=devs, block=100)
    assert np.array_equal(m.x_trange, one.x_trange)
