"""Pins the CPU oracle (oracle/tb_oracle.cpp) against the reference's own goldens.

Each test cites the reference test it restates.  These run without a GPU.
"""
import numpy as np
import pytest

import oracle
import taylorintegration_b200 as tb
from taylorintegration_b200 import _lib as L, problems as P
from taylorintegration_b200 import cos, sin


def xsq(dx, x, p, t):  # test/one_ode.jl: eqs_mov(x, p, t) = x^2
    dx[0] = x[0] ** 2


XSQ = tb.taylorize(xsq, 1, 0)


def test_xsquared_jet_is_powers_of_three_exactly():
    # test/one_ode.jl:203-205, test/many_ode.jl:35-38
    xj, _, dt = oracle.jet(XSQ, 28, [3.0], abstol=1e-20)
    assert np.array_equal(xj[0], 3.0 ** np.arange(1, 30))
    for math in (oracle.MATH_TB, oracle.MATH_LIBM):
        _, _, dt = oracle.jet(XSQ, 28, [3.0], abstol=1e-20, math=math)
        expect = (1e-20 / xj[0, 27]) ** (1 / 27)
        assert abs(dt - expect) <= 2 * np.spacing(expect)


@pytest.mark.parametrize("tmax,npoints,tol", [(0.3, 14, 5e-14), (0.33, 28, 1e-11)])
def test_xsquared_step_counts(tmax, npoints, tol):
    # test/one_ode.jl:219-224, 239-244: length(tv) == 14 / 28, tv[end] == tmax, error bounds
    s = oracle.integrate(XSQ, 28, [3.0], 0.0, tmax, abstol=1e-20, steps=True)
    assert int(s.nsteps[0]) + 1 == npoints
    assert s.tfinal[0] == tmax
    exact = 3.0 / (1.0 - 3.0 * tmax)
    assert abs(s.xfinal[0, 0] - exact) / exact < tol or abs(s.xfinal[0, 0] - exact) < tol * exact


def test_xsquared_pole_gives_zero_step_abort():
    # test/one_ode.jl:26-48: x0 = 1, integrate to t = 1 hits the pole: coefficients all 1, tv[end] < 1
    xj, _, dt = oracle.jet(XSQ, 28, [1.0], abstol=1e-20)
    assert np.array_equal(xj[0], np.ones(29))
    assert abs(dt - 1e-20 ** (1 / 27)) <= 2 * np.spacing(dt)
    s = oracle.integrate(XSQ, 28, [1.0], 0.0, 1.0, abstol=1e-20, maxsteps=500)
    assert s.tfinal[0] < 1.0
    assert s.status[0] & (L.ST_ZERO_STEP | L.ST_MAXSTEPS)


def test_minus_xsquared_dense_polynomial():
    # test/taylorize.jl:282-299: xdot = p - x^2, x0 = 1, p = 0, order 20: psol[1] == [(-1)^i]
    t = tb.taylorize(lambda dx, x, p, t: dx.__setitem__(0, p[0] - x[0] ** 2), 1, 1)
    s = oracle.integrate(t, 20, [1.0], 0.0, 0.5, [0.0], steps=True, dense=True, abstol=1e-20)
    assert np.array_equal(s.p[0, :, 0, 0], (-1.0) ** np.arange(21))


def test_tanh_closed_form():
    # test/taylorize.jl:207-210: xdot = b - x^2, b = 3, x0 = 1... closed form to 1e-15 at t = 1000
    t = tb.taylorize(lambda dx, x, p, t: dx.__setitem__(0, p[0] - x[0] ** 2), 1, 1)
    s = oracle.integrate(t, 20, [1.0], 0.0, 1000.0, [3.0], abstol=1e-20, maxsteps=5000)
    assert s.tfinal[0] == 1000.0
    assert abs(s.xfinal[0, 0] - np.sqrt(3.0)) < 1e-15 * 2


def test_nonautonomous_cos_t():
    # test/one_ode.jl:247-270: xdot = cos(t), order 25, 10.25 periods, |sin(tmax) - x| < 1e-14
    t = tb.taylorize(lambda dx, x, p, t: dx.__setitem__(0, cos(t)), 1, 0)
    tmax = 10.25 * 2 * np.pi
    s = oracle.integrate(t, 25, [0.0], 0.0, tmax, abstol=1e-20, maxsteps=5000)
    assert s.tfinal[0] == tmax
    assert abs(np.sin(tmax) - s.xfinal[0, 0]) < 1e-14
    # backward (test/one_ode.jl:272-286)
    s = oracle.integrate(t, 25, [np.sin(tmax)], tmax, 0.0, abstol=1e-20, maxsteps=5000)
    assert s.tfinal[0] == 0.0 and abs(s.xfinal[0, 0]) < 5e-14


def test_harmonic_oscillator():
    # test/taylorize.jl:878-923: x1' = x2, x2' = -(w^2 x1), w = 2 (as a product of numbers), order 20, t = 300
    def ho(dx, x, p, t):
        dx[0] = x[1]
        dx[1] = -(4.0 * x[0])
    t = tb.taylorize(ho, 2, 0)
    s = oracle.integrate(t, 20, [1.0, 0.0], 0.0, 300.0, abstol=1e-20, maxsteps=1000)
    assert s.tfinal[0] == 300.0 and s.nsteps[0] < 1000
    assert abs(s.xfinal[0, 0] - np.cos(600.0)) < 1e-12
    assert abs(s.xfinal[1, 0] + 2 * np.sin(600.0)) < 3e-12


def test_constant_rhs_uses_second_stepsize():
    # test/taylorize.jl:303-424: xdot = -10 from x(1) = 10 to t = 20: exact line; all high coefficients vanish,
    # which exercises the isinf(h) -> _second_stepsize branch (stepsize.jl:91-97, 119-132)
    t = tb.taylorize(lambda dx, x, p, t: dx.__setitem__(0, -10.0), 1, 0)
    s = oracle.integrate(t, 20, [10.0], 1.0, 20.0, abstol=1e-20)
    assert s.tfinal[0] == 20.0
    assert abs(s.xfinal[0, 0] - (10.0 - 10.0 * 19.0)) < 10 * np.finfo(float).eps * 190


def test_falling_ball():
    # test/many_ode.jl:337-362: x'' = -g: exact parabola at order 10
    def ball(dx, x, p, t):
        dx[0] = x[1]
        dx[1] = -9.8 + 0.0 * x[0]
    t = tb.taylorize(ball, 2, 0)
    s = oracle.integrate(t, 10, [10.0, 0.0], 0.0, 1.0, abstol=1e-20)
    assert s.tfinal[0] == 1.0
    assert abs(s.xfinal[0, 0] - (10.0 - 4.9)) < 1e-14 and abs(s.xfinal[1, 0] + 9.8) < 1e-14


def test_kepler_energy_and_step_density():
    # e = 0.8 ([0.2,0,0,3.0], kepler_benchmarks.jl:81), order 28, abstol 1e-20: 43.6 steps/period
    # (examples/Kepler-problem.ipynb: 436152 steps / 1e4 periods); energy drift < 1e-12 over 100 periods
    # (test/taylorize.jl:2430-2433)
    tape = P.kepler_tape()
    q0 = np.array([0.2, 0.0, 0.0, 3.0])
    s = oracle.integrate(tape, 28, q0, 0.0, 2 * np.pi * 100, [-1.0], maxsteps=10000)
    assert s.status[0] == 0 and s.tfinal[0] == 2 * np.pi * 100
    assert abs(s.nsteps[0] / 100 - 43.6) < 0.5

    def energy(q):
        return 0.5 * (q[2] ** 2 + q[3] ** 2) - 1.0 / np.hypot(q[0], q[1])
    assert abs(energy(s.xfinal[:, 0]) - energy(q0)) < 1e-12
    assert np.max(np.abs(s.xfinal[:, 0] - q0)) < 1e-9  # closed orbit


def test_pendulum_event_times():
    # test/rootfinding.jl:30, 106-122: x0 = [1.3, 0], g = x[2], order 28, to 3T: 5 events at k*T/2 within 1e-13
    T = 7.019250311844546
    tape = P.pendulum_tape()
    s = oracle.integrate(tape, 28, [1.3, 0.0], 0.0, 3 * T, kind=L.KIND_EVENTS, abstol=1e-20, maxsteps=1000)
    assert s.nevents[0] == 5
    te = s.tevents[:5, 0]
    assert np.max(np.abs(te - np.arange(1, 6) * T / 2)) < 1e-13
    assert np.max(np.abs(s.gresids[:5, 0])) < np.finfo(float).eps
    assert np.max(np.abs(s.xevents[1, :5, 0])) < 1e-14


def test_pendulum_gated_events():
    # test/taylorize.jl:2549-2603: gate x[1] > 0 keeps the crossings at T and 2T only
    T = 7.019250311844546
    tape = tb.taylorize(P.pendulum, 2, 0, event=lambda dx, x, p, t: (x[0] > 0, x[1]))
    s = oracle.integrate(tape, 28, [1.3, 0.0], 0.0, 3 * T, kind=L.KIND_EVENTS, abstol=1e-20, maxsteps=1000)
    assert s.nevents[0] == 2
    assert np.max(np.abs(s.tevents[:2, 0] - np.array([T, 2 * T]))) < 1e-13


def test_lorenz_lyapunov_spectrum_t20():
    # test/lyapunov.jl:191-215: sigma = 16, rho = 45.92, beta = 4, q0 = [19,20,50], order 28, abstol 1e-20, t = 20
    tape = P.lorenz_tape()
    par = [16.0, 45.92, 4.0]
    s = oracle.integrate(tape, 28, [19.0, 20.0, 50.0], 0.0, 20.0, par, kind=L.KIND_LYAP, maxsteps=2000)
    lam = s.lam[:, 0]
    assert abs(lam[0] - 1.47167) < 1e-4 and abs(lam[1] + 0.00830) < 1e-4 and abs(lam[2] + 22.46336) < 1e-4
    assert abs(lam.sum() + 21.0) < 1e-10  # trace of the Jacobian: -(1 + sigma + beta)
    # orbit identical to the plain integration (test/lyapunov.jl:212-215)
    s2 = oracle.integrate(P.lorenz_tape(jacobian=False), 28, [19.0, 20.0, 50.0], 0.0, 20.0, par, maxsteps=2000)
    assert np.array_equal(s.xfinal, s2.xfinal) and s.nsteps[0] == s2.nsteps[0]


def test_lorenz_tape_jacobian_matches_analytic():
    # test/lyapunov.jl:45-86: AD Jacobian == analytic Jacobian, trace == -(1 + sigma + beta)
    tape = P.lorenz_tape()
    J = np.array(tape.jac, dtype=object).reshape(3, 3, 3)
    kinds = J[:, :, 0].astype(int)
    assert kinds.tolist() == [[L.JAC_PAR, L.JAC_PAR, L.JAC_ZERO], [L.JAC_SERIES, L.JAC_CONST, L.JAC_SERIES],
                              [L.JAC_SERIES, L.JAC_SERIES, L.JAC_PAR]]
    q0, par = [19.0, 20.0, 50.0], [16.0, 45.92, 4.0]
    j10 = oracle.series(tape, 5, q0, par, int(J[1, 0, 1]), with_jac=True)
    j12 = oracle.series(tape, 5, q0, par, int(J[1, 2, 1]), with_jac=True)
    assert j10[0] == 45.92 - 50.0 and j12[0] == -19.0
    assert float(J[0, 0, 2]) == -1.0 and float(J[1, 1, 2]) == -1.0 and float(J[2, 2, 2]) == -1.0


def test_trange_matches_step_output_and_nan_fill():
    # test/one_ode.jl:101: trange output is NaN past the blow-up; test/many_ode.jl:325 consistency
    tr = np.linspace(0.0, 0.34, 18)
    s = oracle.integrate(XSQ, 28, [3.0], trange=tr, abstol=1e-20, maxsteps=100)
    x = s.x_trange[0, :, 0]
    ok = tr < 1 / 3 - 1e-3
    assert np.all(np.abs(x[ok] - 3 / (1 - 3 * tr[ok])) < 1e-9 * np.abs(x[ok]))
    assert np.isnan(x[-1])


def test_events_with_trange_variant():
    # src/rootfinding.jl:485-584: same events as the (t0, tmax) driver, same in-step outputs as the plain trange
    # driver; the step that lands on tmax exits before its event test and before nsteps += 1 (547-550)
    T = 7.019250311844546
    tape = P.pendulum_tape()
    tr = np.linspace(0.0, 3 * T, 25)
    a = oracle.integrate(tape, 28, [1.3, 0.0], kind=L.KIND_EVENTS, trange=tr, abstol=1e-20, maxsteps=1000)
    b = oracle.integrate(tape, 28, [1.3, 0.0], 0.0, 3 * T, kind=L.KIND_EVENTS, abstol=1e-20, maxsteps=1000)
    c = oracle.integrate(tape, 28, [1.3, 0.0], trange=tr, abstol=1e-20, maxsteps=1000)
    assert a.nevents[0] == b.nevents[0] == 5 and np.array_equal(a.tevents[:5], b.tevents[:5])
    assert np.array_equal(a.xevents[:, :5], b.xevents[:, :5]) and np.array_equal(a.gresids[:5], b.gresids[:5])
    assert np.array_equal(a.x_trange, c.x_trange) and not np.isnan(a.x_trange).any()
    assert a.nsteps[0] == c.nsteps[0] == b.nsteps[0] - 1 and a.tfinal[0] == tr[-1]
    assert np.max(np.abs(a.tevents[:5, 0] - np.arange(1, 6) * T / 2)) < 1e-13  # test/rootfinding.jl:120-122


def test_lyapunov_trange_variant():
    # src/lyapunovspectrum.jl:481-596: states and exponents at the output times (NaN-prefilled, first column q0 / 0);
    # the last column carries the exponents of the (t0, tmax) driver at tmax, and the orbit is the plain trange orbit
    tape = P.lorenz_tape()
    par = [16.0, 45.92, 4.0]
    tr = np.linspace(0.0, 2.0, 41)
    a = oracle.integrate(tape, 28, [19.0, 20.0, 50.0], params=par, kind=L.KIND_LYAP, trange=tr, maxsteps=2000)
    b = oracle.integrate(tape, 28, [19.0, 20.0, 50.0], 0.0, 2.0, par, kind=L.KIND_LYAP, maxsteps=2000)
    c = oracle.integrate(P.lorenz_tape(jacobian=False), 28, [19.0, 20.0, 50.0], params=par, trange=tr, maxsteps=2000)
    lam_tr = a.lam_trange if getattr(a, "lam_trange", None) is not None else a.lam_steps
    assert np.array_equal(a.x_trange, c.x_trange) and not np.isnan(a.x_trange).any()
    assert np.array_equal(lam_tr[:, 0, 0], np.zeros(3)) and not np.isnan(lam_tr).any()
    assert np.array_equal(lam_tr[:, -1, 0], b.lam[:, 0]) and np.array_equal(a.lam[:, 0], b.lam[:, 0])
    assert abs(lam_tr[:, -1, 0].sum() + 21.0) < 1e-9 and a.nsteps[0] == b.nsteps[0] - 1
