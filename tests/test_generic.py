"""The reference's strongest invariant, restated: the generic `jetcoeffs!` (RHS re-evaluated on truncated series per
order, src/integrator/jetcoeffs.jl:66-92; oracle/tb_generic.cpp, right-hand sides hand-written from the Julia
sources) and the parsed/tape path (oracle/tb_oracle.cpp interpreting the tape built by tape.py) agree BITWISE:
`iszero(norm(tv - tvp, Inf))`, `iszero(norm(xv - xvp, Inf))` (test/taylorize.jl:1267-1273, 1354-1364, 1817-1819).
The product and the tape oracle share the tape front end; this check does not pass through it."""
import numpy as np

import oracle
import taylorintegration_b200 as tb
from taylorintegration_b200 import problems as P
from taylorintegration_b200 import cos


def _both(name, tape, order, q0, t0, tmax, params=None, maxsteps=500):
    tg, xg, sg = oracle.generic_integrate(name, order, q0, t0, tmax, params, maxsteps=maxsteps)
    o = oracle.integrate(tape, order, q0, t0, tmax, params, steps=True, maxsteps=maxsteps)
    n = int(o.nsteps[0])
    return (tg, xg, sg), (o.t[: n + 1, 0], o.x[:, : n + 1, 0], int(o.status[0]))


def test_kepler_e08_order28_100_periods_generic_equals_parsed_bitwise():
    # test/taylorize.jl:1180-1273: q0 = [0.2, 0, 0, 3], order 28, tf = 2pi*100, abstol 1e-20
    (tg, xg, sg), (tp, xp, sp) = _both("kepler1!", P.kepler_tape(), 28, np.array([0.2, 0.0, 0.0, 3.0]), 0.0,
                                       2 * np.pi * 100, np.array([-1.0]), maxsteps=10_000)
    assert sg == sp == 0 and len(tg) == len(tp) > 4000
    assert np.array_equal(tg, tp) and np.array_equal(xg, xp)
    assert tg[-1] == 2 * np.pi * 100


def test_first_step_jets_generic_equals_parsed_bitwise():
    q0, mu, _ = P.kepler_ensemble(17)
    for i in range(17):
        xj, _, _ = oracle.jet(P.kepler_tape(), 28, q0[:, i], mu)
        assert np.array_equal(oracle.generic_jet("kepler1!", 28, q0[:, i], mu), xj)
    xj, _, _ = oracle.jet(P.pendulum_tape(events=False), 25, np.array([1.3, 0.0]))
    assert np.array_equal(oracle.generic_jet("pendulum!", 25, np.array([1.3, 0.0])), xj)
    par = np.array([16.0, 45.92, 4.0])
    xj, _, _ = oracle.jet(P.lorenz_tape(jacobian=False), 28, np.array([19.0, 20.0, 50.0]), par)
    assert np.array_equal(oracle.generic_jet("lorenz!", 28, np.array([19.0, 20.0, 50.0]), par), xj)
    t = tb.taylorize(lambda dx, x, p, t: dx.__setitem__(0, x[0] ** 2), 1, 0)
    xj, _, _ = oracle.jet(t, 28, np.array([3.0]))
    assert np.array_equal(oracle.generic_jet("xsquared", 28, np.array([3.0])), xj)
    assert np.array_equal(xj[0], 3.0 ** np.arange(1, 30))  # test/one_ode.jl:203


def test_pendulum_lorenz_cos_t_integrations_generic_equals_parsed_bitwise():
    (tg, xg, _), (tp, xp, _) = _both("pendulum!", P.pendulum_tape(events=False), 25, np.array([1.3, 0.0]), 0.0, 50.0,
                                     maxsteps=2000)
    assert np.array_equal(tg, tp) and np.array_equal(xg, xp) and tg[-1] == 50.0
    par = np.array([16.0, 45.92, 4.0])
    (tg, xg, _), (tp, xp, _) = _both("lorenz!", P.lorenz_tape(jacobian=False), 28, np.array([19.0, 20.0, 50.0]), 0.0,
                                     20.0, par, maxsteps=2000)
    assert np.array_equal(tg, tp) and np.array_equal(xg, xp) and len(tg) > 800  # chaotic: any bit would show
    # non-autonomous, backward too (test/many_ode.jl:239-335): the time series is a tape operand
    tape = tb.taylorize(lambda dx, x, p, t: dx.__setitem__(0, cos(t)), 1, 0)
    for tmax in (10.25 * 2 * np.pi, -10.25 * 2 * np.pi):
        (tg, xg, _), (tp, xp, _) = _both("cos_t", tape, 25, np.array([0.0]), 0.0, tmax)
        assert np.array_equal(tg, tp) and np.array_equal(xg, xp)
        assert abs(np.sin(tmax) - xg[0, -1]) < 5e-14


def test_harmonic_oscillator_scalar_times_series():
    # test/taylorize.jl:878-923: dx2 = -(w^2 * x1) with w^2 a plain number
    def f(dx, x, p, t):
        dx[0] = x[1]
        dx[1] = -((p[0] * p[0]) * x[0])
    tape = tb.taylorize(f, 2, 1)
    (tg, xg, _), (tp, xp, _) = _both("harmonic", tape, 20, np.array([1.0, 0.0]), 0.0, 300.0, np.array([2.0]), maxsteps=1000)
    assert np.array_equal(tg, tp) and np.array_equal(xg, xp)
    assert abs(xg[0, -1] - np.cos(600.0)) < 1e-12 and abs(xg[1, -1] + 2 * np.sin(600.0)) < 3e-12


def test_manyspin5_generic_equals_parsed_bitwise():
    # benchmark/manyspin_benchmarks.jl:10-41 with 5 spins (the reference benchmarks 3..7): the coupling loops
    # `sum_s[i] = aux[i] + J[a,i]*x[a]` written out on Taylor1 values against the tape the tracer produces
    nn = 5
    B, J = P.fix_B(nn), P.fix_J(nn)
    tape = P.manyspin_tape(nn, B, J)
    q0 = P.manyspin_ensemble(nn, 3)[:, 1].copy()
    par = np.concatenate([B, J.reshape(-1)])  # J[a, i] at 5 + a*5 + i
    xg = oracle.generic_jet("spin_odes!5", 25, q0, par)
    xp, _, _ = oracle.jet(tape, 25, q0)
    assert np.array_equal(xg, xp)
    tg, xgi, _ = oracle.generic_integrate("spin_odes!5", 25, q0, 0.0, 10.0, par, maxsteps=2000)
    s = oracle.integrate(tape, 25, q0, 0.0, 10.0, steps=True, maxsteps=2000)
    n = int(s.nsteps[0])
    assert len(tg) == n + 1 and np.array_equal(tg, s.t[: n + 1, 0]) and np.array_equal(xgi, s.x[:, : n + 1, 0])
    # |s_i| = 1 is conserved by the equations of motion
    assert np.allclose((xgi[:, -1].reshape(nn, 3) ** 2).sum(1), 1.0, atol=1e-12)


def test_kepler_with_sqrt_and_looped_pendula_generic_equals_parsed_bitwise():
    # test/taylorize.jl:1488-1500 (r = sqrt(q1^2 + q2^2), r^3) and 999-1005 (pendula in a loop): sqrt!, an integer
    # power through pow!, and a RHS written with a loop, against the traced tapes
    from taylorintegration_b200.tape import sqrt, sin

    def kep(dq, q, p, t):
        mu = -1.0
        r = sqrt(q[0] ** 2 + q[1] ** 2)
        r3 = r ** 3
        dq[0] = q[2]
        dq[1] = q[3]
        dq[2] = mu * q[0] / r3
        dq[3] = mu * q[1] / r3
    q0 = np.array([0.2, 0.0, 0.0, 3.0])
    (tg, xg, _), (tp, xp, _) = _both("kepler_sqrt!", tb.taylorize(kep, 4, 0), 28, q0, 0.0, 2 * np.pi * 10, maxsteps=5000)
    assert np.array_equal(tg, tp) and np.array_equal(xg, xp) and len(tg) > 400

    def pend(dx, x, p, t):
        for i in range(3):
            dx[i] = x[3 + i]
            dx[i + 3] = -sin(x[i])
    q0 = np.array([1.3, 0.4, 2.9, 0.0, 0.5, -0.2])
    (tg, xg, _), (tp, xp, _) = _both("multpendula!", tb.taylorize(pend, 6, 0), 25, q0, 0.0, 40.0, maxsteps=3000)
    assert np.array_equal(tg, tp) and np.array_equal(xg, xp)
