"""More oracle known-answer tests against literals the reference holds (round 2): higher-order event detection forward
and backward (test/rootfinding.jl:286-339), the Kepler notebook's long run (examples/Kepler-problem.ipynb cells 3, 9,
23), Julia's NaN-propagating min in the step-size control (src/integrator/stepsize.jl:74, 86), the reltol switch
(stepsize.jl:69-73) and the clamp of taylorstep! (src/integrator/taylorstep.jl:79)."""
import numpy as np
import pytest

import oracle
import taylorintegration_b200 as tb
from taylorintegration_b200 import _lib as L, problems as P
from taylorintegration_b200.jettransport import evaluate

T_PEND = 7.019250311844546  # test/rootfinding.jl:30


def test_eventorder2_newtoniter2_forward_and_backward():
    # test/rootfinding.jl:301-339: zeros of the SECOND derivative of g = x[2] are the zeros of g for this pendulum;
    # two Newton iterations do not reach eps (warning = TB_ST_NR_NOCONV) yet the events are exact to 1e-13
    tape = P.pendulum_tape()
    kw = dict(kind=L.KIND_EVENTS, maxsteps=1000, eventorder=2, newtoniter=2, steps=True)
    o = oracle.integrate(tape, 28, np.array([1.3, 0.0]), 0.0, 3 * T_PEND, **kw)
    n = int(o.nevents[0])
    tvr = np.arange(1, 6) * T_PEND / 2
    assert n == 5 and (o.status[0] & L.ST_NR_NOCONV)
    assert np.max(np.abs(o.tevents[:n, 0] - tvr)) < 1e-13 and np.linalg.norm(o.gresids[:n, 0]) < 1e-15
    assert o.t[0, 0] == 0.0 and np.array_equal(o.x[:, 0, 0], [1.3, 0.0])
    b = oracle.integrate(tape, 28, o.xfinal[:, 0], 3 * T_PEND, 0.0, **kw)
    nb = int(b.nevents[0])
    assert nb == 6 and b.tfinal[0] == 0.0 and b.tevents[0, 0] > b.tevents[nb - 1, 0] and (b.status[0] & L.ST_NR_NOCONV)
    assert np.linalg.norm(b.gresids[:nb, 0]) < 1e-14
    assert np.max(np.abs(b.tevents[1:nb, 0] - tvr[::-1])) < 1e-13
    assert np.max(np.abs(b.xevents[:, 1:nb, 0] - o.xevents[:, :n, 0][:, ::-1])) < 1e-13


def test_eventorder_above_order_is_refused():
    with pytest.raises(AssertionError):  # rootfinding.jl:321 (@assert order >= eventorder), test/rootfinding.jl:287-298
        tb.taylorinteg_events(P.pendulum_tape(), np.array([1.3, 0.0]), 0.0, 1.0, 28, 1e-20, eventorder=29, newtoniter=2)


def test_kepler_notebook_ten_thousand_periods():
    # examples/Kepler-problem.ipynb: q0 = [0.2, 0, 0, 3] (cell 3), order 28, abstol 1e-20, t = 10^4 periods:
    # `436153-element` time vector (cell 23) = 436152 steps; final state (0.2, 2.72854e-9, -2.27378e-8, 3.0) (cell 9).
    # The phase error (components 2, 3) is accumulated rounding (the notebook ran Julia 0.5's pow and the generic
    # path), so those two are compared by magnitude; the step count is reproduced exactly.
    o = oracle.integrate(P.kepler_tape(), 28, np.array([0.2, 0.0, 0.0, 3.0]), 0.0, 2 * np.pi * 1e4, np.array([-1.0]),
                         maxsteps=1_000_000)
    assert abs(int(o.nsteps[0]) - 436152) <= 0.005 * 436152 and int(o.nsteps[0]) == 436152
    x = o.xfinal[:, 0]
    assert o.tfinal[0] == 2 * np.pi * 1e4 and o.status[0] == 0
    assert abs(x[0] - 0.2) < 5e-9 and abs(x[3] - 3.0) < 5e-8 and abs(x[1]) < 1e-8 and abs(x[2]) < 1e-7
    e0, e1 = 0.5 * 9 - 1 / 0.2, 0.5 * (x[2] ** 2 + x[3] ** 2) - 1 / np.hypot(x[0], x[1])
    assert abs((e1 - e0) / e0) < 2e-13  # notebook: about -7e-14


def test_nan_coefficient_makes_the_step_nan_like_julias_min():
    # x' = x / x from x0 = 0: 0/0 = NaN at order 0, every higher coefficient NaN.  Julia's min(h, NaN) = NaN
    # (stepsize.jl:74, 86), so dt, the state and t0 become NaN in ONE step and `sign*t0 < sign*tmax` ends the loop.
    t = tb.taylorize(lambda dx, x, p, t: dx.__setitem__(0, x[0] / x[0]), 1, 0)
    o = oracle.integrate(t, 20, np.array([[0.0, 2.0]]), 0.0, 1.0, maxsteps=50)
    assert o.nsteps[0] == 1 and np.isnan(o.tfinal[0]) and np.isnan(o.xfinal[0, 0]) and (o.status[0] & L.ST_NONFINITE)
    assert o.status[1] == 0 and o.tfinal[1] == 1.0 and o.xfinal[0, 1] == 3.0  # x' = 1 exactly (second stepsize path)


def test_reltol_branch_and_stepsize_clamp():
    # stepsize.jl:69-73: eps = abstol if abstol >= reltol*|x0| else reltol*|x0|; the larger tolerance gives fewer steps
    q0, mu, _ = P.kepler_ensemble(8)
    a = oracle.integrate(P.kepler_tape(), 28, q0, 0.0, 20.0, mu, maxsteps=5000)
    r = oracle.integrate(P.kepler_tape(), 28, q0, 0.0, 20.0, mu, maxsteps=5000, reltol=1e-12)
    assert np.all(r.nsteps < a.nsteps) and np.all(r.tfinal == 20.0)
    assert np.max(np.abs(r.xfinal - a.xfinal)) < 1e-7
    # taylorstep.jl:79 clamp(dt, minstepsize, maxstepsize): a finite maxstepsize bounds every step from above
    c = oracle.integrate(P.kepler_tape(), 28, q0, 0.0, 20.0, mu, maxsteps=5000, maxstepsize=0.05, steps=True)
    assert np.all(c.nsteps >= 400)
    for i in range(8):
        assert np.max(np.diff(c.t[: c.nsteps[i] + 1, i])) <= 0.05 + 1e-15
    m = oracle.integrate(P.kepler_tape(), 28, q0, 0.0, 20.0, mu, maxsteps=5000, minstepsize=0.5, steps=True)
    for i in range(8):
        d = np.diff(m.t[: m.nsteps[i] + 1, i])
        assert np.all(d[:-1] >= 0.5)  # every step but the last (which lands on tmax) is at least minstepsize


def test_monomial_order_is_pinned_by_an_asymmetric_displacement():
    # test/jettransport.jl:705-805: a jet evaluated at xi must equal the run from q0 + xi.  With xi_1 != xi_2 a wrong
    # monomial order (inside the product tables or in the flat layout) breaks the equality: permuting the two
    # variables in the evaluation must fail.
    tape = P.pendulum_tape(events=False)
    mon = oracle.tn_monomials(2, 4)
    q0 = np.zeros((2, len(mon), 1))
    q0[0, 0, 0], q0[0, 1, 0], q0[1, 2, 0] = 1.3, 1.0, 1.0
    r = oracle.integrate_tn(tape, 28, 2, 4, q0, 0.0, T_PEND, maxsteps=500, abstol=1e-20)
    xi = np.array([1e-3, -0.5e-3])
    s2 = oracle.integrate(tape, 28, np.array([1.3, 0.0]) + xi, 0.0, T_PEND, maxsteps=500)
    good = evaluate(r["xfinal"], mon, xi)[:, 0]
    swapped = evaluate(r["xfinal"], mon[:, ::-1], xi)[:, 0]
    assert np.max(np.abs(good - s2.xfinal[:, 0])) < 1e-11
    assert np.max(np.abs(swapped - s2.xfinal[:, 0])) > 1e-5


def test_modified_gram_schmidt_has_the_properties_the_reference_tests():
    """test/lyapunov.jl:88-110 (stated there for classicalGS!, the same contract holds for the modifiedGS! the
    integrator calls, src/lyapunovspectrum.jl:118-146): A = QR to 1e-14, orthonormal columns, R upper triangular with
    a positive diagonal."""
    import ctypes as C
    lib = oracle.lib()
    lib.tbo_modified_gs.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int]
    lib.tbo_modified_gs.restype = None
    rng = np.random.Generator(np.random.PCG64(7))
    for n in (2, 3, 6):
        A = np.asfortranarray(rng.random((n, n)))
        Q, R = np.zeros((n, n), order="F"), np.zeros((n, n), order="F")
        lib.tbo_modified_gs(A.ctypes.data, Q.ctypes.data, R.ctypes.data, n)
        assert np.abs(A - Q @ R).max() < 1e-14
        assert np.allclose(np.diag(Q.T @ Q), 1.0, rtol=0, atol=1e-14)
        assert np.abs(Q.T @ Q - np.eye(n)).max() < 1e-10
        assert np.array_equal(R, np.triu(R)) and np.all(np.diag(R) > 0)
