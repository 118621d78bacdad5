"""Lyapunov "continue from state" (the state of a run is (t0, x0, λtsum, Q); tb_out_t.lyap_state /
lyap_state0).  The reference keeps that state inside lyap_taylorinteg! (src/lyapunovspectrum.jl:340-429: `jt`, `λtsum`,
`t00`) and has no resume call, so the checks are: a resumed run carries on with exactly the arithmetic of the loop
(the frame is the Q of the last modifiedGS!, the sums are the running sums, the exponents are averaged from t00), a run
cut into two legs reproduces the golden exponents of test/lyapunov.jl:297-302, and GPU == oracle bit for bit."""
import numpy as np
import pytest

import oracle
import taylorintegration_b200 as tb
from taylorintegration_b200 import _lib as L, problems as P

Q0 = np.array([[19.0], [20.0], [50.0]])
PAR = np.array([16.0, 45.92, 4.0])


def test_oracle_state_is_the_loop_state_and_two_legs_reach_the_golden():
    tape = P.lorenz_tape()
    full = oracle.integrate(tape, 28, Q0, 0.0, 20.0, PAR, kind=L.KIND_LYAP, maxsteps=2000)
    s = full.lyap_state[:, 0]
    Q = s[:9].reshape(3, 3, order="F")
    assert np.allclose(Q.T @ Q, np.eye(3), atol=1e-13)                 # the frame modifiedGS! left (lyapunovspectrum.jl:118-146)
    assert np.array_equal(s[9:12] / (full.tfinal[0] - s[12]), full.lam[:, 0]) and s[12] == 0.0   # λ = λtsum / (t - t00)
    # two legs: 0 -> 7.5 -> 20.  Leg 1 ends with a truncated step, so the step sequence differs from the single run,
    # the exponents agree to the accuracy of the method and meet the reference's golden (rtol 1e-4)
    a = oracle.integrate(tape, 28, Q0, 0.0, 7.5, PAR, kind=L.KIND_LYAP, maxsteps=2000)
    b = oracle.integrate(tape, 28, a.xfinal, float(a.tfinal[0]), 20.0, PAR, kind=L.KIND_LYAP, maxsteps=2000,
                         lyap_state0=a.lyap_state)
    assert b.tfinal[0] == 20.0 and b.lyap_state[12, 0] == 0.0
    assert abs(b.lam[0, 0] - 1.47167) < 3e-4 and abs(b.lam[2, 0] + 22.46336) < 3e-3
    assert abs(b.lam[:, 0].sum() + 21.0) < 1e-9                          # trace of the Jacobian: exact invariant
    assert np.allclose(b.lam[:, 0], full.lam[:, 0], rtol=0, atol=2e-3)
    # a fresh run from the same point forgets the history: different exponents
    c = oracle.integrate(tape, 28, a.xfinal, float(a.tfinal[0]), 20.0, PAR, kind=L.KIND_LYAP, maxsteps=2000)
    assert abs(c.lam[0, 0] - b.lam[0, 0]) > 1e-3
    # resuming at the final time is a no-op that returns the state
    d = oracle.integrate(tape, 28, b.xfinal, 20.0, 20.0, PAR, kind=L.KIND_LYAP, maxsteps=10, lyap_state0=b.lyap_state)
    assert np.array_equal(d.lyap_state, b.lyap_state) and d.nsteps[0] == 0


@pytest.mark.gpu
@pytest.mark.parametrize("columns", [True, False])
def test_gpu_resume_equals_oracle(columns):
    q0, par = P.lorenz_ensemble(24)
    tape = P.lorenz_tape()
    kw = dict(maxsteps=2000, kernel_opts={"columns": columns})
    a = tb.lyap_taylorinteg(tape, q0, 0.0, 1.5, 28, 1e-20, par, **kw)
    oa = oracle.integrate(tape, 28, q0, 0.0, 1.5, par, kind=L.KIND_LYAP, maxsteps=2000, nthreads=8)
    assert np.array_equal(a.lyap_state, oa.lyap_state) and np.array_equal(a.lam, oa.lam)
    # tfinal is 1.5 for every member: one scalar t0 continues the ensemble
    b = tb.lyap_taylorinteg(tape, a.xfinal, 1.5, 3.0, 28, 1e-20, par, resume=a.lyap_state, **kw)
    ob = oracle.integrate(tape, 28, oa.xfinal, 1.5, 3.0, par, kind=L.KIND_LYAP, maxsteps=2000, nthreads=8,
                          lyap_state0=oa.lyap_state)
    for f in ("xfinal", "tfinal", "nsteps", "status", "lam", "lyap_state"):
        assert np.array_equal(getattr(b, f), getattr(ob, f)), f
    assert np.all(b.lyap_state[12] == 0.0) and np.all(b.tfinal == 3.0)
