"""`lyap_taylorinteg(f!, q0, t0, tmax, order, abstol, params, jacobianfunc!)`: the user-Jacobian path of
src/lyapunovspectrum.jl:65-78.  The reference's own test integrates Lorenz with the automatic (TaylorN) Jacobian and
with `lorenz_jac!` and requires `tv == tv2`, `xv == xv2`, `λv == λv2` (test/lyapunov.jl:270-296): held here on the
oracle (CPU) and on the GPU kernels, plus the golden λ(t = 20) of test/lyapunov.jl:297-302."""
import numpy as np
import pytest

import oracle
import taylorintegration_b200 as tb
from taylorintegration_b200 import _lib as L, problems as P
from taylorintegration_b200.tape import lyap_column_tape, taylorize


def lorenz_jac(jac, x, p, t):
    # test/lyapunov.jl:32-43 (`-σ + zero(x[1])` is the constant series: a number here)
    s, r, b = p[0], p[1], p[2]
    jac[0, 0] = -s
    jac[0, 1] = s
    jac[0, 2] = 0.0
    jac[1, 0] = r - x[2]
    jac[1, 1] = -1.0
    jac[1, 2] = -x[0]
    jac[2, 0] = x[1]
    jac[2, 1] = x[0]
    jac[2, 2] = -b


Q0 = np.array([19.0, 20.0, 50.0])
PAR = np.array([16.0, 45.92, 4.0])  # test/lyapunov.jl:17-21


def test_user_jacobian_tape_and_oracle_equal_the_automatic_one():
    auto = P.lorenz_tape()
    user = taylorize(P.lorenz, 3, 3, jacobian=lorenz_jac, name="lorenz!+jac")
    assert user.jac is not None and len(user.jac) == 9 and user.jac[2][0] == L.JAC_ZERO
    assert [e[0] for e in user.jac].count(L.JAC_SERIES) >= 5
    a = oracle.integrate(auto, 28, Q0, 0.0, 20.0, PAR, kind=L.KIND_LYAP, maxsteps=2000)
    u = oracle.integrate(user, 28, Q0, 0.0, 20.0, PAR, kind=L.KIND_LYAP, maxsteps=2000)
    for f in ("xfinal", "tfinal", "nsteps", "lam"):
        assert np.array_equal(getattr(a, f), getattr(u, f)), f
    # golden of the reference for this run (test/lyapunov.jl:297-302)
    assert abs(u.lam[0, 0] - 1.47167) < 2e-4 and abs(u.lam[2, 0] + 22.46336) < 2e-3
    assert abs(u.lam[:, 0].sum() + (1 + 16.0 + 4.0)) < 1e-9
    with pytest.raises(ValueError):
        taylorize(P.lorenz, 3, 3, jacobian=lambda jac, x, p, t: jac.__setitem__((0, 0), 1.0))
    # the column form (v' = J v per column) exists for the user Jacobian as well
    assert lyap_column_tape(user) is not None


@pytest.mark.gpu
def test_gpu_user_jacobian_equals_automatic_and_oracle():
    q0, par = P.lorenz_ensemble(40)
    user = taylorize(P.lorenz, 3, 3, jacobian=lorenz_jac, name="lorenz!+jac")
    o = oracle.integrate(user, 28, q0, 0.0, 3.0, par, kind=L.KIND_LYAP, maxsteps=2000, nthreads=8)
    a = tb.lyap_taylorinteg(P.lorenz, q0, 0.0, 3.0, 28, 1e-20, par, maxsteps=2000)
    u = tb.lyap_taylorinteg(P.lorenz, q0, 0.0, 3.0, 28, 1e-20, par, lorenz_jac, maxsteps=2000)
    s = tb.lyap_taylorinteg(P.lorenz, q0, 0.0, 3.0, 28, 1e-20, par, lorenz_jac, maxsteps=2000, kernel_opts={"columns": False})
    for f in ("xfinal", "tfinal", "nsteps", "status", "lam"):
        assert np.array_equal(getattr(u, f), getattr(o, f)), f
        assert np.array_equal(getattr(u, f), getattr(a, f)), f
        assert np.array_equal(getattr(s, f), getattr(o, f)), f
