"""SciML common-interface caller (ext/TaylorIntegrationDiffEqExt.jl; test/common.jl:222-300): `solve(prob, TaylorMethod(order))`
gives the very steps of `taylorinteg`, forward and backward, with reltol, with saveat, dense interpolation, and a
batched EnsembleProblem."""
import numpy as np
import pytest

import taylorintegration_b200 as tb
from taylorintegration_b200.common import EnsembleProblem, ODEProblem, TaylorMethod, solve


def harmosc(dx, x, p, t):  # test/common.jl:40-45
    dx[0] = x[1]
    dx[1] = -x[0]


def test_interface_objects_without_gpu():
    prob = ODEProblem(harmosc, [1.0, 0.0], (0.0, 10 * np.pi))
    assert TaylorMethod(25).order == 25 and TaylorMethod(25).parse_eqs and prob.p is None
    ep = EnsembleProblem(prob, prob_func=lambda pr, i: ODEProblem(pr.f, [1.0 + i, 0.0], pr.tspan))
    assert ep.prob_func(prob, 2).u0 == [3.0, 0.0]


@pytest.mark.gpu
def test_solve_is_taylorinteg():
    order, abstol, u0 = 25, 1e-20, np.array([1.0, 0.0])
    for tspan, kw in (((0.0, 10 * np.pi), {}), ((0.0, 10 * np.pi), {"reltol": 1e-18}), ((0.0, -10 * np.pi), {})):
        sol = solve(ODEProblem(harmosc, u0, tspan), TaylorMethod(order), abstol=abstol, **kw)
        ti = tb.taylorinteg(harmosc, u0, tspan[0], tspan[1], order, abstol, **kw)
        assert np.array_equal(ti.t, sol.t) and np.array_equal(ti.x[-1], sol.u[-1]) and sol.retcode == "Success"
        assert sol.interp == "Taylor series polynomial evaluation"
    # dense interpolation inside the last step (test/common.jl:289-300)
    sol = solve(ODEProblem(harmosc, u0, (0.0, 10 * np.pi)), TaylorMethod(order), abstol=abstol)
    tq = sol.t[-2] + 0.37 * (sol.t[-1] - sol.t[-2])
    assert np.allclose(sol(tq)[0], [np.cos(tq), -np.sin(tq)], rtol=0, atol=1e-13)
    # saveat: states on the grid, from the trange driver
    grid = np.linspace(0.0, 10 * np.pi, 21)
    ss = solve(ODEProblem(harmosc, u0, (0.0, 10 * np.pi)), TaylorMethod(order), abstol=abstol, saveat=grid)
    tr = tb.taylorinteg(harmosc, u0, grid, order, abstol)
    assert np.array_equal(ss.t, grid) and np.array_equal(np.array(ss.u), tr.x)
    assert np.allclose(np.array(ss.u)[:, 0], np.cos(grid), atol=1e-12)


@pytest.mark.gpu
def test_ensemble_problem_is_one_launch():
    base = ODEProblem(harmosc, [1.0, 0.0], (0.0, 7.0))
    ep = EnsembleProblem(base, prob_func=lambda pr, i: ODEProblem(pr.f, [1.0 + 0.01 * i, 0.0], pr.tspan))
    n0 = tb._lib.lib().tb_launch_count()
    sols = solve(ep, TaylorMethod(25), abstol=1e-20, trajectories=300)
    assert tb._lib.lib().tb_launch_count() - n0 == 1 and len(sols) == 300
    for i in (0, 17, 299):
        one = solve(ep.prob_func(base, i), TaylorMethod(25), abstol=1e-20)
        assert np.array_equal(sols[i].u[-1], one.u[-1]) and sols[i].t[-1] == 7.0
