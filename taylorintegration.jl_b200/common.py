"""SciML common-interface caller: `solve(prob, TaylorMethod(order); abstol, ...)` routed to the same C ABI.

Mirrors ext/TaylorIntegrationDiffEqExt.jl:62-168, 203-308 at the level the reference tests it (test/common.jl:222-300):
`solve(ODEProblem(f!, u0, tspan, p), TaylorMethod(order), abstol=...)` must return the very steps of
`taylorinteg(f!, u0, tspan[1], tspan[2], order, abstol, p)` -- `sol.t == solti.t`, `sol.u[end] == solti.x[end, :]` --
with `reltol = 0` unless given (DiffEqExt.jl:287-288), dense "Taylor series polynomial evaluation" interpolation, and
`saveat` served by evaluating the step polynomials.  `EnsembleProblem` batches the trajectories into ONE fused launch
(the reference's EnsembleProblem would loop `solve`).  Callbacks are not routed here: continuous callbacks are the
event drivers (`taylorinteg(f, g, ...)`), discrete ones need per-step host control, which this path does not have.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Callable, List, Optional, Sequence, Tuple

import numpy as np

from .api import TaylorSolution, taylorinteg


@dataclass
class TaylorMethod:
    """`TaylorMethod(order; parse_eqs = true)` (src/common.jl:3, DiffEqExt.jl:30-45)."""
    order: int
    parse_eqs: bool = True


@dataclass
class ODEProblem:
    """`ODEProblem(f!, u0, tspan, p)` with the in-place signature `f!(du, u, p, t)` (the order the SciML interface uses;
    it is also the argument order of the reference's own RHS functions, so `f` can be shared with `taylorinteg`)."""
    f: Callable
    u0: Sequence[float]
    tspan: Tuple[float, float]
    p: Optional[Sequence[float]] = None


@dataclass
class EnsembleProblem:
    """`EnsembleProblem(prob; prob_func)`: `prob_func(prob, i)` returns the i-th problem (new `u0` and/or `p`)."""
    prob: ODEProblem
    prob_func: Optional[Callable[[ODEProblem, int], ODEProblem]] = None


@dataclass
class ODESolution:
    """`sol.t`, `sol.u` (list of states), callable `sol(t)`; `retcode` from the per-trajectory status word."""
    t: np.ndarray
    u: List[np.ndarray]
    retcode: str = "Success"
    interp: str = "Taylor series polynomial evaluation"  # DiffEqBase.interp_summary(sol.interp), test/common.jl:238
    dense: Optional[TaylorSolution] = field(default=None, repr=False)

    def __call__(self, t):
        if self.dense is None or self.dense.p is None:
            raise ValueError("solution was computed with dense = False")
        return self.dense(t)


def _retcode(status: int) -> str:
    return "Success" if status == 0 else ("MaxIters" if status & 1 else "Unstable" if status & 8 else "DtLessThanMin")


def solve(prob, alg: TaylorMethod, *, abstol: float = 1e-6, reltol: float = 0.0, saveat=None, dense: bool = True,
          maxiters: int = 1_000_000, dtmin: float = 0.0, dtmax: float = np.inf, trajectories: Optional[int] = None,
          mode="parity"):
    """`solve(prob, TaylorMethod(order); abstol, reltol = 0, saveat, dense, maxiters, dtmin, dtmax)`.

    One `ODEProblem`: the steps of `taylorinteg` (same `t`, same states, bit for bit).  `saveat` (a grid): `sol.t` is the
    grid and `sol.u` the states there (the `trange` driver, taylorinteg.jl:558-642).  An `EnsembleProblem` with
    `trajectories = N`: a list of N solutions from one launch (final state + saveat grid; no per-step storage)."""
    if isinstance(prob, EnsembleProblem):
        return _solve_ensemble(prob, alg, int(trajectories or 1), abstol, reltol, saveat, maxiters, dtmin, dtmax, mode)
    u0 = np.asarray(prob.u0, dtype=np.float64)
    t0, t1 = float(prob.tspan[0]), float(prob.tspan[1])
    kw = dict(maxsteps=int(maxiters), reltol=reltol, minstepsize=dtmin, maxstepsize=dtmax, mode=mode)
    if saveat is not None:
        grid = np.asarray(saveat, dtype=np.float64)
        if grid.ndim == 0:  # a step: t0:saveat:t1
            grid = np.arange(t0, t1 + 0.5 * float(saveat) * np.sign(t1 - t0), float(saveat) * np.sign(t1 - t0))
        if grid[0] != t0:
            grid = np.concatenate([[t0], grid])
        if grid[-1] != t1:
            grid = np.concatenate([grid, [t1]])
        s = taylorinteg(prob.f, u0, grid, alg.order, abstol, prob.p, **kw)
        return ODESolution(t=grid, u=[row.copy() for row in s.x], retcode=_retcode(s.status))
    s = taylorinteg(prob.f, u0, t0, t1, alg.order, abstol, prob.p, dense=dense, **kw)
    return ODESolution(t=s.t, u=[row.copy() for row in s.x], retcode=_retcode(s.status), dense=s if dense else None)


def _solve_ensemble(ep, alg, N, abstol, reltol, saveat, maxiters, dtmin, dtmax, mode):
    probs = [ep.prob_func(ep.prob, i) if ep.prob_func else ep.prob for i in range(N)]
    t0, t1 = float(ep.prob.tspan[0]), float(ep.prob.tspan[1])
    if any(tuple(map(float, p.tspan)) != (t0, t1) or p.f is not ep.prob.f for p in probs):
        raise ValueError("the batched ensemble needs one RHS and one time span (swept: u0 and p)")
    q0 = np.stack([np.asarray(p.u0, dtype=np.float64) for p in probs], axis=1)
    par = None
    if ep.prob.p is not None:
        pm = np.stack([np.atleast_1d(np.asarray(p.p, dtype=np.float64)) for p in probs], axis=1)
        par = pm[:, 0].copy() if np.all(pm == pm[:, :1]) else pm
    kw = dict(maxsteps=int(maxiters), reltol=reltol, minstepsize=dtmin, maxstepsize=dtmax, mode=mode)
    if saveat is not None:
        grid = np.asarray(saveat, dtype=np.float64)
        s = taylorinteg(ep.prob.f, q0, grid, alg.order, abstol, par, **kw)
        return [ODESolution(t=grid, u=[s.x_trange[:, j, i].copy() for j in range(grid.size)], retcode=_retcode(int(s.status[i])))
                for i in range(N)]
    s = taylorinteg(ep.prob.f, q0, t0, t1, alg.order, abstol, par, **kw)
    return [ODESolution(t=np.array([t0, s.tfinal[i]]), u=[q0[:, i].copy(), s.xfinal[:, i].copy()],
                        retcode=_retcode(int(s.status[i]))) for i in range(N)]
