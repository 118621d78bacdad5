"""Jet transport: `taylorinteg` with TaylorN-valued initial conditions, batched over nominal points.

The reference has no separate entry point: `taylorinteg(f!, q0 + xi, ...)` with `xi = set_variables(...)`
dispatches the same driver on `U = TaylorN{Float64}` (docs/src/pendulum.md:92-108, test/jettransport.jl).
Here the polynomial coefficients are explicit arrays `[dof, ncoef, N]` in the flat order of `monomials()`
(ascending total degree; inside a degree, lexicographically descending exponents, TaylorSeries' order).
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass
from typing import Optional

import numpy as np

from . import _lib as L
from .api import _params, _ptr, make_opts
from .tape import Tape, taylorize


@dataclass
class JetSolution:
    """Final jets `x[dof, ncoef, N]` (+ `x_trange[dof, ncoef, nt, N]`), `t[N]`, `nsteps[N]`, `status[N]`."""
    xfinal: np.ndarray
    tfinal: np.ndarray
    nsteps: np.ndarray
    status: np.ndarray
    monomials: np.ndarray
    x_trange: Optional[np.ndarray] = None

    def evaluate(self, xi, at_trange: bool = False) -> np.ndarray:
        """Evaluate every polynomial at the displacement `xi[nvars]` (TaylorSeries `evaluate(x, xi)`)."""
        return evaluate(self.x_trange if at_trange else self.xfinal, self.monomials, xi)


class JetTransport:
    """Compiled jet-transport integrator for one ODE (tape), time order, and TaylorN shape."""

    def __init__(self, tape: Tape, order: int, nvars: int, varorder: int, mode: int = L.MODE_PARITY):
        lib = L.lib()
        self.tape, self.order, self.nvars, self.varorder = tape, int(order), int(nvars), int(varorder)
        desc, keep = tape.c_desc()
        ode = lib.tb_ode_create(C.byref(desc))
        if not ode:
            raise L.TaylorB200Error("tb_ode_create: " + L.last_error())
        self._t = lib.tb_tn_create(ode, self.order, self.nvars, self.varorder, mode)
        err = L.last_error()
        lib.tb_ode_free(ode)
        del keep
        if not self._t:
            raise L.TaylorB200Error("tb_tn_create: " + err)
        self.ncoef = lib.tb_tn_ncoef(self._t)
        self.monomials = np.zeros((self.ncoef, self.nvars), dtype=np.int32)
        lib.tb_tn_monomials(self._t, _ptr(self.monomials))

    def __del__(self):
        try:
            if getattr(self, "_t", None):
                L.lib().tb_tn_free(self._t)
        except Exception:
            pass

    def initial(self, q0: np.ndarray) -> np.ndarray:
        """`q0 + xi`: constant term q0[c, i], and variable c (for c < nvars) added to component c."""
        q0 = np.asarray(q0, dtype=np.float64)
        q0 = q0[:, None] if q0.ndim == 1 else q0
        dof, N = q0.shape
        p = np.zeros((dof, self.ncoef, N))
        p[:, 0, :] = q0
        for c in range(min(dof, self.nvars)):
            e = np.zeros(self.nvars, dtype=np.int32)
            e[c] = 1
            p[c, int(np.flatnonzero((self.monomials == e).all(1))[0]), :] = 1.0
        return p

    def integrate(self, q0poly, t0=None, tmax=None, abstol=1e-20, params=None, *, trange=None, maxsteps=500, reltol=0.0,
                  minstepsize=0.0, maxstepsize=np.inf) -> JetSolution:
        q = np.ascontiguousarray(q0poly, dtype=np.float64)
        if q.ndim == 2:
            q = np.ascontiguousarray(q[:, :, None])
        dof, nc, N = q.shape
        if dof != self.tape.dof or nc != self.ncoef:
            raise ValueError(f"q0 must be [dof={self.tape.dof}, ncoef={self.ncoef}, N]")
        par, per = _params(self.tape, params, N)
        o = make_opts(abstol=abstol, reltol=reltol, minstepsize=minstepsize, maxstepsize=maxstepsize, maxsteps=maxsteps,
                      params_per_traj=per)
        sol = JetSolution(np.zeros((dof, nc, N)), np.zeros(N), np.zeros(N, dtype=np.int64), np.zeros(N, dtype=np.int32),
                          self.monomials)
        out = L.TbTnOut()
        out.xfinal, out.tfinal = sol.xfinal.ctypes.data_as(L._DP), sol.tfinal.ctypes.data_as(L._DP)
        out.nsteps = sol.nsteps.ctypes.data_as(C.POINTER(C.c_int64))
        out.status = sol.status.ctypes.data_as(C.POINTER(C.c_int32))
        tr = None
        if trange is not None:
            tr = np.ascontiguousarray(trange, dtype=np.float64)
            sol.x_trange = np.zeros((dof, nc, tr.size, N))
            out.xv_trange = sol.x_trange.ctypes.data_as(L._DP)
            t0, tmax = tr[0], tr[-1]
        L.check(L.lib().tb_tn_integrate(self._t, N, _ptr(q), _ptr(par), float(t0), float(tmax), _ptr(tr),
                                        0 if tr is None else tr.size, C.byref(o), C.byref(out)), "tb_tn_integrate")
        return sol


def evaluate(coefs: np.ndarray, monomials: np.ndarray, xi) -> np.ndarray:
    """sum_m coefs[:, m, ...] * prod_v xi[v]**monomials[m, v] along axis 1."""
    xi = np.asarray(xi, dtype=np.float64)
    w = np.prod(xi[None, :] ** monomials, axis=1)
    return np.tensordot(w, coefs, axes=([0], [1]))


def taylorinteg_jet(f, q0poly, t0, tmax, order, abstol, params=None, *, nvars: int, varorder: int, npar: int = 0,
                    mode="parity", **kw) -> JetSolution:
    """`taylorinteg(f!, q0 + xi, t0, tmax, order, abstol, params; ...)` for TaylorN `q0` (taylorinteg.jl:306-381)."""
    tape = f if isinstance(f, Tape) else taylorize(f, np.shape(q0poly)[0], npar)
    jt = JetTransport(tape, order, nvars, varorder, L.MODE_PARITY if mode == "parity" else L.MODE_FAST)
    return jt.integrate(q0poly, t0, tmax, abstol, params, **kw)
