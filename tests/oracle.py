"""ctypes wrapper of oracle/libtb_oracle.so -- TEST INFRASTRUCTURE ONLY.

Imported by tests/, bench.py's CPU-baseline legs and __graft_entry__.smoke(); never by the product package.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

import taylorintegration_b200 as tb
from taylorintegration_b200 import _lib as L

_ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
_SO = os.path.join(_ROOT, "oracle", "libtb_oracle.so")
MATH_TB, MATH_LIBM = 0, 1
_lib = None


def build() -> None:
    subprocess.check_call(["make", "-s", "-C", os.path.join(_ROOT, "oracle")])


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        if not os.path.exists(_SO):
            build()
        o = C.CDLL(_SO)
        vp = C.c_void_p
        o.tbo_jet.argtypes = [C.POINTER(L.TbOdeDesc), C.c_int, C.c_int, vp, vp, C.c_double, C.POINTER(L.TbOpts), vp, vp, vp]
        o.tbo_series.argtypes = [C.POINTER(L.TbOdeDesc), C.c_int, C.c_int, vp, vp, C.c_double, C.c_int32, C.c_int, vp]
        o.tbo_integrate.argtypes = [C.POINTER(L.TbOdeDesc), C.c_int, C.c_int, C.c_int, C.c_int, C.c_int64, vp, vp,
                                    C.c_double, C.c_double, vp, C.c_int64, C.POINTER(L.TbOpts), C.POINTER(L.TbOut)]
        for f in ("tbo_math_pow", "tbo_math_exp", "tbo_math_log"):
            getattr(o, f).restype = C.c_double
        o.tbo_math_pow.argtypes = [C.c_double, C.c_double]
        o.tbo_math_exp.argtypes = [C.c_double]
        o.tbo_math_log.argtypes = [C.c_double]
        o.tbo_math_sincos.argtypes = [C.c_double, L._DP, L._DP]
        _lib = o
    return _lib


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def _opts(**kw) -> L.TbOpts:
    o = L.TbOpts()
    o.abstol, o.reltol, o.minstepsize, o.maxstepsize = kw.get("abstol", 1e-20), kw.get("reltol", 0.0), \
        kw.get("minstepsize", 0.0), kw.get("maxstepsize", np.inf)
    o.maxsteps, o.eventorder, o.newtoniter = kw.get("maxsteps", 500), kw.get("eventorder", 0), kw.get("newtoniter", 10)
    o.nrabstol = kw.get("nrabstol", float(np.finfo(float).eps))
    o.event_capacity, o.params_per_traj, o.steps_stride = kw.get("event_capacity", 0), kw.get("params_per_traj", 0), 0
    return o


def jet(tape: tb.Tape, order: int, q0, params=None, t0=0.0, math=MATH_TB, **kw):
    """One taylorstep!: (xjet[dof, order+1], dxjet[dof, order+1], dt)."""
    d, keep = tape.c_desc()
    q0 = np.ascontiguousarray(q0, dtype=np.float64)
    par = np.ascontiguousarray(np.atleast_1d(params if params is not None else [0.0]), dtype=np.float64)
    xj, dxj = np.zeros((tape.dof, order + 1)), np.zeros((tape.dof, order + 1))
    dt = C.c_double(0.0)
    o = _opts(**kw)
    rc = lib().tbo_jet(C.byref(d), order, math, _p(q0), _p(par), t0, C.byref(o), _p(xj), _p(dxj), C.byref(dt))
    assert rc == 0, rc
    del keep
    return xj, dxj, dt.value


def series(tape: tb.Tape, order: int, q0, params, operand: int, with_jac=False, t0=0.0, math=MATH_TB):
    d, keep = tape.c_desc()
    q0 = np.ascontiguousarray(q0, dtype=np.float64)
    par = np.ascontiguousarray(np.atleast_1d(params if params is not None else [0.0]), dtype=np.float64)
    out = np.zeros(order + 1)
    rc = lib().tbo_series(C.byref(d), order, math, _p(q0), _p(par), t0, operand, int(with_jac), _p(out))
    assert rc == 0, rc
    del keep
    return out


def integrate(tape: tb.Tape, order: int, q0, t0=0.0, tmax=None, params=None, *, kind=L.KIND_PLAIN, trange=None,
              steps=False, dense=False, math=MATH_TB, nthreads=1, lyap_state0=None, **kw) -> tb.EnsembleSolution:
    """Ensemble run on the CPU oracle with the layout of tb_integrate."""
    d, keep = tape.c_desc()
    q = np.asarray(q0, dtype=np.float64)
    q = np.ascontiguousarray(q[:, None] if q.ndim == 1 else q)
    dof, N = q.shape
    per = 0
    par = None
    if tape.npar:
        par = np.ascontiguousarray(np.atleast_1d(np.asarray(params, dtype=np.float64)))
        per = 1 if par.ndim == 2 else 0
    else:
        par = np.zeros(1)
    maxsteps = kw.get("maxsteps", 500)
    cap = kw.get("event_capacity", 0) or maxsteps + 1
    o = _opts(**dict(kw, params_per_traj=per, event_capacity=cap))
    stride = maxsteps + 1
    out = L.TbOut()
    sol = tb.EnsembleSolution(np.zeros((dof, N)), np.zeros(N), np.zeros(N, dtype=np.int64), np.zeros(N, dtype=np.int32))
    out.xfinal, out.tfinal = sol.xfinal.ctypes.data_as(L._DP), sol.tfinal.ctypes.data_as(L._DP)
    out.nsteps = sol.nsteps.ctypes.data_as(C.POINTER(C.c_int64))
    out.status = sol.status.ctypes.data_as(C.POINTER(C.c_int32))
    tr = None
    if trange is not None:
        tr = np.ascontiguousarray(trange, dtype=np.float64)
        sol.x_trange = np.zeros((dof, tr.size, N))
        out.xv_trange = sol.x_trange.ctypes.data_as(L._DP)
        t0, tmax = tr[0], tr[-1]
        if kind == L.KIND_LYAP:
            sol.lam_trange = np.zeros((dof, tr.size, N))
            out.lambda_v = sol.lam_trange.ctypes.data_as(L._DP)
    elif steps:
        sol.t, sol.x = np.zeros((stride, N)), np.zeros((dof, stride, N))
        out.tv, out.xv = sol.t.ctypes.data_as(L._DP), sol.x.ctypes.data_as(L._DP)
        if dense:
            sol.p = np.zeros((dof, order + 1, stride, N))
            out.psol = sol.p.ctypes.data_as(L._DP)
        if kind == L.KIND_LYAP:
            sol.lam_steps = np.zeros((dof, stride, N))
            out.lambda_v = sol.lam_steps.ctypes.data_as(L._DP)
    if kind == L.KIND_EVENTS:
        sol.tevents, sol.gresids = np.zeros((cap, N)), np.zeros((cap, N))
        sol.xevents, sol.nevents = np.zeros((dof, cap, N)), np.zeros(N, dtype=np.int32)
        out.tevents, out.gresids = sol.tevents.ctypes.data_as(L._DP), sol.gresids.ctypes.data_as(L._DP)
        out.xevents = sol.xevents.ctypes.data_as(L._DP)
        out.nevents = sol.nevents.ctypes.data_as(C.POINTER(C.c_int32))
    if kind == L.KIND_LYAP:
        sol.lam = np.zeros((dof, N))
        out.lambda_ = sol.lam.ctypes.data_as(L._DP)
        if trange is None:
            sol.lyap_state = np.zeros((dof * dof + dof + 1, N))
            out.lyap_state = sol.lyap_state.ctypes.data_as(L._DP)
            if lyap_state0 is not None:
                ls0 = np.ascontiguousarray(lyap_state0, dtype=np.float64).reshape(dof * dof + dof + 1, N)
                out.lyap_state0 = ls0.ctypes.data_as(L._DP)
    rc = lib().tbo_integrate(C.byref(d), order, kind, math, nthreads, N, _p(q), _p(par), float(t0), float(tmax),
                             _p(tr), 0 if tr is None else tr.size, C.byref(o), C.byref(out))
    assert rc == 0, rc
    del keep
    return sol


# ---- TaylorN jet transport oracle (oracle/libtb_tn_oracle.so) -----------------------------------------------
_SO_TN = os.path.join(_ROOT, "oracle", "libtb_tn_oracle.so")
_lib_tn = None


def lib_tn() -> C.CDLL:
    global _lib_tn
    if _lib_tn is None:
        if not os.path.exists(_SO_TN):
            build()
        o = C.CDLL(_SO_TN)
        vp = C.c_void_p
        o.tbo_tn_monomials.argtypes = [C.c_int, C.c_int, vp]
        o.tbo_tn_integrate.argtypes = [C.POINTER(L.TbOdeDesc), C.c_int, C.c_int, C.c_int, C.c_int64, vp, vp, C.c_double,
                                       C.c_double, vp, C.c_int64, C.POINTER(L.TbOpts), vp, vp, vp, vp, vp]
        _lib_tn = o
    return _lib_tn


def tn_monomials(nvars: int, varorder: int) -> np.ndarray:
    n = lib_tn().tbo_tn_monomials(nvars, varorder, None)
    e = np.zeros((n, nvars), dtype=np.int32)
    lib_tn().tbo_tn_monomials(nvars, varorder, _p(e))
    return e


def integrate_tn(tape: tb.Tape, order: int, nvars: int, varorder: int, q0, t0=0.0, tmax=None, params=None, *,
                 trange=None, **kw):
    """Jet-transport ensemble on the CPU oracle.  q0 [dof, ncoef, N]; returns dict(xfinal, tfinal, nsteps, status[, x_trange])."""
    d, keep = tape.c_desc()
    q = np.ascontiguousarray(q0, dtype=np.float64)
    dof, nc, N = q.shape
    per = 0
    par = np.zeros(1)
    if tape.npar:
        par = np.ascontiguousarray(np.atleast_1d(np.asarray(params, dtype=np.float64)))
        per = 1 if par.ndim == 2 else 0
    o = _opts(**dict(kw, params_per_traj=per))
    res = dict(xfinal=np.zeros((dof, nc, N)), tfinal=np.zeros(N), nsteps=np.zeros(N, dtype=np.int64),
               status=np.zeros(N, dtype=np.int32))
    tr = None
    if trange is not None:
        tr = np.ascontiguousarray(trange, dtype=np.float64)
        res["x_trange"] = np.zeros((dof, nc, tr.size, N))
        t0, tmax = tr[0], tr[-1]
    rc = lib_tn().tbo_tn_integrate(C.byref(d), order, nvars, varorder, N, _p(q), _p(par), float(t0), float(tmax), _p(tr),
                                   0 if tr is None else tr.size, C.byref(o), _p(res["xfinal"]), _p(res["tfinal"]),
                                   _p(res["nsteps"]), _p(res["status"]), _p(res.get("x_trange")))
    assert rc == 0, rc
    del keep
    return res


# ---- generic (parse_eqs = false) path: oracle/libtb_generic.so ---------------------------------------------------
_SO_GEN = os.path.join(_ROOT, "oracle", "libtb_generic.so")
_lib_gen = None


def lib_generic() -> C.CDLL:
    global _lib_gen
    if _lib_gen is None:
        if not os.path.exists(_SO_GEN):
            build()
        o = C.CDLL(_SO_GEN)
        vp = C.c_void_p
        o.tbg_problem_name.restype = C.c_char_p
        o.tbg_jet.argtypes = [C.c_int, C.c_int, vp, vp, C.c_double, vp]
        o.tbg_integrate.argtypes = [C.c_int, C.c_int, vp, vp, C.c_double, C.c_double, C.POINTER(L.TbOpts), vp, vp, vp, vp,
                                    vp, vp]
        _lib_gen = o
    return _lib_gen


def generic_problem(name: str) -> int:
    g = lib_generic()
    names = [g.tbg_problem_name(i).decode() for i in range(g.tbg_problem_count())]
    return names.index(name)


def generic_jet(name: str, order: int, q0, params=None, t0=0.0) -> np.ndarray:
    """jets [dof, order+1] from the reference's generic jetcoeffs! (src/integrator/jetcoeffs.jl:66-92)."""
    pid = generic_problem(name)
    q0 = np.ascontiguousarray(q0, dtype=np.float64)
    par = np.ascontiguousarray(np.atleast_1d(params if params is not None else [0.0]), dtype=np.float64)
    xj = np.zeros((q0.shape[0], order + 1))
    rc = lib_generic().tbg_jet(pid, order, _p(q0), _p(par), float(t0), _p(xj))
    assert rc == 0, rc
    return xj


def generic_integrate(name: str, order: int, q0, t0, tmax, params=None, **kw):
    """One trajectory through taylorinteg with parse_eqs = false; returns (tv[n+1], xv[dof, n+1], status)."""
    pid = generic_problem(name)
    q0 = np.ascontiguousarray(q0, dtype=np.float64)
    par = np.ascontiguousarray(np.atleast_1d(params if params is not None else [0.0]), dtype=np.float64)
    o = _opts(**kw)
    stride = o.maxsteps + 1
    dof = q0.shape[0]
    tv, xv = np.zeros(stride), np.zeros((dof, stride))
    xf, tf = np.zeros(dof), np.zeros(1)
    ns, st = np.zeros(1, dtype=np.int64), np.zeros(1, dtype=np.int32)
    rc = lib_generic().tbg_integrate(pid, order, _p(q0), _p(par), float(t0), float(tmax), C.byref(o), _p(tv), _p(xv),
                                     _p(xf), _p(tf), _p(ns), _p(st))
    assert rc == 0, rc
    n = int(ns[0])
    return tv[: n + 1].copy(), xv[:, : n + 1].copy(), int(st[0])
