"""Public interface: `taylorinteg`, `lyap_taylorinteg`, `TaylorSolution` -- the reference's exports
(src/TaylorIntegration.jl:14-16) with the same positional arguments and keyword names, batched over
trajectories and executed by the fused sm_100a kernels of libtaylorb200.so.

    taylorinteg(f, q0, t0, tmax, order, abstol, params=None; maxsteps=500, dense=True, reltol=0, ...)
        src/integrator/taylorinteg.jl:306-381
    taylorinteg(f, q0, trange, order, abstol, params=None; ...)          taylorinteg.jl:558-642
    taylorinteg(f, g, q0, t0, tmax, order, abstol, params=None; eventorder=0, newtoniter=10, nrabstol=eps)
        src/rootfinding.jl:301-438
    lyap_taylorinteg(f, q0, t0, tmax, order, abstol, params=None; ...)   src/lyapunovspectrum.jl:293-429

`q0` is either one initial condition `[dof]` (the reference's call; returns a `TaylorSolution`) or an ensemble
`[dof, N]` (returns an `EnsembleSolution`).  `params` is `[npar]` (shared) or `[npar, N]` (swept).  `f` is either
a Python function `f(dx, x, p, t)` (traced by `taylorize`, the `@taylorize` counterpart) or a `Tape`.
"""
from __future__ import annotations

import ctypes as C
import warnings
from dataclasses import dataclass
from typing import Callable, Dict, Optional, Tuple, Union

import numpy as np

from . import _lib as L
from .tape import Tape, lyap_column_tape, taylorize

_EPS = float(np.finfo(np.float64).eps)


def _ptr(a: Optional[np.ndarray]):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


class Kernel:
    """A compiled ensemble kernel (one per ODE x order x mode x kind)."""

    def __init__(self, tape: Tape, order: int, kind: int = L.KIND_PLAIN, mode: int = L.MODE_PARITY,
                 block_threads: int = 0, flags: int = 0, reg_budget: int = 0, min_blocks_per_sm: int = 0,
                 tmem_budget: int = 0):
        lib = L.lib()
        self.tape, self.order, self.kind, self.mode = tape, int(order), kind, mode
        desc, keep = tape.c_desc()
        self._ode = lib.tb_ode_create(C.byref(desc))
        if not self._ode:
            raise L.TaylorB200Error("tb_ode_create: " + L.last_error())
        co = L.TbCompileOpts()
        lib.tb_compile_opts_default(C.byref(co))
        co.order, co.mode, co.kind, co.block_threads = self.order, mode, kind, block_threads
        co.flags, co.reg_budget, co.min_blocks_per_sm, co.tmem_budget = flags, reg_budget, min_blocks_per_sm, tmem_budget
        self._k = lib.tb_compile(self._ode, C.byref(co))
        del keep
        if not self._k:
            err = L.last_error()
            lib.tb_ode_free(self._ode)
            self._ode = None
            raise L.TaylorB200Error("tb_compile: " + err)

    def __del__(self):
        try:
            lib = L.lib()
            if getattr(self, "_k", None):
                lib.tb_kernel_free(self._k)
            if getattr(self, "_ode", None):
                lib.tb_ode_free(self._ode)
        except Exception:
            pass

    @property
    def handle(self):
        return self._k

    @property
    def source(self) -> str:
        return L.lib().tb_kernel_source(self._k).decode()

    @property
    def log(self) -> str:
        return L.lib().tb_kernel_log(self._k).decode()

    def info(self) -> Dict[str, int]:
        v = [C.c_int(-1) for _ in range(5)]
        L.lib().tb_kernel_info(self._k, *[C.byref(x) for x in v])
        return dict(zip(["regs", "smem_bytes", "local_bytes", "block_threads", "blocks_per_sm"], [x.value for x in v]))

    def cubin(self) -> bytes:
        n = C.c_size_t(0)
        p = L.lib().tb_kernel_cubin(self._k, C.byref(n))
        return C.string_at(p, n.value)

    def jet(self, q0: np.ndarray, params=None, t0: float = 0.0, abstol: float = 1e-20, reltol: float = 0.0):
        """First `taylorstep!` of every trajectory: jets [dof, order+1, N] and step sizes [N]."""
        q0 = np.ascontiguousarray(q0, dtype=np.float64)
        if q0.ndim == 1:
            q0 = q0[:, None]
        dof, N = q0.shape
        par, per = _params(self.tape, params, N)
        o = make_opts(abstol=abstol, reltol=reltol, params_per_traj=per)
        xj = np.empty((dof, self.order + 1, N))
        dt = np.empty(N)
        L.check(L.lib().tb_jet(self._k, N, _ptr(q0), _ptr(par), float(t0), C.byref(o), _ptr(xj), _ptr(dt)), "tb_jet")
        return xj, dt


class WideKernel:
    """Large-state program (one CTA per trajectory, csrc/tb_wide.cu): tapes with more than 16 components or more
    than 200 entries, e.g. the many-spin system.  Plain drivers only (t0..tmax and trange)."""

    def __init__(self, tape: Tape, order: int, mode: int = L.MODE_PARITY):
        lib = L.lib()
        self.tape, self.order, self.mode = tape, int(order), mode
        desc, keep = tape.c_desc()
        ode = lib.tb_ode_create(C.byref(desc))
        if not ode:
            raise L.TaylorB200Error("tb_ode_create: " + L.last_error())
        self._w = lib.tb_wide_create(ode, self.order, mode)
        err = L.last_error()
        lib.tb_ode_free(ode)
        del keep
        if not self._w:
            raise L.TaylorB200Error("tb_wide_create: " + err)

    def __del__(self):
        try:
            if getattr(self, "_w", None):
                L.lib().tb_wide_free(self._w)
        except Exception:
            pass

    @property
    def handle(self):
        return self._w

    def info(self) -> Dict[str, int]:
        v = [C.c_int(-1) for _ in range(4)]
        L.lib().tb_wide_info(self._w, *[C.byref(x) for x in v])
        d = dict(zip(["entries", "levels", "series", "max_width"], [x.value for x in v]))
        v2 = [C.c_int(-1) for _ in range(7)]
        if L.lib().tb_wide_layout(self._w, *[C.byref(x) for x in v2]) == 0:  # the shared-memory cluster kernel applies
            d.update(zip(["clusters", "stored_series", "scalar_slots", "threads", "smem_bytes", "ctas_per_sm", "groups"],
                         [x.value for x in v2]))
            v3 = [C.c_int(0) for _ in range(4)]
            if L.lib().tb_wide_generated(self._w, *[C.byref(x) for x in v3]) == 0:
                d.update(zip(["generated", "regs", "local_bytes", "coupling_in_smem"], [x.value for x in v3]))
        return d

    @property
    def source(self) -> str:
        """Text of the NVRTC-specialised cluster kernel ('' when the interpreter runs)."""
        s = L.lib().tb_wide_source(self._w)
        return s.decode() if s else ""

    @property
    def log(self) -> str:
        s = L.lib().tb_wide_log(self._w)
        return s.decode() if s else ""


def is_wide(tape: Tape) -> bool:
    return tape.dof > 16 or len(tape.ops) > 200


_WIDE: Dict[Tuple, WideKernel] = {}


def get_wide_kernel(tape: Tape, order: int, mode: int) -> WideKernel:
    key = (id(tape), order, mode)
    k = _WIDE.get(key)
    if k is None:
        k = _WIDE[key] = WideKernel(tape, order, mode)
    return k


_KERNELS: Dict[Tuple, Kernel] = {}
_TAPES: Dict[Tuple, Tape] = {}


_COLUMN_TAPES: Dict[int, Tuple[Tape, Optional[Tape]]] = {}


def _column_tape(tape: Tape) -> Optional[Tape]:
    ent = _COLUMN_TAPES.get(id(tape))
    if ent is None or ent[0] is not tape:
        ent = _COLUMN_TAPES[id(tape)] = (tape, lyap_column_tape(tape))
    return ent[1]


def get_kernel(tape: Tape, order: int, kind: int, mode: int, **kw) -> Kernel:
    key = (id(tape), order, kind, mode, tuple(sorted(kw.items())))
    k = _KERNELS.get(key)
    if k is None:
        k = _KERNELS[key] = Kernel(tape, order, kind, mode, **kw)
    return k


def make_opts(abstol=1e-20, reltol=0.0, minstepsize=0.0, maxstepsize=np.inf, maxsteps=500, eventorder=0,
              newtoniter=10, nrabstol=_EPS, event_capacity=0, params_per_traj=0, steps_stride=0, sort_by_cost=0) -> L.TbOpts:
    o = L.TbOpts()
    L.lib().tb_opts_default(C.byref(o))
    o.abstol, o.reltol, o.minstepsize, o.maxstepsize = float(abstol), float(reltol), float(minstepsize), float(maxstepsize)
    o.maxsteps, o.eventorder, o.newtoniter, o.nrabstol = int(maxsteps), int(eventorder), int(newtoniter), float(nrabstol)
    o.event_capacity, o.params_per_traj, o.steps_stride = int(event_capacity), int(params_per_traj), int(steps_stride)
    o.sort_by_cost = int(sort_by_cost)
    return o


def _params(tape: Tape, params, N: int):
    if tape.npar == 0:
        return None, 0
    if params is None:
        raise ValueError(f"the ODE has {tape.npar} parameter(s) but params is None")
    p = np.ascontiguousarray(np.atleast_1d(np.asarray(params, dtype=np.float64)))
    if p.ndim == 1:
        if p.shape[0] != tape.npar:
            raise ValueError(f"params has {p.shape[0]} entries, the ODE has {tape.npar}")
        return p, 0
    if p.shape != (tape.npar, N):
        raise ValueError(f"params must be [npar] or [npar, N]; got {p.shape}")
    return p, 1


def _as_tape(f, dof: int, params, g=None, jacobian=False) -> Tape:
    if isinstance(f, Tape):
        return f
    npar = 0 if params is None else int(np.atleast_1d(np.asarray(params)).shape[0])
    key = (f, g, dof, npar, jacobian)
    t = _TAPES.get(key)
    if t is None:
        t = _TAPES[key] = taylorize(f, dof, npar, event=g, jacobian=jacobian)
    return t


def _mode(mode) -> int:
    return {"parity": L.MODE_PARITY, "fast": L.MODE_FAST, L.MODE_PARITY: L.MODE_PARITY, L.MODE_FAST: L.MODE_FAST}[mode]


@dataclass
class TaylorSolution:
    """One trajectory, laid out like the reference's `TaylorSolution` (src/integrator/taylorsolution.jl:21-63):
    `t[nsteps+1]`, `x[nsteps+1, dof]`, `p[nsteps, dof, order+1]` (dense), events, Lyapunov spectrum."""
    t: np.ndarray
    x: np.ndarray
    p: Optional[np.ndarray] = None
    tevents: Optional[np.ndarray] = None
    xevents: Optional[np.ndarray] = None
    gresids: Optional[np.ndarray] = None
    lam: Optional[np.ndarray] = None  # the reference's field `λ`
    status: int = 0

    def __call__(self, tq):
        """Dense output `sol(t)` (taylorsolution.jl:337-423), evaluated on the GPU by tb_dense_eval."""
        if self.p is None:
            raise ValueError("dense output was not requested (dense=False)")
        tq = np.atleast_1d(np.asarray(tq, dtype=np.float64))
        ns, dof, np1 = self.p.shape
        stride = ns + 1
        tv = np.ascontiguousarray(self.t.reshape(stride, 1))
        psol = np.zeros((dof, np1, stride, 1))
        psol[:, :, :ns, 0] = np.transpose(self.p, (1, 2, 0))
        nsteps = np.array([ns], dtype=np.int64)
        xq = np.empty((dof, tq.size, 1))
        L.check(L.lib().tb_dense_eval(1, dof, np1 - 1, stride, _ptr(tv), _ptr(psol), _ptr(nsteps), _ptr(tq), tq.size,
                                      _ptr(xq)), "tb_dense_eval")
        return xq[:, :, 0].T


@dataclass
class EnsembleSolution:
    """N trajectories, structure-of-arrays with the trajectory index last (the device layout)."""
    xfinal: np.ndarray            # [dof, N]
    tfinal: np.ndarray            # [N]
    nsteps: np.ndarray            # [N]
    status: np.ndarray            # [N] TB_ST_* flags
    t: Optional[np.ndarray] = None        # [stride, N]
    x: Optional[np.ndarray] = None        # [dof, stride, N]
    p: Optional[np.ndarray] = None        # [dof, order+1, stride, N]
    x_trange: Optional[np.ndarray] = None  # [dof, nt, N]
    tevents: Optional[np.ndarray] = None  # [cap, N]
    xevents: Optional[np.ndarray] = None  # [dof, cap, N]
    gresids: Optional[np.ndarray] = None  # [cap, N]
    nevents: Optional[np.ndarray] = None  # [N]
    lam: Optional[np.ndarray] = None      # [dof, N]
    lam_steps: Optional[np.ndarray] = None  # [dof, stride, N]
    lam_trange: Optional[np.ndarray] = None  # [dof, nt, N]
    lyap_state: Optional[np.ndarray] = None  # [dof*dof + dof + 1, N]: Q (column-major), running sums, t00 (resume)

    def trajectory(self, i: int) -> TaylorSolution:
        ns = int(self.nsteps[i])
        if self.x_trange is not None:
            s = TaylorSolution(t=None, x=self.x_trange[:, :, i].T.copy(), status=int(self.status[i]))
            if self.lam_trange is not None:
                s.lam = self.lam_trange[:, :, i].T.copy()
            if self.nevents is not None:
                ne = int(self.nevents[i])
                s.tevents, s.gresids = self.tevents[:ne, i].copy(), self.gresids[:ne, i].copy()
                s.xevents = self.xevents[:, :ne, i].T.copy()
            return s
        if self.t is None:
            raise ValueError("per-step output was not requested (output='final')")
        s = TaylorSolution(t=self.t[: ns + 1, i].copy(), x=self.x[:, : ns + 1, i].T.copy(), status=int(self.status[i]))
        if self.p is not None:
            s.p = np.transpose(self.p[:, :, :ns, i], (2, 0, 1)).copy()
        if self.nevents is not None:
            ne = int(self.nevents[i])
            s.tevents = self.tevents[:ne, i].copy()
            s.xevents = self.xevents[:, :ne, i].T.copy()
            s.gresids = self.gresids[:ne, i].copy()
        if self.lam_steps is not None:
            s.lam = self.lam_steps[:, : ns + 1, i].T.copy()
        return s


def _warn_status(status: np.ndarray) -> None:
    """Re-emit the reference's warnings (taylorinteg.jl:360-363, 372-377; rootfinding.jl:222-224)."""
    if np.any(status & L.ST_ZERO_STEP):
        warnings.warn("The step-size is zero; aborting integration.")
    if np.any(status & L.ST_MAXSTEPS):
        warnings.warn("Maximum number of integration steps reached; exiting.")
    if np.any(status & L.ST_NR_NOCONV):
        warnings.warn("Newton-Raphson did not converge for prescribed tolerance and maximum allowed iterations.")


def _integrate(tape: Tape, kind: int, q0, t0, tmax, trange, order, abstol, params, *, maxsteps, dense, reltol,
               minstepsize, maxstepsize, eventorder=0, newtoniter=10, nrabstol=_EPS, event_capacity=0, mode="parity",
               output=None, kernel_opts=None, devices=None, block=4096, lyap_state0=None):
    q0 = np.asarray(q0, dtype=np.float64)
    single = q0.ndim == 1
    q = np.ascontiguousarray(q0[:, None] if single else q0)
    dof, N = q.shape
    if dof != tape.dof:
        raise ValueError(f"q0 has {dof} components, the ODE has {tape.dof}")
    if kind == L.KIND_EVENTS and order < eventorder:
        raise AssertionError("order ≥ eventorder")  # rootfinding.jl:321
    par, per = _params(tape, params, N)
    if output is None:
        output = "steps" if single else "final"
    if output not in ("steps", "final"):
        raise ValueError("output must be 'steps' or 'final'")
    kernel_opts = dict(kernel_opts or {})
    # scheduling hint (tb_opts_t.sort_by_cost): 1 = the ensemble's cost grows with the index, hand out the far end first
    sort_by_cost = int(kernel_opts.pop("sort_by_cost", 0))
    wide = bool(kernel_opts.pop("wide", False)) or is_wide(tape)
    kkind = kind
    if wide:
        if kind != L.KIND_PLAIN:
            raise L.TaylorB200Error("large-state tapes support the plain drivers only (no events / Lyapunov yet)")
        if single and output == "steps" and trange is None:
            raise L.TaylorB200Error("large-state tapes return final states or trange outputs; pass q0 as [dof, N] "
                                    "or a trange")
        k = get_wide_kernel(tape, int(order), _mode(mode))
    else:
        ktape, kkind = tape, kind
        columns = kernel_opts.pop("columns", True)
        if kind == L.KIND_LYAP and columns:
            # column-parallel Lyapunov kernel (a team of dof lanes per trajectory, TB_KIND_LYAPCOL) when the
            # Jacobian can be written as a column tape with the same bits; else the single-thread kernel
            col = _column_tape(tape)
            if col is not None:
                ktape, kkind = col, L.KIND_LYAPCOL
                if trange is not None:  # the trange variant is a separate build of the kernel (TB_CF_LYAP_TRANGE)
                    kernel_opts["flags"] = int(kernel_opts.get("flags", 0)) | 131072
        if kind == L.KIND_PLAIN and trange is None and output == "final" and not kernel_opts.pop("full_outputs", False):
            # final states only: the build without output code in the step body (TB_CF_FINAL_ONLY), +4 % on Kepler-28
            kernel_opts["flags"] = int(kernel_opts.get("flags", 0)) | 262144
        k = get_kernel(ktape, int(order), kkind, _mode(mode), **(kernel_opts or {}))
    cap = int(event_capacity) if event_capacity else int(maxsteps) + 1
    o = make_opts(abstol, reltol, minstepsize, maxstepsize, maxsteps, eventorder, newtoniter, nrabstol, cap, per,
                  sort_by_cost=sort_by_cost)
    stride = int(maxsteps) + 1
    out = L.TbOut()
    xfinal, tfinal = np.empty((dof, N)), np.empty(N)
    nsteps, status = np.empty(N, dtype=np.int64), np.empty(N, dtype=np.int32)
    out.xfinal, out.tfinal = xfinal.ctypes.data_as(L._DP), tfinal.ctypes.data_as(L._DP)
    out.nsteps = nsteps.ctypes.data_as(C.POINTER(C.c_int64))
    out.status = status.ctypes.data_as(C.POINTER(C.c_int32))
    sol = EnsembleSolution(xfinal, tfinal, nsteps, status)
    lib = L.lib()
    if trange is not None:
        tr = np.ascontiguousarray(trange, dtype=np.float64)
        d = np.diff(tr)
        assert tr.size >= 2 and (np.all(d > 0) or np.all(d < 0)), "trange must be sorted"  # taylorinteg.jl:489,573
        sol.x_trange = np.empty((dof, tr.size, N))
        out.xv_trange = sol.x_trange.ctypes.data_as(L._DP)
        if kind == L.KIND_LYAP:  # lyap_taylorinteg(f!, q0, trange, ...): exponents at every output time
            if kkind != L.KIND_LYAPCOL:
                raise L.TaylorB200Error("the trange variant of lyap_taylorinteg needs the column-parallel kernel")
            sol.lam_trange, sol.lam = np.empty((dof, tr.size, N)), np.empty((dof, N))
            out.lambda_v, out.lambda_ = sol.lam_trange.ctypes.data_as(L._DP), sol.lam.ctypes.data_as(L._DP)
        if kind == L.KIND_EVENTS:  # taylorinteg(f!, g, q0, trange, ...): src/rootfinding.jl:440-584
            sol.tevents, sol.gresids = np.zeros((cap, N)), np.zeros((cap, N))
            sol.xevents, sol.nevents = np.zeros((dof, cap, N)), np.zeros(N, dtype=np.int32)
            out.tevents, out.gresids = sol.tevents.ctypes.data_as(L._DP), sol.gresids.ctypes.data_as(L._DP)
            out.xevents = sol.xevents.ctypes.data_as(L._DP)
            out.nevents = sol.nevents.ctypes.data_as(C.POINTER(C.c_int32))
        if wide:
            L.check(lib.tb_wide_integrate(k.handle, N, _ptr(q), _ptr(par), float(tr[0]), float(tr[-1]), _ptr(tr), tr.size,
                                          C.byref(o), C.byref(out)), "tb_wide_integrate")
        elif devices is not None:
            dv = np.ascontiguousarray(devices, dtype=np.int32)
            L.check(lib.tb_integrate_multi(k.handle, _ptr(dv), dv.size, int(block), N, _ptr(q), _ptr(par), float(tr[0]),
                                           float(tr[-1]), _ptr(tr), tr.size, C.byref(o), C.byref(out)), "tb_integrate_multi")
        else:
            L.check(lib.tb_integrate_trange(k.handle, N, _ptr(q), _ptr(par), _ptr(tr), tr.size, C.byref(o),
                                            C.byref(out)), "tb_integrate_trange")
    elif wide:
        L.check(lib.tb_wide_integrate(k.handle, N, _ptr(q), _ptr(par), float(t0), float(tmax), None, 0, C.byref(o),
                                      C.byref(out)), "tb_wide_integrate")
    else:
        if output == "steps":
            sol.t, sol.x = np.empty((stride, N)), np.empty((dof, stride, N))
            out.tv, out.xv = sol.t.ctypes.data_as(L._DP), sol.x.ctypes.data_as(L._DP)
            if dense and kind != L.KIND_LYAP:
                sol.p = np.empty((dof, order + 1, stride, N))
                out.psol = sol.p.ctypes.data_as(L._DP)
            if kind == L.KIND_LYAP:
                sol.lam_steps = np.empty((dof, stride, N))
                out.lambda_v = sol.lam_steps.ctypes.data_as(L._DP)
        if kind == L.KIND_EVENTS:
            sol.tevents, sol.gresids = np.empty((cap, N)), np.empty((cap, N))
            sol.xevents, sol.nevents = np.empty((dof, cap, N)), np.empty(N, dtype=np.int32)
            out.tevents, out.gresids = sol.tevents.ctypes.data_as(L._DP), sol.gresids.ctypes.data_as(L._DP)
            out.xevents = sol.xevents.ctypes.data_as(L._DP)
            out.nevents = sol.nevents.ctypes.data_as(C.POINTER(C.c_int32))
        if kind == L.KIND_LYAP:
            sol.lam = np.empty((dof, N))
            out.lambda_ = sol.lam.ctypes.data_as(L._DP)
            if devices is None:  # continue-from-state: the Lyapunov state of this call, and the one to resume from
                sol.lyap_state = np.empty((dof * dof + dof + 1, N))
                out.lyap_state = sol.lyap_state.ctypes.data_as(L._DP)
                if lyap_state0 is not None:
                    ls0 = np.ascontiguousarray(lyap_state0, dtype=np.float64).reshape(dof * dof + dof + 1, N)
                    out.lyap_state0 = ls0.ctypes.data_as(L._DP)
        if devices is not None:  # one ensemble over several GPUs of this process (tb_integrate_multi)
            dv = np.ascontiguousarray(devices, dtype=np.int32)
            L.check(lib.tb_integrate_multi(k.handle, _ptr(dv), dv.size, int(block), N, _ptr(q), _ptr(par), float(t0),
                                           float(tmax), None, 0, C.byref(o), C.byref(out)), "tb_integrate_multi")
        else:
            L.check(lib.tb_integrate(k.handle, N, _ptr(q), _ptr(par), float(t0), float(tmax), C.byref(o), C.byref(out)),
                    "tb_integrate")
    _warn_status(status)
    if single:
        return sol.trajectory(0)
    return sol


def taylorinteg(f, *args, maxsteps: int = 500, parse_eqs: bool = True, dense: bool = True, reltol: float = 0.0,
                minstepsize: float = 0.0, maxstepsize: float = np.inf, eventorder: int = 0, newtoniter: int = 10,
                nrabstol: float = _EPS, event_capacity: int = 0, mode="parity", output=None, kernel_opts=None,
                devices=None, block: int = 4096):
    """See the module docstring.  `devices=[0, 1, ...]` shards ONE ensemble over several GPUs of this process in
    round-robin blocks of `block` trajectories (tb_integrate_multi).  `parse_eqs` is accepted for signature compatibility: the tape path is the only
    path (there is no generic `jetcoeffs!` fallback on the GPU)."""
    g = None
    if len(args) >= 1 and (callable(args[0]) and not isinstance(args[0], (np.ndarray, Tape))):
        g, args = args[0], args[1:]
    trange = None
    if len(args) >= 2 and np.ndim(args[1]) >= 1:  # (q0, trange, order, abstol[, params])
        q0, trange, order, abstol = args[:4]
        params = args[4] if len(args) > 4 else None
        t0 = tmax = None
    else:
        q0, t0, tmax, order, abstol = args[:5]
        params = args[5] if len(args) > 5 else None
    dof = np.asarray(q0).shape[0]
    tape = _as_tape(f, dof, params, g=g)
    kind = L.KIND_EVENTS if g is not None else L.KIND_PLAIN
    return _integrate(tape, kind, q0, t0, tmax, trange, int(order), abstol, params, maxsteps=maxsteps, dense=dense,
                      reltol=reltol, minstepsize=minstepsize, maxstepsize=maxstepsize, eventorder=eventorder,
                      newtoniter=newtoniter, nrabstol=nrabstol, event_capacity=event_capacity, mode=mode, output=output,
                      kernel_opts=kernel_opts, devices=devices, block=block)


def taylorinteg_events(tape: Tape, q0, t0, tmax, order, abstol, params=None, trange=None, **kw):
    """Event-detecting driver for a pre-built `Tape` that already carries its event function; with `trange`
    (t0, tmax ignored) it is the variant of src/rootfinding.jl:440-584."""
    kw.setdefault("maxsteps", 500)
    kw.setdefault("dense", True)
    kw.setdefault("reltol", 0.0)
    kw.setdefault("minstepsize", 0.0)
    kw.setdefault("maxstepsize", np.inf)
    return _integrate(tape, L.KIND_EVENTS, q0, t0, tmax, trange, int(order), abstol, params, **kw)


def lyap_taylorinteg(f, q0, *args, maxsteps: int = 500, parse_eqs: bool = True, reltol: float = 0.0,
                     minstepsize: float = 0.0, maxstepsize: float = np.inf, mode="parity", output=None, kernel_opts=None,
                     devices=None, block: int = 4096, resume=None):
    """Lyapunov spectrum (src/lyapunovspectrum.jl:293-429).  Without `jacobianfunc` the Jacobian comes from first-order
    forward mode on the tape (the reference's TaylorN path, lyapunovspectrum.jl:45-63); a `jacobianfunc(jac, x, p, t)`
    given after `params` is traced into the tape instead (lyapunovspectrum.jl:65-78).  The reference's own test
    requires the two to give identical `t`, `x`, `λ` (test/lyapunov.jl:292-296); tests/test_lyap_userjac.py holds that.
    Continue from state: the solution carries `lyap_state` (frame Q, running sums λtsum, start of the
    averaging span); `lyap_taylorinteg(f, sol.xfinal, sol.tfinal[0], t2, ..., resume=sol.lyap_state)` goes on from there."""
    # (q0, t0, tmax, order, abstol[, params[, jacobianfunc]])  or  (q0, trange, order, abstol[, params[, jacobianfunc]])
    trange = None
    if len(args) >= 1 and np.ndim(args[0]) >= 1:  # src/lyapunovspectrum.jl:431-596
        trange, order, abstol = args[:3]
        params = args[3] if len(args) > 3 else None
        jacobianfunc = args[4] if len(args) > 4 else None
        t0 = tmax = None
    else:
        t0, tmax, order, abstol = args[:4]
        params = args[4] if len(args) > 4 else None
        jacobianfunc = args[5] if len(args) > 5 else None
    dof = np.asarray(q0).shape[0]
    tape = _as_tape(f, dof, params, jacobian=jacobianfunc if callable(jacobianfunc) else True)
    if tape.jac is None:
        raise ValueError("the tape has no Jacobian; build it with taylorize(..., jacobian=True)")
    return _integrate(tape, L.KIND_LYAP, q0, t0, tmax, trange, int(order), abstol, params, maxsteps=maxsteps, dense=False,
                      reltol=reltol, minstepsize=minstepsize, maxstepsize=maxstepsize, mode=mode, output=output,
                      kernel_opts=kernel_opts, devices=devices, block=block, lyap_state0=resume)


def fp64_peak(seconds: float = 0.5) -> Tuple[float, float]:
    """Measured FP64 DFMA peak of the current device in TFLOP/s (tb_fp64_peak)."""
    tf, ms = C.c_double(0.0), C.c_double(0.0)
    L.check(L.lib().tb_fp64_peak(float(seconds), C.byref(tf), C.byref(ms)), "tb_fp64_peak")
    return tf.value, ms.value
