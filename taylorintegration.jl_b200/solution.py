"""Consumers of the dense output: `reverse` / `flipsign` / `join` of a `TaylorSolution`
(src/integrator/taylorsolution.jl:475-511), the flat `TaylorSolutionSerialization` / `TaylorSolutionNSerialization`
layouts (513-657) and a device-resident batched `sol(t)` over the `psol` an ensemble run left in HBM.

The reference writes JLD2 files through `writeas`/`readas`; a JLD2 container cannot be produced without the Julia
package, so the arrays are stored as an `.npz` with the SAME fields (`order`, `dims`, `t`, `p`[, `vars`, `varorder`]):
in Julia `TaylorSolutionSerialization{Float64,2}(order, dims, t, p)` followed by `convert(TaylorSolution, ...)` rebuilds
the solution.
"""
from __future__ import annotations

import ctypes as C
from typing import Dict, Optional

import numpy as np

from . import _lib as L
from .api import EnsembleSolution, TaylorSolution


# ---- Taylor1 helpers with the reference's operation order (host side, one trajectory) --------------------------
def _evaluate_at_taylor(p: np.ndarray, d: float) -> np.ndarray:
    """`p(d + Taylor1(ord))` for one polynomial `p[0..P]`: Horner with Taylor1 arithmetic truncated at order P
    (TaylorSeries `evaluate(a::Taylor1, x::Taylor1)`): s = p[P]; s = s*(d + tau) + p[k].  With the operand
    `[d, 1, 0, ...]` the Cauchy product mul! reduces to c[k] = s[k-1] + s[k]*d (its other terms are exact zeros)."""
    P = p.shape[-1] - 1
    s = np.zeros_like(p)
    s[..., 0] = p[..., P]
    for k in range(P - 1, -1, -1):
        c = np.empty_like(s)
        c[..., 0] = s[..., 0] * d
        c[..., 1:] = s[..., :-1] + s[..., 1:] * d
        c[..., 0] = c[..., 0] + p[..., k]
        s = c
    return s


def _dense_solution(t: np.ndarray, p: np.ndarray) -> TaylorSolution:
    """`_dense_solution(t, p)` (taylorsolution.jl:104-131): x[i] = constant term of p[i], x[end] = p[end](t[end]-t[end-1])."""
    t = np.asarray(t, dtype=np.float64)
    n = p.shape[0]
    assert t.shape[0] == n + 1, "Cannot infer solution values unless length(t) == size(p, 1) + 1"
    x = np.zeros((n + 1,) + p.shape[1:-1])
    if n:
        x[:n] = p[..., 0]
        dt = t[-1] - t[-2]
        s = p[n - 1, ..., -1].copy()
        for k in range(p.shape[-1] - 2, -1, -1):
            s = s * dt + p[n - 1, ..., k]
        x[n] = s
    return TaylorSolution(t=t, x=x, p=p)


def reverse(sol: TaylorSolution) -> TaylorSolution:
    """`reverse(sol)` (taylorsolution.jl:475-495): times reversed; the first polynomial is the last step's expansion
    re-centred at the final time (`sol(sol.t[end] + Taylor1(ord))`), polynomial i >= 2 is `p[end-i+2]`."""
    if sol.p is None:
        raise ValueError("TaylorSolution objects computed with dense=False do not store Taylor polynomials")
    p, t = sol.p, sol.t
    n = p.shape[0]
    prev = np.empty_like(p)
    prev[0] = _evaluate_at_taylor(p[n - 1], t[-1] - t[-2])
    for i in range(1, n):
        prev[i] = p[n - i]
    return _dense_solution(t[::-1].copy(), prev)


def flipsign(sol: TaylorSolution) -> TaylorSolution:
    """`flipsign(sol)` (taylorsolution.jl:497-502): t -> 2 t0 - t, p -> p(-tau): coefficient k changes sign for odd k."""
    if sol.p is None:
        raise ValueError("TaylorSolution objects computed with dense=False do not store Taylor polynomials")
    sign = np.where(np.arange(sol.p.shape[-1]) % 2 == 0, 1.0, -1.0)
    return _dense_solution(2 * sol.t[0] - sol.t, sol.p * sign)


def join(bwd: TaylorSolution, fwd: TaylorSolution) -> TaylorSolution:
    """`join(bwd, fwd)` (taylorsolution.jl:504-511): one solution over [bwd.t[end], fwd.t[end]]."""
    assert bwd.t[0] == fwd.t[0], "Initial time must be the same for both TaylorSolution"
    assert bwd.p.shape[-1] == fwd.p.shape[-1], "Expansion order must be the same for both TaylorSolution"
    r = reverse(bwd)
    return _dense_solution(np.concatenate([r.t, fwd.t[1:]]), np.concatenate([r.p, fwd.p], axis=0))


# ---- serialization layouts --------------------------------------------------------------------------------------
def to_serialization(sol: TaylorSolution) -> Dict[str, np.ndarray]:
    """Fields of `TaylorSolutionSerialization{T,N}` (taylorsolution.jl:513-545): `p[(i-1)*k+1 : i*k]` are the coefficients
    of the i-th polynomial in `eachindex(sol.p)` order, i.e. column-major over the (steps x dof) matrix."""
    if sol.p is None:
        raise ValueError("only dense solutions serialise (the reference dispatches on a non-Nothing `p`)")
    ns, dof, k = sol.p.shape
    flat = np.ascontiguousarray(np.transpose(sol.p, (1, 0, 2))).reshape(-1)  # dof outer, step, coefficient inner
    return {"order": np.int64(k - 1), "dims": np.array([ns, dof], dtype=np.int64), "t": np.asarray(sol.t, dtype=np.float64).copy(),
            "p": flat}


def from_serialization(d) -> TaylorSolution:
    """Inverse (taylorsolution.jl:547-558)."""
    k = int(d["order"]) + 1
    ns, dof = (int(v) for v in d["dims"])
    p = np.transpose(np.asarray(d["p"], dtype=np.float64).reshape(dof, ns, k), (1, 0, 2)).copy()
    return _dense_solution(np.asarray(d["t"], dtype=np.float64), p)


def to_serialization_n(t: np.ndarray, p: np.ndarray, monomials: np.ndarray, names=None) -> Dict[str, np.ndarray]:
    """Fields of `TaylorSolutionNSerialization{T,N}` (taylorsolution.jl:560-614) for jet-transport polynomials
    `p[steps, dof, order+1, ncoef]` whose last axis is in the flat TaylorN order of `JetTransport.monomials` (ascending
    degree, then TaylorSeries' monomial order): polynomial, time coefficient, homogeneous degree, monomial."""
    ns, dof, k, nc = p.shape
    nvars = monomials.shape[1]
    names = names or [f"x{subscript}" for subscript in range(1, nvars + 1)]
    flat = np.ascontiguousarray(np.transpose(p, (1, 0, 2, 3))).reshape(-1)
    return {"vars": np.array(names), "order": np.int64(k - 1), "varorder": np.int64(int(monomials.sum(1).max())),
            "dims": np.array([ns, dof], dtype=np.int64), "t": np.asarray(t, dtype=np.float64).copy(), "p": flat}


def save(path: str, sol: TaylorSolution) -> None:
    np.savez(path, **to_serialization(sol))


def load(path: str) -> TaylorSolution:
    with np.load(path) as d:
        return from_serialization({k: d[k] for k in d.files})


# ---- device-resident dense output of an ensemble ------------------------------------------------------------------
class DenseEnsemble:
    """The dense output of an ensemble kept in HBM: `psol`, `tv`, `nsteps` are uploaded once (or adopted from torch CUDA
    tensors) and then `eval(tq)` (batched `sol(t)`, taylorsolution.jl:337-423) and `serialize()` (the flat
    `TaylorSolutionSerialization` vectors of every trajectory) run on the device without touching the host."""

    def __init__(self, sol: EnsembleSolution, order: int, device: int = 0):
        import torch
        if sol.p is None or sol.t is None:
            raise ValueError("the ensemble was integrated without dense output (output='steps', dense=True)")
        self.torch = torch
        self.dev = torch.device("cuda", device)
        self.order, self.dof, self.stride, self.N = int(order), sol.p.shape[0], sol.p.shape[2], sol.p.shape[3]
        self.psol = torch.from_numpy(np.ascontiguousarray(sol.p)).to(self.dev)
        self.tv = torch.from_numpy(np.ascontiguousarray(sol.t)).to(self.dev)
        self.nsteps = torch.from_numpy(np.ascontiguousarray(sol.nsteps, dtype=np.int64)).to(self.dev)

    def eval(self, tq) -> "object":
        """x(tq) for every trajectory: torch tensor [dof, nq, N] on the device (NaN outside a trajectory's span)."""
        torch = self.torch
        tq = torch.as_tensor(np.atleast_1d(np.asarray(tq, dtype=np.float64))).to(self.dev)
        xq = torch.empty((self.dof, tq.numel(), self.N), dtype=torch.float64, device=self.dev)
        L.check(L.lib().tb_dense_eval_device(self.N, self.dof, self.order, self.stride, self.tv.data_ptr(), self.psol.data_ptr(),
                                             self.nsteps.data_ptr(), tq.data_ptr(), tq.numel(), xq.data_ptr(),
                                             torch.cuda.current_stream().cuda_stream), "tb_dense_eval_device")
        return xq

    def serialize(self):
        """(flat, offsets): `flat[offsets[i] : offsets[i+1]]` is the `p` field of trajectory i's
        `TaylorSolutionSerialization` (device tensors)."""
        torch = self.torch
        k = self.order + 1
        sizes = self.nsteps * (self.dof * k)
        offsets = torch.zeros(self.N + 1, dtype=torch.int64, device=self.dev)
        offsets[1:] = torch.cumsum(sizes, 0)
        total = int(offsets[-1].item())
        flat = torch.empty(total, dtype=torch.float64, device=self.dev)
        mx = int(self.nsteps.max().item()) if self.N else 0
        L.check(L.lib().tb_serialize_device(self.N, self.dof, self.order, self.stride, mx, self.psol.data_ptr(),
                                            self.nsteps.data_ptr(), offsets.data_ptr(), flat.data_ptr(),
                                            torch.cuda.current_stream().cuda_stream), "tb_serialize_device")
        return flat, offsets
