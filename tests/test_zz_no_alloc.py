"""tb_tn_integrate_static: the TaylorN kernel with its scratch (work-queue counter, device copies of the tape and the
index tables, jet workspace) in a static device array of the library -- the entry point for callers that cannot call
cudaMalloc.  It has no counterpart in the reference; what it must equal
is the existing entry point tb_tn_integrate on the same inputs, bit for bit (parity mode: `array_equal`, the tolerance
of every other parity test here).

Collected last on purpose.  Inputs and outputs are device buffers of the caller:
torch tensors here -- the entry point itself allocates nothing.
"""
import ctypes as C

import numpy as np
import pytest

import taylorintegration_b200 as tb
from taylorintegration_b200 import _lib as L, problems as P
from taylorintegration_b200.tape import taylorize

pytestmark = pytest.mark.gpu
T_PEND = 7.019250311844546
SCRATCH_BYTES = 16 << 20  # TB_TN_STATIC_SCRATCH_BYTES of include/taylorb200.h


def _static_call(jt, q0poly, t0, tmax, params=None, trange=None, maxsteps=500, abstol=1e-20, expect_rc=0):
    """One call of tb_tn_integrate_static through the C ABI on caller-owned device buffers; returns host copies."""
    import torch
    dev = torch.device("cuda", 0)
    q = np.ascontiguousarray(q0poly, dtype=np.float64)
    dof, nc, N = q.shape
    d_q = torch.from_numpy(q).to(dev)
    per = 0
    d_p = None
    if params is not None:
        p = np.ascontiguousarray(params, dtype=np.float64)
        per = 1 if p.ndim == 2 else 0
        d_p = torch.from_numpy(p).to(dev)
    # outputs start as garbage: everything compared below must have been written by the kernel
    d_xf = torch.full((dof, nc, N), -7.25, dtype=torch.float64, device=dev)
    d_tf = torch.full((N,), -7.25, dtype=torch.float64, device=dev)
    d_ns = torch.full((N,), -3, dtype=torch.int64, device=dev)
    d_st = torch.full((N,), -3, dtype=torch.int32, device=dev)
    out = L.TbTnOut()
    out.xfinal = C.cast(d_xf.data_ptr(), L._DP)
    out.tfinal = C.cast(d_tf.data_ptr(), L._DP)
    out.nsteps = C.cast(d_ns.data_ptr(), C.POINTER(C.c_int64))
    out.status = C.cast(d_st.data_ptr(), C.POINTER(C.c_int32))
    d_tr = d_xv = None
    nt = 0
    if trange is not None:
        tr = np.ascontiguousarray(trange, dtype=np.float64)
        nt = tr.size
        d_tr = torch.from_numpy(tr).to(dev)
        d_xv = torch.full((dof, nc, nt, N), -7.25, dtype=torch.float64, device=dev)
        out.xv_trange = C.cast(d_xv.data_ptr(), L._DP)
        t0, tmax = tr[0], tr[-1]
    o = tb.make_opts(abstol=abstol, maxsteps=maxsteps, params_per_traj=per)
    torch.cuda.synchronize()
    rc = L.lib().tb_tn_integrate_static(jt._t, N, d_q.data_ptr(), None if d_p is None else d_p.data_ptr(), float(t0),
                                        float(tmax), None if d_tr is None else d_tr.data_ptr(), nt, C.byref(o),
                                        C.byref(out), None)
    assert rc == expect_rc, (rc, L.lib().tb_last_error())
    torch.cuda.synchronize()
    res = dict(xfinal=d_xf.cpu().numpy(), tfinal=d_tf.cpu().numpy(), nsteps=d_ns.cpu().numpy(), status=d_st.cpu().numpy())
    if d_xv is not None:
        res["x_trange"] = d_xv.cpu().numpy()
    return res


def _same(got, ref):
    for f in ("xfinal", "tfinal", "nsteps", "status"):
        assert np.array_equal(got[f], getattr(ref, f), equal_nan=True), f
    if "x_trange" in got:
        assert np.array_equal(got["x_trange"], ref.x_trange, equal_nan=True), "x_trange"


def test_static_scratch_entry_matches_the_allocating_entry_on_calls_in_a_row():
    # pendulum, 2 variables of order 4 (15 coefficients): two different ensembles, back to back, then the first again --
    # the array keeps the previous call's counter, tables and jets, and none of that may leak into the next call
    jt = tb.JetTransport(P.pendulum_tape(events=False), 25, 2, 4)
    qa = jt.initial(P.pendulum_ensemble(48))
    qb = jt.initial(P.pendulum_ensemble(33, seed=7))
    ref_a = jt.integrate(qa, 0.0, T_PEND, 1e-20, maxsteps=500)
    ref_b = jt.integrate(qb, 0.0, 2.5, 1e-20, maxsteps=500)
    assert ref_a.nsteps.min() > 3 and np.all(ref_a.status == 0)
    _same(_static_call(jt, qa, 0.0, T_PEND), ref_a)
    _same(_static_call(jt, qb, 0.0, 2.5), ref_b)
    _same(_static_call(jt, qa, 0.0, T_PEND), ref_a)
    # another program right after (other tables, parameters, more coefficients): Kepler in 4 variables of order 3
    jk = tb.JetTransport(P.kepler_tape(), 28, 4, 3)
    q0, mu, _ = P.kepler_ensemble(12)
    qk = jk.initial(q0)
    ref_k = jk.integrate(qk, 0.0, 3.0, 1e-20, mu, maxsteps=500)
    _same(_static_call(jk, qk, 0.0, 3.0, params=mu), ref_k)
    # and the trange driver of the first program again
    grid = np.linspace(0.0, 4.0, 9)
    ref_t = jt.integrate(qb, trange=grid, abstol=1e-20, maxsteps=500)
    _same(_static_call(jt, qb, None, None, trange=grid), ref_t)


def test_static_scratch_entry_rejects_what_does_not_fit():
    # 16 components with two temporaries each, 1001 coefficients (4 variables of order 10), order 60:
    # 65 jets x 61 x 1001 doubles = 31.8 MB for ONE trajectory > 16 MiB
    def f(dx, x, p, t):
        for i in range(16):
            dx[i] = x[i] * x[(i + 1) % 16] + x[(i + 5) % 16]
    jt = tb.JetTransport(taylorize(f, 16, 0, name="ring16"), 60, 4, 10)
    assert (2 * 16 + jt.tape.ntmp + 1) * 61 * jt.ncoef * 8 > SCRATCH_BYTES
    q = np.zeros((16, jt.ncoef, 1))
    q[:, 0, 0] = 0.1
    _static_call(jt, q, 0.0, 0.1, expect_rc=-4)  # TB_ERR_UNSUPPORTED
    msg = L.lib().tb_last_error()
    msg = msg.decode() if isinstance(msg, bytes) else str(msg)
    assert "TB_TN_STATIC_SCRATCH_BYTES" in msg and "tb_tn_integrate" in msg
    # bad arguments are refused before anything is enqueued
    o = tb.make_opts()
    out = L.TbTnOut()
    assert L.lib().tb_tn_integrate_static(jt._t, 4, None, None, 0.0, 1.0, None, 0, C.byref(o), C.byref(out), None) == -1  # TB_ERR_ARG
