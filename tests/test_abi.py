"""The C-ABI library loads without a GPU, exports every symbol include/taylorb200.h declares, generates and
NVRTC-compiles kernels for sm_100a, and refuses to compute without a device (no CPU fallback)."""
import ctypes as C
import os
import re

import numpy as np
import pytest

import taylorintegration_b200 as tb
from taylorintegration_b200 import _lib as L, problems as P

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_every_declared_symbol_is_exported():
    hdr = open(os.path.join(ROOT, "include", "taylorb200.h")).read()
    declared = set(re.findall(r"\b(tb_[a-z0-9_]+)\s*\(", hdr)) - {"tb_op", "tb_out", "tb_opts"}
    lib = L.lib()
    missing = [s for s in sorted(declared) if not hasattr(lib, s)]
    assert not missing, missing
    assert set(L.EXPORTS) <= declared
    assert lib.tb_abi_version() == 2  # 2: tb_out_t gained lyap_state / lyap_state0


def test_defaults_match_reference_keywords():
    # taylorinteg.jl:313-319, rootfinding.jl:310-318
    o = L.TbOpts()
    L.lib().tb_opts_default(C.byref(o))
    assert (o.maxsteps, o.reltol, o.minstepsize, o.eventorder, o.newtoniter) == (500, 0.0, 0.0, 0, 10)
    assert o.maxstepsize == np.inf and o.nrabstol == np.finfo(float).eps


def test_kepler_tape_is_the_reference_unfolding():
    # two sqr (pow 2), add, pow 1.5, two identity, two mul-by-scalar, two div
    t = P.kepler_tape()
    ops = [o.op for o in t.ops]
    assert ops == [L.OP_POW, L.OP_POW, L.OP_ADD, L.OP_POW, L.OP_IDENTITY, L.OP_IDENTITY, L.OP_MUL, L.OP_DIV, L.OP_MUL, L.OP_DIV]
    assert [o.imm for o in t.ops if o.op == L.OP_POW] == [2.0, 2.0, 1.5]


def test_compiles_for_sm100a_without_gpu():
    k = tb.Kernel(P.kepler_tape(), 28)
    info = k.info()
    assert info["block_threads"] % 32 == 0 and info["smem_bytes"] <= 227 * 1024
    assert 0 <= info["local_bytes"] <= 232 + 256, "kernel spills more than the bound (232 B is the cold-path array)"
    assert "tb_ensemble_kernel" in k.source and "sm_100a" in k.log
    assert k.cubin()[:4] == b"\x7fELF"


def test_invalid_tapes_are_rejected():
    with pytest.raises(TypeError):
        tb.taylorize(lambda dx, x, p, t: dx.__setitem__(0, p[0] ** 2 * x[0]), 1, 1)  # test/taylorize.jl:953-992
    with pytest.raises(ValueError):
        tb.taylorize(lambda dx, x, p, t: None, 1, 0)
    t = P.kepler_tape()
    bad = tb.Tape(t.dof, t.npar, 1, t.ops)  # ntmp too small -> operand out of range
    with pytest.raises(tb.TaylorB200Error):
        tb.Kernel(bad, 28)
    with pytest.raises(tb.TaylorB200Error):
        tb.Kernel(P.lorenz_tape(jacobian=False), 28, kind=L.KIND_LYAP)


@pytest.mark.skipif(L.lib().tb_device_count() > 0, reason="box has a GPU")
def test_no_cpu_fallback():
    q0, mu, _ = P.kepler_ensemble(4)
    with pytest.raises(tb.TaylorB200Error, match="CUDA|device|driver"):
        tb.taylorinteg(P.kepler_tape(), q0, 0.0, 1.0, 28, 1e-20, mu)


def test_entry_points_validate_arguments_without_a_gpu():
    """The argument / build-flag checks run before anything touches the device, and are the same for the host-buffer
    and the device-buffer entry points (tb_integrate_device skipped them)."""
    lib = L.lib()
    q0, mu, _ = P.kepler_ensemble(8)
    xf = np.zeros((4, 8))
    out = L.TbOut()
    out.xfinal = xf.ctypes.data_as(L._DP)
    o = tb.make_opts(maxsteps=50)
    tr = np.linspace(0.0, 1.0, 5)
    p = lambda a: a.ctypes.data_as(C.c_void_p)  # noqa: E731
    kf = tb.Kernel(P.kepler_tape(), 28, flags=262144)  # TB_CF_FINAL_ONLY
    k = tb.Kernel(P.kepler_tape(), 28)
    work = np.zeros(32, dtype=np.int64)
    for dev_entry in (False, True):
        def call(kk, trange=None, nt=0, outp=out, params=p(mu), q=p(q0), opts=o):
            if dev_entry:
                return lib.tb_integrate_device(kk.handle, 8, q, params, 0.0, 1.0, trange, nt, C.byref(opts), C.byref(outp),
                                               p(work), None)
            if trange is not None:
                return lib.tb_integrate_trange(kk.handle, 8, q, params, trange, nt, C.byref(opts), C.byref(outp))
            return lib.tb_integrate(kk.handle, 8, q, params, 0.0, 1.0, C.byref(opts), C.byref(outp))
        assert call(kf, p(tr), 5) == L.TB_ERR_ARG and "TB_CF_FINAL_ONLY" in L.last_error()
        o2 = L.TbOut()
        o2.xfinal, o2.tv = out.xfinal, xf.ctypes.data_as(L._DP)
        assert call(kf, outp=o2) == L.TB_ERR_ARG
        assert call(k, p(tr), 1) == L.TB_ERR_ARG            # trange needs two points
        assert call(k, params=None) == L.TB_ERR_ARG and call(k, q=None) == L.TB_ERR_ARG
        bad = tb.make_opts(maxsteps=50, steps_stride=3)
        assert call(k, outp=o2, opts=bad) == L.TB_ERR_ARG   # per-step output with a stride below maxsteps + 1
    ke = tb.Kernel(P.pendulum_tape(), 25, L.KIND_EVENTS)
    oe = tb.make_opts(maxsteps=50, eventorder=26)            # rootfinding.jl:321
    qp = np.zeros((2, 8))
    assert lib.tb_integrate_device(ke.handle, 8, p(qp), None, 0.0, 1.0, None, 0, C.byref(oe), C.byref(out), p(work), None) == L.TB_ERR_ARG
    assert lib.tb_integrate(ke.handle, 8, p(qp), None, 0.0, 1.0, C.byref(oe), C.byref(out)) == L.TB_ERR_ARG
    from taylorintegration_b200.tape import lyap_column_tape
    kl = tb.Kernel(lyap_column_tape(P.lorenz_tape()), 28, L.KIND_LYAPCOL)  # built without TB_CF_LYAP_TRANGE
    ql, pl = P.lorenz_ensemble(8)
    assert lib.tb_integrate_device(kl.handle, 8, p(ql), p(pl), 0.0, 1.0, p(tr), 5, C.byref(o), C.byref(out), p(work), None) == L.TB_ERR_ARG


def test_kernel_info_reports_the_stack_frame():
    """tb_kernel_info parses ptxas -v (round 1 always reported 0): the parity Kepler kernel carries the 232-byte
    array of the cold second-stepsize path plus a bounded amount of spill; the gate is a bound, not zero."""
    k = tb.Kernel(P.kepler_tape(), 28, flags=262144)
    info = k.info()
    assert "bytes stack frame" in k.log
    assert 0 < info["local_bytes"] <= 232 + 256, info
