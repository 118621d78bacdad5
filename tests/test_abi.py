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
    assert lib.tb_abi_version() == 1


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
    assert info["local_bytes"] == 0, "kernel spills"
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
