"""ctypes binding of libtaylorb200.so (include/taylorb200.h).

The CUDA library is the product: importing this module fails loudly when the
shared object has not been built, and every integration call fails loudly when
no GPU is usable.  There is no CPU fallback here; the CPU oracle lives in
oracle/ and is only ever loaded by tests/, bench.py's baseline legs and
__graft_entry__.smoke().
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libtaylorb200.so")

# ---- enums (keep in sync with include/taylorb200.h) -------------------------
TB_OK, TB_ERR_ARG, TB_ERR_COMPILE, TB_ERR_CUDA, TB_ERR_UNSUPPORTED = 0, -1, -2, -3, -4
ST_MAXSTEPS, ST_ZERO_STEP, ST_NR_NOCONV, ST_NONFINITE, ST_EVENT_OVERFLOW = 1, 2, 4, 8, 16

OP_NOP, OP_IDENTITY, OP_NEG, OP_ADD, OP_SUB, OP_MUL, OP_DIV, OP_SQR, OP_SQRT, OP_POW = range(10)
OP_EXP, OP_LOG, OP_SINCOS, OP_ZERO, OP_ONE, OP_TAN, OP_SINHCOSH, OP_TANH, OP_ATAN, OP_ASIN, OP_ACOS = range(10, 21)
OP_NAMES = {
    OP_IDENTITY: "identity!", OP_NEG: "subst!(neg)", OP_ADD: "add!", OP_SUB: "subst!", OP_MUL: "mul!", OP_DIV: "div!",
    OP_SQR: "sqr!", OP_SQRT: "sqrt!", OP_POW: "pow!", OP_EXP: "exp!", OP_LOG: "log!", OP_SINCOS: "sincos!",
    OP_ZERO: "zero!", OP_ONE: "one!", OP_TAN: "tan!", OP_SINHCOSH: "sinhcosh!", OP_TANH: "tanh!", OP_ATAN: "atan!",
    OP_ASIN: "asin!", OP_ACOS: "acos!",
}
K_NONE, K_X, K_DX, K_TMP, K_PAR, K_TIME, K_IMM = range(7)
JAC_ZERO, JAC_CONST, JAC_SERIES, JAC_PAR = range(4)
GATE_ALWAYS, GATE_GT0, GATE_LT0, GATE_GE0, GATE_LE0 = range(5)
MODE_PARITY, MODE_FAST = 0, 1
KIND_PLAIN, KIND_EVENTS, KIND_LYAP, KIND_LYAPCOL = 0, 1, 2, 3


def operand(kind: int, index: int = 0) -> int:
    v = ((kind & 0xFF) << 24) | (index & 0xFFFFFF)
    return v - (1 << 32) if v >= (1 << 31) else v


def operand_kind(o: int) -> int:
    return (o & 0xFFFFFFFF) >> 24


def operand_index(o: int) -> int:
    return o & 0xFFFFFF


class TbOp(C.Structure):
    _fields_ = [("op", C.c_int32), ("res", C.c_int32), ("a", C.c_int32), ("b", C.c_int32), ("aux", C.c_int32),
                ("flags", C.c_int32), ("imm", C.c_double)]


class TbJacEntry(C.Structure):
    _fields_ = [("kind", C.c_int32), ("operand", C.c_int32), ("c", C.c_double)]


class TbOdeDesc(C.Structure):
    _fields_ = [("dof", C.c_int32), ("npar", C.c_int32), ("ntmp", C.c_int32), ("nops", C.c_int32),
                ("ev_nops", C.c_int32), ("jac_nops", C.c_int32), ("ops", C.POINTER(TbOp)),
                ("ev_g", C.c_int32), ("ev_gate", C.c_int32), ("ev_gate_cmp", C.c_int32), ("has_jac", C.c_int32),
                ("jac", C.POINTER(TbJacEntry))]


class TbCompileOpts(C.Structure):
    _fields_ = [("order", C.c_int32), ("mode", C.c_int32), ("kind", C.c_int32), ("block_threads", C.c_int32),
                ("reg_budget", C.c_int32), ("min_blocks_per_sm", C.c_int32), ("flags", C.c_int32),
                ("tmem_budget", C.c_int32)]


class TbOpts(C.Structure):
    _fields_ = [("abstol", C.c_double), ("reltol", C.c_double), ("minstepsize", C.c_double),
                ("maxstepsize", C.c_double), ("maxsteps", C.c_int64), ("eventorder", C.c_int32),
                ("newtoniter", C.c_int32), ("nrabstol", C.c_double), ("event_capacity", C.c_int32),
                ("params_per_traj", C.c_int32), ("steps_stride", C.c_int64), ("sort_by_cost", C.c_int32),
                ("reserved", C.c_int32)]


_DP = C.POINTER(C.c_double)


class TbOut(C.Structure):
    _fields_ = [("xfinal", _DP), ("tfinal", _DP), ("nsteps", C.POINTER(C.c_int64)), ("status", C.POINTER(C.c_int32)),
                ("xv_trange", _DP), ("tv", _DP), ("xv", _DP), ("psol", _DP), ("tevents", _DP), ("xevents", _DP),
                ("gresids", _DP), ("nevents", C.POINTER(C.c_int32)), ("lambda_", _DP), ("lambda_v", _DP),
                ("lyap_state", _DP), ("lyap_state0", _DP)]


EXPORTS = [
    "tb_abi_version", "tb_last_error", "tb_ode_create", "tb_ode_free", "tb_opts_default", "tb_compile_opts_default",
    "tb_generate_source", "tb_compile", "tb_integrate_multi", "tb_wide_layout", "tb_wide_generated", "tb_wide_source", "tb_wide_log", "tb_dense_eval_device", "tb_serialize_device", "tb_kernel_free", "tb_kernel_source", "tb_kernel_log", "tb_kernel_info",
    "tb_kernel_cubin", "tb_integrate", "tb_integrate_trange", "tb_work_bytes", "tb_integrate_device",
    "tb_launch_count", "tb_jet", "tb_dense_eval", "tb_fp64_peak", "tb_device_count", "tb_set_device",
    "tb_tn_create", "tb_tn_free", "tb_tn_ncoef", "tb_tn_monomials", "tb_tn_schedule", "tb_tn_integrate", "tb_tn_integrate_static",
    "tb_wide_create", "tb_wide_free", "tb_wide_info", "tb_wide_integrate", "tb_last_kernel_ms",
]


class TbTnOut(C.Structure):
    _fields_ = [("xfinal", _DP), ("tfinal", _DP), ("nsteps", C.POINTER(C.c_int64)), ("status", C.POINTER(C.c_int32)),
                ("xv_trange", _DP)]


class TaylorB200Error(RuntimeError):
    pass


_lib = None


def lib() -> C.CDLL:
    """Load libtaylorb200.so (built in-tree by `make -C taylorintegration.jl_b200/csrc`)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise TaylorB200Error(
            f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(there is no CPU fallback)")
    L = C.CDLL(LIB_PATH)
    vp = C.c_void_p
    L.tb_abi_version.restype = C.c_int
    L.tb_last_error.restype = C.c_char_p
    L.tb_ode_create.restype = vp
    L.tb_ode_create.argtypes = [C.POINTER(TbOdeDesc)]
    L.tb_ode_free.argtypes = [vp]
    L.tb_opts_default.argtypes = [C.POINTER(TbOpts)]
    L.tb_compile_opts_default.argtypes = [C.POINTER(TbCompileOpts)]
    L.tb_generate_source.restype = C.c_char_p
    L.tb_generate_source.argtypes = [vp, C.POINTER(TbCompileOpts)]
    L.tb_compile.restype = vp
    L.tb_compile.argtypes = [vp, C.POINTER(TbCompileOpts)]
    L.tb_kernel_free.argtypes = [vp]
    L.tb_kernel_source.restype = C.c_char_p
    L.tb_kernel_source.argtypes = [vp]
    L.tb_kernel_log.restype = C.c_char_p
    L.tb_kernel_log.argtypes = [vp]
    L.tb_kernel_info.argtypes = [vp] + [C.POINTER(C.c_int)] * 5
    L.tb_kernel_cubin.restype = vp
    L.tb_kernel_cubin.argtypes = [vp, C.POINTER(C.c_size_t)]
    L.tb_integrate.argtypes = [vp, C.c_int64, vp, vp, C.c_double, C.c_double, C.POINTER(TbOpts), C.POINTER(TbOut)]
    L.tb_integrate_trange.argtypes = [vp, C.c_int64, vp, vp, vp, C.c_int64, C.POINTER(TbOpts), C.POINTER(TbOut)]
    L.tb_integrate_multi.argtypes = [vp, vp, C.c_int32, C.c_int64, C.c_int64, vp, vp, C.c_double, C.c_double, vp, C.c_int64,
                                     C.POINTER(TbOpts), C.POINTER(TbOut)]
    L.tb_work_bytes.restype = C.c_size_t
    L.tb_integrate_device.argtypes = [vp, C.c_int64, vp, vp, C.c_double, C.c_double, vp, C.c_int64,
                                      C.POINTER(TbOpts), C.POINTER(TbOut), vp, vp]
    L.tb_launch_count.restype = C.c_int64
    L.tb_jet.argtypes = [vp, C.c_int64, vp, vp, C.c_double, C.POINTER(TbOpts), vp, vp]
    L.tb_dense_eval.argtypes = [C.c_int64, C.c_int32, C.c_int32, C.c_int64, vp, vp, vp, vp, C.c_int64, vp]
    L.tb_fp64_peak.argtypes = [C.c_double, _DP, _DP]
    L.tb_dense_eval_device.argtypes = [C.c_int64, C.c_int32, C.c_int32, C.c_int64, vp, vp, vp, vp, C.c_int64, vp, vp]
    L.tb_serialize_device.argtypes = [C.c_int64, C.c_int32, C.c_int32, C.c_int64, C.c_int64, vp, vp, vp, vp, vp]
    L.tb_set_device.argtypes = [C.c_int]
    L.tb_tn_create.restype = vp
    L.tb_tn_create.argtypes = [vp, C.c_int, C.c_int, C.c_int, C.c_int]
    L.tb_tn_free.argtypes = [vp]
    L.tb_tn_ncoef.argtypes = [vp]
    L.tb_tn_monomials.argtypes = [vp, vp]
    L.tb_tn_schedule.argtypes = [vp, vp, C.c_int]
    L.tb_tn_integrate.argtypes = [vp, C.c_int64, vp, vp, C.c_double, C.c_double, vp, C.c_int64, C.POINTER(TbOpts),
                                  C.POINTER(TbTnOut)]
    L.tb_tn_integrate_static.argtypes = [vp, C.c_int64, vp, vp, C.c_double, C.c_double, vp, C.c_int64, C.POINTER(TbOpts),
                                         C.POINTER(TbTnOut), vp]
    L.tb_last_kernel_ms.restype = C.c_double
    L.tb_wide_create.restype = vp
    L.tb_wide_create.argtypes = [vp, C.c_int, C.c_int]
    L.tb_wide_free.argtypes = [vp]
    L.tb_wide_info.argtypes = [vp] + [C.POINTER(C.c_int)] * 4
    L.tb_wide_layout.argtypes = [vp] + [C.POINTER(C.c_int)] * 7
    L.tb_wide_generated.argtypes = [vp] + [C.POINTER(C.c_int)] * 4
    L.tb_wide_source.argtypes = [vp]; L.tb_wide_source.restype = C.c_char_p
    L.tb_wide_log.argtypes = [vp]; L.tb_wide_log.restype = C.c_char_p
    L.tb_wide_integrate.argtypes = [vp, C.c_int64, vp, vp, C.c_double, C.c_double, vp, C.c_int64, C.POINTER(TbOpts),
                                    C.POINTER(TbOut)]
    _lib = L
    return L


def check(rc: int, what: str) -> None:
    if rc != TB_OK:
        msg = lib().tb_last_error()
        raise TaylorB200Error(f"{what} failed ({rc}): {msg.decode() if msg else ''}")


def last_error() -> str:
    msg = lib().tb_last_error()
    return msg.decode() if msg else ""
