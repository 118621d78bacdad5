"""The primitives with an auxiliary series (src/parse_eqs.jl:1013-1046): tan!, sinhcosh!, tanh!, atan!, asin!, acos!.
CPU: the oracle's recurrences against closed-form Taylor coefficients (mpmath, 60 digits) and the seeds of tb_math.h
against mpmath; GPU: the generated kernel against the oracle, bit for bit."""
import mpmath as mp
import numpy as np
import pytest

import oracle
import taylorintegration_b200 as tb
from taylorintegration_b200 import acos, asin, atan, cosh, sinh, tan, tanh
from taylorintegration_b200 import _lib as L

FUNCS = [("tan", tan, mp.tan, 0.3), ("sinh", sinh, mp.sinh, 0.7), ("cosh", cosh, mp.cosh, 0.7), ("tanh", tanh, mp.tanh, 0.4),
         ("atan", atan, mp.atan, 0.6), ("asin", asin, mp.asin, 0.3), ("acos", acos, mp.acos, -0.25)]


@pytest.mark.parametrize("name,f,mpf,t0", FUNCS)
def test_recurrence_against_closed_form_taylor_coefficients(name, f, mpf, t0):
    # x' = f(t): the jet of x is the integrated Taylor series of f about t0 (non-autonomous: the time series is the operand)
    order = 20
    tape = tb.taylorize(lambda dx, x, p, t: dx.__setitem__(0, f(t)), 1, 0)
    assert tape.ops[0].op in (L.OP_TAN, L.OP_SINHCOSH, L.OP_TANH, L.OP_ATAN, L.OP_ASIN, L.OP_ACOS)
    xj, dxj, _ = oracle.jet(tape, order, np.array([0.0]), t0=t0)
    mp.mp.dps = 60
    want = mp.taylor(mpf, mp.mpf(t0), order)
    for k in range(order):
        assert abs(mp.mpf(float(dxj[0, k])) - want[k]) <= 1e-13 * max(abs(want[k]), mp.mpf(1e-3)), (name, k)
        assert xj[0, k + 1] == dxj[0, k] / (k + 1)  # solcoeff!


def test_seed_functions_against_mpmath():
    mp.mp.dps = 40
    o = oracle.lib()
    import ctypes as C
    rng = np.random.default_rng(3)
    xs = np.concatenate([rng.uniform(-3, 3, 400), rng.uniform(-0.01, 0.01, 50), [0.0, 0.5, -0.5, 1.0, -1.0, 20.0, -30.0]])

    def ulps(got, want):
        want_f = float(want)
        return abs(mp.mpf(got) - want) / mp.mpf(float(np.spacing(abs(want_f)) if want_f != 0 else 5e-324))
    for nm, mpf, dom in (("tan", mp.tan, lambda x: abs(np.cos(x)) > 0.05), ("tanh", mp.tanh, None), ("atan", mp.atan, None),
                         ("asin", mp.asin, lambda x: abs(x) <= 1), ("acos", mp.acos, lambda x: abs(x) <= 1)):
        fn = getattr(o, "tbo_math_" + nm)
        fn.restype, fn.argtypes = C.c_double, [C.c_double]
        for x in xs:
            if dom is not None and not dom(x):
                continue
            assert ulps(fn(float(x)), mpf(mp.mpf(float(x)))) <= 4, (nm, x)
    o.tbo_math_sinhcosh.argtypes = [C.c_double, C.POINTER(C.c_double), C.POINTER(C.c_double)]
    for x in xs:
        s, c = C.c_double(), C.c_double()
        o.tbo_math_sinhcosh(float(x), C.byref(s), C.byref(c))
        assert ulps(s.value, mp.sinh(mp.mpf(float(x)))) <= 4 and ulps(c.value, mp.cosh(mp.mpf(float(x)))) <= 4, x


def _rhs(dx, x, p, t):
    a = tan(0.3 * x[0]) + sinh(x[1]) * cosh(0.5 * x[0])
    b = tanh(x[1] + t) - atan(x[0] * x[1])
    dx[0] = asin(0.4 * b) + p[0] * a
    dx[1] = acos(0.3 * tanh(x[0])) - x[1]


def test_generic_jet_structure_and_kernel_builds():
    tape = tb.taylorize(_rhs, 2, 1)
    k = tb.Kernel(tape, 12)           # the rolled generator lowers these (the unrolled one declines)
    assert "tb_sinhcosh" in k.source and "tb_acos" in k.source and "#define TB_VARIANT 1" in k.source or "TB_VARIANT" not in k.source.split("tb_math")[0]
    xj, _, dt = oracle.jet(tape, 12, np.array([0.2, -0.1]), np.array([0.7]), t0=0.1)
    assert np.all(np.isfinite(xj)) and dt > 0


@pytest.mark.gpu
def test_gpu_matches_oracle_bitwise():
    tape = tb.taylorize(_rhs, 2, 1)
    rng = np.random.default_rng(11)
    q0 = rng.uniform(-0.5, 0.5, (2, 70))
    par = np.array([0.7])
    for order in (6, 17):
        xj, dt = tb.Kernel(tape, order).jet(q0, par, t0=0.1)
        for i in range(0, 70, 5):
            xo, _, dto = oracle.jet(tape, order, q0[:, i], par, t0=0.1)
            assert np.array_equal(xj[:, :, i], xo) and dt[i] == dto
    import warnings
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")  # some members run into the branch point of asin (zero step): same on both sides
        g = tb.taylorinteg(tape, q0, 0.1, 3.0, 17, 1e-18, par, maxsteps=2000)
    o = oracle.integrate(tape, 17, q0, 0.1, 3.0, par, abstol=1e-18, maxsteps=2000)
    assert np.array_equal(g.xfinal, o.xfinal) and np.array_equal(g.nsteps, o.nsteps) and np.array_equal(g.status, o.status)
    assert (g.status == 0).sum() >= 5 and np.array_equal(g.tfinal, o.tfinal)
