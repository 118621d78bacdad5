"""Accuracy of the deterministic elementary functions shared by the device kernels and the oracle
(taylorintegration.jl_b200/csrc/tb_math.h), measured against mpmath."""
import ctypes as C

import mpmath as mp
import numpy as np

import oracle

mp.mp.prec = 200


def _ulp_err(got, exact):
    return float(abs(mp.mpf(got) - exact) / mp.mpf(float(np.spacing(abs(float(exact))))))


def test_sincos_ulp():
    o = oracle.lib()
    s, c = C.c_double(), C.c_double()
    rng = np.random.default_rng(1)
    xs = np.concatenate([rng.uniform(-100, 100, 4000), rng.uniform(-1e5, 1e5, 1000), rng.uniform(-1, 1, 1000) * 1e-3,
                         [0.0, 0.1, 0.5, 0.78, 0.8, 1.0, 1.3, 1.5, 2.0, 3.0, 4.0, 10.0]])
    ms = mc = 0.0
    for x in xs:
        o.tbo_math_sincos(float(x), C.byref(s), C.byref(c))
        ms = max(ms, _ulp_err(s.value, mp.sin(mp.mpf(float(x)))))
        mc = max(mc, _ulp_err(c.value, mp.cos(mp.mpf(float(x)))))
    assert ms < 1.0 and mc < 1.0, (ms, mc)


def test_pow_exp_log_ulp():
    o = oracle.lib()
    rng = np.random.default_rng(2)
    m = 0.0
    for x in 10 ** rng.uniform(-40, 40, 3000):
        for y in (1 / 27, 1 / 28, 1.5, 1 / 25, 1 / 24):
            m = max(m, _ulp_err(o.tbo_math_pow(float(x), y), mp.power(mp.mpf(float(x)), mp.mpf(y))))
    assert m < 0.6, m
    assert o.tbo_math_pow(-0.9, 3.0) == -o.tbo_math_pow(0.9, 3.0) and o.tbo_math_pow(-2.0, 2.0) == 4.0
    assert np.isnan(o.tbo_math_pow(-2.0, 0.5))
    m = 0.0
    for x in rng.uniform(-700, 700, 3000):
        m = max(m, _ulp_err(o.tbo_math_exp(float(x)), mp.exp(mp.mpf(float(x)))))
    assert m < 0.6, m
    m = 0.0
    for x in 10 ** rng.uniform(-300, 300, 3000):
        m = max(m, _ulp_err(o.tbo_math_log(float(x)), mp.log(mp.mpf(float(x)))))
    assert m < 0.6, m


def test_pow_pair_is_bitwise_tb_pow():
    # tb_pow_pair (the interleaved step-size pair of the parity kernel) must return tb_pow's bits, special cases included
    o = oracle.lib()
    o.tbo_math_pow_pair.argtypes = [C.c_double] * 4 + [C.c_void_p]
    rng = np.random.default_rng(3)
    xs = list(10 ** rng.uniform(-320, 308, 4000)) + [0.0, 1.0, -2.0, np.inf, np.nan, 5e-324, 2.2250738585072014e-308, 1e-310]
    r = (C.c_double * 2)()
    for i, x1 in enumerate(xs):
        x2 = xs[(i * 7 + 3) % len(xs)]
        for y1, y2 in ((1 / 27, 1 / 28), (1 / 24, 1 / 25), (300.0, -300.0), (0.0, 1.5)):
            o.tbo_math_pow_pair(float(x1), y1, float(x2), y2, r)
            e1, e2 = o.tbo_math_pow(float(x1), y1), o.tbo_math_pow(float(x2), y2)
            assert np.array_equal(np.array([r[0], r[1]]).view(np.uint64), np.array([e1, e2]).view(np.uint64)), (x1, y1, x2, y2)
