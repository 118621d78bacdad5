"""A second, tape-free evaluator for arbitrary right-hand sides (test infrastructure, CPU only).

The reference's GENERIC `jetcoeffs!` (src/integrator/jetcoeffs.jl:66-92) re-evaluates the user's RHS on truncated
Taylor1 values at every order; its tests require that to equal the parsed path bit for bit.  oracle/tb_generic.cpp does
that for a fixed list of ODEs compiled into it; this module does it in Python for ANY expression, so that randomly
generated right-hand sides can be run through (a) the tracer + tape oracle and (b) plain series arithmetic that never
sees the tape.  The recurrences are the ones of tb_generic.cpp, same operation order; the order-0
seeds of pow / exp / log / sincos come from the oracle's tb_math.h probes, sqrt is IEEE."""
import ctypes as C
import math

import oracle

_lib = None


def _m():
    global _lib
    if _lib is None:
        _lib = oracle.lib()
        for f in ("tbo_math_pow",):
            getattr(_lib, f).argtypes = [C.c_double, C.c_double]
            getattr(_lib, f).restype = C.c_double
        for f in ("tbo_math_exp", "tbo_math_log"):
            getattr(_lib, f).argtypes = [C.c_double]
            getattr(_lib, f).restype = C.c_double
        _lib.tbo_math_sincos.argtypes = [C.c_double, C.POINTER(C.c_double), C.POINTER(C.c_double)]
        _lib.tbo_math_sincos.restype = None
    return _lib


class T1:
    """Truncated Taylor series with float coefficients c[0..order]."""
    __slots__ = ("c",)

    def __init__(self, c):
        self.c = list(c)

    @property
    def order(self):
        return len(self.c) - 1

    # ---- series (+-*/) series, series (+-*/) number: tb_generic.cpp:45-76 ----
    def __add__(self, o):
        if isinstance(o, T1):
            return T1([a + b for a, b in zip(self.c, o.c)])
        return T1([self.c[0] + o] + self.c[1:])

    __radd__ = lambda self, o: T1([o + self.c[0]] + self.c[1:])

    def __sub__(self, o):
        if isinstance(o, T1):
            return T1([a - b for a, b in zip(self.c, o.c)])
        return T1([self.c[0] - o] + self.c[1:])

    def __rsub__(self, o):
        return T1([o - self.c[0]] + [-v for v in self.c[1:]])

    def __neg__(self):
        return T1([-v for v in self.c])

    def __mul__(self, o):
        if not isinstance(o, T1):
            return T1([v * o for v in self.c])
        r = []
        for k in range(len(self.c)):
            s = self.c[0] * o.c[k]
            for i in range(1, k + 1):
                s += self.c[i] * o.c[k - i]
            r.append(s)
        return T1(r)

    def __rmul__(self, o):
        return T1([o * v for v in self.c])

    def __truediv__(self, o):
        if not isinstance(o, T1):
            return T1([v / o for v in self.c])
        return _div(self.c, o.c)

    def __rtruediv__(self, o):  # number / series: the numerator is the constant series
        return _div([o] + [0.0] * self.order, self.c, scalar_num=True)

    def __pow__(self, r):
        r = float(r)
        if r == 0.0:
            return T1([1.0] + [0.0] * self.order)
        if r == 1.0:
            return T1(self.c)
        if r == 2.0:
            return square(self)
        if r == 0.5:
            return sqrt(self)
        a = self.c
        c = [_m().tbo_math_pow(a[0], r)]
        for k in range(1, len(a)):
            s = r * k * c[0] * a[k]
            for i in range(1, k):
                rr = r * (k - i) - i
                s += rr * c[i] * a[k - i]
            c.append(s / (k * a[0]))
        return T1(c)


def _div(a, b, scalar_num=False):
    r = [a[0] / b[0]]
    for k in range(1, len(a)):
        s = r[0] * b[k]
        for i in range(1, k):
            s += r[i] * b[k - i]
        r.append((-s if scalar_num else a[k] - s) / b[0])
    return T1(r)


def square(x):
    a = x.c
    r = [a[0] * a[0]]
    for k in range(1, len(a)):
        kodd = k % 2
        kend = (k - 2 + kodd) // 2
        s = a[0] * a[k]
        for i in range(1, kend + 1):
            s += a[i] * a[k - i]
        s = 2 * s
        if kodd == 0:
            s += a[k // 2] * a[k // 2]
        r.append(s)
    return T1(r)


def sqrt(x):
    a = x.c
    r = [math.sqrt(a[0])]
    for k in range(1, len(a)):
        kodd = k % 2
        kend = (k - 2 + kodd) // 2
        s = 0.0
        for i in range(1, kend + 1):
            s = r[i] * r[k - i] if i == 1 else s + r[i] * r[k - i]
        aux = a[k] - 2 * s
        if kodd == 0:
            aux -= r[kend + 1] * r[kend + 1]
        r.append(aux / (2 * r[0]))
    return T1(r)


def sincos(x):
    a = x.c
    s0, c0 = C.c_double(), C.c_double()
    _m().tbo_math_sincos(a[0], C.byref(s0), C.byref(c0))
    s, c = [s0.value], [c0.value]
    for k in range(1, len(a)):
        xx = a[1]
        sk, ck = xx * c[k - 1], -xx * s[k - 1]
        for i in range(2, k + 1):
            xx = i * a[i]
            sk += xx * c[k - i]
            ck -= xx * s[k - i]
        s.append(sk / k)
        c.append(ck / k)
    return T1(s), T1(c)


def sin(x):
    return sincos(x)[0]


def cos(x):
    return sincos(x)[1]


def exp(x):
    a = x.c
    c = [_m().tbo_math_exp(a[0])]
    for k in range(1, len(a)):
        s = k * a[k] * c[0]
        for i in range(1, k):
            s += (k - i) * a[k - i] * c[i]
        c.append(s / k)
    return T1(c)


def log(x):
    # log!(c, a, k): c_1 = a_1 / a_0; c_k = (a_k - (sum_{i=1}^{k-1} (k-i) a_i c_{k-i}) / k) / a_0
    a = x.c
    c = [_m().tbo_math_log(a[0])]
    for k in range(1, len(a)):
        if k == 1:
            c.append(a[1] / a[0])
            continue
        s = (k - 1) * a[1] * c[k - 1]
        for i in range(2, k):
            s += (k - i) * a[i] * c[k - i]
        c.append((a[k] - s / k) / a[0])
    return T1(c)


def jetcoeffs(f, q0, params, t0, order):
    """jetcoeffs!(eqsdiff!, t, x, dx, xaux, params), src/integrator/jetcoeffs.jl:66-92: returns x[dof][order+1]."""
    dof = len(q0)
    x = [[float(v)] + [0.0] * order for v in q0]
    t = [float(t0), 1.0] + [0.0] * (order - 1)
    for o in range(order):
        taux = T1(t[: o + 1])
        xaux = [T1(xi[: o + 1]) for xi in x]
        dx = [None] * dof
        f(dx, xaux, params, taux)
        for j in range(dof):
            d = dx[j]
            dj = d.c[o] if isinstance(d, T1) else (float(d) if o == 0 else 0.0)
            x[j][o + 1] = dj / (o + 1)
    return x


# ---- primitives with an auxiliary series (src/parse_eqs.jl:1013-1046; recurrences as oracle/tb_oracle.cpp p_tan ...) ----
def _m1(name, x):
    f = getattr(_m(), name)
    f.argtypes = [C.c_double]
    f.restype = C.c_double
    return f(x)


def _sqr_k(a, k):
    """coefficient k of a^2 (sqr!)"""
    if k == 0:
        return a[0] * a[0]
    kodd = k % 2
    kend = (k - 2 + kodd) // 2
    s = a[0] * a[k]
    for i in range(1, kend + 1):
        s += a[i] * a[k - i]
    s = 2 * s
    if kodd == 0:
        s += a[k // 2] * a[k // 2]
    return s


def _tan(x, hyp):
    a = x.c
    c = [_m1("tbo_math_tanh" if hyp else "tbo_math_tan", a[0])]
    c2 = [c[0] * c[0]]
    for k in range(1, len(a)):
        s = k * a[k] * c2[0]
        for i in range(1, k):
            s += (k - i) * a[k - i] * c2[i]
        c.append(a[k] - s / k if hyp else a[k] + s / k)
        c2.append(_sqr_k(c, k))
    return T1(c)


def tan(x):
    return _tan(x, False)


def tanh(x):
    return _tan(x, True)


def sinhcosh(x):
    a = x.c
    s0, c0 = C.c_double(), C.c_double()
    f = _m().tbo_math_sinhcosh
    f.argtypes = [C.c_double, C.POINTER(C.c_double), C.POINTER(C.c_double)]
    f.restype = None
    f(a[0], C.byref(s0), C.byref(c0))
    s, c = [s0.value], [c0.value]
    for k in range(1, len(a)):
        xx = a[1]
        sk, ck = xx * c[k - 1], xx * s[k - 1]
        for i in range(2, k + 1):
            xx = i * a[i]
            sk += xx * c[k - i]
            ck += xx * s[k - i]
        s.append(sk / k)
        c.append(ck / k)
    return T1(s), T1(c)


def sinh(x):
    return sinhcosh(x)[0]


def cosh(x):
    return sinhcosh(x)[1]


def atan(x):
    a = x.c
    c, r = [_m1("tbo_math_atan", a[0])], [1 + a[0] * a[0]]
    for k in range(1, len(a)):
        s = 0.0
        if k >= 2:
            s = (k - 1) * r[1] * c[k - 1]
            for i in range(2, k):
                s += (k - i) * r[i] * c[k - i]
        r.append(_sqr_k(a, k))
        c.append((a[k] - s / k) / r[0])
    return T1(c)


def _asin(x, is_acos):
    a = x.c
    c = [_m1("tbo_math_acos" if is_acos else "tbo_math_asin", a[0])]
    r = [math.sqrt(1 - a[0] * a[0])]
    for k in range(1, len(a)):
        s = 0.0
        if k >= 2:
            s = (k - 1) * r[1] * c[k - 1]
            for i in range(2, k):
                s += (k - i) * r[i] * c[k - i]
        ak = -_sqr_k(a, k)
        kodd = k % 2
        kend = (k - 2 + kodd) // 2
        q = 0.0
        for i in range(1, kend + 1):
            q = r[i] * r[k - i] if i == 1 else q + r[i] * r[k - i]
        aux = ak - 2 * q
        if kodd == 0:
            aux -= r[kend + 1] * r[kend + 1]
        r.append(aux / (2 * r[0]))
        c.append(-(a[k] + s / k) / r[0] if is_acos else (a[k] - s / k) / r[0])
    return T1(c)


def asin(x):
    return _asin(x, False)


def acos(x):
    return _asin(x, True)
