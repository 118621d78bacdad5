/*
 * tb_math.h -- deterministic FP64 elementary functions shared by the device
 * kernels (parity mode) and the CPU oracle.
 *
 * Why: the reference seeds its recurrences with Julia's `^`, `sincos`, `exp`,
 * `log` (TaylorSeries pow!/sincos!/exp!/log! at k = 0) and computes the step
 * size with `aux^kinv` (src/integrator/stepsize.jl:107-111).  To make
 * "GPU == CPU oracle, bit for bit" a testable statement, both sides must use
 * the SAME elementary functions, built only from IEEE-754 +,-,*,/,fma with a
 * fixed operation order.  On the device every operation is an explicit
 * round-to-nearest intrinsic (never contracted); on the host the translation
 * unit is compiled with -ffp-contract=off.
 *
 * Accuracy (checked against mpmath in tests/test_math.py):
 *   tb_log, tb_exp, tb_pow : < 0.6 ulp     tb_sincos : < 1 ulp for |x| < 1e5
 */
#ifndef TB_MATH_H
#define TB_MATH_H

#if defined(__CUDACC_RTC__) || defined(__CUDA_ARCH__)
#define TB_DEVICE_MATH 1
#else
#define TB_DEVICE_MATH 0
#include <math.h>
#include <string.h>
#include <stdint.h>
#endif

#if defined(__CUDACC__) || defined(__CUDACC_RTC__)
#define TB_HD __host__ __device__ __forceinline__
#else
#define TB_HD static inline
#endif

#if TB_DEVICE_MATH
#define TB_ADD(a, b) __dadd_rn((a), (b))
#define TB_SUB(a, b) __dsub_rn((a), (b))
#define TB_MUL(a, b) __dmul_rn((a), (b))
#define TB_DIV(a, b) __ddiv_rn((a), (b))
#define TB_FMA(a, b, c) __fma_rn((a), (b), (c))
#define TB_SQRT(a) __dsqrt_rn((a))
#define TB_ABS(a) fabs((a))
#define TB_TO_BITS(x) ((unsigned long long)__double_as_longlong((x)))
#define TB_FROM_BITS(u) __longlong_as_double((long long)(u))
#else
#define TB_ADD(a, b) ((a) + (b))
#define TB_SUB(a, b) ((a) - (b))
#define TB_MUL(a, b) ((a) * (b))
#define TB_DIV(a, b) ((a) / (b))
#define TB_FMA(a, b, c) fma((a), (b), (c))
#define TB_SQRT(a) sqrt((a))
#define TB_ABS(a) fabs((a))
static inline unsigned long long tb_to_bits_(double x) { unsigned long long u; memcpy(&u, &x, 8); return u; }
static inline double tb_from_bits_(unsigned long long u) { double x; memcpy(&x, &u, 8); return x; }
#define TB_TO_BITS(x) tb_to_bits_((x))
#define TB_FROM_BITS(u) tb_from_bits_((u))
#endif

/* ---- double-double helpers ------------------------------------------------ */
typedef struct { double hi, lo; } tb_dd;

TB_HD tb_dd tb_two_sum(double a, double b) {
    tb_dd r;
    r.hi = TB_ADD(a, b);
    double bb = TB_SUB(r.hi, a);
    r.lo = TB_ADD(TB_SUB(a, TB_SUB(r.hi, bb)), TB_SUB(b, bb));
    return r;
}
/* requires |a| >= |b| (or a == 0) */
TB_HD tb_dd tb_fast_two_sum(double a, double b) {
    tb_dd r;
    r.hi = TB_ADD(a, b);
    r.lo = TB_SUB(b, TB_SUB(r.hi, a));
    return r;
}
TB_HD tb_dd tb_two_prod(double a, double b) {
    tb_dd r;
    r.hi = TB_MUL(a, b);
    r.lo = TB_FMA(a, b, -r.hi);
    return r;
}
TB_HD tb_dd tb_dd_add(tb_dd a, tb_dd b) {
    tb_dd s = tb_two_sum(a.hi, b.hi);
    double t = TB_ADD(TB_ADD(a.lo, b.lo), s.lo);
    return tb_fast_two_sum(s.hi, t);
}
TB_HD tb_dd tb_dd_add_d(tb_dd a, double b) {
    tb_dd s = tb_two_sum(a.hi, b);
    double t = TB_ADD(a.lo, s.lo);
    return tb_fast_two_sum(s.hi, t);
}
TB_HD tb_dd tb_dd_mul(tb_dd a, tb_dd b) {
    tb_dd p = tb_two_prod(a.hi, b.hi);
    double t = TB_ADD(p.lo, TB_ADD(TB_MUL(a.hi, b.lo), TB_MUL(a.lo, b.hi)));
    return tb_fast_two_sum(p.hi, t);
}
TB_HD tb_dd tb_dd_mul_d(tb_dd a, double b) {
    tb_dd p = tb_two_prod(a.hi, b);
    double t = TB_ADD(p.lo, TB_MUL(a.lo, b));
    return tb_fast_two_sum(p.hi, t);
}

/* 2^k for integer k in [-1022, 1023] */
TB_HD double tb_pow2i(int k) { return TB_FROM_BITS(((unsigned long long)(k + 1023)) << 52); }

/* ---- log in double-double --------------------------------------------------
 * x = 2^e * m, m in [sqrt(1/2), sqrt(2));  log m = 2 atanh(s), s = (m-1)/(m+1)
 * = 2s + 2 s z (1/3 + z/5 + z^2/7 + ...), z = s^2 <= 0.02944.
 * s, z and the first two series terms are carried in double-double, the tail
 * (1/7 ... 1/29) in double.  Relative error of the pair ~ 2^-66.
 * Domain: finite x > 0 (callers handle the rest).                            */
TB_HD tb_dd tb_log_dd_norm(double x, int e);
TB_HD tb_dd tb_log_dd(double x) {
    int e = 0;
    if ((TB_TO_BITS(x) >> 52) == 0ull) { /* subnormal */
        x = TB_MUL(x, 18014398509481984.0); /* 2^54 */
        e = -54;
    }
    return tb_log_dd_norm(x, e);
}
/* the branch-free body: x normal, positive, finite; result log(x) + e*ln2 */
TB_HD tb_dd tb_log_dd_norm(double x, int e) {
    unsigned long long u = TB_TO_BITS(x);
    e += (int)((u >> 52) & 0x7FFull) - 1023;
    double m = TB_FROM_BITS((u & 0x000FFFFFFFFFFFFFull) | 0x3FF0000000000000ull); /* [1,2) */
    if (m > 1.4142135623730951) { m = TB_MUL(m, 0.5); e += 1; }
    double f = TB_SUB(m, 1.0);            /* exact */
    tb_dd den = tb_two_sum(m, 1.0);
    /* s = f / den in double-double */
    tb_dd s;
    s.hi = TB_DIV(f, den.hi);
    double rem = TB_FMA(-s.hi, den.hi, f);
    rem = TB_SUB(rem, TB_MUL(s.hi, den.lo));
    s.lo = TB_DIV(rem, den.hi);
    s = tb_fast_two_sum(s.hi, s.lo);
    tb_dd z = tb_dd_mul(s, s);
    double zh = z.hi;
    /* tail t = z*(1/7 + z*(1/9 + ... + z/29)) in double */
    double t = 1.0 / 29.0;
    t = TB_FMA(t, zh, 1.0 / 27.0);
    t = TB_FMA(t, zh, 1.0 / 25.0);
    t = TB_FMA(t, zh, 1.0 / 23.0);
    t = TB_FMA(t, zh, 1.0 / 21.0);
    t = TB_FMA(t, zh, 1.0 / 19.0);
    t = TB_FMA(t, zh, 1.0 / 17.0);
    t = TB_FMA(t, zh, 1.0 / 15.0);
    t = TB_FMA(t, zh, 1.0 / 13.0);
    t = TB_FMA(t, zh, 1.0 / 11.0);
    t = TB_FMA(t, zh, 1.0 / 9.0);
    t = TB_FMA(t, zh, 1.0 / 7.0);
    t = TB_MUL(t, zh);
    /* u5 = 1/5 + t ; q = 1/3 + z*u5 (double-double) */
    tb_dd c5; c5.hi = 0.2; c5.lo = -1.1102230246251566e-17;               /* 1/5 */
    tb_dd c3; c3.hi = 0.3333333333333333; c3.lo = 1.850371707708594e-17;  /* 1/3 */
    tb_dd u5 = tb_dd_add_d(c5, t);
    tb_dd q = tb_dd_add(c3, tb_dd_mul(z, u5));
    /* logm = 2s + 2 s z q */
    tb_dd sz = tb_dd_mul(s, z);
    tb_dd corr = tb_dd_mul(sz, q);
    tb_dd logm = tb_dd_add(s, corr);
    logm.hi = TB_MUL(logm.hi, 2.0);
    logm.lo = TB_MUL(logm.lo, 2.0);
    /* + e*ln2 ; ln2_hi has 32 trailing zero bits so e*ln2_hi is exact */
    double de = (double)e;
    tb_dd eln2;
    eln2.hi = TB_MUL(de, 6.93147180369123816490e-01);
    eln2.lo = TB_MUL(de, 1.90821492927058770002e-10);
    tb_dd r = tb_two_sum(eln2.hi, logm.hi);
    double lo = TB_ADD(TB_ADD(eln2.lo, logm.lo), r.lo);
    return tb_fast_two_sum(r.hi, lo);
}

/* exp(h + l), |l| << |h|, result rounded to double; error < 0.6 ulp.         */
TB_HD double tb_exp_dd_core(double h, double l);
TB_HD double tb_exp_dd(double h, double l) {
    if (h != h) return h;
    if (h > 709.782712893384) return TB_FROM_BITS(0x7FF0000000000000ull);
    if (h < -745.2) return 0.0;
    return tb_exp_dd_core(h, l);
}
/* the branch-free body: -745.2 <= h <= 709.78 */
TB_HD double tb_exp_dd_core(double h, double l) {
    /* k = round(h / ln2) by the magic-number trick */
    double fn = TB_SUB(TB_FMA(h, 1.4426950408889634, 6755399441055744.0), 6755399441055744.0);
    int k = (int)fn;
    double rh = TB_FMA(-fn, 6.93147180369123816490e-01, h); /* exact */
    double rl = TB_FMA(-fn, 1.90821492927058770002e-10, l);
    tb_dd r = tb_fast_two_sum(rh, rl);
    double x = r.hi;
    /* p = sum_{n=2..15} x^n/n!  = x^2 * (1/2 + x*(1/6 + ...)) */
    double p = 1.0 / 1307674368000.0;           /* 1/15! */
    p = TB_FMA(p, x, 1.0 / 87178291200.0);      /* 1/14! */
    p = TB_FMA(p, x, 1.0 / 6227020800.0);       /* 1/13! */
    p = TB_FMA(p, x, 1.0 / 479001600.0);        /* 1/12! */
    p = TB_FMA(p, x, 1.0 / 39916800.0);         /* 1/11! */
    p = TB_FMA(p, x, 1.0 / 3628800.0);          /* 1/10! */
    p = TB_FMA(p, x, 1.0 / 362880.0);           /* 1/9!  */
    p = TB_FMA(p, x, 1.0 / 40320.0);            /* 1/8!  */
    p = TB_FMA(p, x, 1.0 / 5040.0);             /* 1/7!  */
    p = TB_FMA(p, x, 1.0 / 720.0);              /* 1/6!  */
    p = TB_FMA(p, x, 1.0 / 120.0);              /* 1/5!  */
    p = TB_FMA(p, x, 1.0 / 24.0);               /* 1/4!  */
    p = TB_FMA(p, x, 1.0 / 6.0);                /* 1/3!  */
    p = TB_FMA(p, x, 0.5);                      /* 1/2!  */
    p = TB_MUL(p, TB_MUL(x, x));
    /* exp(r) = 1 + x + (lo*(1+x) + p) */
    tb_dd s = tb_two_sum(1.0, x);
    double tail = TB_ADD(TB_ADD(s.lo, TB_FMA(r.lo, x, r.lo)), p);
    double y = TB_ADD(s.hi, tail);
    /* scale by 2^k in two exact steps to survive the subnormal range */
    int k1 = k / 2, k2 = k - k1;
    return TB_MUL(TB_MUL(y, tb_pow2i(k1)), tb_pow2i(k2));
}

TB_HD double tb_exp(double x) { return tb_exp_dd(x, 0.0); }

TB_HD double tb_log(double x) {
    if (x != x || x < 0.0) return TB_FROM_BITS(0x7FF8000000000000ull);
    if (x == 0.0) return TB_FROM_BITS(0xFFF0000000000000ull);
    if (TB_TO_BITS(x) == 0x7FF0000000000000ull) return x;
    tb_dd L = tb_log_dd(x);
    return L.hi;
}

/* x^y; x < 0 only with integer y (the cases the integrator meets: (eps/|x_k|)^(1/k), r2^1.5, x^3.0) */
TB_HD double tb_pow(double x, double y) {
    if (x != x || y != y) return TB_ADD(x, y);
    if (y == 0.0) return 1.0;
    if (x == 0.0) return (y > 0.0) ? 0.0 : TB_FROM_BITS(0x7FF0000000000000ull);
    double sgn = 1.0;
    if (x < 0.0) { /* negative base: defined for integer exponents only (Julia: DomainError otherwise) */
        if (TB_ABS(y) >= 9007199254740992.0) { x = -x; }
        else {
            const long long n = (long long)y;
            if ((double)n != y) return TB_FROM_BITS(0x7FF8000000000000ull);
            if (n & 1) sgn = -1.0;
            x = -x;
        }
    }
    if (TB_TO_BITS(x) == 0x7FF0000000000000ull) return (y > 0.0) ? TB_MUL(sgn, x) : 0.0;
    if (x == 1.0) return sgn;
    tb_dd L = tb_log_dd(x);
    tb_dd p = tb_two_prod(y, L.hi);
    p.lo = TB_FMA(y, L.lo, p.lo);
    return TB_MUL(sgn, tb_exp_dd(p.hi, p.lo));
}

/* Two independent powers x1^y1, x2^y2 in one straight-line body, so that the two dependent
 * chains interleave on the FP64 pipe (8-cycle dependent-issue latency on B200; the step-size
 * control calls pow twice per step, stepsize.jl:83-86).  Same operations as tb_pow, same bits;
 * anything outside the plain case (x normal, positive, finite; moderate result) goes to tb_pow. */
typedef struct { double a, b; } tb_pair;
TB_HD tb_pair tb_pow_pair(double x1, double y1, double x2, double y2) {
    tb_pair r;
    const unsigned long long e1 = (TB_TO_BITS(x1) >> 52), e2 = (TB_TO_BITS(x2) >> 52); /* sign + exponent fields */
    const int plain = e1 >= 1ull && e1 <= 0x7FEull && e2 >= 1ull && e2 <= 0x7FEull && x1 != 1.0 && x2 != 1.0 &&
                      y1 == y1 && y2 == y2 && y1 != 0.0 && y2 != 0.0;
    if (plain) {
        tb_dd L1 = tb_log_dd_norm(x1, 0), L2 = tb_log_dd_norm(x2, 0);
        tb_dd p1 = tb_two_prod(y1, L1.hi), p2 = tb_two_prod(y2, L2.hi);
        p1.lo = TB_FMA(y1, L1.lo, p1.lo);
        p2.lo = TB_FMA(y2, L2.lo, p2.lo);
        if (TB_ABS(p1.hi) < 700.0 && TB_ABS(p2.hi) < 700.0) {
            r.a = TB_MUL(1.0, tb_exp_dd_core(p1.hi, p1.lo));
            r.b = TB_MUL(1.0, tb_exp_dd_core(p2.hi, p2.lo));
            return r;
        }
    }
    r.a = tb_pow(x1, y1);
    r.b = tb_pow(x2, y2);
    return r;
}

/* ---- sin / cos --------------------------------------------------------------
 * Cody-Waite reduction by pi/2 in three 33-bit pieces (exact products for
 * |n| < 2^20), then Taylor kernels on |r| <= pi/4 with the reduced argument
 * carried as a pair (r, rl).                                                  */
TB_HD double tb_ksin(double x, double y) { /* sin(x + y), |x| <= pi/4 */
    double z = TB_MUL(x, x);
    double p = 1.0 / 51090942171709440000.0;         /*  1/21! */
    p = TB_FMA(p, z, -1.0 / 121645100408832000.0);   /* -1/19! */
    p = TB_FMA(p, z, 1.0 / 355687428096000.0);       /*  1/17! */
    p = TB_FMA(p, z, -1.0 / 1307674368000.0);        /* -1/15! */
    p = TB_FMA(p, z, 1.0 / 6227020800.0);            /*  1/13! */
    p = TB_FMA(p, z, -1.0 / 39916800.0);             /* -1/11! */
    p = TB_FMA(p, z, 1.0 / 362880.0);                /*  1/9!  */
    p = TB_FMA(p, z, -1.0 / 5040.0);                 /* -1/7!  */
    p = TB_FMA(p, z, 1.0 / 120.0);                   /*  1/5!  */
    /* sin = x + (x*z*(-1/6 + z*p) + y*(1 - z/2)) */
    double v = TB_MUL(z, x);
    double w = TB_FMA(z, p, -1.0 / 6.0);
    double c = TB_FMA(v, w, TB_MUL(y, TB_FMA(-0.5, z, 1.0)));
    return TB_ADD(x, c);
}
TB_HD double tb_kcos(double x, double y) { /* cos(x + y), |x| <= pi/4 */
    double z = TB_MUL(x, x);
    double p = 1.0 / 2432902008176640000.0;          /*  1/20! */
    p = TB_FMA(p, z, -1.0 / 6402373705728000.0);     /* -1/18! */
    p = TB_FMA(p, z, 1.0 / 20922789888000.0);        /*  1/16! */
    p = TB_FMA(p, z, -1.0 / 87178291200.0);          /* -1/14! */
    p = TB_FMA(p, z, 1.0 / 479001600.0);             /*  1/12! */
    p = TB_FMA(p, z, -1.0 / 3628800.0);              /* -1/10! */
    p = TB_FMA(p, z, 1.0 / 40320.0);                 /*  1/8!  */
    p = TB_FMA(p, z, -1.0 / 720.0);                  /* -1/6!  */
    p = TB_FMA(p, z, 1.0 / 24.0);                    /*  1/4!  */
    double hz = TB_MUL(0.5, z);
    double w = TB_SUB(1.0, hz);
    double zz = TB_MUL(z, z);
    double t = TB_FMA(zz, p, -TB_MUL(x, y));
    return TB_ADD(w, TB_ADD(TB_SUB(TB_SUB(1.0, w), hz), t));
}
TB_HD void tb_sincos(double x, double *sn, double *cs) {
    if (x != x || TB_ABS(x) > 1.0e300) { *sn = *cs = TB_FROM_BITS(0x7FF8000000000000ull); return; }
    double fn = TB_SUB(TB_FMA(x, 6.36619772367581382433e-01, 6755399441055744.0), 6755399441055744.0);
    /* pi/2 = p1 + p2 + p3 + p3t, p1,p2,p3 with 33 significant bits */
    double r = TB_FMA(-fn, 1.57079632673412561417e+00, x);   /* exact for |fn| < 2^20 */
    double w = TB_MUL(fn, 6.07710050630396597660e-11);       /* exact */
    tb_dd a = tb_two_sum(r, -w);
    double w2 = TB_MUL(fn, 2.02226624871116645580e-21);      /* exact */
    tb_dd b = tb_two_sum(a.hi, -w2);
    double lo = TB_ADD(TB_ADD(a.lo, b.lo), -TB_MUL(fn, 8.47842766036889956997e-32));
    tb_dd rr = tb_fast_two_sum(b.hi, lo);
    double s = tb_ksin(rr.hi, rr.lo);
    double c = tb_kcos(rr.hi, rr.lo);
    long long n = (long long)fn;
    switch ((int)(n & 3)) {
        case 0: *sn = s; *cs = c; break;
        case 1: *sn = c; *cs = -s; break;
        case 2: *sn = -s; *cs = -c; break;
        default: *sn = -c; *cs = s; break;
    }
}


/* ---- tan, sinh/cosh, tanh, atan, asin, acos: seeds of tan!, sinhcosh!, tanh!, atan!, asin!, acos! at k = 0 -------
 * Same contract as above (IEEE operations in a fixed order, identical on host and device); accuracy a few ulp
 * (tests/test_math.py), which is what the order-0 coefficient of these recurrences inherits.                    */
TB_HD double tb_tan(double x) { double s, c; tb_sincos(x, &s, &c); return TB_DIV(s, c); }
TB_HD void tb_sinhcosh(double x, double *sh, double *ch) {
    const double ax = TB_ABS(x);
    const double e = tb_exp(ax), ei = TB_DIV(1.0, e);
    *ch = TB_MUL(0.5, TB_ADD(e, ei));
    double s;
    if (ax < 0.5) { /* odd Taylor polynomial: no cancellation */
        const double z = TB_MUL(ax, ax);
        double p = 1.0 / 1307674368000.0;
        p = TB_FMA(p, z, 1.0 / 6227020800.0);
        p = TB_FMA(p, z, 1.0 / 39916800.0);
        p = TB_FMA(p, z, 1.0 / 362880.0);
        p = TB_FMA(p, z, 1.0 / 5040.0);
        p = TB_FMA(p, z, 1.0 / 120.0);
        p = TB_FMA(p, z, 1.0 / 6.0);
        s = TB_FMA(TB_MUL(ax, z), p, ax);
    } else s = TB_MUL(0.5, TB_SUB(e, ei));
    *sh = x < 0.0 ? -s : s;
}
TB_HD double tb_tanh(double x) {
    if (TB_ABS(x) > 22.0) return x < 0.0 ? -1.0 : 1.0;
    double s, c;
    tb_sinhcosh(x, &s, &c);
    return TB_DIV(s, c);
}
/* atan: |x| > 1 -> pi/2 - atan(1/x); t = c + u with c = j/8: atan(t) = atan(c) + atan((t - c)/(1 + t c)), |u'| <= 1/16 */
TB_HD double tb_atan(double x) {
    if (x != x) return x;
    const double ax = TB_ABS(x);
    const bool inv = ax > 1.0;
    const double t = inv ? TB_DIV(1.0, ax) : ax;
    const int j = (int)TB_ADD(TB_MUL(8.0, t), 0.5);
    const double c = TB_MUL(0.125, (double)j);
    const double u = TB_DIV(TB_SUB(t, c), TB_FMA(t, c, 1.0));
    const double z = TB_MUL(u, u);
    double p = 1.0 / 15.0;
    p = TB_FMA(TB_MUL(-1.0, p), z, 1.0 / 13.0);
    p = TB_FMA(TB_MUL(-1.0, p), z, 1.0 / 11.0);
    p = TB_FMA(TB_MUL(-1.0, p), z, 1.0 / 9.0);
    p = TB_FMA(TB_MUL(-1.0, p), z, 1.0 / 7.0);
    p = TB_FMA(TB_MUL(-1.0, p), z, 1.0 / 5.0);
    p = TB_FMA(TB_MUL(-1.0, p), z, 1.0 / 3.0);
    const double au = TB_FMA(TB_MUL(TB_MUL(-1.0, u), z), p, u); /* u - u^3/3 + u^5/5 - ... */
    double hi, lo;
    switch (j) {
        case 0: hi = TB_FROM_BITS(0x0000000000000000ULL); lo = TB_FROM_BITS(0x0000000000000000ULL); break;
        case 1: hi = TB_FROM_BITS(0x3fbfd5ba9aac2f6eULL); lo = TB_FROM_BITS(0xbc4cd37686760c17ULL); break;
        case 2: hi = TB_FROM_BITS(0x3fcf5b75f92c80ddULL); lo = TB_FROM_BITS(0x3c68ab6e3cf7afbdULL); break;
        case 3: hi = TB_FROM_BITS(0x3fd6f61941e4def1ULL); lo = TB_FROM_BITS(0xbc7c63aae6f6e918ULL); break;
        case 4: hi = TB_FROM_BITS(0x3fddac670561bb4fULL); lo = TB_FROM_BITS(0x3c7a2b7f222f65e2ULL); break;
        case 5: hi = TB_FROM_BITS(0x3fe1e00babdefeb4ULL); lo = TB_FROM_BITS(0xbc5928df287a668fULL); break;
        case 6: hi = TB_FROM_BITS(0x3fe4978fa3269ee1ULL); lo = TB_FROM_BITS(0x3c72419a87f2a458ULL); break;
        case 7: hi = TB_FROM_BITS(0x3fe700a7c5784634ULL); lo = TB_FROM_BITS(0xbc78c34d25aadef6ULL); break;
        case 8: hi = TB_FROM_BITS(0x3fe921fb54442d18ULL); lo = TB_FROM_BITS(0x3c81a62633145c07ULL); break;
        default: hi = 0.0; lo = 0.0; break;
    }
    double r = TB_ADD(hi, TB_ADD(lo, au));
    if (inv) r = TB_ADD(TB_FROM_BITS(0x3ff921fb54442d18ULL), TB_SUB(TB_FROM_BITS(0x3c91a62633145c07ULL), r));
    return x < 0.0 ? -r : r;
}
TB_HD double tb_asin(double x) {
    const double ax = TB_ABS(x);
    if (ax == 1.0) return x < 0.0 ? -TB_FROM_BITS(0x3ff921fb54442d18ULL) : TB_FROM_BITS(0x3ff921fb54442d18ULL);
    return tb_atan(TB_DIV(x, TB_SQRT(TB_MUL(TB_SUB(1.0, ax), TB_ADD(1.0, ax)))));
}
TB_HD double tb_acos(double x) {
    if (x == -1.0) return TB_MUL(2.0, TB_FROM_BITS(0x3ff921fb54442d18ULL));
    return TB_MUL(2.0, tb_atan(TB_SQRT(TB_DIV(TB_SUB(1.0, x), TB_ADD(1.0, x)))));
}

#endif /* TB_MATH_H */
