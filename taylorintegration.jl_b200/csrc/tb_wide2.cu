/*
 * tb_wide2.cu -- large-state ODEs with the jets in SHARED MEMORY (the many-spin system of
 * benchmark/manyspin_benchmarks.jl:10-41 at N = 100 spins: BASELINE.json config 4, "large-state jetcoeffs! in shared
 * memory").  Same drivers as tb_wide.cu (taylorinteg! src/integrator/taylorinteg.jl:329-381, 582-642; taylorstep!
 * src/integrator/taylorstep.jl:59-82; stepsize src/integrator/stepsize.jl:58-132; solcoeff! src/TaylorIntegration.jl:
 * 33-40) and the same recurrences in the oracle's order, so parity mode is bit-identical to tb_wide.cu and the oracle.
 *
 * What is different from tb_wide.cu (one thread per tape entry per dependency level, jets in a global workspace,
 * ncu: 21 stall cycles per issue on L1/L2 loads, 13 on barriers, FP64 pipe 1.7 % busy):
 *   - CLUSTERS.  Entries of one order that depend on each other (through a series computed in that same order) form a
 *     cluster; clusters only read each other's STATE jets, which solcoeff! finished one order earlier.  One thread
 *     runs a whole cluster in tape order, so an order needs ONE barrier (many-spin: cluster i = everything of spin i,
 *     100 clusters: the 100-term sum over the spins, two Cauchy products, a few scalings, three solcoeff!).
 *   - PERSISTENT vs TRANSIENT series.  Only series read at another index than the current order (Cauchy-product
 *     operands, state jets) are stored whole: [k][series] in shared memory, series numbered role-major so that the
 *     threads of a warp touch consecutive words (no bank conflicts); the coupling sum reads x_a[k] as a broadcast.
 *     Every other series is a scalar of the current order in a [slot][thread] scratch.  A negated series that only
 *     feeds Cauchy products is not stored: the product negates its operand (exact).  Many-spin 100: 400 series x 26
 *     doubles = 83 KB per CTA, two CTAs per SM.
 *   - GROUP PROGRAMS.  Isomorphic clusters share ONE copy of their entries (operands shift by the cluster's rank),
 *     held in shared memory with the per-cluster literals: every lane reads the same entry (broadcast).
 *   - coupling constants come from global memory term-major ([term][cluster]: coalesced, L1/L2 resident).
 * Bound: FP64 pipe; the coupling sum is one dependent DADD chain per thread (parity order).
 */
#include <cuda_runtime.h>

#include <algorithm>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <numeric>
#include <string>
#include <vector>

#include "tb_internal.h"
#include "tb_math.h"
#include "tb_wide_prog.h"

namespace tb {
void count_launch(long long n);
void set_last_kernel_ms(double ms);
}

namespace tbw {

#define TB_CUDA2(call)                                                         \
    do {                                                                       \
        cudaError_t e_ = (call);                                               \
        if (e_ != cudaSuccess) {                                               \
            tb::set_error(std::string(#call) + ": " + cudaGetErrorString(e_)); \
            return TB_ERR_CUDA;                                                \
        }                                                                      \
    } while (0)

/* location of a series: >= 0 persistent id (shared memory W[k*NS + id]); -1 a number; <= -2 transient slot -2 - loc */
struct C2Op {
    int op, res, a, b, aux;
    int flags; /* bit 0: operand a enters negated, bit 1: operand b */
    int para, parb;
    int lc_off, lc_n, lc_stride, lc_soff, lc_sstride;
    int pad;
    double imma, immb, imm;
};

struct W2Args {
    long long N;
    int dof, npar, P, NS, ncl, maxlen, ntr, tsid;
    const C2Op *ops; /* [maxlen x ncl], op-major */
    const double *lc_coef;
    const int *lc_ser;
    const int *xsid; /* [dof] persistent id of x[c] */
    const int *cl_len; /* [ncl] entries of cluster c (0: a padding slot) */
    int lc_smem, lc_count; /* coupling constants copied to shared memory (lc_count doubles) */
    int ngroups;           /* > 0: group programs */
    int g_smem, ngops, ngnum; /* group programs copied to shared memory */
    int g_first[8], g_count[8], g_off[8], g_len[8];
    const C2Op *gops;
    const double *gnum;
    const double *q0, *params, *trange;
    long long nt;
    double t0, tmax, abstol, reltol, minstep, maxstep;
    long long maxsteps;
    int params_per_traj;
    double *xfinal, *tfinal;
    long long *nsteps;
    int *status;
    double *xv_trange;
    unsigned long long *counter;
};

#define XM(a, b) __dmul_rn((a), (b))
#define XA(a, b) __dadd_rn((a), (b))
#define XS(a, b) __dsub_rn((a), (b))
#define XD(a, b) __ddiv_rn((a), (b))
template <bool FAST> __device__ __forceinline__ double mad2(double a, double b, double c) {
    return FAST ? fma(a, b, c) : __dadd_rn(c, __dmul_rn(a, b));
}

/* one entry of a cluster at order k: the recurrences of SURVEY App. B in the oracle's order (as tb_wide.cu exec_op) */
template <bool FAST>
__device__ __forceinline__ void exec2(const W2Args &A, double *__restrict__ W, double *__restrict__ tr, const double *par,
                                      const double *__restrict__ lcc, const bool lc_smem, const C2Op &o, const int k) {
    const int NS = A.NS;
    const bool sa = o.a == -1, sb = o.b == -1;
    const double na = o.para >= 0 ? par[o.para] : o.imma, nb = o.parb >= 0 ? par[o.parb] : o.immb;
    /* value of operand series `loc` at index j (transients exist at j == k only) */
    auto rd = [&](int loc, int j) -> double { return loc >= 0 ? W[j * NS + loc] : tr[(-2 - loc) * blockDim.x]; };
    int op = o.op;
    if (op == TB_OP_POW) {
        if (o.imm == 0.0) op = TB_OP_ONE;
        else if (o.imm == 1.0) op = TB_OP_IDENTITY;
        else if (o.imm == 2.0) op = TB_OP_SQR;
        else if (o.imm == 0.5) op = TB_OP_SQRT;
    }
    double r = 0.0;
    switch (op) {
        case W_NOP: return;
        case W_SOLCOEFF: /* x[k+1] = dx[k] / (k+1): TaylorIntegration.jl:33-40 */
            W[(k + 1) * NS + o.res] = FAST ? rd(o.a, k) * (1.0 / (double)(k + 1)) : XD(rd(o.a, k), (double)(k + 1));
            return;
        case TB_OP_IDENTITY: r = sa ? (k == 0 ? na : 0.0) : rd(o.a, k); break;
        case TB_OP_NEG: r = -rd(o.a, k); break;
        case TB_OP_ZERO: r = 0.0; break;
        case TB_OP_ONE: r = k == 0 ? 1.0 : 0.0; break;
        case TB_OP_ADD:
            if (sa && sb) r = k == 0 ? XA(na, nb) : 0.0;
            else if (sa) r = k == 0 ? XA(na, rd(o.b, 0)) : rd(o.b, k);
            else if (sb) r = k == 0 ? XA(rd(o.a, 0), nb) : rd(o.a, k);
            else r = XA(rd(o.a, k), rd(o.b, k));
            break;
        case TB_OP_SUB:
            if (sa && sb) r = k == 0 ? XS(na, nb) : 0.0;
            else if (sa) r = k == 0 ? XS(na, rd(o.b, 0)) : -rd(o.b, k);
            else if (sb) r = k == 0 ? XS(rd(o.a, 0), nb) : rd(o.a, k);
            else r = XS(rd(o.a, k), rd(o.b, k));
            break;
        case TB_OP_MUL:
            if (sa && sb) r = k == 0 ? XM(na, nb) : 0.0;
            else if (sa) r = XM(na, rd(o.b, k));
            else if (sb) r = XM(rd(o.a, k), nb);
            else { /* Cauchy product; a negated operand is negated on the fly (exact: a sign bit on the DMUL operand) */
                const double *pa = W + o.a, *pb = W + o.b + k * NS;
                const bool neg = ((o.flags & 1) != 0) != ((o.flags & 2) != 0);
                r = XM(neg ? -pa[0] : pa[0], pb[0]);
                if (FAST) {
                    for (int i = 1; i <= k; i++) { pa += NS; pb -= NS; r = fma(neg ? -pa[0] : pa[0], pb[0], r); }
                } else {
                    int i = 1;
                    for (; i + 2 <= k + 1; i += 2) { /* two products ahead of the dependent adds */
                        const double q0 = XM(pa[NS], pb[-NS]), q1 = XM(pa[2 * NS], pb[-2 * NS]);
                        pa += 2 * NS; pb -= 2 * NS;
                        r = XA(r, neg ? -q0 : q0); r = XA(r, neg ? -q1 : q1);
                    }
                    for (; i <= k; i++) { pa += NS; pb -= NS; const double q = XM(pa[0], pb[0]); r = XA(r, neg ? -q : q); }
                }
            }
            break;
        case TB_OP_SQR: {
            const double *a = W + o.a;
            if (k == 0) r = XM(a[0], a[0]);
            else {
                const int kodd = k & 1, kend = (k - 2 + kodd) / 2;
                r = XM(a[0], a[k * NS]);
                for (int i = 1; i <= kend; i++) r = mad2<FAST>(a[i * NS], a[(k - i) * NS], r);
                r = XM(2.0, r);
                if (!kodd) r = mad2<FAST>(a[(k / 2) * NS], a[(k / 2) * NS], r);
            }
            break;
        }
        case TB_OP_DIV:
            if (sa && sb) r = k == 0 ? XD(na, nb) : 0.0;
            else if (sb) r = XD(rd(o.a, k), nb);
            else if (k == 0) r = XD(sa ? na : rd(o.a, 0), W[o.b]);
            else {
                const double *b = W + o.b, *c = W + o.res;
                double s = XM(c[0], b[k * NS]);
                for (int i = 1; i <= k - 1; i++) s = mad2<FAST>(c[i * NS], b[(k - i) * NS], s);
                r = sa ? XD(-s, b[0]) : XD(XS(rd(o.a, k), s), b[0]);
            }
            break;
        case TB_OP_SQRT: {
            const double *c = W + o.res;
            if (k == 0) r = __dsqrt_rn(rd(o.a, 0));
            else {
                const int kodd = k & 1, kend = (k - 2 + kodd) / 2;
                double s = 0.0;
                for (int i = 1; i <= kend; i++) s = i == 1 ? XM(c[i * NS], c[(k - i) * NS]) : mad2<FAST>(c[i * NS], c[(k - i) * NS], s);
                double aux = XS(rd(o.a, k), XM(2.0, s));
                if (!kodd) aux = XS(aux, XM(c[(kend + 1) * NS], c[(kend + 1) * NS]));
                r = XD(aux, XM(2.0, c[0]));
            }
            break;
        }
        case TB_OP_POW: {
            const double *a = W + o.a, *c = W + o.res;
            if (k == 0) r = tb_pow(a[0], o.imm);
            else {
                double s = XM(XM(XM(o.imm, (double)k), c[0]), a[k * NS]);
                for (int i = 1; i <= k - 1; i++) {
                    const double rr = XS(XM(o.imm, (double)(k - i)), (double)i);
                    s = mad2<FAST>(XM(rr, c[i * NS]), a[(k - i) * NS], s);
                }
                r = XD(s, XM((double)k, a[0]));
            }
            break;
        }
        case TB_OP_EXP: {
            const double *a = W + o.a, *c = W + o.res;
            if (k == 0) r = tb_exp(a[0]);
            else {
                double s = XM(XM((double)k, a[k * NS]), c[0]);
                for (int i = 1; i <= k - 1; i++) s = mad2<FAST>(XM((double)(k - i), a[(k - i) * NS]), c[i * NS], s);
                r = XD(s, (double)k);
            }
            break;
        }
        case TB_OP_LOG: {
            const double *a = W + o.a, *c = W + o.res;
            if (k == 0) r = tb_log(a[0]);
            else if (k == 1) r = XD(a[NS], a[0]);
            else {
                double s = XM(XM((double)(k - 1), a[NS]), c[(k - 1) * NS]);
                for (int i = 2; i <= k - 1; i++) s = mad2<FAST>(XM((double)(k - i), a[i * NS]), c[(k - i) * NS], s);
                r = XD(XS(a[k * NS], XD(s, (double)k)), a[0]);
            }
            break;
        }
        case TB_OP_SINCOS: {
            const double *a = W + o.a;
            double *sn = W + o.res, *cs = W + o.aux;
            if (k == 0) { double s, c; tb_sincos(a[0], &s, &c); sn[0] = s; cs[0] = c; return; }
            double x = a[NS];
            double sk = XM(x, cs[(k - 1) * NS]), ck = XM(-x, sn[(k - 1) * NS]);
            for (int i = 2; i <= k; i++) {
                x = XM((double)i, a[i * NS]);
                sk = mad2<FAST>(x, cs[(k - i) * NS], sk);
                ck = FAST ? fma(-x, sn[(k - i) * NS], ck) : XS(ck, XM(x, sn[(k - i) * NS]));
            }
            sn[k * NS] = XD(sk, (double)k);
            cs[k * NS] = XD(ck, (double)k);
            return;
        }
        case W_LINCOMB: { /* ((start + c1 x1[k]) + c2 x2[k]) + ... : the fused add!/mul! chain, in tape order */
            const double *cf = lcc + o.lc_off;
            const double *Wk = W + k * NS;
            const int ls = o.lc_stride, n = o.lc_n;
            int j = 0;
            if (o.lc_sstride == 0) {
                /* the terms are consecutive stored series (x_1[k], x_2[k], ...: the state jets of the other clusters):
                 * no id table, the operand pointer just advances; products are formed one batch ahead of the adds */
                const double *xs = Wk + o.lc_soff;
                const int xstep = o.pad;
                if (sa) { const double m0 = XM(cf[0], xs[0]); r = k == 0 ? XA(na, m0) : m0; j = 1; cf += ls; xs += xstep; }
                else r = rd(o.a, k);
                if (FAST) {
                    for (; j < n; j++, cf += ls, xs += xstep) r = fma(cf[0], xs[0], r);
                } else {
                    double p0 = 0.0, p1 = 0.0, p2 = 0.0, p3 = 0.0;
                    bool have = false;
                    for (; j + 4 <= n; j += 4) {
                        const double q0 = XM(cf[0], xs[0]), q1 = XM(cf[ls], xs[xstep]);
                        const double q2 = XM(cf[2 * ls], xs[2 * xstep]), q3 = XM(cf[3 * ls], xs[3 * xstep]);
                        cf += 4 * ls; xs += 4 * xstep;
                        if (have) { r = XA(r, p0); r = XA(r, p1); r = XA(r, p2); r = XA(r, p3); }
                        p0 = q0; p1 = q1; p2 = q2; p3 = q3; have = true;
                    }
                    if (have) { r = XA(r, p0); r = XA(r, p1); r = XA(r, p2); r = XA(r, p3); }
                    for (; j < n; j++, cf += ls, xs += xstep) r = XA(r, XM(cf[0], xs[0]));
                }
            } else {
                const int *sr = A.lc_ser + o.lc_soff;
                const int ss = o.lc_sstride;
                if (sa) { const double m0 = XM(cf[0], Wk[__ldg(sr)]); r = k == 0 ? XA(na, m0) : m0; j = 1; cf += ls; sr += ss; }
                else r = rd(o.a, k);
                for (; j < n; j++, cf += ls, sr += ss) r = mad2<FAST>(cf[0], Wk[__ldg(sr)], r);
            }
            break;
        }
        default: break;
    }
    if (o.res >= 0) W[k * NS + o.res] = r;
    else tr[(-2 - o.res) * blockDim.x] = r;
}

__device__ __forceinline__ double w2_horner(const double *X, int NS, int P, double dt) {
    double s = X[P * NS];
    for (int k = P - 1; k >= 0; k--) s = XA(XM(s, dt), X[k * NS]);
    return s;
}

/* block-wide min (or max) that PROPAGATES NaN like Julia's min / max (stepsize.jl:74, 86) */
__device__ double w2_block_reduce(double v, double *red, bool take_max) {
    bool nan = v != v;
    for (int o = 16; o > 0; o >>= 1) {
        const double u = __shfl_xor_sync(0xffffffffu, v, o);
        v = take_max ? fmax(v, u) : fmin(v, u);
    }
    const int w = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
    nan = __syncthreads_or(nan);
    if ((threadIdx.x & 31) == 0) red[w] = v;
    __syncthreads();
    double r = red[0];
    for (int i = 1; i < nw; i++) r = take_max ? fmax(r, red[i]) : fmin(r, red[i]);
    return nan ? __longlong_as_double(0x7FF8000000000000LL) : r;
}

template <bool FAST> __global__ void __launch_bounds__(256, 1) wide2_kernel(const W2Args a) {
    extern __shared__ double sm2[];
    __shared__ double red[8];
    __shared__ unsigned long long s_idx;
    const int tid = threadIdx.x, nth = blockDim.x, P = a.P, NS = a.NS;
    double *W = sm2;                              /* [P+1][NS] persistent series */
    double *tr = W + (size_t)(P + 1) * NS + tid;  /* [ntr][nth] transient scalars of the current order */
    double *xnew = W + (size_t)(P + 1) * NS + (size_t)a.ntr * nth; /* [dof] */
    const double *lcc = a.lc_coef;
    double *extra = xnew + a.dof + 8;
    if (a.lc_smem) { /* loop-invariant coupling constants: once per CTA */
        for (int j = tid; j < a.lc_count; j += nth) extra[j] = a.lc_coef[j];
        lcc = extra;
        extra += a.lc_count;
    }
    const C2Op *gops = a.gops;
    const double *gnum = a.gnum;
    if (a.g_smem) {
        double *gn = extra;
        for (int j = tid; j < a.ngnum; j += nth) gn[j] = a.gnum[j];
        int *go = (int *)(gn + a.ngnum);
        const int *src = (const int *)a.gops;
        for (int j = tid; j < a.ngops * (int)(sizeof(C2Op) / 4); j += nth) go[j] = src[j];
        gnum = gn;
        gops = (const C2Op *)go;
    }
    __syncthreads();
    const double inf = __longlong_as_double(0x7FF0000000000000LL), nan = __longlong_as_double(0x7FF8000000000000LL);
    for (;;) {
        if (tid == 0) s_idx = atomicAdd(a.counter, 1ull);
        __syncthreads();
        const long long i = (long long)s_idx;
        __syncthreads();
        if (i >= a.N) break;
        double par[8];
        for (int j = 0; j < a.npar && j < 8; j++) par[j] = a.params_per_traj ? a.params[(long long)j * a.N + i] : a.params[j];
        for (int c = tid; c < a.dof; c += nth) W[a.xsid[c]] = a.q0[(long long)c * a.N + i];
        if (a.tsid >= 0) for (int k = tid; k <= P; k += nth) W[k * NS + a.tsid] = k == 1 ? 1.0 : 0.0;
        double t0 = a.trange ? a.trange[0] : a.t0;
        const double tmax = a.trange ? a.trange[a.nt - 1] : a.tmax;
        const double sign = copysign(1.0, XS(tmax, t0));
        long long nsteps = 1, iter = 2;
        double t1 = a.trange ? a.trange[1] : 0.0;
        int st = 0;
        if (a.trange && a.xv_trange)
            for (int c = tid; c < a.dof; c += nth) {
                a.xv_trange[((long long)c * a.nt) * a.N + i] = a.q0[(long long)c * a.N + i];
                for (long long j = 1; j < a.nt; j++) a.xv_trange[((long long)c * a.nt + j) * a.N + i] = nan;
            }
        __syncthreads();
        while (XM(sign, t0) < XM(sign, tmax)) {
            if (a.tsid >= 0 && tid == 0) W[a.tsid] = t0;
            __syncthreads();
            /* jetcoeffs!: for ord = 0:order-1 (parse_eqs.jl:259-264); every cluster in tape order, then ONE barrier */
            for (int k = 0; k < P; k++) {
                if (a.ngroups > 0) {
                    for (int g = 0; g < a.ngroups; g++)
                        for (int r = tid; r < a.g_count[g]; r += nth)
                            for (int j = 0; j < a.g_len[g]; j++) {
                                C2Op o = gops[a.g_off[g] + j]; /* the same entry for every lane: a broadcast from shared memory */
                                long long no;
                                memcpy(&no, &o.immb, sizeof no);
                                const double lit = ((o.a == -1 && o.para < 0) || (o.b == -1 && o.parb < 0)) ? gnum[no + r] : 0.0;
                                o.imma = lit; o.immb = lit;
                                if (o.flags & 16) o.res += r;
                                if (o.flags & 32) o.a += r;
                                if (o.flags & 64) o.b += r;
                                if (o.flags & 128) o.aux += r;
                                o.lc_off += r;
                                if (o.flags & 256) o.lc_soff += r;
                                exec2<FAST>(a, W, tr, par, lcc, a.lc_smem != 0, o, k);
                            }
                } else
                for (int c = tid; c < a.ncl; c += nth) {
                    const int len = __ldg(a.cl_len + c);
                    for (int j = 0; j < len; j++) exec2<FAST>(a, W, tr, par, lcc, a.lc_smem != 0, a.ops[(size_t)j * a.ncl + c], k);
                }
                __syncthreads();
            }
            /* stepsize: stepsize.jl:58-99 */
            double h = inf;
            for (int c = tid; c < a.dof; c += nth) {
                const double *X = W + a.xsid[c];
                const double x0norm = fabs(X[0]);
                for (int k = P - 1; k <= P; k++) {
                    const double ak = fabs(X[k * NS]);
                    if (ak == 0.0) continue;
                    const double rel = XM(a.reltol, x0norm);
                    const double eps = a.abstol >= rel ? a.abstol : rel;
                    const double hk = tb_pow(XD(eps, ak), XD(1.0, (double)k));
                    h = (h != h || hk != hk) ? nan : fmin(h, hk);
                }
            }
            h = w2_block_reduce(h, red, false);
            if (h == inf) { /* _second_stepsize: stepsize.jl:119-132, max over components */
                double h2 = 0.0;
                for (int c = tid; c < a.dof; c += nth) {
                    const double *X = W + a.xsid[c];
                    bool allzero = true;
                    for (int k = 0; k <= P; k++) if (X[k * NS] != 0.0) { allzero = false; break; }
                    double hc = inf;
                    if (!allzero) {
                        hc = 0.0;
                        for (int k = 1; k <= P - 2; k++) {
                            const double ak = fabs(X[k * NS]);
                            if (ak == 0.0) continue;
                            const double hk = tb_pow(XD(1.0, ak), XD(1.0, (double)k));
                            hc = (hc != hc || hk != hk) ? nan : fmax(hc, hk);
                        }
                    }
                    h2 = (h2 != h2 || hc != hc) ? nan : fmax(h2, hc);
                }
                h = w2_block_reduce(h2, red, true);
            }
            double dt = h < a.minstep ? a.minstep : (h > a.maxstep ? a.maxstep : h);
            if (dt == 0.0) { st |= TB_ST_ZERO_STEP; break; }
            const double rem = XS(tmax, t0);
            const double srem = XM(sign, rem);
            dt = XM(sign, (dt != dt || srem != srem) ? nan : fmin(dt, srem));
            bool fin = true;
            for (int c = tid; c < a.dof; c += nth) {
                const double s = w2_horner(W + a.xsid[c], NS, P, dt);
                xnew[c] = s;
                fin = fin && isfinite(s);
            }
            bool done = false;
            if (a.trange) {
                const double tnext = XA(t0, dt);
                while (XM(sign, t1) < XM(sign, tnext)) {
                    const double d1 = XS(t1, t0);
                    if (a.xv_trange)
                        for (int c = tid; c < a.dof; c += nth)
                            a.xv_trange[((long long)c * a.nt + (iter - 1)) * a.N + i] = w2_horner(W + a.xsid[c], NS, P, d1);
                    iter += 1;
                    t1 = a.trange[iter - 1];
                }
                if (dt == rem) {
                    if (a.xv_trange)
                        for (int c = tid; c < a.dof; c += nth) a.xv_trange[((long long)c * a.nt + (iter - 1)) * a.N + i] = xnew[c];
                    done = true;
                }
                t0 = tnext;
            } else
                t0 = XA(t0, dt);
            __syncthreads();
            for (int c = tid; c < a.dof; c += nth) W[a.xsid[c]] = xnew[c];
            if (done) break;
            nsteps += 1;
            if (nsteps > a.maxsteps) { st |= TB_ST_MAXSTEPS; break; }
            if (__syncthreads_or(!fin)) { st |= TB_ST_NONFINITE; break; }
        }
        __syncthreads();
        if (a.xfinal) for (int c = tid; c < a.dof; c += nth) a.xfinal[(long long)c * a.N + i] = W[a.xsid[c]];
        if (tid == 0) {
            if (a.tfinal) a.tfinal[i] = t0;
            if (a.nsteps) a.nsteps[i] = nsteps - 1;
            if (a.status) a.status[i] = st;
        }
        __syncthreads();
    }
}

/* ---- host: fused program -> clusters, persistent / transient placement ------------------------------------- */
struct Wide2 {
    int order = 0, mode = 0, dof = 0, npar = 0, NS = 0, ncl = 0, maxlen = 0, ntr = 0, tsid = -1, threads = 0, ctas = 1;
    size_t smem = 0;
    bool lc_smem = false, g_smem = false;
    std::vector<C2Op> ops;
    std::vector<double> lc_coef;
    std::vector<int> lc_ser, xsid, cl_len;
    std::vector<C2Op> gops;      /* group programs (empty: per-cluster entries) */
    std::vector<double> gnum;    /* literal of entry j for cluster rank r */
    std::vector<int> g_first, g_count, g_off, g_len;
    C2Op *d_gops = nullptr;
    double *d_gnum = nullptr;
    int device = -1;
    C2Op *d_ops = nullptr;
    double *d_lcc = nullptr;
    int *d_lcs = nullptr, *d_xs = nullptr, *d_len = nullptr;
};

static bool cauchy_op(const WOp &o, int which /*0 a, 1 b, 2 res, 3 aux*/) {
    const bool sa = o.a < 0, sb = o.b < 0;
    int op = o.op;
    if (op == TB_OP_POW) {
        if (o.imm == 0.0 || o.imm == 1.0) return false;
        if (o.imm == 2.0) op = TB_OP_SQR;
        else if (o.imm == 0.5) op = TB_OP_SQRT;
    }
    switch (op) {
        case TB_OP_MUL: return !sa && !sb && which <= 1;
        case TB_OP_SQR: return which == 0;
        case TB_OP_DIV: return !sb && (which == 1 || which == 2);
        case TB_OP_SQRT: return which == 2;
        case TB_OP_POW: case TB_OP_EXP: case TB_OP_LOG: return which == 0 || which == 2;
        case TB_OP_SINCOS: return which == 0 || which == 2 || which == 3;
        default: return false;
    }
}

Wide2 *wide2_build(const Program &p, std::string &why) {
    const int n = (int)p.ops.size(), nser = p.nser;
    if (p.npar > 8) { why = "more than 8 parameters"; return nullptr; }
    std::vector<int> producer(nser, -1), is_state(nser, 0), is_dx(nser, 0);
    for (int c = 0; c < p.dof; c++) { is_state[p.xser[c]] = 1; is_dx[p.dxser[c]] = 1; }
    for (int j = 0; j < n; j++) {
        producer[p.ops[j].res] = j;
        if (p.ops[j].op == TB_OP_SINCOS && p.ops[j].aux >= 0) producer[p.ops[j].aux] = j;
    }
    /* uses of every series: total, and as operand of a series x series product */
    std::vector<int> uses(nser, 0), mul_uses(nser, 0);
    for (int j = 0; j < n; j++) {
        const WOp &o = p.ops[j];
        const bool cm = o.op == TB_OP_MUL && o.a >= 0 && o.b >= 0;
        if (o.a >= 0) { uses[o.a]++; if (cm) mul_uses[o.a]++; }
        if (o.b >= 0) { uses[o.b]++; if (cm) mul_uses[o.b]++; }
        if (o.op == W_LINCOMB) for (int t = 0; t < o.lc_n; t++) uses[p.lc_ser[o.lc_off + t]]++;
    }
    /* a negation that only feeds Cauchy products is folded into them (-(y) * s: the operand is negated on the fly) */
    std::vector<int> alias(nser, -1), alias_neg(nser, 0), dead(n, 0);
    for (int j = 0; j < n; j++) {
        const WOp &o = p.ops[j];
        if (o.op != TB_OP_NEG || o.a < 0 || is_dx[o.res] || uses[o.res] == 0 || uses[o.res] != mul_uses[o.res]) continue;
        int src = o.a, neg = 1;
        if (alias[src] >= 0) { neg ^= alias_neg[src]; src = alias[src]; }
        alias[o.res] = src; alias_neg[o.res] = neg; dead[j] = 1;
    }
    auto resolve = [&](int s, int *neg) { *neg = 0; if (s >= 0 && alias[s] >= 0) { *neg = alias_neg[s]; return alias[s]; } return s; };
    /* persistent series: state jets, the time series, operands / results of recurrences */
    std::vector<int> persistent(nser, 0);
    for (int s = 0; s < nser; s++) if (is_state[s]) persistent[s] = 1;
    if (p.tser >= 0) persistent[p.tser] = 1;
    for (int j = 0; j < n; j++) {
        if (dead[j]) continue;
        const WOp &o = p.ops[j];
        int ng;
        if (o.a >= 0 && cauchy_op(o, 0)) persistent[resolve(o.a, &ng)] = 1;
        if (o.b >= 0 && cauchy_op(o, 1)) persistent[resolve(o.b, &ng)] = 1;
        if (cauchy_op(o, 2)) persistent[o.res] = 1;
        if (o.aux >= 0 && cauchy_op(o, 3)) persistent[o.aux] = 1;
    }
    /* clusters: union-find over the entries along same-order dependencies */
    std::vector<int> uf(n);
    std::iota(uf.begin(), uf.end(), 0);
    auto find = [&](int x) { while (uf[x] != x) { uf[x] = uf[uf[x]]; x = uf[x]; } return x; };
    auto link = [&](int j, int s) {
        int ng;
        s = resolve(s, &ng);
        if (s < 0 || producer[s] < 0) return;
        uf[find(j)] = find(producer[s]);
    };
    for (int j = 0; j < n; j++) {
        if (dead[j]) continue;
        const WOp &o = p.ops[j];
        link(j, o.a); link(j, o.b);
        if (o.op == W_LINCOMB) for (int t = 0; t < o.lc_n; t++) link(j, p.lc_ser[o.lc_off + t]);
    }
    std::map<int, int> cl_of_root;
    std::vector<std::vector<int>> cl_ops; /* tape indices, ascending */
    std::vector<int> cl_of_op(n, -1);
    for (int j = 0; j < n; j++) {
        if (dead[j]) continue;
        const int r = find(j);
        auto it = cl_of_root.find(r);
        if (it == cl_of_root.end()) { it = cl_of_root.emplace(r, (int)cl_ops.size()).first; cl_ops.emplace_back(); }
        cl_ops[it->second].push_back(j);
        cl_of_op[j] = it->second;
    }
    int ncl = (int)cl_ops.size();
    int threads_pick = 32;
    std::vector<int> grp_first, grp_count; /* isomorphic clusters, contiguous in the layout */
    if (ncl == 0) { why = "empty program"; return nullptr; }
    /* solcoeff! of component c runs in the cluster that produces dx[c] */
    std::vector<std::vector<int>> cl_sol(ncl);
    for (int c = 0; c < p.dof; c++) {
        const int pj = producer[p.dxser[c]];
        if (pj < 0 || dead[pj]) { why = "dx without a producer"; return nullptr; }
        cl_sol[cl_of_op[pj]].push_back(c);
    }
    /* Thread assignment.  Clusters with the same opcode sequence (the same role in the system: "spin i") are laid out
     * next to each other so that the lanes of a warp run the same entries; the groups are ordered by cost, the CTA gets
     * one thread per member of the most expensive group, and every group starts at a multiple of the CTA size, so that
     * thread t runs the t-th member of EVERY group (many-spin: the coupling cluster of spin t and its small
     * dq[3t-2] = -B y cluster).  Slots that pad a group hold an empty cluster.                                        */
    {
        std::vector<double> cost(ncl, 0.0);
        std::vector<std::string> sig(ncl);
        for (int c = 0; c < ncl; c++) {
            for (int j : cl_ops[c]) {
                const WOp &o = p.ops[j];
                sig[c] += std::to_string(o.op) + (o.a < 0 ? "n" : "s") + (o.b < 0 ? "n" : "s") + ",";
                cost[c] += o.op == W_LINCOMB ? o.lc_n : ((cauchy_op(o, 0) || cauchy_op(o, 1) || cauchy_op(o, 2)) ? p.order / 2.0 : 1.0);
            }
            sig[c] += "|" + std::to_string(cl_sol[c].size());
        }
        std::map<std::string, std::vector<int>> groups;
        for (int c = 0; c < ncl; c++) groups[sig[c]].push_back(c);
        std::vector<std::pair<double, std::string>> order;
        for (auto &kv : groups) {
            double w = 0.0;
            for (int c : kv.second) w = std::max(w, cost[c]);
            order.push_back({-w, kv.first});
        }
        std::sort(order.begin(), order.end());
        const int lead = (int)groups[order[0].second].size();
        threads_pick = std::max(32, std::min(256, ((lead + 31) / 32) * 32));
        std::vector<std::vector<int>> new_ops, new_sol;
        for (auto &og : order) {
            const std::vector<int> &mem = groups[og.second];
            grp_first.push_back((int)new_ops.size());
            grp_count.push_back((int)mem.size());
            for (int c : mem) { new_ops.push_back(cl_ops[c]); new_sol.push_back(cl_sol[c]); }
            if (&og != &order.back())
                while (new_ops.size() % (size_t)threads_pick) { new_ops.emplace_back(); new_sol.emplace_back(); }
        }
        cl_ops.swap(new_ops);
        cl_sol.swap(new_sol);
    }
    ncl = (int)cl_ops.size(); /* with the padding slots */
    /* role-major numbering of the persistent series: role r of the clusters that have it, in layout order */
    std::vector<int> pid(nser, -1), tslot(nser, -1);
    int nroles = 0, ntr = 0;
    std::vector<std::vector<int>> cl_pers(ncl);
    for (int c = 0; c < ncl; c++) {
        int nt = 0;
        auto own = [&](int s) {
            if (s < 0) return;
            if (persistent[s]) { if (std::find(cl_pers[c].begin(), cl_pers[c].end(), s) == cl_pers[c].end()) cl_pers[c].push_back(s); }
            else if (tslot[s] < 0) tslot[s] = nt++;
        };
        for (int comp : cl_sol[c]) own(p.xser[comp]);
        for (int j : cl_ops[c]) { own(p.ops[j].res); if (p.ops[j].op == TB_OP_SINCOS) own(p.ops[j].aux); }
        nroles = std::max(nroles, (int)cl_pers[c].size());
        ntr = std::max(ntr, nt);
    }
    int NS = 0, tsid = -1;
    for (int r = 0; r < nroles; r++)
        for (int c = 0; c < ncl; c++)
            if ((int)cl_pers[c].size() > r) pid[cl_pers[c][r]] = NS++;
    if (p.tser >= 0) { tsid = NS++; pid[p.tser] = tsid; }
    for (int s = 0; s < nser; s++)
        if (persistent[s] && pid[s] < 0) { why = "a stored series has no owner cluster"; return nullptr; }
    Wide2 *w = new Wide2;
    w->order = p.order; w->mode = p.mode; w->dof = p.dof; w->npar = p.npar; w->NS = NS; w->ncl = ncl; w->ntr = std::max(ntr, 1); w->tsid = tsid;
    for (int c = 0; c < p.dof; c++) w->xsid.push_back(pid[p.xser[c]]);
    int maxlen = 0;
    for (int c = 0; c < ncl; c++) {
        w->cl_len.push_back((int)(cl_ops[c].size() + cl_sol[c].size()));
        maxlen = std::max(maxlen, w->cl_len.back());
    }
    w->maxlen = maxlen;
    w->ops.assign((size_t)maxlen * ncl, C2Op{});
    auto loc = [&](int s, int *neg) -> int {
        *neg = 0;
        if (s < 0) return -1;
        s = resolve(s, neg);
        if (persistent[s]) return pid[s];
        return -2 - tslot[s];
    };
    std::string bad;
    /* linear-combination terms: coefficients term-major [t][cluster] per entry position, series ids shared when uniform */
    for (int pos = 0; pos < maxlen; pos++) {
        int longest = 0;
        bool uniform = true;
        const WOp *first = nullptr;
        for (int c = 0; c < ncl; c++) {
            if (pos >= (int)cl_ops[c].size()) continue;
            const WOp &o = p.ops[cl_ops[c][pos]];
            if (o.op != W_LINCOMB) continue;
            longest = std::max(longest, o.lc_n);
            if (!first) first = &o;
            else if (o.lc_n != first->lc_n) uniform = false;
            else for (int t = 0; t < o.lc_n && uniform; t++) uniform = p.lc_ser[o.lc_off + t] == p.lc_ser[first->lc_off + t];
        }
        /* clusters with a linear combination at this position: coefficients term-major over exactly those clusters */
        std::vector<int> lc_rank(ncl, -1);
        int nlc = 0;
        for (int c = 0; c < ncl; c++)
            if (pos < (int)cl_ops[c].size() && p.ops[cl_ops[c][pos]].op == W_LINCOMB) lc_rank[c] = nlc++;
        const size_t cbase = w->lc_coef.size(), sbase = w->lc_ser.size();
        if (longest > 0) {
            w->lc_coef.resize(cbase + (size_t)longest * nlc, 0.0);
            w->lc_ser.resize(sbase + (uniform ? (size_t)longest : (size_t)longest * ncl), 0);
        }
        for (int c = 0; c < ncl; c++) {
            C2Op &d = w->ops[(size_t)pos * ncl + c];
            d.op = W_NOP;
            const int nops_c = (int)cl_ops[c].size();
            if (pos >= nops_c + (int)cl_sol[c].size()) continue;
            if (pos >= nops_c) { /* solcoeff! */
                const int comp = cl_sol[c][pos - nops_c];
                int ng;
                d.op = W_SOLCOEFF; d.res = pid[p.xser[comp]]; d.a = loc(p.dxser[comp], &ng); d.b = -1; d.aux = -1; d.para = d.parb = -1;
                continue;
            }
            const WOp &o = p.ops[cl_ops[c][pos]];
            int na = 0, nb = 0, nx = 0;
            d.op = o.op; d.res = loc(o.res, &nx); d.a = loc(o.a, &na); d.b = loc(o.b, &nb); d.aux = o.aux >= 0 ? loc(o.aux, &nx) : -1;
            d.flags = (na ? 1 : 0) | (nb ? 2 : 0);
            d.para = o.para; d.parb = o.parb; d.imma = o.imma; d.immb = o.immb; d.imm = o.imm;
            if ((na || nb) && !(o.op == TB_OP_MUL && o.a >= 0 && o.b >= 0)) bad = "internal: negated operand outside a product";
            if (cauchy_op(o, 0) && d.a < -1) bad = "internal: recurrence operand is not stored";
            if (cauchy_op(o, 1) && d.b < -1) bad = "internal: recurrence operand is not stored";
            if (o.op == W_LINCOMB) {
                d.lc_n = o.lc_n; d.lc_off = (int)(cbase + lc_rank[c]); d.lc_stride = nlc;
                d.lc_soff = (int)(uniform ? sbase : sbase + c); d.lc_sstride = uniform ? 1 : ncl;
                d.pad = 0;
                for (int t = 0; t < o.lc_n; t++) {
                    int ng;
                    const int l = loc(p.lc_ser[o.lc_off + t], &ng);
                    if (l < 0 || ng) bad = "a linear-combination term is not a stored series"; /* terms are state jets in practice */
                    w->lc_coef[cbase + (size_t)t * nlc + lc_rank[c]] = p.lc_coef[o.lc_off + t];
                    w->lc_ser[uniform ? sbase + t : sbase + (size_t)t * ncl + c] = l;
                }
                {   /* equally spaced stored series (or a single term)? then the kernel walks a pointer, no id table */
                    auto sid_at = [&](int t) { return w->lc_ser[uniform ? sbase + t : sbase + (size_t)t * ncl + c]; };
                    const int s0 = sid_at(0), ds = o.lc_n >= 2 ? sid_at(1) - s0 : 0;
                    bool affine = true;
                    for (int t = 2; t < o.lc_n && affine; t++) affine = sid_at(t) == s0 + t * ds;
                    if (affine) { d.lc_sstride = 0; d.lc_soff = s0; d.pad = ds; }
                }
            }
        }
    }
    if (!bad.empty()) { why = bad; delete w; return nullptr; }
    /* GROUP PROGRAMS.  The clusters of a group run the same entries; with the role-major numbering their stored
     * operands differ by the cluster's rank in the group (or not at all), their linear-combination columns by the
     * rank, and their literal by a per-cluster table.  Then ONE copy of the entries serves the whole group: every
     * lane of a warp reads the same 80 bytes (a broadcast that stays in L1) instead of its own copy from L2 --
     * the per-cluster copies are 182 KB for many-spin 100 and their fetch latency was half of the step time.    */
    {
        bool ok = grp_count.size() <= 8;
        std::vector<C2Op> gops;
        std::vector<double> gnum;
        std::vector<int> goff, glen;
        for (size_t g = 0; g < grp_count.size() && ok; g++) {
            const int c0 = grp_first[g], cnt = grp_count[g], len = w->cl_len[c0];
            goff.push_back((int)gops.size());
            glen.push_back(len);
            for (int j = 0; j < len && ok; j++) {
                C2Op u = w->ops[(size_t)j * ncl + c0];
                int dflags = 0;
                const int num_off = (int)gnum.size();
                const bool sa = u.a == -1 && u.para < 0, sb = u.b == -1 && u.parb < 0;
                for (int r = 0; r < cnt && ok; r++) {
                    const C2Op &v = w->ops[(size_t)j * ncl + c0 + r];
                    auto delta = [&](int lu, int lv, int bit) {
                        if (lu < 0) { if (lv != lu) ok = false; return; }
                        const int dlt = lv - lu;
                        if (r == 1) { if (dlt == 1) dflags |= bit; else if (dlt != 0) ok = false; }
                        else if (dlt != ((dflags & bit) ? r : 0)) ok = false;
                    };
                    if (getenv("TB_WIDE_DEBUG") && r == 1) fprintf(stderr, "g%zu j%d op %d/%d fl %d/%d par %d/%d %d/%d n %d/%d ls %d/%d ss %d/%d pad %d/%d res %d/%d a %d/%d b %d/%d aux %d/%d off %d/%d soff %d/%d\n", g, j, u.op, v.op, u.flags, v.flags, u.para, v.para, u.parb, v.parb, u.lc_n, v.lc_n, u.lc_stride, v.lc_stride, u.lc_sstride, v.lc_sstride, u.pad, v.pad, u.res, v.res, u.a, v.a, u.b, v.b, u.aux, v.aux, u.lc_off, v.lc_off, u.lc_soff, v.lc_soff);
                    if (v.op != u.op || v.flags != u.flags || v.para != u.para || v.parb != u.parb || (u.op == TB_OP_POW && v.imm != u.imm) || v.lc_n != u.lc_n ||
                        v.lc_stride != u.lc_stride || v.lc_sstride != u.lc_sstride || v.pad != u.pad) { ok = false; break; }
                    delta(u.res, v.res, 16); delta(u.a, v.a, 32); delta(u.b, v.b, 64); delta(u.aux, v.aux, 128);
                    if (u.op == W_LINCOMB) {
                        if (v.lc_off != u.lc_off + r) ok = false;
                        if (u.lc_sstride == 0) delta(u.lc_soff, v.lc_soff, 256);          /* first operand series: fixed or by rank */
                        else if (u.lc_sstride == 1) { if (v.lc_soff != u.lc_soff) ok = false; } /* shared id table */
                        else ok = false;                                                  /* per-cluster id tables: keep per-cluster entries */
                    }
                    if (!ok && getenv("TB_WIDE_DEBUG")) fprintf(stderr, "FAIL g%zu j%d r%d: res %d/%d a %d/%d b %d/%d aux %d/%d dflags %d off %d/%d soff %d/%d\n", g, j, r, u.res, v.res, u.a, v.a, u.b, v.b, u.aux, v.aux, dflags, u.lc_off, v.lc_off, u.lc_soff, v.lc_soff);
                    gnum.push_back(sa ? v.imma : (sb ? v.immb : 0.0));
                }
                if (cnt == 1) dflags = 0;
                u.flags |= dflags;
                u.lc_soff = u.lc_soff;
                u.pad = u.pad;
                /* literal of cluster r at gnum[num_off + r]; stash the offset in the (unused here) immb slot as bits */
                long long no = num_off;
                memcpy(&u.immb, &no, sizeof no);
                gops.push_back(u);
            }
        }
        if (ok && !getenv("TB_WIDE_NO_GROUPS")) {
            w->gops.swap(gops); w->gnum.swap(gnum);
            for (size_t g = 0; g < grp_count.size(); g++) { w->g_first.push_back(grp_first[g]); w->g_count.push_back(grp_count[g]); }
            w->g_off = goff; w->g_len = glen;
        }
    }
    const int threads = threads_pick;
    w->threads = threads;
    w->smem = ((size_t)(p.order + 1) * NS + (size_t)w->ntr * threads + p.dof + 16) * sizeof(double);
    const size_t smem_max = 227 * 1024;
    /* the coupling constants are loop-invariant (every order of every step of every trajectory reads all of them):
     * they are copied to shared memory once per CTA when they fit beside the jets.  From global memory the dependent
     * add chain waits on L2 (measured on many-spin 100: 9.1e5 steps/s with 2 CTAs per SM and the table in L1/L2).  */
    w->lc_smem = !w->lc_coef.empty() && w->smem + w->lc_coef.size() * sizeof(double) + 1024 <= smem_max;
    /* ncu (profiles/r02_ncu_wide2_summary.txt): with the table in L1/L2 the multiplies of the coupling sum wait ~300
     * cycles per batch of loads, 25 batches per order; TB_WIDE_LC_GLOBAL=1 keeps it in global memory (2 CTAs per SM) */
    if (!getenv("TB_WIDE_LC_SMEM")) w->lc_smem = false; /* measured (many-spin 100, 2^14 members): 1.17e6 steps/s with the table in
                                                          * shared memory and 1 CTA per SM, 1.75e6 from L1/L2 with 2 CTAs per SM */
    if (w->lc_smem) w->smem += w->lc_coef.size() * sizeof(double);
    /* group programs and their literals live in shared memory as well (a few KB) */
    w->g_smem = !w->gops.empty() && w->smem + w->gops.size() * sizeof(C2Op) + w->gnum.size() * sizeof(double) + 1024 <= smem_max;
    if (w->g_smem) w->smem += w->gops.size() * sizeof(C2Op) + w->gnum.size() * sizeof(double);
    if (w->smem > smem_max) { why = "stored series do not fit shared memory (" + std::to_string(w->smem) + " B)"; delete w; return nullptr; }
    w->ctas = (int)std::max<size_t>(1, std::min<size_t>(smem_max / (w->smem + 1024), 2048 / threads));
    return w;
}

void wide2_free(Wide2 *w) {
    if (!w) return;
    if (w->d_ops) { cudaFree(w->d_ops); cudaFree(w->d_lcc); cudaFree(w->d_lcs); cudaFree(w->d_xs); cudaFree(w->d_len); cudaFree(w->d_gops); cudaFree(w->d_gnum); }
    delete w;
}

void wide2_info(const Wide2 *w, int *nclusters, int *npersist, int *ntransient, int *threads, int *smem_bytes, int *ctas_per_sm, int *ngroups) {
    if (!w) return;
    if (nclusters) *nclusters = w->ncl;
    if (npersist) *npersist = w->NS;
    if (ntransient) *ntransient = w->ntr;
    if (threads) *threads = w->threads;
    if (smem_bytes) *smem_bytes = (int)w->smem;
    if (ctas_per_sm) *ctas_per_sm = w->ctas;
    if (ngroups) *ngroups = w->gops.empty() ? 0 : (int)w->g_count.size();
}

int wide2_integrate(Wide2 *w, long long N, const double *q0, const double *params, double t0, double tmax,
                    const double *trange, long long nt, const tb_opts_t *o, const tb_out_t *out) {
    int dev = 0, nsm = 148;
    TB_CUDA2(cudaGetDevice(&dev));
    TB_CUDA2(cudaDeviceGetAttribute(&nsm, cudaDevAttrMultiProcessorCount, dev));
    const int dof = w->dof;
    const bool fast = w->mode == TB_MODE_FAST;
    if (!w->d_ops || w->device != dev) {
        if (w->d_ops) { cudaFree(w->d_ops); cudaFree(w->d_lcc); cudaFree(w->d_lcs); cudaFree(w->d_xs); cudaFree(w->d_len); cudaFree(w->d_gops); cudaFree(w->d_gnum); w->d_ops = nullptr; }
        w->device = dev;
        auto up = [&](void **dst, const void *src, size_t bytes) -> cudaError_t {
            cudaError_t e = cudaMalloc(dst, bytes ? bytes : 8);
            if (e != cudaSuccess) return e;
            return bytes ? cudaMemcpy(*dst, src, bytes, cudaMemcpyHostToDevice) : cudaSuccess;
        };
        TB_CUDA2(up((void **)&w->d_ops, w->ops.data(), sizeof(C2Op) * w->ops.size()));
        TB_CUDA2(up((void **)&w->d_lcc, w->lc_coef.data(), sizeof(double) * w->lc_coef.size()));
        TB_CUDA2(up((void **)&w->d_lcs, w->lc_ser.data(), sizeof(int) * w->lc_ser.size()));
        TB_CUDA2(up((void **)&w->d_xs, w->xsid.data(), sizeof(int) * dof));
        TB_CUDA2(up((void **)&w->d_len, w->cl_len.data(), sizeof(int) * w->cl_len.size()));
        TB_CUDA2(up((void **)&w->d_gops, w->gops.data(), sizeof(C2Op) * w->gops.size()));
        TB_CUDA2(up((void **)&w->d_gnum, w->gnum.data(), sizeof(double) * w->gnum.size()));
        if (fast) TB_CUDA2(cudaFuncSetAttribute(wide2_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)w->smem));
        else TB_CUDA2(cudaFuncSetAttribute(wide2_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)w->smem));
    }
    int bps = 1;
    if (fast) TB_CUDA2(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bps, wide2_kernel<true>, w->threads, w->smem));
    else TB_CUDA2(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bps, wide2_kernel<false>, w->threads, w->smem));
    bps = std::max(1, bps);
    const long long blocks = std::min<long long>((long long)nsm * bps, N);
    W2Args a{};
    a.N = N; a.dof = dof; a.npar = w->npar; a.P = w->order; a.NS = w->NS; a.ncl = w->ncl; a.maxlen = w->maxlen; a.ntr = w->ntr; a.tsid = w->tsid;
    a.ops = w->d_ops; a.lc_coef = w->d_lcc; a.lc_ser = w->d_lcs; a.xsid = w->d_xs; a.cl_len = w->d_len; a.lc_smem = w->lc_smem ? 1 : 0; a.lc_count = (int)w->lc_coef.size();
    a.ngroups = w->gops.empty() ? 0 : (int)w->g_count.size(); a.gops = w->d_gops; a.gnum = w->d_gnum;
    a.g_smem = w->g_smem ? 1 : 0; a.ngops = (int)w->gops.size(); a.ngnum = (int)w->gnum.size();
    for (int g = 0; g < a.ngroups; g++) { a.g_first[g] = w->g_first[g]; a.g_count[g] = w->g_count[g]; a.g_off[g] = w->g_off[g]; a.g_len[g] = w->g_len[g]; }
    a.nt = trange ? nt : 0; a.t0 = t0; a.tmax = tmax; a.abstol = o->abstol; a.reltol = o->reltol; a.minstep = o->minstepsize;
    a.maxstep = o->maxstepsize; a.maxsteps = o->maxsteps; a.params_per_traj = o->params_per_traj;
    const size_t nq = (size_t)dof * N, npar = (size_t)(w->npar > 0 ? w->npar : 1) * (o->params_per_traj ? N : 1);
    double *dq = nullptr, *dp = nullptr, *dtr = nullptr, *dxf = nullptr, *dtf = nullptr, *dxv = nullptr;
    long long *dns = nullptr; int *dst = nullptr; unsigned long long *dctr = nullptr;
    int rc = TB_OK;
    auto fail = [&](cudaError_t e, const char *what) { if (e != cudaSuccess && rc == TB_OK) { tb::set_error(std::string(what) + ": " + cudaGetErrorString(e)); rc = TB_ERR_CUDA; } };
    fail(cudaMalloc(&dq, 8 * nq), "cudaMalloc q0");
    fail(cudaMalloc(&dp, 8 * npar), "cudaMalloc params");
    fail(cudaMalloc(&dxf, 8 * nq), "cudaMalloc xfinal");
    fail(cudaMalloc(&dtf, 8 * N), "cudaMalloc tfinal");
    fail(cudaMalloc(&dns, 8 * N), "cudaMalloc nsteps");
    fail(cudaMalloc(&dst, 4 * N), "cudaMalloc status");
    fail(cudaMalloc(&dctr, 8), "cudaMalloc counter");
    if (trange) {
        fail(cudaMalloc(&dtr, 8 * nt), "cudaMalloc trange");
        if (out->xv_trange) fail(cudaMalloc(&dxv, 8 * nq * nt), "cudaMalloc xv_trange");
    }
    if (rc == TB_OK) {
        fail(cudaMemcpy(dq, q0, 8 * nq, cudaMemcpyHostToDevice), "copy q0");
        if (w->npar > 0 && params) fail(cudaMemcpy(dp, params, 8 * npar, cudaMemcpyHostToDevice), "copy params");
        if (trange) fail(cudaMemcpy(dtr, trange, 8 * nt, cudaMemcpyHostToDevice), "copy trange");
        fail(cudaMemset(dctr, 0, 8), "memset counter");
        a.q0 = dq; a.params = dp; a.trange = trange ? dtr : nullptr; a.xfinal = dxf; a.tfinal = dtf;
        a.nsteps = dns; a.status = dst; a.xv_trange = dxv; a.counter = dctr;
    }
    if (rc == TB_OK) {
        cudaEvent_t e0, e1;
        cudaEventCreate(&e0);
        cudaEventCreate(&e1);
        cudaEventRecord(e0, 0);
        if (fast) wide2_kernel<true><<<(unsigned)blocks, w->threads, w->smem>>>(a);
        else wide2_kernel<false><<<(unsigned)blocks, w->threads, w->smem>>>(a);
        cudaEventRecord(e1, 0);
        tb::count_launch(1);
        fail(cudaGetLastError(), "wide2_kernel launch");
        fail(cudaDeviceSynchronize(), "wide2_kernel");
        float ms = -1.f;
        cudaEventElapsedTime(&ms, e0, e1);
        tb::set_last_kernel_ms(ms);
        cudaEventDestroy(e0);
        cudaEventDestroy(e1);
    }
    if (rc == TB_OK) {
        if (out->xfinal) fail(cudaMemcpy(out->xfinal, dxf, 8 * nq, cudaMemcpyDeviceToHost), "copy xfinal");
        if (out->tfinal) fail(cudaMemcpy(out->tfinal, dtf, 8 * N, cudaMemcpyDeviceToHost), "copy tfinal");
        if (out->nsteps) fail(cudaMemcpy(out->nsteps, dns, 8 * N, cudaMemcpyDeviceToHost), "copy nsteps");
        if (out->status) fail(cudaMemcpy(out->status, dst, 4 * N, cudaMemcpyDeviceToHost), "copy status");
        if (out->xv_trange && dxv) fail(cudaMemcpy(out->xv_trange, dxv, 8 * nq * nt, cudaMemcpyDeviceToHost), "copy xv_trange");
    }
    cudaFree(dq); cudaFree(dp); cudaFree(dtr); cudaFree(dxf); cudaFree(dtf); cudaFree(dxv);
    cudaFree(dns); cudaFree(dst); cudaFree(dctr);
    return rc;
}

} // namespace tbw
