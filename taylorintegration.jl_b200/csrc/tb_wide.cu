/*
 * tb_wide.cu -- the ensemble hot path for LARGE-STATE ODEs (hundreds of components, tapes of 10^4
 * entries: the many-spin system of benchmark/manyspin_benchmarks.jl:10-41 at N = 100 spins).
 * Same driver as tb_device_drv.cuh (taylorinteg! src/integrator/taylorinteg.jl:329-381, 582-642;
 * taylorstep! src/integrator/taylorstep.jl:59-82; stepsize src/integrator/stepsize.jl:58-132), but
 * the mapping is one CTA per trajectory with the THREADS SPREAD OVER THE TAPE: the host sorts the
 * entries of one order into dependency levels, the threads of the CTA take the entries of a level
 * in turn and meet at a barrier.  Chains `s = s + c_a * x_a` (the inner `for a = 1:N` loop of
 * spin_odes!) are fused into one LINCOMB entry that adds its terms in tape order, so parity mode
 * performs exactly the oracle's operations.  Jets live in a per-CTA global-memory workspace
 * [series][k] that stays L1/L2 resident (many-spin N = 100: ~600 series x 26 doubles = 125 KB).
 */
#include <cuda_runtime.h>

#include <algorithm>
#include <cstdint>
#include <cstdlib>
#include <map>
#include <string>
#include <vector>

#include "tb_internal.h"
#include "tb_math.h"
#include "tb_wide_prog.h"

namespace tb {
void count_launch(long long n);
void set_last_kernel_ms(double ms);
}
struct tb_ode {
    tb::Ode o;
};

namespace {

#define TB_CUDA(call)                                                          \
    do {                                                                       \
        cudaError_t e_ = (call);                                               \
        if (e_ != cudaSuccess) {                                               \
            tb::set_error(std::string(#call) + ": " + cudaGetErrorString(e_)); \
            return TB_ERR_CUDA;                                                \
        }                                                                      \
    } while (0)

using tbw::WOp;
using tbw::W_LINCOMB;

struct WArgs {
    long long N;
    int dof, npar, nops, nlev, P, nser, tser;
    long long wstride; /* doubles of workspace per CTA: nser*(P+1) + dof */
    const WOp *ops;       /* sorted by level */
    const int *lev_off;   /* [nlev+1] */
    const double *lc_coef;
    const int *lc_ser;
    const int *xser, *dxser; /* [dof] series ids of x[c], dx[c] */
    const double *q0, *params, *trange;
    long long nt;
    double t0, tmax, abstol, reltol, minstep, maxstep;
    long long maxsteps;
    int params_per_traj;
    double *work;
    double *xfinal, *tfinal;
    long long *nsteps;
    int *status;
    double *xv_trange;
    unsigned long long *counter;
};

template <bool FAST> __device__ __forceinline__ double mad(double a, double b, double c) {
    return FAST ? fma(a, b, c) : __dadd_rn(c, __dmul_rn(a, b));
}
#define WM(a, b) __dmul_rn((a), (b))
#define WA(a, b) __dadd_rn((a), (b))
#define WS(a, b) __dsub_rn((a), (b))
#define WD(a, b) __ddiv_rn((a), (b))

/* one entry at order k, executed by ONE thread: the recurrences of SURVEY App. B in the oracle's order */
template <bool FAST>
__device__ void exec_op(const WArgs &A, double *__restrict__ W, const double *par, const WOp &o, int k) {
    const int np1 = A.P + 1;
    double *res = W + (size_t)o.res * np1;
    const bool sa = o.a < 0, sb = o.b < 0;
    const double *a = sa ? nullptr : W + (size_t)o.a * np1;
    const double *b = sb ? nullptr : W + (size_t)o.b * np1;
    const double na = o.para >= 0 ? par[o.para] : o.imma, nb = o.parb >= 0 ? par[o.parb] : o.immb;
    int op = o.op;
    if (op == TB_OP_POW) {
        if (o.imm == 0.0) op = TB_OP_ONE;
        else if (o.imm == 1.0) op = TB_OP_IDENTITY;
        else if (o.imm == 2.0) op = TB_OP_SQR;
        else if (o.imm == 0.5) op = TB_OP_SQRT;
    }
    double r = 0.0;
    switch (op) {
        case TB_OP_IDENTITY: r = sa ? (k == 0 ? na : 0.0) : a[k]; break;
        case TB_OP_NEG: r = -a[k]; break;
        case TB_OP_ZERO: r = 0.0; break;
        case TB_OP_ONE: r = k == 0 ? 1.0 : 0.0; break;
        case TB_OP_ADD:
            if (sa && sb) r = k == 0 ? WA(na, nb) : 0.0;
            else if (sa) r = k == 0 ? WA(na, b[0]) : b[k];
            else if (sb) r = k == 0 ? WA(a[0], nb) : a[k];
            else r = WA(a[k], b[k]);
            break;
        case TB_OP_SUB:
            if (sa && sb) r = k == 0 ? WS(na, nb) : 0.0;
            else if (sa) r = k == 0 ? WS(na, b[0]) : -b[k];
            else if (sb) r = k == 0 ? WS(a[0], nb) : a[k];
            else r = WS(a[k], b[k]);
            break;
        case TB_OP_MUL:
            if (sa && sb) r = k == 0 ? WM(na, nb) : 0.0;
            else if (sa) r = WM(na, b[k]);
            else if (sb) r = WM(a[k], nb);
            else {
                r = WM(a[0], b[k]);
                for (int i = 1; i <= k; i++) r = mad<FAST>(a[i], b[k - i], r);
            }
            break;
        case TB_OP_SQR:
            if (k == 0) r = WM(a[0], a[0]);
            else {
                const int kodd = k & 1, kend = (k - 2 + kodd) / 2;
                r = WM(a[0], a[k]);
                for (int i = 1; i <= kend; i++) r = mad<FAST>(a[i], a[k - i], r);
                r = WM(2.0, r);
                if (!kodd) r = mad<FAST>(a[k / 2], a[k / 2], r);
            }
            break;
        case TB_OP_DIV:
            if (sa && sb) r = k == 0 ? WD(na, nb) : 0.0;
            else if (sb) r = WD(a[k], nb);
            else if (k == 0) r = WD(sa ? na : a[0], b[0]);
            else {
                double s = WM(res[0], b[k]);
                for (int i = 1; i <= k - 1; i++) s = mad<FAST>(res[i], b[k - i], s);
                r = sa ? WD(-s, b[0]) : WD(WS(a[k], s), b[0]);
            }
            break;
        case TB_OP_SQRT:
            if (k == 0) r = __dsqrt_rn(a[0]);
            else {
                const int kodd = k & 1, kend = (k - 2 + kodd) / 2;
                double s = 0.0;
                for (int i = 1; i <= kend; i++) s = i == 1 ? WM(res[i], res[k - i]) : mad<FAST>(res[i], res[k - i], s);
                double aux = WS(a[k], WM(2.0, s));
                if (!kodd) aux = WS(aux, WM(res[kend + 1], res[kend + 1]));
                r = WD(aux, WM(2.0, res[0]));
            }
            break;
        case TB_OP_POW:
            if (k == 0) r = tb_pow(a[0], o.imm);
            else {
                double s = WM(WM(WM(o.imm, (double)k), res[0]), a[k]);
                for (int i = 1; i <= k - 1; i++) {
                    const double rr = WS(WM(o.imm, (double)(k - i)), (double)i);
                    s = mad<FAST>(WM(rr, res[i]), a[k - i], s);
                }
                r = WD(s, WM((double)k, a[0]));
            }
            break;
        case TB_OP_EXP:
            if (k == 0) r = tb_exp(a[0]);
            else {
                double s = WM(WM((double)k, a[k]), res[0]);
                for (int i = 1; i <= k - 1; i++) s = mad<FAST>(WM((double)(k - i), a[k - i]), res[i], s);
                r = WD(s, (double)k);
            }
            break;
        case TB_OP_LOG:
            if (k == 0) r = tb_log(a[0]);
            else if (k == 1) r = WD(a[1], a[0]);
            else {
                double s = WM(WM((double)(k - 1), a[1]), res[k - 1]);
                for (int i = 2; i <= k - 1; i++) s = mad<FAST>(WM((double)(k - i), a[i]), res[k - i], s);
                r = WD(WS(a[k], WD(s, (double)k)), a[0]);
            }
            break;
        case TB_OP_SINCOS: {
            double *cs = W + (size_t)o.aux * np1;
            if (k == 0) { double s, c; tb_sincos(a[0], &s, &c); res[0] = s; cs[0] = c; return; }
            double x = a[1];
            double sk = WM(x, cs[k - 1]), ck = WM(-x, res[k - 1]);
            for (int i = 2; i <= k; i++) {
                x = WM((double)i, a[i]);
                sk = mad<FAST>(x, cs[k - i], sk);
                ck = FAST ? fma(-x, res[k - i], ck) : WS(ck, WM(x, res[k - i]));
            }
            res[k] = WD(sk, (double)k);
            cs[k] = WD(ck, (double)k);
            return;
        }
        case W_LINCOMB: { /* ((start + c1 x1[k]) + c2 x2[k]) + ... : the fused add!/mul! chain, in tape order */
            const double *cf = A.lc_coef + o.lc_off;
            const int *sr = A.lc_ser + o.lc_off;
            const int ls = o.lc_stride;
            int j0 = 0;
            if (sa) { /* numeric start: add!(c, a::Number, b, k) touches k = 0 only */
                const double m0 = WM(__ldg(cf), W[(size_t)__ldg(sr) * np1 + k]);
                r = k == 0 ? WA(na, m0) : m0;
                j0 = 1;
            } else r = a[k];
            for (int j = j0; j < o.lc_n; j++) r = mad<FAST>(__ldg(cf + (size_t)j * ls), W[(size_t)__ldg(sr + (size_t)j * ls) * np1 + k], r);
            break;
        }
        default: break;
    }
    res[k] = r;
}

__device__ double wide_block_min(double v, double *red, bool take_max) {
    for (int o = 16; o > 0; o >>= 1) {
        const double u = __shfl_xor_sync(0xffffffffu, v, o);
        v = take_max ? fmax(v, u) : fmin(v, u);
    }
    const int w = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
    __syncthreads();
    if ((threadIdx.x & 31) == 0) red[w] = v;
    __syncthreads();
    double r = red[0];
    for (int i = 1; i < nw; i++) r = take_max ? fmax(r, red[i]) : fmin(r, red[i]);
    return r;
}

__device__ __forceinline__ double wide_horner(const double *X, int P, double dt) {
    double s = X[P];
    for (int k = P - 1; k >= 0; k--) s = WA(WM(s, dt), X[k]);
    return s;
}

template <bool FAST, int MINB> __global__ void __launch_bounds__(256, MINB) wide_kernel(const WArgs a) {
    __shared__ double red[8];
    __shared__ unsigned long long s_idx;
    const int tid = threadIdx.x, nth = blockDim.x, P = a.P, np1 = P + 1;
    double *W = a.work + (size_t)blockIdx.x * a.wstride;
    double *xnew = W + (size_t)a.nser * np1; /* [dof] scratch for the new state */
    const double inf = __longlong_as_double(0x7FF0000000000000LL), nan = __longlong_as_double(0x7FF8000000000000LL);
    for (;;) {
        if (tid == 0) s_idx = atomicAdd(a.counter, 1ull);
        __syncthreads();
        const long long i = (long long)s_idx;
        __syncthreads();
        if (i >= a.N) break;
        double par[8];
        for (int j = 0; j < a.npar && j < 8; j++) par[j] = a.params_per_traj ? a.params[(long long)j * a.N + i] : a.params[j];
        for (int c = tid; c < a.dof; c += nth) W[(size_t)a.xser[c] * np1] = a.q0[(long long)c * a.N + i];
        if (a.tser >= 0) for (int k = tid; k <= P; k += nth) W[(size_t)a.tser * np1 + k] = k == 1 ? 1.0 : 0.0;
        double t0 = a.trange ? a.trange[0] : a.t0;
        const double tmax = a.trange ? a.trange[a.nt - 1] : a.tmax;
        const double sign = copysign(1.0, WS(tmax, t0));
        long long nsteps = 1, iter = 2;
        double t1 = a.trange ? a.trange[1] : 0.0;
        int st = 0;
        if (a.trange && a.xv_trange)
            for (int c = tid; c < a.dof; c += nth) {
                a.xv_trange[((long long)c * a.nt) * a.N + i] = a.q0[(long long)c * a.N + i];
                for (long long j = 1; j < a.nt; j++) a.xv_trange[((long long)c * a.nt + j) * a.N + i] = nan;
            }
        __syncthreads();
        while (WM(sign, t0) < WM(sign, tmax)) {
            if (a.tser >= 0 && tid == 0) W[(size_t)a.tser * np1] = t0;
            __syncthreads();
            /* jetcoeffs!: for ord = 0:order-1 (parse_eqs.jl:259-264), entries by dependency level */
            for (int k = 0; k < P; k++) {
                for (int l = 0; l < a.nlev; l++) {
                    for (int j = a.lev_off[l] + tid; j < a.lev_off[l + 1]; j += nth) exec_op<FAST>(a, W, par, a.ops[j], k);
                    __syncthreads();
                }
                for (int c = tid; c < a.dof; c += nth) /* solcoeff!: TaylorIntegration.jl:33-40 */
                    W[(size_t)a.xser[c] * np1 + k + 1] = WD(W[(size_t)a.dxser[c] * np1 + k], (double)(k + 1));
                __syncthreads();
            }
            /* stepsize: stepsize.jl:58-99 */
            double h = inf;
            for (int c = tid; c < a.dof; c += nth) {
                const double *X = W + (size_t)a.xser[c] * np1;
                const double x0norm = fabs(X[0]);
                for (int k = P - 1; k <= P; k++) {
                    const double ak = fabs(X[k]);
                    if (ak == 0.0) continue;
                    const double rel = WM(a.reltol, x0norm);
                    const double eps = a.abstol >= rel ? a.abstol : rel;
                    h = fmin(h, tb_pow(WD(eps, ak), WD(1.0, (double)k)));
                }
            }
            h = wide_block_min(h, red, false);
            if (h == inf) { /* _second_stepsize: stepsize.jl:119-132, max over components */
                double h2 = 0.0;
                for (int c = tid; c < a.dof; c += nth) {
                    const double *X = W + (size_t)a.xser[c] * np1;
                    bool allzero = true;
                    for (int k = 0; k <= P; k++) if (X[k] != 0.0) { allzero = false; break; }
                    double hc = inf;
                    if (!allzero) {
                        hc = 0.0;
                        for (int k = 1; k <= P - 2; k++) {
                            const double ak = fabs(X[k]);
                            if (ak == 0.0) continue;
                            hc = fmax(hc, tb_pow(WD(1.0, ak), WD(1.0, (double)k)));
                        }
                    }
                    h2 = fmax(h2, hc);
                }
                h = wide_block_min(h2, red, true);
            }
            double dt = h < a.minstep ? a.minstep : (h > a.maxstep ? a.maxstep : h);
            if (dt == 0.0) { st |= TB_ST_ZERO_STEP; break; }
            const double rem = WS(tmax, t0);
            dt = WM(sign, fmin(dt, WM(sign, rem)));
            bool fin = true;
            for (int c = tid; c < a.dof; c += nth) {
                const double s = wide_horner(W + (size_t)a.xser[c] * np1, P, dt);
                xnew[c] = s;
                fin = fin && isfinite(s);
            }
            bool done = false;
            if (a.trange) {
                const double tnext = WA(t0, dt);
                while (WM(sign, t1) < WM(sign, tnext)) {
                    const double d1 = WS(t1, t0);
                    if (a.xv_trange)
                        for (int c = tid; c < a.dof; c += nth)
                            a.xv_trange[((long long)c * a.nt + (iter - 1)) * a.N + i] = wide_horner(W + (size_t)a.xser[c] * np1, P, d1);
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
                t0 = WA(t0, dt);
            __syncthreads();
            for (int c = tid; c < a.dof; c += nth) W[(size_t)a.xser[c] * np1] = xnew[c];
            if (done) break;
            nsteps += 1;
            if (nsteps > a.maxsteps) { st |= TB_ST_MAXSTEPS; break; }
            if (__syncthreads_or(!fin)) { st |= TB_ST_NONFINITE; break; }
        }
        __syncthreads();
        if (a.xfinal) for (int c = tid; c < a.dof; c += nth) a.xfinal[(long long)c * a.N + i] = W[(size_t)a.xser[c] * np1];
        if (tid == 0) {
            if (a.tfinal) a.tfinal[i] = t0;
            if (a.nsteps) a.nsteps[i] = nsteps - 1;
            if (a.status) a.status[i] = st;
        }
        __syncthreads();
    }
}

} // namespace

/* ---- host side: tape -> fused, levelled internal program ------------------------------------------ */
struct tb_wide {
    int order = 0, mode = 0, dof = 0, npar = 0, nser = 0, tser = -1, nlev = 0, maxwidth = 0;
    std::vector<WOp> ops;
    std::vector<int> lev_off, lc_ser, xser, dxser;
    std::vector<double> lc_coef;
    int device = -1;
    WOp *d_ops = nullptr;
    int *d_lev = nullptr, *d_lcs = nullptr, *d_xs = nullptr, *d_dxs = nullptr;
    double *d_lcc = nullptr;
    tbw::Wide2 *v2 = nullptr;  /* cluster kernel with jets in shared memory (tb_wide2.cu); null when it does not apply */
    std::string v2_why;
};

static bool w_num(int32_t o) { int k = TB_OPERAND_KIND(o); return k == TB_K_PAR || k == TB_K_IMM; }
static bool w_ser(int32_t o) { int k = TB_OPERAND_KIND(o); return k == TB_K_X || k == TB_K_DX || k == TB_K_TMP || k == TB_K_TIME; }

extern "C" {

tb_wide_t *tb_wide_create(const tb_ode_t *ode, int order, int mode) {
    if (!ode || order < 1 || order > 60) { tb::set_error("tb_wide_create: bad argument"); return nullptr; }
    const tb_ode_desc_t &d = ode->o.desc;
    const std::vector<tb_op_t> &T = ode->o.ops;
    if (d.npar > 8) { tb::set_error("tb_wide_create: npar <= 8 supported (use immediates for coupling constants)"); return nullptr; }
    const int n = d.nops;
    /* use counts and producers */
    std::map<int32_t, int> uses, producer;
    for (int j = 0; j < n; j++) {
        const tb_op_t &op = T[j];
        if (w_ser(op.a)) uses[op.a]++;
        if (w_ser(op.b)) uses[op.b]++;
        producer[op.res] = j;
        if (op.op == TB_OP_SINCOS) producer[op.aux] = j;
    }
    /* fuse s_a = s_{a-1} + c_a * x_a chains (the `for a` loop of spin_odes!, manyspin_benchmarks.jl:22-30) */
    struct Lin { int32_t start; bool start_num; int start_par; double start_imm; std::vector<double> c; std::vector<int32_t> x; std::vector<int> cpar; };
    std::vector<char> dead(n, 0);
    std::map<int, Lin> lin; /* tape index of the (current) last ADD of a chain -> chain */
    auto scaled = [&](int32_t o, double &c, int32_t &x) -> int { /* o = MUL(number, series) used once? returns its tape index */
        if (TB_OPERAND_KIND(o) != TB_K_TMP || uses[o] != 1 || !producer.count(o)) return -1;
        const tb_op_t &m = T[producer[o]];
        if (m.op != TB_OP_MUL) return -1;
        if (TB_OPERAND_KIND(m.a) == TB_K_IMM && w_ser(m.b)) { c = m.imm; x = m.b; return producer[o]; }
        if (TB_OPERAND_KIND(m.b) == TB_K_IMM && w_ser(m.a)) { c = m.imm; x = m.a; return producer[o]; }
        return -1;
    };
    for (int j = 0; j < n; j++) {
        const tb_op_t &op = T[j];
        if (op.op != TB_OP_ADD) continue;
        double c = 0.0; int32_t x = 0;
        int32_t prev = op.a;
        int mj = scaled(op.b, c, x);
        if (mj < 0) { mj = scaled(op.a, c, x); prev = op.b; }
        if (mj < 0) continue;
        /* extend the chain that produced `prev` (if prev is used only here), else start one */
        int pj = (w_ser(prev) && TB_OPERAND_KIND(prev) == TB_K_TMP && uses[prev] == 1 && producer.count(prev)) ? producer[prev] : -1;
        Lin L;
        if (pj >= 0 && lin.count(pj)) { L = std::move(lin[pj]); lin.erase(pj); dead[pj] = 1; }
        else {
            L.start = prev; L.start_num = w_num(prev);
            L.start_par = TB_OPERAND_KIND(prev) == TB_K_PAR ? TB_OPERAND_INDEX(prev) : -1;
            L.start_imm = op.imm;
        }
        L.c.push_back(c); L.x.push_back(x);
        dead[mj] = 1;
        lin[j] = std::move(L);
    }
    /* compact series ids over what the surviving program touches */
    std::map<int32_t, int> sid;
    auto id = [&](int32_t o) -> int {
        if (!w_ser(o)) return -1;
        auto it = sid.find(o);
        if (it != sid.end()) return it->second;
        int v = (int)sid.size();
        sid[o] = v;
        return v;
    };
    tb_wide *w = new tb_wide;
    w->order = order; w->mode = mode; w->dof = d.dof; w->npar = d.npar;
    for (int c = 0; c < d.dof; c++) { w->xser.push_back(id(TB_OPERAND(TB_K_X, c))); }
    for (int c = 0; c < d.dof; c++) { w->dxser.push_back(id(TB_OPERAND(TB_K_DX, c))); }
    std::vector<WOp> prog;
    std::vector<int> level;
    std::map<int, int> lev_of_ser; /* series id -> level of its producer within an order (inputs: -1) */
    auto lv = [&](int s) { auto it = lev_of_ser.find(s); return it == lev_of_ser.end() ? -1 : it->second; };
    std::string bad;
    for (int j = 0; j < n && bad.empty(); j++) {
        if (dead[j]) continue;
        const tb_op_t &op = T[j];
        WOp o{};
        o.op = op.op; o.res = id(op.res); o.a = id(op.a); o.b = id(op.b); o.aux = id(op.aux);
        o.para = TB_OPERAND_KIND(op.a) == TB_K_PAR ? TB_OPERAND_INDEX(op.a) : -1;
        o.parb = TB_OPERAND_KIND(op.b) == TB_K_PAR ? TB_OPERAND_INDEX(op.b) : -1;
        o.imma = o.immb = o.imm = op.imm;
        int L = -1;
        if (lin.count(j)) {
            const Lin &c = lin[j];
            o.op = W_LINCOMB;
            o.a = c.start_num ? -1 : id(c.start);
            o.para = c.start_par; o.imma = c.start_imm; o.b = -1;
            o.lc_off = (int)w->lc_coef.size(); o.lc_n = (int)c.c.size();
            for (size_t t = 0; t < c.c.size(); t++) {
                w->lc_coef.push_back(c.c[t]);
                const int s = id(c.x[t]);
                w->lc_ser.push_back(s);
                L = std::max(L, lv(s));
            }
        }
        switch (o.op) {
            case TB_OP_IDENTITY: case TB_OP_NEG: case TB_OP_ZERO: case TB_OP_ONE: case TB_OP_ADD: case TB_OP_SUB: case TB_OP_MUL:
            case TB_OP_DIV: case TB_OP_SQR: case TB_OP_SQRT: case TB_OP_POW: case TB_OP_EXP: case TB_OP_LOG: case TB_OP_SINCOS:
            case W_LINCOMB: break;
            default: bad = "tape entry " + std::to_string(j) + " (opcode " + std::to_string(op.op) + ") has no wide-kernel implementation"; break;
        }
        if (o.a >= 0) L = std::max(L, lv(o.a));
        if (o.b >= 0) L = std::max(L, lv(o.b));
        L += 1;
        lev_of_ser[o.res] = L;
        if (o.op == TB_OP_SINCOS) lev_of_ser[o.aux] = L;
        prog.push_back(o);
        level.push_back(L);
    }
    if (!bad.empty()) { tb::set_error("tb_wide_create: " + bad); delete w; return nullptr; }
    for (int c = 0; c < d.dof; c++)
        if (!lev_of_ser.count(w->dxser[c])) { tb::set_error("tb_wide_create: tape does not define dx[" + std::to_string(c) + "]"); delete w; return nullptr; }
    w->tser = sid.count(TB_OPERAND(TB_K_TIME, 0)) ? sid[TB_OPERAND(TB_K_TIME, 0)] : -1;
    w->nser = (int)sid.size();
    int nlev = 0;
    for (int L : level) nlev = std::max(nlev, L + 1);
    w->nlev = nlev;
    w->lev_off.assign(nlev + 1, 0);
    for (int L : level) w->lev_off[L + 1]++;
    for (int l = 0; l < nlev; l++) { w->maxwidth = std::max(w->maxwidth, w->lev_off[l + 1]); w->lev_off[l + 1] += w->lev_off[l]; }
    {   /* the same fused program, in tape order, for the shared-memory cluster kernel */
        tbw::Program pg;
        pg.order = order; pg.mode = mode; pg.dof = d.dof; pg.npar = d.npar; pg.nser = w->nser; pg.tser = w->tser;
        pg.ops = prog;
        for (tbw::WOp &o : pg.ops) o.lc_stride = 1;
        pg.lc_coef = w->lc_coef; pg.lc_ser = w->lc_ser; pg.xser = w->xser; pg.dxser = w->dxser;
        w->v2 = tbw::wide2_build(pg, w->v2_why);
    }
    w->ops.resize(prog.size());
    std::vector<int> fill(w->lev_off.begin(), w->lev_off.end() - 1);
    for (size_t j = 0; j < prog.size(); j++) w->ops[fill[level[j]]++] = prog[j]; /* stable inside a level */
    /* interleave the chains of one level: term j of the p-th chain at base + j*W + p, so the threads of a warp
     * (consecutive entries of the level) read consecutive coefficients                                       */
    std::vector<double> cT;
    std::vector<int> sT;
    for (int l = 0; l < nlev; l++) {
        std::vector<int> chains;
        int longest = 0;
        for (int j = w->lev_off[l]; j < w->lev_off[l + 1]; j++)
            if (w->ops[j].op == W_LINCOMB) { chains.push_back(j); longest = std::max(longest, w->ops[j].lc_n); }
        const int Wd = (int)chains.size();
        const size_t base = cT.size();
        cT.resize(base + (size_t)Wd * longest, 0.0);
        sT.resize(base + (size_t)Wd * longest, 0);
        for (int pidx = 0; pidx < Wd; pidx++) {
            WOp &o = w->ops[chains[pidx]];
            for (int t = 0; t < o.lc_n; t++) {
                cT[base + (size_t)t * Wd + pidx] = w->lc_coef[o.lc_off + t];
                sT[base + (size_t)t * Wd + pidx] = w->lc_ser[o.lc_off + t];
            }
            o.lc_off = (int)(base + pidx);
            o.lc_stride = Wd;
        }
    }
    w->lc_coef.swap(cT);
    w->lc_ser.swap(sT);
    return w;
}

void tb_wide_free(tb_wide_t *w) {
    if (!w) return;
    if (w->d_ops) { cudaFree(w->d_ops); cudaFree(w->d_lev); cudaFree(w->d_lcs); cudaFree(w->d_xs); cudaFree(w->d_dxs); cudaFree(w->d_lcc); }
    tbw::wide2_free(w->v2);
    delete w;
}

/* program statistics: entries after fusion, dependency levels per order, stored series, widest level */
int tb_wide_info(const tb_wide_t *w, int *nops, int *nlev, int *nser, int *maxwidth) {
    if (!w) return TB_ERR_ARG;
    if (nops) *nops = (int)w->ops.size();
    if (nlev) *nlev = w->nlev;
    if (nser) *nser = w->nser;
    if (maxwidth) *maxwidth = w->maxwidth;
    return TB_OK;
}

/* layout of the shared-memory cluster kernel; TB_ERR_UNSUPPORTED (reason in tb_last_error) when it does not apply */
int tb_wide_layout(const tb_wide_t *w, int *nclusters, int *npersistent, int *ntransient, int *threads, int *smem_bytes, int *ctas_per_sm,
                   int *ngroups) {
    if (!w) return TB_ERR_ARG;
    if (!w->v2) { tb::set_error("tb_wide_layout: " + w->v2_why); return TB_ERR_UNSUPPORTED; }
    tbw::wide2_info(w->v2, nclusters, npersistent, ntransient, threads, smem_bytes, ctas_per_sm, ngroups);
    return TB_OK;
}

/* the NVRTC-specialised build of the cluster kernel: 1 / registers / stack bytes / coupling constants in shared memory,
 * or generated = 0 when the interpreter runs (no groups of isomorphic clusters, or TB_WIDE_INTERP=1) */
int tb_wide_generated(const tb_wide_t *w, int *generated, int *regs, int *local_bytes, int *lc_in_smem) {
    if (!w) return TB_ERR_ARG;
    if (!w->v2) { tb::set_error("tb_wide_generated: " + w->v2_why); return TB_ERR_UNSUPPORTED; }
    tbw::wide2_gen_info(w->v2, generated, regs, local_bytes, lc_in_smem);
    return TB_OK;
}
const char *tb_wide_source(const tb_wide_t *w) { return (w && w->v2) ? tbw::wide2_gen_source(w->v2) : nullptr; }
const char *tb_wide_log(const tb_wide_t *w) { return (w && w->v2) ? tbw::wide2_gen_log(w->v2) : nullptr; }

int tb_wide_integrate(tb_wide_t *w, int64_t N, const double *q0, const double *params, double t0, double tmax,
                      const double *trange, int64_t nt, const tb_opts_t *o, const tb_out_t *out) {
    if (!w || !q0 || !o || !out || N < 0 || (trange && nt < 2)) { tb::set_error("tb_wide_integrate: bad argument"); return TB_ERR_ARG; }
    if (out->tv || out->xv || out->psol || out->tevents || out->lambda) { tb::set_error("tb_wide_integrate: only xfinal/tfinal/nsteps/status/xv_trange outputs"); return TB_ERR_UNSUPPORTED; }
    if (N == 0) return TB_OK;
    if (w->npar > 0 && !params) { tb::set_error("tb_wide_integrate: params is NULL"); return TB_ERR_ARG; }
    /* jets in shared memory when the stored series fit (tb_wide2.cu); TB_WIDE_V1=1 forces the global-workspace kernel */
    const char *v1 = getenv("TB_WIDE_V1");
    if (w->v2 && !(v1 && v1[0] == '1')) return tbw::wide2_integrate(w->v2, N, q0, params, t0, tmax, trange, nt, o, out);
    int dev = 0, nsm = 148;
    TB_CUDA(cudaGetDevice(&dev));
    TB_CUDA(cudaDeviceGetAttribute(&nsm, cudaDevAttrMultiProcessorCount, dev));
    const int P = w->order, np1 = P + 1, dof = w->dof;
    if (!w->d_ops || w->device != dev) {
        w->device = dev;
        auto up = [&](void **dst, const void *src, size_t bytes) -> cudaError_t {
            cudaError_t e = cudaMalloc(dst, bytes ? bytes : 8);
            if (e != cudaSuccess) return e;
            return bytes ? cudaMemcpy(*dst, src, bytes, cudaMemcpyHostToDevice) : cudaSuccess;
        };
        TB_CUDA(up((void **)&w->d_ops, w->ops.data(), sizeof(WOp) * w->ops.size()));
        TB_CUDA(up((void **)&w->d_lev, w->lev_off.data(), sizeof(int) * w->lev_off.size()));
        TB_CUDA(up((void **)&w->d_lcs, w->lc_ser.data(), sizeof(int) * w->lc_ser.size()));
        TB_CUDA(up((void **)&w->d_lcc, w->lc_coef.data(), sizeof(double) * w->lc_coef.size()));
        TB_CUDA(up((void **)&w->d_xs, w->xser.data(), sizeof(int) * dof));
        TB_CUDA(up((void **)&w->d_dxs, w->dxser.data(), sizeof(int) * dof));
    }
    const bool fast = w->mode == TB_MODE_FAST;
    int threads = ((std::max(w->maxwidth, 32) + 31) / 32) * 32;
    if (threads > 256) threads = 256;
    /* the kernel waits on L1/L2 loads of its global-memory workspace (ncu: long_scoreboard 21 per issue at 2 CTAs
     * per SM), so the register budget is capped to fit 8 CTAs (64 warps) per SM: many-spin 1.6e5 -> 3.2e5 (4 CTAs)
     * -> 4.3e5 (8 CTAs) trajectory-steps/s at 2048 members.  TB_WIDE_MINB=2|4 select the wider register budgets. */
    int minb = 8;
    if (const char *e = getenv("TB_WIDE_MINB")) { const int v = atoi(e); minb = v == 2 ? 2 : (v == 4 ? 4 : 8); }
    int bps = 1;
    if (fast && minb == 8) TB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bps, wide_kernel<true, 8>, threads, 0));
    else if (minb == 8) TB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bps, wide_kernel<false, 8>, threads, 0));
    else if (fast && minb == 4) TB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bps, wide_kernel<true, 4>, threads, 0));
    else if (fast) TB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bps, wide_kernel<true, 2>, threads, 0));
    else if (minb == 4) TB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bps, wide_kernel<false, 4>, threads, 0));
    else TB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bps, wide_kernel<false, 2>, threads, 0));
    bps = std::max(1, std::min(bps, 8));
    long long blocks = std::min<long long>((long long)nsm * bps, N);
    WArgs a{};
    a.N = N; a.dof = dof; a.npar = w->npar; a.nops = (int)w->ops.size(); a.nlev = w->nlev; a.P = P; a.nser = w->nser; a.tser = w->tser;
    a.wstride = (long long)w->nser * np1 + dof;
    a.ops = w->d_ops; a.lev_off = w->d_lev; a.lc_coef = w->d_lcc; a.lc_ser = w->d_lcs; a.xser = w->d_xs; a.dxser = w->d_dxs;
    a.nt = trange ? nt : 0; a.t0 = t0; a.tmax = tmax; a.abstol = o->abstol; a.reltol = o->reltol; a.minstep = o->minstepsize;
    a.maxstep = o->maxstepsize; a.maxsteps = o->maxsteps; a.params_per_traj = o->params_per_traj;
    const size_t nq = (size_t)dof * N, npar = (size_t)(w->npar > 0 ? w->npar : 1) * (o->params_per_traj ? N : 1);
    const size_t wbytes = 8 * (size_t)blocks * a.wstride;
    double *dq = nullptr, *dp = nullptr, *dtr = nullptr, *dwork = nullptr, *dxf = nullptr, *dtf = nullptr, *dxv = nullptr;
    long long *dns = nullptr; int *dst = nullptr; unsigned long long *dctr = nullptr;
    int rc = TB_OK;
    auto fail = [&](cudaError_t e, const char *what) { if (e != cudaSuccess && rc == TB_OK) { tb::set_error(std::string(what) + ": " + cudaGetErrorString(e)); rc = TB_ERR_CUDA; } };
    fail(cudaMalloc(&dq, 8 * nq), "cudaMalloc q0");
    fail(cudaMalloc(&dp, 8 * npar), "cudaMalloc params");
    fail(cudaMalloc(&dwork, wbytes), "cudaMalloc workspace");
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
        fail(cudaMemset(dwork, 0, wbytes), "memset workspace");
        a.q0 = dq; a.params = dp; a.trange = trange ? dtr : nullptr; a.work = dwork; a.xfinal = dxf; a.tfinal = dtf;
        a.nsteps = dns; a.status = dst; a.xv_trange = dxv; a.counter = dctr;
    }
    if (rc == TB_OK) {
        cudaEvent_t e0, e1;
        cudaEventCreate(&e0);
        cudaEventCreate(&e1);
        cudaEventRecord(e0, 0);
        if (fast && minb == 8) wide_kernel<true, 8><<<(unsigned)blocks, threads>>>(a);
        else if (minb == 8) wide_kernel<false, 8><<<(unsigned)blocks, threads>>>(a);
        else if (fast && minb == 4) wide_kernel<true, 4><<<(unsigned)blocks, threads>>>(a);
        else if (fast) wide_kernel<true, 2><<<(unsigned)blocks, threads>>>(a);
        else if (minb == 4) wide_kernel<false, 4><<<(unsigned)blocks, threads>>>(a);
        else wide_kernel<false, 2><<<(unsigned)blocks, threads>>>(a);
        cudaEventRecord(e1, 0);
        tb::count_launch(1);
        fail(cudaGetLastError(), "wide_kernel launch");
        fail(cudaDeviceSynchronize(), "wide_kernel");
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
    cudaFree(dq); cudaFree(dp); cudaFree(dtr); cudaFree(dwork); cudaFree(dxf); cudaFree(dtf); cudaFree(dxv);
    cudaFree(dns); cudaFree(dst); cudaFree(dctr);
    return rc;
}

} /* extern "C" */
