/*
 * tb_lib.cpp -- host side of libtaylorb200.so: tape validation, NVRTC
 * compilation for sm_100a, module loading and launches through the CUDA
 * driver API (resolved with dlopen so that the library also loads, generates
 * source and compiles cubins on a machine without a GPU), and the host-buffer
 * convenience entry points.  See include/taylorb200.h for the contract.
 *
 * There is deliberately NO CPU execution path in this library: without a
 * usable device every tb_integrate* / tb_jet call fails with TB_ERR_CUDA.
 */
#include "tb_internal.h"

#include <cuda.h>
#include <cuda_runtime.h>
#include <dlfcn.h>
#include <nvrtc.h>
#include <dlfcn.h>
#include <sys/stat.h>
#include <unistd.h>

#include <atomic>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <limits>
#include <mutex>
#include <thread>

namespace tb {

static thread_local std::string g_err;
void set_error(const std::string &m) { g_err = m; }
static std::atomic<long long> g_launches{0};
void count_launch(long long n) { g_launches += n; }
static thread_local double g_last_ms = -1.0;
void set_last_kernel_ms(double ms) { g_last_ms = ms; }

/* ---- driver API, resolved lazily ------------------------------------------ */
struct Driver {
    void *h = nullptr;
    CUresult (*ModuleLoadData)(CUmodule *, const void *) = nullptr;
    CUresult (*ModuleUnload)(CUmodule) = nullptr;
    CUresult (*ModuleGetFunction)(CUfunction *, CUmodule, const char *) = nullptr;
    CUresult (*FuncSetAttribute)(CUfunction, CUfunction_attribute, int) = nullptr;
    CUresult (*FuncGetAttribute)(int *, CUfunction_attribute, CUfunction) = nullptr;
    CUresult (*LaunchKernel)(CUfunction, unsigned, unsigned, unsigned, unsigned, unsigned, unsigned, unsigned, CUstream,
                             void **, void **) = nullptr;
    CUresult (*GetErrorString)(CUresult, const char **) = nullptr;
    bool ok = false;
    std::string why;
};
static Driver &driver() {
    static Driver d;
    static std::once_flag once;
    std::call_once(once, [] {
        d.h = dlopen("libcuda.so.1", RTLD_NOW | RTLD_GLOBAL);
        if (!d.h) { d.why = "libcuda.so.1 not found: no NVIDIA driver on this machine"; return; }
#define TB_SYM(field, name)                                                   \
    *(void **)(&d.field) = dlsym(d.h, name);                                  \
    if (!d.field) { d.why = std::string("missing driver symbol ") + name; return; }
        TB_SYM(ModuleLoadData, "cuModuleLoadData")
        TB_SYM(ModuleUnload, "cuModuleUnload")
        TB_SYM(ModuleGetFunction, "cuModuleGetFunction")
        TB_SYM(FuncSetAttribute, "cuFuncSetAttribute")
        TB_SYM(FuncGetAttribute, "cuFuncGetAttribute")
        TB_SYM(LaunchKernel, "cuLaunchKernel")
        TB_SYM(GetErrorString, "cuGetErrorString")
#undef TB_SYM
        d.ok = true;
    });
    return d;
}
static std::string cu_err(CUresult r) {
    const char *s = nullptr;
    if (driver().ok) driver().GetErrorString(r, &s);
    return s ? s : ("CUresult " + std::to_string((int)r));
}
#define TB_CUDA(call)                                                                                   \
    do {                                                                                                \
        cudaError_t e_ = (call);                                                                        \
        if (e_ != cudaSuccess) {                                                                        \
            set_error(std::string(#call) + ": " + cudaGetErrorString(e_));                              \
            return TB_ERR_CUDA;                                                                         \
        }                                                                                               \
    } while (0)

/* mirrors struct TbArgs of tb_device_pre.cuh */
struct TbArgs {
    long long N;
    const double *q0;
    const double *params;
    const double *trange;
    long long nt;
    double t0, tmax;
    double abstol, reltol, minstep, maxstep, nrabstol;
    long long maxsteps;
    long long stride;
    int params_per_traj;
    int eventorder, newtoniter, event_cap;
    int lockstep;
    int queue_reverse;
    double *xfinal;
    double *tfinal;
    long long *nsteps;
    int *status;
    double *xv_trange;
    double *tv;
    double *xv;
    double *psol;
    double *tevents;
    double *xevents;
    double *gresids;
    int *nevents;
    double *lambda;
    double *lambda_v;
    unsigned long long *counter;
    double *xjet;
    double *dtout;
    double *lyap_state;
    const double *lyap_state0;
};
static_assert(sizeof(TbArgs) == 288, "TbArgs layout must match tb_device_pre.cuh");

struct Loaded {
    CUmodule mod = nullptr;
    CUfunction ensemble = nullptr, jet = nullptr;
    int sm_count = 0;
};

} // namespace tb

struct tb_ode {
    tb::Ode o;
};

struct tb_kernel {
    tb::Ode ode;
    tb_compile_opts_t co;
    tb::Layout L;
    std::string source, log;
    std::vector<char> cubin;
    int regs = -1, local_bytes = -1;
    std::mutex mu;
    std::map<int, std::mutex> call_mu; /* per device: serialises the host-buffer entry points (they share that device's arena) */
    std::map<int, tb::Loaded> loaded; /* per device */
    /* grow-only device arena for the host-buffer entry points, per device */
    std::map<int, std::pair<char *, size_t>> arena;
};

namespace tb {

static int arity_ok(const tb_op_t &op, std::string &why) {
    auto kind = [](int32_t o) { return TB_OPERAND_KIND(o); };
    auto series = [&](int32_t o) { int k = kind(o); return k == TB_K_X || k == TB_K_DX || k == TB_K_TMP || k == TB_K_TIME; };
    auto number = [&](int32_t o) { int k = kind(o); return k == TB_K_PAR || k == TB_K_IMM; };
    if (kind(op.res) != TB_K_DX && kind(op.res) != TB_K_TMP) { why = "result must be a dx or tmp series"; return 0; }
    switch (op.op) {
        case TB_OP_IDENTITY: if (!series(op.a) && !number(op.a)) { why = "identity! needs an argument"; return 0; } break;
        case TB_OP_ZERO: case TB_OP_ONE: break;
        case TB_OP_NEG: case TB_OP_SQR: case TB_OP_SQRT: case TB_OP_EXP: case TB_OP_LOG: case TB_OP_POW:
            /* pow!/sqr!/... need a series base: a plain number makes the generated
             * method throw in the reference too (test/taylorize.jl:953-992)       */
            if (!series(op.a)) { why = "unary Taylor primitive needs a series argument"; return 0; }
            break;
        case TB_OP_SINCOS: case TB_OP_TAN: case TB_OP_SINHCOSH: case TB_OP_TANH: case TB_OP_ATAN: case TB_OP_ASIN: case TB_OP_ACOS:
            if (!series(op.a)) { why = "primitive with an auxiliary series needs a series argument"; return 0; }
            if (kind(op.aux) != TB_K_TMP && kind(op.aux) != TB_K_DX) { why = "sincos!/tan!/tanh!/atan!/asin!/acos!/sinhcosh! need an auxiliary series"; return 0; }
            break;
        case TB_OP_ADD: case TB_OP_SUB: case TB_OP_MUL: case TB_OP_DIV:
            if (!(series(op.a) || number(op.a)) || !(series(op.b) || number(op.b))) { why = "binary op needs two arguments"; return 0; }
            break;
        default: why = "opcode " + std::to_string(op.op) + " is not implemented"; return 0;
    }
    return 1;
}

static int validate(const tb_ode_desc_t &d, std::string &why) {
    if (d.dof < 1 || d.dof > 4096) { why = "dof out of range"; return 0; }
    if (d.npar < 0 || d.ntmp < 0 || d.nops < 1 || d.ev_nops < 0 || d.jac_nops < 0) { why = "negative count"; return 0; }
    if (!d.ops) { why = "ops is NULL"; return 0; }
    const int total = d.nops + d.ev_nops + d.jac_nops;
    for (int j = 0; j < total; j++) {
        const tb_op_t &op = d.ops[j];
        std::string w;
        if (!arity_ok(op, w)) { why = "op " + std::to_string(j) + ": " + w; return 0; }
        const int32_t all[4] = {op.res, op.a, op.b, op.aux};
        for (int32_t o : all) {
            int k = TB_OPERAND_KIND(o), i = TB_OPERAND_INDEX(o);
            bool ok = true;
            switch (k) {
                case TB_K_NONE: case TB_K_IMM: case TB_K_TIME: break;
                case TB_K_X: case TB_K_DX: ok = i < d.dof; break;
                case TB_K_TMP: ok = i < d.ntmp; break;
                case TB_K_PAR: ok = i < d.npar; break;
                default: ok = false;
            }
            if (!ok) { why = "op " + std::to_string(j) + ": operand out of range"; return 0; }
        }
    }
    if (d.has_jac && !d.jac) { why = "has_jac set but jac is NULL"; return 0; }
    return 1;
}

/* parse "Used N registers" / "N bytes stack frame" of the ptxas -v log for the ensemble kernel.  The stack-frame
 * line reads "    360 bytes stack frame, 192 bytes spill stores, ..." (the number precedes the words). */
static void parse_ptxas(const std::string &log, int *regs, int *local_bytes) {
    size_t p = log.find("tb_ensemble_kernel");
    if (p == std::string::npos) return;
    size_t q = log.find("Used ", p);
    if (q != std::string::npos) *regs = atoi(log.c_str() + q + 5);
    size_t s = log.find(" bytes stack frame", p);
    if (s != std::string::npos && (q == std::string::npos || s < q + 400)) {
        size_t b = s;
        while (b > 0 && log[b - 1] >= '0' && log[b - 1] <= '9') b--;
        *local_bytes = atoi(log.c_str() + b);
    }
}

static int get_device(int *dev) {
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n == 0) {
        set_error(std::string("no usable CUDA device: ") + (e != cudaSuccess ? cudaGetErrorString(e) : "device count is 0"));
        return TB_ERR_CUDA;
    }
    TB_CUDA(cudaGetDevice(dev));
    return TB_OK;
}

static int ensure_loaded(tb_kernel *k, Loaded **out) {
    int dev = 0;
    int rc = get_device(&dev);
    if (rc) return rc;
    std::lock_guard<std::mutex> lk(k->mu);
    auto it = k->loaded.find(dev);
    if (it != k->loaded.end()) { *out = &it->second; return TB_OK; }
    Driver &d = driver();
    if (!d.ok) { set_error(d.why); return TB_ERR_CUDA; }
    TB_CUDA(cudaFree(0)); /* make the primary context current */
    Loaded l;
    CUresult r = d.ModuleLoadData(&l.mod, k->cubin.data());
    if (r != CUDA_SUCCESS) { set_error("cuModuleLoadData: " + cu_err(r)); return TB_ERR_CUDA; }
    r = d.ModuleGetFunction(&l.ensemble, l.mod, "tb_ensemble_kernel");
    if (r != CUDA_SUCCESS) { set_error("cuModuleGetFunction(tb_ensemble_kernel): " + cu_err(r)); return TB_ERR_CUDA; }
    r = d.ModuleGetFunction(&l.jet, l.mod, "tb_jet_kernel");
    if (r != CUDA_SUCCESS) { set_error("cuModuleGetFunction(tb_jet_kernel): " + cu_err(r)); return TB_ERR_CUDA; }
    for (CUfunction f : {l.ensemble, l.jet}) {
        r = d.FuncSetAttribute(f, CU_FUNC_ATTRIBUTE_MAX_DYNAMIC_SHARED_SIZE_BYTES, (int)k->L.smem_bytes);
        if (r != CUDA_SUCCESS) { set_error("cuFuncSetAttribute(max dynamic smem): " + cu_err(r)); return TB_ERR_CUDA; }
    }
    int v = 0;
    if (d.FuncGetAttribute(&v, CU_FUNC_ATTRIBUTE_NUM_REGS, l.ensemble) == CUDA_SUCCESS) k->regs = v;
    if (d.FuncGetAttribute(&v, CU_FUNC_ATTRIBUTE_LOCAL_SIZE_BYTES, l.ensemble) == CUDA_SUCCESS) k->local_bytes = v;
    cudaDeviceProp prop;
    TB_CUDA(cudaGetDeviceProperties(&prop, dev));
    l.sm_count = prop.multiProcessorCount;
    k->loaded[dev] = l;
    *out = &k->loaded[dev];
    return TB_OK;
}

static void fill_args(TbArgs &a, int64_t N, const double *q0, const double *params, double t0, double tmax,
                      const double *trange, int64_t nt, const tb_opts_t &o, const tb_out_t &out, void *work) {
    memset(&a, 0, sizeof a);
    a.N = N;
    a.q0 = q0;
    a.params = params;
    a.trange = trange;
    a.nt = nt;
    a.t0 = t0;
    a.tmax = tmax;
    a.abstol = o.abstol;
    a.reltol = o.reltol;
    a.minstep = o.minstepsize;
    a.maxstep = o.maxstepsize;
    a.nrabstol = o.nrabstol;
    a.maxsteps = o.maxsteps;
    a.stride = o.steps_stride ? o.steps_stride : o.maxsteps + 1;
    a.params_per_traj = o.params_per_traj;
    a.eventorder = o.eventorder;
    a.newtoniter = o.newtoniter;
    a.event_cap = o.event_capacity > 0 ? o.event_capacity : (int)(o.maxsteps + 1);
    a.lockstep = (out.tv || out.xv || out.psol || out.lambda_v) ? 1 : 0;
    a.queue_reverse = o.sort_by_cost == 1 ? 1 : 0;
    a.xfinal = out.xfinal;
    a.tfinal = out.tfinal;
    a.nsteps = (long long *)out.nsteps;
    a.status = out.status;
    a.xv_trange = out.xv_trange;
    a.tv = out.tv;
    a.xv = out.xv;
    a.psol = out.psol;
    a.tevents = out.tevents;
    a.xevents = out.xevents;
    a.gresids = out.gresids;
    a.nevents = out.nevents;
    a.lambda = out.lambda;
    a.lambda_v = out.lambda_v;
    a.lyap_state = out.lyap_state;
    a.lyap_state0 = out.lyap_state0;
    a.counter = (unsigned long long *)work;
}

static int launch(tb_kernel *k, Loaded *l, CUfunction f, TbArgs &a, bool persistent, cudaStream_t stream) {
    Driver &d = driver();
    const int bt = k->L.block_threads;
    long long per_block = bt; /* trajectories one block holds at a time */
    if (k->co.kind == TB_KIND_LYAPCOL) { const int nq = k->ode.desc.dof / 2; per_block = (long long)(bt / 32) * (32 / nq); }
    long long want = (a.N + per_block - 1) / per_block;
    long long cap = (long long)l->sm_count * k->L.blocks_per_sm;
    unsigned grid = (unsigned)std::max<long long>(1, persistent ? std::min(want, cap) : std::min<long long>(want, cap * 4));
    void *params[1] = {&a};
    CUresult r = d.LaunchKernel(f, grid, 1, 1, (unsigned)bt, 1, 1, (unsigned)k->L.smem_bytes, (CUstream)stream, params, nullptr);
    if (r != CUDA_SUCCESS) { set_error("cuLaunchKernel: " + cu_err(r)); return TB_ERR_CUDA; }
    count_launch(1);
    return TB_OK;
}

/* bump allocator over the kernel's per-device arena */
struct Arena {
    char *base = nullptr;
    size_t cap = 0, off = 0;
    template <class T> T *take(size_t n) {
        size_t bytes = (n * sizeof(T) + 255) & ~(size_t)255;
        T *p = (T *)(base + off);
        off += bytes;
        return off <= cap ? p : nullptr;
    }
};
static size_t rup(size_t b) { return (b + 255) & ~(size_t)255; }

static int arena_get(tb_kernel *k, size_t bytes, Arena &A) {
    int dev = 0;
    int rc = get_device(&dev);
    if (rc) return rc;
    std::lock_guard<std::mutex> lk(k->mu);
    auto &slot = k->arena[dev];
    if (slot.second < bytes) {
        if (slot.first) cudaFree(slot.first);
        slot = {nullptr, 0};
        char *p = nullptr;
        TB_CUDA(cudaMalloc(&p, bytes));
        slot = {p, bytes};
    }
    A.base = slot.first;
    A.cap = slot.second;
    A.off = 0;
    return TB_OK;
}

/* the mutex that serialises host-buffer calls on the CURRENT device for this kernel handle */
static std::mutex &device_call_mutex(tb_kernel *k) {
    int dev = 0;
    cudaGetDevice(&dev);
    std::lock_guard<std::mutex> lk(k->mu);
    return k->call_mu[dev];
}

/* Argument / build-flag checks shared by the host-buffer and the device-buffer entry points. */
static int check_call(const char *fn, tb_kernel *k, int64_t N, const void *q0, const void *params, const double *trange,
                      int64_t nt, const tb_opts_t *opts, const tb_out_t *out) {
    const std::string f = fn;
    if (!k || !opts || !out || N < 0) { set_error(f + ": bad argument"); return TB_ERR_ARG; }
    if (N > 0 && !q0) { set_error(f + ": q0 is NULL"); return TB_ERR_ARG; }
    if (trange && nt < 2) { set_error(f + ": trange needs at least two points"); return TB_ERR_ARG; }
    if (trange && k->co.kind == TB_KIND_LYAP) { set_error("trange is implemented for TB_KIND_PLAIN, TB_KIND_EVENTS and TB_KIND_LYAPCOL kernels"); return TB_ERR_UNSUPPORTED; }
    const tb_ode_desc_t &d = k->ode.desc;
    if (N > 0 && d.npar > 0 && !params) { set_error(f + ": params is NULL"); return TB_ERR_ARG; }
    if (opts->maxsteps < 0 || opts->event_capacity < 0 || opts->steps_stride < 0) { set_error(f + ": negative maxsteps / event_capacity / steps_stride"); return TB_ERR_ARG; }
    if (k->co.kind == TB_KIND_EVENTS && (opts->eventorder < 0 || opts->eventorder > k->co.order)) { set_error("order must be >= eventorder >= 0 (rootfinding.jl:321)"); return TB_ERR_ARG; }
    if ((k->co.flags & TB_CF_FINAL_ONLY) && (k->co.kind == TB_KIND_PLAIN || k->co.kind == TB_KIND_EVENTS) && (trange || out->tv || out->xv || out->psol || out->xv_trange)) {
        set_error("this kernel was built with TB_CF_FINAL_ONLY: no trange, per-step or dense output");
        return TB_ERR_ARG;
    }
    const bool lyapcol = k->co.kind == TB_KIND_LYAPCOL;
    if (lyapcol && (trange != nullptr) != ((k->co.flags & TB_CF_LYAP_TRANGE) != 0)) {
        set_error("TB_KIND_LYAPCOL: kernels built with TB_CF_LYAP_TRANGE serve trange calls, the others (t0, tmax) calls");
        return TB_ERR_ARG;
    }
    if (lyapcol && out->psol) { set_error("TB_KIND_LYAPCOL has no dense output"); return TB_ERR_UNSUPPORTED; }
    if (out->lyap_state || out->lyap_state0) {
        if (k->co.kind != TB_KIND_LYAP && !lyapcol) { set_error(f + ": lyap_state / lyap_state0 need a Lyapunov kernel"); return TB_ERR_ARG; }
        if (trange) { set_error(f + ": continue-from-state is implemented for the (t0, tmax) Lyapunov drivers"); return TB_ERR_UNSUPPORTED; }
    }
    const int64_t stride = opts->steps_stride ? opts->steps_stride : opts->maxsteps + 1;
    if (!trange && (out->tv || out->xv || out->psol || out->lambda_v) && stride < opts->maxsteps + 1) {
        set_error(f + ": steps_stride must be >= maxsteps + 1");
        return TB_ERR_ARG;
    }
    return TB_OK;
}

struct EventPair { /* RAII: destroyed on every exit path */
    cudaEvent_t e0 = nullptr, e1 = nullptr;
    ~EventPair() { if (e0) cudaEventDestroy(e0); if (e1) cudaEventDestroy(e1); }
};
struct OwnStream {
    cudaStream_t s = nullptr;
    ~OwnStream() { if (s) cudaStreamDestroy(s); }
};

static int integrate_host(tb_kernel *k, int64_t N, const double *q0, const double *params, double t0, double tmax,
                          const double *trange, int64_t nt, const tb_opts_t *opts, tb_out_t *out) {
    int rc = check_call("tb_integrate", k, N, q0, params, trange, nt, opts, out);
    if (rc) return rc;
    const tb_ode_desc_t &d = k->ode.desc;
    if (N == 0) return TB_OK;
    Loaded *l = nullptr;
    rc = ensure_loaded(k, &l);
    if (rc) return rc;
    const bool lyapcol = k->co.kind == TB_KIND_LYAPCOL;
    const size_t dof = lyapcol ? d.dof / 2 : d.dof, n = (size_t)N, np1 = k->co.order + 1;
    const size_t stride = opts->steps_stride ? opts->steps_stride : opts->maxsteps + 1;
    const size_t cap = opts->event_capacity > 0 ? opts->event_capacity : (size_t)(opts->maxsteps + 1);
    const size_t npar_elems = d.npar ? (opts->params_per_traj ? (size_t)d.npar * n : (size_t)d.npar) : 0;
    /* rows: the buffer is [.. x rows x N] and only the first `used` rows (known after the kernel: max nsteps + 1,
     * max nevents) are copied back; planes = leading extent (dof, dof*(order+1)), 1 = a plain copy of everything */
    enum { ROWS_ALL = 0, ROWS_STEPS = 1, ROWS_STEPS_M1 = 2, ROWS_EVENTS = 3 };
    struct Item { void **dptr; void *host; size_t bytes; int rows_kind; size_t planes, rows; };
    tb_out_t dv;
    memset(&dv, 0, sizeof dv);
    std::vector<Item> outs;
    auto want = [&](void *host, void **dptr, size_t bytes, int rk = ROWS_ALL, size_t planes = 1, size_t rows = 0) {
        if (host) outs.push_back({dptr, host, bytes, rk, planes, rows});
    };
    const bool events = k->co.kind == TB_KIND_EVENTS;
    std::vector<int64_t> ns_tmp;
    std::vector<int32_t> ne_tmp;
    want(out->xfinal, (void **)&dv.xfinal, dof * n * 8);
    want(out->tfinal, (void **)&dv.tfinal, n * 8);
    const bool per_step = !trange && (out->tv || out->xv || out->psol || out->lambda_v);
    int64_t *h_ns = out->nsteps;
    if (!h_ns && per_step) { ns_tmp.resize(n); h_ns = ns_tmp.data(); }
    want(h_ns, (void **)&dv.nsteps, n * 8);
    want(out->status, (void **)&dv.status, n * 4);
    if (trange) want(out->xv_trange, (void **)&dv.xv_trange, dof * (size_t)nt * n * 8);
    else {
        want(out->tv, (void **)&dv.tv, stride * n * 8, ROWS_STEPS, 1, stride);
        want(out->xv, (void **)&dv.xv, dof * stride * n * 8, ROWS_STEPS, dof, stride);
        want(out->psol, (void **)&dv.psol, dof * np1 * stride * n * 8, ROWS_STEPS_M1, dof * np1, stride);
    }
    int32_t *h_ne = out->nevents;
    if (events) {
        const bool any_ev = out->tevents || out->xevents || out->gresids;
        if (!h_ne && any_ev) { ne_tmp.resize(n); h_ne = ne_tmp.data(); }
        want(out->tevents, (void **)&dv.tevents, cap * n * 8, ROWS_EVENTS, 1, cap);
        want(out->xevents, (void **)&dv.xevents, dof * cap * n * 8, ROWS_EVENTS, dof, cap);
        want(out->gresids, (void **)&dv.gresids, cap * n * 8, ROWS_EVENTS, 1, cap);
        want(h_ne, (void **)&dv.nevents, n * 4);
    }
    if (k->co.kind == TB_KIND_LYAP || lyapcol) {
        want(out->lambda, (void **)&dv.lambda, dof * n * 8);
        if (!trange) want(out->lambda_v, (void **)&dv.lambda_v, dof * stride * n * 8, ROWS_STEPS, dof, stride);
        else want(out->lambda_v, (void **)&dv.lambda_v, dof * (size_t)nt * n * 8); /* exponents at trange, [dof x nt x N] */
    }
    const size_t lyap_rows = dof * dof + dof + 1; /* Q, running sums, t00 */
    if (k->co.kind == TB_KIND_LYAP || lyapcol) want(out->lyap_state, (void **)&dv.lyap_state, lyap_rows * n * 8);
    size_t total = rup(256) + rup(dof * n * 8) + rup(npar_elems * 8) + rup((size_t)(trange ? nt : 0) * 8);
    if (out->lyap_state0) total += rup(lyap_rows * n * 8);
    for (auto &it : outs) total += rup(it.bytes);
    /* one host-buffer call at a time per kernel handle: the arena and the work counter belong to the call */
    std::lock_guard<std::mutex> call_lock(device_call_mutex(k));
    Arena A;
    rc = arena_get(k, total, A);
    if (rc) return rc;
    OwnStream os;
    TB_CUDA(cudaStreamCreateWithFlags(&os.s, cudaStreamNonBlocking));
    cudaStream_t st = os.s;
    void *work = A.take<char>(256);
    double *d_q0 = A.take<double>(dof * n);
    double *d_par = npar_elems ? A.take<double>(npar_elems) : nullptr;
    double *d_tr = trange ? A.take<double>((size_t)nt) : nullptr;
    for (auto &it : outs) *it.dptr = A.take<char>(it.bytes);
    if (out->lyap_state0) {
        double *d_ls0 = A.take<double>(lyap_rows * n);
        TB_CUDA(cudaMemcpyAsync(d_ls0, out->lyap_state0, lyap_rows * n * 8, cudaMemcpyHostToDevice, st));
        dv.lyap_state0 = d_ls0;
    }
    TB_CUDA(cudaMemsetAsync(work, 0, 256, st));
    TB_CUDA(cudaMemcpyAsync(d_q0, q0, dof * n * 8, cudaMemcpyHostToDevice, st));
    if (d_par) TB_CUDA(cudaMemcpyAsync(d_par, params, npar_elems * 8, cudaMemcpyHostToDevice, st));
    if (d_tr) TB_CUDA(cudaMemcpyAsync(d_tr, trange, (size_t)nt * 8, cudaMemcpyHostToDevice, st));
    /* buffers the kernel fills only partially start from zero (a device memset runs at HBM speed; the expensive part,
     * the copy back over PCIe, is limited to the rows in use below) */
    for (auto &it : outs)
        if (it.rows_kind != ROWS_ALL) TB_CUDA(cudaMemsetAsync(*it.dptr, 0, it.bytes, st));
    TbArgs a;
    fill_args(a, N, d_q0, d_par, t0, tmax, d_tr, nt, *opts, dv, work);
    EventPair ev;
    TB_CUDA(cudaEventCreate(&ev.e0));
    TB_CUDA(cudaEventCreate(&ev.e1));
    TB_CUDA(cudaEventRecord(ev.e0, st));
    rc = launch(k, l, l->ensemble, a, true, st);
    if (rc) return rc;
    TB_CUDA(cudaEventRecord(ev.e1, st));
    bool partial = false;
    for (auto &it : outs) {
        if (it.rows_kind != ROWS_ALL) { partial = true; continue; }
        TB_CUDA(cudaMemcpyAsync(it.host, *it.dptr, it.bytes, cudaMemcpyDeviceToHost, st));
    }
    if (partial) {
        /* the counters are on the host now: copy rows [0, used) of every [planes x rows x N] buffer.  Rows past
         * `used` of the caller's buffers are left untouched (they are past every trajectory's nsteps / nevents). */
        TB_CUDA(cudaStreamSynchronize(st));
        int64_t max_ns = 0;
        int32_t max_ne = 0;
        if (h_ns && per_step) for (size_t i = 0; i < n; i++) max_ns = std::max(max_ns, h_ns[i]);
        if (h_ne) for (size_t i = 0; i < n; i++) max_ne = std::max(max_ne, h_ne[i]);
        for (auto &it : outs) {
            if (it.rows_kind == ROWS_ALL) continue;
            size_t used = it.rows_kind == ROWS_EVENTS ? (size_t)max_ne
                        : it.rows_kind == ROWS_STEPS ? (size_t)max_ns + 1 : (size_t)max_ns;
            used = std::min(used, it.rows);
            if (used == 0) continue;
            TB_CUDA(cudaMemcpy2DAsync(it.host, it.rows * n * 8, *it.dptr, it.rows * n * 8, used * n * 8, it.planes,
                                      cudaMemcpyDeviceToHost, st));
        }
    }
    TB_CUDA(cudaStreamSynchronize(st));
    float ms = -1.f;
    cudaEventElapsedTime(&ms, ev.e0, ev.e1);
    set_last_kernel_ms(ms);
    return TB_OK;
}

} // namespace tb

using namespace tb;

extern "C" {

int tb_abi_version(void) { return TB_ABI_VERSION; }
const char *tb_last_error(void) { return g_err.c_str(); }

tb_ode_t *tb_ode_create(const tb_ode_desc_t *desc) {
    if (!desc) { set_error("tb_ode_create: NULL description"); return nullptr; }
    std::string why;
    if (!validate(*desc, why)) { set_error("tb_ode_create: " + why); return nullptr; }
    tb_ode_t *o = new tb_ode_t;
    o->o.desc = *desc;
    const int total = desc->nops + desc->ev_nops + desc->jac_nops;
    o->o.ops.assign(desc->ops, desc->ops + total);
    if (desc->has_jac) o->o.jac.assign(desc->jac, desc->jac + (size_t)desc->dof * desc->dof);
    o->o.desc.ops = o->o.ops.data();
    o->o.desc.jac = desc->has_jac ? o->o.jac.data() : nullptr;
    return o;
}
void tb_ode_free(tb_ode_t *ode) { delete ode; }

void tb_opts_default(tb_opts_t *o) {
    if (!o) return;
    memset(o, 0, sizeof *o);
    o->abstol = 1e-20;
    o->reltol = 0.0;
    o->minstepsize = 0.0;
    o->maxstepsize = std::numeric_limits<double>::infinity();
    o->maxsteps = 500;
    o->eventorder = 0;
    o->newtoniter = 10;
    o->nrabstol = std::numeric_limits<double>::epsilon();
    o->event_capacity = 0;
    o->params_per_traj = 0;
    o->steps_stride = 0;
}
void tb_compile_opts_default(tb_compile_opts_t *o) {
    if (!o) return;
    memset(o, 0, sizeof *o);
    o->order = 28;
    o->mode = TB_MODE_PARITY;
    o->kind = TB_KIND_PLAIN;
}

static bool read_file(const std::string &path, std::vector<char> &out) {
    FILE *f = fopen(path.c_str(), "rb");
    if (!f) return false;
    fseek(f, 0, SEEK_END);
    long n = ftell(f);
    fseek(f, 0, SEEK_SET);
    out.resize(n > 0 ? (size_t)n : 0);
    const bool ok = n >= 0 && (n == 0 || fread(out.data(), 1, (size_t)n, f) == (size_t)n);
    fclose(f);
    return ok;
}
static void write_file(const std::string &path, const void *p, size_t n) {
    const std::string tmp = path + ".tmp" + std::to_string((long)getpid());
    FILE *f = fopen(tmp.c_str(), "wb");
    if (!f) return;
    const bool ok = fwrite(p, 1, n, f) == n;
    fclose(f);
    if (ok) rename(tmp.c_str(), path.c_str()); else remove(tmp.c_str());
}
static std::string cache_path(const std::string &source, int mode) {
    std::string dir;
    if (const char *e = getenv("TB_CACHE_DIR")) dir = e;
    else {
        Dl_info di;
        if (!dladdr((const void *)&tb_abi_version, &di) || !di.dli_fname) return std::string();
        dir = di.dli_fname;
        const size_t sl = dir.find_last_of('/');
        dir = (sl == std::string::npos ? std::string(".") : dir.substr(0, sl)) + "/kcache";
    }
    mkdir(dir.c_str(), 0755);
    unsigned long long h = 1469598103934665603ull;
    auto mix = [&](const char *p, size_t n) { for (size_t i = 0; i < n; i++) { h ^= (unsigned char)p[i]; h *= 1099511628211ull; } };
    mix(source.data(), source.size());
    /* the NVRTC version is NOT part of the key: a process that imported torch first resolves libnvrtc.so.12 to
     * torch's bundled 12.8, whose ptxas spills where 12.9 does not; cubins prebuilt by build() (CUDA 12.9) win */
    std::string tag = "|sm_100a|mode" + std::to_string(mode);
    if (const char *x = getenv("TB_PTXAS_EXTRA")) tag += std::string("|") + x; /* experiment hook, see tb_compile */
    mix(tag.data(), tag.size());
    char name[64];
    snprintf(name, sizeof name, "/%016llx", h);
    return dir + name;
}

/* NVRTC -> cubin for sm_100a with the in-tree disk cache; shared by the ensemble kernels (tb_compile) and the generated
 * cluster kernels of tb_wide2.cu */
} // extern "C"
namespace tb {
int compile_cubin(const std::string &source, int mode, bool disk_cache, std::vector<char> &cubin, std::string &log) {
    /* on-disk cubin cache (in-tree, next to the library, so build() on the CPU box pre-compiles for the GPU box):
     * key = FNV-1a of the generated source + compile mode.  TB_CF_NO_DISK_CACHE skips it. */
    const std::string cpath = disk_cache ? cache_path(source, mode) : std::string();
    if (!cpath.empty() && read_file(cpath + ".cubin", cubin) && !cubin.empty()) {
        std::vector<char> lg;
        if (read_file(cpath + ".log", lg)) log.assign(lg.begin(), lg.end());
        /* a cubin written by an OLDER NVRTC than the one running is stale (toolkit upgrade): recompile.  One written
         * by a newer NVRTC is kept: that is build()'s CUDA 12.9 cubin met by a process whose libnvrtc.so.12 resolved
         * to torch's bundled 12.8, whose ptxas spills where 12.9 does not.                                         */
        int cmaj = 0, cmin = 0, rmaj = 0, rmin = 0;
        const size_t vp = log.rfind("compiled by NVRTC ");
        const bool stamped = vp != std::string::npos && sscanf(log.c_str() + vp, "compiled by NVRTC %d.%d", &cmaj, &cmin) == 2;
        nvrtcVersion(&rmaj, &rmin);
        if (stamped && (rmaj < cmaj || (rmaj == cmaj && rmin <= cmin))) {
            return TB_OK;
        }
        cubin.clear();
        log.clear();
    }
    nvrtcProgram prog;
    if (nvrtcCreateProgram(&prog, source.c_str(), "tb_kernel.cu", 0, nullptr, nullptr) != NVRTC_SUCCESS) {
        set_error("nvrtcCreateProgram failed");
        return TB_ERR_COMPILE;
    }
    std::vector<const char *> o = {"--gpu-architecture=sm_100a", "-std=c++17", "-lineinfo", "--ptxas-options=-v",
                                   "-default-device"};
    o.push_back(mode == TB_MODE_PARITY ? "--fmad=false" : "--fmad=true");
    o.push_back("--prec-div=true");
    o.push_back("--prec-sqrt=true");
    /* experiment hook: extra ptxas options (e.g. TB_PTXAS_EXTRA=-O2), part of the cache key */
    std::string extra;
    if (const char *x = getenv("TB_PTXAS_EXTRA")) { extra = std::string("--ptxas-options=") + x; o.push_back(extra.c_str()); }
    nvrtcResult r = nvrtcCompileProgram(prog, (int)o.size(), o.data());
    size_t ls = 0;
    nvrtcGetProgramLogSize(prog, &ls);
    log.resize(ls);
    if (ls) nvrtcGetProgramLog(prog, &log[0]);
    if (r != NVRTC_SUCCESS) {
        set_error(std::string("NVRTC: ") + nvrtcGetErrorString(r) + "\n" + log);
        nvrtcDestroyProgram(&prog);
        return TB_ERR_COMPILE;
    }
    size_t cs = 0;
    nvrtcGetCUBINSize(prog, &cs);
    cubin.resize(cs);
    nvrtcGetCUBIN(prog, cubin.data());
    nvrtcDestroyProgram(&prog);
    if (!cpath.empty()) {
        write_file(cpath + ".cubin", cubin.data(), cubin.size());
        int maj = 0, min = 0;
        nvrtcVersion(&maj, &min);
        log += "\n[libtaylorb200] compiled by NVRTC " + std::to_string(maj) + "." + std::to_string(min) + "\n";
        write_file(cpath + ".log", log.data(), log.size());
    }
    return TB_OK;
}
int module_load(const std::vector<char> &cubin, const char *fname, size_t smem, void **mod, void **fn) {
    Driver &d = driver();
    if (!d.ok) { set_error(d.why); return TB_ERR_CUDA; }
    TB_CUDA(cudaFree(0)); /* make the primary context current */
    CUmodule m = nullptr;
    CUfunction f = nullptr;
    CUresult r = d.ModuleLoadData(&m, cubin.data());
    if (r != CUDA_SUCCESS) { set_error("cuModuleLoadData: " + cu_err(r)); return TB_ERR_CUDA; }
    r = d.ModuleGetFunction(&f, m, fname);
    if (r != CUDA_SUCCESS) { set_error(std::string("cuModuleGetFunction(") + fname + "): " + cu_err(r)); d.ModuleUnload(m); return TB_ERR_CUDA; }
    r = d.FuncSetAttribute(f, CU_FUNC_ATTRIBUTE_MAX_DYNAMIC_SHARED_SIZE_BYTES, (int)smem);
    if (r != CUDA_SUCCESS) { set_error("cuFuncSetAttribute(max dynamic smem): " + cu_err(r)); d.ModuleUnload(m); return TB_ERR_CUDA; }
    *mod = m;
    *fn = f;
    return TB_OK;
}
void module_unload(void *mod) {
    if (mod && driver().ok) driver().ModuleUnload((CUmodule)mod);
}
int launch_fn(void *fn, unsigned grid, unsigned block, size_t smem, void *stream, void *arg_struct) {
    void *params[1] = {arg_struct};
    CUresult r = driver().LaunchKernel((CUfunction)fn, grid, 1, 1, block, 1, 1, (unsigned)smem, (CUstream)stream, params, nullptr);
    if (r != CUDA_SUCCESS) { set_error("cuLaunchKernel: " + cu_err(r)); return TB_ERR_CUDA; }
    count_launch(1);
    return TB_OK;
}
} // namespace tb
using namespace tb;
extern "C" {

const char *tb_generate_source(const tb_ode_t *ode, const tb_compile_opts_t *co) {
    static thread_local std::string src;
    if (!ode || !co) { set_error("tb_generate_source: NULL argument"); return nullptr; }
    Layout L;
    std::string err;
    if (generate_any(ode->o, *co, L, src, err) != TB_OK) { set_error("tb_generate_source: " + err); return nullptr; }
    return src.c_str();
}

tb_kernel_t *tb_compile(const tb_ode_t *ode, const tb_compile_opts_t *co) {
    if (!ode || !co) { set_error("tb_compile: NULL argument"); return nullptr; }
    tb_kernel_t *k = new tb_kernel_t;
    k->ode = ode->o;
    k->ode.desc.ops = k->ode.ops.data();
    k->ode.desc.jac = k->ode.desc.has_jac ? k->ode.jac.data() : nullptr;
    k->co = *co;
    std::string err;
    if (generate_any(k->ode, k->co, k->L, k->source, err) != TB_OK) {
        set_error("tb_compile: " + err);
        delete k;
        return nullptr;
    }
    if (compile_cubin(k->source, k->co.mode, !(k->co.flags & TB_CF_NO_DISK_CACHE), k->cubin, k->log) != TB_OK) {
        delete k;
        return nullptr;
    }
    parse_ptxas(k->log, &k->regs, &k->local_bytes);
    return k;
}

void tb_kernel_free(tb_kernel_t *k) {
    if (!k) return;
    for (auto &kv : k->loaded)
        if (kv.second.mod && driver().ok) driver().ModuleUnload(kv.second.mod);
    for (auto &kv : k->arena)
        if (kv.second.first) cudaFree(kv.second.first);
    delete k;
}
const char *tb_kernel_source(const tb_kernel_t *k) { return k ? k->source.c_str() : nullptr; }
const char *tb_kernel_log(const tb_kernel_t *k) { return k ? k->log.c_str() : nullptr; }
int tb_kernel_info(const tb_kernel_t *k, int *regs, int *smem_bytes, int *local_bytes, int *block_threads, int *blocks_per_sm) {
    if (!k) return TB_ERR_ARG;
    if (regs) *regs = k->regs;
    if (smem_bytes) *smem_bytes = (int)k->L.smem_bytes;
    if (local_bytes) *local_bytes = k->local_bytes;
    if (block_threads) *block_threads = k->L.block_threads;
    if (blocks_per_sm) *blocks_per_sm = k->L.blocks_per_sm;
    return TB_OK;
}
const void *tb_kernel_cubin(const tb_kernel_t *k, size_t *nbytes) {
    if (!k) return nullptr;
    if (nbytes) *nbytes = k->cubin.size();
    return k->cubin.data();
}

int tb_integrate(tb_kernel_t *k, int64_t N, const double *q0, const double *params, double t0, double tmax,
                 const tb_opts_t *opts, tb_out_t *out) {
    return integrate_host(k, N, q0, params, t0, tmax, nullptr, 0, opts, out);
}
int tb_integrate_trange(tb_kernel_t *k, int64_t N, const double *q0, const double *params, const double *trange,
                        int64_t nt, const tb_opts_t *opts, tb_out_t *out) {
    if (!trange) { set_error("tb_integrate_trange: trange is NULL"); return TB_ERR_ARG; }
    return integrate_host(k, N, q0, params, trange[0], trange[nt > 0 ? nt - 1 : 0], trange, nt, opts, out);
}

/* ---- one ensemble over several GPUs of one process -------------------------------------------------------------
 * trajectories are independent, so the ensemble is cut into round-robin blocks of `block`
 * trajectories (cost-ordered ensembles: contiguous halves would hand one GPU all the expensive orbits), device d
 * takes blocks d, d + ndev, ...; one host thread per device packs its shard, runs the ordinary host-buffer path on
 * that device (own stream, own arena) and scatters the results straight into the caller's [rows x N] buffers.
 * No collective is needed inside one process.                                                                   */
} /* extern "C" */

struct ShardCopy {
    int64_t N, block;
    int dev_index, ndev;
    int64_t count() const {
        int64_t n = 0;
        for (int64_t b = dev_index; b * block < N; b += ndev) n += std::min(block, N - b * block);
        return n;
    }
    /* global [rows x N] -> local [rows x n] */
    template <class T> void pack(const T *g, T *l, size_t rows, int64_t n) const {
        for (size_t r = 0; r < rows; r++) {
            int64_t o = 0;
            for (int64_t b = dev_index; b * block < N; b += ndev) {
                const int64_t len = std::min(block, N - b * block);
                memcpy(l + r * (size_t)n + o, g + r * (size_t)N + b * block, (size_t)len * sizeof(T));
                o += len;
            }
        }
    }
    template <class T> void unpack(T *g, const T *l, size_t rows, int64_t n) const {
        for (size_t r = 0; r < rows; r++) {
            int64_t o = 0;
            for (int64_t b = dev_index; b * block < N; b += ndev) {
                const int64_t len = std::min(block, N - b * block);
                memcpy(g + r * (size_t)N + b * block, l + r * (size_t)n + o, (size_t)len * sizeof(T));
                o += len;
            }
        }
    }
};

extern "C" {

int tb_integrate_multi(tb_kernel_t *k, const int32_t *devices, int32_t ndev, int64_t block, int64_t N, const double *q0,
                       const double *params, double t0, double tmax, const double *trange, int64_t nt,
                       const tb_opts_t *opts, tb_out_t *out) {
    if (!devices || ndev < 1) { set_error("tb_integrate_multi: no devices"); return TB_ERR_ARG; }
    if (out && (out->lyap_state || out->lyap_state0)) { set_error("tb_integrate_multi: Lyapunov continue-from-state is a one-device call"); return TB_ERR_UNSUPPORTED; }
    if (block <= 0) block = std::max<int64_t>(1, std::min<int64_t>(4096, (N + 8 * (int64_t)ndev - 1) / (8 * (int64_t)ndev))); /* >= ~8 blocks per device */
    int rc = check_call("tb_integrate_multi", k, N, q0, params, trange, nt, opts, out);
    if (rc) return rc;
    if (trange) { t0 = trange[0]; tmax = trange[nt - 1]; }
    if (N == 0) return TB_OK;
    const tb_ode_desc_t &d = k->ode.desc;
    const bool lyapcol = k->co.kind == TB_KIND_LYAPCOL;
    const size_t dof = lyapcol ? d.dof / 2 : d.dof, np1 = k->co.order + 1;
    const size_t stride = opts->steps_stride ? opts->steps_stride : opts->maxsteps + 1;
    const size_t cap = opts->event_capacity > 0 ? opts->event_capacity : (size_t)(opts->maxsteps + 1);
    std::vector<int> rcs(ndev, TB_OK);
    std::vector<std::string> errs(ndev);
    std::vector<double> kms(ndev, 0.0);
    int prev = 0;
    cudaGetDevice(&prev);
    auto worker = [&](int di) {
        if (cudaSetDevice(devices[di]) != cudaSuccess) { rcs[di] = TB_ERR_CUDA; errs[di] = "cudaSetDevice failed"; return; }
        ShardCopy sc{N, block, di, ndev};
        const int64_t n = sc.count();
        if (n == 0) return;
        std::vector<double> q((size_t)dof * n), par;
        sc.pack(q0, q.data(), dof, n);
        const double *pp = params;
        if (d.npar && opts->params_per_traj) { par.resize((size_t)d.npar * n); sc.pack(params, par.data(), d.npar, n); pp = par.data(); }
        /* local output buffers for whatever the caller asked for */
        struct Buf { void *global; void **slot; size_t rows, esz; std::vector<char> mem; };
        tb_out_t lo;
        memset(&lo, 0, sizeof lo);
        std::vector<Buf> bufs;
        auto want = [&](void *g, void **slot, size_t rows, size_t esz) {
            if (!g) return;
            bufs.push_back({g, slot, rows, esz, std::vector<char>(rows * (size_t)n * esz)});
        };
        want(out->xfinal, (void **)&lo.xfinal, dof, 8);
        want(out->tfinal, (void **)&lo.tfinal, 1, 8);
        want(out->nsteps, (void **)&lo.nsteps, 1, 8);
        want(out->status, (void **)&lo.status, 1, 4);
        if (trange) want(out->xv_trange, (void **)&lo.xv_trange, dof * (size_t)nt, 8);
        else {
            want(out->tv, (void **)&lo.tv, stride, 8);
            want(out->xv, (void **)&lo.xv, dof * stride, 8);
            want(out->psol, (void **)&lo.psol, dof * np1 * stride, 8);
        }
        if (k->co.kind == TB_KIND_EVENTS) {
            want(out->tevents, (void **)&lo.tevents, cap, 8);
            want(out->xevents, (void **)&lo.xevents, dof * cap, 8);
            want(out->gresids, (void **)&lo.gresids, cap, 8);
            want(out->nevents, (void **)&lo.nevents, 1, 4);
        }
        if (k->co.kind == TB_KIND_LYAP || lyapcol) {
            want(out->lambda, (void **)&lo.lambda, dof, 8);
            want(out->lambda_v, (void **)&lo.lambda_v, trange ? dof * (size_t)nt : dof * stride, 8);
        }
        for (auto &b : bufs) *b.slot = b.mem.data();
        rcs[di] = integrate_host(k, n, q.data(), pp, t0, tmax, trange, nt, opts, &lo);
        if (rcs[di]) { errs[di] = tb_last_error(); return; }
        kms[di] = tb_last_kernel_ms();
        for (auto &b : bufs) {
            if (b.esz == 8) sc.unpack((double *)b.global, (const double *)b.mem.data(), b.rows, n);
            else sc.unpack((int32_t *)b.global, (const int32_t *)b.mem.data(), b.rows, n);
        }
    };
    std::vector<std::thread> th;
    for (int di = 1; di < ndev; di++) th.emplace_back(worker, di);
    worker(0);
    for (auto &t : th) t.join();
    cudaSetDevice(prev);
    double worst = 0.0;
    for (int di = 0; di < ndev; di++) {
        if (rcs[di]) { set_error("tb_integrate_multi: device " + std::to_string(devices[di]) + ": " + errs[di]); return rcs[di]; }
        worst = std::max(worst, kms[di]);
    }
    set_last_kernel_ms(worst); /* the slowest device's kernel */
    return TB_OK;
}

size_t tb_work_bytes(void) { return 256; }

int tb_integrate_device(tb_kernel_t *k, int64_t N, const double *d_q0, const double *d_params, double t0, double tmax,
                        const double *d_trange, int64_t nt, const tb_opts_t *opts, const tb_out_t *d_out, void *d_work,
                        void *stream) {
    if (!d_work) { set_error("tb_integrate_device: d_work is NULL"); return TB_ERR_ARG; }
    int rc0 = check_call("tb_integrate_device", k, N, d_q0, d_params, d_trange, nt, opts, d_out);
    if (rc0) return rc0;
    if (N == 0) return TB_OK;
    Loaded *l = nullptr;
    int rc = ensure_loaded(k, &l);
    if (rc) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    TB_CUDA(cudaMemsetAsync(d_work, 0, 8, st));
    TbArgs a;
    fill_args(a, N, d_q0, d_params, t0, tmax, d_trange, nt, *opts, *d_out, d_work);
    return launch(k, l, l->ensemble, a, true, st);
}

int64_t tb_launch_count(void) { return g_launches.load(); }
double tb_last_kernel_ms(void) { return g_last_ms; }

int tb_jet(tb_kernel_t *k, int64_t N, const double *q0, const double *params, double t0, const tb_opts_t *opts,
           double *xjet, double *dt) {
    if (!k || !opts || !q0 || N < 0) { set_error("tb_jet: bad argument"); return TB_ERR_ARG; }
    if (k->co.kind == TB_KIND_LYAPCOL) { set_error("tb_jet: not offered for TB_KIND_LYAPCOL kernels"); return TB_ERR_UNSUPPORTED; }
    if (N == 0) return TB_OK;
    Loaded *l = nullptr;
    int rc = ensure_loaded(k, &l);
    if (rc) return rc;
    const tb_ode_desc_t &d = k->ode.desc;
    const size_t dof = d.dof, n = (size_t)N, np1 = k->co.order + 1;
    const size_t npar_elems = d.npar ? (opts->params_per_traj ? (size_t)d.npar * n : (size_t)d.npar) : 0;
    if (d.npar > 0 && !params) { set_error("tb_jet: params is NULL"); return TB_ERR_ARG; }
    std::lock_guard<std::mutex> call_lock(device_call_mutex(k));
    Arena A;
    rc = arena_get(k, rup(dof * n * 8) + rup(npar_elems * 8) + rup(dof * np1 * n * 8) + rup(n * 8) + 256, A);
    if (rc) return rc;
    double *d_q0 = A.take<double>(dof * n);
    double *d_par = npar_elems ? A.take<double>(npar_elems) : nullptr;
    double *d_jet = A.take<double>(dof * np1 * n);
    double *d_dt = A.take<double>(n);
    TB_CUDA(cudaMemcpy(d_q0, q0, dof * n * 8, cudaMemcpyHostToDevice));
    if (d_par) TB_CUDA(cudaMemcpy(d_par, params, npar_elems * 8, cudaMemcpyHostToDevice));
    TbArgs a;
    tb_out_t none;
    memset(&none, 0, sizeof none);
    fill_args(a, N, d_q0, d_par, t0, t0, nullptr, 0, *opts, none, nullptr);
    a.xjet = d_jet;
    a.dtout = d_dt;
    rc = launch(k, l, l->jet, a, false, 0);
    if (rc) return rc;
    TB_CUDA(cudaDeviceSynchronize());
    if (xjet) TB_CUDA(cudaMemcpy(xjet, d_jet, dof * np1 * n * 8, cudaMemcpyDeviceToHost));
    if (dt) TB_CUDA(cudaMemcpy(dt, d_dt, n * 8, cudaMemcpyDeviceToHost));
    return TB_OK;
}

int tb_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) return 0;
    return n;
}
int tb_set_device(int device) {
    TB_CUDA(cudaSetDevice(device));
    return TB_OK;
}

} /* extern "C" */
