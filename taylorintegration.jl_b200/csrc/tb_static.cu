/*
 * tb_static.cu -- the tape-independent kernels of libtaylorb200.so, compiled
 * ahead of time by nvcc for sm_100a:
 *   - tb_fp64_peak   : register-only DFMA chains; the denominator of the
 *                      "% of FP64 peak" metric
 *   - tb_dense_eval  : batched `(sol::TaylorSolution)(t)`
 *                      src/integrator/taylorsolution.jl:337-423 over the psol
 *                      layout written by the ensemble kernel
 */
#include <cuda_runtime.h>

#include <algorithm>
#include <cstdint>
#include <string>

#include "tb_internal.h"

namespace tb {
void count_launch(long long n);
}

namespace {

#define TB_CUDA(call)                                                              \
    do {                                                                           \
        cudaError_t e_ = (call);                                                   \
        if (e_ != cudaSuccess) {                                                   \
            tb::set_error(std::string(#call) + ": " + cudaGetErrorString(e_));     \
            return TB_ERR_CUDA;                                                    \
        }                                                                          \
    } while (0)

/* 16 independent DFMA chains per thread, all operands in registers */
__global__ void __launch_bounds__(256) fp64_peak_kernel(double *out, int iters, double a, double b) {
    double x[16];
#pragma unroll
    for (int j = 0; j < 16; j++) x[j] = (double)(threadIdx.x + j) * 1e-3;
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int r = 0; r < 8; r++) {
#pragma unroll
            for (int j = 0; j < 16; j++) x[j] = fma(x[j], a, b);
        }
    }
    double s = 0.0;
#pragma unroll
    for (int j = 0; j < 16; j++) s += x[j];
    if (s == 12345.678) out[blockIdx.x * blockDim.x + threadIdx.x] = s; /* keeps the chains alive */
}

/* One thread per (query, trajectory).  timeindex() of the reference
 * (taylorsolution.jl:337-360): the step whose interval [tv[s], tv[s+1])
 * contains t (the last step also owns its right end); evaluate that step's
 * polynomial at t - tv[s] (Horner).  Out of range -> NaN.                      */
__global__ void dense_eval_kernel(long long N, int dof, int order, long long stride, const double *__restrict__ tv,
                                  const double *__restrict__ psol, const long long *__restrict__ nsteps,
                                  const double *__restrict__ tq, long long nq, double *__restrict__ xq) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const long long q = blockIdx.y;
    if (i >= N || q >= nq) return;
    const long long ns = nsteps[i];
    const double t = tq[q];
    const double nan = __longlong_as_double(0x7FF8000000000000LL);
    const double tb = tv[i], te = tv[ns * N + i];
    const double sgn = (te >= tb) ? 1.0 : -1.0;
    bool inside = ns > 0 && sgn * t >= sgn * tb && sgn * t <= sgn * te;
    long long s = 0;
    if (inside) { /* binary search for the last s with sgn*tv[s] <= sgn*t, capped at ns-1 */
        long long lo = 0, hi = ns - 1;
        while (lo < hi) {
            long long mid = (lo + hi + 1) >> 1;
            if (sgn * tv[mid * N + i] <= sgn * t) lo = mid; else hi = mid - 1;
        }
        s = lo;
    }
    const int np1 = order + 1;
    for (int c = 0; c < dof; c++) {
        double v = nan;
        if (inside) {
            const double dt = __dsub_rn(t, tv[s * N + i]);
            const double *p = psol + ((long long)c * np1 * stride + s) * N + i;
            /* evaluate (Horner) with separate multiply and add like the reference and the parity kernels */
            v = p[(long long)order * stride * N];
            for (int k = order - 1; k >= 0; k--) v = __dadd_rn(__dmul_rn(v, dt), p[(long long)k * stride * N]);
        }
        xq[((long long)c * nq + q) * N + i] = v;
    }
}

/* TaylorSolutionSerialization layout (src/integrator/taylorsolution.jl:513-557): per trajectory the flat vector
 * p[((c*ns + s)*(order+1)) + k], i.e. component-major, then step, coefficients contiguous.  The ensemble kernel writes
 * psol[((c*(order+1) + k)*stride + s)*N + i] (trajectory fastest).  One warp moves the (order+1) coefficients of 32
 * trajectories of one (c, s) through shared memory: coalesced reads along i, contiguous writes along k.        */
__global__ void __launch_bounds__(256) serialize_kernel(long long N, int dof, int order, long long stride, const double *__restrict__ psol,
                                                        const long long *__restrict__ nsteps, const long long *__restrict__ offs,
                                                        double *__restrict__ out, long long smax) {
    extern __shared__ double tile[]; /* [warps][np1][33] */
    const int np1 = order + 1, lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = blockDim.x >> 5;
    double *T = tile + (size_t)w * np1 * 33;
    const long long i0 = (long long)blockIdx.x * 32;
    const long long i = i0 + lane;
    const long long ns_i = i < N ? nsteps[i] : 0;
    for (long long cs = (long long)blockIdx.y * nw + w; cs < (long long)dof * smax; cs += (long long)gridDim.y * nw) {
        const int c = (int)(cs / smax);
        const long long sidx = cs % smax;
        /* read: lane = trajectory */
        for (int k = 0; k < np1; k++)
            T[k * 33 + lane] = (i < N && sidx < ns_i) ? psol[(((long long)c * np1 + k) * stride + sidx) * N + i] : 0.0;
        __syncwarp();
        /* write: lane = coefficient */
        for (int j = 0; j < 32; j++) {
            const long long tj = i0 + j;
            if (tj >= N) break;
            const long long nsj = nsteps[tj];
            if (sidx >= nsj) continue;
            double *dst = out + offs[tj] + ((long long)c * nsj + sidx) * np1;
            for (int k = lane; k < np1; k += 32) dst[k] = T[k * 33 + j];
        }
        __syncwarp();
    }
}

} // namespace

/* sol(t) on DEVICE buffers: everything stays resident (psol as the ensemble kernel wrote it) */
extern "C" int tb_dense_eval_device(int64_t N, int32_t dof, int32_t order, int64_t stride, const double *d_tv, const double *d_psol,
                                    const int64_t *d_nsteps, const double *d_tq, int64_t nq, double *d_xq, void *stream) {
    if (N <= 0 || nq <= 0) return TB_OK;
    if (!d_tv || !d_psol || !d_nsteps || !d_tq || !d_xq || nq > 65535) { tb::set_error("tb_dense_eval_device: bad argument"); return TB_ERR_ARG; }
    dim3 grid((unsigned)((N + 127) / 128), (unsigned)nq);
    dense_eval_kernel<<<grid, 128, 0, (cudaStream_t)stream>>>(N, dof, order, stride, d_tv, d_psol, (const long long *)d_nsteps, d_tq, nq, d_xq);
    tb::count_launch(1);
    TB_CUDA(cudaGetLastError());
    return TB_OK;
}

/* psol (device, ensemble layout) -> the reference's serialization layout (device), offsets[i] = first double of trajectory i */
extern "C" int tb_serialize_device(int64_t N, int32_t dof, int32_t order, int64_t stride, int64_t max_nsteps, const double *d_psol,
                                   const int64_t *d_nsteps, const int64_t *d_offsets, double *d_out, void *stream) {
    if (N <= 0 || max_nsteps <= 0) return TB_OK;
    if (!d_psol || !d_nsteps || !d_offsets || !d_out || dof < 1 || order < 0) { tb::set_error("tb_serialize_device: bad argument"); return TB_ERR_ARG; }
    const int threads = 256, nw = threads / 32, np1 = order + 1;
    const size_t smem = (size_t)nw * np1 * 33 * sizeof(double);
    if (smem > 200 * 1024) { tb::set_error("tb_serialize_device: order too large"); return TB_ERR_UNSUPPORTED; }
    TB_CUDA(cudaFuncSetAttribute(serialize_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const long long work = (long long)dof * max_nsteps;
    dim3 grid((unsigned)((N + 31) / 32), (unsigned)std::max<long long>(1, std::min<long long>(64, (work + nw - 1) / nw)));
    serialize_kernel<<<grid, threads, smem, (cudaStream_t)stream>>>(N, dof, order, stride, d_psol, (const long long *)d_nsteps,
                                                                     (const long long *)d_offsets, d_out, max_nsteps);
    tb::count_launch(1);
    TB_CUDA(cudaGetLastError());
    return TB_OK;
}

extern "C" int tb_fp64_peak(double seconds, double *tflops, double *ms_out) {
    int dev = 0, n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess || n == 0) { tb::set_error("tb_fp64_peak: no usable CUDA device"); return TB_ERR_CUDA; }
    TB_CUDA(cudaGetDevice(&dev));
    cudaDeviceProp prop;
    TB_CUDA(cudaGetDeviceProperties(&prop, dev));
    const int threads = 256, blocks = prop.multiProcessorCount * 8;
    double *d_out = nullptr;
    TB_CUDA(cudaMalloc(&d_out, (size_t)threads * blocks * sizeof(double)));
    cudaEvent_t e0, e1;
    TB_CUDA(cudaEventCreate(&e0));
    TB_CUDA(cudaEventCreate(&e1));
    int iters = 2000;
    float ms = 0.f;
    double best = 0.0, total_ms = 0.0;
    /* warm-up, then repeat until `seconds` of device time has been spent */
    fp64_peak_kernel<<<blocks, threads>>>(d_out, iters, 1.0000001, 1e-9);
    tb::count_launch(1);
    TB_CUDA(cudaDeviceSynchronize());
    do {
        TB_CUDA(cudaEventRecord(e0));
        fp64_peak_kernel<<<blocks, threads>>>(d_out, iters, 1.0000001, 1e-9);
        tb::count_launch(1);
        TB_CUDA(cudaEventRecord(e1));
        TB_CUDA(cudaEventSynchronize(e1));
        TB_CUDA(cudaEventElapsedTime(&ms, e0, e1));
        const double flops = 2.0 * 16 * 8 * (double)iters * threads * blocks;
        const double tf = flops / (ms * 1e-3) / 1e12;
        if (tf > best) best = tf;
        total_ms += ms;
        if (ms < 50.f) iters *= 2;
    } while (total_ms < seconds * 1e3);
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    cudaFree(d_out);
    if (tflops) *tflops = best;
    if (ms_out) *ms_out = total_ms;
    return TB_OK;
}

extern "C" int tb_dense_eval(int64_t N, int32_t dof, int32_t order, int64_t stride, const double *tv, const double *psol,
                             const int64_t *nsteps, const double *tq, int64_t nq, double *xq) {
    if (N <= 0 || nq <= 0) return TB_OK;
    if (!tv || !psol || !nsteps || !tq || !xq || nq > 65535) { tb::set_error("tb_dense_eval: bad argument"); return TB_ERR_ARG; }
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess || n == 0) { tb::set_error("tb_dense_eval: no usable CUDA device"); return TB_ERR_CUDA; }
    const size_t np1 = order + 1;
    const size_t b_tv = (size_t)stride * N * 8, b_ps = (size_t)dof * np1 * stride * N * 8, b_ns = (size_t)N * 8, b_tq = (size_t)nq * 8,
                 b_xq = (size_t)dof * nq * N * 8;
    double *d_tv = nullptr, *d_ps = nullptr, *d_tq = nullptr, *d_xq = nullptr;
    long long *d_ns = nullptr;
    TB_CUDA(cudaMalloc(&d_tv, b_tv));
    TB_CUDA(cudaMalloc(&d_ps, b_ps));
    TB_CUDA(cudaMalloc(&d_ns, b_ns));
    TB_CUDA(cudaMalloc(&d_tq, b_tq));
    TB_CUDA(cudaMalloc(&d_xq, b_xq));
    TB_CUDA(cudaMemcpy(d_tv, tv, b_tv, cudaMemcpyHostToDevice));
    TB_CUDA(cudaMemcpy(d_ps, psol, b_ps, cudaMemcpyHostToDevice));
    TB_CUDA(cudaMemcpy(d_ns, nsteps, b_ns, cudaMemcpyHostToDevice));
    TB_CUDA(cudaMemcpy(d_tq, tq, b_tq, cudaMemcpyHostToDevice));
    dim3 grid((unsigned)((N + 127) / 128), (unsigned)nq);
    dense_eval_kernel<<<grid, 128>>>(N, dof, order, stride, d_tv, d_ps, d_ns, d_tq, nq, d_xq);
    tb::count_launch(1);
    TB_CUDA(cudaGetLastError());
    TB_CUDA(cudaMemcpy(xq, d_xq, b_xq, cudaMemcpyDeviceToHost));
    cudaFree(d_tv); cudaFree(d_ps); cudaFree(d_ns); cudaFree(d_tq); cudaFree(d_xq);
    return TB_OK;
}
