// tmem_bench.cu -- tensor memory as plain per-thread storage: latency / throughput of tcgen05.ld / tcgen05.st
// (SASS LDTM / STTM) with the 32x32b shape, alone and under an FP64 add chain, against LDS.64 on the same pattern.
// Build: nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -lineinfo -o tools/tmem_bench tools/tmem_bench.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e_), __LINE__); return 1; } } while (0)

__device__ __forceinline__ uint32_t tmem_alloc_all(uint32_t *slot) {
    if ((threadIdx.x >> 5) == 0) {
        uint32_t sa = (uint32_t)__cvta_generic_to_shared(slot);
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(sa) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    return *slot;
}
__device__ __forceinline__ void tmem_free_all(uint32_t base) {
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if ((threadIdx.x >> 5) == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(base) : "memory");
}
__device__ __forceinline__ void tm_st2(uint32_t addr, double v) {
    asm volatile("tcgen05.st.sync.aligned.32x32b.x2.b32 [%0], {%1, %2};" ::"r"(addr), "r"(__double2loint(v)), "r"(__double2hiint(v)) : "memory");
}
__device__ __forceinline__ double tm_ld2(uint32_t addr) {
    uint32_t lo, hi;
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x2.b32 {%0, %1}, [%2];" : "=r"(lo), "=r"(hi) : "r"(addr) : "memory");
    return __hiloint2double(hi, lo);
}
__device__ __forceinline__ void tm_wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void tm_wait_st() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void tm_wait_ld1(double &a) { asm volatile("tcgen05.wait::ld.sync.aligned;" : "+d"(a)::"memory"); }

// this warp's private window: lanes of quadrant (warp & 3), columns [cbase, cbase + ncol)
__device__ __forceinline__ uint32_t my_window(uint32_t base, int ncol_per_warp) {
    const int w = threadIdx.x >> 5;
    const int share = ((blockDim.x >> 5) + 3) / 4;          /* warps per lane quadrant */
    const int per = min(ncol_per_warp, 512 / share) & ~1;
    return base + ((uint32_t)((w & 3) * 32) << 16) + (uint32_t)((w >> 2) * per);
}

// mode 0: store then load back NR doubles per thread, check (correctness)
__global__ void k_check(int *bad) {
    __shared__ uint32_t slot;
    const uint32_t base = tmem_alloc_all(&slot);
    const int nw = blockDim.x >> 5;
    const int per = 512 / ((nw + 3) / 4);
    const uint32_t win = my_window(base, per);
    const int nd = per / 2;
    for (int j = 0; j < nd; j++) tm_st2(win + 2 * j, (double)(threadIdx.x * 1000 + j) + 0.25);
    tm_wait_st();
    int nb = 0;
    for (int j = 0; j < nd; j++) {
        double v = tm_ld2(win + 2 * j);
        tm_wait_ld1(v);
        if (v != (double)(threadIdx.x * 1000 + j) + 0.25) nb++;
    }
    if (nb) atomicAdd(bad, nb);
    tmem_free_all(base);
}

// dependent latency: ld -> wait -> address depends on the value (always +0)
__global__ void k_ld_latency(long long *cycles, int iters) {
    __shared__ uint32_t slot;
    const uint32_t base = tmem_alloc_all(&slot);
    const uint32_t win = my_window(base, 128);
    for (int j = 0; j < 64; j++) tm_st2(win + 2 * j, 0.0);
    tm_wait_st();
    uint32_t off = 0;
    long long t0 = clock64();
    for (int i = 0; i < iters; i++) {
        double v = tm_ld2(win + off);
        tm_wait_ld1(v);
        off = (uint32_t)__double2loint(v) & 62u;
    }
    long long t1 = clock64();
    if (threadIdx.x == 0 && blockIdx.x == 0) cycles[0] = (t1 - t0);
    if (off == 12345) cycles[1] = off;
    tmem_free_all(base);
}

// throughput: each thread reads U doubles (x2 each, static offsets) per round, one wait per round
template <int U>
__global__ void k_ld_tput(long long *cycles, double *sink, int iters) {
    __shared__ uint32_t slot;
    const uint32_t base = tmem_alloc_all(&slot);
    const uint32_t win = my_window(base, 256);
    for (int j = 0; j < U; j++) tm_st2(win + 2 * j, 1.0 + j);
    tm_wait_st();
    double acc = 0.0;
    __syncthreads();
    long long t0 = clock64();
    for (int i = 0; i < iters; i++) {
        double v[U];
#pragma unroll
        for (int u = 0; u < U; u++) v[u] = tm_ld2(win + 2 * u);
        tm_wait_ld();
#pragma unroll
        for (int u = 0; u < U; u++) acc += v[u];
    }
    long long t1 = clock64();
    __syncthreads();
    if (threadIdx.x == 0 && blockIdx.x == 0) cycles[0] = (t1 - t0);
    sink[blockIdx.x * blockDim.x + threadIdx.x] = acc;
    tmem_free_all(base);
}

// wide loads: one tcgen05.ld.32x32b.x32 = 16 doubles per thread
__global__ void k_ld_wide(long long *cycles, double *sink, int iters) {
    __shared__ uint32_t slot;
    const uint32_t base = tmem_alloc_all(&slot);
    const uint32_t win = my_window(base, 256);
    for (int j = 0; j < 64; j++) tm_st2(win + 2 * j, 1.0 + j);
    tm_wait_st();
    double acc = 0.0;
    __syncthreads();
    long long t0 = clock64();
    for (int i = 0; i < iters; i++) {
        uint32_t r[32];
        asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
                     : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]),
                       "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]), "=r"(r[17]), "=r"(r[18]),
                       "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]),
                       "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
                     : "r"(win + (uint32_t)((i & 3) * 32))
                     : "memory");
        tm_wait_ld();
#pragma unroll
        for (int u = 0; u < 16; u++) acc += __hiloint2double(r[2 * u + 1], r[2 * u]);
    }
    long long t1 = clock64();
    __syncthreads();
    if (threadIdx.x == 0 && blockIdx.x == 0) cycles[0] = (t1 - t0);
    sink[blockIdx.x * blockDim.x + threadIdx.x] = acc;
    tmem_free_all(base);
}

template <int U>
__global__ void k_st_tput(long long *cycles, int iters) {
    __shared__ uint32_t slot;
    const uint32_t base = tmem_alloc_all(&slot);
    const uint32_t win = my_window(base, 256);
    __syncthreads();
    long long t0 = clock64();
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int u = 0; u < U; u++) tm_st2(win + 2 * u, (double)(i + u));
    }
    tm_wait_st();
    long long t1 = clock64();
    __syncthreads();
    if (threadIdx.x == 0 && blockIdx.x == 0) cycles[0] = (t1 - t0);
    tmem_free_all(base);
}

// LDS.64 on the [k][thread] layout of the product kernel, same shape as k_ld_tput
template <int U>
__global__ void k_lds_tput(long long *cycles, double *sink, int iters) {
    extern __shared__ double sm[];
    volatile double *smt = sm + threadIdx.x;
    for (int j = 0; j < U; j++) smt[j * blockDim.x] = 1.0 + j;
    double acc = 0.0;
    __syncthreads();
    long long t0 = clock64();
    for (int i = 0; i < iters; i++) {
        double v[U];
#pragma unroll
        for (int u = 0; u < U; u++) v[u] = smt[u * blockDim.x];
#pragma unroll
        for (int u = 0; u < U; u++) acc += v[u];
    }
    long long t1 = clock64();
    __syncthreads();
    if (threadIdx.x == 0 && blockIdx.x == 0) cycles[0] = (t1 - t0);
    sink[blockIdx.x * blockDim.x + threadIdx.x] = acc;
}

// the product kernel's inner pattern: acc = acc + a_reg * b_mem, NCH independent chains, operand b from
// SRC = 0 registers, 1 shared memory, 2 tensor memory (U loads in flight per wait)
template <int SRC, int NCH, int U>
__global__ void k_mix(long long *cycles, double *sink, int iters) {
    extern __shared__ double sm[];
    __shared__ uint32_t slot;
    const uint32_t base = tmem_alloc_all(&slot);
    const uint32_t win = my_window(base, 256);
    volatile double *smt = sm + threadIdx.x;
    for (int j = 0; j < 32; j++) { tm_st2(win + 2 * j, 1.0 + 1e-9 * j); smt[j * blockDim.x] = 1.0 + 1e-9 * j; }
    tm_wait_st();
    double acc[NCH], a[NCH];
#pragma unroll
    for (int c = 0; c < NCH; c++) { acc[c] = 0.0; a[c] = 1.0 + c * 1e-3; }
    __syncthreads();
    long long t0 = clock64();
    for (int i = 0; i < iters; i++) {
        double b[U];
#pragma unroll
        for (int u = 0; u < U; u++) {
            if (SRC == 2) b[u] = tm_ld2(win + 2 * u);
            else if (SRC == 1) b[u] = smt[u * blockDim.x];
            else b[u] = a[u % NCH] + 1.0;
        }
        if (SRC == 2) tm_wait_ld();
#pragma unroll
        for (int u = 0; u < U; u++)
#pragma unroll
            for (int c = 0; c < NCH; c++) acc[c] = __dadd_rn(acc[c], __dmul_rn(a[c], b[u]));
    }
    long long t1 = clock64();
    __syncthreads();
    if (threadIdx.x == 0 && blockIdx.x == 0) cycles[0] = (t1 - t0);
    double s = 0.0;
    for (int c = 0; c < NCH; c++) s += acc[c];
    sink[blockIdx.x * blockDim.x + threadIdx.x] = s;
    tmem_free_all(base);
}

// the intended code shape: batch i+1 is requested before batch i is consumed, one wait per batch
template <int NCH, int U>
__global__ void k_mix_db(long long *cycles, double *sink, int iters) {
    __shared__ uint32_t slot;
    const uint32_t base = tmem_alloc_all(&slot);
    const uint32_t win = my_window(base, 128);
    for (int j = 0; j < 64; j++) tm_st2(win + 2 * j, 1.0 + 1e-9 * j);
    tm_wait_st();
    double acc[NCH], a[NCH];
#pragma unroll
    for (int c = 0; c < NCH; c++) { acc[c] = 0.0; a[c] = 1.0 + c * 1e-3; }
    double b0[U], b1[U];
#pragma unroll
    for (int u = 0; u < U; u++) b0[u] = tm_ld2(win + 2 * u);
    tm_wait_ld();
    __syncthreads();
    long long t0 = clock64();
    for (int i = 0; i < iters; i += 2) {
#pragma unroll
        for (int u = 0; u < U; u++) b1[u] = tm_ld2(win + 2 * (U + u));
#pragma unroll
        for (int u = 0; u < U; u++)
#pragma unroll
            for (int c = 0; c < NCH; c++) acc[c] = __dadd_rn(acc[c], __dmul_rn(a[c], b0[u]));
        tm_wait_ld();
#pragma unroll
        for (int u = 0; u < U; u++) b0[u] = tm_ld2(win + 2 * u);
#pragma unroll
        for (int u = 0; u < U; u++)
#pragma unroll
            for (int c = 0; c < NCH; c++) acc[c] = __dadd_rn(acc[c], __dmul_rn(a[c], b1[u]));
        tm_wait_ld();
    }
    long long t1 = clock64();
    __syncthreads();
    if (threadIdx.x == 0 && blockIdx.x == 0) cycles[0] = (t1 - t0);
    double s = 0.0;
    for (int c = 0; c < NCH; c++) s += acc[c];
    sink[blockIdx.x * blockDim.x + threadIdx.x] = s;
    tmem_free_all(base);
}

int main() {
    long long *cyc;
    double *sink;
    int *bad;
    CK(cudaMallocManaged(&cyc, 64));
    CK(cudaMallocManaged(&bad, 4));
    CK(cudaMalloc(&sink, 148 * 1024 * 8));
    const int iters = 20000;
    *bad = 0;
    for (int bt : {128, 256, 384, 512}) {
        k_check<<<148, bt>>>(bad);
        CK(cudaDeviceSynchronize());
        printf("check block %d: %d mismatches\n", bt, *bad);
    }
    k_ld_latency<<<1, 32>>>(cyc, iters);
    CK(cudaDeviceSynchronize());
    printf("LDTM.x2 dependent round trip (ld, wait::ld, address from value): %.1f cycles\n", (double)cyc[0] / iters);
#define RUN_LD(U, BT) { k_ld_tput<U><<<148, BT>>>(cyc, sink, iters); CK(cudaDeviceSynchronize()); \
    printf("LDTM.x2 x %2d per wait, %3d thr/SM: %.2f cyc/round/warp, %.1f B/clk/SM\n", U, BT, (double)cyc[0] / iters, (double)U * 8 * BT * iters / cyc[0]); }
    RUN_LD(1, 256) RUN_LD(2, 256) RUN_LD(4, 256) RUN_LD(8, 256) RUN_LD(16, 256) RUN_LD(32, 256)
    RUN_LD(8, 128) RUN_LD(8, 384) RUN_LD(8, 512)
    k_ld_wide<<<148, 256>>>(cyc, sink, iters);
    CK(cudaDeviceSynchronize());
    printf("LDTM.x32 (16 doubles) per wait, 256 thr/SM: %.2f cyc/round/warp, %.1f B/clk/SM\n", (double)cyc[0] / iters, 16.0 * 8 * 256 * iters / cyc[0]);
#define RUN_ST(U, BT) { k_st_tput<U><<<148, BT>>>(cyc, iters); CK(cudaDeviceSynchronize()); \
    printf("STTM.x2 x %2d, %3d thr/SM: %.2f cyc/round/warp, %.1f B/clk/SM\n", U, BT, (double)cyc[0] / iters, (double)U * 8 * BT * iters / cyc[0]); }
    RUN_ST(8, 256) RUN_ST(32, 256)
#define RUN_LDS(U, BT) { k_lds_tput<U><<<148, BT, U * BT * 8>>>(cyc, sink, iters); CK(cudaDeviceSynchronize()); \
    printf("LDS.64  x %2d, %3d thr/SM: %.2f cyc/round/warp, %.1f B/clk/SM\n", U, BT, (double)cyc[0] / iters, (double)U * 8 * BT * iters / cyc[0]); }
    RUN_LDS(8, 256) RUN_LDS(32, 256)
#define RUN_MIX(SRC, NCH, U, BT) { k_mix<SRC, NCH, U><<<148, BT, 32 * BT * 8>>>(cyc, sink, iters); CK(cudaDeviceSynchronize()); \
    printf("mul+add chains=%d, operand from %s, %2d loads/round, %3d thr/SM: %.2f cyc/round, FP64 pipe %.1f %%\n", NCH, \
           SRC == 0 ? "regs" : (SRC == 1 ? "smem" : "tmem"), U, BT, (double)cyc[0] / iters, \
           100.0 * (2.0 * NCH * U * (BT / 32) / 4 * 2) * iters / cyc[0]); }
    RUN_MIX(0, 2, 8, 256) RUN_MIX(1, 2, 8, 256) RUN_MIX(2, 2, 8, 256)
    RUN_MIX(0, 1, 8, 256) RUN_MIX(1, 1, 8, 256) RUN_MIX(2, 1, 8, 256)
    RUN_MIX(1, 2, 2, 256) RUN_MIX(2, 2, 2, 256) RUN_MIX(2, 2, 4, 256) RUN_MIX(2, 2, 16, 256)
    RUN_MIX(1, 4, 8, 256) RUN_MIX(2, 4, 8, 256)
    RUN_MIX(2, 2, 8, 384) RUN_MIX(1, 2, 8, 384) RUN_MIX(0, 2, 8, 384)
#define RUN_DB(NCH, U, BT) { k_mix_db<NCH, U><<<148, BT>>>(cyc, sink, iters); CK(cudaDeviceSynchronize()); \
    printf("double-buffered tmem operand, chains=%d, batch %2d, %3d thr/SM: %.2f cyc/round, FP64 pipe %.1f %%\n", NCH, U, BT, \
           (double)cyc[0] / iters, 100.0 * (2.0 * NCH * U * (BT / 32) / 4 * 2) * iters / cyc[0]); }
    RUN_DB(1, 4, 256) RUN_DB(2, 4, 256) RUN_DB(2, 8, 256) RUN_DB(4, 4, 256) RUN_DB(1, 8, 256) RUN_DB(2, 2, 256)
    return 0;
}
