// FP64 pipe microbenchmark for B200 (sm_100a): dependent-issue latency of DADD/DMUL/DFMA and the
// throughput of C independent chains per warp with W warps per SM sub-partition.  Build:
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o tools/fp64_latency tools/fp64_latency.cu
// Output feeds DESIGN.md section 4 (how much ILP / occupancy the jet body needs to fill the pipe).
#include <cstdio>
#include <cuda_runtime.h>

template <int OP, int CH>
__global__ void chains(double *out, double seed, int iters, long long *cyc) {
    double v[CH];
#pragma unroll
    for (int c = 0; c < CH; ++c) v[c] = seed + c + threadIdx.x * 1e-9;
    const double m = 1.0000000001, a = 1e-9;
    long long t0 = clock64();
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int u = 0; u < 16; ++u) {
#pragma unroll
            for (int c = 0; c < CH; ++c) {
                if (OP == 0) v[c] = __dadd_rn(v[c], a);
                if (OP == 1) v[c] = __dmul_rn(v[c], m);
                if (OP == 2) v[c] = __fma_rn(v[c], m, a);
                if (OP == 3) { double p = __dmul_rn(v[(c + 1) % CH], m); v[c] = __dadd_rn(v[c], p); }  // mul feeding an add chain
            }
        }
    }
    long long t1 = clock64();
    double s = 0;
#pragma unroll
    for (int c = 0; c < CH; ++c) s += v[c];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}

template <int OP, int CH>
void run(const char *name, int warps_per_sm, double *out, long long *cyc) {
    const int iters = 2000;
    chains<OP, CH><<<148, warps_per_sm * 32>>>(out, 1.0, 10, cyc);
    cudaDeviceSynchronize();
    chains<OP, CH><<<148, warps_per_sm * 32>>>(out, 1.0, iters, cyc);
    cudaDeviceSynchronize();
    long long c;
    cudaMemcpy(&c, cyc, sizeof c, cudaMemcpyDeviceToHost);
    const double per_round = (double)c / (iters * 16.0);           // cycles for one instruction of every chain
    const int fp = (OP == 3) ? 2 : 1;
    const double warp_inst_per_clk_smsp = (double)CH * fp * (warps_per_sm / 4.0) / per_round;
    printf("%-8s chains/warp %2d warps/SM %2d : %.2f cycles per round, %.3f FP64 warp-inst/clk/sub-partition (pipe max 0.5)\n", name,
           CH, warps_per_sm, per_round, warp_inst_per_clk_smsp);
}

int main() {
    double *out;
    long long *cyc;
    cudaMalloc(&out, 148 * 1024 * sizeof(double));
    cudaMalloc(&cyc, sizeof(long long));
    printf("== dependent-issue latency (1 chain, 1 warp per sub-partition) ==\n");
    run<0, 1>("DADD", 4, out, cyc);
    run<1, 1>("DMUL", 4, out, cyc);
    run<2, 1>("DFMA", 4, out, cyc);
    run<3, 1>("MUL+ADD", 4, out, cyc);
    printf("== ILP needed: C chains per warp, 1 and 2 warps per sub-partition ==\n");
    run<0, 2>("DADD", 4, out, cyc);
    run<0, 3>("DADD", 4, out, cyc);
    run<0, 4>("DADD", 4, out, cyc);
    run<0, 6>("DADD", 4, out, cyc);
    run<0, 8>("DADD", 4, out, cyc);
    run<0, 1>("DADD", 8, out, cyc);
    run<0, 2>("DADD", 8, out, cyc);
    run<0, 3>("DADD", 8, out, cyc);
    run<0, 4>("DADD", 8, out, cyc);
    run<3, 2>("MUL+ADD", 4, out, cyc);
    run<3, 4>("MUL+ADD", 4, out, cyc);
    run<3, 2>("MUL+ADD", 8, out, cyc);
    run<3, 4>("MUL+ADD", 8, out, cyc);
    run<2, 2>("DFMA", 8, out, cyc);
    run<2, 4>("DFMA", 8, out, cyc);
    run<2, 8>("DFMA", 8, out, cyc);
    run<2, 4>("DFMA", 16, out, cyc);
    return 0;
}
