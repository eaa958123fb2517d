// Does a DFMA warp-instruction with only 16 (or 8) active lanes cost less of the FP64 pipe than one with 32?  B200's FP64 pipe is 16
// lanes per SM sub-partition: a full warp takes two passes.  One warp, 8 independent chains (throughput, not latency), lanes masked by
// a divergent branch.   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o halfmask halfmask.cu && ./halfmask
#include <cstdio>
#include <cuda_runtime.h>
#define REP 1024
__global__ void k(double *out, long long *cyc, double m, int active_lanes, int nwarps)
{
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (warp >= nwarps) return;
    double a0 = lane * 1e-3, a1 = a0 + 1, a2 = a0 + 2, a3 = a0 + 3, a4 = a0 + 4, a5 = a0 + 5, a6 = a0 + 6, a7 = a0 + 7;
    long long t0 = 0, t1 = 0;
    __syncthreads();
    if (lane < active_lanes) {
        t0 = clock64();
#pragma unroll 8
        for (int i = 0; i < REP; i++) {
            a0 = fma(a0, m, 1e-9); a1 = fma(a1, m, 1e-9); a2 = fma(a2, m, 1e-9); a3 = fma(a3, m, 1e-9);
            a4 = fma(a4, m, 1e-9); a5 = fma(a5, m, 1e-9); a6 = fma(a6, m, 1e-9); a7 = fma(a7, m, 1e-9);
        }
        t1 = clock64();
    }
    if (lane == 0) cyc[warp] = t1 - t0;
    out[threadIdx.x] = a0 + a1 + a2 + a3 + a4 + a5 + a6 + a7;
}
int main()
{
    double *out; long long *cyc, h[8];
    cudaMalloc(&out, 8 * 256); cudaMalloc(&cyc, 64);
    for (int nw = 1; nw <= 8; nw *= 2)          // warps w, w+4 share a scheduler: 1, 2, 4 warps = one per scheduler, 8 = two per scheduler
        for (int act = 32; act >= 8; act /= 2) {
            k<<<1, 256>>>(out, cyc, 1.0000001, act, nw); k<<<1, 256>>>(out, cyc, 1.0000001, act, nw);
            cudaMemcpy(h, cyc, 64, cudaMemcpyDeviceToHost);
            printf("warps %d  active lanes %2d : %.3f cycles per DFMA warp-instruction (warp 0)\n", nw, act, (double)h[0] / (8.0 * REP));
        }
    return 0;
}
