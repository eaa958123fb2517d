// Is the FP64 tensor path (mma.sync.m8n8k4.f64, "DMMA") a SECOND FP64 resource beside the vector pipe on B200, or the same one?
// One warp per scheduler (4 warps per CTA, one CTA on one SM): cycles per instruction for (a) 8 independent DFMA chains, (b) 8 independent
// DMMA accumulators, (c) both interleaved.  If (c) ~ max(a, b) the two overlap; if (c) ~ a + b they share the pipe.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o dmma dmma.cu && ./dmma
#include <cstdio>
#include <cuda_runtime.h>
#define REP 512
__device__ __forceinline__ void dmma(double &c0, double &c1, double a, double b)
{
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}
__global__ void k(double *out, long long *cyc, double m, int mode)
{
    const int lane = threadIdx.x & 31;
    double a0 = lane * 1e-3, a1 = a0 + 1, a2 = a0 + 2, a3 = a0 + 3, a4 = a0 + 4, a5 = a0 + 5, a6 = a0 + 6, a7 = a0 + 7;
    double c[16];
    for (int i = 0; i < 16; i++) c[i] = i * 1e-3 + lane;
    const double fa = 1.0 + lane * 1e-9, fb = 1e-9;
    __syncthreads();
    const long long t0 = clock64();
#pragma unroll 4
    for (int i = 0; i < REP; i++) {
        if (mode & 1) {
            a0 = fma(a0, m, 1e-9); a1 = fma(a1, m, 1e-9); a2 = fma(a2, m, 1e-9); a3 = fma(a3, m, 1e-9);
            a4 = fma(a4, m, 1e-9); a5 = fma(a5, m, 1e-9); a6 = fma(a6, m, 1e-9); a7 = fma(a7, m, 1e-9);
        }
        if (mode & 2) {
#pragma unroll
            for (int j = 0; j < 8; j++) dmma(c[2 * j], c[2 * j + 1], fa, fb);
        }
    }
    const long long t1 = clock64();
    if (threadIdx.x == 0) cyc[0] = t1 - t0;
    double s = a0 + a1 + a2 + a3 + a4 + a5 + a6 + a7;
    for (int i = 0; i < 16; i++) s += c[i];
    out[threadIdx.x] = s;
}
int main()
{
    double *out; long long *cyc, h;
    cudaMalloc(&out, 8 * 256); cudaMalloc(&cyc, 64);
    const char *names[4] = {"", "8 DFMA", "8 DMMA m8n8k4", "8 DFMA + 8 DMMA interleaved"};
    for (int warps = 1; warps <= 8; warps *= 2)
        for (int mode = 1; mode <= 3; mode++) {
            k<<<1, 32 * warps>>>(out, cyc, 1.0000001, mode); k<<<1, 32 * warps>>>(out, cyc, 1.0000001, mode);
            cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
            printf("warps %d  %-30s: %8.2f cycles per loop trip (warp 0)\n", warps, names[mode], (double)h / REP);
        }
    printf("(a DFMA warp-instruction = 32 FMA; a DMMA m8n8k4 warp-instruction = 8x8x4 = 256 FMA)\n");
    return 0;
}
