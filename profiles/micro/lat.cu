// Latency microbenchmarks on one warp (cycles per dependent op), to ground the solve-kernel design:
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o lat lat.cu && ./lat
#include <cstdio>
#include <cuda_runtime.h>
#include <math.h>
__device__ __forceinline__ double seed(double d) { double x; asm volatile("rcp.approx.ftz.f64 %0, %1;" : "=d"(x) : "d"(d)); return x; }
__device__ __forceinline__ double frcp(double d) { double x = seed(d); double e = fma(-d, x, 1.0); x = fma(x, e, x); e = fma(-d, x, 1.0); return fma(x, e, x); }
__device__ __forceinline__ double frcp3(double d) { double x = seed(d); double e = fma(-d, x, 1.0); double t = fma(e, e, e); return fma(x, t, x); }
#define REP 512
__global__ void k(double *out, long long *cyc, double a0, double m, int nw)
{
    __shared__ double sh[64];
    const int lane = threadIdx.x & 31;
    if (threadIdx.x >= 32 * nw) return;
    double a = a0 + lane * 1e-3; long long t0, t1; int n = 0;
    sh[lane] = a; sh[lane + 32] = a; __syncwarp();
    // 0: dependent DFMA
    t0 = clock64();
#pragma unroll 16
    for (int i = 0; i < REP; i++) a = fma(a, m, 1e-9);
    t1 = clock64(); if (threadIdx.x == 0) cyc[n] = t1 - t0; n++;
    // 1: 4 independent DFMA chains (throughput per warp)
    double b = a + 1, c = a + 2, d = a + 3;
    t0 = clock64();
#pragma unroll 16
    for (int i = 0; i < REP; i++) { a = fma(a, m, 1e-9); b = fma(b, m, 1e-9); c = fma(c, m, 1e-9); d = fma(d, m, 1e-9); }
    t1 = clock64(); if (threadIdx.x == 0) cyc[n] = t1 - t0; n++; a += b + c + d;
    // 2: dependent MUFU seed
    t0 = clock64();
#pragma unroll 16
    for (int i = 0; i < REP; i++) a = seed(a);
    t1 = clock64(); if (threadIdx.x == 0) cyc[n] = t1 - t0; n++;
    // 3: dependent frcp (seed + 4 FMA)
    a = fabs(a) + 1.5;
    t0 = clock64();
#pragma unroll 16
    for (int i = 0; i < REP; i++) a = frcp(a);
    t1 = clock64(); if (threadIdx.x == 0) cyc[n] = t1 - t0; n++;
    // 4: dependent 1.0/x
    t0 = clock64();
#pragma unroll 16
    for (int i = 0; i < REP; i++) a = 1.0 / a;
    t1 = clock64(); if (threadIdx.x == 0) cyc[n] = t1 - t0; n++;
    // 5: dependent 64-bit shuffle
    t0 = clock64();
#pragma unroll 16
    for (int i = 0; i < REP; i++) a = __shfl_sync(0xffffffffu, a, (lane + 1) & 31);
    t1 = clock64(); if (threadIdx.x == 0) cyc[n] = t1 - t0; n++;
    // 6: dependent LDS.64 (pointer chase through values)
    int idx = lane;
    t0 = clock64();
#pragma unroll 16
    for (int i = 0; i < REP; i++) { double v = sh[idx]; idx = (int)(v * 0.0) + ((idx + 1) & 63); }
    t1 = clock64(); if (threadIdx.x == 0) cyc[n] = t1 - t0; n++; a += idx;
    // 7: STS -> syncwarp -> LDS round trip
    t0 = clock64();
#pragma unroll 16
    for (int i = 0; i < REP; i++) { sh[lane] = a; __syncwarp(); a = sh[(lane + 1) & 31] + 1.0; __syncwarp(); }
    t1 = clock64(); if (threadIdx.x == 0) cyc[n] = t1 - t0; n++;
    // 8: dependent frcp3 (seed + 3 FMA)
    a = fabs(a) + 1.5;
    t0 = clock64();
#pragma unroll 16
    for (int i = 0; i < REP; i++) a = frcp3(a);
    t1 = clock64(); if (threadIdx.x == 0) cyc[n] = t1 - t0; n++;
    // 9: dependent DADD / DMUL
    t0 = clock64();
#pragma unroll 16
    for (int i = 0; i < REP; i++) a = a * m;
    t1 = clock64(); if (threadIdx.x == 0) cyc[n] = t1 - t0; n++;
    // 10: 8 independent chains
    double e2 = a + 4, f2 = a + 5, g2 = a + 6, h2 = a + 7;
    t0 = clock64();
#pragma unroll 8
    for (int i = 0; i < REP; i++) { a = fma(a, m, 1e-9); b = fma(b, m, 1e-9); c = fma(c, m, 1e-9); d = fma(d, m, 1e-9);
        e2 = fma(e2, m, 1e-9); f2 = fma(f2, m, 1e-9); g2 = fma(g2, m, 1e-9); h2 = fma(h2, m, 1e-9); }
    t1 = clock64(); if (threadIdx.x == 0) cyc[n] = t1 - t0; n++; a += b + c + d + e2 + f2 + g2 + h2;
    out[threadIdx.x] = a;
}
__global__ void acc(double *err)
{
    // accuracy of seed / frcp / frcp3 over a sweep of inputs
    double es = 0, e4 = 0, e3 = 0;
    for (int i = 0; i < 200000; i++) {
        double d = (1.0 + i * 5e-6) * ((i & 1) ? 1e-7 : 3e5);
        double ex = 1.0 / d;
        es = fmax(es, fabs(seed(d) - ex) / ex); e4 = fmax(e4, fabs(frcp(d) - ex) / ex); e3 = fmax(e3, fabs(frcp3(d) - ex) / ex);
    }
    err[0] = es; err[1] = e4; err[2] = e3;
}
int main()
{
    double *out, *err; long long *cyc;
    cudaMalloc(&out, 1024 * 8); cudaMalloc(&cyc, 64 * 8); cudaMalloc(&err, 64);
    const char *names[] = {"dep DFMA", "4 indep DFMA (per 4)", "dep MUFU.RCP64H", "dep frcp(seed+4fma)", "dep 1.0/x", "dep SHFL64", "dep LDS64", "STS+sync+LDS+sync", "dep frcp3(seed+3fma)", "dep DMUL", "8 indep DFMA (per 8)"};
    for (int nw = 1; nw <= 4; nw *= 2) {
        k<<<1, 128>>>(out, cyc, 1.0000001, 0.9999999, nw); cudaDeviceSynchronize();
        k<<<1, 128>>>(out, cyc, 1.0000001, 0.9999999, nw); cudaDeviceSynchronize();
        long long h[16]; cudaMemcpy(h, cyc, sizeof h, cudaMemcpyDeviceToHost);
        printf("warps in CTA = %d (they share the SM: 1 per SMSP up to 4)\n", nw);
        for (int i = 0; i < 11; i++) printf("  %-24s %7.2f cycles/iter\n", names[i], (double)h[i] / REP);
    }
    acc<<<1, 1>>>(err); double e[3]; cudaMemcpy(e, err, sizeof e, cudaMemcpyDeviceToHost);
    printf("max rel err: seed %.3e  frcp(2 newton) %.3e  frcp3(cubic) %.3e\n", e[0], e[1], e[2]);
    printf("%s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
