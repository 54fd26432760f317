// Dependent-issue latencies of the instructions the Riccati sweeps are made of (one warp, one SM): cycles per operation of a
// dependent chain.  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o lat_microbench lat_microbench.cu
#include <cstdio>
#include <cuda_runtime.h>
#define REP 4096
template <int MODE>
__global__ void k(double *out, long long *cyc, double seed, int idx)
{
    __shared__ double sm[512];
    for (int i = threadIdx.x; i < 512; i += blockDim.x) sm[i] = seed * 1e-9 * i + (MODE == 3 ? 0.0 : 1.0);
    __syncthreads();
    double a = seed + threadIdx.x * 1e-3, b = 1.0000001, c = 1e-9, a2 = a + 1, a3 = a + 2, a4 = a + 3;
    int p = idx;
    long long t0 = clock64();
#pragma unroll 16
    for (int i = 0; i < REP; i++)
    {
        if (MODE == 0) a = fma(a, b, c);                                  // DFMA chain
        if (MODE == 1) { a = fma(a, b, c); a2 = fma(a2, b, c); a3 = fma(a3, b, c); a4 = fma(a4, b, c); } // 4 independent chains
        if (MODE == 2) a = rsqrt(a) + 1.5;                               // rsqrt + add chain
        if (MODE == 3) { p = (int) sm[p]; }                               // LDS pointer chase (double -> int)
        if (MODE == 4) a = 1.0 / a + 0.5;                                 // IEEE division + add
        if (MODE == 5) a = __shfl_xor_sync(0xffffffffu, a, 1) + c;        // 64-bit shuffle + add
        if (MODE == 6) a = fma(sm[(i * 7) & 511], b, a);                  // independent LDS feeding a DFMA chain
        if (MODE == 7) { a = sqrt(a) + 1.5; }                             // sqrt
        if (MODE == 8) { a = a * b; }                                     // DMUL chain
        if (MODE == 9) { a = a + b; }                                     // DADD chain
        if (MODE == 10) { float f = (float) a; f = fmaf(f, 1.0001f, 1e-6f); a = f; } // conversions
    }
    long long t1 = clock64();
    if (threadIdx.x == 0) cyc[MODE] = t1 - t0;
    out[threadIdx.x + 32 * MODE] = a + a2 + a3 + a4 + p;
}
int main()
{
    double *out; long long *cyc;
    cudaMalloc(&out, 8 * 32 * 16); cudaMalloc(&cyc, 8 * 16);
    const char *name[] = {"DFMA dependent", "4 independent DFMA chains (per 4 ops)", "rsqrt(double) + DADD", "LDS pointer chase (+ F2I)", "1.0 / x + DADD",
                          "SHFL (64-bit) + DADD", "LDS + DFMA (loads independent)", "sqrt(double) + DADD", "DMUL dependent", "DADD dependent", "D2F FFMA F2D"};
#define RUN(M) k<M><<<1, 32>>>(out, cyc, 1.25, 0);
    for (int rep = 0; rep < 2; rep++) { RUN(0) RUN(1) RUN(2) RUN(3) RUN(4) RUN(5) RUN(6) RUN(7) RUN(8) RUN(9) RUN(10) }
    cudaDeviceSynchronize();
    long long h[16];
    cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
    for (int m = 0; m <= 10; m++) printf("%-45s %8.1f cycles per iteration\n", name[m], (double) h[m] / REP);
    printf("%s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
