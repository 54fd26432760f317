// Side experiment for DESIGN.md 9(d): DRAM bandwidth of the engine's access pattern under two slab layouts.
// B instances, S stages, A arrays of W doubles per stage.  One group of 16 lanes per instance walks the stages in order and at
// each stage reads every array's W doubles (16 bytes per lane) and rewrites the first AW arrays, as a sweep does.
//   layout 0 (array-major, the engine's): addr(i, a, s) = i*A*S*W + a*S*W + s*W     -> A separate 8*W-byte chunks per stage
//   layout 1 (stage-major):               addr(i, a, s) = i*A*S*W + s*A*W + a*W     -> one contiguous 8*A*W-byte record per stage
// nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o layout_microbench layout_microbench.cu && ./layout_microbench
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>

template <int LAYOUT, int A, int W, int AW>
__global__ void __launch_bounds__(64, 8) walk(double *slab, int B, int S, double *out)
{
    const int grp = threadIdx.x / 16, lane = threadIdx.x % 16;
    const long inst = (long) blockIdx.x * 4 + grp;
    if (inst >= B) return;
    double *base = slab + inst * (long) A * S * W;
    double acc = 0.0;
    for (int s = 0; s < S; s++)
    {
        double2 v[A];
#pragma unroll
        for (int a = 0; a < A; a++)
        {
            const long off = LAYOUT == 0 ? (long) a * S * W + (long) s * W : (long) s * A * W + (long) a * W;
            v[a] = make_double2(0.0, 0.0);
            if (2 * lane < W) v[a] = *reinterpret_cast<const double2 *>(base + off + 2 * lane);
        }
#pragma unroll
        for (int a = 0; a < A; a++)
        {
            acc += v[a].x + v[a].y;
            if (a < AW && 2 * lane < W)
            {
                const long off = LAYOUT == 0 ? (long) a * S * W + (long) s * W : (long) s * A * W + (long) a * W;
                *reinterpret_cast<double2 *>(base + off + 2 * lane) = make_double2(v[a].x + 1.0, v[a].y + 1.0);
            }
        }
    }
    if (acc == 123.456) out[inst] = acc;
}

template <int LAYOUT, int A, int W, int AW>
static void run(double *slab, int B, int S, double *out, const char *name)
{
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    for (int k = 0; k < 2; k++) walk<LAYOUT, A, W, AW><<<(B + 3) / 4, 64>>>(slab, B, S, out);
    cudaEventRecord(e0);
    const int reps = 5;
    for (int k = 0; k < reps; k++) walk<LAYOUT, A, W, AW><<<(B + 3) / 4, 64>>>(slab, B, S, out);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms;
    cudaEventElapsedTime(&ms, e0, e1);
    ms /= reps;
    const double bytes = (double) B * S * W * 8.0 * (A + AW);
    printf("{\"layout\": \"%s\", \"arrays\": %d, \"doubles_per_array\": %d, \"arrays_rewritten\": %d, \"instances\": %d, \"stages\": %d, "
           "\"ms\": %.3f, \"GBs\": %.1f, \"error\": \"%s\"}\n",
           name, A, W, AW, B, S, ms, bytes / (ms * 1e-3) / 1e9, cudaGetErrorString(cudaGetLastError()));
}

int main()
{
    const int B = 4096, S = 100;
    double *slab, *out;
    const size_t n = (size_t) B * S * 16 * 24; // room for the largest case (16 arrays x 24 doubles)
    cudaMalloc(&slab, n * sizeof(double));
    cudaMalloc(&out, B * sizeof(double));
    cudaMemset(slab, 0, n * sizeof(double));
    run<0, 10, 24, 3>(slab, B, S, out, "array-major");
    run<1, 10, 24, 3>(slab, B, S, out, "stage-major");
    run<0, 16, 24, 4>(slab, B, S, out, "array-major");
    run<1, 16, 24, 4>(slab, B, S, out, "stage-major");
    run<0, 10, 12, 3>(slab, B, S, out, "array-major");
    run<1, 10, 12, 3>(slab, B, S, out, "stage-major");
    run<0, 10, 24, 0>(slab, B, S, out, "array-major");
    run<1, 10, 24, 0>(slab, B, S, out, "stage-major");
    return 0;
}
