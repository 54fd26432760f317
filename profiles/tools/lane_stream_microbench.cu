/* profiles/tools/lane_stream_microbench.cu -- can ONE warp per SM, fed by cp.async.bulk (TMA 1-D bulk copies, UBLKCP) into a
 * double-buffered shared-memory stage and completing on an mbarrier, stream the engine's per-stage working set at HBM speed
 * while it computes?  Geometry of the lane-per-instance design: a TILE of 32 instances per warp, stage-major tile
 * records, per stage `in_doubles` doubles per instance staged (tile chunk = in_doubles * 256 B, contiguous), `fma` dependent-ish
 * FP64 FMAs per lane and stage on operands read from the staged buffer, `out_doubles` per instance written back coalesced.
 *   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o lane_stream_microbench lane_stream_microbench.cu && ./lane_stream_microbench */
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <cstdint>

__device__ __forceinline__ void mbar_init(uint64_t *bar, int count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"((unsigned) __cvta_generic_to_shared(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, unsigned bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"((unsigned) __cvta_generic_to_shared(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, unsigned phase)
{
    asm volatile(
        "{\n.reg .pred p;\nWAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE_%=;\nbra WAIT_%=;\nDONE_%=:\n}" ::"r"((unsigned) __cvta_generic_to_shared(bar)), "r"(phase) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void *dst, const void *src, unsigned bytes, uint64_t *bar)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     (unsigned) __cvta_generic_to_shared(dst)),
                 "l"(src), "r"(bytes), "r"((unsigned) __cvta_generic_to_shared(bar))
                 : "memory");
}

__global__ void __launch_bounds__(32, 1) stream_kernel(const double *in, double *out, int stages, int in_doubles, int out_doubles, int nfma,
                                                        int rec_doubles)
{
    extern __shared__ __align__(128) unsigned char smem_raw[];
    double *buf = reinterpret_cast<double *>(smem_raw);
    __shared__ uint64_t bar[2];
    const int lane = threadIdx.x, tile = blockIdx.x;
    const size_t tile_stride = (size_t) stages * rec_doubles * 32;
    const double *tin = in + tile * tile_stride;
    double *tout = out + tile * tile_stride;
    const unsigned bytes = (unsigned) in_doubles * 256u;
    if (lane == 0) { mbar_init(&bar[0], 1); mbar_init(&bar[1], 1); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    __syncwarp();
    if (lane == 0) { mbar_expect_tx(&bar[0], bytes); bulk_g2s(buf, tin, bytes, &bar[0]); }
    double acc = 0.0;
    for (int s = 0; s < stages; s++)
    {
        const int b = s & 1;
        if (s + 1 < stages && lane == 0)
        {
            mbar_expect_tx(&bar[b ^ 1], bytes);
            bulk_g2s(buf + (size_t) (b ^ 1) * in_doubles * 32, tin + (size_t) (s + 1) * rec_doubles * 32, bytes, &bar[b ^ 1]);
        }
        mbar_wait(&bar[b], (s >> 1) & 1);
        const double *g = buf + (size_t) b * in_doubles * 32;
        double a0 = acc, a1 = 1.0, a2 = 2.0, a3 = 3.0;
        for (int k = 0; k < nfma; k += 4)
        {
            const int c = (k >> 2) % (in_doubles - 3);
            a0 = fma(g[c * 32 + lane], a0, 1e-3);
            a1 = fma(g[(c + 1) * 32 + lane], a1, 1e-3);
            a2 = fma(g[(c + 2) * 32 + lane], a2, 1e-3);
            a3 = fma(g[(c + 3) * 32 + lane], a3, 1e-3);
        }
        acc = a0 + a1 + a2 + a3;
        double *o = tout + (size_t) s * rec_doubles * 32 + (size_t) in_doubles * 32;
        for (int k = 0; k < out_doubles; k++) o[k * 32 + lane] = acc + g[(k % in_doubles) * 32 + lane];
        __syncwarp(); /* everyone is done with buffer b before it is refilled two iterations later */
    }
    if (acc == 12345.678) out[0] = acc;
}

int main(int argc, char **argv)
{
    const int tiles = argc > 1 ? atoi(argv[1]) : 128, stages = 100;
    const int rec = 832; /* doubles per stage and instance in the slab */
    struct Cfg { const char *name; int in, out, fma; } cfgs[] = {
        {"factor-like  in162 out202", 162, 202, 1000}, {"forward-like in352 out78 ", 352, 78, 800},
        {"backward-like in274 out42", 274, 42, 400},   {"residual-like in270 out156", 270, 156, 300},
        {"forward-like, no compute ", 352, 78, 0},      {"forward-like, 3000 fma   ", 352, 78, 3000},
        {"factor-like, 4000 fma    ", 162, 202, 4000}};
    const size_t n = (size_t) tiles * stages * rec * 32;
    double *in, *out;
    cudaMalloc(&in, n * 8); cudaMalloc(&out, n * 8);
    cudaMemset(in, 0, n * 8); cudaMemset(out, 0, n * 8);
    cudaFuncSetAttribute(stream_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    for (auto &c : cfgs)
    {
        const size_t sm = (size_t) 2 * c.in * 256;
        cudaEvent_t e0, e1;
        cudaEventCreate(&e0); cudaEventCreate(&e1);
        for (int rep = 0; rep < 3; rep++)
        {
            cudaEventRecord(e0);
            for (int it = 0; it < 5; it++) stream_kernel<<<tiles, 32, sm>>>(in, out, stages, c.in, c.out, c.fma, rec);
            cudaEventRecord(e1);
            cudaEventSynchronize(e1);
        }
        float ms;
        cudaEventElapsedTime(&ms, e0, e1);
        ms /= 5;
        const double bytes = (double) tiles * stages * (c.in + c.out) * 256.0;
        printf("{\"case\": \"%s\", \"tiles\": %d, \"fma_per_stage\": %d, \"ms\": %.4f, \"GBps\": %.1f, \"us_per_stage\": %.3f, \"err\": \"%s\"}\n", c.name, tiles,
               c.fma, ms, bytes / ms / 1e6, ms * 1e3 / stages, cudaGetErrorString(cudaGetLastError()));
    }
    return 0;
}
