/* robotic-mpc_b200/csrc/rmpc_capi.cu -- C ABI (include/rmpc.h) over the sm_100a kernels of rmpc_device.cuh.
 *
 * Launch geometry: one group of RMPC_G lanes per instance (two instances per warp by default), RMPC_WARPS_PER_CTA
 * warps per CTA; the grid covers the batch.  CTAs are small -- 64 threads = 4 instances, 8 CTAs resident per SM -- so
 * that the block scheduler can spread a batch evenly: with 4096 instances, 1024 CTAs land 7 or 6 per SM (28 / 24
 * instances), where 128-thread CTAs gave 4 or 3 (32 / 24) and left more of the slower SMs' tail exposed (measured on
 * B200, 4096 x N=100: 506k MPC steps/s with 2 warps per CTA, 496k with 3, 478k with 4).  The latency of the sequential
 * Riccati sweeps of one instance is hidden by the other instances on the SM.
 * No host fallback: every entry point needs the CUDA device the handle was created on.
 */
#include <cuda_runtime.h>

#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>
#include <vector>

#include "../../include/rmpc.h"
#include "rmpc_device.cuh"
#include "rmpc_layout.h"

#ifndef RMPC_WARPS_PER_CTA
#define RMPC_WARPS_PER_CTA 2
#endif
#define RMPC_GROUPS_PER_CTA (RMPC_WARPS_PER_CTA * 32 / RMPC_G)

static_assert(sizeof(rmpc_dims) == sizeof(RmpcDims), "rmpc_dims / RmpcDims layout mismatch");

/* ------------------------------------------------------------------------------------------------ kernels */
#ifndef RMPC_MIN_CTAS
#define RMPC_MIN_CTAS 8
#endif
#ifdef RMPC_MAXNREG
#define RMPC_KERNEL_BOUNDS __maxnreg__(RMPC_MAXNREG)
#else
#define RMPC_KERNEL_BOUNDS __launch_bounds__(32 * RMPC_WARPS_PER_CTA, RMPC_MIN_CTAS)
#endif
__global__ void RMPC_KERNEL_BOUNDS
rmpc_closed_loop_kernel(const __grid_constant__ RmpcChain ch, const __grid_constant__ RmpcDims D, const __grid_constant__ RmpcWs W, int nsteps,
                        int advance)
{
    extern __shared__ __align__(16) unsigned char smem_raw[]; /* > 48 KB per CTA: dynamic, opted in at rmpc_create */
    GrpSmem *smem = reinterpret_cast<GrpSmem *>(smem_raw);
    const int grp = threadIdx.x / RMPC_G;
    const long inst = (long) blockIdx.x * RMPC_GROUPS_PER_CTA + grp;
    if (inst >= D.B) return; /* a whole group exits together; the barriers of the other group carry its own mask */
    /* groups of this warp that run an instance: they re-converge at fixed points of every closed-loop step */
    const long first = inst - (grp % (32 / RMPC_G));
    unsigned pair = 0;
    for (int g = 0; g < 32 / RMPC_G; g++)
        if (first + g < D.B) pair |= (RMPC_G == 32 ? 0xffffffffu : ((1u << (RMPC_G & 31)) - 1u)) << (g * RMPC_G);
    closed_loop_group(ch, D, W, inst, smem[grp], nsteps, advance, pair);
}

/* x_k = z, u_k = 0 for every stage; flags = {first_call, qp_warm_pending, 0, 0} */
__global__ void rmpc_init_iterate_kernel(const RmpcDims D, const RmpcWs W)
{
    const long S = D.N + 1;
    const long tot = (long) D.B * S * NV;
    for (long k = (long) blockIdx.x * blockDim.x + threadIdx.x; k < tot; k += (long) gridDim.x * blockDim.x)
    {
        const long i = k / (S * NV), r = k % (S * NV);
        const int c = (int) (k % NV);
        W.slab[i * W.slab_stride + RMPC_HEAD + (long) RC_ux * W.Sp + r] = c >= 6 ? W.z[i * NX + c - 6] : 0.0;
    }
    for (long k = (long) blockIdx.x * blockDim.x + threadIdx.x; k < (long) D.B * 4; k += (long) gridDim.x * blockDim.x)
        W.flags[k] = (k % 4) < 2 ? 1 : 0;
    if (W.zlog)
        for (long k = (long) blockIdx.x * blockDim.x + threadIdx.x; k < (long) D.B * NX; k += (long) gridDim.x * blockDim.x)
            W.zlog[(k / NX) * (W.nlog + 1) * NX + (k % NX)] = W.z[k];
}

/* dst[i][k] = src[i*stride + off + k]  (gather == 1)   or   src[i*stride + off + k] = dst[i][k]  (gather == 0) */
__global__ void rmpc_strided_copy_kernel(double *packed, double *strided, long B, long stride, long off, int n, int gather)
{
    const long tot = B * n;
    for (long k = (long) blockIdx.x * blockDim.x + threadIdx.x; k < tot; k += (long) gridDim.x * blockDim.x)
    {
        const long i = k / n, c = k % n;
        if (gather) packed[k] = strided[i * stride + off + c];
        else strided[i * stride + off + c] = packed[k];
    }
}

/* one thread per logged plant sample: pose / rpy / twist logs and tracking errors (R/simulation_model.py:60-77,
 * R/simulator.py:265-344), out = [B][nlog+1][RK_COUNT] */
__global__ void rmpc_plant_logs_kernel(const RmpcChain ch, const double *par, const double *zlog, double *out, long B, int nsamp)
{
    const long tot = B * nsamp;
    for (long k = (long) blockIdx.x * blockDim.x + threadIdx.x; k < tot; k += (long) gridDim.x * blockDim.x)
    {
        const long i = k / nsamp;
        double z[NX], e[5], rec[RK_COUNT];
        for (int c = 0; c < NX; c++) z[c] = zlog[k * NX + c];
        rec[RK_COUNT - 1] = 0.0;
        ur10_plant_record<true>(ch, par + i * RP_COUNT, z, z + 6, e, rec);
        for (int c = 0; c < RK_COUNT; c++) out[k * RK_COUNT + c] = rec[c];
    }
}

/* per-stage table [B][N][RT_COUNT] <- the per-instance values of the parameter block, for every stage, columns
 * c0 .. c0 + nc - 1 of the table only */
__global__ void rmpc_stage_table_fill_kernel(double *stg, const double *par, long B, int N, int c0, int nc)
{
    const long tot = B * N * RT_COUNT;
    for (long k = (long) blockIdx.x * blockDim.x + threadIdx.x; k < tot; k += (long) gridDim.x * blockDim.x)
    {
        const long i = k / ((long) N * RT_COUNT);
        const int c = (int) (k % RT_COUNT);
        const double *p = par + i * RP_COUNT;
        if (c < c0 || c >= c0 + nc) continue;
        double v;
        if (c < RT_LB) v = p[RP_YREF + c];
        else if (c < RT_UB) v = (c - RT_LB) < 6 ? p[RP_UMIN + c - RT_LB] : p[RP_QMIN + c - RT_LB - 6];
        else v = (c - RT_UB) < 6 ? p[RP_UMAX + c - RT_UB] : p[RP_QMAX + c - RT_UB - 6];
        stg[k] = v;
    }
}

__global__ void rmpc_flags_to_double_kernel(const int *flags, double *out, long n)
{
    for (long k = (long) blockIdx.x * blockDim.x + threadIdx.x; k < n; k += (long) gridDim.x * blockDim.x) out[k] = flags[k];
}

/* ------------------------------------------------------------------------------------------------ handle */
struct rmpc_handle
{
    RmpcDims D;
    RmpcChain ch;
    RmpcWs W;
    int device;
    int nlog;
    int steps_done;
    long long bytes;
    long long launches;
    std::vector<void *> owned;
    double *par_dev;
    double *pe_dev; /* [B][RE_COUNT] terminal-set blocks, allocated by the first "P_e" / "alpha_e" / "lh_e" / "term" set */
    double *stg_dev; /* [B][N][RT_COUNT] per-stage references and bounds, allocated by the first per-stage (stage >= 0) set */
    double *sqpstat_dev; /* [B][max_iter + 2][RQ_COUNT], allocated by rmpc_set_option("statistics", 1) */
    double *asrti_dev = nullptr; /* [B][RMPC_ASRTI_STRIDE], allocated by rmpc_set_option("as_rti_level", 0) */
    /* rmpc_step_host: pinned staging of x0 / u0 / step record, the stream used when the caller passes NULL (a captured
     * sequence cannot live on the legacy stream), and the instantiated graph with the launch arguments it was captured with */
    double *pin_z = nullptr, *pin_u = nullptr, *pin_st = nullptr;
    cudaStream_t own_stream = nullptr;
    cudaGraphExec_t gexec = nullptr;
    cudaStream_t gstream = nullptr;
    RmpcWs gW;
    RmpcDims gD;
    RmpcChain gch;
    int use_graph = 1;
    std::string err;
};

static thread_local std::string g_create_err;

static int fail(rmpc_handle *h, int code, const std::string &msg)
{
    if (h) h->err = msg;
    else g_create_err = msg;
    return code;
}

#define CU(call)                                                                                     \
    do                                                                                               \
    {                                                                                                \
        cudaError_t e_ = (call);                                                                     \
        if (e_ != cudaSuccess) return fail(h, RMPC_ECUDA, std::string(#call ": ") + cudaGetErrorString(e_)); \
    } while (0)

/* Every entry point runs on the handle's device and puts the caller's current device back on every exit path (PyTorch
 * reads the current device from the runtime: an Engine on device k must not switch the process to k as a side effect). */
struct DeviceGuard
{
    int prev = -1;
    bool ok = true;
    explicit DeviceGuard(int dev)
    {
        if (cudaGetDevice(&prev) != cudaSuccess) prev = -1;
        if (prev != dev) ok = (cudaSetDevice(dev) == cudaSuccess);
        else prev = -1; /* nothing to restore */
    }
    ~DeviceGuard()
    {
        if (prev >= 0) cudaSetDevice(prev);
    }
};
#define ON_DEVICE(h)                                                                          \
    DeviceGuard dev_guard_((h)->device);                                                      \
    if (!dev_guard_.ok) return fail((h), RMPC_ECUDA, "cudaSetDevice failed for the handle's device")

static int grid_for(long n) { long g = (n + 255) / 256; return (int) (g < 1 ? 1 : (g > 148 * 8 ? 148 * 8 : g)); }

extern "C" const char *rmpc_last_error(const rmpc_handle *h) { return h ? h->err.c_str() : g_create_err.c_str(); }
extern "C" long long rmpc_device_bytes(const rmpc_handle *h) { return h ? h->bytes : 0; }
extern "C" int rmpc_steps_done(const rmpc_handle *h) { return h ? h->steps_done : 0; }
extern "C" long long rmpc_launch_count(const rmpc_handle *h) { return h ? h->launches : 0; }

extern "C" int rmpc_destroy(rmpc_handle *h)
{
    if (!h) return 0;
    DeviceGuard g(h->device);
    for (void *p : h->owned) cudaFree(p);
    if (h->gexec) cudaGraphExecDestroy(h->gexec);
    if (h->own_stream) cudaStreamDestroy(h->own_stream);
    if (h->pin_z) cudaFreeHost(h->pin_z);
    if (h->pin_u) cudaFreeHost(h->pin_u);
    if (h->pin_st) cudaFreeHost(h->pin_st);
    delete h;
    return 0;
}

extern "C" int rmpc_create(const rmpc_dims *dims, int device, int nlog, rmpc_handle **out)
{
    rmpc_handle *h = nullptr;
    if (!dims || !out) return fail(nullptr, RMPC_EINVAL, "rmpc_create: null argument");
    if (dims->B < 1 || dims->N < 2 || dims->N > 4096 || nlog < 0)
        return fail(nullptr, RMPC_EINVAL, "rmpc_create: need B >= 1, 2 <= N <= 4096, nlog >= 0");
    if (dims->solver_type != RMPC_SOLVER_SQP && dims->solver_type != RMPC_SOLVER_SQP_RTI)
        return fail(nullptr, RMPC_EINVAL, "rmpc_create: solver_type must be RMPC_SOLVER_SQP or RMPC_SOLVER_SQP_RTI");
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev < 1 || device < 0 || device >= ndev)
        return fail(nullptr, RMPC_ENODEV, "rmpc_create: no usable CUDA device (this library has no host fallback)");
    DeviceGuard dev_guard_(device);
    if (!dev_guard_.ok) return fail(nullptr, RMPC_ENODEV, "rmpc_create: cudaSetDevice failed");
    if (cudaFuncSetAttribute(rmpc_closed_loop_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                             (int) (sizeof(GrpSmem) * RMPC_GROUPS_PER_CTA)) != cudaSuccess)
        return fail(nullptr, RMPC_ECUDA, "rmpc_create: cannot reserve the shared memory the closed-loop kernel needs");
    h = new rmpc_handle();
    memcpy(&h->D, dims, sizeof(RmpcDims));
    rmpc_chain_init(&h->ch);
    memset(&h->W, 0, sizeof(h->W));
    h->device = device; h->nlog = nlog; h->steps_done = 0; h->bytes = 0; h->launches = 0;
    h->par_dev = nullptr;
    h->pe_dev = nullptr;
    h->stg_dev = nullptr;
    h->sqpstat_dev = nullptr;
    const size_t B = (size_t) dims->B, S = (size_t) dims->N + 1;
    auto dalloc = [&](size_t n_doubles, void **p) -> bool {
        const size_t bytes = n_doubles * sizeof(double);
        if (cudaMalloc(p, bytes) != cudaSuccess) return false;
        cudaMemset(*p, 0, bytes);
        h->owned.push_back(*p);
        h->bytes += (long long) bytes;
        return true;
    };
    bool ok = true;
    (void) S;
    h->W.slab_stride = rmpc_slab_stride(dims->N);
    h->W.Sp = rmpc_slab_sp(dims->N);
    ok = ok && dalloc(B * (size_t) h->W.slab_stride, (void **) &h->W.slab);
    ok = ok && dalloc(B * NX, (void **) &h->W.z);
    ok = ok && dalloc(B * RP_COUNT, (void **) &h->par_dev);
    ok = ok && dalloc(B * RS_COUNT, (void **) &h->W.summary);
    void *fl = nullptr;
    ok = ok && dalloc(B * 2, &fl); /* 4 ints per instance */
    h->W.flags = (int *) fl;
    if (nlog > 0)
    {
        ok = ok && dalloc(B * (size_t) (nlog + 1) * NX, (void **) &h->W.zlog);
        ok = ok && dalloc(B * (size_t) nlog * 6, (void **) &h->W.ulog);
        ok = ok && dalloc(B * (size_t) nlog * RL_COUNT, (void **) &h->W.slog);
    }
    if (!ok)
    {
        cudaGetLastError();
        rmpc_destroy(h);
        return fail(nullptr, RMPC_ENOMEM, "rmpc_create: device allocation failed");
    }
    /* the zero fills above run on the legacy default stream; the caller's streams may be non-blocking ones that are not
     * ordered against it, so they must have landed before the first rmpc_set / rmpc_reset on any stream */
    if (cudaDeviceSynchronize() != cudaSuccess)
    {
        rmpc_destroy(h);
        return fail(nullptr, RMPC_ECUDA, "rmpc_create: zero-initialisation of the workspaces failed");
    }
    h->W.par = h->par_dev;
    h->W.nlog = nlog;
    h->W.step0 = 0;
    h->W.as_rti_iter = 1;
    *out = h;
    return 0;
}

extern "C" int rmpc_set_option(rmpc_handle *h, const char *name, double value)
{
    if (!h || !name) return fail(h, RMPC_EINVAL, "rmpc_set_option: null argument");
    const std::string f(name);
    if (f == "shift")
    {
        if (value != 0.0 && value != 1.0) return fail(h, RMPC_EINVAL, "rmpc_set_option: shift must be 0 or 1");
        h->W.shift = (int) value;
        return 0;
    }
    if (f == "graph")
    {
        if (value != 0.0 && value != 1.0) return fail(h, RMPC_EINVAL, "rmpc_set_option: graph must be 0 or 1");
        h->use_graph = (int) value;
        return 0;
    }
    if (f == "statistics")
    {
        if (value != 0.0 && value != 1.0) return fail(h, RMPC_EINVAL, "rmpc_set_option: statistics must be 0 or 1");
        if (value == 1.0 && !h->sqpstat_dev)
        {
            ON_DEVICE(h);
            const size_t bytes = sizeof(double) * (size_t) h->D.B * (size_t) (h->D.max_iter + 2) * RQ_COUNT;
            void *p = nullptr;
            if (cudaMalloc(&p, bytes) != cudaSuccess) { cudaGetLastError(); return fail(h, RMPC_ENOMEM, "rmpc_set_option: device allocation failed"); }
            if (cudaMemset(p, 0, bytes) != cudaSuccess || cudaDeviceSynchronize() != cudaSuccess) return fail(h, RMPC_ECUDA, "rmpc_set_option: cudaMemset failed");
            h->owned.push_back(p);
            h->bytes += (long long) bytes;
            h->sqpstat_dev = (double *) p;
        }
        h->W.sqpstat = value == 1.0 ? h->sqpstat_dev : nullptr;
        return 0;
    }
    if (f == "as_rti_level")
    { /* acados' as_rti_level: 0 = A, 1 = B, 2 = C, 3 = D, 4 (or -1) = standard RTI */
        if (h->D.solver_type != RMPC_SOLVER_SQP_RTI) return fail(h, RMPC_EINVAL, "rmpc_set_option: as_rti_level needs an SQP_RTI handle");
        const int lv = (int) value;
        if ((double) lv != value || lv < -1 || lv > 4) return fail(h, RMPC_EINVAL, "rmpc_set_option: as_rti_level must be in [0, 1, 2, 3, 4]");
        const int mode = (lv < 0 || lv == 4) ? 0 : lv + 1;
        if (mode == 1 && !h->asrti_dev)
        { /* level A keeps a copy of the iterate (acados: work->tmp_nlp_out) */
            ON_DEVICE(h);
            const size_t bytes = sizeof(double) * (size_t) h->D.B * (size_t) RMPC_ASRTI_STRIDE(h->W.Sp);
            void *p = nullptr;
            if (cudaMalloc(&p, bytes) != cudaSuccess) { cudaGetLastError(); return fail(h, RMPC_ENOMEM, "rmpc_set_option: device allocation failed"); }
            if (cudaMemset(p, 0, bytes) != cudaSuccess || cudaDeviceSynchronize() != cudaSuccess) return fail(h, RMPC_ECUDA, "rmpc_set_option: cudaMemset failed");
            h->owned.push_back(p);
            h->bytes += (long long) bytes;
            h->asrti_dev = (double *) p;
        }
        h->W.as_rti = mode;
        h->W.asrti = mode == 1 ? h->asrti_dev : nullptr;
        if (h->W.as_rti_iter < 1) h->W.as_rti_iter = 1;
        return 0;
    }
    if (f == "qp_cond_N")
    { /* qp_solver_cond_N (T/acados_ocp_options.py:72-84): horizon of the partially condensed QP; N (or 0) = no condensing */
        const int n2 = (int) value;
        if ((double) n2 != value || n2 < 0 || n2 > h->D.N) return fail(h, RMPC_EINVAL, "rmpc_set_option: qp_cond_N must be an integer in [1, N]");
        h->W.cond_N = (n2 == h->D.N) ? 0 : n2;
        return 0;
    }
    if (f == "as_rti_iter")
    {
        const int it = (int) value;
        if ((double) it != value || it < 0) return fail(h, RMPC_EINVAL, "rmpc_set_option: as_rti_iter must be a nonnegative integer");
        h->W.as_rti_iter = it;
        return 0;
    }
    return fail(h, RMPC_EINVAL, "rmpc_set_option: unknown option " + f);
}

extern "C" int rmpc_set_chain(rmpc_handle *h, const double *xyzR, const int *axis)
{
    if (!h) return fail(h, RMPC_EINVAL, "rmpc_set_chain: null handle");
    if (!xyzR || !axis) { rmpc_chain_init(&h->ch); return 0; } /* back to the built-in UR10 */
    for (int k = 0; k < 8 * 12; k++)
        if (!std::isfinite(xyzR[k])) return fail(h, RMPC_EINVAL, "rmpc_set_chain: non-finite entry in the chain table");
    RmpcChain ch;
    if (rmpc_chain_from_table(&ch, xyzR, axis) != 0) return fail(h, RMPC_EINVAL, "rmpc_set_chain: joint axes must be 0 (x), 1 (y) or 2 (z)");
    h->ch = ch;
    return 0;
}

static int strided(rmpc_handle *h, double *packed, double *strided_base, long stride, long off, int n, int gather, cudaStream_t s)
{
    rmpc_strided_copy_kernel<<<grid_for((long) h->D.B * n), 256, 0, s>>>(packed, strided_base, h->D.B, stride, off, n, gather);
    h->launches++;
    CU(cudaGetLastError());
    return 0;
}

/* dst[i * dstride + doff + c] = src[i * sstride + soff + c], c < n */
__global__ void rmpc_strided2_kernel(const double *src, long sstride, long soff, double *dst, long dstride, long doff, long B, int n)
{
    const long tot = B * n;
    for (long k = (long) blockIdx.x * blockDim.x + threadIdx.x; k < tot; k += (long) gridDim.x * blockDim.x)
    {
        const long i = k / n, c = k % n;
        dst[i * dstride + doff + c] = src[i * sstride + soff + c];
    }
}
static int strided2(rmpc_handle *h, const double *src, long sstride, long soff, double *dst, long dstride, long doff, int n, cudaStream_t s)
{
    rmpc_strided2_kernel<<<grid_for((long) h->D.B * n), 256, 0, s>>>(src, sstride, soff, dst, dstride, doff, h->D.B, n);
    h->launches++;
    CU(cudaGetLastError());
    return 0;
}
static int stage_table_fill(rmpc_handle *h, int c0, int nc, cudaStream_t s)
{
    rmpc_stage_table_fill_kernel<<<grid_for((long) h->D.B * h->D.N * RT_COUNT), 256, 0, s>>>(h->stg_dev, h->par_dev, h->D.B, h->D.N, c0, nc);
    h->launches++;
    CU(cudaGetLastError());
    return 0;
}

extern "C" int rmpc_set(rmpc_handle *h, const char *field, int stage, const double *dev, void *stream)
{
    if (!h || !field || !dev) return fail(h, RMPC_EINVAL, "rmpc_set: null argument");
    ON_DEVICE(h);
    cudaStream_t s = (cudaStream_t) stream;
    const long S = h->D.N + 1;
    const std::string f(field);
    if (f == "par")
    {
        CU(cudaMemcpyAsync(h->par_dev, dev, sizeof(double) * h->D.B * RP_COUNT, cudaMemcpyDeviceToDevice, s));
        if (h->stg_dev) return stage_table_fill(h, 0, RT_COUNT, s); /* a new parameter block resets every per-stage override */
        return 0;
    }
    /* ---- named fields of the cost / constraint modules (AcadosOcpSolver.cost_set / constraints_set,
     * T/acados_ocp_solver.py:1863-2035 -> ocp_nlp_cost_model_set / ocp_nlp_constraints_model_set,
     * acados_c/ocp_nlp_interface.c:516-560,637-760).  stage = -1: every stage (a slice of the parameter block);
     * stage >= 0: that stage only (the per-stage table, created on first use from the parameter block). */
    if (f == "yref" || f == "lbu" || f == "ubu" || f == "lbx" || f == "ubx")
    {
        const bool is_y = f == "yref", is_u = f == "lbu" || f == "ubu", lower = f == "lbu" || f == "lbx";
        if (!is_y && !is_u && stage == 0)
        { /* lbx_0 = ubx_0 = x0: the initial-state constraint, [B][12] (the reference sets both to the plant state) */
            CU(cudaMemcpyAsync(h->W.z, dev, sizeof(double) * h->D.B * NX, cudaMemcpyDeviceToDevice, s));
            return 0;
        }
        if (stage < -1 || stage >= h->D.N)
            return fail(h, RMPC_EINVAL, "rmpc_set: " + f + " exists for stages 0..N-1 (stage N has no cost and no box bounds), -1 = every stage");
        const int n = is_y ? 18 : 6;
        if (stage < 0)
        {
            const long off = is_y ? RP_YREF : (is_u ? (lower ? RP_UMIN : RP_UMAX) : (lower ? RP_QMIN : RP_QMAX));
            int rc = strided(h, (double *) dev, h->par_dev, RP_COUNT, off, n, 0, s);
            if (rc) return rc;
            /* only this field's columns: per-stage overrides of the other fields stay (lbu next to lbx in the lower group) */
            if (h->stg_dev) return stage_table_fill(h, is_y ? RT_YREF : ((lower ? RT_LB : RT_UB) + (is_u ? 0 : 6)), n, s);
            return 0;
        }
        if (!h->stg_dev)
        {
            const size_t bytes = sizeof(double) * (size_t) h->D.B * (size_t) h->D.N * RT_COUNT;
            void *p = nullptr;
            if (cudaMalloc(&p, bytes) != cudaSuccess) { cudaGetLastError(); return fail(h, RMPC_ENOMEM, "rmpc_set: device allocation failed"); }
            h->owned.push_back(p);
            h->bytes += (long long) bytes;
            h->stg_dev = (double *) p;
            int rc = stage_table_fill(h, 0, RT_COUNT, s);
            if (rc) return rc;
            h->W.stg = h->stg_dev;
        }
        const long off = (long) stage * RT_COUNT + (is_y ? RT_YREF : ((lower ? RT_LB : RT_UB) + (is_u ? 0 : 6)));
        return strided(h, (double *) dev, h->stg_dev, (long) h->D.N * RT_COUNT, off, n, 0, s);
    }
    if (f == "W_diag" || f == "surf_coeffs" || f == "wcv")
    {
        if (stage != -1) return fail(h, RMPC_EINVAL, "rmpc_set: " + f + " is per instance, not per stage: pass stage = -1");
        if (f == "surf_coeffs") return strided(h, (double *) dev, h->par_dev, RP_COUNT, RP_SURF, 6, 0, s);
        /* both need host arithmetic on the values (and libm's exp for the exact-ZOH coefficients, like the reference's own
         * Python): device -> host, transform, host -> device; synchronises the stream */
        const size_t B = (size_t) h->D.B;
        const int n = f == "wcv" ? 6 : 18;
        std::vector<double> in(B * n), out;
        CU(cudaMemcpyAsync(in.data(), dev, sizeof(double) * B * n, cudaMemcpyDeviceToHost, s));
        CU(cudaStreamSynchronize(s));
        if (f == "wcv")
        { /* R/prediction_model.py:87-115: a22 = exp(-w Ts), a12 = (1 - a22) / w, b1 = Ts - a12, b2 = 1 - a22 */
            out.resize(B * 30);
            const double Ts = h->D.dt;
            for (size_t i = 0; i < B; i++)
                for (int k = 0; k < 6; k++)
                {
                    const double w = in[i * 6 + k], a22 = std::exp(-w * Ts), a12 = (1.0 - a22) / w;
                    double *o = out.data() + i * 30;
                    o[k] = a12; o[6 + k] = a22; o[12 + k] = Ts - a12; o[18 + k] = 1.0 - a22; o[24 + k] = w;
                }
            double *tmp = nullptr;
            if (cudaMalloc((void **) &tmp, sizeof(double) * B * 30) != cudaSuccess) { cudaGetLastError(); return fail(h, RMPC_ENOMEM, "rmpc_set: device allocation failed"); }
            cudaError_t e = cudaMemcpyAsync(tmp, out.data(), sizeof(double) * B * 30, cudaMemcpyHostToDevice, s);
            int rc = e == cudaSuccess ? strided(h, tmp, h->par_dev, RP_COUNT, RP_A12, 30, 0, s) : RMPC_ECUDA;
            cudaStreamSynchronize(s);
            cudaFree(tmp);
            if (e != cudaSuccess) return fail(h, RMPC_ECUDA, std::string("rmpc_set: ") + cudaGetErrorString(e));
            return rc;
        }
        /* W_diag [B][18]: entries 0..4 task weights, 5..10 = 2 w_u per joint, 11..16 = w_qddot per joint, 17 ignored (the
         * manipulability output has weight 0, R/trajectory_optimizer.py:139-152) */
        out.resize(B * 17);
        for (size_t i = 0; i < B; i++)
        {
            const double *w = in.data() + i * 18;
            double *o = out.data() + i * 17;
            for (int k = 0; k < 6; k++) { o[k] = 0.5 * w[5 + k]; o[6 + k] = w[11 + k]; }
            for (int k = 0; k < 5; k++) o[12 + k] = w[k];
        }
        double *tmp = nullptr;
        if (cudaMalloc((void **) &tmp, sizeof(double) * B * 17) != cudaSuccess) { cudaGetLastError(); return fail(h, RMPC_ENOMEM, "rmpc_set: device allocation failed"); }
        cudaError_t e = cudaMemcpyAsync(tmp, out.data(), sizeof(double) * B * 17, cudaMemcpyHostToDevice, s);
        int rc = e == cudaSuccess ? strided2(h, tmp, 17, 0, h->par_dev, RP_COUNT, RP_WU, 12, s) : RMPC_ECUDA;
        if (!rc) rc = strided2(h, tmp, 17, 12, h->par_dev, RP_COUNT, RP_WTASK, 5, s);
        cudaStreamSynchronize(s);
        cudaFree(tmp);
        if (e != cudaSuccess) return fail(h, RMPC_ECUDA, std::string("rmpc_set: ") + cudaGetErrorString(e));
        return rc;
    }
    if (f == "z") { CU(cudaMemcpyAsync(h->W.z, dev, sizeof(double) * h->D.B * NX, cudaMemcpyDeviceToDevice, s)); return 0; }
    if (f == "P_e" || f == "alpha_e" || f == "lh_e" || f == "term")
    {
        /* terminal ellipsoid row of stage N; the first call switches the row on (all-zero block until everything is set) */
        if (!h->pe_dev)
        {
            const size_t bytes = sizeof(double) * (size_t) h->D.B * RE_COUNT;
            void *p = nullptr;
            if (cudaMalloc(&p, bytes) != cudaSuccess) { cudaGetLastError(); return fail(h, RMPC_ENOMEM, "rmpc_set: device allocation failed"); }
            CU(cudaMemsetAsync(p, 0, bytes, s));
            h->owned.push_back(p);
            h->bytes += (long long) bytes;
            h->pe_dev = (double *) p;
            h->W.pe = h->pe_dev;
        }
        if (f == "term") { CU(cudaMemcpyAsync(h->pe_dev, dev, sizeof(double) * h->D.B * RE_COUNT, cudaMemcpyDeviceToDevice, s)); return 0; }
        if (f == "P_e") return strided(h, (double *) dev, h->pe_dev, RE_COUNT, RE_P, 144, 0, s);
        if (f == "alpha_e") return strided(h, (double *) dev, h->pe_dev, RE_COUNT, RE_UH, 1, 0, s);
        return strided(h, (double *) dev, h->pe_dev, RE_COUNT, RE_LH, 1, 0, s);
    }
    if (f == "x" || f == "u" || f == "pi")
    {
        const int smax = (f == "x") ? h->D.N : h->D.N - 1;
        if (stage < 0 || stage > smax) return fail(h, RMPC_EINVAL, "rmpc_set: stage out of range for field " + f);
        const long st = h->W.slab_stride;
        if (f == "x") return strided(h, (double *) dev, h->W.slab, st, rmpc_slab_off(h->D.N, RC_ux, NV, stage) + 6, NX, 0, s);
        if (f == "u") return strided(h, (double *) dev, h->W.slab, st, rmpc_slab_off(h->D.N, RC_ux, NV, stage), NU, 0, s);
        return strided(h, (double *) dev, h->W.slab, st, rmpc_slab_off(h->D.N, RC_pi, NX, stage), NX, 0, s);
    }
    return fail(h, RMPC_EINVAL, "rmpc_set: unknown field '" + f + "'");
}

extern "C" int rmpc_reset(rmpc_handle *h, void *stream)
{
    if (!h) return fail(h, RMPC_EINVAL, "rmpc_reset: null handle");
    ON_DEVICE(h);
    cudaStream_t s = (cudaStream_t) stream;
    const size_t B = (size_t) h->D.B, S = (size_t) h->D.N + 1;
    (void) S;
    /* duals, merit weights, QP memory, summaries: back to the calloc'ed state of a new acados solver */
    CU(cudaMemsetAsync(h->W.slab, 0, sizeof(double) * B * (size_t) h->W.slab_stride, s));
    CU(cudaMemsetAsync(h->W.summary, 0, sizeof(double) * B * RS_COUNT, s));
    if (h->W.slog) CU(cudaMemsetAsync(h->W.slog, 0, sizeof(double) * B * h->nlog * RL_COUNT, s));
    if (h->W.ulog) CU(cudaMemsetAsync(h->W.ulog, 0, sizeof(double) * B * h->nlog * 6, s));
    if (h->W.zlog) CU(cudaMemsetAsync(h->W.zlog, 0, sizeof(double) * B * (h->nlog + 1) * NX, s));
    rmpc_init_iterate_kernel<<<grid_for((long) B * S * NV), 256, 0, s>>>(h->D, h->W);
    h->launches++;
    CU(cudaGetLastError());
    h->steps_done = 0;
    h->W.step0 = 0;
    return 0;
}

/* ---- checkpoint / resume of the complete solver + plant state of the batch (AcadosOcpSolver.store_iterate / load_iterate,
 * T/acados_ocp_solver.py:1158-1440, extended to everything a closed loop carries: iterate, multipliers, merit weights, QP
 * memory, plant state, flags, accumulated summaries, step counter).  Blob = [magic, B, N, steps_done | slab | z | summary | flags]. */
static const double RMPC_STATE_MAGIC = 20261104.0;
#define RMPC_STATE_HDR 8
/* header = [magic, B, N, steps_done, solver_type, terminal row on/off, shift + AS-RTI + cond_N options, checksum of the chain table]: a blob
 * only loads into a handle whose solver semantics match (flags[] and the merit weights mean different things for SQP and
 * SQP_RTI; the iterate of another arm or another shift setting continues a different closed loop).  NOT in the blob: the
 * parameters ("par", terminal set, chain) -- set them again on the new handle -- and the per-step logs, whose rows below
 * the resume point stay zero. */
static double chain_checksum(const RmpcChain &ch)
{
    double s = 0.0;
    const double *p = &ch.Rfix[0][0];
    for (size_t k = 0; k < sizeof(RmpcChain) / sizeof(double); k++) s += (double) (k + 1) * p[k];
    return s;
}
static void state_header(const rmpc_handle *h, double *hdr)
{
    hdr[0] = RMPC_STATE_MAGIC; hdr[1] = h->D.B; hdr[2] = h->D.N; hdr[3] = h->steps_done;
    hdr[4] = h->D.solver_type; hdr[5] = h->pe_dev ? 1.0 : 0.0; hdr[6] = h->W.shift + 10.0 * h->W.as_rti + 1e3 * (h->W.as_rti ? h->W.as_rti_iter : 0) + 1e6 * h->W.cond_N; hdr[7] = chain_checksum(h->ch);
}
extern "C" long long rmpc_state_size(const rmpc_handle *h)
{
    if (!h) return 0;
    const long long B = h->D.B;
    return RMPC_STATE_HDR + B * h->W.slab_stride + B * NX + B * RS_COUNT + B * 2;
}

extern "C" int rmpc_store_state(rmpc_handle *h, double *dev, void *stream)
{
    if (!h || !dev) return fail(h, RMPC_EINVAL, "rmpc_store_state: null argument");
    if (h->W.as_rti == 1) /* the level-A backup of the iterate (acados: work->tmp_nlp_out) is not part of the blob */
        return fail(h, RMPC_EINVAL, "rmpc_store_state: not supported with AS-RTI level A");
    ON_DEVICE(h);
    cudaStream_t s = (cudaStream_t) stream;
    const size_t B = (size_t) h->D.B;
    double hdr[RMPC_STATE_HDR];
    state_header(h, hdr);
    CU(cudaMemcpyAsync(dev, hdr, sizeof(hdr), cudaMemcpyHostToDevice, s)); /* pageable source: staged before the call returns */
    double *p = dev + RMPC_STATE_HDR;
    CU(cudaMemcpyAsync(p, h->W.slab, sizeof(double) * B * (size_t) h->W.slab_stride, cudaMemcpyDeviceToDevice, s));
    p += B * (size_t) h->W.slab_stride;
    CU(cudaMemcpyAsync(p, h->W.z, sizeof(double) * B * NX, cudaMemcpyDeviceToDevice, s));
    p += B * NX;
    CU(cudaMemcpyAsync(p, h->W.summary, sizeof(double) * B * RS_COUNT, cudaMemcpyDeviceToDevice, s));
    p += B * RS_COUNT;
    CU(cudaMemcpyAsync(p, h->W.flags, sizeof(int) * B * 4, cudaMemcpyDeviceToDevice, s));
    return 0;
}

extern "C" int rmpc_load_state(rmpc_handle *h, const double *dev, void *stream)
{
    if (!h || !dev) return fail(h, RMPC_EINVAL, "rmpc_load_state: null argument");
    ON_DEVICE(h);
    cudaStream_t s = (cudaStream_t) stream;
    const size_t B = (size_t) h->D.B;
    double hdr[RMPC_STATE_HDR] = {0}, want[RMPC_STATE_HDR];
    CU(cudaMemcpyAsync(hdr, dev, sizeof(hdr), cudaMemcpyDeviceToHost, s));
    CU(cudaStreamSynchronize(s));
    state_header(h, want);
    if (hdr[0] != RMPC_STATE_MAGIC || hdr[1] != want[1] || hdr[2] != want[2] || hdr[3] < 0)
        return fail(h, RMPC_EINVAL, "rmpc_load_state: the blob was not stored by a handle with this B and N");
    if (hdr[4] != want[4]) return fail(h, RMPC_EINVAL, "rmpc_load_state: the blob comes from a handle with another solver type");
    if (hdr[5] != want[5]) return fail(h, RMPC_EINVAL, "rmpc_load_state: terminal row on in one handle and off in the other (set it before loading)");
    if (hdr[6] != want[6]) return fail(h, RMPC_EINVAL, "rmpc_load_state: the blob comes from a handle with another shift / AS-RTI / partial-condensing option");
    if (hdr[7] != want[7]) return fail(h, RMPC_EINVAL, "rmpc_load_state: the blob comes from a handle with another kinematic chain");
    const double *p = dev + RMPC_STATE_HDR;
    CU(cudaMemcpyAsync(h->W.slab, p, sizeof(double) * B * (size_t) h->W.slab_stride, cudaMemcpyDeviceToDevice, s));
    p += B * (size_t) h->W.slab_stride;
    CU(cudaMemcpyAsync(h->W.z, p, sizeof(double) * B * NX, cudaMemcpyDeviceToDevice, s));
    p += B * NX;
    CU(cudaMemcpyAsync(h->W.summary, p, sizeof(double) * B * RS_COUNT, cudaMemcpyDeviceToDevice, s));
    p += B * RS_COUNT;
    CU(cudaMemcpyAsync(h->W.flags, p, sizeof(int) * B * 4, cudaMemcpyDeviceToDevice, s));
    h->steps_done = (int) hdr[3];
    h->W.step0 = h->steps_done;
    return 0;
}

static int launch_loop(rmpc_handle *h, int nsteps, int advance, cudaStream_t s)
{
    const int grid = (h->D.B + RMPC_GROUPS_PER_CTA - 1) / RMPC_GROUPS_PER_CTA;
    h->W.step0 = h->steps_done;
    rmpc_closed_loop_kernel<<<grid, 32 * RMPC_WARPS_PER_CTA, sizeof(GrpSmem) * RMPC_GROUPS_PER_CTA, s>>>(h->ch, h->D, h->W, nsteps, advance);
    h->launches++;
    CU(cudaGetLastError());
    return 0;
}

extern "C" int rmpc_solve(rmpc_handle *h, void *stream)
{
    if (!h) return fail(h, RMPC_EINVAL, "rmpc_solve: null handle");
    if (h->W.as_rti) /* AC/ocp_nlp/ocp_nlp_sqp_rti.c:1150-1154 */
        return fail(h, RMPC_EINVAL, "rmpc_solve: rti_phase PREPARATION_AND_FEEDBACK is not supported with AS-RTI (as_rti_level != 4)");
    ON_DEVICE(h);
    h->W.rti_phase = 0;
    return launch_loop(h, 1, 0, (cudaStream_t) stream);
}

extern "C" int rmpc_solve_rti_phase(rmpc_handle *h, int phase, void *stream)
{
    if (!h) return fail(h, RMPC_EINVAL, "rmpc_solve_rti_phase: null handle");
    if (h->D.solver_type != RMPC_SOLVER_SQP_RTI) return fail(h, RMPC_EINVAL, "rmpc_solve_rti_phase: rti_phase is only supported for SQP_RTI");
    if (phase < 0 || phase > 2) return fail(h, RMPC_EINVAL, "rmpc_solve_rti_phase: phase must be 0, 1 or 2");
    if (phase == 0 && h->W.as_rti)
        return fail(h, RMPC_EINVAL, "rmpc_solve_rti_phase: rti_phase PREPARATION_AND_FEEDBACK is not supported with AS-RTI (as_rti_level != 4)");
    ON_DEVICE(h);
    h->W.rti_phase = phase;
    int rc = launch_loop(h, 1, 0, (cudaStream_t) stream);
    h->W.rti_phase = 0;
    return rc;
}

extern "C" int rmpc_closed_loop(rmpc_handle *h, int nsteps, void *stream)
{
    if (!h || nsteps < 0) return fail(h, RMPC_EINVAL, "rmpc_closed_loop: bad argument");
    ON_DEVICE(h);
    if (nsteps == 0) return 0;
    /* One launch per closed-loop step by default: a launch boundary puts every instance of the batch back in the same
     * phase (linearise / factorise / substitute ...).  Measured on B200 (4096 x N=100): 447k steps/s at 1 step per
     * launch, 367k at 5, 288k at 600 (r1c); 722k / 704k / 660k at 1 / 2 / 3 steps per launch (r1k) -- instances that drift into different phases compete for the memory system and the
     * one-stage-ahead staging of the Riccati sweeps no longer hides the latency.  Launch overhead is ~0.05% of a step.
     * RMPC_STEPS_PER_LAUNCH overrides (experiments). */
    static const int chunk_env = []() { const char *e = getenv("RMPC_STEPS_PER_LAUNCH"); int v = e ? atoi(e) : 1; return v > 0 ? v : 1; }();
    for (int done = 0; done < nsteps;)
    {
        const int n = (nsteps - done) < chunk_env ? (nsteps - done) : chunk_env;
        int rc = launch_loop(h, n, 1, (cudaStream_t) stream);
        if (rc) return rc;
        h->steps_done += n;
        done += n;
    }
    return 0;
}

extern "C" int rmpc_get(rmpc_handle *h, const char *field, int stage, double *dev, void *stream)
{
    if (!h || !field || !dev) return fail(h, RMPC_EINVAL, "rmpc_get: null argument");
    ON_DEVICE(h);
    cudaStream_t s = (cudaStream_t) stream;
    const long B = h->D.B, S = h->D.N + 1;
    const std::string f(field);
    if (f == "x" || f == "u" || f == "pi" || f == "lam")
    {
        /* "lam" of stage N exists with a terminal row only: entry 0 = lower side, entry 12 = upper side */
        const int smax = (f == "x" || (f == "lam" && h->pe_dev)) ? h->D.N : h->D.N - 1;
        if (stage < 0 || stage > smax) return fail(h, RMPC_EINVAL, "rmpc_get: stage out of range for field " + f);
        const long st = h->W.slab_stride;
        if (f == "x") return strided(h, dev, h->W.slab, st, rmpc_slab_off(h->D.N, RC_ux, NV, stage) + 6, NX, 1, s);
        if (f == "u") return strided(h, dev, h->W.slab, st, rmpc_slab_off(h->D.N, RC_ux, NV, stage), NU, 1, s);
        if (f == "pi") return strided(h, dev, h->W.slab, st, rmpc_slab_off(h->D.N, RC_pi, NX, stage), NX, 1, s);
        return strided(h, dev, h->W.slab, st, rmpc_slab_off(h->D.N, RC_lam, NB, stage), NB, 1, s);
    }
    if (f == "lamx0") return strided(h, dev, h->W.slab, h->W.slab_stride, H_LAMX0, NB, 1, s);
    if (f == "par") { CU(cudaMemcpyAsync(dev, h->par_dev, sizeof(double) * B * RP_COUNT, cudaMemcpyDeviceToDevice, s)); return 0; }
    if (f == "sqp_stat")
    {
        if (!h->sqpstat_dev) return fail(h, RMPC_EINVAL, "rmpc_get: per-iteration statistics are off (rmpc_set_option(h, \"statistics\", 1))");
        CU(cudaMemcpyAsync(dev, h->sqpstat_dev, sizeof(double) * B * (h->D.max_iter + 2) * RQ_COUNT, cudaMemcpyDeviceToDevice, s));
        return 0;
    }
    if (f == "yref" || f == "lbu" || f == "ubu" || f == "lbx" || f == "ubx")
    { /* the value a stage uses: its row of the per-stage table if there is one, else the parameter block */
        const bool is_y = f == "yref", is_u = f == "lbu" || f == "ubu", lower = f == "lbu" || f == "lbx";
        if (stage < 0 || stage >= h->D.N || (!is_y && !is_u && stage == 0))
            return fail(h, RMPC_EINVAL, "rmpc_get: stage out of range for field " + f);
        const int n = is_y ? 18 : 6;
        if (h->stg_dev)
            return strided(h, dev, h->stg_dev, (long) h->D.N * RT_COUNT,
                           (long) stage * RT_COUNT + (is_y ? RT_YREF : ((lower ? RT_LB : RT_UB) + (is_u ? 0 : 6))), n, 1, s);
        return strided(h, dev, h->par_dev, RP_COUNT, is_y ? RP_YREF : (is_u ? (lower ? RP_UMIN : RP_UMAX) : (lower ? RP_QMIN : RP_QMAX)), n, 1, s);
    }
    if (f == "z") { CU(cudaMemcpyAsync(dev, h->W.z, sizeof(double) * B * NX, cudaMemcpyDeviceToDevice, s)); return 0; }
    if (f == "summary") { CU(cudaMemcpyAsync(dev, h->W.summary, sizeof(double) * B * RS_COUNT, cudaMemcpyDeviceToDevice, s)); return 0; }
    if (f == "flags")
    {
        rmpc_flags_to_double_kernel<<<grid_for(B * 4), 256, 0, s>>>(h->W.flags, dev, B * 4);
        h->launches++;
        CU(cudaGetLastError());
        return 0;
    }
    if (f == "stats" || f == "zlog" || f == "ulog" || f == "slog" || f == "plog")
    {
        if (h->nlog <= 0) return fail(h, RMPC_EINVAL, "rmpc_get: handle was created with nlog = 0, no logs");
        if (f == "plog")
        {
            rmpc_plant_logs_kernel<<<grid_for(B * (h->nlog + 1)), 256, 0, s>>>(h->ch, h->par_dev, h->W.zlog, dev, B, h->nlog + 1);
            h->launches++;
            CU(cudaGetLastError());
            return 0;
        }
        if (f == "zlog") { CU(cudaMemcpyAsync(dev, h->W.zlog, sizeof(double) * B * (h->nlog + 1) * NX, cudaMemcpyDeviceToDevice, s)); return 0; }
        if (f == "ulog") { CU(cudaMemcpyAsync(dev, h->W.ulog, sizeof(double) * B * h->nlog * 6, cudaMemcpyDeviceToDevice, s)); return 0; }
        if (f == "slog") { CU(cudaMemcpyAsync(dev, h->W.slog, sizeof(double) * B * h->nlog * RL_COUNT, cudaMemcpyDeviceToDevice, s)); return 0; }
        /* stats: record of the most recent step (rmpc_solve writes at index steps_done, closed loop at steps_done-1) */
        int idx = stage >= 0 ? stage : 0;
        if (idx >= h->nlog) return fail(h, RMPC_EINVAL, "rmpc_get: stats index beyond nlog");
        return strided(h, dev, h->W.slog, (long) h->nlog * RL_COUNT, (long) idx * RL_COUNT, RL_COUNT, 1, s);
    }
    return fail(h, RMPC_EINVAL, "rmpc_get: unknown field '" + f + "'");
}

extern "C" int rmpc_run_host(rmpc_handle *h, const double *par_host, const double *z0_host, int nsteps, double *summary_host,
                             double *u_last_host, void *stream)
{
    if (!h || !par_host || !z0_host || nsteps < 0) return fail(h, RMPC_EINVAL, "rmpc_run_host: bad argument");
    ON_DEVICE(h);
    cudaStream_t s = (cudaStream_t) stream;
    const size_t B = (size_t) h->D.B;
    CU(cudaMemcpyAsync(h->par_dev, par_host, sizeof(double) * B * RP_COUNT, cudaMemcpyHostToDevice, s));
    if (h->stg_dev)
    { /* like rmpc_set "par": a new parameter block discards every per-stage override */
        int rc_fill = stage_table_fill(h, 0, RT_COUNT, s);
        if (rc_fill) return rc_fill;
    }
    CU(cudaMemcpyAsync(h->W.z, z0_host, sizeof(double) * B * NX, cudaMemcpyHostToDevice, s));
    int rc = rmpc_reset(h, stream);
    if (rc) return rc;
    rc = rmpc_closed_loop(h, nsteps, stream);
    if (rc) return rc;
    if (summary_host) CU(cudaMemcpyAsync(summary_host, h->W.summary, sizeof(double) * B * RS_COUNT, cudaMemcpyDeviceToHost, s));
    if (u_last_host)
    {
        /* u0 of the last solve = control iterate of stage 0 */
        CU(cudaMemcpy2DAsync(u_last_host, sizeof(double) * 6, h->W.slab + rmpc_slab_off(h->D.N, RC_ux, NV, 0), sizeof(double) * h->W.slab_stride,
                             sizeof(double) * 6, B, cudaMemcpyDeviceToHost, s));
    }
    CU(cudaStreamSynchronize(s));
    return 0;
}

/* H2D x0 -> one solve -> D2H u0 (+ step record) on stream s, from / to the handle's pinned staging buffers */
static int step_host_enqueue(rmpc_handle *h, cudaStream_t s)
{
    const size_t B = (size_t) h->D.B;
    CU(cudaMemcpyAsync(h->W.z, h->pin_z, sizeof(double) * B * NX, cudaMemcpyHostToDevice, s));
    int rc = launch_loop(h, 1, 0, s);
    if (rc) return rc;
    CU(cudaMemcpy2DAsync(h->pin_u, sizeof(double) * 6, h->W.slab + rmpc_slab_off(h->D.N, RC_ux, NV, 0), sizeof(double) * h->W.slab_stride,
                         sizeof(double) * 6, B, cudaMemcpyDeviceToHost, s));
    if (h->nlog > 0 && h->steps_done < h->nlog)
        CU(cudaMemcpy2DAsync(h->pin_st, sizeof(double) * RL_COUNT, h->W.slog + (size_t) h->steps_done * RL_COUNT,
                             sizeof(double) * h->nlog * RL_COUNT, sizeof(double) * RL_COUNT, B, cudaMemcpyDeviceToHost, s));
    return 0;
}

static void step_graph_drop(rmpc_handle *h)
{
    if (h->gexec) cudaGraphExecDestroy(h->gexec);
    h->gexec = nullptr;
}

extern "C" int rmpc_step_host(rmpc_handle *h, const double *x0_host, double *u0_host, double *stats_host, void *stream)
{
    if (!h || !x0_host || !u0_host) return fail(h, RMPC_EINVAL, "rmpc_step_host: null argument");
    if (stats_host && !(h->nlog > 0 && h->steps_done < h->nlog))
        return fail(h, RMPC_EINVAL, "rmpc_step_host: the step record needs a handle created with nlog > 0");
    if (h->W.as_rti) /* one combined solve per call, which acados refuses with a level set (AC/ocp_nlp/ocp_nlp_sqp_rti.c:1150-1154) */
        return fail(h, RMPC_EINVAL, "rmpc_step_host: rti_phase PREPARATION_AND_FEEDBACK is not supported with AS-RTI (as_rti_level != 4); "
                                    "drive the two phases with rmpc_solve_rti_phase");
    ON_DEVICE(h);
    const size_t B = (size_t) h->D.B;
    if (!h->pin_z)
    {
        CU(cudaHostAlloc((void **) &h->pin_z, sizeof(double) * B * NX, cudaHostAllocDefault));
        CU(cudaHostAlloc((void **) &h->pin_u, sizeof(double) * B * 6, cudaHostAllocDefault));
        CU(cudaHostAlloc((void **) &h->pin_st, sizeof(double) * B * RL_COUNT, cudaHostAllocDefault));
        CU(cudaStreamCreate(&h->own_stream));
    }
    cudaStream_t s = stream ? (cudaStream_t) stream : h->own_stream;
    std::memcpy(h->pin_z, x0_host, sizeof(double) * B * NX);
    h->W.rti_phase = 0;
    h->W.step0 = h->steps_done;
    if (h->use_graph)
    {
        /* the captured sequence is valid as long as the launch arguments are the ones it was captured with */
        if (h->gexec && (h->gstream != s || std::memcmp(&h->gW, &h->W, sizeof(RmpcWs)) != 0 || std::memcmp(&h->gD, &h->D, sizeof(RmpcDims)) != 0 ||
                         std::memcmp(&h->gch, &h->ch, sizeof(RmpcChain)) != 0)) /* the chain is a by-value kernel argument too */
            step_graph_drop(h);
        if (!h->gexec)
        {
            cudaGraph_t g = nullptr;
            CU(cudaStreamBeginCapture(s, cudaStreamCaptureModeThreadLocal));
            int rc = step_host_enqueue(h, s);
            cudaError_t e = cudaStreamEndCapture(s, &g);
            if (rc) { if (g) cudaGraphDestroy(g); return rc; }
            if (e != cudaSuccess) return fail(h, RMPC_ECUDA, std::string("rmpc_step_host: cudaStreamEndCapture: ") + cudaGetErrorString(e));
            e = cudaGraphInstantiate(&h->gexec, g, 0);
            cudaGraphDestroy(g);
            if (e != cudaSuccess) { h->gexec = nullptr; return fail(h, RMPC_ECUDA, std::string("rmpc_step_host: cudaGraphInstantiate: ") + cudaGetErrorString(e)); }
            std::memcpy(&h->gW, &h->W, sizeof(RmpcWs));
            std::memcpy(&h->gD, &h->D, sizeof(RmpcDims));
            std::memcpy(&h->gch, &h->ch, sizeof(RmpcChain));
            h->gstream = s;
        }
        else
            h->launches++; /* the capture counted its own launch */
        CU(cudaGraphLaunch(h->gexec, s));
    }
    else
    {
        int rc = step_host_enqueue(h, s);
        if (rc) return rc;
    }
    CU(cudaStreamSynchronize(s));
    std::memcpy(u0_host, h->pin_u, sizeof(double) * B * 6);
    if (stats_host) std::memcpy(stats_host, h->pin_st, sizeof(double) * B * RL_COUNT);
    return 0;
}

/* ------------------------------------------------------------------------------------------------ rmpc_solve_static
 * One cold-started solve for callers that must not allocate device memory (include/rmpc.h has the contract).  The solver's
 * work set -- the per-instance slabs and the flag words, which rmpc_create would cudaMalloc -- lives in the file-scope array
 * below: it is part of the loaded module, so every GPU that loads librmpc.so carries RMPC_STATIC_SCRATCH_BYTES for it until
 * the process exits, used or not.
 * In HOST code the name rmpc_static_scratch is only a stand-in for the symbol; it is not a device address, and a kernel
 * handed &rmpc_static_scratch[0] from the host would touch the wrong memory.  The device address is looked up once per
 * device with cudaGetSymbolAddress (with that device current) and that pointer is what the kernels receive. */
static __device__ __align__(256) unsigned char rmpc_static_scratch[RMPC_STATIC_SCRATCH_BYTES];

/* rmpc_closed_loop_kernel restricted to one solve without a plant step, with the stage-0 control written straight to the
 * caller's output (the handle path gathers it from the slab with a second launch) */
__global__ void RMPC_KERNEL_BOUNDS
rmpc_solve_static_kernel(const __grid_constant__ RmpcChain ch, const __grid_constant__ RmpcDims D, const __grid_constant__ RmpcWs W, double *u0)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    GrpSmem *smem = reinterpret_cast<GrpSmem *>(smem_raw);
    const int grp = threadIdx.x / RMPC_G;
    const long inst = (long) blockIdx.x * RMPC_GROUPS_PER_CTA + grp;
    if (inst >= D.B) return;
    const long first = inst - (grp % (32 / RMPC_G));
    unsigned pair = 0;
    for (int g = 0; g < 32 / RMPC_G; g++)
        if (first + g < D.B) pair |= (RMPC_G == 32 ? 0xffffffffu : ((1u << (RMPC_G & 31)) - 1u)) << (g * RMPC_G);
    closed_loop_group(ch, D, W, inst, smem[grp], 1, 0, pair); /* ends with a group barrier: the iterate is visible to all lanes */
    const int lane = threadIdx.x % RMPC_G;
    if (lane < NU) u0[inst * NU + lane] = W.slab[inst * W.slab_stride + rmpc_slab_off(D.N, RC_ux, NV, 0) + lane];
}

#define RMPC_STATIC_MAX_DEVICES 64
static std::mutex g_static_mutex;
static unsigned char *g_static_base[RMPC_STATIC_MAX_DEVICES]; /* device address of rmpc_static_scratch on each device, null: not set up */

extern "C" int rmpc_solve_static(const rmpc_dims *dims, int device, const double *par_dev, const double *x0_dev, double *u0_dev,
                                 double *stats_dev, void *stream)
{
    rmpc_handle *h = nullptr; /* no handle: errors go to rmpc_last_error(NULL) */
    if (!dims || !par_dev || !x0_dev || !u0_dev) return fail(h, RMPC_EINVAL, "rmpc_solve_static: null argument");
    if (dims->B < 1 || dims->N < 2 || dims->N > 4096) return fail(h, RMPC_EINVAL, "rmpc_solve_static: need B >= 1, 2 <= N <= 4096");
    if (dims->solver_type != RMPC_SOLVER_SQP && dims->solver_type != RMPC_SOLVER_SQP_RTI)
        return fail(h, RMPC_EINVAL, "rmpc_solve_static: solver_type must be RMPC_SOLVER_SQP or RMPC_SOLVER_SQP_RTI");
    /* scratch = [B slabs | B x 4 flag words]; refuse what the fixed array cannot hold */
    const size_t B = (size_t) dims->B, stride = (size_t) rmpc_slab_stride(dims->N);
    const size_t per_inst = stride * sizeof(double) + 4 * sizeof(int);
    if (B > (size_t) RMPC_STATIC_SCRATCH_BYTES / per_inst)
        return fail(h, RMPC_EINVAL, "rmpc_solve_static: B = " + std::to_string(B) + " instances at N = " + std::to_string(dims->N) + " need " +
                                        std::to_string(B * per_inst) + " bytes of scratch, the static array holds " +
                                        std::to_string((size_t) RMPC_STATIC_SCRATCH_BYTES) + " (at most " +
                                        std::to_string((size_t) RMPC_STATIC_SCRATCH_BYTES / per_inst) + " instances at this N); use rmpc_create for larger batches");
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev < 1 || device < 0 || device >= ndev || device >= RMPC_STATIC_MAX_DEVICES)
        return fail(h, RMPC_ENODEV, "rmpc_solve_static: no usable CUDA device (this library has no host fallback)");
    DeviceGuard dev_guard_(device);
    if (!dev_guard_.ok) return fail(h, RMPC_ENODEV, "rmpc_solve_static: cudaSetDevice failed");
    unsigned char *base = nullptr;
    {
        /* set-up, once per device and with that device current: resolve the array's device address and opt the kernel in
         * to its shared-memory size.  This is the part of the first call that loads the module. */
        std::lock_guard<std::mutex> lock(g_static_mutex);
        if (!g_static_base[device])
        {
            void *p = nullptr;
            cudaError_t e = cudaGetSymbolAddress(&p, rmpc_static_scratch);
            if (e != cudaSuccess || !p)
                return fail(h, RMPC_ECUDA, std::string("rmpc_solve_static: cudaGetSymbolAddress(rmpc_static_scratch): ") + cudaGetErrorString(e));
            e = cudaFuncSetAttribute(rmpc_solve_static_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) (sizeof(GrpSmem) * RMPC_GROUPS_PER_CTA));
            if (e != cudaSuccess)
                return fail(h, RMPC_ECUDA, std::string("rmpc_solve_static: cannot reserve the shared memory the kernel needs: ") + cudaGetErrorString(e));
            g_static_base[device] = (unsigned char *) p;
        }
        base = g_static_base[device];
    }
    cudaStream_t s = (cudaStream_t) stream;
    RmpcDims D;
    memcpy(&D, dims, sizeof(RmpcDims));
    RmpcChain ch;
    rmpc_chain_init(&ch);
    RmpcWs W;
    memset(&W, 0, sizeof(W));
    W.par = par_dev;
    W.slab = (double *) base;
    W.slab_stride = (long) stride;
    W.Sp = rmpc_slab_sp(dims->N);
    W.z = const_cast<double *>(x0_dev); /* read only here: without a plant step no kernel of this path writes z */
    W.flags = (int *) (base + B * stride * sizeof(double));
    W.slog = stats_dev; /* log capacity of one step: the record of this solve lands in the caller's buffer */
    W.nlog = stats_dev ? 1 : 0;
    W.as_rti_iter = 1;
    /* the array still holds the previous call's iterate, multipliers, merit weights and QP memory: a new solver starts
     * from zeros (rmpc_reset), then x_k = x0, u_k = 0 and the first-call flags */
    CU(cudaMemsetAsync(base, 0, B * stride * sizeof(double), s));
    rmpc_init_iterate_kernel<<<grid_for((long) B * (dims->N + 1) * NV), 256, 0, s>>>(D, W);
    CU(cudaGetLastError());
    const int grid = (dims->B + RMPC_GROUPS_PER_CTA - 1) / RMPC_GROUPS_PER_CTA;
    rmpc_solve_static_kernel<<<grid, 32 * RMPC_WARPS_PER_CTA, sizeof(GrpSmem) * RMPC_GROUPS_PER_CTA, s>>>(ch, D, W, u0_dev);
    CU(cudaGetLastError());
    return 0;
}

#ifdef RMPC_SECPROF
/* experiments only (-DRMPC_SECPROF builds): cycles per code section of the sweeps, see SECPROF_* in rmpc_device.cuh */
extern "C" int rmpc_debug_secprof(double *out64, int clear)
{
    unsigned long long v[64];
    if (cudaMemcpyFromSymbol(v, rmpc_secprof, sizeof(v)) != cudaSuccess) return RMPC_ECUDA;
    for (int k = 0; k < 64; k++) out64[k] = (double) v[k];
    if (clear)
    {
        memset(v, 0, sizeof(v));
        if (cudaMemcpyToSymbol(rmpc_secprof, v, sizeof(v)) != cudaSuccess) return RMPC_ECUDA;
    }
    return 0;
}
#endif
