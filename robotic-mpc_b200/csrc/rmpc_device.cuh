/* robotic-mpc_b200/csrc/rmpc_device.cuh -- device code of the batched UR10 surface-tracking MPC engine (sm_100a).
 *
 * A GROUP of GL = RMPC_G lanes (16 by default: two problem instances per warp) owns ONE problem instance for the
 * whole closed loop (R/simulator.py:199-241):
 *     set x0 -> SQP_RTI / SQP solve -> u0 -> plant RK4 -> metrics,   nsteps times, no host round trip.
 * Inside a solve the group alternates between two parallelisation patterns:
 *   (S) stage-parallel  : lane l handles stages l, l+GL, ... (forward kinematics / Jacobian / Gauss-Newton data,
 *                         KKT and QP residuals, merit function, iterate updates) with shuffle reductions;
 *   (R) stage-sequential: the Riccati sweeps of the interior-point method.  Lane l < 6 owns control row u_l and state
 *                         row qd_l, lane 6 + j owns state row q_j of the stage block; the 12 x 12 cost-to-go matrix P
 *                         of the next stage lives in shared memory, the 6 x 6 control block is factorised redundantly
 *                         by every lane (no communication), and everything a later sweep needs is streamed to HBM in
 *                         one 208-double factor block per stage (RF_* in rmpc_types.h).
 * All shuffles / barriers carry the group's own lane mask, so the two instances of a warp may sit in different
 * interior-point or SQP iterations (independent thread scheduling); they run in lock step while they agree.
 *
 * Replaces, for a batch: CasADi-generated model functions, acados ocp_nlp (SQP / SQP_RTI, NLS cost, discrete
 * dynamics, BGH bounds, merit backtracking), acados ocp_qp partial condensing wrapper (x0 elimination) and the
 * HPIPM interior-point solver -- file:line citations are on each function.  R/ = reference root,
 * AC/ = acados/acados, H/ = acados/external/hpipm.
 *
 * Numerical route of the Riccati recursion: HPIPM factorises the whole 19 x 18 stage block and carries the Cholesky
 * factor Lxx of P (square-root form, H/ocp_qp/x_ocp_qp_kkt.c:836-1007).  Here only the 6 control columns are
 * eliminated and P = Mxx - Lxu Lxu' is carried explicitly -- the same Schur complement the reference's Cholesky
 * forms before factorising it -- and A = [[I, diag],[0, diag]], B = [diag; diag] are applied as the sparse row
 * combinations they are.  Same iterates up to rounding; parity with the reference is asserted at 1e-6 relative on
 * u / x trajectories with equal iteration counts (tests/).
 *
 * The file compiles in two ways: with nvcc for the product, and with g++ -DRMPC_EMU against
 * tests/simt_emu/simt_emu.h (a lock-step lane emulator) for CPU-side unit tests of this very source.
 */
#pragma once
#include "rmpc_types.h"

#ifndef RMPC_G
#define RMPC_G 16
#endif
#define GL RMPC_G

#ifdef RMPC_EMU
#include "simt_emu.h"
#define RD static inline
#define RDN static
#define RDM inline
RD int rd_lane() { return simt_emu::lane_id(); }
RD void rd_sync() { simt_emu::syncwarp(); }
RD void rd_sync_mask(unsigned) { simt_emu::syncwarp(); }
RD unsigned rd_mask() { return 0; }
RD double rd_shfl(double v, int src) { return simt_emu::shfl(v, src & (GL - 1)); }
RD double rd_shfl_xor(double v, int m) { return simt_emu::shfl(v, (simt_emu::lane_id() ^ m) & (GL - 1)); }
RD double rd_rsqrt(double x) { return 1.0 / sqrt(x); }
RD void rd_sincos(double a, double *s, double *c) { *s = sin(a); *c = cos(a); }
RD double rd_fmax(double a, double b) { return a > b ? a : b; } /* NaN-ignoring like fmax for our uses */
RD double rd_inf() { return INFINITY; }
RD double rd_time() { return 0.0; }
RD void rd_prefetch_l2(const void *) {}
RD void rd_cp_async16(double *dst_smem, const double *src) { dst_smem[0] = src[0]; dst_smem[1] = src[1]; }
RD void rd_cp_commit() {}
RD void rd_cp_wait1() {}
RD void rd_cp_wait0() {}
RD void rd_bar_init(unsigned long long *) {}
RD void rd_bar_inval(unsigned long long *) {}
RD void rd_bar_expect(unsigned long long *, unsigned) {}
RD void rd_bulk_cp(double *dst_smem, const double *src, unsigned bytes, unsigned long long *) { memcpy(dst_smem, src, bytes); }
RD void rd_bar_wait(unsigned long long *, unsigned) {}
RD void rd_fence_bar_init() {}
#define RD_UNROLL
struct rd_double2 { double x, y; };
#else
#define RD __device__ __forceinline__
/* the large phases are real functions: one copy each (the SQP and SQP_RTI drivers, the predictor and corrector solves
 * share them), which keeps the kernel's code inside the instruction caches */
#define RDN __device__ __noinline__
#define RDM __device__ __forceinline__
RD unsigned rd_mask()
{
    return GL == 32 ? 0xffffffffu : (((1u << (GL & 31)) - 1u) << ((threadIdx.x & 31) & ~(GL - 1)));
}
RD int rd_lane() { return threadIdx.x & (GL - 1); }
RD void rd_sync() { __syncwarp(rd_mask()); }
RD void rd_sync_mask(unsigned m) { __syncwarp(m); }
RD double rd_shfl(double v, int src) { return __shfl_sync(rd_mask(), v, src, GL); }
RD double rd_shfl_xor(double v, int m) { return __shfl_xor_sync(rd_mask(), v, m, GL); }
RD double rd_rsqrt(double x) { return rsqrt(x); }
RD void rd_sincos(double a, double *s, double *c) { sincos(a, s, c); }
RD double rd_fmax(double a, double b) { return a > b ? a : b; }
RD double rd_inf() { return __longlong_as_double(0x7ff0000000000000LL); }
RD double rd_time() /* seconds on the GPU's global nanosecond timer */
{
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return 1e-9 * (double) t;
}
RD void rd_prefetch_l2(const void *p) { asm volatile("prefetch.global.L2 [%0];" ::"l"(p)); }
/* asynchronous 16-byte global -> shared copy (LDGSTS, bypasses L1 and the register file); completion is tracked per
 * thread in commit groups; data of OTHER lanes is visible after wait + rd_sync() */
RD void rd_cp_async16(double *dst_smem, const double *src)
{
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((unsigned) __cvta_generic_to_shared(dst_smem)), "l"(src) : "memory");
}
/* TMA 1-D bulk copy global -> shared (UBLKCP): ONE instruction per contiguous chunk, issued by one lane of the group; the copy
 * engine generates the addresses and writes shared memory itself -- nothing goes through the LSU address / data pipes that the
 * sweeps' own shared-memory loads saturate.  Completion is counted in bytes on an mbarrier in shared memory.
 * src, dst 16-byte aligned, bytes a multiple of 16. */
RD unsigned rd_smem_u32(const void *p) { return (unsigned) __cvta_generic_to_shared(p); }
RD void rd_bar_init(unsigned long long *bar) { asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(rd_smem_u32(bar)) : "memory"); }
RD void rd_bar_inval(unsigned long long *bar) { asm volatile("mbarrier.inval.shared::cta.b64 [%0];" ::"r"(rd_smem_u32(bar)) : "memory"); }
RD void rd_fence_bar_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
RD void rd_bar_expect(unsigned long long *bar, unsigned bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(rd_smem_u32(bar)), "r"(bytes) : "memory");
}
RD void rd_bulk_cp(double *dst_smem, const double *src, unsigned bytes, unsigned long long *bar)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(rd_smem_u32(dst_smem)), "l"(src),
                 "r"(bytes), "r"(rd_smem_u32(bar))
                 : "memory");
}
RD void rd_bar_wait(unsigned long long *bar, unsigned parity)
{
    asm volatile("{\n\t.reg .pred p;\n\tRD_WAIT_%=:\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t@p bra RD_DONE_%=;\n\tbra RD_WAIT_%=;\n\tRD_DONE_%=:\n\t}" ::"r"(
                     rd_smem_u32(bar)),
                 "r"(parity)
                 : "memory");
}
RD void rd_cp_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
RD void rd_cp_wait1() { asm volatile("cp.async.wait_group 1;" ::: "memory"); }
RD void rd_cp_wait0() { asm volatile("cp.async.wait_group 0;" ::: "memory"); }
#define RD_UNROLL _Pragma("unroll")
typedef double2 rd_double2;
#endif

#define NX RMPC_NX
#define NU RMPC_NU
#define NV RMPC_NV
#define NB RMPC_NB
#define NJ RMPC_NJ
#define FS RMPC_FS

RD double grp_sum(double v)
{
    for (int m = GL / 2; m; m >>= 1) v += rd_shfl_xor(v, m);
    return v;
}
/* NaN-propagating, so that every lane of the group ends with the same value whatever the data (a comparison-only butterfly
 * leaves lanes disagreeing once a NaN is involved, and the callers branch on the result assuming it is uniform) */
RD double grp_max(double v)
{
    for (int m = GL / 2; m; m >>= 1) { double o = rd_shfl_xor(v, m); v = (o > v || o != o) ? o : v; }
    return v;
}

/* Section profiler of the sweeps (experiments only, -DRMPC_SECPROF): lane 0 of every group accumulates SM clock cycles per
 * code section of a stage and adds them to rmpc_secprof[base + k] at the end of the sweep; read with rmpc_debug_secprof(). */
#if defined(RMPC_SECPROF) && !defined(RMPC_EMU)
__device__ unsigned long long rmpc_secprof[64];
#define SECPROF_BEGIN long long sp_a[10] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0}, sp_t = clock64()
#define SEC(k) do { const long long t_ = clock64(); sp_a[k] += t_ - sp_t; sp_t = t_; } while (0)
#define SECPROF_END(base) do { if (rd_lane() == 0) for (int k_ = 0; k_ < 10; k_++) atomicAdd(&rmpc_secprof[(base) + k_], (unsigned long long) sp_a[k_]); } while (0)
#else
#define SECPROF_BEGIN do { } while (0)
#define SEC(k) do { } while (0)
#define SECPROF_END(base) do { } while (0)
#endif

/* ---------------------------------------------------------------------------------------------- shared memory
 * one block per group (instance).  `buf` is carved per sweep:
 *   factorisation : P+ (12 x SM_PS) | Lxu (72) | Muu in (28) | (36 free) | two staging buffers of FS_STGN doubles
 *   forward / backward substitution : two staging buffers of SW_STGN doubles
 * The staging buffers are filled one stage AHEAD by cp.async (LDGSTS), so that inside a stage every operand comes from
 * shared memory and the HBM / L2 latency of the streamed factors and residuals is off the critical path. */
#define SM_PS 14 /* row stride of P: even (16-byte aligned pairs), 28 words -> rows 0..5 hit distinct banks */
#define SM_BUF 752
struct GrpSmem
{
    double cst[24];       /* a12(6) a22(6) b1(6) b2(6) of the exact-ZOH model */
    double hc[18];        /* hu(6) hs(6) hq(6): constant diagonal blocks of the Gauss-Newton Hessian (StageConst) */
    double xv[12];        /* running state of the forward sweep */
    double uv[8];
    double tv[12];
    double pv[12];        /* cost-to-go vector p of stage ss+1 */
    double dp[12];        /* multiplier step of the current stage (forward sweep, residual check) */
    double wv[6];         /* Jc dx of the current stage (forward sweep, residual check) */
    unsigned long long bar[2]; /* mbarriers of the two staging buffers (bulk-copy staging) */
    double buf[SM_BUF];
};
#define C_A12(c) sm.cst[(c)]
#define C_A22(c) sm.cst[6 + (c)]
#define C_B1(c) sm.cst[12 + (c)]
#define C_B2(c) sm.cst[18 + (c)]
#define FS_P(sm) ((sm).buf)         /* cost-to-go matrix of stage ss+1, full symmetric, state order [q; qd] */
#define FS_LXU(sm) ((sm).buf + 168) /* Lxu of the current stage (for the Schur complement) */
#define FS_MIN(sm) ((sm).buf + 240) /* lower 6x6 block Muu (21) + m_u (6) */
#define FS_STGN 164
#define FS_STG(sm, k) ((sm).buf + 304 + (k) * FS_STGN)
#define SW_STGN 376
#define SW_STG(sm, k) ((sm).buf + (k) * SW_STGN)

/* group-wide asynchronous copy of n doubles (n even, both sides 16-byte aligned) global -> shared */
RD void stage_cp(double *dst, const double *src, const int n)
{
    /* n is a compile-time constant at every call site: fully unrolled, the chunk offsets become immediates */
    const int l2 = 2 * rd_lane();
    RD_UNROLL
    for (int c = 0; c < (n / 2 + GL - 1) / GL; c++)
        if ((c + 1) * GL <= n / 2 || rd_lane() + c * GL < n / 2) rd_cp_async16(dst + l2 + 2 * GL * c, src + l2 + 2 * GL * c);
}

/* Flat group-wide vector updates over whole arrays (n even, both sides 16-byte aligned): 128-bit accesses, and the loads of
 * four strides -- eight doubles per operand and lane -- are issued before the first use.  These passes are pure streaming;
 * what they cost is round trips to memory, so each trip carries as much as the registers allow. */
#ifdef RMPC_EMU
#define RD_RESTRICT
#else
#define RD_RESTRICT __restrict__
#endif
#define RD_V2(p) reinterpret_cast<rd_double2 *>(p)
#define RD_CV2(p) reinterpret_cast<const rd_double2 *>(p)
template <class F>
RD void grp_map2(double *RD_RESTRICT y, const double *RD_RESTRICT x, int k0, int n, F f) /* y[k] = f(x[k], y[k]) on [k0, n) */
{
    rd_double2 *y2 = RD_V2(y + k0);
    const rd_double2 *x2 = RD_CV2(x + k0);
    const int n2 = (n - k0) / 2;
    int k = rd_lane();
    for (; k + 3 * GL < n2; k += 4 * GL)
    {
        const rd_double2 x0 = x2[k], x1 = x2[k + GL], xx2 = x2[k + 2 * GL], x3 = x2[k + 3 * GL];
        rd_double2 y0 = y2[k], y1 = y2[k + GL], yy2 = y2[k + 2 * GL], y3 = y2[k + 3 * GL];
        y0.x = f(x0.x, y0.x); y0.y = f(x0.y, y0.y); y1.x = f(x1.x, y1.x); y1.y = f(x1.y, y1.y);
        yy2.x = f(xx2.x, yy2.x); yy2.y = f(xx2.y, yy2.y); y3.x = f(x3.x, y3.x); y3.y = f(x3.y, y3.y);
        y2[k] = y0; y2[k + GL] = y1; y2[k + 2 * GL] = yy2; y2[k + 3 * GL] = y3;
    }
    for (; k < n2; k += GL)
    {
        const rd_double2 x0 = x2[k];
        rd_double2 y0 = y2[k];
        y0.x = f(x0.x, y0.x); y0.y = f(x0.y, y0.y);
        y2[k] = y0;
    }
}
RD void grp_axpy(double *RD_RESTRICT y, const double *RD_RESTRICT x, double a, int n) /* y += a x */
{
    grp_map2(y, x, 0, n, [a](double xv, double yv) { return yv + a * xv; });
}
RD void grp_xpay_to(double *RD_RESTRICT z, const double *RD_RESTRICT x, const double *RD_RESTRICT y, double a, int n) /* z = x + a y */
{
    rd_double2 *z2 = RD_V2(z);
    const rd_double2 *x2 = RD_CV2(x), *y2 = RD_CV2(y);
    const int n2 = n / 2;
    int k = rd_lane();
    for (; k + 3 * GL < n2; k += 4 * GL)
    {
        const rd_double2 x0 = x2[k], x1 = x2[k + GL], xx2 = x2[k + 2 * GL], x3 = x2[k + 3 * GL];
        const rd_double2 y0 = y2[k], y1 = y2[k + GL], yy2 = y2[k + 2 * GL], y3 = y2[k + 3 * GL];
        rd_double2 r0, r1, r2, r3;
        r0.x = x0.x + a * y0.x; r0.y = x0.y + a * y0.y; r1.x = x1.x + a * y1.x; r1.y = x1.y + a * y1.y;
        r2.x = xx2.x + a * yy2.x; r2.y = xx2.y + a * yy2.y; r3.x = x3.x + a * y3.x; r3.y = x3.y + a * y3.y;
        z2[k] = r0; z2[k + GL] = r1; z2[k + 2 * GL] = r2; z2[k + 3 * GL] = r3;
    }
    for (; k < n2; k += GL)
    {
        const rd_double2 x0 = x2[k], y0 = y2[k];
        rd_double2 r0;
        r0.x = x0.x + a * y0.x; r0.y = x0.y + a * y0.y;
        z2[k] = r0;
    }
}
RD void grp_lerp(double *RD_RESTRICT y, const double *RD_RESTRICT x, double al, int k0, int n) /* y = (1-al) y + al x on [k0, n) */
{
    grp_map2(y, x, k0, n, [al](double xv, double yv) { return (1.0 - al) * yv + al * xv; });
}
/* Leineweber merit-weight update w = t0 (first SQP iteration) or max(t0, (t0 + w) / 2) with t0 = t or |t|, on [k0, n) */
RD void grp_weights(double *RD_RESTRICT w, const double *RD_RESTRICT t, int first, int use_abs, int k0, int n)
{
    grp_map2(w, t, k0, n, [first, use_abs](double tv, double wv) {
        const double t0 = use_abs ? fabs(tv) : tv;
        if (first) return t0;
        const double t1 = 0.5 * (t0 + wv);
        return t0 > t1 ? t0 : t1;
    });
}

/* per-instance view (layout: rmpc_types.h): parameter row, slab head, first array, stage slots, plant state, flags */
struct Inst
{
    const double *par;
    const double *pe; /* terminal-set block (RE_*), null without a terminal constraint */
    const double *stg; /* [N][RT_COUNT] per-stage references and bounds, null: every stage uses the parameter row */
    double *head, *arr0, *z;
    int *flags;
    int Sp;
    int N2; /* horizon of the partially condensed QP (qp_solver_cond_N), 0: no condensing */
    unsigned pair; /* lane mask of all groups of this warp that run an instance (re-convergence points) */
};
/* Re-convergence point for the groups that share a warp.  They run independent instances and may drift apart inside
 * the interior-point loop (different iteration counts); once apart, the warp issues every instruction twice at half
 * width.  At places that every group passes exactly once per step they wait for each other and run in lock step again. */
#define PAIR_SYNC() rd_sync_mask(I.pair)
#define AR(name) (I.arr0 + (long) RC_##name * I.Sp)

RD Inst inst_view(const RmpcWs &w, int N, long i)
{
    Inst v;
    v.head = w.slab + i * w.slab_stride;
    v.arr0 = v.head + RMPC_HEAD;
    v.Sp = w.Sp;
    v.N2 = (w.cond_N > 0 && w.cond_N < N) ? w.cond_N : 0;
    v.par = w.par + i * RP_COUNT;
    v.pe = w.pe ? w.pe + i * RE_COUNT : nullptr;
    v.stg = w.stg ? w.stg + i * (long) N * RT_COUNT : nullptr;
    v.z = w.z + i * NX;
    v.flags = w.flags + i * 4;
    v.pair = rd_mask();
    return v;
}

/* element (r, c) of a packed-lower symmetric 12 x 12 matrix */
RD double psym(const double *P, int r, int c) { return r >= c ? P[(r * (r + 1)) / 2 + c] : P[(c * (c + 1)) / 2 + r]; }

/* Terminal ellipsoid row h_e(x) = x' P x of stage N, group-wide: lane k < 12 forms row k of P x; returns h_e on every
 * lane and, with `grad`, writes the gradient 2 P x (the stage-N row of DCt).  Replaces the CasADi-generated
 * *_constr_h_e_fun[_jac_uxt_zt] pair that acados' BGH module would call (AC/ocp_nlp/ocp_nlp_constraints_bgh.c:
 * 1219-1445,1563-1619) for con_h_expr_e = x' P x with P from R/_terminal_constraint.py:96-116. */
#define TERM_LO(N) (NB * (N))      /* slot of the lower side lh - h_e in the NB-wide per-stage arrays */
#define TERM_UP(N) (NB * (N) + 12) /* slot of the upper side h_e - uh */
RD double term_eval(const double *pe, const double *x, double *grad)
{
    const int l = rd_lane();
    double px = 0.0, hx = 0.0;
    if (l < NX)
    {
        RD_UNROLL
        for (int k = 0; k < NX; k++) px += pe[RE_P + NX * l + k] * x[k];
        hx = x[l] * px;
        if (grad) grad[l] = 2.0 * px;
    }
    return grp_sum(hx);
}

/* ---------------------------------------------------------------------------------------------- model (K1)
 * UR10 forward kinematics, WORLD-frame geometric Jacobian, task-frame outputs and their derivatives.
 * Replaces the CasADi-generated *_cost_y_fun[_jac_ut_xt] (T/casadi_function_generation.py:555-575) whose
 * expression graph comes from R/prediction_model.py:117-173,226-326 (pinocchio FK of frame ee_link, WORLD
 * Jacobian, task transform, v_task), R/surface.py:21,243-262 (S, unit normal) and
 * R/trajectory_optimizer.py:88-137 (g1..g5).  Chain constants: R/ur_description/urdf/ur10.urdf:50-222.
 *   y5 = [S(p_t) - p_tz, n . R_t[:,2], R_t[0,1], p_tx, (R_t' (v + w x t_w))_y]
 *   J  = d y5 / d x  (5 x 12; rows 0..3 depend on q only, so only 5 x 6 + 6 entries are non-zero)
 * The remaining 13 outputs (u, qddot_k, manipulability with weight 0) are affine/irrelevant and handled by the
 * callers in closed form. */
RD void cross3(const double *a, const double *b, double *c)
{
    c[0] = a[1] * b[2] - a[2] * b[1];
    c[1] = a[2] * b[0] - a[0] * b[2];
    c[2] = a[0] * b[1] - a[1] * b[0];
}
RD double dot3(const double *a, const double *b) { return a[0] * b[0] + a[1] * b[1] + a[2] * b[2]; }

struct NoEmit
{
    RDM void begin(const double *) {}
    RDM void operator()(int, const double *, double) const {}
};

/* emit.begin(y5) is called once the outputs are known; emit(k, Jq[5], Jd) then receives, for joint k, column k of
 * d y5 / d q and the entry d y5_4 / d qd_k */
template <bool JAC, class Emit>
RD void ur10_task(const RmpcChain &ch, const double *par, const double *q, const double *qd, double *y5, Emit &emit)
{
    double R[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1}, p[3] = {0, 0, 0};
    double pj[6][3], aj[6][3];
    RD_UNROLL
    for (int j = 0; j < 6; j++)
    {
        const double *F = ch.Rfix[j], *t = ch.pfix[j];
        double Rj[9];
        RD_UNROLL
        for (int r = 0; r < 3; r++)
        {
            p[r] += R[3 * r] * t[0] + R[3 * r + 1] * t[1] + R[3 * r + 2] * t[2];
            RD_UNROLL
            for (int c = 0; c < 3; c++) Rj[3 * r + c] = R[3 * r] * F[c] + R[3 * r + 1] * F[3 + c] + R[3 * r + 2] * F[6 + c];
        }
        const int ax = (j == 0 || j == 4) ? 2 : 1; /* z y y y z y */
        double s, c;
        rd_sincos(q[j], &s, &c);
        RD_UNROLL
        for (int r = 0; r < 3; r++)
        {
            pj[j][r] = p[r];
            aj[j][r] = Rj[3 * r + ax];
            if (ax == 2)
            { /* Rj * Rz(q) */
                R[3 * r] = c * Rj[3 * r] + s * Rj[3 * r + 1];
                R[3 * r + 1] = c * Rj[3 * r + 1] - s * Rj[3 * r];
                R[3 * r + 2] = Rj[3 * r + 2];
            }
            else
            { /* Rj * Ry(q) */
                R[3 * r] = c * Rj[3 * r] - s * Rj[3 * r + 2];
                R[3 * r + 1] = Rj[3 * r + 1];
                R[3 * r + 2] = s * Rj[3 * r] + c * Rj[3 * r + 2];
            }
        }
    }
    /* frame ee_link, task frame: R_t = R_ee, p_t = p_ee + R_ee t_ee (R/prediction_model.py:256-281) */
    const double *F = ch.Rfix[6], *t = ch.pfix[6], *te = par + RP_TEE;
    double r1[3], r2[3], tw[3], pt[3];
    RD_UNROLL
    for (int r = 0; r < 3; r++)
    {
        double e0 = R[3 * r] * F[0] + R[3 * r + 1] * F[3] + R[3 * r + 2] * F[6];
        r1[r] = R[3 * r] * F[1] + R[3 * r + 1] * F[4] + R[3 * r + 2] * F[7];
        r2[r] = R[3 * r] * F[2] + R[3 * r + 1] * F[5] + R[3 * r + 2] * F[8];
        tw[r] = e0 * te[0] + r1[r] * te[1] + r2[r] * te[2];
        pt[r] = p[r] + R[3 * r] * t[0] + R[3 * r + 1] * t[1] + R[3 * r + 2] * t[2] + tw[r];
    }
    /* wv = v + w x t_w with [v; w] = J_world qd  =  sum_j qd_j a_j x (t_w - p_j) */
    double wv[3] = {0, 0, 0};
    RD_UNROLL
    for (int j = 0; j < 6; j++)
    {
        double e[3] = {tw[0] - pj[j][0], tw[1] - pj[j][1], tw[2] - pj[j][2]}, c3[3];
        cross3(aj[j], e, c3);
        RD_UNROLL
        for (int r = 0; r < 3; r++) wv[r] += qd[j] * c3[r];
    }
    const double sa = par[RP_SURF], sb = par[RP_SURF + 1], sc = par[RP_SURF + 2], sd = par[RP_SURF + 3], se = par[RP_SURF + 4],
                 sf = par[RP_SURF + 5];
    const double X = pt[0], Y = pt[1];
    const double S = sa * X * X + sb * Y * Y + sc * X * Y + sd * X + se * Y + sf;
    const double Sx = 2 * sa * X + sc * Y + sd, Sy = 2 * sb * Y + sc * X + se;
    const double m[3] = {Sx, Sy, -1.0};
    const double irho = rd_rsqrt(Sx * Sx + Sy * Sy + 1.0);
    const double n[3] = {m[0] * irho, m[1] * irho, m[2] * irho};
    y5[0] = S - pt[2];
    y5[1] = dot3(n, r2);
    y5[2] = r1[0];
    y5[3] = pt[0];
    y5[4] = dot3(r1, wv);
    if (!JAC) return;
    emit.begin(y5);
    /* dn/dX, dn/dY */
    const double mX[3] = {2 * sa, sc, 0.0}, mY[3] = {sc, 2 * sb, 0.0};
    double nX[3], nY[3];
    {
        const double mmX = dot3(m, mX), mmY = dot3(m, mY), i3 = irho * irho * irho;
        RD_UNROLL
        for (int r = 0; r < 3; r++)
        {
            nX[r] = mX[r] * irho - m[r] * mmX * i3;
            nY[r] = mY[r] * irho - m[r] * mmY * i3;
        }
    }
    RD_UNROLL
    for (int k = 0; k < 6; k++)
    {
        /* d p_t / d q_k = a_k x (p_t - p_k);  d r / d q_k = a_k x r for EE-fixed directions */
        double dd[3] = {pt[0] - pj[k][0], pt[1] - pj[k][1], pt[2] - pj[k][2]};
        double dpt[3], dr1[3], dr2[3], dtw[3];
        cross3(aj[k], dd, dpt);
        cross3(aj[k], r1, dr1);
        cross3(aj[k], r2, dr2);
        cross3(aj[k], tw, dtw);
        double Jq[5];
        Jq[0] = Sx * dpt[0] + Sy * dpt[1] - dpt[2];
        double dn[3] = {nX[0] * dpt[0] + nY[0] * dpt[1], nX[1] * dpt[0] + nY[1] * dpt[1], nX[2] * dpt[0] + nY[2] * dpt[1]};
        Jq[1] = dot3(dn, r2) + dot3(n, dr2);
        Jq[2] = dr1[0];
        Jq[3] = dpt[0];
        double dwv[3] = {0, 0, 0};
        RD_UNROLL
        for (int j = 0; j < 6; j++)
        {
            double e[3] = {tw[0] - pj[j][0], tw[1] - pj[j][1], tw[2] - pj[j][2]};
            double de[3] = {dtw[0], dtw[1], dtw[2]};
            double t1[3] = {0, 0, 0}, t2[3];
            if (k < j)
            { /* joint k precedes joint j: its rotation moves a_j and p_j */
                double da[3], pk[3] = {pj[j][0] - pj[k][0], pj[j][1] - pj[k][1], pj[j][2] - pj[k][2]}, dpj[3];
                cross3(aj[k], aj[j], da);
                cross3(aj[k], pk, dpj);
                RD_UNROLL
                for (int r = 0; r < 3; r++) de[r] -= dpj[r];
                cross3(da, e, t1);
            }
            cross3(aj[j], de, t2);
            RD_UNROLL
            for (int r = 0; r < 3; r++) dwv[r] += qd[j] * (t1[r] + t2[r]);
        }
        Jq[4] = dot3(dr1, wv) + dot3(r1, dwv);
        double e[3] = {tw[0] - pj[k][0], tw[1] - pj[k][1], tw[2] - pj[k][2]}, ce[3];
        cross3(aj[k], e, ce);
        emit(k, Jq, dot3(r1, ce));
    }
}

/* plant logs: pose of frame ee_link and WORLD twist J qd (R/simulation_model.py:60-77), then the post-hoc tracking
 * errors e1..e5 of R/simulator.py:301-342 -- NB e5 deliberately uses the reference's own formula
 * R_w_t (v + (w . t_w)), not the OCP's g5.  REC = 1 also writes the per-sample record RK_* (rmpc_types.h). */
template <bool REC>
RD void ur10_plant_record(const RmpcChain &ch, const double *par, const double *q, const double *qd, double *e, double *rec)
{
    double R[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1}, p[3] = {0, 0, 0};
    double tws[6] = {0, 0, 0, 0, 0, 0};
    RD_UNROLL
    for (int j = 0; j < 6; j++)
    {
        const double *F = ch.Rfix[j], *t = ch.pfix[j];
        double Rj[9];
        RD_UNROLL
        for (int r = 0; r < 3; r++)
        {
            p[r] += R[3 * r] * t[0] + R[3 * r + 1] * t[1] + R[3 * r + 2] * t[2];
            RD_UNROLL
            for (int c = 0; c < 3; c++) Rj[3 * r + c] = R[3 * r] * F[c] + R[3 * r + 1] * F[3 + c] + R[3 * r + 2] * F[6 + c];
        }
        const int ax = (j == 0 || j == 4) ? 2 : 1;
        double a[3] = {Rj[ax], Rj[3 + ax], Rj[6 + ax]}, c3[3];
        cross3(p, a, c3);
        RD_UNROLL
        for (int r = 0; r < 3; r++) { tws[r] += c3[r] * qd[j]; tws[3 + r] += a[r] * qd[j]; }
        double s, c;
        rd_sincos(q[j], &s, &c);
        RD_UNROLL
        for (int r = 0; r < 3; r++)
        {
            if (ax == 2)
            {
                R[3 * r] = c * Rj[3 * r] + s * Rj[3 * r + 1];
                R[3 * r + 1] = c * Rj[3 * r + 1] - s * Rj[3 * r];
                R[3 * r + 2] = Rj[3 * r + 2];
            }
            else
            {
                R[3 * r] = c * Rj[3 * r] - s * Rj[3 * r + 2];
                R[3 * r + 1] = Rj[3 * r + 1];
                R[3 * r + 2] = s * Rj[3 * r] + c * Rj[3 * r + 2];
            }
        }
    }
    const double *F = ch.Rfix[6], *t = ch.pfix[6], *te = par + RP_TEE;
    double Re[9], pe[3], tw[3], pt[3];
    RD_UNROLL
    for (int r = 0; r < 3; r++)
    {
        RD_UNROLL
        for (int c = 0; c < 3; c++) Re[3 * r + c] = R[3 * r] * F[c] + R[3 * r + 1] * F[3 + c] + R[3 * r + 2] * F[6 + c];
        pe[r] = p[r] + R[3 * r] * t[0] + R[3 * r + 1] * t[1] + R[3 * r + 2] * t[2];
    }
    RD_UNROLL
    for (int r = 0; r < 3; r++)
    {
        tw[r] = Re[3 * r] * te[0] + Re[3 * r + 1] * te[1] + Re[3 * r + 2] * te[2];
        pt[r] = pe[r] + tw[r];
    }
    const double sdot = tws[3] * tw[0] + tws[4] * tw[1] + tws[5] * tw[2];
    const double vv[3] = {tws[0] + sdot, tws[1] + sdot, tws[2] + sdot};
    const double vt1 = Re[3] * vv[0] + Re[4] * vv[1] + Re[5] * vv[2];
    const double sa = par[RP_SURF], sb = par[RP_SURF + 1], sc = par[RP_SURF + 2], sd = par[RP_SURF + 3], se = par[RP_SURF + 4],
                 sf = par[RP_SURF + 5];
    const double X = pt[0], Y = pt[1];
    const double S = sa * X * X + sb * Y * Y + sc * X * Y + sd * X + se * Y + sf;
    const double Sx = 2 * sa * X + sc * Y + sd, Sy = 2 * sb * Y + sc * X + se;
    const double irho = rd_rsqrt(Sx * Sx + Sy * Sy + 1.0);
    e[0] = S - pt[2];
    e[1] = 1.0 - (Sx * irho * Re[2] + Sy * irho * Re[5] - irho * Re[8]);
    e[2] = Re[1];
    e[3] = par[RP_PXREF] - pt[0];
    e[4] = par[RP_VYREF] - vt1;
    if (REC)
    { /* the per-sample logs of R/simulation_model.py:60-77,85-91 and R/simulator.py:344 */
        RD_UNROLL
        for (int r = 0; r < 3; r++) rec[RK_POSE + r] = pe[r];
        RD_UNROLL
        for (int r = 0; r < 9; r++) rec[RK_POSE + 3 + r] = Re[r];
        rec[RK_RPY] = atan2(Re[7], Re[8]);
        rec[RK_RPY + 1] = atan2(-Re[6], sqrt(Re[0] * Re[0] + Re[3] * Re[3]));
        rec[RK_RPY + 2] = atan2(Re[3], Re[0]);
        RD_UNROLL
        for (int r = 0; r < 6; r++) rec[RK_TWIST + r] = tws[r];
        RD_UNROLL
        for (int r = 0; r < 5; r++) rec[RK_ERR + r] = e[r];
        rec[RK_PTZ] = pt[2];
    }
}
RD void ur10_tracking_errors(const RmpcChain &ch, const double *par, const double *q, const double *qd, double *e)
{
    ur10_plant_record<false>(ch, par, q, qd, e, nullptr);
}
/* ---------------------------------------------------------------------------------------------- NLP level (K2, K4)
 * Per-instance constants of the Gauss-Newton Hessian that do not depend on the stage:
 *   y[5+i] = u_i (weight 2 w_u),  y[11+i] = (b2_i u_i + (a22_i - 1) qd_i) / Ts (weight w_qddot)
 * => H_uu = diag(hu), H_{qd,u} = diag(hs), H_{qd,qd} += diag(hq)        (R/trajectory_optimizer.py:121-154) */
struct StageConst
{
    double hu, hs, hq, cu, cq; /* for one joint */
};
RD StageConst stage_const(const double *par, double dt, int i)
{
    StageConst c;
    c.cu = par[RP_B2 + i] / dt;
    c.cq = (par[RP_A22 + i] - 1.0) / dt;
    const double wq = par[RP_WQDD + i], wu = 2.0 * par[RP_WU + i];
    c.hu = dt * (wu + wq * c.cu * c.cu);
    c.hs = dt * (wq * c.cu * c.cq);
    c.hq = dt * (wq * c.cq * c.cq);
    return c;
}

/* stage cost dt * 1/2 |y - yref|^2_W from the five task outputs and (u, qd): ocp_nlp_cost_nls_compute_fun
 * (AC/ocp_nlp/ocp_nlp_cost_nls.c:1045-1133) */
/* yr = the 18 references of the stage: the stage's row of the per-stage table, else RP_YREF of the parameter block
 * (the reference's yref = [0, 1, 0, px_ref, vy_ref, 0 ...]: y - 0.0 is y bit for bit) */
RD const double *yref_of(const double *par, const double *stg, int i) { return stg ? stg + (long) RT_COUNT * i + RT_YREF : par + RP_YREF; }
RD double stage_cost(const double *par, const double *yr, double dt, const double *y5, const double *u, const double *qd)
{
    const double r0 = y5[0] - yr[0], r1 = y5[1] - yr[1], r2 = y5[2] - yr[2], r3 = y5[3] - yr[3], r4 = y5[4] - yr[4];
    double f = par[RP_WTASK] * r0 * r0 + par[RP_WTASK + 1] * r1 * r1 + par[RP_WTASK + 2] * r2 * r2 + par[RP_WTASK + 3] * r3 * r3 +
               par[RP_WTASK + 4] * r4 * r4;
    RD_UNROLL
    for (int i = 0; i < 6; i++)
    {
        const double wq = par[RP_WQDD + i], wu = 2.0 * par[RP_WU + i];
        const double a = (par[RP_B2 + i] * u[i] + (par[RP_A22 + i] - 1.0) * qd[i]) / dt - yr[11 + i];
        const double ru = u[i] - yr[5 + i];
        f += wu * ru * ru + wq * a * a;
    }
    return dt * (0.5 * f);
}

RD void load18(const double *src, double *v)
{
    RD_UNROLL
    for (int k = 0; k < NV; k++) v[k] = src[k];
}

/* lower/upper bound of slot j in [0,12) of a path stage: [u_min|q_min], [u_max|q_max] */
RD double lb_of(const double *par, const double *sp, int j) { return sp ? sp[RT_LB + j] : (j < 6 ? par[RP_UMIN + j] : par[RP_QMIN + j - 6]); }
RD double ub_of(const double *par, const double *sp, int j) { return sp ? sp[RT_UB + j] : (j < 6 ? par[RP_UMAX + j] : par[RP_QMAX + j - 6]); }

struct LinEmit
{
    double *Jc, *g;
    const double *par, *yr;
    double dt, wr[5], sw[5], gqd0[6];
    RDM void begin(const double *y5)
    {
        const double res[5] = {y5[0] - yr[0], y5[1] - yr[1], y5[2] - yr[2], y5[3] - yr[3], y5[4] - yr[4]};
        RD_UNROLL
        for (int m = 0; m < 5; m++)
        {
            wr[m] = dt * par[RP_WTASK + m] * res[m];
            sw[m] = sqrt(dt * par[RP_WTASK + m]);
        }
    }
    RDM void operator()(int k, const double *Jq, double Jd) const
    {
        double gq = 0.0, gqd = gqd0[k];
        RD_UNROLL
        for (int m = 0; m < 5; m++)
        {
            gq += Jq[m] * wr[m];
            Jc[6 * m + k] = sw[m] * Jq[m];
        }
        gqd += Jd * wr[4];
        Jc[30 + k] = sw[4] * Jd;
        g[6 + k] = gq;
        g[12 + k] = gqd;
    }
};

/* Linearisation at the current iterate, stage-parallel: ocp_nlp_approximate_qp_matrices + _vectors_sqp
 * (AC/ocp_nlp/ocp_nlp_common.c:2983-3106) with NLS cost (ocp_nlp_cost_nls.c:830-1032), discrete dynamics
 * (ocp_nlp_dynamics_disc.c:611-708) and box bounds (ocp_nlp_constraints_bgh.c:1563-1619).
 * Writes per stage: Jc = sqrt(dt W_m) * d y5_m / d x (non-zero entries), rqz = cost gradient, b = dynamics gap,
 * d = [lb - v; v - ub]; dx0 = x0 - x_nlp[0] (both bound residuals of the stage-0 state equality).
 * Returns the total cost. */
RDN double nlp_linearise(const RmpcChain &ch, const RmpcDims &D, const Inst &I_, const double *x0)
{
    const Inst I = I_; /* by value: registers, not re-read from the caller's frame after every store */
    const int N = D.N, lane = rd_lane();
    const double dt = D.dt;
    const double *par = I.par;
    double cost = 0.0;
    for (int i = lane; i < N; i += GL)
    {
        double v[NV], y5[5];
        load18(AR(ux) + NV * i, v);
        double *g = AR(rqz) + NV * i;
        LinEmit em;
        em.Jc = AR(Jc) + NJ * i;
        em.g = g;
        const double *sp = I.stg ? I.stg + (long) RT_COUNT * i : nullptr, *yr = yref_of(par, I.stg, i);
        RD_UNROLL
        for (int k = 0; k < 6; k++)
        {
            const double wq = par[RP_WQDD + k], wu = 2.0 * par[RP_WU + k];
            const double cu = par[RP_B2 + k] / dt, cq = (par[RP_A22 + k] - 1.0) / dt;
            const double a = (cu * v[k] + cq * v[12 + k]) - yr[11 + k]; /* qddot_k - its reference */
            g[k] = dt * (wu * (v[k] - yr[5 + k]) + wq * cu * a);
            em.gqd0[k] = dt * (wq * cq * a);
        }
        em.par = par;
        em.yr = yr;
        em.dt = dt;
        ur10_task<true>(ch, par, v + 6, v + 12, y5, em);
        const double ci = stage_cost(par, yr, dt, y5, v, v + 12);
        cost += ci;
        AR(tux)[NV * i] = ci; /* kept for the merit value at this iterate (nlp_merit_lin): tux is free until the line search */
        /* dynamics gap b = Ad x + Bd u - x_next */
        const double *xn = AR(ux) + NV * (i + 1) + 6;
        double *b = AR(b) + NX * i, *d = AR(d) + NB * i;
        RD_UNROLL
        for (int k = 0; k < 6; k++)
        {
            b[k] = (v[6 + k] + par[RP_A12 + k] * v[12 + k]) + par[RP_B1 + k] * v[k] - xn[k];
            b[6 + k] = par[RP_A22 + k] * v[12 + k] + par[RP_B2 + k] * v[k] - xn[6 + k];
        }
        RD_UNROLL
        for (int j = 0; j < 12; j++)
        {
            if (i == 0 && j >= 6) { d[j] = 0.0; d[12 + j] = 0.0; }
            else { d[j] = lb_of(par, sp, j) - v[j]; d[12 + j] = v[j] - ub_of(par, sp, j); }
        }
        if (i == 0)
        {
            RD_UNROLL
            for (int k = 0; k < NX; k++)
            {
                (I.head + H_DX0)[k] = x0[k] - v[6 + k];
                (I.head + H_B0)[k] = b[k];
                (I.head + H_XLIN0)[k] = v[6 + k];
            }
        }
    }
    for (int k = lane; k < NV; k += GL) AR(rqz)[NV * N + k] = 0.0; /* stage N: no cost (ny_e = 0) */
    if (I.pe)
    { /* terminal row: fun = [lh - h_e; h_e - uh], DCt = grad h_e (kept in Jc[N][0..11]) */
        const double h = term_eval(I.pe, AR(ux) + NV * N + 6, AR(Jc) + NJ * N);
        if (lane == 0)
        {
            AR(d)[TERM_LO(N)] = I.pe[RE_LH] - h;
            AR(d)[TERM_UP(N)] = h - I.pe[RE_UH];
        }
    }
    rd_sync();
    return grp_sum(cost);
}

/* Merit function at the trial iterate tux, ocp_nlp_evaluate_merit_fun
 * (AC/ocp_nlp/ocp_nlp_globalization_merit_backtracking.c:606-702):
 *   M = cost + sum |w_pi| |dyn gap| + sum |w_lam| max(0, bound violation)   (incl. the stage-0 x0 equality) */
RDN double nlp_merit(const RmpcChain &ch, const RmpcDims &D, const Inst &I_, const double *tux, const double *x0, double *cost_out)
{
    double csum = 0.0;
    const Inst I = I_; /* by value: registers, not re-read from the caller's frame after every store */
    const int N = D.N, lane = rd_lane();
    const double *par = I.par;
    double m = 0.0;
    for (int i = lane; i < N; i += GL)
    {
        double v[NV], y5[5];
        load18(tux + NV * i, v);
        NoEmit ne;
        ur10_task<false>(ch, par, v + 6, v + 12, y5, ne);
        const double *sp = I.stg ? I.stg + (long) RT_COUNT * i : nullptr;
        const double ci = stage_cost(par, yref_of(par, I.stg, i), D.dt, y5, v, v + 12);
        m += ci;
        csum += ci; /* the cost alone, summed like solver.get_cost() does (closed_loop_group) */
        const double *xn = tux + NV * (i + 1) + 6, *wpi = AR(wpi) + NX * i, *wl = AR(wlam) + NB * i;
        RD_UNROLL
        for (int k = 0; k < 6; k++)
        {
            const double g0 = (v[6 + k] + par[RP_A12 + k] * v[12 + k]) + par[RP_B1 + k] * v[k] - xn[k];
            const double g1 = par[RP_A22 + k] * v[12 + k] + par[RP_B2 + k] * v[k] - xn[6 + k];
            m += fabs(wpi[k]) * fabs(g0) + fabs(wpi[6 + k]) * fabs(g1);
        }
        RD_UNROLL
        for (int j = 0; j < 12; j++)
        {
            if (i == 0 && j >= 6) continue;
            const double fl = lb_of(par, sp, j) - v[j], fu = v[j] - ub_of(par, sp, j);
            if (fl > 0.0) m += fabs(wl[j]) * fl;
            if (fu > 0.0) m += fabs(wl[12 + j]) * fu;
        }
        if (i == 0)
        {
            RD_UNROLL
            for (int k = 0; k < NX; k++)
            {
                const double fl = x0[k] - v[6 + k];
                if (fl > 0.0) m += fabs((I.head + H_WLAMX0)[k]) * fl;
                if (-fl > 0.0) m += fabs((I.head + H_WLAMX0)[12 + k]) * (-fl);
            }
        }
    }
    if (I.pe)
    {
        const double h = term_eval(I.pe, tux + NV * N + 6, nullptr);
        const double fl = I.pe[RE_LH] - h, fu = h - I.pe[RE_UH];
        if (lane == 0)
        {
            if (fl > 0.0) m += fabs(AR(wlam)[TERM_LO(N)]) * fl;
            if (fu > 0.0) m += fabs(AR(wlam)[TERM_UP(N)]) * fu;
        }
    }
    if (cost_out) *cost_out = grp_sum(csum);
    return grp_sum(m);
}

/* The same merit value at the CURRENT iterate, without a kinematics pass: every ingredient was produced by the
 * linearisation of this iterate -- stage costs (kept in tux[.][0]), dynamics gaps b, bound residuals d, dx0 -- and is
 * accumulated in the order nlp_merit uses.  Stage 0's gap is re-formed from the iterate (qp_reduce folded dx0 into b_0). */
RDN double nlp_merit_lin(const RmpcDims &D, const Inst &I_, const double *x0)
{
    const Inst I = I_;
    const int N = D.N, lane = rd_lane();
    const double *par = I.par;
    double m = 0.0;
    for (int i = lane; i < N; i += GL)
    {
        m += AR(tux)[NV * i];
        const double *wpi = AR(wpi) + NX * i, *wl = AR(wlam) + NB * i, *d = AR(d) + NB * i;
        if (i == 0)
        {
            const double *v = AR(ux), *xn = AR(ux) + NV + 6;
            RD_UNROLL
            for (int k = 0; k < 6; k++)
            {
                const double g0 = (v[6 + k] + par[RP_A12 + k] * v[12 + k]) + par[RP_B1 + k] * v[k] - xn[k];
                const double g1 = par[RP_A22 + k] * v[12 + k] + par[RP_B2 + k] * v[k] - xn[6 + k];
                m += fabs(wpi[k]) * fabs(g0) + fabs(wpi[6 + k]) * fabs(g1);
            }
        }
        else
        {
            const double *b = AR(b) + NX * i;
            RD_UNROLL
            for (int k = 0; k < 6; k++) m += fabs(wpi[k]) * fabs(b[k]) + fabs(wpi[6 + k]) * fabs(b[6 + k]);
        }
        RD_UNROLL
        for (int j = 0; j < 12; j++)
        {
            if (i == 0 && j >= 6) continue;
            const double fl = d[j], fu = d[12 + j];
            if (fl > 0.0) m += fabs(wl[j]) * fl;
            if (fu > 0.0) m += fabs(wl[12 + j]) * fu;
        }
        if (i == 0)
        {
            RD_UNROLL
            for (int k = 0; k < NX; k++)
            {
                const double fl = x0[k] - AR(ux)[6 + k];
                if (fl > 0.0) m += fabs((I.head + H_WLAMX0)[k]) * fl;
                if (-fl > 0.0) m += fabs((I.head + H_WLAMX0)[12 + k]) * (-fl);
            }
        }
    }
    if (I.pe && lane == 0)
    {
        const double fl = AR(d)[TERM_LO(N)], fu = AR(d)[TERM_UP(N)];
        if (fl > 0.0) m += fabs(AR(wlam)[TERM_LO(N)]) * fl;
        if (fu > 0.0) m += fabs(AR(wlam)[TERM_UP(N)]) * fu;
    }
    return grp_sum(m);
}

/* KKT residual inf-norms of the NLP, ocp_nlp_res_compute (AC/ocp_nlp/ocp_nlp_common.c:3641-3751); uses the
 * linearisation (rqz, b, d, dx0) of the current iterate.  res = [stat, eq, ineq, comp]. */
RDN void nlp_residuals(const RmpcDims &D, const Inst &I_, double *res)
{
    const Inst I = I_; /* by value: registers, not re-read from the caller's frame after every store */
    const int N = D.N, lane = rd_lane();
    const double *par = I.par;
    double rs = 0, re = 0, ri = 0, rc = 0;
    for (int i = lane; i <= N; i += GL)
    {
        double r[NV];
        RD_UNROLL
        for (int k = 0; k < NV; k++) r[k] = AR(rqz)[NV * i + k];
        if (i < N)
        {
            const double *lam = AR(lam) + NB * i, *pi = AR(pi) + NX * i, *d = AR(d) + NB * i;
            RD_UNROLL
            for (int j = 0; j < 12; j++)
            {
                if (i == 0 && j >= 6) continue;
                r[j] -= lam[j] - lam[12 + j];
                ri = rd_fmax(ri, rd_fmax(d[j], d[12 + j]));
                rc = rd_fmax(rc, rd_fmax(fabs(lam[j] * d[j]), fabs(lam[12 + j] * d[12 + j])));
            }
            RD_UNROLL
            for (int k = 0; k < 6; k++)
            {
                r[k] += par[RP_B1 + k] * pi[k] + par[RP_B2 + k] * pi[6 + k];
                r[6 + k] += pi[k];
                r[12 + k] += par[RP_A12 + k] * pi[k] + par[RP_A22 + k] * pi[6 + k];
                re = rd_fmax(re, rd_fmax(fabs(AR(b)[NX * i + k]), fabs(AR(b)[NX * i + 6 + k])));
            }
        }
        if (i == 0)
        {
            RD_UNROLL
            for (int k = 0; k < NX; k++)
            {
                const double ll = (I.head + H_LAMX0)[k], lu = (I.head + H_LAMX0)[12 + k], f = (I.head + H_DX0)[k];
                r[6 + k] -= ll - lu;
                ri = rd_fmax(ri, rd_fmax(f, -f));
                rc = rd_fmax(rc, rd_fmax(fabs(ll * f), fabs(lu * f)));
            }
        }
        else
        {
            RD_UNROLL
            for (int k = 0; k < NX; k++) r[6 + k] -= AR(pi)[NX * (i - 1) + k];
        }
        if (i == N && I.pe)
        { /* terminal row: adj += DCt (lam_l - lam_u) */
            const double ll = AR(lam)[TERM_LO(N)], lu = AR(lam)[TERM_UP(N)], fl = AR(d)[TERM_LO(N)], fu = AR(d)[TERM_UP(N)];
            RD_UNROLL
            for (int k = 0; k < NX; k++) r[6 + k] -= (AR(Jc) + NJ * N)[k] * (ll - lu);
            ri = rd_fmax(ri, rd_fmax(fl, fu));
            rc = rd_fmax(rc, rd_fmax(fabs(ll * fl), fabs(lu * fu)));
        }
        RD_UNROLL
        for (int k = 0; k < NV; k++)
            if (i < N || k >= 6) rs = rd_fmax(rs, fabs(r[k]));
        if (r[6] != r[6]) rs = r[6]; /* propagate NaN like the reference's isnan test */
    }
    res[0] = grp_max(rs); res[1] = grp_max(re); res[2] = grp_max(ri); res[3] = grp_max(rc);
}

/* ---------------------------------------------------------------------------------------------- QP level (K3)
 * x0 elimination at stage 0, d_ocp_qp_reduce_eq_dof (H/ocp_qp/x_ocp_qp_red.c:278-466), done in place:
 *   b_0 += A dx0,  r_0 += S' dx0 (S = H_{x,u}: only the diagonal qd-u coupling), original r_0 kept in g0u. */
RD void qp_reduce(const RmpcDims &D, const Inst &I)
{
    const int lane = rd_lane();
    const double *par = I.par;
    if (lane < 6)
    {
        const StageConst c = stage_const(par, D.dt, lane);
        const double g = AR(rqz)[lane];
        (I.head + H_G0U)[lane] = g;
        AR(rqz)[lane] = g + c.hs * (I.head + H_DX0)[6 + lane];
        AR(b)[lane] += (I.head + H_DX0)[lane] + par[RP_A12 + lane] * (I.head + H_DX0)[6 + lane];
        AR(b)[6 + lane] += par[RP_A22 + lane] * (I.head + H_DX0)[6 + lane];
    }
    rd_sync();
}

/* OCP_QP_INIT_VAR (H/ocp_qp/x_ocp_qp_ipm.c:1592-1980).  warm = 2: keep pi, clip lam, t at 0.1 (:1638-1654).
 * warm = 0 (every QP but the first, see oracle/rmpc_oracle.c:solve_qp for the reference quirk): ux = 0, pi = 0,
 * t = slack pushed to >= 0.1 by moving ux, lam = mu0 / t with mu0 = 1 (t0_init 2, var_init_scheme 1, :1837-1978).
 * The primal iterate is zeroed first in both cases (AC/ocp_qp/ocp_qp_hpipm.c:309-316). */
/* Partial condensing, qp_solver_cond_N = N2 < N (AC/ocp_qp/ocp_qp_partial_condensing.c:523-559,664-689; H/cond/x_part_cond.c:
 * 486-558, block sizes :36-60): acados hands HPIPM an N2-stage QP whose stage k holds the controls of the stages of block k
 * and the state at its first stage; the states inside a block are eliminated through the dynamics and their bounds become
 * general constraints.  The interior-point iterates of that QP are reproduced here WITHOUT forming its dense blocks.  Newton
 * steps on the condensed KKT system are the Newton steps of the stage-wise system restricted to iterates that (i) satisfy the
 * dynamics inside every block and (ii) carry, inside every block, the multipliers pi given by the adjoint recursion
 * (stationarity of the eliminated states) -- both properties are linear, so every damped Newton step preserves them, the
 * eliminated rows contribute nothing to the residual norms, and the set of inequality multipliers / slacks is the same.  What
 * differs is the START: OCP_QP_INIT_VAR on the condensed QP (H/ocp_qp/x_ocp_qp_ipm.c:1724-1817) zeroes u and the block-start
 * states only and treats the bounds of the eliminated states as GENERAL rows (slack = max(0.1, +-(C x + D u) - d), no move
 * of the variables).  ipm_init_pcond() builds exactly that starting point; the stage-wise Riccati sweeps then walk the
 * condensed problem's iterates (to rounding) and the final iterate is the expanded solution (d_part_cond_qp_expand_sol).
 * Forming the dense blocks would not shorten the dependent chain on this machine: the factorisation of a block of k stages
 * has the same 6 k sequential pivots as k stage-wise steps, at (6 k + 12)^3 / 3 instead of k 18^3 / 3 flops. */
RD void task_hess_mv(const double *Jc, const double *x, double *y);
RD void pc_block(int N, int N2, int k, int *s, int *bs)
{
    const int b0 = N / N2, r = N - N2 * b0;
    *bs = b0 + (k < r ? 1 : 0);
    *s = k < r ? k * (b0 + 1) : r * (b0 + 1) + (k - r) * b0;
}
RD bool pc_is_start(int N, int N2, int i)
{
    const int b0 = N / N2, r = N - N2 * b0, big = r * (b0 + 1);
    return i < big ? (i % (b0 + 1)) == 0 : ((i - big) % b0) == 0;
}

RDN void ipm_init_pcond(const RmpcDims &D, const Inst &I_, const GrpSmem &sm, int warm)
{
    const Inst I = I_;
    const int N = D.N, N2 = I.N2, lane = rd_lane();
    for (int kb = lane; kb < N2; kb += GL)
    {
        int s, bs;
        pc_block(N, N2, kb, &s, &bs);
        if (bs < 2) continue;
        /* (i) states inside the block from the dynamics, bounds of those states as general rows */
        double x[NX];
        RD_UNROLL
        for (int k = 0; k < NX; k++) x[k] = AR(sux)[NV * s + 6 + k];
        for (int i = s; i < s + bs - 1; i++)
        {
            const double *u = AR(sux) + NV * i, *b = AR(b) + NX * i;
            double xn[NX];
            RD_UNROLL
            for (int k = 0; k < 6; k++)
            {
                xn[k] = (x[k] + C_A12(k) * x[6 + k]) + C_B1(k) * u[k] + b[k];
                xn[6 + k] = C_A22(k) * x[6 + k] + C_B2(k) * u[k] + b[6 + k];
            }
            RD_UNROLL
            for (int k = 0; k < NX; k++) { x[k] = xn[k]; AR(sux)[NV * (i + 1) + 6 + k] = xn[k]; }
            if (warm != 2)
            {
                const double *d = AR(d) + NB * (i + 1);
                RD_UNROLL
                for (int j = 0; j < 6; j++)
                {
                    double tl = xn[j] - d[6 + j], tu = -xn[j] - d[18 + j];
                    tl = 1e-1 > tl ? 1e-1 : tl;
                    tu = 1e-1 > tu ? 1e-1 : tu;
                    AR(st)[NB * (i + 1) + 6 + j] = tl;
                    AR(st)[NB * (i + 1) + 18 + j] = tu;
                    AR(slam)[NB * (i + 1) + 6 + j] = 1.0 / tl;
                    AR(slam)[NB * (i + 1) + 18 + j] = 1.0 / tu;
                }
            }
        }
        /* (ii) multipliers of the dynamics inside the block: stationarity of the eliminated states, backwards from the
         *      multiplier of the block's last stage */
        for (int i = s + bs - 1; i > s; i--)
        {
            double v[NV], g[NX];
            load18(AR(sux) + NV * i, v);
            RD_UNROLL
            for (int k = 0; k < NX; k++) g[k] = AR(rqz)[NV * i + 6 + k];
            task_hess_mv(AR(Jc) + NJ * i, v + 6, g);
            const double *pin = AR(spi) + NX * i, *l = AR(slam) + NB * i;
            RD_UNROLL
            for (int k = 0; k < 6; k++)
            {
                g[k] += pin[k];
                g[6 + k] += sm.hc[6 + k] * v[k] + sm.hc[12 + k] * v[12 + k] + C_A12(k) * pin[k] + C_A22(k) * pin[6 + k];
                g[k] += l[18 + k] - l[6 + k];
            }
            RD_UNROLL
            for (int k = 0; k < NX; k++) AR(spi)[NX * (i - 1) + k] = g[k];
        }
    }
    rd_sync();
}

RDN void ipm_init(const RmpcDims &D, const Inst &I_, int warm)
{
    const Inst I = I_; /* by value: registers, not re-read from the caller's frame after every store */
    const int N = D.N, lane = rd_lane();
    {
        rd_double2 zz;
        zz.x = 0.0; zz.y = 0.0;
        for (int k = lane; k < (N + 1) * NV / 2; k += GL) RD_V2(AR(sux))[k] = zz;
        if (warm != 2)
            for (int k = lane; k < N * NX / 2; k += GL) RD_V2(AR(spi))[k] = zz;
    }
    rd_sync();
    /* stage-parallel over the bound slots: lane l takes stages l, l + GL, ..., six slots (lower side j, upper side 12 + j)
     * at a time with 128-bit accesses; stage 0 only has the u half */
    for (int i = lane; i < N; i += GL)
    {
        RD_UNROLL
        for (int h = 0; h < 12; h += 6)
        {
            if (i == 0 && h >= 6) continue;
            if (h >= 6 && warm != 2 && I.N2 && !pc_is_start(N, I.N2, i)) continue; /* general rows of the condensed QP: ipm_init_pcond */
            double la[6], lb[6], ta[6], tb[6];
            double *pl = AR(slam) + NB * i + h, *pt = AR(st) + NB * i + h;
            if (warm == 2)
            {
                RD_UNROLL
                for (int c = 0; c < 6; c += 2)
                {
                    const rd_double2 a0 = *RD_CV2(pl + c), a1 = *RD_CV2(pl + 12 + c), a2 = *RD_CV2(pt + c), a3 = *RD_CV2(pt + 12 + c);
                    la[c] = a0.x; la[c + 1] = a0.y; lb[c] = a1.x; lb[c + 1] = a1.y;
                    ta[c] = a2.x; ta[c + 1] = a2.y; tb[c] = a3.x; tb[c + 1] = a3.y;
                }
                RD_UNROLL
                for (int j = 0; j < 6; j++)
                {
                    if (la[j] < 1e-1) la[j] = 1e-1;
                    if (lb[j] < 1e-1) lb[j] = 1e-1;
                    if (ta[j] < 1e-1) ta[j] = 1e-1;
                    if (tb[j] < 1e-1) tb[j] = 1e-1;
                }
            }
            else
            {
                double da[6], db[6];
                const double *pd = AR(d) + NB * i + h;
                RD_UNROLL
                for (int c = 0; c < 6; c += 2)
                {
                    const rd_double2 a0 = *RD_CV2(pd + c), a1 = *RD_CV2(pd + 12 + c);
                    da[c] = a0.x; da[c + 1] = a0.y; db[c] = a1.x; db[c + 1] = a1.y;
                }
                RD_UNROLL
                for (int j = 0; j < 6; j++)
                {
                    const double thr0 = 1e-1, dl = da[j], du = db[j];
                    double v = 0.0, tl = -dl, tu = -du;
                    if (tl < thr0)
                    {
                        if (tu < thr0) { v = 0.5 * (dl - du); tl = thr0; tu = thr0; }
                        else { tl = thr0; v = dl + thr0; }
                    }
                    else if (tu < thr0) { tu = thr0; v = -du - thr0; }
                    da[j] = v;
                    ta[j] = tl; tb[j] = tu;
                    la[j] = 1.0 / tl; lb[j] = 1.0 / tu;
                }
                RD_UNROLL
                for (int c = 0; c < 6; c += 2)
                {
                    rd_double2 v2;
                    v2.x = da[c]; v2.y = da[c + 1];
                    *RD_V2(AR(sux) + NV * i + h + c) = v2;
                }
            }
            RD_UNROLL
            for (int c = 0; c < 6; c += 2)
            {
                rd_double2 a0, a1, a2, a3;
                a0.x = la[c]; a0.y = la[c + 1]; a1.x = lb[c]; a1.y = lb[c + 1];
                a2.x = ta[c]; a2.y = ta[c + 1]; a3.x = tb[c]; a3.y = tb[c + 1];
                *RD_V2(pl + c) = a0; *RD_V2(pl + 12 + c) = a1; *RD_V2(pt + c) = a2; *RD_V2(pt + 12 + c) = a3;
            }
        }
    }
    if (I.pe && lane < 2)
    { /* general row of stage N (x_ocp_qp_ipm.c:1780-1817): t = +-C ux - d at ux = 0, clipped at 0.1; lam = mu0 / t */
        const int k = lane ? TERM_UP(N) : TERM_LO(N);
        if (warm == 2)
        {
            if (AR(slam)[k] < 1e-1) AR(slam)[k] = 1e-1;
            if (AR(st)[k] < 1e-1) AR(st)[k] = 1e-1;
        }
        else
        {
            double t = -AR(d)[k];
            t = 1e-1 > t ? 1e-1 : t;
            AR(st)[k] = t;
            AR(slam)[k] = 1.0 / t;
        }
    }
    rd_sync();
}

/* y += H_xx x for the task part of the Gauss-Newton Hessian, H_xx = Jc' Jc with the sparse Jc of one stage
 * (rows 0..3 touch q only, row 4 touches q and qd). */
RD void task_hess_mv(const double *Jc, const double *x, double *y)
{
    RD_UNROLL
    for (int m = 0; m < 5; m++)
    {
        double w = 0.0;
        RD_UNROLL
        for (int k = 0; k < 6; k++) w += Jc[6 * m + k] * x[k];
        if (m == 4)
        {
            RD_UNROLL
            for (int k = 0; k < 6; k++) w += Jc[30 + k] * x[6 + k];
        }
        RD_UNROLL
        for (int k = 0; k < 6; k++) y[k] += Jc[6 * m + k] * w;
        if (m == 4)
        {
            RD_UNROLL
            for (int k = 0; k < 6; k++) y[6 + k] += Jc[30 + k] * w;
        }
    }
}

/* QP residuals, stage-parallel.  LIN = 0: OCP_QP_RES_COMPUTE (H/ocp_qp/x_ocp_qp_res.c:345-533) at (ux,pi,lam,t), m = 0;
 * LIN = 1: OCP_QP_RES_COMPUTE_LIN (:535-685), residual of the linearised KKT system at the step (ux,pi,lam,t) around
 * (Lam,T) with right-hand sides (rqz,b,d,m - mshift).  Outputs the four inf-norms (OCP_QP_RES_COMPUTE_INF_NORM
 * :689-726) in nrm[0..3], nrm[4] = |res_m - tau_min|_inf, nrm[5] = NaN flag of rg[0]; returns sum lam*t (LIN = 0).
 * WRITE = 0 computes the norms only. */
template <int LIN, int WRITE>
RDN double ipm_res(const RmpcDims &D, const Inst &I_, const GrpSmem &sm, const double *rqz, const double *b, const double *d, const double *m,
                  double mshift, const double *ux, const double *pi, const double *lam, const double *t, const double *Lam,
                  const double *T, double *rg, double *rb, double *rd, double *rm, double *nrm)
{
    const Inst I = I_; /* by value: registers, not re-read from the caller's frame after every store */
    const int N = D.N, lane = rd_lane();
    const double *par = I.par;
    double mu = 0.0, n0 = 0, n1 = 0, n2 = 0, n3 = 0, n4 = 0, nanflag = 0;
    for (int i = lane; i <= N; i += GL)
    {
        double v[NV], g[NV];
        load18(ux + NV * i, v);
        RD_UNROLL
        for (int k = 0; k < NV; k++) g[k] = rqz[NV * i + k];
        if (i == 0)
        {
            RD_UNROLL
            for (int k = 6; k < NV; k++) { v[k] = 0.0; g[k] = 0.0; }
        }
        if (i < N)
        {
            /* H v: task part on x (stages >= 1), constant diagonal blocks */
            if (i > 0) task_hess_mv(AR(Jc) + NJ * i, v + 6, g + 6);
            const double *pin = pi + NX * i, *l = lam + NB * i, *tt = t + NB * i;
            const double *xn = ux + NV * (i + 1) + 6;
            RD_UNROLL
            for (int k = 0; k < 6; k++)
            {
                const double hu = sm.hc[k], hs = sm.hc[6 + k], hq = sm.hc[12 + k];
                const double a12 = C_A12(k), a22 = C_A22(k), b1 = C_B1(k), b2 = C_B2(k);
                g[k] += hu * v[k] + hs * v[12 + k] + b1 * pin[k] + b2 * pin[6 + k];
                if (i > 0)
                {
                    g[6 + k] += pin[k];
                    g[12 + k] += hs * v[k] + hq * v[12 + k] + a12 * pin[k] + a22 * pin[6 + k];
                }
                const double r0 = b[NX * i + k] - xn[k] + (v[6 + k] + a12 * v[12 + k]) + b1 * v[k];
                const double r1 = b[NX * i + 6 + k] - xn[6 + k] + a22 * v[12 + k] + b2 * v[k];
                if (WRITE) { rb[NX * i + k] = r0; rb[NX * i + 6 + k] = r1; }
                n1 = rd_fmax(n1, rd_fmax(fabs(r0), fabs(r1)));
            }
            RD_UNROLL
            for (int j = 0; j < 12; j++)
            {
                if (i == 0 && j >= 6) continue;
                g[j] += l[12 + j] - l[j];
                const double dl = d[NB * i + j] + tt[j] - v[j], du = d[NB * i + 12 + j] + tt[12 + j] + v[j];
                if (WRITE) { rd[NB * i + j] = dl; rd[NB * i + 12 + j] = du; }
                double ml, mu_;
                if (!LIN) { ml = l[j] * tt[j]; mu_ = l[12 + j] * tt[12 + j]; mu += ml + mu_; }
                else
                {
                    ml = (m[NB * i + j] - mshift) + Lam[NB * i + j] * tt[j] + l[j] * T[NB * i + j];
                    mu_ = (m[NB * i + 12 + j] - mshift) + Lam[NB * i + 12 + j] * tt[12 + j] + l[12 + j] * T[NB * i + 12 + j];
                }
                if (WRITE) { rm[NB * i + j] = ml; rm[NB * i + 12 + j] = mu_; }
                n2 = rd_fmax(n2, rd_fmax(fabs(dl), fabs(du)));
                n3 = rd_fmax(n3, rd_fmax(fabs(ml), fabs(mu_)));
                n4 = rd_fmax(n4, rd_fmax(fabs(ml - 1e-16), fabs(mu_ - 1e-16)));
            }
        }
        if (i > 0)
        {
            RD_UNROLL
            for (int k = 0; k < NX; k++) g[6 + k] -= pi[NX * (i - 1) + k];
        }
        if (i == N && I.pe)
        { /* general row C x of stage N: res_g += C' (lam_u - lam_l), res_d = d + t -+ C x, res_m */
            const double *c = AR(Jc) + NJ * N;
            const int ka = TERM_LO(N), kb = TERM_UP(N);
            double cx = 0.0;
            RD_UNROLL
            for (int k = 0; k < NX; k++) cx += c[k] * v[6 + k];
            RD_UNROLL
            for (int k = 0; k < NX; k++) g[6 + k] += c[k] * (lam[kb] - lam[ka]);
            const double dl = d[ka] + t[ka] - cx, du = d[kb] + t[kb] + cx;
            if (WRITE) { rd[ka] = dl; rd[kb] = du; }
            double ml, mu_;
            if (!LIN) { ml = lam[ka] * t[ka]; mu_ = lam[kb] * t[kb]; mu += ml + mu_; }
            else
            {
                ml = (m[ka] - mshift) + Lam[ka] * t[ka] + lam[ka] * T[ka];
                mu_ = (m[kb] - mshift) + Lam[kb] * t[kb] + lam[kb] * T[kb];
            }
            if (WRITE) { rm[ka] = ml; rm[kb] = mu_; }
            n2 = rd_fmax(n2, rd_fmax(fabs(dl), fabs(du)));
            n3 = rd_fmax(n3, rd_fmax(fabs(ml), fabs(mu_)));
            n4 = rd_fmax(n4, rd_fmax(fabs(ml - 1e-16), fabs(mu_ - 1e-16)));
        }
        RD_UNROLL
        for (int k = 0; k < NV; k++)
        {
            if (WRITE) rg[NV * i + k] = g[k];
            if ((i < N || k >= 6) && (i > 0 || k < 6)) n0 = rd_fmax(n0, fabs(g[k]));
        }
        if (i == 0 && g[0] != g[0]) nanflag = 1.0;
    }
    rd_sync();
    nrm[0] = grp_max(n0); nrm[1] = grp_max(n1); nrm[2] = grp_max(n2); nrm[3] = grp_max(n3); nrm[4] = grp_max(n4);
    nrm[5] = grp_max(nanflag);
    return grp_sum(mu);
}

/* Row roles of the Riccati sweeps: lane l < 6 owns control row u_l and state row qd_l, lane 6 + j owns state row q_j.
 * The rows of [B A]' are ca e_q + cb e_qd (B, A of the exact-ZOH model are built from diagonal blocks). */
struct Role
{
    int j;        /* joint */
    int xk;       /* index of the owned state row in [q; qd] (lanes < 12) */
    bool x, u, q; /* owns a state row / a control row / the state row is a q row */
    double ca, cb;
};
RD Role make_role(const double *par)
{
    Role r;
    const int l = rd_lane();
    r.j = l % 6;
    r.x = l < 12;
    r.u = l < 6;
    r.q = l >= 6 && l < 12;
    r.xk = r.q ? r.j : 6 + r.j;
    r.ca = r.q ? 1.0 : par[RP_A12 + r.j];
    r.cb = r.q ? 0.0 : par[RP_A22 + r.j];
    return r;
}

RD void smem_consts(const Inst &I, GrpSmem &sm, double dt)
{
    for (int k = rd_lane(); k < 24; k += GL) sm.cst[k] = I.par[RP_A12 + k]; /* a12 a22 b1 b2 are contiguous */
    if (rd_lane() < 6)
    {
        const StageConst c = stage_const(I.par, dt, rd_lane());
        sm.hc[rd_lane()] = c.hu; sm.hc[6 + rd_lane()] = c.hs; sm.hc[12 + rd_lane()] = c.hq;
    }
    rd_sync();
}

/* Backward Riccati factorisation, first half of OCP_QP_FACT_SOLVE_KKT_STEP (H/ocp_qp/x_ocp_qp_kkt.c:836-966) incl.
 * COMPUTE_GAMMA_GAMMA_QP (H/ipm_core/x_core_qp_ipm_aux.c:38-86).  For ss = N-1..0, with P+, p+ of stage ss+1:
 *   M = [[R S'],[S Q]] + reg I + diag(Gamma_lb + Gamma_ub) + [B A]' P+ [B A]
 *   m = res_g + gamma_lb - gamma_ub + [B A]' (P+ res_b + p+)
 *   Luu Luu' = Muu, Lxu = Mxu Luu^-T, l_u = Luu^-1 m_u, P = Mxx - Lxu Lxu', p = m_x - Lxu l_u
 * (the reference factorises P = Lxx Lxx' as well and carries Lxx; see the header of this file).  The complementarity
 * residual enters as res_m - tau_min (BACKUP_RES_M + COMPUTE_TAU_MIN_QP, x_core_qp_ipm_aux.c:457-475,533-553). */
/* staging layout of one stage for the factorisation */
enum { FG_LAM = 0, FG_T = 24, FG_RD = 48, FG_RM = 72, FG_RG = 96, FG_RB = 114, FG_JC = 126 };
RD void fac_stage_in(const Inst &I, double *dst, int ss)
{
    stage_cp(dst + FG_LAM, AR(slam) + NB * ss, NB);
    stage_cp(dst + FG_T, AR(st) + NB * ss, NB);
    stage_cp(dst + FG_RD, AR(res_d) + NB * ss, NB);
    stage_cp(dst + FG_RM, AR(res_m) + NB * ss, NB);
    stage_cp(dst + FG_RG, AR(res_g) + NV * ss, NV);
    stage_cp(dst + FG_RB, AR(res_b) + NX * ss, NX);
    if (ss > 0) stage_cp(dst + FG_JC, AR(Jc) + NJ * ss, NJ);
}

/* the same staging with TMA bulk copies: lane 0 of the group arms the buffer's mbarrier with the byte count and issues one
 * copy per array */
RD void fac_stage_in_bulk(const Inst &I, double *dst, int ss, unsigned long long *bar)
{
    if (rd_lane() == 0)
    {
        rd_bar_expect(bar, (unsigned) sizeof(double) * (4 * NB + NV + NX + (ss > 0 ? NJ : 0)));
        rd_bulk_cp(dst + FG_LAM, AR(slam) + NB * ss, sizeof(double) * NB, bar);
        rd_bulk_cp(dst + FG_T, AR(st) + NB * ss, sizeof(double) * NB, bar);
        rd_bulk_cp(dst + FG_RD, AR(res_d) + NB * ss, sizeof(double) * NB, bar);
        rd_bulk_cp(dst + FG_RM, AR(res_m) + NB * ss, sizeof(double) * NB, bar);
        rd_bulk_cp(dst + FG_RG, AR(res_g) + NV * ss, sizeof(double) * NV, bar);
        rd_bulk_cp(dst + FG_RB, AR(res_b) + NX * ss, sizeof(double) * NX, bar);
        if (ss > 0) rd_bulk_cp(dst + FG_JC, AR(Jc) + NJ * ss, sizeof(double) * NJ, bar);
    }
}

/* Shape of the stage body: every section LOADS its operands into registers first, computes, and STORES last, and the lanes
 * run one straight-line instruction stream (roles differ by coefficients, ca / cb / jcol, not by branches) -- shared-memory
 * stores in the middle of a section and divergent blocks both stop the compiler from batching the loads, and a stage then
 * costs one shared-memory round trip per operand (measured: 8650 cycles per stage for one warp alone, profiles/r2k_*).
 * Three group barriers per stage:
 *   S1  [P+, p+ of the previous stage, staged operands]  barrier terms, P+ b, Muu row, m_u, row of [B A]' P+ [B A] + H  -> Min
 *   S2  own Mxx row -> P (the row of P+ is dead now); 6 x 6 Cholesky in every lane; Lxu row, p                        -> Lxu, pv
 *   S3  own row: P = Mxx - Lxu Lxu'                                                                                  -> P, F  */
RDN void ipm_factor(const RmpcDims &D, const Inst &I_, GrpSmem &sm)
{
    const Inst I = I_; /* by value: registers, not re-read from the caller's frame after every store */
    const int N = D.N, lane = rd_lane();
    const double *par = I.par;
    const Role R = make_role(par);
    const int j = R.j, xk = R.xk;
    const StageConst sc = stage_const(par, D.dt, j);
    const double reg = 1e-15, tau_min = 1e-16;
    double *P = FS_P(sm), *Lxu = FS_LXU(sm), *Min = FS_MIN(sm);
    const int slot = lane < 12 ? lane : lane - 12; /* bound slot of the lane (lanes 12.. mirror lanes 0.., results unused) */
#ifdef RMPC_BULK_STAGING
    rd_sync();
    if (lane == 0) { rd_bar_init(&sm.bar[0]); rd_bar_init(&sm.bar[1]); rd_fence_bar_init(); }
    rd_sync();
    unsigned phase = 0; /* bit b: parity of the next completion of buffer b's barrier */
    fac_stage_in_bulk(I, FS_STG(sm, (N - 1) & 1), N - 1, &sm.bar[(N - 1) & 1]);
#else
    fac_stage_in(I, FS_STG(sm, (N - 1) & 1), N - 1);
    rd_cp_commit();
#endif
    /* stage N: zero Hessian + reg => P_N = reg I, p_N = res_g_N  (POTRF_L_MN of :871-879) */
    if (R.x)
    {
        double *FN = AR(F) + (long) FS * N;
        double pv = AR(res_g)[NV * N + 6 + xk];
        double Prow[12];
        RD_UNROLL
        for (int c = 0; c < 12; c++) Prow[c] = (c == xk) ? reg : 0.0;
        if (I.pe)
        { /* general row of stage N (x_ocp_qp_kkt.c:859-868): P_N += C' Gamma C, p_N += C' gamma */
            const double *cv = AR(Jc) + NJ * N;
            const int ka = TERM_LO(N), kb = TERM_UP(N);
            const double til = 1.0 / AR(st)[ka], tiu = 1.0 / AR(st)[kb];
            const double la = AR(slam)[ka], lb = AR(slam)[kb];
            const double Ga = tiu * lb + til * la;
            const double ga = til * ((AR(res_m)[ka] - tau_min) - la * AR(res_d)[ka]) - tiu * ((AR(res_m)[kb] - tau_min) - lb * AR(res_d)[kb]);
            const double cg = cv[xk] * Ga;
            RD_UNROLL
            for (int c = 0; c < 12; c++) Prow[c] += cg * cv[c];
            pv += ga * cv[xk];
        }
        RD_UNROLL
        for (int c = 0; c < 12; c++) P[xk * SM_PS + c] = Prow[c];
        sm.pv[xk] = pv;
        RD_UNROLL
        for (int c = 0; c < 12; c++)
            if (c <= xk) FN[RF_P + (xk * (xk + 1)) / 2 + c] = Prow[c];
        FN[RF_PV + xk] = pv;
    }
    SECPROF_BEGIN;
    for (int ss = N - 1; ss >= 0; ss--)
    {
        const bool s0 = (ss == 0);
        double *F = AR(F) + (long) FS * ss;
        const double *G = FS_STG(sm, ss & 1);
#ifdef RMPC_BULK_STAGING
        if (ss > 0) fac_stage_in_bulk(I, FS_STG(sm, (ss - 1) & 1), ss - 1, &sm.bar[(ss - 1) & 1]);
        SEC(0);
        rd_bar_wait(&sm.bar[ss & 1], (phase >> (ss & 1)) & 1u);
        phase ^= 1u << (ss & 1);
#else
        if (ss > 0) fac_stage_in(I, FS_STG(sm, (ss - 1) & 1), ss - 1);
        rd_cp_commit();
        SEC(0);
        rd_cp_wait1();
#endif
        rd_sync();
        SEC(1);
        /* ------------------------------------------------------------------------------------------------ S1 */
        double pj[12], pd[12]; /* rows q_j and qd_j of P+ */
        RD_UNROLL
        for (int c = 0; c < 12; c += 2)
        {
            const rd_double2 a2 = *RD_CV2(&P[j * SM_PS + c]), b2v = *RD_CV2(&P[(6 + j) * SM_PS + c]);
            pj[c] = a2.x; pj[c + 1] = a2.y; pd[c] = b2v.x; pd[c + 1] = b2v.y;
        }
        const double pvq = sm.pv[j], pvd = sm.pv[6 + j];
        /* (1) barrier terms of the lane's bound slot (slot l: u_l for l < 6, q_{l-6} for 6 <= l < 12) */
        const bool act = lane < 12 && !(s0 && lane >= 6);
        double Gam, gam;
        {
            const double la = G[FG_LAM + slot], lb = G[FG_LAM + 12 + slot], ta = G[FG_T + slot], tb = G[FG_T + 12 + slot];
            const double da = G[FG_RD + slot], db = G[FG_RD + 12 + slot], ma = G[FG_RM + slot], mb = G[FG_RM + 12 + slot];
            const double til = 1.0 / (act ? ta : 1.0), tiu = 1.0 / (act ? tb : 1.0);
            const double Gv = til * la + tiu * lb;
            const double gv = til * ((ma - tau_min) - la * da) - tiu * ((mb - tau_min) - lb * db);
            Gam = act ? Gv : 0.0;
            gam = act ? gv : 0.0;
        }
        /* (2) tv = P+ b + p+ for both rows of the joint; P+ b of the own row is kept for the corrector solves */
        double aq = 0.0, ad = 0.0;
        RD_UNROLL
        for (int c = 0; c < 12; c += 2)
        {
            const rd_double2 b2v = *RD_CV2(&G[FG_RB + c]);
            aq += pj[c] * b2v.x;
            aq += pj[c + 1] * b2v.y;
            ad += pd[c] * b2v.x;
            ad += pd[c + 1] * b2v.y;
        }
        const double tvq = aq + pvq, tvd = ad + pvd;
        const double b1j = C_B1(j), b2j = C_B2(j);
        /* (3a) control row: Muu row j and m_u */
        double mrow[6];
        RD_UNROLL
        for (int c = 0; c < 6; c++)
        {
            const double eq = b1j * pj[c] + b2j * pd[c];
            const double ed = b1j * pj[6 + c] + b2j * pd[6 + c];
            double row = C_B1(c) * eq + C_B2(c) * ed;
            if (c == j) row += reg + Gam + sc.hu;
            mrow[c] = row;
        }
        const double mu_ = G[FG_RG + j] + gam + (b1j * tvq + b2j * tvd);
        /* (3c') state row: e = row of [B A]' P+, v = e B (+ S), Mxx row = e A + H, right-hand side */
        double v[6], Pn[12], pk = 0.0;
        if (!s0)
        {
            const double *Jc = G + FG_JC;
            double e[12], jcol[5];
            RD_UNROLL
            for (int c = 0; c < 12; c++) e[c] = R.ca * pj[c] + R.cb * pd[c];
            RD_UNROLL
            for (int m = 0; m < 5; m++) jcol[m] = R.q ? Jc[6 * m + j] : (m == 4 ? Jc[30 + j] : 0.0);
            RD_UNROLL
            for (int c = 0; c < 6; c++)
            {
                double vv = C_B1(c) * e[c] + C_B2(c) * e[6 + c];
                if (!R.q && c == j) vv += sc.hs;
                v[c] = vv;
                Pn[c] = e[c];
                Pn[6 + c] = C_A12(c) * e[c] + C_A22(c) * e[6 + c];
            }
            RD_UNROLL
            for (int c = 0; c < 6; c++)
            {
                double a = 0.0;
                RD_UNROLL
                for (int m = 0; m < 5; m++) a += jcol[m] * Jc[6 * m + c];
                Pn[c] += a;
                Pn[6 + c] += jcol[4] * Jc[30 + c];
            }
            pk = G[FG_RG + 6 + xk] + (R.q ? gam : 0.0) + (R.ca * tvq + R.cb * tvd);
        }
        if (R.x) F[RF_PB + xk] = R.q ? aq : ad;
        if (R.u)
        {
            RD_UNROLL
            for (int c = 0; c < 6; c++)
                if (c <= j) Min[(j * (j + 1)) / 2 + c] = mrow[c];
            Min[21 + j] = mu_;
        }
        SEC(2);
        rd_sync();
        SEC(3);
        /* ------------------------------------------------------------------------------------------------ S2 */
        if (R.x && !s0)
        { /* every lane has read P+: the own row now holds the Mxx row (diagonal: + reg + Gamma | hq) */
            RD_UNROLL
            for (int c = 0; c < 12; c += 2)
            {
                rd_double2 w2;
                w2.x = Pn[c]; w2.y = Pn[c + 1];
                *RD_V2(&P[xk * SM_PS + c]) = w2;
            }
            P[xk * SM_PS + xk] += reg + (R.q ? Gam : sc.hq);
        }
        /* (3b) 6 x 6 Cholesky + forward substitution of m_u, redundantly in every lane (BLASFEO dpotrf_l: a
         *      non-positive pivot zeroes its column) */
        double Lu[21], lu[6], inv[6];
        {
            double mv[28];
            RD_UNROLL
            for (int k = 0; k < 28; k += 2)
            {
                const rd_double2 m2 = *RD_CV2(&Min[k]);
                mv[k] = m2.x; mv[k + 1] = m2.y;
            }
            RD_UNROLL
            for (int k = 0; k < 21; k++) Lu[k] = mv[k];
            RD_UNROLL
            for (int k = 0; k < 6; k++) lu[k] = mv[21 + k];
        }
#define LU(r, c) Lu[((r) * ((r) + 1)) / 2 + (c)]
        RD_UNROLL
        for (int c = 0; c < 6; c++)
        {
            const double dj = LU(c, c);
            const double iv = dj > 0.0 ? rd_rsqrt(dj) : 0.0;
            inv[c] = iv;
            LU(c, c) = dj * iv;
            RD_UNROLL
            for (int r = c + 1; r < 6; r++) LU(r, c) *= iv;
            lu[c] *= iv;
            RD_UNROLL
            for (int k = c + 1; k < 6; k++)
            {
                RD_UNROLL
                for (int r = k; r < 6; r++) LU(r, k) -= LU(r, c) * LU(k, c);
                lu[k] -= lu[c] * LU(k, c);
            }
        }
        if (lane == 12 % GL)
        { /* one lane publishes 1 / diag, l_u and the strict lower part of Luu: F[RF_INV .. RF_LXU) with 128-bit stores */
            double o[28];
            RD_UNROLL
            for (int k = 0; k < 6; k++) { o[RF_INV + k] = inv[k] > 0.0 ? inv[k] : rd_inf(); o[RF_LU + k] = lu[k]; }
            RD_UNROLL
            for (int r = 1; r < 6; r++)
                RD_UNROLL
                for (int c = 0; c < r; c++) o[RF_LUU + (r * (r - 1)) / 2 + c] = LU(r, c);
            o[27] = 0.0;
            RD_UNROLL
            for (int k = 0; k < 28; k += 2)
            {
                rd_double2 w2;
                w2.x = o[k]; w2.y = o[k + 1];
                *RD_V2(&F[k]) = w2;
            }
        }
        SEC(4);
        if (s0) continue;
        /* (3c) Lxu row = v Luu^-T, p = m_x - Lxu l_u */
        double Lx[6];
        RD_UNROLL
        for (int c = 0; c < 6; c++)
        {
            double vv = v[c];
            RD_UNROLL
            for (int k = 0; k < c; k++) vv -= Lx[k] * LU(c, k);
            Lx[c] = vv * inv[c];
        }
        RD_UNROLL
        for (int c = 0; c < 6; c++) pk -= Lx[c] * lu[c];
#undef LU
        if (R.x)
        {
            RD_UNROLL
            for (int c = 0; c < 6; c += 2)
            {
                rd_double2 w2;
                w2.x = Lx[c]; w2.y = Lx[c + 1];
                *RD_V2(&Lxu[xk * 6 + c]) = w2;
                *RD_V2(&F[RF_LXU + xk * 6 + c]) = w2;
            }
            sm.pv[xk] = pk;
            F[RF_PV + xk] = pk;
        }
        SEC(5);
        rd_sync();
        SEC(6);
        /* ------------------------------------------------------------------------------------------------ S3
         * Schur complement row: P[xk][:] = Mxx[xk][:] - Lxu[xk][:] Lxu'; the rows of Lxu are fetched three at a time */
        if (R.x)
        {
            double Pr[12];
            RD_UNROLL
            for (int c = 0; c < 12; c += 2)
            {
                const rd_double2 a2 = *RD_CV2(&P[xk * SM_PS + c]);
                Pr[c] = a2.x; Pr[c + 1] = a2.y;
            }
            RD_UNROLL
            for (int c0 = 0; c0 < 12; c0 += 3)
            {
                rd_double2 l2[3][3];
                RD_UNROLL
                for (int u = 0; u < 3; u++)
                    RD_UNROLL
                    for (int k = 0; k < 3; k++) l2[u][k] = *RD_CV2(&Lxu[(c0 + u) * 6 + 2 * k]);
                RD_UNROLL
                for (int u = 0; u < 3; u++)
                {
                    double a = Pr[c0 + u];
                    RD_UNROLL
                    for (int k = 0; k < 3; k++)
                    {
                        a -= Lx[2 * k] * l2[u][k].x;
                        a -= Lx[2 * k + 1] * l2[u][k].y;
                    }
                    Pr[c0 + u] = a;
                }
            }
            RD_UNROLL
            for (int c = 0; c < 12; c += 2)
            {
                rd_double2 w2;
                w2.x = Pr[c]; w2.y = Pr[c + 1];
                *RD_V2(&P[xk * SM_PS + c]) = w2;
            }
            RD_UNROLL
            for (int c = 0; c < 12; c++)
                if (c <= xk) F[RF_P + (xk * (xk + 1)) / 2 + c] = Pr[c];
        }
        SEC(7);
    }
#ifdef RMPC_BULK_STAGING
    if (lane == 0) { rd_bar_inval(&sm.bar[0]); rd_bar_inval(&sm.bar[1]); }
#else
    rd_cp_wait0();
#endif
    rd_sync(); /* the factor blocks written by single lanes (stage 0: lane 12) are read by every lane of the next sweep */
    SECPROF_END(0);
}

/* Householder-LQ re-factorisation of one interior-point iteration, OCP_QP_FACT_LQ_SOLVE_KKT_STEP
 * (H/ocp_qp/x_ocp_qp_kkt.c:1201-1541), the fall-back of lq_fact = 1 when the Cholesky route left a linear-system residual
 * above 1e-5 (H/ocp_qp/x_ocp_qp_ipm.c:2246-2276).  The stage factor L (lower, L L' = the Riccati matrix of the stage) is
 * obtained WITHOUT forming that matrix -- no subtraction Mxx - Lxu Lxu', no squared condition number: it is the triangular
 * factor of the LQ decomposition of
 *     [ Lh | diag(sqrt(Gamma_box)) | C' sqrt(Gamma_g) (stage N) | [B A]' Lxx+ ],     Lh Lh' = H + reg I   (:1263-1320),
 * with the diagonal made positive.  BLASFEO's GELQF_PD_LLA is a blocked, structure-exploiting Householder LQ; this is the
 * textbook unblocked one in the operation order of oracle/rmpc_oracle.c (lq_lower) -- same factor up to rounding.
 * What the sweeps with stored factors need is then written in the usual factor block: 1 / diag Luu, Luu, Lxu, P = Lxx Lxx'
 * (no cancellation in that product) and P+ b; the step itself follows from ipm_backward<0, 1> + ipm_forward<1> (:1322,
 * 1469-1512).  A rare path (active terminal row): written for clarity, rows of the stage matrix are owned by
 * lanes (row r: lane r mod GL), operands come straight from HBM, one barrier per Householder step.
 * Shared memory: T = the Lh block, packed lower (column i is only touched by reflector i and becomes column i of L);
 * W = the other blocks, nv x 24, [box 0..11 | BA 12..23] (stage N: column 0 = the general row); LN = Lxx of stage ss + 1. */
#define LQ_T(sm) ((sm).buf)
#define LQ_KW 24
#define LQ_W(sm) ((sm).buf + 172)
#define LQ_LN(sm) ((sm).buf + 172 + 18 * LQ_KW)
#define LQ_D(sm) ((sm).buf + 172 + 18 * LQ_KW + 78) /* 18: positive diagonal of L */
#define TL(r, c) T[((r) * ((r) + 1)) / 2 + (c)]
RDN void ipm_factor_lq(const RmpcDims &D, const Inst &I_, GrpSmem &sm)
{
    const Inst I = I_;
    const int N = D.N, lane = rd_lane();
    const double reg = 1e-15;
    double *T = LQ_T(sm), *W = LQ_W(sm), *LN = LQ_LN(sm), *Dg = LQ_D(sm);
    for (int ss = N; ss >= 0; ss--)
    {
        const int nu = ss < N ? 6 : 0, nx = ss > 0 ? 12 : 0, nv = nu + nx, nb = ss == N ? 0 : (ss == 0 ? 6 : 12);
        const double *Jc = AR(Jc) + NJ * ss;
        double *F = AR(F) + (long) FS * ss;
        /* (a) rows of the compound matrix */
        for (int r = lane; r < nv; r += GL)
        {
            const int vr = r + (6 - nu); /* index in [u; q; qd] */
            for (int c = 0; c <= r; c++)
            {
                const int vc = c + (6 - nu);
                double h = 0.0;
                if (ss < N)
                {
                    if (vr < 6) h = (vc == vr) ? sm.hc[vr] : 0.0;
                    else if (ss > 0)
                    {
                        if (vr < 12)
                        {
                            if (vc >= 6)
                                for (int m = 0; m < 5; m++) h += Jc[6 * m + (vr - 6)] * Jc[6 * m + (vc - 6)];
                        }
                        else if (vc < 6) h = (vc == vr - 12) ? sm.hc[6 + vc] : 0.0;
                        else if (vc < 12) h = Jc[30 + (vr - 12)] * Jc[24 + (vc - 6)];
                        else h = Jc[30 + (vr - 12)] * Jc[30 + (vc - 12)] + (vc == vr ? sm.hc[12 + (vr - 12)] : 0.0);
                    }
                }
                TL(r, c) = h + (c == r ? reg : 0.0);
            }
            double *w = W + LQ_KW * r;
            for (int k = 0; k < LQ_KW; k++) w[k] = 0.0;
            if (r < nb)
            { /* box rows, idxb = 0 .. nb-1: sqrt(Gamma_lb + Gamma_ub) on the diagonal of the block */
                const double tl = 1.0 / AR(st)[NB * ss + r], tu = 1.0 / AR(st)[NB * ss + 12 + r];
                double g = tl * AR(slam)[NB * ss + r] + tu * AR(slam)[NB * ss + 12 + r];
                g = g >= 0.0 ? g : 0.0;
                g = sqrt(g);
                w[r] = g > 0.0 ? g : 0.0;
            }
            if (ss == N && I.pe)
            { /* general row of stage N: DCt sqrt(Gamma_g) */
                const int ka = TERM_LO(N), kb = TERM_UP(N);
                double g = (1.0 / AR(st)[ka]) * AR(slam)[ka] + (1.0 / AR(st)[kb]) * AR(slam)[kb];
                g = g >= 0.0 ? g : 0.0;
                w[0] = Jc[r] * sqrt(g);
            }
            if (ss < N)
            { /* TRMM_RLNN: [B A]' Lxx+, rows of [B A]' = ca e_q + cb e_qd */
                const int jj = vr % 6;
                const double ca = vr < 6 ? C_B1(jj) : (vr < 12 ? 1.0 : C_A12(jj)), cb = vr < 6 ? C_B2(jj) : (vr < 12 ? 0.0 : C_A22(jj));
                for (int c = 0; c < 12; c++)
                {
                    double a = 0.0;
                    if (jj >= c) a += ca * LN[(jj * (jj + 1)) / 2 + c];
                    if ((vr < 6 || vr >= 12) && 6 + jj >= c) a += cb * LN[((6 + jj) * (7 + jj)) / 2 + c];
                    w[12 + c] = a;
                }
            }
        }
        /* P+ b = Lxx+ (Lxx+' res_b), kept for the corrector solves */
        if (ss < N && lane < NX)
        {
            const double *b = AR(res_b) + NX * ss;
            double tmp[NX];
            for (int c = 0; c < NX; c++)
            {
                double a = 0.0;
                for (int k = c; k < NX; k++) a += LN[(k * (k + 1)) / 2 + c] * b[k];
                tmp[c] = a;
            }
            double a = 0.0;
            for (int k = 0; k <= lane; k++) a += LN[(lane * (lane + 1)) / 2 + k] * tmp[k];
            F[RF_PB + lane] = a;
        }
        rd_sync();
        /* (b) Lh = chol(H + reg I) in place, column by column (POTRF_L: a non-positive pivot zeroes its column) */
        for (int jc = 0; jc < nv; jc++)
        {
            double dj = TL(jc, jc);
            for (int k = 0; k < jc; k++) dj -= TL(jc, k) * TL(jc, k);
            double inv;
            if (dj > 0.0) { dj = sqrt(dj); inv = 1.0 / dj; }
            else { dj = 0.0; inv = 0.0; }
            double a0 = 0.0, a1 = 0.0;
            const int r0 = lane, r1 = lane + GL;
            if (r0 > jc && r0 < nv)
            {
                a0 = TL(r0, jc);
                for (int k = 0; k < jc; k++) a0 -= TL(r0, k) * TL(jc, k);
            }
            if (r1 > jc && r1 < nv)
            {
                a1 = TL(r1, jc);
                for (int k = 0; k < jc; k++) a1 -= TL(r1, k) * TL(jc, k);
            }
            rd_sync();
            if (r0 == jc) TL(jc, jc) = dj;
            if (r0 > jc && r0 < nv) TL(r0, jc) = a0 * inv;
            if (r1 == jc) TL(jc, jc) = dj;
            if (r1 > jc && r1 < nv) TL(r1, jc) = a1 * inv;
            rd_sync();
        }
        /* (c) Householder reflectors, row by row */
        for (int i = 0; i < nv; i++)
        {
            const double *wi = W + LQ_KW * i;
            const double d0 = TL(i, i);
            double sigma = 0.0;
            for (int k = 0; k < LQ_KW; k++) sigma += wi[k] * wi[k];
            double diag = d0, sg = 1.0;
            if (sigma > 0.0)
            {
                const double mu = sqrt(d0 * d0 + sigma);
                const double v0 = d0 <= 0.0 ? d0 - mu : -sigma / (d0 + mu);
                const double beta = 2.0 * v0 * v0 / (sigma + v0 * v0);
                for (int r = lane; r < nv; r += GL)
                    if (r > i)
                    {
                        double *wr = W + LQ_KW * r;
                        double w = TL(r, i);
                        for (int k = 0; k < LQ_KW; k++) w += wr[k] * (wi[k] / v0);
                        w *= beta;
                        TL(r, i) -= w;
                        for (int k = 0; k < LQ_KW; k++) wr[k] -= w * (wi[k] / v0);
                    }
                diag = mu;
            }
            else if (d0 < 0.0)
            { /* untouched row with a negative diagonal: the column changes sign */
                sg = -1.0;
                for (int r = lane; r < nv; r += GL)
                    if (r > i) TL(r, i) = -TL(r, i);
            }
            if (lane == i % GL) Dg[i] = sg * diag;
            rd_sync();
        }
        /* (d) factor block of the stage: 1 / diag Luu, Luu, Lxu, P = Lxx Lxx' */
        if (nu)
        {
            if (lane < 6)
            {
                const double dd = Dg[lane];
                F[RF_INV + lane] = dd > 0.0 ? 1.0 / dd : rd_inf();
                F[RF_LU + lane] = 0.0;
                for (int c = 0; c < lane; c++) F[RF_LUU + (lane * (lane - 1)) / 2 + c] = TL(lane, c);
            }
            if (nx && lane < NX)
                for (int c = 0; c < 6; c++) F[RF_LXU + lane * 6 + c] = TL(6 + lane, c);
        }
        if (nx && lane < NX)
        {
            const int r = lane;
            for (int c = 0; c <= r; c++)
            {
                double a = 0.0;
                for (int k = 0; k <= c; k++)
                    a += (k == r ? Dg[nu + r] : TL(nu + r, nu + k)) * (k == c ? Dg[nu + c] : TL(nu + c, nu + k));
                F[RF_P + (r * (r + 1)) / 2 + c] = a;
            }
            F[RF_PV + r] = 0.0;
        }
        rd_sync();
        if (nx && lane < NX) /* Lxx of this stage for the next one */
            for (int c = 0; c <= lane; c++) LN[(lane * (lane + 1)) / 2 + c] = (c == lane) ? Dg[nu + lane] : TL(nu + lane, nu + c);
        rd_sync();
    }
}
#undef TL

/* Forward sweep of the Riccati recursion (H/ocp_qp/x_ocp_qp_kkt.c:970-1005 and :1671-1721) fused with everything that
 * only depends on the step of one stage:
 *   du = -Luu^-T (src_u + Lxu' dx),  dx+ = A dx + B du + rb,  dpi = P+ dx+ + p+,
 *   dt = +-dv - rd, dlam = -(rm + lam dv -+ lam rd) / t        (COMPUTE_LAM_T_QP, H/ipm_core/x_core_qp_ipm_aux.c:164-189)
 *   alpha = largest step in (0,1] keeping lam, t >= 0            (COMPUTE_ALPHA_QP, :193-263, split_step 0)
 *   sums[0] = sum lam dt + dlam t, sums[1] = sum dlam dt          (for COMPUTE_MU_AFF_QP, :424-441, see ipm_mu_aff)
 * STORED = 0: src_u = l_u and p+ from the factorisation (affine step); STORED = 1: src_u = z_u left in X[.][0:6] by
 * ipm_backward() and p+ = pvec.  rm enters as rm - mshift. */
/* staging layout of one stage for the forward sweep: [0,100) = F[ss] (1/diag, l_u, Luu, Lxu), [100,190) = P, p of
 * F[ss+1] at their RF_* offsets, then the bound-slot data, res_b and (STORED) z_u, p of the corrector */
enum { WG_LAM = 190, WG_T = 214, WG_RD = 238, WG_RM = 262, WG_RB = 286, WG_ZU = 298, WG_PVEC = 304, WG_RG = 316, WG_JC = 334 };
template <int STORED>
RD void fwd_stage_in(const Inst &I, double *dst, int ss, const double *rg, const double *rb, const double *rd, const double *rm,
                     const double *X)
{
    stage_cp(dst + WG_RG, rg + NV * ss, NV);
    if (ss > 0) stage_cp(dst + WG_JC, AR(Jc) + NJ * ss, NJ);
    const double *F = AR(F) + (long) FS * ss;
    stage_cp(dst, F, RF_P);
    stage_cp(dst + RF_P, F + FS + RF_P, STORED ? RF_PV - RF_P : RF_PB - RF_P);
    stage_cp(dst + WG_LAM, AR(slam) + NB * ss, NB);
    stage_cp(dst + WG_T, AR(st) + NB * ss, NB);
    stage_cp(dst + WG_RD, rd + NB * ss, NB);
    stage_cp(dst + WG_RM, rm + NB * ss, NB);
    stage_cp(dst + WG_RB, rb + NX * ss, NX);
    if (STORED)
    {
        stage_cp(dst + WG_ZU, X + NV * ss, 6);
        stage_cp(dst + WG_PVEC, AR(pvec) + NX * (ss + 1), NX);
    }
}

/* The sweep also evaluates, at no extra pass over the data, the stationarity residual of the LINEAR system it solves,
 *   r = rg + H dv + [B A]' dpi - [0; dpi-] + dlam_ub - dlam_lb          (OCP_QP_RES_COMPUTE_LIN, H/ocp_qp/x_ocp_qp_res.c:535-685)
 * whose inf-norm drives the LQ-fallback trigger and the iterative refinement (x_ocp_qp_ipm.c:2228-2276,2493-2594):
 * lin[0] = |r|_inf, lin[1] = NaN flag of r (any row: the max skips NaNs like HPIPM's norm does).  The other three residuals of that system (dynamics,
 * bounds, complementarity) are zero up to rounding BY CONSTRUCTION of dx+, dt and dlam in this sweep and are not
 * re-evaluated. */
template <int STORED>
RDN void ipm_forward(const RmpcDims &D, const Inst &I_, GrpSmem &sm, const double *rg, const double *rb, const double *rd,
                    const double *rm, double mshift, double *X, double *PI, double *DL, double *DT, double *alpha_out,
                    double *sums, double *lin)
{
    const Inst I = I_; /* by value: registers, not re-read from the caller's frame after every store */
    const int N = D.N, lane = rd_lane();
    const Role R = make_role(I.par);
    const int j = R.j, xk = R.xk;
    const double a12 = C_A12(j), a22 = C_A22(j), b1 = C_B1(j), b2 = C_B2(j);
    const double hu = sm.hc[j], hs = sm.hc[6 + j], hq = sm.hc[12 + j];
    fwd_stage_in<STORED>(I, SW_STG(sm, 0), 0, rg, rb, rd, rm, X);
    rd_cp_commit();
    if (R.x) { sm.xv[xk] = 0.0; X[6 + xk] = 0.0; }
    double alpha = 1.0, sB = 0.0, sC = 0.0;
    double n0 = 0.0, nanflag = 0.0, dp_prev = 0.0;
    SECPROF_BEGIN;
    for (int ss = 0; ss < N; ss++)
    {
        const double *G = SW_STG(sm, ss & 1);
        if (ss + 1 < N) fwd_stage_in<STORED>(I, SW_STG(sm, (ss + 1) & 1), ss + 1, rg, rb, rd, rm, X);
        rd_cp_commit();
        SEC(0);
        rd_cp_wait1();
        rd_sync();
        SEC(1);
        /* s = src_u + Lxu' dx (lane c < 6) */
        if (R.u)
        {
            double s = STORED ? G[WG_ZU + lane] : G[RF_LU + lane];
            if (ss > 0)
            {
                RD_UNROLL
                for (int k = 0; k < 12; k++) s += G[RF_LXU + k * 6 + lane] * sm.xv[k];
            }
            sm.uv[lane] = s;
        }
        /* w = Jc dx of this stage (lane m < 5), for the residual check */
        double wm = 0.0;
        if (ss > 0 && lane < 5)
        {
            RD_UNROLL
            for (int k = 0; k < 6; k++) wm += G[WG_JC + 6 * lane + k] * sm.xv[k];
            if (lane == 4)
            {
                RD_UNROLL
                for (int k = 0; k < 6; k++) wm += G[WG_JC + 30 + k] * sm.xv[6 + k];
            }
        }
        rd_sync();
        SEC(2);
        /* du = -Luu^-T s, redundantly in every lane */
        double du[6];
        RD_UNROLL
        for (int k = 0; k < 6; k++) du[k] = -sm.uv[k];
        RD_UNROLL
        for (int k = 5; k >= 0; k--)
        {
            du[k] *= G[RF_INV + k];
            RD_UNROLL
            for (int c = 0; c < k; c++) du[c] -= G[RF_LUU + (k * (k - 1)) / 2 + c] * du[k];
        }
        double duj = du[0], dul = du[0];
        RD_UNROLL
        for (int k = 1; k < 6; k++)
        {
            if (k == j) duj = du[k];
            if (k == lane) dul = du[k];
        }
        /* step of the variable of this lane's bound slot: u_l or q_{l-6} of THIS stage */
        const double dv = (lane < 6) ? dul : (lane < 12 ? sm.xv[lane - 6] : 0.0);
        const double xo = R.x ? sm.xv[xk] : 0.0; /* own state row of THIS stage */
        double xn = 0.0;
        if (R.x)
        {
            if (R.q) xn = (sm.xv[j] + a12 * sm.xv[6 + j]) + b1 * duj + G[WG_RB + j];
            else xn = a22 * sm.xv[6 + j] + b2 * duj + G[WG_RB + 6 + j];
        }
        SEC(3);
        rd_sync(); /* every lane has read xv, uv */
        if (R.u) X[NV * ss + lane] = dul;
        if (R.x) { sm.xv[xk] = xn; X[NV * (ss + 1) + 6 + xk] = xn; }
        if (lane < 5) sm.wv[lane] = wm;
        rd_sync();
        SEC(4);
        /* dpi = P+ dx+ + p+ */
        double dpo = 0.0;
        if (R.x)
        {
            double a = STORED ? G[WG_PVEC + xk] : G[RF_PV + xk];
            RD_UNROLL
            for (int c = 0; c < 12; c++) a += psym(G + RF_P, xk, c) * sm.xv[c];
            PI[NX * ss + xk] = a;
            sm.dp[xk] = a;
            dpo = a;
        }
        SEC(5);
        /* multiplier / slack step of this lane's bound slot, step length, second-order sums */
        double dlm = 0.0; /* dlam_ub - dlam_lb of the own slot */
        if (lane < 12 && !(ss == 0 && lane >= 6))
        {
            const int ia = NB * ss + lane, ib = ia + 12;
            const double la = G[WG_LAM + lane], lb = G[WG_LAM + 12 + lane], ta = G[WG_T + lane], tb = G[WG_T + 12 + lane];
            const double da = G[WG_RD + lane], db = G[WG_RD + 12 + lane], ma = G[WG_RM + lane] - mshift, mb = G[WG_RM + 12 + lane] - mshift;
            const double dla = -(ma + (la * dv) - (la * da)) / ta;
            const double dlb = -(mb + (lb * (-dv)) - (lb * db)) / tb;
            const double dta = dv - da, dtb = -dv - db;
            DL[ia] = dla; DL[ib] = dlb; DT[ia] = dta; DT[ib] = dtb;
            if (-alpha * dla > la) alpha = -la / dla;
            if (-alpha * dta > ta) alpha = -ta / dta;
            if (-alpha * dlb > lb) alpha = -lb / dlb;
            if (-alpha * dtb > tb) alpha = -tb / dtb;
            sB += (la * dta + dla * ta) + (lb * dtb + dlb * tb);
            sC += dla * dta + dlb * dtb;
            dlm = dlb - dla;
        }
        SEC(6);
        /* operands of the residual rows that live in the staging buffer (refilled after the barrier) */
        double rgu = 0.0, rgx = 0.0, jq[5] = {0, 0, 0, 0, 0}, jd = 0.0;
        if (R.u) rgu = G[WG_RG + lane];
        if (R.x && ss > 0)
        {
            rgx = G[WG_RG + 6 + xk];
            if (R.q)
            {
                RD_UNROLL
                for (int m = 0; m < 5; m++) jq[m] = G[WG_JC + 6 * m + j];
            }
            else jd = G[WG_JC + 30 + j];
        }
        rd_sync(); /* dp, wv published; the staging buffer of this stage is refilled by the next iteration */
        if (R.u)
        {
            const double r = rgu + hu * dul + hs * xo + b1 * sm.dp[j] + b2 * sm.dp[6 + j] + dlm;
            n0 = rd_fmax(n0, fabs(r));
            if (r != r) nanflag = 1.0;
        }
        if (R.x && ss > 0)
        {
            double r;
            if (R.q)
            {
                r = rgx + dlm + sm.dp[j] - dp_prev;
                RD_UNROLL
                for (int m = 0; m < 5; m++) r += jq[m] * sm.wv[m];
            }
            else r = rgx + jd * sm.wv[4] + hs * dul + hq * xo + a12 * sm.dp[j] + a22 * sm.dp[6 + j] - dp_prev;
            n0 = rd_fmax(n0, fabs(r));
            if (r != r) nanflag = 1.0;
        }
        dp_prev = dpo;
        SEC(7);
    }
    rd_cp_wait0();
    SECPROF_END(STORED ? 20 : 10);
    if (R.u) X[NV * N + lane] = 0.0;
    double dlmN = 0.0;
    if (I.pe)
    { /* general row of stage N: dv = C dx_N, then the same multiplier / slack step, step length and sums */
        const double *cv = AR(Jc) + NJ * N;
        const double dv = grp_sum(R.x ? cv[xk] * sm.xv[xk] : 0.0);
        const int ia = TERM_LO(N), ib = TERM_UP(N);
        const double la = AR(slam)[ia], lb = AR(slam)[ib], ta = AR(st)[ia], tb = AR(st)[ib];
        const double da = rd[ia], db = rd[ib], ma = rm[ia] - mshift, mb = rm[ib] - mshift;
        const double dla = -(ma + (la * dv) - (la * da)) / ta;
        const double dlb = -(mb + (lb * (-dv)) - (lb * db)) / tb;
        const double dta = dv - da, dtb = -dv - db;
        if (-alpha * dla > la) alpha = -la / dla;
        if (-alpha * dta > ta) alpha = -ta / dta;
        if (-alpha * dlb > lb) alpha = -lb / dlb;
        if (-alpha * dtb > tb) alpha = -tb / dtb;
        if (lane == 0)
        {
            DL[ia] = dla; DL[ib] = dlb; DT[ia] = dta; DT[ib] = dtb;
            sB += (la * dta + dla * ta) + (lb * dtb + dlb * tb);
            sC += dla * dta + dlb * dtb;
        }
        if (R.x) dlmN = cv[xk] * (dlb - dla);
    }
    if (R.x) n0 = rd_fmax(n0, fabs(rg[NV * N + 6 + xk] + dlmN - dp_prev)); /* stage N: no cost, no box bounds */
    rd_sync();
    *alpha_out = -grp_max(-alpha);
    sums[0] = grp_sum(sB);
    sums[1] = grp_sum(sC);
    lin[0] = grp_max(n0);
    lin[1] = grp_max(nanflag);
}

/* Backward substitution with stored factors, first half of OCP_QP_SOLVE_KKT_STEP (H/ocp_qp/x_ocp_qp_kkt.c:
 * 1543-1665) incl. COMPUTE_GAMMA_QP (x_core_qp_ipm_aux.c:133-160): leaves z_u = Luu^-1 r_u in X[.][0:6] and the
 * Riccati vectors p_ss in pvec.  The complementarity right-hand side depends on MODE:
 *   0  rm - sigma_mu with rm given (iterative refinement: sigma_mu = 0; LQ route: the affine right-hand side res_m - tau_min);
 *   1  COMPUTE_CENTERING_CORRECTION_QP (:479-505): rm = res_m + dt_aff dlam_aff - sigma_mu, written to res_mc;
 *   2  COMPUTE_CENTERING_QP (:509-529):            rm = res_m - sigma_mu, written to res_mc.
 * USE_PB: P+ rb was stored by the factorisation (rb = res_b); otherwise it is recomputed from P+. */
/* staging layout of one stage for the backward substitution: [0,100) = F[ss] (1/diag, -, Luu, Lxu), then P+ b, res_g,
 * lam, t, res_d and the complementarity data of the MODE */
enum { BG_PB = 100, BG_RG = 112, BG_LAM = 130, BG_T = 154, BG_RD = 178, BG_RM = 202, BG_DT = 226, BG_DLAM = 250 };
template <int MODE, int USE_PB>
RD void bwd_stage_in(const Inst &I, double *dst, int ss, const double *rg, const double *rd, const double *rm)
{
    const double *F = AR(F) + (long) FS * ss;
    stage_cp(dst, F, RF_P);
    if (USE_PB) stage_cp(dst + BG_PB, F + RF_PB, NX);
    stage_cp(dst + BG_RG, rg + NV * ss, NV);
    stage_cp(dst + BG_LAM, AR(slam) + NB * ss, NB);
    stage_cp(dst + BG_T, AR(st) + NB * ss, NB);
    stage_cp(dst + BG_RD, rd + NB * ss, NB);
    if (MODE == 0) stage_cp(dst + BG_RM, rm + NB * ss, NB);
    else
    {
        stage_cp(dst + BG_RM, AR(res_m) + NB * ss, NB);
        if (MODE == 1)
        {
            stage_cp(dst + BG_DT, AR(dt_) + NB * ss, NB);
            stage_cp(dst + BG_DLAM, AR(dlam) + NB * ss, NB);
        }
    }
}

template <int MODE, int USE_PB>
RDN void ipm_backward(const RmpcDims &D, const Inst &I_, GrpSmem &sm, const double *rg, const double *rb, const double *rd,
                     const double *rm, double sigma_mu, double *X)
{
    const Inst I = I_; /* by value: registers, not re-read from the caller's frame after every store */
    const int N = D.N, lane = rd_lane();
    const Role R = make_role(I.par);
    const int j = R.j, xk = R.xk;
    const double b1 = C_B1(j), b2 = C_B2(j);
    bwd_stage_in<MODE, USE_PB>(I, SW_STG(sm, (N - 1) & 1), N - 1, rg, rd, rm);
    rd_cp_commit();
    if (R.x)
    {
        double pv = rg[NV * N + 6 + xk];
        if (I.pe)
        { /* general row of stage N: p_N += C' gamma with the complementarity right-hand side of the MODE */
            const int ia = TERM_LO(N), ib = TERM_UP(N);
            const double la = AR(slam)[ia], lb = AR(slam)[ib], ta = AR(st)[ia], tb = AR(st)[ib];
            double ma, mb;
            if (MODE == 0) { ma = rm[ia] - sigma_mu; mb = rm[ib] - sigma_mu; }
            else
            {
                if (MODE == 1)
                {
                    ma = AR(res_m)[ia] + AR(dt_)[ia] * AR(dlam)[ia] - sigma_mu;
                    mb = AR(res_m)[ib] + AR(dt_)[ib] * AR(dlam)[ib] - sigma_mu;
                }
                else { ma = AR(res_m)[ia] - sigma_mu; mb = AR(res_m)[ib] - sigma_mu; }
                if (xk == 0) { AR(res_mc)[ia] = ma; AR(res_mc)[ib] = mb; }
            }
            pv += (AR(Jc) + NJ * N)[xk] * ((ma - la * rd[ia]) / ta - (mb - lb * rd[ib]) / tb);
        }
        sm.pv[xk] = pv;
        AR(pvec)[NX * N + xk] = pv;
    }
    SECPROF_BEGIN;
    for (int ss = N - 1; ss >= 0; ss--)
    {
        const bool s0 = (ss == 0);
        const double *G = SW_STG(sm, ss & 1);
        if (ss > 0) bwd_stage_in<MODE, USE_PB>(I, SW_STG(sm, (ss - 1) & 1), ss - 1, rg, rd, rm);
        rd_cp_commit();
        SEC(0);
        rd_cp_wait1();
        rd_sync();
        SEC(1);
        if (R.x)
        {
            double pb;
            if (USE_PB) pb = G[BG_PB + xk];
            else
            {
                const double *Fn = AR(F) + (long) FS * (ss + 1);
                pb = 0.0;
                RD_UNROLL
                for (int c = 0; c < 12; c++) pb += psym(Fn + RF_P, xk, c) * rb[NX * ss + c];
            }
            sm.tv[xk] = sm.pv[xk] + pb;
        }
        SEC(2);
        double gam = 0.0;
        if (lane < 12 && !(s0 && lane >= 6))
        {
            const int ia = NB * ss + lane, ib = ia + 12;
            const double la = G[BG_LAM + lane], lb = G[BG_LAM + 12 + lane], ta = G[BG_T + lane], tb = G[BG_T + 12 + lane];
            double ma, mb;
            if (MODE == 0) { ma = G[BG_RM + lane] - sigma_mu; mb = G[BG_RM + 12 + lane] - sigma_mu; }
            else
            {
                if (MODE == 1)
                {
                    ma = G[BG_RM + lane] + G[BG_DT + lane] * G[BG_DLAM + lane] - sigma_mu;
                    mb = G[BG_RM + 12 + lane] + G[BG_DT + 12 + lane] * G[BG_DLAM + 12 + lane] - sigma_mu;
                }
                else { ma = G[BG_RM + lane] - sigma_mu; mb = G[BG_RM + 12 + lane] - sigma_mu; }
                AR(res_mc)[ia] = ma; AR(res_mc)[ib] = mb;
            }
            gam = (ma - la * G[BG_RD + lane]) / ta - (mb - lb * G[BG_RD + 12 + lane]) / tb;
        }
        rd_sync();
        SEC(3);
        double rx = 0.0;
        if (R.u) sm.uv[lane] = G[BG_RG + lane] + gam + (b1 * sm.tv[j] + b2 * sm.tv[6 + j]);
        if (R.x && !s0) rx = G[BG_RG + 6 + xk] + (R.q ? gam : 0.0) + (R.ca * sm.tv[j] + R.cb * sm.tv[6 + j]);
        rd_sync();
        SEC(4);
        /* TRSV_LNN_MN(18, 6): z_u = Luu^-1 r_u (redundantly in every lane), r_x -= Lxu z_u */
        double z[6];
        RD_UNROLL
        for (int k = 0; k < 6; k++) z[k] = sm.uv[k];
        RD_UNROLL
        for (int k = 0; k < 6; k++)
        {
            z[k] *= G[RF_INV + k];
            RD_UNROLL
            for (int r = k + 1; r < 6; r++) z[r] -= G[RF_LUU + (r * (r - 1)) / 2 + k] * z[k];
        }
        if (R.u)
        {
            double zl = z[0];
            RD_UNROLL
            for (int k = 1; k < 6; k++)
                if (k == lane) zl = z[k];
            X[NV * ss + lane] = zl;
        }
        if (R.x && !s0)
        {
            RD_UNROLL
            for (int k = 0; k < 6; k++) rx -= G[RF_LXU + xk * 6 + k] * z[k];
            sm.pv[xk] = rx; /* tv, not pv, is what the other lanes read after the last barrier */
            AR(pvec)[NX * ss + xk] = rx;
        }
        rd_sync();
        SEC(5);
    }
    rd_cp_wait0();
    SECPROF_END(30);
}

/* COMPUTE_ALPHA_QP as a separate pass (only needed after an iterative-refinement correction) */
RD double ipm_alpha(const RmpcDims &D, const Inst &I, const double *dlam, const double *dt)
{
    const int N = D.N, lane = rd_lane();
    double a = 1.0;
    for (int k = lane; k < N * NB; k += GL)
    {
        const int i = k / NB, j = k - NB * i;
        if (i == 0 && (j % 12) >= 6) continue;
        const double dl = dlam[k], dtt = dt[k];
        if (-a * dl > AR(slam)[k]) a = -AR(slam)[k] / dl;
        if (-a * dtt > AR(st)[k]) a = -AR(st)[k] / dtt;
    }
    if (I.pe && lane < 2)
    {
        const int k = lane ? TERM_UP(N) : TERM_LO(N);
        const double dl = dlam[k], dtt = dt[k];
        if (-a * dl > AR(slam)[k]) a = -AR(slam)[k] / dl;
        if (-a * dtt > AR(st)[k]) a = -AR(st)[k] / dtt;
    }
    return -grp_max(-a);
}

/* the two sums of ipm_mu_aff as a separate pass (only needed after a refinement of the predictor step) */
RD void ipm_mu_sums(const RmpcDims &D, const Inst &I, const double *dlam, const double *dt, double *sums)
{
    const int N = D.N, lane = rd_lane();
    double sB = 0.0, sC = 0.0;
    for (int k = lane; k < N * NB; k += GL)
    {
        const int i = k / NB, j = k - NB * i;
        if (i == 0 && (j % 12) >= 6) continue;
        sB += AR(slam)[k] * dt[k] + dlam[k] * AR(st)[k];
        sC += dlam[k] * dt[k];
    }
    if (I.pe && lane < 2)
    {
        const int k = lane ? TERM_UP(N) : TERM_LO(N);
        sB += AR(slam)[k] * dt[k] + dlam[k] * AR(st)[k];
        sC += dlam[k] * dt[k];
    }
    sums[0] = grp_sum(sB);
    sums[1] = grp_sum(sC);
}

/* COMPUTE_MU_AFF_QP (x_core_qp_ipm_aux.c:424-441): sum (lam + alpha dlam)(t + alpha dt) / nc, expanded in powers of
 * alpha so that the sums can be accumulated by the forward sweep before alpha is known */
RD double ipm_mu_aff(double sum_lt, const double *sums, double alpha, double nc_inv)
{
    return (sum_lt + alpha * sums[0] + alpha * alpha * sums[1]) * nc_inv;
}

/* UPDATE_VAR_QP + OCP_QP_RES_COMPUTE in ONE stage-by-stage pass with staged operands (the hot pair of every interior-point
 * iteration).
 * UPD = 1: the iterate is first moved, (ux, pi) += a (dux, dpi), (lam, t) = max((lam, t) + a (dlam, dt), 1e-16), written
 * back, and the residuals are evaluated at the new point; UPD = 0: residuals at the current point (first call of a QP).
 * The 16 lanes of the group work on ONE stage at a time with the row roles of the Riccati sweeps (lane l < 6: rows u_l
 * and qd_l, lane 6 + j: row q_j; lane l < 12 also owns bound slot l), all operands of the stage are fetched one stage
 * ahead by cp.async into shared memory -- coalesced 16-byte copies instead of 16 lanes striding through 16 different
 * stages -- and the multiplier of the previous stage is carried in a register.  Returns sum lam t; norms as ipm_res. */
enum { QG_UX = 0, QG_DUX = 18, QG_XN = 36, QG_DXN = 48, QG_RQZ = 60, QG_PI = 78, QG_DPI = 90, QG_LAM = 102, QG_DLAM = 126,
       QG_T = 150, QG_DT = 174, QG_B = 198, QG_D = 210, QG_JC = 234, QG_END = 270 };
template <int UPD>
RD void res_stage_in(const Inst &I, double *dst, int i, int N)
{
    const bool cons = (i < N) || I.pe; /* stage N carries constraint slots only with a terminal row */
    stage_cp(dst + QG_UX, AR(sux) + NV * i, NV);
    stage_cp(dst + QG_RQZ, AR(rqz) + NV * i, NV);
    if (UPD) stage_cp(dst + QG_DUX, AR(dux) + NV * i, NV);
    if (i < N)
    {
        stage_cp(dst + QG_XN, AR(sux) + NV * (i + 1) + 6, NX);
        stage_cp(dst + QG_PI, AR(spi) + NX * i, NX);
        stage_cp(dst + QG_B, AR(b) + NX * i, NX);
        if (UPD)
        {
            stage_cp(dst + QG_DXN, AR(dux) + NV * (i + 1) + 6, NX);
            stage_cp(dst + QG_DPI, AR(dpi) + NX * i, NX);
        }
    }
    if (cons)
    {
        stage_cp(dst + QG_LAM, AR(slam) + NB * i, NB);
        stage_cp(dst + QG_T, AR(st) + NB * i, NB);
        stage_cp(dst + QG_D, AR(d) + NB * i, NB);
        if (UPD)
        {
            stage_cp(dst + QG_DLAM, AR(dlam) + NB * i, NB);
            stage_cp(dst + QG_DT, AR(dt_) + NB * i, NB);
        }
        if (i > 0) stage_cp(dst + QG_JC, AR(Jc) + NJ * i, NJ);
    }
}

template <int UPD>
RDN double ipm_res_sweep(const RmpcDims &D, const Inst &I_, GrpSmem &sm, const double a, double *nrm)
{
    const Inst I = I_; /* by value: registers, not re-read from the caller's frame after every store */
    const int N = D.N, lane = rd_lane();
    const Role R = make_role(I.par);
    const int j = R.j, xk = R.xk;
    const double a12 = C_A12(j), a22 = C_A22(j), b1 = C_B1(j), b2 = C_B2(j);
    const double hu = sm.hc[j], hs = sm.hc[6 + j], hq = sm.hc[12 + j];
    const double lmin = 1e-16;
    double mu = 0.0, n0 = 0, n1 = 0, n2 = 0, n3 = 0, n4 = 0, nanflag = 0, pprev = 0.0;
    res_stage_in<UPD>(I, SW_STG(sm, 0), 0, N);
    rd_cp_commit();
#define QV(off, doff, c) (UPD ? G[(off) + (c)] + a * G[(doff) + (c)] : G[(off) + (c)])
#define QF(off, doff, c, out)                                                                 \
    do                                                                                        \
    {                                                                                         \
        double q_ = G[(off) + (c)];                                                           \
        if (UPD) { q_ = q_ + a * G[(doff) + (c)]; q_ = q_ <= lmin ? lmin : q_; }              \
        (out) = q_;                                                                           \
    } while (0)
    SECPROF_BEGIN;
    for (int i = 0; i <= N; i++)
    {
        const double *G = SW_STG(sm, i & 1);
        if (i < N) res_stage_in<UPD>(I, SW_STG(sm, (i + 1) & 1), i + 1, N);
        rd_cp_commit();
        SEC(0);
        rd_cp_wait1();
        rd_sync();
        SEC(1);
        const bool mid = i < N, x_on = i > 0;
        /* new values of this lane's joint: u_j, q_j, qd_j (the state is fixed at stage 0 of the reduced QP) */
        const double uj = mid ? QV(QG_UX, QG_DUX, j) : 0.0;
        const double qj = x_on ? QV(QG_UX, QG_DUX, 6 + j) : 0.0;
        const double dj = x_on ? QV(QG_UX, QG_DUX, 12 + j) : 0.0;
        double pq = 0.0, pd = 0.0;
        if (mid) { pq = QV(QG_PI, QG_DPI, j); pd = QV(QG_PI, QG_DPI, 6 + j); }
        /* bound slot of this lane */
        double dlm = 0.0;
        if (mid && lane < 12 && !(i == 0 && lane >= 6))
        {
            const double vl = lane < 6 ? uj : qj;
            double la, lb, ta, tb;
            QF(QG_LAM, QG_DLAM, lane, la); QF(QG_LAM, QG_DLAM, 12 + lane, lb);
            QF(QG_T, QG_DT, lane, ta); QF(QG_T, QG_DT, 12 + lane, tb);
            const int ia = NB * i + lane, ib = ia + 12;
            if (UPD) { AR(slam)[ia] = la; AR(slam)[ib] = lb; AR(st)[ia] = ta; AR(st)[ib] = tb; }
            const double dl = G[QG_D + lane] + ta - vl, du = G[QG_D + 12 + lane] + tb + vl;
            const double ml = la * ta, mu_ = lb * tb;
            mu += ml + mu_;
            AR(res_d)[ia] = dl; AR(res_d)[ib] = du; AR(res_m)[ia] = ml; AR(res_m)[ib] = mu_;
            n2 = rd_fmax(n2, rd_fmax(fabs(dl), fabs(du)));
            n3 = rd_fmax(n3, rd_fmax(fabs(ml), fabs(mu_)));
            n4 = rd_fmax(n4, rd_fmax(fabs(ml - 1e-16), fabs(mu_ - 1e-16)));
            dlm = lb - la;
        }
        SEC(2);
        /* control row u_j */
        if (R.u && mid)
        {
            double g = G[QG_RQZ + lane];
            g += hu * uj + hs * dj + b1 * pq + b2 * pd;
            g += dlm;
            AR(res_g)[NV * i + lane] = g;
            n0 = rd_fmax(n0, fabs(g));
            if (i == 0 && lane == 0 && g != g) nanflag = 1.0;
        }
        SEC(3);
        /* state row (q_j on lanes 6..11, qd_j on lanes 0..5), dynamics residual of the same row */
        if (R.x)
        {
            if (x_on)
            {
                double g = G[QG_RQZ + 6 + xk];
                if (mid)
                {
                    if (R.q)
                    {
                        RD_UNROLL
                        for (int m = 0; m < 5; m++)
                        {
                            double w = 0.0;
                            RD_UNROLL
                            for (int k = 0; k < 6; k++) w += G[QG_JC + 6 * m + k] * QV(QG_UX, QG_DUX, 6 + k);
                            if (m == 4)
                            {
                                RD_UNROLL
                                for (int k = 0; k < 6; k++) w += G[QG_JC + 30 + k] * QV(QG_UX, QG_DUX, 12 + k);
                            }
                            g += G[QG_JC + 6 * m + j] * w;
                        }
                        g += pq;
                        g += dlm;
                    }
                    else
                    {
                        double w = 0.0;
                        RD_UNROLL
                        for (int k = 0; k < 6; k++) w += G[QG_JC + 24 + k] * QV(QG_UX, QG_DUX, 6 + k);
                        RD_UNROLL
                        for (int k = 0; k < 6; k++) w += G[QG_JC + 30 + k] * QV(QG_UX, QG_DUX, 12 + k);
                        g += G[QG_JC + 30 + j] * w;
                        g += hs * uj + hq * dj + a12 * pq + a22 * pd;
                    }
                }
                g -= pprev;
                if (!mid && I.pe)
                { /* general row of stage N: res_g += C' (lam_u - lam_l) */
                    double la, lb;
                    QF(QG_LAM, QG_DLAM, 0, la); QF(QG_LAM, QG_DLAM, 12, lb);
                    g += G[QG_JC + xk] * (lb - la);
                }
                AR(res_g)[NV * i + 6 + xk] = g;
                n0 = rd_fmax(n0, fabs(g));
            }
            if (mid)
            {
                const double xn = QV(QG_XN, QG_DXN, xk);
                double r;
                if (R.q) r = G[QG_B + j] - xn + (qj + a12 * dj) + b1 * uj;
                else r = G[QG_B + 6 + j] - xn + a22 * dj + b2 * uj;
                AR(res_b)[NX * i + xk] = r;
                n1 = rd_fmax(n1, fabs(r));
                pprev = R.q ? pq : pd;
            }
        }
        SEC(4);
        if (!mid && I.pe && lane == 12)
        { /* general row of stage N: res_d = d + t -+ C x, res_m; its two slots are moved here as well */
            double cx = 0.0, la, lb, ta, tb;
            RD_UNROLL
            for (int k = 0; k < NX; k++) cx += G[QG_JC + k] * QV(QG_UX, QG_DUX, 6 + k);
            QF(QG_LAM, QG_DLAM, 0, la); QF(QG_LAM, QG_DLAM, 12, lb);
            QF(QG_T, QG_DT, 0, ta); QF(QG_T, QG_DT, 12, tb);
            const int ia = TERM_LO(N), ib = TERM_UP(N);
            if (UPD) { AR(slam)[ia] = la; AR(slam)[ib] = lb; AR(st)[ia] = ta; AR(st)[ib] = tb; }
            const double dl = G[QG_D] + ta - cx, du = G[QG_D + 12] + tb + cx;
            const double ml = la * ta, mu_ = lb * tb;
            mu += ml + mu_;
            AR(res_d)[ia] = dl; AR(res_d)[ib] = du; AR(res_m)[ia] = ml; AR(res_m)[ib] = mu_;
            n2 = rd_fmax(n2, rd_fmax(fabs(dl), fabs(du)));
            n3 = rd_fmax(n3, rd_fmax(fabs(ml), fabs(mu_)));
            n4 = rd_fmax(n4, rd_fmax(fabs(ml - 1e-16), fabs(mu_ - 1e-16)));
        }
        if (UPD)
        { /* write the moved primal iterate and multipliers of this stage back */
            for (int c = lane; c < NV; c += GL) AR(sux)[NV * i + c] = G[QG_UX + c] + a * G[QG_DUX + c];
            if (mid && lane < NX) AR(spi)[NX * i + lane] = G[QG_PI + lane] + a * G[QG_DPI + lane];
        }
        rd_sync(); /* the buffer of this stage is refilled by the next iteration */
        SEC(5);
    }
    SECPROF_END(UPD ? 50 : 40);
#undef QV
#undef QF
    rd_cp_wait0();
    nrm[0] = grp_max(n0); nrm[1] = grp_max(n1); nrm[2] = grp_max(n2); nrm[3] = grp_max(n3); nrm[4] = grp_max(n4);
    nrm[5] = grp_max(nanflag);
    return grp_sum(mu);
}

/* d_ocp_qp_ipm_solve (H/ocp_qp/x_ocp_qp_ipm.c:2626-3018) with OCP_QP_IPM_DELTA_STEP (:2155-2622) for the option set
 * acados resolves: Mehrotra predictor-corrector, conditional second centering solve, up to two
 * refinement steps on the corrector, single damped step, lam/t floor 1e-16.  Returns the acados QP status
 * (0 ok, 2 max iter, 3 min step, 1 NaN).  The Householder-LQ fall-back of lq_fact = 1 (trigger: linear-system residual
 * > 1e-5, :2246-2276, evaluated by the forward sweep itself) is ipm_factor_lq(); the events are counted in flags[2]. */
RDN int ipm_solve_qp(const RmpcDims &D, const Inst &I, GrpSmem &sm, int warm, int *iters, double *tq)
{
    const int N = D.N, lane = rd_lane();
    const double tol = D.qp_tol, alpha_min = 1e-8, tau_min = 1e-16;
    const double nc_inv = 1.0 / (double) (NB * (N - 1) + 12 + (I.pe ? 2 : 0));
    /* device timers of the phases (tq[0] factorisation, [1] forward sweeps, [2] backward substitutions, [3] initialisation +
     * update / residual sweeps): read on the GPU's global timer at the phase boundaries, a handful of reads per iteration */
    double t0 = rd_time(), t1;
#define TQ(k) do { t1 = rd_time(); tq[k] += t1 - t0; t0 = t1; } while (0)
    ipm_init(D, I, warm);
    if (I.N2) ipm_init_pcond(D, I, sm, warm);
    double nrm[6], inrm[6], alpha = 1.0, sums[2];
    double sum_lt = ipm_res_sweep<0>(D, I, sm, 0.0, nrm);
    double mu = nc_inv * sum_lt;
    TQ(3);
    int kk;
    for (kk = 0; kk < D.qp_iter_max && alpha > alpha_min && (nrm[0] > tol || nrm[1] > tol || nrm[2] > tol || nrm[4] > tol); kk++)
    {
        /* affine (predictor) step */
                ipm_factor(D, I, sm);
        TQ(0);
        double lin[2];
        ipm_forward<0>(D, I, sm, AR(res_g), AR(res_b), AR(res_d), AR(res_m), tau_min, AR(dux), AR(dpi), AR(dlam), AR(dt_), &alpha, sums, lin);
        TQ(1);
        bool lq = lin[1] > 0.5 || lin[0] > 1e-5;
        if (I.pe && !lq)
        { /* with a terminal row the recursion can be ill-conditioned enough for the rows that vanish by construction of the
           * forward sweep (dynamics, bounds, complementarity) to leave rounding residue above HPIPM's threshold: evaluate
           * all four norms of OCP_QP_RES_COMPUTE_LIN like the reference does */
            double pn[6];
            ipm_res<1, 0>(D, I, sm, AR(res_g), AR(res_b), AR(res_d), AR(res_m), tau_min, AR(dux), AR(dpi), AR(dlam), AR(dt_), AR(slam), AR(st), nullptr,
                          nullptr, nullptr, nullptr, pn);
            lq = pn[0] > 1e-5 || pn[1] > 1e-5 || pn[2] > 1e-5 || pn[3] > 1e-5 || pn[5] > 0.5;
        }
#ifdef RMPC_EMU_FORCE_LQ
        lq = true;
#endif
        if (lq)
        {
            /* Inaccurate factorisation (x_ocp_qp_ipm.c:2246-2276: linear-system residual above 1e-5, or NaN): this
             * iteration is re-factorised with Householder LQ and the affine step re-solved with the stored factors
             * (OCP_QP_FACT_LQ_SOLVE_KKT_STEP).  The corrector solves below then run on the LQ factors.  Counted in flags[2]. */
            if (lane == 0) I.flags[2] += 1;
            ipm_factor_lq(D, I, sm);
            TQ(0);
            ipm_backward<0, 1>(D, I, sm, AR(res_g), AR(res_b), AR(res_d), AR(res_m), tau_min, AR(dux));
            TQ(2);
            ipm_forward<1>(D, I, sm, AR(res_g), AR(res_b), AR(res_d), AR(res_m), tau_min, AR(dux), AR(dpi), AR(dlam), AR(dt_), &alpha, sums, lin);
            TQ(1);
        }
        double mu_aff = ipm_mu_aff(sum_lt, sums, alpha, nc_inv);
        const double tmp = mu_aff / mu, sigma = tmp * tmp * tmp;
        double sigma_mu = sigma * mu;
        sigma_mu = sigma_mu > tau_min ? sigma_mu : tau_min;
        /* corrector: centering + second-order correction */
        ipm_backward<1, 1>(D, I, sm, AR(res_g), AR(res_b), AR(res_d), nullptr, sigma_mu, AR(dux));
        TQ(2);
        double clin[2];
        ipm_forward<1>(D, I, sm, AR(res_g), AR(res_b), AR(res_d), AR(res_mc), 0.0, AR(dux), AR(dpi), AR(dlam), AR(dt_), &alpha, sums, clin);
        TQ(1);
        /* conditional predictor-corrector */
        const double mu_aff0 = mu_aff;
        mu_aff = ipm_mu_aff(sum_lt, sums, alpha, nc_inv);
        if (mu_aff > 2.0 * mu_aff0)
        {
            ipm_backward<2, 1>(D, I, sm, AR(res_g), AR(res_b), AR(res_d), nullptr, sigma_mu, AR(dux));
            TQ(2);
            ipm_forward<1>(D, I, sm, AR(res_g), AR(res_b), AR(res_d), AR(res_mc), 0.0, AR(dux), AR(dpi), AR(dlam), AR(dt_), &alpha, sums, clin);
            TQ(1);
        }
        /* iterative refinement of the corrector step, itref_corr_max = 2 */
        int refined = 0;
        for (int it = 0; it < 2; it++)
        {
            /* first check: residual evaluated by the corrector's forward sweep itself; after a refinement: full pass */
            if (it == 0) { inrm[0] = clin[0]; inrm[1] = inrm[2] = inrm[3] = 0.0; }
            else
                ipm_res<1, 0>(D, I, sm, AR(res_g), AR(res_b), AR(res_d), AR(res_mc), 0.0, AR(dux), AR(dpi), AR(dlam), AR(dt_), AR(slam), AR(st), nullptr,
                              nullptr, nullptr, nullptr, inrm);
            if ((inrm[0] < tol || inrm[0] < 1e-3 * nrm[0]) && (inrm[1] < tol || inrm[1] < 1e-3 * nrm[1]) &&
                (inrm[2] < tol || inrm[2] < 1e-3 * nrm[2]) && (inrm[3] < tol || inrm[3] < 1e-3 * nrm[3]))
                break;
            ipm_res<1, 1>(D, I, sm, AR(res_g), AR(res_b), AR(res_d), AR(res_mc), 0.0, AR(dux), AR(dpi), AR(dlam), AR(dt_), AR(slam), AR(st), AR(ig), AR(ib), AR(id),
                          AR(im), inrm);
            double a2, s2[2], l2[2];
            ipm_backward<0, 0>(D, I, sm, AR(ig), AR(ib), AR(id), AR(im), 0.0, AR(iux));
            ipm_forward<1>(D, I, sm, AR(ig), AR(ib), AR(id), AR(im), 0.0, AR(iux), AR(ipi), AR(ilam), AR(it), &a2, s2, l2);
            refined = 1;
            for (int k = lane; k < (N + 1) * NV; k += GL) AR(dux)[k] += AR(iux)[k];
            for (int k = lane; k < N * NX; k += GL) AR(dpi)[k] += AR(ipi)[k];
            for (int k = lane; k < N * NB; k += GL) { AR(dlam)[k] += AR(ilam)[k]; AR(dt_)[k] += AR(it)[k]; }
            if (I.pe && lane < 2)
            {
                const int k = lane ? TERM_UP(N) : TERM_LO(N);
                AR(dlam)[k] += AR(ilam)[k]; AR(dt_)[k] += AR(it)[k];
            }
            rd_sync();
        }
        if (refined) alpha = ipm_alpha(D, I, AR(dlam), AR(dt_));
        double a = alpha;
        if (a < 1.0) a = a * ((1.0 - a) * 0.99 + a * 0.9999999);
        sum_lt = ipm_res_sweep<1>(D, I, sm, a, nrm); /* UPDATE_VAR_QP + OCP_QP_RES_COMPUTE */
        mu = nc_inv * sum_lt;
        TQ(3); /* incl. the rare refinement passes of this iteration */
    }
#undef TQ
    *iters = kk;
    if (kk == D.qp_iter_max) return 2;
    if (alpha <= alpha_min) return 3;
    if (mu != mu) return 1;
    return 0;
}

/* d_ocp_qp_restore_eq_dof (H/ocp_qp/x_ocp_qp_red.c:848-995): put dx0 back into the stage-0 step and recover the
 * multipliers of the 12 eliminated state equalities from stage-0 stationarity (sign decides lower/upper). */
RD void qp_restore(const RmpcDims &D, const Inst &I)
{
    const int lane = rd_lane();
    const double *par = I.par;
    if (lane < 12)
    {
        const double *Jc = AR(Jc); /* stage 0 */
        const int jj = lane % 6;
        const StageConst c = stage_const(par, D.dt, jj);
        double tmp = AR(rqz)[6 + lane] + ((lane < 6) ? AR(spi)[jj] : par[RP_A12 + jj] * AR(spi)[jj] + par[RP_A22 + jj] * AR(spi)[6 + jj]);
        for (int mm = 0; mm < 5; mm++)
        {
            double w = 0.0;
            for (int k = 0; k < 6; k++) w += Jc[6 * mm + k] * (I.head + H_DX0)[k];
            if (mm == 4)
                for (int k = 0; k < 6; k++) w += Jc[30 + k] * (I.head + H_DX0)[6 + k];
            if (lane < 6) tmp += Jc[6 * mm + lane] * w;
            else if (mm == 4) tmp += Jc[30 + jj] * w;
        }
        if (lane >= 6) tmp += c.hs * AR(sux)[jj] + c.hq * (I.head + H_DX0)[lane];
        (I.head + H_QLAMX0)[lane] = tmp >= 0.0 ? tmp : 1e-16;
        (I.head + H_QLAMX0)[12 + lane] = tmp >= 0.0 ? 1e-16 : -tmp;
        AR(sux)[6 + lane] = (I.head + H_DX0)[lane];
    }
    rd_sync();
}

/* ---------------------------------------------------------------------------------------------- SQP / RTI (K4)
 * MERIT_BACKTRACKING line search and iterate update: merit weights (Leineweber update,
 * AC/ocp_nlp/ocp_nlp_globalization_merit_backtracking.c:705-755), ocp_nlp_line_search (:290-408, alpha = 1, 0.7, ...
 * while alpha*0.7 > 0.05; the last alpha is taken untested), ocp_nlp_update_variables_sqp
 * (AC/ocp_nlp/ocp_nlp_common.c:3259-3318, full_step_dual = 0).  Returns 1 on a NaN/inf merit value. */
/* trial: workspace of the trial iterates, null = the slab's tux.  acados uses work->tmp_nlp_out for them -- the SAME buffer
 * AS-RTI level A keeps its iterate backup in (AC/ocp_nlp/ocp_nlp_sqp_rti.c:792,822,1104 vs
 * ocp_nlp_globalization_merit_backtracking.c:308-374), so with level A the 'restored' primal iterate is the last trial point
 * of the last line search; passing the backup's ux block here reproduces that. */
RDN int nlp_line_search_update(const RmpcChain &ch, const RmpcDims &D, const Inst &I_, int nlp_iter, double *alpha_out,
                               double *cost_new, int *cost_valid, const double *x0, int lin_valid, double *trial = nullptr)
{
    *cost_valid = 0; /* set when the accepted trial point's cost is known from its merit evaluation */
    const Inst I = I_; /* by value: registers, not re-read from the caller's frame after every store */
    const int N = D.N, lane = rd_lane();
    grp_weights(AR(wpi), AR(spi), nlp_iter == 0, 1, 0, N * NX);
    if (lane < 12)
    { /* stage 0 has bounds on u only: slots 0..5 and 12..17 */
        const int k = lane < 6 ? lane : lane + 6;
        const double t0 = AR(slam)[k];
        if (nlp_iter == 0) AR(wlam)[k] = t0;
        else { const double t1 = 0.5 * (t0 + AR(wlam)[k]); AR(wlam)[k] = t0 > t1 ? t0 : t1; }
    }
    grp_weights(AR(wlam), AR(slam), nlp_iter == 0, 0, NB, N * NB);
    for (int kk = lane; kk < NB; kk += GL)
    { /* the eliminated stage-0 state equalities */
        const double t0 = (I.head + H_QLAMX0)[kk];
        double *w = (I.head + H_WLAMX0) + kk;
        if (nlp_iter == 0) *w = t0;
        else { const double t1 = 0.5 * (t0 + *w); *w = t0 > t1 ? t0 : t1; }
    }
    if (I.pe && lane < 2)
    {
        const int k = lane ? TERM_UP(N) : TERM_LO(N);
        const double t0 = AR(slam)[k];
        if (nlp_iter == 0) AR(wlam)[k] = t0;
        else { const double t1 = 0.5 * (t0 + AR(wlam)[k]); AR(wlam)[k] = t0 > t1 ? t0 : t1; }
    }
    rd_sync();
    /* merit value at the current iterate: from the linearisation of this iterate where there is one (lin_valid), else by a
     * kinematics pass (AS-RTI levels B, C move the iterate without re-linearising) */
    const double m0 = lin_valid ? nlp_merit_lin(D, I, x0) : nlp_merit(ch, D, I, AR(ux), x0, nullptr);
    if (!trial) trial = AR(tux);
    double alpha = 1.0, m1 = 0.0, c1 = 0.0;
    int accepted = 0;
    while (alpha * 0.7 > 0.05)
    {
        grp_xpay_to(trial, AR(ux), AR(sux), alpha, (N + 1) * NV);
        rd_sync();
        m1 = nlp_merit(ch, D, I, trial, x0, &c1);
        const bool bad = (m1 != m1) || (m1 - m1 != 0.0);
        if (m1 < m0 && !bad) { accepted = 1; *cost_new = c1; *cost_valid = 1; break; }
        alpha *= 0.7;
    }
    *alpha_out = alpha;
    if (!accepted && ((m1 != m1) || (m1 - m1 != 0.0))) return 1;
    grp_axpy(AR(ux), AR(sux), alpha, (N + 1) * NV);
    grp_lerp(AR(pi), AR(spi), alpha, 0, N * NX);
    if (lane < 12)
    {
        const int k = lane < 6 ? lane : lane + 6;
        AR(lam)[k] = (1.0 - alpha) * AR(lam)[k] + alpha * AR(slam)[k];
    }
    grp_lerp(AR(lam), AR(slam), alpha, NB, N * NB);
    if (I.pe && lane >= 12 && lane < 14)
    {
        const int k = lane == 12 ? TERM_LO(N) : TERM_UP(N);
        AR(lam)[k] = (1.0 - alpha) * AR(lam)[k] + alpha * AR(slam)[k];
    }
    for (int k = lane; k < NB; k += GL) (I.head + H_LAMX0)[k] = (1.0 - alpha) * (I.head + H_LAMX0)[k] + alpha * (I.head + H_QLAMX0)[k];
    rd_sync();
    return 0;
}

/* ocp_qp_xcond_solve (AC/ocp_qp/ocp_qp_xcond_solver.c:518-576): reduce -> IPM -> restore.  warm start bookkeeping:
 * flags[1] != 0 until the first QP of the instance's lifetime has been solved (reference quirk, see ipm_init). */
RD int qp_solve(const RmpcDims &D, const Inst &I, GrpSmem &sm, int init_from_nlp, int *iters, double *tq, int keep_warm = 0)
{
    const int N = D.N, lane = rd_lane();
    const int warm = I.flags[1] ? 2 : 0;
    rd_sync();
    /* keep_warm: QPs of an AS-RTI preparation -- the option is only put back to its default by a FEEDBACK phase
     * (AC/ocp_nlp/ocp_nlp_sqp_rti.c:590-593), so the level-D iterations of a first call all run with warm_start 2 */
    if (lane == 0 && !keep_warm) I.flags[1] = 0;
    const double tx0 = rd_time();
    qp_reduce(D, I);
    if (init_from_nlp && warm == 2)
    {
        /* ocp_nlp_initialize_qp_from_nlp (AC/ocp_nlp/ocp_nlp_common.c:3320-3340) + ocp_qp_compute_t at ux = 0 */
        for (int k = lane; k < N * NX; k += GL) AR(spi)[k] = AR(pi)[k];
        for (int k = lane; k < N * NB; k += GL)
        {
            if (k < NB && (k % 12) >= 6) continue;
            AR(slam)[k] = AR(lam)[k];
            AR(st)[k] = -AR(d)[k];
        }
        if (I.pe && lane < 2)
        {
            const int k = lane ? TERM_UP(N) : TERM_LO(N);
            AR(slam)[k] = AR(lam)[k];
            AR(st)[k] = -AR(d)[k];
        }
        rd_sync();
    }
    const double tx1 = rd_time();
    const int st = ipm_solve_qp(D, I, sm, warm, iters, tq);
    const double tx2 = rd_time();
    qp_restore(D, I);
    tq[4] += (tx1 - tx0) + (rd_time() - tx2); /* time_qp_xcond */
    return st;
}

struct StepOut
{
    int status, sqp_iter, qp_iter, qp_status;
    int cost_valid; /* cost holds the NLS cost at the returned iterate (else closed_loop_group evaluates it) */
    double res[4], alpha, cost;
    double t_lin, t_qp, t_glob, tq[5]; /* device timers: linearisation, QP (tq: factor, forward, backward, residual, xcond), globalisation */
};
RD void step_timers_clear(StepOut &o)
{
    o.t_lin = o.t_qp = o.t_glob = 0.0;
    o.tq[0] = o.tq[1] = o.tq[2] = o.tq[3] = o.tq[4] = 0.0;
}

/* ---------------------------------------------------------------------------------------------- AS-RTI
 * Advanced-step real-time iterations, ocp_nlp_sqp_rti_preparation_advanced_step (AC/ocp_nlp/ocp_nlp_sqp_rti.c:773-1100): the
 * PREPARATION phase first solves an ADVANCED problem -- stage-0 state fixed at the prediction phi(x_0, u_0) of the current
 * iterate (as_rti_advance_problem, SIMULATE_ADVANCE, :697-727) -- by
 *   level A  one QP with the matrices AND vectors of the last linearisation (iterate restored to that point);
 *   level B  as_rti_iter zero-order iterations: gaps and bound residuals re-evaluated, gradient corrected by H * last step
 *            (ocp_nlp_zero_order_qp_update, AC/ocp_nlp/ocp_nlp_common.c:3109-3160);
 *   level C  as_rti_iter iterations with exact gaps, bound residuals and cost gradient, matrices kept
 *            (ocp_nlp_level_c_update, :3163-3232; its adjoint correction (BAbt_new - BAbt_old) pi vanishes: linear dynamics);
 *   level D  as_rti_iter full SQP iterations;
 * each followed by the merit line search / iterate update, and then re-linearises (the normal preparation).  The FEEDBACK
 * phase is the standard one.  The very first call only linearises.  iter counts like nlp_mem->iter does: it decides between
 * initialising and updating the merit weights (ocp_nlp_globalization_merit_backtracking.c:321-328). */
RD void asrti_backup(const RmpcDims &D, const Inst &I, double *bk, int restore)
{
    const int lane = rd_lane(), Sp = I.Sp;
    double *bux = bk, *bpi = bk + (long) Sp * NV, *blam = bpi + (long) Sp * NX, *bl0 = blam + (long) Sp * NB;
    const int N = D.N;
    if (restore)
    {
        for (int k = lane; k < (N + 1) * NV; k += GL) AR(ux)[k] = bux[k];
        for (int k = lane; k < N * NX; k += GL) AR(pi)[k] = bpi[k];
        for (int k = lane; k < (N + 1) * NB; k += GL) AR(lam)[k] = blam[k];
        for (int k = lane; k < NB; k += GL) (I.head + H_LAMX0)[k] = bl0[k];
    }
    else
    {
        for (int k = lane; k < (N + 1) * NV; k += GL) bux[k] = AR(ux)[k];
        for (int k = lane; k < N * NX; k += GL) bpi[k] = AR(pi)[k];
        for (int k = lane; k < (N + 1) * NB; k += GL) blam[k] = AR(lam)[k];
        for (int k = lane; k < NB; k += GL) bl0[k] = (I.head + H_LAMX0)[k];
    }
    rd_sync();
}

/* the vectors of the QP as ocp_nlp_approximate_qp_vectors_sqp leaves them: the x0 elimination of the last QP solve is undone
 * (qp_reduce works in place: r_0 and b_0 come back from their copies) */
RD void asrti_unreduce(const Inst &I, int redo_b0)
{
    const int lane = rd_lane();
    if (lane < 6) AR(rqz)[lane] = (I.head + H_G0U)[lane];
    if (redo_b0 && lane < NX) AR(b)[lane] = (I.head + H_B0)[lane];
    rd_sync();
}

struct GradEmit
{ /* cost gradient only: the Gauss-Newton data Jc of the last linearisation stays */
    double *g;
    const double *par, *yr;
    double dt, wr[5], gqd0[6];
    RDM void begin(const double *y5)
    {
        RD_UNROLL
        for (int m = 0; m < 5; m++) wr[m] = dt * par[RP_WTASK + m] * (y5[m] - yr[m]);
    }
    RDM void operator()(int k, const double *Jq, double Jd) const
    {
        double gq = 0.0, gqd = gqd0[k];
        RD_UNROLL
        for (int m = 0; m < 5; m++) gq += Jq[m] * wr[m];
        gqd += Jd * wr[4];
        g[6 + k] = gq;
        g[12 + k] = gqd;
    }
};

/* Vectors of the QP at the CURRENT iterate with the matrices of the last linearisation.  MODE 1 (level B): dynamics gaps,
 * bound residuals, r += H * (last step); MODE 2 (level C): gaps, bound residuals, r = exact cost gradient. */
template <int MODE>
RDN void asrti_update_vectors(const RmpcChain &ch, const RmpcDims &D, const Inst &I_, const double *x0)
{
    const Inst I = I_;
    const int N = D.N, lane = rd_lane();
    const double dt = D.dt;
    const double *par = I.par;
    for (int i = lane; i < N; i += GL)
    {
        double v[NV];
        load18(AR(ux) + NV * i, v);
        double *g = AR(rqz) + NV * i;
        const double *sp = I.stg ? I.stg + (long) RT_COUNT * i : nullptr, *yr = yref_of(par, I.stg, i);
        if (MODE == 1)
        { /* r += RSQ * step (dsymv_l on the Gauss-Newton Hessian of the last linearisation) */
            double s[NV], hv[NV];
            load18(AR(sux) + NV * i, s);
            RD_UNROLL
            for (int k = 0; k < NV; k++) hv[k] = 0.0;
            task_hess_mv(AR(Jc) + NJ * i, s + 6, hv + 6);
            RD_UNROLL
            for (int k = 0; k < 6; k++)
            {
                const StageConst c = stage_const(par, dt, k);
                hv[k] += c.hu * s[k] + c.hs * s[12 + k];
                hv[12 + k] += c.hs * s[k] + c.hq * s[12 + k];
            }
            RD_UNROLL
            for (int k = 0; k < NV; k++) g[k] += hv[k];
        }
        else
        {
            double y5[5];
            GradEmit em;
            em.g = g;
            RD_UNROLL
            for (int k = 0; k < 6; k++)
            {
                const double wq = par[RP_WQDD + k], wu = 2.0 * par[RP_WU + k];
                const double cu = par[RP_B2 + k] / dt, cq = (par[RP_A22 + k] - 1.0) / dt;
                const double a = (cu * v[k] + cq * v[12 + k]) - yr[11 + k];
                g[k] = dt * (wu * (v[k] - yr[5 + k]) + wq * cu * a);
                em.gqd0[k] = dt * (wq * cq * a);
            }
            em.par = par;
            em.yr = yr;
            em.dt = dt;
            ur10_task<true>(ch, par, v + 6, v + 12, y5, em);
        }
        const double *xn = AR(ux) + NV * (i + 1) + 6;
        double *b = AR(b) + NX * i, *d = AR(d) + NB * i;
        RD_UNROLL
        for (int k = 0; k < 6; k++)
        {
            b[k] = (v[6 + k] + par[RP_A12 + k] * v[12 + k]) + par[RP_B1 + k] * v[k] - xn[k];
            b[6 + k] = par[RP_A22 + k] * v[12 + k] + par[RP_B2 + k] * v[k] - xn[6 + k];
        }
        RD_UNROLL
        for (int j = 0; j < 12; j++)
        {
            if (i == 0 && j >= 6) { d[j] = 0.0; d[12 + j] = 0.0; }
            else { d[j] = lb_of(par, sp, j) - v[j]; d[12 + j] = v[j] - ub_of(par, sp, j); }
        }
        if (i == 0)
        {
            RD_UNROLL
            for (int k = 0; k < NX; k++) (I.head + H_DX0)[k] = x0[k] - v[6 + k];
        }
    }
    if (I.pe)
    { /* terminal row: function value at the new x_N, gradient row of the last linearisation */
        const double h = term_eval(I.pe, AR(ux) + NV * N + 6, nullptr);
        if (lane == 0)
        {
            AR(d)[TERM_LO(N)] = I.pe[RE_LH] - h;
            AR(d)[TERM_UP(N)] = h - I.pe[RE_UH];
        }
    }
    rd_sync();
}

/* ocp_nlp_sqp_rti, PREPARATION_AND_FEEDBACK (AC/ocp_nlp/ocp_nlp_sqp_rti.c:455-646,1123-1175) */
/* phase (opts->rti_phase, :1140-1170): 0 = both, 1 = PREPARATION only -- everything that does not depend on x0: model
 * functions, Gauss-Newton data, dynamics gaps, bound residuals (ocp_nlp_approximate_qp_matrices) --, 2 = FEEDBACK only:
 * the x0-dependent right-hand side (here: dx0 of the stage-0 equality), QP solve, globalisation and iterate update.
 * Preparation followed by feedback gives bit-identical results to the combined call for the same x0. */
/* as_level / as_iter / bk: AS-RTI level (-1: standard RTI), iterations and level-A backup of this instance (RmpcWs) */
RD void solve_rti(const RmpcChain &ch, const RmpcDims &D, const Inst &I, GrpSmem &sm, StepOut &o, int phase, int as_level = -1,
                  int as_iter = 0, double *bk = nullptr)
{
    o.sqp_iter = 0; o.qp_iter = 0; o.qp_status = 0; o.alpha = 0.0; o.status = 0;
    o.res[0] = o.res[1] = o.res[2] = o.res[3] = 0.0;
    step_timers_clear(o);
    double tt = rd_time();
    const int lane = rd_lane();
    if (phase != 2)
    {
        int iter = 0; /* nlp_mem->iter, reset by every preparation (reset_stats_and_sub_timers) */
        const int first_call = I.flags[0];
        if (as_level >= 0 && phase == 1 && (!first_call || as_level == 3))
        { /* ocp_nlp_sqp_rti_preparation_advanced_step; the first call only linearises -- except level D, whose iterations are
           * not guarded by is_first_call (:1018) and then run on the stage-0 bounds as they are */
            double *x0a = I.head + H_X0ADV;
            const double *par = I.par;
            rd_sync();
            if (first_call)
            {
                if (lane < NX) x0a[lane] = I.z[lane];
            }
            else if (lane < 6)
            { /* as_rti_advance_problem: lbx = ubx = (phi(x_0, u_0) - x_1) + x_1 of the current iterate */
                const double *v = AR(ux), *xn = AR(ux) + NV + 6;
                const double g0 = (v[6 + lane] + par[RP_A12 + lane] * v[12 + lane]) + par[RP_B1 + lane] * v[lane] - xn[lane];
                const double g1 = par[RP_A22 + lane] * v[12 + lane] + par[RP_B2 + lane] * v[lane] - xn[6 + lane];
                x0a[lane] = g0 + xn[lane];
                x0a[6 + lane] = g1 + xn[6 + lane];
            }
            rd_sync();
            int qi = 0, failed = 0;
            if (as_level == 0)
            { /* level A: the iterate goes back to the last linearisation point; only the stage-0 equality changes */
                asrti_backup(D, I, bk, 1);
                asrti_unreduce(I, 1);
                /* bound residuals are re-formed from the constraint values kept at the linearisation (update_qp_vectors,
                 * AC/ocp_nlp/ocp_nlp_constraints_bgh.c:1564-1586): only the moved stage-0 bounds change anything */
                if (lane < NX) (I.head + H_DX0)[lane] = x0a[lane] - (I.head + H_XLIN0)[lane];
                rd_sync();
                { const double t = rd_time(); o.t_lin += t - tt; tt = t; }
                iter += 1;
                o.qp_status = qp_solve(D, I, sm, 0, &qi, o.tq, 1);
                o.qp_iter += qi;
                { const double t = rd_time(); o.t_qp += t - tt; tt = t; }
                /* no status test here in the reference (:838-866): the iterate is updated whatever the QP returned */
                int cv;
                double cn;
                if (nlp_line_search_update(ch, D, I, iter, &o.alpha, &cn, &cv, x0a, 0, bk)) { o.status = 1; failed = 1; }
                { const double t = rd_time(); o.t_glob += t - tt; tt = t; }
            }
            else
            {
                for (; iter < as_iter && !failed; iter++)
                {
                    if (as_level == 1) { asrti_unreduce(I, 0); asrti_update_vectors<1>(ch, D, I, x0a); }
                    else if (as_level == 2) asrti_update_vectors<2>(ch, D, I, x0a);
                    else nlp_linearise(ch, D, I, x0a);
                    { const double t = rd_time(); o.t_lin += t - tt; tt = t; }
                    o.qp_status = qp_solve(D, I, sm, 0, &qi, o.tq, 1);
                    o.qp_iter += qi;
                    { const double t = rd_time(); o.t_qp += t - tt; tt = t; }
                    if (o.qp_status != 0 && o.qp_status != 2) { o.status = 4; failed = 1; break; }
                    int cv;
                    double cn;
                    if (nlp_line_search_update(ch, D, I, iter, &o.alpha, &cn, &cv, x0a, as_level == 3)) { o.status = 1; failed = 1; }
                    { const double t = rd_time(); o.t_glob += t - tt; tt = t; }
                }
            }
            if (failed)
            { /* the reference returns from the preparation without re-linearising (:851-866,930-945) */
                if (lane == 0) I.flags[3] = iter;
                rd_sync();
                o.sqp_iter = iter;
                o.cost_valid = 0;
                return;
            }
        }
        o.cost = nlp_linearise(ch, D, I, I.z);
        if (lane == 0) { (I.head + H_LINCOST)[0] = o.cost; I.flags[3] = iter; }
        if (as_level == 0 && phase == 1) asrti_backup(D, I, bk, 0); /* level A: tmp_nlp_out <- nlp_out (:1080-1085) */
        o.sqp_iter = iter;
    }
    else
    {
        /* ocp_nlp_approximate_qp_vectors_sqp: only the stage-0 state equality depends on x0 */
        if (lane < NX) (I.head + H_DX0)[lane] = I.z[lane] - AR(ux)[6 + lane];
        rd_sync();
        o.cost = (I.head + H_LINCOST)[0];
    }
    o.cost_valid = 1; /* cost at the iterate the step starts from; replaced below once the iterate moves */
    { const double t = rd_time(); o.t_lin += t - tt; tt = t; }
    if (phase == 1) return;
    const int first = I.flags[0];
    const int iter = (phase == 2 ? I.flags[3] : 0) + 1; /* nlp_mem->iter += 1 (:535) */
    rd_sync();
    int it = 0;
    /* ocp_nlp_initialize_qp_from_nlp on the first call reaches the interior-point start only through
     * ocp_qp_xcond_solve (AC/ocp_qp/ocp_qp_xcond_solver.c:548-559); a FEEDBACK phase solves with
     * ocp_qp_xcond_condense_rhs_and_solve (:614-660), which does not look at that request: the first QP of a split call
     * then starts from the QP memory as it is (warm_start 2: zeros clipped at 0.1) */
    const int qs = qp_solve(D, I, sm, first && phase != 2, &it, o.tq);
    { const double t = rd_time(); o.t_qp += t - tt; tt = t; }
    PAIR_SYNC();
    o.sqp_iter = iter; o.qp_iter = it; o.qp_status = qs;
    o.res[0] = o.res[1] = o.res[2] = o.res[3] = 0.0; /* rti_log_residuals = 0: the reference reports zeros */
    o.alpha = 0.0;
    if (qs != 0 && qs != 2) { o.status = 4; return; } /* ACADOS_QP_FAILURE, iterate untouched (:609-622) */
    /* globalisation status ignored (:627-635); a NaN merit leaves the iterate where it was */
    if (nlp_line_search_update(ch, D, I, 1, &o.alpha, &o.cost, &o.cost_valid, I.z, 1, as_level == 0 ? bk : nullptr)) o.cost_valid = 1;
    if (lane == 0) I.flags[0] = 0;
    rd_sync();
    o.t_glob += rd_time() - tt;
    o.status = 0;
}

/* ocp_nlp_sqp (AC/ocp_nlp/ocp_nlp_sqp.c:474-801) with check_termination (:351-443) */
RD void solve_sqp(const RmpcChain &ch, const RmpcDims &D, const Inst &I, GrpSmem &sm, StepOut &o, double *stat)
{
    o.qp_iter = 0; o.qp_status = 0; o.alpha = 0.0; o.cost = 0.0; o.cost_valid = 0;
    step_timers_clear(o);
    /* stat: rows of acados' mem->stat for this solve (RQ_*, AC/ocp_nlp/ocp_nlp_sqp.c:573-576,700-701,788), null: not kept */
    const bool log = stat && rd_lane() == 0;
    if (log)
        for (int k = 0; k < RQ_COUNT; k++) stat[k] = 0.0;
    for (int it = 0; it <= D.max_iter; it++)
    {
        o.sqp_iter = it;
        double cost = 0.0;
        double tt = rd_time();
        if (it != D.max_iter)
        {
            cost = nlp_linearise(ch, D, I, I.z);
            o.cost = cost; o.cost_valid = 1; /* every exit before the next update returns this iterate */
            nlp_residuals(D, I, o.res);
            if (log)
                for (int k = 0; k < 4; k++) stat[RQ_COUNT * it + RQ_RES + k] = o.res[k];
        }
        { const double t = rd_time(); o.t_lin += t - tt; tt = t; }
        const double *r = o.res;
        if (r[0] != r[0] || r[1] != r[1] || r[2] != r[2] || r[3] != r[3]) { o.status = 1; return; }
        if (it >= D.max_iter) { o.status = 2; return; }
        if (r[0] < D.tol[0] && r[1] < D.tol[1] && r[2] < D.tol[2] && r[3] < D.tol[3]) { o.status = 0; return; }
        if (cost <= -1e10) { o.status = 6; return; }
        int qi = 0;
        const int qs = qp_solve(D, I, sm, 0, &qi, o.tq);
        { const double t = rd_time(); o.t_qp += t - tt; tt = t; }
        o.qp_iter += qi; o.qp_status = qs;
        if (log)
        {
            for (int k = 0; k < RQ_COUNT; k++) stat[RQ_COUNT * (it + 1) + k] = 0.0;
            stat[RQ_COUNT * (it + 1) + RQ_QP_STATUS] = qs;
            stat[RQ_COUNT * (it + 1) + RQ_QP_ITER] = qi;
        }
        if (qs != 0 && qs != 2) { o.sqp_iter = it + 1; o.status = 4; return; }
        const int ls = nlp_line_search_update(ch, D, I, it, &o.alpha, &o.cost, &o.cost_valid, I.z, 1);
        o.t_glob += rd_time() - tt;
        if (log) stat[RQ_COUNT * (it + 1) + RQ_ALPHA] = o.alpha;
        if (ls) { o.cost = cost; o.cost_valid = 1; o.status = 1; return; }
    }
    o.status = 2;
}

/* ---------------------------------------------------------------------------------------------- closed loop (K5)
 * R/simulator.py:199-241: x = plant state; solver.set(0,'lbx'/'ubx',x); solve; u = solver.get(0,'u'); plant RK4.
 * Plant: zdot = [qd; -W qd + W u] (R/simulation_model.py:79-83) decouples per joint, RK4 (:111-117) on lanes 0..5. */
/* Primal shift of the initial guess between two closed-loop steps (RmpcWs.shift; the acados MPC idiom
 * set(i, 'x', get(i+1, 'x')), set(i, 'u', get(i+1, 'u')) for i < N-1, set(N-1, 'x', get(N, 'x')) -- SURVEY.md 8(f) #1; the
 * reference itself never shifts).  Multipliers and merit weights stay.  [u; x] of the stages are contiguous, so the shift
 * is one move by NV doubles towards the front: ascending chunks of GL elements, each lane loads before it stores and no
 * chunk writes what a later one still reads (NV > GL). */
RD void nlp_shift(const RmpcDims &D, const Inst &I)
{
    const int N = D.N, lane = rd_lane();
    double *ux = AR(ux);
    const int n = NV * (N - 1);
    for (int k0 = 0; k0 < n; k0 += GL)
    {
        const int k = k0 + lane;
        double v = 0.0;
        if (k < n) v = ux[k + NV];
        rd_sync();
        if (k < n) ux[k] = v;
    }
    rd_sync();
    if (lane < NX) ux[NV * (N - 1) + 6 + lane] = ux[NV * N + 6 + lane];
    rd_sync();
}

RD void closed_loop_group(const RmpcChain &ch, const RmpcDims &D, const RmpcWs &W, long inst, GrpSmem &sm, int nsteps,
                         int advance = 1, unsigned pairmask = 0)
{
    Inst I_ = inst_view(W, D.N, inst);
    if (pairmask) I_.pair = pairmask;
    const Inst I = I_;
    const int lane = rd_lane(), N = D.N;
    const double *par = I.par;
    smem_consts(I, sm, D.dt);
    /* tracking errors e1..e5 are logged at every plant sample incl. the initial state (Nsim + 1 samples,
     * R/simulation_model.py:24-38, R/simulator.py:265-344); RMSE / ITSE sums of R/simulator.py:347-390 */
    if (W.summary && W.step0 == 0 && lane == 0 && advance)
    {
        double e[5], *sm_ = W.summary + inst * RS_COUNT;
        ur10_tracking_errors(ch, par, I.z, I.z + 6, e);
        for (int m = 0; m < 5; m++) sm_[RS_SSE + m] += e[m] * e[m];
        sm_[RS_NSAMPLES] += 1.0;
    }
    rd_sync();
    for (int s = 0; s < nsteps; s++)
    {
        const int gs = W.step0 + s;
        StepOut o;
        PAIR_SYNC();
        const double t_begin = rd_time();
        if (D.solver_type == RMPC_SQP_RTI)
        {
            double *bk = W.asrti ? W.asrti + inst * RMPC_ASRTI_STRIDE(W.Sp) : nullptr;
            if (advance && W.as_rti)
            { /* AS-RTI closed loop: advanced preparation on the prediction, then feedback on the measured state */
                StepOut op;
                solve_rti(ch, D, I, sm, op, 1, W.as_rti - 1, W.as_rti_iter, bk);
                solve_rti(ch, D, I, sm, o, 2, W.as_rti - 1, W.as_rti_iter, bk);
                o.qp_iter += op.qp_iter;
                if (o.status == 0 && op.status != 0) o.status = op.status; /* a failed advanced step (1 NaN merit, 4 QP failure) reaches slog / RS_NFAIL */
                o.t_lin += op.t_lin; o.t_qp += op.t_qp; o.t_glob += op.t_glob;
                for (int k = 0; k < 5; k++) o.tq[k] += op.tq[k];
            }
            else solve_rti(ch, D, I, sm, o, advance ? 0 : W.rti_phase, W.as_rti - 1, W.as_rti_iter, bk);
        }
        else solve_sqp(ch, D, I, sm, o, W.sqpstat ? W.sqpstat + inst * (long) (D.max_iter + 2) * RQ_COUNT : nullptr);
        const double t_solve = rd_time() - t_begin; /* stands in for get_stats('time_tot') (R/simulator.py:220) */
        /* solver.get_cost(): cost at the returned iterate (T/acados_ocp_solver.py:1664-1679).  The solve usually knows it
         * already (linearisation of an iterate it did not move, or the merit evaluation of the accepted trial point). */
        double cost = o.cost;
        if (!o.cost_valid)
        {
            cost = 0.0;
            for (int i = lane; i < N; i += GL)
            {
                double v[NV], y5[5];
                load18(AR(ux) + NV * i, v);
                NoEmit ne;
                ur10_task<false>(ch, par, v + 6, v + 12, y5, ne);
                cost += stage_cost(par, yref_of(par, I.stg, i), D.dt, y5, v, v + 12);
            }
            cost = grp_sum(cost);
        }
        double zq = 0.0, zqd = 0.0, u = 0.0;
        if (lane < 6)
        {
            u = AR(ux)[lane];
            const double w = par[RP_WCV + lane], h = D.plant_dt;
            const double q = I.z[lane], qd = I.z[6 + lane];
            const double k1q = qd, k1d = -w * qd + w * u;
            const double q2 = q + 0.5 * h * k1q, d2 = qd + 0.5 * h * k1d;
            const double k2q = d2, k2d = -w * d2 + w * u;
            const double q3 = q + 0.5 * h * k2q, d3 = qd + 0.5 * h * k2d;
            const double k3q = d3, k3d = -w * d3 + w * u;
            const double q4 = q + h * k3q, d4 = qd + h * k3d;
            const double k4q = d4, k4d = -w * d4 + w * u;
            zq = q + (h / 6) * k1q + (h / 3) * k2q + (h / 3) * k3q + (h / 6) * k4q;
            zqd = qd + (h / 6) * k1d + (h / 3) * k2d + (h / 3) * k3d + (h / 6) * k4d;
            (void) q2; (void) q3; (void) q4;
        }
        rd_sync();
        if (lane < 6 && advance)
        {
            I.z[lane] = zq; I.z[6 + lane] = zqd;
            if (W.ulog && gs < W.nlog) W.ulog[(inst * W.nlog + gs) * 6 + lane] = u;
            if (W.zlog && gs < W.nlog)
            {
                W.zlog[(inst * (W.nlog + 1) + gs + 1) * NX + lane] = zq;
                W.zlog[(inst * (W.nlog + 1) + gs + 1) * NX + 6 + lane] = zqd;
            }
        }
        if (lane == 0)
        {
            if (W.slog && gs < W.nlog)
            {
                double *sl = W.slog + (inst * W.nlog + gs) * RL_COUNT;
                sl[RL_STATUS] = o.status; sl[RL_SQP_ITER] = o.sqp_iter; sl[RL_QP_ITER] = o.qp_iter; sl[RL_COST] = cost;
                sl[RL_RES] = o.res[0]; sl[RL_RES + 1] = o.res[1]; sl[RL_RES + 2] = o.res[2]; sl[RL_RES + 3] = o.res[3];
                sl[RL_ALPHA] = o.alpha; sl[RL_QP_STATUS] = o.qp_status; sl[RL_TIME] = t_solve;
                sl[RL_TIME_LIN] = o.t_lin; sl[RL_TIME_QP] = o.t_qp; sl[RL_TIME_GLOB] = o.t_glob;
                sl[RL_TIME_QP_FACT] = o.tq[0]; sl[RL_TIME_QP_FWD] = o.tq[1]; sl[RL_TIME_QP_BWD] = o.tq[2]; sl[RL_TIME_QP_RES] = o.tq[3];
                sl[RL_TIME_XCOND] = o.tq[4]; sl[RL_COUNT - 1] = 0.0;
            }
            if (W.summary)
            {
                double *sm_ = W.summary + inst * RS_COUNT;
                sm_[RS_NFAIL] += (o.status != 0);
                sm_[RS_SQP_ITERS] += o.sqp_iter;
                sm_[RS_QP_ITERS] += o.qp_iter;
                for (int k = 0; k < 4; k++) sm_[RS_MAX_RES + k] = rd_fmax(sm_[RS_MAX_RES + k], o.res[k]);
                sm_[RS_COST_LAST] = cost;
            }
        }
        rd_sync();
        if (W.summary && lane == 0 && advance)
        {
            double e[5], *sm_ = W.summary + inst * RS_COUNT;
            ur10_tracking_errors(ch, par, I.z, I.z + 6, e);
            const double tk = (gs + 1) * D.plant_dt;
            for (int m = 0; m < 5; m++)
            {
                sm_[RS_SSE + m] += e[m] * e[m];
                sm_[RS_ITSE + m] += tk * (e[m] * e[m]) * D.plant_dt;
            }
            sm_[RS_NSAMPLES] += 1.0;
        }
        rd_sync();
        if (W.shift && advance) nlp_shift(D, I);
    }
}
