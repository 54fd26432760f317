/* robotic-mpc_b200/csrc/rmpc_device.cuh -- device code of the batched UR10 surface-tracking MPC engine (sm_100a).
 *
 * ONE WARP owns ONE problem instance for the whole closed loop (R/simulator.py:199-241):
 *     set x0 -> SQP_RTI / SQP solve -> u0 -> plant RK4 -> metrics,   nsteps times, no host round trip.
 * Inside a solve the warp alternates between two parallelisation patterns:
 *   (S) stage-parallel  : lane l handles stages l, l+32, ... (forward kinematics / Jacobian / Gauss-Newton data,
 *                         KKT and QP residuals, merit function, dual steps) with warp-shuffle reductions;
 *   (R) stage-sequential: the square-root Riccati recursion of the interior-point method; lane r owns row r of
 *                         the 19x18 stage block [[R S'],[S Q],[r' q']] + [B A b]'P[B A b], the previous stage's
 *                         factor is staged in shared memory.
 * Replaces, for a batch: CasADi-generated model functions, acados ocp_nlp (SQP / SQP_RTI, NLS cost, discrete
 * dynamics, BGH bounds, merit backtracking), acados ocp_qp partial condensing wrapper (x0 elimination) and the
 * HPIPM interior-point solver -- file:line citations are on each function.  R/ = reference root,
 * AC/ = acados/acados, H/ = acados/external/hpipm.
 *
 * The file compiles in two ways: with nvcc for the product, and with g++ -DRMPC_EMU against
 * tests/simt_emu/simt_emu.h (a lock-step warp emulator) for CPU-side unit tests of this very source.
 */
#pragma once
#include "rmpc_types.h"

#ifdef RMPC_EMU
#include "simt_emu.h"
#define RD static inline
RD int rd_lane() { return simt_emu::lane_id(); }
RD void rd_sync() { simt_emu::syncwarp(); }
RD double rd_shfl(double v, int src) { return simt_emu::shfl(v, src); }
RD double rd_shfl_xor(double v, int m) { return simt_emu::shfl_xor(v, m); }
RD double rd_rsqrt(double x) { return 1.0 / sqrt(x); }
RD void rd_sincos(double a, double *s, double *c) { *s = sin(a); *c = cos(a); }
RD double rd_fmax(double a, double b) { return a > b ? a : b; } /* NaN-ignoring like fmax for our uses */
#define RD_UNROLL
#else
#define RD __device__ __forceinline__
RD int rd_lane() { return threadIdx.x & 31; }
RD void rd_sync() { __syncwarp(); }
RD double rd_shfl(double v, int src) { return __shfl_sync(0xffffffffu, v, src); }
RD double rd_shfl_xor(double v, int m) { return __shfl_xor_sync(0xffffffffu, v, m); }
RD double rd_rsqrt(double x) { return rsqrt(x); }
RD void rd_sincos(double a, double *s, double *c) { sincos(a, s, c); }
RD double rd_fmax(double a, double b) { return a > b ? a : b; }
#define RD_UNROLL _Pragma("unroll")
#endif

#define NX RMPC_NX
#define NU RMPC_NU
#define NV RMPC_NV
#define NB RMPC_NB
#define LS RMPC_LS

RD double warp_sum(double v)
{
    for (int m = 16; m; m >>= 1) v += rd_shfl_xor(v, m);
    return v;
}
RD double warp_max(double v)
{
    for (int m = 16; m; m >>= 1) { double o = rd_shfl_xor(v, m); v = o > v ? o : v; }
    return v;
}
/* NaN-propagating flag OR */
RD int warp_any(int f)
{
    double v = f ? 1.0 : 0.0;
    return warp_max(v) > 0.5;
}

/* ---------------------------------------------------------------------------------------------- shared memory */
#define SM_LXS 13 /* row stride of the staged Lxx block (odd: conflict-free row reads) */
#define SM_ALS 18 /* row stride of AL (even: 16-byte aligned pairs) */
struct WarpSmem
{
    double Lx[12 * SM_LXS]; /* Lxx of stage ss+1 (lower), row-major */
    double lx[12];          /* x part of the last row of L[ss+1] */
    double AL[19 * SM_ALS]; /* [B A b]' Lxx' | scaled task Jacobian */
    double col[2][20];      /* potrf column broadcast, double buffered */
    double xv[12];          /* running state vector of the sweeps */
    double uv[8];
    double bv[12], wv[12], tv[18];
};

/* per-instance views */
struct Inst
{
    const double *par;
    double *ux, *pi, *lam, *lamx0, *wpi, *wlam, *wlamx0, *z;
    int *flags;
    double *Jc, *rqz, *b, *d, *dx0, *g0u;
    double *sux, *spi, *slam, *st, *qlamx0, *dux, *dpi, *dlam, *dt_;
    double *res_g, *res_b, *res_d, *res_m, *res_m_bkp, *pvec, *Pb, *L;
    double *ig, *ib, *id, *im, *iux, *ipi, *ilam, *it, *tux;
};

RD Inst inst_view(const RmpcWs &w, int N, long i)
{
    Inst v;
    const long S = N + 1;
    v.par = w.par + i * RP_COUNT;
    v.ux = w.ux + i * S * NV; v.pi = w.pi + i * S * NX; v.lam = w.lam + i * S * NB; v.lamx0 = w.lamx0 + i * NB;
    v.wpi = w.wpi + i * S * NX; v.wlam = w.wlam + i * S * NB; v.wlamx0 = w.wlamx0 + i * NB;
    v.z = w.z + i * NX; v.flags = w.flags + i * 4;
    v.Jc = w.Jc + i * S * RMPC_NJ; v.rqz = w.rqz + i * S * NV; v.b = w.b + i * S * NX; v.d = w.d + i * S * NB;
    v.dx0 = w.dx0 + i * NX; v.g0u = w.g0u + i * 8;
    v.sux = w.sux + i * S * NV; v.spi = w.spi + i * S * NX; v.slam = w.slam + i * S * NB; v.st = w.st + i * S * NB;
    v.qlamx0 = w.qlamx0 + i * NB;
    v.dux = w.dux + i * S * NV; v.dpi = w.dpi + i * S * NX; v.dlam = w.dlam + i * S * NB; v.dt_ = w.dt_ + i * S * NB;
    v.res_g = w.res_g + i * S * NV; v.res_b = w.res_b + i * S * NX; v.res_d = w.res_d + i * S * NB;
    v.res_m = w.res_m + i * S * NB; v.res_m_bkp = w.res_m_bkp + i * S * NB;
    v.pvec = w.pvec + i * S * NX; v.Pb = w.Pb + i * S * NX; v.L = w.L + i * S * LS;
    v.ig = w.ig + i * S * NV; v.ib = w.ib + i * S * NX; v.id = w.id + i * S * NB; v.im = w.im + i * S * NB;
    v.iux = w.iux + i * S * NV; v.ipi = w.ipi + i * S * NX; v.ilam = w.ilam + i * S * NB; v.it = w.it + i * S * NB;
    v.tux = w.tux + i * S * NV;
    return v;
}

/* offset of column c in the packed lower-trapezoidal 19x18 factor: rows c..18 stored contiguously */
RD int lcol(int c) { return 19 * c - (c * (c - 1)) / 2; }
#define LEL(Lp, r, c) ((Lp)[lcol(c) + (r) - (c)])

/* a bound slot (stage i, slot j in [0,24)) takes part in the QP?  stage 0 keeps only the u bounds after the
 * x0 elimination (H/ocp_qp/x_ocp_qp_red.c:278-466); stage N has none */
RD bool slot_on(int i, int j, int N) { return i < N && (i > 0 || (j % 12) < 6); }

/* ---------------------------------------------------------------------------------------------- model (K1)
 * UR10 forward kinematics, WORLD-frame geometric Jacobian, task-frame outputs and their derivatives.
 * Replaces the CasADi-generated *_cost_y_fun[_jac_ut_xt] (T/casadi_function_generation.py:555-575) whose
 * expression graph comes from R/prediction_model.py:117-173,226-326 (pinocchio FK of frame ee_link, WORLD
 * Jacobian, task transform, v_task), R/surface.py:21,243-262 (S, unit normal) and
 * R/trajectory_optimizer.py:88-137 (g1..g5).  Chain constants: R/ur_description/urdf/ur10.urdf:50-222.
 *   y5 = [S(p_t) - p_tz, n . R_t[:,2], R_t[0,1], p_tx, (R_t' (v + w x t_w))_y]
 *   J  = d y5 / d x  (5 x 12; rows 0..3 depend on q only)
 * The remaining 13 outputs (u, qddot_k, manipulability with weight 0) are affine/irrelevant and handled by the
 * callers in closed form. */
RD void cross3(const double *a, const double *b, double *c)
{
    c[0] = a[1] * b[2] - a[2] * b[1];
    c[1] = a[2] * b[0] - a[0] * b[2];
    c[2] = a[0] * b[1] - a[1] * b[0];
}
RD double dot3(const double *a, const double *b) { return a[0] * b[0] + a[1] * b[1] + a[2] * b[2]; }

template <bool JAC>
RD void ur10_task(const RmpcChain &ch, const double *par, const double *q, const double *qd, double *y5, double *J /*[5][12]*/)
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
        J[0 * 12 + k] = Sx * dpt[0] + Sy * dpt[1] - dpt[2];
        double dn[3] = {nX[0] * dpt[0] + nY[0] * dpt[1], nX[1] * dpt[0] + nY[1] * dpt[1], nX[2] * dpt[0] + nY[2] * dpt[1]};
        J[1 * 12 + k] = dot3(dn, r2) + dot3(n, dr2);
        J[2 * 12 + k] = dr1[0];
        J[3 * 12 + k] = dpt[0];
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
        J[4 * 12 + k] = dot3(dr1, wv) + dot3(r1, dwv);
        double e[3] = {tw[0] - pj[k][0], tw[1] - pj[k][1], tw[2] - pj[k][2]}, ce[3];
        cross3(aj[k], e, ce);
        J[4 * 12 + 6 + k] = dot3(r1, ce);
        J[0 * 12 + 6 + k] = 0.0; J[1 * 12 + 6 + k] = 0.0; J[2 * 12 + 6 + k] = 0.0; J[3 * 12 + 6 + k] = 0.0;
    }
}

/* plant logs: pose of frame ee_link and WORLD twist J qd (R/simulation_model.py:60-77), then the post-hoc tracking
 * errors e1..e5 of R/simulator.py:301-342 -- NB e5 deliberately uses the reference's own formula
 * R_w_t (v + (w . t_w)), not the OCP's g5. */
RD void ur10_tracking_errors(const RmpcChain &ch, const double *par, const double *q, const double *qd, double *e)
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
}

/* ---------------------------------------------------------------------------------------------- NLP level (K2, K4)
 * Per-instance constants of the Gauss-Newton Hessian that do not depend on the stage:
 *   y[5+i] = u_i (weight 2 w_u),  y[11+i] = (b2_i u_i + (a22_i - 1) qd_i) / Ts (weight w_qddot)
 * => H_uu = diag(hu), H_{qd,u} = diag(hs), H_{qd,qd} += diag(hq)        (R/trajectory_optimizer.py:121-154) */
struct StageConst
{
    double hu, hs, hq, cu, cq; /* for joint (lane % 6) */
};
RD StageConst stage_const(const double *par, double dt, int i)
{
    StageConst c;
    c.cu = par[RP_B2 + i] / dt;
    c.cq = (par[RP_A22 + i] - 1.0) / dt;
    const double wq = par[RP_WQDD], wu = 2.0 * par[RP_WU];
    c.hu = dt * (wu + wq * c.cu * c.cu);
    c.hs = dt * (wq * c.cu * c.cq);
    c.hq = dt * (wq * c.cq * c.cq);
    return c;
}

/* stage cost dt * 1/2 |y - yref|^2_W from the five task outputs and (u, qd): ocp_nlp_cost_nls_compute_fun
 * (AC/ocp_nlp/ocp_nlp_cost_nls.c:1045-1133) */
RD double stage_cost(const double *par, double dt, const double *y5, const double *u, const double *qd)
{
    const double r0 = y5[0], r1 = y5[1] - 1.0, r2 = y5[2], r3 = y5[3] - par[RP_PXREF], r4 = y5[4] - par[RP_VYREF];
    double f = par[RP_WTASK] * r0 * r0 + par[RP_WTASK + 1] * r1 * r1 + par[RP_WTASK + 2] * r2 * r2 + par[RP_WTASK + 3] * r3 * r3 +
               par[RP_WTASK + 4] * r4 * r4;
    const double wq = par[RP_WQDD], wu = 2.0 * par[RP_WU];
    RD_UNROLL
    for (int i = 0; i < 6; i++)
    {
        const double a = (par[RP_B2 + i] * u[i] + (par[RP_A22 + i] - 1.0) * qd[i]) / dt;
        f += wu * u[i] * u[i] + wq * a * a;
    }
    return dt * (0.5 * f);
}

RD void load18(const double *src, double *v)
{
    RD_UNROLL
    for (int k = 0; k < NV; k++) v[k] = src[k];
}

/* lower/upper bound of slot j in [0,12) of a path stage: [u_min|q_min], [u_max|q_max] */
RD double lb_of(const double *par, int j) { return j < 6 ? par[RP_UMIN + j] : par[RP_QMIN + j - 6]; }
RD double ub_of(const double *par, int j) { return j < 6 ? par[RP_UMAX + j] : par[RP_QMAX + j - 6]; }

/* Linearisation at the current iterate, stage-parallel: ocp_nlp_approximate_qp_matrices + _vectors_sqp
 * (AC/ocp_nlp/ocp_nlp_common.c:2983-3106) with NLS cost (ocp_nlp_cost_nls.c:830-1032), discrete dynamics
 * (ocp_nlp_dynamics_disc.c:611-708) and box bounds (ocp_nlp_constraints_bgh.c:1563-1619).
 * Writes per stage: Jc = sqrt(dt W_m) * d y5_m / d x, rqz = cost gradient, b = dynamics gap, d = [lb - v; v - ub];
 * dx0 = x0 - x_nlp[0] (both bound residuals of the stage-0 state equality).  Returns the total cost. */
RD double nlp_linearise(const RmpcChain &ch, const RmpcDims &D, const Inst &I, const double *x0)
{
    const int N = D.N, lane = rd_lane();
    const double dt = D.dt;
    const double *par = I.par;
    double cost = 0.0;
    for (int i = lane; i < N; i += 32)
    {
        double v[NV], y5[5], J[60];
        load18(I.ux + NV * i, v);
        ur10_task<true>(ch, par, v + 6, v + 12, y5, J);
        cost += stage_cost(par, dt, y5, v, v + 12);
        const double res[5] = {y5[0], y5[1] - 1.0, y5[2], y5[3] - par[RP_PXREF], y5[4] - par[RP_VYREF]};
        double wr[5], sw[5];
        RD_UNROLL
        for (int m = 0; m < 5; m++)
        {
            wr[m] = dt * par[RP_WTASK + m] * res[m];
            sw[m] = sqrt(dt * par[RP_WTASK + m]);
        }
        double *g = I.rqz + NV * i, *Jc = I.Jc + RMPC_NJ * i;
        const double wq = par[RP_WQDD], wu = 2.0 * par[RP_WU];
        RD_UNROLL
        for (int k = 0; k < 6; k++)
        {
            const double cu = par[RP_B2 + k] / dt, cq = (par[RP_A22 + k] - 1.0) / dt;
            const double a = cu * v[k] + cq * v[12 + k]; /* qddot_k */
            g[k] = dt * (wu * v[k] + wq * cu * a);
            double gq = 0.0, gqd = dt * (wq * cq * a);
            RD_UNROLL
            for (int m = 0; m < 5; m++)
            {
                gq += J[12 * m + k] * wr[m];
                gqd += J[12 * m + 6 + k] * wr[m];
                Jc[12 * m + k] = sw[m] * J[12 * m + k];
                Jc[12 * m + 6 + k] = sw[m] * J[12 * m + 6 + k];
            }
            g[6 + k] = gq;
            g[12 + k] = gqd;
        }
        /* dynamics gap b = Ad x + Bd u - x_next */
        const double *xn = I.ux + NV * (i + 1) + 6;
        double *b = I.b + NX * i, *d = I.d + NB * i;
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
            else { d[j] = lb_of(par, j) - v[j]; d[12 + j] = v[j] - ub_of(par, j); }
        }
        if (i == 0)
        {
            RD_UNROLL
            for (int k = 0; k < NX; k++) I.dx0[k] = x0[k] - v[6 + k];
        }
    }
    if (lane < NV) I.rqz[NV * N + lane] = 0.0; /* stage N: no cost (ny_e = 0) */
    rd_sync();
    return warp_sum(cost);
}

/* Merit function at the trial iterate tux, ocp_nlp_evaluate_merit_fun
 * (AC/ocp_nlp/ocp_nlp_globalization_merit_backtracking.c:606-702):
 *   M = cost + sum |w_pi| |dyn gap| + sum |w_lam| max(0, bound violation)   (incl. the stage-0 x0 equality) */
RD double nlp_merit(const RmpcChain &ch, const RmpcDims &D, const Inst &I, const double *tux, const double *x0)
{
    const int N = D.N, lane = rd_lane();
    const double *par = I.par;
    double m = 0.0;
    for (int i = lane; i < N; i += 32)
    {
        double v[NV], y5[5];
        load18(tux + NV * i, v);
        ur10_task<false>(ch, par, v + 6, v + 12, y5, nullptr);
        m += stage_cost(par, D.dt, y5, v, v + 12);
        const double *xn = tux + NV * (i + 1) + 6, *wpi = I.wpi + NX * i, *wl = I.wlam + NB * i;
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
            const double fl = lb_of(par, j) - v[j], fu = v[j] - ub_of(par, j);
            if (fl > 0.0) m += fabs(wl[j]) * fl;
            if (fu > 0.0) m += fabs(wl[12 + j]) * fu;
        }
        if (i == 0)
        {
            RD_UNROLL
            for (int k = 0; k < NX; k++)
            {
                const double fl = x0[k] - v[6 + k];
                if (fl > 0.0) m += fabs(I.wlamx0[k]) * fl;
                if (-fl > 0.0) m += fabs(I.wlamx0[12 + k]) * (-fl);
            }
        }
    }
    return warp_sum(m);
}

/* KKT residual inf-norms of the NLP, ocp_nlp_res_compute (AC/ocp_nlp/ocp_nlp_common.c:3641-3751); uses the
 * linearisation (rqz, b, d, dx0) of the current iterate.  res = [stat, eq, ineq, comp]. */
RD void nlp_residuals(const RmpcDims &D, const Inst &I, double *res)
{
    const int N = D.N, lane = rd_lane();
    const double *par = I.par;
    double rs = 0, re = 0, ri = 0, rc = 0;
    for (int i = lane; i <= N; i += 32)
    {
        double r[NV];
        RD_UNROLL
        for (int k = 0; k < NV; k++) r[k] = I.rqz[NV * i + k];
        if (i < N)
        {
            const double *lam = I.lam + NB * i, *pi = I.pi + NX * i, *d = I.d + NB * i;
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
                re = rd_fmax(re, rd_fmax(fabs(I.b[NX * i + k]), fabs(I.b[NX * i + 6 + k])));
            }
        }
        if (i == 0)
        {
            RD_UNROLL
            for (int k = 0; k < NX; k++)
            {
                const double ll = I.lamx0[k], lu = I.lamx0[12 + k], f = I.dx0[k];
                r[6 + k] -= ll - lu;
                ri = rd_fmax(ri, rd_fmax(f, -f));
                rc = rd_fmax(rc, rd_fmax(fabs(ll * f), fabs(lu * f)));
            }
        }
        else
        {
            RD_UNROLL
            for (int k = 0; k < NX; k++) r[6 + k] -= I.pi[NX * (i - 1) + k];
        }
        RD_UNROLL
        for (int k = 0; k < NV; k++)
            if (i < N || k >= 6) rs = rd_fmax(rs, fabs(r[k]));
        if (r[6] != r[6]) rs = r[6]; /* propagate NaN like the reference's isnan test */
    }
    res[0] = warp_max(rs); res[1] = warp_max(re); res[2] = warp_max(ri); res[3] = warp_max(rc);
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
        const double g = I.rqz[lane];
        I.g0u[lane] = g;
        I.rqz[lane] = g + c.hs * I.dx0[6 + lane];
        I.b[lane] += I.dx0[lane] + par[RP_A12 + lane] * I.dx0[6 + lane];
        I.b[6 + lane] += par[RP_A22 + lane] * I.dx0[6 + lane];
    }
    rd_sync();
}

/* OCP_QP_INIT_VAR (H/ocp_qp/x_ocp_qp_ipm.c:1592-1980).  warm = 2: keep pi, clip lam, t at 0.1 (:1638-1654).
 * warm = 0 (every QP but the first, see rmpc_oracle.c:solve_qp for the reference quirk): ux = 0, pi = 0,
 * t = slack pushed to >= 0.1 by moving ux, lam = mu0 / t with mu0 = 1 (t0_init 2, var_init_scheme 1, :1837-1978).
 * The primal iterate is zeroed first in both cases (AC/ocp_qp/ocp_qp_hpipm.c:309-316). */
RD void ipm_init(const RmpcDims &D, const Inst &I, int warm)
{
    const int N = D.N, lane = rd_lane();
    for (int k = lane; k < (N + 1) * NV; k += 32) I.sux[k] = 0.0;
    if (warm != 2)
        for (int k = lane; k < N * NX; k += 32) I.spi[k] = 0.0;
    rd_sync();
    for (int k = lane; k < N * 12; k += 32)
    {
        const int i = k / 12, j = k - 12 * i;
        if (i == 0 && j >= 6) continue;
        double *lam = I.slam + NB * i, *t = I.st + NB * i;
        if (warm == 2)
        {
            if (lam[j] < 1e-1) lam[j] = 1e-1;
            if (lam[12 + j] < 1e-1) lam[12 + j] = 1e-1;
            if (t[j] < 1e-1) t[j] = 1e-1;
            if (t[12 + j] < 1e-1) t[12 + j] = 1e-1;
        }
        else
        {
            const double thr0 = 1e-1, dl = I.d[NB * i + j], du = I.d[NB * i + 12 + j];
            double v = 0.0, tl = -dl, tu = -du;
            if (tl < thr0)
            {
                if (tu < thr0) { v = 0.5 * (dl - du); tl = thr0; tu = thr0; }
                else { tl = thr0; v = dl + thr0; }
            }
            else if (tu < thr0) { tu = thr0; v = -du - thr0; }
            I.sux[NV * i + j] = v;
            t[j] = tl; t[12 + j] = tu;
            lam[j] = 1.0 / tl; lam[12 + j] = 1.0 / tu;
        }
    }
    rd_sync();
}

/* QP residuals, stage-parallel.  lin = 0: OCP_QP_RES_COMPUTE (H/ocp_qp/x_ocp_qp_res.c:345-533) at (ux,pi,lam,t), m = 0;
 * lin = 1: OCP_QP_RES_COMPUTE_LIN (:535-685), residual of the linearised KKT system at the step (ux,pi,lam,t) around
 * (Lam,T) with right-hand sides (rqz,b,d,m).  Outputs the four inf-norms (OCP_QP_RES_COMPUTE_INF_NORM :689-726) in
 * nrm[0..3], nrm[4] = |res_m - tau_min|_inf, and returns sum lam*t (lin = 0).  rg[0] NaN is reported in nrm[5]. */
RD double ipm_res(const RmpcDims &D, const Inst &I, int lin, const double *rqz, const double *b, const double *d, const double *m,
                  const double *ux, const double *pi, const double *lam, const double *t, const double *Lam, const double *T,
                  double *rg, double *rb, double *rd, double *rm, double *nrm)
{
    const int N = D.N, lane = rd_lane();
    const double *par = I.par;
    double mu = 0.0, n0 = 0, n1 = 0, n2 = 0, n3 = 0, n4 = 0, nanflag = 0;
    for (int i = lane; i <= N; i += 32)
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
            if (i > 0)
            {
                const double *Jc = I.Jc + RMPC_NJ * i;
                RD_UNROLL
                for (int mm = 0; mm < 5; mm++)
                {
                    double w = 0.0;
                    RD_UNROLL
                    for (int k = 0; k < 12; k++) w += Jc[12 * mm + k] * v[6 + k];
                    RD_UNROLL
                    for (int k = 0; k < 12; k++) g[6 + k] += Jc[12 * mm + k] * w;
                }
            }
            const double *pin = pi + NX * i, *l = lam + NB * i, *tt = t + NB * i;
            const double *xn = ux + NV * (i + 1) + 6;
            RD_UNROLL
            for (int k = 0; k < 6; k++)
            {
                const StageConst c = stage_const(par, D.dt, k);
                const double a12 = par[RP_A12 + k], a22 = par[RP_A22 + k], b1 = par[RP_B1 + k], b2 = par[RP_B2 + k];
                g[k] += c.hu * v[k] + c.hs * v[12 + k] + b1 * pin[k] + b2 * pin[6 + k];
                if (i > 0)
                {
                    g[6 + k] += pin[k];
                    g[12 + k] += c.hs * v[k] + c.hq * v[12 + k] + a12 * pin[k] + a22 * pin[6 + k];
                }
                const double r0 = b[NX * i + k] - xn[k] + (v[6 + k] + a12 * v[12 + k]) + b1 * v[k];
                const double r1 = b[NX * i + 6 + k] - xn[6 + k] + a22 * v[12 + k] + b2 * v[k];
                rb[NX * i + k] = r0; rb[NX * i + 6 + k] = r1;
                n1 = rd_fmax(n1, rd_fmax(fabs(r0), fabs(r1)));
            }
            RD_UNROLL
            for (int j = 0; j < 12; j++)
            {
                if (i == 0 && j >= 6) continue;
                g[j] += l[12 + j] - l[j];
                const double dl = d[NB * i + j] + tt[j] - v[j], du = d[NB * i + 12 + j] + tt[12 + j] + v[j];
                rd[NB * i + j] = dl; rd[NB * i + 12 + j] = du;
                double ml, mu_;
                if (!lin) { ml = l[j] * tt[j]; mu_ = l[12 + j] * tt[12 + j]; mu += ml + mu_; }
                else
                {
                    ml = m[NB * i + j] + Lam[NB * i + j] * tt[j] + l[j] * T[NB * i + j];
                    mu_ = m[NB * i + 12 + j] + Lam[NB * i + 12 + j] * tt[12 + j] + l[12 + j] * T[NB * i + 12 + j];
                }
                rm[NB * i + j] = ml; rm[NB * i + 12 + j] = mu_;
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
        RD_UNROLL
        for (int k = 0; k < NV; k++)
        {
            rg[NV * i + k] = g[k];
            if ((i < N || k >= 6) && (i > 0 || k < 6)) n0 = rd_fmax(n0, fabs(g[k]));
        }
        if (i == 0 && g[0] != g[0]) nanflag = 1.0;
    }
    rd_sync();
    nrm[0] = warp_max(n0); nrm[1] = warp_max(n1); nrm[2] = warp_max(n2); nrm[3] = warp_max(n3); nrm[4] = warp_max(n4);
    nrm[5] = warp_max(nanflag);
    return warp_sum(mu);
}

#ifdef RMPC_EMU
struct rd_double2 { double x, y; };
#else
typedef double2 rd_double2;
#endif

/* Backward square-root Riccati factorisation, first half of OCP_QP_FACT_SOLVE_KKT_STEP (H/ocp_qp/x_ocp_qp_kkt.c:
 * 836-966) incl. COMPUTE_GAMMA_GAMMA_QP (H/ipm_core/x_core_qp_ipm_aux.c:38-86): for ss = N..0
 *   L_ss = chol( [RSQ + reg I + diag(Gamma_lb + Gamma_ub), (res_g + gamma_lb - gamma_ub)'] + AL AL' ),
 *   AL = [B A b]' Lxx_{ss+1}  (last row + l_{ss+1,x}),  Pb_ss = P_{ss+1} res_b_ss.
 * Lane r owns row r of the 19 x 18 block (r = 18: the right-hand-side row).  BLASFEO's dtrmm / dsyrk / dpotrf on
 * panel-major AVX2 kernels become: a sparse row combination (A, B are diagonal blocks), a rank-17 update read as
 * broadcast pairs from shared memory (12 columns of AL + the 5 scaled task-Jacobian rows, which carry the whole
 * state Hessian), and a right-looking Cholesky with one column broadcast per step. */
RD void ipm_factor(const RmpcDims &D, const Inst &I, WarpSmem &sm)
{
    const int N = D.N, lane = rd_lane(), r = lane, jj = lane % 6;
    const double *par = I.par;
    const StageConst sc = stage_const(par, D.dt, jj);
    const double reg = 1e-15;
    double ca = 0.0, cb = 0.0; /* row r of [B;A]' = ca * e_{q_jj} + cb * e_{qd_jj} */
    if (r < 6) { ca = par[RP_B1 + jj]; cb = par[RP_B2 + jj]; }
    else if (r < 12) { ca = 1.0; }
    else if (r < 18) { ca = par[RP_A12 + jj]; cb = par[RP_A22 + jj]; }
    /* stage N: zero Hessian + reg => L_N = sqrt(reg) I, l_N = res_g_N / sqrt(reg)  (POTRF_L_MN of :871-879) */
    {
        const double sreg = sqrt(reg), isreg = 1.0 / sreg;
        double *LN = I.L + (long) LS * N;
        if (lane < 12)
        {
            const double lv = I.res_g[NV * N + 6 + lane] * isreg;
            RD_UNROLL
            for (int c = 0; c < 12; c++) sm.Lx[lane * SM_LXS + c] = (c == lane) ? sreg : 0.0;
            sm.lx[lane] = lv;
            for (int c = 0; c <= lane; c++) LEL(LN, 6 + lane, 6 + c) = (c == lane) ? sreg : 0.0;
            LEL(LN, 18, 6 + lane) = lv;
        }
    }
    for (int ss = N - 1; ss >= 0; ss--)
    {
        const bool s0 = (ss == 0);
        /* (1) right-hand sides and barrier terms of this stage */
        double Gam = 0.0, gam = 0.0;
        if (lane < 12 && !(s0 && lane >= 6))
        {
            const double *lam = I.slam + NB * ss, *t = I.st + NB * ss, *rd = I.res_d + NB * ss, *rm = I.res_m + NB * ss;
            const double til = 1.0 / t[lane], tiu = 1.0 / t[12 + lane];
            Gam = til * lam[lane] + tiu * lam[12 + lane];
            gam = til * (rm[lane] - lam[lane] * rd[lane]) - tiu * (rm[12 + lane] - lam[12 + lane] * rd[12 + lane]);
        }
        if (lane < 12) sm.bv[lane] = I.res_b[NX * ss + lane];
        if (lane < 18) sm.tv[lane] = ((s0 && lane >= 6) ? 0.0 : I.res_g[NV * ss + lane]) + gam;
        rd_sync();
        /* (2) w = Lxx' b ; Pb = Lxx w */
        if (lane < 12)
        {
            double a = 0.0;
            for (int k = lane; k < 12; k++) a += sm.Lx[k * SM_LXS + lane] * sm.bv[k];
            sm.wv[lane] = a;
        }
        rd_sync();
        if (lane < 12)
        {
            double a = 0.0;
            for (int k = 0; k <= lane; k++) a += sm.Lx[lane * SM_LXS + k] * sm.wv[k];
            I.Pb[NX * ss + lane] = a;
        }
        /* (3) own row of AL */
        double al[18];
        {
            const double *La = sm.Lx + jj * SM_LXS, *Lb = sm.Lx + (6 + jj) * SM_LXS;
            RD_UNROLL
            for (int c = 0; c < 12; c++) al[c] = ca * La[c] + cb * Lb[c];
            if (r == 18)
            {
                RD_UNROLL
                for (int c = 0; c < 12; c++) al[c] = sm.wv[c] + sm.lx[c];
            }
            const bool task = (r >= 6 && r < 18 && !s0);
            const double *Jc = I.Jc + RMPC_NJ * ss + (task ? r - 6 : 0);
            RD_UNROLL
            for (int mm = 0; mm < 5; mm++) al[12 + mm] = task ? Jc[12 * mm] : 0.0;
            al[17] = 0.0;
            if (r <= 18)
            {
                RD_UNROLL
                for (int c = 0; c < 18; c++) sm.AL[r * SM_ALS + c] = al[c];
            }
        }
        rd_sync();
        /* (4) row r of  H + AL AL'  (only columns <= r are meaningful) */
        double c[18];
        RD_UNROLL
        for (int cc = 0; cc < 18; cc++)
        {
            const rd_double2 *row = reinterpret_cast<const rd_double2 *>(sm.AL + cc * SM_ALS);
            double a0 = 0.0, a1 = 0.0;
            RD_UNROLL
            for (int k = 0; k < 9; k++)
            {
                const rd_double2 p = row[k];
                a0 += al[2 * k] * p.x;
                a1 += al[2 * k + 1] * p.y;
            }
            double a = a0 + a1;
            if (r == cc) a += reg + Gam + (r < 6 ? sc.hu : (r >= 12 ? sc.hq : 0.0));
            if (r == cc + 12 && r < 18) a += sc.hs;
            if (r == 18) a += sm.tv[cc];
            c[cc] = a;
        }
        /* (5) right-looking Cholesky of the 19 x ncol block (BLASFEO dpotrf_l_mn: non-positive pivot -> zero column) */
        const int ncol = s0 ? 6 : 18;
        int pb = 0;
        RD_UNROLL
        for (int j = 0; j < 18; j++)
        {
            if (j < ncol)
            {
                const double dj = rd_shfl(c[j], j);
                const double inv = dj > 0.0 ? rd_rsqrt(dj) : 0.0;
                const double lj = (r == j) ? dj * inv : c[j] * inv;
                c[j] = lj;
                if (r >= j && r <= 18) sm.col[pb][r] = lj;
                rd_sync();
                RD_UNROLL
                for (int k = j + 1; k < 18; k++) c[k] -= lj * sm.col[pb][k];
                pb ^= 1;
            }
        }
        /* (6) store the factor (packed columns: coalesced over rows) and stage Lxx, l_x for the next stage */
        double *Ls = I.L + (long) LS * ss;
        RD_UNROLL
        for (int cc = 0; cc < 18; cc++)
            if (r >= cc && r <= 18) LEL(Ls, r, cc) = (s0 && r >= 6 && r < 18) ? 0.0 : c[cc];
        if (!s0)
        {
            if (r >= 6 && r < 18)
            {
                RD_UNROLL
                for (int cc = 6; cc < 18; cc++) sm.Lx[(r - 6) * SM_LXS + (cc - 6)] = (cc <= r) ? c[cc] : 0.0;
            }
            if (r == 18)
            {
                RD_UNROLL
                for (int cc = 6; cc < 18; cc++) sm.lx[cc - 6] = c[cc];
            }
        }
        rd_sync();
    }
}

/* Forward substitution of the Riccati recursion (H/ocp_qp/x_ocp_qp_kkt.c:970-1005 and :1671-1721):
 *   du = -Luu^-T (src_u + Lxu' dx),  dx+ = A dx + B du + b      with src_u = l_u (row 18 of L, after the
 * factorisation) or the z_u left in dux by ipm_backward().  The multipliers dpi are computed afterwards by
 * ipm_dpi(), stage-parallel, because nothing in this recursion depends on them. */
RD void ipm_forward(const RmpcDims &D, const Inst &I, WarpSmem &sm, const double *b, double *dux, int from_stored)
{
    const int N = D.N, lane = rd_lane(), jj = lane % 6;
    const double *par = I.par;
    const double a12 = par[RP_A12 + jj], a22 = par[RP_A22 + jj], b1 = par[RP_B1 + jj], b2 = par[RP_B2 + jj];
    if (lane < 12) { sm.xv[lane] = 0.0; dux[6 + lane] = 0.0; }
    rd_sync();
    for (int ss = 0; ss < N; ss++)
    {
        const double *Ls = I.L + (long) LS * ss;
        double rhs = 0.0, inv = 0.0;
        if (lane < 6)
        {
            const double *colp = Ls + lcol(lane) - lane; /* colp[r] = L[r][lane] */
            double t = from_stored ? dux[NV * ss + lane] : colp[18];
            if (ss > 0)
                for (int rr = 6; rr < 18; rr++) t += colp[rr] * sm.xv[rr - 6];
            rhs = -t;
            inv = 1.0 / colp[lane];
        }
        double z = 0.0;
        for (int k = 5; k >= 0; k--)
        {
            const double zk = rd_shfl(rhs * inv, k);
            if (lane < k) rhs -= Ls[lcol(lane) + k - lane] * zk;
            if (lane == k) z = zk;
        }
        if (lane < 6) { sm.uv[lane] = z; dux[NV * ss + lane] = z; }
        rd_sync();
        double xn = 0.0;
        if (lane < 6) xn = (sm.xv[jj] + a12 * sm.xv[6 + jj]) + b1 * sm.uv[jj] + b[NX * ss + lane];
        else if (lane < 12) xn = a22 * sm.xv[6 + jj] + b2 * sm.uv[jj] + b[NX * ss + lane];
        rd_sync();
        if (lane < 12) { sm.xv[lane] = xn; dux[NV * (ss + 1) + 6 + lane] = xn; }
        rd_sync();
    }
    if (lane < 6) dux[NV * N + lane] = 0.0;
    rd_sync();
}

/* Backward substitution with stored factors, first half of OCP_QP_SOLVE_KKT_STEP (H/ocp_qp/x_ocp_qp_kkt.c:
 * 1543-1665) incl. COMPUTE_GAMMA_QP (x_core_qp_ipm_aux.c:133-160): leaves z_u = Luu^-1 r_u in dux[.][0:6] and the
 * Riccati vectors p_ss in pvec. */
RD void ipm_backward(const RmpcDims &D, const Inst &I, WarpSmem &sm, const double *rqz, const double *b, const double *d,
                     const double *m, double *dux, int use_Pb)
{
    const int N = D.N, lane = rd_lane(), r = lane, jj = lane % 6;
    const double *par = I.par;
    double ca = 0.0, cb = 0.0;
    if (r < 6) { ca = par[RP_B1 + jj]; cb = par[RP_B2 + jj]; }
    else if (r < 12) { ca = 1.0; }
    else if (r < 18) { ca = par[RP_A12 + jj]; cb = par[RP_A22 + jj]; }
    if (lane < 12)
    {
        const double pv = rqz[NV * N + 6 + lane];
        sm.xv[lane] = pv;
        I.pvec[NX * N + lane] = pv;
    }
    rd_sync();
    for (int ss = N - 1; ss >= 0; ss--)
    {
        const bool s0 = (ss == 0);
        const double *Ls = I.L + (long) LS * ss;
        double rr = 0.0;
        if (r < 18 && !(s0 && r >= 6)) rr = rqz[NV * ss + r];
        if (lane < 12 && !(s0 && lane >= 6))
        {
            const double *lam = I.slam + NB * ss, *t = I.st + NB * ss, *dd = d + NB * ss, *mm = m + NB * ss;
            rr += (mm[lane] - lam[lane] * dd[lane]) / t[lane] - (mm[12 + lane] - lam[12 + lane] * dd[12 + lane]) / t[12 + lane];
        }
        if (use_Pb)
        {
            if (lane < 12) sm.tv[lane] = sm.xv[lane] + I.Pb[NX * ss + lane];
        }
        else
        {
            /* P_{ss+1} b = Lxx (Lxx' b) from the stored factor of stage ss+1 */
            const double *Ln = I.L + (long) LS * (ss + 1);
            if (lane < 12)
            {
                double a = 0.0;
                for (int k = lane; k < 12; k++) a += LEL(Ln, 6 + k, 6 + lane) * b[NX * ss + k];
                sm.wv[lane] = a;
            }
            rd_sync();
            if (lane < 12)
            {
                double a = 0.0;
                for (int k = 0; k <= lane; k++) a += LEL(Ln, 6 + lane, 6 + k) * sm.wv[k];
                sm.tv[lane] = sm.xv[lane] + a;
            }
        }
        rd_sync();
        rr += ca * sm.tv[jj] + cb * sm.tv[6 + jj];
        /* TRSV_LNN_MN(18, 6): z_u = Luu^-1 r_u, r_x -= Lxu z_u */
        const double inv = (lane < 6) ? 1.0 / Ls[lcol(lane)] : 0.0;
        for (int k = 0; k < 6; k++)
        {
            const double zk = rd_shfl(rr * inv, k);
            if (r > k && r < 18) rr -= Ls[lcol(k) + r - k] * zk;
            if (r == k) rr = zk;
        }
        rd_sync(); /* everyone has read tv / xv */
        if (r < 6) dux[NV * ss + r] = rr;
        else if (r < 18 && !s0)
        {
            sm.xv[r - 6] = rr;
            I.pvec[NX * ss + r - 6] = rr;
        }
        rd_sync();
    }
}

/* dpi_i = P_{i+1} dx_{i+1} + p_{i+1}, stage-parallel (H/ocp_qp/x_ocp_qp_kkt.c:978-984 after the factorisation:
 * p = Lxx l_x; :1683-1688 after a re-solve: p = pvec). */
RD void ipm_dpi(const RmpcDims &D, const Inst &I, const double *dux, double *dpi, int from_stored)
{
    const int N = D.N, lane = rd_lane();
    for (int i = lane; i < N; i += 32)
    {
        const double *Ln = I.L + (long) LS * (i + 1), *x = dux + NV * (i + 1) + 6;
        double w[12];
        RD_UNROLL
        for (int c = 0; c < 12; c++)
        {
            double a = from_stored ? 0.0 : LEL(Ln, 18, 6 + c);
            RD_UNROLL
            for (int k = c; k < 12; k++) a += LEL(Ln, 6 + k, 6 + c) * x[k];
            w[c] = a;
        }
        RD_UNROLL
        for (int rr = 0; rr < 12; rr++)
        {
            double a = from_stored ? I.pvec[NX * (i + 1) + rr] : 0.0;
            RD_UNROLL
            for (int k = 0; k <= rr; k++) a += LEL(Ln, 6 + rr, 6 + k) * w[k];
            dpi[NX * i + rr] = a;
        }
    }
    rd_sync();
}

/* dt = +-dux[idxb], then COMPUTE_LAM_T_QP (H/ipm_core/x_core_qp_ipm_aux.c:164-189) */
RD void ipm_dlam_dt(const RmpcDims &D, const Inst &I, const double *d, const double *m, const double *dux, double *dlam, double *dt)
{
    const int N = D.N, lane = rd_lane();
    for (int k = lane; k < N * 12; k += 32)
    {
        const int i = k / 12, j = k - 12 * i;
        if (i == 0 && j >= 6) continue;
        const double du = dux[NV * i + j];
        const int a = NB * i + j, bq = a + 12;
        const double la = I.slam[a], lb = I.slam[bq];
        dlam[a] = -(m[a] + (la * du) - (la * d[a])) / I.st[a];
        dlam[bq] = -(m[bq] + (lb * (-du)) - (lb * d[bq])) / I.st[bq];
        dt[a] = du - d[a];
        dt[bq] = -du - d[bq];
    }
    rd_sync();
}

/* COMPUTE_ALPHA_QP (x_core_qp_ipm_aux.c:193-263): largest step in (0,1] keeping lam, t >= 0 (single step, split_step 0) */
RD double ipm_alpha(const RmpcDims &D, const Inst &I, const double *dlam, const double *dt)
{
    const int N = D.N, lane = rd_lane();
    double a = 1.0;
    for (int k = lane; k < N * NB; k += 32)
    {
        const int i = k / NB, j = k - NB * i;
        if (i == 0 && (j % 12) >= 6) continue;
        const double dl = dlam[k], dtt = dt[k];
        if (-a * dl > I.slam[k]) a = -I.slam[k] / dl;
        if (-a * dtt > I.st[k]) a = -I.st[k] / dtt;
    }
    return -warp_max(-a);
}

/* COMPUTE_MU_AFF_QP (x_core_qp_ipm_aux.c:424-441) */
RD double ipm_mu_aff(const RmpcDims &D, const Inst &I, double alpha, double nc_inv)
{
    const int N = D.N, lane = rd_lane();
    double mu = 0.0;
    for (int k = lane; k < N * NB; k += 32)
    {
        const int i = k / NB, j = k - NB * i;
        if (i == 0 && (j % 12) >= 6) continue;
        mu += (I.slam[k] + alpha * I.dlam[k]) * (I.st[k] + alpha * I.dt_[k]);
    }
    return warp_sum(mu) * nc_inv;
}

/* d_ocp_qp_ipm_solve (H/ocp_qp/x_ocp_qp_ipm.c:2626-3018) with OCP_QP_IPM_DELTA_STEP (:2155-2622) for the option set
 * acados resolves: Mehrotra predictor-corrector, conditional second centering solve, up to two
 * refinement steps on the corrector, single damped step, lam/t floor 1e-16.  Returns the acados QP status
 * (0 ok, 2 max iter, 3 min step, 1 NaN).  The Householder-LQ fall-back of lq_fact = 1 is not implemented: its
 * trigger (linear-system residual > 1e-5, :2246-2276) is evaluated and counted in flags[2] so that any instance
 * that would have taken it is visible to the caller. */
RD int ipm_solve_qp(const RmpcDims &D, const Inst &I, WarpSmem &sm, int warm, int *iters)
{
    const int N = D.N, lane = rd_lane();
    const double tol = D.qp_tol, alpha_min = 1e-8, tau_min = 1e-16, lmin = 1e-16;
    const double nc_inv = 1.0 / (double) (NB * (N - 1) + 12);
    ipm_init(D, I, warm);
    double nrm[6], inrm[6], alpha = 1.0;
    double mu = nc_inv * ipm_res(D, I, 0, I.rqz, I.b, I.d, nullptr, I.sux, I.spi, I.slam, I.st, nullptr, nullptr, I.res_g, I.res_b,
                                 I.res_d, I.res_m, nrm);
    int kk;
    for (kk = 0; kk < D.qp_iter_max && alpha > alpha_min && (nrm[0] > tol || nrm[1] > tol || nrm[2] > tol || nrm[4] > tol); kk++)
    {
        /* BACKUP_RES_M, COMPUTE_TAU_MIN_QP */
        for (int k = lane; k < N * NB; k += 32)
        {
            const double v = I.res_m[k];
            I.res_m_bkp[k] = v;
            I.res_m[k] = v - tau_min;
        }
        rd_sync();
        /* affine (predictor) step */
        ipm_factor(D, I, sm);
        ipm_forward(D, I, sm, I.res_b, I.dux, 0);
        ipm_dpi(D, I, I.dux, I.dpi, 0);
        ipm_dlam_dt(D, I, I.res_d, I.res_m, I.dux, I.dlam, I.dt_);
        ipm_res(D, I, 1, I.res_g, I.res_b, I.res_d, I.res_m, I.dux, I.dpi, I.dlam, I.dt_, I.slam, I.st, I.ig, I.ib, I.id, I.im, inrm);
#ifdef RMPC_EMU_DEBUG
        if (lane == 0 && kk == 0) printf("dbg st1: res_g6 %.6e dux6 %.6e dpi1_0 %.6e dpi0_0 %.6e dlam_lb %.6e dlam_ub %.6e lam %.3e %.3e t %.3e %.3e res_d %.3e %.3e res_m %.3e %.3e dt %.3e %.3e\n", I.res_g[NV+6], I.dux[NV+6], I.dpi[NX], I.dpi[0], I.dlam[NB+6], I.dlam[NB+18], I.slam[NB+6], I.slam[NB+18], I.st[NB+6], I.st[NB+18], I.res_d[NB+6], I.res_d[NB+18], I.res_m[NB+6], I.res_m[NB+18], I.dt_[NB+6], I.dt_[NB+18]);
        if (lane == 0 && kk == 0) for (int i = 0; i <= N; i++) { printf("  ig[%d]:", i); for (int k = 0; k < NV; k++) printf(" %.2e", I.ig[NV*i+k]); printf("\n"); }
        if (lane == 0) printf("ipm it %d: nrm %.3e %.3e %.3e %.3e | lin res after predictor %.3e %.3e %.3e %.3e\n", kk, nrm[0], nrm[1], nrm[2], nrm[4], inrm[0], inrm[1], inrm[2], inrm[3]);
#endif
        if ((inrm[0] == 0.0 && inrm[5] > 0.5) || inrm[0] > 1e-5 || inrm[1] > 1e-5 || inrm[2] > 1e-5 || inrm[3] > 1e-5)
        {
            if (lane == 0) I.flags[2] += 1;
        }
        alpha = ipm_alpha(D, I, I.dlam, I.dt_);
        double mu_aff = ipm_mu_aff(D, I, alpha, nc_inv);
        const double tmp = mu_aff / mu, sigma = tmp * tmp * tmp;
        double sigma_mu = sigma * mu;
        sigma_mu = sigma_mu > tau_min ? sigma_mu : tau_min;
        /* COMPUTE_CENTERING_CORRECTION_QP */
        for (int k = lane; k < N * NB; k += 32) I.res_m[k] = I.res_m_bkp[k] + I.dt_[k] * I.dlam[k] - sigma_mu;
        rd_sync();
        ipm_backward(D, I, sm, I.res_g, I.res_b, I.res_d, I.res_m, I.dux, 1);
        ipm_forward(D, I, sm, I.res_b, I.dux, 1);
        ipm_dpi(D, I, I.dux, I.dpi, 1);
        ipm_dlam_dt(D, I, I.res_d, I.res_m, I.dux, I.dlam, I.dt_);
        alpha = ipm_alpha(D, I, I.dlam, I.dt_);
        /* conditional predictor-corrector */
        const double mu_aff0 = mu_aff;
        mu_aff = ipm_mu_aff(D, I, alpha, nc_inv);
        if (mu_aff > 2.0 * mu_aff0)
        {
            for (int k = lane; k < N * NB; k += 32) I.res_m[k] = I.res_m_bkp[k] - sigma_mu; /* COMPUTE_CENTERING_QP */
            rd_sync();
            ipm_backward(D, I, sm, I.res_g, I.res_b, I.res_d, I.res_m, I.dux, 1);
            ipm_forward(D, I, sm, I.res_b, I.dux, 1);
            ipm_dpi(D, I, I.dux, I.dpi, 1);
            ipm_dlam_dt(D, I, I.res_d, I.res_m, I.dux, I.dlam, I.dt_);
            alpha = ipm_alpha(D, I, I.dlam, I.dt_);
        }
        /* iterative refinement of the corrector step, itref_corr_max = 2 */
        int refined = 0;
        for (int it = 0; it < 2; it++)
        {
            ipm_res(D, I, 1, I.res_g, I.res_b, I.res_d, I.res_m, I.dux, I.dpi, I.dlam, I.dt_, I.slam, I.st, I.ig, I.ib, I.id, I.im,
                    inrm);
            if ((inrm[0] < tol || inrm[0] < 1e-3 * nrm[0]) && (inrm[1] < tol || inrm[1] < 1e-3 * nrm[1]) &&
                (inrm[2] < tol || inrm[2] < 1e-3 * nrm[2]) && (inrm[3] < tol || inrm[3] < 1e-3 * nrm[3]))
                break;
            ipm_backward(D, I, sm, I.ig, I.ib, I.id, I.im, I.iux, 0);
            ipm_forward(D, I, sm, I.ib, I.iux, 1);
            ipm_dpi(D, I, I.iux, I.ipi, 1);
            ipm_dlam_dt(D, I, I.id, I.im, I.iux, I.ilam, I.it);
            refined = 1;
            for (int k = lane; k < (N + 1) * NV; k += 32) I.dux[k] += I.iux[k];
            for (int k = lane; k < N * NX; k += 32) I.dpi[k] += I.ipi[k];
            for (int k = lane; k < N * NB; k += 32) { I.dlam[k] += I.ilam[k]; I.dt_[k] += I.it[k]; }
            rd_sync();
        }
        if (refined) alpha = ipm_alpha(D, I, I.dlam, I.dt_);
        /* UPDATE_VAR_QP (x_core_qp_ipm_aux.c:267-372) */
        double a = alpha;
        if (a < 1.0) a = a * ((1.0 - a) * 0.99 + a * 0.9999999);
        for (int k = lane; k < (N + 1) * NV; k += 32) I.sux[k] += a * I.dux[k];
        for (int k = lane; k < N * NX; k += 32) I.spi[k] += a * I.dpi[k];
        for (int k = lane; k < N * NB; k += 32)
        {
            const int i = k / NB, j = k - NB * i;
            if (i == 0 && (j % 12) >= 6) continue;
            double l = I.slam[k] + a * I.dlam[k], t = I.st[k] + a * I.dt_[k];
            I.slam[k] = l <= lmin ? lmin : l;
            I.st[k] = t <= lmin ? lmin : t;
        }
        rd_sync();
        mu = nc_inv * ipm_res(D, I, 0, I.rqz, I.b, I.d, nullptr, I.sux, I.spi, I.slam, I.st, nullptr, nullptr, I.res_g, I.res_b,
                              I.res_d, I.res_m, nrm);
    }
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
        const double *Jc = I.Jc; /* stage 0 */
        const int jj = lane % 6;
        const StageConst c = stage_const(par, D.dt, jj);
        double tmp = I.rqz[6 + lane] + ((lane < 6) ? I.spi[jj] : par[RP_A12 + jj] * I.spi[jj] + par[RP_A22 + jj] * I.spi[6 + jj]);
        for (int mm = 0; mm < 5; mm++)
        {
            double w = 0.0;
            for (int k = 0; k < 12; k++) w += Jc[12 * mm + k] * I.dx0[k];
            tmp += Jc[12 * mm + lane] * w;
        }
        if (lane >= 6) tmp += c.hs * I.sux[jj] + c.hq * I.dx0[lane];
        I.qlamx0[lane] = tmp >= 0.0 ? tmp : 1e-16;
        I.qlamx0[12 + lane] = tmp >= 0.0 ? 1e-16 : -tmp;
        I.sux[6 + lane] = I.dx0[lane];
    }
    rd_sync();
}

/* ---------------------------------------------------------------------------------------------- SQP / RTI (K4)
 * MERIT_BACKTRACKING line search and iterate update: merit weights (Leineweber update,
 * AC/ocp_nlp/ocp_nlp_globalization_merit_backtracking.c:705-755), ocp_nlp_line_search (:290-408, alpha = 1, 0.7, ...
 * while alpha*0.7 > 0.05; the last alpha is taken untested), ocp_nlp_update_variables_sqp
 * (AC/ocp_nlp/ocp_nlp_common.c:3259-3318, full_step_dual = 0).  Returns 1 on a NaN/inf merit value. */
RD int nlp_line_search_update(const RmpcChain &ch, const RmpcDims &D, const Inst &I, int nlp_iter, double *alpha_out)
{
    const int N = D.N, lane = rd_lane();
    for (int k = lane; k < N * NX; k += 32)
    {
        const double t0 = fabs(I.spi[k]);
        if (nlp_iter == 0) I.wpi[k] = t0;
        else { const double t1 = 0.5 * (t0 + I.wpi[k]); I.wpi[k] = t0 > t1 ? t0 : t1; }
    }
    for (int k = lane; k < N * NB + NB; k += 32)
    {
        const bool x0s = k >= N * NB;
        const int kk = x0s ? k - N * NB : k;
        if (!x0s && kk < NB && (kk % 12) >= 6) continue;
        const double t0 = x0s ? I.qlamx0[kk] : I.slam[kk];
        double *w = x0s ? I.wlamx0 + kk : I.wlam + kk;
        if (nlp_iter == 0) *w = t0;
        else { const double t1 = 0.5 * (t0 + *w); *w = t0 > t1 ? t0 : t1; }
    }
    rd_sync();
    const double m0 = nlp_merit(ch, D, I, I.ux, I.z);
    double alpha = 1.0, m1 = 0.0;
    int accepted = 0;
    while (alpha * 0.7 > 0.05)
    {
        for (int k = lane; k < (N + 1) * NV; k += 32) I.tux[k] = I.ux[k] + alpha * I.sux[k];
        rd_sync();
        m1 = nlp_merit(ch, D, I, I.tux, I.z);
        const bool bad = (m1 != m1) || (m1 - m1 != 0.0);
        if (m1 < m0 && !bad) { accepted = 1; break; }
        alpha *= 0.7;
    }
    *alpha_out = alpha;
    if (!accepted && ((m1 != m1) || (m1 - m1 != 0.0))) return 1;
    for (int k = lane; k < (N + 1) * NV; k += 32) I.ux[k] = I.ux[k] + alpha * I.sux[k];
    for (int k = lane; k < N * NX; k += 32) I.pi[k] = (1.0 - alpha) * I.pi[k] + alpha * I.spi[k];
    for (int k = lane; k < N * NB; k += 32)
    {
        if (k < NB && (k % 12) >= 6) continue;
        I.lam[k] = (1.0 - alpha) * I.lam[k] + alpha * I.slam[k];
    }
    if (lane < NB) I.lamx0[lane] = (1.0 - alpha) * I.lamx0[lane] + alpha * I.qlamx0[lane];
    rd_sync();
    return 0;
}

/* ocp_qp_xcond_solve (AC/ocp_qp/ocp_qp_xcond_solver.c:518-576): reduce -> IPM -> restore.  warm start bookkeeping:
 * flags[1] != 0 until the first QP of the instance's lifetime has been solved (reference quirk, see ipm_init). */
RD int qp_solve(const RmpcDims &D, const Inst &I, WarpSmem &sm, int init_from_nlp, int *iters)
{
    const int N = D.N, lane = rd_lane();
    const int warm = I.flags[1] ? 2 : 0;
    rd_sync();
    if (lane == 0) I.flags[1] = 0;
    qp_reduce(D, I);
    if (init_from_nlp && warm == 2)
    {
        /* ocp_nlp_initialize_qp_from_nlp (AC/ocp_nlp/ocp_nlp_common.c:3320-3340) + ocp_qp_compute_t at ux = 0 */
        for (int k = lane; k < N * NX; k += 32) I.spi[k] = I.pi[k];
        for (int k = lane; k < N * NB; k += 32)
        {
            if (k < NB && (k % 12) >= 6) continue;
            I.slam[k] = I.lam[k];
            I.st[k] = -I.d[k];
        }
        rd_sync();
    }
    const int st = ipm_solve_qp(D, I, sm, warm, iters);
    qp_restore(D, I);
    return st;
}

struct StepOut
{
    int status, sqp_iter, qp_iter, qp_status;
    double res[4], alpha;
};

/* ocp_nlp_sqp_rti, PREPARATION_AND_FEEDBACK (AC/ocp_nlp/ocp_nlp_sqp_rti.c:455-646,1123-1175) */
RD void solve_rti(const RmpcChain &ch, const RmpcDims &D, const Inst &I, WarpSmem &sm, StepOut &o)
{
    nlp_linearise(ch, D, I, I.z);
    const int first = I.flags[0];
    rd_sync();
    int it = 0;
    const int qs = qp_solve(D, I, sm, first, &it);
    o.sqp_iter = 1; o.qp_iter = it; o.qp_status = qs;
    o.res[0] = o.res[1] = o.res[2] = o.res[3] = 0.0; /* rti_log_residuals = 0: the reference reports zeros */
    o.alpha = 0.0;
    if (qs != 0 && qs != 2) { o.status = 4; return; } /* ACADOS_QP_FAILURE, iterate untouched (:609-622) */
    nlp_line_search_update(ch, D, I, 1, &o.alpha);    /* globalisation status ignored (:627-635) */
    if (rd_lane() == 0) I.flags[0] = 0;
    rd_sync();
    o.status = 0;
}

/* ocp_nlp_sqp (AC/ocp_nlp/ocp_nlp_sqp.c:474-801) with check_termination (:351-443) */
RD void solve_sqp(const RmpcChain &ch, const RmpcDims &D, const Inst &I, WarpSmem &sm, StepOut &o)
{
    o.qp_iter = 0; o.qp_status = 0; o.alpha = 0.0;
    for (int it = 0; it <= D.max_iter; it++)
    {
        o.sqp_iter = it;
        double cost = 0.0;
        if (it != D.max_iter)
        {
            cost = nlp_linearise(ch, D, I, I.z);
            nlp_residuals(D, I, o.res);
        }
        const double *r = o.res;
        if (r[0] != r[0] || r[1] != r[1] || r[2] != r[2] || r[3] != r[3]) { o.status = 1; return; }
        if (it >= D.max_iter) { o.status = 2; return; }
        if (r[0] < D.tol[0] && r[1] < D.tol[1] && r[2] < D.tol[2] && r[3] < D.tol[3]) { o.status = 0; return; }
        if (cost <= -1e10) { o.status = 6; return; }
        int qi = 0;
        const int qs = qp_solve(D, I, sm, 0, &qi);
        o.qp_iter += qi; o.qp_status = qs;
        if (qs != 0 && qs != 2) { o.sqp_iter = it + 1; o.status = 4; return; }
        if (nlp_line_search_update(ch, D, I, it, &o.alpha)) { o.status = 1; return; }
    }
    o.status = 2;
}

/* ---------------------------------------------------------------------------------------------- closed loop (K5)
 * R/simulator.py:199-241: x = plant state; solver.set(0,'lbx'/'ubx',x); solve; u = solver.get(0,'u'); plant RK4.
 * Plant: zdot = [qd; -W qd + W u] (R/simulation_model.py:79-83) decouples per joint, RK4 (:111-117) on lanes 0..5. */
RD void closed_loop_warp(const RmpcChain &ch, const RmpcDims &D, const RmpcWs &W, long inst, WarpSmem &sm, int nsteps,
                         int advance = 1)
{
    const Inst I = inst_view(W, D.N, inst);
    const int lane = rd_lane(), N = D.N;
    const double *par = I.par;
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
        if (D.solver_type == RMPC_SQP_RTI) solve_rti(ch, D, I, sm, o);
        else solve_sqp(ch, D, I, sm, o);
        /* solver.get_cost(): cost at the returned iterate (T/acados_ocp_solver.py:1664-1679) */
        double cost = 0.0;
        for (int i = lane; i < N; i += 32)
        {
            double v[NV], y5[5];
            load18(I.ux + NV * i, v);
            ur10_task<false>(ch, par, v + 6, v + 12, y5, nullptr);
            cost += stage_cost(par, D.dt, y5, v, v + 12);
        }
        cost = warp_sum(cost);
        double zq = 0.0, zqd = 0.0, u = 0.0;
        if (lane < 6)
        {
            u = I.ux[lane];
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
                sl[RL_ALPHA] = o.alpha; sl[RL_QP_STATUS] = o.qp_status;
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
    }
}
