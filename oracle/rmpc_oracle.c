/* oracle/rmpc_oracle.c -- plain-C restatement ("port") of the reference's hot path (TEST INFRASTRUCTURE).
 *
 * Nothing under oracle/ is on the product path: only tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline / --impl reference legs may use it, and only as the checker / reported CPU baseline.
 *
 * What is restated: the acados SQP / SQP_RTI closed loop on the UR10 surface-tracking OCP with
 * PARTIAL_CONDENSING_HPIPM (cond_N = N), for the ONE option set the reference uses.
 * Parity of this file is PINNED against the reference's own compiled code: tests/test_oracle_port.py
 * runs it next to oracle/_ref/librmpc_ref.so (unmodified acados + HPIPM + BLASFEO built by oracle/ref.mk,
 * driven by oracle/ref_driver.c) on the same inputs, and against the committed tests/golden vectors that
 * the reference leg generated.  The model functions (ur10_model.c) are the only part with no reference
 * artefact to compare with -- see ur10_model.h.
 *
 * Shorthand in citations: AC/ = acados/acados, H/ = acados/external/hpipm, R/ = reference root.
 *
 * Stage layout (as in acados/HPIPM): stage variable v_i = [u_i(6); x_i(12)], i < N;  v_N = x_N.
 * QP handed to the IPM = "reduced" QP after x0 elimination (H/ocp_qp/x_ocp_qp_red.c:278-466):
 *   stage 0: nu=6 nx=0 nb=6 ; stages 1..N-1: nu=6 nx=12 nb=12 (u, q) ; stage N: nu=0 nx=12 nb=0.
 */
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "rmpc_problem.h"

#define NX 12
#define NU 6
#define NV 18
#define NY 18
#define LD 19 /* leading dimension of the (nv+1) x nv stage blocks */

/* ------------------------------------------------------------------------------------------------ */
typedef struct
{
    /* problem */
    rmpc_problem prob;
    int N;
    double Wd[NY], yref[NY];
    double BAbt[NV * NX]; /* (nu+nx) x nx, column-major ld 18, identical for all stages (LTI) */
    double x0[NX];        /* lbx_0 = ubx_0 */
    double term_c[NX];    /* gradient 2 P x_N of the terminal row h_e = x_N' P x_N at the current iterate (DCt of stage N) */

    /* NLP iterate (ocp_nlp_out): ux[i] = [u;x] (stage N: x at offset 0), pi[i], lam[i] = [lower; upper] */
    double *ux, *pi, *lam;
    /* linearisation (ocp_nlp_memory / qp_in) */
    double *RSQ;  /* (N+1) x 18x18 lower, col-major ld 18   qp_in->RSQrq */
    double *rqz;  /* (N+1) x 18                              qp_in->rqz = cost_grad */
    double *b;    /* N x 12                                  qp_in->b = dyn_fun */
    double *d;    /* (N+1) x 36                              qp_in->d = ineq_fun */
    double *cost_fun; /* N+1 */
    /* full-space QP solution (ocp_qp_out) */
    double *qux, *qpi, *qlam, *qt;
    /* reduced QP data + IPM iterate that persists between calls (xcond_qp_out: warm start) */
    double *rRSQ, *rrqz, *rb, *rd; /* stage 0 differs from the full QP, the rest are copies */
    double *sux, *spi, *slam, *st;
    /* IPM workspace */
    double *dux, *dpi, *dlam, *dt, *res_g, *res_b, *res_d, *res_m, *res_m_bkp, *Gamma, *gamma, *tinv;
    double *L, *Pb;                 /* (N+1) x 19x18 ; N x 12 */
    double *ig, *ib, *id, *im;      /* res_itref */
    double *iux, *ipi, *ilam, *it;  /* sol_itref */
    /* merit weights (persist for the solver's lifetime) and trial iterate */
    double *wpi, *wlam, *tux;
    /* flags, stats */
    int first_call;
    int qp_warm_start; /* 2 for the first QP of the solver's lifetime, 0 afterwards -- see solve_qp() */
    int sqp_iter, qp_iter_sum, qp_iter_last, qp_status_last, n_lq_needed;
    double res[4], cost_value, alpha_ls;
    /* last-QP statistics for parity debugging: per IPM iteration alpha, mu, sigma */
    double stat[64 * 8];
} rmpc_oracle;

static int nu_of(const rmpc_oracle *s, int i) { return i < s->N ? NU : 0; }
static int nb_full(const rmpc_oracle *s, int i) { return i == 0 ? 18 : (i < s->N ? 12 : 0); }
/* reduced QP dims */
static int rnu(const rmpc_oracle *s, int i) { return i < s->N ? NU : 0; }
static int rnx(const rmpc_oracle *s, int i) { (void) s; return i == 0 ? 0 : NX; }
static int rnb(const rmpc_oracle *s, int i) { return i == 0 ? 6 : (i < s->N ? 12 : 0); }
/* general rows of the QP: the linearised terminal ellipsoid (nh_e = 1 -> ng = 1 at stage N); per-stage constraint vectors
 * are laid out [lb(nb), lg(ng); ub(nb), ug(ng)] like acados / HPIPM do */
static int ng_of(const rmpc_oracle *s, int i) { return (i == s->N && s->prob.term) ? 1 : 0; }
static int nc_full(const rmpc_oracle *s, int i) { return nb_full(s, i) + ng_of(s, i); }
static int rnc(const rmpc_oracle *s, int i) { return rnb(s, i) + ng_of(s, i); }
static double term_h(const rmpc_oracle *s, const double *x, double *grad)
{
    double h = 0.0;
    for (int r = 0; r < NX; r++)
    {
        double a = 0.0;
        for (int k = 0; k < NX; k++) a += s->prob.term_P[NX * r + k] * x[k];
        if (grad) grad[r] = 2.0 * a;
        h += x[r] * a;
    }
    return h;
}

void rmpc_oracle_destroy(rmpc_oracle *s)
{
    if (!s) return;
    free(s->ux);
    free(s);
}

/* One allocation for everything, like an acados capsule. */
rmpc_oracle *rmpc_oracle_create(const rmpc_problem *prob, const double x0[12])
{
    rmpc_oracle *s = calloc(1, sizeof(*s));
    s->prob = *prob;
    const int N = s->N = prob->N;
    const size_t S = (size_t) N + 1;
    size_t tot = 0;
#define TAKE(n) (tot += (n), tot - (n))
    size_t o_ux = TAKE(S * NV), o_pi = TAKE(S * NX), o_lam = TAKE(S * 36);
    size_t o_RSQ = TAKE(S * 324), o_rqz = TAKE(S * NV), o_b = TAKE(S * NX), o_d = TAKE(S * 36), o_cf = TAKE(S);
    size_t o_qux = TAKE(S * NV), o_qpi = TAKE(S * NX), o_qlam = TAKE(S * 36), o_qt = TAKE(S * 36);
    size_t o_rRSQ = TAKE(S * 324), o_rrqz = TAKE(S * NV), o_rb = TAKE(S * NX), o_rd = TAKE(S * 24);
    size_t o_sux = TAKE(S * NV), o_spi = TAKE(S * NX), o_slam = TAKE(S * 24), o_st = TAKE(S * 24);
    size_t o_dux = TAKE(S * NV), o_dpi = TAKE(S * NX), o_dlam = TAKE(S * 24), o_dt = TAKE(S * 24);
    size_t o_rg = TAKE(S * NV), o_rbb = TAKE(S * NX), o_rdd = TAKE(S * 24), o_rm = TAKE(S * 24), o_rmb = TAKE(S * 24);
    size_t o_Ga = TAKE(S * 24), o_ga = TAKE(S * 24), o_ti = TAKE(S * 24);
    size_t o_L = TAKE(S * LD * NV), o_Pb = TAKE(S * NX);
    size_t o_ig = TAKE(S * NV), o_ib = TAKE(S * NX), o_id = TAKE(S * 24), o_im = TAKE(S * 24);
    size_t o_iux = TAKE(S * NV), o_ipi = TAKE(S * NX), o_ilam = TAKE(S * 24), o_it = TAKE(S * 24);
    size_t o_wpi = TAKE(S * NX), o_wlam = TAKE(S * 36), o_tux = TAKE(S * NV);
#undef TAKE
    double *m = calloc(tot, sizeof(double));
    s->ux = m + o_ux; s->pi = m + o_pi; s->lam = m + o_lam;
    s->RSQ = m + o_RSQ; s->rqz = m + o_rqz; s->b = m + o_b; s->d = m + o_d; s->cost_fun = m + o_cf;
    s->qux = m + o_qux; s->qpi = m + o_qpi; s->qlam = m + o_qlam; s->qt = m + o_qt;
    s->rRSQ = m + o_rRSQ; s->rrqz = m + o_rrqz; s->rb = m + o_rb; s->rd = m + o_rd;
    s->sux = m + o_sux; s->spi = m + o_spi; s->slam = m + o_slam; s->st = m + o_st;
    s->dux = m + o_dux; s->dpi = m + o_dpi; s->dlam = m + o_dlam; s->dt = m + o_dt;
    s->res_g = m + o_rg; s->res_b = m + o_rbb; s->res_d = m + o_rdd; s->res_m = m + o_rm; s->res_m_bkp = m + o_rmb;
    s->Gamma = m + o_Ga; s->gamma = m + o_ga; s->tinv = m + o_ti;
    s->L = m + o_L; s->Pb = m + o_Pb;
    s->ig = m + o_ig; s->ib = m + o_ib; s->id = m + o_id; s->im = m + o_im;
    s->iux = m + o_iux; s->ipi = m + o_ipi; s->ilam = m + o_ilam; s->it = m + o_it;
    s->wpi = m + o_wpi; s->wlam = m + o_wlam; s->tux = m + o_tux;

    /* W, yref: R/trajectory_optimizer.py:109-115,142-155 */
    for (int i = 0; i < 5; i++) s->Wd[i] = prob->w_task[i];
    for (int i = 0; i < 6; i++) s->Wd[5 + i] = 2.0 * prob->w_u;
    for (int i = 0; i < 6; i++) s->Wd[11 + i] = prob->w_qddot;
    s->Wd[17] = 0.0;
    s->yref[1] = 1.0;
    s->yref[3] = prob->px_ref;
    s->yref[4] = prob->vy_ref;
    ur10_dyn_jac_BAbt(&s->prob.model, s->BAbt);
    memcpy(s->x0, x0, sizeof(s->x0));
    /* initial guess x_k = x0, u_k = 0, duals 0 (T/c_templates_tera/acados_solver.in.c:2744-2777) */
    for (int i = 0; i < N; i++) memcpy(s->ux + NV * i + NU, x0, NX * sizeof(double));
    memcpy(s->ux + NV * N, x0, NX * sizeof(double));
    s->first_call = 1;
    s->qp_warm_start = 2; /* qp_solver_warm_start = 2, R/trajectory_optimizer.py:66 */
    return s;
}

/* ------------------------------------------------------------------------------------------------
 * Linearisation: ocp_nlp_approximate_qp_matrices + _vectors_sqp (AC/ocp_nlp/ocp_nlp_common.c:2983-3106)
 *   cost    AC/ocp_nlp/ocp_nlp_cost_nls.c:830-1032   res=y-yref, grad=s*Jt*W*res, RSQ=s*(Jt W^1/2)(..)^T, fun
 *   dyn     AC/ocp_nlp/ocp_nlp_dynamics_disc.c:611-708   b = phi(x,u) - x_next
 *   bounds  AC/ocp_nlp/ocp_nlp_constraints_bgh.c:1563-1619   d = [lb - v; v - ub]
 */
static void bounds_of(const rmpc_oracle *s, int i, double *lb, double *ub)
{
    const rmpc_problem *p = &s->prob;
    if (i < s->N)
        for (int j = 0; j < 6; j++) { lb[j] = p->u_min[j]; ub[j] = p->u_max[j]; }
    if (i == 0)
        for (int j = 0; j < 12; j++) { lb[6 + j] = s->x0[j]; ub[6 + j] = s->x0[j]; }
    else if (i < s->N)
        for (int j = 0; j < 6; j++) { lb[6 + j] = p->q_min[j]; ub[6 + j] = p->q_max[j]; }
}

static void eval_ineq(const rmpc_oracle *s, int i, const double *v, double *fun)
{
    double lb[18], ub[18];
    int nb = nb_full(s, i);
    bounds_of(s, i, lb, ub);
    for (int j = 0; j < nb; j++)
    {
        fun[j] = lb[j] - v[j];
        fun[nb + j] = v[j] - ub[j];
    }
}

static double eval_cost_stage(const rmpc_oracle *s, const double *v)
{
    double y[NY], f = 0.0;
    ur10_cost_y(&s->prob.model, v + NU, v, y);
    for (int k = 0; k < NY; k++)
    {
        double r = y[k] - s->yref[k];
        f += (s->Wd[k] * r) * r;
    }
    return s->prob.dt * (0.5 * f);
}

static void linearise(rmpc_oracle *s)
{
    const int N = s->N;
    const double sc = s->prob.dt;
    for (int i = 0; i < N; i++)
    {
        const double *v = s->ux + NV * i;
        double y[NY], Jt[NV * NY], res[NY], wr[NY];
        ur10_cost_y_jac(&s->prob.model, v + NU, v, y, Jt);
        double f = 0.0;
        for (int k = 0; k < NY; k++)
        {
            res[k] = y[k] - s->yref[k];
            wr[k] = s->Wd[k] * res[k];
            f += wr[k] * res[k];
        }
        s->cost_fun[i] = sc * (0.5 * f);
        double *g = s->rqz + NV * i, *H = s->RSQ + 324 * i;
        for (int r = 0; r < NV; r++)
        {
            double a = 0.0;
            for (int k = 0; k < NY; k++) a += Jt[r + NV * k] * wr[k];
            g[r] = sc * a;
        }
        for (int c = 0; c < NV; c++)
            for (int r = c; r < NV; r++)
            {
                double a = 0.0;
                for (int k = 0; k < NY; k++) a += (Jt[r + NV * k] * sqrt(s->Wd[k])) * (Jt[c + NV * k] * sqrt(s->Wd[k]));
                H[r + NV * c] = sc * a;
            }
        double xn[NX];
        ur10_dyn(&s->prob.model, v + NU, v, xn);
        const double *x1 = (i + 1 < N) ? s->ux + NV * (i + 1) + NU : s->ux + NV * N;
        for (int k = 0; k < NX; k++) s->b[NX * i + k] = xn[k] - x1[k];
        eval_ineq(s, i, v, s->d + 36 * i);
    }
    /* stage N: LINEAR_LS with ny_e = 0: zero cost, zero gradient, zero Hessian; no constraints */
    s->cost_fun[N] = 0.0;
    memset(s->rqz + NV * N, 0, NV * sizeof(double));
    memset(s->RSQ + 324 * N, 0, 324 * sizeof(double));
    if (s->prob.term)
    { /* BGH nh_e = 1 (AC/ocp_nlp/ocp_nlp_constraints_bgh.c:1219-1445,1563-1619): fun = [lh - h; h - uh], DCt = grad h */
        const double h = term_h(s, s->ux + NV * N, s->term_c);
        s->d[36 * N] = s->prob.term_lh - h;
        s->d[36 * N + 1] = h - s->prob.term_uh;
    }
}

/* KKT residuals of the NLP, ocp_nlp_res_compute (AC/ocp_nlp/ocp_nlp_common.c:3641-3751):
 * res_stat = cost_grad - ineq_adj - dyn_adj with ineq_adj[idxb] = lam_lb - lam_ub (bgh.c:1397-1404),
 * dyn_adj_i = -BAbt pi_i (+ pi_{i-1} on the x rows)  (ocp_nlp_dynamics_disc.c:700-704, common.c:3022-3042) */
static void nlp_residuals(rmpc_oracle *s)
{
    const int N = s->N;
    double rs = 0, re = 0, ri = 0, rc = 0;
    for (int i = 0; i <= N; i++)
    {
        int nu = nu_of(s, i), nv = nu + NX, nb = nb_full(s, i);
        const double *lam = s->lam + 36 * i;
        double r[NV];
        for (int k = 0; k < nv; k++) r[k] = s->rqz[NV * i + k];
        for (int k = 0; k < nb; k++) r[k] -= lam[k] - lam[nb + k];
        if (ng_of(s, i))
            for (int k = 0; k < NX; k++) r[k] -= s->term_c[k] * (lam[0] - lam[1]); /* adj += DCt (lam_l - lam_u) */
        if (i < N)
            for (int k = 0; k < nv; k++)
            {
                double a = 0.0;
                for (int j = 0; j < NX; j++) a += s->BAbt[k + NV * j] * s->pi[NX * i + j];
                r[k] += a;
            }
        if (i > 0)
            for (int j = 0; j < NX; j++) r[nu + j] -= s->pi[NX * (i - 1) + j];
        for (int k = 0; k < nv; k++) if (fabs(r[k]) > rs) rs = fabs(r[k]);
        if (i < N)
            for (int j = 0; j < NX; j++) if (fabs(s->b[NX * i + j]) > re) re = fabs(s->b[NX * i + j]);
        for (int k = 0; k < 2 * nc_full(s, i); k++)
        {
            double f = s->d[36 * i + k];
            if (f > ri) ri = f;
            if (fabs(lam[k] * f) > rc) rc = fabs(lam[k] * f);
        }
    }
    s->res[0] = rs; s->res[1] = re; s->res[2] = ri; s->res[3] = rc;
}

/* ------------------------------------------------------------------------------------------------
 * x0 elimination: d_ocp_qp_reduce_eq_dof (H/ocp_qp/x_ocp_qp_red.c:278-466); N2 = N so the "partial
 * condensing" that follows is a per-stage copy (H/cond/x_part_cond.c:486-558).
 */
static void qp_reduce(rmpc_oracle *s)
{
    const int N = s->N;
    memcpy(s->rRSQ, s->RSQ, sizeof(double) * 324 * (N + 1));
    memcpy(s->rrqz, s->rqz, sizeof(double) * NV * (N + 1));
    memcpy(s->rb, s->b, sizeof(double) * NX * (size_t) N);
    for (int i = 1; i <= N; i++) memcpy(s->rd + 24 * i, s->d + 36 * i, 24 * sizeof(double));
    /* stage 0: fixed x = d_lb (= x0 - x_nlp0) */
    double v[NV];
    memset(v, 0, sizeof(v));
    for (int j = 0; j < NX; j++) v[NU + j] = s->d[NU + j];
    for (int j = 0; j < NX; j++)
    {
        double a = 0.0;
        for (int k = 0; k < NV; k++) a += s->BAbt[k + NV * j] * v[k];
        s->rb[j] = s->b[j] + a;
    }
    const double *H = s->RSQ;
    for (int r = 0; r < NU; r++)
    {
        double a = 0.0;
        for (int k = 0; k < NV; k++) a += (k <= r ? H[r + NV * k] : H[k + NV * r]) * v[k];
        s->rrqz[r] = s->rqz[r] + a;
    }
    for (int j = 0; j < 6; j++) { s->rd[j] = s->d[j]; s->rd[6 + j] = s->d[18 + j]; }
}

/* OCP_QP_RES_COMPUTE / _LIN (H/ocp_qp/x_ocp_qp_res.c:345-533,535-685).  lin = 0: residual of the QP at (ux,pi,lam,t)
 * with m = 0.  lin = 1: residual of the linearised KKT system at step (ux,pi,lam,t) around (Lam,T). */
static double ipm_res(const rmpc_oracle *s, int lin, const double *rqz, const double *b, const double *d,
                      const double *m, const double *ux, const double *pi, const double *lam, const double *t,
                      const double *Lam, const double *T, double *rg, double *rbb, double *rd, double *rm)
{
    const int N = s->N;
    double mu = 0.0;
    for (int i = 0; i <= N; i++)
    {
        const int nu = rnu(s, i), nx = rnx(s, i), nv = nu + nx, nb = rnb(s, i);
        const double *H = s->rRSQ + 324 * i, *v = ux + NV * i;
        double *g = rg + NV * i;
        for (int r = 0; r < nv; r++)
        {
            double a = 0.0;
            for (int k = 0; k < nv; k++) a += (k <= r ? H[r + NV * k] : H[k + NV * r]) * v[k];
            g[r] = a + rqz[NV * i + r];
        }
        if (i > 0)
            for (int j = 0; j < nx; j++) g[nu + j] -= pi[NX * (i - 1) + j];
        const double *l = lam + 24 * i, *tt = t + 24 * i;
        for (int j = 0; j < nb; j++)
        {
            g[j] += l[nb + j] - l[j];
            rd[24 * i + j] = d[24 * i + j] + tt[j] - v[j];
            rd[24 * i + nb + j] = d[24 * i + nb + j] + tt[nb + j] + v[j];
        }
        if (ng_of(s, i))
        { /* general row C x of stage N (nb = 0): res_g += C' (lam_ug - lam_lg), res_d = d + t -+ C x */
            double cx = 0.0;
            for (int k = 0; k < NX; k++) cx += s->term_c[k] * v[k];
            for (int k = 0; k < NX; k++) g[k] += s->term_c[k] * (l[1] - l[0]);
            rd[24 * i] = d[24 * i] + tt[0] - cx;
            rd[24 * i + 1] = d[24 * i + 1] + tt[1] + cx;
        }
        if (i < N)
        {
            const double *x1 = ux + NV * (i + 1) + rnu(s, i + 1);
            for (int j = 0; j < NX; j++)
            {
                double a = 0.0;
                for (int k = 0; k < nv; k++) a += s->BAbt[k + NV * j] * v[k];
                rbb[NX * i + j] = b[NX * i + j] - x1[j] + a;
            }
            for (int k = 0; k < nv; k++)
            {
                double a = 0.0;
                for (int j = 0; j < NX; j++) a += s->BAbt[k + NV * j] * pi[NX * i + j];
                g[k] += a;
            }
        }
        for (int j = 0; j < 2 * rnc(s, i); j++)
        {
            if (!lin)
            {
                rm[24 * i + j] = l[j] * tt[j];
                mu += rm[24 * i + j];
            }
            else
                rm[24 * i + j] = m[24 * i + j] + Lam[24 * i + j] * tt[j] + l[j] * T[24 * i + j];
        }
    }
    return mu;
}

static void ipm_res_norm(const rmpc_oracle *s, const double *rg, const double *rbb, const double *rd, const double *rm,
                         double nrm[4])
{
    const int N = s->N;
    nrm[0] = nrm[1] = nrm[2] = nrm[3] = 0.0;
    for (int i = 0; i <= N; i++)
    {
        const int nv = rnu(s, i) + rnx(s, i), nb = rnc(s, i);
        for (int k = 0; k < nv; k++) if (fabs(rg[NV * i + k]) > nrm[0]) nrm[0] = fabs(rg[NV * i + k]);
        if (i < N)
            for (int k = 0; k < NX; k++) if (fabs(rbb[NX * i + k]) > nrm[1]) nrm[1] = fabs(rbb[NX * i + k]);
        for (int k = 0; k < 2 * nb; k++)
        {
            if (fabs(rd[24 * i + k]) > nrm[2]) nrm[2] = fabs(rd[24 * i + k]);
            if (fabs(rm[24 * i + k]) > nrm[3]) nrm[3] = fabs(rm[24 * i + k]);
        }
    }
}

/* BLASFEO dpotrf_l_mn semantics (non-positive pivot -> zero column), (nv+1) x nv block, ld LD */
static void potrf_mn(int m, int n, double *C)
{
    for (int j = 0; j < n; j++)
    {
        double dj = C[j + LD * j];
        for (int k = 0; k < j; k++) dj -= C[j + LD * k] * C[j + LD * k];
        double inv;
        if (dj > 0.0) { dj = sqrt(dj); inv = 1.0 / dj; }
        else { dj = 0.0; inv = 0.0; }
        C[j + LD * j] = dj;
        for (int i = j + 1; i < m; i++)
        {
            double a = C[i + LD * j];
            for (int k = 0; k < j; k++) a -= C[i + LD * k] * C[j + LD * k];
            C[i + LD * j] = a * inv;
        }
    }
}

/* P' v = Lxx (Lxx^T v) with Lxx the x-block of L[ss+1] */
static void apply_P(const double *Ln, int nu1, const double *v, double *out)
{
    double tmp[NX];
    for (int c = 0; c < NX; c++)
    {
        double a = 0.0;
        for (int k = c; k < NX; k++) a += Ln[(nu1 + k) + LD * (nu1 + c)] * v[k];
        tmp[c] = a;
    }
    for (int r = 0; r < NX; r++)
    {
        double a = 0.0;
        for (int k = 0; k <= r; k++) a += Ln[(nu1 + r) + LD * (nu1 + k)] * tmp[k];
        out[r] = a;
    }
}

/* solve L[0:n,0:n]^T z = x[0:n] - L[n:m,0:n]^T x[n:m]   (TRSV_LTN_MN) */
static void trsv_ltn_mn(int m, int n, const double *L, double *x)
{
    for (int j = n - 1; j >= 0; j--)
    {
        double a = x[j];
        for (int k = j + 1; k < m; k++) a -= L[k + LD * j] * x[k];
        x[j] = a / L[j + LD * j];
    }
}
/* z[0:n] = L[0:n,0:n]^-1 x[0:n]; z[n:m] = x[n:m] - L[n:m,0:n] z[0:n]   (TRSV_LNN_MN) */
static void trsv_lnn_mn(int m, int n, const double *L, double *x)
{
    for (int j = 0; j < n; j++)
    {
        double a = x[j];
        for (int k = 0; k < j; k++) a -= L[j + LD * k] * x[k];
        x[j] = a / L[j + LD * j];
    }
    for (int r = n; r < m; r++)
    {
        double a = x[r];
        for (int k = 0; k < n; k++) a -= L[r + LD * k] * x[k];
        x[r] = a;
    }
}

static void ipm_dt_dlam(rmpc_oracle *s, const double *res_d, const double *res_m, double *dux, double *dlam, double *dt)
{
    /* tail of OCP_QP_FACT_SOLVE_KKT_STEP + COMPUTE_LAM_T_QP (H/ipm_core/x_core_qp_ipm_aux.c:164-189) */
    for (int i = 0; i <= s->N; i++)
    {
        const int nb = rnb(s, i);
        for (int j = 0; j < nb; j++)
        {
            dt[24 * i + j] = dux[NV * i + j];
            dt[24 * i + nb + j] = -dux[NV * i + j];
        }
        if (ng_of(s, i))
        {
            double cx = 0.0;
            for (int k = 0; k < NX; k++) cx += s->term_c[k] * dux[NV * i + k];
            dt[24 * i] = cx;
            dt[24 * i + 1] = -cx;
        }
        for (int j = 0; j < 2 * rnc(s, i); j++)
        {
            const int k = 24 * i + j;
            double lam0 = s->slam[k];
            dlam[k] = -s->tinv[k] * (res_m[k] + (lam0 * dt[k]) - (lam0 * res_d[k]));
            dt[k] -= res_d[k];
        }
    }
}

/* OCP_QP_FACT_SOLVE_KKT_STEP, square-root branch (H/ocp_qp/x_ocp_qp_kkt.c:836-1007) on qp_step
 * (rqz = res_g, b = res_b, d = res_d, m = res_m). */
static void ipm_fact_solve(rmpc_oracle *s)
{
    const int N = s->N;
    const double reg = 1e-15;
    /* COMPUTE_GAMMA_GAMMA_QP, t_lam_min = 2 branch (x_core_qp_ipm_aux.c:38-86) */
    for (int i = 0; i <= N; i++)
        for (int j = 0; j < 2 * rnc(s, i); j++)
        {
            const int k = 24 * i + j;
            s->tinv[k] = 1.0 / s->st[k];
            s->Gamma[k] = s->tinv[k] * s->slam[k];
            s->gamma[k] = s->tinv[k] * (s->res_m[k] - s->slam[k] * s->res_d[k]);
        }
    for (int ss = N; ss >= 0; ss--)
    {
        const int nu = rnu(s, ss), nx = rnx(s, ss), nv = nu + nx, nb = rnb(s, ss);
        double *L = s->L + LD * NV * ss;
        const double *H = s->rRSQ + 324 * ss;
        memset(L, 0, sizeof(double) * LD * NV);
        for (int c = 0; c < nv; c++)
        {
            for (int r = c; r < nv; r++) L[r + LD * c] = H[r + NV * c];
            L[c + LD * c] += reg;
            L[nv + LD * c] = s->res_g[NV * ss + c];
        }
        for (int j = 0; j < nb; j++)
        {
            L[j + LD * j] += s->Gamma[24 * ss + nb + j] + s->Gamma[24 * ss + j];
            L[nv + LD * j] += -s->gamma[24 * ss + nb + j] + s->gamma[24 * ss + j];
        }
        if (ng_of(s, ss))
        { /* general rows (x_ocp_qp_kkt.c:859-868): AL = [DCt diag(Gamma); gamma'], L += AL DCt' (lower part) */
            const double Ga = s->Gamma[24 * ss + 1] + s->Gamma[24 * ss], ga = -s->gamma[24 * ss + 1] + s->gamma[24 * ss];
            for (int c = 0; c < nv; c++)
            {
                for (int r = c; r < nv; r++) L[r + LD * c] += (s->term_c[r] * Ga) * s->term_c[c];
                L[nv + LD * c] += ga * s->term_c[c];
            }
        }
        if (ss < N)
        {
            const int nu1 = rnu(s, ss + 1);
            const double *Ln = s->L + LD * NV * (ss + 1);
            double AL[LD * NX];
            for (int r = 0; r <= nv; r++)
                for (int c = 0; c < NX; c++)
                {
                    double a = 0.0;
                    for (int k = c; k < NX; k++)
                    {
                        double ark = r < nv ? s->BAbt[r + NV * k] : s->res_b[NX * ss + k];
                        a += ark * Ln[(nu1 + k) + LD * (nu1 + c)];
                    }
                    AL[r + LD * c] = a;
                }
            for (int r = 0; r < NX; r++)
            {
                double a = 0.0;
                for (int k = 0; k <= r; k++) a += Ln[(nu1 + r) + LD * (nu1 + k)] * AL[nv + LD * k];
                s->Pb[NX * ss + r] = a;
            }
            for (int c = 0; c < NX; c++) AL[nv + LD * c] += Ln[(nu1 + NX) + LD * (nu1 + c)];
            for (int c = 0; c < nv; c++)
                for (int r = c; r <= nv; r++)
                {
                    double a = 0.0;
                    for (int k = 0; k < NX; k++) a += AL[r + LD * k] * AL[c + LD * k];
                    L[r + LD * c] += a;
                }
        }
        potrf_mn(nv + 1, nv, L);
    }
    /* forward substitution */
    for (int ss = 0; ss <= N; ss++)
    {
        const int nu = rnu(s, ss), nx = rnx(s, ss), nv = nu + nx;
        const double *L = s->L + LD * NV * ss;
        double *dv = s->dux + NV * ss;
        for (int k = 0; k < nu; k++) dv[k] = -L[nv + LD * k];
        trsv_ltn_mn(nv, nu, L, dv);
        if (ss < N)
        {
            const int nu1 = rnu(s, ss + 1);
            const double *Ln = s->L + LD * NV * (ss + 1);
            double *x1 = s->dux + NV * (ss + 1) + nu1;
            for (int j = 0; j < NX; j++)
            {
                double a = 0.0;
                for (int k = 0; k < nv; k++) a += s->BAbt[k + NV * j] * dv[k];
                x1[j] = a + s->res_b[NX * ss + j];
            }
            /* dpi = Lxx (Lxx^T dx1 + l1x) */
            double tmp[NX];
            for (int c = 0; c < NX; c++)
            {
                double a = 0.0;
                for (int k = c; k < NX; k++) a += Ln[(nu1 + k) + LD * (nu1 + c)] * x1[k];
                tmp[c] = a + Ln[(nu1 + NX) + LD * (nu1 + c)];
            }
            for (int r = 0; r < NX; r++)
            {
                double a = 0.0;
                for (int k = 0; k <= r; k++) a += Ln[(nu1 + r) + LD * (nu1 + k)] * tmp[k];
                s->dpi[NX * ss + r] = a;
            }
        }
    }
    ipm_dt_dlam(s, s->res_d, s->res_m, s->dux, s->dlam, s->dt);
}

/* OCP_QP_SOLVE_KKT_STEP, square-root branch (H/ocp_qp/x_ocp_qp_kkt.c:1543-1727): re-solve with stored factors. */
static void ipm_solve(rmpc_oracle *s, const double *rqz, const double *b, const double *d, const double *m, double *dux,
                      double *dpi, double *dlam, double *dt, int use_Pb)
{
    const int N = s->N;
    for (int i = 0; i <= N; i++) /* COMPUTE_GAMMA_QP */
        for (int j = 0; j < 2 * rnc(s, i); j++)
        {
            const int k = 24 * i + j;
            s->gamma[k] = s->tinv[k] * (m[k] - s->slam[k] * d[k]);
        }
    for (int ss = N; ss >= 0; ss--)
    {
        const int nu = rnu(s, ss), nx = rnx(s, ss), nv = nu + nx, nb = rnb(s, ss);
        const double *L = s->L + LD * NV * ss;
        double *dv = dux + NV * ss;
        for (int k = 0; k < nv; k++) dv[k] = rqz[NV * ss + k];
        for (int j = 0; j < nb; j++) dv[j] += -s->gamma[24 * ss + nb + j] + s->gamma[24 * ss + j];
        if (ng_of(s, ss))
            for (int k = 0; k < NX; k++) dv[k] += s->term_c[k] * (-s->gamma[24 * ss + 1] + s->gamma[24 * ss]);
        if (ss < N)
        {
            const int nu1 = rnu(s, ss + 1);
            const double *Ln = s->L + LD * NV * (ss + 1);
            double tmp[NX];
            if (use_Pb)
                for (int j = 0; j < NX; j++) tmp[j] = dux[NV * (ss + 1) + nu1 + j] + s->Pb[NX * ss + j];
            else
            {
                apply_P(Ln, nu1, b + NX * ss, tmp);
                for (int j = 0; j < NX; j++) tmp[j] += dux[NV * (ss + 1) + nu1 + j];
            }
            for (int k = 0; k < nv; k++)
            {
                double a = 0.0;
                for (int j = 0; j < NX; j++) a += s->BAbt[k + NV * j] * tmp[j];
                dv[k] += a;
            }
        }
        trsv_lnn_mn(nv, nu, L, dv);
    }
    for (int ss = 0; ss < N; ss++)
    {
        const int nu = rnu(s, ss), nx = rnx(s, ss), nv = nu + nx, nu1 = rnu(s, ss + 1);
        const double *L = s->L + LD * NV * ss, *Ln = s->L + LD * NV * (ss + 1);
        double *dv = dux + NV * ss, *x1 = dux + NV * (ss + 1) + nu1;
        for (int j = 0; j < NX; j++) dpi[NX * ss + j] = x1[j];
        for (int k = 0; k < nu; k++) dv[k] = -dv[k];
        trsv_ltn_mn(nv, nu, L, dv);
        for (int j = 0; j < NX; j++)
        {
            double a = 0.0;
            for (int k = 0; k < nv; k++) a += s->BAbt[k + NV * j] * dv[k];
            x1[j] = a + b[NX * ss + j];
        }
        double tmp[NX];
        apply_P(Ln, nu1, x1, tmp);
        for (int j = 0; j < NX; j++) dpi[NX * ss + j] += tmp[j];
    }
    ipm_dt_dlam(s, d, m, dux, dlam, dt);
}

/* OCP_QP_FACT_LQ_SOLVE_KKT_STEP (H/ocp_qp/x_ocp_qp_kkt.c:1201-1541): the fall-back of lq_fact = 1 when the Cholesky step
 * left a linear-system residual above 1e-5 (x_ocp_qp_ipm.c:2246-2276).  The stage factor L_ss (lower, L L' = the
 * Riccati matrix M_ss) is obtained WITHOUT forming M: it is the triangular factor of the Householder LQ decomposition
 * of   [ Lh | diag(sqrt(Gamma_b)) | DCt sqrt(Gamma_g) | BAbt L_{ss+1,xx} ],   Lh Lh' = RSQ + reg I   (:1263-1320),
 * with the diagonal made positive.  The vectors then follow the stored-factor solve (ipm_solve, = :1322,1469-1512).
 * BLASFEO's GELQF_PD_LLA is a blocked, structure-exploiting Householder LQ; this is the textbook unblocked one --
 * same factor up to rounding. */
#define LQ_KMAX (2 * NV + 1 + NX)
static void lq_lower(int m, int K, double *A /* m x K row-major, ld LQ_KMAX */, double *L /* ld LD, col-major */)
{
    for (int i = 0; i < m; i++)
    {
        double *v = A + (size_t) LQ_KMAX * i + i;
        const int n = K - i;
        double sigma = 0.0;
        for (int k = 1; k < n; k++) sigma += v[k] * v[k];
        double diag = v[0];
        if (sigma > 0.0)
        {
            const double mu = sqrt(v[0] * v[0] + sigma);
            const double v0 = v[0] <= 0.0 ? v[0] - mu : -sigma / (v[0] + mu);
            const double beta = 2.0 * v0 * v0 / (sigma + v0 * v0);
            for (int k = 1; k < n; k++) v[k] /= v0;
            v[0] = 1.0;
            for (int j = i + 1; j < m; j++)
            {
                double *r = A + (size_t) LQ_KMAX * j + i, w = 0.0;
                for (int k = 0; k < n; k++) w += r[k] * v[k];
                w *= beta;
                for (int k = 0; k < n; k++) r[k] -= w * v[k];
            }
            diag = mu;
        }
        v[0] = diag;
    }
    for (int c = 0; c < m; c++)
    {
        const double sg = A[(size_t) LQ_KMAX * c + c] < 0.0 ? -1.0 : 1.0;
        for (int r = c; r < m; r++) L[r + LD * c] = sg * A[(size_t) LQ_KMAX * r + c];
    }
}

static void ipm_fact_lq_solve(rmpc_oracle *s)
{
    const int N = s->N;
    const double reg = 1e-15;
    for (int i = 0; i <= N; i++) /* COMPUTE_GAMMA_GAMMA_QP */
        for (int j = 0; j < 2 * rnc(s, i); j++)
        {
            const int k = 24 * i + j;
            s->tinv[k] = 1.0 / s->st[k];
            s->Gamma[k] = s->tinv[k] * s->slam[k];
            s->gamma[k] = s->tinv[k] * (s->res_m[k] - s->slam[k] * s->res_d[k]);
        }
    for (int ss = N; ss >= 0; ss--)
    {
        const int nu = rnu(s, ss), nx = rnx(s, ss), nv = nu + nx, nb = rnb(s, ss), ng = ng_of(s, ss);
        double *L = s->L + LD * NV * ss;
        const double *H = s->rRSQ + 324 * ss;
        double A[NV * LQ_KMAX], Lh[LD * NV];
        memset(A, 0, sizeof(A));
        memset(Lh, 0, sizeof(Lh));
        for (int c = 0; c < nv; c++)
        {
            for (int r = c; r < nv; r++) Lh[r + LD * c] = H[r + NV * c];
            Lh[c + LD * c] += reg;
        }
        potrf_mn(nv, nv, Lh); /* POTRF_L of RSQ + reg I (use_hess_fact) */
        int K = 0;
        for (int c = 0; c < nv; c++)
            for (int r = c; r < nv; r++) A[(size_t) LQ_KMAX * r + c] = Lh[r + LD * c];
        K = nv;
        for (int j = 0; j < nb; j++)
        { /* box rows: idxb = 0..nb-1 */
            double t = s->Gamma[24 * ss + j] + s->Gamma[24 * ss + nb + ng + j];
            t = t >= 0.0 ? t : 0.0;
            t = sqrt(t);
            A[(size_t) LQ_KMAX * j + K + j] = t > 0.0 ? t : 0.0;
        }
        K += nv;
        if (ng)
        {
            double t = s->Gamma[24 * ss] + s->Gamma[24 * ss + 1];
            t = t >= 0.0 ? t : 0.0;
            t = sqrt(t);
            for (int r = 0; r < nv; r++) A[(size_t) LQ_KMAX * r + K] = s->term_c[r] * t;
            K += 1;
        }
        if (ss < N)
        {
            const int nu1 = rnu(s, ss + 1);
            const double *Ln = s->L + LD * NV * (ss + 1);
            for (int r = 0; r < nv; r++) /* TRMM_RLNN: BAbt * Lxx */
                for (int c = 0; c < NX; c++)
                {
                    double a = 0.0;
                    for (int k = c; k < NX; k++) a += s->BAbt[r + NV * k] * Ln[(nu1 + k) + LD * (nu1 + c)];
                    A[(size_t) LQ_KMAX * r + K + c] = a;
                }
            K += NX;
            double tmp[NX]; /* Pb = Lxx (Lxx' res_b) */
            for (int c = 0; c < NX; c++)
            {
                double a = 0.0;
                for (int k = c; k < NX; k++) a += Ln[(nu1 + k) + LD * (nu1 + c)] * s->res_b[NX * ss + k];
                tmp[c] = a;
            }
            for (int r = 0; r < NX; r++)
            {
                double a = 0.0;
                for (int k = 0; k <= r; k++) a += Ln[(nu1 + r) + LD * (nu1 + k)] * tmp[k];
                s->Pb[NX * ss + r] = a;
            }
        }
        memset(L, 0, sizeof(double) * LD * NV);
        lq_lower(nv, K, A, L);
    }
    ipm_solve(s, s->res_g, s->res_b, s->res_d, s->res_m, s->dux, s->dpi, s->dlam, s->dt, 1);
}

/* COMPUTE_ALPHA_QP (x_core_qp_ipm_aux.c:193-263) */
static void ipm_alpha(const rmpc_oracle *s, double *alpha, double *alpha_prim, double *alpha_dual)
{
    double ap = -1.0, ad = -1.0;
    for (int i = 0; i <= s->N; i++)
        for (int j = 0; j < 2 * rnc(s, i); j++)
        {
            const int k = 24 * i + j;
            if (ad * s->dlam[k] > s->slam[k]) ad = s->slam[k] / s->dlam[k];
            if (ap * s->dt[k] > s->st[k]) ap = s->st[k] / s->dt[k];
        }
    double a = ap > ad ? ap : ad;
    *alpha_prim = -ap; *alpha_dual = -ad; *alpha = -a;
}

static double ipm_mu_aff(const rmpc_oracle *s, double alpha, double nc_inv)
{
    double mu = 0.0;
    for (int i = 0; i <= s->N; i++)
        for (int j = 0; j < 2 * rnc(s, i); j++)
        {
            const int k = 24 * i + j;
            mu += (s->slam[k] + alpha * s->dlam[k]) * (s->st[k] + alpha * s->dt[k]);
        }
    return mu * nc_inv;
}

static int any_gt(const double n[4], double thr) { return n[0] > thr || n[1] > thr || n[2] > thr || n[3] > thr; }

/* d_ocp_qp_ipm_solve (H/ocp_qp/x_ocp_qp_ipm.c:2626-3018) with OCP_QP_IPM_DELTA_STEP (:2155-2622) inlined, for the
 * argument set acados resolves: warm_start 2, pred_corr, cond_pred_corr, itref_pred_max 0,
 * itref_corr_max 2, lq_fact 1, split_step 0, t_lam_min 2, abs_form 0.  Returns the HPIPM status
 * (0 SUCCESS, 1 MAX_ITER, 2 MIN_STEP, 3 NAN_SOL).  The LQ (Householder) fall-back factorisation of lq_fact = 1 is
 * ipm_fact_lq_solve(); every time its trigger fires the event is also counted in n_lq_needed. */
static int ipm_solve_qp(rmpc_oracle *s, int warm)
{
    const int N = s->N, iter_max = s->prob.qp_iter_max;
    const double tol = s->prob.qp_tol, alpha_min = 1e-8, tau_min = 1e-16, lam_min = 1e-16, t_min = 1e-16;
    int nc = 0;
    for (int i = 0; i <= N; i++) nc += 2 * rnc(s, i);
    const double nc_inv = 1.0 / nc;
    const size_t nc_pad = 24 * (size_t) (N + 1);

    if (warm == 2)
    {
        /* OCP_QP_INIT_VAR, warm_start == 2 (:1638-1654): keep ux (zeroed by acados), pi; clip lam, t */
        for (int i = 0; i <= N; i++)
            for (int j = 0; j < 2 * rnc(s, i); j++)
            {
                const int k = 24 * i + j;
                if (s->slam[k] < 1e-1) s->slam[k] = 1e-1;
                if (s->st[k] < 1e-1) s->st[k] = 1e-1;
            }
    }
    else
    {
        /* OCP_QP_INIT_VAR, warm_start == 0, t0_init == 2, var_init_scheme == 1 (:1656-1672,1837-1978), mu0 = 1:
         * ux = 0, pi = 0, t = slack of the bounds at ux pushed to >= 0.1 (moving ux), lam = mu0 / t */
        const double thr0 = 1e-1, mu0 = 1.0;
        memset(s->sux, 0, sizeof(double) * NV * (N + 1));
        memset(s->spi, 0, sizeof(double) * NX * (size_t) N);
        for (int i = 0; i <= N; i++)
        {
            const int nb = rnb(s, i);
            double *v = s->sux + NV * i, *tl = s->st + 24 * i, *tu = tl + nb;
            const double *dl = s->rd + 24 * i, *du = dl + nb;
            for (int j = 0; j < nb; j++)
            {
                tl[j] = v[j] - dl[j];
                tu[j] = -v[j] - du[j];
                if (tl[j] < thr0)
                {
                    if (tu[j] < thr0) { v[j] = 0.5 * (dl[j] - du[j]); tl[j] = thr0; tu[j] = thr0; }
                    else { tl[j] = thr0; v[j] = dl[j] + thr0; }
                }
                else if (tu[j] < thr0) { tu[j] = thr0; v[j] = -du[j] - thr0; }
            }
            for (int j = 0; j < 2 * nb; j++) s->slam[24 * i + j] = mu0 / tl[j];
            if (ng_of(s, i))
            { /* general rows (:1780-1817): t = +-C ux - d clipped at thr0 (ux is not moved), lam = mu0 / t */
                double cx = 0.0;
                for (int k = 0; k < NX; k++) cx += s->term_c[k] * v[k];
                tl[0] = cx - dl[0];
                tl[1] = -cx - dl[1];
                for (int j = 0; j < 2; j++)
                {
                    tl[j] = thr0 > tl[j] ? thr0 : tl[j];
                    s->slam[24 * i + j] = mu0 / tl[j];
                }
            }
        }
    }
    double alpha = 1.0, nrm[4], mu, res_m_tau;
    mu = nc_inv * ipm_res(s, 0, s->rrqz, s->rb, s->rd, NULL, s->sux, s->spi, s->slam, s->st, NULL, NULL, s->res_g, s->res_b,
                          s->res_d, s->res_m);
    ipm_res_norm(s, s->res_g, s->res_b, s->res_d, s->res_m, nrm);
    res_m_tau = 0.0;
    for (int i = 0; i <= N; i++)
        for (int j = 0; j < 2 * rnc(s, i); j++)
            if (fabs(s->res_m[24 * i + j] - tau_min) > res_m_tau) res_m_tau = fabs(s->res_m[24 * i + j] - tau_min);
    int kk;
    for (kk = 0; kk < iter_max && alpha > alpha_min &&
                 (nrm[0] > tol || nrm[1] > tol || nrm[2] > tol || res_m_tau > tol); kk++)
    {
        /* ---- OCP_QP_IPM_DELTA_STEP ---- */
        memcpy(s->res_m_bkp, s->res_m, nc_pad * sizeof(double));               /* BACKUP_RES_M */
        for (size_t k = 0; k < nc_pad; k++) s->res_m[k] = s->res_m_bkp[k] - tau_min; /* COMPUTE_TAU_MIN_QP */
        ipm_fact_solve(s);
        double inrm[4];
        ipm_res(s, 1, s->res_g, s->res_b, s->res_d, s->res_m, s->dux, s->dpi, s->dlam, s->dt, s->slam, s->st, s->ig, s->ib,
                s->id, s->im);
        ipm_res_norm(s, s->ig, s->ib, s->id, s->im, inrm);
        if (getenv("RMPC_ORACLE_DEBUG") && kk == 0) fprintf(stderr, "port st1: res_g6 %.6e dux6 %.6e dpi1_0 %.6e dpi0_0 %.6e dlam_lb %.6e dlam_ub %.6e\n", s->res_g[NV+6], s->dux[NV+6], s->dpi[NX], s->dpi[0], s->dlam[24+6], s->dlam[24+18]);
        /* HPIPM's own NaN test looks at the first residual entry only (its norms skip NaNs).  A numerically singular
         * stage-N block (rank-one terminal row + reg) can make the dense Cholesky produce NaNs elsewhere, depending on
         * rounding noise that differs between BLASFEO's blocked kernels and this scalar code; any NaN in the step
         * therefore also takes the LQ route here. */
        int step_nan = 0;
        for (size_t k = 0; k < (size_t) NV * (N + 1) && !step_nan; k++) step_nan = isnan(s->dux[k]) || isinf(s->dux[k]);
        for (size_t k = 0; k < (size_t) NX * N && !step_nan; k++) step_nan = isnan(s->dpi[k]) || isinf(s->dpi[k]);
        if ((inrm[0] == 0.0 && isnan(s->ig[0])) || any_gt(inrm, 1e-5) || step_nan || getenv("RMPC_ORACLE_FORCE_LQ"))
        { /* inaccurate factorisation: refactorise this iteration with LQ (:2246-2276) */
            s->n_lq_needed++;
            ipm_fact_lq_solve(s);
        }
        double alpha_prim, alpha_dual;
        ipm_alpha(s, &alpha, &alpha_prim, &alpha_dual);
        double alpha_aff = alpha;
        /* Mehrotra predictor-corrector */
        double mu_aff = ipm_mu_aff(s, alpha, nc_inv);
        double tmp = mu_aff / mu, sigma = tmp * tmp * tmp;
        double sigma_mu = sigma * mu;
        sigma_mu = sigma_mu > tau_min ? sigma_mu : tau_min;
        for (size_t k = 0; k < nc_pad; k++) /* COMPUTE_CENTERING_CORRECTION_QP */
            s->res_m[k] = s->res_m_bkp[k] + s->dt[k] * s->dlam[k] - sigma_mu;
        ipm_solve(s, s->res_g, s->res_b, s->res_d, s->res_m, s->dux, s->dpi, s->dlam, s->dt, 1);
        ipm_alpha(s, &alpha, &alpha_prim, &alpha_dual);
        /* conditional predictor-corrector */
        double mu_aff0 = mu_aff;
        mu_aff = ipm_mu_aff(s, alpha, nc_inv);
        if (mu_aff > 2.0 * mu_aff0)
        {
            for (size_t k = 0; k < nc_pad; k++) s->res_m[k] = s->res_m_bkp[k] - sigma_mu; /* COMPUTE_CENTERING_QP */
            ipm_solve(s, s->res_g, s->res_b, s->res_d, s->res_m, s->dux, s->dpi, s->dlam, s->dt, 1);
            ipm_alpha(s, &alpha, &alpha_prim, &alpha_dual);
        }
        /* iterative refinement on the corrector, itref_corr_max = 2 */
        int refined = 0, itref1;
        for (itref1 = 0; itref1 < 2; itref1++)
        {
            ipm_res(s, 1, s->res_g, s->res_b, s->res_d, s->res_m, s->dux, s->dpi, s->dlam, s->dt, s->slam, s->st, s->ig,
                    s->ib, s->id, s->im);
            ipm_res_norm(s, s->ig, s->ib, s->id, s->im, inrm);
            if ((inrm[0] < tol || inrm[0] < 1e-3 * nrm[0]) && (inrm[1] < tol || inrm[1] < 1e-3 * nrm[1]) &&
                (inrm[2] < tol || inrm[2] < 1e-3 * nrm[2]) && (inrm[3] < tol || inrm[3] < 1e-3 * nrm[3]))
                break;
            ipm_solve(s, s->ig, s->ib, s->id, s->im, s->iux, s->ipi, s->ilam, s->it, 0);
            refined = 1;
            for (size_t k = 0; k < (size_t) NV * (N + 1); k++) s->dux[k] += s->iux[k];
            for (size_t k = 0; k < (size_t) NX * N; k++) s->dpi[k] += s->ipi[k];
            for (size_t k = 0; k < nc_pad; k++) { s->dlam[k] += s->ilam[k]; s->dt[k] += s->it[k]; }
        }
        if (refined) ipm_alpha(s, &alpha, &alpha_prim, &alpha_dual);
        /* UPDATE_VAR_QP (x_core_qp_ipm_aux.c:267-372), split_step = 0, t_lam_min = 2 */
        double a = alpha;
        if (a < 1.0) a = a * ((1.0 - a) * 0.99 + a * 0.9999999);
        for (size_t k = 0; k < (size_t) NV * (N + 1); k++) s->sux[k] += a * s->dux[k];
        for (size_t k = 0; k < (size_t) NX * N; k++) s->spi[k] += a * s->dpi[k];
        for (int i = 0; i <= N; i++)
            for (int j = 0; j < 2 * rnc(s, i); j++)
            {
                const int k = 24 * i + j;
                s->slam[k] += a * s->dlam[k];
                s->slam[k] = s->slam[k] <= lam_min ? lam_min : s->slam[k];
                s->st[k] += a * s->dt[k];
                s->st[k] = s->st[k] <= t_min ? t_min : s->st[k];
            }
        /* ---- end delta step ---- */
        mu = nc_inv * ipm_res(s, 0, s->rrqz, s->rb, s->rd, NULL, s->sux, s->spi, s->slam, s->st, NULL, NULL, s->res_g,
                              s->res_b, s->res_d, s->res_m);
        ipm_res_norm(s, s->res_g, s->res_b, s->res_d, s->res_m, nrm);
        res_m_tau = 0.0;
        for (int i = 0; i <= N; i++)
            for (int j = 0; j < 2 * rnc(s, i); j++)
                if (fabs(s->res_m[24 * i + j] - tau_min) > res_m_tau) res_m_tau = fabs(s->res_m[24 * i + j] - tau_min);
        if (kk < 63)
        {
            double *st = s->stat + 8 * (kk + 1);
            st[0] = alpha_aff; st[1] = mu_aff0; st[2] = sigma; st[3] = alpha; st[4] = mu; st[5] = nrm[0]; st[6] = nrm[1];
            st[7] = nrm[2];
        }
    }
    s->qp_iter_last = kk;
    if (kk == iter_max) return 1;
    if (alpha <= alpha_min) return 2;
    if (isnan(mu)) return 3;
    return 0;
}

/* d_ocp_qp_restore_eq_dof (H/ocp_qp/x_ocp_qp_red.c:848-995): reduced solution -> full-space ocp_qp_out */
static void qp_restore(rmpc_oracle *s)
{
    const int N = s->N;
    for (int i = 1; i <= N; i++)
    {
        memcpy(s->qux + NV * i, s->sux + NV * i, NV * sizeof(double));
        memset(s->qlam + 36 * i, 0, 36 * sizeof(double));
        memset(s->qt + 36 * i, 0, 36 * sizeof(double));
        memcpy(s->qlam + 36 * i, s->slam + 24 * i, 24 * sizeof(double));
        memcpy(s->qt + 36 * i, s->st + 24 * i, 24 * sizeof(double));
    }
    memcpy(s->qpi, s->spi, sizeof(double) * NX * (size_t) N);
    /* stage 0 */
    double *v = s->qux, *lam = s->qlam, *t = s->qt;
    for (int j = 0; j < NU; j++) v[j] = s->sux[j];
    for (int j = 0; j < NX; j++) v[NU + j] = s->d[NU + j];
    for (int j = 0; j < 18; j++)
    {
        if (j < 6)
        {
            lam[j] = s->slam[j]; lam[18 + j] = s->slam[6 + j];
            t[j] = s->st[j]; t[18 + j] = s->st[6 + j];
        }
        else
        {
            lam[j] = 1e-16; lam[18 + j] = 1e-16;
            t[j] = 1e-16; t[18 + j] = 1e-16;
        }
    }
    double tmp[NV];
    const double *H = s->RSQ;
    for (int r = 0; r < NV; r++)
    {
        double a = 0.0;
        for (int k = 0; k < NV; k++) a += (k <= r ? H[r + NV * k] : H[k + NV * r]) * v[k];
        tmp[r] = s->rqz[r] + (lam[18 + r] - lam[r]) + a;
        double c = 0.0;
        for (int j = 0; j < NX; j++) c += s->BAbt[r + NV * j] * s->spi[j];
        tmp[r] += c;
    }
    for (int j = 6; j < 18; j++)
    {
        if (tmp[j] >= 0) lam[j] = tmp[j];
        else lam[18 + j] = -tmp[j];
    }
}

/* ocp_qp_xcond_solve (AC/ocp_qp/ocp_qp_xcond_solver.c:518-576) + ocp_qp_hpipm (AC/ocp_qp/ocp_qp_hpipm.c:293-378).
 * Returns the acados status of the QP.
 * REFERENCE QUIRK reproduced on purpose: after the first QP, SQP (AC/ocp_nlp/ocp_nlp_sqp.c:680-683) and SQP_RTI
 * (AC/ocp_nlp/ocp_nlp_sqp_rti.c:590-593) "restore" the QP solver's warm_start option from nlp_opts->qp_warm_start,
 * a field that ocp_nlp_opts_set never writes for "qp_warm_start" (AC/ocp_nlp/ocp_nlp_common.c:1346-1355 forwards
 * the value to the QP solver only; the field keeps its default 0, :1269).  So qp_solver_warm_start = 2 holds for the
 * first QP of the solver's lifetime only and every later QP is cold-started (warm_start 0). */
static int solve_qp(rmpc_oracle *s, int init_from_qp_out)
{
    const int N = s->N;
    qp_reduce(s);
    if (init_from_qp_out)
    {
        /* initialize_next_xcond_qp_from_qp_out: d_ocp_qp_reduce_eq_dof_sol + part_cond_qp_cond_sol (copy) */
        memcpy(s->spi, s->qpi, sizeof(double) * NX * (size_t) N);
        for (int i = 1; i <= N; i++)
        {
            memcpy(s->slam + 24 * i, s->qlam + 36 * i, 24 * sizeof(double));
            memcpy(s->st + 24 * i, s->qt + 36 * i, 24 * sizeof(double));
        }
        for (int j = 0; j < 6; j++)
        {
            s->slam[j] = s->qlam[j]; s->slam[6 + j] = s->qlam[18 + j];
            s->st[j] = s->qt[j]; s->st[6 + j] = s->qt[18 + j];
        }
    }
    memset(s->sux, 0, sizeof(double) * NV * (N + 1)); /* ocp_qp_hpipm.c:309-316 */
    int hs = ipm_solve_qp(s, s->qp_warm_start);
    s->qp_warm_start = 0;
    qp_restore(s);
    s->qp_iter_sum += s->qp_iter_last;
    static const int map[4] = {0 /*SUCCESS*/, 2 /*MAXITER*/, 3 /*MINSTEP*/, 1 /*NAN*/};
    return s->qp_status_last = map[hs];
}

/* ------------------------------------------------------------------------------------------------
 * Globalisation: MERIT_BACKTRACKING (AC/ocp_nlp/ocp_nlp_globalization_merit_backtracking.c)
 */
/* ocp_nlp_evaluate_merit_fun (:606-702) at the trial iterate tux */
static double merit_fun(rmpc_oracle *s)
{
    const int N = s->N;
    double cost = 0.0, dyn = 0.0, con = 0.0;
    for (int i = 0; i < N; i++) cost += eval_cost_stage(s, s->tux + NV * i);
    cost += 0.0; /* stage N */
    for (int i = 0; i < N; i++)
    {
        const double *v = s->tux + NV * i;
        const double *x1 = (i + 1 < N) ? s->tux + NV * (i + 1) + NU : s->tux + NV * N;
        double xn[NX];
        ur10_dyn(&s->prob.model, v + NU, v, xn);
        for (int j = 0; j < NX; j++) dyn += fabs(s->wpi[NX * i + j]) * fabs(xn[j] - x1[j]);
    }
    for (int i = 0; i < N; i++)
    {
        double fun[36];
        int nb = nb_full(s, i);
        eval_ineq(s, i, s->tux + NV * i, fun);
        for (int j = 0; j < 2 * nb; j++)
            if (fun[j] > 0.0) con += fabs(s->wlam[36 * i + j]) * fun[j];
    }
    if (s->prob.term)
    {
        const double h = term_h(s, s->tux + NV * N, NULL);
        const double fl = s->prob.term_lh - h, fu = h - s->prob.term_uh;
        if (fl > 0.0) con += fabs(s->wlam[36 * N]) * fl;
        if (fu > 0.0) con += fabs(s->wlam[36 * N + 1]) * fu;
    }
    return cost + dyn + con;
}

/* ocp_nlp_line_search (:290-408) followed by ocp_nlp_update_variables_sqp (AC/ocp_nlp/ocp_nlp_common.c:3259-3318);
 * returns 1 if the merit function went NaN/inf (ACADOS_NAN_DETECTED), else 0. */
static int line_search_and_update(rmpc_oracle *s, int nlp_iter)
{
    const int N = s->N;
    const size_t nux = (size_t) NV * (N + 1);
    /* weights (:705-755); Leineweber update unless this is iteration 0 of an SQP call */
    if (nlp_iter == 0)
    {
        for (size_t k = 0; k < (size_t) NX * N; k++) s->wpi[k] = fabs(s->qpi[k]);
        memcpy(s->wlam, s->qlam, sizeof(double) * 36 * (N + 1));
    }
    else
    {
        for (size_t k = 0; k < (size_t) NX * N; k++)
        {
            double t0 = fabs(s->qpi[k]), t1 = 0.5 * (t0 + s->wpi[k]);
            s->wpi[k] = t0 > t1 ? t0 : t1;
        }
        for (int i = 0; i <= N; i++)
            for (int j = 0; j < 2 * nc_full(s, i); j++)
            {
                double t0 = s->qlam[36 * i + j], t1 = 0.5 * (t0 + s->wlam[36 * i + j]);
                s->wlam[36 * i + j] = t0 > t1 ? t0 : t1;
            }
    }
    memcpy(s->tux, s->ux, nux * sizeof(double));
    const double m0 = merit_fun(s);
    double alpha = 1.0, m1 = 0.0;
    int accepted = 0;
    for (; alpha * 0.7 > 0.05;)
    {
        for (size_t k = 0; k < nux; k++) s->tux[k] = s->ux[k] + alpha * s->qux[k];
        m1 = merit_fun(s);
        if (m1 < m0 && !isnan(m1) && !isinf(m1)) { accepted = 1; break; }
        alpha *= 0.7;
    }
    s->alpha_ls = alpha;
    if (!accepted && (isnan(m1) || isinf(m1))) return 1;
    /* step: w += alpha dw; duals = (1-alpha) old + alpha qp  (full_step_dual = 0) */
    for (size_t k = 0; k < nux; k++) s->ux[k] = s->ux[k] + alpha * s->qux[k];
    for (size_t k = 0; k < (size_t) NX * N; k++) s->pi[k] = (1.0 - alpha) * s->pi[k] + alpha * s->qpi[k];
    for (size_t k = 0; k < (size_t) 36 * (N + 1); k++) s->lam[k] = (1.0 - alpha) * s->lam[k] + alpha * s->qlam[k];
    return 0;
}

/* ocp_nlp_initialize_qp_from_nlp (AC/ocp_nlp/ocp_nlp_common.c:3320-3340) + ocp_qp_compute_t (AC/ocp_qp/ocp_qp_common.c:875-921) */
static void init_qp_from_nlp(rmpc_oracle *s)
{
    const int N = s->N;
    memset(s->qux, 0, sizeof(double) * NV * (N + 1));
    memcpy(s->qlam, s->lam, sizeof(double) * 36 * (N + 1));
    memcpy(s->qpi, s->pi, sizeof(double) * NX * (size_t) N);
    for (int i = 0; i <= N; i++)
        for (int j = 0; j < 2 * nc_full(s, i); j++) s->qt[36 * i + j] = -s->d[36 * i + j];
}

/* ocp_nlp_sqp_rti, rti_phase = PREPARATION_AND_FEEDBACK (AC/ocp_nlp/ocp_nlp_sqp_rti.c:455-646,1123-1175) */
static int solve_rti(rmpc_oracle *s)
{
    s->qp_iter_sum = 0;
    linearise(s);
    s->sqp_iter = 1;
    int init = 0;
    if (s->first_call) { init_qp_from_nlp(s); init = 1; } /* warm_start_first_qp && warm_start_first_qp_from_nlp (:561-575) */
    int qs = solve_qp(s, init);
    if (qs != 0 && qs != 2) return 4; /* ACADOS_QP_FAILURE; is_first_call stays set (:609-622) */
    line_search_and_update(s, 1);      /* globalisation status ignored (:627-635) */
    s->first_call = 0;
    return 0;
}

/* ocp_nlp_sqp (AC/ocp_nlp/ocp_nlp_sqp.c:474-801), check_termination (:351-443) */
static int solve_sqp(rmpc_oracle *s)
{
    const rmpc_problem *p = &s->prob;
    s->qp_iter_sum = 0;
    for (int it = 0; it <= p->max_iter; it++)
    {
        s->sqp_iter = it;
        if (it != p->max_iter)
        {
            linearise(s);
            s->cost_value = 0.0;
            for (int i = 0; i <= s->N; i++) s->cost_value += s->cost_fun[i];
            nlp_residuals(s);
        }
        const double *r = s->res;
        if (isnan(r[0]) || isnan(r[1]) || isnan(r[2]) || isnan(r[3])) return 1;
        if (it >= p->max_iter) return 2;
        if (r[0] < p->tol[0] && r[1] < p->tol[1] && r[2] < p->tol[2] && r[3] < p->tol[3]) return 0;
        if (s->cost_value <= -1e10) return 6; /* tol_unbounded default, T/acados_ocp_options.py */
        int qs = solve_qp(s, 0); /* warm_start_first_qp, not from nlp: QP memory keeps its previous solution (:652-666) */
        if (qs != 0 && qs != 2) { s->sqp_iter = it + 1; return 4; }
        if (line_search_and_update(s, it)) return 1;
    }
    return 2;
}

int rmpc_oracle_step(rmpc_oracle *s, const double x[12], double u0[6], int *sqp_iter, int *qp_iter, double res[4],
                     double *cost)
{
    memcpy(s->x0, x, sizeof(s->x0)); /* solver.set(0,'lbx',x); solver.set(0,'ubx',x)  R/simulator.py:210-211 */
    int status = s->prob.solver_type == RMPC_SQP_RTI ? solve_rti(s) : solve_sqp(s);
    memcpy(u0, s->ux, NU * sizeof(double));
    if (sqp_iter) *sqp_iter = s->sqp_iter;
    if (qp_iter) *qp_iter = s->qp_iter_sum;
    if (res)
        for (int k = 0; k < 4; k++) res[k] = s->prob.solver_type == RMPC_SQP_RTI ? 0.0 : s->res[k];
    if (cost)
    {
        /* solver.get_cost(): ocp_nlp_eval_cost at the current iterate (T/acados_ocp_solver.py:1664-1679) */
        double c = 0.0;
        for (int i = 0; i < s->N; i++) c += eval_cost_stage(s, s->ux + NV * i);
        *cost = c;
    }
    return status;
}

void rmpc_oracle_get_iterate(const rmpc_oracle *s, double *xtraj, double *utraj, double *pitraj)
{
    const int N = s->N;
    for (int i = 0; i <= N; i++)
        if (xtraj) memcpy(xtraj + NX * i, s->ux + NV * i + (i < N ? NU : 0), NX * sizeof(double));
    for (int i = 0; i < N; i++)
    {
        if (utraj) memcpy(utraj + NU * i, s->ux + NV * i, NU * sizeof(double));
        if (pitraj) memcpy(pitraj + NX * i, s->pi + NX * i, NX * sizeof(double));
    }
}

/* primal shift of the initial guess (rmpc_problem.shift): x_k <- x_{k+1}, u_k <- u_{k+1} for k < N-1, x_{N-1} <- x_N */
void rmpc_oracle_shift(rmpc_oracle *s)
{
    const int N = s->N;
    for (int i = 0; i + 1 < N; i++) memcpy(s->ux + NV * i, s->ux + NV * (i + 1), NV * sizeof(double));
    memcpy(s->ux + NV * (N - 1) + NU, s->ux + NV * N, NX * sizeof(double));
}

int rmpc_oracle_n_lq_needed(const rmpc_oracle *s) { return s->n_lq_needed; }
/* Debug/parity hook: per-iteration table of the LAST QP (8 columns: alpha_aff mu_aff sigma alpha mu res_stat res_eq
 * res_ineq; row k+1 = iteration k); returns the iteration count. */
int rmpc_oracle_last_qp_stat(const rmpc_oracle *s, double *out, int max_rows)
{
    int rows = s->qp_iter_last + 1 < max_rows ? s->qp_iter_last + 1 : max_rows;
    if (rows > 64) rows = 64;
    memcpy(out, s->stat, sizeof(double) * 8 * (size_t) rows);
    return s->qp_iter_last;
}

/* Whole closed loop, R/simulator.py:199-241 without sleeps / visualisation. Returns #failed steps. */
int rmpc_oracle_closed_loop(const rmpc_problem *prob, const double z0[12], int nsteps, rmpc_trace *tr)
{
    rmpc_oracle *s = rmpc_oracle_create(prob, z0);
    if (!s) return -1;
    const int N = prob->N;
    double z[NX], zn[NX], u0[NU];
    memcpy(z, z0, sizeof(z));
    if (tr->z) memcpy(tr->z, z, sizeof(z));
    int nfail = 0;
    for (int k = 0; k < nsteps; k++)
    {
        int sqp_iter = 0, qp_iter = 0;
        double res[4] = {0, 0, 0, 0}, cost = 0;
        int status = rmpc_oracle_step(s, z, u0, &sqp_iter, &qp_iter, res, &cost);
        nfail += status != 0;
        ur10_plant_rk4(prob->model.wcv, prob->dt, z, u0, zn);
        memcpy(z, zn, sizeof(z));
        if (tr->z) memcpy(tr->z + NX * (k + 1), z, sizeof(z));
        if (tr->u) memcpy(tr->u + NU * k, u0, sizeof(u0));
        if (tr->status) tr->status[k] = status;
        if (tr->sqp_iter) tr->sqp_iter[k] = sqp_iter;
        if (tr->qp_iter) tr->qp_iter[k] = qp_iter;
        if (tr->res) memcpy(tr->res + 4 * k, res, sizeof(res));
        if (tr->cost) tr->cost[k] = cost;
        if (tr->time_tot) tr->time_tot[k] = 0.0;
        if (tr->xtraj || tr->utraj || tr->pitraj)
            rmpc_oracle_get_iterate(s, tr->xtraj ? tr->xtraj + (size_t) k * (N + 1) * NX : NULL,
                                    tr->utraj ? tr->utraj + (size_t) k * N * NU : NULL,
                                    tr->pitraj ? tr->pitraj + (size_t) k * N * NX : NULL);
        if (prob->shift) rmpc_oracle_shift(s);
    }
    if (s->n_lq_needed) fprintf(stderr, "rmpc_oracle: LQ fall-back factorisation used %d times\n", s->n_lq_needed);
    rmpc_oracle_destroy(s);
    return nfail;
}
