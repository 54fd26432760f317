/* oracle/rmpc_problem.h -- problem description shared by the two oracle legs (TEST INFRASTRUCTURE):
 *   ref_driver.c   : drives the UNMODIFIED compiled acados/HPIPM/BLASFEO (oracle/_ref)
 *   rmpc_oracle.c  : plain-C restatement ("port") of the same algorithm
 * Field values = the resolved OCP data / options of SURVEY.md 8(a-0); each cites where the reference sets it.
 */
#ifndef RMPC_ORACLE_PROBLEM_H
#define RMPC_ORACLE_PROBLEM_H

#include "ur10_model.h"

#ifdef __cplusplus
extern "C" {
#endif

#define RMPC_SQP 0
#define RMPC_SQP_RTI 1

typedef struct
{
    int N;            /* prediction_horizon                         R/trajectory_optimizer.py:69 */
    int solver_type;  /* RMPC_SQP (default) | RMPC_SQP_RTI          R/trajectory_optimizer.py:60, R/simulator.py:129-135 */
    int max_iter;     /* nlp_solver_max_iter = 100                  R/trajectory_optimizer.py:67 */
    int qp_iter_max;  /* qp_solver_iter_max = 50                    T/acados_ocp_options.py:77 */
    double tol[4];    /* nlp tol stat, eq, ineq, comp = 1e-6        T/acados_ocp_options.py:50-53 */
    double qp_tol;    /* qp_tol = 1e-8 (all four QP tolerances)     R/trajectory_optimizer.py:63 */
    double dt;        /* Tf/N = shooting interval = cost scaling    R/simulator.py:101, T/acados_ocp.py:1122-1123 */
    double w_task[5]; /* 50,50,50,50,50                             R/trajectory_optimizer.py:44-48 */
    double w_u;       /* W entries 5..10 are 2*w_u                  R/trajectory_optimizer.py:148 */
    double w_qddot;   /* W entries 11..16                           R/trajectory_optimizer.py:150 */
    double px_ref;    /* yref[3]                                    R/trajectory_optimizer.py:112 */
    double vy_ref;    /* yref[4]                                    R/trajectory_optimizer.py:113 */
    double q_min[6], q_max[6]; /* lbx/ubx on q, stages 1..N-1       R/trajectory_optimizer.py:169-171 */
    double u_min[6], u_max[6]; /* lbu/ubu, stages 0..N-1            R/trajectory_optimizer.py:164-166 */
    ur10_params model;
    /* terminal ellipsoid h_e(x_N) = x_N' P x_N, lh_e <= h_e <= uh_e = alpha at stage N (BASELINE config 4).  P, alpha come
     * from R/_terminal_constraint.py:96-116 (terminal_ellipsoid_from_lqr); the reference MPC never wires them into its
     * OCP, so the wiring is the standard acados one: BGH constraint module with nh_e = 1
     * (T/acados_ocp_constraints.py lh_e/uh_e, AC/ocp_nlp/ocp_nlp_constraints_bgh.c:1219-1445). */
    int term;          /* 0: no terminal constraint (the reference MPC as shipped) */
    double term_P[144]; /* symmetric 12 x 12, row-major, state order [q; qd] */
    double term_lh, term_uh;
    /* 1: after every closed-loop step the primal iterate is shifted by one stage before the next solve, the usual acados MPC
     * idiom (x_k <- x_{k+1}, u_k <- u_{k+1} for k < N-1, x_{N-1} <- x_N; multipliers and merit weights stay).  The reference
     * never shifts (R/simulator.py:199-241): 0 is the reference behaviour.  SURVEY.md 8(f) #1. */
    int shift;
    /* reference leg only (the port and the device implement cond_N = N, plain RTI): qp_solver_cond_N (0 = N: no condensing;
     * SURVEY.md 8(f) #2, H/cond/x_part_cond.c) and the AS-RTI options as_rti_level (0 = 4 = STANDARD_RTI; levels A..D are
     * 0..3 in acados' enum, offset by one here so that 0 can mean "default": 1 = LEVEL A ... 4 = LEVEL D) / as_rti_iter. */
    int cond_N;
    int as_rti_level_p1;
    int as_rti_iter;
} rmpc_problem;

/* Fill every field except model.surf / model.wcv / N / dt / solver_type with the reference defaults. */
static inline void rmpc_problem_defaults(rmpc_problem *p)
{
    p->N = 20;
    p->solver_type = RMPC_SQP;
    p->max_iter = 100;
    p->qp_iter_max = 50;
    for (int i = 0; i < 4; i++) p->tol[i] = 1e-6;
    p->qp_tol = 1e-8;
    p->dt = 0.01;
    for (int i = 0; i < 5; i++) p->w_task[i] = 50.0;
    p->w_u = 0.01;
    p->w_qddot = 0.02;
    p->px_ref = 0.40;
    p->vy_ref = 0.05;
    const double qdmax[6] = {2.16, 2.16, 3.141592653589793, 3.20, 3.20, 3.20};
    for (int i = 0; i < 6; i++)
    {
        p->q_min[i] = -2.0 * 3.141592653589793;
        p->q_max[i] = +2.0 * 3.141592653589793;
        p->u_min[i] = -qdmax[i];
        p->u_max[i] = +qdmax[i];
        p->model.wcv[i] = 200.0;
    }
    const double s[6] = {-0.15, 0.15, -0.01, 0.01, 0.01, 0.0};
    for (int i = 0; i < 6; i++) p->model.surf[i] = s[i];
    p->model.Ts = 0.01;
    p->model.t_ee[0] = 0.0;
    p->model.t_ee[1] = 0.0;
    p->model.t_ee[2] = 0.1;
    p->term = 0;
    for (int i = 0; i < 144; i++) p->term_P[i] = 0.0;
    p->term_lh = 0.0;
    p->term_uh = 0.0;
    p->shift = 0;
    p->cond_N = 0;
    p->as_rti_level_p1 = 0;
    p->as_rti_iter = 1;
}

/* Closed-loop trace written by both legs (any pointer may be NULL). Layouts are row-major.
 *   z        (nsteps+1) x 12   plant state log            R/simulation_model.py:24,88
 *   u        nsteps x 6        applied control u0         R/simulator.py:213
 *   status   nsteps            solver return code         R/simulator.py:217
 *   sqp_iter nsteps            get_stats('sqp_iter')      R/simulator.py:218
 *   qp_iter  nsteps            QP iterations summed over the SQP iterations of the step
 *   res      nsteps x 4        get_residuals()            R/simulator.py:219  (zeros for SQP_RTI)
 *   cost     nsteps            get_cost()                 R/simulator.py:221
 *   time_tot nsteps            get_stats('time_tot')      R/simulator.py:220
 *   xtraj    nsteps x (N+1) x 12, utraj nsteps x N x 6    OCP primal iterate after each solve
 *   pitraj   nsteps x N x 12                              OCP dynamics multipliers after each solve
 */
typedef struct
{
    double *z, *u;
    int *status, *sqp_iter, *qp_iter;
    double *res, *cost, *time_tot;
    double *xtraj, *utraj, *pitraj;
} rmpc_trace;

#ifdef __cplusplus
}
#endif
#endif
