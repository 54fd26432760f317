/* include/rmpc_fields.h -- layouts of the blocks that cross the C ABI of librmpc.so (include/rmpc.h) as plain arrays of
 * doubles.  A C caller fills / reads them with these offsets; the Python host side mirrors them in
 * robotic-mpc_b200/params.py (RP, RL, RK, RS, RE, RT dictionaries) and the kernels include this very file.
 *
 * R/ = reference root, T/ = acados_template, AC/ = acados/acados.
 */
#ifndef RMPC_FIELDS_H
#define RMPC_FIELDS_H

/* ---- "par": per-instance parameter block [B][RP_COUNT]  (rmpc_set "par"; the named per-instance fields of rmpc_set write
 * single slices of it).  What the reference bakes into one generated solver: surface, exact-ZOH model, weights W, reference
 * yref, box bounds. */
enum
{
    RP_SURF = 0,    /* 6 : a b c d e f of z = a x^2 + b y^2 + c x y + d x + e y + f      R/surface.py:14-30 */
    RP_A12 = 6,     /* 6 : (1 - a22) / w     exact-ZOH prediction model, per joint        R/prediction_model.py:87-115 */
    RP_A22 = 12,    /* 6 : exp(-w Ts) */
    RP_B1 = 18,     /* 6 : Ts - a12 */
    RP_B2 = 24,     /* 6 : 1 - a22 */
    RP_WCV = 30,    /* 6 : velocity-loop bandwidths w (prediction model and plant)        R/simulation_model.py:79-83 */
    RP_WU = 36,     /* 6 : w_u per joint; W entries 5..10 are 2 w_u                       R/trajectory_optimizer.py:148 */
    RP_WQDD = 42,   /* 6 : w_qddot per joint; W entries 11..16                            R/trajectory_optimizer.py:150 */
    RP_YREF = 48,   /* 18: yref = [0, 1, 0, px_ref, vy_ref, 0 (6), 0 (6), 0]              R/trajectory_optimizer.py:108-114,153 */
    RP_PXREF = 51,  /*     = RP_YREF + 3 */
    RP_VYREF = 52,  /*     = RP_YREF + 4 */
    RP_WTASK = 66,  /* 5 : task weights, W entries 0..4 (50 each)                         R/trajectory_optimizer.py:44-48 */
    RP_UMIN = 72,   /* 6 : lbu   (stages 0..N-1)                                          R/trajectory_optimizer.py:164-166 */
    RP_UMAX = 78,   /* 6 : ubu */
    RP_QMIN = 84,   /* 6 : lbx on q (stages 1..N-1)                                       R/trajectory_optimizer.py:169-171 */
    RP_QMAX = 90,   /* 6 : ubx on q */
    RP_TEE = 96,    /* 3 : task-frame offset in the end-effector frame                    R/prediction_model.py:9 */
    RP_COUNT = 100
};

/* ---- per-stage table [B][N][RT_COUNT], optional: exists once a per-stage field (stage >= 0) has been set through
 * rmpc_set ("yref", "lbu", "ubu", "lbx", "ubx"); until then every stage uses the per-instance values of "par".
 * AcadosOcpSolver.cost_set(stage, 'yref', .) / constraints_set(stage, 'lbu' | 'ubu' | 'lbx' | 'ubx', .)
 * (T/acados_ocp_solver.py:1863-2035 -> ocp_nlp_cost_model_set / ocp_nlp_constraints_model_set). */
enum
{
    RT_YREF = 0,  /* 18: yref of the stage (entry 17, the manipulability output, has weight 0) */
    RT_LB = 18,   /* 12: [lbu(6); lbx on q(6)] */
    RT_UB = 30,   /* 12: [ubu(6); ubx on q(6)] */
    RT_COUNT = 42
};

/* ---- per-step log record [RL_COUNT] written by the closed loop / a solve ("stats", "slog") */
enum
{
    RL_STATUS = 0, RL_SQP_ITER = 1, RL_QP_ITER = 2, RL_COST = 3,
    RL_RES = 4,          /* 4: stat eq ineq comp (zeros for SQP_RTI, like the reference) */
    RL_ALPHA = 8,        /* accepted line-search step of the last SQP iteration */
    RL_QP_STATUS = 9,
    /* device timers of this instance's solve, seconds on the GPU's global timer; the names are acados'
     * (AC/ocp_nlp/ocp_nlp_common.c:4610-4681, T/acados_ocp_solver.py get_stats) */
    RL_TIME = 10,        /* time_tot */
    RL_TIME_LIN = 11,    /* time_lin: model functions, Gauss-Newton data, KKT residuals */
    RL_TIME_QP = 12,     /* time_qp (= time_qp_solver_call + time_qp_xcond): reduce, interior point, restore */
    RL_TIME_GLOB = 13,   /* time_glob: merit line search and iterate update */
    RL_TIME_QP_FACT = 14,  /* inside time_qp: Riccati factorisations */
    RL_TIME_QP_FWD = 15,   /* forward sweeps (affine step, corrector) */
    RL_TIME_QP_BWD = 16,   /* backward substitutions */
    RL_TIME_QP_RES = 17,   /* update + residual sweeps, initialisation */
    RL_TIME_XCOND = 18,    /* time_qp_xcond: x0 elimination and restore */
    RL_COUNT = 20
};

/* ---- per-sample plant record [RK_COUNT] computed from the logged plant states ("plog";
 * R/simulation_model.py:60-77, R/simulator.py:265-344) */
enum
{
    RK_POSE = 0,    /* 12: p_ee(3), R_ee row-major(9) of the end-effector frame */
    RK_RPY = 12,    /* 3 : roll pitch yaw of R_ee */
    RK_TWIST = 15,  /* 6 : WORLD twist J qd */
    RK_ERR = 21,    /* 5 : tracking errors e1..e5 */
    RK_PTZ = 26,    /* 1 : z of the task-frame origin */
    RK_COUNT = 28
};

/* ---- summary row per instance accumulated on the device ("summary"; R/simulator.py:347-390 metrics + solver_stats) */
enum
{
    RS_SSE = 0,      /* 5: sum e_k^2 over the Nsim+1 logged samples */
    RS_ITSE = 5,     /* 5: sum t * e_k^2 * dt */
    RS_NFAIL = 10,   /* steps with status != 0 */
    RS_SQP_ITERS = 11,
    RS_QP_ITERS = 12,
    RS_MAX_RES = 13, /* 4: max over steps of each KKT residual */
    RS_COST_LAST = 17,
    RS_NSAMPLES = 18,
    RS_COUNT = 20
};

/* ---- per-instance terminal-set block [RE_COUNT] ("term"): h_e(x_N) = x_N' P x_N with lh <= h_e <= uh at stage N
 * (R/_terminal_constraint.py:96-116 gives P, alpha = uh; acados BGH nh_e = 1) */
enum
{
    RE_P = 0,     /* 144: P, symmetric, row-major, state order [q; qd] */
    RE_LH = 144,
    RE_UH = 145,
    RE_COUNT = 146
};

/* ---- "sqp_stat": per-iteration statistics of the last full-SQP solve, [B][max_iter + 2][RQ_COUNT] -- the rows of
 * acados' mem->stat (AC/ocp_nlp/ocp_nlp_sqp.c:573-576,700-701,788): row k = KKT residuals at iterate k, and the QP status,
 * QP iterations and accepted step length of the step that LED to iterate k.  get_stats('statistics') transposes it. */
enum
{
    RQ_RES = 0,  /* 4: res_stat res_eq res_ineq res_comp */
    RQ_QP_STATUS = 4,
    RQ_QP_ITER = 5,
    RQ_ALPHA = 6,
    RQ_COUNT = 7
};

#endif
