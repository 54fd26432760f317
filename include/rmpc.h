/* include/rmpc.h -- C ABI of librmpc.so, the batched UR10 surface-tracking MPC engine for B200 (sm_100a).
 *
 * Drop-in boundary.  The reference reaches its solver through ctypes into a per-instance
 * generated shared library plus libacados:
 *     <name>_acados_create_capsule / _create / _solve / _reset / _free      T/acados_ocp_solver.py:288-295,374-444
 *     ocp_nlp_constraints_model_set, ocp_nlp_out_set / ocp_nlp_out_get, ocp_nlp_get, ocp_nlp_eval_cost
 *                                                                          T/acados_ocp_solver.py:322-372
 *     (implementations: acados/interfaces/acados_c/ocp_nlp_interface.c:516-560,637-760,1376-1425)
 * one capsule per problem instance, one call per MPC step.  This library exposes the same operations for a BATCH of
 * B instances that live on one GPU; each entry point names the reference call it replaces.
 *
 * Conventions
 *   - plain C, no exceptions cross the boundary, never exit(): every function returns 0 on success or a negative
 *     RMPC_E* code; rmpc_last_error() gives the message.  Per-instance solver status codes are acados' own
 *     (0 SUCCESS, 1 NAN_DETECTED, 2 MAXITER, 3 MINSTEP, 4 QP_FAILURE, 6 UNBOUNDED; acados/utils/types.h:76-86).
 *   - "dev" pointers are device pointers (e.g. torch tensor.data_ptr()), FP64, contiguous, instance-major
 *     [B][n]; the callee copies, the caller keeps ownership (as acados' setters copy numpy buffers).
 *     "host" pointers are ordinary host memory.
 *   - all work is enqueued on the given stream (a cudaStream_t passed as void*, NULL = default stream) and is
 *     asynchronous with respect to the host unless stated otherwise.
 *   - a handle belongs to one device and is not re-entrant; different handles may be driven from different threads
 *     (like acados capsules, T/c_templates_tera/acados_solver.in.c:3008-3028).
 *   - the library needs a CUDA device: there is no host fallback, rmpc_create fails with RMPC_ENODEV without one.
 */
#ifndef RMPC_H
#define RMPC_H

#include "rmpc_fields.h" /* RP_* RT_* RL_* RK_* RS_* RE_* RQ_*: layouts of the blocks exchanged as arrays of doubles */

#ifdef __cplusplus
extern "C" {
#endif

#define RMPC_EINVAL (-1)  /* bad argument / unknown field */
#define RMPC_ENODEV (-2)  /* no usable CUDA device */
#define RMPC_ENOMEM (-3)  /* device allocation failed */
#define RMPC_ECUDA  (-4)  /* CUDA runtime error (message in rmpc_last_error) */

#define RMPC_SOLVER_SQP 0
#define RMPC_SOLVER_SQP_RTI 1

/* problem dimensions and solver options common to the batch (per-instance data goes through rmpc_set "par") */
typedef struct
{
    int B;            /* number of instances */
    int N;            /* prediction horizon (shooting intervals)             R/trajectory_optimizer.py:69 */
    int solver_type;  /* RMPC_SOLVER_SQP | RMPC_SOLVER_SQP_RTI               R/trajectory_optimizer.py:60 */
    int max_iter;     /* nlp_solver_max_iter (100)                           R/trajectory_optimizer.py:67 */
    int qp_iter_max;  /* qp_solver_iter_max (50)                             T/acados_ocp_options.py:77 */
    double tol[4];    /* nlp tol stat / eq / ineq / comp (1e-6)              T/acados_ocp_options.py:50-53 */
    double qp_tol;    /* qp_tol (1e-8)                                       R/trajectory_optimizer.py:63 */
    double dt;        /* shooting interval = Tf / N = model sample time      R/simulator.py:88-101 */
    double plant_dt;  /* plant integration step (the reference uses dt)      R/simulator.py:78-86 */
} rmpc_dims;

typedef struct rmpc_handle rmpc_handle;

/* Replaces <name>_acados_create_capsule + _create for B instances: allocates every workspace on `device`, zeroes the
 * iterate.  nlog = number of closed-loop steps the on-device logs can hold (0: no per-step logs). */
int rmpc_create(const rmpc_dims *dims, int device, int nlog, rmpc_handle **out);

/* Kinematic chain of the arm, shared by all instances of the handle (default: the UR10 of R/ur_description/urdf/ur10.urdf,
 * frame ee_link).  Replaces the pinocchio model the reference builds from a URDF and bakes into the generated functions
 * (R/loader.py:24-36, R/prediction_model.py:22-31: robot_name 'ur5' -> frame tool0, 'ur10' -> frame ee_link).
 *   xyzR  HOST pointer, [8][12] doubles = {xyz(3), rotation row-major(9)} of: row 0 the fixed transform between the base
 *         link and the parent of joint 1, rows 1..6 the <origin> of the six revolute joints, row 7 joint 6 -> end-effector
 *         frame (fixed joints folded)
 *   axis  HOST pointer, [6] ints: joint axis 0 = +x, 1 = +y, 2 = +z
 * Either pointer NULL restores the built-in UR10.  Call it before the first solve; no device work is done here (the table
 * travels as a kernel argument). */
int rmpc_set_chain(rmpc_handle *h, const double *xyzR, const int *axis);

/* Runtime options of the handle (ocp_nlp_solver_opts_set-style, by name).  Options:
 *   "shift"  0 | 1   closed loop only: after every step shift the primal iterate by one stage before the next solve --
 *                    x_k <- x_{k+1}, u_k <- u_{k+1} for k < N-1, x_{N-1} <- x_N; multipliers and merit weights stay -- i.e. the
 *                    loop `for i < N-1: solver.set(i,'x',solver.get(i+1,'x')); solver.set(i,'u',solver.get(i+1,'u'));
 *                    solver.set(N-1,'x',solver.get(N,'x'))` an acados MPC runs between two solves.  Default 0: the reference's
 *                    loop keeps the unshifted iterate (R/simulator.py:199-241).
 *   "graph"  0 | 1   rmpc_step_host replays a captured CUDA graph (default 1) or enqueues copy, kernel, copy one by one
 *   "statistics" 0 | 1  keep the per-iteration statistics table of the last full-SQP solve (rmpc_get "sqp_stat").
 *   "as_rti_level" 0..4  SQP_RTI only: advanced-step RTI, acados' as_rti_level (0 = A, 1 = B, 2 = C, 3 = D, 4 = standard RTI,
 *                    the default; AC/ocp_nlp/ocp_nlp_sqp_rti.c:654-1100, T/acados_ocp_options.py:1189-1208).  A PREPARATION
 *                    phase (rmpc_solve_rti_phase(h, 1)) then first works on the ADVANCED problem -- stage-0 state fixed at
 *                    the prediction phi(x_0, u_0) of the current iterate -- with one QP on the last linearisation (A),
 *                    as_rti_iter zero-order (B) / exact-gradient (C) / full SQP (D) iterations, and re-linearises;
 *                    rmpc_solve_rti_phase(h, 2) is the usual feedback.  Like acados, the combined rmpc_solve is refused
 *                    while a level is set; rmpc_closed_loop runs preparation + feedback per step and, when the feedback
 *                    succeeds but the advanced step failed (1 NaN merit, 4 QP failure), reports that status in the step
 *                    record and counts it in RS_NFAIL.  Level C has no
 *                    counterpart to check against: the reference aborts in it for a NONLINEAR_LS cost
 *                    ("ocp_nlp_cost_nls_compute_gradient not implemented").
 *   "as_rti_iter" n >= 0  iterations of the levels B, C, D (default 1)
 *   "qp_cond_N" N2 in [1, N]  qp_solver_cond_N of PARTIAL_CONDENSING_HPIPM (AC/ocp_qp/ocp_qp_partial_condensing.c:523-559,
 *                    H/cond/x_part_cond.c:36-60,486-558): the interior point works on the N2-stage QP whose stages hold the
 *                    controls of a block of stages and the block's first state.  Default N: no condensing (the reference's
 *                    setting).  The device reproduces the condensed problem's interior-point iterates -- same QP iteration
 *                    counts and solution as acados with that cond_N -- without forming the dense blocks. */
int rmpc_set_option(rmpc_handle *h, const char *name, double value);

/* Replaces <name>_acados_free + _free_capsule. */
int rmpc_destroy(rmpc_handle *h);

/* Message of the last error on this handle (h may be NULL for rmpc_create failures). */
const char *rmpc_last_error(const rmpc_handle *h);

/* Bytes of device memory owned by the handle. */
long long rmpc_device_bytes(const rmpc_handle *h);

/* Setters (copy [B][n] doubles from a device pointer).  Replaces AcadosOcpSolver.set / constraints_set / cost_set
 * (T/acados_ocp_solver.py:1743-1819).  Fields:
 *   "par"   n = RP_COUNT  per-instance parameter block (layout RP_* in rmpc_fields.h; Python: params.pack_params).  Setting
 *                         it discards every per-stage override made before.
 *   "z"     n = 12   plant state = x0 of the next solve (the reference sets lbx_0 = ubx_0 = x, R/simulator.py:210-211)
 *   "x"     n = 12   state iterate of `stage` (0..N)        solver.set(stage, 'x', .)
 *   "u"     n = 6    control iterate of `stage` (0..N-1)    solver.set(stage, 'u', .)
 *   "pi"    n = 12   dynamics multipliers of `stage`        solver.set(stage, 'pi', .)
 * Named fields of the cost and constraint modules -- AcadosOcpSolver.cost_set / constraints_set
 * (T/acados_ocp_solver.py:1863-2035) -> ocp_nlp_cost_model_set / ocp_nlp_constraints_model_set
 * (acados_c/ocp_nlp_interface.c:516-560,637-760).  `stage` = -1: every stage (writes a slice of "par" and drops that field's
 * per-stage overrides); `stage` in 0..N-1: that stage only (kept in a per-stage table [B][N][RT_COUNT] the first such call
 * creates; time-varying references and bounds):
 *   "yref"  n = 18   reference of the 18 outputs y = [g1..g5, u(6), qddot(6), manip]  cost_set(stage, 'yref', .)
 *   "lbu", "ubu"  n = 6   control bounds                                             constraints_set(stage, 'lbu' | 'ubu', .)
 *   "lbx", "ubx"  n = 6   bounds on q at stages 1..N-1; at stage 0: n = 12, the initial state (= "z"; the reference sets
 *                         both to the plant state)                                   constraints_set(stage, 'lbx' | 'ubx', .)
 * Per instance only (`stage` must be -1; they change what the reference bakes into the generated functions / keeps equal on
 * all stages, R/trajectory_optimizer.py:139-156), synchronise the stream where marked (s):
 *   "W_diag"      n = 18  diagonal of W: 5 task weights, 6 x (2 w_u), 6 x w_qddot, entry 17 ignored (weight 0)   (s)
 *   "surf_coeffs" n = 6   a..f of the target surface (R/surface.py:14-30)
 *   "wcv"         n = 6   velocity-loop bandwidths; the exact-ZOH coefficients a12, a22, b1, b2 are recomputed on the host
 *                         with libm's exp, as the reference computes them (R/prediction_model.py:87-115)          (s)
 * Terminal ellipsoid lh_e <= x_N' P_e x_N <= alpha_e at stage N (BASELINE config 4; P_e, alpha_e from
 * R/_terminal_constraint.py:96-116 terminal_ellipsoid_from_lqr; the acados equivalent is constraints.lh_e / uh_e with
 * con_h_expr_e, i.e. ocp_nlp_constraints_model_set(.., N, "lh" / "uh", ..), acados_solver.in.c:2109-2116).  The first of
 * these calls switches the row on for the handle; set all three (or "term") before the first solve:
 *   "P_e"     n = 144  symmetric 12 x 12, row-major, state order [q; qd]
 *   "alpha_e" n = 1    upper bound uh_e
 *   "lh_e"    n = 1    lower bound (keep it finite and negative, e.g. -alpha_e: the row is two-sided like acados')
 *   "term"    n = 146  [P_e(144), lh_e, alpha_e] in one call
 */
int rmpc_set(rmpc_handle *h, const char *field, int stage, const double *dev, void *stream);

/* Re-initialise every instance as a freshly created solver started at the current plant state "z":
 * x_k = z, u_k = 0, multipliers and merit weights 0, logs and summaries cleared
 * (T/c_templates_tera/acados_solver.in.c:2744-2777 initial guess; :2883-2932 reset). */
int rmpc_reset(rmpc_handle *h, void *stream);

/* One SQP / SQP_RTI solve for all instances with x0 = "z", without advancing the plant.
 * Replaces solver.solve() -> <name>_acados_solve -> ocp_nlp_solve (T/acados_ocp_solver.py:468-476). */
int rmpc_solve(rmpc_handle *h, void *stream);

/* SQP_RTI only: the solve split in the two phases of acados' real-time iteration (opts 'rti_phase',
 * AC/ocp_nlp/ocp_nlp_sqp_rti.c:455-646,1123-1175; T/acados_ocp_solver.py options_set('rti_phase', .)):
 *   phase 1 PREPARATION  everything that does not depend on x0 (model functions, Gauss-Newton data, gaps, bound residuals)
 *   phase 2 FEEDBACK     x0-dependent right-hand side from the current "z", QP solve, globalisation, iterate update
 *   phase 0              both (= rmpc_solve)
 * Phase 1 followed by phase 2 gives bit-identical results to phase 0 for the same "z" (in acados: equal within the QP tolerance,
 * tests/test_oracle_port.py); set "z" between the two calls. */
int rmpc_solve_rti_phase(rmpc_handle *h, int phase, void *stream);

/* Checkpoint / resume of the whole batch: iterate, multipliers, merit weights, QP memory, plant state, flags, accumulated
 * summaries and the step counter in one device blob of rmpc_state_size(h) doubles (the per-step logs are not part of it).
 * A closed loop continued after rmpc_load_state is bit-identical to one that was never interrupted.  Batch counterpart of
 * AcadosOcpSolver.store_iterate / load_iterate (T/acados_ocp_solver.py:1158-1440).  rmpc_load_state checks that the blob
 * comes from a handle with the same B and N (RMPC_EINVAL otherwise) and synchronises the stream once to do so. */
long long rmpc_state_size(const rmpc_handle *h);
int rmpc_store_state(rmpc_handle *h, double *dev, void *stream);
int rmpc_load_state(rmpc_handle *h, const double *dev, void *stream);

/* nsteps closed-loop steps for all instances entirely on the device:
 *   x0 <- z; solve; u0; z <- RK4(z, u0); log; accumulate tracking-error sums.     R/simulator.py:199-241 */
int rmpc_closed_loop(rmpc_handle *h, int nsteps, void *stream);

/* Getters (copy to a device pointer).  Replaces AcadosOcpSolver.get / get_stats / get_residuals / get_cost
 * (T/acados_ocp_solver.py:939-989,1461-1723).  Fields, with the shape written to `dev`:
 *   "x" [B][12], "u" [B][6], "pi" [B][12], "lam" [B][24] ([lower(12); upper(12)] over [u; q]) of `stage`;
 *       with a terminal row, "lam" of stage N holds its multipliers in entries 0 (lower) and 12 (upper)
 *   "z" [B][12] plant state
 *   "stats" [B][RL_COUNT]  record of the last step: status, sqp_iter, qp_iter, cost, res_stat, res_eq, res_ineq, res_comp,
 *                    alpha, qp_status, device times of the solve in seconds (RL_* in rmpc_fields.h); needs nlog > 0
 *   "zlog" [B][nlog+1][12], "ulog" [B][nlog][6], "slog" [B][nlog][RL_COUNT]   closed-loop logs
 *   "plog" [B][nlog+1][28]  per-sample plant record computed from zlog: pose of frame ee_link (12), rpy (3), WORLD
 *                    twist (6), tracking errors e1..e5 (5), task-frame z (RK_* in rmpc_fields.h)
 *                    -- the logs of R/simulation_model.py:60-77,85-91 and the error series of R/simulator.py:265-344
 *   "summary" [B][20]  RS_* in rmpc_fields.h: sum e^2 (5), ITSE (5), failures, sqp/qp iteration totals,
 *                      max KKT residuals (4), last cost, number of samples
 *   "flags" [B][4] as doubles: first_call, qp_warm_pending, LQ-fallback triggers, reserved
 *   "par" [B][RP_COUNT]; "yref" [B][18], "lbu" / "ubu" / "lbx" / "ubx" [B][6] of `stage`: the values that stage uses
 *   "sqp_stat" [B][max_iter + 2][RQ_COUNT]  per-iteration table of the last full-SQP solve (acados' mem->stat rows: KKT
 *                    residuals at iterate k; QP status, QP iterations, step length of the step that led to it); needs
 *                    rmpc_set_option(h, "statistics", 1)                      get_stats('statistics')
 * The step record ("stats", rows of "slog"; RL_*) carries acados' timing breakdown of each instance's own solve, measured on
 * the GPU's global timer: time_tot, time_lin, time_qp (= solver call + xcond), time_glob, plus the interior-point phases
 * (factorisation, forward sweeps, backward substitutions, update + residuals) -- AC/ocp_nlp/ocp_nlp_common.c:4610-4681.
 */
int rmpc_get(rmpc_handle *h, const char *field, int stage, double *dev, void *stream);

/* Number of closed-loop steps taken since create / reset. */
int rmpc_steps_done(const rmpc_handle *h);

/* Host-buffer convenience used for end-to-end measurements: upload par [B][RP_COUNT] and z0 [B][12] from host memory,
 * reset, run nsteps closed-loop steps, download the summary [B][20] and the last applied control [B][6] to host
 * memory; synchronises the stream before returning.  Any output pointer may be NULL.  Uploading the parameter block
 * discards every per-stage override made before, like rmpc_set "par". */
int rmpc_run_host(rmpc_handle *h, const double *par_host, const double *z0_host, int nsteps, double *summary_host,
                  double *u_last_host, void *stream);

/* One step of an EXTERNALLY driven MPC loop with host buffers -- the reference's per-step sequence
 *     solver.set(0,'lbx',x); solver.set(0,'ubx',x); status = solver.solve(); u = solver.get(0,'u')      R/simulator.py:209-217
 * for the whole batch in one call: copy x0 [B][12] from host memory, solve, copy u0 [B][6] (and, when stats_host is not
 * NULL and the handle has nlog > 0, the step record [B][RL_COUNT]) back to host memory; synchronises the stream before
 * returning.  The plant is the caller's (the on-device plant does not advance).  The copy-solve-copy sequence is captured
 * once into a CUDA graph and replayed (rmpc_set_option "graph" 0 turns that off; a change of the options, the kinematic chain
 * or the stream re-captures it); stream NULL = a stream owned by the handle.  Refused (RMPC_EINVAL) while an AS-RTI level is
 * set, like rmpc_solve: it is one combined preparation + feedback solve. */
int rmpc_step_host(rmpc_handle *h, const double *x0_host, double *u0_host, double *stats_host, void *stream);

/* One cold-started solve WITHOUT any device allocation.  No counterpart in the reference: acados allocates each capsule's
 * memory when the capsule is created and has no allocation-free mode; this entry point exists for hosts of this library
 * that are not free to allocate GPU memory themselves --
 *   - a controller thread with a hard per-cycle deadline (the 1 kHz robot loop the MPC feeds): cudaMalloc and cudaFree take
 *     driver-wide locks, may wait for the whole device to go idle and have no bounded run time, so they cannot sit inside
 *     or next to the cycle;
 *   - a process whose device memory is already committed to somebody else's pool (a framework's caching allocator, a
 *     fixed per-client budget on a shared GPU): a cudaMalloc from a plug-in either fails or takes memory the owner counted on.
 * Every other path of this library creates a handle first, and rmpc_create allocates the solver's work set; here the work
 * set lives in a fixed array inside librmpc.so's device image instead.  The price: each GPU that loads the library sets
 * aside RMPC_STATIC_SCRATCH_BYTES for that array until the process exits, whether this function is ever called or not.
 *
 * What it computes: exactly what  rmpc_create(dims, device, 1) ; rmpc_set "par" ; rmpc_set "z" = x0 ; rmpc_reset ; rmpc_solve ;
 * rmpc_get "u" of stage 0 (and rmpc_get "stats")  computes, for the built-in UR10 chain and default options -- a NEW solver
 * per call (iterate x_k = x0, u_k = 0, multipliers, merit weights and QP memory zero), one SQP / SQP_RTI solve, no plant step.
 *   dims      batch and solver settings, as for rmpc_create
 *   device    CUDA device ordinal (below 64); the caller's current device is put back before returning
 *   par_dev   in   [B][RP_COUNT]  parameter blocks (device pointer, read only)
 *   x0_dev    in   [B][12]        initial states (device pointer, read only)
 *   u0_dev    out  [B][6]         control of stage 0
 *   stats_dev out  [B][RL_COUNT]  step record (status, iteration counts, cost, KKT residuals, timers); may be NULL
 * All four buffers are the caller's; the work is enqueued on `stream` and is asynchronous like rmpc_solve.
 *
 * Limit: B * (8 * S(N) + 16) <= RMPC_STATIC_SCRATCH_BYTES, where S(N) = 144 + 832 * (N + 1 rounded up to even), rounded up to
 * a multiple of 16, is the number of doubles one instance's solver data takes -- 98 instances at N = 100, 454 at N = 20,
 * 49 at N = 200.  A larger request returns RMPC_EINVAL with the numbers in rmpc_last_error(NULL); nothing is launched.
 *
 * Things a caller must know:
 *   - the FIRST call on each device does a one-time set-up (looks up the device address of the array, raises the kernel's
 *     shared-memory limit), which makes the CUDA runtime load the module on that device.  Make that first call during
 *     start-up -- ahead of the time-critical loop and before any stream capture is begun -- e.g. with B = 1.
 *   - it calls no cudaMalloc / cudaFree / cudaHostAlloc and no library routine that allocates device memory, on the first
 *     call or any later one.
 *   - there is ONE array per device and every call uses it from the start.  It still holds the previous call's data, so
 *     each call first zeroes the part it is going to use; nothing is carried from call to call (no warm start -- callers
 *     that want one need a handle).  Two calls whose GPU work can overlap (different streams, different threads) write the
 *     same memory and corrupt each other: order them on one stream or with events.  Handles are not affected.
 */
#define RMPC_STATIC_SCRATCH_BYTES (64u << 20)
int rmpc_solve_static(const rmpc_dims *dims, int device, const double *par_dev, const double *x0_dev, double *u0_dev,
                      double *stats_dev, void *stream);

/* Kernels launched by this handle so far (for bench accounting). */
long long rmpc_launch_count(const rmpc_handle *h);

#ifdef __cplusplus
}
#endif
#endif
