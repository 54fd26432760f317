/* oracle/ur10_model.h -- CPU restatement of the reference's model functions (TEST INFRASTRUCTURE).
 *
 * Nothing under oracle/ is on the product path: only tests/, __graft_entry__.smoke() and
 * bench.py's cpu_baseline / --impl reference legs may use it, and only as the checker.
 *
 * What is restated here is the arithmetic the reference obtains from two third-party packages that
 * are NOT vendored in /root/reference and cannot be installed here:
 *   - pinocchio (R/requirements.txt:3, unpinned): URDF forward kinematics and WORLD-frame frame
 *     Jacobian of frame `ee_link` (call sites R/prediction_model.py:126-132,163-164,
 *     R/simulation_model.py:60-77), chain constants from R/ur_description/urdf/ur10.urdf:50-222.
 *   - CasADi >= 3.6 (R/requirements.txt:1): symbolic outputs and their Jacobian, C-generated as
 *     *_cost_y_fun, *_cost_y_fun_jac_ut_xt, *_dyn_disc_phi_fun[_jac]
 *     (T/casadi_function_generation.py:244-255,555-575).
 * PARITY OF THESE FUNCTIONS IS UNPINNED by the reference (it has no tests / golden data); they are
 * pinned here by finite differences, UR10 DH known answers and the exact-ZOH matrix exponential
 * (tests/test_oracle_model.py) and by an independent PyTorch FP64 restatement differentiated with
 * autograd, agreeing to 1e-12 (tests/test_oracle_model_torch.py).
 *
 * Conventions: x = [q(6); qdot(6)], u = qdot_cmd(6), stage variable order [u; x] (AC/ocp_nlp/ocp_nlp_common.h:281).
 */
#ifndef RMPC_ORACLE_UR10_MODEL_H
#define RMPC_ORACLE_UR10_MODEL_H

#ifdef __cplusplus
extern "C" {
#endif

#define UR10_NQ 6
#define UR10_NX 12
#define UR10_NU 6
#define UR10_NY 18
#define UR10_NV 18 /* nu + nx */

/* Per-instance problem data (what differs between batched instances). */
typedef struct
{
    double surf[6];  /* a,b,c,d,e,f of S(x,y)=a x^2+b y^2+c xy+d x+e y+f   (R/surface.py:14-30) */
    double wcv[6];   /* velocity-loop bandwidths w                          (R/prediction_model.py:90) */
    double Ts;       /* sample time dt                                      (R/prediction_model.py:91) */
    double t_ee[3];  /* task-frame offset in EE frame, default (0,0,0.1)    (R/prediction_model.py:9) */
} ur10_params;

/* Forward kinematics of the UR10 chain (pinocchio oMi / oMf for frame ee_link).
 * p_joint[j] = world position of joint j origin, a_joint[j] = world axis of joint j,
 * R_ee (row-major 3x3) / p_ee = placement of frame ee_link. */
void ur10_fk(const double q[6], double p_joint[6][3], double a_joint[6][3], double R_ee[9],
             double p_ee[3]);

/* Replace the built-in UR10 chain by a table of rigid transforms [xyz(3); R row-major(9)] x 8 (root, joints 1..6, joint 6 ->
 * end-effector frame) and joint axes (0 x, 1 y, 2 z); NULL restores UR10.  Process-global (test infrastructure). */
void ur10_model_set_chain(const double *xyzR, const int *axis);

/* pinocchio computeFrameJacobian(..., ee_link, WORLD): J (row-major 6x6), column j = [p_j x a_j ; a_j]. */
void ur10_jacobian_world(const double q[6], double J[36]);

/* Exact-ZOH discrete LTI coefficients (R/prediction_model.py:87-115): per joint
 * a22=exp(-w Ts), a12=(1-a22)/w, b1=Ts-a12, b2=1-a22. */
void ur10_lti_coeffs(const ur10_params *P, double a12[6], double a22[6], double b1[6], double b2[6]);

/* x+ = Ad x + Bd u (R/prediction_model.py:316-320). */
void ur10_dyn(const ur10_params *P, const double x[12], const double u[6], double xn[12]);

/* [B;A]^T as a column-major (nu+nx) x nx matrix (ld = 18): what *_dyn_disc_phi_fun_jac returns. */
void ur10_dyn_jac_BAbt(const ur10_params *P, double BAbt[18 * 12]);

/* Cost output y(x,u) in R^18 (R/trajectory_optimizer.py:88-154). */
void ur10_cost_y(const ur10_params *P, const double x[12], const double u[6], double y[18]);

/* y and Jt = (dy/d[u;x])^T, column-major (nu+nx) x ny with ld = 18: Jt[r + 18*c] = d y_c / d v_r.
 * Column 17 (derivative of the manipulability output exp(-2 sqrt(det(J J^T)+0.01))) is returned as
 * ZERO: that output has weight 0 in W (R/trajectory_optimizer.py:133,152), so acados multiplies the
 * column by W^{1/2}[17]=0 before every use (AC/ocp_nlp/ocp_nlp_cost_nls.c:897,936,961) and it
 * contributes exactly 0.0 to cost, gradient and Gauss-Newton Hessian.  y[17] itself is exact. */
void ur10_cost_y_jac(const ur10_params *P, const double x[12], const double u[6], double y[18],
                     double Jt[18 * 18]);

/* Plant (R/simulation_model.py:79-83,111-117): one RK4 step of zdot=[qdot; -W qdot + W u]. */
void ur10_plant_rk4(const double wcv[6], double dt, const double z[12], const double u[6],
                    double zn[12]);

/* Plant logs (R/simulation_model.py:60-77): pose = [p_ee(3); R_ee row-major(9)], twist = J_world qdot. */
void ur10_plant_logs(const double z[12], double pose[12], double rpy[3], double twist[6]);

/* Post-hoc tracking errors e1..e5 at one logged sample (R/simulator.py:301-342), incl. p_task_z. */
void ur10_tracking_errors(const ur10_params *P, const double pose[12], const double twist[6],
                          double px_ref, double vy_ref, double e[5], double *p_task_z);

#ifdef __cplusplus
}
#endif
#endif
