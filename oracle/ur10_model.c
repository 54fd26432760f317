/* oracle/ur10_model.c -- CPU restatement of the reference's model functions (TEST INFRASTRUCTURE).
 * See ur10_model.h for scope, provenance (pinocchio / CasADi call sites) and the parity statement.
 * Plain C99, no dependencies.  Every function cites the reference lines it follows.
 */
#include "ur10_model.h"

#include <math.h>
#include <string.h>

/* ---------------------------------------------------------------- small 3-vector helpers */
static inline void cross3(const double a[3], const double b[3], double c[3])
{
    c[0] = a[1] * b[2] - a[2] * b[1];
    c[1] = a[2] * b[0] - a[0] * b[2];
    c[2] = a[0] * b[1] - a[1] * b[0];
}
static inline double dot3(const double a[3], const double b[3])
{
    return a[0] * b[0] + a[1] * b[1] + a[2] * b[2];
}
/* C = A*B, row-major 3x3 */
static void mat3_mul(const double A[9], const double B[9], double C[9])
{
    for (int i = 0; i < 3; i++)
        for (int j = 0; j < 3; j++)
            C[3 * i + j] = A[3 * i] * B[j] + A[3 * i + 1] * B[3 + j] + A[3 * i + 2] * B[6 + j];
}
static inline void mat3_vec(const double A[9], const double v[3], double r[3])
{
    for (int i = 0; i < 3; i++) r[i] = A[3 * i] * v[0] + A[3 * i + 1] * v[1] + A[3 * i + 2] * v[2];
}

/* URDF rpy -> rotation the way pinocchio's URDF parser does it: urdfdom Rotation::setFromRPY builds a
 * (normalised) quaternion from half angles, pinocchio converts it with Eigen's toRotationMatrix. */
static void rpy_to_rot(double roll, double pitch, double yaw, double R[9])
{
    double phi = roll / 2.0, the = pitch / 2.0, psi = yaw / 2.0;
    double x = sin(phi) * cos(the) * cos(psi) - cos(phi) * sin(the) * sin(psi);
    double y = cos(phi) * sin(the) * cos(psi) + sin(phi) * cos(the) * sin(psi);
    double z = cos(phi) * cos(the) * sin(psi) - sin(phi) * sin(the) * cos(psi);
    double w = cos(phi) * cos(the) * cos(psi) + sin(phi) * sin(the) * sin(psi);
    double s = sqrt(x * x + y * y + z * z + w * w);
    x /= s; y /= s; z /= s; w /= s;
    double tx = 2 * x, ty = 2 * y, tz = 2 * z;
    double twx = tx * w, twy = ty * w, twz = tz * w;
    double txx = tx * x, txy = ty * x, txz = tz * x;
    double tyy = ty * y, tyz = tz * y, tzz = tz * z;
    R[0] = 1 - (tyy + tzz); R[1] = txy - twz;       R[2] = txz + twy;
    R[3] = txy + twz;       R[4] = 1 - (txx + tzz); R[5] = tyz - twx;
    R[6] = txz - twy;       R[7] = tyz + twx;       R[8] = 1 - (txx + tyy);
}

/* UR10 chain constants, R/ur_description/urdf/ur10.urdf:50-55,78-83,106-111,134-139,162-167,190-195
 * (revolute joints) and :218-222 (fixed ee_fixed_joint -> frame ee_link); world_joint :286-290 is identity. */
static const double UR10_XYZ[6][3] = {
    {0.0, 0.0, 0.1273}, {0.0, 0.220941, 0.0}, {0.0, -0.1719, 0.612},
    {0.0, 0.0, 0.5723}, {0.0, 0.1149, 0.0},   {0.0, 0.0, 0.1157}};
static const double UR10_RPY[6][3] = {
    {0, 0, 0}, {0.0, 1.570796325, 0.0}, {0, 0, 0}, {0.0, 1.570796325, 0.0}, {0, 0, 0}, {0, 0, 0}};
static const int UR10_AXIS[6] = {2, 1, 1, 1, 2, 1}; /* z y y y z y */
static const double UR10_EE_XYZ[3] = {0.0, 0.0922, 0.0};
static const double UR10_EE_RPY[3] = {-1.5707963267948966, 0.0, 0.0};

/* Another serial 6R chain (e.g. the reference's UR5 with frame tool0, R/prediction_model.py:28-31,
 * R/ur_description/urdf/ur5.urdf) as a table of rigid transforms [xyz(3); R row-major(9)]: row 0 = root (fixed joints
 * between the base link and joint 1's parent), rows 1..6 = joint origins, row 7 = joint 6 -> end-effector frame (fixed
 * joints folded), plus the joint axes (0 x, 1 y, 2 z).  Process-global: the oracle is test infrastructure and a test
 * uses one robot at a time.  NULL restores the built-in UR10. */
static int g_custom = 0;
static double g_chain[8][12];
static int g_axis[6];
void ur10_model_set_chain(const double *xyzR /* [8][12] or NULL */, const int *axis /* [6] */)
{
    g_custom = xyzR != NULL;
    if (!g_custom) return;
    memcpy(g_chain, xyzR, sizeof(g_chain));
    memcpy(g_axis, axis, sizeof(g_axis));
}

/* pinocchio forwardKinematics + updateFramePlacements (R/prediction_model.py:126-132). */
void ur10_fk(const double q[6], double p_joint[6][3], double a_joint[6][3], double R_ee[9],
             double p_ee[3])
{
    double R[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1}; /* parent orientation */
    double p[3] = {0, 0, 0};                   /* parent origin */
    if (g_custom)
    {
        memcpy(p, g_chain[0], sizeof(p));
        memcpy(R, g_chain[0] + 3, sizeof(R));
    }
    for (int j = 0; j < 6; j++)
    {
        double Ro[9], Rj[9], t[3], Rq[9], Rn[9];
        if (g_custom) memcpy(Ro, g_chain[1 + j] + 3, sizeof(Ro));
        else rpy_to_rot(UR10_RPY[j][0], UR10_RPY[j][1], UR10_RPY[j][2], Ro);
        mat3_mul(R, Ro, Rj);         /* joint frame orientation before its own rotation */
        mat3_vec(R, g_custom ? g_chain[1 + j] : UR10_XYZ[j], t); /* joint origin */
        for (int k = 0; k < 3; k++) p[k] += t[k];
        for (int k = 0; k < 3; k++) p_joint[j][k] = p[k];
        int ax = g_custom ? g_axis[j] : UR10_AXIS[j];
        for (int k = 0; k < 3; k++) a_joint[j][k] = Rj[3 * k + ax]; /* Rj * e_ax */
        double c = cos(q[j]), s = sin(q[j]);
        if (ax == 2) { /* Rz */
            Rq[0] = c; Rq[1] = -s; Rq[2] = 0; Rq[3] = s; Rq[4] = c; Rq[5] = 0; Rq[6] = 0; Rq[7] = 0; Rq[8] = 1;
        } else if (ax == 1) { /* Ry */
            Rq[0] = c; Rq[1] = 0; Rq[2] = s; Rq[3] = 0; Rq[4] = 1; Rq[5] = 0; Rq[6] = -s; Rq[7] = 0; Rq[8] = c;
        } else { /* Rx */
            Rq[0] = 1; Rq[1] = 0; Rq[2] = 0; Rq[3] = 0; Rq[4] = c; Rq[5] = -s; Rq[6] = 0; Rq[7] = s; Rq[8] = c;
        }
        mat3_mul(Rj, Rq, Rn);
        memcpy(R, Rn, sizeof(Rn));
    }
    double Re[9], t[3];
    if (g_custom) memcpy(Re, g_chain[7] + 3, sizeof(Re));
    else rpy_to_rot(UR10_EE_RPY[0], UR10_EE_RPY[1], UR10_EE_RPY[2], Re);
    mat3_mul(R, Re, R_ee);
    mat3_vec(R, g_custom ? g_chain[7] : UR10_EE_XYZ, t);
    for (int k = 0; k < 3; k++) p_ee[k] = p[k] + t[k];
}

/* pinocchio computeFrameJacobian(model, data, q, ee_link, WORLD) (R/prediction_model.py:163-164):
 * spatial Jacobian referred to the world origin, column j = [p_j x a_j ; a_j]. */
void ur10_jacobian_world(const double q[6], double J[36])
{
    double pj[6][3], aj[6][3], R[9], p[3], c[3];
    ur10_fk(q, pj, aj, R, p);
    for (int j = 0; j < 6; j++)
    {
        cross3(pj[j], aj[j], c);
        for (int k = 0; k < 3; k++)
        {
            J[6 * k + j] = c[k];
            J[6 * (3 + k) + j] = aj[j][k];
        }
    }
}

/* R/prediction_model.py:87-115 */
void ur10_lti_coeffs(const ur10_params *P, double a12[6], double a22[6], double b1[6], double b2[6])
{
    for (int i = 0; i < 6; i++)
    {
        a22[i] = exp(-P->wcv[i] * P->Ts);
        a12[i] = (1.0 - a22[i]) / P->wcv[i];
        b2[i] = 1.0 - a22[i];
        b1[i] = P->Ts - a12[i];
    }
}

/* R/prediction_model.py:316-320: x+ = Ad x + Bd u */
void ur10_dyn(const ur10_params *P, const double x[12], const double u[6], double xn[12])
{
    double a12[6], a22[6], b1[6], b2[6];
    ur10_lti_coeffs(P, a12, a22, b1, b2);
    for (int i = 0; i < 6; i++)
    {
        xn[i] = (x[i] + a12[i] * x[6 + i]) + b1[i] * u[i];
        xn[6 + i] = a22[i] * x[6 + i] + b2[i] * u[i];
    }
}

/* jac_ux_t of T/casadi_function_generation.py:244-255: (nu+nx) x nx, rows [u; x], column-major */
void ur10_dyn_jac_BAbt(const ur10_params *P, double BAbt[18 * 12])
{
    double a12[6], a22[6], b1[6], b2[6];
    ur10_lti_coeffs(P, a12, a22, b1, b2);
    memset(BAbt, 0, sizeof(double) * 18 * 12);
    for (int i = 0; i < 6; i++)
    {
        BAbt[(i) + 18 * (i)] = b1[i];          /* d q+_i / d u_i     */
        BAbt[(i) + 18 * (6 + i)] = b2[i];      /* d qd+_i / d u_i    */
        BAbt[(6 + i) + 18 * (i)] = 1.0;        /* d q+_i / d q_i     */
        BAbt[(12 + i) + 18 * (i)] = a12[i];    /* d q+_i / d qd_i    */
        BAbt[(12 + i) + 18 * (6 + i)] = a22[i];/* d qd+_i / d qd_i   */
    }
}

/* 6x6 determinant by Gaussian elimination with partial pivoting */
static double det6(const double Jin[36])
{
    double M[36];
    memcpy(M, Jin, sizeof(M));
    double det = 1.0;
    for (int c = 0; c < 6; c++)
    {
        int piv = c;
        for (int r = c + 1; r < 6; r++)
            if (fabs(M[6 * r + c]) > fabs(M[6 * piv + c])) piv = r;
        if (M[6 * piv + c] == 0.0) return 0.0;
        if (piv != c)
        {
            for (int k = 0; k < 6; k++) { double t = M[6 * c + k]; M[6 * c + k] = M[6 * piv + k]; M[6 * piv + k] = t; }
            det = -det;
        }
        det *= M[6 * c + c];
        for (int r = c + 1; r < 6; r++)
        {
            double f = M[6 * r + c] / M[6 * c + c];
            for (int k = c; k < 6; k++) M[6 * r + k] -= f * M[6 * c + k];
        }
    }
    return det;
}

/* Shared evaluation.  If Jt != NULL also the transposed Jacobian.
 * Follows R/prediction_model.py:226-326 (task frame, v_task, qddot_k), R/surface.py:21,243-262 (S, n),
 * R/trajectory_optimizer.py:88-154 (g1..g5, y). */
static void cost_eval(const ur10_params *P, const double x[12], const double u[6], double y[18],
                      double *Jt)
{
    const double *q = x, *qd = x + 6;
    double pj[6][3], aj[6][3], R[9], pe[3];
    ur10_fk(q, pj, aj, R, pe);

    /* task frame: R_w_t = R_w_ee (the "rotation of pi about x" is the identity, :264-271);
       translation_w = R_w_ee t ; p_t = p_ee + translation_w (:272-273) */
    double tw[3], pt[3];
    mat3_vec(R, P->t_ee, tw);
    for (int k = 0; k < 3; k++) pt[k] = pe[k] + tw[k];
    /* columns of R_w_t: y[3:6]=col0, y[6:9]=col1, y[9:12]=col2 (:275-279) */
    double r1[3] = {R[1], R[4], R[7]}; /* R_task_y */
    double r2[3] = {R[2], R[5], R[8]}; /* R_task_z */

    /* twist [v; w] = J qd, then v_task = R_w_t^T (v + w x translation_w)  (:303-313) */
    double vw[3] = {0, 0, 0}, om[3] = {0, 0, 0}, c[3];
    for (int j = 0; j < 6; j++)
    {
        cross3(pj[j], aj[j], c);
        for (int k = 0; k < 3; k++) { vw[k] += c[k] * qd[j]; om[k] += aj[j][k] * qd[j]; }
    }
    double wv[3]; /* v + w x translation_w */
    cross3(om, tw, c);
    for (int k = 0; k < 3; k++) wv[k] = vw[k] + c[k];

    /* surface */
    const double sa = P->surf[0], sb = P->surf[1], sc = P->surf[2], sd = P->surf[3], se = P->surf[4],
                 sf = P->surf[5];
    double X = pt[0], Y = pt[1];
    double S = sa * X * X + sb * Y * Y + sc * X * Y + sd * X + se * Y + sf;
    double Sx = 2 * sa * X + sc * Y + sd, Sy = 2 * sb * Y + sc * X + se;
    double m[3] = {Sx, Sy, -1.0};
    double rho = sqrt(Sx * Sx + Sy * Sy + 1.0);
    double n[3] = {m[0] / rho, m[1] / rho, m[2] / rho};

    double a12[6], a22[6], b1[6], b2[6];
    ur10_lti_coeffs(P, a12, a22, b1, b2);

    y[0] = S - pt[2];       /* g1 */
    y[1] = dot3(n, r2);     /* g2 */
    y[2] = r1[0];           /* g3 = R_task_y[0] */
    y[3] = pt[0];           /* g4 */
    y[4] = dot3(r1, wv);    /* g5 = v_t_y */
    for (int i = 0; i < 6; i++) y[5 + i] = u[i];
    for (int i = 0; i < 6; i++)
    {
        double qdn = a22[i] * qd[i] + b2[i] * u[i];
        y[11 + i] = (qdn - qd[i]) / P->Ts; /* qddot_k (:322-326) */
    }
    {
        double J[36];
        for (int j = 0; j < 6; j++)
        {
            cross3(pj[j], aj[j], c);
            for (int k = 0; k < 3; k++) { J[6 * k + j] = c[k]; J[6 * (3 + k) + j] = aj[j][k]; }
        }
        double d = det6(J); /* det(J J^T) = det(J)^2 */
        y[17] = exp(-2.0 * sqrt(d * d + 0.01));
    }
    if (!Jt) return;

    memset(Jt, 0, sizeof(double) * 18 * 18);
#define JT(r, cc) Jt[(r) + 18 * (cc)]
    /* dn/dX, dn/dY */
    double mX[3] = {2 * sa, sc, 0.0}, mY[3] = {sc, 2 * sb, 0.0};
    double nX[3], nY[3];
    {
        double mmX = dot3(m, mX), mmY = dot3(m, mY), r3 = rho * rho * rho;
        for (int k = 0; k < 3; k++)
        {
            nX[k] = mX[k] / rho - m[k] * mmX / r3;
            nY[k] = mY[k] / rho - m[k] * mmY / r3;
        }
    }
    for (int k = 0; k < 6; k++)
    {
        /* d p_t / d q_k = a_k x (p_t - p_k);  d r / d q_k = a_k x r for any EE-fixed direction r */
        double d[3] = {pt[0] - pj[k][0], pt[1] - pj[k][1], pt[2] - pj[k][2]};
        double dpt[3], dr1[3], dr2[3], dtw[3];
        cross3(aj[k], d, dpt);
        cross3(aj[k], r1, dr1);
        cross3(aj[k], r2, dr2);
        cross3(aj[k], tw, dtw);
        int row = 6 + k;
        JT(row, 0) = Sx * dpt[0] + Sy * dpt[1] - dpt[2];
        double dn[3] = {nX[0] * dpt[0] + nY[0] * dpt[1], nX[1] * dpt[0] + nY[1] * dpt[1],
                        nX[2] * dpt[0] + nY[2] * dpt[1]};
        JT(row, 1) = dot3(dn, r2) + dot3(n, dr2);
        JT(row, 2) = dr1[0];
        JT(row, 3) = dpt[0];
        /* g5 = r1 . wv,  wv = sum_j qd_j a_j x (tw - p_j) */
        double dwv[3] = {0, 0, 0};
        for (int j = 0; j < 6; j++)
        {
            double e[3] = {tw[0] - pj[j][0], tw[1] - pj[j][1], tw[2] - pj[j][2]};
            double de[3] = {dtw[0], dtw[1], dtw[2]};
            double da[3] = {0, 0, 0};
            if (k < j)
            { /* joint k precedes joint j: its rotation moves a_j and p_j */
                cross3(aj[k], aj[j], da);
                double pk[3] = {pj[j][0] - pj[k][0], pj[j][1] - pj[k][1], pj[j][2] - pj[k][2]};
                double dpj[3];
                cross3(aj[k], pk, dpj);
                for (int t = 0; t < 3; t++) de[t] -= dpj[t];
            }
            double t1[3], t2[3];
            cross3(da, e, t1);
            cross3(aj[j], de, t2);
            for (int t = 0; t < 3; t++) dwv[t] += qd[j] * (t1[t] + t2[t]);
        }
        JT(row, 4) = dot3(dr1, wv) + dot3(r1, dwv);
        /* d g5 / d qd_k = r1 . (a_k x (tw - p_k)) */
        double e[3] = {tw[0] - pj[k][0], tw[1] - pj[k][1], tw[2] - pj[k][2]}, ce[3];
        cross3(aj[k], e, ce);
        JT(12 + k, 4) = dot3(r1, ce);
    }
    for (int i = 0; i < 6; i++)
    {
        JT(i, 5 + i) = 1.0;
        JT(i, 11 + i) = b2[i] / P->Ts;
        JT(12 + i, 11 + i) = (a22[i] - 1.0) / P->Ts;
    }
    /* column 17: zero, see header */
#undef JT
}

void ur10_cost_y(const ur10_params *P, const double x[12], const double u[6], double y[18])
{
    cost_eval(P, x, u, y, 0);
}

void ur10_cost_y_jac(const ur10_params *P, const double x[12], const double u[6], double y[18],
                     double Jt[18 * 18])
{
    cost_eval(P, x, u, y, Jt);
}

/* R/simulation_model.py:79-83 */
static void plant_f(const double wcv[6], const double z[12], const double u[6], double zd[12])
{
    for (int i = 0; i < 6; i++)
    {
        zd[i] = z[6 + i];
        zd[6 + i] = -wcv[i] * z[6 + i] + wcv[i] * u[i];
    }
}

/* R/simulation_model.py:111-117 */
void ur10_plant_rk4(const double wcv[6], double dt, const double z[12], const double u[6],
                    double zn[12])
{
    double k1[12], k2[12], k3[12], k4[12], t[12];
    plant_f(wcv, z, u, k1);
    for (int i = 0; i < 12; i++) t[i] = z[i] + 0.5 * dt * k1[i];
    plant_f(wcv, t, u, k2);
    for (int i = 0; i < 12; i++) t[i] = z[i] + 0.5 * dt * k2[i];
    plant_f(wcv, t, u, k3);
    for (int i = 0; i < 12; i++) t[i] = z[i] + dt * k3[i];
    plant_f(wcv, t, u, k4);
    for (int i = 0; i < 12; i++)
        zn[i] = z[i] + (dt / 6) * k1[i] + (dt / 3) * k2[i] + (dt / 3) * k3[i] + (dt / 6) * k4[i];
}

/* R/simulation_model.py:60-77 */
void ur10_plant_logs(const double z[12], double pose[12], double rpy[3], double twist[6])
{
    double pj[6][3], aj[6][3], R[9], p[3], c[3];
    ur10_fk(z, pj, aj, R, p);
    for (int k = 0; k < 3; k++) pose[k] = p[k];
    for (int k = 0; k < 9; k++) pose[3 + k] = R[k];
    rpy[0] = atan2(R[7], R[8]);
    rpy[1] = atan2(-R[6], sqrt(R[0] * R[0] + R[3] * R[3]));
    rpy[2] = atan2(R[3], R[0]);
    for (int k = 0; k < 6; k++) twist[k] = 0.0;
    for (int j = 0; j < 6; j++)
    {
        cross3(pj[j], aj[j], c);
        for (int k = 0; k < 3; k++)
        {
            twist[k] += c[k] * z[6 + j];
            twist[3 + k] += aj[j][k] * z[6 + j];
        }
    }
}

/* R/simulator.py:301-342.  NB: e5 uses the post-hoc formula R_w_t @ (v + (w . t_w)) -- rotation NOT
 * transposed and a scalar dot product broadcast onto v -- which differs from the OCP's g5 on purpose. */
void ur10_tracking_errors(const ur10_params *P, const double pose[12], const double twist[6],
                          double px_ref, double vy_ref, double e[5], double *p_task_z)
{
    const double *pe = pose, *R = pose + 3;
    double tw[3], pt[3];
    mat3_vec(R, P->t_ee, tw);
    for (int k = 0; k < 3; k++) pt[k] = pe[k] + tw[k];
    double s = twist[3] * tw[0] + twist[4] * tw[1] + twist[5] * tw[2];
    double vv[3] = {twist[0] + s, twist[1] + s, twist[2] + s}, vt[3];
    mat3_vec(R, vv, vt);
    const double sa = P->surf[0], sb = P->surf[1], sc = P->surf[2], sd = P->surf[3], se = P->surf[4],
                 sf = P->surf[5];
    double X = pt[0], Y = pt[1];
    double S = sa * X * X + sb * Y * Y + sc * X * Y + sd * X + se * Y + sf;
    double Sx = 2 * sa * X + sc * Y + sd, Sy = 2 * sb * Y + sc * X + se;
    double rho = sqrt(Sx * Sx + Sy * Sy + 1.0);
    double n[3] = {Sx / rho, Sy / rho, -1.0 / rho};
    double g1 = S - pt[2];
    double g2 = n[0] * R[2] + n[1] * R[5] + n[2] * R[8];
    double g3 = R[1];
    double g4 = pt[0];
    double g5 = vt[1];
    e[0] = g1;
    e[1] = 1.0 - g2;
    e[2] = g3;
    e[3] = px_ref - g4;
    e[4] = vy_ref - g5;
    if (p_task_z) *p_task_z = pt[2];
}
