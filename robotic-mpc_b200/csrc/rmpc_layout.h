/* robotic-mpc_b200/csrc/rmpc_layout.h -- host-side constants shared by the CUDA allocator in rmpc_capi.cu and by the
 * host-side test harness (the per-instance array list itself, RMPC_WS_ARRAYS, is in rmpc_types.h). */
#ifndef RMPC_LAYOUT_H
#define RMPC_LAYOUT_H
#include "rmpc_types.h"

/* Fixed transforms of the UR10 chain: URDF joint origins (R/ur_description/urdf/ur10.urdf:50-55,78-83,106-111,
 * 134-139,162-167,190-195) and the fixed ee_fixed_joint -> frame ee_link (:218-222).  rpy -> rotation goes through a
 * normalised quaternion, as urdfdom/pinocchio's URDF parser does (R/loader.py:24-28 builds the model from it). */
#ifdef __cplusplus
#include <cmath>
static inline void rmpc_chain_init(RmpcChain *ch)
{
    static const double xyz[7][3] = {{0.0, 0.0, 0.1273}, {0.0, 0.220941, 0.0}, {0.0, -0.1719, 0.612}, {0.0, 0.0, 0.5723},
                                     {0.0, 0.1149, 0.0}, {0.0, 0.0, 0.1157},  {0.0, 0.0922, 0.0}};
    static const double rpy[7][3] = {{0, 0, 0}, {0.0, 1.570796325, 0.0}, {0, 0, 0}, {0.0, 1.570796325, 0.0},
                                     {0, 0, 0}, {0, 0, 0}, {-1.5707963267948966, 0.0, 0.0}};
    for (int j = 0; j < 7; j++)
    {
        const double phi = rpy[j][0] / 2.0, the = rpy[j][1] / 2.0, psi = rpy[j][2] / 2.0;
        double x = std::sin(phi) * std::cos(the) * std::cos(psi) - std::cos(phi) * std::sin(the) * std::sin(psi);
        double y = std::cos(phi) * std::sin(the) * std::cos(psi) + std::sin(phi) * std::cos(the) * std::sin(psi);
        double z = std::cos(phi) * std::cos(the) * std::sin(psi) - std::sin(phi) * std::sin(the) * std::cos(psi);
        double w = std::cos(phi) * std::cos(the) * std::cos(psi) + std::sin(phi) * std::sin(the) * std::sin(psi);
        const double s = std::sqrt(x * x + y * y + z * z + w * w);
        x /= s; y /= s; z /= s; w /= s;
        const double tx = 2 * x, ty = 2 * y, tz = 2 * z;
        const double twx = tx * w, twy = ty * w, twz = tz * w, txx = tx * x, txy = ty * x, txz = tz * x;
        const double tyy = ty * y, tyz = tz * y, tzz = tz * z;
        double *R = ch->Rfix[j];
        R[0] = 1 - (tyy + tzz); R[1] = txy - twz;       R[2] = txz + twy;
        R[3] = txy + twz;       R[4] = 1 - (txx + tzz); R[5] = tyz - twx;
        R[6] = txz - twy;       R[7] = tyz + twx;       R[8] = 1 - (txx + tyy);
        for (int k = 0; k < 3; k++) ch->pfix[j][k] = xyz[j][k];
    }
}

/* Any serial 6R chain with joint axes along +x / +y / +z (UR5: R/ur_description/urdf/ur5.urdf:274-357, all about z; frame
 * tool0, R/prediction_model.py:28-29) expressed in the kernels' compile-time axis pattern z y y y z y.
 * xyzR[8][12] = {xyz(3), R row-major(9)} of: root (fixed joints in front of joint 1), the six joint origins, joint 6 -> EE.
 * With C_j e_pattern_j = e_axis_j (a signed permutation, det +1):  R_axis(q) = C_j R_pattern(q) C_j', so
 *   Rfix'_j = C_{j-1}' R_j C_j,  pfix'_j = C_{j-1}' t_j   (j = 0: the root transform in place of C'),  EE row = C_5' (R_ee, t_ee).
 * Returns 0, or -1 for an axis outside 0..2. */
static inline int rmpc_chain_from_table(RmpcChain *ch, const double *xyzR, const int *axis)
{
    static const int pattern[6] = {2, 1, 1, 1, 2, 1};
    double L[9], lt[3]; /* transform pending in front of the next row */
    for (int k = 0; k < 9; k++) L[k] = xyzR[3 + k];
    for (int k = 0; k < 3; k++) lt[k] = xyzR[k];
    for (int j = 0; j < 7; j++)
    {
        const double *row = xyzR + 12 * (j + 1), *R = row + 3;
        double C[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1};
        if (j < 6)
        {
            const int a = axis[j], p = pattern[j];
            if (a < 0 || a > 2) return -1;
            if (a != p)
            {
                for (int k = 0; k < 9; k++) C[k] = 0.0;
                C[3 * a + p] = 1.0;
                C[3 * p + a] = -1.0;
                C[3 * (3 - a - p) + (3 - a - p)] = 1.0;
            }
        }
        double LR[9];
        for (int r = 0; r < 3; r++)
        {
            ch->pfix[j][r] = lt[r] + L[3 * r] * row[0] + L[3 * r + 1] * row[1] + L[3 * r + 2] * row[2];
            for (int c = 0; c < 3; c++) LR[3 * r + c] = L[3 * r] * R[c] + L[3 * r + 1] * R[3 + c] + L[3 * r + 2] * R[6 + c];
        }
        for (int r = 0; r < 3; r++)
            for (int c = 0; c < 3; c++) ch->Rfix[j][3 * r + c] = LR[3 * r] * C[c] + LR[3 * r + 1] * C[3 + c] + LR[3 * r + 2] * C[6 + c];
        for (int r = 0; r < 3; r++)
        {
            lt[r] = 0.0;
            for (int c = 0; c < 3; c++) L[3 * r + c] = C[3 * c + r]; /* C' */
        }
    }
    return 0;
}
#endif
#endif
