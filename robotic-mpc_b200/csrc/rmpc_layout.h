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
#endif
#endif
