/* robotic-mpc_b200/csrc/rmpc_types.h -- plain-data structures shared by the host side of the C-ABI and the kernels.
 *
 * Layout of a batch in HBM ("instance-major"): every array below is [B][...] with one contiguous chunk per
 * instance; a GROUP of RMPC_G lanes (half a warp by default) owns one instance for the whole closed loop, so inside
 * a chunk the fastest index is the one the lanes sweep (component within a stage block for the Riccati sweeps,
 * stage index for the stage-parallel passes).  Stage blocks use fixed strides:
 *     NV = 18 ([u(6); x(12)], stage N keeps the same slot layout with u unused)
 *     NX = 12,  NB = 24 ([lower(12); upper(12)] over [u(6); q(6)], stage 0 uses only the u slots in the QP)
 *     NJ = 36 (scaled task Jacobian: 5 rows x 6 joint-angle columns + the 6 joint-velocity entries of row 5)
 *     FS = 208 (Riccati factor block of one stage, RF_* offsets below)
 */
#ifndef RMPC_TYPES_H
#define RMPC_TYPES_H

#define RMPC_NX 12
#define RMPC_NU 6
#define RMPC_NV 18
#define RMPC_NB 24
#define RMPC_NJ 36
#define RMPC_FS 208

#define RMPC_SQP 0
#define RMPC_SQP_RTI 1

/* Riccati factor block of one stage (doubles).  With M = [[R S'],[S Q]] + diag(Gamma) + reg I + [B A]' P+ [B A]:
 *   Luu Luu' = Muu,  Lxu = Mxu Luu^-T,  l_u = Luu^-1 m_u,  P = Mxx - Lxu Lxu',  p = m_x - Lxu l_u            */
enum
{
    RF_INV = 0,   /* 6 : 1 / diag(Luu) (+inf where the pivot was not positive, as a division by 0 would give) */
    RF_LU = 6,    /* 6 : l_u */
    RF_LUU = 12,  /* 15: strict lower part of Luu, row-major packed (row r at r(r-1)/2) */
    RF_LXU = 28,  /* 72: Lxu, 12 x 6 row-major, rows in state order [q; qd] */
    RF_P = 100,   /* 78: P, packed lower, row r at r(r+1)/2 */
    RF_PV = 178,  /* 12: p */
    RF_PB = 190,  /* 12: P+ b  (P of the NEXT stage times this stage's res_b) */
    RF_END = 202
};

/* RP_* (parameter block), RT_* (per-stage table), RL_* (step record), RK_* (plant record), RS_* (summary), RE_* (terminal
 * set; in the per-stage solver arrays the terminal row uses slot 0 (lower side) and slot 12 (upper side) of stage N -- free
 * there, stage N has no box bounds -- and its gradient 2 P x_N, the row of DCt, is kept in Jc[N][0..11]), RQ_* (SQP
 * statistics): public, part of the C ABI */
#include "../../include/rmpc_fields.h"

typedef struct
{
    double Rfix[7][9]; /* fixed rotation of joint j's origin (j = 6: ee_fixed_joint), row-major */
    double pfix[7][3]; /* fixed translation */
} RmpcChain;

typedef struct
{
    int B, N;
    int solver_type;   /* RMPC_SQP | RMPC_SQP_RTI */
    int max_iter;      /* nlp_solver_max_iter */
    int qp_iter_max;   /* qp_solver_iter_max */
    double tol[4];     /* nlp tolerances stat eq ineq comp */
    double qp_tol;     /* all four HPIPM tolerances */
    double dt;         /* shooting interval = Ts = cost scaling */
    double plant_dt;   /* plant RK4 step (= dt in the reference) */
} RmpcDims;

/* Per-instance "slab": all solver data of ONE instance is contiguous in HBM,
 *     slab = [ head (RMPC_HEAD doubles) | array 0 | array 1 | ... ],   array k = [Sp stage slots][width_k doubles],
 * Sp = number of stage slots N + 1 rounded up to even (16-byte alignment of every array).  Array k starts at
 * RMPC_HEAD + RC_k * Sp with RC_k a COMPILE-TIME constant (the summed widths of the arrays before it), so the kernels
 * address everything from one base pointer and the register holding Sp -- no table of 40 pointers to keep alive.
 * Within an array the stage index is the slow one: lanes that sweep stages in parallel (stage-parallel passes) and
 * lanes that sweep the components of one stage (Riccati sweeps) both touch neighbouring 128-byte lines. */
#define RMPC_WS_ARRAYS(X)                                                                                             \
    X(ux, RMPC_NV) X(pi, RMPC_NX) X(lam, RMPC_NB) X(wpi, RMPC_NX) X(wlam, RMPC_NB)                                    \
    X(Jc, RMPC_NJ) X(rqz, RMPC_NV) X(b, RMPC_NX) X(d, RMPC_NB)                                                        \
    X(sux, RMPC_NV) X(spi, RMPC_NX) X(slam, RMPC_NB) X(st, RMPC_NB)                                                   \
    X(dux, RMPC_NV) X(dpi, RMPC_NX) X(dlam, RMPC_NB) X(dt_, RMPC_NB)                                                  \
    X(res_g, RMPC_NV) X(res_b, RMPC_NX) X(res_d, RMPC_NB) X(res_m, RMPC_NB) X(res_mc, RMPC_NB) X(pvec, RMPC_NX)       \
    X(F, RMPC_FS)                                                                                                     \
    X(ig, RMPC_NV) X(ib, RMPC_NX) X(id, RMPC_NB) X(im, RMPC_NB) X(iux, RMPC_NV) X(ipi, RMPC_NX) X(ilam, RMPC_NB)      \
    X(it, RMPC_NB) X(tux, RMPC_NV)

/* RC_<name> = doubles per stage slot of all arrays before <name>; RC_TOTAL = doubles per stage slot of a slab */
enum
{
#define X(name, w) RCW_##name = (w),
    RMPC_WS_ARRAYS(X)
#undef X
};
struct rmpc_rc_helper_
{
#define X(name, w) char name[w];
    RMPC_WS_ARRAYS(X)
#undef X
};
enum
{
#define X(name, w) RC_##name = (int) __builtin_offsetof(struct rmpc_rc_helper_, name),
    RMPC_WS_ARRAYS(X)
#undef X
    RC_TOTAL = (int) sizeof(struct rmpc_rc_helper_)
};

/* small per-instance arrays at the head of the slab */
enum
{
    H_LAMX0 = 0,    /* 24: multipliers of the stage-0 state equality [lower(12); upper(12)] (NLP iterate) */
    H_WLAMX0 = 24,  /* 24: their merit weights */
    H_DX0 = 48,     /* 12: x0 - x_nlp[0] */
    H_G0U = 60,     /* 8 : stage-0 control gradient before the x0 elimination */
    H_QLAMX0 = 68,  /* 24: multipliers of the eliminated equalities recovered from the QP */
    H_LINCOST = 92, /* 1 : cost at the linearisation point, carried from an RTI preparation phase to its feedback phase */
    H_X0ADV = 96,   /* 12: AS-RTI: predicted initial state of the advanced problem (lbx = ubx of stage 0 during a preparation) */
    H_B0 = 108,     /* 12: dynamics gap of stage 0 at the last linearisation, before the x0 elimination folds dx0 into it */
    H_XLIN0 = 120,  /* 12: x_0 of the iterate the last linearisation was made at (acados: constr_eval_no_bounds of stage 0) */
    RMPC_HEAD = 144
};

typedef struct
{
    const double *par; /* [B][RP_COUNT] */
    double *slab;      /* [B][slab_stride] per-instance slabs (layout above) */
    long slab_stride;  /* doubles per instance = RMPC_HEAD + RC_TOTAL * Sp, rounded up to a 128-byte line */
    int Sp;            /* stage slots per array: N + 1 rounded up to even */
    double *z;         /* [B][12] plant state */
    int *flags;        /* [B][4]: first_call, qp_warm2_pending, n_lq_needed, reserved */
    /* logs (may be null) */
    double *zlog;      /* [B][nlog+1][12] */
    double *ulog;      /* [B][nlog][6] */
    double *slog;      /* [B][nlog][RL_COUNT] */
    double *summary;   /* [B][RS_COUNT] */
    int nlog;          /* capacity of the logs in steps */
    int step0;         /* index of the first step of this launch */
    const double *pe;  /* [B][RE_COUNT] terminal-set blocks, null: no terminal constraint (the reference MPC as shipped) */
    int rti_phase;     /* SQP_RTI only: 0 preparation + feedback in one solve, 1 preparation, 2 feedback (acados rti_phase) */
    int shift;         /* closed loop: shift the primal iterate by one stage after every step (0: like the reference) */
    const double *stg; /* [B][N][RT_COUNT] per-stage references and bounds, null: every stage uses "par" */
    double *sqpstat;   /* [B][max_iter + 2][RQ_COUNT] per-iteration statistics of the last SQP solve, null: not kept */
    int as_rti;        /* SQP_RTI only: 0 standard RTI, else 1 + acados' as_rti_level (1 = A, 2 = B, 3 = C, 4 = D) */
    int as_rti_iter;   /* iterations of the levels B, C, D inside a preparation (acados as_rti_iter) */
    double *asrti;     /* [B][RMPC_ASRTI_STRIDE(Sp)] level A: backup of the iterate at the last linearisation point */
    int cond_N;        /* qp_solver_cond_N: horizon N2 of the partially condensed QP the interior point works on; 0 or N: none */
} RmpcWs;
/* backup of one iterate: ux [Sp][18] | pi [Sp][12] | lam [Sp][24] | lamx0 [24] */
#define RMPC_ASRTI_STRIDE(Sp) ((long) (Sp) * (RMPC_NV + RMPC_NX + RMPC_NB) + 24)

#ifdef __CUDACC__
#define RMPC_HD __host__ __device__
#else
#define RMPC_HD
#endif
RMPC_HD static inline int rmpc_slab_sp(int N) { return (N + 2) & ~1; }
RMPC_HD static inline long rmpc_slab_stride(int N) { return ((long) RMPC_HEAD + (long) RC_TOTAL * rmpc_slab_sp(N) + 15) / 16 * 16; }
/* offset inside a slab of element 0 of stage s of the array whose RC_ constant is rc and width is w */
RMPC_HD static inline long rmpc_slab_off(int N, int rc, int w, int s) { return (long) RMPC_HEAD + (long) rc * rmpc_slab_sp(N) + (long) w * s; }

#endif
