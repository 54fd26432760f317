/* tests/simt_emu/emu_driver.cpp -- TEST HARNESS ONLY: runs the device code of robotic-mpc_b200/csrc/rmpc_device.cuh
 * on the host lane emulator (simt_emu.h), one emulated group of RMPC_G lanes per instance, so tests/ can check the kernels' logic
 * against the oracle without a GPU.  Not part of the product library. */
#define RMPC_EMU 1
#include <cstdlib>
#include <cstring>
#include <vector>

#include "rmpc_device.cuh"
#include "rmpc_layout.h"

namespace
{
struct Job
{
    const RmpcChain *ch;
    const RmpcDims *D;
    const RmpcWs *W;
    long inst;
    int nsteps;
    GrpSmem *sm;
};
void body(void *p, int)
{
    Job *j = (Job *) p;
    closed_loop_group(*j->ch, *j->D, *j->W, j->inst, *j->sm, j->nsteps);
}
int g_shift = 0;
int g_as_rti = 0, g_as_rti_iter = 1, g_cond_N = 0;
bool g_custom_chain = false;
RmpcChain g_chain;
void current_chain(RmpcChain *ch)
{
    if (g_custom_chain) *ch = g_chain;
    else rmpc_chain_init(ch);
}
}  // namespace

/* same table as rmpc_set_chain (include/rmpc.h), process-global here; null restores the UR10 */
extern "C" int rmpc_emu_set_chain(const double *xyzR, const int *axis)
{
    if (!xyzR || !axis) { g_custom_chain = false; return 0; }
    if (rmpc_chain_from_table(&g_chain, xyzR, axis) != 0) return -1;
    g_custom_chain = true;
    return 0;
}

extern "C" int rmpc_emu_set_shift(int on) { g_shift = on ? 1 : 0; return 0; }
extern "C" int rmpc_emu_set_cond_N(int n2) { g_cond_N = n2; return 0; }
/* level: -1 standard RTI, 0..3 = A..D (acados as_rti_level); the emulated closed loop then runs preparation + feedback per step */
extern "C" int rmpc_emu_set_as_rti(int level, int iter) { g_as_rti = level < 0 ? 0 : level + 1; g_as_rti_iter = iter; return 0; }

/* the per-sample plant record (RK_*, 28 doubles) and the task outputs y5 of one state, with the current chain */
extern "C" int rmpc_emu_plant_record(const double *par, const double *z, double *rec, double *y5)
{
    RmpcChain ch;
    current_chain(&ch);
    double e[5];
    ur10_plant_record<true>(ch, par, z, z + 6, e, rec);
    if (y5)
    {
        NoEmit ne;
        ur10_task<false>(ch, par, z, z + 6, y5, ne);
    }
    return 0;
}

extern "C" int rmpc_emu_closed_loop_term(const RmpcDims *D, const double *par, const double *pe, const double *z0, int nsteps,
                                         double *zlog, double *ulog, double *slog, double *summary, int *flags_out,
                                         unsigned long long seed);

extern "C" int rmpc_emu_closed_loop(const RmpcDims *D, const double *par, const double *z0, int nsteps, double *zlog,
                                    double *ulog, double *slog, double *summary, int *flags_out, unsigned long long seed)
{
    return rmpc_emu_closed_loop_term(D, par, nullptr, z0, nsteps, zlog, ulog, slog, summary, flags_out, seed);
}

/* pe: [B][RE_COUNT] terminal-set blocks or null */
extern "C" int rmpc_emu_closed_loop_term(const RmpcDims *D, const double *par, const double *pe, const double *z0, int nsteps,
                                         double *zlog, double *ulog, double *slog, double *summary, int *flags_out,
                                         unsigned long long seed)
{
    const long B = D->B, S = D->N + 1;
    RmpcChain ch;
    current_chain(&ch);
    RmpcWs W;
    memset(&W, 0, sizeof(W));
    std::vector<double *> owned;
    W.slab_stride = rmpc_slab_stride(D->N);
    W.Sp = rmpc_slab_sp(D->N);
    W.slab = (double *) calloc((size_t) B * (size_t) W.slab_stride, sizeof(double));
    W.z = (double *) calloc((size_t) B * NX, sizeof(double));
    owned.push_back(W.slab);
    owned.push_back(W.z);
    W.par = par;
    W.pe = pe;
    W.flags = (int *) calloc((size_t) B * 4, sizeof(int));
    W.zlog = zlog; W.ulog = ulog; W.slog = slog; W.summary = summary;
    W.nlog = nsteps; W.step0 = 0;
    W.shift = g_shift;
    W.cond_N = g_cond_N;
    W.as_rti = g_as_rti;
    W.as_rti_iter = g_as_rti_iter;
    if (g_as_rti)
    {
        W.asrti = (double *) calloc((size_t) B * (size_t) RMPC_ASRTI_STRIDE(W.Sp), sizeof(double));
        owned.push_back(W.asrti);
    }
    for (long i = 0; i < B; i++)
    {
        memcpy(W.z + i * NX, z0 + i * NX, NX * sizeof(double));
        if (zlog) memcpy(zlog + i * (nsteps + 1) * NX, z0 + i * NX, NX * sizeof(double));
        /* initial guess x_k = x0, u_k = 0 (T/c_templates_tera/acados_solver.in.c:2744-2777) */
        for (long k = 0; k < S; k++) memcpy(W.slab + i * W.slab_stride + rmpc_slab_off(D->N, RC_ux, NV, (int) k) + 6, z0 + i * NX, NX * sizeof(double));
        W.flags[i * 4 + 0] = 1;
        W.flags[i * 4 + 1] = 1;
    }
    GrpSmem sm;
    for (long i = 0; i < B; i++)
    {
        memset(&sm, 0, sizeof(sm));
        Job j{&ch, D, &W, i, nsteps, &sm};
        simt_emu::run_warp(body, &j, seed, GL);
    }
    if (flags_out) memcpy(flags_out, W.flags, (size_t) B * 4 * sizeof(int));
    for (double *p : owned) free(p);
    free(W.flags);
    return 0;
}
