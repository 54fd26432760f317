/* tests/simt_emu/simt_emu.h -- TEST HARNESS ONLY: a lock-step SIMT emulator for ONE warp on the host.
 *
 * There is no GPU on the build machine, so the warp-cooperative device code in
 * robotic-mpc_b200/csrc/rmpc_device.cuh is ALSO compiled by g++ against this header (RMPC_EMU defined) and run
 * by tests/ to check its indexing, synchronisation and arithmetic against the oracle before any GPU time is spent.
 * It is NOT a CPU fallback: the product library (librmpc.so) contains no host implementation of the solver and
 * fails loudly without a CUDA device; nothing outside tests/ links this file.
 *
 * Model: a group of up to 32 lanes = that many fibers of one OS thread.  A lane runs until it reaches a warp-wide primitive
 * (shuffle / syncwarp), then yields to the scheduler, which resumes the lanes of each epoch in a pseudo-random
 * order (seeded) -- so code that forgets a __syncwarp between a write by one lane and a read by another fails
 * here with high probability instead of passing by luck.
 */
#pragma once
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>

namespace simt_emu
{
constexpr int kLanes = 32;
constexpr size_t kStack = 1 << 20;

extern "C" void simt_emu_switch(void **save_sp, void *load_sp);

struct Warp
{
    void *lane_sp[kLanes];
    void *main_sp;
    char *stacks;
    bool done[kLanes];
    int cur;
    uint64_t rng;
    double shfl_d[2][kLanes];
    int shfl_parity[kLanes];
    unsigned ballot_bits[2];
    void (*entry)(void *, int);
    void *arg;
};
extern thread_local Warp *g_warp;

inline int lane_id() { return g_warp->cur; }
inline void yield_to_scheduler()
{
    Warp *w = g_warp;
    simt_emu_switch(&w->lane_sp[w->cur], w->main_sp);
}
inline void syncwarp() { yield_to_scheduler(); }
inline double shfl(double v, int src)
{
    Warp *w = g_warp;
    int me = w->cur, p = w->shfl_parity[me];
    w->shfl_parity[me] = 1 - p;
    w->shfl_d[p][me] = v;
    yield_to_scheduler();
    return w->shfl_d[p][src & 31];
}
inline double shfl_xor(double v, int m) { return shfl(v, lane_id() ^ m); }
inline int shfl_i(int v, int src) { return (int) shfl((double) v, src); }

/* run fn(arg, lane) on nlanes (<= 32) lock-step lanes */
void run_warp(void (*fn)(void *, int), void *arg, uint64_t seed, int nlanes = kLanes);
}  // namespace simt_emu
