/* tests/simt_emu/simt_emu.cpp -- TEST HARNESS ONLY (see simt_emu.h). */
#include "simt_emu.h"

namespace simt_emu
{
thread_local Warp *g_warp = nullptr;

/* minimal x86-64 SysV context switch: callee-saved registers + stack pointer */
asm(R"(
.text
.globl simt_emu_switch
.type simt_emu_switch,@function
simt_emu_switch:
    pushq %rbp
    pushq %rbx
    pushq %r12
    pushq %r13
    pushq %r14
    pushq %r15
    movq %rsp, (%rdi)
    movq %rsi, %rsp
    popq %r15
    popq %r14
    popq %r13
    popq %r12
    popq %rbx
    popq %rbp
    ret
.size simt_emu_switch,.-simt_emu_switch
)");

static void trampoline()
{
    Warp *w = g_warp;
    int me = w->cur;
    w->entry(w->arg, me);
    w->done[me] = true;
    simt_emu_switch(&w->lane_sp[me], w->main_sp);
    abort(); /* never resumed */
}

void run_warp(void (*fn)(void *, int), void *arg, uint64_t seed, int nlanes)
{
    Warp w;
    memset(&w, 0, sizeof(w));
    w.entry = fn;
    w.arg = arg;
    w.rng = seed * 6364136223846793005ull + 1442695040888963407ull;
    w.stacks = (char *) aligned_alloc(64, kStack * kLanes);
    for (int l = 0; l < kLanes; l++)
    {
        /* initial frame: 6 callee-saved slots + return address = trampoline; rsp = 8 mod 16 on entry */
        uintptr_t top = (uintptr_t) (w.stacks + kStack * (l + 1));
        top &= ~(uintptr_t) 15;
        void **sp = (void **) top;
        *(--sp) = nullptr;               /* fake return address of trampoline (alignment) */
        *(--sp) = (void *) &trampoline;  /* 'ret' target */
        for (int k = 0; k < 6; k++) *(--sp) = nullptr;
        w.lane_sp[l] = sp;
    }
    for (int l = nlanes; l < kLanes; l++) w.done[l] = true; /* lanes beyond the group never run */
    Warp *prev = g_warp;
    g_warp = &w;
    int order[kLanes];
    for (;;)
    {
        int n = 0;
        for (int l = 0; l < kLanes; l++)
            if (!w.done[l]) order[n++] = l;
        if (n == 0) break;
        if (seed != 0) /* seed 0 = plain ascending order */
            for (int i = n - 1; i > 0; i--)
            {
                w.rng = w.rng * 6364136223846793005ull + 1442695040888963407ull;
                int j = (int) ((w.rng >> 33) % (uint64_t) (i + 1));
                int t = order[i]; order[i] = order[j]; order[j] = t;
            }
        for (int i = 0; i < n; i++)
        {
            w.cur = order[i];
            simt_emu_switch(&w.main_sp, w.lane_sp[order[i]]);
        }
    }
    g_warp = prev;
    free(w.stacks);
}
}  // namespace simt_emu
