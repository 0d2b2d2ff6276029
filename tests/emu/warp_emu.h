/*
 * warp_emu.h — TEST INFRASTRUCTURE: runs the product's CUDA kernels (raytracer3_b200/csrc/trace_kernel.cuh, pass_kernels.cuh: the very source
 * nvcc compiles into librt3.so) on the CPU over a lock-step warp emulation, so the `-m "not gpu"` suite can check the kernel's control
 * flow — dynamic ray fetch, the two-phase loop, the stack protocol, the optional warp-balanced record phase — against a brute-force
 * loop on a machine without a GPU.  It is not a fallback: librt3.so neither contains nor calls this code.
 *
 * Model: one CUDA block = 4 warps x 32 lanes, every lane a fiber (ucontext) running the kernel function on its own stack.  A warp
 * collective (__ballot_sync, __shfl_sync, __shfl_up_sync, __syncwarp — always with the full mask in this kernel) is a rendezvous: the lane
 * deposits its operand and yields until all 32 lanes of its warp have arrived; a lane that returns from the kernel while its warp
 * waits at a collective is reported as a divergence error (on hardware that is a hang).  Fibers are resumed round-robin over all
 * lanes of all warps of the block, so warps interleave at collective granularity and share the global work counter like resident
 * warps do; blocks run one after the other.  __shared__ becomes a static array (one block at a time), atomics are plain operations
 * (one OS thread).  Arithmetic goes through the host branches of bvh_logic.h / rt3_shared_math.h, which the determinism contract
 * makes bit-identical to the device branches.
 *
 * This header is the emulation itself (CUDA surface + fibers + grid runner); warp_emu.cpp drives the traversal kernel with it, pass_emu.cpp
 * whole frames of the pass kernels.
 */
#ifndef RT3_WARP_EMU_H
#define RT3_WARP_EMU_H
#include <ucontext.h>

#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#define RT3_WARP_EMU 1
static inline uint32_t __ballot_sync(uint32_t mask, bool pred);
struct EmuDim { unsigned x; };
static EmuDim threadIdx;             /* set by the scheduler whenever a fiber is resumed */
static EmuDim blockIdx, blockDim, gridDim;
/* phase statistics: [id][0] = times a warp ran the phase with at least one lane, [id][1] = lanes that took part, [id][2] = times reached */
static uint64_t g_phase[8][3];
#define RT3_EMU_PHASE(id, pred) do { const uint32_t m_ = __ballot_sync(0xFFFFFFFFu, (pred)); if ((threadIdx.x & 31u) == 0u) { g_phase[id][0] += m_ != 0u; g_phase[id][1] += (uint64_t)__builtin_popcount(m_); g_phase[id][2]++; } } while (0)
#include "../../raytracer3_b200/csrc/bvh_logic.h"

/* ------------------------------------------------------------------ CUDA surface used by trace_kernel.cuh ---------------- */
struct float4 { float x, y, z, w; };
struct uint2 { uint32_t x, y; };
static inline float4 make_float4(float x, float y, float z, float w) { float4 r = {x, y, z, w}; return r; }
static inline uint2 make_uint2(uint32_t x, uint32_t y) { uint2 r = {x, y}; return r; }
#define __global__
#define __device__
#define __host__
#define __forceinline__ inline
#define __launch_bounds__(...)
#define __shared__ static
static inline float4 __ldcs(const float4* p) { return *p; }
static inline void __stcs(float4* p, float4 v) { *p = v; }
static inline float __uint_as_float(uint32_t u) { float f; memcpy(&f, &u, 4); return f; }
static inline uint32_t __float_as_uint(float f) { uint32_t u; memcpy(&u, &f, 4); return u; }
static inline int __popc(uint32_t x) { return __builtin_popcount(x); }
static inline int __ffs(uint32_t x) { return __builtin_ffs((int)x); }
static inline unsigned long long atomicAdd(unsigned long long* p, unsigned long long v) { unsigned long long o = *p; *p += v; return o; }
static inline uint32_t atomicOr(uint32_t* p, uint32_t v) { uint32_t o = *p; *p |= v; return o; }
static inline uint32_t atomicAdd(uint32_t* p, uint32_t v) { uint32_t o = *p; *p += v; return o; }
struct uint4 { uint32_t x, y, z, w; };
static inline uint4 make_uint4(uint32_t x, uint32_t y, uint32_t z, uint32_t w) { uint4 r = {x, y, z, w}; return r; }

namespace wemu {

constexpr int LANES = 32;
constexpr size_t FIBER_STACK = 256 * 1024;

struct Warp {
    uint64_t slot[2][LANES];
    uint64_t slot_gen[2] = {~0ull, ~0ull};
    int kind[2] = {0, 0};
    int arrived[2] = {0, 0};
    uint64_t lane_gen[LANES];
    int live = LANES;                 /* lanes that have not returned from the kernel */
    int waiting = 0;                  /* lanes parked at a collective                 */
};
struct Fiber {
    ucontext_t ctx;
    std::vector<char> stack;
    Warp* warp = nullptr;
    int lane = 0, tid = 0;
    bool done = false;
};

static ucontext_t g_sched;
static Fiber* g_cur = nullptr;
static const char* g_error = nullptr;
static uint64_t g_collectives = 0;

static void yield() { swapcontext(&g_cur->ctx, &g_sched); }

/* every lane of the warp contributes v; returns after all have, with all[] = the 32 contributions */
enum { K_BALLOT = 1, K_SHFL32, K_SHFL64, K_SHFL_UP, K_SYNCWARP };
static void gather(int kind, uint64_t v, uint64_t all[LANES]) {
    Fiber* f = g_cur;
    Warp* w = f->warp;
    const uint64_t g = w->lane_gen[f->lane]++;
    const int b = (int)(g & 1u);
    if (w->slot_gen[b] != g) { w->slot_gen[b] = g; w->arrived[b] = 0; w->kind[b] = kind; }     /* first lane of this generation: the buffer was last used by g - 2 */
    else if (w->kind[b] != kind) g_error = "lanes of one warp met at different collectives (divergent code around a *_sync)";
    w->slot[b][f->lane] = v;
    w->arrived[b]++;
    w->waiting++;
    while (!(w->slot_gen[b] == g && w->arrived[b] == LANES)) {
        if (w->live < LANES) { g_error = "a lane left the kernel while its warp waits at a collective"; }
        yield();
    }
    w->waiting--;
    memcpy(all, w->slot[b], sizeof(uint64_t) * LANES);
    if (f->lane == 0) g_collectives++;
}

}  // namespace wemu

static inline uint32_t __ballot_sync(uint32_t mask, bool pred) {
    (void)mask;
    uint64_t all[wemu::LANES];
    wemu::gather(wemu::K_BALLOT, pred ? 1u : 0u, all);
    uint32_t r = 0;
    for (int i = 0; i < wemu::LANES; ++i) if (all[i]) r |= 1u << i;
    return r;
}
static inline void __syncwarp() { uint64_t all[wemu::LANES]; wemu::gather(wemu::K_SYNCWARP, 0, all); }
template <class T>
static inline T emu_shfl(T v, int src, int kind = 0) {
    static_assert(sizeof(T) <= 8, "shuffle operand");
    uint64_t bits = 0, all[wemu::LANES];
    memcpy(&bits, &v, sizeof(T));
    wemu::gather(kind ? kind : (sizeof(T) == 8 ? wemu::K_SHFL64 : wemu::K_SHFL32), bits, all);
    T r;
    memcpy(&r, &all[src & 31], sizeof(T));
    return r;
}
static inline unsigned long long __shfl_sync(uint32_t, unsigned long long v, int src) { return emu_shfl(v, src); }
static inline uint32_t __shfl_sync(uint32_t, uint32_t v, int src) { return emu_shfl(v, src); }
static inline float __shfl_sync(uint32_t, float v, int src) { return emu_shfl(v, src); }
static inline unsigned long long __shfl_xor_sync(uint32_t, unsigned long long v, int mask) { return emu_shfl(v, (int)((threadIdx.x & 31u) ^ (unsigned)mask)); }
static inline uint32_t __shfl_up_sync(uint32_t, uint32_t v, int delta) {
    const int lane = (int)(threadIdx.x & 31u);
    const uint32_t r = emu_shfl(v, lane >= delta ? lane - delta : lane, wemu::K_SHFL_UP);     /* lanes below delta keep their own value */
    return r;
}


/* ------------------------------------------------------------------ grid runner ------------------------------------------ */
namespace wemu {

static void (*g_body)() = nullptr;           /* the kernel call; reads its arguments from globals of the including file */

static void fiber_main() {
    g_body();
    g_cur->done = true;
    g_cur->warp->live--;
    swapcontext(&g_cur->ctx, &g_sched);
}

/* one block of `threads` threads (a multiple of 32 is not required: the last warp's missing lanes count as lanes that left at once —
 * which the kernels here never rely on, all block sizes are multiples of 32); returns 0, or -1 with g_error set */
static int run_block(int threads) {
    const int n_warps = (threads + LANES - 1) / LANES;
    std::vector<Warp> warps(n_warps);
    static std::vector<Fiber> fibers;
    if ((int)fibers.size() < threads) fibers.resize(threads);
    for (int t = 0; t < threads; ++t) {
        Fiber& f = fibers[t];
        f.warp = &warps[t / LANES]; f.lane = t % LANES; f.tid = t; f.done = false;
        f.warp->lane_gen[f.lane] = 0;
        if (f.stack.size() != FIBER_STACK) f.stack.resize(FIBER_STACK);
        getcontext(&f.ctx);
        f.ctx.uc_stack.ss_sp = f.stack.data();
        f.ctx.uc_stack.ss_size = f.stack.size();
        f.ctx.uc_link = &g_sched;
        makecontext(&f.ctx, fiber_main, 0);
    }
    int remaining = threads;
    uint64_t idle_rounds = 0;
    while (remaining && !g_error) {
        const uint64_t before = g_collectives;
        int finished_now = 0;
        for (int t = 0; t < threads && !g_error; ++t) {
            Fiber& f = fibers[t];
            if (f.done) continue;
            g_cur = &f;
            threadIdx.x = (unsigned)f.tid;
            swapcontext(&g_sched, &f.ctx);
            if (f.done) { remaining--; finished_now++; }
        }
        /* a full round in which no collective completed and no lane finished means the warps can make no progress */
        if (g_collectives == before && finished_now == 0) { if (++idle_rounds > 4) g_error = "deadlock: lanes of a warp wait at different collectives"; }
        else idle_rounds = 0;
    }
    return g_error ? -1 : 0;
}

/* kernel<<<grid, block>>>: blocks one after the other (no kernel here synchronises across blocks) */
static int run_grid(void (*body)(), unsigned grid, unsigned block) {
    g_body = body;
    gridDim.x = grid; blockDim.x = block;
    for (unsigned b = 0; b < grid; ++b) {
        blockIdx.x = b;
        if (run_block((int)block) != 0) return -1;
    }
    return 0;
}

}  // namespace wemu
#endif
