/*
 * warp_emu.cpp — TEST INFRASTRUCTURE: the production traversal kernel (raytracer3_b200/csrc/trace_kernel.cuh) on the CPU warp emulation of
 * warp_emu.h, compared with brute force by tests/test_emu_warp.py; phase statistics for tools/emu_warp_stats.py.  Not a fallback:
 * librt3.so neither contains nor calls this code.
 */
#include "warp_emu.h"

using namespace rt3b;
#include "../../raytracer3_b200/csrc/trace_kernel.cuh"

/* Static assignment, as k_gbuffer (passes.cu) and k_trace (trace.cu) use traverse_wide: a warp takes 32 consecutive rays and keeps them
 * until the slowest lane is done.  Statistics only (phases 5 / 6 of traverse_wide). */
namespace {
void k_static(const WideNode* nodes, const Tri48* tris, const rt3_ray* rays, uint64_t n, rt3_hit* hits, unsigned long long* work_counter) {
    const uint32_t lane = threadIdx.x & 31u;
    for (;;) {
        unsigned long long base = 0;
        if (lane == 0) base = atomicAdd(work_counter, 32ull);
        base = __shfl_sync(0xFFFFFFFFu, base, 0);
        if (base >= n) break;
        const uint64_t i = base + lane;
        const bool valid = i < n;
        rt3_ray r;
        if (valid) r = rays[i];
        else { r.o[0] = r.o[1] = r.o[2] = 0.f; r.d[0] = r.d[1] = 0.f; r.d[2] = 1.f; r.tmin = 0.f; r.tmax = 0.f; }
        rt3_hit h; h.t = r.tmax; h.u = 0.f; h.v = 0.f; h.prim = RT3_MISS;
        traverse_wide<false>(nodes, tris, r, h, nullptr, valid, nullptr);
        if (valid) hits[i] = h;
    }
}
}  // namespace

namespace wemu {

struct LaunchArgs {
    const WideNode* nodes; const Tri48* tris; const rt3_ray* rays; const uint32_t* d_count; uint64_t n; rt3_hit* hits; uint32_t* bits;
    unsigned long long* work_counter; const InstRec* insts; int any, two;
};
static LaunchArgs g_args;

#ifndef RT3_EMU_KEEP
#define RT3_EMU_KEEP 28          /* the product's KEEP (launch_dyn in trace.cu); other values for tools/emu_warp_stats.py experiments */
#endif
static void trace_body() {
    const LaunchArgs& a = g_args;
    if (a.any == 2) k_static(a.nodes, a.tris, a.rays, a.n, a.hits, a.work_counter);
    else if (a.any) {
        if (a.two) k_trace_dyn<true, RT3_EMU_KEEP, true>(a.nodes, a.tris, a.rays, a.d_count, a.n, a.hits, a.bits, a.work_counter, 0x47000000u, a.insts);
        else k_trace_dyn<true, RT3_EMU_KEEP, false>(a.nodes, a.tris, a.rays, a.d_count, a.n, a.hits, a.bits, a.work_counter, 0x47000000u, a.insts);
    } else {
        if (a.two) k_trace_dyn<false, RT3_EMU_KEEP, true>(a.nodes, a.tris, a.rays, a.d_count, a.n, a.hits, a.bits, a.work_counter, 0x47000000u, a.insts);
        else k_trace_dyn<false, RT3_EMU_KEEP, false>(a.nodes, a.tris, a.rays, a.d_count, a.n, a.hits, a.bits, a.work_counter, 0x47000000u, a.insts);
    }
}

}  // namespace wemu

extern "C" {

/* any = 2: the static-assignment loop over traverse_wide (closest hit) instead of the production kernel.
 * Runs k_trace_dyn<ANY, 28, TWO> with `blocks` blocks of 128 threads over the ray queue.  d_count_mode != 0 passes the queue length
 * through device memory (as the bounce loop does) instead of by value.  Returns 0, or -1 (message via wemu_last_error). */
int wemu_trace(const uint32_t* nodes20, const uint32_t* tris12, const void* insts, const rt3_ray* rays, uint64_t n, int any, int blocks,
               int d_count_mode, rt3_hit* hits, uint32_t* bits, uint64_t* collectives) {
    using namespace wemu;
    g_error = nullptr; g_collectives = 0;
    memset(g_phase, 0, sizeof(g_phase));
    unsigned long long counter = 0;
    uint32_t count32 = (uint32_t)n;
    g_args.nodes = (const WideNode*)nodes20; g_args.tris = (const Tri48*)tris12; g_args.rays = rays;
    g_args.d_count = d_count_mode ? &count32 : nullptr; g_args.n = d_count_mode ? n + 1000 : n;   /* n_static is the capacity when d_count is given */
    g_args.hits = hits; g_args.bits = bits; g_args.work_counter = &counter; g_args.insts = (const InstRec*)insts;
    g_args.any = any; g_args.two = insts != nullptr;
    if (any == 1) memset(bits, 0, ((n + 31) / 32) * 4);        /* launch_dyn zeroes the result bits before the launch */
    if (run_grid(trace_body, (unsigned)blocks, TRACE_THREADS) != 0) return -1;
    if (collectives) *collectives = g_collectives;
    return 0;
}
/* out[id][3] for id = 0 node step, 1 record step (one per lane), 2 balanced record step: lanes holding triangles, 3 balanced: lanes testing, 4 refill,
 * 5 / 6 node / record step of traverse_wide (static assignment) */
void wemu_phase_stats(uint64_t* out24) { memcpy(out24, g_phase, sizeof(g_phase)); }
const char* wemu_last_error() { return wemu::g_error ? wemu::g_error : ""; }
int wemu_tri_balance() { return RT3_TRI_BALANCE; }

}
