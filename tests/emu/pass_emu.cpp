/*
 * pass_emu.cpp — TEST INFRASTRUCTURE: whole frames of the product's pass kernels (raytracer3_b200/csrc/pass_kernels.cuh: k_gbuffer,
 * k_ref_init / k_ref_shade / k_ref_shadow / k_ref_finish, k_postprocess) and of its traversal kernel (trace_kernel.cuh) on the CPU warp
 * emulation of warp_emu.h.  The launch sequences below restate rt3_launch_gbuffer / rt3_launch_reference / rt3_launch_postprocess
 * (passes.cu) and launch_dyn (trace.cu) with host memory in place of device memory; tests/test_emu_passes.py compares the frames bit for
 * bit with the oracle.  Not a fallback: librt3.so neither contains nor calls this code.
 */
#include "warp_emu.h"

#include <algorithm>
#include <functional>

#include "../../raytracer3_b200/csrc/pass_types.h"

using namespace rt3b;
using namespace rt3m;
#include "../../raytracer3_b200/csrc/trace_kernel.cuh"
#include "../../raytracer3_b200/csrc/pass_kernels.cuh"

namespace {

std::function<void()> g_call;
void call_body() { g_call(); }

/* kernel<<<grid, block>>>(...) */
int launch(unsigned grid, unsigned block, std::function<void()> f) {
    g_call = std::move(f);
    return wemu::run_grid(call_body, grid, block);
}
unsigned div_up(size_t a, size_t b) { return (unsigned)((a + b - 1) / b); }

const int SM_COUNT = 148;
int g_shade64 = 0; uint32_t g_n_prims = 0;   /* pemu_set_shade64: also build and use the 64-byte shading records */
int g_strips = 1;                         /* pemu_set_strips: how many strips rt3_launch_reference would use */
uint64_t* g_round_lengths = nullptr;      /* pemu_set_round_lengths: queue length at the start of every round (regeneration statistics) */

/* launch_dyn (trace.cu): persistent warps over a queue whose length lives in memory */
int trace_closest(const WideNode* nodes, const Tri48* tris, const rt3_ray* rays, const uint32_t* d_count, uint64_t n_max, rt3_hit* hits) {
    if (n_max == 0) return 0;
    unsigned long long counter = 0;
    const unsigned grid = (unsigned)std::min<uint64_t>(4, (n_max + TRACE_THREADS - 1) / TRACE_THREADS);      /* a few resident blocks are enough here */
    return launch(grid, TRACE_THREADS, [&]() { k_trace_dyn<false, 28, false>(nodes, tris, rays, d_count, n_max, hits, nullptr, &counter, 0x47000000u, nullptr); });
}
int trace_any(const WideNode* nodes, const Tri48* tris, const rt3_ray* rays, const uint32_t* d_count, uint64_t n_max, uint32_t* bits) {
    if (n_max == 0) return 0;
    unsigned long long counter = 0;
    memset(bits, 0, ((n_max + 31) / 32) * 4);
    const unsigned grid = (unsigned)std::min<uint64_t>(4, (n_max + TRACE_THREADS - 1) / TRACE_THREADS);
    return launch(grid, TRACE_THREADS, [&]() { k_trace_dyn<true, 28, false>(nodes, tris, rays, d_count, n_max, nullptr, bits, &counter, 0x47000000u, nullptr); });
}

PixelMap full_frame(const rt3_params& p) {
    const uint32_t W = (uint32_t)p.window_size[0], H = (uint32_t)p.window_size[1];
    PixelMap m;
    memset(&m, 0, sizeof(m));
    m.single.x0 = 0; m.single.y0 = 0; m.single.w = W; m.single.h = H;
    m.rects = nullptr; m.n_tiles = 1; m.tile_w = W; m.tile_h = H;
    m.bw = (W + 7) / 8; m.bh = (H + 3) / 4; m.pitch = W; m.slab = 0;
    m.work = m.bw * m.bh * 32u;
    finish_pixel_map(m);
    return m;
}

}  // namespace

extern "C" {

/* fastdiv (pass_types.h) against the integer division on random and edge operands; returns the number of mismatches */
uint32_t pemu_check_fastdiv(uint32_t tries) {
    uint32_t bad = 0, rng = 2463534242u;
    auto next = [&]() { rng ^= rng << 13; rng ^= rng >> 17; rng ^= rng << 5; return rng; };
    const uint32_t edge_d[] = {1u, 2u, 3u, 5u, 7u, 8u, 240u, 255u, 256u, 257u, 65535u, 65536u, 0x7FFFFFFEu, 0x7FFFFFFFu};
    const uint32_t edge_n[] = {0u, 1u, 2u, 0x7FFFFFFEu, 0x7FFFFFFFu - 1u, 0x40000000u, 0x3FFFFFFFu};
    for (uint32_t d : edge_d) {
        const FastDiv f = make_fastdiv(d);
        for (uint32_t n : edge_n) if (fastdiv(n, f) != n / d) bad++;
        for (uint32_t q = 0; q < 64; ++q) {                                    /* around multiples of d */
            const unsigned long long base = (unsigned long long)(next() % (0x7FFFFFFFu / d + 1u)) * d;
            for (long long off = -2; off <= 2; ++off) { const long long n = (long long)base + off; if (n >= 0 && n < 0x7FFFFFFFll && fastdiv((uint32_t)n, f) != (uint32_t)n / d) bad++; }
        }
    }
    for (uint32_t i = 0; i < tries; ++i) {
        uint32_t d = next() >> (next() % 31u); if (d == 0) d = 1; if (d > 0x7FFFFFFFu) d >>= 1;
        const uint32_t n = next() & 0x7FFFFFFFu;
        if (n == 0x7FFFFFFFu) continue;
        if (fastdiv(n, make_fastdiv(d)) != n / d) bad++;
    }
    return bad;
}

/* SceneView with the per-triangle shading records the device build adds (k_shade_records, build.cu) */
void pemu_shade_records(const void* scene_view, uint32_t n_prims, uint32_t* rec4) {
    const SceneView& s = *(const SceneView*)scene_view;
    for (uint32_t i = 0; i < n_prims; ++i) {
        const uint32_t seg = find_segment(s, i);
        const rt3_geometry_info& g = s.geometries[s.seg_geometry[seg]];
        const uint32_t io = g.index_offset + 3u * (i - s.seg_first_prim[seg]);
        rec4[4 * i] = g.vertex_offset + s.indices[io]; rec4[4 * i + 1] = g.vertex_offset + s.indices[io + 1];
        rec4[4 * i + 2] = g.vertex_offset + s.indices[io + 2]; rec4[4 * i + 3] = seg;
    }
}

/* rt3_launch_gbuffer + rt3_launch_reference (+ rt3_launch_postprocess when post != NULL) for one full frame.
 * counters5: rays_primary, rays_secondary, rays_shadow, paths_started, paths_escaped.  Returns 0, or -1 (wemu::g_error). */
int pemu_frame(const void* scene_view, const uint32_t* shade_rec, const uint32_t* nodes20, const uint32_t* tris12, const rt3_params* pp,
               uint32_t* gbuffer, float* depth, uint32_t* ids, float* bary, const float* prev, float* light, float* post, uint64_t* counters5) {
    wemu::g_error = nullptr; wemu::g_collectives = 0;
    SceneView view = *(const SceneView*)scene_view;
    view.shade_rec = shade_rec;
    std::vector<float> shade64;
    if (g_shade64 && shade_rec) {                       /* k_shade_records64 (build.cu): normals x 3, texture coordinates x 3, segment */
        shade64.resize(16 * (size_t)g_n_prims);
        for (uint32_t i = 0; i < g_n_prims; ++i) {
            const rt3_vertex &a = view.vertices[shade_rec[4 * i]], &b = view.vertices[shade_rec[4 * i + 1]], &c = view.vertices[shade_rec[4 * i + 2]];
            float* o = shade64.data() + 16 * (size_t)i;
            o[0] = a.n[0]; o[1] = a.n[1]; o[2] = a.n[2]; o[3] = b.n[0]; o[4] = b.n[1]; o[5] = b.n[2]; o[6] = c.n[0]; o[7] = c.n[1]; o[8] = c.n[2];
            o[9] = a.t[0]; o[10] = a.t[1]; o[11] = b.t[0]; o[12] = b.t[1]; o[13] = c.t[0]; o[14] = c.t[1];
            memcpy(o + 15, &shade_rec[4 * i + 3], 4);
        }
        view.shade64 = shade64.data();
    }
    const rt3_params p = *pp;
    const WideNode* nodes = (const WideNode*)nodes20;
    const Tri48* tris = (const Tri48*)tris12;
    const PixelMap m = full_frame(p);
    DeviceCounters dc;
    memset(&dc, 0, sizeof(dc));
    if (!m.work) return 0;
    /* ---- rt3_launch_gbuffer */
    if (launch(div_up(m.work, 128), 128, [&]() { k_gbuffer(view, nodes, tris, p, m, gbuffer, depth, ids, bary, &dc, nullptr); })) return -1;
    /* ---- rt3_launch_reference: n_strips wavefronts (run one after the other here; on the device they run on streams of their own) */
    const bool nee = nee_enabled(p.flags), layered = (p.flags & RT3_FLAG_SPECULAR) != 0;
    std::vector<float4> px_radiance(m.work);
    std::vector<uint2> px_path(m.work);
    uint32_t S = g_strips < 1 ? 1u : (uint32_t)g_strips;
    for (uint32_t strip = 0; strip < S; ++strip) {
        const StripMap sm = make_strip(m.work, strip, S);
        const size_t W = sm.local_work;
        if (!W) continue;
        std::vector<rt3_ray> q_rays[2] = {std::vector<rt3_ray>(W), std::vector<rt3_ray>(W)}, srays(nee ? W : 1);
        std::vector<float4> q_thr[2] = {std::vector<float4>(W), std::vector<float4>(W)}, scontrib(nee ? W : 1);
        std::vector<rt3_hit> q_hits(W);
        std::vector<uint32_t> sbits((W + 31) / 32 + 1), q_regen(W);
        std::vector<float> q_pdf[2] = {std::vector<float>(nee ? W : 1), std::vector<float>(nee ? W : 1)};
        uint32_t counts[4] = {0, 0, 0, 0};
        NeeQueues nq = {nullptr, nullptr, nullptr, nullptr, nullptr};
        if (nee) { nq.srays = srays.data(); nq.scontrib = scontrib.data(); nq.scount = counts + 2; }
        const unsigned grid = std::min<unsigned>(div_up(W, PASS_THREADS), (unsigned)SM_COUNT * 8u);
        const RegenInputs gin = {gbuffer, depth};
        auto resolve_shadows = [&]() -> int {
            if (trace_any(nodes, tris, nq.srays, nq.scount, W, sbits.data())) return -1;
            return launch(grid, PASS_THREADS, [&]() { k_ref_shadow(nq.scontrib, sbits.data(), nq.scount, px_radiance.data()); });
        };
        nq.pdf_in = nullptr; nq.pdf_out = nee ? q_pdf[0].data() : nullptr;
        if (launch(grid, PASS_THREADS, [&]() {
                if (layered) k_ref_init<true>(view, p, m, sm, gbuffer, depth, px_radiance.data(), px_path.data(), q_rays[0].data(), q_thr[0].data(), counts, nq, &dc);
                else k_ref_init<false>(view, p, m, sm, gbuffer, depth, px_radiance.data(), px_path.data(), q_rays[0].data(), q_thr[0].data(), counts, nq, &dc);
            })) return -1;
        const uint64_t rounds = p.bounces > 1 ? (uint64_t)p.samples * (p.bounces - 1) : 0;
        if (nee && rounds && resolve_shadows()) return -1;
        int cur = 0;
        for (uint64_t r = 0; r < rounds; ++r) {
            if (g_round_lengths && r < 4096) g_round_lengths[r] += counts[cur];
            if (trace_closest(nodes, tris, q_rays[cur].data(), counts + cur, W, q_hits.data())) return -1;
            counts[1 - cur] = 0;
            counts[2] = 0; counts[3] = 0;
            nq.pdf_in = nee ? q_pdf[cur].data() : nullptr; nq.pdf_out = nee ? q_pdf[1 - cur].data() : nullptr;
            if (launch(grid, PASS_THREADS, [&]() {
#define PEMU_SHADE(L, R) k_ref_shade<L, R>(view, p, m, q_regen.data(), counts + 3, q_rays[cur].data(), q_hits.data(), q_thr[cur].data(), counts + cur, q_rays[1 - cur].data(), \
                                           q_thr[1 - cur].data(), counts + (1 - cur), px_radiance.data(), px_path.data(), nq, &dc)
                    if (layered) { if (p.samples > 1) PEMU_SHADE(true, true); else PEMU_SHADE(true, false); }
                    else { if (p.samples > 1) PEMU_SHADE(false, true); else PEMU_SHADE(false, false); }
#undef PEMU_SHADE
                })) return -1;
            if (p.samples > 1 && launch(grid, PASS_THREADS, [&]() {                  /* regeneration: the next samples of the pixels whose path ended */
                    if (layered) k_ref_regen<true>(view, p, m, gin, q_regen.data(), counts + 3, q_rays[1 - cur].data(), q_thr[1 - cur].data(), counts + (1 - cur),
                                                   px_radiance.data(), px_path.data(), nq, &dc);
                    else k_ref_regen<false>(view, p, m, gin, q_regen.data(), counts + 3, q_rays[1 - cur].data(), q_thr[1 - cur].data(), counts + (1 - cur),
                                            px_radiance.data(), px_path.data(), nq, &dc);
                })) return -1;
            if (nee && r + 1 != rounds && resolve_shadows()) return -1;
            cur = 1 - cur;
        }
        if (launch(div_up(W, PASS_THREADS), PASS_THREADS, [&]() { k_ref_finish(p, m, sm, px_radiance.data(), prev, light); })) return -1;
    }
    const size_t W = m.work;
    /* ---- rt3_launch_postprocess */
    if (post && launch(div_up(W, PASS_THREADS), PASS_THREADS, [&]() { k_postprocess(view, p, m, depth, light, post); })) return -1;
    if (counters5) { counters5[0] = dc.rays_primary; counters5[1] = dc.rays_secondary; counters5[2] = dc.rays_shadow; counters5[3] = dc.paths_started; counters5[4] = dc.paths_escaped; }
    return 0;
}
void pemu_set_strips(int n) { g_strips = n; }
void pemu_set_shade64(int on, uint32_t n_prims) { g_shade64 = on; g_n_prims = n_prims; }
void pemu_set_round_lengths(uint64_t* out4096) { g_round_lengths = out4096; }
const char* pemu_last_error() { return wemu::g_error ? wemu::g_error : ""; }

}
