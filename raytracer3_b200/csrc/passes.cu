/*
 * passes.cu — the ray passes as CUDA kernels.
 *
 *   k_gbuffer      <- /root/reference/shaders/old/gbuffer.slang:8-20        (primary visibility, one launch)
 *   k_ref_init / k_ref_shade / k_ref_finish + k_trace
 *                  <- /root/reference/shaders/old/refrence_mode.slang:14-66 (spp x bounce loop)
 *   k_postprocess  <- /root/reference/shaders/old/postprocess.slang:90-112
 *
 * The reference runs one raygen thread per pixel that loops over samples and bounces around a
 * black-box TraceRay.  Here the loop is turned inside out (wavefront): all live paths advance one
 * bounce per kernel, terminated paths are compacted away with warp-aggregated atomics, and the
 * traversal kernel sees a dense queue.  Per-pixel state that must survive between kernels is exactly
 * what the raygen shader keeps in registers: the radiance sum, the RNG counter (random.slang:62: never
 * reset between samples), the throughput and the current ray.  Each pixel has at most one path in
 * flight, so the per-pixel sums are updated without atomics and in the reference's order, which keeps
 * the result bit-identical to the per-pixel loop of the oracle.
 */
#include "common.cuh"

using namespace rt3b;
using namespace rt3m;

#include "pass_kernels.cuh"

#include <cstdio>
#include <cstdlib>
#include <mutex>
static int env_int_passes(const char* name, int dflt) { const char* v = getenv(name); return v ? atoi(v) : dflt; }

void rt3_launch_gbuffer(rt3_context* c, const rt3_params& p, const PixelMap& m, uint32_t* gbuffer, float* depth, uint32_t* ids, float* bary,
                        cudaStream_t st) {
    if (!m.work) return;
    RT3_LAUNCH(c, k_gbuffer, div_up(m.work, 128), 128, 0, st, c->view, c->bvh_nodes.as<WideNode>(), c->bvh_tris.as<Tri48>(), p, m, gbuffer, depth, ids, bary,
               c->d_counters, c->two_level ? c->d_insts.as<InstRec>() : (const InstRec*)nullptr);
}

/* One wavefront (a strip of the frame, pass_types.h) on stream `st` with the lane's own queues and counters. */
static void run_reference_strip(rt3_context* c, RefLane& ln, const rt3_params& p, const PixelMap& m, const StripMap& sm, const uint32_t* gbuffer, const float* depth,
                                const float* prev, float* light, float4* px_radiance, uint2* px_path, cudaStream_t st) {
    const size_t W = sm.local_work;                      /* at most one path per pixel of the strip is in flight */
    const bool nee = nee_enabled(p.flags);
    const bool layered = (p.flags & RT3_FLAG_SPECULAR) != 0;
    rt3_ray* q_rays[2] = {ensure<rt3_ray>(c, ln.q_rays[0], W), ensure<rt3_ray>(c, ln.q_rays[1], W)};
    float4* q_thr[2] = {ensure<float4>(c, ln.q_thr[0], W), ensure<float4>(c, ln.q_thr[1], W)};
    rt3_hit* q_hits = ensure<rt3_hit>(c, ln.q_hits, W);
    const bool regen = p.samples > 1;
    uint32_t* q_regen = regen ? ensure<uint32_t>(c, ln.q_regen, W) : nullptr;
    unsigned long long* cnt64 = ensure<unsigned long long>(c, ln.counts, 64);
    uint32_t* counts = (uint32_t*)cnt64;                  /* [0],[1] path queues, [2] shadow queue, [3] regeneration list; cnt64[32..] the traversal kernel's work counter */
    float* q_pdf[2] = {nullptr, nullptr};
    NeeQueues nq = {nullptr, nullptr, nullptr, nullptr, nullptr};
    uint32_t* sbits = nullptr;
    if (nee) {
        q_pdf[0] = ensure<float>(c, ln.q_pdf[0], W); q_pdf[1] = ensure<float>(c, ln.q_pdf[1], W);
        nq.srays = ensure<rt3_ray>(c, ln.q_srays, W); nq.scontrib = ensure<float4>(c, ln.q_scontrib, W); nq.scount = counts + 2;
        sbits = ensure<uint32_t>(c, ln.q_sbits, (W + 31) / 32);
    }
    const unsigned grid = std::min<unsigned>(div_up(W, PASS_THREADS), (unsigned)c->sm_count * 8u);
    const RegenInputs gin = {gbuffer, depth};
    auto resolve_shadows = [&]() {                      /* any-hit pass over the shadow queue, then add the unoccluded contributions */
        rt3_launch_trace_any(c, nq.srays, nq.scount, W, sbits, cnt64 + 32, st);
        RT3_LAUNCH(c, k_ref_shadow, grid, PASS_THREADS, 0, st, nq.scontrib, sbits, nq.scount, px_radiance);
    };
    RT3_CUDA(c, cudaMemsetAsync(counts, 0, 12, st));
    nq.pdf_in = nullptr; nq.pdf_out = q_pdf[0];
    if (layered)
        RT3_LAUNCH(c, k_ref_init<true>, grid, PASS_THREADS, 0, st, c->view, p, m, sm, gbuffer, depth, px_radiance, px_path, q_rays[0], q_thr[0], counts, nq, c->d_counters);
    else
        RT3_LAUNCH(c, k_ref_init<false>, grid, PASS_THREADS, 0, st, c->view, p, m, sm, gbuffer, depth, px_radiance, px_path, q_rays[0], q_thr[0], counts, nq, c->d_counters);
    /* Rounds: every unfinished pixel advances by one traced ray per round, and a pixel traces at most bounces - 1 rays per sample.  The
     * queue lengths live on the device, so the host enqueues the worst case; rounds past the last live path see an empty queue. */
    const uint64_t rounds = p.bounces > 1 ? (uint64_t)p.samples * (p.bounces - 1) : 0;
    if (nee && rounds) resolve_shadows();
    int cur = 0;
    for (uint64_t r = 0; r < rounds; ++r) {
        rt3_launch_trace_closest(c, q_rays[cur], counts + cur, W, q_hits, cnt64 + 32, st);
        RT3_CUDA(c, cudaMemsetAsync(counts + (1 - cur), 0, 4, st));
        if (nee || regen) RT3_CUDA(c, cudaMemsetAsync(counts + 2, 0, 8, st));
        nq.pdf_in = q_pdf[cur]; nq.pdf_out = q_pdf[1 - cur];
#define RT3_SHADE(L, R) RT3_LAUNCH(c, (k_ref_shade<L, R>), grid, PASS_THREADS, 0, st, c->view, p, m, q_regen, counts + 3, q_rays[cur], q_hits, q_thr[cur], counts + cur, \
                                   q_rays[1 - cur], q_thr[1 - cur], counts + (1 - cur), px_radiance, px_path, nq, c->d_counters)
#define RT3_REGEN(L) RT3_LAUNCH(c, (k_ref_regen<L>), grid, PASS_THREADS, 0, st, c->view, p, m, gin, q_regen, counts + 3, q_rays[1 - cur], q_thr[1 - cur], counts + (1 - cur), \
                                px_radiance, px_path, nq, c->d_counters)
        if (layered) { if (regen) { RT3_SHADE(true, true); RT3_REGEN(true); } else RT3_SHADE(true, false); }
        else { if (regen) { RT3_SHADE(false, true); RT3_REGEN(false); } else RT3_SHADE(false, false); }
#undef RT3_SHADE
#undef RT3_REGEN
        if (nee && r + 1 != rounds) resolve_shadows();
        cur = 1 - cur;
    }
    RT3_LAUNCH(c, k_ref_finish, div_up(W, PASS_THREADS), PASS_THREADS, 0, st, p, m, sm, px_radiance, prev, light);
}

/* The reference pass: S strips (pass_types.h), each a wavefront of its own on its own stream, forked from and joined back into the caller's
 * stream with events.  Why: a persistent traversal launch ends with a tail in which the last long rays keep a shrinking number of warps
 * busy, and the small kernels between the launches (shade, memsets) leave the GPU half empty; with two or more strips in flight the tail
 * of one strip's launch is filled by the other strips' kernels.  The strips touch disjoint pixels, so the image does not depend on S. */
void rt3_launch_reference(rt3_context* c, const rt3_params& p, const PixelMap& m, const uint32_t* gbuffer, const float* depth, const float* prev,
                          float* light, cudaStream_t st) {
    if (!m.work) return;
    static const int env_strips = env_int_passes("RT3_STRIPS", 2);
    uint32_t S = (uint32_t)std::max(1, std::min(c->opt_strips > 0 ? c->opt_strips : env_strips, RT3_MAX_LANES));
    while (S > 1 && (m.work + STRIP_UNIT - 1) / STRIP_UNIT < 16u * S) --S;        /* small frames: one wavefront */
    float4* px_radiance = ensure<float4>(c, c->px_radiance, m.work);
    uint2* px_path = ensure<uint2>(c, c->px_path, m.work);
    if (S == 1) {
        run_reference_strip(c, c->lanes[0], p, m, make_strip(m.work, 0, 1), gbuffer, depth, prev, light, px_radiance, px_path, st);
        return;
    }
    for (uint32_t k = 0; k < S; ++k)                                               /* allocate before forking: ensure() may cudaFree (synchronising) */
        if (!c->lanes[k].stream) {
            RT3_CUDA(c, cudaStreamCreateWithFlags(&c->lanes[k].stream, cudaStreamNonBlocking));
            RT3_CUDA(c, cudaEventCreateWithFlags(&c->lanes[k].done, cudaEventDisableTiming));
        }
    RT3_CUDA(c, cudaEventRecord(c->ev_fork, st));
    for (uint32_t k = 0; k < S; ++k) {
        RefLane& ln = c->lanes[k];
        RT3_CUDA(c, cudaStreamWaitEvent(ln.stream, c->ev_fork, 0));
        run_reference_strip(c, ln, p, m, make_strip(m.work, k, S), gbuffer, depth, prev, light, px_radiance, px_path, ln.stream);
        RT3_CUDA(c, cudaEventRecord(ln.done, ln.stream));
    }
    for (uint32_t k = 0; k < S; ++k) RT3_CUDA(c, cudaStreamWaitEvent(st, c->lanes[k].done, 0));
}

void rt3_launch_postprocess(rt3_context* c, const rt3_params& p, const PixelMap& m, const float* depth, const float* in, float* out, cudaStream_t st) {
    if (!m.work) return;
    RT3_LAUNCH(c, k_postprocess, div_up(m.work, PASS_THREADS), PASS_THREADS, 0, st, c->view, p, m, depth, in, out);
}

void rt3_launch_present(rt3_context* c, const rt3_params& p, const PixelMap& m, const float* depth, const float* in, uint32_t* out, cudaStream_t st) {
    if (!m.work) return;
    RT3_LAUNCH(c, k_present_bgra8, div_up(m.work, PASS_THREADS), PASS_THREADS, 0, st, c->view, p, m, depth, in, out);
}

void rt3_launch_probe_trace(rt3_context* c, const rt3_params& p, const float* depth, const uint16_t* directions, const float* prev_atlas, float* atlas,
                            cudaStream_t st) {
    const uint32_t aw = probe_count_x(p) * 8u, ah = probe_count_y(p) * 8u;
    const size_t n = (size_t)aw * ah;
    if (!n) return;
    rt3_ray* rays = ensure<rt3_ray>(c, c->lanes[0].q_rays[0], std::max<size_t>(n, 1));
    rt3_hit* hits = ensure<rt3_hit>(c, c->lanes[0].q_hits, std::max<size_t>(n, 1));
    RT3_LAUNCH(c, k_probe_raygen, div_up(n, PASS_THREADS), PASS_THREADS, 0, st, p, aw, ah, depth, directions, rays);
    rt3_launch_trace_closest(c, rays, nullptr, n, hits, nullptr, st);
    float4* staged = nullptr; uint32_t* staged_target = nullptr;
    if (directions) {                                   /* several texels of a probe may aim at one target: staged, then resolved in the defined order */
        staged = ensure<float4>(c, c->lanes[0].q_thr[0], n);
        staged_target = ensure<uint32_t>(c, c->lanes[0].q_sbits, n);
    }
    RT3_LAUNCH(c, k_probe_texel, div_up(n, PASS_THREADS), PASS_THREADS, 0, st, c->view, p, aw, ah, depth, directions, rays, hits, (const float4*)prev_atlas,
               (float4*)atlas, staged, staged_target, c->d_counters);
    if (directions) RT3_LAUNCH(c, k_probe_resolve, div_up(n, PASS_THREADS), PASS_THREADS, 0, st, aw, ah, staged, staged_target, (float4*)atlas);
}
/* ---- rt3_pass_probe_trace_static: the probe-trace pass with NO allocation anywhere on its path ----------------------------------------
 * (no counterpart in the reference; added for callers that cannot call cudaMalloc: code under CUDA
 * graph capture, tight serving loops, harnesses that forbid it).  The pass needs scratch per atlas texel — the probe ray (32 B), its hit
 * (16 B) and, when a direction image is bound, the staged texel value (16 B) and its target (4 B) — plus the persistent traversal kernel's
 * ray-fetch counter.  Here that scratch is ONE file-scope device array sized for RT3_STATIC_PROBE_MAX_TEXELS texels (a 1920x1088 window:
 * 960x544 = 522,240): 68 B x 524,288 + 256 B = 34 MiB that every GPU which loads librt3.so holds until the process exits.
 * In host code `g_static_probe_scratch` is only a stand-in for the symbol, not a device address: its address per device comes from
 * cudaGetSymbolAddress, asked once per device with that device current (rt3_static_scratch_setup, called from rt3_create), and that
 * pointer is what the kernels get.  All calls on a device share the array: it keeps the last call's contents, every call overwrites what it
 * reads (rays, hits, staged values and targets of all n texels) and zeroes the ray-fetch counter itself; two calls that overlap on the GPU
 * (different streams or contexts without ordering) race. */
constexpr size_t STATIC_PROBE_N = RT3_STATIC_PROBE_MAX_TEXELS;
constexpr size_t STATIC_OFF_RAYS = 0;
constexpr size_t STATIC_OFF_HITS = STATIC_OFF_RAYS + STATIC_PROBE_N * sizeof(rt3_ray);
constexpr size_t STATIC_OFF_STAGED = STATIC_OFF_HITS + STATIC_PROBE_N * sizeof(rt3_hit);
constexpr size_t STATIC_OFF_TARGET = STATIC_OFF_STAGED + STATIC_PROBE_N * sizeof(float4);
constexpr size_t STATIC_OFF_WORK = STATIC_OFF_TARGET + STATIC_PROBE_N * sizeof(uint32_t);
constexpr size_t STATIC_BYTES = STATIC_OFF_WORK + 256;
static __device__ __align__(256) unsigned char g_static_probe_scratch[STATIC_BYTES];

constexpr int STATIC_MAX_DEVICES = 64;
static unsigned char* g_static_probe_ptr[STATIC_MAX_DEVICES];         /* device address of the array on each device, 0 = not asked yet */
static std::mutex g_static_probe_mutex;

/* Once per device, with that device current (the caller, rt3_create, has set it).  Returns cudaSuccess or the error of the query. */
cudaError_t rt3_static_scratch_setup(int device) {
    if (device < 0 || device >= STATIC_MAX_DEVICES) return cudaErrorInvalidDevice;
    std::lock_guard<std::mutex> lock(g_static_probe_mutex);
    if (g_static_probe_ptr[device]) return cudaSuccess;
    void* ptr = nullptr;
    cudaError_t e = cudaGetSymbolAddress(&ptr, g_static_probe_scratch);
    if (e != cudaSuccess) return e;
    g_static_probe_ptr[device] = (unsigned char*)ptr;
    return cudaSuccess;
}

void rt3_launch_probe_trace_static(rt3_context* c, const rt3_params& p, const float* depth, const uint16_t* directions, const float* prev_atlas, float* atlas,
                                   cudaStream_t st) {
    const uint32_t aw = probe_count_x(p) * 8u, ah = probe_count_y(p) * 8u;
    const size_t n = (size_t)aw * ah;
    if (n > STATIC_PROBE_N) {
        char msg[200];
        snprintf(msg, sizeof(msg), "probe atlas %u x %u = %zu texels exceeds the static scratch of RT3_STATIC_PROBE_MAX_TEXELS = %zu (use rt3_pass_probe_trace)",
                 aw, ah, n, STATIC_PROBE_N);
        RT3_FAIL(c, RT3_ERR_INVALID, msg);
    }
    unsigned char* base = (c->device >= 0 && c->device < STATIC_MAX_DEVICES) ? g_static_probe_ptr[c->device] : nullptr;
    if (!base) RT3_FAIL(c, RT3_ERR_STATE, "static scratch address unknown on this device (cudaGetSymbolAddress failed at rt3_create)");
    if (!n) return;
    rt3_ray* rays = (rt3_ray*)(base + STATIC_OFF_RAYS);
    rt3_hit* hits = (rt3_hit*)(base + STATIC_OFF_HITS);
    float4* staged = directions ? (float4*)(base + STATIC_OFF_STAGED) : nullptr;
    uint32_t* staged_target = directions ? (uint32_t*)(base + STATIC_OFF_TARGET) : nullptr;
    unsigned long long* work = (unsigned long long*)(base + STATIC_OFF_WORK);      /* zeroed by the traversal launcher (cudaMemsetAsync) */
    RT3_LAUNCH(c, k_probe_raygen, div_up(n, PASS_THREADS), PASS_THREADS, 0, st, p, aw, ah, depth, directions, rays);
    rt3_launch_trace_closest(c, rays, nullptr, n, hits, work, st);
    RT3_LAUNCH(c, k_probe_texel, div_up(n, PASS_THREADS), PASS_THREADS, 0, st, c->view, p, aw, ah, depth, directions, rays, hits, (const float4*)prev_atlas,
               (float4*)atlas, staged, staged_target, c->d_counters);
    if (directions) RT3_LAUNCH(c, k_probe_resolve, div_up(n, PASS_THREADS), PASS_THREADS, 0, st, aw, ah, staged, staged_target, (float4*)atlas);
}

void rt3_launch_probe_sis(rt3_context* c, const rt3_params& p, const uint32_t* gbuffer, uint16_t* directions, cudaStream_t st) {
    const uint32_t probes = probe_count_x(p) * probe_count_y(p);
    if (!probes) return;
    RT3_LAUNCH(c, k_probe_sis, probes, 64, 0, st, p, gbuffer, directions);
}
void rt3_launch_probe_sh(rt3_context* c, const rt3_params& p, const float* atlas, float* sh, cudaStream_t st) {
    const size_t n = (size_t)probe_count_x(p) * probe_count_y(p) * 3u;
    if (!n) return;
    RT3_LAUNCH(c, k_probe_sh, div_up(n, PASS_THREADS), PASS_THREADS, 0, st, p, atlas, sh);
}
void rt3_launch_probe_interpolate(rt3_context* c, const rt3_params& p, const uint32_t* gbuffer, const float* depth, const float* sh, float* light,
                                  cudaStream_t st) {
    const size_t n = (size_t)p.window_size[0] * (size_t)p.window_size[1];
    if (!n) return;
    uint32_t* failed = ensure<uint32_t>(c, c->px_failed, n);
    RT3_LAUNCH(c, k_probe_interpolate, div_up(n, PASS_THREADS), PASS_THREADS, 0, st, p, gbuffer, depth, sh, (float4*)light, failed);
    RT3_LAUNCH(c, k_probe_failed, div_up(n, PASS_THREADS), PASS_THREADS, 0, st, (uint32_t)n, failed, (float4*)light);
}

void rt3_launch_untile(rt3_context* c, const rt3_rect* d_rects, uint32_t n_tiles, uint32_t tile_w, uint32_t tile_h, const float* slab, float* frame,
                       uint32_t frame_w, uint32_t frame_h, cudaStream_t st) {
    size_t total = (size_t)n_tiles * tile_w * tile_h;
    if (!total) return;
    RT3_LAUNCH(c, k_untile, div_up(total, PASS_THREADS), PASS_THREADS, 0, st, d_rects, n_tiles, tile_w, tile_h, (const float4*)slab, (float4*)frame, frame_w,
               frame_h);
}

void rt3_launch_scatter_tiles(rt3_context* c, const rt3_rect* d_rects, const unsigned long long* d_dst, uint32_t n_tiles, uint32_t tile_w, uint32_t tile_h,
                              const float* slab, uint32_t dst_pitch, cudaStream_t st) {
    size_t total = (size_t)n_tiles * tile_w * tile_h;
    if (!total) return;
    RT3_LAUNCH(c, k_scatter_tiles, div_up(total, PASS_THREADS), PASS_THREADS, 0, st, d_rects, d_dst, n_tiles, tile_w, tile_h, (const float4*)slab, dst_pitch);
}
