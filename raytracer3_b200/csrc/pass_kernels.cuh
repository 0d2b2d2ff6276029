/*
 * pass_kernels.cuh — the device kernels of the ray passes (G-buffer, reference-mode wavefront, post-process, probes, un-tiling), in a header
 * of their own so that two translation units see the very same source: passes.cu (nvcc, the product) and tests/emu/pass_emu.cpp, which
 * compiles it with g++ over the lock-step warp emulation of tests/emu/warp_emu.h and runs whole frames on the CPU against the oracle.
 * The emulation is test infrastructure; librt3.so contains the device build only.  What each kernel restates is listed in passes.cu.
 */
#ifndef RT3_PASS_KERNELS_CUH
#define RT3_PASS_KERNELS_CUH

namespace {

constexpr int PASS_THREADS = 256;

struct WarpCounters {
    unsigned long long a = 0, b = 0;
};
__device__ __forceinline__ void flush_counter(unsigned long long v, unsigned long long* dst) {
    for (int off = 16; off; off >>= 1) v += __shfl_xor_sync(0xFFFFFFFFu, v, off);
    if ((threadIdx.x & 31) == 0 && v) atomicAdd(dst, v);
}

/* ---- pass 1 -------------------------------------------------------------------------------- */
__global__ void __launch_bounds__(128, 8) k_gbuffer(SceneView s, const WideNode* __restrict__ nodes, const Tri48* __restrict__ tris, rt3_params p, PixelMap m,
                                                 uint32_t* __restrict__ gbuffer, float* __restrict__ depth, uint32_t* __restrict__ ids,
                                                 float* __restrict__ bary, DeviceCounters* counters, const InstRec* __restrict__ insts) {
    RT3_PERM_LUT_SETUP(lut_base, 128);                                           /* octant permutation table of the node test (bvh_logic.h) */
    const uint32_t idx = blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t x, y; size_t px;
    const bool valid = idx < m.work && pixel_of(m, idx, x, y, px);
    rt3_ray ray;
    if (valid) ray = setup_primary_ray(x, y, p);                            /* gbuffer.slang:10 */
    else { ray.o[0] = ray.o[1] = ray.o[2] = 0.f; ray.d[0] = ray.d[1] = 0.f; ray.d[2] = 1.f; ray.tmin = 0.f; ray.tmax = 0.f; }
    rt3_hit h; h.t = ray.tmax; h.u = 0.f; h.v = 0.f; h.prim = RT3_MISS;
    traverse_wide<false>(nodes, tris, ray, h, nullptr, valid, insts, lut_base);  /* gbuffer.slang:12 (warp-synchronous: all lanes call) */
    if (valid) {
        if (h.prim == RT3_MISS) {                                           /* gbuffer.slang:14-16 */
            depth[px] = RT3_BACKGROUND_DEPTH;
            if (ids) { ids[2 * px] = RT3_MISS; ids[2 * px + 1] = RT3_MISS; }
            if (bary) { bary[2 * px] = 0.f; bary[2 * px + 1] = 0.f; }
        } else {                                                            /* gbuffer.slang:17-19 */
            uint32_t geo;
            uint4_ g = gbuffer_pack(hit_info(s, h.prim, h.u, h.v, &geo));
            ((uint4*)gbuffer)[px] = make_uint4(g.x, g.y, g.z, g.w);
            depth[px] = h.t;
            if (ids) { ids[2 * px] = h.prim; ids[2 * px + 1] = geo; }
            if (bary) { bary[2 * px] = h.u; bary[2 * px + 1] = h.v; }
        }
    }
    flush_counter(valid ? 1ull : 0ull, &counters->rays_primary);
}

/* ---- pass 2, wavefront --------------------------------------------------------------------- */
/* RT3_FLAG_NEE plumbing: shadow-ray queue (ray, contribution + pixel slot) and the BRDF pdf carried with every live path. */
struct NeeQueues {
    rt3_ray* srays; float4* scontrib; uint32_t* scount;
    const float* pdf_in; float* pdf_out;
};
__device__ __forceinline__ uint32_t warp_enqueue(bool alive, uint32_t* count) {
    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t mask = __ballot_sync(0xFFFFFFFFu, alive);
    uint32_t base = 0;
    if (mask) {
        const int leader = __ffs(mask) - 1;
        if ((int)lane == leader) base = atomicAdd(count, (uint32_t)__popc(mask));
        base = __shfl_sync(0xFFFFFFFFu, base, leader);
    }
    return base + __popc(mask & ((1u << lane) - 1u));
}
__device__ __forceinline__ void store_ray(rt3_ray* q, uint32_t pos, vec3 o, vec3 d, float tmin, float tmax) {
    float4* p = (float4*)(q + pos);
    p[0] = make_float4(o.x, o.y, o.z, tmin);
    p[1] = make_float4(d.x, d.y, d.z, tmax);
}

/* ---- the reference-mode loop as ONE wavefront per frame call, with ray REGENERATION ---------------------------------------------------
 * refrence_mode.slang:28-57 is `for i < samples { for b < bounces { ... } }` per pixel.  Here all live paths advance one bounce per round
 * (k_trace_dyn, then k_ref_shade); a path that ends — miss, last bounce, invalid lobe — while its pixel has samples left is REPLACED before the
 * next round by the first bounce of the pixel's next sample (the surface comes from the G-buffer again, :23,:30-34): k_ref_shade compacts the
 * survivors into the next queue and the pixels to regenerate into a second list (one ballot + one atomic per warp each), k_ref_regen starts
 * their next samples with all lanes busy and appends them to the same queue.  The queue therefore stays as long as the number of unfinished pixels
 * instead of shrinking bounce by bounce and restarting per sample.  Every pixel still has exactly one path in flight and runs its samples in
 * order, so its RNG stream and the order of its radiance additions are those of the per-pixel loop: bit-identical to the oracle.
 * Per-pixel state: px_radiance (sum, w = 1 for non-background pixels) and px_path = {RNG counter (random.slang:62: never reset between
 * samples), sample index << 8 | loop index b of the ray in flight}. */
struct RegenInputs { const uint32_t* gbuffer; const float* depth; };

__device__ __forceinline__ bool all_finite(vec3 v) { return abs_(v.x) <= 3.4028235e38f && abs_(v.y) <= 3.4028235e38f && abs_(v.z) <= 3.4028235e38f; }
__device__ __forceinline__ bool is_zero(vec3 v) { return v.x == 0.0f && v.y == 0.0f && v.z == 0.0f; }

/* Loop index 0 of one sample (refrence_mode.slang:30-34 + :42-53 with the G-buffer surface).  Returns whether a ray is to be traced. */
template <bool LAYERED>
__device__ __forceinline__ bool start_sample(const SceneView& s, const rt3_params& p, bool nee, uint32_t x, uint32_t y, const GbufferData& surface, float gdepth,
                                             RandomSamplerState& rng, vec3& radiance, vec3& o, vec3& d, vec3& thr, float& pdf_brdf, bool& shadow, vec3& sdir,
                                             vec3& scon) {
    shadow = false;
    if (p.bounces == 0) return false;
    rt3_ray pr = setup_primary_ray(x, y, p);                           /* :30 */
    vec3 ray_color = v3(1.f, 1.f, 1.f);
    vec2 urand = sample_2d(s, p.flags, x, y, p.frame, rng);            /* :43 */
    LayeredMaterial lm;
    bool lobe_ok = true;
    const vec3 thr_before = ray_color;
    PathVertexOut nv = LAYERED ? bounce_step_layered(surface, v3(pr.o), v3(pr.d), gdepth, urand, radiance, ray_color, lm, lobe_ok, &pdf_brdf)
                               : bounce_step(surface, v3(pr.o), v3(pr.d), gdepth, urand, radiance, ray_color, &pdf_brdf);   /* :44-51 */
    o = nv.origin; d = nv.dir; thr = ray_color;
    const bool alive = lobe_ok && p.bounces > 1;                       /* :53 */
    if (nee && alive) {
        vec2 ul = sample_2d(s, p.flags, x, y, p.frame, rng);
        NeeSample ns = LAYERED ? nee_sample_layered(s, lm, thr_before, ul) : nee_sample(s, surface, ray_color, ul);
        shadow = ns.valid; sdir = ns.dir; scon = ns.contribution;
    }
    return alive;
}

/* refrence_mode.slang:18-34: every non-background pixel of the strip starts its first sample (and, while a sample ends at loop index 0,
 * the next one).  LAYERED = RT3_FLAG_SPECULAR: DiffuseBrdf + SpecularBrdf (brdf.slang:95-318) instead of DiffuseBrdf only. */
template <bool LAYERED>
__global__ void __launch_bounds__(PASS_THREADS) k_ref_init(SceneView s, rt3_params p, PixelMap m, StripMap sm, const uint32_t* __restrict__ gbuffer,
                                                           const float* __restrict__ depth, float4* __restrict__ px_radiance, uint2* __restrict__ px_path,
                                                           rt3_ray* __restrict__ q_rays, float4* __restrict__ q_thr, uint32_t* q_count,
                                                           NeeQueues nq, DeviceCounters* counters) {
    unsigned long long started = 0, shadows = 0;
    const bool nee = nee_enabled(p.flags);
    const uint32_t stride = gridDim.x * blockDim.x;
    for (uint32_t base = blockIdx.x * blockDim.x; base < sm.local_work; base += stride) {   /* block-uniform trip count */
        const uint32_t idx = strip_to_work(sm, base + threadIdx.x);
        uint32_t x, y; size_t px;
        bool alive = false, shadow = false;
        float pdf_brdf = 0.f;
        vec3 o = v3(0, 0, 0), d = v3(0, 0, 0), thr = v3(0, 0, 0), sdir = v3(0, 0, 0), scon = v3(0, 0, 0);
        if (idx < m.work && pixel_of(m, idx, x, y, px)) {
            const float gdepth = depth[px];
            if (gdepth != RT3_BACKGROUND_DEPTH) {                                     /* :18-21 */
                uint4 g4 = ((const uint4*)gbuffer)[px];
                uint4_ packed = {g4.x, g4.y, g4.z, g4.w};
                GbufferData surface = gbuffer_unpack(packed);                          /* :23 */
                RandomSamplerState rng = rng_init(x, y, p.frame);                      /* :25 */
                vec3 radiance = v3(0, 0, 0);
                uint32_t sample = 0;
                for (;;) {                                                             /* :28 */
                    started++;
                    alive = start_sample<LAYERED>(s, p, nee, x, y, surface, gdepth, rng, radiance, o, d, thr, pdf_brdf, shadow, sdir, scon);
                    if (alive || ++sample >= p.samples) break;
                }
                px_radiance[idx] = make_float4(radiance.x, radiance.y, radiance.z, 1.0f);   /* w = 1: pixel is not background */
                px_path[idx] = make_uint2(rng.index, (sample << 8) | 1u);
            } else {
                px_radiance[idx] = make_float4(0.f, 0.f, 0.f, 0.f);
            }
        } else if (idx < m.work) {
            px_radiance[idx] = make_float4(0.f, 0.f, 0.f, 0.f);
        }
        const uint32_t pos = warp_enqueue(alive, q_count);
        if (alive) {
            store_ray(q_rays, pos, o, d, 0.001f, RT3_BACKGROUND_DEPTH);               /* :31, TMax of setupPrimaryRay */
            q_thr[pos] = make_float4(thr.x, thr.y, thr.z, __uint_as_float(idx));
            if (nee) nq.pdf_out[pos] = pdf_brdf;
        }
        if (nee) {
            const uint32_t sp = warp_enqueue(shadow, nq.scount);
            if (shadow) {
                shadows++;
                store_ray(nq.srays, sp, o, sdir, 0.001f, RT3_BACKGROUND_DEPTH);
                nq.scontrib[sp] = make_float4(scon.x, scon.y, scon.z, __uint_as_float(idx));
            }
        }
    }
    flush_counter(started, &counters->paths_started);
    if (nee) flush_counter(shadows, &counters->rays_shadow);
}

/* One round: loop iteration b >= 1 of refrence_mode.slang:36-57 for every path whose ray was traced in this round, then regeneration. */
/* experiments only: -DRT3_SHADE_MIN_BLOCKS=5 / 6 trades registers for resident warps; measured on B200 (profiles/r2a_ab_experiments.log):
 * 2967 / 2907 against 3011 Mrays/s, so the default keeps 64 registers */
#ifdef RT3_SHADE_MIN_BLOCKS
#define RT3_SHADE_BOUNDS __launch_bounds__(PASS_THREADS, RT3_SHADE_MIN_BLOCKS)
#else
#define RT3_SHADE_BOUNDS __launch_bounds__(PASS_THREADS)
#endif
/* REGEN = the frame call has more than one sample per pixel: paths that end are handed to k_ref_regen through regen_idx / regen_count */
template <bool LAYERED, bool REGEN>
__global__ void RT3_SHADE_BOUNDS k_ref_shade(SceneView s, rt3_params p, PixelMap m, uint32_t* __restrict__ regen_idx, uint32_t* regen_count,
                                                            const rt3_ray* __restrict__ rays_in,
                                                            const rt3_hit* __restrict__ hits, const float4* __restrict__ thr_in,
                                                            const uint32_t* __restrict__ count_in, rt3_ray* __restrict__ rays_out,
                                                            float4* __restrict__ thr_out, uint32_t* count_out, float4* __restrict__ px_radiance,
                                                            uint2* __restrict__ px_path, NeeQueues nq, DeviceCounters* counters) {
    const uint32_t n = *count_in;
    unsigned long long traced = 0, escaped = 0, shadows = 0;
    const bool nee = nee_enabled(p.flags);
    const uint32_t stride = gridDim.x * blockDim.x;
    for (uint32_t base = blockIdx.x * blockDim.x; base < n; base += stride) {
        const uint32_t i = base + threadIdx.x;
        bool alive = false, shadow = false, regen = false;
        float pdf_brdf = 0.f;
        vec3 o = v3(0, 0, 0), d = v3(0, 0, 0), thr = v3(0, 0, 0), sdir = v3(0, 0, 0), scon = v3(0, 0, 0);
        uint32_t idx = 0;
        if (i < n) {
            traced++;
            const float4 r1 = ((const float4*)(rays_in + i))[1];
            const float4 h4 = ((const float4*)hits)[i];
            const float4 t4 = thr_in[i];
            idx = __float_as_uint(t4.w);
            const uint32_t prim = __float_as_uint(h4.w);
            vec3 ray_color = v3(t4.x, t4.y, t4.z);
            const vec3 rd = v3(r1.x, r1.y, r1.z);
            /* The pixel's radiance sum is read (and written back) only when this thread adds to it — x + 0 == x bit for bit, the sums are
             * never -0 — and its path word only when the path goes on or the pixel has samples left. */
            vec3 radiance = v3(0, 0, 0);
            bool have_radiance = false;
#define RT3_NEED_RADIANCE() do { if (!have_radiance) { const float4 rad4_ = px_radiance[idx]; radiance = v3(rad4_.x, rad4_.y, rad4_.z); have_radiance = true; } } while (0)
            uint32_t rng_index = 0, sample = 0, b_next = 0;
            if (prim == RT3_MISS) {                                                    /* :37-40 */
                escaped++;
                if (REGEN) { const uint2 path = px_path[idx]; rng_index = path.x; sample = path.y >> 8; }
                if (p.flags & RT3_FLAG_ENV) {
                    const float wgt = escape_weight(s, p.flags, rd, nee ? nq.pdf_in[i] : 0.f);       /* 1 unless RT3_FLAG_NEE */
                    RT3_NEED_RADIANCE();
                    radiance = radiance + ray_color * (sample_environment(s, rd) * wgt);   /* :38 (commented out in the reference) */
                }
            } else {
                const uint2 path = px_path[idx];
                const float4 r0 = ((const float4*)(rays_in + i))[0];
                const vec3 ro = v3(r0.x, r0.y, r0.z);
                uint32_t x, y; size_t px;
                pixel_of(m, idx, x, y, px);
                GbufferData surface = hit_info(s, prim, h4.y, h4.z);                   /* :55 */
                RandomSamplerState rng = rng_init(x, y, p.frame);
                rng.index = path.x;
                sample = path.y >> 8;
                vec2 urand = sample_2d(s, p.flags, x, y, p.frame, rng);               /* :43 */
                LayeredMaterial lm;
                bool lobe_ok = true;
                const vec3 thr_before = ray_color;
                if (!is_zero(surface.emissive) || !all_finite(ray_color)) {            /* the emission term of :44-51 (bounce_step adds it to a dummy below) */
                    RT3_NEED_RADIANCE();
                    radiance = radiance + ray_color * surface.emissive;
                }
                vec3 dummy = v3(0, 0, 0);
                PathVertexOut nv = LAYERED ? bounce_step_layered(surface, ro, rd, h4.x, urand, dummy, ray_color, lm, lobe_ok, &pdf_brdf)
                                           : bounce_step(surface, ro, rd, h4.x, urand, dummy, ray_color, &pdf_brdf);   /* :44-51 */
                o = nv.origin; d = nv.dir; thr = ray_color;
                alive = lobe_ok && ((path.y & 0xFFu) != p.bounces - 1);                /* :53 */
                if (nee && alive) {
                    vec2 ul = sample_2d(s, p.flags, x, y, p.frame, rng);
                    NeeSample ns = LAYERED ? nee_sample_layered(s, lm, thr_before, ul) : nee_sample(s, surface, ray_color, ul);
                    shadow = ns.valid; sdir = ns.dir; scon = ns.contribution;
                }
                rng_index = rng.index;
                b_next = (path.y & 0xFFu) + 1u;
            }
            regen = REGEN && !alive && sample + 1u < p.samples;                        /* the pixel has samples left: k_ref_regen starts the next one (:28) */
            if (regen) px_path[idx] = make_uint2(rng_index, sample << 8);
#undef RT3_NEED_RADIANCE
            if (have_radiance) px_radiance[idx] = make_float4(radiance.x, radiance.y, radiance.z, 1.0f);
            if (alive) px_path[idx] = make_uint2(rng_index, (sample << 8) | b_next);
        }
        const uint32_t pos = warp_enqueue(alive, count_out);
        if (alive) {
            store_ray(rays_out, pos, o, d, 0.001f, RT3_BACKGROUND_DEPTH);
            thr_out[pos] = make_float4(thr.x, thr.y, thr.z, __uint_as_float(idx));
            if (nee) nq.pdf_out[pos] = pdf_brdf;
        }
        if (REGEN) {
            const uint32_t rp = warp_enqueue(regen, regen_count);
            if (regen) regen_idx[rp] = idx;
        }
        if (nee) {
            const uint32_t sp = warp_enqueue(shadow, nq.scount);
            if (shadow) {
                shadows++;
                store_ray(nq.srays, sp, o, sdir, 0.001f, RT3_BACKGROUND_DEPTH);
                nq.scontrib[sp] = make_float4(scon.x, scon.y, scon.z, __uint_as_float(idx));
            }
        }
    }
    if (nee) flush_counter(shadows, &counters->rays_shadow);
    flush_counter(traced, &counters->rays_secondary);
    flush_counter(escaped, &counters->paths_escaped);
}

/* Ray regeneration, second half: every pixel on the regeneration queue starts its next sample (loop index 0 from the G-buffer, as in
 * k_ref_init) and the new paths are appended to the queue the survivors of this round were compacted into.  A kernel of its own so that the
 * work runs with all lanes busy: inside k_ref_shade the ~30 % of lanes whose path ended would execute it at a third of the warp width. */
template <bool LAYERED>
__global__ void __launch_bounds__(PASS_THREADS) k_ref_regen(SceneView s, rt3_params p, PixelMap m, RegenInputs gin, const uint32_t* __restrict__ regen_idx,
                                                            const uint32_t* __restrict__ regen_count, rt3_ray* __restrict__ rays_out,
                                                            float4* __restrict__ thr_out, uint32_t* count_out, float4* __restrict__ px_radiance,
                                                            uint2* __restrict__ px_path, NeeQueues nq, DeviceCounters* counters) {
    const uint32_t n = *regen_count;
    unsigned long long started = 0, shadows = 0;
    const bool nee = nee_enabled(p.flags);
    const uint32_t stride = gridDim.x * blockDim.x;
    for (uint32_t base = blockIdx.x * blockDim.x; base < n; base += stride) {
        const uint32_t i = base + threadIdx.x;
        bool alive = false, shadow = false;
        float pdf_brdf = 0.f;
        vec3 o = v3(0, 0, 0), d = v3(0, 0, 0), thr = v3(0, 0, 0), sdir = v3(0, 0, 0), scon = v3(0, 0, 0);
        uint32_t idx = 0;
        if (i < n) {
            idx = regen_idx[i];
            uint32_t x, y; size_t px;
            pixel_of(m, idx, x, y, px);
            const uint2 path = px_path[idx];
            uint32_t sample = path.y >> 8;
            const float gdepth = gin.depth[px];
            uint4 g4 = ((const uint4*)gin.gbuffer)[px];
            uint4_ packed = {g4.x, g4.y, g4.z, g4.w};
            const GbufferData surface = gbuffer_unpack(packed);                    /* :23 */
            RandomSamplerState rng = rng_init(x, y, p.frame);
            rng.index = path.x;
            const float4 rad4 = px_radiance[idx];
            vec3 radiance = v3(rad4.x, rad4.y, rad4.z);
            do {
                ++sample;
                started++;
                alive = start_sample<LAYERED>(s, p, nee, x, y, surface, gdepth, rng, radiance, o, d, thr, pdf_brdf, shadow, sdir, scon);
            } while (!alive && sample + 1u < p.samples);
            px_radiance[idx] = make_float4(radiance.x, radiance.y, radiance.z, 1.0f);
            if (alive) px_path[idx] = make_uint2(rng.index, (sample << 8) | 1u);
        }
        const uint32_t pos = warp_enqueue(alive, count_out);
        if (alive) {
            store_ray(rays_out, pos, o, d, 0.001f, RT3_BACKGROUND_DEPTH);
            thr_out[pos] = make_float4(thr.x, thr.y, thr.z, __uint_as_float(idx));
            if (nee) nq.pdf_out[pos] = pdf_brdf;
        }
        if (nee) {
            const uint32_t sp = warp_enqueue(shadow, nq.scount);
            if (shadow) {
                shadows++;
                store_ray(nq.srays, sp, o, sdir, 0.001f, RT3_BACKGROUND_DEPTH);
                nq.scontrib[sp] = make_float4(scon.x, scon.y, scon.z, __uint_as_float(idx));
            }
        }
    }
    flush_counter(started, &counters->paths_started);
    if (nee) flush_counter(shadows, &counters->rays_shadow);
}

/* RT3_FLAG_NEE: add the light-sample contribution of every unoccluded shadow ray (one per pixel and bounce at most, so
 * no atomics; it lands between the emission of this bounce and whatever the next bounce adds, as in the oracle's loop). */
__global__ void __launch_bounds__(PASS_THREADS) k_ref_shadow(const float4* __restrict__ scontrib, const uint32_t* __restrict__ occluded_bits,
                                                             const uint32_t* __restrict__ scount, float4* __restrict__ px_radiance) {
    const uint32_t n = *scount;
    const uint32_t stride = gridDim.x * blockDim.x;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        if ((occluded_bits[i >> 5] >> (i & 31u)) & 1u) continue;
        const float4 c4 = scontrib[i];
        const uint32_t idx = __float_as_uint(c4.w);
        float4 r = px_radiance[idx];
        r.x = r.x + c4.x; r.y = r.y + c4.y; r.z = r.z + c4.z;
        px_radiance[idx] = r;
    }
}

/* refrence_mode.slang:59-65 */
__global__ void __launch_bounds__(PASS_THREADS) k_ref_finish(rt3_params p, PixelMap m, StripMap sm, const float4* __restrict__ px_radiance,
                                                             const float* __restrict__ prev, float* __restrict__ light) {
    const uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= sm.local_work) return;
    const uint32_t idx = strip_to_work(sm, j);
    uint32_t x, y; size_t px;
    if (idx >= m.work || !pixel_of(m, idx, x, y, px)) return;
    const float4 r4 = px_radiance[idx];
    if (r4.w == 0.0f) return;                                   /* background: Light is not written (:18-21) */
    const float ns = (float)p.samples;
    vec3 radiance = v3(r4.x / ns, r4.y / ns, r4.z / ns);       /* :59 */
    vec3 pv = v3(0, 0, 0);
    if (p.blendfactor < 1.0f && prev) { float4 q = ((const float4*)prev)[px]; pv = v3(q.x, q.y, q.z); }
    vec3 o = temporal_blend(pv, radiance, p.blendfactor);       /* :61-65 */
    ((float4*)light)[px] = make_float4(o.x, o.y, o.z, 0.0f);
}

/* postprocess.slang:90-112 */
__global__ void __launch_bounds__(PASS_THREADS) k_postprocess(SceneView s, rt3_params p, PixelMap m, const float* __restrict__ depth,
                                                              const float* __restrict__ in, float* __restrict__ out) {
    const uint32_t idx = blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t x, y; size_t px;
    if (idx >= m.work || !pixel_of(m, idx, x, y, px)) return;
    float4 c = ((const float4*)in)[px];
    vec3 r = postprocess_pixel(s, p, x, y, depth[px], v3(c.x, c.y, c.z));
    ((float4*)out)[px] = make_float4(r.x, r.y, r.z, 1.0f);
}

/* postprocess.slang:90-112 fused with the conversion to the swapchain format (B8G8R8A8_UNORM, swapchain.rs:27-48): 4 bytes per pixel out
 * instead of 16 — what a host without CUDA-Vulkan interop reads back per presented frame */
__global__ void __launch_bounds__(PASS_THREADS) k_present_bgra8(SceneView s, rt3_params p, PixelMap m, const float* __restrict__ depth,
                                                                const float* __restrict__ in, uint32_t* __restrict__ out) {
    const uint32_t idx = blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t x, y; size_t px;
    if (idx >= m.work || !pixel_of(m, idx, x, y, px)) return;
    float4 c = ((const float4*)in)[px];
    out[px] = pack_bgra8(postprocess_pixel(s, p, x, y, depth[px], v3(c.x, c.y, c.z)));
}

/* ---- screen-space probes ------------------------------------------------------------------------
 * trace_probes.slang as raygen -> k_trace_dyn -> texel kernel; spherical_harmonic_conversion.slang; interpolate_probes.slang */
__global__ void __launch_bounds__(PASS_THREADS) k_probe_raygen(rt3_params p, uint32_t aw, uint32_t ah, const float* __restrict__ depth,
                                                               const uint16_t* __restrict__ directions, rt3_ray* __restrict__ rays) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= aw * ah) return;
    const uint32_t ax = i % aw, ay = i / aw, W = (uint32_t)p.window_size[0];
    const float pd = depth[(size_t)(ay / 8u * 16u) * W + ax / 8u * 16u];
    ProbeRay pr = probe_setup_ray(p, ax, ay, pd, directions ? (uint32_t)directions[i] : probe_uniform_direction(ax, ay));
    if (pr.background) { pr.ray.tmin = 0.f; pr.ray.tmax = 0.f; }                 /* nothing to trace (trace_probes.slang:28-31) */
    store_ray(rays, i, v3(pr.ray.o), v3(pr.ray.d), pr.ray.tmin, pr.ray.tmax);
}
/* staged != NULL (a direction image is bound): the texel value and its target are staged per source texel and k_probe_resolve applies
 * them in the defined order (rt3_shared_math.h: clears first, then target writes, highest source index wins) */
__global__ void __launch_bounds__(PASS_THREADS) k_probe_texel(SceneView s, rt3_params p, uint32_t aw, uint32_t ah, const float* __restrict__ depth,
                                                              const uint16_t* __restrict__ directions, const rt3_ray* __restrict__ rays,
                                                              const rt3_hit* __restrict__ hits, const float4* __restrict__ prev_atlas,
                                                              float4* __restrict__ atlas, float4* __restrict__ staged, uint32_t* __restrict__ staged_target,
                                                              DeviceCounters* counters) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    unsigned long long traced = 0;
    if (i < aw * ah) {
        const uint32_t ax = i % aw, ay = i / aw, W = (uint32_t)p.window_size[0];
        const float pd = depth[(size_t)(ay / 8u * 16u) * W + ax / 8u * 16u];
        if (pd == RT3_BACKGROUND_DEPTH) {
            atlas[i] = make_float4(0.f, 0.f, 0.f, RT3_BACKGROUND_DEPTH);
            if (staged) staged_target[i] = 0xFFFFFFFFu;
        } else {
            traced = 1;
            const ProbeRay pr = probe_setup_ray(p, ax, ay, pd, directions ? (uint32_t)directions[i] : probe_uniform_direction(ax, ay));
            const float4 h4 = ((const float4*)hits)[i];
            rt3_hit h; h.t = h4.x; h.u = h4.y; h.v = h4.z; h.prim = __float_as_uint(h4.w);
            vec3 prev = v3(0.f, 0.f, 0.f);
            if (p.blendfactor < 1.0f && prev_atlas) { const float4 q = prev_atlas[i]; prev = v3(q.x, q.y, q.z); }
            const vec4 t = probe_texel(s, p, pr.ray, h, prev);
            if (staged) { staged[i] = make_float4(t.x, t.y, t.z, t.w); staged_target[i] = pr.target_y * aw + pr.target_x; }
            else atlas[(size_t)pr.target_y * aw + pr.target_x] = make_float4(t.x, t.y, t.z, t.w);
        }
    }
    flush_counter(traced, &counters->rays_secondary);
}
/* one thread per atlas texel of a non-background probe: the value of the highest source texel of the probe that targets it, else the clear */
__global__ void __launch_bounds__(PASS_THREADS) k_probe_resolve(uint32_t aw, uint32_t ah, const float4* __restrict__ staged, const uint32_t* __restrict__ staged_target,
                                                                float4* __restrict__ atlas) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= aw * ah) return;
    if (staged_target[i] == 0xFFFFFFFFu) return;                         /* background probe: k_probe_texel wrote the marker */
    const uint32_t ax = i % aw, ay = i / aw, bx = ax & ~7u, by = ay & ~7u;
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);                          /* trace_probes.slang:32 */
    for (int k = 63; k >= 0; --k) {
        const uint32_t src = (by + (uint32_t)k / 8u) * aw + bx + (uint32_t)k % 8u;
        if (staged_target[src] == i) { v = staged[src]; break; }
    }
    atlas[i] = v;
}
__global__ void __launch_bounds__(PASS_THREADS) k_probe_sh(rt3_params p, const float* __restrict__ atlas, float* __restrict__ sh) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t pw = probe_count_x(p), ph = probe_count_y(p);
    if (i >= pw * ph * 3u) return;
    const uint32_t c = i % 3u, px = (i / 3u) % pw, py = i / (3u * pw);
    const sh3 r = probe_project_sh(atlas, pw * 8u, px, py, c);
    float* o = sh + 9u * (size_t)probe_sh_index(px, py, c);
#pragma unroll
    for (int k = 0; k < 9; ++k) o[k] = r.c[k];
}
__global__ void __launch_bounds__(PASS_THREADS) k_probe_interpolate(rt3_params p, const uint32_t* __restrict__ gbuffer, const float* __restrict__ depth,
                                                                    const float* __restrict__ sh, float4* __restrict__ light, uint32_t* __restrict__ failed_target) {
    const uint32_t W = (uint32_t)p.window_size[0], H = (uint32_t)p.window_size[1];
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= W * H) return;
    uint32_t ox, oy; vec4 v;
    const int k = probe_interpolate_pixel(p, i % W, i / W, gbuffer, depth, sh, ox, oy, v);
    failed_target[i] = (k == 2) ? oy * W + ox : 0xFFFFFFFFu;
    if (k == 1) light[i] = make_float4(v.x, v.y, v.z, v.w);
}
/* the "interpolation failed" markers (interpolate_probes.slang:71-76) land on other pixels: applied after every regular write */
__global__ void __launch_bounds__(PASS_THREADS) k_probe_failed(uint32_t n, const uint32_t* __restrict__ failed_target, float4* __restrict__ light) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint32_t t = failed_target[i];
    if (t != 0xFFFFFFFFu) light[t] = make_float4(1.f, 0.f, 0.f, 1.f);
}

#ifdef __CUDACC__
/* structured_importance_sampling.slang:19-72 (the live lines, rt3_shared_math.h): one block of 64 threads per probe, the probe's 256
 * G-buffer normals and the 64 pdfs in shared memory, the bitonic network of math.slang:142-163 with a block barrier after every step.
 * The 256-term pdf sum reads normals[i] as a broadcast (every thread the same address): conflict-free. */
__global__ void __launch_bounds__(64) k_probe_sis(rt3_params p, const uint32_t* __restrict__ gbuffer, uint16_t* __restrict__ out) {
    __shared__ vec3 normals[256];
    __shared__ float pdfs[64];
    const uint32_t pw = probe_count_x(p), W = (uint32_t)p.window_size[0], aw = pw * 8u;
    const uint32_t pcx = blockIdx.x % pw, pcy = blockIdx.x / pw, ti = threadIdx.x;
#pragma unroll
    for (uint32_t y = 0; y < 2u; ++y)
#pragma unroll
        for (uint32_t x = 0; x < 2u; ++x) {                               /* :25-31 */
            uint32_t px, py;
            probe_sis_pixel(pcx, pcy, ti, x, y, px, py);
            normals[ti * 4u + y * 2u + x] = unpack_normal_11_10_11(gbuffer[4u * ((size_t)py * W + px) + 1u]);
        }
    __syncthreads();                                                      /* :33 */
    const float pdf = probe_sis_pdf(normals, probe_sis_direction(ti));   /* :34-46 */
    pdfs[ti] = pdf;
    __syncthreads();
    for (uint32_t k = 2; k <= 64u; k *= 2u)                               /* :47, math.slang:142-163 */
        for (uint32_t j = k / 2u; j > 0u; j /= 2u) {
            const uint32_t l = ti ^ j;
            if (l > ti) {
                const float a = pdfs[ti], b = pdfs[l];
                if (bitonic_exchange(ti, k, j, a, b)) { pdfs[ti] = b; pdfs[l] = a; }
            }
            __syncthreads();
        }
    out[(size_t)(pcy * 8u + ti / 8u) * aw + pcx * 8u + ti % 8u] = (uint16_t)probe_sis_word(pdf, pdfs[0], ti);   /* :55-70 */
}
#endif

__global__ void __launch_bounds__(PASS_THREADS) k_untile(const rt3_rect* __restrict__ rects, uint32_t n_tiles, uint32_t tile_w, uint32_t tile_h,
                                                         const float4* __restrict__ slab, float4* __restrict__ frame, uint32_t frame_w, uint32_t frame_h) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const size_t per = (size_t)tile_w * tile_h;
    if (i >= per * n_tiles) return;
    const uint32_t k = (uint32_t)(i / per), rem = (uint32_t)(i - (size_t)k * per);
    const uint32_t ly = rem / tile_w, lx = rem - ly * tile_w;
    const rt3_rect r = rects[k];
    if (lx >= r.w || ly >= r.h) return;
    const uint32_t x = r.x0 + lx, y = r.y0 + ly;
    if (x < frame_w && y < frame_h) frame[(size_t)y * frame_w + x] = slab[i];
}

/* Peer-memory delivery of finished tiles (multi-GPU, SURVEY section 8e "zero-copy variant"): tile k of the rank-local slab is stored row by
 * row at dst[k], the address of the tile's top-left pixel inside the band buffer of the rank that presents those rows — its own memory or a
 * peer GPU's, mapped over NVLink.  One kernel replaces the framebuffer collective; the owners only need a barrier afterwards. */
__global__ void __launch_bounds__(PASS_THREADS) k_scatter_tiles(const rt3_rect* __restrict__ rects, const unsigned long long* __restrict__ dst, uint32_t n_tiles,
                                                                uint32_t tile_w, uint32_t tile_h, const float4* __restrict__ slab, uint32_t dst_pitch) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const size_t per = (size_t)tile_w * tile_h;
    if (i >= per * n_tiles) return;
    const uint32_t k = (uint32_t)(i / per), rem = (uint32_t)(i - (size_t)k * per);
    const uint32_t ly = rem / tile_w, lx = rem - ly * tile_w;
    const rt3_rect r = rects[k];
    if (lx >= r.w || ly >= r.h) return;
    float4* d = (float4*)dst[k];
    d[(size_t)ly * dst_pitch + lx] = slab[i];
}

}  // namespace
#endif
