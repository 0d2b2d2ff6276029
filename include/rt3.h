/*
 * rt3.h — C ABI of the B200-native path-tracing core (librt3.so).
 *
 * This is the drop-in boundary for ONE hot path of DerEchteKarsten/RayTracer3: the ray-traced
 * G-buffer pass + multi-bounce "reference mode" path tracer, together with the acceleration-
 * structure build and the scene upload they depend on.  The reference has no FFI of its own; the
 * entry points below are what a Rust `extern "C"` block would bind from the three internal seams
 * (all paths under /root/reference):
 *
 *   seam                         reference call                                   file:line
 *   ---------------------------  -----------------------------------------------  ---------------------------------------------
 *   scene upload                 DynamicBuffer::push / RenderWorld global bufs    src/renderer/vulkan/buffer.rs:406-420,
 *                                                                                 src/renderer/world/mod.rs:83-117
 *   acceleration-structure build RayTracingContext::create_acceleration_structure src/renderer/vulkan/raytracing.rs:88-148
 *   ray pass launch              RayTracingPass … .launch(WorkSize2D) →           src/renderer/render_graph/executions.rs:80-121,
 *                                RayTracingPipelineHandle::launch → cmd_trace_rays src/renderer/pipeline_cache/mod.rs:43-76
 *   per-frame constants          GConst                                           src/renderer/mod.rs:47-63,
 *                                                                                 shaders/include/datatypes.slang:28-43
 *
 * Conventions
 *   - Plain C: pointers + sizes, no C++ / torch types.  Every function returns rt3_status
 *     (0 = OK, <0 = error) and never throws or aborts across the boundary; the message of the
 *     last error on a context is available from rt3_last_error().
 *   - All image / ray / hit pointers passed to rt3_pass_* and rt3_trace_* are CALLER-OWNED DEVICE
 *     pointers (the analogue of render-graph resources); the context owns the scene, the BVH and
 *     the internal wavefront queues.  `*_host` variants take HOST pointers and do the copies
 *     themselves (the end-to-end path a non-CUDA host such as the Rust renderer would use).
 *   - Pass calls are asynchronous on `cuda_stream` (a cudaStream_t cast to void*, NULL = default
 *     stream): the analogue of recording into the reference's command buffer.  The pass / trace calls
 *     of ONE context share its internal queues, so they execute in CALL ORDER whatever streams they are
 *     given: a call on another stream than the previous call's first waits, on the device, for the
 *     previous call's work (one event per call; no host synchronisation).  Use one context per stream
 *     for work that is meant to overlap.  Internally rt3_pass_reference splits the frame into strips that
 *     run as concurrent wavefronts on the context's own streams and joins them back into `cuda_stream`.
 *   - A context is bound to one CUDA device; it is thread-compatible, not thread-safe.
 *   - There is NO CPU fallback: without a CUDA device rt3_create fails with RT3_ERR_NO_DEVICE.
 */
#ifndef RT3_H
#define RT3_H

#include <stdint.h>
#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

#define RT3_API __attribute__((visibility("default")))

typedef struct rt3_context rt3_context;
typedef int rt3_status;

enum {
    RT3_OK = 0,
    RT3_ERR_INVALID = -1,     /* bad argument                                   */
    RT3_ERR_NO_DEVICE = -2,   /* no usable CUDA device (no CPU fallback exists) */
    RT3_ERR_CUDA = -3,        /* a CUDA runtime call failed                     */
    RT3_ERR_STATE = -4,       /* call order (e.g. pass before rt3_accel_build)  */
    RT3_ERR_OOM = -5
};

/* ---- records (layout-identical to the reference's) ------------------------------------------ */

/* Vertex: src/assets/mod.rs:127-133 == shaders/include/datatypes.slang:5-9 (32 B). */
typedef struct { float p[3], n[3], t[2]; } rt3_vertex;

/* GeometryInfo: shaders/include/datatypes.slang:11-19, plus the `transform` that
 * shaders/include/hit_logic.slang:23 reads but the struct lacks, plus index_count
 * (the BuildRangeInfo.primitive_count of raytracing.rs:88-100).  xform is a row-major 3x4. */
typedef struct {
    float    base_color[4];
    int32_t  base_color_tex;      /* -1 = none (hit_logic.slang:30)  */
    float    metallic;
    uint32_t index_offset;        /* first index in the global index buffer */
    uint32_t vertex_offset;       /* added to every index value              */
    float    emission[4];
    float    roughness;
    uint32_t index_count;         /* 3 * triangles */
    float    _pad[2];
    float    xform[12];
} rt3_geometry_info;              /* 112 B */

/* GConst: src/renderer/mod.rs:47-63 (304 B, repr(C)); matrices column-major (glam).
 * `flags` reuses pad[0]; 0 == reference-faithful behaviour. */
typedef struct {
    float    proj[16], view[16], proj_inv[16], view_inv[16];
    float    window_size[2];
    uint32_t frame;
    float    blendfactor;
    uint32_t bounces, samples;          /* limits of this implementation: bounces <= 255, 1 <= samples < 2^24 */
    uint32_t proberng;
    float    cell_size;
    uint32_t mouse[2];
    uint32_t flags;
    uint32_t pad;
} rt3_params;

enum {
    RT3_FLAG_ENV       = 1u << 0, /* miss adds throughput * environment(dir): the term commented out at
                                     shaders/old/refrence_mode.slang:38 */
    RT3_FLAG_BLUENOISE = 1u << 1, /* blue-noise tile + R2 offsets instead of the murmur3 stream */
    RT3_FLAG_NEE       = 1u << 2, /* environment importance sampling with any-hit shadow rays */
    RT3_FLAG_SPECULAR  = 1u << 3  /* material evaluation = DiffuseBrdf + SpecularBrdf (GGX VNDF, Smith, Schlick:
                                     shaders/include/brdf.slang:95-318) layered by metalness, instead of DiffuseBrdf only */
};

typedef struct { uint32_t x0, y0, w, h; } rt3_rect;                /* tile; full frame = {0,0,W,H} */
typedef struct { float o[3], tmin, d[3], tmax; } rt3_ray;           /* 32 B == RayDesc             */
typedef struct { float t, u, v; uint32_t prim; } rt3_hit;           /* 16 B; prim = global triangle id, 0xFFFFFFFF = miss */

#define RT3_MISS 0xFFFFFFFFu
#define RT3_BACKGROUND_DEPTH 100000.0f                              /* datatypes.slang:3 */

/* RT3_BUILD_TWO_LEVEL: one bottom-level tree per distinct mesh (= distinct geometry range of rt3_scene_push_instance) and a
 * top-level tree over the instances — the BLAS / TLAS split of VK_KHR_acceleration_structure (raytracing.rs:88-148 builds
 * either level).  Default: instances are flattened into one world-space tree.  Hits are bit-identical in both modes. */
/* Quality levels of the binary tree under the 8-wide collapse (hits are bit-identical in all of them):
 *   RT3_BUILD_FAST_TRACE (default, also flags == 0; the reference's PREFER_FAST_TRACE, raytracing.rs:121): treelet restructuring (Karras &
 *     Aila 2013: the best topology of every 7-subtree treelet by dynamic programming, two bottom-up passes) and the children of each wide
 *     node chosen by the SAH dynamic programme of Ylitie et al. 2017.  Measured on B200, 263 k triangles: 4.9 ms build.  Trees of more than
 *     4 M records (a flattened 16.8 M-triangle scene would take 146 ms) get the rotations below instead, unless RT3_BUILD_TREELET is set.
 *   RT3_BUILD_ROTATIONS: three bottom-up passes of SAH tree rotations instead of the treelets (1.9 ms build, traversal 4 % slower).
 *   RT3_BUILD_FAST_BUILD: plain LBVH, children chosen greedily by surface area (1.4 ms build, traversal 11 % slower).
 * RT3_BUILD_TREELET: treelet restructuring whatever the size. */
enum { RT3_BUILD_FAST_TRACE = 1u, RT3_BUILD_FAST_BUILD = 2u, RT3_BUILD_TWO_LEVEL = 4u, RT3_BUILD_TREELET = 8u, RT3_BUILD_ROTATIONS = 16u };

typedef struct {
    double   ms_flatten, ms_morton, ms_sort, ms_hierarchy, ms_refit, ms_collapse, ms_total;
    uint64_t node_bytes, tri_bytes;
    uint32_t triangles, wide_nodes, bvh2_nodes, levels;
    float    sah_cost;
    float    scene_lo[3], scene_hi[3];
} rt3_build_stats;

typedef struct {
    uint64_t rays_primary, rays_secondary, rays_shadow;             /* rays issued by the last pass call(s) since reset */
    uint64_t paths_started, paths_escaped;
} rt3_counters;

/* ---- context ------------------------------------------------------------------------------- */
RT3_API rt3_status  rt3_create(int cuda_device, rt3_context** out);
RT3_API void        rt3_destroy(rt3_context*);
RT3_API const char* rt3_last_error(const rt3_context*);            /* ctx may be NULL: error of a failed rt3_create */
RT3_API uint32_t    rt3_abi_version(void);
RT3_API size_t      rt3_sizeof(int which);                         /* 0 params, 1 vertex, 2 geometry_info, 3 ray, 4 hit */

/* ---- scene == RenderWorld's global append-only buffers (world/mod.rs:103-117) --------------- */
RT3_API rt3_status rt3_scene_clear(rt3_context*);
RT3_API rt3_status rt3_scene_push_vertices(rt3_context*, const rt3_vertex* host, uint32_t n, uint32_t* out_first);
RT3_API rt3_status rt3_scene_push_indices(rt3_context*, const uint32_t* host, uint32_t n, uint32_t* out_first);
RT3_API rt3_status rt3_scene_push_geometry(rt3_context*, const rt3_geometry_info* host, uint32_t* out_geometry_index);
/* InstanceInfo{mesh_index, transform} (world/mod.rs:32-37): an instance places geometries
 * [first_geometry, first_geometry+geometry_count) with a row-major 3x4 transform.  With no
 * instance pushed, every geometry is instanced once with the identity. */
RT3_API rt3_status rt3_scene_push_instance(rt3_context*, uint32_t first_geometry, uint32_t geometry_count,
                                           const float xform3x4[12], uint32_t* out_instance_index);
/* Textures[] of the missing bindless_descriptor.slang: RGBA8, sRGB-encoded colour, sampled bilinear / repeat / mip 0. */
RT3_API rt3_status rt3_scene_set_texture(rt3_context*, uint32_t slot, uint32_t w, uint32_t h, const uint8_t* rgba8);
/* Skybox (resources/skybox2.exr role): equirect RGB float, bilinear, repeat-u / clamp-v. */
RT3_API rt3_status rt3_scene_set_environment(rt3_context*, uint32_t w, uint32_t h, const float* rgb);
/* resources/bluenoise.png role: RGBA8 tile, w and h powers of two. */
RT3_API rt3_status rt3_scene_set_bluenoise(rt3_context*, uint32_t w, uint32_t h, const uint8_t* rgba8);

/* ---- accel == create_acceleration_structure (raytracing.rs:88-148) -------------------------- */
RT3_API rt3_status rt3_accel_build(rt3_context*, uint32_t flags, void* cuda_stream, rt3_build_stats* out);
/* global triangle id -> (instance, geometry, primitive): GeometryIndex()/PrimitiveIndex() of default_hit.slang:7-8 */
RT3_API rt3_status rt3_resolve_prim(rt3_context*, uint32_t prim, uint32_t* instance, uint32_t* geometry, uint32_t* primitive);

/* ---- passes == RayTracingPass(...).launch(WorkSize2D) --------------------------------------- */
/* shaders/old/gbuffer.slang:8-20.  Images are full-frame, pitch = window_size.x elements; only
 * pixels inside `rect` are touched.  opt_ids (2 x u32 per pixel: global prim id, geometry index;
 * RT3_MISS on miss) and opt_bary (2 x f32: u, v) may be NULL. */
RT3_API rt3_status rt3_pass_gbuffer(rt3_context*, const rt3_params*, rt3_rect, uint32_t* gbuffer_rgba32ui, float* depth_r32f,
                                    uint32_t* opt_ids_rg32ui, float* opt_bary_rg32f, void* cuda_stream);
/* shaders/old/refrence_mode.slang:14-66.  prev_light may be NULL when blendfactor >= 1. */
RT3_API rt3_status rt3_pass_reference(rt3_context*, const rt3_params*, rt3_rect, const uint32_t* gbuffer_rgba32ui,
                                      const float* depth_r32f, const float* prev_light_rgba32f, float* light_rgba32f,
                                      void* cuda_stream);
/* shaders/old/postprocess.slang:90-112 (AgX tonemap + sky background). */
RT3_API rt3_status rt3_pass_postprocess(rt3_context*, const rt3_params*, rt3_rect, const float* depth_r32f,
                                        const float* in_rgba32f, float* out_rgba32f, void* cuda_stream);

/* postprocess.slang:90-112 fused with the conversion to the reference's swapchain format (B8G8R8A8_UNORM, swapchain.rs:27-48): one 32-bit
 * word per pixel (A << 24 | R << 16 | G << 8 | B), a quarter of the bytes of the RGBA32F image for a host that reads the presented frame back. */
RT3_API rt3_status rt3_pass_present_bgra8(rt3_context*, const rt3_params*, rt3_rect, const float* depth_r32f, const float* in_rgba32f,
                                          uint32_t* out_b8g8r8a8, void* cuda_stream);

/* Screen-space probe GI: shaders/old/trace_probes.slang, spherical_harmonic_conversion.slang,
 * interpolate_probes.slang.  One probe per 16x16 pixels, 8x8 octahedral directions per probe.
 *   atlas       (W/16*8) x (H/16*8) RGBA32F: rgb = radiance along the direction (blended with prev_atlas by blendfactor), a = hit distance
 *   directions  optional u16 per atlas texel (structured_importance_sampling.slang:55-69: bit 15 = mip, low bits = direction index), as written
 *               by rt3_pass_probe_sis; NULL = the uniform assignment (texel k of a probe traces octahedral cell k)
 *   sh          rt3_probe_layout().sh_records records of 9 floats (float3x3 = SH L2), Z-curve indexed by (probe.x * 3 + channel, probe.y)
 * The probe shaders of the reference are in a debug state; include/rt3_shared_math.h documents line by line what is restated. */
RT3_API rt3_status rt3_probe_layout(const rt3_params*, uint32_t* atlas_w, uint32_t* atlas_h, uint32_t* sh_records);
/* shaders/old/structured_importance_sampling.slang:19-72: one u16 direction word per atlas texel from the probe's 16x16 G-buffer normals
 * (64 threads per probe, shared-memory bitonic sort; the LIVE lines of the shader, include/rt3_shared_math.h).  The G-buffer image must be
 * defined on every pixel the probes cover (background pixels keep whatever the image held: clear it once). */
RT3_API rt3_status rt3_pass_probe_sis(rt3_context*, const rt3_params*, const uint32_t* gbuffer_rgba32ui, uint16_t* directions_r16ui, void* cuda_stream);
RT3_API rt3_status rt3_pass_probe_trace(rt3_context*, const rt3_params*, const uint32_t* gbuffer_rgba32ui, const float* depth_r32f,
                                        const uint16_t* opt_directions_r16ui, const float* prev_atlas_rgba32f, float* atlas_rgba32f, void* cuda_stream);
/* rt3_pass_probe_trace that ALLOCATES NOTHING.  No counterpart in the reference: added for callers that
 * cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  Same arguments, same results as
 * rt3_pass_probe_trace; the per-texel scratch (probe ray, hit, staged value + target) and the traversal kernel's ray-fetch counter live in
 * one file-scope `static __device__` array of 34 MiB that every GPU which loads the library holds until exit.
 *   - limit: atlas_w * atlas_h <= RT3_STATIC_PROBE_MAX_TEXELS (a 1920x1088 window); larger frames return RT3_ERR_INVALID with a message;
 *   - the scene tables must be current (RT3_ERR_STATE otherwise: this call never refreshes them) and rt3_profile_enable must be off;
 *   - every call on a device shares the array: it keeps the previous call's contents, each call overwrites all it reads and zeroes its
 *     counter itself; calls that OVERLAP on the GPU (other streams or contexts, no ordering between them) race — serialise them. */
#define RT3_STATIC_PROBE_MAX_TEXELS 524288u
RT3_API rt3_status rt3_pass_probe_trace_static(rt3_context*, const rt3_params*, const uint32_t* gbuffer_rgba32ui, const float* depth_r32f,
                                               const uint16_t* opt_directions_r16ui, const float* prev_atlas_rgba32f, float* atlas_rgba32f,
                                               void* cuda_stream);
RT3_API rt3_status rt3_pass_probe_sh(rt3_context*, const rt3_params*, const float* atlas_rgba32f, float* sh, void* cuda_stream);
RT3_API rt3_status rt3_pass_probe_interpolate(rt3_context*, const rt3_params*, const uint32_t* gbuffer_rgba32ui, const float* depth_r32f,
                                              const float* sh, float* light_rgba32f, void* cuda_stream);

/* Multi-tile variants used for screen-tile sharding across GPUs.  `tiles` is a HOST array of
 * n_tiles rects (all inside the frame).  Images are rank-local tile-major SLABS: tile k occupies
 * elements [k*tile_w*tile_h, (k+1)*tile_w*tile_h) row-major inside the tile, where tile_w/tile_h
 * are the maximum w/h over the list. */
RT3_API rt3_status rt3_pass_gbuffer_tiles(rt3_context*, const rt3_params*, const rt3_rect* tiles, uint32_t n_tiles,
                                          uint32_t* gbuffer_slab, float* depth_slab, void* cuda_stream);
RT3_API rt3_status rt3_pass_reference_tiles(rt3_context*, const rt3_params*, const rt3_rect* tiles, uint32_t n_tiles,
                                            const uint32_t* gbuffer_slab, const float* depth_slab,
                                            const float* prev_light_slab, float* light_slab, void* cuda_stream);
/* Scatter a gathered slab (n_tiles tiles of tile_w x tile_h RGBA32F) into a full frame of pitch frame_w. */
RT3_API rt3_status rt3_untile_rgba32f(rt3_context*, const rt3_rect* tiles, uint32_t n_tiles, uint32_t tile_w, uint32_t tile_h,
                                      const float* slab, float* frame, uint32_t frame_w, uint32_t frame_h, void* cuda_stream);

/* Peer-memory delivery (multi-GPU): tile k of `slab` is stored at dst_top_left[k], the DEVICE address of the tile's top-left pixel in the
 * frame (or band of rows) of whoever presents it — this GPU's memory or a peer's mapped over NVLink (cudaIpc / VMM / torch symmetric
 * memory) — rows dst_pitch_px pixels apart.  Empty rects are skipped.  Both arrays are HOST arrays of n_tiles entries.  The caller
 * synchronises with the destination's owner afterwards (a barrier); see raytracer3_b200/host.py TiledRenderer.exchange. */
RT3_API rt3_status rt3_scatter_tiles_rgba32f(rt3_context*, const rt3_rect* tiles, const uint64_t* dst_top_left, uint32_t n_tiles, uint32_t tile_w,
                                             uint32_t tile_h, const float* slab, uint32_t dst_pitch_px, void* cuda_stream);

/* ---- raw queries == trace() / TraceRay ------------------------------------------------------ */
RT3_API rt3_status rt3_trace_closest(rt3_context*, const rt3_ray* rays, uint64_t n, rt3_hit* out, void* cuda_stream);
RT3_API rt3_status rt3_trace_any(rt3_context*, const rt3_ray* rays, uint64_t n, uint32_t* out_bits, void* cuda_stream);

/* ---- end-to-end host-buffer entry points ---------------------------------------------------- */
/* One frame of both passes with HOST output: params are copied in, light (W*H*4 floats) is copied out
 * to `light_host` (pinned or pageable).  prev_light_host may be NULL (blendfactor >= 1).  Synchronous. */
RT3_API rt3_status rt3_render_frame_host(rt3_context*, const rt3_params*, const float* prev_light_host, float* light_host,
                                         float* opt_depth_host, uint32_t* opt_gbuffer_host);
/* Frames in flight — the reference keeps FRAMES_IN_FLIGHT = 3 frames queued behind a timeline semaphore
 * (src/renderer/vulkan/swapchain.rs:8, src/renderer/render_graph/mod.rs:656-665).  rt3_frame_submit_host enqueues one
 * frame of both passes on the context's own compute stream and its read-back into `light_host` on a copy stream, and
 * returns at once with a ticket; rt3_frame_wait blocks until that frame (and every older one) is in host memory.  At most
 * 3 frames are in flight: a 4th submit first waits for the oldest.  `light_host` (and prev_light_host) should be pinned
 * (rt3_host_alloc or cudaHostRegister) or the copy degrades to a synchronous one; they must stay valid until the wait.
 * Do not interleave with other pass calls on the same context without waiting for the outstanding tickets. */
RT3_API rt3_status rt3_frame_submit_host(rt3_context*, const rt3_params*, const float* prev_light_host, float* light_host,
                                         uint64_t* out_ticket);
RT3_API rt3_status rt3_frame_wait(rt3_context*, uint64_t ticket);
RT3_API rt3_status rt3_host_alloc(rt3_context*, size_t bytes, void** out);        /* pinned host memory */
RT3_API rt3_status rt3_host_free(rt3_context*, void* p);
RT3_API rt3_status rt3_trace_closest_host(rt3_context*, const rt3_ray* rays_host, uint64_t n, rt3_hit* out_host);
RT3_API rt3_status rt3_trace_any_host(rt3_context*, const rt3_ray* rays_host, uint64_t n, uint32_t* out_bits_host);

/* ---- tuning ------------------------------------------------------------------------------------ */
/* RT3_OPT_STRIPS: number of concurrent wavefronts rt3_pass_reference splits a frame into (1..4; default 2, or $RT3_STRIPS).  The image does
 * not depend on it.  bench.py sets 1 for the pass in which single kernels are timed (kernels of concurrent strips overlap). */
enum { RT3_OPT_STRIPS = 1 };
RT3_API rt3_status rt3_set_option(rt3_context*, int option, int value);

/* ---- instrumentation ------------------------------------------------------------------------ */
RT3_API rt3_status rt3_counters_read(rt3_context*, rt3_counters* out, int reset);   /* synchronises the device */
RT3_API uint64_t   rt3_kernel_launches(const rt3_context*);                          /* kernels launched by this context so far */
/* L2-resident read sweep (the L2 roofline denominator of SURVEY §8d): returns GB/s, best of `iters`. */
RT3_API rt3_status rt3_measure_l2_bandwidth(rt3_context*, size_t bytes, int iters, double* out_gbs);
/* Traversal statistics for `n` rays (instrumented kernel, not the timed one): averages per ray. */
typedef struct { double nodes_per_ray, tris_per_ray, node_bytes_per_ray, tri_bytes_per_ray; } rt3_trace_stats;
RT3_API rt3_status rt3_trace_closest_stats(rt3_context*, const rt3_ray* rays, uint64_t n, rt3_trace_stats* out);

/* The ramp-up and tail of a persistent traversal launch, measured by an instrumented instantiation of the kernel (three timer reads per warp):
 * kernel_us = first warp in to last warp out; mean_warp_busy_us = a warp's time in the kernel; mean_warp_tail_us = its time after it first
 * found the queue exhausted (it finishes the rays it still holds, with fewer and fewer lanes busy); tail_fraction = their ratio over all warps. */
typedef struct { double kernel_us, mean_warp_busy_us, mean_warp_tail_us, tail_fraction; uint32_t warps, _pad; } rt3_tail_stats;
RT3_API rt3_status rt3_trace_closest_tail_stats(rt3_context*, const rt3_ray* rays, uint64_t n, rt3_tail_stats* out);

/* Per-kernel timing of the closest-hit traversal launches made by the passes: while enabled, every launch is bracketed by
 * CUDA events on its own stream; rt3_profile_read synchronises, returns the summed duration and launch count, and resets. */
RT3_API rt3_status rt3_profile_enable(rt3_context*, int enable);
RT3_API rt3_status rt3_profile_read(rt3_context*, double* trace_ms, uint64_t* trace_launches);
/* Copy the built acceleration structure to host memory (80-byte wide nodes, 48-byte triangle records; sizes from
 * rt3_build_stats.node_bytes / tri_bytes).  Lets tests prove with an order-independent check that the build kept every
 * triangle exactly once, and replay the traversal on the CPU. */
RT3_API rt3_status rt3_accel_download(rt3_context*, void* nodes_host, size_t nodes_bytes, void* tris_host, size_t tris_bytes);

#ifdef __cplusplus
}
#endif
#endif /* RT3_H */
