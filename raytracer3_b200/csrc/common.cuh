/*
 * common.cuh — context object, device buffers and launch helpers of librt3.so.
 *
 * The context plays the role of the reference's RenderWorld + RayTracingContext
 * (/root/reference/src/renderer/world/mod.rs:103-117, src/renderer/vulkan/raytracing.rs:56-86): it owns
 * the global append-only scene buffers, the acceleration structure and the scratch memory of the passes.
 */
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string>
#include <vector>

#include "rt3.h"
#include "rt3_shared_math.h"
#include "bvh_logic.h"

#define RT3_ABI_VERSION 1u

struct rt3_error_exit { rt3_status code; };

#define RT3_CUDA(ctx, call)                                                                                   \
    do {                                                                                                      \
        cudaError_t e__ = (call);                                                                             \
        if (e__ != cudaSuccess) {                                                                             \
            (ctx)->set_error(std::string(#call) + ": " + cudaGetErrorString(e__) + " (" __FILE__ ":" + std::to_string(__LINE__) + ")"); \
            throw rt3_error_exit{e__ == cudaErrorMemoryAllocation ? RT3_ERR_OOM : RT3_ERR_CUDA};             \
        }                                                                                                     \
    } while (0)

#define RT3_FAIL(ctx, code, msg)            \
    do {                                    \
        (ctx)->set_error(msg);              \
        throw rt3_error_exit{code};         \
    } while (0)

/* Growable device buffer: the analogue of DynamicBuffer (src/renderer/vulkan/buffer.rs:297-421):
 * starts small, doubles on demand, append-only `push` from host memory. */
struct DynBuf {
    void* ptr = nullptr;
    size_t size = 0, capacity = 0;   /* bytes */
    void release() { if (ptr) cudaFree(ptr); ptr = nullptr; size = capacity = 0; }
};

/* Plain owned device allocation, resized (not preserved) on demand. */
struct Scratch {
    void* ptr = nullptr;
    size_t capacity = 0;
    void release() { if (ptr) cudaFree(ptr); ptr = nullptr; capacity = 0; }
    template <class T> T* as() const { return (T*)ptr; }
};

struct InstanceRec { uint32_t first_geometry, geometry_count; float xform[12]; };
struct TextureRec { uint8_t* data = nullptr; uint32_t w = 0, h = 0; };

#include "pass_types.h"


/* One wavefront pipeline of the reference pass: its queues, its queue / work counters and (lanes > 0) its stream.  rt3_launch_reference
 * runs one strip of the frame per lane (passes.cu). */
constexpr int RT3_MAX_LANES = 4;
struct RefLane {
    Scratch q_rays[2], q_hits, q_thr[2], q_pdf[2], q_srays, q_scontrib, q_sbits, q_regen, counts;
    cudaStream_t stream = nullptr;
    cudaEvent_t done = nullptr;
};

struct rt3_context {
    int device = 0;
    int sm_count = 148;
    std::string error;
    uint64_t launches = 0;
    void set_error(const std::string& m) { error = m; }

    /* scene (host mirrors of the small tables + device global buffers) */
    DynBuf vertices, indices;
    std::vector<rt3_geometry_info> h_geometries;
    std::vector<InstanceRec> h_instances;
    std::vector<TextureRec> textures;
    float* d_env = nullptr; uint32_t env_w = 0, env_h = 0;
    float *d_env_func = nullptr, *d_env_marginal = nullptr, *d_env_conditional = nullptr; float env_total = 0.f;   /* RT3_FLAG_NEE tables */
    uint8_t* d_bn = nullptr; uint32_t bn_w = 0, bn_h = 0;
    float* d_srgb_lut = nullptr;
    bool scene_dirty = true;

    /* tables rebuilt by rt3_accel_build */
    Scratch d_geometries, d_seg_first, d_seg_geo, d_seg_inst, d_inst_xform, d_tex_ptr, d_tex_dim;
    std::vector<uint32_t> h_seg_first, h_seg_geo, h_seg_inst;
    uint32_t n_segments = 0, n_triangles = 0;
    rt3m::SceneView view;           /* device pointers */

    /* acceleration structure */
    Scratch bvh_nodes, bvh_tris, d_insts, d_shade_rec, d_shade64;   /* d_insts: rt3b::InstRec table of the two-level structure */
    uint32_t n_wide_nodes = 0, n_leaf_records = 0;
    bool two_level = false;
    bool accel_valid = false;
    rt3_build_stats last_build;

    /* pass scratch */
    RefLane lanes[RT3_MAX_LANES];
    int opt_strips = 0;              /* 0 = default ($RT3_STRIPS or 2); rt3_set_option(RT3_OPT_STRIPS) */
    Scratch px_radiance, px_path, px_failed, d_counts, d_rects, d_untile_rects, d_params;   /* d_untile_rects: un-tiling may run on another stream than the passes */
    cudaEvent_t ev_fork = nullptr;
    /* Stream ordering between calls: all pass / trace calls of one context share the scratch above, so a call on another stream than
     * the previous one first waits for the previous call's work (order_call_begin / order_call_end, api.cu). */
    cudaEvent_t ev_last = nullptr; cudaStream_t last_stream = nullptr; bool have_last = false;
    Scratch f_gbuffer, f_depth, f_light, f_prev, f_rays, f_hits, f_bits, l2_buf;
    DeviceCounters* d_counters = nullptr;
    cudaEvent_t ev[8];
    std::vector<rt3_rect> rects_host; const rt3_rect* rects_dev = nullptr; cudaStream_t rects_stream = nullptr;   /* tile list last uploaded to d_rects */
    std::vector<rt3_rect> urects_host; const rt3_rect* urects_dev = nullptr; cudaStream_t urects_stream = nullptr;   /* ... and to d_untile_rects */
    Scratch d_scatter; std::vector<uint64_t> scatter_host; cudaStream_t scatter_stream = nullptr;                    /* rt3_scatter_tiles_rgba32f: rects + destinations last uploaded */

    /* frames in flight (rt3_frame_submit_host / rt3_frame_wait): FRAMES_IN_FLIGHT = 3 of the reference's swapchain.rs:8 */
    struct FrameSlot { Scratch gb, dp, lt, pv; cudaEvent_t rendered = nullptr, copied = nullptr; bool busy = false; };
    FrameSlot slots[3];
    cudaStream_t s_compute = nullptr, s_copy = nullptr;
    uint64_t next_ticket = 0;

    /* optional per-kernel timing of the traversal launches (rt3_profile_*): event pairs on the launch stream */
    bool prof_on = false;
    std::vector<cudaEvent_t> prof_events;
    size_t prof_used = 0;
    double prof_trace_ms = 0; uint64_t prof_trace_launches = 0;
};

template <class T>
static inline T* ensure(rt3_context* c, Scratch& s, size_t count) {
    size_t bytes = count * sizeof(T);
    if (bytes > s.capacity) {
        if (s.ptr) cudaFree(s.ptr);
        s.ptr = nullptr; s.capacity = 0;
        size_t want = bytes + bytes / 8 + 256;
        RT3_CUDA(c, cudaMalloc(&s.ptr, want));
        s.capacity = want;
    }
    return (T*)s.ptr;
}

static inline unsigned div_up(size_t a, size_t b) { return (unsigned)((a + b - 1) / b); }

#define RT3_LAUNCH(ctx, kernel, grid, block, smem, stream, ...)                         \
    do {                                                                                \
        kernel<<<(grid), (block), (smem), (stream)>>>(__VA_ARGS__);                     \
        (ctx)->launches++;                                                              \
        RT3_CUDA(ctx, cudaGetLastError());                                              \
    } while (0)

/* implemented in build.cu / trace.cu / passes.cu */
void rt3_build_accel(rt3_context* c, uint32_t flags, cudaStream_t st, rt3_build_stats* out);
/* work_counter: 4 x u64 of scratch for the persistent kernel's ray counter (zeroed by the launcher); NULL = the context's own (d_counts) */
void rt3_launch_trace_closest(rt3_context* c, const rt3_ray* rays, const uint32_t* d_count, uint64_t n_max, rt3_hit* hits, unsigned long long* work_counter,
                              cudaStream_t st);
void rt3_launch_trace_any(rt3_context* c, const rt3_ray* rays, const uint32_t* d_count, uint64_t n_max, uint32_t* bits, unsigned long long* work_counter,
                          cudaStream_t st);
void rt3_launch_trace_stats(rt3_context* c, const rt3_ray* rays, uint64_t n, rt3_trace_stats* out);
void rt3_launch_trace_tail_stats(rt3_context* c, const rt3_ray* rays, uint64_t n, rt3_tail_stats* out);
void rt3_launch_gbuffer(rt3_context* c, const rt3_params& p, const PixelMap& m, uint32_t* gbuffer, float* depth, uint32_t* ids, float* bary,
                        cudaStream_t st);
void rt3_launch_reference(rt3_context* c, const rt3_params& p, const PixelMap& m, const uint32_t* gbuffer, const float* depth, const float* prev,
                          float* light, cudaStream_t st);
void rt3_launch_postprocess(rt3_context* c, const rt3_params& p, const PixelMap& m, const float* depth, const float* in, float* out, cudaStream_t st);
void rt3_launch_present(rt3_context* c, const rt3_params& p, const PixelMap& m, const float* depth, const float* in, uint32_t* out, cudaStream_t st);
void rt3_launch_untile(rt3_context* c, const rt3_rect* d_rects, uint32_t n_tiles, uint32_t tile_w, uint32_t tile_h, const float* slab, float* frame,
                       uint32_t frame_w, uint32_t frame_h, cudaStream_t st);
void rt3_launch_probe_trace(rt3_context* c, const rt3_params& p, const float* depth, const uint16_t* directions, const float* prev_atlas, float* atlas,
                            cudaStream_t st);
cudaError_t rt3_static_scratch_setup(int device);
void rt3_launch_probe_trace_static(rt3_context* c, const rt3_params& p, const float* depth, const uint16_t* directions, const float* prev_atlas, float* atlas,
                                   cudaStream_t st);
void rt3_launch_probe_sis(rt3_context* c, const rt3_params& p, const uint32_t* gbuffer, uint16_t* directions, cudaStream_t st);
void rt3_launch_probe_sh(rt3_context* c, const rt3_params& p, const float* atlas, float* sh, cudaStream_t st);
void rt3_launch_probe_interpolate(rt3_context* c, const rt3_params& p, const uint32_t* gbuffer, const float* depth, const float* sh, float* light,
                                  cudaStream_t st);
void rt3_launch_scatter_tiles(rt3_context* c, const rt3_rect* d_rects, const unsigned long long* d_dst, uint32_t n_tiles, uint32_t tile_w, uint32_t tile_h,
                              const float* slab, uint32_t dst_pitch, cudaStream_t st);
double rt3_run_l2_sweep(rt3_context* c, size_t bytes, int iters);
