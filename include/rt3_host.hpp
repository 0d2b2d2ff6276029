/*
 * rt3_host.hpp — C++ host side above the C ABI (include/rt3.h): the reference's render-pass and scene API for the ray-tracing path,
 * restated for a host that links librt3.so directly.  Header only; needs the CUDA runtime for the device memory behind the render
 * graph's images (cudaMalloc / cudaMemcpy), nothing else.
 *
 * The reference's host is Rust (no toolchain in this image); its builder API is mirrored name by name:
 *
 *   RenderWorld global append-only buffers + accel build       /root/reference/src/renderer/world/mod.rs:83-152, vulkan/raytracing.rs:88-148
 *   RenderGraph::image / buffer (named, typed resources)        src/renderer/render_graph/mod.rs:440-483
 *   RayTracingPass::new(rg, name).shader(p).entry(e).constants(&c).read(slot, res).write(slot, res).launch(WorkSize2D)
 *                                                               src/renderer/render_graph/executions.rs:80-121, build.rs:66-193,371-384
 *   ComputePass::new(rg, name) ... .dispatch(WorkSize2D)        executions.rs:21-36
 *   WorkSize2D::{FullScreen, FractionalFullScreen, X, XY}       executions.rs:57-78
 *
 * (`new` is a C++ keyword: the constructors are called `create`.)  Error behaviour: the reference unwraps every result and panics
 * (src/renderer/mod.rs:30); here every failing call throws rt3h::Error carrying the status and rt3_last_error().  Passes execute in
 * declaration order on one CUDA stream, so the barrier synthesis of the Vulkan render graph has no counterpart.
 * The Python mirror raytracer3_b200/host.py has the same shape; tests/cpp/host_demo.cpp drives this one against the oracle.
 */
#pragma once
#include <cuda_runtime_api.h>

#include <cstring>
#include <map>
#include <stdexcept>
#include <string>
#include <vector>

#include "rt3.h"

namespace rt3h {

struct Error : std::runtime_error {
    int status;
    Error(int st, const std::string& what) : std::runtime_error(what), status(st) {}
};

inline void cuda_check(cudaError_t e, const char* what) {
    if (e != cudaSuccess) throw Error(RT3_ERR_CUDA, std::string(what) + ": " + cudaGetErrorString(e));
}

/* rt3_context: RenderWorld + RayTracingContext of the reference (one per CUDA device) */
class Context {
  public:
    explicit Context(int cuda_device = 0) {
        rt3_context* c = nullptr;
        const rt3_status st = rt3_create(cuda_device, &c);
        if (st != RT3_OK) {
            const std::string msg = std::string("rt3_create: ") + rt3_last_error(c);
            if (c) rt3_destroy(c);
            throw Error(st, msg);
        }
        ctx_ = c;
        cuda_check(cudaSetDevice(cuda_device), "cudaSetDevice");
        cuda_check(cudaStreamCreate(&stream_), "cudaStreamCreate");
    }
    ~Context() {
        if (stream_) cudaStreamDestroy(stream_);
        if (ctx_) rt3_destroy(ctx_);
    }
    Context(const Context&) = delete;
    Context& operator=(const Context&) = delete;
    rt3_context* handle() const { return ctx_; }
    cudaStream_t stream() const { return stream_; }
    void check(rt3_status st, const char* call) const {
        if (st != RT3_OK) throw Error(st, std::string(call) + ": " + rt3_last_error(ctx_));
    }
    void synchronize() const { cuda_check(cudaStreamSynchronize(stream_), "cudaStreamSynchronize"); }
    rt3_counters counters(bool reset = false) const {
        rt3_counters c;
        check(rt3_counters_read(ctx_, &c, reset ? 1 : 0), "rt3_counters_read");
        return c;
    }

  private:
    rt3_context* ctx_ = nullptr;
    cudaStream_t stream_ = nullptr;
};

/* world/mod.rs:103-117: global append-only vertex / index / geometry / instance buffers */
class RenderWorld {
  public:
    explicit RenderWorld(Context& ctx) : ctx_(ctx) {}
    uint32_t push_vertices(const rt3_vertex* v, uint32_t n) {
        uint32_t first = 0;
        ctx_.check(rt3_scene_push_vertices(ctx_.handle(), v, n, &first), "rt3_scene_push_vertices");
        return first;
    }
    uint32_t push_indices(const uint32_t* i, uint32_t n) {
        uint32_t first = 0;
        ctx_.check(rt3_scene_push_indices(ctx_.handle(), i, n, &first), "rt3_scene_push_indices");
        return first;
    }
    uint32_t push_geometry(const rt3_geometry_info& g) {
        uint32_t index = 0;
        ctx_.check(rt3_scene_push_geometry(ctx_.handle(), &g, &index), "rt3_scene_push_geometry");
        return index;
    }
    uint32_t push_instance(uint32_t first_geometry, uint32_t geometry_count, const float xform3x4[12]) {   /* InstanceInfo, world/mod.rs:32-37 */
        uint32_t index = 0;
        ctx_.check(rt3_scene_push_instance(ctx_.handle(), first_geometry, geometry_count, xform3x4, &index), "rt3_scene_push_instance");
        return index;
    }
    void set_texture(uint32_t slot, uint32_t w, uint32_t h, const uint8_t* rgba8) {
        ctx_.check(rt3_scene_set_texture(ctx_.handle(), slot, w, h, rgba8), "rt3_scene_set_texture");
    }
    void set_environment(uint32_t w, uint32_t h, const float* rgb) { ctx_.check(rt3_scene_set_environment(ctx_.handle(), w, h, rgb), "rt3_scene_set_environment"); }
    void set_bluenoise(uint32_t w, uint32_t h, const uint8_t* rgba8) { ctx_.check(rt3_scene_set_bluenoise(ctx_.handle(), w, h, rgba8), "rt3_scene_set_bluenoise"); }
    void clear() { ctx_.check(rt3_scene_clear(ctx_.handle()), "rt3_scene_clear"); }
    /* RayTracingContext::create_acceleration_structure (raytracing.rs:88-148), PREFER_FAST_TRACE by default */
    rt3_build_stats create_acceleration_structure(uint32_t flags = RT3_BUILD_FAST_TRACE) {
        rt3_build_stats bs;
        ctx_.check(rt3_accel_build(ctx_.handle(), flags, ctx_.stream(), &bs), "rt3_accel_build");
        return bs;
    }
    Context& context() { return ctx_; }

  private:
    Context& ctx_;
};

/* executions.rs:57-78 */
struct WorkSize2D {
    enum Kind { kFullScreen, kFractionalFullScreen, kX, kXY } kind;
    uint32_t x, y;
    static WorkSize2D FullScreen() { return {kFullScreen, 0, 0}; }
    static WorkSize2D FractionalFullScreen(uint32_t x, uint32_t y) { return {kFractionalFullScreen, x, y}; }
    static WorkSize2D X(uint32_t x) { return {kX, x, 1}; }
    static WorkSize2D XY(uint32_t x, uint32_t y) { return {kXY, x, y}; }
    void size(uint32_t w, uint32_t h, uint32_t& ox, uint32_t& oy) const {
        if (kind == kFullScreen) { ox = w; oy = h; }
        else if (kind == kFractionalFullScreen) { ox = (w + x - 1) / x; oy = (h + y - 1) / y; }
        else { ox = x; oy = y; }
    }
};

enum class Format { R32G32B32A32_UINT, R32_SFLOAT, R32G32B32A32_SFLOAT, R32G32_UINT, R32G32_SFLOAT };
inline size_t format_bytes(Format f) {
    switch (f) {
        case Format::R32_SFLOAT: return 4;
        case Format::R32G32_UINT: case Format::R32G32_SFLOAT: return 8;
        default: return 16;
    }
}

/* Named resources (render_graph/mod.rs:440-483): window-sized images and free-size buffers, zero-initialised device memory */
class RenderGraph {
  public:
    RenderGraph(RenderWorld& world, uint32_t width, uint32_t height) : world_(world), w_(width), h_(height) {}
    ~RenderGraph() { for (auto& kv : res_) cudaFree(kv.second.ptr); }
    RenderGraph(const RenderGraph&) = delete;
    RenderGraph& operator=(const RenderGraph&) = delete;
    const std::string& image(Format f, const std::string& name) { return buffer(name, (size_t)w_ * h_ * format_bytes(f)); }
    const std::string& buffer(const std::string& name, size_t bytes) {
        auto it = res_.find(name);
        if (it == res_.end() || it->second.bytes != bytes) {
            if (it != res_.end()) cudaFree(it->second.ptr);
            Resource r;
            r.bytes = bytes;
            cuda_check(cudaMalloc(&r.ptr, bytes ? bytes : 1), "cudaMalloc");
            cuda_check(cudaMemsetAsync(r.ptr, 0, bytes, world_.context().stream()), "cudaMemsetAsync");
            res_[name] = r;
            it = res_.find(name);
        }
        return it->first;
    }
    void* device_ptr(const std::string& name) const {
        auto it = res_.find(name);
        if (it == res_.end()) throw Error(RT3_ERR_INVALID, "render graph has no resource '" + name + "'");
        return it->second.ptr;
    }
    size_t bytes(const std::string& name) const { device_ptr(name); return res_.at(name).bytes; }
    void download(const std::string& name, void* host) const {      /* waits for the passes recorded so far */
        cuda_check(cudaMemcpyAsync(host, device_ptr(name), bytes(name), cudaMemcpyDeviceToHost, world_.context().stream()), "cudaMemcpyAsync");
        world_.context().synchronize();
    }
    uint32_t width() const { return w_; }
    uint32_t height() const { return h_; }
    RenderWorld& world() { return world_; }
    uint64_t frame_number = 0;

  private:
    struct Resource { void* ptr = nullptr; size_t bytes = 0; };
    RenderWorld& world_;
    uint32_t w_, h_;
    std::map<std::string, Resource> res_;
};

/* build.rs:66-193: the builder verbs shared by the pass kinds */
template <class Derived>
class PassBuilder {
  public:
    Derived& shader(const std::string& path) { shader_ = path; return self(); }
    Derived& entry(const std::string& name) { entry_ = name; return self(); }
    Derived& constants(const rt3_params& gconst) { constants_ = gconst; has_constants_ = true; return self(); }     /* copied, as build.rs:66-94 does */
    Derived& read(const std::string& slot, const std::string& resource) { bind_[slot] = resource; return self(); }
    Derived& write(const std::string& slot, const std::string& resource) { return read(slot, resource); }
    Derived& read_write(const std::string& slot, const std::string& resource) { return read(slot, resource); }

  protected:
    PassBuilder(RenderGraph& rg, const std::string& name) : rg_(rg), name_(name) {}
    Derived& self() { return static_cast<Derived&>(*this); }
    void* ptr(const std::string& slot, bool required = true) const {
        auto it = bind_.find(slot);
        if (it == bind_.end()) {
            if (required) throw Error(RT3_ERR_INVALID, "pass '" + name_ + "' (" + shader_ + "): resource slot '" + slot + "' not bound");
            return nullptr;
        }
        return rg_.device_ptr(it->second);
    }
    const rt3_params* params() const {
        if (!has_constants_) throw Error(RT3_ERR_INVALID, "pass '" + name_ + "': constants(&gconst) not set");
        return &constants_;
    }
    RenderGraph& rg_;
    std::string name_, shader_, entry_ = "main";
    rt3_params constants_;
    bool has_constants_ = false;
    std::map<std::string, std::string> bind_;
};

/* executions.rs:80-121: RayTracingPass ... .launch(WorkSize2D) -> vkCmdTraceRaysKHR(x, y, 1) */
class RayTracingPass : public PassBuilder<RayTracingPass> {
  public:
    static RayTracingPass create(RenderGraph& rg, const std::string& name) { return RayTracingPass(rg, name); }
    const std::string& launch(WorkSize2D work = WorkSize2D::FullScreen()) {
        Context& ctx = rg_.world().context();
        uint32_t x, y;
        work.size(rg_.width(), rg_.height(), x, y);
        const rt3_rect rect = {0, 0, x, y};
        if (shader_ == "gbuffer")                                   /* shaders/old/gbuffer.slang */
            ctx.check(rt3_pass_gbuffer(ctx.handle(), params(), rect, (uint32_t*)ptr("gbuffer"), (float*)ptr("gbuffer_depth"), (uint32_t*)ptr("ids", false),
                                       (float*)ptr("bary", false), ctx.stream()), "rt3_pass_gbuffer");
        else if (shader_ == "refrence_mode")                        /* shaders/old/refrence_mode.slang (the reference's spelling) */
            ctx.check(rt3_pass_reference(ctx.handle(), params(), rect, (const uint32_t*)ptr("gbuffer"), (const float*)ptr("gbuffer_depth"),
                                         (const float*)ptr("PrevLight", false), (float*)ptr("Light"), ctx.stream()), "rt3_pass_reference");
        else if (shader_ == "trace_probes")                         /* shaders/old/trace_probes.slang, launched over the probe atlas */
            ctx.check(rt3_pass_probe_trace(ctx.handle(), params(), (const uint32_t*)ptr("gbuffer"), (const float*)ptr("gbuffer_depth"),
                                           (const uint16_t*)ptr("directions", false), (const float*)ptr("prev_probe_atlas", false),
                                           (float*)ptr("probe_atlas"), ctx.stream()), "rt3_pass_probe_trace");
        else
            throw Error(RT3_ERR_INVALID, "unknown ray-tracing shader '" + shader_ + "' (known: gbuffer, refrence_mode, trace_probes)");
        return name_;
    }

  private:
    RayTracingPass(RenderGraph& rg, const std::string& name) : PassBuilder(rg, name) {}
};

/* executions.rs:21-36: ComputePass ... .dispatch(WorkSize2D) */
class ComputePass : public PassBuilder<ComputePass> {
  public:
    static ComputePass create(RenderGraph& rg, const std::string& name) { return ComputePass(rg, name); }
    const std::string& dispatch(WorkSize2D work = WorkSize2D::FullScreen()) {
        Context& ctx = rg_.world().context();
        uint32_t x, y;
        work.size(rg_.width(), rg_.height(), x, y);
        const rt3_rect rect = {0, 0, x, y};
        if (shader_ == "postprocess")                               /* shaders/old/postprocess.slang */
            ctx.check(rt3_pass_postprocess(ctx.handle(), params(), rect, (const float*)ptr("Depth"), (const float*)ptr("In"), (float*)ptr("Out"), ctx.stream()),
                      "rt3_pass_postprocess");
        else if (shader_ == "spherical_harmonic_conversion")
            ctx.check(rt3_pass_probe_sh(ctx.handle(), params(), (const float*)ptr("probe_atlas"), (float*)ptr("out"), ctx.stream()), "rt3_pass_probe_sh");
        else if (shader_ == "present")                              /* postprocess + the swapchain blit's conversion to B8G8R8A8_UNORM */
            ctx.check(rt3_pass_present_bgra8(ctx.handle(), params(), rect, (const float*)ptr("Depth"), (const float*)ptr("In"), (uint32_t*)ptr("Out"),
                                             ctx.stream()), "rt3_pass_present_bgra8");
        else if (shader_ == "structured_importance_sampling")       /* shaders/old/structured_importance_sampling.slang, 8x8 threads per probe */
            ctx.check(rt3_pass_probe_sis(ctx.handle(), params(), (const uint32_t*)ptr("gbuffer"), (uint16_t*)ptr("out"), ctx.stream()), "rt3_pass_probe_sis");
        else if (shader_ == "interpolate_probes")
            ctx.check(rt3_pass_probe_interpolate(ctx.handle(), params(), (const uint32_t*)ptr("gbuffer"), (const float*)ptr("gbuffer_depth"),
                                                 (const float*)ptr("sh_coeficents"), (float*)ptr("Light"), ctx.stream()), "rt3_pass_probe_interpolate");
        else
            throw Error(RT3_ERR_INVALID, "unknown compute shader '" + shader_ + "' (known: postprocess, structured_importance_sampling, spherical_harmonic_conversion, interpolate_probes)");
        return name_;
    }

  private:
    ComputePass(RenderGraph& rg, const std::string& name) : PassBuilder(rg, name) {}
};

/* The frame of the reference's old graph: gbuffer -> refrence_mode (-> postprocess), Light / PrevLight ping-pong */
class Renderer {
  public:
    Renderer(RenderWorld& world, uint32_t width, uint32_t height) : rg(world, width, height) {
        rg.image(Format::R32G32B32A32_UINT, "gbuffer");
        rg.image(Format::R32_SFLOAT, "gbuffer_depth");
        rg.image(Format::R32G32B32A32_SFLOAT, "light0");
        rg.image(Format::R32G32B32A32_SFLOAT, "light1");
        rg.image(Format::R32G32B32A32_SFLOAT, "out");
    }
    /* returns the name of the Light image this frame wrote */
    std::string frame(const rt3_params& gconst, bool postprocess = false) {
        const std::string light = cur_ ? "light1" : "light0", prev = cur_ ? "light0" : "light1";
        RayTracingPass::create(rg, "gbuffer").shader("gbuffer").constants(gconst).write("gbuffer", "gbuffer").write("gbuffer_depth", "gbuffer_depth")
            .launch(WorkSize2D::FullScreen());
        RayTracingPass::create(rg, "reference").shader("refrence_mode").constants(gconst).read("gbuffer", "gbuffer").read("gbuffer_depth", "gbuffer_depth")
            .read("PrevLight", prev).write("Light", light).launch(WorkSize2D::FullScreen());
        if (postprocess)
            ComputePass::create(rg, "postprocess").shader("postprocess").constants(gconst).read("Depth", "gbuffer_depth").read("In", light).write("Out", "out")
                .dispatch(WorkSize2D::FullScreen());
        cur_ ^= 1;
        rg.frame_number++;
        return light;
    }
    RenderGraph rg;

  private:
    int cur_ = 0;
};

}  // namespace rt3h
