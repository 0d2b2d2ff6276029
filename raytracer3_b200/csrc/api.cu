/*
 * api.cu — the extern "C" boundary of librt3.so (declared in include/rt3.h, which cites the reference
 * interface each entry point replaces).  No CPU fallback: rt3_create fails without a CUDA device.
 */
#include "common.cuh"

#include <algorithm>
#include <mutex>

namespace {

thread_local std::string g_create_error;

template <class F>
rt3_status guarded(rt3_context* c, F f) {
    if (!c) return RT3_ERR_INVALID;
    try {
        cudaError_t e = cudaSetDevice(c->device);
        if (e != cudaSuccess) { c->set_error(std::string("cudaSetDevice: ") + cudaGetErrorString(e)); return RT3_ERR_CUDA; }
        f();
        return RT3_OK;
    } catch (const rt3_error_exit& x) {
        return x.code;
    } catch (const std::bad_alloc&) {
        c->set_error("host allocation failed");
        return RT3_ERR_OOM;
    } catch (...) {
        c->set_error("unexpected exception");
        return RT3_ERR_CUDA;
    }
}

void dyn_push(rt3_context* c, DynBuf& b, const void* host, size_t bytes) {
    if (b.size + bytes > b.capacity) {
        size_t cap = b.capacity ? b.capacity : 65536;   /* initial capacities of world/mod.rs:23-30 */
        while (cap < b.size + bytes) cap *= 2;
        void* np = nullptr;
        RT3_CUDA(c, cudaMalloc(&np, cap));
        if (b.size) RT3_CUDA(c, cudaMemcpy(np, b.ptr, b.size, cudaMemcpyDeviceToDevice));
        if (b.ptr) cudaFree(b.ptr);
        b.ptr = np; b.capacity = cap;
    }
    RT3_CUDA(c, cudaMemcpy((char*)b.ptr + b.size, host, bytes, cudaMemcpyHostToDevice));
    b.size += bytes;
}

template <class T>
T* upload(rt3_context* c, Scratch& s, const std::vector<T>& v) {
    T* d = ensure<T>(c, s, v.size() ? v.size() : 1);
    if (!v.empty()) RT3_CUDA(c, cudaMemcpy(d, v.data(), v.size() * sizeof(T), cudaMemcpyHostToDevice));
    return d;
}

static void free_environment(rt3_context* c) {
    float** ps[] = {&c->d_env, &c->d_env_func, &c->d_env_marginal, &c->d_env_conditional};
    for (float** p : ps) { if (*p) cudaFree(*p); *p = nullptr; }
    c->env_w = c->env_h = 0; c->env_total = 0.f;
}
static void publish_environment(rt3_context* c) {
    rt3m::SceneView& v = c->view;
    v.env = c->d_env; v.env_w = c->env_w; v.env_h = c->env_h;
    v.env_func = c->d_env_func; v.env_marginal = c->d_env_marginal; v.env_conditional = c->d_env_conditional; v.env_total = c->env_total;
}

/* Segment table + SceneView (device pointers) from the current scene state. */
void refresh_scene_tables(rt3_context* c) {
    const uint32_t ng = (uint32_t)c->h_geometries.size();
    const uint32_t n_vertices = (uint32_t)(c->vertices.size / sizeof(rt3_vertex));
    const uint32_t n_indices = (uint32_t)(c->indices.size / sizeof(uint32_t));
    std::vector<InstanceRec> inst = c->h_instances;
    if (inst.empty()) {
        InstanceRec r; r.first_geometry = 0; r.geometry_count = ng;
        const float id[12] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0};
        memcpy(r.xform, id, sizeof(id));
        inst.push_back(r);
    }
    for (uint32_t g = 0; g < ng; ++g) {
        const rt3_geometry_info& gi = c->h_geometries[g];
        if (gi.index_count % 3) RT3_FAIL(c, RT3_ERR_INVALID, "geometry index_count is not a multiple of 3");
        if ((uint64_t)gi.index_offset + gi.index_count > n_indices) RT3_FAIL(c, RT3_ERR_INVALID, "geometry index range exceeds the index buffer");
        if (gi.vertex_offset > n_vertices) RT3_FAIL(c, RT3_ERR_INVALID, "geometry vertex_offset exceeds the vertex buffer");
    }
    c->h_seg_first.clear(); c->h_seg_geo.clear(); c->h_seg_inst.clear();
    std::vector<float> xf;
    uint64_t total = 0;
    for (uint32_t i = 0; i < inst.size(); ++i) {
        if ((uint64_t)inst[i].first_geometry + inst[i].geometry_count > ng) RT3_FAIL(c, RT3_ERR_INVALID, "instance geometry range out of bounds");
        for (uint32_t g = inst[i].first_geometry; g < inst[i].first_geometry + inst[i].geometry_count; ++g) {
            c->h_seg_first.push_back((uint32_t)total); c->h_seg_geo.push_back(g); c->h_seg_inst.push_back(i);
            total += c->h_geometries[g].index_count / 3;
        }
        xf.insert(xf.end(), inst[i].xform, inst[i].xform + 12);
    }
    if (total >= 0x7FFFFFFFull) RT3_FAIL(c, RT3_ERR_INVALID, "too many triangles (limit 2^31 - 1)");
    c->h_seg_first.push_back((uint32_t)total);
    c->n_segments = (uint32_t)c->h_seg_geo.size();
    c->n_triangles = (uint32_t)total;
    /* index values must stay inside the vertex buffer: checked on the device would cost a pass; the
     * reference does not check either (hit_logic.slang:10-19), so only the ranges above are validated. */
    rt3m::SceneView& v = c->view;
    v.vertices = (const rt3_vertex*)c->vertices.ptr;
    v.indices = (const uint32_t*)c->indices.ptr;
    v.geometries = upload(c, c->d_geometries, c->h_geometries);
    v.seg_first_prim = upload(c, c->d_seg_first, c->h_seg_first);
    v.seg_geometry = upload(c, c->d_seg_geo, c->h_seg_geo);
    v.seg_instance = upload(c, c->d_seg_inst, c->h_seg_inst);
    v.inst_xform = upload(c, c->d_inst_xform, xf);
    v.n_segments = c->n_segments;
    std::vector<const uint8_t*> tp; std::vector<uint32_t> td;
    for (const TextureRec& t : c->textures) { tp.push_back(t.data); td.push_back(t.w); td.push_back(t.h); }
    v.tex_data = (const uint8_t* const*)upload(c, c->d_tex_ptr, tp);
    v.tex_dim = upload(c, c->d_tex_dim, td);
    v.n_textures = (uint32_t)c->textures.size();
    v.srgb_lut = c->d_srgb_lut;
    publish_environment(c);
    v.bn = c->d_bn; v.bn_w = c->bn_w; v.bn_h = c->bn_h;
    c->scene_dirty = false;
}

void require_accel(rt3_context* c) {
    if (!c->accel_valid) RT3_FAIL(c, RT3_ERR_STATE, "rt3_accel_build has not been called since the scene changed");
}

PixelMap single_rect_map(rt3_context* c, const rt3_params* p, rt3_rect r) {
    const uint32_t W = (uint32_t)p->window_size[0], H = (uint32_t)p->window_size[1];
    if (W == 0 || H == 0 || (uint64_t)r.x0 + r.w > W || (uint64_t)r.y0 + r.h > H) RT3_FAIL(c, RT3_ERR_INVALID, "rect outside the frame given by window_size");
    PixelMap m;
    m.single = r; m.rects = nullptr; m.n_tiles = 1; m.tile_w = r.w; m.tile_h = r.h;
    m.bw = (r.w + 7) / 8; m.bh = (r.h + 3) / 4; m.pitch = W; m.slab = 0;
    uint64_t work = (uint64_t)m.bw * m.bh * 32u;
    if (work >= 0x7FFFFFFFull) RT3_FAIL(c, RT3_ERR_INVALID, "rect too large");
    m.work = (r.w && r.h) ? (uint32_t)work : 0u;
    finish_pixel_map(m);
    return m;
}

PixelMap tiles_map(rt3_context* c, const rt3_params* p, const rt3_rect* tiles, uint32_t n, cudaStream_t st) {
    const uint32_t W = (uint32_t)p->window_size[0], H = (uint32_t)p->window_size[1];
    PixelMap m;
    memset(&m, 0, sizeof(m));
    uint32_t tw = 0, th = 0;
    for (uint32_t i = 0; i < n; ++i) {
        if ((uint64_t)tiles[i].x0 + tiles[i].w > W || (uint64_t)tiles[i].y0 + tiles[i].h > H) RT3_FAIL(c, RT3_ERR_INVALID, "tile outside the frame");
        tw = std::max(tw, tiles[i].w); th = std::max(th, tiles[i].h);
    }
    rt3_rect* d = ensure<rt3_rect>(c, c->d_rects, n ? n : 1);
    /* the same list on the same stream as the last upload (every frame of a tiled renderer): the device copy is still valid, skip the
     * pageable H2D copy, which costs a staging round trip per pass call */
    const bool same = n && d == c->rects_dev && st == c->rects_stream && c->rects_host.size() == n &&
                      memcmp(c->rects_host.data(), tiles, n * sizeof(rt3_rect)) == 0;
    if (n && !same) {
        RT3_CUDA(c, cudaMemcpyAsync(d, tiles, n * sizeof(rt3_rect), cudaMemcpyHostToDevice, st));
        c->rects_host.assign(tiles, tiles + n); c->rects_dev = d; c->rects_stream = st;
    }
    m.rects = d; m.n_tiles = n; m.tile_w = tw; m.tile_h = th;
    m.bw = (tw + 7) / 8; m.bh = (th + 3) / 4; m.pitch = W; m.slab = 1;
    if (n == 1) m.single = tiles[0];
    uint64_t work = (uint64_t)n * m.bw * m.bh * 32u;
    if (work >= 0x7FFFFFFFull) RT3_FAIL(c, RT3_ERR_INVALID, "tile list too large");
    m.work = (uint32_t)work;
    finish_pixel_map(m);
    return m;
}

/* All pass / trace calls of one context share its scratch (wavefront queues, per-pixel path state, counters, tile lists), so they must
 * execute in call order even when the caller records them on different streams: a call whose stream differs from the previous call's first
 * waits (on the device, cudaStreamWaitEvent) for the event recorded at the end of the previous call.  Same stream: stream order suffices. */
struct CallOrder {
    rt3_context* c; cudaStream_t st;
    CallOrder(rt3_context* c_, cudaStream_t st_) : c(c_), st(st_) {
        if (c->have_last && c->last_stream != st) RT3_CUDA(c, cudaStreamWaitEvent(st, c->ev_last, 0));
    }
    ~CallOrder() {
        if (cudaEventRecord(c->ev_last, st) == cudaSuccess) { c->last_stream = st; c->have_last = true; }
    }
};

void check_params(rt3_context* c, const rt3_params* p) {
    if (!p) RT3_FAIL(c, RT3_ERR_INVALID, "params is NULL");
    if (p->samples == 0) RT3_FAIL(c, RT3_ERR_INVALID, "samples must be >= 1");
    if (p->samples >= (1u << 24) || p->bounces > 255u) RT3_FAIL(c, RT3_ERR_INVALID, "samples must be < 2^24 and bounces <= 255 (per-pixel path state is one 32-bit word)");
    if ((p->flags & (RT3_FLAG_ENV | RT3_FLAG_NEE)) && !c->d_env) RT3_FAIL(c, RT3_ERR_STATE, "RT3_FLAG_ENV / RT3_FLAG_NEE without rt3_scene_set_environment");
    if ((p->flags & RT3_FLAG_NEE) && !(p->flags & RT3_FLAG_ENV)) RT3_FAIL(c, RT3_ERR_INVALID, "RT3_FLAG_NEE requires RT3_FLAG_ENV");
    if ((p->flags & RT3_FLAG_BLUENOISE) && !c->d_bn) RT3_FAIL(c, RT3_ERR_STATE, "RT3_FLAG_BLUENOISE without rt3_scene_set_bluenoise");
}

}  // namespace

extern "C" {

uint32_t rt3_abi_version(void) { return RT3_ABI_VERSION; }
size_t rt3_sizeof(int which) {
    switch (which) {
        case 0: return sizeof(rt3_params);
        case 1: return sizeof(rt3_vertex);
        case 2: return sizeof(rt3_geometry_info);
        case 3: return sizeof(rt3_ray);
        case 4: return sizeof(rt3_hit);
        case 5: return sizeof(rt3_build_stats);
        case 6: return sizeof(rt3_counters);
        default: return 0;
    }
}
const char* rt3_last_error(const rt3_context* c) { return c ? c->error.c_str() : g_create_error.c_str(); }

rt3_status rt3_create(int cuda_device, rt3_context** out) {
    if (!out) return RT3_ERR_INVALID;
    *out = nullptr;
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n == 0) {
        g_create_error = std::string("no CUDA device: ") + (e != cudaSuccess ? cudaGetErrorString(e) : "device count is 0") +
                         " (librt3 has no CPU fallback)";
        return RT3_ERR_NO_DEVICE;
    }
    if (cuda_device < 0 || cuda_device >= n) { g_create_error = "cuda_device out of range"; return RT3_ERR_INVALID; }
    rt3_context* c = new rt3_context();
    c->device = cuda_device;
    rt3_status s = guarded(c, [&]() {
        cudaDeviceProp prop;
        RT3_CUDA(c, cudaGetDeviceProperties(&prop, cuda_device));
        c->sm_count = prop.multiProcessorCount;
        cudaMemPool_t pool;                                   /* keep build temporaries cached in the stream-ordered pool */
        RT3_CUDA(c, cudaDeviceGetDefaultMemPool(&pool, cuda_device));
        uint64_t keep_all = UINT64_MAX;
        RT3_CUDA(c, cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep_all));
        for (int i = 0; i < 8; ++i) RT3_CUDA(c, cudaEventCreate(&c->ev[i]));
        RT3_CUDA(c, cudaEventCreateWithFlags(&c->ev_fork, cudaEventDisableTiming));
        RT3_CUDA(c, cudaEventCreateWithFlags(&c->ev_last, cudaEventDisableTiming));
        (void)rt3_static_scratch_setup(cuda_device);     /* address of the static scratch of rt3_pass_probe_trace_static; a failure is reported by that call */
        RT3_CUDA(c, cudaMalloc(&c->d_counters, sizeof(DeviceCounters)));
        RT3_CUDA(c, cudaMemset(c->d_counters, 0, sizeof(DeviceCounters)));
        float lut[256];
        rt3m::build_srgb_lut(lut);
        RT3_CUDA(c, cudaMalloc(&c->d_srgb_lut, sizeof(lut)));
        RT3_CUDA(c, cudaMemcpy(c->d_srgb_lut, lut, sizeof(lut), cudaMemcpyHostToDevice));
        ensure<unsigned long long>(c, c->d_counts, 64);
        RT3_CUDA(c, cudaMemset(c->d_counts.ptr, 0, 64 * 8));
        RT3_CUDA(c, cudaStreamCreateWithFlags(&c->s_compute, cudaStreamNonBlocking));
        RT3_CUDA(c, cudaStreamCreateWithFlags(&c->s_copy, cudaStreamNonBlocking));
        for (auto& fs : c->slots) {
            RT3_CUDA(c, cudaEventCreateWithFlags(&fs.rendered, cudaEventDisableTiming));
            RT3_CUDA(c, cudaEventCreateWithFlags(&fs.copied, cudaEventDisableTiming));
        }
    });
    if (s != RT3_OK) { g_create_error = c->error; delete c; return s; }
    *out = c;
    return RT3_OK;
}

void rt3_destroy(rt3_context* c) {
    if (!c) return;
    cudaSetDevice(c->device);
    cudaDeviceSynchronize();
    c->vertices.release(); c->indices.release();
    for (TextureRec& t : c->textures) if (t.data) cudaFree(t.data);
    free_environment(c);
    if (c->d_bn) cudaFree(c->d_bn);
    if (c->d_srgb_lut) cudaFree(c->d_srgb_lut);
    if (c->d_counters) cudaFree(c->d_counters);
    Scratch* all[] = {&c->d_geometries, &c->d_seg_first, &c->d_seg_geo, &c->d_seg_inst, &c->d_inst_xform, &c->d_tex_ptr, &c->d_tex_dim, &c->bvh_nodes,
                      &c->bvh_tris, &c->d_insts, &c->d_shade_rec, &c->d_shade64, &c->px_radiance, &c->px_path, &c->px_failed, &c->d_counts, &c->d_rects, &c->d_untile_rects, &c->d_params, &c->f_gbuffer, &c->f_depth, &c->f_light, &c->f_prev, &c->f_rays, &c->f_hits,
                      &c->f_bits, &c->l2_buf};
    for (Scratch* s : all) s->release();
    for (RefLane& ln : c->lanes) {
        Scratch* q[] = {&ln.q_rays[0], &ln.q_rays[1], &ln.q_hits, &ln.q_thr[0], &ln.q_thr[1], &ln.q_pdf[0], &ln.q_pdf[1], &ln.q_srays, &ln.q_scontrib, &ln.q_sbits, &ln.q_regen, &ln.counts};
        for (Scratch* s : q) s->release();
        if (ln.stream) cudaStreamDestroy(ln.stream);
        if (ln.done) cudaEventDestroy(ln.done);
    }
    if (c->ev_fork) cudaEventDestroy(c->ev_fork);
    if (c->ev_last) cudaEventDestroy(c->ev_last);
    for (auto& fs : c->slots) {
        fs.gb.release(); fs.dp.release(); fs.lt.release(); fs.pv.release();
        if (fs.rendered) cudaEventDestroy(fs.rendered);
        if (fs.copied) cudaEventDestroy(fs.copied);
    }
    if (c->s_compute) cudaStreamDestroy(c->s_compute);
    if (c->s_copy) cudaStreamDestroy(c->s_copy);
    for (int i = 0; i < 8; ++i) cudaEventDestroy(c->ev[i]);
    for (cudaEvent_t e : c->prof_events) cudaEventDestroy(e);
    delete c;
}

/* ---- scene ------------------------------------------------------------------------------------ */
rt3_status rt3_scene_clear(rt3_context* c) {
    return guarded(c, [&]() {
        RT3_CUDA(c, cudaDeviceSynchronize());
        c->vertices.size = 0; c->indices.size = 0;
        c->h_geometries.clear(); c->h_instances.clear();
        for (TextureRec& t : c->textures) if (t.data) cudaFree(t.data);
        c->textures.clear();
        free_environment(c);
        if (c->d_bn) { cudaFree(c->d_bn); c->d_bn = nullptr; } c->bn_w = c->bn_h = 0;
        c->scene_dirty = true; c->accel_valid = false; c->n_triangles = 0;
    });
}
rt3_status rt3_scene_push_vertices(rt3_context* c, const rt3_vertex* host, uint32_t n, uint32_t* out_first) {
    return guarded(c, [&]() {
        if (n && !host) RT3_FAIL(c, RT3_ERR_INVALID, "vertices is NULL");
        if (out_first) *out_first = (uint32_t)(c->vertices.size / sizeof(rt3_vertex));
        if (n) dyn_push(c, c->vertices, host, (size_t)n * sizeof(rt3_vertex));
        c->scene_dirty = true; c->accel_valid = false;
    });
}
rt3_status rt3_scene_push_indices(rt3_context* c, const uint32_t* host, uint32_t n, uint32_t* out_first) {
    return guarded(c, [&]() {
        if (n && !host) RT3_FAIL(c, RT3_ERR_INVALID, "indices is NULL");
        if (out_first) *out_first = (uint32_t)(c->indices.size / sizeof(uint32_t));
        if (n) dyn_push(c, c->indices, host, (size_t)n * sizeof(uint32_t));
        c->scene_dirty = true; c->accel_valid = false;
    });
}
rt3_status rt3_scene_push_geometry(rt3_context* c, const rt3_geometry_info* host, uint32_t* out_geometry_index) {
    return guarded(c, [&]() {
        if (!host) RT3_FAIL(c, RT3_ERR_INVALID, "geometry is NULL");
        if (out_geometry_index) *out_geometry_index = (uint32_t)c->h_geometries.size();
        c->h_geometries.push_back(*host);
        c->scene_dirty = true; c->accel_valid = false;
    });
}
rt3_status rt3_scene_push_instance(rt3_context* c, uint32_t first_geometry, uint32_t geometry_count, const float xform3x4[12], uint32_t* out_instance_index) {
    return guarded(c, [&]() {
        if (!xform3x4) RT3_FAIL(c, RT3_ERR_INVALID, "xform is NULL");
        InstanceRec r; r.first_geometry = first_geometry; r.geometry_count = geometry_count;
        memcpy(r.xform, xform3x4, sizeof(r.xform));
        if (out_instance_index) *out_instance_index = (uint32_t)c->h_instances.size();
        c->h_instances.push_back(r);
        c->scene_dirty = true; c->accel_valid = false;
    });
}
rt3_status rt3_scene_set_texture(rt3_context* c, uint32_t slot, uint32_t w, uint32_t h, const uint8_t* rgba8) {
    return guarded(c, [&]() {
        if (!rgba8 || !w || !h || slot > 4095) RT3_FAIL(c, RT3_ERR_INVALID, "bad texture arguments");
        if (c->textures.size() <= slot) c->textures.resize(slot + 1);
        TextureRec& t = c->textures[slot];
        if (t.data) { RT3_CUDA(c, cudaDeviceSynchronize()); cudaFree(t.data); t.data = nullptr; }
        RT3_CUDA(c, cudaMalloc((void**)&t.data, 4 * (size_t)w * h));
        RT3_CUDA(c, cudaMemcpy(t.data, rgba8, 4 * (size_t)w * h, cudaMemcpyHostToDevice));
        t.w = w; t.h = h;
        c->scene_dirty = true;
    });
}
rt3_status rt3_scene_set_environment(rt3_context* c, uint32_t w, uint32_t h, const float* rgb) {
    return guarded(c, [&]() {
        if (!rgb || !w || !h) RT3_FAIL(c, RT3_ERR_INVALID, "bad environment arguments");
        RT3_CUDA(c, cudaDeviceSynchronize());
        free_environment(c);
        const size_t px = (size_t)w * h;
        RT3_CUDA(c, cudaMalloc((void**)&c->d_env, 12 * px));
        RT3_CUDA(c, cudaMemcpy(c->d_env, rgb, 12 * px, cudaMemcpyHostToDevice));
        /* importance tables for RT3_FLAG_NEE: host-side build (same function as the oracle), uploaded once */
        std::vector<float> func(px), marg(h + 1), cond((size_t)h * (w + 1));
        rt3m::build_env_tables(rgb, w, h, func.data(), marg.data(), cond.data(), &c->env_total);
        RT3_CUDA(c, cudaMalloc((void**)&c->d_env_func, 4 * func.size()));
        RT3_CUDA(c, cudaMalloc((void**)&c->d_env_marginal, 4 * marg.size()));
        RT3_CUDA(c, cudaMalloc((void**)&c->d_env_conditional, 4 * cond.size()));
        RT3_CUDA(c, cudaMemcpy(c->d_env_func, func.data(), 4 * func.size(), cudaMemcpyHostToDevice));
        RT3_CUDA(c, cudaMemcpy(c->d_env_marginal, marg.data(), 4 * marg.size(), cudaMemcpyHostToDevice));
        RT3_CUDA(c, cudaMemcpy(c->d_env_conditional, cond.data(), 4 * cond.size(), cudaMemcpyHostToDevice));
        c->env_w = w; c->env_h = h;
        publish_environment(c);
    });
}
rt3_status rt3_scene_set_bluenoise(rt3_context* c, uint32_t w, uint32_t h, const uint8_t* rgba8) {
    return guarded(c, [&]() {
        if (!rgba8 || !w || !h || (w & (w - 1)) || (h & (h - 1))) RT3_FAIL(c, RT3_ERR_INVALID, "blue-noise tile must have power-of-two size");
        if (c->d_bn) { RT3_CUDA(c, cudaDeviceSynchronize()); cudaFree(c->d_bn); c->d_bn = nullptr; }
        RT3_CUDA(c, cudaMalloc((void**)&c->d_bn, 4 * (size_t)w * h));
        RT3_CUDA(c, cudaMemcpy(c->d_bn, rgba8, 4 * (size_t)w * h, cudaMemcpyHostToDevice));
        c->bn_w = w; c->bn_h = h;
        c->view.bn = c->d_bn; c->view.bn_w = w; c->view.bn_h = h;
    });
}

/* ---- accel ------------------------------------------------------------------------------------ */
rt3_status rt3_accel_build(rt3_context* c, uint32_t flags, void* cuda_stream, rt3_build_stats* out) {
    return guarded(c, [&]() {
        refresh_scene_tables(c);
        rt3_build_accel(c, flags, (cudaStream_t)cuda_stream, out);
    });
}
rt3_status rt3_resolve_prim(rt3_context* c, uint32_t prim, uint32_t* instance, uint32_t* geometry, uint32_t* primitive) {
    return guarded(c, [&]() {
        if (c->scene_dirty || prim >= c->n_triangles) RT3_FAIL(c, RT3_ERR_INVALID, "prim out of range (or scene changed since rt3_accel_build)");
        uint32_t seg = (uint32_t)(std::upper_bound(c->h_seg_first.begin(), c->h_seg_first.end(), prim) - c->h_seg_first.begin()) - 1;
        if (instance) *instance = c->h_seg_inst[seg];
        if (geometry) *geometry = c->h_seg_geo[seg];
        if (primitive) *primitive = prim - c->h_seg_first[seg];
    });
}

/* ---- passes ----------------------------------------------------------------------------------- */
rt3_status rt3_pass_gbuffer(rt3_context* c, const rt3_params* p, rt3_rect r, uint32_t* gbuffer, float* depth, uint32_t* ids, float* bary, void* st) {
    return guarded(c, [&]() {
        check_params(c, p); require_accel(c);
        if (!gbuffer || !depth) RT3_FAIL(c, RT3_ERR_INVALID, "gbuffer/depth is NULL");
        if (c->scene_dirty) refresh_scene_tables(c);
        CallOrder order(c, (cudaStream_t)st);
        rt3_launch_gbuffer(c, *p, single_rect_map(c, p, r), gbuffer, depth, ids, bary, (cudaStream_t)st);
    });
}
rt3_status rt3_pass_reference(rt3_context* c, const rt3_params* p, rt3_rect r, const uint32_t* gbuffer, const float* depth, const float* prev, float* light,
                              void* st) {
    return guarded(c, [&]() {
        check_params(c, p); require_accel(c);
        if (!gbuffer || !depth || !light) RT3_FAIL(c, RT3_ERR_INVALID, "gbuffer/depth/light is NULL");
        if (p->blendfactor < 1.0f && !prev) RT3_FAIL(c, RT3_ERR_INVALID, "prev_light is NULL with blendfactor < 1");
        if (c->scene_dirty) refresh_scene_tables(c);
        CallOrder order(c, (cudaStream_t)st);
        rt3_launch_reference(c, *p, single_rect_map(c, p, r), gbuffer, depth, prev, light, (cudaStream_t)st);
    });
}
rt3_status rt3_pass_postprocess(rt3_context* c, const rt3_params* p, rt3_rect r, const float* depth, const float* in, float* out, void* st) {
    return guarded(c, [&]() {
        if (!p || !depth || !in || !out) RT3_FAIL(c, RT3_ERR_INVALID, "NULL argument");
        if (c->scene_dirty) refresh_scene_tables(c);
        CallOrder order(c, (cudaStream_t)st);
        rt3_launch_postprocess(c, *p, single_rect_map(c, p, r), depth, in, out, (cudaStream_t)st);
    });
}
rt3_status rt3_pass_present_bgra8(rt3_context* c, const rt3_params* p, rt3_rect r, const float* depth, const float* in, uint32_t* out, void* st) {
    return guarded(c, [&]() {
        if (!p || !depth || !in || !out) RT3_FAIL(c, RT3_ERR_INVALID, "NULL argument");
        if (c->scene_dirty) refresh_scene_tables(c);
        CallOrder order(c, (cudaStream_t)st);
        rt3_launch_present(c, *p, single_rect_map(c, p, r), depth, in, out, (cudaStream_t)st);
    });
}
/* ---- screen-space probes (row F4) ---- */
rt3_status rt3_probe_layout(const rt3_params* p, uint32_t* atlas_w, uint32_t* atlas_h, uint32_t* sh_records) {
    if (!p) return RT3_ERR_INVALID;
    if (atlas_w) *atlas_w = rt3m::probe_count_x(*p) * 8u;
    if (atlas_h) *atlas_h = rt3m::probe_count_y(*p) * 8u;
    if (sh_records) *sh_records = rt3m::probe_sh_capacity(*p);
    return RT3_OK;
}
rt3_status rt3_pass_probe_trace(rt3_context* c, const rt3_params* p, const uint32_t* gbuffer, const float* depth, const uint16_t* directions,
                                const float* prev_atlas, float* atlas, void* st) {
    return guarded(c, [&]() {
        check_params(c, p); require_accel(c);
        (void)gbuffer;
        if (!depth || !atlas) RT3_FAIL(c, RT3_ERR_INVALID, "depth/atlas is NULL");
        if (p->blendfactor < 1.0f && !prev_atlas) RT3_FAIL(c, RT3_ERR_INVALID, "prev_atlas is NULL with blendfactor < 1");
        if (c->scene_dirty) refresh_scene_tables(c);
        CallOrder order(c, (cudaStream_t)st);
        rt3_launch_probe_trace(c, *p, depth, directions, prev_atlas, atlas, (cudaStream_t)st);
    });
}
/* No allocation on this path: no cudaMalloc / cudaFree, no ensure(), no table refresh (a changed scene is an error here, not a rebuild). */
rt3_status rt3_pass_probe_trace_static(rt3_context* c, const rt3_params* p, const uint32_t* gbuffer, const float* depth, const uint16_t* directions,
                                       const float* prev_atlas, float* atlas, void* st) {
    return guarded(c, [&]() {
        check_params(c, p); require_accel(c);
        (void)gbuffer;
        if (!depth || !atlas) RT3_FAIL(c, RT3_ERR_INVALID, "depth/atlas is NULL");
        if (p->blendfactor < 1.0f && !prev_atlas) RT3_FAIL(c, RT3_ERR_INVALID, "prev_atlas is NULL with blendfactor < 1");
        if (c->scene_dirty) RT3_FAIL(c, RT3_ERR_STATE, "scene tables changed since the last pass call: call an allocating pass (or rt3_accel_build) first");
        if (c->prof_on) RT3_FAIL(c, RT3_ERR_STATE, "rt3_profile_enable creates events per launch: disable it for the allocation-free entry point");
        CallOrder order(c, (cudaStream_t)st);
        rt3_launch_probe_trace_static(c, *p, depth, directions, prev_atlas, atlas, (cudaStream_t)st);
    });
}
rt3_status rt3_pass_probe_sis(rt3_context* c, const rt3_params* p, const uint32_t* gbuffer, uint16_t* directions, void* st) {
    return guarded(c, [&]() {
        if (!p || !gbuffer || !directions) RT3_FAIL(c, RT3_ERR_INVALID, "NULL argument");
        CallOrder order(c, (cudaStream_t)st);
        rt3_launch_probe_sis(c, *p, gbuffer, directions, (cudaStream_t)st);
    });
}
rt3_status rt3_pass_probe_sh(rt3_context* c, const rt3_params* p, const float* atlas, float* sh, void* st) {
    return guarded(c, [&]() {
        if (!p || !atlas || !sh) RT3_FAIL(c, RT3_ERR_INVALID, "NULL argument");
        CallOrder order(c, (cudaStream_t)st);
        rt3_launch_probe_sh(c, *p, atlas, sh, (cudaStream_t)st);
    });
}
rt3_status rt3_pass_probe_interpolate(rt3_context* c, const rt3_params* p, const uint32_t* gbuffer, const float* depth, const float* sh, float* light, void* st) {
    return guarded(c, [&]() {
        if (!p || !gbuffer || !depth || !sh || !light) RT3_FAIL(c, RT3_ERR_INVALID, "NULL argument");
        CallOrder order(c, (cudaStream_t)st);
        rt3_launch_probe_interpolate(c, *p, gbuffer, depth, sh, light, (cudaStream_t)st);
    });
}
rt3_status rt3_pass_gbuffer_tiles(rt3_context* c, const rt3_params* p, const rt3_rect* tiles, uint32_t n, uint32_t* gbuffer, float* depth, void* st) {
    return guarded(c, [&]() {
        check_params(c, p); require_accel(c);
        if ((n && !tiles) || !gbuffer || !depth) RT3_FAIL(c, RT3_ERR_INVALID, "NULL argument");
        if (c->scene_dirty) refresh_scene_tables(c);
        CallOrder order(c, (cudaStream_t)st);
        rt3_launch_gbuffer(c, *p, tiles_map(c, p, tiles, n, (cudaStream_t)st), gbuffer, depth, nullptr, nullptr, (cudaStream_t)st);
    });
}
rt3_status rt3_pass_reference_tiles(rt3_context* c, const rt3_params* p, const rt3_rect* tiles, uint32_t n, const uint32_t* gbuffer, const float* depth,
                                    const float* prev, float* light, void* st) {
    return guarded(c, [&]() {
        check_params(c, p); require_accel(c);
        if ((n && !tiles) || !gbuffer || !depth || !light) RT3_FAIL(c, RT3_ERR_INVALID, "NULL argument");
        if (p->blendfactor < 1.0f && !prev) RT3_FAIL(c, RT3_ERR_INVALID, "prev_light is NULL with blendfactor < 1");
        if (c->scene_dirty) refresh_scene_tables(c);
        CallOrder order(c, (cudaStream_t)st);
        rt3_launch_reference(c, *p, tiles_map(c, p, tiles, n, (cudaStream_t)st), gbuffer, depth, prev, light, (cudaStream_t)st);
    });
}
rt3_status rt3_untile_rgba32f(rt3_context* c, const rt3_rect* tiles, uint32_t n, uint32_t tile_w, uint32_t tile_h, const float* slab, float* frame,
                              uint32_t frame_w, uint32_t frame_h, void* st) {
    return guarded(c, [&]() {
        if ((n && !tiles) || !slab || !frame) RT3_FAIL(c, RT3_ERR_INVALID, "NULL argument");
        rt3_rect* d = ensure<rt3_rect>(c, c->d_untile_rects, n ? n : 1);
        const bool same = n && d == c->urects_dev && (cudaStream_t)st == c->urects_stream && c->urects_host.size() == n &&
                          memcmp(c->urects_host.data(), tiles, n * sizeof(rt3_rect)) == 0;       /* as in tiles_map */
        if (n && !same) {
            RT3_CUDA(c, cudaMemcpyAsync(d, tiles, n * sizeof(rt3_rect), cudaMemcpyHostToDevice, (cudaStream_t)st));
            c->urects_host.assign(tiles, tiles + n); c->urects_dev = d; c->urects_stream = (cudaStream_t)st;
        }
        rt3_launch_untile(c, d, n, tile_w, tile_h, slab, frame, frame_w, frame_h, (cudaStream_t)st);
    });
}

rt3_status rt3_scatter_tiles_rgba32f(rt3_context* c, const rt3_rect* tiles, const uint64_t* dst_top_left, uint32_t n, uint32_t tile_w, uint32_t tile_h,
                                     const float* slab, uint32_t dst_pitch_px, void* st) {
    return guarded(c, [&]() {
        if ((n && (!tiles || !dst_top_left)) || !slab) RT3_FAIL(c, RT3_ERR_INVALID, "NULL argument");
        if (!n) return;
        for (uint32_t i = 0; i < n; ++i)
            if (tiles[i].w && tiles[i].h && (!dst_top_left[i] || (dst_top_left[i] & 15u))) RT3_FAIL(c, RT3_ERR_INVALID, "tile destination is NULL or not 16-byte aligned");
        /* one device buffer: n rects (16 B each) followed by n destination addresses; re-uploaded only when the lists change */
        std::vector<uint64_t> packed(3 * (size_t)n);
        memcpy(packed.data(), tiles, (size_t)n * sizeof(rt3_rect));
        memcpy(packed.data() + 2 * (size_t)n, dst_top_left, (size_t)n * sizeof(uint64_t));
        uint64_t* d = ensure<uint64_t>(c, c->d_scatter, packed.size());
        if (packed != c->scatter_host || (cudaStream_t)st != c->scatter_stream) {
            RT3_CUDA(c, cudaMemcpyAsync(d, packed.data(), packed.size() * 8, cudaMemcpyHostToDevice, (cudaStream_t)st));
            RT3_CUDA(c, cudaStreamSynchronize((cudaStream_t)st));          /* `packed` is pageable and local: the copy must have left it */
            c->scatter_host = packed; c->scatter_stream = (cudaStream_t)st;
        }
        rt3_launch_scatter_tiles(c, (const rt3_rect*)d, (const unsigned long long*)(d + 2 * (size_t)n), n, tile_w, tile_h, slab, dst_pitch_px, (cudaStream_t)st);
    });
}

/* ---- raw queries ------------------------------------------------------------------------------ */
rt3_status rt3_trace_closest(rt3_context* c, const rt3_ray* rays, uint64_t n, rt3_hit* out, void* st) {
    return guarded(c, [&]() {
        require_accel(c);
        if (n && (!rays || !out)) RT3_FAIL(c, RT3_ERR_INVALID, "NULL argument");
        CallOrder order(c, (cudaStream_t)st);
        rt3_launch_trace_closest(c, rays, nullptr, n, out, nullptr, (cudaStream_t)st);
    });
}
rt3_status rt3_trace_any(rt3_context* c, const rt3_ray* rays, uint64_t n, uint32_t* out_bits, void* st) {
    return guarded(c, [&]() {
        require_accel(c);
        if (n && (!rays || !out_bits)) RT3_FAIL(c, RT3_ERR_INVALID, "NULL argument");
        CallOrder order(c, (cudaStream_t)st);
        rt3_launch_trace_any(c, rays, nullptr, n, out_bits, nullptr, (cudaStream_t)st);
    });
}

/* ---- end-to-end host-buffer entry points ------------------------------------------------------ */
rt3_status rt3_render_frame_host(rt3_context* c, const rt3_params* p, const float* prev_light_host, float* light_host, float* opt_depth_host,
                                 uint32_t* opt_gbuffer_host) {
    return guarded(c, [&]() {
        check_params(c, p); require_accel(c);
        if (!light_host) RT3_FAIL(c, RT3_ERR_INVALID, "light_host is NULL");
        if (p->blendfactor < 1.0f && !prev_light_host) RT3_FAIL(c, RT3_ERR_INVALID, "prev_light_host is NULL with blendfactor < 1");
        if (c->scene_dirty) refresh_scene_tables(c);
        const uint32_t W = (uint32_t)p->window_size[0], H = (uint32_t)p->window_size[1];
        const size_t px = (size_t)W * H;
        rt3_rect full = {0, 0, W, H};
        PixelMap m = single_rect_map(c, p, full);
        uint32_t* gb = ensure<uint32_t>(c, c->f_gbuffer, 4 * px);
        float* dp = ensure<float>(c, c->f_depth, px);
        float* lt = ensure<float>(c, c->f_light, 4 * px);
        float* pv = nullptr;
        cudaStream_t st = 0;
        CallOrder order(c, st);
        if (p->blendfactor < 1.0f) {
            pv = ensure<float>(c, c->f_prev, 4 * px);
            RT3_CUDA(c, cudaMemcpyAsync(pv, prev_light_host, 16 * px, cudaMemcpyHostToDevice, st));
        }
        /* Light is only written for non-background pixels (refrence_mode.slang:18-21); the host frame gets zeros there. */
        RT3_CUDA(c, cudaMemsetAsync(lt, 0, 16 * px, st));
        rt3_launch_gbuffer(c, *p, m, gb, dp, nullptr, nullptr, st);
        rt3_launch_reference(c, *p, m, gb, dp, pv, lt, st);
        RT3_CUDA(c, cudaMemcpyAsync(light_host, lt, 16 * px, cudaMemcpyDeviceToHost, st));
        if (opt_depth_host) RT3_CUDA(c, cudaMemcpyAsync(opt_depth_host, dp, 4 * px, cudaMemcpyDeviceToHost, st));
        if (opt_gbuffer_host) RT3_CUDA(c, cudaMemcpyAsync(opt_gbuffer_host, gb, 16 * px, cudaMemcpyDeviceToHost, st));
        RT3_CUDA(c, cudaStreamSynchronize(st));
    });
}
/* Frames in flight: render on the compute stream, read back on the copy stream, up to 3 frames deep. */
rt3_status rt3_frame_submit_host(rt3_context* c, const rt3_params* p, const float* prev_light_host, float* light_host, uint64_t* out_ticket) {
    return guarded(c, [&]() {
        check_params(c, p); require_accel(c);
        if (!light_host || !out_ticket) RT3_FAIL(c, RT3_ERR_INVALID, "light_host / out_ticket is NULL");
        if (p->blendfactor < 1.0f && !prev_light_host) RT3_FAIL(c, RT3_ERR_INVALID, "prev_light_host is NULL with blendfactor < 1");
        if (c->scene_dirty) refresh_scene_tables(c);
        const uint32_t W = (uint32_t)p->window_size[0], H = (uint32_t)p->window_size[1];
        const size_t px = (size_t)W * H;
        rt3_rect full = {0, 0, W, H};
        PixelMap m = single_rect_map(c, p, full);
        const uint64_t ticket = c->next_ticket;
        rt3_context::FrameSlot& fs = c->slots[ticket % 3];
        if (fs.busy) { RT3_CUDA(c, cudaEventSynchronize(fs.copied)); fs.busy = false; }   /* the slot's previous frame (ticket - 3) must have left */
        uint32_t* gb = ensure<uint32_t>(c, fs.gb, 4 * px);
        float* dp = ensure<float>(c, fs.dp, px);
        float* lt = ensure<float>(c, fs.lt, 4 * px);
        float* pv = nullptr;
        cudaStream_t st = c->s_compute;
        CallOrder order(c, st);
        if (p->blendfactor < 1.0f) {
            pv = ensure<float>(c, fs.pv, 4 * px);
            RT3_CUDA(c, cudaMemcpyAsync(pv, prev_light_host, 16 * px, cudaMemcpyHostToDevice, st));
        }
        RT3_CUDA(c, cudaMemsetAsync(lt, 0, 16 * px, st));
        rt3_launch_gbuffer(c, *p, m, gb, dp, nullptr, nullptr, st);
        rt3_launch_reference(c, *p, m, gb, dp, pv, lt, st);
        RT3_CUDA(c, cudaEventRecord(fs.rendered, st));
        RT3_CUDA(c, cudaStreamWaitEvent(c->s_copy, fs.rendered, 0));
        RT3_CUDA(c, cudaMemcpyAsync(light_host, lt, 16 * px, cudaMemcpyDeviceToHost, c->s_copy));
        RT3_CUDA(c, cudaEventRecord(fs.copied, c->s_copy));
        fs.busy = true;
        c->next_ticket = ticket + 1;
        *out_ticket = ticket;
    });
}
rt3_status rt3_frame_wait(rt3_context* c, uint64_t ticket) {
    return guarded(c, [&]() {
        if (ticket >= c->next_ticket) RT3_FAIL(c, RT3_ERR_INVALID, "ticket was never issued");
        if (ticket + 3 < c->next_ticket) return;                    /* its slot has been reused: that frame completed long ago */
        /* frames complete in submission order on both streams: waiting for `ticket` covers every older frame */
        for (uint64_t t = (c->next_ticket > 3 ? c->next_ticket - 3 : 0); t <= ticket; ++t) {
            rt3_context::FrameSlot& fs = c->slots[t % 3];
            if (fs.busy) { RT3_CUDA(c, cudaEventSynchronize(fs.copied)); fs.busy = false; }
        }
    });
}
rt3_status rt3_host_alloc(rt3_context* c, size_t bytes, void** out) {
    return guarded(c, [&]() {
        if (!out) RT3_FAIL(c, RT3_ERR_INVALID, "out is NULL");
        *out = nullptr;
        RT3_CUDA(c, cudaHostAlloc(out, bytes ? bytes : 1, cudaHostAllocDefault));
    });
}
rt3_status rt3_host_free(rt3_context* c, void* p) {
    return guarded(c, [&]() { if (p) RT3_CUDA(c, cudaFreeHost(p)); });
}
rt3_status rt3_trace_closest_host(rt3_context* c, const rt3_ray* rays_host, uint64_t n, rt3_hit* out_host) {
    return guarded(c, [&]() {
        require_accel(c);
        if (n && (!rays_host || !out_host)) RT3_FAIL(c, RT3_ERR_INVALID, "NULL argument");
        if (!n) return;
        rt3_ray* r = ensure<rt3_ray>(c, c->f_rays, n);
        rt3_hit* h = ensure<rt3_hit>(c, c->f_hits, n);
        CallOrder order(c, 0);
        RT3_CUDA(c, cudaMemcpyAsync(r, rays_host, n * sizeof(rt3_ray), cudaMemcpyHostToDevice, 0));
        rt3_launch_trace_closest(c, r, nullptr, n, h, nullptr, 0);
        RT3_CUDA(c, cudaMemcpyAsync(out_host, h, n * sizeof(rt3_hit), cudaMemcpyDeviceToHost, 0));
        RT3_CUDA(c, cudaStreamSynchronize(0));
    });
}
rt3_status rt3_trace_any_host(rt3_context* c, const rt3_ray* rays_host, uint64_t n, uint32_t* out_bits_host) {
    return guarded(c, [&]() {
        require_accel(c);
        if (n && (!rays_host || !out_bits_host)) RT3_FAIL(c, RT3_ERR_INVALID, "NULL argument");
        if (!n) return;
        const size_t words = (n + 31) / 32;
        rt3_ray* r = ensure<rt3_ray>(c, c->f_rays, n);
        uint32_t* b = ensure<uint32_t>(c, c->f_bits, words);
        CallOrder order(c, 0);
        RT3_CUDA(c, cudaMemcpyAsync(r, rays_host, n * sizeof(rt3_ray), cudaMemcpyHostToDevice, 0));
        rt3_launch_trace_any(c, r, nullptr, n, b, nullptr, 0);
        RT3_CUDA(c, cudaMemcpyAsync(out_bits_host, b, words * 4, cudaMemcpyDeviceToHost, 0));
        RT3_CUDA(c, cudaStreamSynchronize(0));
    });
}

rt3_status rt3_set_option(rt3_context* c, int option, int value) {
    return guarded(c, [&]() {
        if (option == RT3_OPT_STRIPS) {
            if (value < 0 || value > RT3_MAX_LANES) RT3_FAIL(c, RT3_ERR_INVALID, "RT3_OPT_STRIPS: 0 (default) .. 4");
            c->opt_strips = value;
        } else RT3_FAIL(c, RT3_ERR_INVALID, "unknown option");
    });
}

/* ---- instrumentation -------------------------------------------------------------------------- */
rt3_status rt3_counters_read(rt3_context* c, rt3_counters* out, int reset) {
    return guarded(c, [&]() {
        if (!out) RT3_FAIL(c, RT3_ERR_INVALID, "out is NULL");
        DeviceCounters h;
        RT3_CUDA(c, cudaDeviceSynchronize());
        RT3_CUDA(c, cudaMemcpy(&h, c->d_counters, sizeof(h), cudaMemcpyDeviceToHost));
        out->rays_primary = h.rays_primary; out->rays_secondary = h.rays_secondary; out->rays_shadow = h.rays_shadow;
        out->paths_started = h.paths_started; out->paths_escaped = h.paths_escaped;
        if (reset) RT3_CUDA(c, cudaMemset(c->d_counters, 0, sizeof(DeviceCounters)));
    });
}
uint64_t rt3_kernel_launches(const rt3_context* c) { return c ? c->launches : 0; }
rt3_status rt3_measure_l2_bandwidth(rt3_context* c, size_t bytes, int iters, double* out_gbs) {
    return guarded(c, [&]() {
        if (!out_gbs || bytes < 4096 || iters < 1) RT3_FAIL(c, RT3_ERR_INVALID, "bad arguments");
        *out_gbs = rt3_run_l2_sweep(c, bytes, iters);
    });
}
rt3_status rt3_trace_closest_stats(rt3_context* c, const rt3_ray* rays, uint64_t n, rt3_trace_stats* out) {
    return guarded(c, [&]() {
        require_accel(c);
        if (!out || (n && !rays)) RT3_FAIL(c, RT3_ERR_INVALID, "NULL argument");
        CallOrder order(c, 0);
        rt3_launch_trace_stats(c, rays, n, out);
    });
}

rt3_status rt3_trace_closest_tail_stats(rt3_context* c, const rt3_ray* rays, uint64_t n, rt3_tail_stats* out) {
    return guarded(c, [&]() {
        require_accel(c);
        if (!out || (n && !rays)) RT3_FAIL(c, RT3_ERR_INVALID, "NULL argument");
        CallOrder order(c, 0);
        rt3_launch_trace_tail_stats(c, rays, n, out);
    });
}

rt3_status rt3_profile_enable(rt3_context* c, int enable) {
    return guarded(c, [&]() {
        RT3_CUDA(c, cudaDeviceSynchronize());
        c->prof_on = enable != 0; c->prof_used = 0; c->prof_trace_ms = 0; c->prof_trace_launches = 0;
    });
}
rt3_status rt3_profile_read(rt3_context* c, double* trace_ms, uint64_t* trace_launches) {
    return guarded(c, [&]() {
        RT3_CUDA(c, cudaDeviceSynchronize());
        for (size_t i = 0; i + 1 < c->prof_used; i += 2) {
            float ms = 0;
            RT3_CUDA(c, cudaEventElapsedTime(&ms, c->prof_events[i], c->prof_events[i + 1]));
            c->prof_trace_ms += ms; c->prof_trace_launches++;
        }
        c->prof_used = 0;
        if (trace_ms) *trace_ms = c->prof_trace_ms;
        if (trace_launches) *trace_launches = c->prof_trace_launches;
        c->prof_trace_ms = 0; c->prof_trace_launches = 0;
    });
}
rt3_status rt3_accel_download(rt3_context* c, void* nodes_host, size_t nodes_bytes, void* tris_host, size_t tris_bytes) {
    return guarded(c, [&]() {
        require_accel(c);
        if (nodes_bytes > (size_t)c->n_wide_nodes * sizeof(rt3b::WideNode) || tris_bytes > (size_t)c->n_leaf_records * sizeof(rt3b::Tri48))
            RT3_FAIL(c, RT3_ERR_INVALID, "download larger than the acceleration structure");
        RT3_CUDA(c, cudaDeviceSynchronize());
        if (nodes_bytes) RT3_CUDA(c, cudaMemcpy(nodes_host, c->bvh_nodes.ptr, nodes_bytes, cudaMemcpyDeviceToHost));
        if (tris_bytes) RT3_CUDA(c, cudaMemcpy(tris_host, c->bvh_tris.ptr, tris_bytes, cudaMemcpyDeviceToHost));
    });
}

}  /* extern "C" */
