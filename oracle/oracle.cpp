/*
 * oracle.cpp — CPU oracle for the ray-traced G-buffer + reference-mode path tracer.
 *
 * TEST INFRASTRUCTURE ONLY (see oracle.h).  PARITY UNPINNED at the driver boundary (see oracle.h).
 *
 * Structure
 *   - per-pixel passes follow the reference's raygen shaders statement by statement
 *       orc_pass_gbuffer    <- shaders/old/gbuffer.slang:8-20
 *       orc_pass_reference  <- shaders/old/refrence_mode.slang:14-66
 *       orc_pass_postprocess<- shaders/old/postprocess.slang:90-112
 *     with the shader math taken from include/rt3_shared_math.h (each function there cites its source);
 *   - trace() (the missing bindless_descriptor.slang wrapper over TraceRay) is a closest-hit query
 *     over world-space triangles, implemented three times: brute force over every triangle (the definition),
 *     a binned-SAH BVH2 written for this oracle, and that BVH collapsed to 4 children per node with SSE slab /
 *     triangle tests (the default and the timed CPU arm); all proven equal bit for bit in tests.
 *   - deliberately single-ray, one-pixel-at-a-time recursion-free loops: nothing here is shared with
 *     the CUDA implementation except the math header.
 */
#include "oracle.h"
#include "rt3_shared_math.h"

#include <immintrin.h>

#include <algorithm>
#include <atomic>
#include <chrono>
#include <cstdlib>
#include <limits>
#include <thread>
#include <vector>

using namespace rt3m;

namespace {

struct WorldTri { vec3 v0, e1, e2; };
struct Box {
    vec3 lo, hi;
    void reset() { lo = v3(3.4e38f, 3.4e38f, 3.4e38f); hi = v3(-3.4e38f, -3.4e38f, -3.4e38f); }
    void grow(vec3 p) {
        lo = v3(min_(lo.x, p.x), min_(lo.y, p.y), min_(lo.z, p.z));
        hi = v3(max_(hi.x, p.x), max_(hi.y, p.y), max_(hi.z, p.z));
    }
    void grow(const Box& b) { grow(b.lo); grow(b.hi); }
    float area() const { vec3 e = hi - lo; return 2.0f * (e.x * e.y + e.y * e.z + e.z * e.x); }
};
struct Node2 { Box box; uint32_t left, count; /* count>0: leaf, tris [left,left+count); else children left,left+1 */ };
/* The TIMED CPU arm's structure (bench.py cpu_baseline / --impl reference; also the default of the passes): the BVH2 above collapsed to four
 * children per node, boxes in SoA for one SSE slab test per node, and every leaf's <= 4 triangles in SoA for one SSE Moeller-Trumbore test
 * whose operations mirror rt3m::ray_triangle one for one (same t, u, v bits).  Results do not depend on the structure (closer_hit is order
 * independent); the BVH2 and brute force remain as cross-checks (tests/test_oracle_kats.py). */
constexpr uint32_t REF_LEAF = 0x80000000u, REF_NONE = 0xFFFFFFFFu;
struct alignas(16) Node4 { float lo[3][4], hi[3][4]; uint32_t child[4]; uint32_t valid, pad_[3]; };   /* child: REF_NONE, REF_LEAF | leaf block, or node index; valid: bit k = slot k is used */
struct alignas(16) Leaf4 { float v0[3][4], e1[3][4], e2[3][4]; uint32_t id[4]; };    /* unused lanes: e1 = e2 = 0 (det = 0: rejected), id = RT3_MISS */

}  // namespace

struct orc_scene {
    std::vector<rt3_vertex> vertices;
    std::vector<uint32_t> indices;
    std::vector<rt3_geometry_info> geometries;
    std::vector<uint32_t> inst_first, inst_count;
    std::vector<float> inst_xform;
    std::vector<uint32_t> seg_first_prim, seg_geometry, seg_instance;
    std::vector<std::vector<uint8_t>> textures;
    std::vector<const uint8_t*> tex_ptr;
    std::vector<uint32_t> tex_dim;
    std::vector<float> env, env_func, env_marginal, env_conditional;
    float env_total = 0.f;
    uint32_t env_w = 0, env_h = 0;
    std::vector<uint8_t> bn;
    uint32_t bn_w = 0, bn_h = 0;
    float srgb_lut[256];
    /* built */
    std::vector<WorldTri> tris;      /* global id order */
    std::vector<uint32_t> order;     /* BVH leaf order -> global id */
    std::vector<Node2> nodes;
    std::vector<Node4> nodes4;
    std::vector<Leaf4> leaves4;
    uint32_t root4 = REF_NONE;
    Box bounds;
    float pad = 0.f;
    SceneView view() const {
        SceneView s;
        s.vertices = vertices.data(); s.indices = indices.data(); s.geometries = geometries.data();
        s.seg_first_prim = seg_first_prim.data(); s.seg_geometry = seg_geometry.data(); s.seg_instance = seg_instance.data();
        s.inst_xform = inst_xform.data(); s.n_segments = (uint32_t)seg_geometry.size();
        s.tex_data = tex_ptr.data(); s.tex_dim = tex_dim.data(); s.n_textures = (uint32_t)tex_ptr.size();
        s.srgb_lut = srgb_lut;
        s.env = env.empty() ? nullptr : env.data(); s.env_w = env_w; s.env_h = env_h;
        s.env_func = env_func.empty() ? nullptr : env_func.data(); s.env_marginal = env_marginal.data(); s.env_conditional = env_conditional.data();
        s.env_total = env_total;
        s.bn = bn.empty() ? nullptr : bn.data(); s.bn_w = bn_w; s.bn_h = bn_h;
        s.shade_rec = nullptr; s.shade64 = nullptr;   /* the oracle keeps the reference's own fetch chain (hit_logic.slang:10-19) */
        return s;
    }
};

namespace {

template <class F>
void parallel_for(uint64_t n, int threads, uint64_t grain, F f) {
    if (threads <= 1 || n <= grain) { f(0, n); return; }
    std::atomic<uint64_t> next(0);
    std::vector<std::thread> pool;
    for (int t = 0; t < threads; ++t)
        pool.emplace_back([&]() {
            for (;;) {
                uint64_t b = next.fetch_add(grain);
                if (b >= n) break;
                f(b, std::min(n, b + grain));
            }
        });
    for (auto& th : pool) th.join();
}

/* ---- BVH2: binned SAH (16 bins), leaves of <= 4 triangles --------------------------------- */
struct Builder {
    orc_scene& s;
    std::vector<Box> tbox;
    std::vector<vec3> cen;
    explicit Builder(orc_scene& sc) : s(sc) {}
    static float axis(vec3 v, int a) { return a == 0 ? v.x : (a == 1 ? v.y : v.z); }
    void build() {
        uint32_t n = (uint32_t)s.tris.size();
        tbox.resize(n); cen.resize(n); s.order.resize(n);
        for (uint32_t i = 0; i < n; ++i) {
            const WorldTri& t = s.tris[i];
            Box b; b.reset(); b.grow(t.v0); b.grow(t.v0 + t.e1); b.grow(t.v0 + t.e2);
            b.lo = b.lo - v3(s.pad, s.pad, s.pad); b.hi = b.hi + v3(s.pad, s.pad, s.pad);
            tbox[i] = b; cen[i] = (b.lo + b.hi) * 0.5f; s.order[i] = i;
        }
        s.nodes.clear(); s.nodes.reserve(2 * (size_t)n + 2);
        if (n == 0) return;
        s.nodes.push_back(Node2());
        struct Job { uint32_t node, first, count; };
        std::vector<Job> stack; stack.push_back({0, 0, n});
        while (!stack.empty()) {
            Job j = stack.back(); stack.pop_back();
            Box nb, cb; nb.reset(); cb.reset();
            for (uint32_t i = j.first; i < j.first + j.count; ++i) { nb.grow(tbox[s.order[i]]); cb.grow(cen[s.order[i]]); }
            Node2& nd = s.nodes[j.node]; nd.box = nb;
            if (j.count <= 2) { nd.left = j.first; nd.count = j.count; continue; }
            const int NB = 16;
            float best = 3.4e38f; int baxis = -1, bsplit = -1;
            for (int a = 0; a < 3; ++a) {
                float lo = axis(cb.lo, a), hi = axis(cb.hi, a);
                if (!(hi > lo)) continue;
                Box bb[NB]; uint32_t bc[NB];
                for (int k = 0; k < NB; ++k) { bb[k].reset(); bc[k] = 0; }
                float scale = NB / (hi - lo);
                for (uint32_t i = j.first; i < j.first + j.count; ++i) {
                    uint32_t id = s.order[i];
                    int k = std::min(NB - 1, (int)((axis(cen[id], a) - lo) * scale));
                    bb[k].grow(tbox[id]); bc[k]++;
                }
                float la[NB], ra[NB]; uint32_t lc[NB], rc[NB];
                Box acc; acc.reset(); uint32_t c = 0;
                for (int k = 0; k < NB - 1; ++k) { if (bc[k]) acc.grow(bb[k]); c += bc[k]; la[k] = c ? acc.area() : 0.f; lc[k] = c; }
                acc.reset(); c = 0;
                for (int k = NB - 1; k > 0; --k) { if (bc[k]) acc.grow(bb[k]); c += bc[k]; ra[k - 1] = c ? acc.area() : 0.f; rc[k - 1] = c; }
                for (int k = 0; k < NB - 1; ++k) {
                    if (!lc[k] || !rc[k]) continue;
                    float cost = la[k] * lc[k] + ra[k] * rc[k];
                    if (cost < best) { best = cost; baxis = a; bsplit = k; }
                }
            }
            float leaf_cost = nb.area() * j.count;
            if (j.count <= 4 && (baxis < 0 || best >= leaf_cost)) { nd.left = j.first; nd.count = j.count; continue; }
            uint32_t mid;
            if (baxis >= 0) {
                float lo = axis(cb.lo, baxis), hi = axis(cb.hi, baxis);
                float scale = NB / (hi - lo);
                auto it = std::partition(s.order.begin() + j.first, s.order.begin() + j.first + j.count, [&](uint32_t id) {
                    int k = std::min(NB - 1, (int)((axis(cen[id], baxis) - lo) * scale));
                    return k <= bsplit;
                });
                mid = (uint32_t)(it - s.order.begin());
            } else {
                mid = j.first + j.count / 2; /* all centroids coincide: median split */
            }
            if (mid == j.first || mid == j.first + j.count) mid = j.first + j.count / 2;
            uint32_t l = (uint32_t)s.nodes.size();
            s.nodes.push_back(Node2()); s.nodes.push_back(Node2());
            s.nodes[j.node].left = l; s.nodes[j.node].count = 0;
            stack.push_back({l, j.first, mid - j.first});
            stack.push_back({l + 1, mid, j.first + j.count - mid});
        }
    }
};

inline float safe_inv(float d) {
    const float tiny = 1e-20f;
    if (abs_(d) < tiny) d = (as_uint(d) >> 31) ? -tiny : tiny;
    return 1.0f / d;
}
/* Returns the entry distance, or a negative value when the (padded) slab test fails. */
inline float slab(const Box& b, vec3 o, vec3 id, float tmin, float tmax) {
    float x1 = (b.lo.x - o.x) * id.x, x2 = (b.hi.x - o.x) * id.x;
    float y1 = (b.lo.y - o.y) * id.y, y2 = (b.hi.y - o.y) * id.y;
    float z1 = (b.lo.z - o.z) * id.z, z2 = (b.hi.z - o.z) * id.z;
    float tn = max_(max_(min_(x1, x2), min_(y1, y2)), max_(min_(z1, z2), tmin));
    float tf = min_(min_(max_(x1, x2), max_(y1, y2)), min_(max_(z1, z2), tmax));
    return tn <= tf * 1.0000005f ? tn : -1.0f;
}

/* trace(): closest hit.  The payload semantics follow default_hit.slang:4-8 / default_miss.slang:4-7. */
rt3_hit closest_brute(const orc_scene& s, const rt3_ray& r) {
    rt3_hit h; h.t = r.tmax; h.u = 0.f; h.v = 0.f; h.prim = RT3_MISS;
    vec3 o = v3(r.o), d = v3(r.d);
    uint32_t n = (uint32_t)s.tris.size();
    for (uint32_t i = 0; i < n; ++i) {
        float t, u, v;
        const WorldTri& tr = s.tris[i];
        if (ray_triangle(o, d, tr.v0, tr.e1, tr.e2, t, u, v) && t > r.tmin && t < r.tmax && closer_hit(t, i, h.t, h.prim)) {
            h.t = t; h.u = u; h.v = v; h.prim = i;
        }
    }
    if (h.prim == RT3_MISS) h.t = r.tmax;
    return h;
}
rt3_hit closest_bvh2(const orc_scene& s, const rt3_ray& r) {
    rt3_hit h; h.t = r.tmax; h.u = 0.f; h.v = 0.f; h.prim = RT3_MISS;
    if (s.nodes.empty()) return h;
    vec3 o = v3(r.o), d = v3(r.d);
    vec3 id = v3(safe_inv(d.x), safe_inv(d.y), safe_inv(d.z));
    if (slab(s.nodes[0].box, o, id, r.tmin, h.t) < 0.0f) return h;
    struct Entry { uint32_t node; float tn; };
    Entry stack[128]; int sp = 0; stack[sp++] = {0, 0.0f};
    while (sp) {
        Entry e = stack[--sp];
        if (e.tn > h.t) continue;                     /* entered farther than the current hit (ties are kept: tn <= t) */
        const Node2& nd = s.nodes[e.node];
        if (nd.count) {
            for (uint32_t k = nd.left; k < nd.left + nd.count; ++k) {
                uint32_t i = s.order[k];
                float t, u, v;
                const WorldTri& tr = s.tris[i];
                if (ray_triangle(o, d, tr.v0, tr.e1, tr.e2, t, u, v) && t > r.tmin && t < r.tmax && closer_hit(t, i, h.t, h.prim)) {
                    h.t = t; h.u = u; h.v = v; h.prim = i;
                }
            }
        } else {
            if (sp + 2 > 128) abort();
            float ta = slab(s.nodes[nd.left].box, o, id, r.tmin, h.t), tb = slab(s.nodes[nd.left + 1].box, o, id, r.tmin, h.t);
            /* near child on top of the stack; traversal order does not change the result (closer_hit is order independent) */
            if (ta >= 0.0f && tb >= 0.0f) {
                if (ta <= tb) { stack[sp++] = {nd.left + 1, tb}; stack[sp++] = {nd.left, ta}; }
                else { stack[sp++] = {nd.left, ta}; stack[sp++] = {nd.left + 1, tb}; }
            } else if (ta >= 0.0f) stack[sp++] = {nd.left, ta};
            else if (tb >= 0.0f) stack[sp++] = {nd.left + 1, tb};
        }
    }
    return h;
}
bool any_brute(const orc_scene& s, const rt3_ray& r) {
    vec3 o = v3(r.o), d = v3(r.d);
    for (const WorldTri& tr : s.tris) {
        float t, u, v;
        if (ray_triangle(o, d, tr.v0, tr.e1, tr.e2, t, u, v) && t > r.tmin && t < r.tmax) return true;
    }
    return false;
}
bool any_bvh2(const orc_scene& s, const rt3_ray& r) {
    if (s.nodes.empty()) return false;
    vec3 o = v3(r.o), d = v3(r.d);
    vec3 id = v3(safe_inv(d.x), safe_inv(d.y), safe_inv(d.z));
    uint32_t stack[128]; int sp = 0; stack[sp++] = 0;
    while (sp) {
        const Node2& nd = s.nodes[stack[--sp]];
        if (slab(nd.box, o, id, r.tmin, r.tmax) < 0.0f) continue;
        if (nd.count) {
            for (uint32_t k = nd.left; k < nd.left + nd.count; ++k) {
                float t, u, v;
                const WorldTri& tr = s.tris[s.order[k]];
                if (ray_triangle(o, d, tr.v0, tr.e1, tr.e2, t, u, v) && t > r.tmin && t < r.tmax) return true;
            }
        } else { stack[sp++] = nd.left; stack[sp++] = nd.left + 1; }
    }
    return false;
}

/* ---- BVH4 + SSE (the timed CPU arm) ------------------------------------------------------------------------------------------------ */
void build_bvh4(orc_scene& s) {
    s.nodes4.clear(); s.leaves4.clear(); s.root4 = REF_NONE;
    if (s.nodes.empty()) return;
    const float inf = std::numeric_limits<float>::infinity();
    auto make_leaf = [&](const Node2& nd) -> uint32_t {
        Leaf4 lf;
        for (int k = 0; k < 4; ++k) {
            for (int a = 0; a < 3; ++a) { lf.v0[a][k] = 0.f; lf.e1[a][k] = 0.f; lf.e2[a][k] = 0.f; }
            lf.id[k] = RT3_MISS;
        }
        for (uint32_t k = 0; k < nd.count && k < 4u; ++k) {
            const uint32_t i = s.order[nd.left + k];
            const WorldTri& t = s.tris[i];
            lf.v0[0][k] = t.v0.x; lf.v0[1][k] = t.v0.y; lf.v0[2][k] = t.v0.z;
            lf.e1[0][k] = t.e1.x; lf.e1[1][k] = t.e1.y; lf.e1[2][k] = t.e1.z;
            lf.e2[0][k] = t.e2.x; lf.e2[1][k] = t.e2.y; lf.e2[2][k] = t.e2.z;
            lf.id[k] = i;
        }
        s.leaves4.push_back(lf);
        return REF_LEAF | (uint32_t)(s.leaves4.size() - 1);
    };
    if (s.nodes[0].count) {                                  /* the whole scene is one leaf: a node with one child */
        if (s.nodes[0].count > 4) return;                     /* cannot happen (the builder splits above 4); the BVH2 stays in charge */
        Node4 n4;
        for (int k = 0; k < 4; ++k) { for (int a = 0; a < 3; ++a) { n4.lo[a][k] = inf; n4.hi[a][k] = inf; } n4.child[k] = REF_NONE; }
        n4.valid = 1u; n4.pad_[0] = n4.pad_[1] = n4.pad_[2] = 0u;
        const Box& b = s.nodes[0].box;
        n4.lo[0][0] = b.lo.x; n4.lo[1][0] = b.lo.y; n4.lo[2][0] = b.lo.z; n4.hi[0][0] = b.hi.x; n4.hi[1][0] = b.hi.y; n4.hi[2][0] = b.hi.z;
        n4.child[0] = make_leaf(s.nodes[0]);
        s.nodes4.push_back(n4); s.root4 = 0;
        return;
    }
    struct Job { uint32_t node2, node4; };
    std::vector<Job> stack;
    s.nodes4.push_back(Node4()); s.root4 = 0;
    stack.push_back({0u, 0u});
    while (!stack.empty()) {
        const Job j = stack.back(); stack.pop_back();
        uint32_t cand[4]; int nc = 0;
        cand[nc++] = s.nodes[j.node2].left; cand[nc++] = s.nodes[j.node2].left + 1;
        while (nc < 4) {                                      /* open the inner candidate of largest surface area */
            int best = -1; float ba = -1.f;
            for (int k = 0; k < nc; ++k) if (!s.nodes[cand[k]].count && s.nodes[cand[k]].box.area() > ba) { ba = s.nodes[cand[k]].box.area(); best = k; }
            if (best < 0) break;
            const uint32_t l = s.nodes[cand[best]].left;
            cand[best] = l; cand[nc++] = l + 1;
        }
        Node4 n4;
        for (int k = 0; k < 4; ++k) { for (int a = 0; a < 3; ++a) { n4.lo[a][k] = inf; n4.hi[a][k] = inf; } n4.child[k] = REF_NONE; }
        n4.valid = (1u << nc) - 1u; n4.pad_[0] = n4.pad_[1] = n4.pad_[2] = 0u;
        for (int k = 0; k < nc; ++k) {
            const Node2& c = s.nodes[cand[k]];
            n4.lo[0][k] = c.box.lo.x; n4.lo[1][k] = c.box.lo.y; n4.lo[2][k] = c.box.lo.z;
            n4.hi[0][k] = c.box.hi.x; n4.hi[1][k] = c.box.hi.y; n4.hi[2][k] = c.box.hi.z;
            if (c.count) {
                if (c.count > 4) { s.nodes4.clear(); s.leaves4.clear(); s.root4 = REF_NONE; return; }
                n4.child[k] = make_leaf(c);
            } else {
                n4.child[k] = (uint32_t)s.nodes4.size();
                s.nodes4.push_back(Node4());
                stack.push_back({cand[k], n4.child[k]});
            }
        }
        s.nodes4[j.node4] = n4;
    }
}

struct Ray4 { __m128 ox, oy, oz, dx, dy, dz, idx, idy, idz; };
inline Ray4 splat_ray(vec3 o, vec3 d, vec3 id) {
    Ray4 r;
    r.ox = _mm_set1_ps(o.x); r.oy = _mm_set1_ps(o.y); r.oz = _mm_set1_ps(o.z);
    r.dx = _mm_set1_ps(d.x); r.dy = _mm_set1_ps(d.y); r.dz = _mm_set1_ps(d.z);
    r.idx = _mm_set1_ps(id.x); r.idy = _mm_set1_ps(id.y); r.idz = _mm_set1_ps(id.z);
    return r;
}
/* slab() for the four children of a node: the same expression per lane (_mm_min_ps(a, b) = a < b ? a : b = rt3m::min_, likewise max);
 * returns the hit mask, tn4 = entry distances */
inline int slab4(const Node4& n, const Ray4& r, float tmin, float tmax, __m128& tn4) {
    const __m128 x1 = _mm_mul_ps(_mm_sub_ps(_mm_load_ps(n.lo[0]), r.ox), r.idx), x2 = _mm_mul_ps(_mm_sub_ps(_mm_load_ps(n.hi[0]), r.ox), r.idx);
    const __m128 y1 = _mm_mul_ps(_mm_sub_ps(_mm_load_ps(n.lo[1]), r.oy), r.idy), y2 = _mm_mul_ps(_mm_sub_ps(_mm_load_ps(n.hi[1]), r.oy), r.idy);
    const __m128 z1 = _mm_mul_ps(_mm_sub_ps(_mm_load_ps(n.lo[2]), r.oz), r.idz), z2 = _mm_mul_ps(_mm_sub_ps(_mm_load_ps(n.hi[2]), r.oz), r.idz);
    const __m128 tn = _mm_max_ps(_mm_max_ps(_mm_min_ps(x1, x2), _mm_min_ps(y1, y2)), _mm_max_ps(_mm_min_ps(z1, z2), _mm_set1_ps(tmin)));
    const __m128 tf = _mm_min_ps(_mm_min_ps(_mm_max_ps(x1, x2), _mm_max_ps(y1, y2)), _mm_min_ps(_mm_max_ps(z1, z2), _mm_set1_ps(tmax)));
    tn4 = tn;
    return _mm_movemask_ps(_mm_cmple_ps(tn, _mm_mul_ps(tf, _mm_set1_ps(1.0000005f)))) & (int)n.valid;        /* unused slots never hit */
}
/* rt3m::ray_triangle for the four triangles of a leaf block, operation for operation:
 *   cross(a, b) = (fma(a.y, b.z, -(a.z b.y)), fma(a.z, b.x, -(a.x b.z)), fma(a.x, b.y, -(a.y b.x))),  dot(a, b) = fma(a.z, b.z, fma(a.y, b.y, a.x b.x)) */
inline __m128 neg4(__m128 a) { return _mm_xor_ps(a, _mm_set1_ps(-0.0f)); }
inline __m128 dot4(__m128 ax, __m128 ay, __m128 az, __m128 bx, __m128 by, __m128 bz) { return _mm_fmadd_ps(az, bz, _mm_fmadd_ps(ay, by, _mm_mul_ps(ax, bx))); }
inline int ray_triangle4(const Leaf4& lf, const Ray4& r, float t[4], float u[4], float v[4]) {
    const __m128 e1x = _mm_load_ps(lf.e1[0]), e1y = _mm_load_ps(lf.e1[1]), e1z = _mm_load_ps(lf.e1[2]);
    const __m128 e2x = _mm_load_ps(lf.e2[0]), e2y = _mm_load_ps(lf.e2[1]), e2z = _mm_load_ps(lf.e2[2]);
    const __m128 px = _mm_fmadd_ps(r.dy, e2z, neg4(_mm_mul_ps(r.dz, e2y))), py = _mm_fmadd_ps(r.dz, e2x, neg4(_mm_mul_ps(r.dx, e2z))),
                 pz = _mm_fmadd_ps(r.dx, e2y, neg4(_mm_mul_ps(r.dy, e2x)));                                   /* pvec = cross(d, e2) */
    const __m128 det = dot4(e1x, e1y, e1z, px, py, pz);
    const __m128 nonzero = _mm_cmpneq_ps(det, _mm_setzero_ps());
    const __m128 inv = _mm_div_ps(_mm_set1_ps(1.0f), det);
    const __m128 tx = _mm_sub_ps(r.ox, _mm_load_ps(lf.v0[0])), ty = _mm_sub_ps(r.oy, _mm_load_ps(lf.v0[1])), tz = _mm_sub_ps(r.oz, _mm_load_ps(lf.v0[2]));
    const __m128 uu = _mm_mul_ps(dot4(tx, ty, tz, px, py, pz), inv);
    const __m128 qx = _mm_fmadd_ps(ty, e1z, neg4(_mm_mul_ps(tz, e1y))), qy = _mm_fmadd_ps(tz, e1x, neg4(_mm_mul_ps(tx, e1z))),
                 qz = _mm_fmadd_ps(tx, e1y, neg4(_mm_mul_ps(ty, e1x)));                                       /* qvec = cross(tvec, e1) */
    const __m128 vv = _mm_mul_ps(dot4(r.dx, r.dy, r.dz, qx, qy, qz), inv);
    const __m128 tt = _mm_mul_ps(dot4(e2x, e2y, e2z, qx, qy, qz), inv);
    const __m128 zero = _mm_setzero_ps();
    const __m128 ok = _mm_and_ps(_mm_and_ps(nonzero, _mm_cmpge_ps(uu, zero)), _mm_and_ps(_mm_cmpge_ps(vv, zero), _mm_cmple_ps(_mm_add_ps(uu, vv), _mm_set1_ps(1.0f))));
    _mm_storeu_ps(t, tt); _mm_storeu_ps(u, uu); _mm_storeu_ps(v, vv);
    return _mm_movemask_ps(ok);
}

rt3_hit closest_bvh4(const orc_scene& s, const rt3_ray& r) {
    rt3_hit h; h.t = r.tmax; h.u = 0.f; h.v = 0.f; h.prim = RT3_MISS;
    const vec3 o = v3(r.o), d = v3(r.d);
    const Ray4 r4 = splat_ray(o, d, v3(safe_inv(d.x), safe_inv(d.y), safe_inv(d.z)));
    struct Entry { uint32_t ref; float tn; };
    Entry stack[192]; int sp = 0; stack[sp++] = {s.root4, 0.0f};
    while (sp) {
        const Entry e = stack[--sp];
        if (e.tn > h.t) continue;                             /* ties are kept: tn <= t */
        if (e.ref & REF_LEAF) {
            const Leaf4& lf = s.leaves4[e.ref & ~REF_LEAF];
            float t[4], u[4], v[4];
            int m = ray_triangle4(lf, r4, t, u, v);
            while (m) {
                const int k = __builtin_ctz(m); m &= m - 1;
                if (t[k] > r.tmin && t[k] < r.tmax && closer_hit(t[k], lf.id[k], h.t, h.prim)) { h.t = t[k]; h.u = u[k]; h.v = v[k]; h.prim = lf.id[k]; }
            }
            continue;
        }
        const Node4& n = s.nodes4[e.ref];
        __m128 tn4;
        int m = slab4(n, r4, r.tmin, h.t, tn4);
        if (!m) continue;
        float tn[4]; _mm_storeu_ps(tn, tn4);
        if (sp + 4 > 192) abort();
        const int base = sp;
        while (m) { const int k = __builtin_ctz(m); m &= m - 1; stack[sp++] = {n.child[k], tn[k]}; }
        for (int i = base + 1; i < sp; ++i) {                 /* farthest first on the stack: the nearest child is popped next */
            const Entry x = stack[i]; int j = i;
            while (j > base && stack[j - 1].tn < x.tn) { stack[j] = stack[j - 1]; --j; }
            stack[j] = x;
        }
    }
    return h;
}
bool any_bvh4(const orc_scene& s, const rt3_ray& r) {
    const vec3 o = v3(r.o), d = v3(r.d);
    const Ray4 r4 = splat_ray(o, d, v3(safe_inv(d.x), safe_inv(d.y), safe_inv(d.z)));
    uint32_t stack[192]; int sp = 0; stack[sp++] = s.root4;
    while (sp) {
        const uint32_t ref = stack[--sp];
        if (ref & REF_LEAF) {
            const Leaf4& lf = s.leaves4[ref & ~REF_LEAF];
            float t[4], u[4], v[4];
            int m = ray_triangle4(lf, r4, t, u, v);
            while (m) { const int k = __builtin_ctz(m); m &= m - 1; if (t[k] > r.tmin && t[k] < r.tmax) return true; }
            continue;
        }
        const Node4& n = s.nodes4[ref];
        __m128 tn4;
        int m = slab4(n, r4, r.tmin, r.tmax, tn4);
        if (sp + 4 > 192) abort();
        while (m) { const int k = __builtin_ctz(m); m &= m - 1; stack[sp++] = n.child[k]; }
    }
    return false;
}
/* the passes and the default query mode: BVH4 + SSE when it was built, else the scalar BVH2 */
rt3_hit closest_bvh(const orc_scene& s, const rt3_ray& r) { return s.root4 != REF_NONE ? closest_bvh4(s, r) : closest_bvh2(s, r); }
bool any_bvh(const orc_scene& s, const rt3_ray& r) { return s.root4 != REF_NONE ? any_bvh4(s, r) : any_bvh2(s, r); }

inline rt3_ray make_ray(vec3 o, vec3 d, float tmin, float tmax) {
    rt3_ray r; r.o[0] = o.x; r.o[1] = o.y; r.o[2] = o.z; r.d[0] = d.x; r.d[1] = d.y; r.d[2] = d.z; r.tmin = tmin; r.tmax = tmax; return r;
}

}  // namespace

extern "C" {

orc_scene* orc_scene_create(const rt3_vertex* v, uint32_t nv, const uint32_t* idx, uint32_t ni, const rt3_geometry_info* g, uint32_t ng,
                            const uint32_t* inst_first_geo, const uint32_t* inst_geo_count, const float* inst_xform, uint32_t n_inst) {
    orc_scene* s = new orc_scene();
    s->vertices.assign(v, v + nv); s->indices.assign(idx, idx + ni); s->geometries.assign(g, g + ng);
    if (n_inst == 0) {
        s->inst_first = {0}; s->inst_count = {ng};
        s->inst_xform = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0};
    } else {
        s->inst_first.assign(inst_first_geo, inst_first_geo + n_inst);
        s->inst_count.assign(inst_geo_count, inst_geo_count + n_inst);
        s->inst_xform.assign(inst_xform, inst_xform + 12 * (size_t)n_inst);
    }
    build_srgb_lut(s->srgb_lut);
    return s;
}
void orc_scene_destroy(orc_scene* s) { delete s; }
void orc_scene_set_texture(orc_scene* s, uint32_t slot, uint32_t w, uint32_t h, const uint8_t* rgba8) {
    if (s->textures.size() <= slot) { s->textures.resize(slot + 1); s->tex_dim.resize(2 * (slot + 1), 0); }
    s->textures[slot].assign(rgba8, rgba8 + 4 * (size_t)w * h);
    s->tex_dim[2 * slot] = w; s->tex_dim[2 * slot + 1] = h;
    s->tex_ptr.resize(s->textures.size());
    for (size_t i = 0; i < s->textures.size(); ++i) s->tex_ptr[i] = s->textures[i].empty() ? nullptr : s->textures[i].data();
}
void orc_scene_set_environment(orc_scene* s, uint32_t w, uint32_t h, const float* rgb) {
    s->env.assign(rgb, rgb + 3 * (size_t)w * h); s->env_w = w; s->env_h = h;
    s->env_func.resize((size_t)w * h); s->env_marginal.resize(h + 1); s->env_conditional.resize((size_t)h * (w + 1));
    build_env_tables(rgb, w, h, s->env_func.data(), s->env_marginal.data(), s->env_conditional.data(), &s->env_total);
}
void orc_scene_set_bluenoise(orc_scene* s, uint32_t w, uint32_t h, const uint8_t* rgba8) { s->bn.assign(rgba8, rgba8 + 4 * (size_t)w * h); s->bn_w = w; s->bn_h = h; }

double orc_scene_build(orc_scene* s) {
    auto t0 = std::chrono::steady_clock::now();
    /* segment table: instances in push order, geometries in range order */
    s->seg_first_prim.clear(); s->seg_geometry.clear(); s->seg_instance.clear();
    uint32_t total = 0;
    for (uint32_t i = 0; i < s->inst_first.size(); ++i)
        for (uint32_t g = s->inst_first[i]; g < s->inst_first[i] + s->inst_count[i]; ++g) {
            s->seg_first_prim.push_back(total); s->seg_geometry.push_back(g); s->seg_instance.push_back(i);
            total += s->geometries[g].index_count / 3;
        }
    s->seg_first_prim.push_back(total);
    SceneView sv = s->view();
    s->tris.resize(total);
    s->bounds.reset();
    for (uint32_t seg = 0; seg < s->seg_geometry.size(); ++seg)
        for (uint32_t p = s->seg_first_prim[seg]; p < s->seg_first_prim[seg + 1]; ++p) {
            vec3 a, b, c;
            world_triangle(sv, seg, p - s->seg_first_prim[seg], a, b, c);
            WorldTri t; t.v0 = a; t.e1 = b - a; t.e2 = c - a;
            s->tris[p] = t;
            s->bounds.grow(a); s->bounds.grow(b); s->bounds.grow(c);
        }
    vec3 e = s->bounds.hi - s->bounds.lo;
    s->pad = total ? max_(e.x, max_(e.y, e.z)) * (1.0f / 65536.0f) : 0.f;
    Builder b(*s); b.build();
    if (!getenv("RT3_ORACLE_BVH2")) build_bvh4(*s);       /* RT3_ORACLE_BVH2=1: A/B of the timed CPU arm against the scalar BVH2 */
    return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
}
uint32_t orc_num_triangles(const orc_scene* s) { return (uint32_t)s->tris.size(); }
void orc_scene_bounds(const orc_scene* s, float lo[3], float hi[3]) {
    lo[0] = s->bounds.lo.x; lo[1] = s->bounds.lo.y; lo[2] = s->bounds.lo.z;
    hi[0] = s->bounds.hi.x; hi[1] = s->bounds.hi.y; hi[2] = s->bounds.hi.z;
}
void orc_resolve_prim(const orc_scene* s, uint32_t prim, uint32_t* instance, uint32_t* geometry, uint32_t* primitive) {
    SceneView sv = s->view();
    uint32_t seg = find_segment(sv, prim);
    *instance = s->seg_instance[seg]; *geometry = s->seg_geometry[seg]; *primitive = prim - s->seg_first_prim[seg];
}
/* the records the traversal tests: (v0, e1, e2) of every triangle in global id order, and the oracle's scene view (host pointers, valid
 * while the scene lives) — for tests/emu/pass_emu.cpp, which runs the product's kernels on the CPU against this oracle */
void orc_world_edges(const orc_scene* s, float* out9) {
    for (size_t i = 0; i < s->tris.size(); ++i) {
        const WorldTri& t = s->tris[i];
        float* o = out9 + 9 * i;
        o[0] = t.v0.x; o[1] = t.v0.y; o[2] = t.v0.z; o[3] = t.e1.x; o[4] = t.e1.y; o[5] = t.e1.z; o[6] = t.e2.x; o[7] = t.e2.y; o[8] = t.e2.z;
    }
}
uint64_t orc_scene_view(const orc_scene* s, void* out, uint64_t capacity) {
    SceneView v = s->view();
    if (capacity >= sizeof(v)) memcpy(out, &v, sizeof(v));
    return sizeof(v);
}
void orc_world_triangles(const orc_scene* s, float* out9) {
    for (size_t i = 0; i < s->tris.size(); ++i) {
        const WorldTri& t = s->tris[i];
        vec3 b = t.v0 + t.e1, c = t.v0 + t.e2;
        float* o = out9 + 9 * i;
        o[0] = t.v0.x; o[1] = t.v0.y; o[2] = t.v0.z; o[3] = b.x; o[4] = b.y; o[5] = b.z; o[6] = c.x; o[7] = c.y; o[8] = c.z;
    }
}

void orc_trace_closest(const orc_scene* s, const rt3_ray* rays, uint64_t n, rt3_hit* out, int mode, int threads) {
    parallel_for(n, threads, mode ? 4096 : 64, [&](uint64_t b, uint64_t e) {
        for (uint64_t i = b; i < e; ++i) out[i] = mode == 2 ? closest_bvh2(*s, rays[i]) : (mode ? closest_bvh(*s, rays[i]) : closest_brute(*s, rays[i]));
    });
}
void orc_trace_any(const orc_scene* s, const rt3_ray* rays, uint64_t n, uint32_t* out_bits, int mode, int threads) {
    uint64_t words = (n + 31) / 32;
    parallel_for(words, threads, mode ? 128 : 2, [&](uint64_t b, uint64_t e) {
        for (uint64_t w = b; w < e; ++w) {
            uint32_t bits = 0;
            for (uint64_t i = w * 32; i < std::min(n, w * 32 + 32); ++i)
                if (mode == 2 ? any_bvh2(*s, rays[i]) : (mode ? any_bvh(*s, rays[i]) : any_brute(*s, rays[i]))) bits |= 1u << (i & 31);
            out_bits[w] = bits;
        }
    });
}

/* shaders/old/gbuffer.slang:8-20 */
void orc_pass_gbuffer(const orc_scene* s, const rt3_params* p, rt3_rect rc, uint32_t* gbuffer, float* depth, uint32_t* opt_ids, float* opt_bary,
                      int threads) {
    SceneView sv = s->view();
    uint32_t pitch = (uint32_t)p->window_size[0];
    parallel_for(rc.h, threads, 4, [&](uint64_t y0, uint64_t y1) {
        for (uint32_t y = rc.y0 + (uint32_t)y0; y < rc.y0 + (uint32_t)y1; ++y)
            for (uint32_t x = rc.x0; x < rc.x0 + rc.w; ++x) {
                size_t px = (size_t)y * pitch + x;
                rt3_ray ray = setup_primary_ray(x, y, *p);     /* gbuffer.slang:10 */
                rt3_hit h = closest_bvh(*s, ray);               /* gbuffer.slang:12 */
                if (h.prim == RT3_MISS) {                       /* gbuffer.slang:14-16: gbuffer colour left untouched */
                    depth[px] = RT3_BACKGROUND_DEPTH;
                    if (opt_ids) { opt_ids[2 * px] = RT3_MISS; opt_ids[2 * px + 1] = RT3_MISS; }
                    if (opt_bary) { opt_bary[2 * px] = 0.f; opt_bary[2 * px + 1] = 0.f; }
                } else {                                        /* gbuffer.slang:17-19 */
                    uint32_t geo;
                    uint4_ g = gbuffer_pack(hit_info(sv, h.prim, h.u, h.v, &geo));
                    gbuffer[4 * px] = g.x; gbuffer[4 * px + 1] = g.y; gbuffer[4 * px + 2] = g.z; gbuffer[4 * px + 3] = g.w;
                    depth[px] = h.t;
                    if (opt_ids) { opt_ids[2 * px] = h.prim; opt_ids[2 * px + 1] = geo; }
                    if (opt_bary) { opt_bary[2 * px] = h.u; opt_bary[2 * px + 1] = h.v; }
                }
            }
    });
}

/* Optional capture of the secondary rays of orc_pass_reference (for tools/emu_warp_stats.py --frame: the warp emulation of the product's
 * traversal kernel is then fed the frame's own bounce rays): rays[b * pixels + px] = the ray traced after loop index b, valid[...] = 1. */
static rt3_ray* g_cap_rays = nullptr; static uint8_t* g_cap_valid = nullptr; static uint64_t g_cap_pixels = 0; static uint32_t g_cap_bounces = 0;
void orc_set_ray_capture(rt3_ray* rays, uint8_t* valid, uint64_t pixels, uint32_t bounces) { g_cap_rays = rays; g_cap_valid = valid; g_cap_pixels = pixels; g_cap_bounces = bounces; }

/* shaders/old/refrence_mode.slang:14-66 */
void orc_pass_reference(const orc_scene* s, const rt3_params* p, rt3_rect rc, const uint32_t* gbuffer, const float* depth, const float* prev_light,
                        float* light, int threads, rt3_counters* opt_counters) {
    SceneView sv = s->view();
    uint32_t pitch = (uint32_t)p->window_size[0];
    std::atomic<uint64_t> n_secondary(0), n_started(0), n_escaped(0), n_shadow(0);
    parallel_for(rc.h, threads, 4, [&](uint64_t y0, uint64_t y1) {
        uint64_t secondary = 0, started = 0, escaped = 0, shadow = 0;
        for (uint32_t y = rc.y0 + (uint32_t)y0; y < rc.y0 + (uint32_t)y1; ++y)
            for (uint32_t x = rc.x0; x < rc.x0 + rc.w; ++x) {
                size_t px = (size_t)y * pitch + x;
                float gdepth = depth[px];
                if (gdepth == RT3_BACKGROUND_DEPTH) continue;                          /* :18-21 */
                uint4_ packed = {gbuffer[4 * px], gbuffer[4 * px + 1], gbuffer[4 * px + 2], gbuffer[4 * px + 3]};
                GbufferData gdata = gbuffer_unpack(packed);                             /* :23 */
                RandomSamplerState rng = rng_init(x, y, p->frame);                      /* :25 */
                vec3 radiance = v3(0.f, 0.f, 0.f);
                for (uint32_t i = 0; i < p->samples; ++i) {                             /* :28 */
                    vec3 ray_color = v3(1.f, 1.f, 1.f);
                    rt3_ray pr = setup_primary_ray(x, y, *p);                           /* :30 */
                    vec3 ro = v3(pr.o), rd = v3(pr.d);
                    const float tmin = 0.001f, tmax = pr.tmax;                          /* :31 */
                    GbufferData surface = gdata;
                    float hit_t = gdepth;
                    bool missed = false;                                                /* payload.geometryIndex = 0, :34 */
                    started++;
                    float pdf_brdf = 0.f;                                               /* pdf of the direction the current ray was sampled with */
                    for (uint32_t b = 0; b < p->bounces; ++b) {                         /* :36 */
                        if (missed) {                                                   /* :37-40 */
                            if (p->flags & RT3_FLAG_ENV) radiance = radiance + ray_color * (sample_environment(sv, rd) * escape_weight(sv, p->flags, rd, pdf_brdf));
                            break;
                        }
                        vec2 urand = sample_2d(sv, p->flags, x, y, p->frame, rng);     /* :43 */
                        const bool layered = (p->flags & RT3_FLAG_SPECULAR) != 0;         /* DiffuseBrdf + SpecularBrdf (brdf.slang:95-318) */
                        LayeredMaterial lm = LayeredMaterial();
                        bool lobe_ok = true;
                        const vec3 thr_before = ray_color;
                        PathVertexOut nv = layered ? bounce_step_layered(surface, ro, rd, hit_t, urand, radiance, ray_color, lm, lobe_ok, &pdf_brdf)
                                                   : bounce_step(surface, ro, rd, hit_t, urand, radiance, ray_color, &pdf_brdf); /* :44-51 */
                        ro = nv.origin; rd = nv.dir;
                        if (!lobe_ok) break;                                            /* invalid specular sample (brdf.slang:226-228): path ends */
                        if (nee_enabled(p->flags) && b != p->bounces - 1) {             /* next-event estimation (no reference counterpart) */
                            vec2 ul = sample_2d(sv, p->flags, x, y, p->frame, rng);
                            NeeSample ns = layered ? nee_sample_layered(sv, lm, thr_before, ul) : nee_sample(sv, surface, ray_color, ul);
                            if (ns.valid) {
                                shadow++;
                                if (!any_bvh(*s, make_ray(ro, ns.dir, tmin, tmax))) radiance = radiance + ns.contribution;
                            }
                        }
                        if (b != p->bounces - 1) {                                      /* :53-56 */
                            rt3_hit h = closest_bvh(*s, make_ray(ro, rd, tmin, tmax));
                            if (g_cap_rays && b < g_cap_bounces && px < g_cap_pixels) { g_cap_rays[b * g_cap_pixels + px] = make_ray(ro, rd, tmin, tmax); g_cap_valid[b * g_cap_pixels + px] = 1; }
                            secondary++;
                            if (h.prim == RT3_MISS) { missed = true; escaped++; }
                            else { surface = hit_info(sv, h.prim, h.u, h.v); hit_t = h.t; }
                        }
                    }
                }
                float ns = float(p->samples);                                           /* :59 (radiance /= samples) */
                radiance = v3(radiance.x / ns, radiance.y / ns, radiance.z / ns);
                vec3 prev = v3(0.f, 0.f, 0.f);
                if (p->blendfactor < 1.0f && prev_light) prev = v3(prev_light[4 * px], prev_light[4 * px + 1], prev_light[4 * px + 2]);
                vec3 o = temporal_blend(prev, radiance, p->blendfactor);                /* :61-65 */
                light[4 * px] = o.x; light[4 * px + 1] = o.y; light[4 * px + 2] = o.z; light[4 * px + 3] = 0.f;
            }
        n_secondary += secondary; n_started += started; n_escaped += escaped; n_shadow += shadow;
    });
    if (opt_counters) {
        opt_counters->rays_secondary += n_secondary; opt_counters->paths_started += n_started; opt_counters->paths_escaped += n_escaped;
        opt_counters->rays_shadow += n_shadow;
    }
}

/* shaders/old/postprocess.slang:90-112 */
void orc_pass_postprocess(const orc_scene* s, const rt3_params* p, rt3_rect rc, const float* depth, const float* in, float* out, int threads) {
    SceneView sv = s->view();
    uint32_t pitch = (uint32_t)p->window_size[0];
    parallel_for(rc.h, threads, 8, [&](uint64_t y0, uint64_t y1) {
        for (uint32_t y = rc.y0 + (uint32_t)y0; y < rc.y0 + (uint32_t)y1; ++y)
            for (uint32_t x = rc.x0; x < rc.x0 + rc.w; ++x) {
                size_t px = (size_t)y * pitch + x;
                vec3 c = postprocess_pixel(sv, *p, x, y, depth[px], v3(in[4 * px], in[4 * px + 1], in[4 * px + 2]));
                out[4 * px] = c.x; out[4 * px + 1] = c.y; out[4 * px + 2] = c.z; out[4 * px + 3] = 1.0f;
            }
    });
}

/* postprocess.slang:90-112 + the conversion to B8G8R8A8_UNORM of the swapchain blit (swapchain.rs:27-48) */
void orc_pass_present_bgra8(const orc_scene* s, const rt3_params* p, rt3_rect rc, const float* depth, const float* in, uint32_t* out, int threads) {
    SceneView sv = s->view();
    uint32_t pitch = (uint32_t)p->window_size[0];
    parallel_for(rc.h, threads, 8, [&](uint64_t y0, uint64_t y1) {
        for (uint32_t y = rc.y0 + (uint32_t)y0; y < rc.y0 + (uint32_t)y1; ++y)
            for (uint32_t x = rc.x0; x < rc.x0 + rc.w; ++x) {
                size_t px = (size_t)y * pitch + x;
                out[px] = pack_bgra8(postprocess_pixel(sv, *p, x, y, depth[px], v3(in[4 * px], in[4 * px + 1], in[4 * px + 2])));
            }
    });
}

/* ---- screen-space probes (row F4) ----------------------------------------------------------------
 * shaders/old/trace_probes.slang: one ray per atlas texel; directions == NULL selects the uniform assignment. */
void orc_pass_probe_trace(const orc_scene* s, const rt3_params* p, const uint32_t* gbuffer, const float* depth, const uint16_t* directions,
                          const float* prev_atlas, float* atlas, int threads) {
    (void)gbuffer;
    SceneView sv = s->view();
    const uint32_t W = (uint32_t)p->window_size[0], aw = probe_count_x(*p) * 8u, ah = probe_count_y(*p) * 8u;
    parallel_for(ah, threads, 8, [&](uint64_t y0, uint64_t y1) {
        for (uint32_t ay = (uint32_t)y0; ay < (uint32_t)y1; ++ay)
            for (uint32_t ax = 0; ax < aw; ++ax) {
                const size_t ai = (size_t)ay * aw + ax;
                const float pd = depth[(size_t)(ay / 8u * 16u) * W + ax / 8u * 16u];
                const uint32_t word = directions ? directions[ai] : probe_uniform_direction(ax, ay);
                ProbeRay pr = probe_setup_ray(*p, ax, ay, pd, word);
                if (pr.background) { atlas[4 * ai] = 0.f; atlas[4 * ai + 1] = 0.f; atlas[4 * ai + 2] = 0.f; atlas[4 * ai + 3] = RT3_BACKGROUND_DEPTH; continue; }   /* :28-31 */
                if (directions && (ay & 7u) == 0u && (ax & 7u) == 0u)                  /* :32, all clears of the probe before its first target write */
                    for (uint32_t k = 0; k < 64u; ++k) {
                        float* c = atlas + 4 * ((size_t)(ay + k / 8u) * aw + ax + k % 8u);
                        c[0] = c[1] = c[2] = c[3] = 0.f;
                    }
                rt3_hit h = closest_bvh(*s, pr.ray);
                const size_t ti = (size_t)pr.target_y * aw + pr.target_x;
                vec3 prev = v3(0.f, 0.f, 0.f);
                if (p->blendfactor < 1.0f && prev_atlas) prev = v3(prev_atlas[4 * ai], prev_atlas[4 * ai + 1], prev_atlas[4 * ai + 2]);
                vec4 t = probe_texel(sv, *p, pr.ray, h, prev);
                atlas[4 * ti] = t.x; atlas[4 * ti + 1] = t.y; atlas[4 * ti + 2] = t.z; atlas[4 * ti + 3] = t.w;
            }
    });
}
/* shaders/old/structured_importance_sampling.slang:19-72 (the live lines; see rt3_shared_math.h): one direction word per atlas texel */
void orc_pass_probe_sis(const rt3_params* p, const uint32_t* gbuffer, uint16_t* out, int threads) {
    const uint32_t W = (uint32_t)p->window_size[0], pw = probe_count_x(*p), ph = probe_count_y(*p), aw = pw * 8u;
    parallel_for(ph, threads, 1, [&](uint64_t y0, uint64_t y1) {
        for (uint32_t pcy = (uint32_t)y0; pcy < (uint32_t)y1; ++pcy)
            for (uint32_t pcx = 0; pcx < pw; ++pcx) {
                vec3 normals[256];
                float pdf[64], sorted[64];
                for (uint32_t ti = 0; ti < 64u; ++ti)                                  /* :25-31 */
                    for (uint32_t y = 0; y < 2u; ++y)
                        for (uint32_t x = 0; x < 2u; ++x) {
                            uint32_t px, py;
                            probe_sis_pixel(pcx, pcy, ti, x, y, px, py);
                            normals[ti * 4u + y * 2u + x] = unpack_normal_11_10_11(gbuffer[4u * ((size_t)py * W + px) + 1u]);
                        }
                for (uint32_t ti = 0; ti < 64u; ++ti) sorted[ti] = pdf[ti] = probe_sis_pdf(normals, probe_sis_direction(ti));   /* :34-46 */
                for (uint32_t k = 2; k <= 64u; k *= 2u)                                /* math.slang:142-163, the steps in lock step */
                    for (uint32_t j = k / 2u; j > 0u; j /= 2u)
                        for (uint32_t i = 0; i < 64u; ++i) {
                            const uint32_t l = i ^ j;
                            if (l > i && bitonic_exchange(i, k, j, sorted[i], sorted[l])) { const float t = sorted[i]; sorted[i] = sorted[l]; sorted[l] = t; }
                        }
                for (uint32_t ti = 0; ti < 64u; ++ti)                                  /* :55-70 */
                    out[(size_t)(pcy * 8u + ti / 8u) * aw + pcx * 8u + ti % 8u] = (uint16_t)probe_sis_word(pdf[ti], sorted[0], ti);
            }
    });
}
/* shaders/old/spherical_harmonic_conversion.slang */
void orc_pass_probe_sh(const rt3_params* p, const float* atlas, float* sh, int threads) {
    const uint32_t pw = probe_count_x(*p), ph = probe_count_y(*p), aw = pw * 8u;
    parallel_for(ph, threads, 1, [&](uint64_t y0, uint64_t y1) {
        for (uint32_t py = (uint32_t)y0; py < (uint32_t)y1; ++py)
            for (uint32_t px = 0; px < pw; ++px)
                for (uint32_t c = 0; c < 3; ++c) {
                    sh3 r = probe_project_sh(atlas, aw, px, py, c);
                    memcpy(sh + 9u * (size_t)probe_sh_index(px, py, c), r.c, sizeof(r.c));
                }
    });
}
/* shaders/old/interpolate_probes.slang */
void orc_pass_probe_interpolate(const rt3_params* p, const uint32_t* gbuffer, const float* depth, const float* sh, float* light, int threads) {
    const uint32_t W = (uint32_t)p->window_size[0], H = (uint32_t)p->window_size[1];
    std::vector<uint8_t> failed((size_t)W * H, 0);
    std::vector<uint32_t> fx((size_t)W * H), fy((size_t)W * H);
    parallel_for(H, threads, 4, [&](uint64_t y0, uint64_t y1) {
        for (uint32_t y = (uint32_t)y0; y < (uint32_t)y1; ++y)
            for (uint32_t x = 0; x < W; ++x) {
                uint32_t ox, oy; vec4 v;
                int k = probe_interpolate_pixel(*p, x, y, gbuffer, depth, sh, ox, oy, v);
                const size_t i = (size_t)y * W + x;
                if (k == 1) { light[4 * i] = v.x; light[4 * i + 1] = v.y; light[4 * i + 2] = v.z; light[4 * i + 3] = v.w; }
                else if (k == 2) { failed[i] = 1; fx[i] = ox; fy[i] = oy; }
            }
    });
    for (size_t i = 0; i < failed.size(); ++i)
        if (failed[i]) { const size_t t = (size_t)fy[i] * W + fx[i]; light[4 * t] = 1.f; light[4 * t + 1] = 0.f; light[4 * t + 2] = 0.f; light[4 * t + 3] = 1.f; }
}

/* ---- known-answer hooks -------------------------------------------------------------------- */
uint32_t orc_hash_jenkins(uint32_t a) { return hash_jenkins(a); }
uint32_t orc_zcurve(uint32_t x, uint32_t y) { return zcurve_to_linear_index(x, y); }
uint32_t orc_murmur3(uint32_t seed, uint32_t index) { return murmur3_at(seed, index); }
float orc_uniform_float(uint32_t seed, uint32_t index) { return u32_to_unit_float(murmur3_at(seed, index)); }
uint32_t orc_rng_seed(uint32_t px, uint32_t py, uint32_t frame) { return rng_init(px, py, frame).seed; }
void orc_onb(const float n[3], float out9[9]) {
    onb t = build_orthonormal_basis(v3(n));
    out9[0] = t.b1.x; out9[1] = t.b1.y; out9[2] = t.b1.z; out9[3] = t.b2.x; out9[4] = t.b2.y; out9[5] = t.b2.z; out9[6] = t.n.x; out9[7] = t.n.y; out9[8] = t.n.z;
}
void orc_diffuse_sample(float ux, float uy, float out3[3]) { vec3 w = diffuse_sample_wi(ux, uy); out3[0] = w.x; out3[1] = w.y; out3[2] = w.z; }
void orc_sincos_2pi(float u, float out2[2]) { sincos_2pi(u, out2[0], out2[1]); }
float orc_atan2(float y, float x) { return atan2_(y, x); }
float orc_asin(float y) { return asin_(y); }
float orc_log2(float x) { return log2_(x); }
float orc_exp2(float x) { return exp2_(x); }
void orc_equirect_uv(const float d[3], float out2[2]) { vec2 r = direction_to_equirect_uv(v3(d)); out2[0] = r.x; out2[1] = r.y; }
uint32_t orc_f32_to_f16(float f) { return f32_to_f16(f); }
float orc_f16_to_f32(uint32_t h) { return f16_to_f32(h); }
uint32_t orc_rgb9e5_pack(const float rgb[3]) { return float3_to_rgb9e5(v3(rgb)); }
void orc_rgb9e5_unpack(uint32_t v, float out3[3]) { vec3 r = rgb9e5_to_float3(v); out3[0] = r.x; out3[1] = r.y; out3[2] = r.z; }
uint32_t orc_pack_normal(const float n[3]) { return pack_normal_11_10_11(v3(n)); }
void orc_unpack_normal(uint32_t p, float out3[3]) { vec3 r = unpack_normal_11_10_11(p); out3[0] = r.x; out3[1] = r.y; out3[2] = r.z; }
uint32_t orc_pack_color(const float c[3]) { return pack_color_888(v3(c)); }
void orc_unpack_color(uint32_t p, float out3[3]) { vec3 r = unpack_color_888(p); out3[0] = r.x; out3[1] = r.y; out3[2] = r.z; }
void orc_gbuffer_pack(const float albedo[3], const float emissive[3], const float normal[3], float roughness, float metalness, uint32_t out4[4]) {
    GbufferData g; g.albedo = v3(albedo); g.emissive = v3(emissive); g.normal = v3(normal); g.roughness = roughness; g.metalness = metalness;
    uint4_ r = gbuffer_pack(g); out4[0] = r.x; out4[1] = r.y; out4[2] = r.z; out4[3] = r.w;
}
void orc_gbuffer_unpack(const uint32_t in4[4], float out11[11]) {
    uint4_ d = {in4[0], in4[1], in4[2], in4[3]};
    GbufferData g = gbuffer_unpack(d);
    out11[0] = g.albedo.x; out11[1] = g.albedo.y; out11[2] = g.albedo.z; out11[3] = g.emissive.x; out11[4] = g.emissive.y; out11[5] = g.emissive.z;
    out11[6] = g.normal.x; out11[7] = g.normal.y; out11[8] = g.normal.z; out11[9] = g.roughness; out11[10] = g.metalness;
}
void orc_primary_ray(const rt3_params* p, uint32_t px, uint32_t py, rt3_ray* out) { *out = setup_primary_ray(px, py, *p); }
int orc_ray_triangle(const float o[3], const float d[3], const float v0[3], const float v1[3], const float v2[3], float out_tuv[3]) {
    vec3 a = v3(v0);
    return ray_triangle(v3(o), v3(d), a, v3(v1) - a, v3(v2) - a, out_tuv[0], out_tuv[1], out_tuv[2]) ? 1 : 0;
}
void orc_r2(uint32_t i, float out2[2]) { vec2 r = r2_sequence(i); out2[0] = r.x; out2[1] = r.y; }
float orc_vdc(uint32_t bits) { return radical_inverse_vdc(bits); }
void orc_sample_2d(const orc_scene* s, uint32_t flags, uint32_t px, uint32_t py, uint32_t frame, uint32_t index, float out2[2]) {
    SceneView sv = s->view();
    RandomSamplerState r = rng_init(px, py, frame); r.index = index;
    vec2 u = sample_2d(sv, flags, px, py, frame, r); out2[0] = u.x; out2[1] = u.y;
}
void orc_agx(const float in3[3], float out3[3]) { vec3 c = agx_eotf(agx_look(agx(v3(in3)))); out3[0] = c.x; out3[1] = c.y; out3[2] = c.z; }
float orc_contract_probe(float a, float b, float c) { return contract_probe(a, b, c); }
void orc_octa_decode(float fx, float fy, float out3[3]) { vec2 f; f.x = fx; f.y = fy; vec3 d = octa_decode(f); out3[0] = d.x; out3[1] = d.y; out3[2] = d.z; }
void orc_octa_encode(const float n[3], float out2[2]) { vec2 e = octa_encode(v3(n)); out2[0] = e.x; out2[1] = e.y; }
void orc_sh3_evaluate(const float d[3], float out9[9]) { sh3 s = sh3_evaluate(v3(d)); memcpy(out9, s.c, sizeof(s.c)); }
void orc_sh3_unproject_cos_lobe(const float sh27[27], const float n[3], float out3[3]) {
    sh3 r, g, b; memcpy(r.c, sh27, 36); memcpy(g.c, sh27 + 9, 36); memcpy(b.c, sh27 + 18, 36);
    vec3 v = sh3_unproject_cos_lobe(r, g, b, v3(n)); out3[0] = v.x; out3[1] = v.y; out3[2] = v.z;
}
void orc_world_pos_from_depth(const rt3_params* p, uint32_t px, uint32_t py, float depth, float out3[3]) {
    vec3 v = world_pos_from_depth(depth, px, py, *p); out3[0] = v.x; out3[1] = v.y; out3[2] = v.z;
}
uint32_t orc_probe_sh_capacity(const rt3_params* p) { return probe_sh_capacity(*p); }
/* SpecularBrdf::sample / ::evaluate (brdf.slang:220-314).  out12 = value_over_pdf[3], value[3], pdf, transmission[3], (sample only) wi... */
int orc_specular_sample(float roughness, const float albedo[3], const float wo[3], float ux, float uy, float out14[14]) {
    vec2 u; u.x = ux; u.y = uy;
    BrdfSample r = specular_sample(roughness, v3(albedo), v3(wo), u);
    const float o[14] = {r.value_over_pdf.x, r.value_over_pdf.y, r.value_over_pdf.z, r.value.x, r.value.y, r.value.z, r.pdf, r.transmission_fraction.x,
                         r.transmission_fraction.y, r.transmission_fraction.z, r.wi.x, r.wi.y, r.wi.z, r.approx_roughness};
    memcpy(out14, o, sizeof(o));
    return brdf_sample_is_valid(r) ? 1 : 0;
}
void orc_specular_evaluate(float roughness, const float albedo[3], const float wo[3], const float wi[3], float out10[10]) {
    BrdfValue r = specular_evaluate(roughness, v3(albedo), v3(wo), v3(wi));
    const float o[10] = {r.value_over_pdf.x, r.value_over_pdf.y, r.value_over_pdf.z, r.value.x, r.value.y, r.value.z, r.pdf, r.transmission_fraction.x,
                         r.transmission_fraction.y, r.transmission_fraction.z};
    memcpy(out10, o, sizeof(o));
}
void orc_specular_dominant_direction(const float n[3], const float v[3], float roughness, float out3[3]) {
    vec3 r = specular_dominant_direction(v3(n), v3(v), roughness); out3[0] = r.x; out3[1] = r.y; out3[2] = r.z;
}
/* Layered material (RT3_FLAG_SPECULAR): one sample.  out8 = world wi[3], weight[3], solid-angle pdf of the mixture, p_spec.  Returns valid | specular << 1. */
int orc_layered_sample(const float albedo[3], const float normal[3], float roughness, float metalness, const float ray_d[3], float ux, float uy, float out8[8]) {
    GbufferData g; g.albedo = v3(albedo); g.normal = v3(normal); g.roughness = roughness; g.metalness = metalness; g.emissive = v3(0.f, 0.f, 0.f);
    LayeredMaterial m = layered_setup(g, v3(ray_d));
    vec2 u; u.x = ux; u.y = uy;
    LayeredSample ls = layered_sample(m, u);
    vec3 w = onb_to_world(m.t, ls.wi), val; float pdf = 0.f;
    if (ls.valid) layered_evaluate(m, ls.wi, val, pdf);
    const float o[8] = {w.x, w.y, w.z, ls.weight.x, ls.weight.y, ls.weight.z, pdf, m.p_spec};
    memcpy(out8, o, sizeof(o));
    return (ls.valid ? 1 : 0) | (ls.specular ? 2 : 0);
}
/* Layered material evaluated for a world-space direction: out4 = BRDF value[3] (projected-solid-angle convention), solid-angle pdf. */
void orc_layered_evaluate(const float albedo[3], const float normal[3], float roughness, float metalness, const float ray_d[3], const float wi_world[3], float out4[4]) {
    GbufferData g; g.albedo = v3(albedo); g.normal = v3(normal); g.roughness = roughness; g.metalness = metalness; g.emissive = v3(0.f, 0.f, 0.f);
    LayeredMaterial m = layered_setup(g, v3(ray_d));
    vec3 val; float pdf;
    layered_evaluate(m, world_to_onb(m.t, v3(wi_world)), val, pdf);
    out4[0] = val.x; out4[1] = val.y; out4[2] = val.z; out4[3] = pdf;
}

}  // extern "C"
