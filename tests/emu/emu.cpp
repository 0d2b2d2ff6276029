/*
 * emu.cpp — TEST INFRASTRUCTURE: runs the per-thread device logic of raytracer3_b200/csrc/bvh_logic.h
 * sequentially on the CPU (Karras hierarchy, refit, wide collapse + quantised encoding, wide traversal)
 * and compares it with a brute-force loop over the same 48-byte triangle records.  It lets the
 * `-m "not gpu"` suite debug the builder/traversal logic on a machine without a GPU.  It is not a
 * fallback: librt3.so neither contains nor calls this code.
 */
#include <algorithm>
#include <cstdio>
#include <cstring>
#include <numeric>
#include <vector>
#include "bvh_experiments.h"

using namespace rt3b;

struct EmuBvh {
    std::vector<Tri48> tris_in, tris_sorted, tris_out;
    std::vector<WideNode> nodes;
    uint32_t n_nodes = 0, n_tris = 0, levels = 0, n_refs = 0, ploc_rounds = 0;
    float sah[2] = {0, 0};
    float pad = 0;
    float lo[3], hi[3];
};

static int g_input_edges = 0;
static inline uint32_t mem_slot(uint32_t n, uint32_t ref) { return (ref & LEAF_FLAG) ? n - 1 + (ref & ~LEAF_FLAG) : ref; }

extern "C" {

/* split_beta > 0: triangle splitting (bvh_logic.h split_*): the hierarchy is built over references, n below is then their number */
static EmuBvh* build_lbvh(const float* tris9, uint32_t n, bool use_dp, float split_beta = 0.0f, uint32_t ploc_radius = 0, int treelet_passes = 0) {
    EmuBvh* b = new EmuBvh();
    b->tris_in.resize(n);
    for (int a = 0; a < 3; ++a) { b->lo[a] = 3.4e38f; b->hi[a] = -3.4e38f; }
    for (uint32_t i = 0; i < n; ++i) {
        const float* p = tris9 + 9 * (size_t)i;
        Tri48& t = b->tris_in[i];
        for (int a = 0; a < 3; ++a) {
            if (g_input_edges) {           /* input is (v0, e1, e2) already: the records of another builder, bit for bit */
                t.v0[a] = p[a]; t.e1[a] = p[3 + a]; t.e2[a] = p[6 + a];
                const float q[3] = {p[a], p[a] + p[3 + a], p[a] + p[6 + a]};
                for (int k = 0; k < 3; ++k) { b->lo[a] = std::min(b->lo[a], q[k]); b->hi[a] = std::max(b->hi[a], q[k]); }
                continue;
            }
            t.v0[a] = p[a]; t.e1[a] = p[3 + a] - p[a]; t.e2[a] = p[6 + a] - p[a];
            for (int k = 0; k < 3; ++k) { b->lo[a] = std::min(b->lo[a], p[3 * k + a]); b->hi[a] = std::max(b->hi[a], p[3 * k + a]); }
        }
        t.prim = i; t.pad0 = t.pad1 = 0;
    }
    float ext = n ? std::max(b->hi[0] - b->lo[0], std::max(b->hi[1] - b->lo[1], b->hi[2] - b->lo[2])) : 0.f;
    b->pad = ext * RT3_PAD_FRACTION;
    if (n == 0) {
        b->nodes.resize(1);
        memset(&b->nodes[0], 0, sizeof(WideNode));
        b->n_nodes = 1;
        return b;
    }
    /* references: one per triangle, or several with clipped boxes */
    std::vector<RefBox> refs;
    if (split_beta > 0.0f) {
        SplitGrid g;
        for (int a = 0; a < 3; ++a) { g.lo[a] = b->lo[a]; g.ext[a] = b->hi[a] - b->lo[a]; }
        std::vector<float> prio(n);
        const float inv_norm = split_inv_norm(g);
        uint64_t total = 0;                                    /* fixed point, as the device sums it with integer atomics */
        for (uint32_t i = 0; i < n; ++i) { prio[i] = split_priority(b->tris_in[i], REC_TRI_EDGES, g); total += split_priority_fixed(prio[i], inv_norm); }
        const float scale = split_scale(total, n, split_beta, inv_norm);
        for (uint32_t i = 0; i < n; ++i) {
            uint32_t s = split_count(prio[i], scale);
            split_triangle(b->tris_in[i], REC_TRI_EDGES, g, s, [&](const float* lo, const float* hi) {
                RefBox r; for (int a = 0; a < 3; ++a) { r.lo[a] = lo[a]; r.hi[a] = hi[a]; } r.tri = i; r.pad = 0; refs.push_back(r); });
        }
    } else {
        refs.resize(n);
        for (uint32_t i = 0; i < n; ++i) { tri_bounds(b->tris_in[i], 0.f, refs[i].lo, refs[i].hi); refs[i].tri = i; refs[i].pad = 0; }
    }
    b->n_refs = n = (uint32_t)refs.size();
    /* morton + sort */
    std::vector<uint64_t> keys(n);
    std::vector<uint32_t> order(n);
    for (uint32_t i = 0; i < n; ++i) {
        const float *l = refs[i].lo, *h = refs[i].hi;
        float c[3];
        for (int a = 0; a < 3; ++a) { float e = b->hi[a] - b->lo[a]; c[a] = e > 0 ? (0.5f * (l[a] + h[a]) - b->lo[a]) / e : 0.f; }
        keys[i] = morton63(c[0], c[1], c[2]);
    }
    std::iota(order.begin(), order.end(), 0u);
    std::stable_sort(order.begin(), order.end(), [&](uint32_t x, uint32_t y) { return keys[x] < keys[y]; });
    std::vector<uint64_t> skeys(n);
    b->tris_sorted.resize(n);
    for (uint32_t i = 0; i < n; ++i) { skeys[i] = keys[order[i]]; b->tris_sorted[i] = b->tris_in[refs[order[i]].tri]; }
    /* hierarchy */
    std::vector<Bvh2Node> n2(2 * (size_t)n - 1);
    std::vector<uint32_t> parent(2 * (size_t)n - 1, 0xFFFFFFFFu), counts(n > 1 ? n - 1 : 1, 0), flags(n > 1 ? n - 1 : 1, 0);
    for (uint32_t k = 0; k < n; ++k) {
        Bvh2Node& leaf = n2[n - 1 + k];
        for (int a = 0; a < 3; ++a) { leaf.lo[a] = refs[order[k]].lo[a] - b->pad; leaf.hi[a] = refs[order[k]].hi[a] + b->pad; }
        leaf.left = leaf.right = 0;
    }
    if (ploc_radius == 0 || n < 2) {
        for (int i = 0; i + 1 < (int)n; ++i) lbvh_build_node(skeys.data(), (int)n, i, n2.data(), parent.data());
    } else {
        /* PLOC rounds, as the device runs them: nearest neighbours, roles, ids by the prefix sum of the merges (handed out downwards
         * so that the last merge, the root, is node 0), compaction in order */
        std::vector<Bvh2Node> cl(n), nx; std::vector<uint32_t> nn(n);
        for (uint32_t k = 0; k < n; ++k) { cl[k] = n2[n - 1 + k]; cl[k].left = LEAF_FLAG | k; }
        uint32_t m = n, next_id = n - 1;
        while (m > 1) {
            for (uint32_t i = 0; i < m; ++i) nn[i] = ploc_nearest(cl.data(), m, i, ploc_radius);
            nx.clear(); uint32_t rank = 0;
            for (uint32_t i = 0; i < m; ++i) {
                const int role = ploc_role(nn.data(), i);
                if (role == 0) nx.push_back(cl[i]);
                else if (role == 1) { const uint32_t id = next_id - 1 - rank++; nx.push_back(ploc_merge(cl[i], cl[nn[i]], id, n2[id])); }
            }
            next_id -= rank; cl.swap(nx); m = (uint32_t)cl.size(); b->ploc_rounds++;
        }
        for (uint32_t i = 0; i + 1 < n; ++i) { parent[mem_slot(n, n2[i].left)] = i; parent[mem_slot(n, n2[i].right)] = i; }
        parent[0] = 0xFFFFFFFFu;
    }
    /* bottom-up passes: the same walk the kernel does (the second arrival at a node runs refit_visit).  Greedy build: one pass,
     * bounds and counts only.  Default build: REFIT_PASSES passes with tree rotations, the collapse programme in the last one. */
    struct HostMem {
        Bvh2Node* n2; uint32_t* counts; DpNode* dp; uint32_t n;
        uint32_t slot(uint32_t ref) const { return (ref & LEAF_FLAG) ? n - 1 + (ref & ~LEAF_FLAG) : ref; }
        Bvh2Node node(uint32_t i) const { return n2[i]; }
        void store_node(uint32_t i, const Bvh2Node& m) { n2[i] = m; }
        uint32_t count(uint32_t ref) const { return (ref & LEAF_FLAG) ? 1u : counts[ref]; }
        void store_count(uint32_t i, uint32_t c) { counts[i] = c; }
        void dp_costs(uint32_t ref, float out[7]) const { for (int i = 0; i < 7; ++i) out[i] = dp[ref].cost[i]; }
        void store_dp(uint32_t i, const DpNode& d) { dp[i] = d; }
    };
    std::vector<DpNode> dp(use_dp && n > 1 ? n - 1 : 0);
    HostMem mem{n2.data(), counts.data(), dp.data(), n};
    const int passes = treelet_passes > 0 ? treelet_passes : (use_dp ? RT3_REFIT_PASSES : 1);
    for (int pass = 0; pass < passes; ++pass) {
        std::fill(flags.begin(), flags.end(), 0u);
        if (pass) for (uint32_t i = 0; i + 1 < n; ++i) { parent[mem.slot(n2[i].left)] = i; parent[mem.slot(n2[i].right)] = i; }   /* rotations moved subtrees */
        for (uint32_t k = 0; k < n; ++k) {
            uint32_t cur = parent[n - 1 + k];
            while (cur != 0xFFFFFFFFu) {
                if (flags[cur]++ == 0) break;
                if (treelet_passes > 0) treelet_visit(mem, n, cur, use_dp && pass == passes - 1);
                else refit_visit(mem, n, cur, use_dp, use_dp && pass == passes - 1);
                cur = parent[cur];
            }
        }
    }
    /* collapse, level by level */
    b->nodes.resize(n);  /* upper bound: every wide node has >= 2 children unless n == 1 */
    b->tris_out.resize(n);
    std::vector<uint32_t> q(2 * (size_t)n + 2), qn(2 * (size_t)n + 2);
    uint32_t node_counter = 1, tri_counter = 0, q_count = 1, next_count = 0;
    q[0] = (n == 1) ? (LEAF_FLAG | 0u) : 0u; q[1] = 0;
    CollapseCtx c;
    c.nodes = n2.data(); c.counts = counts.data(); c.tris_sorted = b->tris_sorted.data(); c.n_leaves = n;
    c.out_nodes = b->nodes.data(); c.out_tris = b->tris_out.data(); c.node_counter = &node_counter; c.tri_counter = &tri_counter;
    c.next_count = &next_count; c.sah_accum = b->sah; c.dp = dp.empty() ? nullptr : dp.data();
    while (q_count) {
        next_count = 0; c.next_items = qn.data();
        for (uint32_t i = 0; i < q_count; ++i) collapse_one(c, q[2 * i], q[2 * i + 1]);
        q.swap(qn); q_count = next_count; b->levels++;
    }
    b->n_nodes = node_counter; b->n_tris = tri_counter;
    return b;
}
EmuBvh* emu_build(const float* tris9, uint32_t n) { return build_lbvh(tris9, n, false); }
EmuBvh* emu_build_dp(const float* tris9, uint32_t n) { return build_lbvh(tris9, n, true); }
EmuBvh* emu_build_ploc(const float* tris9, uint32_t n, int use_dp, uint32_t radius) { return build_lbvh(tris9, n, use_dp != 0, 0.0f, radius); }
uint32_t emu_ploc_rounds(const EmuBvh* b) { return b->ploc_rounds; }
void emu_set_input_edges(int on) { g_input_edges = on; }
EmuBvh* emu_build_ex(const float* tris9, uint32_t n, int use_dp, float split_beta, uint32_t ploc_radius, int treelet_passes) {
    return build_lbvh(tris9, n, use_dp != 0, split_beta, ploc_radius, treelet_passes);
}
EmuBvh* emu_build_treelet(const float* tris9, uint32_t n, int passes) { return build_lbvh(tris9, n, true, 0.0f, 0, passes); }
EmuBvh* emu_build_split(const float* tris9, uint32_t n, int use_dp, float beta) { return build_lbvh(tris9, n, use_dp != 0, beta); }
/* wrap an acceleration structure built elsewhere (downloaded from the GPU) so that it can be traversed here */
EmuBvh* emu_wrap(const uint32_t* nodes20, uint32_t n_nodes, const uint32_t* tris12, uint32_t n_tris) {
    EmuBvh* b = new EmuBvh();
    b->nodes.resize(n_nodes ? n_nodes : 1);
    memcpy(b->nodes.data(), nodes20, (size_t)n_nodes * sizeof(WideNode));
    b->tris_out.resize(n_tris);
    memcpy(b->tris_out.data(), tris12, (size_t)n_tris * sizeof(Tri48));
    b->tris_in.resize(n_tris);
    for (uint32_t i = 0; i < n_tris; ++i) if (b->tris_out[i].prim < n_tris) b->tris_in[b->tris_out[i].prim] = b->tris_out[i];
    b->n_nodes = n_nodes; b->n_tris = n_tris;
    return b;
}
/* the wide nodes and leaf records as the device holds them (for tests/emu/warp_emu.cpp) */
void emu_export(const EmuBvh* b, uint32_t* nodes20, uint32_t* tris12) {
    memcpy(nodes20, b->nodes.data(), (size_t)b->n_nodes * sizeof(WideNode));
    memcpy(tris12, b->tris_out.data(), (size_t)b->n_tris * sizeof(Tri48));
}
void emu_destroy(EmuBvh* b) { delete b; }
void emu_info(const EmuBvh* b, uint32_t* out4, float* sah2) {
    out4[0] = b->n_nodes; out4[1] = b->n_tris; out4[2] = b->levels; out4[3] = (uint32_t)b->tris_in.size();
    sah2[0] = b->sah[0]; sah2[1] = b->sah[1];
}
/* order-independent check that the collapse kept every triangle exactly once: returns number of mismatches */
/* Self-check of the hit-byte-first node format (bvh_logic.h, RT3_NODE_V2): the octant permutation table, record_index / dense_records on
 * random masks, and every node of a built tree: valid24 and the inner mask own disjoint slots, the record counts add up.  Returns the
 * number of violations. */
uint32_t emu_check_node_v2(const EmuBvh* b) {
    uint32_t bad = 0;
#if RT3_NODE_V2
    const PermLut lut = make_perm_lut();
    for (uint32_t o = 0; o < 8; ++o)
        for (uint32_t x = 0; x < 256; ++x) {
            const uint32_t y = lut.v[o * 256 + x];
            if (__builtin_popcount(y) != __builtin_popcount(x)) bad++;
            if (lut.v[o * 256 + y] != x) bad++;                                   /* XOR permutations are involutions */
            for (uint32_t s = 0; s < 8; ++s) if (((x >> s) & 1u) != ((y >> (s ^ o)) & 1u)) bad++;
            if (perm_lookup((o << 8) | x) != y) bad++;
        }
    uint32_t rng = 12345u;
    for (int it = 0; it < 20000; ++it) {
        uint32_t valid = 0;
        for (int s = 0; s < 8; ++s) { rng = rng * 1664525u + 1013904223u; const uint32_t c = (rng >> 24) & 3u; valid |= ((1u << c) - 1u) << (3 * s); }
        rng = rng * 1664525u + 1013904223u;
        const uint32_t hits = valid & (rng >> 8);
        const uint32_t dense = dense_records(hits, valid);
        if (__builtin_popcount(dense) != __builtin_popcount(hits)) bad++;
        uint32_t k = 0;
        for (int bit = 0; bit < 24; ++bit)
            if ((valid >> bit) & 1u) {
                if (record_index(valid, bit) != k) bad++;
                if ((((hits >> bit) & 1u) != ((dense >> k) & 1u))) bad++;
                if (((dense >> k) & 1u) && record_index(0x00FFFFFFu, (int)k) != k) bad++;   /* a dense group addresses itself */
                k++;
            }
    }
    if (b) {
        uint64_t records = 0;
        for (uint32_t ni = 0; ni < b->n_nodes; ++ni) {
            const WideNode& n = b->nodes[ni];
            const uint32_t imask = n.w[3] >> 24, valid = n.w[6];
            if (valid >> 24) bad++;
            if (n.w[7] != 0u) bad++;
            for (int s = 0; s < 8; ++s) {
                const uint32_t f = (valid >> (3 * s)) & 7u;
                if (f != 0u && f != 1u && f != 3u && f != 7u) bad++;             /* unary count */
                if (f != 0u && ((imask >> s) & 1u)) bad++;                       /* a slot is a leaf or an inner child, never both */
            }
            records += (uint64_t)__builtin_popcount(valid);
        }
        if (records != b->n_tris) bad++;
    }
#else
    (void)b;
#endif
    return bad;
}
uint32_t emu_check_triangles(const EmuBvh* b) {
    std::vector<uint32_t> seen(b->tris_in.size(), 0);
    uint32_t bad = 0;
    for (uint32_t i = 0; i < b->n_tris; ++i) {
        const Tri48& t = b->tris_out[i];
        if (t.prim >= seen.size()) { bad++; continue; }
        seen[t.prim]++;
        if (memcmp(&t, &b->tris_in[t.prim], sizeof(Tri48)) != 0) bad++;
    }
    const bool split = b->n_refs > b->tris_in.size();      /* with triangle splitting a triangle is kept once per reference */
    for (uint32_t s : seen) if (split ? s < 1 : s != 1) bad++;
    return bad;
}
void emu_trace_closest(const EmuBvh* b, const rt3_ray* rays, uint64_t n, rt3_hit* out, int brute, double* stats2) {
    uint64_t nodes = 0, tris = 0;
    for (uint64_t i = 0; i < n; ++i) {
        rt3_hit h; h.t = rays[i].tmax; h.u = h.v = 0.f; h.prim = RT3_MISS;
        if (brute) {
            rt3m::vec3 o = rt3m::v3(rays[i].o), d = rt3m::v3(rays[i].d);
            for (const Tri48& t : b->tris_in) {
                float tt, u, v;
                if (rt3m::ray_triangle(o, d, rt3m::v3(t.v0), rt3m::v3(t.e1), rt3m::v3(t.e2), tt, u, v) && tt > rays[i].tmin && tt < rays[i].tmax &&
                    rt3m::closer_hit(tt, t.prim, h.t, h.prim)) { h.t = tt; h.u = u; h.v = v; h.prim = t.prim; }
            }
        } else {
            TraverseStats st = {0, 0};
            traverse_wide<false>(b->nodes.data(), b->tris_out.data(), rays[i], h, &st);
            nodes += st.nodes; tris += st.tris;
        }
        out[i] = h;
    }
    if (stats2 && n) { stats2[0] = (double)nodes / n; stats2[1] = (double)tris / n; }
}
void emu_trace_any(const EmuBvh* b, const rt3_ray* rays, uint64_t n, uint32_t* bits, int brute) {
    for (uint64_t w = 0; w < (n + 31) / 32; ++w) bits[w] = 0;
    for (uint64_t i = 0; i < n; ++i) {
        bool hit = false;
        if (brute) {
            rt3m::vec3 o = rt3m::v3(rays[i].o), d = rt3m::v3(rays[i].d);
            for (const Tri48& t : b->tris_in) {
                float tt, u, v;
                if (rt3m::ray_triangle(o, d, rt3m::v3(t.v0), rt3m::v3(t.e1), rt3m::v3(t.e2), tt, u, v) && tt > rays[i].tmin && tt < rays[i].tmax) { hit = true; break; }
            }
        } else {
            rt3_hit h; h.t = rays[i].tmax; h.u = h.v = 0.f; h.prim = RT3_MISS;
            hit = traverse_wide<true>(b->nodes.data(), b->tris_out.data(), rays[i], h, nullptr);
        }
        if (hit) bits[i >> 5] |= 1u << (i & 31);
    }
}
}

/* EXPERIMENT: top-down binned-SAH BVH2 (CPU) fed to the same collapse, to measure how much traversal work the LBVH topology costs. */
namespace {
struct SahBuilder {
    std::vector<Bvh2Node>& n2; std::vector<uint32_t>& counts; const std::vector<Tri48>& tris; float pad; uint32_t n;
    std::vector<uint32_t> order; std::vector<float> lo, hi, cen;
    uint32_t next_internal = 0;
    std::vector<uint32_t> leaf_of;   /* position in order -> leaf index (sorted position) */
    SahBuilder(std::vector<Bvh2Node>& a, std::vector<uint32_t>& c, const std::vector<Tri48>& t, float p) : n2(a), counts(c), tris(t), pad(p), n((uint32_t)t.size()) {
        order.resize(n); lo.resize(3 * n); hi.resize(3 * n); cen.resize(3 * n);
        for (uint32_t i = 0; i < n; ++i) { order[i] = i; tri_bounds(tris[i], pad, &lo[3 * i], &hi[3 * i]); for (int a2 = 0; a2 < 3; ++a2) cen[3 * i + a2] = 0.5f * (lo[3 * i + a2] + hi[3 * i + a2]); }
    }
    /* returns ref; leaves reference position in `order` (the caller reorders triangles accordingly) */
    uint32_t build(uint32_t first, uint32_t count) {
        if (count == 1) {
            Bvh2Node& leaf = n2[n - 1 + first];
            for (int a = 0; a < 3; ++a) { leaf.lo[a] = lo[3 * order[first] + a]; leaf.hi[a] = hi[3 * order[first] + a]; }
            leaf.left = leaf.right = 0;
            return LEAF_FLAG | first;
        }
        float blo[3] = {3e38f, 3e38f, 3e38f}, bhi[3] = {-3e38f, -3e38f, -3e38f}, clo[3] = {3e38f, 3e38f, 3e38f}, chi[3] = {-3e38f, -3e38f, -3e38f};
        for (uint32_t i = first; i < first + count; ++i) for (int a = 0; a < 3; ++a) {
            uint32_t id = order[i];
            blo[a] = std::min(blo[a], lo[3 * id + a]); bhi[a] = std::max(bhi[a], hi[3 * id + a]);
            clo[a] = std::min(clo[a], cen[3 * id + a]); chi[a] = std::max(chi[a], cen[3 * id + a]);
        }
        const int NB = 32;
        float best = 3e38f; int baxis = -1, bsplit = -1;
        for (int a = 0; a < 3; ++a) {
            if (!(chi[a] > clo[a])) continue;
            float l[NB][3], h[NB][3]; uint32_t c[NB];
            for (int k = 0; k < NB; ++k) { c[k] = 0; for (int x = 0; x < 3; ++x) { l[k][x] = 3e38f; h[k][x] = -3e38f; } }
            float scale = NB / (chi[a] - clo[a]);
            for (uint32_t i = first; i < first + count; ++i) {
                uint32_t id = order[i]; int k = std::min(NB - 1, (int)((cen[3 * id + a] - clo[a]) * scale));
                c[k]++; for (int x = 0; x < 3; ++x) { l[k][x] = std::min(l[k][x], lo[3 * id + x]); h[k][x] = std::max(h[k][x], hi[3 * id + x]); }
            }
            float la[NB], ra[NB]; uint32_t lc[NB], rc[NB];
            float al[3] = {3e38f, 3e38f, 3e38f}, ah[3] = {-3e38f, -3e38f, -3e38f}; uint32_t cc = 0;
            for (int k = 0; k < NB - 1; ++k) { for (int x = 0; x < 3; ++x) { al[x] = std::min(al[x], l[k][x]); ah[x] = std::max(ah[x], h[k][x]); } cc += c[k]; lc[k] = cc; la[k] = cc ? box_area(al, ah) : 0; }
            for (int x = 0; x < 3; ++x) { al[x] = 3e38f; ah[x] = -3e38f; } cc = 0;
            for (int k = NB - 1; k > 0; --k) { for (int x = 0; x < 3; ++x) { al[x] = std::min(al[x], l[k][x]); ah[x] = std::max(ah[x], h[k][x]); } cc += c[k]; rc[k - 1] = cc; ra[k - 1] = cc ? box_area(al, ah) : 0; }
            for (int k = 0; k < NB - 1; ++k) if (lc[k] && rc[k]) { float cost = la[k] * lc[k] + ra[k] * rc[k]; if (cost < best) { best = cost; baxis = a; bsplit = k; } }
        }
        uint32_t mid;
        if (baxis >= 0) {
            float scale = NB / (chi[baxis] - clo[baxis]);
            auto it = std::partition(order.begin() + first, order.begin() + first + count, [&](uint32_t id) { return std::min(NB - 1, (int)((cen[3 * id + baxis] - clo[baxis]) * scale)) <= bsplit; });
            mid = (uint32_t)(it - order.begin());
        } else mid = first + count / 2;
        if (mid == first || mid == first + count) mid = first + count / 2;
        uint32_t me = next_internal++;
        uint32_t l = build(first, mid - first), r = build(mid, first + count - mid);
        Bvh2Node& nd = n2[me];
        nd.left = l; nd.right = r;
        for (int a = 0; a < 3; ++a) { nd.lo[a] = blo[a]; nd.hi[a] = bhi[a]; }
        counts[me] = count;
        return me;
    }
};
}
extern "C" EmuBvh* emu_build_sah(const float* tris9, uint32_t n) {
    EmuBvh* b = new EmuBvh();
    b->tris_in.resize(n);
    for (int a = 0; a < 3; ++a) { b->lo[a] = 3.4e38f; b->hi[a] = -3.4e38f; }
    for (uint32_t i = 0; i < n; ++i) {
        const float* p = tris9 + 9 * (size_t)i; Tri48& t = b->tris_in[i];
        for (int a = 0; a < 3; ++a) { t.v0[a] = p[a]; t.e1[a] = p[3 + a] - p[a]; t.e2[a] = p[6 + a] - p[a];
            for (int k = 0; k < 3; ++k) { b->lo[a] = std::min(b->lo[a], p[3 * k + a]); b->hi[a] = std::max(b->hi[a], p[3 * k + a]); } }
        t.prim = i; t.pad0 = t.pad1 = 0;
    }
    float ext = std::max(b->hi[0] - b->lo[0], std::max(b->hi[1] - b->lo[1], b->hi[2] - b->lo[2]));
    b->pad = ext * RT3_PAD_FRACTION;
    std::vector<Bvh2Node> n2(2 * (size_t)n - 1); std::vector<uint32_t> counts(n, 0);
    SahBuilder sb(n2, counts, b->tris_in, b->pad);
    uint32_t root = sb.build(0, n);
    b->tris_sorted.resize(n);
    for (uint32_t i = 0; i < n; ++i) b->tris_sorted[i] = b->tris_in[sb.order[i]];
    b->nodes.resize(n); b->tris_out.resize(n);
    std::vector<uint32_t> q(2 * (size_t)n + 2), qn(2 * (size_t)n + 2);
    uint32_t node_counter = 1, tri_counter = 0, q_count = 1, next_count = 0;
    q[0] = root; q[1] = 0;
    CollapseCtx c; c.nodes = n2.data(); c.counts = counts.data(); c.tris_sorted = b->tris_sorted.data(); c.n_leaves = n;
    c.out_nodes = b->nodes.data(); c.out_tris = b->tris_out.data(); c.node_counter = &node_counter; c.tri_counter = &tri_counter; c.next_count = &next_count; c.sah_accum = b->sah; c.dp = nullptr;
    while (q_count) { next_count = 0; c.next_items = qn.data(); for (uint32_t i = 0; i < q_count; ++i) collapse_one(c, q[2 * i], q[2 * i + 1]); q.swap(qn); q_count = next_count; b->levels++; }
    b->n_nodes = node_counter; b->n_tris = tri_counter;
    return b;
}
