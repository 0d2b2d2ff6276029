/*
 * bvh_experiments.h — TEST INFRASTRUCTURE: builder experiments that were measured in the CPU emulator and NOT taken into the device
 * build: a PLOC hierarchy in place of the Karras hierarchy, and triangle splitting.  They are written against
 * the product's per-thread logic (raytracer3_b200/csrc/bvh_logic.h) so that the rest of the pipeline — rotations / treelets, collapse
 * programme, wide collapse, traversal — is the product's own; tests/test_emu_bvh.py checks that neither changes a hit, and
 * tools/emu_topology.py prints the traversal work they lead to.
 */
#ifndef RT3_BVH_EXPERIMENTS_H
#define RT3_BVH_EXPERIMENTS_H

#include "../../raytracer3_b200/csrc/bvh_logic.h"

namespace rt3b {

/* ---------------------------------------------------------------- PLOC hierarchy --------------- */
/* Parallel locally-ordered clustering (Meister & Bittner 2018) as an alternative to the Karras hierarchy: the clusters, kept in
 * Morton order, each look `radius` places to both sides for the neighbour whose union box has the least surface area; mutual
 * nearest neighbours merge, the array is compacted in order, and the round repeats until one cluster is left.  Distances are
 * symmetric bit for bit (min / max commute) and ties go to the lower index, so every round has at least one mutual pair and the
 * tree is a function of the input alone.  A cluster is a Bvh2Node whose `left` holds its BVH2 reference. */
RT3_HD uint32_t ploc_nearest(const Bvh2Node* cl, uint32_t m, uint32_t i, uint32_t radius) {
    const uint32_t j0 = i > radius ? i - radius : 0u, j1 = (i + radius < m - 1u) ? i + radius : m - 1u;
    const Bvh2Node me = cl[i];
    float best = 3.4e38f; uint32_t bj = i;
    for (uint32_t j = j0; j <= j1; ++j) {
        if (j == i) continue;
        const float a = union_area(me, cl[j]);
        if (a < best) { best = a; bj = j; }
    }
    return bj;
}
/* what cluster i does in the merge step: 0 = stays, 1 = merges with nn[i] and keeps this place, 2 = absorbed by nn[i] */
RT3_HD int ploc_role(const uint32_t* nn, uint32_t i) {
    const uint32_t j = nn[i];
    if (j == i || nn[j] != i) return 0;
    return i < j ? 1 : 2;
}
RT3_HD Bvh2Node ploc_merge(const Bvh2Node& a, const Bvh2Node& b, uint32_t id, Bvh2Node& node_out) {
    Bvh2Node m;
    box_union(a, b, m.lo, m.hi);
    node_out = m; node_out.left = a.left; node_out.right = b.left;
    m.left = id; m.right = 0;
    return m;
}

/* ---------------------------------------------------------------- triangle splitting ----------- */
/* Early split clipping in the form of Karras & Aila 2013 (section 4): before the Morton sort a triangle whose box is much larger
 * than the triangle itself is replaced by several REFERENCES to it, each with the tight box of the part of the triangle inside one
 * cell of the Morton grid.  The hierarchy is then built over references; a leaf record is a copy of the triangle, so a ray may test
 * one triangle through two references — the same (t, u, v, id) both times, and closer_hit() is idempotent — while long thin
 * triangles no longer inflate every node above them.  Hits do not change (the union of a triangle's reference boxes covers the
 * triangle; each is padded like a whole-triangle box), only the work to find them.
 *
 *   priority(t) = cbrt(2^-level * (A_box - A_ideal)),  level = level of the most important Morton-grid plane the box straddles
 *                 (1 = the scene's median plane), A = half surface area, A_ideal = |e1 x e2|_1 = the least a box of t can have
 *   splits(t)   = floor(priority(t) * budget / sum of priorities)                      (budget = beta * triangles)
 *   recursion   : split the box at the most important plane it straddles, clip the triangle for the tight boxes of both sides,
 *                 share the remaining splits by the sides' longest extents.
 * Everything is a function of the triangle and the scene box alone, so the reference list is deterministic. */
struct SplitGrid { float lo[3], ext[3]; };                    /* scene box (unpadded) */
struct RefBox { float lo[3]; uint32_t tri; float hi[3]; uint32_t pad; };   /* 32 B */
#define RT3_SPLIT_MAX_PER_TRI 4096u
#define RT3_SPLIT_STACK 14

RT3_HD void tri_vertices(const Tri48& t, int kind, float p[3][3]) {
    for (int a = 0; a < 3; ++a) {
        p[0][a] = t.v0[a];
        p[1][a] = kind == REC_TRI_VERTS ? t.e1[a] : t.v0[a] + t.e1[a];
        p[2][a] = kind == REC_TRI_VERTS ? t.e2[a] : t.v0[a] + t.e2[a];
    }
}
/* most important grid plane strictly inside (lo, hi): axis, position; returns its level (1 = median of the scene box), 0 = none */
RT3_HD int split_plane(const SplitGrid& g, const float lo[3], const float hi[3], int& axis, float& pos) {
    int best = 0; axis = -1; pos = 0.0f;
    float best_ext = -1.0f;
    for (int a = 0; a < 3; ++a) {
        if (!(g.ext[a] > 0.0f) || !(hi[a] > lo[a])) continue;
        const float S = 2097152.0f;
        float fl = rt3m::min_(rt3m::max_((lo[a] - g.lo[a]) / g.ext[a] * S, 0.0f), S - 1.0f);
        float fh = rt3m::min_(rt3m::max_((hi[a] - g.lo[a]) / g.ext[a] * S, 0.0f), S - 1.0f);
        uint32_t xl = (uint32_t)fl, xh = (uint32_t)fh;
        if (xl == xh) continue;
        int b = bfind32(xl ^ xh);                             /* the plane is the lowest multiple of 2^b above xl */
        uint32_t plane = (xh >> b) << b;
        float pp = rt3m::fma_((float)plane * (1.0f / S), g.ext[a], g.lo[a]);
        if (!(pp > lo[a] && pp < hi[a])) continue;
        int level = 21 - b;
        float e = hi[a] - lo[a];
        if (best == 0 || level < best || (level == best && e > best_ext)) { best = level; axis = a; pos = pp; best_ext = e; }
    }
    return best;
}
RT3_HD float half_area(const float lo[3], const float hi[3]) {
    float ex = hi[0] - lo[0], ey = hi[1] - lo[1], ez = hi[2] - lo[2];
    return ex * ey + ey * ez + ez * ex;
}
/* cube root of w > 0 by Newton from an exponent guess: +, *, / only, so host and device agree bit for bit */
RT3_HD float cbrt_newton(float w) {
    float x = rt3m::as_float(rt3m::as_uint(w) / 3u + 0x2A514067u);
    for (int i = 0; i < 4; ++i) x = x - (x * x * x - w) / (3.0f * x * x);
    return x;
}
RT3_HD float split_priority(const Tri48& t, int kind, const SplitGrid& g) {
    float p[3][3], lo[3], hi[3];
    tri_vertices(t, kind, p);
    for (int a = 0; a < 3; ++a) { lo[a] = rt3m::min_(p[0][a], rt3m::min_(p[1][a], p[2][a])); hi[a] = rt3m::max_(p[0][a], rt3m::max_(p[1][a], p[2][a])); }
    int axis; float pos;
    int level = split_plane(g, lo, hi, axis, pos);
    if (level == 0) return 0.0f;
    float d1[3], d2[3];
    for (int a = 0; a < 3; ++a) { d1[a] = p[1][a] - p[0][a]; d2[a] = p[2][a] - p[0][a]; }
    float cx = d1[1] * d2[2] - d1[2] * d2[1], cy = d1[2] * d2[0] - d1[0] * d2[2], cz = d1[0] * d2[1] - d1[1] * d2[0];
    float ideal = 0.5f * (rt3m::abs_(cx) + rt3m::abs_(cy) + rt3m::abs_(cz));
    float w = (half_area(lo, hi) - ideal) * rt3m::exp2i(-level);
    if (!(w > 0.0f)) return 0.0f;
    return cbrt_newton(w);
}
/* The priorities are summed in fixed point (integer atomics on the device: the sum must not depend on the order of the threads).
 * inv_norm = 1 / cbrt(extent^2) brings them to [0, ~1.2]. */
RT3_HD float split_inv_norm(const SplitGrid& g) {
    const float e = rt3m::max_(g.ext[0], rt3m::max_(g.ext[1], g.ext[2]));
    return e > 0.0f ? 1.0f / cbrt_newton(e * e) : 0.0f;
}
RT3_HD uint64_t split_priority_fixed(float prio, float inv_norm) { return (uint64_t)(prio * inv_norm * 16777216.0f); }
RT3_HD float split_scale(uint64_t total_fixed, uint32_t n, float beta, float inv_norm) {
    return total_fixed ? (beta * (float)n) * (16777216.0f / (float)total_fixed) * inv_norm : 0.0f;
}
RT3_HD uint32_t split_count(float prio, float scale) {
    const float c = prio * scale;
    return c >= (float)RT3_SPLIT_MAX_PER_TRI ? RT3_SPLIT_MAX_PER_TRI : (uint32_t)c;
}
/* boxes of the triangle's parts on both sides of the plane x[axis] = pos, each clamped to the parent box (lo, hi) */
RT3_HD void split_clip(const float p[3][3], int axis, float pos, const float lo[3], const float hi[3], float llo[3], float lhi[3], float rlo[3], float rhi[3]) {
    for (int a = 0; a < 3; ++a) { llo[a] = rlo[a] = 3.4e38f; lhi[a] = rhi[a] = -3.4e38f; }
    for (int i = 0; i < 3; ++i) {
        const float* v0 = p[i]; const float* v1 = p[(i + 1) % 3];
        const float a0 = v0[axis], a1 = v1[axis];
        if (a0 <= pos) for (int a = 0; a < 3; ++a) { llo[a] = rt3m::min_(llo[a], v0[a]); lhi[a] = rt3m::max_(lhi[a], v0[a]); }
        if (a0 >= pos) for (int a = 0; a < 3; ++a) { rlo[a] = rt3m::min_(rlo[a], v0[a]); rhi[a] = rt3m::max_(rhi[a], v0[a]); }
        if ((a0 < pos && a1 > pos) || (a0 > pos && a1 < pos)) {
            const float f = rt3m::min_(rt3m::max_((pos - a0) / (a1 - a0), 0.0f), 1.0f);
            for (int a = 0; a < 3; ++a) {
                const float x = a == axis ? pos : rt3m::fma_(f, v1[a] - v0[a], v0[a]);
                llo[a] = rt3m::min_(llo[a], x); lhi[a] = rt3m::max_(lhi[a], x);
                rlo[a] = rt3m::min_(rlo[a], x); rhi[a] = rt3m::max_(rhi[a], x);
            }
        }
    }
    lhi[axis] = rt3m::min_(lhi[axis], pos); rlo[axis] = rt3m::max_(rlo[axis], pos);
    for (int a = 0; a < 3; ++a) {
        llo[a] = rt3m::max_(llo[a], lo[a]); lhi[a] = rt3m::min_(lhi[a], hi[a]);
        rlo[a] = rt3m::max_(rlo[a], lo[a]); rhi[a] = rt3m::min_(rhi[a], hi[a]);
    }
}
RT3_HD bool box_valid(const float lo[3], const float hi[3]) { return lo[0] <= hi[0] && lo[1] <= hi[1] && lo[2] <= hi[2]; }
/* References of one triangle given `splits` plane cuts; emit(lo, hi) is called once per reference, the count is returned.
 * Depth first, the side with fewer cuts first and the other one parked, so the stack never holds more than log2(splits) + 1 boxes. */
template <class Emit>
RT3_HD uint32_t split_triangle(const Tri48& t, int kind, const SplitGrid& g, uint32_t splits, Emit&& emit) {
    float p[3][3];
    tri_vertices(t, kind, p);
    float st_lo[RT3_SPLIT_STACK][3], st_hi[RT3_SPLIT_STACK][3]; uint32_t st_s[RT3_SPLIT_STACK];
    int sp = 0;
    float lo[3], hi[3];
    for (int a = 0; a < 3; ++a) { lo[a] = rt3m::min_(p[0][a], rt3m::min_(p[1][a], p[2][a])); hi[a] = rt3m::max_(p[0][a], rt3m::max_(p[1][a], p[2][a])); }
    uint32_t s = splits < RT3_SPLIT_MAX_PER_TRI ? splits : RT3_SPLIT_MAX_PER_TRI, count = 0;
    for (;;) {
        int axis = -1; float pos = 0.0f;
        bool cut = s > 0u && split_plane(g, lo, hi, axis, pos) != 0;
        float llo[3], lhi[3], rlo[3], rhi[3];
        if (cut) {
            split_clip(p, axis, pos, lo, hi, llo, lhi, rlo, rhi);
            const bool lv = box_valid(llo, lhi), rv = box_valid(rlo, rhi);
            if (!lv && !rv) cut = false;
            else if (!lv || !rv) {                               /* everything on one side: tighter box, no cut used */
                for (int a = 0; a < 3; ++a) { lo[a] = lv ? llo[a] : rlo[a]; hi[a] = lv ? lhi[a] : rhi[a]; }
                continue;
            }
        }
        if (!cut) {
            emit(lo, hi); count++;
            if (sp == 0) break;
            --sp;
            for (int a = 0; a < 3; ++a) { lo[a] = st_lo[sp][a]; hi[a] = st_hi[sp][a]; }
            s = st_s[sp];
            continue;
        }
        const float wl = rt3m::max_(lhi[0] - llo[0], rt3m::max_(lhi[1] - llo[1], lhi[2] - llo[2]));
        const float wr = rt3m::max_(rhi[0] - rlo[0], rt3m::max_(rhi[1] - rlo[1], rhi[2] - rlo[2]));
        const uint32_t rest = s - 1u;
        uint32_t sl = (wl + wr) > 0.0f ? (uint32_t)((float)rest * (wl / (wl + wr)) + 0.5f) : rest / 2u;
        if (sl > rest) sl = rest;
        const uint32_t sr = rest - sl;
        const bool left_first = sl <= sr;
        if (sp < RT3_SPLIT_STACK) {
            for (int a = 0; a < 3; ++a) { st_lo[sp][a] = left_first ? rlo[a] : llo[a]; st_hi[sp][a] = left_first ? rhi[a] : lhi[a]; }
            st_s[sp] = left_first ? sr : sl; sp++;
            for (int a = 0; a < 3; ++a) { lo[a] = left_first ? llo[a] : rlo[a]; hi[a] = left_first ? lhi[a] : rhi[a]; }
            s = left_first ? sl : sr;
        } else {                                                  /* cannot park: keep the box whole (never reached with the caps above) */
            s = 0u;
        }
    }
    return count;
}

}  /* namespace rt3b */
#endif
