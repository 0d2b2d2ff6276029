/*
 * bvh_logic.h — per-thread logic of the GPU BVH builder and of the wide-BVH traversal.
 *
 * Replaces what the reference delegates to the Vulkan driver / RT cores:
 *   vkCmdBuildAccelerationStructuresKHR   /root/reference/src/renderer/vulkan/raytracing.rs:106-141
 *   TraceRay inside vkCmdTraceRaysKHR     /root/reference/src/renderer/pipeline_cache/mod.rs:64-73
 *
 * Everything here is `RT3_HD` (host + device): the CUDA kernels in build.cu / trace.cu call these
 * functions one thread at a time, and tests/emu compiles the very same functions with g++ to check the
 * algorithmic logic (Karras ranges, wide collapse, node encoding, traversal order) on a machine
 * without a GPU.  The emulation is test infrastructure; librt3.so contains device code only.
 *
 * Design (B200 has no RT cores, so traversal is SM software):
 *   LBVH   : 63-bit Morton codes -> radix sort -> Karras 2012 hierarchy -> atomic bottom-up refit
 *   wide   : BVH2 collapsed top-down into an 8-wide node with child boxes quantised to 8 bits on a
 *            per-node power-of-two grid (80 bytes per node = five 16-byte loads), children placed in
 *            octant slots so that near-to-far order is a bit trick on the ray's sign octant
 *            (after Ylitie, Karras, Laine 2017), triangles stored next to the leaves as 48-byte
 *            (v0, e1, e2, id) records = three 16-byte loads.
 */
#ifndef RT3_BVH_LOGIC_H
#define RT3_BVH_LOGIC_H

#include "rt3_shared_math.h"

namespace rt3b {

using rt3m::vec3;

/* ------------------------------------------------------------------------------------------------ */
struct Bvh2Node {            /* 32 B */
    float lo[3]; uint32_t left;
    float hi[3]; uint32_t right;
};
struct Tri48 {               /* 48 B: three float4 */
    float v0[3]; uint32_t prim;
    float e1[3]; uint32_t pad0;
    float e2[3]; uint32_t pad1;
};
struct WideNode { uint32_t w[20]; };   /* 80 B: see encode_wide_node() */

static const uint32_t LEAF_FLAG = 0x80000000u;
/* Traversal stack.  A node step pushes at most ONE entry (the rest of the node group it leaves behind) and then descends one level, so
 * the stack never holds more than one entry per level of the path from the root: sp <= levels of the wide tree.  Two levels: entering an
 * instance pushes up to three entries (pending node group, pending record group, sentinel) on top of the top-level path, then the
 * bottom-level path: sp <= levels_top + 3 + levels_bottom.  rt3_build_accel refuses trees for which these bounds reach RT3_MAX_STACK
 * (build.cu: "deeper than the traversal stack"), so the `sp < RT3_MAX_STACK` tests at the push sites can never fail on a structure the
 * builder handed out; they only keep a corrupted structure from writing out of bounds. */
#define RT3_MAX_STACK 64
/* Padding of every triangle box, as a fraction of the scene extent E.  It has to cover (a) the acceptance slop of the ray/triangle test,
 * (b) the rounding of the quantised plane distances: b = fma(-32768, s, (p - o) * idir) carries up to 2^-24 * (|p - o| + 257 * node extent)
 * in units of distance, i.e. 2^-24 * (|p - o| + 257 E) at the root.  With E / 32768 = 512 * 2^-24 E the culling stays conservative for ray
 * origins within ~250 E of the scene (round 1 used E / 65536, which left no headroom at the root).  Rays that start farther
 * away than that should be clipped to the scene box by the caller. */
#ifndef RT3_PAD_FRACTION
#define RT3_PAD_FRACTION (1.0f / 32768.0f)
#endif

/* RT3_NODE_V2 (default): the node test forms the HIT BYTE of the eight children first and derives both halves of the hit mask from it —
 * one predicated add of an immediate per child (7 << 3 slot | 1 << 24 + slot), the record bits by one AND with the node's valid24, the inner
 * bits in this ray's near-to-far order by one shared-memory table look-up (perm[7 - octant][byte]: bit s of the byte moves to bit
 * s ^ (7 - octant)) — 22 instructions where the per-child (count << offset) assembly of round 1's format took about 60 (SASS counts in
 * DESIGN.md section 4.3).  0 keeps round 1's node format and test. */
#ifndef RT3_NODE_V2
#define RT3_NODE_V2 1
#endif
struct u32x4 { uint32_t x, y, z, w; };

#ifdef __CUDA_ARCH__
#define RT3_ATOMIC_ADD_U32(p, v) atomicAdd((p), (v))
#define RT3_ATOMIC_ADD_F32(p, v) atomicAdd((p), (v))
RT3_HD u32x4 ld_u32x4(const void* p) { uint4 v = __ldg((const uint4*)p); u32x4 r; r.x = v.x; r.y = v.y; r.z = v.z; r.w = v.w; return r; }
RT3_HD int clz32(uint32_t x) { return __clz((int)x); }
RT3_HD int clz64(uint64_t x) { return __clzll((long long)x); }
RT3_HD int popc32(uint32_t x) { return __popc(x); }
RT3_HD int bfind32(uint32_t x) { int r; asm("bfind.u32 %0, %1;" : "=r"(r) : "r"(x)); return r; }   /* one FLO; -1 for x = 0, like 31 - clz */
/* per-byte sign fill: 0x80 -> 0xFF, else 0x00 */
/* prmt generic mode: a selector nibble with its msb set replicates the sign bit of the selected byte
 * (the __byte_perm intrinsic only honours 3 selector bits, so the PTX form is used directly). */
RT3_HD uint32_t sign_extend_s8x4(uint32_t x) {
    uint32_t r;
    asm("prmt.b32 %0, %1, %2, %3;" : "=r"(r) : "r"(x), "r"(0u), "r"(0x0000BA98u));
    return r;
}
#else
inline uint32_t host_atomic_add(uint32_t* p, uint32_t v) { uint32_t o = *p; *p += v; return o; }
inline float host_atomic_add(float* p, float v) { float o = *p; *p += v; return o; }
#define RT3_ATOMIC_ADD_U32(p, v) rt3b::host_atomic_add((p), (v))
#define RT3_ATOMIC_ADD_F32(p, v) rt3b::host_atomic_add((p), (v))
RT3_HD u32x4 ld_u32x4(const void* p) { u32x4 r; memcpy(&r, p, 16); return r; }
RT3_HD int clz32(uint32_t x) { return x ? __builtin_clz(x) : 32; }
RT3_HD int clz64(uint64_t x) { return x ? __builtin_clzll(x) : 64; }
RT3_HD int popc32(uint32_t x) { return __builtin_popcount(x); }
RT3_HD int bfind32(uint32_t x) { return 31 - __builtin_clz(x); }
RT3_HD uint32_t sign_extend_s8x4(uint32_t x) {
    uint32_t r = 0;
    for (int i = 0; i < 4; ++i) if (x & (0x80u << (8 * i))) r |= 0xFFu << (8 * i);
    return r;
}
#endif

/* ---------------------------------------------------------------- Morton codes ----------------- */
RT3_HD uint64_t expand_bits_21(uint64_t v) {
    v &= 0x1FFFFFull;
    v = (v | (v << 32)) & 0x001F00000000FFFFull;
    v = (v | (v << 16)) & 0x001F0000FF0000FFull;
    v = (v | (v << 8)) & 0x100F00F00F00F00Full;
    v = (v | (v << 4)) & 0x10C30C30C30C30C3ull;
    v = (v | (v << 2)) & 0x1249249249249249ull;
    return v;
}
/* x, y, z in [0,1] */
RT3_HD uint64_t morton63(float x, float y, float z) {
    const float S = 2097152.0f; /* 2^21 */
    float fx = rt3m::min_(rt3m::max_(x * S, 0.0f), S - 1.0f);
    float fy = rt3m::min_(rt3m::max_(y * S, 0.0f), S - 1.0f);
    float fz = rt3m::min_(rt3m::max_(z * S, 0.0f), S - 1.0f);
    return (expand_bits_21((uint64_t)fx) << 2) | (expand_bits_21((uint64_t)fy) << 1) | expand_bits_21((uint64_t)fz);
}

/* ---------------------------------------------------------------- Karras 2012 hierarchy -------- */
/* Common-prefix length of keys i and j, with the index as tie-breaker for duplicate keys. */
RT3_HD int lbvh_delta(const uint64_t* keys, int n, int i, int j) {
    if (j < 0 || j >= n) return -1;
    uint64_t a = keys[i], b = keys[j];
    if (a == b) return 64 + clz32((uint32_t)i ^ (uint32_t)j);
    return clz64(a ^ b);
}
/* Internal node i of n-1; leaves are referenced as LEAF_FLAG | sorted index.  parent[] has 2n-1
 * entries: internal nodes first, then leaves at n-1+k. */
RT3_HD void lbvh_build_node(const uint64_t* keys, int n, int i, Bvh2Node* nodes, uint32_t* parent) {
    int d = (lbvh_delta(keys, n, i, i + 1) - lbvh_delta(keys, n, i, i - 1)) >= 0 ? 1 : -1;
    int dmin = lbvh_delta(keys, n, i, i - d);
    int lmax = 2;
    while (lbvh_delta(keys, n, i, i + lmax * d) > dmin) lmax *= 2;
    int l = 0;
    for (int t = lmax / 2; t >= 1; t /= 2)
        if (lbvh_delta(keys, n, i, i + (l + t) * d) > dmin) l += t;
    int j = i + l * d;
    int dnode = lbvh_delta(keys, n, i, j);
    int s = 0;
    int t = l;
    do {
        t = (t + 1) >> 1;
        if (lbvh_delta(keys, n, i, i + (s + t) * d) > dnode) s += t;
    } while (t > 1);
    int gamma = i + s * d + (d < 0 ? -1 : 0);
    int lo = i < j ? i : j, hi = i < j ? j : i;
    uint32_t left = (lo == gamma) ? (LEAF_FLAG | (uint32_t)gamma) : (uint32_t)gamma;
    uint32_t right = (hi == gamma + 1) ? (LEAF_FLAG | (uint32_t)(gamma + 1)) : (uint32_t)(gamma + 1);
    nodes[i].left = left;
    nodes[i].right = right;
    parent[(left & LEAF_FLAG) ? (uint32_t)(n - 1) + (left & ~LEAF_FLAG) : left] = (uint32_t)i;
    parent[(right & LEAF_FLAG) ? (uint32_t)(n - 1) + (right & ~LEAF_FLAG) : right] = (uint32_t)i;
    if (i == 0) parent[0] = 0xFFFFFFFFu;
}

/* Leaf records are 48 bytes in three interpretations (the builder only needs their bounds):
 *   REC_TRI_EDGES  (v0, e1, e2, global id)            single-level BVH over world-space triangles
 *   REC_TRI_VERTS  (a, b, c, mesh-local id)           bottom level of the two-level BVH: object-space VERTICES, so that the
 *                                                     traversal can rebuild the world-space triangle with exactly the arithmetic
 *                                                     of the flattening kernel (bit-identical hits in both modes)
 *   REC_BOX        (lo, instance id, hi)              top level: one record per instance (world-space box of its BLAS root) */
enum { REC_TRI_EDGES = 0, REC_TRI_VERTS = 1, REC_BOX = 2 };
RT3_HD void tri_bounds(const Tri48& t, float pad, float lo[3], float hi[3], int kind = REC_TRI_EDGES) {
    for (int a = 0; a < 3; ++a) {
        if (kind == REC_BOX) { lo[a] = t.v0[a] - pad; hi[a] = t.e1[a] + pad; continue; }
        float p0 = t.v0[a], p1 = kind == REC_TRI_VERTS ? t.e1[a] : t.v0[a] + t.e1[a], p2 = kind == REC_TRI_VERTS ? t.e2[a] : t.v0[a] + t.e2[a];
        lo[a] = rt3m::min_(p0, rt3m::min_(p1, p2)) - pad;
        hi[a] = rt3m::max_(p0, rt3m::max_(p1, p2)) + pad;
    }
}
/* Instance table entry of the two-level BVH (InstanceInfo of world/mod.rs:32-37 + what the traversal needs). */
struct InstRec {
    float xform[12];        /* object -> world, row-major 3x4 (as pushed)           */
    float inv[12];          /* world -> object                                       */
    uint32_t root;          /* wide-node index of the instance's BLAS root           */
    uint32_t first_prim;    /* global id of the instance's first triangle            */
    uint32_t pad0, pad1;
};
RT3_HD float box_area(const float lo[3], const float hi[3]) {
    float ex = hi[0] - lo[0], ey = hi[1] - lo[1], ez = hi[2] - lo[2];
    return 2.0f * (ex * ey + ey * ez + ez * ex);
}

RT3_HD void box_union(const Bvh2Node& a, const Bvh2Node& b, float lo[3], float hi[3]) {
    for (int q = 0; q < 3; ++q) { lo[q] = rt3m::min_(a.lo[q], b.lo[q]); hi[q] = rt3m::max_(a.hi[q], b.hi[q]); }
}
RT3_HD float union_area(const Bvh2Node& a, const Bvh2Node& b) { float lo[3], hi[3]; box_union(a, b, lo, hi); return box_area(lo, hi); }

/* ---------------------------------------------------------------- BVH2 -> 8-wide collapse ------ */
/* BVH2 storage for the collapse: nodes[0 .. n-2] internal, nodes[n-1 .. 2n-2] leaves (box only);
 * counts[i] = triangles below internal node i. */
struct CollapseCtx {
    const Bvh2Node* nodes;
    const uint32_t* counts;
    const Tri48* tris_sorted;     /* leaf k -> triangle */
    uint32_t n_leaves;
    WideNode* out_nodes;
    Tri48* out_tris;
    uint32_t* node_counter;       /* next free wide node  */
    uint32_t* tri_counter;        /* next free triangle   */
    uint32_t* next_items;         /* pairs (bvh2 ref, wide index) for the next level */
    uint32_t* next_count;
    float* sah_accum;             /* [0] node term, [1] triangle term (area-weighted)   */
    const struct DpNode* dp;      /* optional: cost-optimal child selection (dp_node); nullptr = greedy by surface area */
};

RT3_HD uint32_t bvh2_slot(const CollapseCtx& c, uint32_t ref) {
    return (ref & LEAF_FLAG) ? (c.n_leaves - 1u) + (ref & ~LEAF_FLAG) : ref;
}
RT3_HD uint32_t bvh2_count(const CollapseCtx& c, uint32_t ref) { return (ref & LEAF_FLAG) ? 1u : c.counts[ref]; }

/* Smallest e with extent * 2^-e <= 255 (then widened by the caller if rounding pushes a child past 255). */
RT3_HD int quant_exponent(float extent) {
    if (!(extent > 0.0f)) return -100;
    float q = extent / 255.0f;
    int e = (int)((rt3m::as_uint(q) >> 23) & 0xFF) - 127;   /* floor(log2 q) for normal q */
    if (rt3m::exp2i(e < -126 ? -126 : e) < q) e += 1;
    if (e < -100) e = -100;
    if (e > 100) e = 100;
    return e;
}

/* ---- cost-optimal choice of the children of a wide node (Ylitie, Karras, Laine 2017, section 3.1) -----------------------------
 * For every BVH2 node n and i = 1..7, C(n, i) = least SAH cost of representing the subtree of n by at most i wide-BVH children:
 *   C(n, 1)      = leaf cost A(n) * P(n) * c_prim           if P(n) <= 3 (our leaf limit)
 *                = A(n) * c_node + D(n, 8)                  otherwise (n becomes an inner wide node)
 *   C(n, i)      = min(D(n, i), C(n, i - 1))                i = 2..7
 *   D(n, j)      = min over 0 < k < j of C(left, k) + C(right, j - k)
 * dec[i] records the argmin: 0 = "same as i - 1", else k = children given to the left subtree; dec[0] is the k of D(n, 8).
 * Computed bottom-up (dp_node, after the refit), consumed top-down by collapse_one. */
struct DpNode { float cost[7]; uint8_t dec[8]; uint32_t pad; };            /* 40 B per BVH2 internal node */
#define RT3_REFIT_PASSES 3      /* bottom-up passes of the default build (rotations in all, the programme in the last) */
#define RT3_DP_C_NODE 1.0f     /* cost of one wide-node test ...                                        */
#define RT3_DP_C_PRIM 0.3f     /* ... and of one triangle test (the result is flat between 0.2 and 1.2) */
RT3_HD void dp_leaf_costs(float area, float cost[7]) {
    for (int i = 0; i < 7; ++i) cost[i] = area * RT3_DP_C_PRIM;
}
RT3_HD void dp_node(const float cl[7], const float cr[7], float area, uint32_t count, DpNode& out) {
    /* D(n, j), j = 2..8 */
    float D[9]; uint8_t K[9];
    for (int j = 2; j <= 8; ++j) {
        float best = 3.4e38f; int bk = 1;
        for (int k = 1; k < j; ++k) {
            if (k > 7 || j - k > 7) continue;
            float v = cl[k - 1] + cr[j - k - 1];
            if (v < best) { best = v; bk = k; }
        }
        D[j] = best; K[j] = (uint8_t)bk;
    }
    out.dec[0] = K[8];
    out.cost[0] = count <= 3u ? area * (float)count * RT3_DP_C_PRIM : area * RT3_DP_C_NODE + D[8];
    for (int i = 2; i <= 7; ++i) {
        if (D[i] < out.cost[i - 2]) { out.cost[i - 1] = D[i]; out.dec[i] = K[i]; }
        else { out.cost[i - 1] = out.cost[i - 2]; out.dec[i] = 0; }
    }
    out.dec[1] = 0; out.pad = 0;
}
/* children of the wide node rooted at BVH2 internal node `ref`, from the decisions */
RT3_HD int dp_select_children(const Bvh2Node* nodes, const DpNode* dp, uint32_t ref, uint32_t ch[8]) {
    uint32_t st_ref[8]; int st_i[8]; int sp = 0, nc = 0;
    const int k8 = dp[ref].dec[0];
    st_ref[sp] = nodes[ref].right; st_i[sp++] = 8 - k8;
    st_ref[sp] = nodes[ref].left; st_i[sp++] = k8;
    while (sp) {
        uint32_t m = st_ref[--sp]; int i = st_i[sp];
        if (m & LEAF_FLAG) { ch[nc++] = m; continue; }
        while (i > 1 && dp[m].dec[i] == 0) --i;
        if (i <= 1) { ch[nc++] = m; continue; }
        const int k = dp[m].dec[i];
        st_ref[sp] = nodes[m].right; st_i[sp++] = i - k;
        st_ref[sp] = nodes[m].left; st_i[sp++] = k;
    }
    return nc;
}

/* ---- one step of the bottom-up pass over the BVH2 (the thread that arrives second at `cur` runs it; both subtrees are final) ------
 * 1. optional tree rotations (Kensler 2008): exchange a child of `cur` with a grandchild on the other side, or two grandchildren of
 *    different sides, when that shrinks the surface area of the rebuilt child(ren) — the SAH term those nodes contribute.  Only
 *    nodes below `cur` change, so threads still walking elsewhere are not affected; the choice depends on the tree alone, not on
 *    which thread got here, so the build stays deterministic.
 * 2. bounds and triangle count of `cur`, 3. optionally its row of the collapse programme (dp_node).
 * `Mem` hides how nodes written by OTHER threads of the same launch are read (device: ld.global.cg past the non-coherent L1;
 * emulator: plain arrays).  Leaves are referenced as LEAF_FLAG | k and stored at slot n - 1 + k (box only). */
template <class Mem>
RT3_HD void refit_visit(Mem& mem, uint32_t n, uint32_t cur, bool rotate, bool with_dp) {
    Bvh2Node P = mem.node(cur);
    uint32_t L = P.left, R = P.right;
    Bvh2Node nl = mem.node(mem.slot(L)), nr = mem.node(mem.slot(R));
    /* (re)build internal node `id` from two finished subtrees; ka / kb: programme rows of a / b when they were rebuilt by this very
     * visit (otherwise they are read from memory), out: the row of `id` */
    auto rebuild = [&](uint32_t id, uint32_t a_ref, const Bvh2Node& a, uint32_t na, const float* ka, uint32_t b_ref, const Bvh2Node& b, uint32_t nb, const float* kb,
                       float* out) {
        Bvh2Node m;
        box_union(a, b, m.lo, m.hi);
        m.left = a_ref; m.right = b_ref;
        mem.store_node(id, m);
        const uint32_t cnt = na + nb;
        mem.store_count(id, cnt);
        if (with_dp) {
            float ca[7], cb[7];
            if (a_ref & LEAF_FLAG) dp_leaf_costs(box_area(a.lo, a.hi), ca); else if (ka) { for (int i = 0; i < 7; ++i) ca[i] = ka[i]; } else mem.dp_costs(a_ref, ca);
            if (b_ref & LEAF_FLAG) dp_leaf_costs(box_area(b.lo, b.hi), cb); else if (kb) { for (int i = 0; i < 7; ++i) cb[i] = kb[i]; } else mem.dp_costs(b_ref, cb);
            DpNode d;
            dp_node(ca, cb, box_area(m.lo, m.hi), cnt, d);
            mem.store_dp(id, d);
            if (out) for (int i = 0; i < 7; ++i) out[i] = d.cost[i];
        }
        return m;
    };
    float kl[7], kr[7];
    const float *pkl = nullptr, *pkr = nullptr;
    uint32_t cL = mem.count(L), cR = mem.count(R);
    if (rotate) {
        const bool li = !(L & LEAF_FLAG), ri = !(R & LEAF_FLAG);
        Bvh2Node ll = nl, lr = nl, rl = nr, rr = nr;
        if (li) { ll = mem.node(mem.slot(nl.left)); lr = mem.node(mem.slot(nl.right)); }
        if (ri) { rl = mem.node(mem.slot(nr.left)); rr = mem.node(mem.slot(nr.right)); }
        const float aL = box_area(nl.lo, nl.hi), aR = box_area(nr.lo, nr.hi);
        float best = 0.0f; int opt = 0;
        if (ri) {
            const float g1 = aR - union_area(nl, rr);        /* L <-> R.left  : R' = (L, R.right)  */
            const float g2 = aR - union_area(rl, nl);        /* L <-> R.right : R' = (R.left, L)   */
            if (g1 > best) { best = g1; opt = 1; }
            if (g2 > best) { best = g2; opt = 2; }
        }
        if (li) {
            const float g3 = aL - union_area(nr, lr);        /* R <-> L.left  : L' = (R, L.right)  */
            const float g4 = aL - union_area(ll, nr);        /* R <-> L.right : L' = (L.left, R)   */
            if (g3 > best) { best = g3; opt = 3; }
            if (g4 > best) { best = g4; opt = 4; }
        }
        if (li && ri) {
            const float g5 = (aL + aR) - (union_area(rl, lr) + union_area(ll, rr));   /* L.left <-> R.left  */
            const float g6 = (aL + aR) - (union_area(rr, lr) + union_area(rl, ll));   /* L.left <-> R.right */
            if (g5 > best) { best = g5; opt = 5; }
            if (g6 > best) { best = g6; opt = 6; }
        }
        const uint32_t ll_ref = nl.left, lr_ref = nl.right, rl_ref = nr.left, rr_ref = nr.right;
        const uint32_t cll = li ? mem.count(ll_ref) : 0u, clr = li ? mem.count(lr_ref) : 0u, crl = ri ? mem.count(rl_ref) : 0u, crr = ri ? mem.count(rr_ref) : 0u;
        if (opt == 1) { nr = rebuild(R, L, nl, cL, nullptr, rr_ref, rr, crr, nullptr, kr); pkr = kr; cR = cL + crr; nl = rl; L = rl_ref; cL = crl; }
        else if (opt == 2) { nr = rebuild(R, rl_ref, rl, crl, nullptr, L, nl, cL, nullptr, kr); pkr = kr; cR = crl + cL; nl = rr; L = rr_ref; cL = crr; }
        else if (opt == 3) { nl = rebuild(L, R, nr, cR, nullptr, lr_ref, lr, clr, nullptr, kl); pkl = kl; cL = cR + clr; nr = ll; R = ll_ref; cR = cll; }
        else if (opt == 4) { nl = rebuild(L, ll_ref, ll, cll, nullptr, R, nr, cR, nullptr, kl); pkl = kl; cL = cll + cR; nr = lr; R = lr_ref; cR = clr; }
        else if (opt == 5) {
            nl = rebuild(L, rl_ref, rl, crl, nullptr, lr_ref, lr, clr, nullptr, kl); nr = rebuild(R, ll_ref, ll, cll, nullptr, rr_ref, rr, crr, nullptr, kr);
            pkl = kl; pkr = kr; cL = crl + clr; cR = cll + crr;
        } else if (opt == 6) {
            nl = rebuild(L, rr_ref, rr, crr, nullptr, lr_ref, lr, clr, nullptr, kl); nr = rebuild(R, rl_ref, rl, crl, nullptr, ll_ref, ll, cll, nullptr, kr);
            pkl = kl; pkr = kr; cL = crr + clr; cR = crl + cll;
        }
    }
    rebuild(cur, L, nl, cL, pkl, R, nr, cR, pkr, nullptr);
}

/* ---- treelet restructuring (Karras & Aila 2013, section 3) as an alternative to the rotations of refit_visit ---------------------
 * The second thread to arrive at `cur` grows a treelet below it — up to RT3_TREELET_LEAVES subtrees, always opening the subtree of
 * largest surface area — and finds, by dynamic programming over the subsets of those subtrees, the binary topology of least summed
 * surface area:  copt(S) = A(S) + min over partitions P of S of copt(P) + copt(S \ P),  copt({l}) = 0.  The treelet's internal nodes
 * (cur and the nodes that were opened) are re-linked in place when that beats the topology found; subtrees below the treelet's leaves
 * are untouched, so, as with the rotations, only nodes below `cur` change and the result depends on the tree alone.
 * Rotations are the special case of 3- and 4-leaf treelets. */
#ifndef RT3_TREELET_LEAVES
#define RT3_TREELET_LEAVES 7      /* 5 / 7 / 9 leaves: 8.82 / 8.53 / 8.54 work units per incoherent ray on the bench scene (rotations: 9.12) */
#endif
#ifndef RT3_TREELET_AUTO_MAX
#define RT3_TREELET_AUTO_MAX (1u << 22)   /* FAST_TRACE restructures treelets in trees of up to 4 M records by default (build.cu) */
#endif
#define RT3_TREELET_PASSES 2       /* 1 / 2 / 3 / 5 passes: 8.92 / 8.55 / 8.53 / 8.52 */
/* Subtrees of fewer than RT3_TREELET_MIN_LEAVES triangles are not restructured on their own (Karras & Aila 2013 section 3.4, gamma): they are
 * covered by the treelets of their ancestors, and most internal nodes of a binary tree are that small, so the threshold removes most of the
 * dynamic-programming runs.  Such nodes get the cheap visit (RT3_TREELET_SMALL_ROTATE: with the tree rotations, else bounds only). */
#ifndef RT3_TREELET_MIN_LEAVES
#define RT3_TREELET_MIN_LEAVES 7      /* 0 / 7 / 16 / 32: 8.558 / 8.558 / 8.602 / 8.646 work units per incoherent ray on the bench scene */
#endif
#ifndef RT3_TREELET_SMALL_ROTATE
#define RT3_TREELET_SMALL_ROTATE 1
#endif
/* The dynamic programme over the subsets of a treelet's leaves, serial form (the DEFINITION: the CPU emulation runs it, and the device's
 * warp-cooperative solver in build.cu must return the same copt / part tables, ties included).  copt[S] = least summed area of the internal
 * nodes of a binary tree over the leaves in S (the treelet root's own area excluded), part[S] = the child set of that tree's root which
 * holds S's lowest member. */
struct SerialTreeletDp {
    float area[1 << RT3_TREELET_LEAVES], copt[1 << RT3_TREELET_LEAVES];
    uint16_t part[1 << RT3_TREELET_LEAVES];
    RT3_HD void solve(const Bvh2Node* leaf_box, int nl) {
        const int K = RT3_TREELET_LEAVES;
        const uint32_t full = (1u << nl) - 1u;
        /* box of a subset = union of the boxes of its members in the low three and in the remaining places: 8 + 2^(K-3) partial unions,
         * then one union per subset (min / max are exact, so the order of the unions does not matter) */
        {
            const int KH = K > 3 ? K - 3 : 0;
            float plo[8][3], phi[8][3], qlo[1 << KH][3], qhi[1 << KH][3];
            for (int g = 0; g < 2; ++g) {
                const int first = g ? 3 : 0, cnt = g ? KH : 3;
                for (uint32_t sub = 0; sub < (1u << cnt); ++sub) {
                    float lo[3] = {3.4e38f, 3.4e38f, 3.4e38f}, hi[3] = {-3.4e38f, -3.4e38f, -3.4e38f};
                    for (int k = 0; k < cnt; ++k)
                        if ((sub & (1u << k)) && first + k < nl)
                            for (int q = 0; q < 3; ++q) { lo[q] = rt3m::min_(lo[q], leaf_box[first + k].lo[q]); hi[q] = rt3m::max_(hi[q], leaf_box[first + k].hi[q]); }
                    for (int q = 0; q < 3; ++q) { if (g) { qlo[sub][q] = lo[q]; qhi[sub][q] = hi[q]; } else { plo[sub][q] = lo[q]; phi[sub][q] = hi[q]; } }
                }
            }
            for (uint32_t sset = 1; sset <= full; ++sset) {
                float lo[3], hi[3];
                for (int q = 0; q < 3; ++q) { lo[q] = rt3m::min_(plo[sset & 7u][q], qlo[sset >> 3][q]); hi[q] = rt3m::max_(phi[sset & 7u][q], qhi[sset >> 3][q]); }
                area[sset] = box_area(lo, hi);
                copt[sset] = 0.0f; part[sset] = 0;
            }
        }
        /* subsets in increasing numeric order: every proper subset of S is smaller than S, so its copt is final when S is reached */
        for (uint32_t sset = 3; sset <= full; ++sset) {
            if ((sset & (sset - 1u)) == 0u) continue;                                  /* singleton */
            float bc = 3.4e38f; uint32_t bp = 0;
            const uint32_t low = sset & (0u - sset);                                    /* the partition holding the lowest member is P: halves the search */
            const uint32_t rest = sset ^ low;
            for (uint32_t sub = (rest - 1u) & rest;; sub = (sub - 1u) & rest) {        /* P = low | sub for every proper subset sub of rest (incl. empty) */
                const uint32_t pp = low | sub;
                const float c = copt[pp] + copt[sset ^ pp];
                if (c < bc) { bc = c; bp = pp; }
                if (sub == 0u) break;
            }
            copt[sset] = (sset == full ? 0.0f : area[sset]) + bc;                       /* the root's own area is the same for every topology */
            part[sset] = (uint16_t)bp;
        }
    }
    RT3_HD float cost_of(uint32_t sset) const { return copt[sset]; }
    RT3_HD uint32_t part_of(uint32_t sset) const { return part[sset]; }
};

/* Dp: SerialTreeletDp, or the device's warp-cooperative solver (then every lane of the warp runs this function with the same arguments and
 * only the treelet's owner lane stores: Mem's store_* are predicated, build.cu). */
template <class Mem, class Dp>
RT3_HD void treelet_visit_t(Mem& mem, uint32_t n, uint32_t cur, bool with_dp, Dp& dp) {
    const int K = RT3_TREELET_LEAVES;
    if (RT3_TREELET_MIN_LEAVES > 0) {
        const Bvh2Node P0 = mem.node(cur);
        if (mem.count(P0.left) + mem.count(P0.right) < (uint32_t)RT3_TREELET_MIN_LEAVES) { refit_visit(mem, n, cur, RT3_TREELET_SMALL_ROTATE != 0, with_dp); return; }
    }
    uint32_t leaf_ref[K], inner_id[K];
    Bvh2Node leaf_box[K];
    float leaf_area[K];
    int nl = 0, ni = 0;
    {
        const Bvh2Node P = mem.node(cur);
        leaf_ref[0] = P.left; leaf_ref[1] = P.right; nl = 2;
        inner_id[ni++] = cur;
        for (int k = 0; k < 2; ++k) { leaf_box[k] = mem.node(mem.slot(leaf_ref[k])); leaf_area[k] = box_area(leaf_box[k].lo, leaf_box[k].hi); }
    }
    float found_cost = 0.0f;                       /* summed area of the internal nodes below cur as they are now */
    while (nl < K) {
        int best = -1; float ba = -1.0f;
        for (int k = 0; k < nl; ++k) if (!(leaf_ref[k] & LEAF_FLAG) && leaf_area[k] > ba) { ba = leaf_area[k]; best = k; }
        if (best < 0) break;
        const Bvh2Node open = leaf_box[best];
        found_cost += ba;
        inner_id[ni++] = leaf_ref[best];
        leaf_ref[best] = open.left; leaf_ref[nl] = open.right;
        leaf_box[best] = mem.node(mem.slot(open.left)); leaf_box[nl] = mem.node(mem.slot(open.right));
        leaf_area[best] = box_area(leaf_box[best].lo, leaf_box[best].hi); leaf_area[nl] = box_area(leaf_box[nl].lo, leaf_box[nl].hi);
        nl++;
    }
    const uint32_t full = (1u << nl) - 1u;
    dp.solve(leaf_box, nl);
    const bool restructure = nl > 2 && dp.cost_of(full) < found_cost * 0.99999f;
    /* the topology to write: node list in pre-order (set, id), children resolved after; written in reverse (children first) */
    uint32_t t_set[K], t_id[K]; int t_n = 0;
    uint16_t t_l[K], t_r[K];                                                        /* child sets */
    if (restructure) {
        uint32_t st[K]; int sp = 0, next_inner = 1;
        st[sp++] = full; t_id[0] = cur;
        /* pre-order with explicit stack; ids handed out in visit order (inner_id[0] = cur is the root) */
        uint8_t idx_of_set[1 << K];
        while (sp) {
            const uint32_t sset = st[--sp];
            const int me = t_n++;
            t_set[me] = sset;
            if (me > 0) t_id[me] = inner_id[next_inner++];
            idx_of_set[sset] = (uint8_t)me;
            const uint32_t pl = dp.part_of(sset), pr = sset ^ pl;
            t_l[me] = (uint16_t)pl; t_r[me] = (uint16_t)pr;
            if (pr & (pr - 1u)) st[sp++] = pr;
            if (pl & (pl - 1u)) st[sp++] = pl;
        }
        /* children first: a node's children that are treelet-internal come later in pre-order, so walk backwards */
        Bvh2Node nb[K]; uint32_t ncount[K]; float nrow[K][7];
        for (int me = t_n - 1; me >= 0; --me) {
            Bvh2Node cb[2]; uint32_t cref[2], cc[2]; float crow[2][7];
            for (int side = 0; side < 2; ++side) {
                const uint32_t cs = side ? t_r[me] : t_l[me];
                if ((cs & (cs - 1u)) == 0u) {
                    const int k = bfind32(cs);
                    cref[side] = leaf_ref[k]; cb[side] = leaf_box[k]; cc[side] = mem.count(leaf_ref[k]);
                    if (with_dp) { if (leaf_ref[k] & LEAF_FLAG) dp_leaf_costs(leaf_area[k], crow[side]); else mem.dp_costs(leaf_ref[k], crow[side]); }
                } else {
                    const int ci = idx_of_set[cs];
                    cref[side] = t_id[ci]; cb[side] = nb[ci]; cc[side] = ncount[ci];
                    if (with_dp) for (int i = 0; i < 7; ++i) crow[side][i] = nrow[ci][i];
                }
            }
            Bvh2Node m;
            box_union(cb[0], cb[1], m.lo, m.hi);
            m.left = cref[0]; m.right = cref[1];
            mem.store_node(t_id[me], m);
            mem.store_count(t_id[me], cc[0] + cc[1]);
            nb[me] = m; ncount[me] = cc[0] + cc[1];
            if (with_dp) {
                DpNode d;
                dp_node(crow[0], crow[1], box_area(m.lo, m.hi), cc[0] + cc[1], d);
                mem.store_dp(t_id[me], d);
                for (int i = 0; i < 7; ++i) nrow[me][i] = d.cost[i];
            }
        }
    } else {
        refit_visit(mem, n, cur, false, with_dp);                                   /* topology kept: bounds, count and programme row of cur */
    }
}
template <class Mem>
RT3_HD void treelet_visit(Mem& mem, uint32_t n, uint32_t cur, bool with_dp) {
    SerialTreeletDp dp;
    treelet_visit_t(mem, n, cur, with_dp, dp);
}

/*
 * Node layout (20 words):
 *   w0..w2  p = node box min (float)            w3  ex | ey<<8 | ez<<16 | imask<<24   (e biased by 127)
 *   w4      child_base (first inner child)      w5  tri_base
 *   w6      valid24: leaf child in slot s with c records (1..3) owns bits 3 s .. 3 s + c - 1; inner and empty slots own none.  The records
 *           of a node lie consecutively from tri_base in slot order, so record bit b has index tri_base + popc(valid24 & ((1 << b) - 1)).
 *   w7      0 (reserved)
 *           (RT3_NODE_V2 = 0, round 1's format: w6,w7 = meta[8]: 0 = empty; inner: 0x20 | (24 + slot); leaf: unary(count) << 5 | tri offset)
 *   w8,w9   qlo.x[8]   w10,w11 qlo.y[8]   w12,w13 qlo.z[8]
 *   w14,w15 qhi.x[8]   w16,w17 qhi.y[8]   w18,w19 qhi.z[8]
 * child box = p + q * 2^e, conservative (floor / ceil).
 */
RT3_HD void collapse_one(const CollapseCtx& c, uint32_t ref, uint32_t wide_index) {
    uint32_t ch[8];
    int nc;
    uint32_t self = bvh2_slot(c, ref);
    if (ref & LEAF_FLAG) { ch[0] = ref; nc = 1; }
    else if (c.dp != nullptr && c.counts[ref] > 3u) nc = dp_select_children(c.nodes, c.dp, ref, ch);
    else { ch[0] = c.nodes[ref].left; ch[1] = c.nodes[ref].right; nc = 2; }
    /* greedy: open the inner child of largest surface area until 8 slots are used */
    while (c.dp == nullptr && nc < 8) {
        int best = -1; float best_a = -1.0f;
        for (int k = 0; k < nc; ++k) {
            if (ch[k] & LEAF_FLAG) continue;
            const Bvh2Node& nd = c.nodes[ch[k]];
            float a = box_area(nd.lo, nd.hi);
            if (a > best_a) { best_a = a; best = k; }
        }
        if (best < 0) break;
        uint32_t r = ch[best];
        ch[best] = c.nodes[r].left;
        ch[nc++] = c.nodes[r].right;
    }
    const Bvh2Node& me = c.nodes[self];
    float cx = 0.5f * (me.lo[0] + me.hi[0]), cy = 0.5f * (me.lo[1] + me.hi[1]), cz = 0.5f * (me.lo[2] + me.hi[2]);
    /* octant slot assignment: greedy maximisation of sum dot(child centre - node centre, corner direction) */
    float cost[8][8];
    for (int k = 0; k < nc; ++k) {
        const Bvh2Node& nd = c.nodes[bvh2_slot(c, ch[k])];
        float dx = 0.5f * (nd.lo[0] + nd.hi[0]) - cx, dy = 0.5f * (nd.lo[1] + nd.hi[1]) - cy, dz = 0.5f * (nd.lo[2] + nd.hi[2]) - cz;
        for (int s = 0; s < 8; ++s) cost[k][s] = ((s & 1) ? dx : -dx) + ((s & 2) ? dy : -dy) + ((s & 4) ? dz : -dz);
    }
    int slot_child[8];
    for (int s = 0; s < 8; ++s) slot_child[s] = -1;
    uint32_t child_done = 0;
    for (int it = 0; it < nc; ++it) {
        int bk = -1, bs = -1; float bc = -3.4e38f;
        for (int k = 0; k < nc; ++k) {
            if (child_done & (1u << k)) continue;
            for (int s = 0; s < 8; ++s) {
                if (slot_child[s] >= 0) continue;
                if (cost[k][s] > bc) { bc = cost[k][s]; bk = k; bs = s; }
            }
        }
        slot_child[bs] = bk; child_done |= 1u << bk;
    }
    /* classify, count */
    uint32_t n_inner = 0, n_tris = 0, imask = 0;
    for (int s = 0; s < 8; ++s) {
        int k = slot_child[s];
        if (k < 0) continue;
        uint32_t cnt = bvh2_count(c, ch[k]);
        if (cnt <= 3u) n_tris += cnt; else { n_inner++; imask |= 1u << s; }
    }
    uint32_t child_base = n_inner ? RT3_ATOMIC_ADD_U32(c.node_counter, n_inner) : 0u;
    uint32_t tri_base = n_tris ? RT3_ATOMIC_ADD_U32(c.tri_counter, n_tris) : 0u;
    /* quantisation grid */
    int e[3];
    for (int a = 0; a < 3; ++a) e[a] = quant_exponent(me.hi[a] - me.lo[a]);
    uint32_t qlo[3][8], qhi[3][8];
    for (int a = 0; a < 3; ++a) {
        for (;;) {
            float inv = rt3m::exp2i(-e[a]);
            bool ok = true;
            for (int s = 0; s < 8; ++s) {
                int k = slot_child[s];
                if (k < 0) { qlo[a][s] = 255; qhi[a][s] = 0; continue; }
                const Bvh2Node& nd = c.nodes[bvh2_slot(c, ch[k])];
                float fl = rt3m::floor_((nd.lo[a] - me.lo[a]) * inv);
                float fh = -rt3m::floor_(-((nd.hi[a] - me.lo[a]) * inv));
                if (fl < 0.0f) fl = 0.0f;
                if (fh > 255.0f) { ok = false; break; }
                if (fh < fl) fh = fl;
                /* make sure the decoded planes really enclose the child (guards the rounding of the subtraction) */
                float scale = rt3m::exp2i(e[a]);
                if (rt3m::fma_(fl, scale, me.lo[a]) > nd.lo[a] && fl > 0.0f) fl -= 1.0f;
                if (rt3m::fma_(fh, scale, me.lo[a]) < nd.hi[a]) { fh += 1.0f; if (fh > 255.0f) { ok = false; break; } }
                qlo[a][s] = (uint32_t)fl; qhi[a][s] = (uint32_t)fh;
            }
            if (ok) break;
            e[a] += 1;
        }
    }
    WideNode out;
    out.w[0] = rt3m::as_uint(me.lo[0]); out.w[1] = rt3m::as_uint(me.lo[1]); out.w[2] = rt3m::as_uint(me.lo[2]);
    out.w[3] = (uint32_t)(e[0] + 127) | ((uint32_t)(e[1] + 127) << 8) | ((uint32_t)(e[2] + 127) << 16) | (imask << 24);
    out.w[4] = child_base; out.w[5] = tri_base;
    uint32_t meta[8];
    uint32_t tri_off = 0, inner_rank = 0;
    float my_area = box_area(me.lo, me.hi);
    float tri_term = 0.0f;
    for (int s = 0; s < 8; ++s) {
        int k = slot_child[s];
        if (k < 0) { meta[s] = 0; continue; }
        uint32_t cnt = bvh2_count(c, ch[k]);
        if (cnt > 3u) {
            meta[s] = 0x20u | (24u + (uint32_t)s);
            uint32_t q = RT3_ATOMIC_ADD_U32(c.next_count, 1u);
            c.next_items[2 * q] = ch[k];
            c.next_items[2 * q + 1] = child_base + inner_rank;
            inner_rank++;
        } else {
            meta[s] = ((cnt == 1u ? 1u : (cnt == 2u ? 3u : 7u)) << 5) | tri_off;
            /* gather the (<= 3) triangles below ch[k] */
            uint32_t st[4]; int sp = 0; st[sp++] = ch[k];
            while (sp) {
                uint32_t r = st[--sp];
                if (r & LEAF_FLAG) { c.out_tris[tri_base + tri_off] = c.tris_sorted[r & ~LEAF_FLAG]; tri_off++; }
                else { st[sp++] = c.nodes[r].right; st[sp++] = c.nodes[r].left; }
            }
            const Bvh2Node& nd = c.nodes[bvh2_slot(c, ch[k])];
            tri_term += box_area(nd.lo, nd.hi) * (float)cnt;
        }
    }
#if RT3_NODE_V2
    uint32_t valid24 = 0;
    for (int s = 0; s < 8; ++s)
        if (meta[s] != 0u && (meta[s] & 0x1Fu) < 24u) valid24 |= (meta[s] >> 5) << (3 * s);       /* leaf: unary(count) at 3 s */
    out.w[6] = valid24;
    out.w[7] = 0u;
#else
    out.w[6] = meta[0] | (meta[1] << 8) | (meta[2] << 16) | (meta[3] << 24);
    out.w[7] = meta[4] | (meta[5] << 8) | (meta[6] << 16) | (meta[7] << 24);
#endif
    for (int a = 0; a < 3; ++a) {
        out.w[8 + 2 * a] = qlo[a][0] | (qlo[a][1] << 8) | (qlo[a][2] << 16) | (qlo[a][3] << 24);
        out.w[9 + 2 * a] = qlo[a][4] | (qlo[a][5] << 8) | (qlo[a][6] << 16) | (qlo[a][7] << 24);
        out.w[14 + 2 * a] = qhi[a][0] | (qhi[a][1] << 8) | (qhi[a][2] << 16) | (qhi[a][3] << 24);
        out.w[15 + 2 * a] = qhi[a][4] | (qhi[a][5] << 8) | (qhi[a][6] << 16) | (qhi[a][7] << 24);
    }
    c.out_nodes[wide_index] = out;
    if (c.sah_accum) {
        RT3_ATOMIC_ADD_F32(c.sah_accum, my_area);
        RT3_ATOMIC_ADD_F32(c.sah_accum + 1, tri_term);
    }
}

/* ---------------------------------------------------------------- traversal -------------------- */
struct PermLut { uint8_t v[2048]; };
constexpr PermLut make_perm_lut() {
    PermLut t{};
    for (int o = 0; o < 8; ++o)
        for (int b = 0; b < 256; ++b) {
            unsigned r = 0;
            for (int s = 0; s < 8; ++s) if ((b >> s) & 1) r |= 1u << (s ^ o);
            t.v[o * 256 + b] = (uint8_t)r;
        }
    return t;
}
#if RT3_NODE_V2
#ifdef __CUDACC__
static __device__ const PermLut g_perm_lut = make_perm_lut();
#endif
#ifdef __CUDA_ARCH__
/* every kernel that traverses copies the table into shared memory once per CTA (2 KB: one 16-byte word per thread of a 128-thread CTA;
 * NT = the kernel's CTA size, a compile-time constant) */
#define RT3_PERM_LUT_SETUP(base, NT)                                                                                               \
    __shared__ __align__(2048) uint8_t s_perm_lut_[2048];                                                                          \
    _Pragma("unroll") for (uint32_t i_ = 0; i_ < 128u; i_ += (NT))                                                                 \
        if ((NT) >= 128 ? threadIdx.x < 128u : true) ((uint4*)s_perm_lut_)[i_ + threadIdx.x] = __ldg((const uint4*)g_perm_lut.v + i_ + threadIdx.x); \
    __syncthreads();                                                                                                               \
    const uint32_t base = (uint32_t)__cvta_generic_to_shared(s_perm_lut_)
RT3_HD uint32_t perm_lookup(uint32_t addr) {
    uint32_t v;
    asm("ld.shared.u8 %0, [%1];" : "=r"(v) : "r"(addr));
    return v;
}
#else
#define RT3_PERM_LUT_SETUP(base, NT) const uint32_t base = 0u
inline uint32_t perm_lookup(uint32_t addr) {
    static const PermLut lut = make_perm_lut();
    return lut.v[addr & 2047u];
}
#endif
#else
#define RT3_PERM_LUT_SETUP(base, NT) const uint32_t base = 0u
#endif
struct RayPrep {
    float ox, oy, oz, dx, dy, dz;     /* origin, direction                        */
    float idx, idy, idz;               /* safe reciprocal direction                */
    uint32_t octinv4;                  /* (7 - octant) replicated in 4 bytes       */
#if RT3_NODE_V2
    uint32_t lutrow;                   /* address of row (7 - octant) of the permutation table (shared memory; host: index) */
#endif
    float tmin;
    uint32_t magic;                    /* 0x47000000 held in a REGISTER: ptxas then encodes the PRMT selector as the immediate
                                          (with both constant it materialises each selector with an extra move per plane) */
};
RT3_HD float safe_rcp_dir(float d) {
    const float tiny = 1e-20f;
    if (rt3m::abs_(d) < tiny) d = (rt3m::as_uint(d) >> 31) ? -tiny : tiny;
    return 1.0f / d;
}
/* FAST (k_trace_dyn only): the reciprocal as MUFU.RCP + one Newton step, 3 instructions and no branch instead of the 15 of an IEEE division with
 * its range check; |d| is in [1e-20, ~1], so no special case can occur.  The result is within 1 ulp of 1 / d (an IEEE division: 1/2 ulp).  It
 * only feeds the slab tests, which stay conservative: the extra 2^-24 relative error of a plane distance is one more |p - o| term in the bound
 * that RT3_PAD_FRACTION covers (origins within ~125 scene extents instead of ~250).  Hits do not depend on it; the warp emulation and
 * traverse_wide (whose visit counts the tests compare with the emulator's) keep the exact division. */
template <bool FAST>
RT3_HD float safe_rcp_dir_t(float d) {
#ifdef __CUDA_ARCH__
    if (FAST) {
        const float tiny = 1e-20f;
        if (rt3m::abs_(d) < tiny) d = (rt3m::as_uint(d) >> 31) ? -tiny : tiny;
        float r;
        asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(d));
        return rt3m::fma_(r, rt3m::fma_(-d, r, 1.0f), r);
    }
#endif
    return safe_rcp_dir(d);
}
template <bool FAST = false>
RT3_HD RayPrep prepare_ray(const rt3_ray& r, uint32_t magic = 0x47000000u, uint32_t lut_base = 0u) {
    RayPrep p;
    p.magic = magic;
    p.ox = r.o[0]; p.oy = r.o[1]; p.oz = r.o[2];
    p.dx = r.d[0]; p.dy = r.d[1]; p.dz = r.d[2];
    p.idx = safe_rcp_dir_t<FAST>(p.dx); p.idy = safe_rcp_dir_t<FAST>(p.dy); p.idz = safe_rcp_dir_t<FAST>(p.dz);
    uint32_t oct = (p.idx < 0.0f ? 1u : 0u) | (p.idy < 0.0f ? 2u : 0u) | (p.idz < 0.0f ? 4u : 0u);
    p.octinv4 = (7u - oct) * 0x01010101u;
#if RT3_NODE_V2
    p.lutrow = lut_base + ((7u - oct) << 8);
#else
    (void)lut_base;
#endif
    p.tmin = r.tmin;
    return p;
}

RT3_HD uint32_t byte_of(uint32_t w, int j) { return (w >> (8 * j)) & 0xFFu; }

/* Byte j of w as the float 32768 + q, built with one byte permute instead of an integer-to-float conversion
 * (I2F runs at a quarter of the FMA rate): bits = 0x47000000 | q << 8, i.e. exponent 2^15 with q in mantissa bits 8..15. */
template <int J>
RT3_HD float byte_as_biased_float(uint32_t w, uint32_t magic) {
#ifdef __CUDA_ARCH__
    uint32_t r;
    asm("prmt.b32 %0, %1, %2, %3;" : "=r"(r) : "r"(w), "r"(magic), "n"(0x7604 | (J << 4)));
    return __uint_as_float(r);
#else
    return rt3m::as_float(magic | (((w >> (8 * J)) & 0xFFu) << 8));
#endif
}
/* min / max for the slab test only: no NaN can occur there (directions are clamped away from 0), so the
 * single-instruction FMNMX forms are exact equivalents of the a<b?a:b forms used by the shader math. */
RT3_HD float bmin(float a, float b) {
#ifdef __CUDA_ARCH__
    return fminf(a, b);
#else
    return a < b ? a : b;
#endif
}
RT3_HD float bmax(float a, float b) {
#ifdef __CUDA_ARCH__
    return fmaxf(a, b);
#else
    return a > b ? a : b;
#endif
}

template <int J>
RT3_HD uint32_t slab_child(uint32_t nx, uint32_t fx, uint32_t ny, uint32_t fy, uint32_t nz, uint32_t fz, float sx, float sy, float sz, float bx,
                           float by, float bz, float tmin, float tmax, uint32_t child_bits4, uint32_t bit_index4, uint32_t mg) {
    const float t0x = rt3m::fma_(byte_as_biased_float<J>(nx, mg), sx, bx), t1x = rt3m::fma_(byte_as_biased_float<J>(fx, mg), sx, bx);
    const float t0y = rt3m::fma_(byte_as_biased_float<J>(ny, mg), sy, by), t1y = rt3m::fma_(byte_as_biased_float<J>(fy, mg), sy, by);
    const float t0z = rt3m::fma_(byte_as_biased_float<J>(nz, mg), sz, bz), t1z = rt3m::fma_(byte_as_biased_float<J>(fz, mg), sz, bz);
    const float tn = bmax(bmax(t0x, t0y), bmax(t0z, tmin));
    const float tf = bmin(bmin(t1x, t1y), bmin(t1z, tmax));
    return (tn <= tf) ? (((child_bits4 >> (8 * J)) & 0xFFu) << ((bit_index4 >> (8 * J)) & 0xFFu)) : 0u;
}

template <int J>
RT3_HD bool slab_hit(uint32_t nx, uint32_t fx, uint32_t ny, uint32_t fy, uint32_t nz, uint32_t fz, float sx, float sy, float sz, float bx,
                     float by, float bz, float tmin, float tmax, uint32_t mg) {
    const float t0x = rt3m::fma_(byte_as_biased_float<J>(nx, mg), sx, bx), t1x = rt3m::fma_(byte_as_biased_float<J>(fx, mg), sx, bx);
    const float t0y = rt3m::fma_(byte_as_biased_float<J>(ny, mg), sy, by), t1y = rt3m::fma_(byte_as_biased_float<J>(fy, mg), sy, by);
    const float t0z = rt3m::fma_(byte_as_biased_float<J>(nz, mg), sz, bz), t1z = rt3m::fma_(byte_as_biased_float<J>(fz, mg), sz, bz);
    const float tn = bmax(bmax(t0x, t0y), bmax(t0z, tmin));
    const float tf = bmin(bmin(t1x, t1y), bmin(t1z, tmax));
    return tn <= tf;
}
/* index of record bit `bit` of a node: the records lie consecutively in bit order, valid = the node's valid24 (all ones for a group that
 * was made dense, see dense_records) */
RT3_HD uint32_t record_index(uint32_t valid, int bit) {
#if RT3_NODE_V2
    return (uint32_t)popc32(valid & ~(0xFFFFFFFFu << bit));
#else
    (void)valid;
    return (uint32_t)bit;
#endif
}
/* a pending record group in self-contained form (two-level mode parks groups on the stack, which has no room for valid24): bit k of the
 * result stands for record tri_base + k */
RT3_HD uint32_t dense_records(uint32_t hits, uint32_t valid) {
#if RT3_NODE_V2
    uint32_t d = 0;
    while (hits) { const int b = bfind32(hits); hits &= ~(1u << b); d |= 1u << record_index(valid, b); }
    return d;
#else
    (void)valid;
    return hits;
#endif
}
/* Intersect the 8 quantised child boxes of one node.  Returns the hit mask: bits 24..31 = inner
 * children in near-to-far priority order for this ray's octant, bits 0..23 = triangles of hit leaves.
 * Plane distance = (p + q 2^e - o) / d = (32768 + q) * s + (b - 32768 s) with s = 2^e / d, b = (p - o) / d: one PRMT and
 * one FFMA per plane.  The rounding of (b - 32768 s) costs at most 1/512 of a quantisation step, which the
 * builder's triangle-box padding (scene extent x RT3_PAD_FRACTION) covers. */
RT3_HD uint32_t intersect_wide_node(const RayPrep& r, float tmax, const u32x4& n0, const u32x4& n1, const u32x4& n2,
                                    const u32x4& n3, const u32x4& n4) {
    const uint32_t e = n0.w;
    const float sx = rt3m::as_float((e & 0xFFu) << 23) * r.idx;
    const float sy = rt3m::as_float(((e >> 8) & 0xFFu) << 23) * r.idy;
    const float sz = rt3m::as_float(((e >> 16) & 0xFFu) << 23) * r.idz;
    const float bx = rt3m::fma_(-32768.0f, sx, (rt3m::as_float(n0.x) - r.ox) * r.idx);
    const float by = rt3m::fma_(-32768.0f, sy, (rt3m::as_float(n0.y) - r.oy) * r.idy);
    const float bz = rt3m::fma_(-32768.0f, sz, (rt3m::as_float(n0.z) - r.oz) * r.idz);
#if RT3_NODE_V2
    uint32_t acc = 0;
#pragma unroll
    for (int half = 0; half < 2; ++half) {
        const uint32_t qlx = half ? n2.y : n2.x, qly = half ? n2.w : n2.z, qlz = half ? n3.y : n3.x;
        const uint32_t qhx = half ? n3.w : n3.z, qhy = half ? n4.y : n4.x, qhz = half ? n4.w : n4.z;
        const uint32_t nx = r.idx < 0.0f ? qhx : qlx, fx = r.idx < 0.0f ? qlx : qhx;
        const uint32_t ny = r.idy < 0.0f ? qhy : qly, fy = r.idy < 0.0f ? qly : qhy;
        const uint32_t nz = r.idz < 0.0f ? qhz : qlz, fz = r.idz < 0.0f ? qlz : qhz;
        /* the constants are disjoint, so the compiler turns every line into one predicated add of an immediate */
        if (slab_hit<0>(nx, fx, ny, fy, nz, fz, sx, sy, sz, bx, by, bz, r.tmin, tmax, r.magic)) acc |= (7u << (12 * half + 0)) | (1u << (24 + 4 * half));
        if (slab_hit<1>(nx, fx, ny, fy, nz, fz, sx, sy, sz, bx, by, bz, r.tmin, tmax, r.magic)) acc |= (7u << (12 * half + 3)) | (2u << (24 + 4 * half));
        if (slab_hit<2>(nx, fx, ny, fy, nz, fz, sx, sy, sz, bx, by, bz, r.tmin, tmax, r.magic)) acc |= (7u << (12 * half + 6)) | (4u << (24 + 4 * half));
        if (slab_hit<3>(nx, fx, ny, fy, nz, fz, sx, sy, sz, bx, by, bz, r.tmin, tmax, r.magic)) acc |= (7u << (12 * half + 9)) | (8u << (24 + 4 * half));
    }
    /* empty slots never hit (their planes are lo = 255 > hi = 0) and own no bit of imask / valid24 anyway */
    const uint32_t inner = perm_lookup(((acc >> 24) & (n0.w >> 24)) | r.lutrow);
    return (inner << 24) | (acc & n1.z & 0x00FFFFFFu);
}
#else
    uint32_t hitmask = 0;
#pragma unroll
    for (int half = 0; half < 2; ++half) {
        const uint32_t meta4 = half ? n1.w : n1.z;
        const uint32_t is_inner4 = (meta4 & (meta4 << 1)) & 0x10101010u;
        const uint32_t inner_mask4 = sign_extend_s8x4(is_inner4 << 3);
        const uint32_t bit_index4 = (meta4 ^ (r.octinv4 & inner_mask4)) & 0x1F1F1F1Fu;
        const uint32_t child_bits4 = (meta4 >> 5) & 0x07070707u;
        const uint32_t qlx = half ? n2.y : n2.x, qly = half ? n2.w : n2.z, qlz = half ? n3.y : n3.x;
        const uint32_t qhx = half ? n3.w : n3.z, qhy = half ? n4.y : n4.x, qhz = half ? n4.w : n4.z;
        const uint32_t nx = r.idx < 0.0f ? qhx : qlx, fx = r.idx < 0.0f ? qlx : qhx;
        const uint32_t ny = r.idy < 0.0f ? qhy : qly, fy = r.idy < 0.0f ? qly : qhy;
        const uint32_t nz = r.idz < 0.0f ? qhz : qlz, fz = r.idz < 0.0f ? qlz : qhz;
        hitmask |= slab_child<0>(nx, fx, ny, fy, nz, fz, sx, sy, sz, bx, by, bz, r.tmin, tmax, child_bits4, bit_index4, r.magic);
        hitmask |= slab_child<1>(nx, fx, ny, fy, nz, fz, sx, sy, sz, bx, by, bz, r.tmin, tmax, child_bits4, bit_index4, r.magic);
        hitmask |= slab_child<2>(nx, fx, ny, fy, nz, fz, sx, sy, sz, bx, by, bz, r.tmin, tmax, child_bits4, bit_index4, r.magic);
        hitmask |= slab_child<3>(nx, fx, ny, fy, nz, fz, sx, sy, sz, bx, by, bz, r.tmin, tmax, child_bits4, bit_index4, r.magic);
    }
    return hitmask;
}
#endif

struct TraverseStats { uint32_t nodes, tris; };

/* World -> object ray for the bottom level: same parameter t (the direction is not renormalised). */
RT3_HD rt3_ray object_ray(const InstRec& in, const rt3_ray& w) {
    rt3_ray o;
    const vec3 p = rt3m::xform_point(in.inv, rt3m::v3(w.o)), d = rt3m::xform_dir(in.inv, rt3m::v3(w.d));
    o.o[0] = p.x; o.o[1] = p.y; o.o[2] = p.z; o.d[0] = d.x; o.d[1] = d.y; o.d[2] = d.z; o.tmin = w.tmin; o.tmax = w.tmax;
    return o;
}
/* World-space (v0, e1, e2) of a bottom-level record: the instance transform applied to the stored vertices, then the edge
 * subtractions — the operations of k_flatten (build.cu) / world_triangle (rt3_shared_math.h) in the same order. */
RT3_HD void instance_triangle(const InstRec& in, const u32x4& a, const u32x4& b, const u32x4& c, vec3& v0, vec3& e1, vec3& e2) {
    vec3 pa = rt3m::v3(rt3m::as_float(a.x), rt3m::as_float(a.y), rt3m::as_float(a.z));
    vec3 pb = rt3m::v3(rt3m::as_float(b.x), rt3m::as_float(b.y), rt3m::as_float(b.z));
    vec3 pc = rt3m::v3(rt3m::as_float(c.x), rt3m::as_float(c.y), rt3m::as_float(c.z));
    if (!rt3m::xform_is_identity(in.xform)) { pa = rt3m::xform_point(in.xform, pa); pb = rt3m::xform_point(in.xform, pb); pc = rt3m::xform_point(in.xform, pc); }
    v0 = pa; e1 = pb - pa; e2 = pc - pa;
}
#define RT3_STACK_SENTINEL 0xFFFFFFFFu

/* Closest hit (ANY = false) or first hit found (ANY = true).  `hit.t` must come in as the ray's tmax and
 * hit.prim as RT3_MISS.  Returns true if something was hit.
 * insts == nullptr: single-level BVH over world-space triangles.  insts != nullptr: two-level — the top level's leaf records
 * are instances; entering one pushes the pending node group, the pending record group and a sentinel, switches to the
 * object-space ray and restarts at the instance's root; popping the sentinel restores the world-space ray. */
template <bool ANY>
RT3_HD bool traverse_wide(const WideNode* nodes, const Tri48* tris, const rt3_ray& ray, rt3_hit& hit, TraverseStats* stats, bool active = true,
                          const InstRec* insts = nullptr, uint32_t lut_base = 0u) {
    RayPrep r = prepare_ray(ray, 0x47000000u, lut_base);
    const vec3 o = rt3m::v3(ray.o), d = rt3m::v3(ray.d);
    uint32_t stack_base[RT3_MAX_STACK], stack_hits[RT3_MAX_STACK];
    int sp = 0;
    uint32_t ng_base = 0, ng_hits = 0x80000000u;   /* node group: root */
    uint32_t tg_base = 0, tg_hits = 0;             /* record group     */
    uint32_t tg_valid = 0x00FFFFFFu;               /* ... and its node's valid24 (record_index) */
    uint32_t cur_inst = RT3_STACK_SENTINEL;        /* two-level: instance being traversed, or none (top level) */
    /* Two-phase iteration (same shape as k_trace_dyn): a node step for rays with no record pending, then ONE record
     * per ray that has one pending.  On the device the loop is warp-synchronous (every lane of the warp must call this
     * function; lanes without a ray pass active = false) so that both phases run converged. */
#ifdef __CUDA_ARCH__
#define RT3_WARP_ANY(x) __any_sync(0xFFFFFFFFu, (x))
#elif defined(RT3_WARP_EMU)
#define RT3_WARP_ANY(x) (__ballot_sync(0xFFFFFFFFu, (x)) != 0u)      /* tests/emu/warp_emu.cpp: 32 lanes in lock step on the CPU */
#else
#define RT3_WARP_ANY(x) (x)
#endif
#ifndef RT3_EMU_PHASE
#define RT3_EMU_PHASE(id, pred) do { } while (0)                      /* phase statistics of the warp emulation; empty everywhere else */
#endif
    while (RT3_WARP_ANY(active)) {
        RT3_EMU_PHASE(5, active && tg_hits == 0u && ng_hits > 0x00FFFFFFu);
        if (active && tg_hits == 0u && ng_hits > 0x00FFFFFFu) {
            const int bit = bfind32(ng_hits);
            const uint32_t imask = ng_hits & 0xFFu;
            ng_hits &= ~(1u << bit);
            if (ng_hits > 0x00FFFFFFu && sp < RT3_MAX_STACK) { stack_base[sp] = ng_base; stack_hits[sp] = ng_hits; sp++; }
            const uint32_t slot = ((uint32_t)(bit - 24) ^ (r.octinv4 & 0xFFu)) & 7u;
            const uint32_t rel = (uint32_t)popc32(imask & ~(0xFFFFFFFFu << slot));
            const WideNode* n = nodes + (ng_base + rel);
            const u32x4 n0 = ld_u32x4(n->w), n1 = ld_u32x4(n->w + 4), n2 = ld_u32x4(n->w + 8), n3 = ld_u32x4(n->w + 12), n4 = ld_u32x4(n->w + 16);
            if (stats) stats->nodes++;
            const uint32_t hm = intersect_wide_node(r, hit.t, n0, n1, n2, n3, n4);
            ng_base = n1.x; ng_hits = (hm & 0xFF000000u) | (n0.w >> 24);
            tg_base = n1.y; tg_hits = hm & 0x00FFFFFFu; tg_valid = n1.z;
        }
        bool found_any = false;
        for (int rep = 0; rep < 2; ++rep) {   /* up to two record steps per node step, as in k_trace_dyn */
        if (rep && !RT3_WARP_ANY(active && tg_hits != 0u && !found_any)) break;
        RT3_EMU_PHASE(6, active && tg_hits && !found_any);
        if (active && tg_hits && !found_any) {
            const int bit = bfind32(tg_hits);
            tg_hits &= ~(1u << bit);
            const Tri48* tp = tris + (tg_base + record_index(tg_valid, bit));
            const u32x4 a = ld_u32x4(tp), b = ld_u32x4((const uint32_t*)tp + 4), c = ld_u32x4((const uint32_t*)tp + 8);
            if (insts != nullptr && cur_inst == RT3_STACK_SENTINEL) {
                /* top-level record = instance: save what is pending here, dive into its bottom-level tree */
                if (ng_hits > 0x00FFFFFFu && sp < RT3_MAX_STACK) { stack_base[sp] = ng_base; stack_hits[sp] = ng_hits; sp++; }
                if (tg_hits && sp < RT3_MAX_STACK) { stack_base[sp] = tg_base; stack_hits[sp] = dense_records(tg_hits, tg_valid); sp++; }
                if (sp < RT3_MAX_STACK) { stack_base[sp] = RT3_STACK_SENTINEL; stack_hits[sp] = 0u; sp++; }
                cur_inst = a.w;
                r = prepare_ray(object_ray(insts[cur_inst], ray), 0x47000000u, lut_base);
                ng_base = insts[cur_inst].root; ng_hits = 0x80000000u; tg_hits = 0u;
            } else {
                if (stats) stats->tris++;
                float t, u, v;
                vec3 v0, e1, e2;
                uint32_t prim = a.w;
                if (insts != nullptr) { instance_triangle(insts[cur_inst], a, b, c, v0, e1, e2); prim += insts[cur_inst].first_prim; }
                else {
                    v0 = rt3m::v3(rt3m::as_float(a.x), rt3m::as_float(a.y), rt3m::as_float(a.z));
                    e1 = rt3m::v3(rt3m::as_float(b.x), rt3m::as_float(b.y), rt3m::as_float(b.z));
                    e2 = rt3m::v3(rt3m::as_float(c.x), rt3m::as_float(c.y), rt3m::as_float(c.z));
                }
                if (rt3m::ray_triangle(o, d, v0, e1, e2, t, u, v) && t > ray.tmin && t < ray.tmax) {
                    if (ANY) { hit.t = t; hit.u = u; hit.v = v; hit.prim = prim; found_any = true; }
                    else if (rt3m::closer_hit(t, prim, hit.t, hit.prim)) { hit.t = t; hit.u = u; hit.v = v; hit.prim = prim; }
                }
            }
        }
        }
        if (active && found_any) active = false;
        while (active && tg_hits == 0u && ng_hits <= 0x00FFFFFFu) {       /* nothing pending here: pop (more than once only across a sentinel) */
            if (sp == 0) { active = false; break; }
            sp--;
            const uint32_t eb = stack_base[sp], eh = stack_hits[sp];
            if (eb == RT3_STACK_SENTINEL && eh == 0u) { cur_inst = RT3_STACK_SENTINEL; r = prepare_ray(ray, 0x47000000u, lut_base); continue; }
            if (eh > 0x00FFFFFFu) { ng_base = eb; ng_hits = eh; } else { tg_base = eb; tg_hits = eh; tg_valid = 0x00FFFFFFu; ng_hits = 0u; }
        }
    }
#undef RT3_WARP_ANY
    return hit.prim != RT3_MISS;
}

}  /* namespace rt3b */
#endif
