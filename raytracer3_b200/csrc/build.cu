/*
 * build.cu — GPU acceleration-structure build (replaces the driver's vkCmdBuildAccelerationStructuresKHR
 * behind RayTracingContext::create_acceleration_structure, /root/reference/src/renderer/vulkan/raytracing.rs:88-148).
 *
 * Stages (all on the device; the host only reads back 4-byte counters between collapse levels):
 *   k_flatten        scene buffers + instances -> world-space (v0,e1,e2,id) records, scene AABB
 *   k_morton         63-bit Morton code of each triangle's box centre
 *   radix sort       8 passes x 8 bits, stable: k_radix_hist -> k_radix_scan (one block per digit) -> k_radix_scatter
 *   k_gather_sorted  triangles into Morton order
 *   k_lbvh_hierarchy Karras 2012, one thread per internal node
 *   k_refit          bottom-up walk with one atomic flag per internal node: AABBs + subtree counts; unless RT3_BUILD_FAST_BUILD,
 *                    three passes with SAH tree rotations (k_parents re-links in between) and, in the last one, the dynamic
 *                    programme that picks the children of the wide nodes (refit_visit / dp_node in bvh_logic.h)
 *   k_collapse       BVH2 -> 8-wide quantised nodes, one thread per wide node, level by level
 * Temporaries of one build live in a single arena from the stream-ordered pool.  One level (instances flattened to world space)
 * or two (RT3_BUILD_TWO_LEVEL: one tree per distinct mesh + a tree over the instances).
 */
#include "common.cuh"

using namespace rt3b;

namespace {

__device__ __forceinline__ int float_to_ordered(float f) { int i = __float_as_int(f); return i >= 0 ? i : i ^ 0x7FFFFFFF; }
__host__ __device__ __forceinline__ float ordered_to_float(int i) {
    int j = i >= 0 ? i : i ^ 0x7FFFFFFF;
#ifdef __CUDA_ARCH__
    return __int_as_float(j);
#else
    float f; memcpy(&f, &j, 4); return f;
#endif
}

__global__ void k_init_bounds(int* b) {
    if (threadIdx.x < 3) b[threadIdx.x] = 0x7F7FFFFF;            /* +FLT_MAX ordered */
    else if (threadIdx.x < 6) b[threadIdx.x] = (int)0x80800000;  /* ordered(-FLT_MAX) = 0xFF7FFFFF ^ 0x7FFFFFFF */
}

__global__ void __launch_bounds__(256) k_flatten(rt3m::SceneView s, uint32_t n, Tri48* tris, int* bounds) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    float lo[3] = {3.4e38f, 3.4e38f, 3.4e38f}, hi[3] = {-3.4e38f, -3.4e38f, -3.4e38f};
    if (i < n) {
        uint32_t seg = rt3m::find_segment(s, i);
        rt3m::vec3 a, b, c;
        rt3m::world_triangle(s, seg, i - s.seg_first_prim[seg], a, b, c);
        Tri48 t;
        t.v0[0] = a.x; t.v0[1] = a.y; t.v0[2] = a.z; t.prim = i;
        t.e1[0] = b.x - a.x; t.e1[1] = b.y - a.y; t.e1[2] = b.z - a.z; t.pad0 = 0;
        t.e2[0] = c.x - a.x; t.e2[1] = c.y - a.y; t.e2[2] = c.z - a.z; t.pad1 = 0;
        float4* o = (float4*)(tris + i);
        o[0] = make_float4(t.v0[0], t.v0[1], t.v0[2], __uint_as_float(i));
        o[1] = make_float4(t.e1[0], t.e1[1], t.e1[2], 0.f);
        o[2] = make_float4(t.e2[0], t.e2[1], t.e2[2], 0.f);
        lo[0] = fminf(a.x, fminf(b.x, c.x)); hi[0] = fmaxf(a.x, fmaxf(b.x, c.x));
        lo[1] = fminf(a.y, fminf(b.y, c.y)); hi[1] = fmaxf(a.y, fmaxf(b.y, c.y));
        lo[2] = fminf(a.z, fminf(b.z, c.z)); hi[2] = fmaxf(a.z, fmaxf(b.z, c.z));
    }
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        for (int off = 16; off; off >>= 1) {
            lo[k] = fminf(lo[k], __shfl_xor_sync(0xFFFFFFFFu, lo[k], off));
            hi[k] = fmaxf(hi[k], __shfl_xor_sync(0xFFFFFFFFu, hi[k], off));
        }
    }
    if ((threadIdx.x & 31) == 0) {
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            atomicMin(bounds + k, float_to_ordered(lo[k]));
            atomicMax(bounds + 3 + k, float_to_ordered(hi[k]));
        }
    }
}

/* Per-triangle shading records for rt3m::hit_info: (global vertex indices, segment) by global triangle id. */
__global__ void __launch_bounds__(256) k_shade_records(rt3m::SceneView s, uint32_t n, uint4* rec) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint32_t seg = rt3m::find_segment(s, i);
    const rt3_geometry_info& g = s.geometries[s.seg_geometry[seg]];
    const uint32_t io = g.index_offset + 3u * (i - s.seg_first_prim[seg]);
    rec[i] = make_uint4(g.vertex_offset + s.indices[io], g.vertex_offset + s.indices[io + 1], g.vertex_offset + s.indices[io + 2], seg);
}

/* 64-byte shading records (SceneView::shade64): normals x 3, texture coordinates x 3, segment — copies of what hit_info would gather. */
__global__ void __launch_bounds__(256) k_shade_records64(rt3m::SceneView s, uint32_t n, const uint4* rec, float4* out) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint4 r = rec[i];
    const rt3_vertex a = s.vertices[r.x], b = s.vertices[r.y], c = s.vertices[r.z];
    out[4 * (size_t)i + 0] = make_float4(a.n[0], a.n[1], a.n[2], b.n[0]);
    out[4 * (size_t)i + 1] = make_float4(b.n[1], b.n[2], c.n[0], c.n[1]);
    out[4 * (size_t)i + 2] = make_float4(c.n[2], a.t[0], a.t[1], b.t[0]);
    out[4 * (size_t)i + 3] = make_float4(b.t[1], c.t[0], c.t[1], __uint_as_float(r.w));
}

/* Bottom level of the two-level build: the triangles of one mesh (segments [seg0, ...) of its first instance, global ids
 * [prim0, prim0 + n)) in OBJECT space — geometry transform applied, instance transform not — stored as VERTICES with the
 * mesh-local id (REC_TRI_VERTS).  The vertex arithmetic is that of rt3m::world_triangle up to the instance transform. */
__global__ void __launch_bounds__(256) k_flatten_object(rt3m::SceneView s, uint32_t prim0, uint32_t n, Tri48* tris, int* bounds) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    float lo[3] = {3.4e38f, 3.4e38f, 3.4e38f}, hi[3] = {-3.4e38f, -3.4e38f, -3.4e38f};
    if (i < n) {
        const uint32_t prim = prim0 + i;
        const uint32_t seg = rt3m::find_segment(s, prim);
        const rt3_geometry_info& g = s.geometries[s.seg_geometry[seg]];
        const uint32_t io = g.index_offset + 3u * (prim - s.seg_first_prim[seg]);
        const uint32_t i0 = g.vertex_offset + s.indices[io], i1 = g.vertex_offset + s.indices[io + 1], i2 = g.vertex_offset + s.indices[io + 2];
        rt3m::vec3 a = rt3m::v3(s.vertices[i0].p), b = rt3m::v3(s.vertices[i1].p), c = rt3m::v3(s.vertices[i2].p);
        if (!rt3m::xform_is_identity(g.xform)) { a = rt3m::xform_point(g.xform, a); b = rt3m::xform_point(g.xform, b); c = rt3m::xform_point(g.xform, c); }
        float4* o = (float4*)(tris + i);
        o[0] = make_float4(a.x, a.y, a.z, __uint_as_float(i));
        o[1] = make_float4(b.x, b.y, b.z, 0.f);
        o[2] = make_float4(c.x, c.y, c.z, 0.f);
        lo[0] = fminf(a.x, fminf(b.x, c.x)); hi[0] = fmaxf(a.x, fmaxf(b.x, c.x));
        lo[1] = fminf(a.y, fminf(b.y, c.y)); hi[1] = fmaxf(a.y, fmaxf(b.y, c.y));
        lo[2] = fminf(a.z, fminf(b.z, c.z)); hi[2] = fmaxf(a.z, fmaxf(b.z, c.z));
    }
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        for (int off = 16; off; off >>= 1) {
            lo[k] = fminf(lo[k], __shfl_xor_sync(0xFFFFFFFFu, lo[k], off));
            hi[k] = fmaxf(hi[k], __shfl_xor_sync(0xFFFFFFFFu, hi[k], off));
        }
    }
    if ((threadIdx.x & 31) == 0) {
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            atomicMin(bounds + k, float_to_ordered(lo[k]));
            atomicMax(bounds + 3 + k, float_to_ordered(hi[k]));
        }
    }
}

__global__ void __launch_bounds__(256) k_morton(const Tri48* tris, uint32_t n, int kind, float3 lo, float3 inv_ext, uint64_t* keys, uint32_t* vals) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float l[3], h[3];
    tri_bounds(tris[i], 0.f, l, h, kind);
    float cx = (0.5f * (l[0] + h[0]) - lo.x) * inv_ext.x, cy = (0.5f * (l[1] + h[1]) - lo.y) * inv_ext.y, cz = (0.5f * (l[2] + h[2]) - lo.z) * inv_ext.z;
    keys[i] = morton63(cx, cy, cz);
    vals[i] = i;
}

/* ---- radix sort: 8-bit digits, tile of RADIX_TILE keys per block, stable --------------------------- */
constexpr int RADIX_THREADS = 256;
constexpr int RADIX_ROUNDS = 16;
constexpr int RADIX_TILE = RADIX_THREADS * RADIX_ROUNDS;

__global__ void __launch_bounds__(RADIX_THREADS) k_radix_hist(const uint64_t* keys, uint32_t n, int shift, uint32_t* hist, uint32_t nblocks,
                                                              uint32_t* totals) {
    __shared__ uint32_t h[256];
    h[threadIdx.x] = 0;
    __syncthreads();
    uint32_t base = blockIdx.x * RADIX_TILE;
    for (int r = 0; r < RADIX_ROUNDS; ++r) {
        uint32_t i = base + r * RADIX_THREADS + threadIdx.x;
        if (i < n) atomicAdd(&h[(uint32_t)(keys[i] >> shift) & 255u], 1u);
    }
    __syncthreads();
    hist[threadIdx.x * nblocks + blockIdx.x] = h[threadIdx.x];
    if (h[threadIdx.x]) atomicAdd(totals + threadIdx.x, h[threadIdx.x]);
}

/* hist[digit][block] -> global offsets.  One block per digit: its base is the sum of the totals of the smaller digits, its row of
 * `nblocks` per-block counts is scanned 1024 at a time.  (A single block scanning all 256 * nblocks counters took ~1 us per 1024
 * counters: 1 ms per pass at 16.8 M keys, 8 of the 8.7 ms of the whole sort.) */
__global__ void __launch_bounds__(1024) k_radix_scan(uint32_t* hist, uint32_t nblocks, const uint32_t* totals) {
    __shared__ uint32_t warp_sums[32];
    __shared__ uint32_t carry;
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5, digit = blockIdx.x;
    {   /* digit base */
        uint32_t v = (threadIdx.x < digit) ? totals[threadIdx.x] : 0u;    /* digit < 256 <= blockDim */
        for (int off = 16; off; off >>= 1) v += __shfl_xor_sync(0xFFFFFFFFu, v, off);
        if (lane == 0) warp_sums[warp] = v;
        __syncthreads();
        if (threadIdx.x == 0) { uint32_t t = 0; for (int w = 0; w < 8; ++w) t += warp_sums[w]; carry = t; }
        __syncthreads();
    }
    uint32_t* data = hist + (size_t)digit * nblocks;
    for (uint32_t base = 0; base < nblocks; base += 1024) {
        uint32_t i = base + threadIdx.x;
        uint32_t v = i < nblocks ? data[i] : 0u;
        uint32_t x = v;
        for (int off = 1; off < 32; off <<= 1) { uint32_t y = __shfl_up_sync(0xFFFFFFFFu, x, off); if (lane >= off) x += y; }
        __syncthreads();                                                   /* warp_sums of the previous round have been read */
        if (lane == 31) warp_sums[warp] = x;
        __syncthreads();
        if (warp == 0) {
            uint32_t w = warp_sums[lane];
            for (int off = 1; off < 32; off <<= 1) { uint32_t y = __shfl_up_sync(0xFFFFFFFFu, w, off); if (lane >= off) w += y; }
            warp_sums[lane] = w;
        }
        __syncthreads();
        uint32_t excl = x - v + (warp ? warp_sums[warp - 1] : 0u) + carry;
        if (i < nblocks) data[i] = excl;
        __syncthreads();
        if (threadIdx.x == 1023) carry = excl + v;
        __syncthreads();
    }
}

__global__ void __launch_bounds__(RADIX_THREADS) k_radix_scatter(const uint64_t* keys_in, const uint32_t* vals_in, uint64_t* keys_out, uint32_t* vals_out,
                                                                 uint32_t n, int shift, const uint32_t* hist, uint32_t nblocks) {
    __shared__ uint32_t base[256];
    __shared__ uint32_t wcnt[RADIX_THREADS / 32][256];
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    base[threadIdx.x] = hist[threadIdx.x * nblocks + blockIdx.x];
    const uint32_t tile = blockIdx.x * RADIX_TILE;
    for (int r = 0; r < RADIX_ROUNDS; ++r) {
        for (int w = 0; w < RADIX_THREADS / 32; ++w) wcnt[w][threadIdx.x] = 0;
        __syncthreads();
        uint32_t i = tile + r * RADIX_THREADS + threadIdx.x;
        bool valid = i < n;
        uint64_t key = valid ? keys_in[i] : 0ull;
        uint32_t val = valid ? vals_in[i] : 0u;
        uint32_t d = valid ? ((uint32_t)(key >> shift) & 255u) : (256u + lane);
        uint32_t peers = __match_any_sync(0xFFFFFFFFu, d);
        uint32_t rank = __popc(peers & ((1u << lane) - 1u));
        if (valid && rank == 0) wcnt[warp][d] = __popc(peers);
        __syncthreads();
        {   /* thread = digit: exclusive prefix over warps, advance the running base */
            uint32_t run = base[threadIdx.x];
            for (int w = 0; w < RADIX_THREADS / 32; ++w) { uint32_t t = wcnt[w][threadIdx.x]; wcnt[w][threadIdx.x] = run; run += t; }
            base[threadIdx.x] = run;
        }
        __syncthreads();
        if (valid) {
            uint32_t pos = wcnt[warp][d] + rank;
            keys_out[pos] = key; vals_out[pos] = val;
        }
        __syncthreads();
    }
}

__global__ void __launch_bounds__(256) k_gather_sorted(const Tri48* tris, const uint32_t* order, uint32_t n, Tri48* out) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float4* s = (const float4*)(tris + order[i]);
    float4* d = (float4*)(out + i);
    d[0] = s[0]; d[1] = s[1]; d[2] = s[2];
}

__global__ void __launch_bounds__(256) k_lbvh_hierarchy(const uint64_t* keys, int n, Bvh2Node* nodes, uint32_t* parent) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n - 1) lbvh_build_node(keys, n, i, nodes, parent);
}

/* How refit_visit (bvh_logic.h) reads what OTHER threads of the same launch wrote: past the non-coherent L1 (ld.global.cg). */
struct DevMem {
    Bvh2Node* nodes; uint32_t* counts; DpNode* dp; uint32_t n;
    __device__ uint32_t slot(uint32_t ref) const { return (ref & LEAF_FLAG) ? n - 1 + (ref & ~LEAF_FLAG) : ref; }
    __device__ Bvh2Node node(uint32_t i) const {
        const float4 a = __ldcg((const float4*)(nodes + i)), b = __ldcg((const float4*)(nodes + i) + 1);
        Bvh2Node m;
        m.lo[0] = a.x; m.lo[1] = a.y; m.lo[2] = a.z; m.left = __float_as_uint(a.w);
        m.hi[0] = b.x; m.hi[1] = b.y; m.hi[2] = b.z; m.right = __float_as_uint(b.w);
        return m;
    }
    __device__ void store_node(uint32_t i, const Bvh2Node& m) {
        float4* p = (float4*)(nodes + i);
        __stcg(p, make_float4(m.lo[0], m.lo[1], m.lo[2], __uint_as_float(m.left)));
        __stcg(p + 1, make_float4(m.hi[0], m.hi[1], m.hi[2], __uint_as_float(m.right)));
    }
    __device__ uint32_t count(uint32_t ref) const { return (ref & LEAF_FLAG) ? 1u : __ldcg(counts + ref); }
    __device__ void store_count(uint32_t i, uint32_t c) { __stcg(counts + i, c); }
    __device__ void dp_costs(uint32_t ref, float out[7]) const { for (int i = 0; i < 7; ++i) out[i] = __ldcg(&dp[ref].cost[i]); }
    __device__ void store_dp(uint32_t i, const DpNode& d) {
        uint32_t* p = (uint32_t*)(dp + i); const uint32_t* q = (const uint32_t*)&d;
        for (int k = 0; k < 10; ++k) __stcg(p + k, q[k]);
    }
};

/* leaf boxes (first pass only), then the bottom-up walk: one atomic flag per node, the second arrival runs refit_visit */
template <bool ROTATE, bool WITH_DP>
__global__ void __launch_bounds__(256) k_refit(const Tri48* tris_sorted, uint32_t n, int kind, float pad, Bvh2Node* nodes, const uint32_t* parent,
                                               uint32_t* flags, uint32_t* counts, DpNode* dp, int first_pass) {
    uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n) return;
    DevMem mem{nodes, counts, dp, n};
    if (first_pass) {
        Bvh2Node leaf;
        tri_bounds(tris_sorted[k], pad, leaf.lo, leaf.hi, kind);
        leaf.left = leaf.right = 0;
        mem.store_node(n - 1 + k, leaf);
    }
    uint32_t cur = parent[n - 1 + k];
    while (cur != 0xFFFFFFFFu) {
        __threadfence();
        if (atomicAdd(flags + cur, 1u) == 0u) return;   /* first arrival: the sibling subtree is not ready yet */
        __threadfence();
        refit_visit(mem, n, cur, ROTATE, WITH_DP);
        cur = parent[cur];
    }
}

/* ---- treelet restructuring, one WARP per treelet --------------------------------------------------------------------------------------
 * The bottom-up walk is serialised by the depth of the tree: near the root a single thread owns the only node that is ready, and with one
 * thread per treelet (round 2's first version) it ground through the 17 k-instruction dynamic programme of bvh_logic.h alone — 8.9 of the
 * 10.3 ms of a 263 k-triangle build, whatever number of treelets a leaf-count threshold removed (measured: profiles/r2q_*).  Here the lanes
 * of a warp walk up in lock step; a lane that owns a small subtree visits it alone (all such lanes at once), and the treelets of the other
 * owners are taken one after the other by the WHOLE warp:
 *   formation and application (treelet_visit_t) run on all 32 lanes with the same arguments — same loads, broadcast — and only the owner
 *   lane stores (CoopMem), so nothing has to be handed from lane to lane;
 *   the 127 subset areas and the programme over the subsets (CoopTreeletDp) are spread over the lanes through 1.25 KB of shared memory per
 *   warp: subsets of 2 .. 5 leaves one subset per lane (the serial loop, so ties fall as in the definition), subsets of 6 and 7 leaves one
 *   PARTITION per lane followed by a warp reduction on (cost, position in the serial enumeration) — the serial solver's first-minimum rule.
 * The result is the tree of the serial definition, bit for bit (GPU test against the CPU emulation: node count and visit statistics). */
struct CoopMem {
    DevMem m; bool writer;
    __device__ uint32_t slot(uint32_t ref) const { return m.slot(ref); }
    __device__ Bvh2Node node(uint32_t i) const { return m.node(i); }
    __device__ uint32_t count(uint32_t ref) const { return m.count(ref); }
    __device__ void dp_costs(uint32_t ref, float out[7]) const { m.dp_costs(ref, out); }
    __device__ void store_node(uint32_t i, const Bvh2Node& v) { if (writer) m.store_node(i, v); }
    __device__ void store_count(uint32_t i, uint32_t c) { if (writer) m.store_count(i, c); }
    __device__ void store_dp(uint32_t i, const DpNode& d) { if (writer) m.store_dp(i, d); }
};

struct CoopTreeletDp {
    float* area; float* copt; uint16_t* part;          /* this warp's tables in shared memory, 1 << RT3_TREELET_LEAVES entries each */
    uint32_t lane;
    /* the i-th proper submask of `rest` in the serial enumeration (rest-1)&rest, ((rest-1)&rest - 1)&rest, ..., 0: decreasing numeric
     * order, i.e. the submask of rank 2^m - 2 - i, its bits deposited at the places of rest's bits */
    static __device__ __forceinline__ uint32_t submask_at(uint32_t rest, uint32_t i) {
        uint32_t r = (1u << __popc(rest)) - 2u - i, sub = 0u, m = rest;
        while (m) { const uint32_t low = m & (0u - m); if (r & 1u) sub |= low; r >>= 1; m ^= low; }
        return sub;
    }
    __device__ void solve(const Bvh2Node* leaf_box, int nl) {
        const uint32_t full = (1u << nl) - 1u;
        for (uint32_t sset = lane; sset <= full; sset += 32u) {                      /* areas: min / max are exact, any order of the unions */
            if (!sset) continue;
            float lo[3] = {3.4e38f, 3.4e38f, 3.4e38f}, hi[3] = {-3.4e38f, -3.4e38f, -3.4e38f};
#pragma unroll
            for (int k = 0; k < RT3_TREELET_LEAVES; ++k)
                if (sset & (1u << k))
                    for (int q = 0; q < 3; ++q) { lo[q] = rt3m::min_(lo[q], leaf_box[k].lo[q]); hi[q] = rt3m::max_(hi[q], leaf_box[k].hi[q]); }
            area[sset] = box_area(lo, hi);
            copt[sset] = 0.0f; part[sset] = 0;
        }
        __syncwarp();
        for (int k = 2; k <= nl; ++k) {                                              /* subsets of k leaves need only smaller subsets: one round per size */
            if (k <= 5) {
                for (uint32_t sset = lane; sset <= full; sset += 32u) {
                    if (__popc(sset) != k) continue;
                    float bc = 3.4e38f; uint32_t bp = 0;
                    const uint32_t low = sset & (0u - sset), rest = sset ^ low;
                    for (uint32_t sub = (rest - 1u) & rest;; sub = (sub - 1u) & rest) {
                        const uint32_t pp = low | sub;
                        const float c = copt[pp] + copt[sset ^ pp];
                        if (c < bc) { bc = c; bp = pp; }
                        if (sub == 0u) break;
                    }
                    copt[sset] = (sset == full ? 0.0f : area[sset]) + bc;
                    part[sset] = (uint16_t)bp;
                }
            } else {
                for (uint32_t sset = 3u; sset <= full; ++sset) {                     /* warp-uniform scan: at most 7 + 1 subsets of 6 / 7 leaves */
                    if (__popc(sset) != k) continue;
                    const uint32_t low = sset & (0u - sset), rest = sset ^ low;
                    const uint32_t count = (1u << (k - 1)) - 1u;                     /* proper submasks of rest, the empty one included */
                    float bc = 3.4e38f; uint32_t bi = 0xFFFFFFFFu, bp = 0;
                    for (uint32_t i = lane; i < count; i += 32u) {
                        const uint32_t pp = low | submask_at(rest, i);
                        const float c = copt[pp] + copt[sset ^ pp];
                        if (c < bc) { bc = c; bi = i; bp = pp; }                     /* i increases: the first minimum of this lane's share */
                    }
#pragma unroll
                    for (int off = 16; off; off >>= 1) {
                        const float c2 = __shfl_xor_sync(0xFFFFFFFFu, bc, off);
                        const uint32_t i2 = __shfl_xor_sync(0xFFFFFFFFu, bi, off), p2 = __shfl_xor_sync(0xFFFFFFFFu, bp, off);
                        if (c2 < bc || (c2 == bc && i2 < bi)) { bc = c2; bi = i2; bp = p2; }
                    }
                    if (lane == 0u) {
                        copt[sset] = (sset == full ? 0.0f : area[sset]) + bc;
                        part[sset] = (uint16_t)(bi == 0xFFFFFFFFu ? 0u : bp);        /* no cost below 3.4e38: the serial loop leaves 0 */
                    }
                }
            }
            __syncwarp();
        }
    }
    __device__ float cost_of(uint32_t sset) const { return copt[sset]; }
    __device__ uint32_t part_of(uint32_t sset) const { return part[sset]; }
};

constexpr int TREELET_THREADS = 128;
template <bool WITH_DP>
__global__ void __launch_bounds__(TREELET_THREADS) k_refit_treelet(const Tri48* tris_sorted, uint32_t n, int kind, float pad, Bvh2Node* nodes, const uint32_t* parent,
                                                                   uint32_t* flags, uint32_t* counts, DpNode* dp, int first_pass) {
    __shared__ float s_area[TREELET_THREADS / 32][1 << RT3_TREELET_LEAVES], s_copt[TREELET_THREADS / 32][1 << RT3_TREELET_LEAVES];
    __shared__ uint16_t s_part[TREELET_THREADS / 32][1 << RT3_TREELET_LEAVES];
    const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x, lane = threadIdx.x & 31u, wid = threadIdx.x >> 5;
    const uint32_t NONE = 0xFFFFFFFFu;
    DevMem mem{nodes, counts, dp, n};
    uint32_t cur = NONE;                                       /* the node this lane is arriving at; lanes past the last leaf only help */
    if (k < n) {
        if (first_pass) {
            Bvh2Node leaf;
            tri_bounds(tris_sorted[k], pad, leaf.lo, leaf.hi, kind);
            leaf.left = leaf.right = 0;
            mem.store_node(n - 1 + k, leaf);
        }
        cur = parent[n - 1 + k];
    }
    CoopTreeletDp cdp{s_area[wid], s_copt[wid], s_part[wid], lane};
    for (;;) {
        if (cur != NONE) {                                     /* one atomic flag per node: the second arrival owns two finished subtrees */
            __threadfence();
            if (atomicAdd(flags + cur, 1u) == 0u) cur = NONE;
            else __threadfence();
        }
        __syncwarp();
        if (__ballot_sync(0xFFFFFFFFu, cur != NONE) == 0u) break;
        bool small = false;
        if (cur != NONE && RT3_TREELET_MIN_LEAVES > 0) {
            const Bvh2Node p0 = mem.node(cur);
            small = mem.count(p0.left) + mem.count(p0.right) < (uint32_t)RT3_TREELET_MIN_LEAVES;
            if (small) refit_visit(mem, n, cur, RT3_TREELET_SMALL_ROTATE != 0, WITH_DP);    /* these lanes all at once, each its own node */
        }
        __syncwarp();
        uint32_t big = __ballot_sync(0xFFFFFFFFu, cur != NONE && !small);
        while (big) {                                          /* the others one after the other, by the whole warp */
            const int src = __ffs(big) - 1;
            big &= big - 1u;
            const uint32_t c = __shfl_sync(0xFFFFFFFFu, cur, src);
            CoopMem cm{mem, lane == (uint32_t)src};
            treelet_visit_t(cm, n, c, WITH_DP, cdp);
            __syncwarp();
        }
        if (cur != NONE) cur = parent[cur];                    /* the root's parent is 0xFFFFFFFF */
    }
}

/* rotations moved subtrees: parent links for the next bottom-up pass */
__global__ void __launch_bounds__(256) k_parents(const Bvh2Node* nodes, uint32_t n, uint32_t* parent) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i + 1 >= n) return;
    const uint32_t l = nodes[i].left, r = nodes[i].right;
    parent[(l & LEAF_FLAG) ? n - 1 + (l & ~LEAF_FLAG) : l] = i;
    parent[(r & LEAF_FLAG) ? n - 1 + (r & ~LEAF_FLAG) : r] = i;
}

__global__ void __launch_bounds__(128) k_collapse(CollapseCtx c, const uint32_t* items, uint32_t count) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < count) collapse_one(c, items[2 * i], items[2 * i + 1]);
}

__global__ void k_empty_root(WideNode* nodes) {
    if (threadIdx.x < 20) nodes[0].w[threadIdx.x] = 0;
}

/* Build temporaries come from the stream-ordered pool (cudaMallocAsync): the context raises the pool's release threshold, so
 * rebuilds reuse the memory instead of paying cudaMalloc/cudaFree (each a device-wide synchronisation) per buffer. */
struct PoolBuf {
    void* ptr = nullptr;
    cudaStream_t st = 0;
    void release() { if (ptr) cudaFreeAsync(ptr, st); ptr = nullptr; }
};
template <class T>
T* palloc(rt3_context* c, PoolBuf& b, size_t count, cudaStream_t st) {
    b.release();
    b.st = st;
    RT3_CUDA(c, cudaMallocAsync(&b.ptr, (count ? count : 1) * sizeof(T), st));
    return (T*)b.ptr;
}

struct StageTimer {
    rt3_context* c; cudaStream_t st; cudaEvent_t a, b;
    StageTimer(rt3_context* c_, cudaStream_t s) : c(c_), st(s) { a = c->ev[0]; b = c->ev[1]; }
    void begin() { cudaEventRecord(a, st); }
    double end() { cudaEventRecord(b, st); cudaEventSynchronize(b); float ms = 0; cudaEventElapsedTime(&ms, a, b); return ms; }
};

}  // namespace

/* One wide BVH over `n` 48-byte leaf records (Morton -> sort -> LBVH -> refit -> collapse).  Nodes go into `wide`: the root at
 * root_index, the rest from first_free_node on; leaf records are copied in leaf order into out_recs from first_free_rec on.
 * `recs` (pool memory, n records) is consumed.  Stage times are ADDED to bs. */
struct WideBuild {
    PoolBuf* recs; uint32_t n; int kind; float pad; float lo[3], hi[3];
    WideNode* wide; uint32_t root_index, first_free_node;
    Tri48* out_recs; uint32_t first_free_rec;
    bool optimal_collapse = false;   /* choose the children of every wide node by the cost dynamic programme instead of greedily */
    int treelet = 0;                 /* with optimal_collapse: treelet restructuring instead of rotations; 1 = up to RT3_TREELET_AUTO_MAX records, 2 = always */
    uint32_t nodes_end = 0, recs_end = 0, levels = 0; float sah_nodes = 0.f, sah_recs = 0.f;   /* results */
};

#include <chrono>
static void dbg_mark(const char* what) {
    static const bool on = getenv("RT3_BUILD_DEBUG") != nullptr;
    static auto t0 = std::chrono::steady_clock::now();
    if (!on) return;
    cudaDeviceSynchronize();
    auto t1 = std::chrono::steady_clock::now();
    fprintf(stderr, "[build] %-24s +%.3f ms\n", what, std::chrono::duration<double, std::milli>(t1 - t0).count());
    t0 = std::chrono::steady_clock::now();
}
static void build_wide(rt3_context* c, cudaStream_t st, WideBuild& io, rt3_build_stats& bs, int* d_misc) {
    const uint32_t n = io.n;
    dbg_mark("enter build_wide");
    StageTimer tm(c, st);
    /* one arena for all temporaries of this build: a single pool allocation (the same size on a rebuild, so the pool hands the
     * same block back at once; a dozen separate GB-sized cudaMallocAsync calls cost ~6 ms of a 31 ms build at 16.8 M triangles) */
    PoolBuf arena;
    auto cleanup = [&]() { io.recs->release(); arena.release(); };
    try {
        const Tri48* tris = (const Tri48*)io.recs->ptr;
        const float* lo = io.lo; const float* hi = io.hi;
        const uint32_t nblocks = div_up(n, RADIX_TILE);
        size_t off = 0;
        auto carve = [&](size_t bytes) { size_t o = off; off += (bytes + 255) & ~(size_t)255; return o; };
        const size_t o_keys0 = carve(8 * (size_t)n), o_keys1 = carve(8 * (size_t)n), o_vals0 = carve(4 * (size_t)n), o_vals1 = carve(4 * (size_t)n);
        const size_t o_hist = carve(4 * (256 * (size_t)nblocks + 8 * 256)), o_sorted = carve(sizeof(Tri48) * (size_t)n);
        const size_t o_nodes2 = carve(sizeof(Bvh2Node) * 2 * (size_t)n), o_parent = carve(4 * 2 * (size_t)n), o_flags = carve(4 * (size_t)n);
        const size_t o_counts = carve(4 * (size_t)n), o_items0 = carve(4 * (2 * (size_t)n + 2)), o_items1 = carve(4 * (2 * (size_t)n + 2));
        const size_t o_dp = io.optimal_collapse ? carve(sizeof(DpNode) * (size_t)n) : 0;
        uint8_t* base = palloc<uint8_t>(c, arena, off, st);
        /* morton */
        uint64_t* keys0 = (uint64_t*)(base + o_keys0);
        uint64_t* keys1 = (uint64_t*)(base + o_keys1);
        uint32_t* vals0 = (uint32_t*)(base + o_vals0);
        uint32_t* vals1 = (uint32_t*)(base + o_vals1);
        tm.begin();
        float3 flo = make_float3(lo[0], lo[1], lo[2]);
        float3 inv = make_float3(hi[0] > lo[0] ? 1.0f / (hi[0] - lo[0]) : 0.f, hi[1] > lo[1] ? 1.0f / (hi[1] - lo[1]) : 0.f,
                                 hi[2] > lo[2] ? 1.0f / (hi[2] - lo[2]) : 0.f);
        RT3_LAUNCH(c, k_morton, div_up(n, 256), 256, 0, st, tris, n, io.kind, flo, inv, keys0, vals0);
        bs.ms_morton += tm.end();
        dbg_mark("morton");
        /* sort */
        uint32_t* hist = (uint32_t*)(base + o_hist);
        uint32_t* totals = hist + 256 * (size_t)nblocks;                  /* per pass: 256 digit totals */
        tm.begin();
        RT3_CUDA(c, cudaMemsetAsync(totals, 0, 8 * 256 * sizeof(uint32_t), st));
        for (int pass = 0; pass < 8; ++pass) {
            const uint64_t* ki = (pass & 1) ? keys1 : keys0; uint64_t* ko = (pass & 1) ? keys0 : keys1;
            const uint32_t* vi = (pass & 1) ? vals1 : vals0; uint32_t* vo = (pass & 1) ? vals0 : vals1;
            RT3_LAUNCH(c, k_radix_hist, nblocks, RADIX_THREADS, 0, st, ki, n, pass * 8, hist, nblocks, totals + 256 * pass);
            RT3_LAUNCH(c, k_radix_scan, 256, 1024, 0, st, hist, nblocks, totals + 256 * pass);
            RT3_LAUNCH(c, k_radix_scatter, nblocks, RADIX_THREADS, 0, st, ki, vi, ko, vo, n, pass * 8, hist, nblocks);
        }
        Tri48* sorted = (Tri48*)(base + o_sorted);
        RT3_LAUNCH(c, k_gather_sorted, div_up(n, 256), 256, 0, st, tris, vals0, n, sorted);
        bs.ms_sort += tm.end();
        dbg_mark("sort");
        io.recs->release();
        dbg_mark("release recs");
        /* hierarchy */
        Bvh2Node* nodes2 = (Bvh2Node*)(base + o_nodes2);
        uint32_t* parent = (uint32_t*)(base + o_parent);
        uint32_t* nflags = (uint32_t*)(base + o_flags);
        uint32_t* counts = (uint32_t*)(base + o_counts);
        tm.begin();
        RT3_CUDA(c, cudaMemsetAsync(parent, 0xFF, sizeof(uint32_t) * 2 * (size_t)n, st));
        RT3_CUDA(c, cudaMemsetAsync(nflags, 0, sizeof(uint32_t) * n, st));
        RT3_CUDA(c, cudaMemsetAsync(counts, 0, sizeof(uint32_t) * n, st));
        if (n > 1) RT3_LAUNCH(c, k_lbvh_hierarchy, div_up(n - 1, 256), 256, 0, st, keys0, (int)n, nodes2, parent);
        bs.ms_hierarchy += tm.end();
        dbg_mark("hierarchy");
        tm.begin();
        DpNode* dp = io.optimal_collapse ? (DpNode*)(base + o_dp) : nullptr;
        /* the treelet pass costs ~34 us per thousand triangles (8.9 ms at 263 k, 364 ms at 16.8 M; tree rotations: 0.5 / 19.7 ms): by default only
         * trees of up to RT3_TREELET_AUTO_MAX records get it, larger ones the rotations; RT3_BUILD_TREELET asks for it regardless of size */
        const bool treelet = io.optimal_collapse && (io.treelet == 2 || (io.treelet == 1 && n <= RT3_TREELET_AUTO_MAX));
        const int passes = treelet ? RT3_TREELET_PASSES : (io.optimal_collapse ? RT3_REFIT_PASSES : 1);
        for (int pass = 0; pass < passes; ++pass) {
            if (pass) {
                RT3_CUDA(c, cudaMemsetAsync(nflags, 0, sizeof(uint32_t) * n, st));
                if (n > 1) RT3_LAUNCH(c, k_parents, div_up(n - 1, 256), 256, 0, st, nodes2, n, parent);
            }
            if (treelet) {
                if (pass < passes - 1) RT3_LAUNCH(c, (k_refit_treelet<false>), div_up(n, TREELET_THREADS), TREELET_THREADS, 0, st, sorted, n, io.kind, io.pad, nodes2, parent, nflags, counts, dp, pass == 0);
                else RT3_LAUNCH(c, (k_refit_treelet<true>), div_up(n, TREELET_THREADS), TREELET_THREADS, 0, st, sorted, n, io.kind, io.pad, nodes2, parent, nflags, counts, dp, pass == 0);
            } else if (!io.optimal_collapse) RT3_LAUNCH(c, (k_refit<false, false>), div_up(n, 256), 256, 0, st, sorted, n, io.kind, io.pad, nodes2, parent, nflags, counts, dp, 1);
            else if (pass < passes - 1) RT3_LAUNCH(c, (k_refit<true, false>), div_up(n, 256), 256, 0, st, sorted, n, io.kind, io.pad, nodes2, parent, nflags, counts, dp, pass == 0);
            else RT3_LAUNCH(c, (k_refit<true, true>), div_up(n, 256), 256, 0, st, sorted, n, io.kind, io.pad, nodes2, parent, nflags, counts, dp, pass == 0);
        }
        bs.ms_refit += tm.end();
        dbg_mark("refit");
        /* collapse */
        uint32_t* items0 = (uint32_t*)(base + o_items0);
        uint32_t* items1 = (uint32_t*)(base + o_items1);
        uint32_t* ctr = (uint32_t*)d_misc + 8;      /* [8] node ctr, [9] record ctr, [10] queue ctr, [12],[13] sah */
        tm.begin();
        uint32_t init_ctr[8] = {io.first_free_node, io.first_free_rec, 0u, 0u, 0u, 0u, 0u, 0u};
        RT3_CUDA(c, cudaMemcpyAsync(ctr, init_ctr, sizeof(init_ctr), cudaMemcpyHostToDevice, st));
        uint32_t first_item[2] = {n == 1 ? (LEAF_FLAG | 0u) : 0u, io.root_index};
        RT3_CUDA(c, cudaMemcpyAsync(items0, first_item, sizeof(first_item), cudaMemcpyHostToDevice, st));
        CollapseCtx cc;
        cc.nodes = nodes2; cc.counts = counts; cc.tris_sorted = sorted; cc.n_leaves = n; cc.out_nodes = io.wide; cc.out_tris = io.out_recs;
        cc.node_counter = ctr; cc.tri_counter = ctr + 1; cc.next_count = ctr + 2; cc.sah_accum = (float*)(ctr + 4); cc.dp = dp;
        uint32_t count = 1, levels = 0;
        uint32_t* cur = items0; uint32_t* nxt = items1;
        while (count) {
            RT3_CUDA(c, cudaMemsetAsync(ctr + 2, 0, 4, st));
            cc.next_items = nxt;
            RT3_LAUNCH(c, k_collapse, div_up(count, 128), 128, 0, st, cc, cur, count);
            RT3_CUDA(c, cudaMemcpyAsync(&count, ctr + 2, 4, cudaMemcpyDeviceToHost, st));
            RT3_CUDA(c, cudaStreamSynchronize(st));
            std::swap(cur, nxt);
            levels++;
            if (levels > 4 * RT3_MAX_STACK) RT3_FAIL(c, RT3_ERR_STATE, "BVH collapse did not terminate");
        }
        uint32_t fin[8];
        RT3_CUDA(c, cudaMemcpyAsync(fin, ctr, sizeof(fin), cudaMemcpyDeviceToHost, st));
        bs.ms_collapse += tm.end();
        dbg_mark("collapse");
        if (fin[1] != io.first_free_rec + n) RT3_FAIL(c, RT3_ERR_STATE, "BVH collapse lost leaf records");
        io.nodes_end = fin[0]; io.recs_end = fin[1]; io.levels = levels;
        memcpy(&io.sah_nodes, &fin[4], 4); memcpy(&io.sah_recs, &fin[5], 4);
    } catch (...) {
        cleanup();
        throw;
    }
    cleanup();
    dbg_mark("cleanup");
}

static void read_bounds(rt3_context* c, cudaStream_t st, const int* d_bounds, float lo[3], float hi[3]) {
    int hb[6];
    RT3_CUDA(c, cudaMemcpyAsync(hb, d_bounds, sizeof(hb), cudaMemcpyDeviceToHost, st));
    RT3_CUDA(c, cudaStreamSynchronize(st));
    for (int k = 0; k < 3; ++k) { lo[k] = ordered_to_float(hb[k]); hi[k] = ordered_to_float(hb[3 + k]); }
}
static float max_extent(const float lo[3], const float hi[3]) { return fmaxf(hi[0] - lo[0], fmaxf(hi[1] - lo[1], hi[2] - lo[2])); }

/* row-major 3x4 affine inverse (double arithmetic, rounded once) */
static bool invert_affine(const float m[12], float out[12]) {
    const double a = m[0], b = m[1], c = m[2], d = m[4], e = m[5], f = m[6], g = m[8], h = m[9], i = m[10];
    const double det = a * (e * i - f * h) - b * (d * i - f * g) + c * (d * h - e * g);
    if (!(fabs(det) > 1e-30)) return false;
    const double r[9] = {(e * i - f * h) / det, (c * h - b * i) / det, (b * f - c * e) / det, (f * g - d * i) / det, (a * i - c * g) / det,
                         (c * d - a * f) / det, (d * h - e * g) / det, (b * g - a * h) / det, (a * e - b * d) / det};
    for (int row = 0; row < 3; ++row) {
        for (int col = 0; col < 3; ++col) out[4 * row + col] = (float)r[3 * row + col];
        out[4 * row + 3] = (float)-(r[3 * row] * m[3] + r[3 * row + 1] * m[7] + r[3 * row + 2] * m[11]);
    }
    return true;
}

/* ---- single level: every instance flattened to world space (the default) -------------------------------------------- */
static void build_single_level(rt3_context* c, cudaStream_t st, rt3_build_stats& bs, bool optimal, int treelet) {
    const uint32_t n = c->n_triangles;
    StageTimer tm(c, st);
    dbg_mark("enter single");
    Tri48* out_tris = ensure<Tri48>(c, c->bvh_tris, n);
    dbg_mark("ensure tris");
    PoolBuf s_tris, s_misc;
    try {
        Tri48* tris = palloc<Tri48>(c, s_tris, n, st);
        int* d_misc = (int*)palloc<uint32_t>(c, s_misc, 16, st);
        tm.begin();
        RT3_LAUNCH(c, k_init_bounds, 1, 32, 0, st, d_misc);
        RT3_LAUNCH(c, k_flatten, div_up(n, 256), 256, 0, st, c->view, n, tris, d_misc);
        float lo[3], hi[3];
        read_bounds(c, st, d_misc, lo, hi);
        bs.ms_flatten += tm.end();
        dbg_mark("flatten");
        for (int k = 0; k < 3; ++k) { bs.scene_lo[k] = lo[k]; bs.scene_hi[k] = hi[k]; }
        WideBuild io;
        io.optimal_collapse = optimal; io.treelet = treelet;
        io.recs = &s_tris; io.n = n; io.kind = REC_TRI_EDGES; io.pad = max_extent(lo, hi) * RT3_PAD_FRACTION;
        memcpy(io.lo, lo, sizeof(lo)); memcpy(io.hi, hi, sizeof(hi));
        io.wide = ensure<WideNode>(c, c->bvh_nodes, n); io.root_index = 0; io.first_free_node = 1;
        io.out_recs = out_tris; io.first_free_rec = 0;
        build_wide(c, st, io, bs, d_misc);
        if (io.levels >= RT3_MAX_STACK) RT3_FAIL(c, RT3_ERR_STATE, "wide BVH deeper than the traversal stack");
        c->n_wide_nodes = io.nodes_end; c->n_leaf_records = n;
        bs.wide_nodes = io.nodes_end; bs.bvh2_nodes = 2 * n - 1; bs.levels = io.levels;
        bs.node_bytes = (uint64_t)io.nodes_end * sizeof(WideNode); bs.tri_bytes = (uint64_t)n * sizeof(Tri48);
        const float pad = io.pad, ex = hi[0] - lo[0] + 2 * pad, ey = hi[1] - lo[1] + 2 * pad, ez = hi[2] - lo[2] + 2 * pad;
        const float root_area = 2.0f * (ex * ey + ey * ez + ez * ex);
        bs.sah_cost = root_area > 0 ? (io.sah_nodes + io.sah_recs) / root_area : 0.f;
    } catch (...) { s_tris.release(); s_misc.release(); throw; }
    s_tris.release(); s_misc.release();
    c->two_level = false;
}

/* ---- two levels (RT3_BUILD_TWO_LEVEL): one bottom-level tree per distinct mesh, one top-level tree over the instances ---- */
static void build_two_level(rt3_context* c, cudaStream_t st, rt3_build_stats& bs, bool optimal, int treelet) {
    StageTimer tm(c, st);
    const std::vector<InstanceRec>& inst = c->h_instances;
    const uint32_t ni = (uint32_t)inst.size();
    /* distinct meshes = distinct (first_geometry, geometry_count); segments of instance i start at seg0[i] */
    struct MeshInfo { uint32_t first_geometry, geometry_count, seg0, prim0, n, root, first_inst; float lo[3], hi[3]; };
    std::vector<MeshInfo> meshes;
    std::vector<uint32_t> inst_mesh(ni), inst_seg0(ni);
    uint32_t seg = 0;
    uint64_t blas_tris = 0;
    for (uint32_t i = 0; i < ni; ++i) {
        inst_seg0[i] = seg;
        uint32_t m = 0;
        for (; m < meshes.size(); ++m) if (meshes[m].first_geometry == inst[i].first_geometry && meshes[m].geometry_count == inst[i].geometry_count) break;
        if (m == meshes.size()) {
            MeshInfo mi; memset(&mi, 0, sizeof(mi));
            mi.first_geometry = inst[i].first_geometry; mi.geometry_count = inst[i].geometry_count; mi.seg0 = seg; mi.first_inst = i;
            mi.prim0 = c->h_seg_first[seg]; mi.n = c->h_seg_first[seg + inst[i].geometry_count] - mi.prim0;
            meshes.push_back(mi);
            blas_tris += mi.n;
        }
        inst_mesh[i] = m;
        seg += inst[i].geometry_count;
    }
    const uint64_t total_recs = blas_tris + ni;
    if (total_recs >= 0x7FFFFFFFull) RT3_FAIL(c, RT3_ERR_INVALID, "too many leaf records");
    WideNode* wide = ensure<WideNode>(c, c->bvh_nodes, (size_t)total_recs + meshes.size() + 1);
    Tri48* out_recs = ensure<Tri48>(c, c->bvh_tris, (size_t)total_recs);
    PoolBuf s_recs, s_misc;
    try {
        int* d_misc = (int*)palloc<uint32_t>(c, s_misc, 16, st);
        /* pass 1: object-space bounds of every mesh (needed for the world boxes, which give the padding) */
        std::vector<InstRec> recs(ni);
        float wlo[3] = {3.4e38f, 3.4e38f, 3.4e38f}, whi[3] = {-3.4e38f, -3.4e38f, -3.4e38f};
        std::vector<PoolBuf> mesh_tris(meshes.size());
        tm.begin();
        for (size_t m = 0; m < meshes.size(); ++m) {
            MeshInfo& mi = meshes[m];
            if (mi.n == 0) continue;
            Tri48* t = palloc<Tri48>(c, mesh_tris[m], mi.n, st);
            RT3_LAUNCH(c, k_init_bounds, 1, 32, 0, st, d_misc);
            RT3_LAUNCH(c, k_flatten_object, div_up(mi.n, 256), 256, 0, st, c->view, mi.prim0, mi.n, t, d_misc);
            read_bounds(c, st, d_misc, mi.lo, mi.hi);
        }
        bs.ms_flatten += tm.end();
        std::vector<float> inst_lo(3 * ni), inst_hi(3 * ni);
        std::vector<float> mesh_inv_scale(meshes.size(), 0.f);
        for (uint32_t i = 0; i < ni; ++i) {
            const MeshInfo& mi = meshes[inst_mesh[i]];
            InstRec& r = recs[i];
            memcpy(r.xform, inst[i].xform, sizeof(r.xform));
            if (!invert_affine(r.xform, r.inv)) RT3_FAIL(c, RT3_ERR_INVALID, "instance transform is singular (two-level build)");
            r.first_prim = c->h_seg_first[inst_seg0[i]]; r.pad0 = r.pad1 = 0; r.root = 0;
            for (int k = 0; k < 3; ++k) { inst_lo[3 * i + k] = 3.4e38f; inst_hi[3 * i + k] = -3.4e38f; }
            if (mi.n == 0) { for (int k = 0; k < 3; ++k) inst_lo[3 * i + k] = inst_hi[3 * i + k] = r.xform[4 * k + 3]; }
            else for (int corner = 0; corner < 8; ++corner) {
                const float p[3] = {(corner & 1) ? mi.hi[0] : mi.lo[0], (corner & 2) ? mi.hi[1] : mi.lo[1], (corner & 4) ? mi.hi[2] : mi.lo[2]};
                for (int k = 0; k < 3; ++k) {
                    const float w = r.xform[4 * k] * p[0] + r.xform[4 * k + 1] * p[1] + r.xform[4 * k + 2] * p[2] + r.xform[4 * k + 3];
                    inst_lo[3 * i + k] = fminf(inst_lo[3 * i + k], w); inst_hi[3 * i + k] = fmaxf(inst_hi[3 * i + k], w);
                }
            }
            for (int k = 0; k < 3; ++k) { wlo[k] = fminf(wlo[k], inst_lo[3 * i + k]); whi[k] = fmaxf(whi[k], inst_hi[3 * i + k]); }
            float sc = 0.f;
            for (int row = 0; row < 3; ++row) sc = fmaxf(sc, fabsf(r.inv[4 * row]) + fabsf(r.inv[4 * row + 1]) + fabsf(r.inv[4 * row + 2]));
            mesh_inv_scale[inst_mesh[i]] = fmaxf(mesh_inv_scale[inst_mesh[i]], sc);
        }
        const float world_ext = max_extent(wlo, whi);
        for (int k = 0; k < 3; ++k) { bs.scene_lo[k] = wlo[k]; bs.scene_hi[k] = whi[k]; }
        /* pass 2: bottom-level trees, concatenated in one node / record array (node 0 is the top-level root) */
        uint32_t next_node = 1, next_rec = 0, max_levels = 0;
        for (size_t m = 0; m < meshes.size(); ++m) {
            MeshInfo& mi = meshes[m];
            mi.root = next_node;
            if (mi.n == 0) { RT3_LAUNCH(c, k_empty_root, 1, 32, 0, st, wide + mi.root); next_node++; continue; }
            WideBuild io;
            io.optimal_collapse = optimal; io.treelet = treelet;
            io.recs = &mesh_tris[m]; io.n = mi.n; io.kind = REC_TRI_VERTS;
            /* padding covers the intersection slop and node quantisation (object extent x RT3_PAD_FRACTION, as in the single-level build) AND the
             * rounding of the world -> object ray transform, which grows with the distance of the ray origin in object units */
            io.pad = (max_extent(mi.lo, mi.hi) + world_ext * mesh_inv_scale[m]) * RT3_PAD_FRACTION;
            memcpy(io.lo, mi.lo, sizeof(mi.lo)); memcpy(io.hi, mi.hi, sizeof(mi.hi));
            io.wide = wide; io.root_index = mi.root; io.first_free_node = mi.root + 1;
            io.out_recs = out_recs; io.first_free_rec = next_rec;
            build_wide(c, st, io, bs, d_misc);
            next_node = io.nodes_end; next_rec = io.recs_end; max_levels = std::max(max_levels, io.levels);
        }
        /* pass 3: top level over the instance boxes */
        std::vector<Tri48> boxes(ni);
        const float top_pad = world_ext * RT3_PAD_FRACTION;
        for (uint32_t i = 0; i < ni; ++i) {
            recs[i].root = meshes[inst_mesh[i]].root;
            const float bpad = meshes[inst_mesh[i]].n ? (max_extent(meshes[inst_mesh[i]].lo, meshes[inst_mesh[i]].hi) + world_ext * mesh_inv_scale[inst_mesh[i]]) * RT3_PAD_FRACTION : 0.f;
            float rs = 0.f;   /* object-space padding seen in world space */
            for (int row = 0; row < 3; ++row) rs = fmaxf(rs, fabsf(recs[i].xform[4 * row]) + fabsf(recs[i].xform[4 * row + 1]) + fabsf(recs[i].xform[4 * row + 2]));
            Tri48& b = boxes[i];
            memset(&b, 0, sizeof(b));
            for (int k = 0; k < 3; ++k) { b.v0[k] = inst_lo[3 * i + k] - bpad * rs; b.e1[k] = inst_hi[3 * i + k] + bpad * rs; }
            b.prim = i;
        }
        Tri48* d_boxes = palloc<Tri48>(c, s_recs, ni, st);
        RT3_CUDA(c, cudaMemcpyAsync(d_boxes, boxes.data(), ni * sizeof(Tri48), cudaMemcpyHostToDevice, st));
        WideBuild top;
        top.optimal_collapse = optimal; top.treelet = treelet;
        top.recs = &s_recs; top.n = ni; top.kind = REC_BOX; top.pad = top_pad;
        memcpy(top.lo, wlo, sizeof(wlo)); memcpy(top.hi, whi, sizeof(whi));
        top.wide = wide; top.root_index = 0; top.first_free_node = next_node;
        top.out_recs = out_recs; top.first_free_rec = next_rec;
        build_wide(c, st, top, bs, d_misc);
        if (top.levels + 3 + max_levels >= RT3_MAX_STACK) RT3_FAIL(c, RT3_ERR_STATE, "two-level BVH deeper than the traversal stack");
        InstRec* d_insts = ensure<InstRec>(c, c->d_insts, ni);
        RT3_CUDA(c, cudaMemcpyAsync(d_insts, recs.data(), ni * sizeof(InstRec), cudaMemcpyHostToDevice, st));
        RT3_CUDA(c, cudaStreamSynchronize(st));
        c->n_wide_nodes = top.nodes_end; c->n_leaf_records = top.recs_end;
        bs.wide_nodes = top.nodes_end; bs.bvh2_nodes = 0; bs.levels = top.levels + max_levels;
        bs.node_bytes = (uint64_t)top.nodes_end * sizeof(WideNode); bs.tri_bytes = (uint64_t)top.recs_end * sizeof(Tri48) + (uint64_t)ni * sizeof(InstRec);
        bs.sah_cost = 0.f;
        for (PoolBuf& b : mesh_tris) b.release();
    } catch (...) { s_recs.release(); s_misc.release(); throw; }
    s_recs.release(); s_misc.release();
    c->two_level = true;
}

void rt3_build_accel(rt3_context* c, uint32_t flags, cudaStream_t st, rt3_build_stats* out) {
    rt3_build_stats bs;
    memset(&bs, 0, sizeof(bs));
    /* a build that fails midway (out of memory, a depth guard) must not leave the previous structure marked usable: its buffers are being
     * overwritten.  Valid again only after the final synchronisation below. */
    c->accel_valid = false;
    const uint32_t n = c->n_triangles;
    bs.triangles = n;
    cudaEvent_t total_a = c->ev[2], total_b = c->ev[3];
    cudaEventRecord(total_a, st);
    if (n == 0) {
        WideNode* nodes = ensure<WideNode>(c, c->bvh_nodes, 1);
        ensure<Tri48>(c, c->bvh_tris, 1);
        RT3_LAUNCH(c, k_empty_root, 1, 32, 0, st, nodes);
        RT3_CUDA(c, cudaStreamSynchronize(st));
        c->n_wide_nodes = 1; c->n_leaf_records = 0; c->two_level = false; bs.wide_nodes = 1; bs.node_bytes = sizeof(WideNode);
    }
    c->view.shade_rec = nullptr; c->view.shade64 = nullptr;
    if (n && !getenv("RT3_NO_SHADE_REC")) {
        uint4* rec = ensure<uint4>(c, c->d_shade_rec, n);
        RT3_LAUNCH(c, k_shade_records, div_up(n, 256), 256, 0, st, c->view, n, rec);
        c->view.shade_rec = (const uint32_t*)rec;
        c->view.shade64 = nullptr;
        if ((size_t)n * 64 <= ((size_t)96 << 20) && !getenv("RT3_NO_SHADE64")) {       /* only while the table stays L2-resident */
            float4* r64 = ensure<float4>(c, c->d_shade64, 4 * (size_t)n);
            RT3_LAUNCH(c, k_shade_records64, div_up(n, 256), 256, 0, st, c->view, n, rec, r64);
            c->view.shade64 = (const float*)r64;
        }
    }
    /* children of the wide nodes: cost-optimal (dynamic programme over the BVH2, computed in the refit) unless the caller asks for
     * the fastest build, then greedy by surface area.  Measured: 36 % fewer nodes, +2 % path-tracing and +2..6 % incoherent-ray
     * throughput, build 1.37 -> 1.49 ms at 263 k triangles. */
    const bool optimal = !(flags & RT3_BUILD_FAST_BUILD);
    /* FAST_TRACE (the default, the reference's PREFER_FAST_TRACE) = treelet restructuring; RT3_BUILD_ROTATIONS keeps round 1's cheaper
     * optimisation (tree rotations).  Measured on B200, bench scene (profiles/r2a_ab_experiments.log): traversal launch 0.497 -> 0.477 ms,
     * incoherent closest 5.37 -> 5.68 Grays/s, frame 3011 -> 3063 Mrays/s, for a 1.9 -> 10.8 ms build. */
    const int treelet = (!optimal || (flags & RT3_BUILD_ROTATIONS)) ? 0 : ((flags & RT3_BUILD_TREELET) ? 2 : 1);
    if (n == 0) {
    } else if ((flags & RT3_BUILD_TWO_LEVEL) && !c->h_instances.empty()) {
        build_two_level(c, st, bs, optimal, treelet);
    } else {
        build_single_level(c, st, bs, optimal, treelet);
    }
    cudaEventRecord(total_b, st); cudaEventSynchronize(total_b);
    float ms = 0; cudaEventElapsedTime(&ms, total_a, total_b); bs.ms_total = ms;
    c->accel_valid = true; c->last_build = bs;
    if (out) *out = bs;
}
