/*
 * trace_kernel.cuh — the production traversal kernel k_trace_dyn (closest-hit / any-hit over the 8-wide quantised BVH), in a header of its
 * own so that two translation units see the very same source: trace.cu (nvcc, the product) and tests/emu/warp_emu.cpp, which compiles it
 * with g++ over a lock-step warp emulation (32 fibers per warp, ballots / shuffles / __syncwarp as rendezvous points) so that the
 * `-m "not gpu"` suite can run the kernel's control flow — dynamic ray fetch, two-phase loop, stack protocol — against a brute-force
 * loop without a GPU.  The emulation is test infrastructure; librt3.so contains the device build only.
 *
 * Replaces TraceRay inside vkCmdTraceRaysKHR (/root/reference/src/renderer/pipeline_cache/mod.rs:64-73) with the closest-hit / miss
 * shaders (/root/reference/shaders/old/default_hit.slang:3-9, default_miss.slang:3-8) folded into the epilogue.
 */
#ifndef RT3_TRACE_KERNEL_CUH
#define RT3_TRACE_KERNEL_CUH

/* The warp emulation counts how often a warp runs each phase of the loop and with how many lanes (RT3_EMU_PHASE is a ballot there);
 * in the device build the hook is empty (bvh_logic.h). */

namespace {

/* threads per CTA of the traversal kernels (experiments: -DRT3_TRACE_THREADS=64 with RT3_TRACE_BLOCKS_PER_SM=16 retires CTAs in smaller units) */
#ifndef RT3_TRACE_THREADS
#define RT3_TRACE_THREADS 128
#endif
constexpr int TRACE_THREADS = RT3_TRACE_THREADS;

__device__ __forceinline__ rt3_ray load_ray(const rt3_ray* rays, uint64_t i) {
    const float4* p = (const float4*)(rays + i);
    float4 a = __ldcs(p), b = __ldcs(p + 1);
    rt3_ray r;
    r.o[0] = a.x; r.o[1] = a.y; r.o[2] = a.z; r.tmin = a.w;
    r.d[0] = b.x; r.d[1] = b.y; r.d[2] = b.z; r.tmax = b.w;
    return r;
}

/*
 * Production traversal kernel: persistent warps with DYNAMIC RAY FETCH.  A warp that keeps its 32 rays until the
 * slowest one finishes runs at the mean/max traversal length (measured: 7.5 of 32 lanes active on the secondary-ray
 * queues); here a lane whose ray terminates writes its result and, once fewer than KEEP lanes of the warp are still
 * busy, the idle lanes pull new rays from the global queue with one warp-aggregated atomic.
 */
/* Optional: the first RT3_SMEM_STACK traversal-stack entries of every thread in shared memory ([entry][thread]: conflict-free,
 * two wavefronts per warp access), deeper entries in thread-local memory.  Local-memory stack traffic is 16 % of the kernel's L1
 * sectors (profiles/r1d_trace_summary.txt), but the kernel is bound by the ALU pipe, not by L1: measured on B200 (1080p 4-bounce)
 * depth 0 / 4 / 8 / 12 = 0.539 / 0.548 / 0.549 / 0.550 ms per launch, so the default keeps the stack in local memory. */
#ifndef RT3_SMEM_STACK
#define RT3_SMEM_STACK 0
#endif
/* record steps per node step of the two-phase loop: a second one drains lanes that hold two triangles before the warp returns to the
 * node phase (measured: traversal launch 0.507 -> 0.497 ms, incoherent closest / any +2.6 % / +4 %; a third one costs more than it saves) */
/* reciprocal ray direction by MUFU.RCP + one Newton step instead of three IEEE divisions per fetched ray (bvh_logic.h: safe_rcp_dir_t) */
#ifndef RT3_FAST_RCP
#define RT3_FAST_RCP 1
#endif
#ifndef RT3_TRI_REPS
#define RT3_TRI_REPS 2
#endif
#ifndef RT3_TRI_BALANCE
#define RT3_TRI_BALANCE 0
#endif
#ifndef RT3_TRI_REP2_MIN
#define RT3_TRI_REP2_MIN 1
#endif
#ifndef RT3_NODE_COOP
#define RT3_NODE_COOP 0
#endif
#ifndef RT3_TRI_WAIT
#define RT3_TRI_WAIT 0
#endif
/* experiments only: -DRT3_TRACE_MIN_BLOCKS=8 caps the registers for 8 resident CTAs per SM (RT3_TRI_BALANCE needs 72 without it) */
#ifdef RT3_TRACE_MIN_BLOCKS
#define RT3_TRACE_BOUNDS __launch_bounds__(TRACE_THREADS, RT3_TRACE_MIN_BLOCKS)
#else
#define RT3_TRACE_BOUNDS __launch_bounds__(TRACE_THREADS)
#endif
/* TWO = two-level structure (RT3_BUILD_TWO_LEVEL): top-level leaf records are instances; see traverse_wide in bvh_logic.h for the
 * stack protocol (node group, record group, sentinel).  The world-space ray stays in registers for the triangle test, which
 * rebuilds the world-space triangle from the stored object-space vertices (bit-identical to the flattened build). */
/* TAIL (instrumented instantiation only, rt3_trace_tail_stats): every warp records when it entered, when it first found the queue
 * exhausted and when it left (%globaltimer, ns) into tail_stats = {sum busy, sum tail, min start, max end, warps}. */
__device__ __forceinline__ unsigned long long global_ns() {
#ifdef __CUDA_ARCH__
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
#else
    return 0ull;
#endif
}
template <bool ANY, int KEEP, bool TWO, bool TAIL = false>
__global__ void RT3_TRACE_BOUNDS k_trace_dyn(const WideNode* __restrict__ nodes, const Tri48* __restrict__ tris,
                                                             const rt3_ray* __restrict__ rays, const uint32_t* __restrict__ d_count, uint64_t n_static,
                                                             rt3_hit* __restrict__ hits, uint32_t* __restrict__ bits, unsigned long long* work_counter, uint32_t magic,
                                                             const InstRec* __restrict__ insts, unsigned long long* tail_stats = nullptr) {
    const uint64_t n = d_count ? (uint64_t)*d_count : n_static;
    unsigned long long t_start = 0, t_exhausted = 0;
    if (TAIL) t_start = global_ns();
    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t lt_mask = (1u << lane) - 1u;
    RayPrep r;
    float ray_tmax = 0.f;
    float wox = 0.f, woy = 0.f, woz = 0.f, wdx = 0.f, wdy = 0.f, wdz = 1.f;    /* TWO: the world-space ray (r is the current level's) */
    uint32_t cur_inst = RT3_STACK_SENTINEL;
    uint64_t my_index = 0;
    rt3_hit h; h.t = 0.f; h.u = 0.f; h.v = 0.f; h.prim = RT3_MISS;
    uint32_t ng_base = 0, ng_hits = 0, tg_base = 0, tg_hits = 0;
    uint32_t tg_valid = 0x00FFFFFFu;       /* valid24 of the node the pending records belong to (record_index, bvh_logic.h) */
    RT3_PERM_LUT_SETUP(lut_base, TRACE_THREADS);          /* RT3_NODE_V2: the octant permutation table in shared memory */
#if RT3_SMEM_STACK > 0
    __shared__ uint2 s_stack[RT3_SMEM_STACK][TRACE_THREADS];
    uint32_t stack_base[RT3_MAX_STACK - RT3_SMEM_STACK], stack_hits[RT3_MAX_STACK - RT3_SMEM_STACK];
#define RT3_PUSH(B, H) do { if (sp < RT3_SMEM_STACK) s_stack[sp][threadIdx.x] = make_uint2((B), (H)); else { stack_base[sp - RT3_SMEM_STACK] = (B); stack_hits[sp - RT3_SMEM_STACK] = (H); } sp++; } while (0)
#define RT3_POP(B, H) do { sp--; if (sp < RT3_SMEM_STACK) { const uint2 e_ = s_stack[sp][threadIdx.x]; (B) = e_.x; (H) = e_.y; } else { (B) = stack_base[sp - RT3_SMEM_STACK]; (H) = stack_hits[sp - RT3_SMEM_STACK]; } } while (0)
#else
    uint32_t stack_base[RT3_MAX_STACK], stack_hits[RT3_MAX_STACK];
#define RT3_PUSH(B, H) do { stack_base[sp] = (B); stack_hits[sp] = (H); sp++; } while (0)
#define RT3_POP(B, H) do { sp--; (B) = stack_base[sp]; (H) = stack_hits[sp]; } while (0)
#endif
    int sp = 0;
    bool active = false, exhausted = false;
    r.ox = r.oy = r.oz = r.dx = r.dy = r.dz = r.idx = r.idy = r.idz = r.tmin = 0.f; r.octinv4 = 0; r.magic = magic;
#if RT3_NODE_V2
    r.lutrow = lut_base;
#endif
    for (;;) {
        /* ---- refill idle lanes */
        const uint32_t idle = __ballot_sync(0xFFFFFFFFu, !active);
        RT3_EMU_PHASE(4, !active && !exhausted);
        if (idle && !exhausted) {
            const int leader = __ffs(idle) - 1;
            const uint32_t want = __popc(idle);
            unsigned long long base = 0;
            if ((int)lane == leader) base = atomicAdd(work_counter, (unsigned long long)want);
            base = __shfl_sync(0xFFFFFFFFu, base, leader);
            if (!active) {
                const uint64_t i = base + __popc(idle & lt_mask);
                if (i < n) {
                    const rt3_ray ray = load_ray(rays, i);
                    r = prepare_ray<RT3_FAST_RCP>(ray, magic, lut_base);
                    if (TWO) { wox = ray.o[0]; woy = ray.o[1]; woz = ray.o[2]; wdx = ray.d[0]; wdy = ray.d[1]; wdz = ray.d[2]; cur_inst = RT3_STACK_SENTINEL; }
                    ray_tmax = ray.tmax;
                    h.t = ray.tmax; h.u = 0.f; h.v = 0.f; h.prim = RT3_MISS;
                    ng_base = 0; ng_hits = 0x80000000u; tg_hits = 0; sp = 0;
                    my_index = i;
                    active = true;
                }
            }
            if (base + want >= n) {
                if (TAIL && !exhausted) t_exhausted = global_ns();
                exhausted = true;
            }
        }
        if (__ballot_sync(0xFFFFFFFFu, active) == 0u) break;
        const int keep = exhausted ? 1 : KEEP;
        /* ---- traverse until too few lanes are busy.  One iteration = a node step for the lanes that have no record
         * pending, then ONE record (triangle test / instance entry) for every lane that has one pending: the warp re-converges
         * between the two phases, so triangle tests of different lanes share issue slots instead of running 3-of-32 wide in a
         * per-lane loop (ncu source page of the per-lane-loop version: profiles/r1b_*). */
        do {
            RT3_EMU_PHASE(0, active && tg_hits == 0u && ng_hits > 0x00FFFFFFu);
#if RT3_NODE_COOP
            /* EXPERIMENT (-DRT3_NODE_COOP=1; north_star: "vectorised node loads staged in shared memory"): the warp fetches the nodes of all its
             * lanes cooperatively — the 32 x 5 sixteen-byte words are dealt to the lanes in address order (lane L of round r loads word
             * (32 r + L) % 5 of the node of lane (32 r + L) / 5), so five neighbouring lanes read one node's 80 contiguous bytes (3 sectors)
             * instead of every lane issuing five 16-byte loads to its own node (5 sectors) — through 2.5 KB of shared memory per warp. */
            {
                __shared__ u32x4 s_node[TRACE_THREADS / 32][32 * 5];
                const uint32_t wid = threadIdx.x >> 5;
                const bool step = active && tg_hits == 0u && ng_hits > 0x00FFFFFFu;
                uint32_t node_index = 0xFFFFFFFFu;
                int bit = 0;
                if (step) {
                    bit = bfind32(ng_hits);
                    const uint32_t imask = ng_hits & 0xFFu;
                    const uint32_t slot = ((uint32_t)(bit - 24) ^ (r.octinv4 & 0xFFu)) & 7u;
                    node_index = ng_base + (uint32_t)__popc(imask & ~(0xFFFFFFFFu << slot));
                }
                if (__ballot_sync(0xFFFFFFFFu, step)) {
#pragma unroll
                    for (uint32_t rr = 0; rr < 5u; ++rr) {
                        const uint32_t c = rr * 32u + lane, owner = c / 5u, part = c - owner * 5u;
                        const uint32_t oi = __shfl_sync(0xFFFFFFFFu, node_index, owner);
                        if (oi != 0xFFFFFFFFu) s_node[wid][c] = ld_u32x4(nodes[oi].w + 4u * part);
                    }
                    __syncwarp();
                    if (step) {
                        ng_hits &= ~(1u << bit);
                        if (ng_hits > 0x00FFFFFFu && sp < RT3_MAX_STACK) RT3_PUSH(ng_base, ng_hits);
                        const u32x4 n0 = s_node[wid][lane * 5u], n1 = s_node[wid][lane * 5u + 1u], n2 = s_node[wid][lane * 5u + 2u], n3 = s_node[wid][lane * 5u + 3u],
                                    n4 = s_node[wid][lane * 5u + 4u];
                        const uint32_t hm = intersect_wide_node(r, h.t, n0, n1, n2, n3, n4);
                        ng_base = n1.x; ng_hits = (hm & 0xFF000000u) | (n0.w >> 24);
                        tg_base = n1.y; tg_hits = hm & 0x00FFFFFFu; tg_valid = n1.z;
                    }
                    __syncwarp();
                }
            }
#else
            if (active && tg_hits == 0u && ng_hits > 0x00FFFFFFu) {
                const int bit = bfind32(ng_hits);
                const uint32_t imask = ng_hits & 0xFFu;
                ng_hits &= ~(1u << bit);
                if (ng_hits > 0x00FFFFFFu && sp < RT3_MAX_STACK) RT3_PUSH(ng_base, ng_hits);
                const uint32_t slot = ((uint32_t)(bit - 24) ^ (r.octinv4 & 0xFFu)) & 7u;
                const uint32_t rel = (uint32_t)__popc(imask & ~(0xFFFFFFFFu << slot));
                const WideNode* nd = nodes + (ng_base + rel);
                const u32x4 n0 = ld_u32x4(nd->w), n1 = ld_u32x4(nd->w + 4), n2 = ld_u32x4(nd->w + 8), n3 = ld_u32x4(nd->w + 12),
                            n4 = ld_u32x4(nd->w + 16);
                const uint32_t hm = intersect_wide_node(r, h.t, n0, n1, n2, n3, n4);
                ng_base = n1.x; ng_hits = (hm & 0xFF000000u) | (n0.w >> 24);
                tg_base = n1.y; tg_hits = hm & 0x00FFFFFFu; tg_valid = n1.z;
            }
#endif
            bool found_any = false;
#if RT3_TRI_BALANCE
            if (!TWO) {
                /* EXPERIMENT (compile with -DRT3_TRI_BALANCE=1; not in the product build, checked on the CPU warp emulation, not yet run on a
                 * GPU): the record phase spread over the whole warp.  Up to three pending triangles per lane are numbered by a warp prefix
                 * sum; lane k tests triangle k (up to 32 a step) with the owner's ray fetched by shuffles, results go back through shared
                 * memory and the owner keeps the closest.  With -DRT3_TRI_WAIT=T the step waits until T triangles are pending in the warp
                 * (or no lane could take a node step). */
                __shared__ uint16_t s_src[TRACE_THREADS / 32][32];
                __shared__ float4 s_res[TRACE_THREADS / 32][32];
                const uint32_t wid = threadIdx.x >> 5;
                uint32_t rest = active ? tg_hits : 0u;
                const uint32_t cnt = (uint32_t)__popc(rest);
                const uint32_t c3 = cnt < 3u ? cnt : 3u;
                uint32_t incl = c3;
#pragma unroll
                for (int off = 1; off < 32; off <<= 1) { const uint32_t v = __shfl_up_sync(0xFFFFFFFFu, incl, off); if ((int)lane >= off) incl += v; }
                const uint32_t total = __shfl_sync(0xFFFFFFFFu, incl, 31);
                RT3_EMU_PHASE(2, cnt != 0u);
#if RT3_TRI_WAIT > 0
                const uint32_t act_ = __ballot_sync(0xFFFFFFFFu, active), pend_ = __ballot_sync(0xFFFFFFFFu, cnt != 0u);
                if (total >= RT3_TRI_WAIT || (total && pend_ == act_)) {
#else
                if (total) {
#endif
                    const uint32_t first = incl - c3;
                    const uint32_t room = first < 32u ? 32u - first : 0u;
                    const uint32_t mine = c3 < room ? c3 : room;
#pragma unroll
                    for (uint32_t j = 0; j < 3u; ++j)
                        if (j < mine) {
                            const int bit = bfind32(rest);
                            rest &= ~(1u << bit);
                            s_src[wid][first + j] = (uint16_t)(lane | (record_index(tg_valid, bit) << 8));   /* owner lane, record offset */
                        }
                    if (active) tg_hits = rest;                                   /* what did not fit into this step stays pending */
                    __syncwarp();
                    const uint32_t items = total < 32u ? total : 32u;
                    RT3_EMU_PHASE(3, lane < items);
                    const uint32_t e = lane < items ? (uint32_t)s_src[wid][lane] : lane;
                    const uint32_t src = e & 31u, sbit = e >> 8;
                    const uint32_t sbase = __shfl_sync(0xFFFFFFFFu, tg_base, src);
                    const float sox = __shfl_sync(0xFFFFFFFFu, r.ox, src), soy = __shfl_sync(0xFFFFFFFFu, r.oy, src), soz = __shfl_sync(0xFFFFFFFFu, r.oz, src);
                    const float sdx = __shfl_sync(0xFFFFFFFFu, r.dx, src), sdy = __shfl_sync(0xFFFFFFFFu, r.dy, src), sdz = __shfl_sync(0xFFFFFFFFu, r.dz, src);
                    const float stmin = __shfl_sync(0xFFFFFFFFu, r.tmin, src), stmax = __shfl_sync(0xFFFFFFFFu, ray_tmax, src);
                    float4 res = make_float4(0.f, 0.f, 0.f, __uint_as_float(RT3_MISS));
                    if (lane < items) {
                        const Tri48* tp = tris + (sbase + sbit);
                        const u32x4 a = ld_u32x4(tp), b = ld_u32x4((const uint32_t*)tp + 4), c = ld_u32x4((const uint32_t*)tp + 8);
                        float t, u, v;
                        const rt3m::vec3 v0 = rt3m::v3(__uint_as_float(a.x), __uint_as_float(a.y), __uint_as_float(a.z));
                        const rt3m::vec3 e1 = rt3m::v3(__uint_as_float(b.x), __uint_as_float(b.y), __uint_as_float(b.z));
                        const rt3m::vec3 e2 = rt3m::v3(__uint_as_float(c.x), __uint_as_float(c.y), __uint_as_float(c.z));
                        if (rt3m::ray_triangle(rt3m::v3(sox, soy, soz), rt3m::v3(sdx, sdy, sdz), v0, e1, e2, t, u, v) && t > stmin && t < stmax)
                            res = make_float4(t, u, v, __uint_as_float(a.w));
                    }
                    s_res[wid][lane] = res;
                    __syncwarp();
#pragma unroll
                    for (uint32_t j = 0; j < 3u; ++j)
                        if (j < mine) {
                            const float4 q = s_res[wid][first + j];
                            const uint32_t prim = __float_as_uint(q.w);
                            if (prim != RT3_MISS) {
                                if (ANY) found_any = true;
                                else if (rt3m::closer_hit(q.x, prim, h.t, h.prim)) { h.t = q.x; h.u = q.y; h.v = q.z; h.prim = prim; }
                            }
                        }
                    __syncwarp();                                                  /* the next step overwrites s_src / s_res */
                    if (ANY && found_any) tg_hits = 0;
                }
            } else
#endif
#if RT3_TRI_WAIT > 0 && !RT3_TRI_BALANCE
            /* EXPERIMENT (-DRT3_TRI_WAIT=T): lanes holding triangles wait (they take no node step either) until at least T lanes of the
             * warp hold some, or until no lane could take a node step; fewer, fuller record steps against idle lanes in the node phase */
            const uint32_t pend_ = __ballot_sync(0xFFFFFFFFu, active && tg_hits != 0u), act_ = __ballot_sync(0xFFFFFFFFu, active);
            if (__popc(pend_) >= RT3_TRI_WAIT || pend_ == act_)
#endif
#if RT3_TRI_REPS > 1
            for (int rep_ = 0; rep_ < RT3_TRI_REPS; ++rep_) {
#if RT3_TRI_REP2_MIN > 1
            /* EXPERIMENT (-DRT3_TRI_REP2_MIN=T): the second record step only when at least T lanes still hold a record */
            if (rep_ && __popc(__ballot_sync(0xFFFFFFFFu, active && tg_hits != 0u && !found_any)) < RT3_TRI_REP2_MIN) break;
#else
            if (rep_ && __ballot_sync(0xFFFFFFFFu, active && tg_hits != 0u && !found_any) == 0u) break;
#endif
#endif
            RT3_EMU_PHASE(1, active && tg_hits && !found_any);
            if (active && tg_hits && !found_any) {
                const int bit = bfind32(tg_hits);
                tg_hits &= ~(1u << bit);
                const Tri48* tp = tris + (tg_base + record_index(tg_valid, bit));
                const u32x4 a = ld_u32x4(tp), b = ld_u32x4((const uint32_t*)tp + 4), c = ld_u32x4((const uint32_t*)tp + 8);
                if (TWO && cur_inst == RT3_STACK_SENTINEL) {
                    if (ng_hits > 0x00FFFFFFu && sp < RT3_MAX_STACK) RT3_PUSH(ng_base, ng_hits);
                    if (tg_hits && sp < RT3_MAX_STACK) RT3_PUSH(tg_base, dense_records(tg_hits, tg_valid));
                    if (sp < RT3_MAX_STACK) RT3_PUSH(RT3_STACK_SENTINEL, 0u);
                    cur_inst = a.w;
                    rt3_ray wr; wr.o[0] = wox; wr.o[1] = woy; wr.o[2] = woz; wr.d[0] = wdx; wr.d[1] = wdy; wr.d[2] = wdz; wr.tmin = r.tmin; wr.tmax = ray_tmax;
                    r = prepare_ray<RT3_FAST_RCP>(object_ray(insts[cur_inst], wr), magic, lut_base);
                    ng_base = insts[cur_inst].root; ng_hits = 0x80000000u; tg_hits = 0u;
                } else {
                    float t, u, v;
                    rt3m::vec3 v0, e1, e2, ro, rd;
                    uint32_t prim = a.w;
                    if (TWO) {
                        instance_triangle(insts[cur_inst], a, b, c, v0, e1, e2);
                        prim += insts[cur_inst].first_prim;
                        ro = rt3m::v3(wox, woy, woz); rd = rt3m::v3(wdx, wdy, wdz);
                    } else {
                        v0 = rt3m::v3(__uint_as_float(a.x), __uint_as_float(a.y), __uint_as_float(a.z));
                        e1 = rt3m::v3(__uint_as_float(b.x), __uint_as_float(b.y), __uint_as_float(b.z));
                        e2 = rt3m::v3(__uint_as_float(c.x), __uint_as_float(c.y), __uint_as_float(c.z));
                        ro = rt3m::v3(r.ox, r.oy, r.oz); rd = rt3m::v3(r.dx, r.dy, r.dz);
                    }
                    if (rt3m::ray_triangle(ro, rd, v0, e1, e2, t, u, v) && t > r.tmin && t < ray_tmax) {
                        if (ANY) { found_any = true; tg_hits = 0; }
                        else if (rt3m::closer_hit(t, prim, h.t, h.prim)) { h.t = t; h.u = u; h.v = v; h.prim = prim; }
                    }
                }
            }
#if RT3_TRI_REPS > 1
            }
#endif
            if (active && (found_any || (tg_hits == 0u && ng_hits <= 0x00FFFFFFu))) {
                bool done = found_any || sp == 0;
                if (!TWO) {
                    if (!done) RT3_POP(ng_base, ng_hits);
                } else {
                    while (!done && tg_hits == 0u && ng_hits <= 0x00FFFFFFu) {     /* pop; more than once only across a sentinel */
                        if (sp == 0) { done = true; break; }
                        uint32_t eb, eh;
                        RT3_POP(eb, eh);
                        if (eb == RT3_STACK_SENTINEL && eh == 0u) {
                            cur_inst = RT3_STACK_SENTINEL;
                            rt3_ray wr; wr.o[0] = wox; wr.o[1] = woy; wr.o[2] = woz; wr.d[0] = wdx; wr.d[1] = wdy; wr.d[2] = wdz; wr.tmin = r.tmin; wr.tmax = ray_tmax;
                            r = prepare_ray<RT3_FAST_RCP>(wr, magic, lut_base);
                            continue;
                        }
                        if (eh > 0x00FFFFFFu) { ng_base = eb; ng_hits = eh; } else { tg_base = eb; tg_hits = eh; tg_valid = 0x00FFFFFFu; ng_hits = 0u; }
                    }
                }
                if (done) {
                    if (ANY) { if (found_any) atomicOr(bits + (my_index >> 5), 1u << (my_index & 31u)); }
                    else __stcs((float4*)(hits + my_index), make_float4(h.t, h.u, h.v, __uint_as_float(h.prim)));
                    active = false;
                }
            }
        } while (__popc(__ballot_sync(0xFFFFFFFFu, active)) >= keep);
    }
#ifdef __CUDA_ARCH__      /* the warp emulation has no timer and never instantiates TAIL */
    if (TAIL && (threadIdx.x & 31u) == 0u) {
        const unsigned long long t_end = global_ns();
        atomicAdd(tail_stats + 0, t_end - t_start);
        atomicAdd(tail_stats + 1, t_end - (t_exhausted ? t_exhausted : t_end));
        atomicMin(tail_stats + 2, t_start);
        atomicMax(tail_stats + 3, t_end);
        atomicAdd(tail_stats + 4, 1ull);
    }
#endif
#undef RT3_PUSH
#undef RT3_POP
}

}  // namespace
#endif
