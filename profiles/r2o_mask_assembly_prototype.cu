#include <cuda_runtime.h>
#include <stdint.h>
#include "rt3.h"
#include "rt3_shared_math.h"
#include "bvh_logic.h"
using namespace rt3b;

template <int J>
__device__ __forceinline__ float slab_gap(uint32_t nx, uint32_t fx, uint32_t ny, uint32_t fy, uint32_t nz, uint32_t fz, float sx, float sy, float sz, float bx,
                           float by, float bz, float tmin, float tmax, uint32_t mg) {
    const float t0x = rt3m::fma_(byte_as_biased_float<J>(nx, mg), sx, bx), t1x = rt3m::fma_(byte_as_biased_float<J>(fx, mg), sx, bx);
    const float t0y = rt3m::fma_(byte_as_biased_float<J>(ny, mg), sy, by), t1y = rt3m::fma_(byte_as_biased_float<J>(fy, mg), sy, by);
    const float t0z = rt3m::fma_(byte_as_biased_float<J>(nz, mg), sz, bz), t1z = rt3m::fma_(byte_as_biased_float<J>(fz, mg), sz, bz);
    const float tn = bmax(bmax(t0x, t0y), bmax(t0z, tmin));
    const float tf = bmin(bmin(t1x, t1y), bmin(t1z, tmax));
    return tf - tn;       /* sign bit set <=> miss */
}
__device__ __forceinline__ uint32_t prmt(uint32_t a, uint32_t b, uint32_t s) { uint32_t r; asm("prmt.b32 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(s)); return r; }

/* new: hit byte first.  n1.z = tri_valid24, lut = shared table perm[oct][byte] */
__device__ __forceinline__ uint32_t intersect_new(const RayPrep& r, float tmax, const u32x4& n0, const u32x4& n1, const u32x4& n2, const u32x4& n3, const u32x4& n4, const uint8_t* lut_oct) {
    const uint32_t e = n0.w;
    const float sx = rt3m::as_float((e & 0xFFu) << 23) * r.idx;
    const float sy = rt3m::as_float(((e >> 8) & 0xFFu) << 23) * r.idy;
    const float sz = rt3m::as_float(((e >> 16) & 0xFFu) << 23) * r.idz;
    const float bx = rt3m::fma_(-32768.0f, sx, (rt3m::as_float(n0.x) - r.ox) * r.idx);
    const float by = rt3m::fma_(-32768.0f, sy, (rt3m::as_float(n0.y) - r.oy) * r.idy);
    const float bz = rt3m::fma_(-32768.0f, sz, (rt3m::as_float(n0.z) - r.oz) * r.idz);
    uint32_t nib[2];
#pragma unroll
    for (int half = 0; half < 2; ++half) {
        const uint32_t qlx = half ? n2.y : n2.x, qly = half ? n2.w : n2.z, qlz = half ? n3.y : n3.x;
        const uint32_t qhx = half ? n3.w : n3.z, qhy = half ? n4.y : n4.x, qhz = half ? n4.w : n4.z;
        const uint32_t nx = r.idx < 0.0f ? qhx : qlx, fx = r.idx < 0.0f ? qlx : qhx;
        const uint32_t ny = r.idy < 0.0f ? qhy : qly, fy = r.idy < 0.0f ? qly : qhy;
        const uint32_t nz = r.idz < 0.0f ? qhz : qlz, fz = r.idz < 0.0f ? qlz : qhz;
        const float d0 = slab_gap<0>(nx, fx, ny, fy, nz, fz, sx, sy, sz, bx, by, bz, r.tmin, tmax, r.magic);
        const float d1 = slab_gap<1>(nx, fx, ny, fy, nz, fz, sx, sy, sz, bx, by, bz, r.tmin, tmax, r.magic);
        const float d2 = slab_gap<2>(nx, fx, ny, fy, nz, fz, sx, sy, sz, bx, by, bz, r.tmin, tmax, r.magic);
        const float d3 = slab_gap<3>(nx, fx, ny, fy, nz, fz, sx, sy, sz, bx, by, bz, r.tmin, tmax, r.magic);
        const uint32_t a = prmt(__float_as_uint(d0), __float_as_uint(d1), 0x0073u), b = prmt(__float_as_uint(d2), __float_as_uint(d3), 0x0073u);
        const uint32_t w = prmt(a, b, 0x5410u);
        nib[half] = ((w & 0x80808080u) * 0x00204081u) >> 28;       /* miss bits of children 4 half .. 4 half + 3 */
    }
    const uint32_t imask = n0.w >> 24;
    const uint32_t hb = ~(nib[0] | (nib[1] << 4)) & 0xFFu;
    const uint32_t inner = (uint32_t)lut_oct[hb & imask] << 24;
    uint32_t x = hb & ~imask;
    x = (x | (x << 8)) & 0x00F00Fu; x = (x | (x << 4)) & 0x0C30C3u; x = (x | (x << 2)) & 0x249249u;
    const uint32_t tri = (x * 7u) & n1.z;
    return inner | tri;
}

__global__ void k_old(const WideNode* nodes, const float* rays, uint32_t* out) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    rt3_ray ray; for (int k = 0; k < 3; ++k) { ray.o[k] = rays[8 * i + k]; ray.d[k] = rays[8 * i + 4 + k]; } ray.tmin = rays[8 * i + 3]; ray.tmax = rays[8 * i + 7];
    RayPrep r = prepare_ray(ray, out[0]);
    const WideNode* nd = nodes + i;
    const u32x4 n0 = ld_u32x4(nd->w), n1 = ld_u32x4(nd->w + 4), n2 = ld_u32x4(nd->w + 8), n3 = ld_u32x4(nd->w + 12), n4 = ld_u32x4(nd->w + 16);
    out[i] = intersect_wide_node(r, ray.tmax, n0, n1, n2, n3, n4);
}
__global__ void k_new(const WideNode* nodes, const float* rays, uint32_t* out, const uint8_t* lut) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    rt3_ray ray; for (int k = 0; k < 3; ++k) { ray.o[k] = rays[8 * i + k]; ray.d[k] = rays[8 * i + 4 + k]; } ray.tmin = rays[8 * i + 3]; ray.tmax = rays[8 * i + 7];
    RayPrep r = prepare_ray(ray, out[0]);
    const WideNode* nd = nodes + i;
    const u32x4 n0 = ld_u32x4(nd->w), n1 = ld_u32x4(nd->w + 4), n2 = ld_u32x4(nd->w + 8), n3 = ld_u32x4(nd->w + 12), n4 = ld_u32x4(nd->w + 16);
    out[i] = intersect_new(r, ray.tmax, n0, n1, n2, n3, n4, lut + ((r.octinv4 & 7u) << 8));
}
