/*
 * rt3_shared_math.h — the reference's Slang shader math restated as plain C++ that compiles both
 * as host code (g++, used by oracle/) and as device code (nvcc, used by the CUDA kernels).
 *
 * Determinism contract.  Every function below uses only IEEE-754 binary32 operations that are
 * correctly rounded on x86-64 and on sm_100a ( + - * / sqrt fma, int<->float conversion ) in a
 * fixed order.  Transcendentals (sin, cos, atan2, asin, exp2, log2) are replaced by fixed
 * polynomial kernels built from those operations.  Build flags that make this hold:
 *     nvcc:  -fmad=false   (no implicit contraction; explicit fmaf() still maps to FFMA)
 *            default -prec-div=true -prec-sqrt=true -ftz=false, never --use_fast_math
 *     g++ :  -ffp-contract=off -mfma  (fmaf inlines to vfmadd), never -ffast-math
 * Under that contract the CPU oracle and the GPU kernels produce BIT-IDENTICAL results for every
 * function in this file; rt3m::contract_probe() is a runtime check of the flags.
 *
 * Each function cites the reference file:line it follows (paths under /root/reference).
 */
#ifndef RT3_SHARED_MATH_H
#define RT3_SHARED_MATH_H

#include <stdint.h>
#include <string.h>
#include <math.h>
#include "rt3.h"

#if defined(__FAST_MATH__)
#error "rt3_shared_math.h must not be compiled with -ffast-math"
#endif

#ifdef __CUDACC__
#define RT3_HD __host__ __device__ __forceinline__
#else
#define RT3_HD inline
#endif

namespace rt3m {

/* ---------------------------------------------------------------- scalar helpers ------------- */
RT3_HD float as_float(uint32_t u) {
#ifdef __CUDA_ARCH__
    return __uint_as_float(u);
#else
    float f; memcpy(&f, &u, 4); return f;
#endif
}
RT3_HD uint32_t as_uint(float f) {
#ifdef __CUDA_ARCH__
    return __float_as_uint(f);
#else
    uint32_t u; memcpy(&u, &f, 4); return u;
#endif
}
RT3_HD float fma_(float a, float b, float c) {
#ifdef __CUDA_ARCH__
    return __fmaf_rn(a, b, c);
#else
    return __builtin_fmaf(a, b, c);
#endif
}
RT3_HD float sqrt_(float a) {
#ifdef __CUDA_ARCH__
    return __fsqrt_rn(a);
#else
    return __builtin_sqrtf(a);
#endif
}
RT3_HD float min_(float a, float b) { return a < b ? a : b; }
RT3_HD float max_(float a, float b) { return a > b ? a : b; }
RT3_HD float clamp_(float v, float lo, float hi) { return min_(max_(v, lo), hi); }
RT3_HD float floor_(float a) {
#ifdef __CUDA_ARCH__
    return floorf(a);
#else
    return __builtin_floorf(a);
#endif
}
RT3_HD float abs_(float a) { return as_float(as_uint(a) & 0x7FFFFFFFu); }
/* a*b+c evaluated with two roundings; used to detect accidental contraction. */
RT3_HD float contract_probe(float a, float b, float c) { return a * b + c; }

/* ---------------------------------------------------------------- vectors -------------------- */
struct vec2 { float x, y; };
struct vec3 { float x, y, z; };
struct vec4 { float x, y, z, w; };

RT3_HD vec3 v3(float x, float y, float z) { vec3 r; r.x = x; r.y = y; r.z = z; return r; }
RT3_HD vec3 v3(const float* p) { return v3(p[0], p[1], p[2]); }
RT3_HD vec3 operator+(vec3 a, vec3 b) { return v3(a.x + b.x, a.y + b.y, a.z + b.z); }
RT3_HD vec3 operator-(vec3 a, vec3 b) { return v3(a.x - b.x, a.y - b.y, a.z - b.z); }
RT3_HD vec3 operator*(vec3 a, vec3 b) { return v3(a.x * b.x, a.y * b.y, a.z * b.z); }
RT3_HD vec3 operator*(vec3 a, float s) { return v3(a.x * s, a.y * s, a.z * s); }
RT3_HD float dot(vec3 a, vec3 b) { return fma_(a.z, b.z, fma_(a.y, b.y, a.x * b.x)); }
RT3_HD vec3 cross(vec3 a, vec3 b) {
    return v3(fma_(a.y, b.z, -(a.z * b.y)), fma_(a.z, b.x, -(a.x * b.z)), fma_(a.x, b.y, -(a.y * b.x)));
}
/* Slang normalize(v) = v * rsqrt(dot(v,v)); restated with a correctly rounded 1/sqrt. */
RT3_HD vec3 normalize(vec3 a) { float s = 1.0f / sqrt_(dot(a, a)); return a * s; }
/* a + t*d */
RT3_HD vec3 madd(vec3 a, float t, vec3 d) { return v3(fma_(t, d.x, a.x), fma_(t, d.y, a.y), fma_(t, d.z, a.z)); }

/* column-major 4x4 (glam::Mat4, `column_major float4x4`) times (x,y,z,w): HLSL mul(M, v). */
RT3_HD vec4 mat4_mul(const float* m, float x, float y, float z, float w) {
    vec4 r;
    r.x = fma_(m[12], w, fma_(m[8], z, fma_(m[4], y, m[0] * x)));
    r.y = fma_(m[13], w, fma_(m[9], z, fma_(m[5], y, m[1] * x)));
    r.z = fma_(m[14], w, fma_(m[10], z, fma_(m[6], y, m[2] * x)));
    r.w = fma_(m[15], w, fma_(m[11], z, fma_(m[7], y, m[3] * x)));
    return r;
}
/* row-major 3x4 affine transform of a point / of a direction (w = 0). */
RT3_HD vec3 xform_point(const float* m, vec3 p) {
    return v3(fma_(m[2], p.z, fma_(m[1], p.y, fma_(m[0], p.x, m[3]))),
              fma_(m[6], p.z, fma_(m[5], p.y, fma_(m[4], p.x, m[7]))),
              fma_(m[10], p.z, fma_(m[9], p.y, fma_(m[8], p.x, m[11]))));
}
RT3_HD vec3 xform_dir(const float* m, vec3 d) {
    return v3(fma_(m[2], d.z, fma_(m[1], d.y, m[0] * d.x)),
              fma_(m[6], d.z, fma_(m[5], d.y, m[4] * d.x)),
              fma_(m[10], d.z, fma_(m[9], d.y, m[8] * d.x)));
}
RT3_HD bool xform_is_identity(const float* m) {
    return m[0] == 1.f && m[1] == 0.f && m[2] == 0.f && m[3] == 0.f && m[4] == 0.f && m[5] == 1.f && m[6] == 0.f &&
           m[7] == 0.f && m[8] == 0.f && m[9] == 0.f && m[10] == 1.f && m[11] == 0.f;
}

/* ---------------------------------------------------------------- RNG ------------------------ */
/* shaders/include/random.slang:5-15 (Bob Jenkins' integer hash). */
RT3_HD uint32_t hash_jenkins(uint32_t a) {
    a = (a + 0x7ed55d16u) + (a << 12);
    a = (a ^ 0xc761c23cu) ^ (a >> 19);
    a = (a + 0x165667b1u) + (a << 5);
    a = (a + 0xd3a2646cu) ^ (a << 9);
    a = (a + 0xfd7046c5u) + (a << 3);
    a = (a ^ 0xb55a4f09u) ^ (a >> 16);
    return a;
}
/* shaders/include/math.slang:105-112 */
RT3_HD uint32_t integer_explode(uint32_t x) {
    x = (x | (x << 8)) & 0x00FF00FFu;
    x = (x | (x << 4)) & 0x0F0F0F0Fu;
    x = (x | (x << 2)) & 0x33333333u;
    x = (x | (x << 1)) & 0x55555555u;
    return x;
}
/* shaders/include/math.slang:114-117 */
RT3_HD uint32_t zcurve_to_linear_index(uint32_t x, uint32_t y) { return integer_explode(x) | (integer_explode(y) << 1); }

/* shaders/include/random.slang:18-24 */
RT3_HD float radical_inverse_vdc(uint32_t bits) {
    bits = (bits << 16) | (bits >> 16);
    bits = ((bits & 0x55555555u) << 1) | ((bits & 0xAAAAAAAAu) >> 1);
    bits = ((bits & 0x33333333u) << 2) | ((bits & 0xCCCCCCCCu) >> 2);
    bits = ((bits & 0x0F0F0F0Fu) << 4) | ((bits & 0xF0F0F0F0u) >> 4);
    bits = ((bits & 0x00FF00FFu) << 8) | ((bits & 0xFF00FF00u) >> 8);
    return float(bits) * 2.3283064365386963e-10f;
}
RT3_HD float frac_(float x) { return x - floor_(x); }
/* shaders/include/random.slang:30-35 (R2 / plastic-number sequence). */
RT3_HD vec2 r2_sequence(uint32_t i) {
    const float a1 = (float)(1.0 / 1.32471795724474602596);
    const float a2 = (float)(1.0 / (1.32471795724474602596 * 1.32471795724474602596));
    vec2 r;
    r.x = frac_(a1 * float(i) + 0.5f);
    r.y = frac_(a2 * float(i) + 0.5f);
    return r;
}

/* The same sequence evaluated in 0.32 fixed point (exact for every i): frac(a*i + 0.5) with
 * a1 = 0xC13FA9A9 / 2^32, a2 = 0x91E10DA5 / 2^32, reduced to 24 bits.  Used by the blue-noise sampler,
 * where the float form above loses its fraction bits as i grows. */
RT3_HD vec2 r2_sequence_fixed(uint32_t i) {
    vec2 r;
    r.x = float((i * 0xC13FA9A9u + 0x80000000u) >> 8) * 5.9604644775390625e-08f;
    r.y = float((i * 0x91E10DA5u + 0x80000000u) >> 8) * 5.9604644775390625e-08f;
    return r;
}

/* shaders/include/random.slang:37-90 */
struct RandomSamplerState {
    uint32_t seed, index;
};
RT3_HD RandomSamplerState rng_init(uint32_t px, uint32_t py, uint32_t frame_index) {
    RandomSamplerState s;
    s.index = 0;
    s.seed = hash_jenkins(zcurve_to_linear_index(px, py)) + frame_index;
    return s;
}
RT3_HD uint32_t rotl32(uint32_t x, uint32_t r) { return (x << r) | (x >> (32 - r)); }
RT3_HD uint32_t murmur3_at(uint32_t seed, uint32_t index) {
    uint32_t k = index;
    k *= 0xcc9e2d51u;
    k = rotl32(k, 15);
    k *= 0x1b873593u;
    uint32_t h = seed ^ k;
    h = rotl32(h, 13) * 5u + 0xe6546b64u;
    h ^= 4u;
    h ^= h >> 16;
    h *= 0x85ebca6bu;
    h ^= h >> 13;
    h *= 0xc2b2ae35u;
    h ^= h >> 16;
    return h;
}
RT3_HD uint32_t rng_murmur3(RandomSamplerState& s) { return murmur3_at(s.seed, s.index++); }
RT3_HD float u32_to_unit_float(uint32_t v) { return as_float((v & 0x007FFFFFu) | 0x3F800000u) - 1.0f; }
RT3_HD float rng_uniform_float(RandomSamplerState& s) { return u32_to_unit_float(rng_murmur3(s)); }

/* ---------------------------------------------------------------- deterministic trig --------- */
/* sin and cos of 2*pi*u for u in [0,1): exact quadrant reduction (x4 is exact, x - round(x) is
 * exact), then the Cephes sinf/cosf minimax kernels on [-pi/4, pi/4].  Replaces sin(phi)/cos(phi)
 * with phi = u*TAU (shaders/include/brdf.slang:57-62); differs from a correctly rounded sin(phi)
 * by a few 1e-7, i.e. by less than the rounding of phi itself. */
RT3_HD void sincos_2pi(float u, float& s, float& c) {
    float x = u * 4.0f;
    int q = (int)(x + 0.5f);
    float r = x - (float)q;
    float a = r * 1.57079632679489661923f;
    float a2 = a * a;
    float sp = fma_(fma_(fma_(-1.9515295891e-4f, a2, 8.3321608736e-3f), a2, -1.6666654611e-1f), a2 * a, a);
    float cp = fma_(fma_(fma_(2.443315711809948e-5f, a2, -1.388731625493765e-3f), a2, 4.166664568298827e-2f), a2 * a2,
                    fma_(-0.5f, a2, 1.0f));
    switch (q & 3) {
        case 0: s = sp; c = cp; break;
        case 1: s = cp; c = -sp; break;
        case 2: s = -sp; c = -cp; break;
        default: s = -cp; c = sp; break;
    }
}
/* atan(x) for x in [0,1]: odd minimax polynomial (max error ~1e-6 rad). */
RT3_HD float atan01(float x) {
    float x2 = x * x;
    float p = fma_(fma_(fma_(fma_(fma_(-0.0117212f, x2, 0.05265332f), x2, -0.11643287f), x2, 0.19354346f), x2, -0.33262347f), x2,
                   0.99997726f);
    return p * x;
}
RT3_HD float atan2_(float y, float x) {
    float ax = abs_(x), ay = abs_(y);
    float mx = max_(ax, ay), mn = min_(ax, ay);
    float a = (mx == 0.0f) ? 0.0f : atan01(mn / mx);
    if (ay > ax) a = 1.57079632679489661923f - a;
    if (x < 0.0f) a = 3.14159265358979323846f - a;
    return (y < 0.0f) ? -a : a;
}
RT3_HD float asin_(float y) {
    float yc = clamp_(y, -1.0f, 1.0f);
    return atan2_(yc, sqrt_(max_(0.0f, fma_(-yc, yc, 1.0f))));
}
/* 2^e for integer e in [-126,127], exact. */
RT3_HD float exp2i(int e) { return as_float((uint32_t)(e + 127) << 23); }
/* log2(x), x > 0 normal: exponent + log2(m), m in [sqrt(1/2), sqrt(2)), by a degree-9 polynomial in t = m - 1 for log2(1 + t) / t
 * (Chebyshev fit in float64, max error of the float32 evaluation 4.9e-8; no division — the atanh series this replaces in round 2 cost one
 * IEEE division per call, six per post-processed pixel). */
RT3_HD float log2_(float x) {
    uint32_t b = as_uint(x);
    int e = (int)((b >> 23) & 0xFF) - 127;
    float m = as_float((b & 0x007FFFFFu) | 0x3F800000u); /* [1,2) */
    if (m > 1.41421356f) { m = m * 0.5f; e += 1; }
    const float t = m - 1.0f;
    float p = fma_(fma_(fma_(fma_(-1.102015898e-01f, t, 1.863120943e-01f), t, -1.910249740e-01f), t, 2.045752853e-01f), t, -2.396190464e-01f);
    p = fma_(fma_(fma_(fma_(fma_(p, t, 2.885688841e-01f), t, -3.606966436e-01f), t, 4.808982015e-01f), t, -7.213473320e-01f), t, 1.442695022e+00f);
    return fma_(p, t, (float)e);
}
/* 2^x for x in [-126, 126]: integer / fraction split, degree-7 Taylor polynomial of 2^f on [-0.5,0.5]
 * (truncation error < 2e-9 relative). */
RT3_HD float exp2_(float x) {
    float xc = clamp_(x, -126.0f, 126.0f);
    float fl = floor_(xc + 0.5f);
    float f = xc - fl;
    float p = fma_(fma_(fma_(fma_(fma_(fma_(1.52527338e-5f, f, 1.54035304e-4f), f, 1.33335581e-3f), f, 9.61812911e-3f), f,
                             5.55041087e-2f), f, 2.40226507e-1f), f, 6.93147181e-1f);
    p = fma_(p, f, 1.0f);
    return p * exp2i((int)fl);
}
RT3_HD float pow_(float x, float y) { return (x <= 0.0f) ? 0.0f : exp2_(y * log2_(x)); }

/* ---------------------------------------------------------------- math.slang ----------------- */
/* shaders/include/math.slang:6-12 */
RT3_HD vec2 direction_to_equirect_uv(vec3 d) {
    vec2 r;
    r.x = 0.5f + atan2_(d.z, d.x) * 0.15915494309189533577f;
    r.y = 0.5f - asin_(d.y) * 0.31830988618379067154f;
    return r;
}
/* shaders/include/math.slang:29-50 (Duff et al. 2017); columns (b1, b2, n). */
struct onb { vec3 b1, b2, n; };
RT3_HD onb build_orthonormal_basis(vec3 n) {
    onb r;
    if (n.z < 0.0f) {
        const float a = 1.0f / (1.0f - n.z);
        const float b = n.x * n.y * a;
        r.b1 = v3(1.0f - n.x * n.x * a, -b, n.x);
        r.b2 = v3(b, n.y * n.y * a - 1.0f, -n.y);
    } else {
        const float a = 1.0f / (1.0f + n.z);
        const float b = -n.x * n.y * a;
        r.b1 = v3(1.0f - n.x * n.x * a, b, -n.x);
        r.b2 = v3(b, 1.0f - n.y * n.y * a, -n.y);
    }
    r.n = n;
    return r;
}
/* mul(tangent_to_world, wi) with the float3x3 of math.slang:45-49: row i = (b1[i], b2[i], n[i]). */
RT3_HD vec3 onb_to_world(const onb& t, vec3 wi) {
    return v3(fma_(t.n.x, wi.z, fma_(t.b2.x, wi.y, t.b1.x * wi.x)), fma_(t.n.y, wi.z, fma_(t.b2.y, wi.y, t.b1.y * wi.x)),
              fma_(t.n.z, wi.z, fma_(t.b2.z, wi.y, t.b1.z * wi.x)));
}
/* shaders/include/math.slang:120-123 */
RT3_HD float luminance(vec3 c) { return fma_(c.z, 0.114f, fma_(c.y, 0.587f, c.x * 0.299f)); }

/* shaders/include/brdf.slang:56-73 — DiffuseBrdf::sample: wi in tangent space; value_over_pdf = albedo. */
RT3_HD vec3 diffuse_sample_wi(float ux, float uy) {
    float cos_theta = sqrt_(max_(0.0f, 1.0f - uy));
    float sin_theta = sqrt_(max_(0.0f, 1.0f - cos_theta * cos_theta));
    float sp, cp;
    sincos_2pi(ux, sp, cp);
    return v3(cp * sin_theta, sp * sin_theta, cos_theta);
}

/* ---------------------------------------------------------------- packing.slang -------------- */
/* shaders/include/packing.slang:2-10 */
RT3_HD float unpack_unorm(uint32_t pckd, uint32_t bits) {
    uint32_t maxv = (1u << bits) - 1u;
    return float(pckd & maxv) / float(maxv);
}
RT3_HD uint32_t pack_unorm(float val, uint32_t bits) {
    uint32_t maxv = (1u << bits) - 1u;
    return (uint32_t)(clamp_(val, 0.0f, 1.0f) * float(maxv) + 0.5f);
}
/* shaders/include/packing.slang:12-27 */
RT3_HD uint32_t pack_normal_11_10_11(vec3 n) {
    uint32_t p = 0;
    p += pack_unorm(n.x * 0.5f + 0.5f, 11);
    p += pack_unorm(n.y * 0.5f + 0.5f, 10) << 11;
    p += pack_unorm(n.z * 0.5f + 0.5f, 11) << 21;
    return p;
}
RT3_HD vec3 unpack_normal_11_10_11(uint32_t p) {
    vec3 v = v3(unpack_unorm(p, 11) * 2.0f - 1.0f, unpack_unorm(p >> 11, 10) * 2.0f - 1.0f, unpack_unorm(p >> 21, 11) * 2.0f - 1.0f);
    return normalize(v);
}
/* shaders/include/packing.slang:46-62 */
RT3_HD uint32_t pack_color_888(vec3 c) {
    uint32_t p = 0;
    p += pack_unorm(sqrt_(c.x), 8);
    p += pack_unorm(sqrt_(c.y), 8) << 8;
    p += pack_unorm(sqrt_(c.z), 8) << 16;
    return p;
}
RT3_HD vec3 unpack_color_888(uint32_t p) {
    vec3 c = v3(unpack_unorm(p, 8), unpack_unorm(p >> 8, 8), unpack_unorm(p >> 16, 8));
    return c * c;
}
/* f32tof16 / f16tof32 (packing.slang:88-97 uses the HLSL intrinsics): software IEEE round-to-nearest-even. */
RT3_HD uint32_t f32_to_f16(float f) {
    uint32_t x = as_uint(f);
    uint32_t sign = (x >> 16) & 0x8000u;
    uint32_t absx = x & 0x7FFFFFFFu;
    if (absx >= 0x7F800000u) return sign | (absx > 0x7F800000u ? 0x7E00u : 0x7C00u);
    if (absx >= 0x477FF000u) return sign | 0x7C00u; /* rounds to >= 65520 -> inf */
    if (absx < 0x33000001u) return sign;            /* <= 2^-25 -> 0 */
    int e = (int)(absx >> 23) - 127;
    uint32_t m = (absx & 0x007FFFFFu) | 0x00800000u;
    if (e < -14) { /* subnormal half */
        int shift = (-14 - e) + 13;
        uint32_t q = m >> shift;
        uint32_t rem = m & ((1u << shift) - 1u);
        uint32_t half = 1u << (shift - 1);
        if (rem > half || (rem == half && (q & 1u))) q++;
        return sign | q;
    }
    uint32_t q = ((uint32_t)(e + 15) << 10) | ((m >> 13) & 0x3FFu);
    uint32_t rem = m & 0x1FFFu;
    if (rem > 0x1000u || (rem == 0x1000u && (q & 1u))) q++;
    return sign | q;
}
RT3_HD float f16_to_f32(uint32_t h) {
    uint32_t sign = (h & 0x8000u) << 16;
    uint32_t e = (h >> 10) & 0x1Fu;
    uint32_t m = h & 0x3FFu;
    if (e == 0) {
        if (m == 0) return as_float(sign);
        float v = float(m) * 5.9604644775390625e-08f; /* 2^-24 */
        return as_float(as_uint(v) | sign);
    }
    if (e == 31) return as_float(sign | 0x7F800000u | (m << 13));
    return as_float(sign | ((e + 112u) << 23) | (m << 13));
}
RT3_HD uint32_t pack_2x16f_uint(float a, float b) { return f32_to_f16(a) | (f32_to_f16(b) << 16); }

/* shaders/include/packing.slang:99-162 (EXT_texture_shared_exponent). */
RT3_HD int floor_log2(float x) { return (int)((as_uint(x) & 0x7F800000u) >> 23) - 127; }
RT3_HD uint32_t float3_to_rgb9e5(vec3 rgb) {
    const float MAX_RGB9E5 = (511.0f / 512.0f) * 65536.0f;
    float rc = clamp_(rgb.x, 0.0f, MAX_RGB9E5);
    float gc = clamp_(rgb.y, 0.0f, MAX_RGB9E5);
    float bc = clamp_(rgb.z, 0.0f, MAX_RGB9E5);
    float maxrgb = max_(rc, max_(gc, bc));
    int fl = floor_log2(maxrgb);
    int exp_shared = (fl > -16 ? fl : -16) + 1 + 15;
    float denom = exp2i(exp_shared - 15 - 9);
    int maxm = (int)floor_(maxrgb / denom + 0.5f);
    if (maxm == 512) { denom *= 2.0f; exp_shared += 1; }
    int rm = (int)floor_(rc / denom + 0.5f);
    int gm = (int)floor_(gc / denom + 0.5f);
    int bm = (int)floor_(bc / denom + 0.5f);
    return ((uint32_t)rm << 23) | ((uint32_t)gm << 14) | ((uint32_t)bm << 5) | (uint32_t)exp_shared;
}
RT3_HD vec3 rgb9e5_to_float3(uint32_t v) {
    int exponent = (int)(v & 31u) - 15 - 9;
    float scale = exp2i(exponent);
    return v3(float((v >> 23) & 511u) * scale, float((v >> 14) & 511u) * scale, float((v >> 5) & 511u) * scale);
}

/* ---------------------------------------------------------------- gbuffer_helpers.slang ------ */
struct GbufferData {
    vec3 albedo, emissive, normal;
    float roughness, metalness;
};
struct uint4_ { uint32_t x, y, z, w; };
/* shaders/include/gbuffer_helpers.slang:22-34 */
RT3_HD uint4_ gbuffer_pack(const GbufferData& g) {
    uint4_ r;
    r.x = pack_color_888(g.albedo);
    r.y = pack_normal_11_10_11(g.normal);
    r.z = pack_2x16f_uint(sqrt_(g.roughness), g.metalness);
    r.w = float3_to_rgb9e5(g.emissive);
    return r;
}
/* shaders/include/gbuffer_helpers.slang:59-70 */
RT3_HD GbufferData gbuffer_unpack(uint4_ d) {
    GbufferData r;
    r.albedo = unpack_color_888(d.x);
    r.normal = unpack_normal_11_10_11(d.y);
    float pr = f16_to_f32(d.z & 0xFFFFu);
    r.roughness = pr * pr;
    r.metalness = f16_to_f32((d.z >> 16) & 0xFFFFu);
    r.emissive = rgb9e5_to_float3(d.w);
    return r;
}

/* shaders/include/gbuffer_helpers.slang:85-103 with PlanarViewConstants mapped onto GConst:
 * matClipToView = proj_inverse, matViewToWorld = view_inverse, viewportSize = window_size,
 * cameraDirectionOrPosition = translation column of view_inverse. */
RT3_HD vec3 view_dir(uint32_t px, uint32_t py, const rt3_params& p) {
    float ux = (float(px) + 0.5f) / p.window_size[0];
    float uy = (float(py) + 0.5f) / p.window_size[1];
    float dx = ux * 2.0f - 1.0f, dy = uy * 2.0f - 1.0f;
    vec4 target = mat4_mul(p.proj_inv, dx, dy, 1.0f, 1.0f);
    vec3 nt = normalize(v3(target.x, target.y, target.z));
    vec4 d = mat4_mul(p.view_inv, nt.x, nt.y, nt.z, 0.0f);
    return v3(d.x, d.y, d.z);
}
RT3_HD vec3 camera_position(const rt3_params& p) { return v3(p.view_inv[12], p.view_inv[13], p.view_inv[14]); }
RT3_HD rt3_ray setup_primary_ray(uint32_t px, uint32_t py, const rt3_params& p) {
    rt3_ray r;
    vec3 o = camera_position(p), d = view_dir(px, py, p);
    r.o[0] = o.x; r.o[1] = o.y; r.o[2] = o.z;
    r.d[0] = d.x; r.d[1] = d.y; r.d[2] = d.z;
    r.tmin = 0.0f;
    r.tmax = RT3_BACKGROUND_DEPTH;
    return r;
}

/* ---------------------------------------------------------------- ray / triangle -------------- */
/* The reference delegates this to the Vulkan driver (TraceRay; src/renderer/pipeline_cache/mod.rs:64-73),
 * so the arithmetic is DEFINED here: Moeller-Trumbore on (v0, e1 = v1-v0, e2 = v2-v0), no culling,
 * barycentrics (u,v) weight v1 and v2 (BuiltInTriangleIntersectionAttributes), hit iff
 * u >= 0, v >= 0, u+v <= 1 and tmin < t < tmax.  Closest hit = smallest t; equal t resolves to the
 * smaller global triangle id (a traversal-order-independent rule). */
RT3_HD bool ray_triangle(vec3 o, vec3 d, vec3 v0, vec3 e1, vec3 e2, float& t, float& u, float& v) {
    vec3 pvec = cross(d, e2);
    float det = dot(e1, pvec);
    if (det == 0.0f) return false;
    float inv = 1.0f / det;
    vec3 tvec = o - v0;
    u = dot(tvec, pvec) * inv;
    vec3 qvec = cross(tvec, e1);
    v = dot(d, qvec) * inv;
    t = dot(e2, qvec) * inv;
    return (u >= 0.0f) & (v >= 0.0f) & (u + v <= 1.0f);
}
RT3_HD bool closer_hit(float t, uint32_t prim, float best_t, uint32_t best_prim) {
    return t < best_t || (t == best_t && prim < best_prim);
}

/* ---------------------------------------------------------------- scene views ---------------- */
struct SceneView {
    const rt3_vertex* vertices;
    const uint32_t* indices;
    const rt3_geometry_info* geometries;
    /* segment table: one entry per (instance, geometry) in flattening order */
    const uint32_t* seg_first_prim; /* n_segments + 1 (prefix of triangle counts) */
    const uint32_t* seg_geometry;   /* n_segments */
    const uint32_t* seg_instance;   /* n_segments */
    const float* inst_xform;        /* n_instances x 12 (row-major 3x4) */
    uint32_t n_segments;
    /* textures: slot table */
    const uint8_t* const* tex_data; /* RGBA8 */
    const uint32_t* tex_dim;        /* 2 per slot */
    uint32_t n_textures;
    const float* srgb_lut;          /* 256 floats: sRGB byte -> linear */
    /* environment (+ importance tables built by build_env_tables: RT3_FLAG_NEE) */
    const float* env; uint32_t env_w, env_h;
    const float* env_func;          /* w*h: luminance * sin(theta_row)                  */
    const float* env_marginal;      /* h+1: CDF over rows                                */
    const float* env_conditional;   /* h*(w+1): CDF over columns inside each row         */
    float env_total;                /* sum of env_func                                   */
    /* blue noise */
    const uint8_t* bn; uint32_t bn_w, bn_h;
    /* optional per-triangle shading records (built by rt3_accel_build, NULL in the oracle): (global vertex index x 3, segment)
     * indexed by the global triangle id — replaces the binary search + geometry + index fetches in front of the vertex fetch by
     * one 16-byte load; the values fetched afterwards, and so the results, are the same */
    const uint32_t* shade_rec;
    /* optional 64-byte shading records (rt3_accel_build, when the table fits the L2: <= 96 MB): the three vertex normals, the three
     * texture coordinates and the segment of a triangle, COPIED next to each other — one gather of two sectors in place of the 16-byte
     * record above followed by three dependent 32-byte vertex fetches.  Same floats, same results. */
    const float* shade64;
};

RT3_HD uint32_t find_segment(const SceneView& s, uint32_t prim) {
    uint32_t lo = 0, hi = s.n_segments; /* seg_first_prim[lo] <= prim < seg_first_prim[hi] */
    while (hi - lo > 1) {
        uint32_t mid = (lo + hi) >> 1;
        if (s.seg_first_prim[mid] <= prim) lo = mid; else hi = mid;
    }
    return lo;
}

/* World-space triangle of (segment, primitive): geometry transform then instance transform. */
RT3_HD void world_triangle(const SceneView& s, uint32_t seg, uint32_t prim_in_geo, vec3& a, vec3& b, vec3& c) {
    const rt3_geometry_info& g = s.geometries[s.seg_geometry[seg]];
    uint32_t io = g.index_offset + 3u * prim_in_geo;
    uint32_t i0 = g.vertex_offset + s.indices[io], i1 = g.vertex_offset + s.indices[io + 1], i2 = g.vertex_offset + s.indices[io + 2];
    a = v3(s.vertices[i0].p); b = v3(s.vertices[i1].p); c = v3(s.vertices[i2].p);
    if (!xform_is_identity(g.xform)) { a = xform_point(g.xform, a); b = xform_point(g.xform, b); c = xform_point(g.xform, c); }
    const float* ix = s.inst_xform + 12u * s.seg_instance[seg];
    if (!xform_is_identity(ix)) { a = xform_point(ix, a); b = xform_point(ix, b); c = xform_point(ix, c); }
}

/* Repeat addressing of the two texel columns (rows) of a bilinear footprint: x0 = floor(frac(u) * n - 0.5) lies in [-1, n - 1] and
 * x1 = x0 + 1, so (x mod n + n) mod n is a conditional wrap; the general formula — four integer divisions per footprint axis pair, two
 * thirds of the instructions of the whole lookup — is kept for coordinates outside that range (non-finite uv).  Same result for every input. */
RT3_HD void wrap_texel_pair(int& x0, int& x1, int n) {
    if (x0 >= -1 && x0 < n) {
        x0 = x0 < 0 ? n - 1 : x0;
        x1 = x1 >= n ? 0 : x1;
    } else {
        x0 = (x0 % n + n) % n; x1 = (x1 % n + n) % n;
    }
}
/* Textures[i].SampleLevel(uv, 0) as DEFINED for the missing bindless_descriptor.slang: RGBA8 sRGB,
 * bilinear, repeat addressing, texel centres at (i + 0.5) / size. */
RT3_HD vec3 sample_texture(const SceneView& s, int slot, float u, float v) {
    if (slot < 0 || (uint32_t)slot >= s.n_textures || s.tex_data[slot] == nullptr) return v3(1.f, 1.f, 1.f);
    uint32_t w = s.tex_dim[2 * slot], h = s.tex_dim[2 * slot + 1];
    const uint8_t* t = s.tex_data[slot];
    float fx = frac_(u) * float(w) - 0.5f, fy = frac_(v) * float(h) - 0.5f;
    float flx = floor_(fx), fly = floor_(fy);
    float ax = fx - flx, ay = fy - fly;
    int x0 = (int)flx, y0 = (int)fly;
    int x1 = x0 + 1, y1 = y0 + 1;
    wrap_texel_pair(x0, x1, (int)w); wrap_texel_pair(y0, y1, (int)h);
    const uint8_t* p00 = t + 4u * ((uint32_t)y0 * w + (uint32_t)x0);
    const uint8_t* p10 = t + 4u * ((uint32_t)y0 * w + (uint32_t)x1);
    const uint8_t* p01 = t + 4u * ((uint32_t)y1 * w + (uint32_t)x0);
    const uint8_t* p11 = t + 4u * ((uint32_t)y1 * w + (uint32_t)x1);
    float w00 = (1.0f - ax) * (1.0f - ay), w10 = ax * (1.0f - ay), w01 = (1.0f - ax) * ay, w11 = ax * ay;
    vec3 r;
    r.x = fma_(s.srgb_lut[p11[0]], w11, fma_(s.srgb_lut[p01[0]], w01, fma_(s.srgb_lut[p10[0]], w10, s.srgb_lut[p00[0]] * w00)));
    r.y = fma_(s.srgb_lut[p11[1]], w11, fma_(s.srgb_lut[p01[1]], w01, fma_(s.srgb_lut[p10[1]], w10, s.srgb_lut[p00[1]] * w00)));
    r.z = fma_(s.srgb_lut[p11[2]], w11, fma_(s.srgb_lut[p01[2]], w01, fma_(s.srgb_lut[p10[2]], w10, s.srgb_lut[p00[2]] * w00)));
    return r;
}

/* Skybox.SampleLevel(direction_to_equirect_uv(d), 0) — shaders/old/refrence_mode.slang:38 (commented
 * out there), shaders/old/postprocess.slang:101-105.  RGB float equirect image, bilinear, repeat-u, clamp-v. */
RT3_HD vec3 sample_environment(const SceneView& s, vec3 dir) {
    if (s.env == nullptr) return v3(0.f, 0.f, 0.f);
    vec2 uv = direction_to_equirect_uv(dir);
    float fx = uv.x * float(s.env_w) - 0.5f, fy = uv.y * float(s.env_h) - 0.5f;
    float flx = floor_(fx), fly = floor_(fy);
    float ax = fx - flx, ay = fy - fly;
    int x0 = (int)flx, y0 = (int)fly, x1 = x0 + 1, y1 = y0 + 1;
    int w = (int)s.env_w, h = (int)s.env_h;
    wrap_texel_pair(x0, x1, w);
    y0 = y0 < 0 ? 0 : (y0 > h - 1 ? h - 1 : y0);
    y1 = y1 < 0 ? 0 : (y1 > h - 1 ? h - 1 : y1);
    const float* p00 = s.env + 3u * ((uint32_t)y0 * s.env_w + (uint32_t)x0);
    const float* p10 = s.env + 3u * ((uint32_t)y0 * s.env_w + (uint32_t)x1);
    const float* p01 = s.env + 3u * ((uint32_t)y1 * s.env_w + (uint32_t)x0);
    const float* p11 = s.env + 3u * ((uint32_t)y1 * s.env_w + (uint32_t)x1);
    float w00 = (1.0f - ax) * (1.0f - ay), w10 = ax * (1.0f - ay), w01 = (1.0f - ax) * ay, w11 = ax * ay;
    vec3 r;
    r.x = fma_(p11[0], w11, fma_(p01[0], w01, fma_(p10[0], w10, p00[0] * w00)));
    r.y = fma_(p11[1], w11, fma_(p01[1], w01, fma_(p10[1], w10, p00[1] * w00)));
    r.z = fma_(p11[2], w11, fma_(p01[2], w01, fma_(p10[2], w10, p00[2] * w00)));
    return r;
}

/* shaders/include/hit_logic.slang:5-40.  `prim` is the global triangle id (the payload's
 * geometryIndex/primitiveId pair, default_hit.slang:7-8, is recovered through the segment table).
 * The per-vertex colour the old shader reads (hit_logic.slang:27) does not exist in the 32-byte
 * Vertex (assets/mod.rs:127-133) and is taken as (1,1,1). */
RT3_HD GbufferData hit_info(const SceneView& s, uint32_t prim, float bu, float bv, uint32_t* out_geometry = nullptr) {
    uint32_t seg, i0 = 0, i1 = 0, i2 = 0;
    float an[3], bn[3], cn[3], at[2], bt[2], ct[2];                /* normals and texture coordinates of the three vertices */
    bool have_attributes = false;
#ifndef RT3_SHADE64
#define RT3_SHADE64 1          /* compile-time A/B switch of the 64-byte shading records (0: the gather chain of round 1) */
#endif
    if (RT3_SHADE64 && s.shade64 != nullptr) {
        const float* r = s.shade64 + 16u * (size_t)prim;
#ifdef __CUDA_ARCH__
        const float4 q0 = __ldg((const float4*)r), q1 = __ldg((const float4*)r + 1), q2 = __ldg((const float4*)r + 2), q3 = __ldg((const float4*)r + 3);
        an[0] = q0.x; an[1] = q0.y; an[2] = q0.z; bn[0] = q0.w; bn[1] = q1.x; bn[2] = q1.y; cn[0] = q1.z; cn[1] = q1.w; cn[2] = q2.x;
        at[0] = q2.y; at[1] = q2.z; bt[0] = q2.w; bt[1] = q3.x; ct[0] = q3.y; ct[1] = q3.z; seg = as_uint(q3.w);
#else
        an[0] = r[0]; an[1] = r[1]; an[2] = r[2]; bn[0] = r[3]; bn[1] = r[4]; bn[2] = r[5]; cn[0] = r[6]; cn[1] = r[7]; cn[2] = r[8];
        at[0] = r[9]; at[1] = r[10]; bt[0] = r[11]; bt[1] = r[12]; ct[0] = r[13]; ct[1] = r[14]; seg = as_uint(r[15]);
#endif
        have_attributes = true;
    } else if (s.shade_rec != nullptr) {
        const uint32_t* r = s.shade_rec + 4u * (size_t)prim;
#ifdef __CUDA_ARCH__
        const uint4 q = __ldg((const uint4*)r);
        i0 = q.x; i1 = q.y; i2 = q.z; seg = q.w;
#else
        i0 = r[0]; i1 = r[1]; i2 = r[2]; seg = r[3];
#endif
    } else {
        seg = find_segment(s, prim);
        const rt3_geometry_info& g0 = s.geometries[s.seg_geometry[seg]];
        uint32_t io = g0.index_offset + 3u * (prim - s.seg_first_prim[seg]);
        i0 = g0.vertex_offset + s.indices[io]; i1 = g0.vertex_offset + s.indices[io + 1]; i2 = g0.vertex_offset + s.indices[io + 2];
    }
    uint32_t gi = s.seg_geometry[seg];
    if (out_geometry) *out_geometry = gi;
    const rt3_geometry_info& g = s.geometries[gi];
    if (!have_attributes) {
        const rt3_vertex& a = s.vertices[i0];
        const rt3_vertex& b = s.vertices[i1];
        const rt3_vertex& c = s.vertices[i2];
        for (int k = 0; k < 3; ++k) { an[k] = a.n[k]; bn[k] = b.n[k]; cn[k] = c.n[k]; }
        at[0] = a.t[0]; at[1] = a.t[1]; bt[0] = b.t[0]; bt[1] = b.t[1]; ct[0] = c.t[0]; ct[1] = c.t[1];
    }
    float w0 = 1.0f - bu - bv, w1 = bu, w2 = bv;
    vec3 n = v3(fma_(cn[0], w2, fma_(bn[0], w1, an[0] * w0)), fma_(cn[1], w2, fma_(bn[1], w1, an[1] * w0)),
                fma_(cn[2], w2, fma_(bn[2], w1, an[2] * w0)));
    n = normalize(n);
    n = normalize(xform_dir(g.xform, n));
    const float* ix = s.inst_xform + 12u * s.seg_instance[seg];
    if (!xform_is_identity(ix)) n = normalize(xform_dir(ix, n));
    float tu = fma_(ct[0], w2, fma_(bt[0], w1, at[0] * w0));
    float tv = fma_(ct[1], w2, fma_(bt[1], w1, at[1] * w0));
    vec3 color = v3(g.base_color[0], g.base_color[1], g.base_color[2]);
    if (g.base_color_tex > -1) color = color * sample_texture(s, g.base_color_tex, tu, tv);
    GbufferData o;
    o.metalness = g.metallic;
    o.roughness = g.roughness;
    o.emissive = v3(g.emission[0] * 12.0f, g.emission[1] * 12.0f, g.emission[2] * 12.0f);
    o.normal = n;
    o.albedo = color;
    return o;
}

/* ---------------------------------------------------------------- environment importance sampling
 * No reference counterpart (the reference has no importance sampling, no CDF, no shadow rays); this
 * is the north star's "skybox2.exr environment importance lookup" + any-hit shadow rays, enabled by RT3_FLAG_NEE.
 * Piecewise-constant density over texels proportional to luminance * sin(theta), sampled by two binary searches. */
RT3_HD vec3 equirect_uv_to_direction(float u, float v) {   /* inverse of direction_to_equirect_uv (math.slang:6-12) */
    float s2, c2, se, ce;
    sincos_2pi(frac_(u), s2, c2);                          /* phi = 2 pi u - pi: cos(phi) = -cos(2 pi u), sin(phi) = -sin(2 pi u) */
    sincos_2pi(clamp_(v, 0.0f, 1.0f) * 0.5f, se, ce);      /* elevation = (0.5 - v) pi: sin(el) = cos(pi v), cos(el) = sin(pi v) */
    return v3(-c2 * se, ce, -s2 * se);
}
RT3_HD uint32_t cdf_find(const float* cdf, uint32_t n, float u) {   /* largest k in [0,n) with cdf[k] <= u; cdf has n+1 entries */
    uint32_t lo = 0, hi = n;
    while (hi - lo > 1) {
        uint32_t mid = (lo + hi) >> 1;
        if (cdf[mid] <= u) lo = mid; else hi = mid;
    }
    return lo;
}
/* pdf with respect to solid angle of the texel density at texel (i, j), for a direction whose |cos(elevation)| is sin_theta */
RT3_HD float env_pdf_texel(const SceneView& s, uint32_t i, uint32_t j, float sin_theta) {
    if (!(s.env_total > 0.0f)) return 0.0f;
    float f = s.env_func[(size_t)j * s.env_w + i];
    return (f / s.env_total) * (float(s.env_w) * float(s.env_h)) / (19.7392088021787172f * max_(sin_theta, 1e-6f));   /* 2 pi^2 */
}
RT3_HD float env_pdf(const SceneView& s, vec3 dir) {
    if (s.env_func == nullptr) return 0.0f;
    vec2 uv = direction_to_equirect_uv(dir);
    uint32_t i = (uint32_t)min_(max_(uv.x * float(s.env_w), 0.0f), float(s.env_w - 1));
    uint32_t j = (uint32_t)min_(max_(uv.y * float(s.env_h), 0.0f), float(s.env_h - 1));
    return env_pdf_texel(s, i, j, sqrt_(max_(0.0f, fma_(-dir.y, dir.y, 1.0f))));
}
RT3_HD vec3 env_sample(const SceneView& s, float u1, float u2, float& pdf) {
    pdf = 0.0f;
    if (s.env_func == nullptr || !(s.env_total > 0.0f)) return v3(0.f, 1.f, 0.f);
    uint32_t j = cdf_find(s.env_marginal, s.env_h, u2);
    float m0 = s.env_marginal[j], m1 = s.env_marginal[j + 1];
    float dv = (m1 > m0) ? (u2 - m0) / (m1 - m0) : 0.5f;
    const float* row = s.env_conditional + (size_t)j * (s.env_w + 1);
    uint32_t i = cdf_find(row, s.env_w, u1);
    float c0 = row[i], c1 = row[i + 1];
    float du = (c1 > c0) ? (u1 - c0) / (c1 - c0) : 0.5f;
    float u = (float(i) + clamp_(du, 0.0f, 0.99999994f)) / float(s.env_w);
    float v = (float(j) + clamp_(dv, 0.0f, 0.99999994f)) / float(s.env_h);
    vec3 d = equirect_uv_to_direction(u, v);
    pdf = env_pdf_texel(s, i, j, sqrt_(max_(0.0f, fma_(-d.y, d.y, 1.0f))));
    return d;
}
/* Host-side table build (double accumulation, rounded once); identical code runs in librt3 and in the oracle. */
inline void build_env_tables(const float* rgb, uint32_t w, uint32_t h, float* func, float* marginal, float* conditional, float* total) {
    double sum = 0.0;
    for (uint32_t j = 0; j < h; ++j) {
        double st = sin(3.14159265358979323846 * (j + 0.5) / h);
        for (uint32_t i = 0; i < w; ++i) {
            const float* p = rgb + 3 * ((size_t)j * w + i);
            double lum = 0.299 * p[0] + 0.587 * p[1] + 0.114 * p[2];
            float f = (float)((lum > 0 ? lum : 0) * st);
            func[(size_t)j * w + i] = f;
            sum += f;
        }
    }
    *total = (float)sum;
    double acc = 0.0;
    for (uint32_t j = 0; j < h; ++j) {
        double rs = 0.0;
        for (uint32_t i = 0; i < w; ++i) rs += func[(size_t)j * w + i];
        marginal[j] = sum > 0 ? (float)(acc / sum) : (float)j / h;
        float* row = conditional + (size_t)j * (w + 1);
        double ra = 0.0;
        for (uint32_t i = 0; i < w; ++i) { row[i] = rs > 0 ? (float)(ra / rs) : (float)i / w; ra += func[(size_t)j * w + i]; }
        row[w] = 1.0f;
        acc += rs;
    }
    marginal[h] = 1.0f;
}

/* ---------------------------------------------------------------- samplers ------------------- */
/* The 2 random numbers of bounce `dim_pair` (refrence_mode.slang:43: x drawn first).  Faithful mode:
 * the murmur3 counter stream.  RT3_FLAG_BLUENOISE (no reference counterpart; resources/bluenoise.png
 * ships unused): blue-noise tile value of the pixel (channel pair by dimension) Cranley-Patterson
 * rotated by the R2 sequence (random.slang:30-35) indexed by the running draw counter and frame. */
RT3_HD vec2 sample_2d(const SceneView& s, uint32_t flags, uint32_t px, uint32_t py, uint32_t frame, RandomSamplerState& rng) {
    vec2 u;
    if ((flags & RT3_FLAG_BLUENOISE) && s.bn != nullptr) {
        uint32_t k = rng.index >> 1; /* 2-D draw number within the pixel's stream */
        rng.index += 2;
        /* decorrelate successive dimensions by shifting the tile with a per-dimension Jenkins hash */
        uint32_t h = hash_jenkins(k * 0x9E3779B9u + 0x85EBCA6Bu);
        uint32_t tx = (px + (h & 0xFFFFu)) & (s.bn_w - 1u), ty = (py + (h >> 16)) & (s.bn_h - 1u);
        const uint8_t* t = s.bn + 4u * (ty * s.bn_w + tx);
        uint32_t ch = (k & 1u) * 2u;
        vec2 r = r2_sequence_fixed(frame * 64u + k);
        u.x = frac_((float(t[ch]) + 0.5f) * (1.0f / 256.0f) + r.x);
        u.y = frac_((float(t[ch + 1]) + 0.5f) * (1.0f / 256.0f) + r.y);
        /* keep strictly below 1 */
        u.x = min_(u.x, 0.99999994f); u.y = min_(u.y, 0.99999994f);
    } else {
        u.x = rng_uniform_float(rng);
        u.y = rng_uniform_float(rng);
    }
    return u;
}

/* ---------------------------------------------------------------- brdf.slang: specular lobe ---
 * shaders/include/brdf.slang:95-322 restated: Schlick Fresnel, height-correlated Smith masking, GGX NDF, VNDF sampling
 * (Heitz 2018 via Falcor, USE_GGX_VNDF_SAMPLING = 1, USE_GGX_CORRELATED_MASKING = 1).  All quantities are in tangent
 * space (z = normal) and "defined wrt the projected solid angle metric" (brdf.slang:30) exactly as in the reference.
 * pow(x, 5) is DEFINED as ((x*x)*(x*x))*x; sin/cos(2 pi u) are the shared deterministic kernels. */
static const float BRDF_SAMPLING_MIN_COS = 1e-5f;                                /* brdf.slang:8 */
static const float RT3_PI = 3.14159265358979323846f;

RT3_HD vec3 eval_fresnel_schlick(vec3 f0, vec3 f90, float cos_theta) {           /* brdf.slang:95-97 */
    float x = max_(0.0f, 1.0f - cos_theta);
    float x2 = x * x;
    float x5 = (x2 * x2) * x;
    return v3(f0.x + (f90.x - f0.x) * x5, f0.y + (f90.y - f0.y) * x5, f0.z + (f90.z - f0.z) * x5);
}
RT3_HD float g_smith_ggx_correlated(float ndotv, float ndotl, float a2) {        /* brdf.slang:108-113 */
    float lambda_v = ndotl * sqrt_((-ndotv * a2 + ndotv) * ndotv + a2);
    float lambda_l = ndotv * sqrt_((-ndotl * a2 + ndotl) * ndotl + a2);
    return 2.0f * ndotl * ndotv / (lambda_v + lambda_l);
}
RT3_HD float g_smith_ggx1(float ndotv, float a2) {                               /* brdf.slang:115-118 */
    float tan2_v = (1.0f - ndotv * ndotv) / (ndotv * ndotv);
    return 2.0f / (1.0f + sqrt_(1.0f + a2 * tan2_v));
}
struct SmithShadowingMasking { float g, g_over_g1_wo; };
RT3_HD SmithShadowingMasking smith_eval(float ndotv, float ndotl, float a2) {    /* brdf.slang:128-139 */
    SmithShadowingMasking r;
    r.g = g_smith_ggx_correlated(ndotv, ndotl, a2);
    r.g_over_g1_wo = r.g / g_smith_ggx1(ndotv, a2);
    return r;
}
RT3_HD float ggx_ndf(float a2, float cos_theta) {                                /* brdf.slang:147-150 */
    float denom_sqrt = cos_theta * cos_theta * (a2 - 1.0f) + 1.0f;
    return a2 / (RT3_PI * denom_sqrt * denom_sqrt);
}
RT3_HD float pdf_ggx_vn(float a2, vec3 wo, vec3 h) {                             /* brdf.slang:162-166 */
    float g1 = g_smith_ggx1(wo.z, a2);
    float d = ggx_ndf(a2, h.z);
    return g1 * d * max_(0.0f, dot(wo, h)) / wo.z;
}
struct NdfSample { vec3 m; float pdf; };
RT3_HD NdfSample sample_vndf(float alpha, vec3 wo, vec2 urand) {                 /* brdf.slang:186-218 */
    float a2 = alpha * alpha;
    vec3 Vh = normalize(v3(alpha * wo.x, alpha * wo.y, wo.z));
    vec3 T1 = (Vh.z < 0.9999f) ? normalize(cross(v3(0.f, 0.f, 1.f), Vh)) : v3(1.f, 0.f, 0.f);
    vec3 T2 = cross(Vh, T1);
    float r = sqrt_(urand.x);
    float sp, cp;
    sincos_2pi(urand.y, sp, cp);
    float t1 = r * cp;
    float t2 = r * sp;
    float sc = 0.5f * (1.0f + Vh.z);
    t2 = (1.0f - sc) * sqrt_(max_(0.0f, 1.0f - t1 * t1)) + sc * t2;
    float nz = sqrt_(max_(0.0f, 1.0f - t1 * t1 - t2 * t2));
    vec3 Nh = v3(t1 * T1.x + t2 * T2.x + nz * Vh.x, t1 * T1.y + t2 * T2.y + nz * Vh.y, t1 * T1.z + t2 * T2.z + nz * Vh.z);
    vec3 h = normalize(v3(alpha * Nh.x, alpha * Nh.y, max_(0.0f, Nh.z)));
    NdfSample res;
    res.m = h;
    res.pdf = pdf_ggx_vn(a2, wo, h);
    return res;
}
/* BrdfValue / BrdfSample of brdf.slang:14-50 */
struct BrdfValue { vec3 value_over_pdf, value; float pdf; vec3 transmission_fraction; };
struct BrdfSample { vec3 value_over_pdf, value; float pdf; vec3 transmission_fraction; vec3 wi; float approx_roughness; };
RT3_HD BrdfSample brdf_sample_invalid() {                                        /* brdf.slang:36-44 */
    BrdfSample r;
    r.value_over_pdf = v3(0.f, 0.f, 0.f); r.value = v3(0.f, 0.f, 0.f); r.pdf = 0.0f; r.wi = v3(0.f, 0.f, -1.f);
    r.transmission_fraction = v3(0.f, 0.f, 0.f); r.approx_roughness = 0.0f;
    return r;
}
RT3_HD bool brdf_sample_is_valid(const BrdfSample& s) { return s.wi.z > 1e-6f; } /* brdf.slang:46-48 */
RT3_HD vec3 reflect_(vec3 i, vec3 n) { float k = 2.0f * dot(i, n); return v3(i.x - k * n.x, i.y - k * n.y, i.z - k * n.z); }

RT3_HD BrdfSample specular_sample(float roughness, vec3 albedo, vec3 wo, vec2 urand) {   /* SpecularBrdf::sample, brdf.slang:220-267 */
    NdfSample ndf = sample_vndf(roughness, wo, urand);
    vec3 wi = reflect_(v3(-wo.x, -wo.y, -wo.z), ndf.m);
    if (ndf.m.z <= BRDF_SAMPLING_MIN_COS || wi.z <= BRDF_SAMPLING_MIN_COS || wo.z <= BRDF_SAMPLING_MIN_COS) return brdf_sample_invalid();
    float jacobian = 1.0f / (4.0f * dot(wi, ndf.m));
    vec3 fresnel = eval_fresnel_schlick(albedo, v3(1.f, 1.f, 1.f), dot(ndf.m, wi));
    float a2 = roughness * roughness;
    float cos_theta = ndf.m.z;
    SmithShadowingMasking sm = smith_eval(wo.z, wi.z, a2);
    BrdfSample res;
    res.pdf = ndf.pdf * jacobian / wi.z;
    res.wi = wi;
    res.transmission_fraction = v3(1.0f - fresnel.x, 1.0f - fresnel.y, 1.0f - fresnel.z);
    res.approx_roughness = roughness;
    res.value_over_pdf = fresnel * sm.g_over_g1_wo;
    float k = sm.g * ggx_ndf(a2, cos_theta) / (4.0f * wo.z * wi.z);
    res.value = fresnel * k;
    return res;
}
RT3_HD BrdfValue specular_evaluate(float roughness, vec3 albedo, vec3 wo, vec3 wi) {     /* SpecularBrdf::evaluate, brdf.slang:269-314 */
    BrdfValue res;
    res.value_over_pdf = v3(0.f, 0.f, 0.f); res.value = v3(0.f, 0.f, 0.f); res.pdf = 0.0f; res.transmission_fraction = v3(0.f, 0.f, 0.f);
    if (wi.z <= 0.0f || wo.z <= 0.0f) return res;
    float a2 = roughness * roughness;
    vec3 m = normalize(wo + wi);
    float cos_theta = m.z;
    float pdf_h = pdf_ggx_vn(a2, wo, m);
    float jacobian = 1.0f / (4.0f * dot(wi, m));
    vec3 fresnel = eval_fresnel_schlick(albedo, v3(1.f, 1.f, 1.f), dot(m, wi));
    SmithShadowingMasking sm = smith_eval(wo.z, wi.z, a2);
    res.pdf = pdf_h * jacobian / wi.z;
    res.transmission_fraction = v3(1.0f - fresnel.x, 1.0f - fresnel.y, 1.0f - fresnel.z);
    res.value_over_pdf = fresnel * sm.g_over_g1_wo;
    float k = sm.g * ggx_ndf(a2, cos_theta) / (4.0f * wo.z * wi.z);
    res.value = fresnel * k;
    return res;
}
RT3_HD vec3 specular_dominant_direction(vec3 n, vec3 v, float roughness) {       /* brdf.slang:318-322 */
    vec3 r = reflect_(v3(-v.x, -v.y, -v.z), n);
    float f = (1.0f - roughness) * (sqrt_(1.0f - roughness) + roughness);
    return normalize(v3(n.x + (r.x - n.x) * f, n.y + (r.y - n.y) * f, n.z + (r.z - n.z) * f));
}

/* ---- layered material (RT3_FLAG_SPECULAR) -----------------------------------------------------------------------------
 * The reference's path tracer samples DiffuseBrdf only (refrence_mode.slang:44-51); its brdf.slang also carries the
 * SpecularBrdf above, and roughness / metalness already flow through hit_info and the G-buffer, but no shader in the tree
 * combines the two lobes.  The combination is therefore DEFINED here (metalness workflow, one-sample lobe selection):
 *   f0 = lerp(0.04, albedo, metalness)        specular reflectance at normal incidence (SpecularBrdf.albedo)
 *   kd = albedo * (1 - metalness)             DiffuseBrdf.albedo
 *   alpha = max(roughness, 0.001)             SpecularBrdf.roughness (GbufferData.roughness is already alpha:
 *                                             gbuffer_helpers.slang:27,64 store its square root as "perceptual")
 *   p_spec = lum(f0) / (lum(f0) + lum(kd)), 0 when the surface is seen from behind (wo.z <= MIN_COS)
 * urand.x selects the lobe and is rescaled to [0,1) inside the chosen one; the path weight is the chosen lobe's
 * value_over_pdf / p_lobe; an invalid specular sample (brdf.slang:226-228) ends the path. */
struct LayeredMaterial {
    onb t;
    vec3 wo;            /* tangent-space direction towards the previous vertex */
    vec3 f0, kd;
    float alpha, p_spec;
};
RT3_HD vec3 world_to_onb(const onb& t, vec3 w) { return v3(dot(w, t.b1), dot(w, t.b2), dot(w, t.n)); }
RT3_HD LayeredMaterial layered_setup(const GbufferData& surface, vec3 ray_d) {
    LayeredMaterial m;
    m.t = build_orthonormal_basis(surface.normal);
    m.wo = world_to_onb(m.t, v3(-ray_d.x, -ray_d.y, -ray_d.z));
    float mt = clamp_(surface.metalness, 0.0f, 1.0f);
    m.f0 = v3(0.04f + (surface.albedo.x - 0.04f) * mt, 0.04f + (surface.albedo.y - 0.04f) * mt, 0.04f + (surface.albedo.z - 0.04f) * mt);
    m.kd = surface.albedo * (1.0f - mt);
    m.alpha = max_(surface.roughness, 0.001f);
    float ls = luminance(m.f0), ld = luminance(m.kd);
    m.p_spec = (m.wo.z > BRDF_SAMPLING_MIN_COS && (ls + ld) > 0.0f) ? ls / (ls + ld) : 0.0f;
    return m;
}
/* BRDF value (projected-solid-angle convention: multiply by cos for the rendering equation) and solid-angle pdf of the
 * lobe mixture for a tangent-space direction wi. */
RT3_HD void layered_evaluate(const LayeredMaterial& m, vec3 wi, vec3& value, float& pdf_solid) {
    value = v3(0.f, 0.f, 0.f); pdf_solid = 0.0f;
    if (wi.z <= 0.0f) return;
    const float inv_pi = 0.31830988618379067154f;
    value = m.kd * inv_pi;
    pdf_solid = (1.0f - m.p_spec) * (wi.z * inv_pi);
    if (m.p_spec > 0.0f) {
        BrdfValue sv = specular_evaluate(m.alpha, m.f0, m.wo, wi);
        value = value + sv.value;
        pdf_solid = pdf_solid + m.p_spec * (sv.pdf * wi.z);
    }
}
struct LayeredSample { vec3 wi; vec3 weight; bool valid; bool specular; };
RT3_HD LayeredSample layered_sample(const LayeredMaterial& m, vec2 urand) {
    LayeredSample r;
    r.valid = true;
    if (urand.x < m.p_spec) {
        vec2 u; u.x = min_(urand.x / m.p_spec, 0.99999994f); u.y = urand.y;
        BrdfSample sp = specular_sample(m.alpha, m.f0, m.wo, u);
        r.specular = true;
        r.valid = brdf_sample_is_valid(sp);
        r.wi = sp.wi;
        r.weight = sp.value_over_pdf * (1.0f / m.p_spec);
    } else {
        float q = 1.0f - m.p_spec;
        float ux = min_((urand.x - m.p_spec) / q, 0.99999994f);
        r.specular = false;
        r.wi = diffuse_sample_wi(ux, urand.y);
        r.weight = m.kd * (1.0f / q);
    }
    return r;
}

/* ---------------------------------------------------------------- path tracing step ---------- */
/* One bounce of shaders/old/refrence_mode.slang:42-51: sample the diffuse lobe at `surface`,
 * advance the ray to the hit point, gather emission, update throughput. */
struct PathVertexOut { vec3 origin, dir; };
RT3_HD PathVertexOut bounce_step(const GbufferData& surface, vec3 ray_o, vec3 ray_d, float depth, vec2 urand, vec3& radiance,
                                 vec3& ray_color, float* out_pdf = nullptr) {
    onb t = build_orthonormal_basis(surface.normal);
    vec3 wi = diffuse_sample_wi(urand.x, urand.y);
    PathVertexOut o;
    o.origin = v3(ray_o.x + depth * ray_d.x, ray_o.y + depth * ray_d.y, ray_o.z + depth * ray_d.z);
    o.dir = onb_to_world(t, wi);
    radiance = radiance + ray_color * surface.emissive;
    ray_color = ray_color * surface.albedo;
    if (out_pdf) *out_pdf = wi.z * 0.31830988618379067154f;   /* cos(theta) / pi: solid-angle pdf of the cosine lobe */
    return o;
}

/* RT3_FLAG_SPECULAR variant of bounce_step: same bookkeeping, layered material instead of the diffuse lobe.  Returns
 * false in `alive` when the sampled lobe yields no direction (the path ends after its emission has been gathered). */
RT3_HD PathVertexOut bounce_step_layered(const GbufferData& surface, vec3 ray_o, vec3 ray_d, float depth, vec2 urand, vec3& radiance,
                                         vec3& ray_color, LayeredMaterial& m, bool& alive, float* out_pdf = nullptr) {
    m = layered_setup(surface, ray_d);
    LayeredSample ls = layered_sample(m, urand);
    PathVertexOut o;
    o.origin = v3(ray_o.x + depth * ray_d.x, ray_o.y + depth * ray_d.y, ray_o.z + depth * ray_d.z);
    o.dir = onb_to_world(m.t, ls.wi);
    radiance = radiance + ray_color * surface.emissive;
    alive = ls.valid;
    if (out_pdf) {
        vec3 val; float pdf = 0.0f;
        if (ls.valid) layered_evaluate(m, ls.wi, val, pdf);
        *out_pdf = pdf;
    }
    ray_color = ls.valid ? ray_color * ls.weight : v3(0.f, 0.f, 0.f);
    return o;
}

/* ---- next-event estimation of the environment (RT3_FLAG_NEE; needs RT3_FLAG_ENV and the importance tables) ----------
 * At every vertex from which a BRDF ray is traced, one environment direction is drawn from the luminance tables and a
 * shadow (any-hit) ray is cast; the BRDF-sampled escape and the light sample are combined with the balance heuristic. */
RT3_HD bool nee_enabled(uint32_t flags) { return (flags & (RT3_FLAG_NEE | RT3_FLAG_ENV)) == (RT3_FLAG_NEE | RT3_FLAG_ENV); }
struct NeeSample { vec3 dir, contribution; bool valid; };
/* `throughput_after` = ray_color AFTER bounce_step (already multiplied by the albedo). */
RT3_HD NeeSample nee_sample(const SceneView& s, const GbufferData& surface, vec3 throughput_after, vec2 u) {
    NeeSample r;
    float pl;
    r.dir = env_sample(s, u.x, u.y, pl);
    float cosl = dot(surface.normal, r.dir);
    r.valid = (pl > 0.0f) && (cosl > 0.0f);
    r.contribution = v3(0.f, 0.f, 0.f);
    if (r.valid) {
        float pb = cosl * 0.31830988618379067154f;
        float w = pl / (pl + pb);
        float k = (cosl * 0.31830988618379067154f / pl) * w;
        r.contribution = throughput_after * (sample_environment(s, r.dir) * k);
    }
    return r;
}
/* Layered-material variant: `throughput_before` = ray_color BEFORE bounce_step_layered. */
RT3_HD NeeSample nee_sample_layered(const SceneView& s, const LayeredMaterial& m, vec3 throughput_before, vec2 u) {
    NeeSample r;
    float pl;
    r.dir = env_sample(s, u.x, u.y, pl);
    vec3 wl = world_to_onb(m.t, r.dir);
    r.valid = (pl > 0.0f) && (wl.z > 0.0f);
    r.contribution = v3(0.f, 0.f, 0.f);
    if (r.valid) {
        vec3 f; float pb;
        layered_evaluate(m, wl, f, pb);
        float w = pl / (pl + pb);
        float k = (wl.z / pl) * w;
        r.contribution = throughput_before * (f * sample_environment(s, r.dir)) * k;
    }
    return r;
}
/* MIS weight of an environment hit reached by BRDF sampling with solid-angle pdf `pdf_brdf` (1 when NEE is off). */
RT3_HD float escape_weight(const SceneView& s, uint32_t flags, vec3 dir, float pdf_brdf) {
    if (!nee_enabled(flags)) return 1.0f;
    float pl = env_pdf(s, dir);
    float d = pdf_brdf + pl;
    return d > 0.0f ? pdf_brdf / d : 0.0f;
}
/* shaders/old/refrence_mode.slang:59-65 */
RT3_HD vec3 temporal_blend(vec3 prev, vec3 radiance, float blendfactor) {
    if (blendfactor >= 1.0f) return radiance;
    return v3(prev.x + (radiance.x - prev.x) * blendfactor, prev.y + (radiance.y - prev.y) * blendfactor,
              prev.z + (radiance.z - prev.z) * blendfactor);
}

/* ---------------------------------------------------------------- screen-space probe GI (row F4) ----
 * shaders/old/{structured_importance_sampling,trace_probes,spherical_harmonic_conversion,interpolate_probes}.slang and
 * shaders/include/spherical_harmonics.slang.  One probe per 16x16 pixels (anchored at the tile's first pixel), 8x8 octahedral
 * directions per probe in an atlas of (W/16*8) x (H/16*8) RGBA32F texels, projected to 3 x SH-L2 per probe, interpolated per
 * pixel from the 4 surrounding probes.  The reference's probe shaders are in a debug state (trace_probes.slang:75 writes the
 * direction as a colour; the importance-sampling pass refines nothing because its culling test `brdf_pdf < 0` never holds):
 * where a line is debug output, the intended expression that stands commented beside it is the one restated, and it is
 * said so at the line. */
RT3_HD float sign_step(float v) { return (v >= 0.0f) ? 1.0f : -1.0f; }               /* step(0, v) * 2 - 1 */
RT3_HD vec2 octa_wrap(vec2 v) {                                                      /* packing.slang:64-66 */
    vec2 r; r.x = (1.0f - abs_(v.y)) * sign_step(v.x); r.y = (1.0f - abs_(v.x)) * sign_step(v.y); return r;
}
RT3_HD vec2 octa_encode(vec3 n) {                                                    /* packing.slang:68-75 */
    float k = abs_(n.x) + abs_(n.y) + abs_(n.z);
    vec2 e; e.x = n.x / k; e.y = n.y / k;
    if (n.z / k < 0.0f) e = octa_wrap(e);
    e.x = e.x * 0.5f + 0.5f; e.y = e.y * 0.5f + 0.5f;
    return e;
}
RT3_HD vec3 octa_decode(vec2 f) {                                                    /* packing.slang:77-86 */
    float fx = f.x * 2.0f - 1.0f, fy = f.y * 2.0f - 1.0f;
    vec3 n = v3(fx, fy, 1.0f - abs_(fx) - abs_(fy));
    float t = clamp_(-n.z, 0.0f, 1.0f);
    n.x = n.x - sign_step(n.x) * t;
    n.y = n.y - sign_step(n.y) * t;
    return normalize(n);
}
RT3_HD vec3 world_pos_from_depth(float depth, uint32_t px, uint32_t py, const rt3_params& p) {   /* gbuffer_helpers.slang:81-83 */
    vec3 c = camera_position(p), d = view_dir(px, py, p);
    return v3(c.x + d.x * depth, c.y + d.y * depth, c.z + d.z * depth);
}
/* sh3 = float3x3 (spherical_harmonics.slang:9), row-major c[3*row + col] */
struct sh3 { float c[9]; };
RT3_HD sh3 sh3_evaluate(vec3 d) {                                                    /* spherical_harmonics.slang:32-46 */
    sh3 r;
    r.c[0] = 0.28209479177387814347403972578039f;
    r.c[1] = -0.48860251190291992158638462283836f * d.y;
    r.c[2] = 0.48860251190291992158638462283836f * d.z;
    r.c[3] = -0.48860251190291992158638462283836f * d.x;
    r.c[4] = 1.09254843059207907054338570580268f * d.x * d.y;
    r.c[5] = 1.09254843059207907054338570580268f * d.y * d.z;
    r.c[6] = 0.31539156525252000603089369029571f * (3.0f * d.z * d.z - 1.0f);
    r.c[7] = 1.09254843059207907054338570580268f * d.x * d.z;
    r.c[8] = 0.54627421529603953527169285290134f * (d.x * d.x - d.y * d.y);
    return r;
}
RT3_HD float sh3_dot(const sh3& a, const sh3& b) {                                   /* matrix_dot, spherical_harmonics.slang:61-65 (left to right) */
    float r = a.c[0] * b.c[0];
    for (int i = 1; i < 9; ++i) r = r + a.c[i] * b.c[i];
    return r;
}
RT3_HD sh3 sh3_transform_cos_lobe(vec3 normal) {                                     /* spherical_harmonics.slang:74-91 */
    sh3 s = sh3_evaluate(normal);
    s.c[0] *= 3.1415926536f;
    s.c[1] *= 2.0943951023931954923f; s.c[2] *= 2.0943951023931954923f; s.c[3] *= 2.0943951023931954923f;
    for (int i = 4; i < 9; ++i) s.c[i] *= 0.7853981633974483096f;
    return s;
}
RT3_HD vec3 sh3_unproject_cos_lobe(const sh3& r, const sh3& g, const sh3& b, vec3 normal) {   /* spherical_harmonics.slang:106-110 */
    sh3 s = sh3_transform_cos_lobe(normal);
    return v3(sh3_dot(r, s), sh3_dot(g, s), sh3_dot(b, s));
}
/* Probe-atlas geometry for a W x H frame. */
RT3_HD uint32_t probe_count_x(const rt3_params& p) { return (uint32_t)p.window_size[0] / 16u; }
RT3_HD uint32_t probe_count_y(const rt3_params& p) { return (uint32_t)p.window_size[1] / 16u; }
/* Direction word of an atlas texel (structured_importance_sampling.slang:55-69): bit 15 = mip, low 15 bits = direction index on
 * a grid of (8 << mip)^2 cells.  probe_uniform_direction is the assignment of the shader's else-branch (:69): mip 0, index = thread
 * index inside the probe; rt3_pass_probe_trace uses it when no direction image is bound. */
RT3_HD uint32_t probe_uniform_direction(uint32_t ax, uint32_t ay) { return (ay & 7u) * 8u + (ax & 7u); }
/* ---- structured_importance_sampling.slang:19-72, the LIVE lines --------------------------------------------------------------------
 * One 8x8 thread group per probe; thread ti = ly * 8 + lx owns octahedral cell ti and 2x2 of the probe's 16x16 pixels.
 *   :25-32  normals[ti * 4 + y * 2 + x] = unpack_normal(gbuffer[(probe * 8 + (lx, ly)) * 2 + (x, y)])
 *   :35-41  brdf_pdf = sum_{i < 256, in index order} max(dot(normals[i], dir(ti)), 0) / 256,  dir(ti) = octa_decode((cell + 0.5) / 8)
 *   :44-47  pdf = brdf_pdf (the `* light_pdf` factor is commented out); pdfs[ti] = pdf; bitonic sort of pdfs, ASCENDING (math.slang:142-163)
 *   :49-52  `brdf_pdf < 0` never holds (a sum of max(., 0)): index stays -1, culled_rays stays 0
 *   :55     the test is therefore `pdfs[0] < pdf`, pdfs[0] being the MINIMUM of the probe's 64 pdfs
 *   :57-66  true (every thread whose pdf exceeds the minimum): out = 0x8000 | ti * 4   (mip 1: cell ti * 4 of the 16x16 grid)
 *   :68-70  false (the minimum-pdf thread(s)):                 out = ti              (mip 0: the thread's own cell)
 * The index array passed to the sort (pdf_directions, never initialised) is only read in the branch that cannot be taken.  What the
 * shader seems to INTEND (refine where the pdf is high, `+ index % 3 + 1`) is commented out in the reference and is not restated. */
RT3_HD vec3 probe_sis_direction(uint32_t ti) {
    vec2 f; f.x = (float(ti % 8u) + 0.5f) / 8.0f; f.y = (float(ti / 8u) + 0.5f) / 8.0f;     /* :34-35 */
    return octa_decode(f);
}
RT3_HD float probe_sis_pdf(const vec3* normals256, vec3 dir) {                             /* :37-40 */
    float pdf = 0.0f;
    for (int i = 0; i < 256; ++i) pdf = pdf + max_(dot(normals256[i], dir), 0.0f) / 256.0f;
    return pdf;
}
RT3_HD uint32_t probe_sis_word(float pdf, float min_pdf, uint32_t ti) {                    /* :55-70 with index == -1, culled_rays == 0 */
    return (min_pdf < pdf) ? (0x8000u | (ti * 4u)) : ti;
}
/* pixel whose G-buffer normal lands in normals[ti * 4 + y * 2 + x] (:25-31) */
RT3_HD void probe_sis_pixel(uint32_t pcx, uint32_t pcy, uint32_t ti, uint32_t x, uint32_t y, uint32_t& px, uint32_t& py) {
    px = (pcx * 8u + ti % 8u) * 2u + x; py = (pcy * 8u + ti / 8u) * 2u + y;
}
/* One step (k, j) of the bitonic network of math.slang:142-163 for element i of a 64-element array: returns true when i and l = i ^ j
 * are to be exchanged by the thread that owns the smaller index. */
RT3_HD bool bitonic_exchange(uint32_t i, uint32_t k, uint32_t j, float a_i, float a_l) {
    return ((i & k) == 0u) ? (a_i > a_l) : (a_i < a_l);
}

struct ProbeRay { rt3_ray ray; uint32_t target_x, target_y; bool background; };
/* trace_probes.slang:16-56: the ray of atlas texel (ax, ay). */
RT3_HD ProbeRay probe_setup_ray(const rt3_params& p, uint32_t ax, uint32_t ay, float probe_pixel_depth, uint32_t direction_word) {
    ProbeRay o;
    const uint32_t pcx = ax / 8u, pcy = ay / 8u;                                     /* :18 */
    const uint32_t px = pcx * 16u, py = pcy * 16u;                                   /* :23 */
    o.background = probe_pixel_depth == RT3_BACKGROUND_DEPTH;                        /* :28 */
    RandomSamplerState rng = rng_init(ax, ay, p.frame);                              /* :21 ray_rng */
    const uint32_t index = direction_word & 0x7FFFu, mip = direction_word >> 15;     /* :40-42 */
    const uint32_t size = (1u << mip) * 8u;
    const float dx = float(index % size), dy = float(index / size);                  /* :44 */
    const float u0 = rng_uniform_float(rng), u1 = rng_uniform_float(rng);            /* :46, x drawn first */
    vec2 f; f.x = (dx + u0) / float(size); f.y = (dy + u1) / float(size);
    const vec3 dir = octa_decode(f);
    const vec3 org = world_pos_from_depth(probe_pixel_depth, px, py, p);             /* :52 */
    o.ray.o[0] = org.x; o.ray.o[1] = org.y; o.ray.o[2] = org.z; o.ray.tmin = 0.0005f;   /* :54 */
    o.ray.d[0] = dir.x; o.ray.d[1] = dir.y; o.ray.d[2] = dir.z; o.ray.tmax = RT3_BACKGROUND_DEPTH;
    o.target_x = pcx * 8u + (uint32_t)(dx / float(size) * 8.0f);                     /* :75 (target texel of the debug line) */
    o.target_y = pcy * 8u + (uint32_t)(dy / float(size) * 8.0f);
    return o;
}
/* Order of the atlas writes.  Every thread first clears ITS texel (:28-32) and later writes the texel its direction word points at (:75);
 * with mip-1 words several threads of a probe aim at the same texel and others are aimed at by nobody — the shader races.  DEFINED here
 * as if the threads of a probe ran in index order: all clears first, then the target writes, the highest thread index winning.  With the
 * uniform assignment every thread writes its own texel and none of this matters. */
/* trace_probes.slang:58-75: the texel value.  :75 writes the direction as a debug colour; restated is the expression
 * commented beside it: float4(lerp(prev.xyz, radiance, blendfactor), probe_depth) with radiance = the hit's emission (:63).
 * A miss (the reference reads hit_info of an unwritten payload there: undefined) is DEFINED as radiance 0, or the sky with
 * RT3_FLAG_ENV, at depth BACKGROUND_DEPTH. */
RT3_HD vec4 probe_texel(const SceneView& s, const rt3_params& p, const rt3_ray& ray, const rt3_hit& h, vec3 prev) {
    vec3 radiance = v3(0.f, 0.f, 0.f);
    float depth = RT3_BACKGROUND_DEPTH;
    if (h.prim != RT3_MISS) { radiance = hit_info(s, h.prim, h.u, h.v).emissive; depth = h.t; }
    else if (p.flags & RT3_FLAG_ENV) radiance = sample_environment(s, v3(ray.d));
    vec3 o = temporal_blend(prev, radiance, p.blendfactor);
    vec4 r; r.x = o.x; r.y = o.y; r.z = o.z; r.w = depth;
    return r;
}
/* spherical_harmonic_conversion.slang:11-32 for one (probe, colour channel): sum over the 64 directions in index order (the
 * shader's WaveActiveSum leaves the order to the hardware and covers 64 lanes only on wave64 parts), times 4 pi / 64. */
RT3_HD sh3 probe_project_sh(const float* atlas_rgba, uint32_t atlas_w, uint32_t pcx, uint32_t pcy, uint32_t channel) {
    sh3 acc;
    for (int i = 0; i < 9; ++i) acc.c[i] = 0.0f;
    for (uint32_t k = 0; k < 64u; ++k) {
        vec2 f; f.x = (float(k % 8u) + 0.5f) / 8.0f; f.y = (float(k / 8u) + 0.5f) / 8.0f;       /* :13-14 */
        const sh3 b = sh3_evaluate(octa_decode(f));
        const float v = atlas_rgba[4u * ((size_t)(pcy * 8u + k / 8u) * atlas_w + (pcx * 8u + k % 8u)) + channel];   /* :16 */
        for (int i = 0; i < 9; ++i) acc.c[i] = acc.c[i] + b.c[i] * v;
    }
    const float factor = 4.0f * 3.1415926536f / 64.0f;                                            /* :25 */
    for (int i = 0; i < 9; ++i) acc.c[i] = acc.c[i] * factor;
    return acc;
}
RT3_HD uint32_t probe_sh_index(uint32_t pcx, uint32_t pcy, uint32_t channel) { return zcurve_to_linear_index(pcx * 3u + channel, pcy); }   /* :29-31 */
RT3_HD uint32_t probe_sh_capacity(const rt3_params& p) {       /* number of sh3 records addressed by probe_sh_index */
    uint32_t m = probe_count_x(p) * 3u; if (probe_count_y(p) > m) m = probe_count_y(p);
    uint32_t n = 1; while (n < m) n <<= 1;
    return n * n;
}
/* interpolate_probes.slang:13-103 for pixel (x, y).  Returns 0 when nothing is written, 1 for a radiance written at the pixel itself,
 * 2 for the "interpolation failed" marker, which the shader writes at the JITTERED pixel.  In the shader those stray writes race with
 * the regular write of the pixel they land on; here the markers are applied after all regular writes (oracle and kernels alike). */
RT3_HD int probe_interpolate_pixel(const rt3_params& p, uint32_t x, uint32_t y, const uint32_t* gbuffer, const float* depth_img,
                                    const float* sh, uint32_t& out_x, uint32_t& out_y, vec4& value) {
    const uint32_t W = (uint32_t)p.window_size[0], H = (uint32_t)p.window_size[1];
    if (float(x) > p.window_size[0] || float(y) > p.window_size[1] || x >= W || y >= H) return 0;   /* :14-16 (+ image bounds) */
    uint32_t pcx_ = x, pcy_ = y;
    const float pixel_depth = depth_img[(size_t)y * W + x];
    if (pixel_depth == RT3_BACKGROUND_DEPTH) return 0;                                                   /* :20-22 */
    RandomSamplerState rng = rng_init(x, y, p.frame);                                                    /* :24 */
    const uint32_t* g = gbuffer + 4u * ((size_t)y * W + x);
    uint4_ packed = {g[0], g[1], g[2], g[3]};
    const GbufferData surface = gbuffer_unpack(packed);                                                  /* :26 */
    const vec3 pixel_pos = world_pos_from_depth(pixel_depth, x, y, p);                                   /* :28 */
    const float u0 = rng_uniform_float(rng), u1 = rng_uniform_float(rng);
    const int jx = (int)((2.0f * u0 - 1.0f) * 16.0f), jy = (int)((2.0f * u1 - 1.0f) * 16.0f);          /* :31 */
    int qx = (int)x + jx, qy = (int)y + jy;
    qx = qx < 0 ? 0 : (qx > (int)W - 1 ? (int)W - 1 : qx); qy = qy < 0 ? 0 : (qy > (int)H - 1 ? (int)H - 1 : qy);   /* :32 */
    const float jdepth = depth_img[(size_t)qy * W + qx];
    const vec3 jpos = world_pos_from_depth(jdepth, (uint32_t)qx, (uint32_t)qy, p);                       /* :34 */
    if (abs_(dot(normalize(jpos - pixel_pos), surface.normal)) < 0.01f) { pcx_ = (uint32_t)qx; pcy_ = (uint32_t)qy; }   /* :36-38 */
    const uint32_t lx = pcx_ / 16u, ly = pcy_ / 16u;                                                     /* :40 */
    const uint32_t pxs[4] = {lx, lx + 1u, lx, lx + 1u}, pys[4] = {ly, ly, ly + 1u, ly + 1u};             /* :42-47 */
    float wgt[4] = {0.f, 0.f, 0.f, 0.f};
    for (int i = 0; i < 4; ++i) {                                                                        /* :51-68 */
        if (pxs[i] >= W / 16u || pys[i] >= H / 16u) continue;                                            /* :56 (bounds first: the depth read must stay inside the image) */
        const uint32_t ppx = pxs[i] * 16u, ppy = pys[i] * 16u;
        const float pd = depth_img[(size_t)ppy * W + ppx];
        if (pd == RT3_BACKGROUND_DEPTH) continue;
        const vec3 ppos = world_pos_from_depth(pd, ppx, ppy, p);                                         /* :60 */
        if (abs_(dot(normalize(ppos - pixel_pos), surface.normal)) > 0.01f) wgt[i] = 0.0f;               /* :62-63 */
        else {
            float w = clamp_(1.0f - abs_(pd - pixel_depth) / pixel_depth, 0.0f, 1.0f);                   /* :65 */
            const vec3 pn = unpack_normal_11_10_11(gbuffer[4u * ((size_t)ppy * W + ppx) + 1u]);
            w = w * max_(dot(surface.normal, pn), 0.0f);                                                 /* :66 */
            const float w2 = w * w, w4 = w2 * w2;
            wgt[i] = w4 * w4;                                                                            /* :67 pow(w, 8) */
        }
    }
    const float ww = wgt[0] * wgt[0] + wgt[1] * wgt[1] + wgt[2] * wgt[2] + wgt[3] * wgt[3];
    if (ww == 0.0f) {                                                                                    /* :71-76: "Interpolation failed", written at the JITTERED pixel */
        out_x = pcx_; out_y = pcy_; value.x = 1.0f; value.y = 0.0f; value.z = 0.0f; value.w = 1.0f;
        return 2;
    }
    const float wsum = wgt[0] + wgt[1] + wgt[2] + wgt[3];                                                /* :78 */
    vec3 radiance = v3(0.f, 0.f, 0.f);
    for (int i = 0; i < 4; ++i) {                                                                        /* :81-98 */
        const float w = wgt[i] / wsum;
        vec3 pr;
        if (w == 0.0f) continue;                         /* a probe outside the grid has weight 0; its coefficients are not addressed */
        if (p.proberng == 1u) {                                                                          /* :84-88: debug view, probe normal as colour */
            const vec3 pn = unpack_normal_11_10_11(gbuffer[4u * ((size_t)(pys[i] * 16u) * W + pxs[i] * 16u) + 1u]);
            pr = v3((pn.x + 1.0f) / 2.0f, (pn.y + 1.0f) / 2.0f, (pn.z + 1.0f) / 2.0f);
        } else {                                                                                         /* :90-95 */
            sh3 r, gg, b;
            const float* sr = sh + 9u * (size_t)probe_sh_index(pxs[i], pys[i], 0u);
            const float* sg = sh + 9u * (size_t)probe_sh_index(pxs[i], pys[i], 1u);
            const float* sb = sh + 9u * (size_t)probe_sh_index(pxs[i], pys[i], 2u);
            for (int k = 0; k < 9; ++k) { r.c[k] = sr[k]; gg.c[k] = sg[k]; b.c[k] = sb[k]; }
            pr = sh3_unproject_cos_lobe(r, gg, b, surface.normal);
        }
        radiance = v3(radiance.x + w * max_(0.0f, pr.x), radiance.y + w * max_(0.0f, pr.y), radiance.z + w * max_(0.0f, pr.z));   /* :97 */
    }
    const float inv_pi = 0.31830988618379067153f;
    radiance = v3(radiance.x * (surface.albedo.x * inv_pi), radiance.y * (surface.albedo.y * inv_pi), radiance.z * (surface.albedo.z * inv_pi));   /* :100 */
    radiance = radiance + surface.emissive;                                                              /* :101 */
    out_x = x; out_y = y; value.x = radiance.x; value.y = radiance.y; value.z = radiance.z; value.w = 1.0f;   /* :103: Light[threadId.xy] */
    return 1;
}

/* ---------------------------------------------------------------- postprocess.slang ---------- */
/* shaders/old/postprocess.slang:13-24 */
RT3_HD float agx_contrast(float x) {
    float x2 = x * x, x4 = x2 * x2;
    return 15.5f * x4 * x2 - 40.14f * x4 * x + 31.96f * x4 - 6.868f * x2 * x + 0.4298f * x2 + 0.1191f * x - 0.00232f;
}
/* shaders/old/postprocess.slang:26-46: HLSL mul(v, M) with the row-major literal = row vector times matrix. */
RT3_HD vec3 agx(vec3 val) {
    const float m[9] = {0.842479062253094f, 0.0423282422610123f, 0.0423756549057051f, 0.0784335999999992f, 0.878468636469772f,
                        0.0784336f,         0.0792237451477643f, 0.0791661274605434f, 0.879142973793104f};
    const float min_ev = -12.47393f, max_ev = 4.026069f;
    vec3 r = v3(fma_(val.z, m[6], fma_(val.y, m[3], val.x * m[0])), fma_(val.z, m[7], fma_(val.y, m[4], val.x * m[1])),
                fma_(val.z, m[8], fma_(val.y, m[5], val.x * m[2])));
    float range = max_ev - min_ev;
    r.x = (clamp_(r.x > 0.f ? log2_(r.x) : min_ev, min_ev, max_ev) - min_ev) / range;
    r.y = (clamp_(r.y > 0.f ? log2_(r.y) : min_ev, min_ev, max_ev) - min_ev) / range;
    r.z = (clamp_(r.z > 0.f ? log2_(r.z) : min_ev, min_ev, max_ev) - min_ev) / range;
    return v3(agx_contrast(r.x), agx_contrast(r.y), agx_contrast(r.z));
}
/* shaders/old/postprocess.slang:63-88 with AGX_LOOK == 2 ("punchy"). */
RT3_HD vec3 agx_look(vec3 val) {
    float luma = fma_(val.z, 0.0722f, fma_(val.y, 0.7152f, val.x * 0.2126f));
    vec3 p = v3(pow_(val.x, 1.1f), pow_(val.y, 1.1f), pow_(val.z, 1.1f));
    const float sat = 1.1f;
    return v3(luma + sat * (p.x - luma), luma + sat * (p.y - luma), luma + sat * (p.z - luma));
}
/* shaders/old/postprocess.slang:48-61 (the 2.2 power is commented out there). */
RT3_HD vec3 agx_eotf(vec3 val) {
    const float m[9] = {1.19687900512017f,  -0.0528968517574562f, -0.0529716355144438f, -0.0980208811401368f, 1.15190312990417f,
                        -0.0980434501171241f, -0.0990297440797205f, -0.0989611768448433f, 1.15107367264116f};
    return v3(fma_(val.z, m[6], fma_(val.y, m[3], val.x * m[0])), fma_(val.z, m[7], fma_(val.y, m[4], val.x * m[1])),
              fma_(val.z, m[8], fma_(val.y, m[5], val.x * m[2])));
}
RT3_HD vec3 postprocess_pixel(const SceneView& s, const rt3_params& p, uint32_t px, uint32_t py, float depth, vec3 in_color) {
    vec3 color = in_color;
    if (depth == RT3_BACKGROUND_DEPTH) color = sample_environment(s, view_dir(px, py, p));
    return agx_eotf(agx_look(agx(color)));
}

/* The presented pixel: postprocess.slang's output stored in the reference's swapchain format, B8G8R8A8_UNORM (swapchain.rs:27-48).  The
 * float -> UNORM8 conversion is Vulkan's (spec "Conversion from Normalized Fixed-Point..."): clamp to [0, 1], times 255, round to nearest
 * (half up here; the spec allows either neighbour).  Memory order B, G, R, A = one little-endian word A << 24 | R << 16 | G << 8 | B. */
RT3_HD uint32_t unorm8(float x) { return (uint32_t)(clamp_(x, 0.0f, 1.0f) * 255.0f + 0.5f); }
RT3_HD uint32_t pack_bgra8(vec3 c) { return 0xFF000000u | (unorm8(c.x) << 16) | (unorm8(c.y) << 8) | unorm8(c.z); }

/* Host-side table: sRGB byte -> linear float (IEC 61966-2-1), evaluated in double and rounded once. */
inline void build_srgb_lut(float* lut256) {
    for (int i = 0; i < 256; ++i) {
        double c = i / 255.0;
        lut256[i] = (float)(c <= 0.04045 ? c / 12.92 : pow((c + 0.055) / 1.055, 2.4));
    }
}

} /* namespace rt3m */
#endif
