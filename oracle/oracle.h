/*
 * oracle.h — C interface of the CPU oracle (liboracle.so).
 *
 * TEST INFRASTRUCTURE ONLY.  This is a CPU restatement of the reference's algorithm for the path
 * (shaders/old/gbuffer.slang, shaders/old/refrence_mode.slang and their includes).  Only tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load it; the
 * product (librt3.so) never links, loads or calls anything under oracle/.
 *
 * PARITY UNPINNED at the traversal boundary: the reference has no tests, no golden vectors and no
 * CPU implementation; BVH build, traversal and ray/triangle arithmetic live in the Vulkan driver
 * (src/renderer/vulkan/raytracing.rs:106-141, src/renderer/pipeline_cache/mod.rs:64-73) and the
 * reference cannot be built or run here (no rustc, no slangc, no Vulkan; shaders include a file that
 * is not in the tree).  The oracle is pinned instead by: published known answers for the integer
 * hashes (MurmurHash3 x86_32, Jenkins), IEEE half/RGB9E5 conversions checked against numpy,
 * brute-force == BVH closest-hit equivalence, and an analytic furnace test (tests/).
 *
 * Query structures (orc_trace_* `mode`): 0 = every triangle (the definition), 2 = a scalar binned-SAH BVH2, 1 = the default of the passes and
 * of the TIMED CPU arm: that BVH collapsed to four children per node with SSE 4-wide slab and Moeller-Trumbore tests which mirror the scalar
 * arithmetic operation for operation.  All three return the same bits (tests/test_oracle_kats.py).
 */
#ifndef RT3_ORACLE_H
#define RT3_ORACLE_H
#include "rt3.h"
#ifdef __cplusplus
extern "C" {
#endif

typedef struct orc_scene orc_scene;

orc_scene* orc_scene_create(const rt3_vertex* v, uint32_t nv, const uint32_t* idx, uint32_t ni, const rt3_geometry_info* g, uint32_t ng,
                            const uint32_t* inst_first_geo, const uint32_t* inst_geo_count, const float* inst_xform, uint32_t n_inst);
void orc_scene_destroy(orc_scene*);
void orc_scene_set_texture(orc_scene*, uint32_t slot, uint32_t w, uint32_t h, const uint8_t* rgba8);
void orc_scene_set_environment(orc_scene*, uint32_t w, uint32_t h, const float* rgb);
void orc_scene_set_bluenoise(orc_scene*, uint32_t w, uint32_t h, const uint8_t* rgba8);
/* Flatten to world-space triangles and build the binned-SAH BVH2.  Returns build milliseconds. */
double orc_scene_build(orc_scene*);
uint32_t orc_num_triangles(const orc_scene*);
void orc_scene_bounds(const orc_scene*, float lo[3], float hi[3]);
void orc_resolve_prim(const orc_scene*, uint32_t prim, uint32_t* instance, uint32_t* geometry, uint32_t* primitive);
/* World-space triangles in global id order: 9 floats each (v0, v1, v2). */
void orc_world_triangles(const orc_scene*, float* out9);
void orc_world_edges(const orc_scene*, float* out9);                       /* (v0, e1, e2) per triangle */
uint64_t orc_scene_view(const orc_scene*, void* out, uint64_t capacity);   /* rt3m::SceneView of host pointers; returns its size */

/* mode 0 = brute force over all triangles, 1 = BVH2 */
/* mode: 0 brute force, 1 BVH4 + SSE (default), 2 scalar BVH2 */
void orc_trace_closest(const orc_scene*, const rt3_ray* rays, uint64_t n, rt3_hit* out, int mode, int threads);
void orc_trace_any(const orc_scene*, const rt3_ray* rays, uint64_t n, uint32_t* out_bits, int mode, int threads);

void orc_pass_gbuffer(const orc_scene*, const rt3_params*, rt3_rect, uint32_t* gbuffer, float* depth, uint32_t* opt_ids, float* opt_bary,
                      int threads);
void orc_set_ray_capture(rt3_ray* rays, uint8_t* valid, uint64_t pixels, uint32_t bounces);   /* see oracle.cpp; NULL switches it off */
void orc_pass_reference(const orc_scene*, const rt3_params*, rt3_rect, const uint32_t* gbuffer, const float* depth, const float* prev_light,
                        float* light, int threads, rt3_counters* opt_counters);
void orc_pass_postprocess(const orc_scene*, const rt3_params*, rt3_rect, const float* depth, const float* in, float* out, int threads);

/* Screen-space probe passes (shaders/old/trace_probes, spherical_harmonic_conversion, interpolate_probes .slang).
 * atlas: (W/16*8) x (H/16*8) RGBA32F; sh: orc_probe_sh_capacity() records of 9 floats, Z-curve indexed; directions may be NULL. */
void orc_pass_probe_trace(const orc_scene*, const rt3_params*, const uint32_t* gbuffer, const float* depth, const uint16_t* directions,
                          const float* prev_atlas, float* atlas, int threads);
void orc_pass_present_bgra8(const orc_scene*, const rt3_params*, rt3_rect, const float* depth, const float* in, uint32_t* out_b8g8r8a8, int threads);
void orc_pass_probe_sis(const rt3_params*, const uint32_t* gbuffer, uint16_t* directions_out, int threads);   /* structured_importance_sampling.slang */
void orc_pass_probe_sh(const rt3_params*, const float* atlas, float* sh, int threads);
void orc_pass_probe_interpolate(const rt3_params*, const uint32_t* gbuffer, const float* depth, const float* sh, float* light, int threads);

/* Known-answer hooks into rt3_shared_math.h (one call evaluates one function). */
uint32_t orc_hash_jenkins(uint32_t a);
uint32_t orc_zcurve(uint32_t x, uint32_t y);
uint32_t orc_murmur3(uint32_t seed, uint32_t index);
float orc_uniform_float(uint32_t seed, uint32_t index);
uint32_t orc_rng_seed(uint32_t px, uint32_t py, uint32_t frame);
void orc_onb(const float n[3], float out9[9]);
void orc_diffuse_sample(float ux, float uy, float out3[3]);
void orc_sincos_2pi(float u, float out2[2]);
float orc_atan2(float y, float x);
float orc_asin(float y);
float orc_log2(float x);
float orc_exp2(float x);
void orc_equirect_uv(const float d[3], float out2[2]);
uint32_t orc_f32_to_f16(float f);
float orc_f16_to_f32(uint32_t h);
uint32_t orc_rgb9e5_pack(const float rgb[3]);
void orc_rgb9e5_unpack(uint32_t v, float out3[3]);
uint32_t orc_pack_normal(const float n[3]);
void orc_unpack_normal(uint32_t p, float out3[3]);
uint32_t orc_pack_color(const float c[3]);
void orc_unpack_color(uint32_t p, float out3[3]);
void orc_gbuffer_pack(const float albedo[3], const float emissive[3], const float normal[3], float roughness, float metalness, uint32_t out4[4]);
void orc_gbuffer_unpack(const uint32_t in4[4], float out11[11]);
void orc_primary_ray(const rt3_params*, uint32_t px, uint32_t py, rt3_ray* out);
int orc_ray_triangle(const float o[3], const float d[3], const float v0[3], const float v1[3], const float v2[3], float out_tuv[3]);
void orc_r2(uint32_t i, float out2[2]);
float orc_vdc(uint32_t bits);
void orc_sample_2d(const orc_scene*, uint32_t flags, uint32_t px, uint32_t py, uint32_t frame, uint32_t index, float out2[2]);
void orc_agx(const float in3[3], float out3[3]);
float orc_contract_probe(float a, float b, float c);
void orc_octa_decode(float fx, float fy, float out3[3]);
void orc_octa_encode(const float n[3], float out2[2]);
void orc_sh3_evaluate(const float d[3], float out9[9]);
void orc_sh3_unproject_cos_lobe(const float sh27[27], const float n[3], float out3[3]);
void orc_world_pos_from_depth(const rt3_params*, uint32_t px, uint32_t py, float depth, float out3[3]);
uint32_t orc_probe_sh_capacity(const rt3_params*);
int orc_specular_sample(float roughness, const float albedo[3], const float wo[3], float ux, float uy, float out14[14]);
void orc_specular_evaluate(float roughness, const float albedo[3], const float wo[3], const float wi[3], float out10[10]);
void orc_specular_dominant_direction(const float n[3], const float v[3], float roughness, float out3[3]);
int orc_layered_sample(const float albedo[3], const float normal[3], float roughness, float metalness, const float ray_d[3], float ux, float uy, float out8[8]);
void orc_layered_evaluate(const float albedo[3], const float normal[3], float roughness, float metalness, const float ray_d[3], const float wi_world[3], float out4[4]);

#ifdef __cplusplus
}
#endif
#endif
