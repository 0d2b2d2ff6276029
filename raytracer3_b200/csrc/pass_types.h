/*
 * pass_types.h — what the pass kernels (pass_kernels.cuh) share with their launchers: the pixel work list and the device counters.
 * No CUDA runtime in here: the CPU warp emulation (tests/emu/pass_emu.cpp) includes it too.
 */
#ifndef RT3_PASS_TYPES_H
#define RT3_PASS_TYPES_H
#include <stddef.h>
#include <stdint.h>

#include "rt3.h"

/* Division of a work index (< 2^31) by a launch constant as one multiply-high and a shift: q = (n * m) >> (31 + s) with s = ceil(log2 d),
 * m = floor(2^(31 + s) / d) + 1 (Granlund & Montgomery 1994: exact for every n < 2^31; m fits 32 bits because d > 2^(s - 1)).  pixel_of ran
 * two 32-bit integer divisions per thread of every pass kernel, ~20 instructions each. */
struct FastDiv { uint32_t m, s; };
static inline FastDiv make_fastdiv(uint32_t d) {
    if (d == 0) d = 1;
    uint32_t s = 0;
    while ((1ull << s) < d) ++s;
    FastDiv f;
    f.s = s; f.m = (uint32_t)((1ull << (31 + s)) / d + 1ull);
    return f;
}
__host__ __device__ __forceinline__ uint32_t fastdiv(uint32_t n, const FastDiv& f) { return (uint32_t)(((unsigned long long)n * f.m) >> (31u + f.s)); }

/* Pixel work list: one or many rects, frame-pitch or tile-slab addressing (see rt3.h). */
struct PixelMap {
    rt3_rect single;          /* used when n_tiles == 1          */
    const rt3_rect* rects;    /* device array when n_tiles > 1   */
    uint32_t n_tiles;
    uint32_t tile_w, tile_h;  /* max rect size; slab stride      */
    uint32_t bw, bh;          /* 8x4-pixel blocks per tile       */
    uint32_t pitch;           /* frame width (frame addressing)  */
    uint32_t slab;            /* 1 = tile-major slab addressing  */
    uint32_t work;            /* n_tiles * bw * bh * 32          */
    FastDiv div_tile, div_bw; /* by bw * bh * 32 and by bw (finish_pixel_map) */
};
static inline void finish_pixel_map(PixelMap& m) { m.div_tile = make_fastdiv(m.bw * m.bh * 32u); m.div_bw = make_fastdiv(m.bw); }

/* A frame's work indices are dealt to S strips in units of STRIP_UNIT indices (unit u goes to strip u mod S): every strip sees the whole
 * image at a coarser grain, so the strips carry equal work.  The reference pass runs one wavefront per strip, each on its own stream
 * (passes.cu): the tail of one strip's traversal launch overlaps the other strips' kernels, and a strip's queues are S times smaller. */
constexpr uint32_t STRIP_UNIT = 2048u;
struct StripMap {
    uint32_t strip, n_strips;
    uint32_t local_work;      /* number of local indices j of this strip (a multiple of STRIP_UNIT; indices that map past `work` are padding) */
};
__host__ __device__ __forceinline__ uint32_t strip_to_work(const StripMap& sm, uint32_t j) {
    return ((j / STRIP_UNIT) * sm.n_strips + sm.strip) * STRIP_UNIT + (j % STRIP_UNIT);
}
static inline StripMap make_strip(uint32_t work, uint32_t strip, uint32_t n_strips) {
    const uint32_t units = (work + STRIP_UNIT - 1) / STRIP_UNIT;
    StripMap sm;
    sm.strip = strip; sm.n_strips = n_strips;
    sm.local_work = ((units + n_strips - 1 - strip) / n_strips) * STRIP_UNIT;
    return sm;
}

struct DeviceCounters {
    unsigned long long rays_primary, rays_secondary, rays_shadow, paths_started, paths_escaped;
};

/* work index -> pixel.  Returns false for padding lanes. */
__device__ __forceinline__ bool pixel_of(const PixelMap& m, uint32_t idx, uint32_t& x, uint32_t& y, size_t& out_index) {
    const uint32_t per_tile = m.bw * m.bh * 32u;
    const uint32_t k = m.n_tiles == 1 ? 0u : fastdiv(idx, m.div_tile), rem = idx - k * per_tile;
    const uint32_t blk = rem >> 5, lane = rem & 31u;
    const uint32_t brow = fastdiv(blk, m.div_bw), bcol = blk - brow * m.bw;
    const uint32_t lx = bcol * 8u + (lane & 7u), ly = brow * 4u + (lane >> 3);
    rt3_rect r;
    if (m.n_tiles == 1) r = m.single; else r = m.rects[k];
    if (lx >= r.w || ly >= r.h) return false;
    x = r.x0 + lx; y = r.y0 + ly;
    out_index = m.slab ? ((size_t)k * m.tile_w * m.tile_h + (size_t)ly * m.tile_w + lx) : ((size_t)y * m.pitch + x);
    return true;
}

#endif
