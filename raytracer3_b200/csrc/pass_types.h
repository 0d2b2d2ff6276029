/*
 * pass_types.h — what the pass kernels (pass_kernels.cuh) share with their launchers: the pixel work list and the device counters.
 * No CUDA runtime in here: the CPU warp emulation (tests/emu/pass_emu.cpp) includes it too.
 */
#ifndef RT3_PASS_TYPES_H
#define RT3_PASS_TYPES_H
#include <stddef.h>
#include <stdint.h>

#include "rt3.h"

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
};

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
    const uint32_t k = idx / per_tile, rem = idx - k * per_tile;
    const uint32_t blk = rem >> 5, lane = rem & 31u;
    const uint32_t lx = (blk % m.bw) * 8u + (lane & 7u), ly = (blk / m.bw) * 4u + (lane >> 3);
    rt3_rect r;
    if (m.n_tiles == 1) r = m.single; else r = m.rects[k];
    if (lx >= r.w || ly >= r.h) return false;
    x = r.x0 + lx; y = r.y0 + ly;
    out_index = m.slab ? ((size_t)k * m.tile_w * m.tile_h + (size_t)ly * m.tile_w + lx) : ((size_t)y * m.pitch + x);
    return true;
}

#endif
