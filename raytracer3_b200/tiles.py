"""Screen-tile sharding across GPUs (no reference counterpart: the reference is single-GPU).

Pixels are independent (the RNG is keyed by pixel coordinate and frame, shaders/include/random.slang:42-46), so a frame
shards by screen tile with the scene and BVH replicated and NO data-path collective while samples accumulate; the only
exchange is one gather of the finished rank-local slabs per output image.

  tile order   row-major over ceil(W/T) x ceil(H/T) tiles of T x T pixels (edge tiles are clipped)
  ownership    owner(tx, ty) = (tx + ty * rotation) mod world_size: interleaved, rotated per row so that vertical
               structures do not land on one rank
  slab         each rank stores its tiles contiguously, tile k at [k*T*T, (k+1)*T*T) row-major inside the tile
               (the layout rt3_pass_*_tiles write and rt3_untile_rgba32f reads); lists are padded with empty rects to
               the same length on every rank so that the gather is a plain all_gather of equal slabs
  bands        the finished image is DELIVERED in bands of whole tile rows: rank r ends up with pixel rows band_rows()[r] of the
               frame (the part it presents / reads back to the host).  A rank's tile list is ordered by destination band, so that
               what it owes every other rank is one contiguous piece of its slab: the exchange is ONE all_to_all in which every rank
               sends (N-1)/N of its slab and receives its band — 1/N of the traffic of an all_gather to everyone — or, over NVLink peer
               memory, one kernel that stores the tiles straight into the owners' band buffers (rt3_scatter_tiles_rgba32f)
"""
import numpy as np

RECT = np.dtype([("x0", "<u4"), ("y0", "<u4"), ("w", "<u4"), ("h", "<u4")])


def tile_grid(width, height, tile=64):
    return -(-width // tile), -(-height // tile)


def owner(tx, ty, tiles_x, world_size):
    rot = max(1, world_size // 2 + 1) if world_size > 2 else 1
    return (tx + ty * rot) % world_size


def band_rows(height, world_size, tile=64):
    """[(row0, row1)] per rank: contiguous bands of whole tile rows, as even as the tile grid allows (a band may be empty when there
    are fewer tile rows than ranks)."""
    ny = -(-height // tile)
    return [(min(height, (r * ny // world_size) * tile), min(height, ((r + 1) * ny // world_size) * tile)) for r in range(world_size)]


def band_of_row(y0, height, world_size, tile=64):
    ny = -(-height // tile)
    ty = y0 // tile
    for r in range(world_size):
        if r * ny // world_size <= ty < (r + 1) * ny // world_size:
            return r
    raise ValueError("row outside the frame")


def rank_tiles(width, height, rank, world_size, tile=64):
    """RECT array of the tiles `rank` owns, ordered by destination band (then row-major), padded with empty rects to the common
    per-rank length."""
    nx, ny = tile_grid(width, height, tile)
    per_rank = [[] for _ in range(world_size)]
    for ty in range(ny):                                   # row-major == ascending destination band
        for tx in range(nx):
            x0, y0 = tx * tile, ty * tile
            per_rank[owner(tx, ty, nx, world_size)].append((x0, y0, min(tile, width - x0), min(tile, height - y0)))
    n = max(len(p) for p in per_rank)
    mine = per_rank[rank] + [(0, 0, 0, 0)] * (n - len(per_rank[rank]))
    return np.array(mine, RECT)


def exchange_plan(width, height, world_size, tile=64):
    """counts[s][d] = number of (real) tiles rank s owns inside rank d's band.  Because rank_tiles() is row-major, the tiles a rank owes
    band d are the contiguous run [sum(counts[s][:d]), sum(counts[s][:d + 1])) of its list."""
    counts = np.zeros((world_size, world_size), np.int64)
    for s in range(world_size):
        for r in rank_tiles(width, height, s, world_size, tile):
            if r["w"] and r["h"]:
                counts[s][band_of_row(int(r["y0"]), height, world_size, tile)] += 1
    return counts


def band_tiles(width, height, rank, world_size, tile=64):
    """What rank `rank` RECEIVES in the exchange, in arrival order (source rank major): RECTs with y0 relative to the band's first row."""
    counts = exchange_plan(width, height, world_size, tile)
    row0 = band_rows(height, world_size, tile)[rank][0]
    out = []
    for s in range(world_size):
        lst = rank_tiles(width, height, s, world_size, tile)
        a = int(counts[s][:rank].sum())
        part = lst[a:a + int(counts[s][rank])].copy()
        part["y0"] -= row0
        out.append(part)
    return np.concatenate(out) if out else np.zeros(0, RECT)


def all_tiles(width, height, world_size, tile=64):
    """Concatenation of every rank's padded list, in rank order == the layout of the all_gather'ed slab."""
    return np.concatenate([rank_tiles(width, height, r, world_size, tile) for r in range(world_size)])


def untile_host(slab, rects, tile, frame):
    """numpy mirror of rt3_untile_rgba32f (used by the CPU tests of the sharding logic)."""
    slab = slab.reshape(len(rects), tile, tile, -1)
    for k, r in enumerate(rects):
        if r["w"] and r["h"]:
            frame[r["y0"]:r["y0"] + r["h"], r["x0"]:r["x0"] + r["w"]] = slab[k, :r["h"], :r["w"]]
    return frame


def tile_host(frame, rects, tile):
    """Inverse of untile_host: cut a full frame into a slab."""
    slab = np.zeros((len(rects), tile, tile, frame.shape[-1]), frame.dtype)
    for k, r in enumerate(rects):
        if r["w"] and r["h"]:
            slab[k, :r["h"], :r["w"]] = frame[r["y0"]:r["y0"] + r["h"], r["x0"]:r["x0"] + r["w"]]
    return slab
