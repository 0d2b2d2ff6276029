"""Host-side logic that needs no GPU: camera / GConst fill, WorkSize2D, tile sharding, and the world_size-2 gather path on gloo."""
import os
import subprocess
import sys

import numpy as np
import pytest

from raytracer3_b200 import tiles
from raytracer3_b200.camera import Camera, look_at_rh, make_params, perspective_rh
from raytracer3_b200.host import WorkSize2D
from raytracer3_b200.types import PARAMS


def test_camera_matches_glam_conventions():
    # reference defaults: pos (0,0,-1), dir (0,0,1), fov 65 deg, near 0.1, far 1000 (src/main.rs:69-76)
    cam = Camera()
    v = cam.view_matrix()
    assert np.allclose(v @ np.array([0, 0, -1, 1]), [0, 0, 0, 1])            # eye maps to the origin
    assert np.allclose((v @ np.array([0, 0, 1, 0]))[:3], [0, 0, -1])         # right-handed: forward is -z in view space
    p = cam.projection_matrix()
    near = p @ np.array([0, 0, -0.1, 1]); far = p @ np.array([0, 0, -1000, 1])
    assert abs(near[2] / near[3]) < 1e-9 and abs(far[2] / far[3] - 1) < 1e-9  # depth 0..1
    assert np.isclose(p[1, 1], 1 / np.tan(np.deg2rad(65) / 2))
    g = make_params(cam, 1920, 1088, frame=7, bounces=4, samples=2)
    assert g.dtype == PARAMS and g.nbytes == 304
    proj = g["proj"][0].reshape(4, 4).T                                        # column-major storage (glam::Mat4)
    assert np.allclose(proj, p, atol=1e-6)
    assert np.allclose(g["proj_inv"][0].reshape(4, 4).T @ proj, np.eye(4), atol=1e-4)
    assert np.allclose(g["view_inv"][0][12:15], cam.position)                  # camera position = translation column of view^-1
    assert tuple(g["window_size"][0]) == (1920.0, 1088.0) and g["frame"][0] == 7


def test_work_size_2d():
    win = (1920, 1088)                                                         # WINDOW_SIZE, src/main.rs:61
    assert WorkSize2D.FullScreen().size(win) == (1920, 1088)
    assert WorkSize2D.FractionalFullScreen(8, 8).size(win) == (240, 136)
    assert WorkSize2D.FractionalFullScreen(16, 7).size(win) == (120, 156)      # div_ceil (executions.rs:66-70)
    assert WorkSize2D.X(5).size(win) == (5, 1) and WorkSize2D.XY(3, 9).size(win) == (3, 9)


@pytest.mark.parametrize("W,H,N", [(1920, 1080, 1), (3840, 1080, 2), (3840, 2160, 4), (7680, 2160, 8), (200, 130, 3)])
def test_tiles_partition_the_frame(W, H, N):
    cover = np.zeros((H, W), np.int32)
    counts = []
    for r in range(N):
        rects = tiles.rank_tiles(W, H, r, N, 64)
        counts.append(int((rects["w"].astype(np.int64) * rects["h"]).sum()))
        for t in rects:
            cover[t["y0"]:t["y0"] + t["h"], t["x0"]:t["x0"] + t["w"]] += 1
    assert (cover == 1).all()                                                  # every pixel owned exactly once
    assert len({len(tiles.rank_tiles(W, H, r, N, 64)) for r in range(N)}) == 1  # equal (padded) slab sizes
    assert max(counts) - min(counts) <= 0.05 * max(counts) + 64 * 64 * 2       # balanced
    counts = tiles.exchange_plan(W, H, N, 64)
    bands = tiles.band_rows(H, N, 64)
    assert bands[0][0] == 0 and bands[-1][1] == H and all(bands[i][1] == bands[i + 1][0] for i in range(N - 1))
    for r in range(N):
        bt = tiles.band_tiles(W, H, r, N, 64)
        assert int((bt["w"].astype(np.int64) * bt["h"]).sum()) == (bands[r][1] - bands[r][0]) * W          # a band is covered exactly once
        assert int(counts[:, r].sum()) == len(bt)
        mine = tiles.rank_tiles(W, H, r, N, 64)
        dest = [tiles.band_of_row(int(t["y0"]), H, N, 64) for t in mine if t["w"] and t["h"]]
        assert dest == sorted(dest)                                                                        # what a rank owes a band is one contiguous run
    allr = tiles.all_tiles(W, H, N, 64)
    frame = np.arange(H * W * 4, dtype=np.float32).reshape(H, W, 4)
    back = tiles.untile_host(tiles.tile_host(frame, allr, 64), allr, 64, np.zeros_like(frame))
    assert (back == frame).all()


WORKER = r'''
import os, sys
import numpy as np, torch, torch.distributed as dist
sys.path.insert(0, sys.argv[1])
from raytracer3_b200 import tiles
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
dist.init_process_group("gloo")
W, H, T = 200, 136, 64
frame = (np.arange(H * W * 4, dtype=np.float32).reshape(H, W, 4) * 0.5)       # what a 1-rank render would produce
mine = tiles.rank_tiles(W, H, rank, world, T)
slab = torch.from_numpy(tiles.tile_host(frame, mine, T))                        # this rank "renders" only its tiles
out = torch.zeros((world * len(mine), T, T, 4))
dist.all_gather_into_tensor(out, slab)                                          # the one collective of the path
got = tiles.untile_host(out.numpy(), tiles.all_tiles(W, H, world, T), T, np.zeros_like(frame))
assert (got == frame).all(), "gathered frame differs from the single-rank frame"
# delivery in bands (TiledRenderer.exchange, 'a2a' path): ONE all_to_all in which every rank sends what it owes each band and receives its own
counts = tiles.exchange_plan(W, H, world, T)
n_real = int(counts[rank].sum())
per = T * T * 4
send = slab.reshape(-1)[: n_real * per].contiguous()
in_splits = [int(c) * per for c in counts[rank]]
out_splits = [int(counts[s][rank]) * per for s in range(world)]
recv = torch.zeros(sum(out_splits))
dist.all_to_all_single(recv, send, out_splits, in_splits)
row0, row1 = tiles.band_rows(H, world, T)[rank]
band = tiles.untile_host(recv.numpy(), tiles.band_tiles(W, H, rank, world, T), T, np.zeros((row1 - row0, W, 4), np.float32))
assert (band == frame[row0:row1]).all(), "band assembled from the all_to_all differs from the single-rank frame's rows"
assert sum(b - a for a, b in tiles.band_rows(H, world, T)) == H
dist.barrier()
sys.stdout.write("rank%dok\n" % rank); sys.stdout.flush()
'''


def test_two_rank_gather_on_gloo(tmp_path):
    """N > 1 host path with the gloo backend: each rank fills its slab; (a) one all_gather + un-tile == the single-rank frame on every rank,
    (b) one all_to_all + un-tile == the rank's band of rows of that frame (the delivery bench.py uses)."""
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    script = tmp_path / "worker.py"
    script.write_text(WORKER)
    env = dict(os.environ, MASTER_ADDR="127.0.0.1", OMP_NUM_THREADS="1")
    import socket
    with socket.socket() as s:                       # a free port: a fixed one fails when two runs follow each other within TIME_WAIT
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
                        "--master-port", str(port), str(script), root], capture_output=True, text=True, env=env, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "rank0ok" in r.stdout and "rank1ok" in r.stdout
