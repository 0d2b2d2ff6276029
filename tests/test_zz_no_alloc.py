"""rt3_pass_probe_trace_static — the allocation-free variant of the probe-trace pass (no counterpart in the reference; added
for callers that cannot call cudaMalloc).  Its scratch is a file-scope `static __device__` array inside librt3.so.

Checked through the C ABI against the existing entry point rt3_pass_probe_trace: bit-identical atlases (the tolerance of every parity test in
this repo) on two calls in a row — the second with the temporal blend against the first call's atlas, so stale scratch from call one would
show — with and without a direction image, plus the size limit.  This file sorts last on purpose; it loads
its own scene, so it passes alone in a fresh process and after the other tests."""
import ctypes as C

import numpy as np
import pytest

from raytracer3_b200 import scenes
from raytracer3_b200.camera import camera_for, make_params
from raytracer3_b200.types import FLAG_ENV, ptr

pytestmark = pytest.mark.gpu

MAX_TEXELS = 524288          # RT3_STATIC_PROBE_MAX_TEXELS (include/rt3.h)


def _setup(gpu_ctx, W, H):
    from raytracer3_b200.host import Renderer, RenderWorld
    sc = scenes.cornell_box()
    sc["environment"] = scenes.analytic_sky(64, 32)
    w = RenderWorld(gpu_ctx)
    w.clear()
    w.load_scene(sc)
    w.create_acceleration_structure()
    gpu_ctx._scene_token = None                  # other test modules cache "their scene is loaded" on the shared context
    r = Renderer(gpu_ctx, w, W, H)
    return sc, w, r


def _atlas_shape(gpu_ctx, p):
    aw, ah, nsh = C.c_uint32(), C.c_uint32(), C.c_uint32()
    gpu_ctx.check(gpu_ctx.L.rt3_probe_layout(ptr(p), C.byref(aw), C.byref(ah), C.byref(nsh)))
    return ah.value, aw.value


@pytest.mark.parametrize("with_directions", [False, True])
def test_static_probe_trace_equals_the_allocating_entry_point_twice_in_a_row(gpu_ctx, with_directions):
    import torch
    W, H = 640, 368
    sc, w, r = _setup(gpu_ctx, W, H)
    L, h = gpu_ctx.L, gpu_ctx.h
    cam = camera_for(sc, "cam0", W, H)
    p0 = make_params(cam, W, H, frame=0, bounces=1, samples=1, flags=FLAG_ENV)
    r.frame(p0)                                                                       # G-buffer + depth of the frame (also refreshes the scene tables)
    gb, dp = r.rg.tensor("gbuffer"), r.rg.tensor("gbuffer_depth")
    ah, aw = _atlas_shape(gpu_ctx, p0)
    assert 0 < ah * aw <= MAX_TEXELS
    directions = None
    if with_directions:
        directions = torch.zeros((ah, aw), dtype=torch.int16, device="cuda")
        gpu_ctx.check(L.rt3_pass_probe_sis(h, ptr(p0), gb.data_ptr(), directions.data_ptr(), gpu_ctx.stream()))
    dptr = None if directions is None else directions.data_ptr()
    new = lambda: torch.full((ah, aw, 4), -7.0, dtype=torch.float32, device="cuda")
    prev_ref = prev_sta = None
    for call in range(2):                                                             # two calls in a row; the second blends with the first
        p = make_params(cam, W, H, frame=call, bounces=1, samples=1, blendfactor=1.0 if call == 0 else 0.5, flags=FLAG_ENV)
        ref, sta = new(), new()
        gpu_ctx.check(L.rt3_pass_probe_trace(h, ptr(p), gb.data_ptr(), dp.data_ptr(), dptr, None if prev_ref is None else prev_ref.data_ptr(),
                                             ref.data_ptr(), gpu_ctx.stream()))
        gpu_ctx.check(L.rt3_pass_probe_trace_static(h, ptr(p), gb.data_ptr(), dp.data_ptr(), dptr, None if prev_sta is None else prev_sta.data_ptr(),
                                                    sta.data_ptr(), gpu_ctx.stream()))
        torch.cuda.synchronize()
        assert torch.equal(ref.view(torch.int32), sta.view(torch.int32)), f"call {call}: {(ref != sta).sum().item()} values differ"
        a = ref.cpu().numpy()
        assert (a != -7.0).all() and np.isfinite(a).all()                             # every texel written
        assert (a[..., 3] > 0).any() and a[..., :3].max() > 0                         # hits with radiance: the pass did trace something
        if call == 0:
            first = ref.clone()
        prev_ref, prev_sta = ref, sta
    # the second call's atlas differs from the first's (other frame index -> other jitter, blend): not two copies of one stale result
    assert not torch.equal(first, prev_ref)


def test_static_probe_trace_rejects_frames_beyond_its_scratch(gpu_ctx):
    import torch
    from raytracer3_b200 import _lib
    W, H = 4096, 2304                                                                 # atlas 2048 x 1152 = 2.36 M texels > the limit
    sc, w, r = _setup(gpu_ctx, 64, 48)
    cam = camera_for(sc, "cam0", W, H)
    p = make_params(cam, W, H, frame=0, bounces=1, samples=1)
    ah, aw = _atlas_shape(gpu_ctx, p)
    assert ah * aw > MAX_TEXELS
    one = torch.zeros(16, dtype=torch.float32, device="cuda")                         # never touched: the call must fail before any launch
    st = gpu_ctx.L.rt3_pass_probe_trace_static(gpu_ctx.h, ptr(p), one.data_ptr(), one.data_ptr(), None, None, one.data_ptr(), gpu_ctx.stream())
    assert st == _lib.ERR_INVALID
    msg = gpu_ctx.L.rt3_last_error(gpu_ctx.h).decode()
    assert "exceeds" in msg and "RT3_STATIC_PROBE_MAX_TEXELS" in msg, msg
    torch.cuda.synchronize()
    assert float(one.abs().sum()) == 0.0
