"""Reduces the ONE image the reference ships for this path — /root/reference/resources/refrence.png, a 1920x1080 Cornell-box render
(SURVEY.md section 8c "(d) visual Cornell-box check") — to a small committed fixture: the lit region (bounding box of the non-black pixels)
area-averaged to 64x64 display-referred RGB, its bounding box and the whole-image mean.  Runs only where /root/reference is mounted:

    python tests/golden/make_refimage_stats.py        # writes tests/golden/ref_image_64.npz

`crop64` is shared with tests/test_reference_image.py, which reduces our own render the same way and compares region statistics
(a QUALITATIVE check: the PNG comes from an unknown revision / renderer, so per-pixel parity with it is not defined)."""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
OUT = os.path.join(HERE, "ref_image_64.npz")


def crop64(rgb, threshold=8.0 / 255.0, n=64):
    """rgb: (H, W, 3) float in 0..1, row 0 = top.  Returns (n x n x 3 area means over the bounding box of pixels brighter than threshold, bbox)."""
    lit = rgb.max(-1) > threshold
    ys, xs = np.where(lit.any(1))[0], np.where(lit.any(0))[0]
    y0, y1, x0, x1 = int(ys[0]), int(ys[-1]) + 1, int(xs[0]), int(xs[-1]) + 1
    c = rgb[y0:y1, x0:x1].astype(np.float64)
    ye = np.linspace(0, c.shape[0], n + 1).round().astype(int)
    xe = np.linspace(0, c.shape[1], n + 1).round().astype(int)
    out = np.zeros((n, n, 3))
    for j in range(n):
        for i in range(n):
            out[j, i] = c[ye[j]:max(ye[j + 1], ye[j] + 1), xe[i]:max(xe[i + 1], xe[i] + 1)].mean((0, 1))
    return out.astype(np.float32), np.array([x0, y0, x1, y1], np.int32)


def main():
    sys.path.insert(0, ROOT)
    from raytracer3_b200 import assets
    with open("/root/reference/resources/refrence.png", "rb") as f:
        img = assets.decode_image(f.read())
    rgb = img[..., :3].astype(np.float32) / 255.0
    small, bbox = crop64(rgb)
    np.savez_compressed(OUT, rgb64=small, bbox=bbox, size=np.array(rgb.shape[:2][::-1], np.int32), mean_rgb=rgb.mean((0, 1)).astype(np.float32))
    print("wrote", OUT, "bbox", bbox, "mean", rgb.mean((0, 1)) * 255)


if __name__ == "__main__":
    main()
