"""Camera and per-frame constants — host mirror of the reference's Camera and GConst fill.

Reference (paths under /root/reference):
  Camera{position,direction,fov,aspect_ratio,z_near,z_far}, view_matrix = Mat4::look_at_rh(pos, pos+dir, UP),
  projection_matrix = Mat4::perspective_rh(fov, aspect, near, far)        src/components/camera.rs:21-58
  defaults: pos (0,0,-1), dir (0,0,1), fov 65 deg, near 0.1, far 1000        src/main.rs:69-76
  GConst fill: proj, proj_inverse, view, view_inverse, window_size, frame   src/renderer/mod.rs:65-78
glam 0.29 conventions: column-major storage, right-handed, depth 0..1.
"""
from dataclasses import dataclass, field

import numpy as np

from .types import PARAMS

UP = np.array([0.0, 1.0, 0.0])  # src/components/camera.rs:21


def look_at_rh(eye, center, up=UP):
    """glam Mat4::look_at_rh -> 4x4 (row, col) float64 matrix."""
    eye, center, up = (np.asarray(a, np.float64) for a in (eye, center, up))
    f = center - eye
    f /= np.linalg.norm(f)
    s = np.cross(f, up)
    s /= np.linalg.norm(s)
    u = np.cross(s, f)
    m = np.eye(4)
    m[0, :3], m[1, :3], m[2, :3] = s, u, -f
    m[0, 3], m[1, 3], m[2, 3] = -s @ eye, -u @ eye, f @ eye
    return m


def perspective_rh(fov_y, aspect, z_near, z_far):
    """glam Mat4::perspective_rh (depth range 0..1)."""
    h = np.cos(0.5 * fov_y) / np.sin(0.5 * fov_y)
    w = h / aspect
    r = z_far / (z_near - z_far)
    m = np.zeros((4, 4))
    m[0, 0], m[1, 1], m[2, 2], m[3, 2], m[2, 3] = w, h, r, -1.0, r * z_near
    return m


@dataclass
class Camera:
    position: tuple = (0.0, 0.0, -1.0)
    direction: tuple = (0.0, 0.0, 1.0)
    fov_deg: float = 65.0
    aspect_ratio: float = 1920.0 / 1088.0
    z_near: float = 0.1
    z_far: float = 1000.0

    def view_matrix(self):
        d = np.asarray(self.direction, np.float64)
        d = d / np.linalg.norm(d)  # Camera::new normalises (camera.rs:44)
        p = np.asarray(self.position, np.float64)
        return look_at_rh(p, p + d)

    def projection_matrix(self):
        return perspective_rh(np.deg2rad(self.fov_deg), self.aspect_ratio, self.z_near, self.z_far)


def make_params(camera, width, height, frame=0, bounces=4, samples=1, blendfactor=1.0, flags=0):
    """Fill a GConst record exactly as `commands` does (src/renderer/mod.rs:72-78) plus the knobs the
    (now unwired) constants editor set: bounces, samples, blendfactor."""
    p = np.zeros(1, PARAMS)
    proj, view = camera.projection_matrix(), camera.view_matrix()
    col = lambda m: np.asarray(m, np.float64).T.reshape(16).astype(np.float32)  # column-major storage
    p["proj"], p["view"] = col(proj), col(view)
    p["proj_inv"], p["view_inv"] = col(np.linalg.inv(proj)), col(np.linalg.inv(view))
    p["window_size"] = (float(width), float(height))
    p["frame"], p["blendfactor"], p["bounces"], p["samples"], p["flags"] = frame, blendfactor, bounces, samples, flags
    return p


def camera_for(scene, name="cam0", width=1920, height=1080):
    kw = dict(scene["cameras"][name])
    return Camera(aspect_ratio=width / height, **kw)
