"""Seeded synthetic Gaussians and cameras for the parity tests and the benchmark.

Cameras follow the reference's conventions exactly — W2C from (R, T) as
utils/graphics_utils.py:38-50 builds it, the projection of :52-71 with znear 0.01 / zfar 100
(scene/cameras.py:48-49), both stored TRANSPOSED, full_proj = view @ proj, camera centre = last
row of the inverse view (scene/cameras.py:54-57) — so the flat 16 floats the rasterizer reads are
what the reference would hand it.  Everything is generated with numpy's PCG64 on the host
(bit-reproducible across machines) in float32.
"""
from __future__ import annotations

import math
from dataclasses import dataclass

import numpy as np

ZNEAR, ZFAR = 0.01, 100.0


@dataclass
class Camera:
    width: int
    height: int
    fovx: float
    fovy: float
    world_view_transform: np.ndarray   # [4,4] f32, transposed W2C (row-vector convention)
    projection_matrix: np.ndarray      # [4,4] f32, transposed
    full_proj_transform: np.ndarray    # [4,4] f32 = world_view_transform @ projection_matrix
    camera_center: np.ndarray          # [3] f32

    @property
    def tanfovx(self) -> float:
        return math.tan(self.fovx * 0.5)

    @property
    def tanfovy(self) -> float:
        return math.tan(self.fovy * 0.5)


def world_to_view(R: np.ndarray, t: np.ndarray) -> np.ndarray:
    """W2C = [[R^T, t],[0,1]] (R is the camera-to-world rotation), float32."""
    m = np.zeros((4, 4), np.float64)
    m[:3, :3] = R.T
    m[:3, 3] = t
    m[3, 3] = 1.0
    return m.astype(np.float32)


def projection(znear: float, zfar: float, fovx: float, fovy: float) -> np.ndarray:
    ty, tx = math.tan(fovy / 2), math.tan(fovx / 2)
    top, right = ty * znear, tx * znear
    p = np.zeros((4, 4), np.float32)
    p[0, 0] = 2.0 * znear / (2.0 * right)
    p[1, 1] = 2.0 * znear / (2.0 * top)
    p[3, 2] = 1.0
    p[2, 2] = zfar / (zfar - znear)
    p[2, 3] = -(zfar * znear) / (zfar - znear)
    return p


def make_camera(R: np.ndarray, t: np.ndarray, fovx: float, width: int, height: int) -> Camera:
    focal = width / (2.0 * math.tan(fovx / 2))
    fovy = 2.0 * math.atan(height / (2.0 * focal))
    view_t = np.ascontiguousarray(world_to_view(R, t).T)
    proj_t = np.ascontiguousarray(projection(ZNEAR, ZFAR, fovx, fovy).T)
    full = (view_t @ proj_t).astype(np.float32)
    center = np.linalg.inv(view_t.astype(np.float64))[3, :3].astype(np.float32)
    return Camera(width, height, fovx, fovy, view_t, proj_t, np.ascontiguousarray(full), center)


def look_at(eye, target, up=(0.0, -1.0, 0.0)):
    """(R, t) of a camera at `eye` looking at `target`; +z forward, +x right, +y down."""
    eye, target, up = (np.asarray(v, np.float64) for v in (eye, target, up))
    fwd = target - eye
    fwd /= np.linalg.norm(fwd)
    right = np.cross(-up, fwd)
    right /= np.linalg.norm(right)
    down = np.cross(fwd, right)
    R = np.stack([right, down, fwd], axis=1)  # columns = camera axes in world
    t = -R.T @ eye
    return R, t


def front_camera(width: int, height: int, fovx_deg: float = 60.0) -> Camera:
    """Camera at the origin looking down +z."""
    return make_camera(np.eye(3), np.zeros(3), math.radians(fovx_deg), width, height)


def ring_cameras(n: int, width: int, height: int, radius: float = 4.0, centre=(0.0, 0.0, 10.5),
                 fovx_deg: float = 60.0, seed: int = 7):
    """n cameras on a ring of `radius` in the z=0 plane, all looking at the scene centre."""
    rng = np.random.Generator(np.random.PCG64(seed))
    cams = []
    for k in range(n):
        a = 2 * math.pi * k / n + rng.normal(0, 0.05)
        eye = np.array([radius * math.cos(a), radius * math.sin(a), rng.normal(0, 0.1)])
        R, t = look_at(eye, centre)
        cams.append(make_camera(R, t, math.radians(fovx_deg), width, height))
    return cams


def batch_cameras(n: int, width: int, height: int):
    """Cameras of an n-view training batch over the scene `make_gaussians` builds for the front camera: view 0
    alone is the front camera (the single-view workload); a batch of n > 1 views sits on SURVEY §8(d)'s ring
    of radius 4 around the scene axis, every camera looking at the scene centre — distinct views, different
    instance counts per view."""
    if n == 1:
        return [front_camera(width, height)]
    return ring_cameras(n, width, height)


def make_gaussians(P: int, cam: Camera, sh_coeffs: int = 16, seed: int = 0, sigma_px: float = 0.62,
                   huge_frac: float = 0.0005, needle_frac: float = 0.005, behind_frac: float = 0.10):
    """SURVEY §8d scene: 90 % of the means uniform in the view-frustum slab z in [1, 20] (x, y
    scaled by 1.1 so some fall outside the image), 10 % in a cube around the camera (exercises the
    near-plane cull); log-normal scales sized so the median splat radius is ~4 px, plus a few
    screen-filling splats (150-600 px radius) and 50:1 needles; random unit quaternions;
    opacity = sigmoid(N(0, 2^2)); SH DC ~ N(0,1), higher bands ~ N(0, 0.15^2).
    Returns float32 numpy arrays in the layouts the reference passes."""
    g = lambda s: np.random.Generator(np.random.PCG64(seed * 16 + s))
    r0, r1, r2, r3, r4 = g(0), g(1), g(2), g(3), g(4)
    focal = cam.width / (2.0 * cam.tanfovx)
    n_behind = int(P * behind_frac)
    n_front = P - n_behind
    z = r0.uniform(1.0, 20.0, n_front)
    x = r0.uniform(-1.0, 1.0, n_front) * z * cam.tanfovx * 1.1
    y = r0.uniform(-1.0, 1.0, n_front) * z * cam.tanfovy * 1.1
    front = np.stack([x, y, z], 1)
    behind = r0.uniform(-3.0, 3.0, (n_behind, 3))
    # camera-space points -> world space through the inverse view (row-vector convention)
    pts_cam = np.concatenate([front, behind], 0)
    perm = r0.permutation(P)
    pts_cam = pts_cam[perm]
    c2w = np.linalg.inv(cam.world_view_transform.astype(np.float64))
    means = (np.concatenate([pts_cam, np.ones((P, 1))], 1) @ c2w)[:, :3]

    depth = np.maximum(np.abs(pts_cam[:, 2]), 0.5)
    base = sigma_px * depth / focal                       # world size of sigma_px pixels at that depth
    scales = base[:, None] * np.exp(r1.normal(0.0, 0.7, (P, 3)))
    kind = r1.uniform(0, 1, P)
    huge = kind < huge_frac
    needle = (kind >= huge_frac) & (kind < huge_frac + needle_frac)
    huge_sigma = r1.uniform(50.0, 200.0, P)               # radius = 3 sigma -> 150..600 px
    scales[huge] = (huge_sigma[huge] * depth[huge] / focal)[:, None] * np.exp(r1.normal(0, 0.1, (int(huge.sum()), 3)))
    ax = r1.integers(0, 3, P)
    needle_idx = np.nonzero(needle)[0]
    scales[needle_idx] = (base[needle_idx] * 0.4)[:, None]
    scales[needle_idx, ax[needle_idx]] *= 50.0

    q = r2.normal(0, 1, (P, 4))
    q /= np.linalg.norm(q, axis=1, keepdims=True)
    opac = 1.0 / (1.0 + np.exp(-r3.normal(0.0, 2.0, (P, 1))))
    shs = r4.normal(0.0, 0.15, (P, sh_coeffs, 3))
    shs[:, 0, :] = r4.normal(0.0, 1.0, (P, 3))
    f32 = lambda a: np.ascontiguousarray(a, dtype=np.float32)
    return dict(means3D=f32(means), scales=f32(scales), rotations=f32(q), opacities=f32(opac), shs=f32(shs))


def settings_kwargs(cam: Camera, bg, sh_degree: int, scale_modifier: float = 1.0, debug: bool = False):
    """Keyword arguments of GaussianRasterizationSettings as gaussian_renderer/__init__.py:42-55
    fills them (numpy values; callers turn the arrays into tensors on their device)."""
    return dict(image_height=cam.height, image_width=cam.width, tanfovx=cam.tanfovx, tanfovy=cam.tanfovy,
                bg=np.asarray(bg, np.float32), scale_modifier=scale_modifier,
                viewmatrix=cam.world_view_transform, projmatrix=cam.full_proj_transform,
                sh_degree=sh_degree, campos=cam.camera_center, prefiltered=False, debug=debug)
