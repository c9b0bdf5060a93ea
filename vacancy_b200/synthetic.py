"""Deterministic synthetic scenes for the benchmark and the parity tests:
analytic-shape silhouettes seen by pinhole cameras on a circle.  No RNG, numpy float64 only; the
masks are INPUTS shared by every implementation under comparison, so how they are rasterised does
not matter for parity as long as it is fixed.
"""
import math

import numpy as np

from .carver import PinholeCamera, VoxelCarverOption, VoxelUpdate, VoxelUpdateOption

SPHERE = ((0.0, 0.0, 0.0, 0.7),)
DUMBBELL = ((-0.35, 0.1, 0.0, 0.38), (0.4, -0.15, 0.1, 0.3))


def look_at_c2w(position, target=(0.0, 0.0, 0.0), up=(0.0, -1.0, 0.0)):
    """common.h:46-76 (c2w from position/target/up): z = forward, x = z cross up, y = z cross x."""
    p, t, u = (np.asarray(v, np.float64) for v in (position, target, up))
    z = (t - p) / np.linalg.norm(t - p)
    x = np.cross(z, u)
    x /= np.linalg.norm(x)
    y = np.cross(z, x)
    m = np.eye(4)
    m[:3, 0], m[:3, 1], m[:3, 2], m[:3, 3] = x, y, z, p
    return m


def circle_cameras(n_views, size, radius=3.0, tilt=0.0):
    """Pinhole cameras of size x size pixels, fx = fy = 1.2 * size, cx = cy = size/2 - 0.5, centres on
    a circle of `radius` in the XZ plane (optionally lifted by `tilt` in y), looking at the origin."""
    cams = []
    f = 1.2 * size
    c = size / 2 - 0.5
    for v in range(n_views):
        a = 2.0 * math.pi * v / n_views
        pos = (radius * math.sin(a), tilt * math.sin(3 * a), -radius * math.cos(a))
        cams.append(PinholeCamera(size, size, look_at_c2w(pos), (c, c), (f, f)))
    return cams


def silhouette(camera, spheres=SPHERE):
    """uint8 mask: 255 where the pixel's viewing ray hits any sphere (cx, cy, cz, r), else 0."""
    w, h = camera.width(), camera.height()
    fx, fy = camera.focal_length()
    cx, cy = camera.principal_point()
    c2w = camera.c2w()
    R, t = c2w[:3, :3], c2w[:3, 3]
    dx = (np.arange(w, dtype=np.float64) - cx) / fx
    dy = (np.arange(h, dtype=np.float64) - cy) / fy
    DX, DY = np.meshgrid(dx, dy)
    dd = DX * DX + DY * DY + 1.0
    mask = np.zeros((h, w), bool)
    for sx, sy, sz, r in spheres:
        c = R.T @ (np.array([sx, sy, sz]) - t)  # sphere centre in camera coordinates
        b = DX * c[0] + DY * c[1] + c[2]
        disc = b * b - dd * (c @ c - r * r)
        mask |= (disc >= 0.0) & (b > 0.0)
    return (mask * np.uint8(255)).astype(np.uint8)


def grid_option(n, weighted=False, truncation=None, band=0.1, max_update_num=255):
    """bbox [-1, 1]^3 with resolution 2/n (n a power of two => exactly n voxels per axis)."""
    truncation = weighted if truncation is None else truncation
    upd = VoxelUpdateOption(
        voxel_update=VoxelUpdate.kWeightedAverage if weighted else VoxelUpdate.kMax,
        use_truncation=truncation, truncation_band=band, voxel_max_update_num=max_update_num)
    return VoxelCarverOption(bb_max=(1.0, 1.0, 1.0), bb_min=(-1.0, -1.0, -1.0),
                             resolution=float(np.float32(2.0) / np.float32(n)), update_option=upd)


# BASELINE.json configs 2-5 (config 1 is the bundled bunny, tests/golden/)
CONFIGS = {
    "cfg2": dict(grid=256, views=36, image=1024, weighted=False),
    "cfg3": dict(grid=512, views=100, image=2048, weighted=True),
    "cfg4": dict(grid=1024, views=360, image=2048, weighted=True),
    "cfg5": dict(grid=2048, views=720, image=2048, weighted=True),
    # cfg4's geometry without truncation (kMax): no kVal / 1.0 plateaus in the SDF images, so every
    # tile that projects into the image takes a sampling path - the fusion kernel's floor
    "cfg4_nomask": dict(grid=1024, views=360, image=2048, weighted=False),
}
