"""Shared scenario definitions for the parity tests (inputs only - no implementation logic)."""
import json
import os

import numpy as np

from oracle import bindings as B
from vacancy_b200 import synthetic as S

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")

# examples.cc:87-115
BUNNY_RESOLUTION = 10.0
BUNNY_INTR = dict(fx=258.65, fy=258.25, cx=159.3, cy=127.65)


def bunny_bbox():
    mn = np.array([-250.0, -344.586151, -129.982697], np.float32) - np.float32(20.0)
    mx = np.array([250.0, 150.542343, 257.329224], np.float32) + np.float32(20.0)
    return mn, mx


class Bunny:
    pass


def load_bunny():
    b = Bunny()
    d = np.load(os.path.join(GOLDEN, "bunny_inputs.npz"))
    b.masks, b.c2w, b.w2c = d["masks"], d["c2w"], d["w2c"]
    with open(os.path.join(GOLDEN, "bunny_golden.json")) as f:
        b.golden = json.load(f)
    b.final = np.load(os.path.join(GOLDEN, "bunny_final.npz"))
    b.bb_min, b.bb_max = bunny_bbox()
    return b


def load_synth_golden():
    with open(os.path.join(GOLDEN, "synth_golden.json")) as f:
        return json.load(f)


MODES = {"max": dict(), "tsdf": dict(update=1, truncation=True, band=0.1)}


def oracle_options(bb_min, bb_max, resolution, **kw):
    return B.make_options(bb_min, bb_max, resolution, **kw)


def bunny_camera(bunny, i):
    return B.Camera(320, 240, bunny.c2w[i], BUNNY_INTR["fx"], BUNNY_INTR["fy"], BUNNY_INTR["cx"],
                    BUNNY_INTR["cy"])


def to_oracle_camera(cam):
    """vacancy_b200 PinholeCamera/OrthoCamera -> oracle.bindings.Camera (same numbers)."""
    return B.Camera(cam.width(), cam.height(), cam.c2w(), *cam._intrinsics(), model=cam.model)


def product_options(opt):
    """oracle.bindings.Options -> vacancy_b200.VoxelCarverOption (field for field)."""
    from vacancy_b200 import (SdfInterpolation, UpdateOutsideImage, VoxelCarverOption, VoxelUpdate,
                              VoxelUpdateOption)
    u = VoxelUpdateOption(VoxelUpdate(opt.voxel_update), SdfInterpolation(opt.sdf_interp),
                          UpdateOutsideImage(opt.update_outside), opt.voxel_max_update_num,
                          opt.voxel_update_weight, bool(opt.use_truncation), opt.truncation_band)
    return VoxelCarverOption(tuple(opt.bb_max), tuple(opt.bb_min), opt.resolution,
                             bool(opt.sdf_minmax_normalize), u)


def product_camera(ocam, w2c=None):
    from vacancy_b200 import OrthoCamera, PinholeCamera
    if ocam.model == 0:
        cam = PinholeCamera(ocam.width, ocam.height, ocam.c2w, (ocam.cx, ocam.cy),
                            (ocam.fx, ocam.fy))
    else:
        cam = OrthoCamera(ocam.width, ocam.height, ocam.c2w)
    if w2c is not None:
        cam.set_w2c_f32(w2c)
    return cam


def random_mask(rng, h, w, kind):
    """Fuzz masks: blobs, noise, all-foreground, all-background, non-{0,255} greys."""
    if kind == "blobs":
        yy, xx = np.mgrid[0:h, 0:w]
        m = np.zeros((h, w), bool)
        for _ in range(rng.integers(1, 5)):
            cy, cx = rng.uniform(0, h), rng.uniform(0, w)
            ry, rx = rng.uniform(1, max(2.0, h / 2)), rng.uniform(1, max(2.0, w / 2))
            m |= ((yy - cy) / ry) ** 2 + ((xx - cx) / rx) ** 2 < 1.0
        return (m * 255).astype(np.uint8)
    if kind == "noise":
        return (rng.random((h, w)) < rng.uniform(0.05, 0.95)).astype(np.uint8) * 255
    if kind == "grey":  # only ==255 counts as foreground (voxel_carver.cc:109)
        return rng.choice(np.array([0, 1, 128, 254, 255], np.uint8), size=(h, w))
    if kind == "full":
        return np.full((h, w), 255, np.uint8)
    if kind == "empty":
        return np.zeros((h, w), np.uint8)
    raise ValueError(kind)


def assert_mesh_equal(got, want, bb_min, bb_max, exact=True):
    """north_star: identical topology and count; vertices within 1e-5 of the bbox diagonal.
    With exact=True the vertex buffers must in addition be bitwise equal."""
    gv, gf = got
    wv, wf = want
    assert gv.shape == wv.shape, (gv.shape, wv.shape)
    assert gf.shape == wf.shape, (gf.shape, wf.shape)
    assert np.array_equal(gf, wf), "face index buffers differ"
    diag = float(np.linalg.norm(np.asarray(bb_max, np.float64) - np.asarray(bb_min, np.float64)))
    if len(gv):
        assert np.abs(gv.astype(np.float64) - wv.astype(np.float64)).max() <= 1e-5 * diag
    if exact:
        assert np.array_equal(gv.view(np.uint32), wv.view(np.uint32)), "vertex bits differ"
