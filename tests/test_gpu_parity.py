"""GPU parity tests: every result of the CUDA path, obtained through the C ABI, is compared with
the oracle (plain-C restatement; plus the compiled reference when oracle/_ref travelled along) on
the same inputs.  Bar: BIT-EXACT for the 2D SDF, the fused grid, occupancy and mesh indices;
mesh vertices bit-exact as well (north_star only asks for 1e-5 of the bbox diagonal)."""
import hashlib

import numpy as np
import pytest

from oracle import bindings as B
from tests import helpers as H
from vacancy_b200 import (VoxelCarver, distance_transform_l1, make_signed_distance_field,
                          synthetic as S)

pytestmark = pytest.mark.gpu


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()[:16]


def bits_equal(a, b):
    return np.array_equal(np.ascontiguousarray(a).view(np.uint32),
                          np.ascontiguousarray(b).view(np.uint32))


def make_carver(opt):
    car = VoxelCarver(H.product_options(opt))
    assert car.Init(), car.last_error
    return car


# ---- K0: 2D SDF ---------------------------------------------------------------------------------
@pytest.mark.parametrize("kind", ["blobs", "noise", "grey", "full", "empty"])
def test_sdf2d_fuzz(cuda_lib, port, kind):
    rng = np.random.default_rng(sum(map(ord, kind)) + 1)
    for t in range(10):
        h, w = int(rng.integers(1, 150)), int(rng.integers(1, 150))
        mask = H.random_mask(rng, h, w, kind)
        roi = None
        if t % 3 == 1:
            x0, y0 = int(rng.integers(0, w)), int(rng.integers(0, h))
            roi = (x0, y0, int(rng.integers(x0, w)), int(rng.integers(y0, h)))
        rmin, rmax = (roi[:2], roi[2:]) if roi else (None, None)
        assert bits_equal(distance_transform_l1(mask, rmin, rmax),
                          port.distance_transform_l1(mask, roi)), (kind, t, "dt")
        for norm in (True, False):
            for trunc in (True, False):
                got = make_signed_distance_field(mask, rmin, rmax, norm, trunc, 0.1)
                want = port.make_sdf(mask, roi, norm, trunc, 0.1)
                assert bits_equal(got, want), (kind, t, norm, trunc, h, w, roi)


def test_sdf2d_bunny_and_large(cuda_lib, port, bunny):
    for i in range(6):
        got = make_signed_distance_field(bunny.masks[i])
        assert sha(got) == bunny.golden["modes"]["max"]["views"][i]["sdf2d"]["sha"]
    # widths that are not multiples of 32, more than one block of row segments
    cam = S.circle_cameras(3, 1000)[1]
    mask = S.silhouette(cam, S.DUMBBELL)[:777, :]
    for trunc in (False, True):
        assert bits_equal(make_signed_distance_field(mask, use_truncation=trunc),
                          port.make_sdf(mask, None, True, trunc, 0.1))


@pytest.mark.parametrize("shape,roi", [
    ((64, 16384), None),                       # wide: width + height > 16000 -> strip-in-shared-memory column pass
    ((16384, 64), (5, 3000, 60, 6100)),        # tall, ROI of 3101 rows (<= 3168): same kernel, partial ROI
    ((5000, 12000), None),                     # 5000 rows > 3168: the three-sweep column pass
    ((5000, 12000), (100, 200, 11000, 4900)),  # same, partial ROI (4701 rows)
    ((3500, 300), None),                       # packed pass, but too tall for the on-chip strip
    ((1400, 500), (7, 9, 450, 1390)),          # packed pass, 64-column strip, odd ROI origin
])
def test_sdf2d_every_column_pass_variant(cuda_lib, port, shape, roi):
    """Images with width + height > 16000 leave the packed 16-bit column pass for
    sdf_col_smem_kernel (ROI height <= 3168) or sdf_col_kernel; smaller ones take the packed pass
    with its forward strip in shared memory (64 or 32 columns per block) or, when too tall for that,
    through global memory (sdf_col_pack_kernel).  All must stay bit-exact, with columns that are all
    foreground / all background inside the ROI (no seed in the column)."""
    h, w = shape
    rng = np.random.default_rng(h * 7 + w)
    yy, xx = np.mgrid[0:h, 0:w]
    mask = (((yy * 3 + xx) // 97) % 2 == 0)
    mask &= rng.random((h, w)) < 0.995        # a few isolated background pixels
    mask[:, : w // 9] = True                   # all-foreground columns
    mask[:, w - w // 11:] = False              # all-background columns
    mask[h // 3: h // 3 + max(1, h // 50), :] = True  # all-foreground rows
    mask = (mask * 255).astype(np.uint8)
    rmin, rmax = (roi[:2], roi[2:]) if roi else (None, None)
    assert bits_equal(distance_transform_l1(mask, rmin, rmax), port.distance_transform_l1(mask, roi))
    for norm, trunc in ((True, False), (True, True), (False, True)):
        got = make_signed_distance_field(mask, rmin, rmax, norm, trunc, 0.1)
        want = port.make_sdf(mask, roi, norm, trunc, 0.1)
        assert bits_equal(got, want), (shape, roi, norm, trunc)
    # no pixel of the other class anywhere: +-FLT_MAX distances on both passes
    full = np.full((h, w), 255, np.uint8)
    assert bits_equal(make_signed_distance_field(full, rmin, rmax, True, False, 0.1),
                      port.make_sdf(full, roi, True, False, 0.1))


# ---- K1 + K2 + K3 on the bundled bunny (config 1) -----------------------------------------------
@pytest.mark.parametrize("mode", ["max", "tsdf"])
def test_bunny_view_by_view(cuda_lib, bunny, mode):
    """examples.cc's loop: carve one view, extract voxel + both iso surfaces, repeat."""
    opt = H.oracle_options(bunny.bb_min, bunny.bb_max, H.BUNNY_RESOLUTION, **H.MODES[mode])
    car = make_carver(opt)
    assert list(car.dims) == bunny.golden["dims"]
    rows = bunny.golden["modes"][mode]["views"]
    for i in range(6):
        cam = H.product_camera(H.bunny_camera(bunny, i), bunny.w2c[i])
        ok, img = car.Carve(cam, bunny.masks[i], return_sdf=True)
        assert ok
        want = rows[i]
        assert sha(img) == want["sdf2d"]["sha"]
        sdf, un = car.download_grid()
        assert sha(un) == want["un_sha"], i
        assert sha(sdf) == want["sdf_sha"], i
        v, f = car.ExtractIsoSurface(0.0, True)
        assert [len(v), len(f)] == want["mc"]
        assert [sha(v), sha(f)] == want["mc_sha"], i
        assert car.ExtractVoxel(False, copy=False) == want["voxel_cubes"]
        assert car.ExtractVoxel(True, copy=False) == want["surface_cubes"]
    for key, want in bunny.golden["modes"][mode]["extra"].items():
        iso, interp = key[3:].split("_interp")
        v, f = car.ExtractIsoSurface(float(iso), bool(int(interp)))
        assert [len(v), len(f), sha(v), sha(f)] == want, key
    v, f = car.ExtractIsoSurface(0.0, True)
    assert bits_equal(v, bunny.final[mode + "_verts"]) and np.array_equal(f, bunny.final[mode + "_faces"])


def test_bunny_batched_equals_sequential(cuda_lib, bunny):
    """All six views in ONE C-ABI call (one fused launch) give the same grid as six calls."""
    opt = H.oracle_options(bunny.bb_min, bunny.bb_max, H.BUNNY_RESOLUTION, **H.MODES["tsdf"])
    car = make_carver(opt)
    cams = [H.product_camera(H.bunny_camera(bunny, i), bunny.w2c[i]) for i in range(6)]
    ok, sdfs = car.carve_batch(cams, list(bunny.masks), return_sdf=True)
    assert ok
    sdf, un = car.download_grid()
    assert bits_equal(sdf, bunny.final["tsdf_sdf"]) and np.array_equal(un, bunny.final["tsdf_un"])
    # feeding the returned float SDFs through the Image1f overload reproduces it again
    car2 = make_carver(opt)
    for cam, img in zip(cams, sdfs):
        assert car2.Carve(cam, img)
    sdf2, un2 = car2.download_grid()
    assert bits_equal(sdf2, sdf) and np.array_equal(un2, un)


# ---- synthetic scenes, every option combination ---------------------------------------------------
@pytest.mark.parametrize("cfg", [
    dict(update=0, interp=1), dict(update=0, interp=0), dict(update=1, interp=1, truncation=True),
    dict(update=1, interp=1, truncation=True, weight=0.7, max_update_num=3),
    dict(update=0, interp=1, outside=1), dict(update=1, interp=0, outside=1, normalize=False),
])
def test_options_matrix(cuda_lib, port, cfg):
    o = S.grid_option(40)
    opt = B.make_options(o.bb_min, o.bb_max, o.resolution, **cfg)
    car, orc = make_carver(opt), port.carver(opt)
    assert tuple(car.dims) == tuple(orc.dims)
    for a, b in zip(car.positions(), orc.positions()):
        assert bits_equal(a, b)
    pcams = S.circle_cameras(9, 200, tilt=0.5)
    masks = [S.silhouette(c, S.DUMBBELL) for c in pcams]
    ocams = [H.to_oracle_camera(c) for c in pcams]
    ortho_c2w = np.eye(4)
    ortho_c2w[:3, 3] = (-0.9, -0.8, -3.0)
    ocams.append(B.Camera(200, 200, ortho_c2w, model=1))
    pcams.append(H.product_camera(ocams[-1]))
    masks.append(np.zeros((200, 200), np.uint8))
    masks[-1][:2, :2] = 255
    for i, (pc, oc, mask) in enumerate(zip(pcams, ocams, masks)):
        roi = (7, 5, 190, 180) if i == 2 else None
        kw = dict(roi_min=roi[:2], roi_max=roi[2:]) if roi else {}
        ok, got = car.Carve(pc, mask, return_sdf=True, **kw)
        want = orc.carve_mask(oc, mask, roi)
        assert ok and bits_equal(got, want), i
    (gs, gn), (ws, wn) = car.download_grid(), orc.download()
    assert np.array_equal(gn, wn)
    assert bits_equal(gs, ws)
    for iso, interp in ((0.0, True), (0.0, False), (0.03, True)):
        H.assert_mesh_equal(car.ExtractIsoSurface(iso, interp), orc.extract_iso(iso, interp),
                            o.bb_min, o.bb_max)
    for inside_empty in (False, True):
        flags, n = car.occupancy(inside_empty)
        wflags, wn_ = orc.occupancy(inside_empty)
        assert n == wn_ and np.array_equal(flags, wflags)
        gv, gf = car.ExtractVoxel(inside_empty)
        wv, wf = orc.extract_voxel(inside_empty)
        assert np.array_equal(gf, wf) and bits_equal(gv, wv)


def test_synthetic_golden_counts(cuda_lib):
    """Counts recorded from the compiled reference (tests/golden/synth_golden.json)."""
    want_all = H.load_synth_golden()
    for name, n, views, size, weighted, shape in (
            ("sphere32_max", 32, 6, 128, False, S.SPHERE),
            ("dumbbell48_tsdf", 48, 10, 160, True, S.DUMBBELL),
            ("dumbbell64_max", 64, 12, 256, False, S.DUMBBELL)):
        want = want_all[name]
        car = VoxelCarver(S.grid_option(n, weighted=weighted))
        assert car.Init()
        cams = S.circle_cameras(views, size, tilt=0.4)
        assert car.carve_batch(cams, [S.silhouette(c, shape) for c in cams])
        sdf, un = car.download_grid()
        assert sha(sdf) == want["sdf_sha"] and sha(un) == want["un_sha"], name
        v, f = car.ExtractIsoSurface(0.0, True)
        assert [len(v), len(f)] == want["mc"] and [sha(v), sha(f)] == want["mc_sha"], name
        assert car.ExtractVoxel(False, copy=False) == want["voxel_cubes"]
        assert car.ExtractVoxel(True, copy=False) == want["surface_cubes"]


def test_more_views_than_one_launch_and_update_cap(cuda_lib, port):
    """150 views > 64 per launch; voxel_max_update_num = 40 makes the `update_num > max` gate
    (voxel_carver.cc:447-450) bite in weighted mode; mixed image sizes in one batch."""
    o = S.grid_option(32, weighted=True, max_update_num=40)
    opt = B.make_options(o.bb_min, o.bb_max, o.resolution, update=1, truncation=True,
                         max_update_num=40)
    car, orc = make_carver(opt), port.carver(opt)
    cams = S.circle_cameras(75, 96, tilt=0.3) + S.circle_cameras(75, 128, tilt=-0.2)
    masks = [S.silhouette(c, S.DUMBBELL) for c in cams]
    assert car.carve_batch(cams, masks)
    for c, m in zip(cams, masks):
        orc.carve_mask(H.to_oracle_camera(c), m)
    (gs, gn), (ws, wn) = car.download_grid(), orc.download()
    assert gn.max() == 41 and np.array_equal(gn, wn) and bits_equal(gs, ws)
    H.assert_mesh_equal(car.ExtractIsoSurface(), orc.extract_iso(), o.bb_min, o.bb_max)


def test_marching_cubes_handmade_grids(cuda_lib, port):
    """Arbitrary uploaded states: skipped cubes, kVal corners, near-iso values, non-multiple-of-32
    rows - the ownership / numbering rule away from the easy interior."""
    rng = np.random.default_rng(11)
    for bb_max, res in (((1.3, 1.1, 0.9), 0.1), ((7.0, 0.5, 0.45), 0.1), ((0.25, 0.25, 6.7), 0.1)):
        opt = B.make_options((0, 0, 0), bb_max, res)
        car, orc = make_carver(opt), port.carver(opt)
        kval = port.invalid_sdf()
        for t in range(5):
            sdf = rng.normal(size=orc.count).astype(np.float32)
            un = rng.integers(0, 3, orc.count).astype(np.int32)
            sdf[rng.random(orc.count) < 0.15] = kval
            sdf[rng.random(orc.count) < 0.05] = np.float32(1e-6) * rng.integers(-3, 4)
            if t == 0:
                un[:] = 1
            car.upload_grid(sdf, un)
            orc.upload(sdf, un)
            for iso, interp in ((0.0, True), (0.25, True), (0.0, False)):
                H.assert_mesh_equal(car.ExtractIsoSurface(iso, interp),
                                    orc.extract_iso(iso, interp), (0, 0, 0), bb_max)
            for inside_empty in (False, True):
                flags, n = car.occupancy(inside_empty)
                wflags, wn = orc.occupancy(inside_empty)
                assert n == wn and np.array_equal(flags, wflags)


def test_empty_and_degenerate(cuda_lib, port):
    o = S.grid_option(16)
    opt = B.make_options(o.bb_min, o.bb_max, o.resolution)
    car, orc = make_carver(opt), port.carver(opt)
    # nothing carved yet: empty meshes, no cubes
    v, f = car.ExtractIsoSurface()
    assert v.shape == (0, 3) and f.shape == (0, 3)
    assert car.ExtractVoxel(copy=False) == 0
    # Carve before Init returns False like the shim of a null grid
    fresh = VoxelCarver(H.product_options(opt))
    assert fresh.Carve(S.circle_cameras(1, 32)[0], np.zeros((32, 32), np.uint8)) is False
    # all-background and all-foreground silhouettes (FLT_MAX distances, voxel_carver.cc:107-114)
    cam = S.circle_cameras(2, 64)[0]
    for val in (0, 255):
        mask = np.full((64, 64), val, np.uint8)
        ok, got = car.Carve(cam, mask, return_sdf=True)
        want = orc.carve_mask(H.to_oracle_camera(cam), mask)
        assert ok and bits_equal(got, want)
    (gs, gn), (ws, wn) = car.download_grid(), orc.download()
    assert np.array_equal(gn, wn) and bits_equal(gs, ws)


def test_against_compiled_reference(cuda_lib, ref):
    """Same scenario through the UNMODIFIED reference (oracle/_ref) when it travelled along."""
    o = S.grid_option(48, weighted=True)
    opt = B.make_options(o.bb_min, o.bb_max, o.resolution, update=1, truncation=True)
    car, rc = make_carver(opt), ref.carver(opt)
    cams = S.circle_cameras(8, 160, tilt=0.4)
    for c in cams:
        m = S.silhouette(c, S.DUMBBELL)
        ok, got = car.Carve(c, m, return_sdf=True)
        want = rc.carve_mask(H.to_oracle_camera(c), m)
        assert ok and bits_equal(got, want)
    (gs, gn), (ws, wn) = car.download_grid(), rc.download()
    assert np.array_equal(gn, wn) and bits_equal(gs, ws)
    H.assert_mesh_equal(car.ExtractIsoSurface(), rc.extract_iso(), o.bb_min, o.bb_max)
    gv, gf = car.ExtractVoxel()
    wv, wf = rc.extract_voxel()
    assert np.array_equal(gf, wf)
    assert np.abs(gv - wv).max() <= 1e-5 * np.linalg.norm(np.array(o.bb_max) - o.bb_min)


# ---- z-slab sharding: N virtual slabs on one GPU must equal the unsharded result ------------------
@pytest.mark.parametrize("n_slabs", [2, 3, 5])
def test_slabs_equal_unsharded(cuda_lib, bunny, n_slabs):
    from vacancy_b200 import sharded
    opt = H.oracle_options(bunny.bb_min, bunny.bb_max, H.BUNNY_RESOLUTION, **H.MODES["tsdf"])
    cams = [H.product_camera(H.bunny_camera(bunny, i), bunny.w2c[i]) for i in range(6)]
    slabs = sharded.LocalSlabs(H.product_options(opt), n_slabs)
    slabs.carve_batch(cams, list(bunny.masks))
    sdf, un = slabs.download_grid()
    assert bits_equal(sdf, bunny.final["tsdf_sdf"]) and np.array_equal(un, bunny.final["tsdf_un"])
    v, f = slabs.extract_iso(0.0, True)
    assert np.array_equal(f, bunny.final["tsdf_faces"]) and bits_equal(v, bunny.final["tsdf_verts"])
    for key, want in bunny.golden["modes"]["tsdf"]["extra"].items():
        iso, interp = key[3:].split("_interp")
        v, f = slabs.extract_iso(float(iso), bool(int(interp)))
        assert [len(v), len(f), sha(v), sha(f)] == want, key
    flags, n = slabs.occupancy(False)
    assert n == bunny.golden["modes"]["tsdf"]["views"][5]["voxel_cubes"]


def test_slabs_synthetic_vs_oracle(cuda_lib, port):
    from vacancy_b200 import sharded
    o = S.grid_option(40, weighted=True)
    opt = B.make_options(o.bb_min, o.bb_max, o.resolution, update=1, truncation=True)
    orc = port.carver(opt)
    cams = S.circle_cameras(9, 200, tilt=0.5)
    masks = [S.silhouette(c, S.DUMBBELL) for c in cams]
    for c, m in zip(cams, masks):
        orc.carve_mask(H.to_oracle_camera(c), m)
    for n_slabs in (2, 4, 7):
        slabs = sharded.LocalSlabs(o, n_slabs)
        slabs.carve_batch(cams, masks)
        for iso in (0.0, 0.03):
            H.assert_mesh_equal(slabs.extract_iso(iso, True), orc.extract_iso(iso, True), o.bb_min,
                                o.bb_max)


def test_cfg2_full_size_against_compiled_reference(cuda_lib, ref):
    """BASELINE.json config 2 at full size (256^3, 36 views of 1024^2, kMax): 0.6 G voxel-views,
    every one of them through the shared-reciprocal projection - the fused grid must still equal
    the UNMODIFIED reference bit for bit, and the mesh must be identical."""
    cfg = S.CONFIGS["cfg2"]
    o = S.grid_option(cfg["grid"])
    opt = B.make_options(o.bb_min, o.bb_max, o.resolution)
    car, rc = make_carver(opt), ref.carver(opt)
    cams = S.circle_cameras(cfg["views"], cfg["image"], tilt=0.35)
    masks = [S.silhouette(c, S.DUMBBELL) for c in cams]
    assert car.carve_batch(cams, masks)
    for c, m in zip(cams, masks):
        rc.carve_mask(H.to_oracle_camera(c), m)
    (gs, gn), (ws, wn) = car.download_grid(), rc.download()
    assert np.array_equal(gn, wn) and bits_equal(gs, ws)
    H.assert_mesh_equal(car.ExtractIsoSurface(), rc.extract_iso(), o.bb_min, o.bb_max)
    flags, n = car.occupancy(False)
    assert n == int(((ws <= 0) & (wn >= 1)).sum())


def test_cfg3_slab_weighted_against_compiled_reference(cuda_lib, ref):
    """A 512 x 512 x 48 slab of config 3's grid with 2048^2 views in weighted TSDF mode, 70 views
    (more than one launch) against the UNMODIFIED reference."""
    n, planes = 512, 48
    half = planes / n
    res = float(np.float32(2.0) / np.float32(n))
    opt = B.make_options((-1.0, -1.0, -half), (1.0, 1.0, half), res, update=1, truncation=True)
    car, rc = make_carver(opt), ref.carver(opt)
    assert car.dims == (n, n, planes)
    cams = S.circle_cameras(70, 2048, tilt=0.2)
    mask_cache = {}
    for i, c in enumerate(cams):
        m = S.silhouette(c, S.DUMBBELL)
        assert car.Carve(c, m)
        rc.carve_mask(H.to_oracle_camera(c), m)
    (gs, gn), (ws, wn) = car.download_grid(), rc.download()
    assert np.array_equal(gn, wn) and bits_equal(gs, ws)
    H.assert_mesh_equal(car.ExtractIsoSurface(), rc.extract_iso(), opt.bb_min, opt.bb_max)


def test_cfg4_slab_all_360_views_against_compiled_reference(cuda_lib, ref):
    """The benchmarked configuration itself (BASELINE.json config 4: 1024^2 planes, 2048^2 images,
    the sphere scene, weighted TSDF with truncation, ALL 360 views so that the 256-update cap fires
    and the fusion kernel's cap variant runs) on a 1024 x 1024 x 32 slab around z = 0, against the
    UNMODIFIED reference: grid, update_num and the marching-cubes mesh of the slab, bit for bit.
    The sphere looks the same from every camera of the circle (bench.py uses one rasterised mask
    for all views, so does this test); the reference fuses the SDF image it computed itself."""
    cfg = S.CONFIGS["cfg4"]
    n, V, size, planes = cfg["grid"], cfg["views"], cfg["image"], 32
    half = planes / n
    res = float(np.float32(2.0) / np.float32(n))
    opt = B.make_options((-1.0, -1.0, -half), (1.0, 1.0, half), res, update=1, truncation=True,
                         band=0.1)
    car, rc = make_carver(opt), ref.carver(opt)
    assert car.dims == (n, n, planes) and rc.dims == (n, n, planes)
    cams = S.circle_cameras(V, size)
    mask = S.silhouette(cams[0])
    assert np.array_equal(mask, S.silhouette(cams[V // 3]))  # what bench.py relies on
    assert car.carve_batch(cams, [mask] * V)
    ref_sdf = ref.make_sdf(mask, None, True, True, 0.1)
    for c in cams:
        rc.carve_sdf(H.to_oracle_camera(c), ref_sdf)
    (gs, gn), (ws, wn) = car.download_grid(), rc.download()
    assert int(wn.max()) == 256  # voxel_max_update_num = 255: a voxel takes at most 256 updates
    assert np.array_equal(gn, wn) and bits_equal(gs, ws)
    H.assert_mesh_equal(car.ExtractIsoSurface(), rc.extract_iso(), opt.bb_min, opt.bb_max)
    flags, cnt = car.occupancy(False)
    assert cnt == int(((ws <= 0) & (wn >= 1)).sum())


def test_cfg5_slab_with_linear_indices_beyond_2_31(cuda_lib, port):
    """BASELINE.json config 5 (2048^3, 2048^2 views, weighted TSDF) cannot run on the reference: its
    voxel ids are int (voxel_carver.cc:298-301,307,333).  The checker there is the 64-bit slab
    restatement (oracle slab_carver: the same C functions as the whole-grid oracle, which is pinned
    to the compiled reference, on planes addressed by their GLOBAL linear index).  Planes
    [1100, 1116): the first voxel's linear id is 1100 * 2048^2 = 4.6e9 > 2^32, so positions, fusion
    tables and the marching-cubes edge keys all work on 64-bit indices.  The slab's halo is left at
    kVal (no lower neighbour), which makes its extraction the marching cubes of the slab alone."""
    n, size, z0, z1 = 2048, 2048, 1100, 1116
    assert z0 * n * n > 2 ** 32
    o = S.grid_option(n, weighted=True)
    opt = B.make_options(o.bb_min, o.bb_max, o.resolution, update=1, truncation=True, band=0.1)
    car = VoxelCarver(o, z_range=(z0, z1))
    assert car.Init(), car.last_error
    assert car.dims == (n, n, n) and car.z_range == (z0, z1)
    all_cams = S.circle_cameras(S.CONFIGS["cfg5"]["views"], size)
    cams = [all_cams[i] for i in (0, 97, 250, 431, 600)]
    mask = S.silhouette(cams[0])
    assert car.carve_batch(cams, [mask] * len(cams))
    orc = port.slab_carver(opt, z0, z1)
    sdf = port.make_sdf(mask, None, True, True, 0.1)
    for c in cams:
        orc.carve_sdf(H.to_oracle_camera(c), sdf)
    px, py, pz = car.positions()
    assert bits_equal(pz, orc.pz) and bits_equal(px, orc.px)
    (gs, gn), (ws, wn) = car.download_grid(), orc.download()
    assert int(wn.max()) == len(cams) and np.array_equal(gn, wn) and bits_equal(gs, ws)
    gv, gf = car.ExtractIsoSurface()
    assert len(gv) > 10000
    H.assert_mesh_equal((gv, gf), orc.extract_iso(), o.bb_min, o.bb_max)
    flags, cnt = car.occupancy(False)
    wflags, wcnt = orc.occupancy(False)
    assert cnt == wcnt and np.array_equal(flags, wflags)


def test_extract_voxel_when_a_touched_voxel_holds_kval(cuda_lib, port):
    """A caller's own SDF image may contain kVal; fused without truncation (nearest-neighbour
    sampling, first touch assigns the sample, voxel_carver.cc:482-486) a voxel then holds kVal WITH
    update_num = 1 - and ExtractVoxel emits it (sdf <= 0, extract_voxel.cc:285).  The shortcut that
    derives "updated" from sdf != kVal must not be taken for such a grid."""
    o = S.grid_option(32)
    opt = B.make_options(o.bb_min, o.bb_max, o.resolution, interp=0)
    cam = S.circle_cameras(4, 96)[1]
    sdf = port.make_sdf(S.silhouette(cam, S.DUMBBELL), None, True, False, 0.1)
    sdf[30:60, 20:70] = port.invalid_sdf()
    car, orc = make_carver(opt), port.carver(opt)
    assert car.Carve(cam, sdf)
    orc.carve_sdf(H.to_oracle_camera(cam), sdf)
    (gs, gn), (ws, wn) = car.download_grid(), orc.download()
    assert np.array_equal(gn, wn) and bits_equal(gs, ws)
    assert int(((ws == port.invalid_sdf()) & (wn >= 1)).sum()) > 0  # the case in question
    for inside_empty in (False, True):
        flags, n = car.occupancy(inside_empty)
        wflags, wcount = orc.occupancy(inside_empty)
        assert n == wcount and np.array_equal(flags, wflags), inside_empty
    H.assert_mesh_equal(car.ExtractIsoSurface(), orc.extract_iso(), o.bb_min, o.bb_max)


def test_sdf_color_map(cuda_lib, port, bunny):
    """SignedDistance2Color on the device, byte-exact against the oracle."""
    from vacancy_b200 import signed_distance_to_color
    rng = np.random.default_rng(4)
    imgs = [bunny.final["sdf2d_view0"], rng.normal(scale=0.7, size=(61, 47)).astype(np.float32),
            make_signed_distance_field(bunny.masks[3], use_truncation=True)]
    for img in imgs:
        for lo, hi in ((-1.0, 1.0), (-0.3, 2.5)):
            assert np.array_equal(signed_distance_to_color(img, lo, hi),
                                  port.signed_distance_to_color(img, lo, hi))


def test_error_paths_through_the_abi(cuda_lib, port):
    """Argument errors are reported through status codes + vcb_last_error, never by a fallback."""
    import ctypes as C
    from vacancy_b200 import capi, Camera
    from vacancy_b200.carver import make_view
    o = S.grid_option(16)
    car = VoxelCarver(o)
    assert car.Init()
    cam = S.circle_cameras(1, 32)[0]
    mask = np.zeros((32, 32), np.uint8)
    L, h = car._lib, car._h

    def carve(view):
        arr = (C.c_void_p * 1)(mask.ctypes.data)
        return L.vcb_carve_masks(h, 1, C.byref(view), arr, None)

    v = make_view(cam, 32, 32, (0, 0, 32, 31))  # ROI one column beyond the image
    assert carve(v) == -1 and b"roi" in L.vcb_last_error(h)
    v = make_view(cam, 32, 32, (5, 0, 4, 31))  # empty ROI
    assert carve(v) == -1
    v = make_view(cam, 32, 32)
    v.model = 7  # not one of the reference's two concrete cameras
    assert carve(v) == -4 and b"pinhole and ortho" in L.vcb_last_error(h)
    v = make_view(cam, 32, 32)
    v.width = 20000  # beyond the int16 distance range of the 2D SDF passes
    assert carve(v) == -1
    # an abstract Camera (no projection model) cannot be carved with
    with pytest.raises(capi.VcbError):
        car.Carve(Camera(32, 32), mask)
    # fill before count
    buf = np.empty(16, np.float32)
    rc = L.vcb_extract_iso_fill(h, buf.ctypes.data_as(C.POINTER(C.c_float)), None)
    assert rc == -3 and b"has not run" in L.vcb_last_error(h)
    # nothing above left a trace in the grid
    sdf, un = car.download_grid()
    assert (un == 0).all() and (sdf == port.invalid_sdf()).all()
    # slab outside the grid
    bad = VoxelCarver(o, z_range=(10, 40))
    assert bad.Init() is False and "slab" in bad.last_error


def test_euclidean_metric_option(cuda_lib, port):
    """VCB_METRIC_L2 (an addition; the reference is L1): the 2D field equals the Euclidean checker
    bit for bit, and a carve with that option equals the oracle's fusion of those images."""
    from vacancy_b200 import SdfMetric
    rng = np.random.default_rng(33)
    for kind in ("blobs", "noise", "grey", "full", "empty"):
        for t in range(6):
            h, w = int(rng.integers(1, 140)), int(rng.integers(1, 140))
            mask = H.random_mask(rng, h, w, kind)
            roi = None
            if t % 3 == 1:
                x0, y0 = int(rng.integers(0, w)), int(rng.integers(0, h))
                roi = (x0, y0, int(rng.integers(x0, w)), int(rng.integers(y0, h)))
            rmin, rmax = (roi[:2], roi[2:]) if roi else (None, None)
            for norm, trunc in ((True, True), (False, False), (True, False)):
                got = make_signed_distance_field(mask, rmin, rmax, norm, trunc, 0.1,
                                                 metric=SdfMetric.kL2)
                want = port.make_sdf_l2(mask, roi, norm, trunc, 0.1)
                assert bits_equal(got, want), (kind, t, norm, trunc, h, w, roi)
    cam = S.circle_cameras(3, 1000)[1]
    big = S.silhouette(cam, S.DUMBBELL)[:777, :]
    assert bits_equal(make_signed_distance_field(big, use_truncation=True, metric=SdfMetric.kL2),
                      port.make_sdf_l2(big, None, True, True, 0.1))
    # end to end: carve with the option == oracle fusion fed with the Euclidean images
    o = S.grid_option(40, weighted=True)
    o.sdf_metric = SdfMetric.kL2
    car = VoxelCarver(o)
    assert car.Init()
    opt = B.make_options(o.bb_min, o.bb_max, o.resolution, update=1, truncation=True)
    orc = port.carver(opt)
    for c in S.circle_cameras(7, 200, tilt=0.4):
        m = S.silhouette(c, S.DUMBBELL)
        ok, got = car.Carve(c, m, return_sdf=True)
        want = port.make_sdf_l2(m, None, True, True, 0.1)
        assert ok and bits_equal(got, want)
        orc.carve_sdf(H.to_oracle_camera(c), want)
    (gs, gn), (ws, wn) = car.download_grid(), orc.download()
    assert np.array_equal(gn, wn) and bits_equal(gs, ws)
    H.assert_mesh_equal(car.ExtractIsoSurface(), orc.extract_iso(), o.bb_min, o.bb_max)


def test_random_cameras_fuzz(cuda_lib, port):
    """Random poses (any roll, cameras inside / outside the box, some behind it), fx != fy, ROI
    sub-windows, both update modes: exercises the tile classification, the shared-reciprocal
    projection and the fallback paths against the oracle, bit for bit."""
    rng = np.random.default_rng(2024)
    for trial in range(10):
        n = int(rng.choice([24, 40, 56]))
        weighted = bool(trial % 2)
        o = S.grid_option(n, weighted=weighted, band=float(rng.choice([0.05, 0.1, 0.3])))
        opt = B.make_options(o.bb_min, o.bb_max, o.resolution, update=int(weighted),
                             truncation=weighted, band=o.update_option.truncation_band,
                             outside=int(trial % 5 == 4))
        car, orc = make_carver(opt), port.carver(opt)
        size = int(rng.choice([96, 160, 256]))
        for v in range(6):
            pos = rng.normal(size=3)
            pos *= rng.uniform(0.2, 4.0) / np.linalg.norm(pos)
            target = rng.uniform(-0.3, 0.3, 3)
            up = rng.normal(size=3)
            c2w = S.look_at_c2w(pos, target, up)
            fx, fy = rng.uniform(0.5, 2.0, 2) * size
            cx, cy = size / 2 + rng.uniform(-20, 20, 2)
            ocam = B.Camera(size, size, c2w, fx, fy, cx, cy)
            pcam = H.product_camera(ocam)
            ref_cam = S.PinholeCamera(size, size, c2w, (cx, cy), (fx, fy))
            mask = S.silhouette(ref_cam, S.DUMBBELL)
            roi = None
            if v % 3 == 2:
                roi = (int(rng.integers(0, 20)), int(rng.integers(0, 20)),
                       size - 1 - int(rng.integers(0, 20)), size - 1 - int(rng.integers(0, 20)))
            kw = dict(roi_min=roi[:2], roi_max=roi[2:]) if roi else {}
            ok, got = car.Carve(pcam, mask, return_sdf=True, **kw)
            want = orc.carve_mask(ocam, mask, roi)
            assert ok and bits_equal(got, want), (trial, v)
        (gs, gn), (ws, wn) = car.download_grid(), orc.download()
        assert np.array_equal(gn, wn), trial
        assert bits_equal(gs, ws), trial
        H.assert_mesh_equal(car.ExtractIsoSurface(), orc.extract_iso(), o.bb_min, o.bb_max)


def test_long_weight_table_beyond_shared_copy(cuda_lib, port):
    """700 weighted views with voxel_max_update_num = 5000: update counts pass 512, so the launches
    switch to the kernel variant that reads the weighted-average table from global memory."""
    o = S.grid_option(24, weighted=True, max_update_num=5000)
    opt = B.make_options(o.bb_min, o.bb_max, o.resolution, update=1, truncation=True,
                         max_update_num=5000, weight=0.37)
    car, orc = make_carver(opt), port.carver(opt)
    base = S.circle_cameras(7, 64, tilt=0.3)
    base_masks = [S.silhouette(c, S.SPHERE) for c in base]
    cams = [base[i % 7] for i in range(700)]
    masks = [base_masks[i % 7] for i in range(700)]
    assert car.carve_batch(cams, masks)
    for c, m in zip(cams, masks):
        orc.carve_mask(H.to_oracle_camera(c), m)
    (gs, gn), (ws, wn) = car.download_grid(), orc.download()
    assert gn.max() > 600 and np.array_equal(gn, wn) and bits_equal(gs, ws)


@pytest.mark.parametrize("outside", [0, 1])
def test_zoomed_cameras_tiles_outside_the_image(cuda_lib, port, outside):
    """Long focal lengths: most voxel tiles project outside the image (the kernel's "outside"
    tile class), with UpdateOutsideImage::kNone and ::kMax, both update rules."""
    for weighted in (False, True):
        o = S.grid_option(64, weighted=weighted)
        opt = B.make_options(o.bb_min, o.bb_max, o.resolution, update=int(weighted),
                             truncation=weighted, outside=outside)
        car, orc = make_carver(opt), port.carver(opt)
        size = 128
        for v in range(8):
            a = 2.0 * np.pi * v / 8
            pos = (2.5 * np.sin(a), 0.4 * np.cos(2 * a), -2.5 * np.cos(a))
            c2w = S.look_at_c2w(pos, (0.2 * np.cos(a), 0.1, -0.1))
            f = (3.0 + v) * size  # field of view far narrower than the grid
            ocam = B.Camera(size, size, c2w, f, f, size / 2 - 0.5, size / 2 - 0.5)
            mask = S.silhouette(S.PinholeCamera(size, size, c2w, (size / 2 - 0.5,) * 2, (f, f)),
                                ((0.2, 0.1, -0.1, 0.08),))
            ok, got = car.Carve(H.product_camera(ocam), mask, return_sdf=True)
            assert ok and bits_equal(got, orc.carve_mask(ocam, mask)), v
        (gs, gn), (ws, wn) = car.download_grid(), orc.download()
        assert np.array_equal(gn, wn) and bits_equal(gs, ws), weighted


def test_huge_image_coordinates(cuda_lib, port):
    """Cameras inside the grid with absurd focal lengths: image coordinates beyond +-2^22 (where the
    add-magic floor of the sampling path is not valid), infinities from pc.z == 0 planes and
    depths next to zero must all end up on the literal path and match the reference bit for bit."""
    for weighted in (False, True):
        o = S.grid_option(32, weighted=weighted)
        opt = B.make_options(o.bb_min, o.bb_max, o.resolution, update=int(weighted),
                             truncation=weighted, outside=int(weighted))
        car, orc = make_carver(opt), port.carver(opt)
        size = 64
        centre = float(np.float32(2.0) / np.float32(32)) * 0.5  # a voxel centre plane: pc.z == 0 there
        # (pc.z == 0 together with pc.x == 0 would be NaN coordinates - undefined behaviour in the
        # reference - so no camera sits exactly on a voxel centre in x and y)
        for v, (pos, f) in enumerate([((0.0, 0.0, 0.0), 1.0e6),
                                      ((centre + 0.013, centre - 0.021, centre), 3.0e7),
                                      ((0.3, -0.2, 0.1), 8.0e9), ((0.0, 0.0, -0.5), 2.0e5),
                                      ((centre, 0.0, -3.0), 70.0)]):
            c2w = np.eye(4)
            c2w[:3, 3] = pos
            ocam = B.Camera(size, size, c2w, f, f * 1.5, size / 2 - 0.5, size / 2 - 0.5)
            mask = np.zeros((size, size), np.uint8)
            mask[20:44, 16:40] = 255
            ok, got = car.Carve(H.product_camera(ocam), mask, return_sdf=True)
            assert ok and bits_equal(got, orc.carve_mask(ocam, mask)), v
        (gs, gn), (ws, wn) = car.download_grid(), orc.download()
        assert np.array_equal(gn, wn) and bits_equal(gs, ws), weighted
