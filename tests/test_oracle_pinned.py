"""Pins the oracle (CPU only): the plain-C restatement must reproduce (a) the committed golden
fixtures that were generated from the compiled reference and (b) the compiled reference itself
(oracle/_ref) wherever that library is present, bit for bit."""
import hashlib

import numpy as np
import pytest

from oracle import bindings as B
from tests import helpers as H
from vacancy_b200 import synthetic as S
from vacancy_b200 import w2c_from_c2w


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()[:16]


def grid_stats(sdf, un):
    upd = un >= 1
    return dict(updated=int(upd.sum()), occupied=int(((sdf <= 0) & upd).sum()),
                max_update_num=int(un.max()), sum_sdf=float(sdf[upd].astype(np.float64).sum()),
                sdf_sha=sha(sdf), un_sha=sha(un))


@pytest.mark.parametrize("mode", ["max", "tsdf"])
def test_port_reproduces_bunny_golden(port, bunny, mode):
    """examples.cc's scenario, view by view, against tests/golden/bunny_golden.json."""
    opt = H.oracle_options(bunny.bb_min, bunny.bb_max, H.BUNNY_RESOLUTION, **H.MODES[mode])
    car = port.carver(opt)
    assert list(car.dims) == bunny.golden["dims"] == [54, 53, 42]
    rows = bunny.golden["modes"][mode]["views"]
    for i in range(6):
        img = car.carve_mask(H.bunny_camera(bunny, i), bunny.masks[i])
        want = rows[i]
        assert sha(img) == want["sdf2d"]["sha"]
        sdf, un = car.download()
        got = grid_stats(sdf, un)
        for k in ("updated", "occupied", "max_update_num", "sdf_sha", "un_sha"):
            assert got[k] == want[k], (i, k)
        v, f = car.extract_iso(0.0, True)
        assert [len(v), len(f)] == want["mc"]
        assert [sha(v), sha(f)] == want["mc_sha"]
        assert car.occupancy(False)[1] == want["voxel_cubes"]
        assert car.occupancy(True)[1] == want["surface_cubes"]
    for key, want in bunny.golden["modes"][mode]["extra"].items():
        iso, interp = key[3:].split("_interp")
        v, f = car.extract_iso(float(iso), bool(int(interp)))
        assert [len(v), len(f), sha(v), sha(f)] == want, key
    sdf, un = car.download()
    assert np.array_equal(sdf.view(np.uint32), bunny.final[mode + "_sdf"].view(np.uint32))
    assert np.array_equal(un, bunny.final[mode + "_un"])


def test_survey_table(bunny):
    """The fixtures themselves carry the numbers of SURVEY.md section 4 (reference run)."""
    rows = bunny.golden["modes"]["max"]["views"]
    assert [r["updated"] for r in rows] == [119556] + [120204] * 5
    assert [r["occupied"] for r in rows] == [69504, 48789, 34622, 30463, 29065, 28475]
    assert [r["mc"] for r in rows] == [[9469, 18305], [11136, 21657], [9885, 19564], [9293, 18498],
                                       [8822, 17560], [8672, 17270]]
    assert [r["surface_cubes"] for r in rows] == [7228, 8689, 7712, 7238, 6878, 6742]
    assert [r["mc"] for r in bunny.golden["modes"]["tsdf"]["views"]] == [
        [9469, 18305], [13970, 27191], [15818, 30902], [15651, 30686], [14662, 28956],
        [14245, 28130]]
    assert abs(rows[5]["sum_sdf"] - 16349.3455) < 1e-3


def test_port_synthetic_golden(port):
    want_all = H.load_synth_golden()
    for name, n, views, size, weighted, shape in (
            ("sphere32_max", 32, 6, 128, False, S.SPHERE),
            ("dumbbell48_tsdf", 48, 10, 160, True, S.DUMBBELL)):
        want = want_all[name]
        o = S.grid_option(n, weighted=weighted)
        car = port.carver(B.make_options(o.bb_min, o.bb_max, o.resolution, update=int(weighted),
                                         truncation=weighted))
        assert list(car.dims) == want["dims"]
        for cam in S.circle_cameras(views, size, tilt=0.4):
            car.carve_mask(H.to_oracle_camera(cam), S.silhouette(cam, shape))
        sdf, un = car.download()
        got = grid_stats(sdf, un)
        for k in ("updated", "occupied", "sdf_sha", "un_sha"):
            assert got[k] == want[k], (name, k)
        v, f = car.extract_iso(0.0, True)
        assert [sha(v), sha(f)] == want["mc_sha"]
        assert car.occupancy(False)[1] == want["voxel_cubes"]
        assert car.occupancy(True)[1] == want["surface_cubes"]


def test_python_w2c_matches_port_and_golden(port, bunny):
    """camera.cc:39-42 restated twice (C oracle, product host code); both equal the reference's."""
    for i in range(6):
        assert np.array_equal(port.w2c_from_c2w(bunny.c2w[i]).view(np.uint32),
                              bunny.w2c[i].view(np.uint32))
        assert np.array_equal(w2c_from_c2w(bunny.c2w[i]).view(np.uint32),
                              bunny.w2c[i].view(np.uint32))


# ---- against the compiled reference itself -------------------------------------------------------
def test_w2c_vs_reference(ref, port):
    rng = np.random.default_rng(0)
    for _ in range(300):
        q = rng.normal(size=4)
        q /= np.linalg.norm(q)
        c2w = ref.tum_pose_to_c2w(rng.uniform(-5, 5, 3), q)
        if rng.random() < 0.3:  # not only rigid poses
            c2w[:3, :3] *= rng.uniform(0.5, 2.0)
        want = ref.w2c_from_c2w(c2w)
        assert np.array_equal(port.w2c_from_c2w(c2w).view(np.uint32), want.view(np.uint32))
        assert np.array_equal(w2c_from_c2w(c2w).view(np.uint32), want.view(np.uint32))


def test_mc_tables_vs_reference(ref, port):
    re, rt = ref.mc_tables()
    pe, pt = port.mc_tables()
    assert np.array_equal(re, pe), "derived edge table differs from marching_cubes_lut.cc"
    assert np.array_equal(rt, pt)


@pytest.mark.parametrize("kind", ["blobs", "noise", "grey", "full", "empty"])
def test_sdf2d_vs_reference(ref, port, kind):
    rng = np.random.default_rng(sum(map(ord, kind)))
    for t in range(12):
        h, w = int(rng.integers(3, 90)), int(rng.integers(3, 90))
        mask = H.random_mask(rng, h, w, kind)
        roi = None
        if t % 3 == 1:
            x0, y0 = int(rng.integers(0, w - 1)), int(rng.integers(0, h - 1))
            roi = (x0, y0, int(rng.integers(x0, w)), int(rng.integers(y0, h)))
        a = ref.distance_transform_l1(mask, roi)
        b = port.distance_transform_l1(mask, roi)
        assert np.array_equal(a.view(np.uint32), b.view(np.uint32))
        for norm in (True, False):
            for trunc in (True, False):
                a = ref.make_sdf(mask, roi, norm, trunc, 0.1)
                b = port.make_sdf(mask, roi, norm, trunc, 0.1)
                assert np.array_equal(a.view(np.uint32), b.view(np.uint32)), (kind, t, norm, trunc)


@pytest.mark.parametrize("cfg", [
    dict(update=0, interp=1), dict(update=0, interp=0), dict(update=1, interp=1, truncation=True),
    dict(update=1, interp=1, truncation=True, weight=0.7, max_update_num=3),
    dict(update=0, interp=1, outside=1), dict(update=1, interp=0, outside=1, normalize=False),
])
def test_fusion_and_extraction_vs_reference(ref, port, cfg):
    """Synthetic dumbbell, pinhole + one ortho view, every option combination, ROI sub-window."""
    o = S.grid_option(24)
    opt = B.make_options(o.bb_min, o.bb_max, o.resolution, **cfg)
    cr, cp = ref.carver(opt), port.carver(opt)
    assert cr.dims == cp.dims
    for a, b in zip(cr.positions(), cp.positions()):
        assert np.array_equal(a.view(np.uint32), b.view(np.uint32))
    cams = [H.to_oracle_camera(c) for c in S.circle_cameras(7, 96, tilt=0.5)]
    masks = [S.silhouette(c, S.DUMBBELL) for c in S.circle_cameras(7, 96, tilt=0.5)]
    # an orthographic view: image coordinates are camera coordinates (camera.cc:201-205)
    ortho_c2w = np.eye(4)
    ortho_c2w[:3, 3] = (-0.9, -0.8, -3.0)
    cams.append(B.Camera(96, 96, ortho_c2w, model=1))
    masks.append(np.zeros((96, 96), np.uint8))
    masks[-1][:2, :2] = 255
    for i, (cam, mask) in enumerate(zip(cams, masks)):
        roi = (5, 3, 90, 88) if i == 2 else None
        a = cr.carve_mask(cam, mask, roi)
        b = cp.carve_mask(cam, mask, roi)
        assert np.array_equal(a.view(np.uint32), b.view(np.uint32))
        (sa, na), (sb, nb) = cr.download(), cp.download()
        assert np.array_equal(na, nb), i
        assert np.array_equal(sa.view(np.uint32), sb.view(np.uint32)), i
    for iso, interp in ((0.0, True), (0.0, False), (0.03, True)):
        H.assert_mesh_equal(cp.extract_iso(iso, interp), cr.extract_iso(iso, interp), o.bb_min,
                            o.bb_max)
    for inside_empty in (False, True):
        rv, rf = cr.extract_voxel(inside_empty)
        pv, pf = cp.extract_voxel(inside_empty)
        assert np.array_equal(rf, pf)
        # the reference translates one cube back and forth (extract_voxel.cc:292-311): tolerance
        assert np.abs(rv - pv).max() <= 1e-5 * np.linalg.norm(np.array(o.bb_max) - o.bb_min)


def test_marching_cubes_on_handmade_grids_vs_reference(ref, port):
    """Random sdf / update_num states exercise skipped cubes, kVal corners and near-iso values,
    i.e. the map-ordering paths the carving scenarios rarely reach."""
    rng = np.random.default_rng(7)
    opt = B.make_options((0, 0, 0), (1.3, 1.1, 0.9), 0.1)
    cr, cp = ref.carver(opt), port.carver(opt)
    kval = ref.invalid_sdf()
    for t in range(6):
        sdf = rng.normal(size=cr.count).astype(np.float32)
        un = rng.integers(0, 3, cr.count).astype(np.int32)
        sdf[rng.random(cr.count) < 0.15] = kval
        sdf[rng.random(cr.count) < 0.05] = np.float32(1e-6) * rng.integers(-3, 4)
        if t == 0:
            un[:] = 1
        cr.upload(sdf, un)
        cp.upload(sdf, un)
        for iso, interp in ((0.0, True), (0.25, True), (0.0, False)):
            H.assert_mesh_equal(cp.extract_iso(iso, interp), cr.extract_iso(iso, interp),
                                (0, 0, 0), (1.3, 1.1, 0.9))
        for inside_empty in (False, True):
            assert np.array_equal(cr.extract_voxel(inside_empty)[1],
                                  cp.extract_voxel(inside_empty)[1])


def test_sdf_color_map_vs_reference(ref, port, bunny):
    """SignedDistance2Color (voxel_carver.cc:239-267): oracle restatement == compiled reference."""
    rng = np.random.default_rng(3)
    imgs = [bunny.final["sdf2d_view0"], rng.normal(scale=0.7, size=(37, 53)).astype(np.float32),
            np.zeros((4, 5), np.float32)]
    for img in imgs:
        for lo, hi in ((-1.0, 1.0), (-0.3, 2.5)):
            assert np.array_equal(ref.signed_distance_to_color(img, lo, hi),
                                  port.signed_distance_to_color(img, lo, hi))


def test_euclidean_sdf_oracle_vs_scipy_and_brute_force(port):
    """The opt-in Euclidean metric has no counterpart in the reference; its checker is pinned against
    scipy.ndimage.distance_transform_edt (full-image ROI) and an exhaustive search (ROI windows)."""
    from scipy import ndimage
    rng = np.random.default_rng(21)
    for kind in ("blobs", "noise", "grey"):
        for _ in range(6):
            h, w = int(rng.integers(2, 70)), int(rng.integers(2, 70))
            mask = H.random_mask(rng, h, w, kind)
            fg = mask == 255
            if fg.all() or not fg.any():
                continue
            want = np.where(fg, -ndimage.distance_transform_edt(fg),
                            ndimage.distance_transform_edt(~fg)).astype(np.float32)
            got = port.make_sdf_l2(mask, None, False, False, 0.1)
            assert np.array_equal(got.view(np.uint32), want.view(np.uint32)), kind
    for _ in range(10):  # ROI windows, exhaustive
        h, w = int(rng.integers(3, 24)), int(rng.integers(3, 24))
        mask = H.random_mask(rng, h, w, "noise")
        x0, y0 = int(rng.integers(0, w - 1)), int(rng.integers(0, h - 1))
        x1, y1 = int(rng.integers(x0, w)), int(rng.integers(y0, h))
        got = port.make_sdf_l2(mask, (x0, y0, x1, y1), False, False, 0.1)
        want = np.zeros((h, w), np.float32)
        ys, xs = np.mgrid[y0:y1 + 1, x0:x1 + 1]
        fg = mask[y0:y1 + 1, x0:x1 + 1] == 255
        for y in range(y0, y1 + 1):
            for x in range(x0, x1 + 1):
                me = mask[y, x] == 255
                other = fg != me
                if other.any():
                    d2 = ((ys - y) ** 2 + (xs - x) ** 2)[other].min()
                    e = np.float32(np.sqrt(np.float64(d2)))
                else:
                    e = np.finfo(np.float32).max
                want[y, x] = -e if me else e
        assert np.array_equal(got.view(np.uint32), want.view(np.uint32))


def test_slab_oracle_equals_the_planes_of_the_whole_grid_oracle(port):
    """The 64-bit slab restatement used where the reference overflows (config 5): on a grid small
    enough for the whole-grid oracle - itself bit-identical to the compiled reference, above - the
    slab carver must hold exactly the planes it covers, for grid state and occupancy."""
    from vacancy_b200 import synthetic as S
    o = S.grid_option(40, weighted=True)
    opt = B.make_options(o.bb_min, o.bb_max, o.resolution, update=1, truncation=True, band=0.1)
    cams = S.circle_cameras(7, 120, tilt=0.3)
    whole, slab = port.carver(opt), port.slab_carver(opt, 13, 29)
    for c in cams:
        m = S.silhouette(c, S.DUMBBELL)
        whole.carve_mask(H.to_oracle_camera(c), m)
        slab.carve_mask(H.to_oracle_camera(c), m)
    (ws, wn), (ss, sn) = whole.download(), slab.download()
    pl = 40 * 40
    assert np.array_equal(ws[13 * pl:29 * pl].view(np.uint32), ss.view(np.uint32))
    assert np.array_equal(wn[13 * pl:29 * pl], sn)
    assert len(slab.extract_iso()[0]) > 0
