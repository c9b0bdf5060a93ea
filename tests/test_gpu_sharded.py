"""GPU tests of the z-slab engine behind vcb_sharded_* (csrc/sharded.cu) on ONE GPU: the device is
listed several times, so every slab, every split of the 2D SDF work, every gather and every
halo / count / vertex-id exchange of the multi-GPU algorithm runs - with peer copies that happen to
stay on one device.  The checker is the CPU oracle on the whole, unsharded grid: the slabs together
must reproduce its grid, its occupancy and its mesh bit for bit (same vertex numbering, same face
order).  The NCCL transport of the one-process-per-GPU mode shares all of this code except the
exchange calls; bench.py checks it by content (mesh sha256) on real multi-GPU runs."""
import numpy as np
import pytest

from oracle import bindings as B
from tests import helpers as H
from vacancy_b200 import VoxelCarver, sharded
from vacancy_b200 import synthetic as S

pytestmark = pytest.mark.gpu


def bits_equal(a, b):
    return a.shape == b.shape and np.array_equal(a.view(np.uint32), b.view(np.uint32))


def scene(n=96, views=70, size=256, weighted=True, tilt=0.3):
    o = S.grid_option(n, weighted=weighted)
    opt = B.make_options(o.bb_min, o.bb_max, o.resolution, update=int(weighted),
                         truncation=weighted, band=0.1)
    cams = S.circle_cameras(views, size, tilt=tilt)
    masks = [S.silhouette(c, S.DUMBBELL) for c in cams]
    return o, opt, cams, masks


def oracle_carve(port, opt, cams, masks):
    orc = port.carver(opt)
    for c, m in zip(cams, masks):
        orc.carve_mask(H.to_oracle_camera(c), m)
    return orc


@pytest.mark.parametrize("n_slabs", [1, 2, 3, 8])
def test_engine_equals_oracle_more_views_than_one_chunk(cuda_lib, port, n_slabs):
    """70 views = two fusion launches (the second chunk's transforms run under the first chunk's
    fusion), slab counts that do and do not divide the chunk of 64, host masks."""
    o, opt, cams, masks = scene()
    orc = oracle_carve(port, opt, cams, masks)
    eng = sharded.LocalSlabs(o, n_slabs)
    assert eng.carve_batch(cams, masks)
    gs, gn = eng.download_grid()
    ws, wn = orc.download()
    assert np.array_equal(gn, wn) and bits_equal(gs, ws)
    for iso, interp in ((0.0, True), (0.02, False)):
        H.assert_mesh_equal(eng.extract_iso(iso, interp), orc.extract_iso(iso, interp), o.bb_min,
                            o.bb_max)
    # a second pass after reset reuses both halves of every pool
    eng.reset()
    assert eng.carve_batch(cams[:10], masks[:10])
    orc2 = oracle_carve(port, opt, cams[:10], masks[:10])
    gs, gn = eng.download_grid()
    ws, wn = orc2.download()
    assert np.array_equal(gn, wn) and bits_equal(gs, ws)


def test_engine_view_by_view_with_sdf_out_and_mixed_sizes(cuda_lib, port):
    """Views arriving one per call (the reference's Carve(camera, mask, &sdf) pattern), some asking
    for their 2D SDF back, with image sizes that grow mid-stream (the pools are re-allocated), on
    unequal slabs."""
    import ctypes as C
    from vacancy_b200 import capi
    o = S.grid_option(64, weighted=True)
    opt = B.make_options(o.bb_min, o.bb_max, o.resolution, update=1, truncation=True, band=0.1)
    eng = sharded.LocalSlabs(o, 3, ranges=[(0, 10), (10, 40), (40, 64)])
    orc = port.carver(opt)
    sizes = [128] * 5 + [200] * 4 + [160] * 3
    for i, size in enumerate(sizes):
        cam = S.circle_cameras(len(sizes), size, tilt=0.25)[i]
        mask = S.silhouette(cam, S.DUMBBELL)
        view = sharded.make_view(cam, size, size)
        want_sdf = i % 3 == 0
        out = np.full((size, size), np.nan, np.float32) if want_sdf else None
        views = (capi.View * 1)(view)
        masks = (C.c_void_p * 1)(mask.ctypes.data)
        outs = (C.c_void_p * 1)(out.ctypes.data) if want_sdf else None
        eng._check(eng._lib.vcb_sharded_carve_masks(eng._h, 1, views, masks, outs))
        orc.carve_mask(H.to_oracle_camera(cam), mask)
        if want_sdf:
            assert bits_equal(out, port.make_sdf(mask, None, True, True, 0.1))
    gs, gn = eng.download_grid()
    ws, wn = orc.download()
    assert np.array_equal(gn, wn) and bits_equal(gs, ws)
    H.assert_mesh_equal(eng.extract_iso(), orc.extract_iso(), o.bb_min, o.bb_max)


@pytest.mark.parametrize("n_slabs", [2, 5])
def test_engine_extract_voxel_and_inside_empty_across_slabs(cuda_lib, port, n_slabs):
    """ExtractVoxel over slabs: the default rule and the inside_empty surface rule (which looks one
    plane down, i.e. into the halo below every slab) give the oracle's emitted set; the cube mesh
    equals the one a single carver writes (numbered over the whole grid)."""
    o, opt, cams, masks = scene(n=64, views=9, size=192, weighted=False)
    orc = oracle_carve(port, opt, cams, masks)
    eng = sharded.LocalSlabs(o, n_slabs)
    assert eng.carve_batch(cams, masks)
    one = VoxelCarver(o)
    assert one.Init() and one.carve_batch(cams, masks)
    for inside_empty in (False, True):
        flags, n = eng.occupancy(inside_empty)
        wflags, wn = orc.occupancy(inside_empty)
        assert n == wn and np.array_equal(flags, wflags), inside_empty
        gv, gf = eng.extract_voxel(inside_empty)
        sv, sf = one.ExtractVoxel(inside_empty)
        assert np.array_equal(gf, sf) and bits_equal(gv, sv)


def test_engine_repartition_and_errors(cuda_lib, port):
    o, opt, cams, masks = scene(n=64, views=6, size=128)
    orc = oracle_carve(port, opt, cams, masks)
    eng = sharded.LocalSlabs(o, 4)
    assert eng.ranges == [(0, 16), (16, 32), (32, 48), (48, 64)]
    assert eng.carve_batch(cams[:2], masks[:2])
    import ctypes as C
    new = [(0, 8), (8, 40), (40, 42), (42, 64)]
    eng.slabs = []
    eng._check(eng._lib.vcb_sharded_repartition(eng._h, sharded._bounds_array(new)))
    eng._refresh()
    assert eng.ranges == new
    assert eng.carve_batch(cams, masks)  # the grid started over, like Init()
    gs, gn = eng.download_grid()
    ws, wn = orc.download()
    assert np.array_equal(gn, wn) and bits_equal(gs, ws)
    H.assert_mesh_equal(eng.extract_iso(), orc.extract_iso(), o.bb_min, o.bb_max)
    # a slab of one plane cannot serve its neighbour's halo
    bad = (C.c_int32 * 5)(0, 1, 32, 48, 64)
    assert eng._lib.vcb_sharded_repartition(eng._h, bad) == -1
    assert b"at least 2" in eng._lib.vcb_sharded_last_error(eng._h)
    # fill before count
    eng2 = sharded.LocalSlabs(o, 2)
    assert eng2._lib.vcb_sharded_extract_iso_fill(eng2._h, None, None) == -3
    with pytest.raises(Exception):
        sharded.LocalSlabs(o, 2, devices=[0, 99])


def test_checkpoint_restore_then_carve_equals_uninterrupted(cuda_lib, port, tmp_path):
    """vcb_save_grid / vcb_load_grid (and the vcb_sharded_* pair): carve 40 views on 3 slabs, save,
    restore into (a) a single carver and (b) 2 slabs on other boundaries, carve 35 more views -
    grid, update_num and mesh must equal the oracle's uninterrupted 75-view carve bit for bit.  In
    weighted mode update_num is the running weight, so a wrong restore would show."""
    o, opt, cams, masks = scene(n=64, views=75, size=160)
    orc = oracle_carve(port, opt, cams, masks)
    ws, wn = orc.download()
    first = sharded.LocalSlabs(o, 3)
    assert first.carve_batch(cams[:40], masks[:40])
    paths = first.save_grid(tmp_path / "ckpt")
    assert len(paths) == 3
    first.close()
    one = VoxelCarver(o)
    assert one.Init()
    assert sum(one.load_grid(p) for p in paths) == 64
    assert one.carve_batch(cams[40:], masks[40:])
    gs, gn = one.download_grid()
    assert np.array_equal(gn, wn) and bits_equal(gs, ws)
    H.assert_mesh_equal(one.ExtractIsoSurface(), orc.extract_iso(), o.bb_min, o.bb_max)
    two = sharded.LocalSlabs(o, 2, ranges=[(0, 27), (27, 64)])
    two.load_grid(paths)
    assert two.carve_batch(cams[40:], masks[40:])
    gs, gn = two.download_grid()
    assert np.array_equal(gn, wn) and bits_equal(gs, ws)
    H.assert_mesh_equal(two.extract_iso(), orc.extract_iso(), o.bb_min, o.bb_max)
    # files of another grid / other carving options, and incomplete sets, are refused
    other = VoxelCarver(S.grid_option(32, weighted=True))
    assert other.Init()
    with pytest.raises(Exception, match="another grid"):
        other.load_grid(paths[0])
    kmax = VoxelCarver(S.grid_option(64, weighted=False))
    assert kmax.Init()
    with pytest.raises(Exception, match="another grid"):
        kmax.load_grid(paths[0])
    with pytest.raises(Exception, match="cover"):
        sharded.LocalSlabs(o, 2).load_grid(paths[:2])


def test_fusion_class_counts(cuda_lib):
    """vcb_fusion_class_counts: every (tile, view) pair of every launch lands in exactly one class."""
    o, opt, cams, masks = scene(n=64, views=70, size=256)
    car = VoxelCarver(o)
    assert car.Init()
    car.set_statistics(True)
    assert car.carve_batch(cams, masks)
    car.synchronize()
    c = car.fusion_class_counts()
    tiles = (64 // 32) * (64 // 8) * (64 // 8)
    assert sum(c[k] for k in VoxelCarver.TILE_CLASSES) == tiles * 70
    assert c["voxels_per_tile"] == 2048
    assert c["all_ones"] + c["all_kval"] + c["outside"] > 0 and c["inside"] + c["general"] > 0
    assert c["warp_iterations"]["all_ones"] + c["warp_iterations"]["sampling"] > 0
    assert sum(car.fusion_class_counts()[k] for k in VoxelCarver.TILE_CLASSES) == 0  # cleared


def test_carve_mask_files_single_and_slabs(cuda_lib, port, tmp_path):
    """vcb_carve_mask_files / vcb_sharded_carve_mask_files: silhouettes arriving as PNG and PGM files
    (decoded on host threads inside the library, 70 files = two batches) must give the grid and the
    mesh of the same masks handed over as arrays - here checked against the oracle."""
    from tests.test_ingest_host import write_png
    o, opt, cams, masks = scene(n=64, views=70, size=160)
    orc = oracle_carve(port, opt, cams, masks)
    ws, wn = orc.download()
    paths = []
    for i, m in enumerate(masks):
        if i % 5 == 4:
            p = tmp_path / ("m%03d.pgm" % i)
            with open(p, "wb") as f:
                f.write(b"P5\n160 160\n255\n" + m.tobytes())
        else:
            p = tmp_path / ("m%03d.png" % i)
            write_png(p, m, filters=(i % 5,), level=1 + i % 9)
        paths.append(p)
    one = VoxelCarver(o)
    assert one.Init() and one.carve_files(cams, paths, (160, 160))
    gs, gn = one.download_grid()
    assert np.array_equal(gn, wn) and bits_equal(gs, ws)
    eng = sharded.LocalSlabs(o, 3)
    assert eng.carve_files(cams, paths, (160, 160), n_threads=3)
    gs, gn = eng.download_grid()
    assert np.array_equal(gn, wn) and bits_equal(gs, ws)
    H.assert_mesh_equal(eng.extract_iso(), orc.extract_iso(), o.bb_min, o.bb_max)
    # a file whose size differs from what the view says is refused
    with pytest.raises(Exception, match="the view says"):
        one.carve_files(cams[:1], paths[:1], (200, 200))
    with pytest.raises(Exception, match="larger than the buffer"):
        one.carve_files(cams[:1], paths[:1], (128, 128))


def test_engine_on_two_real_devices_when_present(cuda_lib, port):
    """ONE process driving two different GPUs (peer copies over NVLink between the devices' pools,
    cross-device event waits): same checks as with virtual slabs.  Skipped on a single-GPU box."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two CUDA devices")
    o, opt, cams, masks = scene(n=96, views=70, size=256)
    orc = oracle_carve(port, opt, cams, masks)
    ws, wn = orc.download()
    for devices in ([0, 1], [1, 0, 1]):
        eng = sharded.LocalSlabs(o, len(devices), devices=devices)
        assert eng.carve_batch(cams, masks)
        gs, gn = eng.download_grid()
        assert np.array_equal(gn, wn) and bits_equal(gs, ws)
        H.assert_mesh_equal(eng.extract_iso(), orc.extract_iso(), o.bb_min, o.bb_max)
        flags, n = eng.occupancy(True)
        wflags, wcount = orc.occupancy(True)
        assert n == wcount and np.array_equal(flags, wflags)
        eng.close()


def test_engine_auto_balance_places_boundaries_and_keeps_the_result(cuda_lib, port, tmp_path):
    """vcb_sharded_set_auto_balance: the boundaries move (whole fusion tiles of 8 planes, the ends of
    the grid get thicker slabs than the middle, which sees the object) before the first chunk is
    fused - and the result is still the oracle's, bit for bit."""
    from vacancy_b200 import PinholeCamera
    n = 256
    o = S.grid_option(n, weighted=True)
    opt = B.make_options(o.bb_min, o.bb_max, o.resolution, update=1, truncation=True, band=0.1)
    # long lenses on a turntable: the images cover only the middle of the grid, so the tiles of the
    # first and last z layers project outside the image in most views (skipped, cheap) and the
    # middle layers carry the work - the end slabs must come out thicker
    cams = [PinholeCamera(512, 512, c.c2w(), c.principal_point(), (3.5 * 512, 3.5 * 512))
            for c in S.circle_cameras(70, 512, tilt=0.2)]
    masks = [S.silhouette(c, S.DUMBBELL) for c in cams]
    orc = oracle_carve(port, opt, cams, masks)
    ws, wn = orc.download()
    eng = sharded.LocalSlabs(o, 4)
    equal = [(0, 64), (64, 128), (128, 192), (192, 256)]
    assert eng.ranges == equal
    eng.set_auto_balance(True)
    assert eng.carve_batch(cams, masks)
    moved = eng.ranges
    assert moved != equal
    assert all(a % 8 == 0 and b % 8 == 0 and b - a >= 8 for a, b in moved)
    assert moved[0][0] == 0 and moved[-1][1] == n
    assert moved[0][1] - moved[0][0] > moved[1][1] - moved[1][0]  # the cheap end gets more planes
    assert moved[3][1] - moved[3][0] > moved[2][1] - moved[2][0]
    gs, gn = eng.download_grid()
    assert np.array_equal(gn, wn) and bits_equal(gs, ws)
    H.assert_mesh_equal(eng.extract_iso(), orc.extract_iso(), o.bb_min, o.bb_max)
    # a later pass keeps the partition
    eng.reset()
    assert eng.carve_batch(cams[:5], masks[:5]) and eng.ranges == moved
    # the same through the file path (the sample files are decoded first)
    from tests.test_ingest_host import write_png
    paths = []
    for i, m in enumerate(masks):
        paths.append(tmp_path / ("m%03d.png" % i))
        write_png(paths[-1], m, filters=(0,), level=1)
    eng2 = sharded.LocalSlabs(o, 4)
    eng2.set_auto_balance(True)
    assert eng2.carve_files(cams, paths, (512, 512)) and eng2.ranges == moved
    gs, gn = eng2.download_grid()
    assert np.array_equal(gn, wn) and bits_equal(gs, ws)
