"""GPU parity of the PLY text path (F2) through the C ABI: bytes equal to the reference's
Mesh::WritePly (mesh.cc:583-631) - checked against the oracle's restatement and, when present, the
compiled reference itself."""
import os
import tempfile

import numpy as np
import pytest

from tests import helpers as H
from oracle import bindings as B
from tests.test_ply_format_host import _mesh

pytestmark = pytest.mark.gpu


def test_ply_format_free_function(cuda_lib, port):
    from vacancy_b200.carver import ply_format
    v, f = _mesh(seed=11, n=200000)
    assert ply_format(v, f) == port.ply_text(v, f)
    empty_v, empty_f = np.zeros((0, 3), np.float32), np.zeros((0, 3), np.int32)
    assert ply_format(empty_v, empty_f) == port.ply_text(empty_v, empty_f)
    assert ply_format(v[:5], empty_f) == port.ply_text(v[:5], empty_f)


def test_every_exponent_and_rounding_ties(cuda_lib, port):
    """All 256 binary exponents x sparse / dense mantissas, both signs, plus the neighbours of the
    six-digit rounding boundaries."""
    from vacancy_b200.carver import ply_format
    rng = np.random.default_rng(5)
    expo = np.repeat(np.arange(256, dtype=np.uint32), 3000)
    mant = rng.integers(0, 1 << 23, expo.size, dtype=np.uint32)
    mant[::3] &= np.uint32(0x7FF000)
    mant[1::7] |= np.uint32(0x7FFFF0)
    sign = rng.integers(0, 2, expo.size, dtype=np.uint32)
    bits = (sign << np.uint32(31)) | (expo << np.uint32(23)) | mant
    ties = []
    for j in range(-40, 38):
        for k in (1, 5, 15, 25, 99999, 100000, 999999, 123456):
            with np.errstate(over="ignore"):
                centre = np.float32((k + 0.5) * 10.0 ** j)
            b = int(centre.view(np.uint32)) if np.isfinite(centre) else 0
            ties.extend(b + d for d in range(-2, 3) if 0 <= b + d < (1 << 31))
    bits = np.concatenate([bits, np.array(ties, np.uint32)])
    bits = np.concatenate([bits, np.zeros((-bits.size) % 3, np.uint32)])
    v = bits.view(np.float32).reshape(-1, 3)
    f = np.zeros((1, 3), np.int32)
    assert ply_format(v, f) == port.ply_text(v, f)


def test_bunny_iso_surface_ply(cuda_lib, port, bunny):
    """examples.cc:143-144: ExtractIsoSurface + WritePly of the carved bunny."""
    from tests.test_gpu_parity import make_carver
    opt = H.oracle_options(bunny.bb_min, bunny.bb_max, H.BUNNY_RESOLUTION)
    car = make_carver(opt)
    cams = [H.product_camera(H.bunny_camera(bunny, i), bunny.w2c[i]) for i in range(6)]
    assert car.carve_batch(cams, [bunny.masks[i] for i in range(6)])
    text = car.ExtractIsoSurfacePly(0.0, True)
    v, f = car.ExtractIsoSurface(0.0, True)
    assert len(v) > 1000
    assert text == port.ply_text(v, f)
    if os.path.exists(B.REF_SO):
        ref = B.RefLib()
        rc = ref.carver(opt)
        for i in range(6):
            rc.carve_mask(H.bunny_camera(bunny, i), bunny.masks[i])
        rc.extract_iso(0.0, True)
        with tempfile.TemporaryDirectory() as d:
            path = os.path.join(d, "surface.ply")
            assert rc.write_ply(path) == 0
            with open(path, "rb") as fh:
                assert text == fh.read()


def test_slab_text_blocks_assemble_to_the_unsharded_file(cuda_lib, port, bunny):
    """z slabs: header for the global counts + every slab's vertex block + every slab's face block
    is the file of the unsharded mesh."""
    from vacancy_b200 import sharded
    opt = H.product_options(H.oracle_options(bunny.bb_min, bunny.bb_max, H.BUNNY_RESOLUTION))
    cams = [H.product_camera(H.bunny_camera(bunny, i), bunny.w2c[i]) for i in range(6)]
    slabs = sharded.LocalSlabs(opt, 3)
    slabs.carve_batch(cams, [bunny.masks[i] for i in range(6)])
    v, f = slabs.extract_iso(0.0, True)
    assert slabs.ply_text() == port.ply_text(v, f)
