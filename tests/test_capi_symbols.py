"""CPU-only: the C-ABI library builds for sm_100a, loads without a GPU and exports every symbol that
include/vacancy_b200.h declares; option validation (no device needed) follows the reference."""
import ctypes as C
import os
import re

import pytest

from vacancy_b200 import build, capi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib():
    build.build_library()
    return capi.load_library()


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "vacancy_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(vcb_[a-z0-9_]+)\s*\(", text)))


def test_header_symbols_exported(lib):
    names = declared_symbols()
    assert len(names) >= 27
    raw = C.CDLL(capi.LIB_PATH)
    for n in names:
        assert hasattr(raw, n), "missing export: " + n
    assert sorted(capi.EXPORTS) == [n for n in names if n in capi.EXPORTS]
    assert set(names) == set(capi.EXPORTS), set(names) ^ set(capi.EXPORTS)


def test_struct_layout(lib):
    # sizeof(vcb_options) == 76, sizeof(vcb_view) == 92 (checked against a C compile of the header)
    assert C.sizeof(capi.Options) == 76
    assert C.sizeof(capi.View) == 92
    o = capi.Options()
    lib.vcb_options_default(C.byref(o))
    # voxel_carver.h:43-60 defaults
    assert (o.voxel_update, o.sdf_interp, o.update_outside, o.voxel_max_update_num) == (0, 1, 0, 255)
    assert abs(o.resolution - 0.1) < 1e-7 and o.sdf_minmax_normalize == 1
    assert abs(o.voxel_update_weight - 1.0) < 1e-7 and o.use_truncation == 0
    assert abs(o.truncation_band - 0.1) < 1e-7
    assert lib.vcb_invalid_sdf() == -3.4028234663852886e+38


@pytest.mark.parametrize("field,value", [("voxel_max_update_num", 0), ("voxel_update_weight", 0.0),
                                         ("truncation_band", -1.0), ("resolution", 0.0)])
def test_init_rejects_what_the_reference_rejects(lib, field, value):
    """voxel_carver.cc:375-389 and :277-286: these fail before any device is touched."""
    o = capi.Options()
    lib.vcb_options_default(C.byref(o))
    o.bb_max[:] = [1, 1, 1]
    o.bb_min[:] = [-1, -1, -1]
    setattr(o, field, value)
    h = C.c_void_p()
    assert lib.vcb_create(C.byref(o), C.byref(h)) == -1
    assert lib.vcb_last_error(None)


def test_empty_bbox_rejected(lib):
    o = capi.Options()
    lib.vcb_options_default(C.byref(o))
    o.bb_max[:] = [1, 1, 1]
    o.bb_min[:] = [1, -1, -1]
    h = C.c_void_p()
    assert lib.vcb_create(C.byref(o), C.byref(h)) == -1


def test_no_cpu_fallback_without_device(lib):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    o = capi.Options()
    lib.vcb_options_default(C.byref(o))
    o.bb_max[:] = [1, 1, 1]
    o.bb_min[:] = [-1, -1, -1]
    h = C.c_void_p()
    assert lib.vcb_create(C.byref(o), C.byref(h)) == -2  # VCB_ERR_CUDA, not a silent CPU path
    assert b"no CPU fallback" in lib.vcb_last_error(None)
