"""GPU test of vcb_make_signed_distance_field_noalloc, the one entry point that allocates nothing.

It exists for hosts that cannot let a library allocate device memory while they run: programs
whose own memory arena already owns the GPU (a driver allocation could fail or fragment what is
left) and capture loops that cannot tolerate the device-wide synchronisation an allocation brings.
Such a caller hands in device buffers for the mask and the result; the intermediate images live in
a fixed array inside the library (2048 x 2048 pixels at most, the width counted rounded up to 32).

Checked here, through the C ABI, on calls in a row that reuse that array: the result must be
bit-identical to the allocating entry point vcb_make_signed_distance_field and to the CPU oracle's
MakeSignedDistanceField (the tolerance of every other parity test in this suite is exact equality
as well), and the amount of free device memory must be the same before and after a call.  The second call is smaller than the
first and uses other options, so stale contents of the shared array would show.  Runs last in the
suite (file name) and also passes alone in a fresh process, where its first call performs the
per-device set-up.
"""
import ctypes as C

import numpy as np
import pytest

from vacancy_b200 import capi, make_signed_distance_field
from vacancy_b200 import synthetic as S

pytestmark = pytest.mark.gpu


def noalloc_sdf(lib, torch, mask, roi, normalize, truncate, band):
    h, w = mask.shape
    d_mask = torch.from_numpy(mask).cuda()
    d_out = torch.full((h, w), float("nan"), dtype=torch.float32, device="cuda")
    rmin = (C.c_int32 * 2)(roi[0], roi[1])
    rmax = (C.c_int32 * 2)(roi[2], roi[3])
    stream = torch.cuda.current_stream().cuda_stream
    rc = lib.vcb_make_signed_distance_field_noalloc(
        0, C.c_void_p(d_mask.data_ptr()), w, h, rmin, rmax, int(normalize), int(truncate),
        C.c_float(band), C.c_void_p(d_out.data_ptr()), C.c_void_p(stream))
    assert rc == 0, lib.vcb_last_error(None)
    torch.cuda.synchronize()
    return d_out.cpu().numpy()


def test_noalloc_sdf_matches_allocating_entry_point_twice(port):
    import torch
    lib = capi.load_library()
    cam = S.circle_cameras(5, 900)[2]
    big = S.silhouette(cam, S.DUMBBELL)[:700, :]
    rng = np.random.default_rng(5)
    small = (rng.random((77, 123)) < 0.4).astype(np.uint8) * 255
    cases = [
        (big, (0, 0, big.shape[1] - 1, big.shape[0] - 1), True, True, 0.1),
        (small, (3, 2, 110, 70), True, False, 0.1),
        (big, (0, 0, big.shape[1] - 1, big.shape[0] - 1), False, True, 0.25),
    ]
    for mask, roi, normalize, truncate, band in cases:
        got = noalloc_sdf(lib, torch, mask, roi, normalize, truncate, band)
        want = make_signed_distance_field(mask, roi[:2], roi[2:], normalize, truncate, band)
        assert np.array_equal(got.view(np.uint32), want.view(np.uint32))
        # and directly against the checker, not only against the library's other entry point
        oracle = port.make_sdf(mask, roi, normalize, truncate, band)
        assert np.array_equal(got.view(np.uint32), oracle.view(np.uint32))


def test_noalloc_sdf_leaves_free_device_memory_unchanged():
    """The point of the entry point: a call (after the first one on the device, which resolves the
    array's address) must not change what cudaMemGetInfo reports as free, and must leave the
    caller's current device alone."""
    import torch
    lib = capi.load_library()
    rng = np.random.default_rng(11)
    mask = (rng.random((480, 640)) < 0.5).astype(np.uint8) * 255
    d_mask = torch.from_numpy(mask).cuda()
    d_out = torch.empty((480, 640), dtype=torch.float32, device="cuda")
    rmin, rmax = (C.c_int32 * 2)(0, 0), (C.c_int32 * 2)(639, 479)
    stream = torch.cuda.current_stream().cuda_stream

    def call():
        rc = lib.vcb_make_signed_distance_field_noalloc(
            0, C.c_void_p(d_mask.data_ptr()), 640, 480, rmin, rmax, 1, 1, C.c_float(0.1),
            C.c_void_p(d_out.data_ptr()), C.c_void_p(stream))
        assert rc == 0, lib.vcb_last_error(None)
        torch.cuda.synchronize()

    call()  # set-up call
    free_before, _ = torch.cuda.mem_get_info()
    dev_before = torch.cuda.current_device()
    call()
    call()
    free_after, _ = torch.cuda.mem_get_info()
    assert free_after == free_before
    assert torch.cuda.current_device() == dev_before


def test_noalloc_sdf_rejects_images_beyond_the_fixed_scratch():
    import torch
    lib = capi.load_library()
    h, w = 2048, 2049  # one column more than the 2048 x 2048 pixel limit
    d_mask = torch.zeros((h, w), dtype=torch.uint8, device="cuda")
    d_out = torch.zeros((h, w), dtype=torch.float32, device="cuda")
    rmin, rmax = (C.c_int32 * 2)(0, 0), (C.c_int32 * 2)(w - 1, h - 1)
    rc = lib.vcb_make_signed_distance_field_noalloc(
        0, C.c_void_p(d_mask.data_ptr()), w, h, rmin, rmax, 1, 0, C.c_float(0.1),
        C.c_void_p(d_out.data_ptr()), None)
    assert rc == -1
    assert b"exceeds the fixed scratch" in lib.vcb_last_error(None)
    # the limit itself is accepted
    d_mask = torch.zeros((2048, 2048), dtype=torch.uint8, device="cuda")
    d_mask[100:900, 300:1500] = 255
    got = noalloc_sdf(lib, torch, d_mask.cpu().numpy(), (0, 0, 2047, 2047), True, False, 0.1)
    want = make_signed_distance_field(d_mask.cpu().numpy())
    assert np.array_equal(got.view(np.uint32), want.view(np.uint32))
