"""Python mirror of the reference's carving interface (include/vacancy/voxel_carver.h:20-128,
include/vacancy/camera.h) on top of the C ABI.  Same names, argument meaning and error behaviour
as the reference: `Init()` and `Carve()` return bool, options carry the reference's defaults,
`Carve` before `Init` returns False.  numpy arrays stand in for Image1b / Image1f / Mesh.

This is host-side glue only: every computation happens in libvacancy_b200.so on the GPU.
"""
import ctypes as C
import enum
from dataclasses import dataclass, field

import numpy as np

from . import capi


class VoxelUpdate(enum.IntEnum):  # voxel_carver.h:20-24
    kMax = 0
    kWeightedAverage = 1


class SdfInterpolation(enum.IntEnum):  # voxel_carver.h:27-30
    kNn = 0
    kBilinear = 1


class UpdateOutsideImage(enum.IntEnum):  # voxel_carver.h:33-37
    kNone = 0
    kMax = 1


class SdfMetric(enum.IntEnum):
    """Metric of the 2D silhouette distance: kL1 is the reference's (voxel_carver.cc:102-167) and
    the default; kL2 (exact Euclidean) is an addition of this library."""
    kL1 = 0
    kL2 = 1


@dataclass
class VoxelUpdateOption:  # voxel_carver.h:43-52
    voxel_update: VoxelUpdate = VoxelUpdate.kMax
    sdf_interp: SdfInterpolation = SdfInterpolation.kBilinear
    update_outside: UpdateOutsideImage = UpdateOutsideImage.kNone
    voxel_max_update_num: int = 255
    voxel_update_weight: float = 1.0
    use_truncation: bool = False
    truncation_band: float = 0.1


@dataclass
class VoxelCarverOption:  # voxel_carver.h:54-60
    bb_max: tuple = (0.0, 0.0, 0.0)
    bb_min: tuple = (0.0, 0.0, 0.0)
    resolution: float = 0.1
    sdf_minmax_normalize: bool = True
    update_option: VoxelUpdateOption = field(default_factory=VoxelUpdateOption)
    sdf_metric: SdfMetric = SdfMetric.kL1  # not in the reference; kL1 reproduces it


def w2c_from_c2w(c2w):
    """camera.cc:39-42 `w2c_ = c2w_.inverse()` in float64 followed by voxel_carver.cc:438
    `.cast<float>()`.  Follows Eigen 3.4.0's Affine inverse (cofactor 3x3 inverse of the linear
    part, translation = -inv * t) operation for operation, in Python floats (IEEE double, no FMA),
    so the float32 result equals what the reference's Camera hands to the fusion loop."""
    m = [[float(c2w[r][c]) for c in range(4)] for r in range(4)]

    def cof(i, j):
        i1, i2, j1, j2 = (i + 1) % 3, (i + 2) % 3, (j + 1) % 3, (j + 2) % 3
        return m[i1][j1] * m[i2][j2] - m[i1][j2] * m[i2][j1]

    c0 = [cof(0, 0), cof(1, 0), cof(2, 0)]
    det = (c0[0] * m[0][0] + c0[1] * m[1][0]) + c0[2] * m[2][0]
    invdet = 1.0 / det
    inv = [[c0[0] * invdet, c0[1] * invdet, c0[2] * invdet],
           [cof(0, 1) * invdet, cof(1, 1) * invdet, cof(2, 1) * invdet],
           [cof(0, 2) * invdet, cof(1, 2) * invdet, cof(2, 2) * invdet]]
    out = np.empty(12, np.float32)
    for r in range(3):
        t = ((-inv[r][0]) * m[0][3] + (-inv[r][1]) * m[1][3]) + (-inv[r][2]) * m[2][3]
        out[r * 4:r * 4 + 3] = inv[r]
        out[r * 4 + 3] = t
    return out


class Camera:
    """camera.h:17-61: size + pose; `set_c2w` keeps `w2c` in sync (camera.cc:39-42)."""
    model = None

    def __init__(self, width=-1, height=-1, c2w=None):
        self._width, self._height = int(width), int(height)
        self.set_c2w(np.eye(4) if c2w is None else c2w)

    def width(self):
        return self._width

    def height(self):
        return self._height

    def set_size(self, width, height):
        self._width, self._height = int(width), int(height)

    def c2w(self):
        return self._c2w

    def w2c_f32(self):
        return self._w2c

    def set_c2w(self, c2w):
        self._c2w = np.ascontiguousarray(c2w, np.float64).reshape(4, 4)
        self._w2c = w2c_from_c2w(self._c2w)

    def set_w2c_f32(self, w2c12):
        """Use a float32 w2c obtained elsewhere (e.g. from the reference's own Camera)."""
        self._w2c = np.ascontiguousarray(w2c12, np.float32).reshape(12)

    def _intrinsics(self):
        return (0.0, 0.0, 0.0, 0.0)


class PinholeCamera(Camera):  # camera.h:66-111
    model = 0

    def __init__(self, width=-1, height=-1, c2w=None, principal_point=(-1.0, -1.0),
                 focal_length=(-1.0, -1.0)):
        super().__init__(width, height, c2w)
        self._pp = tuple(float(np.float32(v)) for v in principal_point)
        self._fl = tuple(float(np.float32(v)) for v in focal_length)

    def principal_point(self):
        return self._pp

    def focal_length(self):
        return self._fl

    def set_principal_point(self, pp):
        self._pp = tuple(float(np.float32(v)) for v in pp)

    def set_focal_length(self, fl):
        self._fl = tuple(float(np.float32(v)) for v in fl)

    def _intrinsics(self):
        return (self._fl[0], self._fl[1], self._pp[0], self._pp[1])


class OrthoCamera(Camera):  # camera.h:116-135
    model = 1


def make_view(camera, width, height, roi=None):
    if camera.model is None:
        # the reference calls the virtual Camera::Project; only its two concrete models exist on GPU
        raise capi.VcbError(-4, "unsupported camera type %s" % type(camera).__name__)
    v = capi.View()
    v.w2c[:] = [float(x) for x in camera.w2c_f32()]
    v.model = camera.model
    v.fx, v.fy, v.cx, v.cy = camera._intrinsics()
    v.width, v.height = int(width), int(height)
    roi = roi if roi is not None else (0, 0, width - 1, height - 1)
    v.roi_min[:] = [int(roi[0]), int(roi[1])]
    v.roi_max[:] = [int(roi[2]), int(roi[3])]
    return v


def _ptr_array(arrays):
    arr = (C.c_void_p * len(arrays))()
    for i, a in enumerate(arrays):
        arr[i] = None if a is None else (a if isinstance(a, int) else a.ctypes.data)
    return arr


def _fill_options(o, opt, device=0, z_range=None):
    o.bb_max[:] = [float(np.float32(v)) for v in opt.bb_max]
    o.bb_min[:] = [float(np.float32(v)) for v in opt.bb_min]
    o.resolution = opt.resolution
    o.sdf_minmax_normalize = int(bool(opt.sdf_minmax_normalize))
    u = opt.update_option
    o.voxel_update = int(u.voxel_update)
    o.sdf_interp = int(u.sdf_interp)
    o.update_outside = int(u.update_outside)
    o.voxel_max_update_num = int(u.voxel_max_update_num)
    o.voxel_update_weight = u.voxel_update_weight
    o.use_truncation = int(bool(u.use_truncation))
    o.truncation_band = u.truncation_band
    o.device = int(device)
    o.sdf_metric = int(getattr(opt, "sdf_metric", 0))
    if z_range is not None:
        o.z_begin, o.z_end = int(z_range[0]), int(z_range[1])


class VoxelCarver:
    """voxel_carver.h:95-118.  Extra keyword arguments place the carver: `device` (CUDA ordinal)
    and `z_range` (the z slab of the global grid this carver owns, for multi-GPU sharding)."""

    def __init__(self, option=None, device=0, z_range=None):
        self._lib = capi.load_library()
        self._h = None
        self._option = option if option is not None else VoxelCarverOption()
        self._device = device
        self._z_range = z_range
        self.last_error = ""

    @classmethod
    def borrowed(cls, handle, option=None):
        """Wraps a carver handle that somebody else owns (a slab of a ShardedCarver): the state
        access / extraction helpers work on it, close() leaves it alone."""
        self = cls(option)
        self._h, self._owned = C.c_void_p(handle), False
        self._query_geometry()
        return self

    def __del__(self):
        self.close()

    def close(self):
        if getattr(self, "_h", None) and getattr(self, "_owned", True):
            self._lib.vcb_destroy(self._h)
        self._h = None

    def set_option(self, option):  # voxel_carver.cc:373
        self._option = option

    def Init(self):  # voxel_carver.cc:375-392
        self.close()
        o = capi.Options()
        self._lib.vcb_options_default(C.byref(o))
        _fill_options(o, self._option, self._device, self._z_range)
        h = C.c_void_p()
        rc = self._lib.vcb_create(C.byref(o), C.byref(h))
        if rc != 0:
            self.last_error = self._lib.vcb_last_error(None).decode()
            if rc == capi.VcbError(-2, "").code:  # CUDA trouble is never a soft failure
                raise capi.VcbError(rc, self.last_error)
            return False
        self._h = h
        self._owned = True
        self._query_geometry()
        return True

    def _query_geometry(self):
        d = (C.c_int32 * 3)()
        self._lib.vcb_grid_dims(self._h, d)
        self.dims = (d[0], d[1], d[2])
        z0, z1 = C.c_int32(), C.c_int32()
        self._lib.vcb_slab_range(self._h, C.byref(z0), C.byref(z1))
        self.z_range = (z0.value, z1.value)
        self.slab_voxels = self.dims[0] * self.dims[1] * (z1.value - z0.value)

    def _check(self, rc):
        capi.check(self._lib, self._h, rc)

    # ---- Carve overloads (voxel_carver.h:104-114) --------------------------------------------
    def Carve(self, camera, image=None, roi_min=None, roi_max=None, sdf=None, return_sdf=False):
        """Carve(camera, silhouette[, roi_min, roi_max]) with a uint8 image builds the 2D SDF and
        fuses it (voxel_carver.cc:394-413); with a float32 image it fuses the SDF as given
        (:415-496).  `return_sdf=True` returns (ok, sdf_image) like the reference's `Image1f* sdf`
        out-parameter.  Lists of cameras / silhouettes are the batch overload (:516-528)."""
        if isinstance(camera, (list, tuple)):
            return self.carve_batch(camera, image)
        if self._h is None:
            self.last_error = "VoxelCarver::Carve voxel grid has not been initialized"
            return (False, None) if return_sdf else False
        img = np.ascontiguousarray(image)
        h, w = img.shape
        roi = None
        if roi_min is not None:
            roi = (roi_min[0], roi_min[1], roi_max[0], roi_max[1])
        view = make_view(camera, w, h, roi)
        if img.dtype == np.uint8:
            out = np.empty((h, w), np.float32) if return_sdf else None
            self._check(self._lib.vcb_carve_masks(self._h, 1, C.byref(view), _ptr_array([img]),
                                                  _ptr_array([out]) if return_sdf else None))
            return (True, out) if return_sdf else True
        if img.dtype != np.float32:
            raise TypeError("Carve needs a uint8 silhouette or a float32 SDF image")
        self._check(self._lib.vcb_carve_sdfs(self._h, 1, C.byref(view), _ptr_array([img])))
        return (True, img) if return_sdf else True

    def carve_batch(self, cameras, images, rois=None, return_sdf=False):
        """Carve(cameras, silhouettes) (voxel_carver.cc:516-528): all views in one C-ABI call."""
        if self._h is None:
            self.last_error = "VoxelCarver::Carve voxel grid has not been initialized"
            return False
        assert len(cameras) == len(images)
        n = len(cameras)
        if n == 0:
            return True
        imgs = [np.ascontiguousarray(i) for i in images]
        views = (capi.View * n)()
        for i in range(n):
            h, w = imgs[i].shape
            views[i] = make_view(cameras[i], w, h, None if rois is None else rois[i])
        if imgs[0].dtype == np.uint8:
            outs = [np.empty(i.shape, np.float32) for i in imgs] if return_sdf else None
            self._check(self._lib.vcb_carve_masks(self._h, n, views, _ptr_array(imgs),
                                                  _ptr_array(outs) if outs else None))
            return (True, outs) if return_sdf else True
        self._check(self._lib.vcb_carve_sdfs(self._h, n, views, _ptr_array(imgs)))
        return True

    def carve_files(self, cameras, paths, size, n_threads=0):
        """Image1b::Load + Carve for silhouette files (8-bit grey PNG / binary PGM) of `size` =
        (width, height): decoded on host threads inside the library (vcb_carve_mask_files)."""
        n = len(paths)
        views = (capi.View * n)(*[make_view(c, size[0], size[1]) for c in cameras])
        arr = (C.c_char_p * n)(*[str(p).encode() for p in paths])
        self._check(self._lib.vcb_carve_mask_files(self._h, n, views, arr, int(n_threads)))
        return True

    def carve_device(self, views, device_ptrs, masks=True):
        """Stream-ordered fusion of images already resident on the carver's GPU (ints = device
        pointers, e.g. torch tensor.data_ptr())."""
        n = len(device_ptrs)
        arr = (capi.View * n)(*views)
        fn = self._lib.vcb_carve_masks_device if masks else self._lib.vcb_carve_sdfs_device
        if masks:
            self._check(fn(self._h, n, arr, _ptr_array(list(device_ptrs)), None))
        else:
            self._check(fn(self._h, n, arr, _ptr_array(list(device_ptrs))))

    # ---- extraction (voxel_carver.h:115-117) -----------------------------------------------------
    def ExtractIsoSurface(self, iso_level=0.0, linear_interp=True, copy=True):
        """Returns (vertices float32 [V,3], faces int32 [F,3]) numbered as the reference's Mesh."""
        nv, nf = C.c_int64(), C.c_int64()
        self._check(self._lib.vcb_extract_iso_count(self._h, float(iso_level), int(linear_interp),
                                                    C.byref(nv), C.byref(nf)))
        if not copy:
            return nv.value, nf.value
        v = np.empty((nv.value, 3), np.float32)
        f = np.empty((nf.value, 3), np.int32)
        self._check(self._lib.vcb_extract_iso_fill(
            self._h, v.ctypes.data_as(C.POINTER(C.c_float)), f.ctypes.data_as(C.POINTER(C.c_int32))))
        return v, f

    def ExtractIsoSurfacePly(self, iso_level=0.0, linear_interp=True):
        """ExtractIsoSurface + Mesh::WritePly (mesh.cc:583-631) in one go: the bytes of the ASCII PLY
        file the reference would write for this mesh, formatted on the GPU (examples.cc:143-144)."""
        nv, nf = self.ExtractIsoSurface(iso_level, linear_interp, copy=False)
        return self.ply_text(nv, nf)

    def ply_text(self, n_vertices=None, n_faces=None, header=True, split=False):
        """PLY text of the mesh left on the device by the last extraction; with z slabs pass the
        GLOBAL counts for the header (or header=False) - see vcb_ply_body_size.  split=True returns
        (vertex_text, face_text) instead of one string (slabs: the blocks interleave)."""
        vb, fb = C.c_size_t(), C.c_size_t()
        self._check(self._lib.vcb_ply_body_size(self._h, C.byref(vb), C.byref(fb)))
        hb = C.c_size_t(0)
        if header:
            self._lib.vcb_ply_header(int(n_vertices), int(n_faces), None, 0, C.byref(hb))
        out = np.empty(hb.value + vb.value + fb.value, np.uint8)
        if header:
            self._lib.vcb_ply_header(int(n_vertices), int(n_faces),
                                     out.ctypes.data_as(C.c_char_p), out.size, None)
        base = out.ctypes.data
        self._check(self._lib.vcb_ply_body_fill(self._h, C.c_void_p(base + hb.value),
                                                C.c_void_p(base + hb.value + vb.value)))
        if split:
            return out[:hb.value + vb.value].tobytes(), out[hb.value + vb.value:].tobytes()
        return out.tobytes()

    def ExtractVoxel(self, inside_empty=False, copy=True):
        n = C.c_int64()
        self._check(self._lib.vcb_extract_voxel_count(self._h, int(inside_empty), C.byref(n)))
        if not copy:
            return n.value
        v = np.empty((n.value * 24, 3), np.float32)
        f = np.empty((n.value * 12, 3), np.int32)
        if n.value:
            self._check(self._lib.vcb_extract_voxel_fill(
                self._h, v.ctypes.data_as(C.POINTER(C.c_float)),
                f.ctypes.data_as(C.POINTER(C.c_int32))))
        return v, f

    def occupancy(self, inside_empty=False):
        n = C.c_int64()
        self._check(self._lib.vcb_extract_voxel_count(self._h, int(inside_empty), C.byref(n)))
        flags = np.empty(self.slab_voxels, np.uint8)
        self._check(self._lib.vcb_extract_voxel_occupancy(
            self._h, flags.ctypes.data_as(C.POINTER(C.c_uint8))))
        return flags, n.value

    # ---- state access / plumbing --------------------------------------------------------------
    def download_grid(self):
        sdf = np.empty(self.slab_voxels, np.float32)
        un = np.empty(self.slab_voxels, np.int32)
        self._check(self._lib.vcb_download_grid(self._h, sdf.ctypes.data_as(C.POINTER(C.c_float)),
                                                un.ctypes.data_as(C.POINTER(C.c_int32))))
        return sdf, un

    def upload_grid(self, sdf, update_num):
        sdf = np.ascontiguousarray(sdf, np.float32).reshape(-1)
        un = np.ascontiguousarray(update_num, np.int32).reshape(-1)
        assert sdf.size == self.slab_voxels and un.size == self.slab_voxels
        self._check(self._lib.vcb_upload_grid(self._h, sdf.ctypes.data_as(C.POINTER(C.c_float)),
                                              un.ctypes.data_as(C.POINTER(C.c_int32))))

    def positions(self):
        px, py, pz = (np.empty(n, np.float32) for n in self.dims)
        self._check(self._lib.vcb_grid_positions(
            self._h, *(a.ctypes.data_as(C.POINTER(C.c_float)) for a in (px, py, pz))))
        return px, py, pz

    def reset(self):
        """Init() again on the same carver: grid back to its initial state, buffers kept."""
        self._check(self._lib.vcb_reset(self._h))

    def set_stream(self, cuda_stream):
        self._check(self._lib.vcb_set_stream(self._h, C.c_void_p(cuda_stream)))

    def flush(self):
        self._check(self._lib.vcb_flush(self._h))

    def synchronize(self):
        self._check(self._lib.vcb_synchronize(self._h))

    def stage_times(self):
        ms = (C.c_float * 4)()
        n = (C.c_int64 * 4)()
        self._check(self._lib.vcb_stage_times(self._h, ms, n))
        names = ("sdf2d", "fusion", "extract_voxel", "marching_cubes")
        return {k: (ms[i], n[i]) for i, k in enumerate(names)}

    TILE_CLASSES = ("general", "all_ones", "all_kval", "inside", "outside")

    def set_statistics(self, enabled):
        """Counts the view iterations the fusion warps execute (fusion_class_counts) while on."""
        self._check(self._lib.vcb_set_statistics(self._h, int(bool(enabled))))

    def fusion_class_counts(self):
        """How the fusion kernel treated the (voxel tile, view) pairs since the last call:
        {class: pairs} plus "voxels_per_tile" (see vcb_fusion_class_counts)."""
        pairs = (C.c_uint64 * 5)()
        iters = (C.c_uint64 * 2)()
        vpt = C.c_int64()
        self._check(self._lib.vcb_fusion_class_counts(self._h, pairs, C.byref(vpt), iters))
        out = {k: int(pairs[i]) for i, k in enumerate(self.TILE_CLASSES)}
        out["voxels_per_tile"] = vpt.value
        # view iterations the warps executed (256 voxel-views each): load-free path, sampling paths
        out["warp_iterations"] = {"all_ones": int(iters[0]), "sampling": int(iters[1])}
        return out

    def save_grid(self, path):
        """Checkpoint of this carver's slab (vcb_save_grid)."""
        self._check(self._lib.vcb_save_grid(self._h, str(path).encode()))

    def load_grid(self, path):
        """Restores the planes this slab shares with the file; returns how many."""
        n = C.c_int32()
        self._check(self._lib.vcb_load_grid(self._h, str(path).encode(), C.byref(n)))
        return n.value

    def grid_device(self):
        a, b = C.c_void_p(), C.c_void_p()
        self._check(self._lib.vcb_grid_device(self._h, C.byref(a), C.byref(b)))
        return a.value, b.value


def distance_transform_l1(mask, roi_min=None, roi_max=None, device=0):
    """DistanceTransformL1 (voxel_carver.h:120-121) on the GPU."""
    L = capi.load_library()
    mask = np.ascontiguousarray(mask, np.uint8)
    h, w = mask.shape
    rmin = (C.c_int32 * 2)(*(roi_min if roi_min is not None else (0, 0)))
    rmax = (C.c_int32 * 2)(*(roi_max if roi_max is not None else (w - 1, h - 1)))
    out = np.empty((h, w), np.float32)
    capi.check(L, None, L.vcb_distance_transform_l1(
        device, mask.ctypes.data_as(C.POINTER(C.c_uint8)), w, h, rmin, rmax,
        out.ctypes.data_as(C.POINTER(C.c_float))))
    return out


def make_signed_distance_field(mask, roi_min=None, roi_max=None, minmax_normalize=True,
                               use_truncation=False, truncation_band=0.1, device=0,
                               metric=SdfMetric.kL1):
    """MakeSignedDistanceField (voxel_carver.h:122-126) on the GPU; metric=kL2 selects the exact
    Euclidean variant (no reference counterpart)."""
    L = capi.load_library()
    mask = np.ascontiguousarray(mask, np.uint8)
    h, w = mask.shape
    rmin = (C.c_int32 * 2)(*(roi_min if roi_min is not None else (0, 0)))
    rmax = (C.c_int32 * 2)(*(roi_max if roi_max is not None else (w - 1, h - 1)))
    out = np.empty((h, w), np.float32)
    fn = (L.vcb_make_signed_distance_field_l2 if int(metric) == int(SdfMetric.kL2)
          else L.vcb_make_signed_distance_field)
    capi.check(L, None, fn(
        device, mask.ctypes.data_as(C.POINTER(C.c_uint8)), w, h, rmin, rmax,
        int(minmax_normalize), int(use_truncation), C.c_float(truncation_band),
        out.ctypes.data_as(C.POINTER(C.c_float))))
    return out


def load_mask_file(path):
    """One silhouette file -> uint8 [H, W] (vcb_load_mask_file; the decoder of carve_files)."""
    L = capi.load_library()
    w, h = C.c_int32(), C.c_int32()
    capi.check(L, None, L.vcb_load_mask_file(str(path).encode(), None, 0, C.byref(w), C.byref(h)))
    out = np.empty((h.value, w.value), np.uint8)
    capi.check(L, None, L.vcb_load_mask_file(str(path).encode(), C.c_void_p(out.ctypes.data),
                                             out.size, C.byref(w), C.byref(h)))
    return out


def ply_format(vertices, faces, device=0):
    """Mesh::WritePly (mesh.cc:583-631) for any mesh without vertex colours: the bytes of the ASCII
    PLY file, formatted on the GPU."""
    L = capi.load_library()
    v = np.ascontiguousarray(vertices, np.float32).reshape(-1, 3)
    f = np.ascontiguousarray(faces, np.int32).reshape(-1, 3)
    n = C.c_size_t()
    vp_, fp_ = v.ctypes.data_as(C.POINTER(C.c_float)), f.ctypes.data_as(C.POINTER(C.c_int32))
    rc = L.vcb_ply_format(device, vp_, len(v), fp_, len(f), None, 0, C.byref(n))
    if rc != 0:
        raise capi.VcbError(rc, L.vcb_last_error(None).decode())
    out = np.empty(n.value, np.uint8)
    rc = L.vcb_ply_format(device, vp_, len(v), fp_, len(f), C.c_void_p(out.ctypes.data), out.size,
                          C.byref(n))
    if rc != 0:
        raise capi.VcbError(rc, L.vcb_last_error(None).decode())
    return out.tobytes()


def signed_distance_to_color(sdf, min_negative_d, max_positive_d, device=0):
    """SignedDistance2Color (voxel_carver.h:127-128) on the GPU: float32 [H, W] -> uint8 [H, W, 3]."""
    L = capi.load_library()
    sdf = np.ascontiguousarray(sdf, np.float32)
    h, w = sdf.shape
    out = np.empty((h, w, 3), np.uint8)
    capi.check(L, None, L.vcb_signed_distance_to_color(
        device, sdf.ctypes.data_as(C.POINTER(C.c_float)), w, h, C.c_float(min_negative_d),
        C.c_float(max_positive_d), out.ctypes.data_as(C.POINTER(C.c_uint8))))
    return out
