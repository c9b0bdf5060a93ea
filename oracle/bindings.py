"""TEST INFRASTRUCTURE ONLY - ctypes bindings for the two parity checkers.

* ``RefLib``  - oracle/_ref/libvacancy_ref.so: the UNMODIFIED reference (unclearness/vacancy) compiled
  from /root/reference by oracle/Makefile, behind the C harness oracle/ref_harness.cc.
* ``PortLib`` - oracle/_build/libvacancy_oracle.so: the plain-C restatement oracle/vacancy_oracle.c.

Both expose the same small ``Carver``-like interface over numpy arrays so a test can run the same
scenario through either and through the CUDA library.  Nothing in vacancy_b200/ imports this module.
"""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REF_SO = os.path.join(HERE, "_ref", "libvacancy_ref.so")
PORT_SO = os.path.join(HERE, "_build", "libvacancy_oracle.so")

c_float_p = C.POINTER(C.c_float)
c_int_p = C.POINTER(C.c_int)
c_u8_p = C.POINTER(C.c_uint8)
c_double_p = C.POINTER(C.c_double)


def build(ref=True, port=True):
    """Compile the checkers (no-op for the reference when /root/reference is absent)."""
    targets = []
    if port:
        targets.append("port")
    if ref:
        targets.append("ref")
    subprocess.run(["make", "-C", HERE, "--no-print-directory"] + targets, check=True,
                   stdout=subprocess.DEVNULL)


def _p(a, t):
    return a.ctypes.data_as(t)


class Options(C.Structure):
    """vo_options / VoxelCarverOption + VoxelUpdateOption (voxel_carver.h:43-60)."""
    _fields_ = [("bb_max", C.c_float * 3), ("bb_min", C.c_float * 3), ("resolution", C.c_float),
                ("sdf_minmax_normalize", C.c_int32), ("voxel_update", C.c_int32),
                ("sdf_interp", C.c_int32), ("update_outside", C.c_int32),
                ("voxel_max_update_num", C.c_int32), ("voxel_update_weight", C.c_float),
                ("use_truncation", C.c_int32), ("truncation_band", C.c_float)]


class View(C.Structure):
    _fields_ = [("w2c", C.c_float * 12), ("model", C.c_int32), ("fx", C.c_float),
                ("fy", C.c_float), ("cx", C.c_float), ("cy", C.c_float), ("width", C.c_int32),
                ("height", C.c_int32), ("roi_min", C.c_int32 * 2), ("roi_max", C.c_int32 * 2)]


def make_options(bb_min, bb_max, resolution, normalize=True, update=0, interp=1, outside=0,
                 max_update_num=255, weight=1.0, truncation=False, band=0.1):
    o = Options()
    o.bb_max[:] = [float(v) for v in bb_max]
    o.bb_min[:] = [float(v) for v in bb_min]
    o.resolution = resolution
    o.sdf_minmax_normalize = int(normalize)
    o.voxel_update = int(update)
    o.sdf_interp = int(interp)
    o.update_outside = int(outside)
    o.voxel_max_update_num = int(max_update_num)
    o.voxel_update_weight = weight
    o.use_truncation = int(truncation)
    o.truncation_band = band
    return o


class Camera:
    """Host-side camera description: c2w is a 4x4 float64 (camera.h:22-23)."""

    def __init__(self, width, height, c2w, fx=1.0, fy=1.0, cx=0.0, cy=0.0, model=0):
        self.width, self.height = int(width), int(height)
        self.c2w = np.ascontiguousarray(c2w, dtype=np.float64).reshape(4, 4)
        self.fx, self.fy, self.cx, self.cy = (float(np.float32(v)) for v in (fx, fy, cx, cy))
        self.model = int(model)

    def full_roi(self, w=None, h=None):
        w = self.width if w is None else w
        h = self.height if h is None else h
        return (0, 0, w - 1, h - 1)


# ------------------------------------------------------------------------------------------------
class RefLib:
    def __init__(self, path=REF_SO):
        if not os.path.exists(path):
            raise FileNotFoundError(path)
        L = self.L = C.CDLL(path)
        L.ref_invalid_sdf.restype = C.c_float
        L.ref_carver_create.restype = C.c_void_p
        L.ref_carver_create.argtypes = [c_float_p, c_float_p, C.c_float, C.c_int, C.c_int, C.c_int,
                                        C.c_int, C.c_int, C.c_float, C.c_int, C.c_float]
        L.ref_carver_destroy.argtypes = [C.c_void_p]
        L.ref_carver_dims.argtypes = [C.c_void_p, c_int_p]
        L.ref_carver_carve_mask.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, c_double_p,
                                            C.c_float, C.c_float, C.c_float, C.c_float, c_u8_p,
                                            C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
                                            c_float_p]
        L.ref_carver_carve_sdf.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, c_double_p,
                                           C.c_float, C.c_float, C.c_float, C.c_float, c_float_p,
                                           C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
                                           c_double_p]
        L.ref_carver_download.argtypes = [C.c_void_p, c_float_p, c_int_p]
        L.ref_carver_upload.argtypes = [C.c_void_p, c_float_p, c_int_p]
        L.ref_carver_positions.argtypes = [C.c_void_p, c_float_p, c_float_p, c_float_p]
        L.ref_carver_extract_iso.argtypes = [C.c_void_p, C.c_double, C.c_int,
                                             C.POINTER(C.c_longlong), C.POINTER(C.c_longlong),
                                             c_double_p]
        L.ref_carver_extract_voxel.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_longlong),
                                               C.POINTER(C.c_longlong), c_double_p]
        L.ref_carver_mesh_copy.argtypes = [C.c_void_p, c_float_p, c_int_p]
        L.ref_carver_mesh_write_ply.argtypes = [C.c_void_p, C.c_char_p]
        L.ref_write_ply.argtypes = [C.c_char_p, c_float_p, C.c_longlong, c_int_p, C.c_longlong]
        L.ref_distance_transform_l1.argtypes = [c_u8_p] + [C.c_int] * 6 + [c_float_p]
        L.ref_make_sdf.argtypes = [c_u8_p] + [C.c_int] * 8 + [C.c_float, c_float_p]
        L.ref_time_make_sdf.restype = C.c_double
        L.ref_time_make_sdf.argtypes = [c_u8_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_float,
                                        C.c_int]
        L.ref_signed_distance_to_color.argtypes = [c_float_p, C.c_int, C.c_int, C.c_float, C.c_float,
                                                   c_u8_p]
        L.ref_w2c_from_c2w.argtypes = [c_double_p, c_float_p]
        L.ref_tum_pose_to_c2w.argtypes = [c_double_p, c_double_p, c_double_p]
        L.ref_pinhole_ray_c.argtypes = [C.c_int, C.c_int] + [C.c_float] * 6 + [c_float_p]
        L.ref_load_png_gray.argtypes = [C.c_char_p, c_u8_p, c_int_p, c_int_p]
        L.ref_mc_tables.argtypes = [c_int_p, c_int_p]
        L.ref_set_quiet(1)
        self.kind = "reference"

    # free functions
    def invalid_sdf(self):
        return np.float32(self.L.ref_invalid_sdf())

    def threads(self):
        return int(self.L.ref_omp_threads())

    def set_threads(self, n=-1):
        """OpenMP threads of the reference's parallel loops; n < 0 = every core of the machine
        (torch.distributed.run exports OMP_NUM_THREADS=1 to its workers)."""
        return int(self.L.ref_set_omp_threads(int(n)))

    def distance_transform_l1(self, mask, roi=None):
        mask = np.ascontiguousarray(mask, dtype=np.uint8)
        h, w = mask.shape
        roi = roi or (0, 0, w - 1, h - 1)
        out = np.empty((h, w), np.float32)
        self.L.ref_distance_transform_l1(_p(mask, c_u8_p), w, h, *roi, _p(out, c_float_p))
        return out

    def make_sdf(self, mask, roi=None, normalize=True, truncate=False, band=0.1):
        mask = np.ascontiguousarray(mask, dtype=np.uint8)
        h, w = mask.shape
        roi = roi or (0, 0, w - 1, h - 1)
        out = np.empty((h, w), np.float32)
        self.L.ref_make_sdf(_p(mask, c_u8_p), w, h, *roi, int(normalize), int(truncate),
                            C.c_float(band), _p(out, c_float_p))
        return out

    def signed_distance_to_color(self, sdf, min_negative_d, max_positive_d):
        sdf = np.ascontiguousarray(sdf, np.float32)
        h, w = sdf.shape
        out = np.empty((h, w, 3), np.uint8)
        self.L.ref_signed_distance_to_color(_p(sdf, c_float_p), w, h, C.c_float(min_negative_d),
                                            C.c_float(max_positive_d), _p(out, c_u8_p))
        return out

    def time_make_sdf(self, mask, normalize, truncate, band, reps):
        mask = np.ascontiguousarray(mask, dtype=np.uint8)
        h, w = mask.shape
        return float(self.L.ref_time_make_sdf(_p(mask, c_u8_p), w, h, int(normalize),
                                              int(truncate), C.c_float(band), reps))

    def w2c_from_c2w(self, c2w):
        c2w = np.ascontiguousarray(c2w, dtype=np.float64).reshape(16)
        out = np.empty(12, np.float32)
        self.L.ref_w2c_from_c2w(_p(c2w, c_double_p), _p(out, c_float_p))
        return out

    def tum_pose_to_c2w(self, t, q):
        t = np.ascontiguousarray(t, dtype=np.float64)
        q = np.ascontiguousarray(q, dtype=np.float64)
        out = np.empty(16, np.float64)
        self.L.ref_tum_pose_to_c2w(_p(t, c_double_p), _p(q, c_double_p), _p(out, c_double_p))
        return out.reshape(4, 4)

    def pinhole_ray_c(self, cam, x, y):
        out = np.empty(3, np.float32)
        self.L.ref_pinhole_ray_c(cam.width, cam.height, cam.fx, cam.fy, cam.cx, cam.cy,
                                 C.c_float(x), C.c_float(y), _p(out, c_float_p))
        return out

    def load_png_gray(self, path):
        w, h = C.c_int(), C.c_int()
        if self.L.ref_load_png_gray(path.encode(), None, C.byref(w), C.byref(h)) != 0:
            raise IOError(path)
        out = np.empty((h.value, w.value), np.uint8)
        self.L.ref_load_png_gray(path.encode(), _p(out, c_u8_p), C.byref(w), C.byref(h))
        return out

    def write_ply(self, path, vertices, faces):
        """Mesh::WritePly of the UNMODIFIED reference on an arbitrary mesh."""
        v = np.ascontiguousarray(vertices, np.float32).reshape(-1, 3)
        f = np.ascontiguousarray(faces, np.int32).reshape(-1, 3)
        return self.L.ref_write_ply(path.encode(), _p(v, c_float_p), len(v), _p(f, c_int_p), len(f))

    def mc_tables(self):
        e = np.empty(256, np.int32)
        t = np.empty((256, 16), np.int32)
        self.L.ref_mc_tables(_p(e, c_int_p), _p(t, c_int_p))
        return e, t

    def carver(self, opt):
        return _RefCarver(self, opt)


class _RefCarver:
    def __init__(self, lib, opt):
        self.lib, self.L, self.opt = lib, lib.L, opt
        bmin = np.array(list(opt.bb_min), np.float32)
        bmax = np.array(list(opt.bb_max), np.float32)
        self.h = self.L.ref_carver_create(_p(bmin, c_float_p), _p(bmax, c_float_p),
                                          C.c_float(opt.resolution), opt.sdf_minmax_normalize,
                                          opt.voxel_update, opt.sdf_interp, opt.update_outside,
                                          opt.voxel_max_update_num,
                                          C.c_float(opt.voxel_update_weight), opt.use_truncation,
                                          C.c_float(opt.truncation_band))
        if not self.h:
            raise ValueError("reference Init() returned false")
        d = np.empty(3, np.int32)
        self.L.ref_carver_dims(self.h, _p(d, c_int_p))
        self.dims = tuple(int(v) for v in d)
        self.count = int(np.prod(np.array(self.dims, np.int64)))

    def close(self):
        if self.h:
            self.L.ref_carver_destroy(self.h)
            self.h = None

    def __del__(self):
        self.close()

    def _cam(self, cam):
        c2w = np.ascontiguousarray(cam.c2w, np.float64).reshape(16)
        return (cam.model, cam.width, cam.height, _p(c2w, c_double_p), C.c_float(cam.fx),
                C.c_float(cam.fy), C.c_float(cam.cx), C.c_float(cam.cy)), c2w

    def carve_mask(self, cam, mask, roi=None):
        mask = np.ascontiguousarray(mask, np.uint8)
        h, w = mask.shape
        roi = roi or (0, 0, w - 1, h - 1)
        out = np.empty((h, w), np.float32)
        args, keep = self._cam(cam)
        self.L.ref_carver_carve_mask(self.h, *args, _p(mask, c_u8_p), w, h, *roi,
                                     _p(out, c_float_p))
        return out

    def carve_sdf(self, cam, sdf, roi=None):
        sdf = np.ascontiguousarray(sdf, np.float32)
        h, w = sdf.shape
        roi = roi or (0, 0, w - 1, h - 1)
        sec = C.c_double()
        args, keep = self._cam(cam)
        self.L.ref_carver_carve_sdf(self.h, *args, _p(sdf, c_float_p), w, h, *roi, C.byref(sec))
        return sec.value

    def download(self):
        sdf = np.empty(self.count, np.float32)
        un = np.empty(self.count, np.int32)
        self.L.ref_carver_download(self.h, _p(sdf, c_float_p), _p(un, c_int_p))
        return sdf, un

    def upload(self, sdf, un):
        sdf = np.ascontiguousarray(sdf, np.float32).reshape(-1)
        un = np.ascontiguousarray(un, np.int32).reshape(-1)
        assert sdf.size == self.count and un.size == self.count
        self.L.ref_carver_upload(self.h, _p(sdf, c_float_p), _p(un, c_int_p))

    def positions(self):
        px, py, pz = (np.empty(n, np.float32) for n in self.dims)
        self.L.ref_carver_positions(self.h, _p(px, c_float_p), _p(py, c_float_p),
                                    _p(pz, c_float_p))
        return px, py, pz

    def _mesh(self, nv, nf):
        v = np.empty((nv, 3), np.float32)
        f = np.empty((nf, 3), np.int32)
        self.L.ref_carver_mesh_copy(self.h, _p(v, c_float_p), _p(f, c_int_p))
        return v, f

    def extract_iso(self, iso=0.0, interp=True, timing=None):
        nv, nf, sec = C.c_longlong(), C.c_longlong(), C.c_double()
        self.L.ref_carver_extract_iso(self.h, iso, int(interp), C.byref(nv), C.byref(nf),
                                      C.byref(sec))
        if timing is not None:
            timing.append(sec.value)
        return self._mesh(nv.value, nf.value)

    def extract_voxel(self, inside_empty=False, timing=None, copy=True):
        nv, nf, sec = C.c_longlong(), C.c_longlong(), C.c_double()
        self.L.ref_carver_extract_voxel(self.h, int(inside_empty), C.byref(nv), C.byref(nf),
                                        C.byref(sec))
        if timing is not None:
            timing.append(sec.value)
        if not copy:
            return nv.value, nf.value
        return self._mesh(nv.value, nf.value)

    def write_ply(self, path):
        return self.L.ref_carver_mesh_write_ply(self.h, path.encode())


# ------------------------------------------------------------------------------------------------
class PortLib:
    def __init__(self, path=PORT_SO):
        if not os.path.exists(path):
            build(ref=False, port=True)
        L = self.L = C.CDLL(path)
        OP, VP = C.POINTER(Options), C.POINTER(View)
        L.vo_invalid_sdf.restype = C.c_float
        L.vo_w2c_from_c2w.argtypes = [c_double_p, c_float_p]
        L.vo_distance_transform_l1.argtypes = [c_u8_p] + [C.c_int] * 6 + [c_float_p]
        L.vo_make_sdf.argtypes = [c_u8_p] + [C.c_int] * 8 + [C.c_float, c_float_p]
        L.vo_signed_distance_to_color.argtypes = [c_float_p, C.c_int, C.c_int, C.c_float, C.c_float,
                                                  c_u8_p]
        L.vo_make_sdf_l2.argtypes = [c_u8_p] + [C.c_int] * 8 + [C.c_float, c_float_p]
        L.vo_grid_dims.argtypes = [OP, c_int_p]
        L.vo_grid_positions.argtypes = [OP, c_float_p, c_float_p, c_float_p]
        L.vo_grid_reset.argtypes = [c_float_p, c_int_p, C.c_int64]
        L.vo_carve_sdf.argtypes = [OP, c_int_p, c_float_p, c_float_p, c_float_p, C.c_int, C.c_int,
                                   VP, c_float_p, c_float_p, c_int_p]
        L.vo_extract_voxel_flags.restype = C.c_int64
        L.vo_extract_voxel_flags.argtypes = [c_int_p, c_float_p, c_int_p, C.c_int, c_u8_p]
        L.vo_extract_voxel_mesh.argtypes = [c_int_p, c_float_p, c_float_p, c_float_p, C.c_float,
                                            c_u8_p, c_float_p, c_int_p]
        L.vo_marching_cubes.argtypes = [c_int_p, c_float_p, c_float_p, c_float_p, c_float_p,
                                        c_int_p, C.c_double, C.c_int, C.POINTER(c_float_p),
                                        C.POINTER(C.c_int64), C.POINTER(c_int_p),
                                        C.POINTER(C.c_int64)]
        L.vo_free.argtypes = [C.c_void_p]
        L.vo_mc_tables.argtypes = [c_int_p, c_int_p]
        self.kind = "port"

    def set_threads(self, n=-1):
        return int(self.L.vo_set_threads(int(n)))

    def threads(self):
        return int(self.L.vo_set_threads(0))

    def invalid_sdf(self):
        return np.float32(self.L.vo_invalid_sdf())

    def distance_transform_l1(self, mask, roi=None):
        mask = np.ascontiguousarray(mask, dtype=np.uint8)
        h, w = mask.shape
        roi = roi or (0, 0, w - 1, h - 1)
        out = np.empty((h, w), np.float32)
        self.L.vo_distance_transform_l1(_p(mask, c_u8_p), w, h, *roi, _p(out, c_float_p))
        return out

    def make_sdf(self, mask, roi=None, normalize=True, truncate=False, band=0.1):
        mask = np.ascontiguousarray(mask, dtype=np.uint8)
        h, w = mask.shape
        roi = roi or (0, 0, w - 1, h - 1)
        out = np.empty((h, w), np.float32)
        self.L.vo_make_sdf(_p(mask, c_u8_p), w, h, *roi, int(normalize), int(truncate),
                           C.c_float(band), _p(out, c_float_p))
        return out

    def signed_distance_to_color(self, sdf, min_negative_d, max_positive_d):
        sdf = np.ascontiguousarray(sdf, np.float32)
        h, w = sdf.shape
        out = np.empty((h, w, 3), np.uint8)
        self.L.vo_signed_distance_to_color(_p(sdf, c_float_p), w, h, C.c_float(min_negative_d),
                                           C.c_float(max_positive_d), _p(out, c_u8_p))
        return out

    def make_sdf_l2(self, mask, roi=None, normalize=True, truncate=False, band=0.1):
        """Euclidean variant (not in the reference)."""
        mask = np.ascontiguousarray(mask, dtype=np.uint8)
        h, w = mask.shape
        roi = roi or (0, 0, w - 1, h - 1)
        out = np.empty((h, w), np.float32)
        self.L.vo_make_sdf_l2(_p(mask, c_u8_p), w, h, *roi, int(normalize), int(truncate),
                              C.c_float(band), _p(out, c_float_p))
        return out

    def w2c_from_c2w(self, c2w):
        c2w = np.ascontiguousarray(c2w, dtype=np.float64).reshape(16)
        out = np.empty(12, np.float32)
        self.L.vo_w2c_from_c2w(_p(c2w, c_double_p), _p(out, c_float_p))
        return out

    def ply_text(self, vertices, faces):
        """Mesh::WritePly restated (vo_ply_text): the bytes of the ASCII PLY file."""
        v = np.ascontiguousarray(vertices, np.float32).reshape(-1, 3)
        f = np.ascontiguousarray(faces, np.int32).reshape(-1, 3)
        self.L.vo_ply_text.restype = C.c_int64
        self.L.vo_ply_text.argtypes = [C.c_void_p, C.c_int64, C.c_void_p, C.c_int64, C.c_void_p,
                                       C.c_int64]
        n = self.L.vo_ply_text(v.ctypes.data, len(v), f.ctypes.data, len(f), None, 0)
        out = np.empty(n, np.uint8)
        self.L.vo_ply_text(v.ctypes.data, len(v), f.ctypes.data, len(f), out.ctypes.data, n)
        return out.tobytes()

    def mc_tables(self):
        e = np.empty(256, np.int32)
        t = np.empty((256, 16), np.int32)
        self.L.vo_mc_tables(_p(e, c_int_p), _p(t, c_int_p))
        return e, t

    def carver(self, opt):
        return _PortCarver(self, opt)

    def slab_carver(self, opt, z_begin, z_end):
        """Planes [z_begin, z_end) of the grid `opt` describes, with GLOBAL positions and 64-bit
        linear indices: the checker for grids the reference cannot hold (its ids are int,
        voxel_carver.cc:298-301,307,333).  SURVEY.md section 8(c), last row."""
        return _PortSlabCarver(self, opt, z_begin, z_end)


def _shifted(arr, elements, ptr_type):
    """Pointer `elements` items BEFORE arr[0]: lets C code that indexes with global linear ids
    (size_t, z * nx * ny + ...) work on an array that holds only the planes from some z on."""
    return C.cast(C.c_void_p(arr.ctypes.data - elements * arr.itemsize), ptr_type)


class _PortCarver:
    def __init__(self, lib, opt):
        self.lib, self.L, self.opt = lib, lib.L, opt
        d = np.zeros(3, np.int32)
        if not self.L.vo_grid_dims(C.byref(opt), _p(d, c_int_p)):
            raise ValueError("oracle: Init() would return false")
        self._d = d
        self.dims = tuple(int(v) for v in d)
        self.count = int(np.prod(np.array(self.dims, np.int64)))
        self.px, self.py, self.pz = (np.zeros(n, np.float32) for n in self.dims)
        self.L.vo_grid_positions(C.byref(opt), _p(self.px, c_float_p), _p(self.py, c_float_p),
                                 _p(self.pz, c_float_p))
        self.sdf = np.empty(self.count, np.float32)
        self.un = np.empty(self.count, np.int32)
        self.L.vo_grid_reset(_p(self.sdf, c_float_p), _p(self.un, c_int_p), self.count)

    def close(self):
        pass

    def view(self, cam, roi):
        v = View()
        v.w2c[:] = [float(x) for x in self.lib.w2c_from_c2w(cam.c2w)]
        v.model = cam.model
        v.fx, v.fy, v.cx, v.cy = cam.fx, cam.fy, cam.cx, cam.cy
        v.width, v.height = cam.width, cam.height
        v.roi_min[:] = roi[:2]
        v.roi_max[:] = roi[2:]
        return v

    def carve_mask(self, cam, mask, roi=None):
        mask = np.ascontiguousarray(mask, np.uint8)
        h, w = mask.shape
        roi = roi or (0, 0, w - 1, h - 1)
        sdf = self.lib.make_sdf(mask, roi, bool(self.opt.sdf_minmax_normalize),
                                bool(self.opt.use_truncation), self.opt.truncation_band)
        self.carve_sdf(cam, sdf, roi)
        return sdf

    def carve_sdf(self, cam, sdf, roi=None, z_range=None):
        sdf = np.ascontiguousarray(sdf, np.float32)
        h, w = sdf.shape
        roi = roi or (0, 0, w - 1, h - 1)
        v = self.view(cam, roi)
        v.width, v.height = w, h
        z0, z1 = z_range or (0, self.dims[2])
        self.L.vo_carve_sdf(C.byref(self.opt), _p(self._d, c_int_p), _p(self.px, c_float_p),
                            _p(self.py, c_float_p), _p(self.pz, c_float_p), z0, z1, C.byref(v),
                            _p(sdf, c_float_p), _p(self.sdf, c_float_p), _p(self.un, c_int_p))

    def download(self):
        return self.sdf.copy(), self.un.copy()

    def upload(self, sdf, un):
        self.sdf[:] = np.asarray(sdf, np.float32).reshape(-1)
        self.un[:] = np.asarray(un, np.int32).reshape(-1)

    def positions(self):
        return self.px.copy(), self.py.copy(), self.pz.copy()

    def extract_iso(self, iso=0.0, interp=True, timing=None):
        vp, fp = c_float_p(), c_int_p()
        nv, nf = C.c_int64(), C.c_int64()
        self.L.vo_marching_cubes(_p(self._d, c_int_p), _p(self.px, c_float_p),
                                 _p(self.py, c_float_p), _p(self.pz, c_float_p),
                                 _p(self.sdf, c_float_p), _p(self.un, c_int_p), iso, int(interp),
                                 C.byref(vp), C.byref(nv), C.byref(fp), C.byref(nf))
        v = np.ctypeslib.as_array(vp, shape=(max(nv.value, 1), 3))[:nv.value].copy()
        f = np.ctypeslib.as_array(fp, shape=(max(nf.value, 1), 3))[:nf.value].copy()
        self.L.vo_free(vp)
        self.L.vo_free(fp)
        return v, f

    def occupancy(self, inside_empty=False):
        flags = np.empty(self.count, np.uint8)
        n = self.L.vo_extract_voxel_flags(_p(self._d, c_int_p), _p(self.sdf, c_float_p),
                                          _p(self.un, c_int_p), int(inside_empty),
                                          _p(flags, c_u8_p))
        return flags, int(n)

    def extract_voxel(self, inside_empty=False, timing=None):
        flags, n = self.occupancy(inside_empty)
        v = np.empty((n * 24, 3), np.float32)
        f = np.empty((n * 12, 3), np.int32)
        self.L.vo_extract_voxel_mesh(_p(self._d, c_int_p), _p(self.px, c_float_p),
                                     _p(self.py, c_float_p), _p(self.pz, c_float_p),
                                     C.c_float(self.opt.resolution), _p(flags, c_u8_p),
                                     _p(v, c_float_p), _p(f, c_int_p))
        return v, f


class _PortSlabCarver(_PortCarver):
    """_PortCarver restricted to the planes [z_begin, z_end): same C functions, same global
    positions; the state arrays hold only the slab and are addressed through shifted pointers."""

    def __init__(self, lib, opt, z_begin, z_end):
        self.lib, self.L, self.opt = lib, lib.L, opt
        d = np.zeros(3, np.int32)
        if not self.L.vo_grid_dims(C.byref(opt), _p(d, c_int_p)):
            raise ValueError("oracle: Init() would return false")
        self._d = d
        self.dims = tuple(int(v) for v in d)
        assert 0 <= z_begin < z_end <= self.dims[2]
        self.z_begin, self.z_end = int(z_begin), int(z_end)
        self.plane = self.dims[0] * self.dims[1]
        self.count = self.plane * (self.z_end - self.z_begin)
        self.px, self.py, self.pz = (np.zeros(n, np.float32) for n in self.dims)
        self.L.vo_grid_positions(C.byref(opt), _p(self.px, c_float_p), _p(self.py, c_float_p),
                                 _p(self.pz, c_float_p))
        self.sdf = np.empty(self.count, np.float32)
        self.un = np.empty(self.count, np.int32)
        self.L.vo_grid_reset(_p(self.sdf, c_float_p), _p(self.un, c_int_p), self.count)

    def carve_sdf(self, cam, sdf, roi=None, z_range=None):
        sdf = np.ascontiguousarray(sdf, np.float32)
        h, w = sdf.shape
        roi = roi or (0, 0, w - 1, h - 1)
        v = self.view(cam, roi)
        v.width, v.height = w, h
        shift = self.z_begin * self.plane  # first global linear id of the slab (may exceed 2^31)
        self.L.vo_carve_sdf(C.byref(self.opt), _p(self._d, c_int_p), _p(self.px, c_float_p),
                            _p(self.py, c_float_p), _p(self.pz, c_float_p), self.z_begin,
                            self.z_end, C.byref(v), _p(sdf, c_float_p),
                            _shifted(self.sdf, shift, c_float_p), _shifted(self.un, shift, c_int_p))

    def _sub(self):
        """(dims, pz) of the slab seen as a grid of its own (positions stay the global ones)."""
        d = np.array([self.dims[0], self.dims[1], self.z_end - self.z_begin], np.int32)
        return d, np.ascontiguousarray(self.pz[self.z_begin:self.z_end])

    def extract_iso(self, iso=0.0, interp=True, timing=None):
        """Marching cubes over the slab alone (no planes below it): what a z-slab carver whose halo
        holds kVal extracts, numbered from 0."""
        d, pz = self._sub()
        vp, fp = c_float_p(), c_int_p()
        nv, nf = C.c_int64(), C.c_int64()
        self.L.vo_marching_cubes(_p(d, c_int_p), _p(self.px, c_float_p), _p(self.py, c_float_p),
                                 _p(pz, c_float_p), _p(self.sdf, c_float_p), _p(self.un, c_int_p),
                                 iso, int(interp), C.byref(vp), C.byref(nv), C.byref(fp),
                                 C.byref(nf))
        v = np.ctypeslib.as_array(vp, shape=(max(nv.value, 1), 3))[:nv.value].copy()
        f = np.ctypeslib.as_array(fp, shape=(max(nf.value, 1), 3))[:nf.value].copy()
        self.L.vo_free(vp)
        self.L.vo_free(fp)
        return v, f

    def occupancy(self, inside_empty=False):
        d, _ = self._sub()
        flags = np.empty(self.count, np.uint8)
        n = self.L.vo_extract_voxel_flags(_p(d, c_int_p), _p(self.sdf, c_float_p),
                                          _p(self.un, c_int_p), int(inside_empty),
                                          _p(flags, c_u8_p))
        return flags, int(n)
