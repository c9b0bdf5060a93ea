"""ctypes binding of the C ABI in include/vacancy_b200.h (one-to-one, no logic)."""
import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("VACANCY_B200_LIB", os.path.join(HERE, "libvacancy_b200.so"))


ABI_VERSION = 2  # VCB_ABI_VERSION of include/vacancy_b200.h


class LibraryMissing(RuntimeError):
    pass


class VcbError(RuntimeError):
    def __init__(self, code, text):
        super().__init__("vacancy_b200 error %d: %s" % (code, text))
        self.code = code


class Options(C.Structure):
    """vcb_options (include/vacancy_b200.h) = VoxelCarverOption + VoxelUpdateOption + placement."""
    _fields_ = [("bb_max", C.c_float * 3), ("bb_min", C.c_float * 3), ("resolution", C.c_float),
                ("sdf_minmax_normalize", C.c_int32), ("voxel_update", C.c_int32),
                ("sdf_interp", C.c_int32), ("update_outside", C.c_int32),
                ("voxel_max_update_num", C.c_int32), ("voxel_update_weight", C.c_float),
                ("use_truncation", C.c_int32), ("truncation_band", C.c_float),
                ("device", C.c_int32), ("z_begin", C.c_int32), ("z_end", C.c_int32),
                ("sdf_metric", C.c_int32)]


class View(C.Structure):
    """vcb_view."""
    _fields_ = [("w2c", C.c_float * 12), ("model", C.c_int32), ("fx", C.c_float),
                ("fy", C.c_float), ("cx", C.c_float), ("cy", C.c_float), ("width", C.c_int32),
                ("height", C.c_int32), ("roi_min", C.c_int32 * 2), ("roi_max", C.c_int32 * 2)]


EXPORTS = [
    "vcb_last_error", "vcb_abi_version", "vcb_invalid_sdf", "vcb_options_default", "vcb_create", "vcb_reset", "vcb_destroy",
    "vcb_grid_dims", "vcb_slab_range", "vcb_grid_positions", "vcb_set_stream", "vcb_synchronize",
    "vcb_flush", "vcb_carve_masks", "vcb_carve_masks_device", "vcb_carve_sdfs",
    "vcb_carve_sdfs_device", "vcb_distance_transform_l1", "vcb_make_signed_distance_field",
    "vcb_make_signed_distance_field_noalloc", "vcb_signed_distance_to_color",
    "vcb_make_signed_distance_field_l2",
    "vcb_extract_iso_count", "vcb_extract_iso_fill", "vcb_extract_iso_device",
    "vcb_extract_iso_begin", "vcb_extract_iso_vertices", "vcb_extract_iso_export_top",
    "vcb_extract_iso_import_bottom", "vcb_extract_iso_faces", "vcb_halo_device",
    "vcb_extract_voxel_count", "vcb_extract_voxel_fill", "vcb_extract_voxel_occupancy",
    "vcb_download_grid", "vcb_upload_grid", "vcb_grid_device", "vcb_stage_times",
    "vcb_fusion_class_counts", "vcb_set_statistics", "vcb_save_grid", "vcb_load_grid", "vcb_sharded_save_grid",
    "vcb_sharded_load_grid", "vcb_carve_mask_files", "vcb_load_mask_file",
    "vcb_sharded_carve_mask_files",
    "vcb_ply_header", "vcb_ply_body_size", "vcb_ply_body_fill", "vcb_ply_format",
    "vcb_sharded_create", "vcb_sharded_unique_id", "vcb_sharded_create_rank", "vcb_sharded_destroy",
    "vcb_sharded_last_error", "vcb_sharded_info", "vcb_sharded_bounds", "vcb_sharded_repartition",
    "vcb_sharded_local", "vcb_sharded_stream", "vcb_sharded_set_auto_balance", "vcb_balance_bounds", "vcb_sharded_reset", "vcb_sharded_carve_masks",
    "vcb_sharded_carve_masks_device", "vcb_sharded_carve_sdfs", "vcb_sharded_flush", "vcb_sharded_synchronize",
    "vcb_sharded_extract_iso_count", "vcb_sharded_extract_iso_fill", "vcb_sharded_local_mesh",
    "vcb_sharded_extract_voxel_count", "vcb_sharded_extract_voxel_fill", "vcb_sharded_stage_times",
]

UNIQUE_ID_BYTES = 128  # VCB_UNIQUE_ID_BYTES

_lib = None


def load_library(path=LIB_PATH):
    """Loads libvacancy_b200.so; raises LibraryMissing (never falls back) when it is not built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(path):
        raise LibraryMissing(
            "%s not found: build it with `python -m vacancy_b200.build` "
            "(there is no CPU fallback)" % path)
    L = C.CDLL(path)
    vp, i32, i64 = C.c_void_p, C.c_int32, C.c_int64
    fp, ip, u8p = C.POINTER(C.c_float), C.POINTER(C.c_int32), C.POINTER(C.c_uint8)
    L.vcb_last_error.restype = C.c_char_p
    L.vcb_last_error.argtypes = [vp]
    L.vcb_invalid_sdf.restype = C.c_float
    if L.vcb_abi_version() != ABI_VERSION:
        raise LibraryMissing("%s has ABI version %d, this package expects %d: rebuild it"
                             % (path, L.vcb_abi_version(), ABI_VERSION))
    L.vcb_options_default.argtypes = [C.POINTER(Options)]
    L.vcb_options_default.restype = None
    L.vcb_create.argtypes = [C.POINTER(Options), C.POINTER(vp)]
    L.vcb_reset.argtypes = [vp]
    L.vcb_destroy.argtypes = [vp]
    L.vcb_destroy.restype = None
    L.vcb_grid_dims.argtypes = [vp, ip]
    L.vcb_slab_range.argtypes = [vp, ip, ip]
    L.vcb_grid_positions.argtypes = [vp, fp, fp, fp]
    L.vcb_set_stream.argtypes = [vp, vp]
    L.vcb_synchronize.argtypes = [vp]
    L.vcb_flush.argtypes = [vp]
    pp = C.POINTER(vp)
    L.vcb_carve_masks.argtypes = [vp, i32, C.POINTER(View), pp, pp]
    L.vcb_carve_masks_device.argtypes = [vp, i32, C.POINTER(View), pp, pp]
    L.vcb_carve_sdfs.argtypes = [vp, i32, C.POINTER(View), pp]
    L.vcb_carve_sdfs_device.argtypes = [vp, i32, C.POINTER(View), pp]
    L.vcb_distance_transform_l1.argtypes = [i32, u8p, i32, i32, ip, ip, fp]
    L.vcb_make_signed_distance_field.argtypes = [i32, u8p, i32, i32, ip, ip, i32, i32, C.c_float,
                                                 fp]
    L.vcb_make_signed_distance_field_noalloc.argtypes = [i32, vp, i32, i32, ip, ip, i32, i32,
                                                         C.c_float, vp, vp]
    L.vcb_make_signed_distance_field_l2.argtypes = [i32, u8p, i32, i32, ip, ip, i32, i32, C.c_float,
                                                    fp]
    L.vcb_signed_distance_to_color.argtypes = [i32, fp, i32, i32, C.c_float, C.c_float, u8p]
    L.vcb_extract_iso_count.argtypes = [vp, C.c_double, i32, C.POINTER(i64), C.POINTER(i64)]
    L.vcb_extract_iso_begin.argtypes = [vp, C.c_double, i32, C.POINTER(i64), C.POINTER(i64)]
    L.vcb_extract_iso_vertices.argtypes = [vp, i64]
    L.vcb_extract_iso_export_top.argtypes = [vp, vp]
    L.vcb_extract_iso_import_bottom.argtypes = [vp, vp]
    L.vcb_extract_iso_faces.argtypes = [vp]
    L.vcb_halo_device.argtypes = [vp, C.POINTER(vp), ip]
    L.vcb_extract_iso_fill.argtypes = [vp, fp, ip]
    L.vcb_extract_iso_device.argtypes = [vp, C.POINTER(vp), C.POINTER(vp)]
    L.vcb_extract_voxel_count.argtypes = [vp, i32, C.POINTER(i64)]
    L.vcb_extract_voxel_fill.argtypes = [vp, fp, ip]
    L.vcb_extract_voxel_occupancy.argtypes = [vp, u8p]
    L.vcb_download_grid.argtypes = [vp, fp, ip]
    L.vcb_upload_grid.argtypes = [vp, fp, ip]
    L.vcb_grid_device.argtypes = [vp, C.POINTER(vp), C.POINTER(vp)]
    L.vcb_stage_times.argtypes = [vp, fp, C.POINTER(i64)]
    L.vcb_set_statistics.argtypes = [vp, i32]
    L.vcb_fusion_class_counts.argtypes = [vp, C.POINTER(C.c_uint64), C.POINTER(i64),
                                          C.POINTER(C.c_uint64)]
    L.vcb_carve_mask_files.argtypes = [vp, i32, C.POINTER(View), C.POINTER(C.c_char_p), i32]
    L.vcb_sharded_carve_mask_files.argtypes = [vp, i32, C.POINTER(View), C.POINTER(C.c_char_p), i32]
    L.vcb_load_mask_file.argtypes = [C.c_char_p, vp, C.c_size_t, ip, ip]
    L.vcb_save_grid.argtypes = [vp, C.c_char_p]
    L.vcb_load_grid.argtypes = [vp, C.c_char_p, ip]
    L.vcb_sharded_save_grid.argtypes = [vp, C.c_char_p]
    L.vcb_sharded_load_grid.argtypes = [vp, C.POINTER(C.c_char_p), i32]
    szp = C.POINTER(C.c_size_t)
    L.vcb_ply_header.argtypes = [i64, i64, C.c_char_p, C.c_size_t, szp]
    L.vcb_ply_body_size.argtypes = [vp, szp, szp]
    L.vcb_ply_body_fill.argtypes = [vp, vp, vp]
    L.vcb_ply_format.argtypes = [i32, fp, i64, ip, i64, vp, C.c_size_t, szp]
    L.vcb_sharded_create.argtypes = [C.POINTER(Options), ip, i32, ip, C.POINTER(vp)]
    L.vcb_sharded_unique_id.argtypes = [vp]
    L.vcb_sharded_create_rank.argtypes = [C.POINTER(Options), i32, i32, i32, vp, ip, C.POINTER(vp)]
    L.vcb_sharded_destroy.argtypes = [vp]
    L.vcb_sharded_destroy.restype = None
    L.vcb_sharded_last_error.argtypes = [vp]
    L.vcb_sharded_last_error.restype = C.c_char_p
    L.vcb_sharded_info.argtypes = [vp, ip, ip, ip]
    L.vcb_sharded_bounds.argtypes = [vp, ip]
    L.vcb_sharded_repartition.argtypes = [vp, ip]
    L.vcb_sharded_local.argtypes = [vp, i32, C.POINTER(vp), ip]
    L.vcb_sharded_stream.argtypes = [vp, i32, C.POINTER(vp)]
    L.vcb_sharded_set_auto_balance.argtypes = [vp, i32]
    L.vcb_balance_bounds.argtypes = [C.POINTER(C.c_uint64), i32, i32, i32, i32, ip]
    L.vcb_sharded_reset.argtypes = [vp]
    L.vcb_sharded_carve_masks.argtypes = [vp, i32, C.POINTER(View), pp, pp]
    L.vcb_sharded_carve_masks_device.argtypes = [vp, i32, C.POINTER(View), pp, pp]
    L.vcb_sharded_carve_sdfs.argtypes = [vp, i32, C.POINTER(View), pp]
    L.vcb_sharded_flush.argtypes = [vp]
    L.vcb_sharded_synchronize.argtypes = [vp]
    L.vcb_sharded_extract_iso_count.argtypes = [vp, C.c_double, i32, C.POINTER(i64), C.POINTER(i64)]
    L.vcb_sharded_extract_iso_fill.argtypes = [vp, fp, ip]
    L.vcb_sharded_local_mesh.argtypes = [vp, i32] + [C.POINTER(i64)] * 4
    L.vcb_sharded_extract_voxel_count.argtypes = [vp, i32, C.POINTER(i64)]
    L.vcb_sharded_extract_voxel_fill.argtypes = [vp, fp, ip]
    L.vcb_sharded_stage_times.argtypes = [vp, i32, fp, C.POINTER(i64)]
    _lib = L
    return L


def check(L, handle, rc):
    if rc != 0:
        text = L.vcb_last_error(handle)
        raise VcbError(rc, text.decode() if text else "")


def check_sharded(L, handle, rc):
    if rc != 0:
        text = L.vcb_sharded_last_error(handle)
        raise VcbError(rc, text.decode() if text else "")
