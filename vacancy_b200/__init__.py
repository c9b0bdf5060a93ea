"""vacancy_b200 - B200-native voxel carving behind the Vacancy `VoxelCarver` interface.

The product is libvacancy_b200.so (hand-written sm_100a CUDA behind the C ABI declared in
include/vacancy_b200.h).  This package holds its sources (csrc/), the C++ shim that implements
vacancy::VoxelCarver on top of it (shim/), and a thin ctypes mirror of the same interface for
Python hosts (tests, bench.py).  There is no CPU fallback: importing works anywhere, but creating
a carver without the built library or without a CUDA device raises.
"""
from .capi import LibraryMissing, VcbError, load_library  # noqa: F401
from .carver import (Camera, OrthoCamera, PinholeCamera, SdfInterpolation, SdfMetric,  # noqa: F401
                     UpdateOutsideImage, VoxelCarver, VoxelCarverOption, VoxelUpdate,
                     VoxelUpdateOption, distance_transform_l1, make_signed_distance_field,
                     load_mask_file, signed_distance_to_color, w2c_from_c2w)
