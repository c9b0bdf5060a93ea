/*
 * Batch entry point for C++ hosts that use the reference's classes.
 *
 * The reference's only batch overload, VoxelCarver::Carve(const std::vector<Camera>&,
 * const std::vector<Image1b>&) (voxel_carver.h:113-114), takes a vector of the ABSTRACT class Camera
 * and can therefore never be called with elements.  This free function is the usable form: pointers
 * to cameras and silhouettes, all views handed to the library in one call (one upload stream, one
 * fused kernel per 64 views).  Same result as calling Carve(camera, silhouette) view by view.
 * Implemented in vacancy_b200/shim/voxel_carver_shim.cc.
 */
#pragma once

#include <vector>

#include "vacancy/voxel_carver.h"

namespace vacancy_b200 {

bool CarveBatch(vacancy::VoxelCarver* carver, const std::vector<const vacancy::Camera*>& cameras,
                const std::vector<const vacancy::Image1b*>& silhouettes);

}  // namespace vacancy_b200
