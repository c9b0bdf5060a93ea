/*
 * Extensions for C++ hosts that use the reference's classes: a usable batch overload and the
 * GPU-side PLY writer.
 *
 * The reference's only batch overload, VoxelCarver::Carve(const std::vector<Camera>&,
 * const std::vector<Image1b>&) (voxel_carver.h:113-114), takes a vector of the ABSTRACT class Camera
 * and can therefore never be called with elements.  This free function is the usable form: pointers
 * to cameras and silhouettes, all views handed to the library in one call (one upload stream, one
 * fused kernel per 64 views).  Same result as calling Carve(camera, silhouette) view by view.
 * Implemented in vacancy_b200/shim/voxel_carver_shim.cc.
 */
#pragma once

#include <string>
#include <vector>

#include "vacancy/voxel_carver.h"

namespace vacancy_b200 {

bool CarveBatch(vacancy::VoxelCarver* carver, const std::vector<const vacancy::Camera*>& cameras,
                const std::vector<const vacancy::Image1b*>& silhouettes);

/* ExtractIsoSurface(&mesh, iso_level, linear_interp) followed by mesh.WritePly(path)
 * (examples.cc:143-144, mesh.cc:583-631) without the mesh ever visiting the host as vectors: the
 * marching-cubes result stays on the GPU, the ASCII text is formatted there (byte for byte what
 * WritePly writes) and lands in the file with one write.  Returns false on any error. */
bool ExtractIsoSurfaceToPly(vacancy::VoxelCarver* carver, const std::string& ply_path,
                            double iso_level = 0.0, bool linear_interp = true);

/* Mesh::WritePly for any mesh without vertex colours, formatted on GPU `device`. */
bool WritePly(const vacancy::Mesh& mesh, const std::string& ply_path, int device = 0);

}  // namespace vacancy_b200
