// Small host program against the reference's headers + vacancy_b200_batch.h: the bunny scenario of
// examples.cc (same bounding box, intrinsics and poses) carved with ONE CarveBatch call, then
// the iso surface; prints "vertices faces occupied_cubes" for tests/test_gpu_shim_demo.py.  With a
// second argument (an output directory) it also writes the iso surface three ways - the reference's
// Mesh::WritePly, vacancy_b200::WritePly and vacancy_b200::ExtractIsoSurfaceToPly - so that the test
// can compare the files byte for byte.
#include <cstdio>
#include <fstream>
#include <memory>
#include <sstream>

#include "vacancy_b200_batch.h"

int main(int argc, char* argv[]) {
  const std::string data_dir = argc > 1 ? argv[1] : "../data/";
  vacancy::set_log_level(vacancy::LogLevel::kError);
  std::vector<Eigen::Affine3d> poses;
  std::ifstream ifs(data_dir + "tumpose.txt");
  std::string line;
  while (std::getline(ifs, line)) {
    std::istringstream ss(line);
    double id, t[3], q[4];
    ss >> id >> t[0] >> t[1] >> t[2] >> q[0] >> q[1] >> q[2] >> q[3];
    Eigen::Quaterniond quat;
    quat.x() = q[0];
    quat.y() = q[1];
    quat.z() = q[2];
    quat.w() = q[3];
    poses.push_back(Eigen::Translation3d(t[0], t[1], t[2]) * quat);
  }
  vacancy::VoxelCarverOption option;
  option.bb_min = Eigen::Vector3f(-250.000000f, -344.586151f, -129.982697f);
  option.bb_max = Eigen::Vector3f(250.000000f, 150.542343f, 257.329224f);
  for (int a = 0; a < 3; a++) {
    option.bb_min[a] -= 20.0f;
    option.bb_max[a] += 20.0f;
  }
  option.resolution = 10.0f;
  vacancy::VoxelCarver carver(option);
  if (!carver.Init()) return 1;
  std::vector<std::unique_ptr<vacancy::PinholeCamera>> cams;
  std::vector<vacancy::Image1b> masks(poses.size());
  std::vector<const vacancy::Camera*> cam_ptrs;
  std::vector<const vacancy::Image1b*> mask_ptrs;
  for (size_t i = 0; i < poses.size(); i++) {
    cams.emplace_back(new vacancy::PinholeCamera(320, 240, poses[i], Eigen::Vector2f(159.3f, 127.65f),
                                                 Eigen::Vector2f(258.65f, 258.25f)));
    if (!masks[i].Load(data_dir + "mask_" + vacancy::zfill(i) + ".png")) return 2;
    cam_ptrs.push_back(cams.back().get());
    mask_ptrs.push_back(&masks[i]);
  }
  if (!vacancy_b200::CarveBatch(&carver, cam_ptrs, mask_ptrs)) return 3;
  vacancy::Mesh mesh, voxels;
  carver.ExtractIsoSurface(&mesh, 0.0);
  carver.ExtractVoxel(&voxels);
  std::printf("%zu %zu %zu\n", mesh.vertices().size(), mesh.vertex_indices().size(),
              voxels.vertices().size() / 24);
  if (argc > 2) {
    const std::string out = argv[2];
    if (!mesh.WritePly(out + "/surface_reference_writer.ply")) return 4;
    if (!vacancy_b200::WritePly(mesh, out + "/surface_gpu_writer.ply")) return 5;
    if (!vacancy_b200::ExtractIsoSurfaceToPly(&carver, out + "/surface_gpu_direct.ply", 0.0)) return 6;
  }
  return 0;
}
