// TEST INFRASTRUCTURE ONLY - never linked into, imported by or called from the product path.
//
// C-linkage harness over the UNMODIFIED reference (unclearness/vacancy), compiled where the
// sources lie under /root/reference by oracle/Makefile into oracle/_ref/libvacancy_ref.so.
// It lets the tests (ctypes) and bench.py's cpu_baseline / --impl reference leg drive the
// reference's own VoxelCarver (include/vacancy/voxel_carver.h:95-128) and free functions
// (voxel_carver.h:120-126) on plain arrays, and read back the private voxel grid.
//
// The private grid is reached with `#define class struct` around the reference header
// (SURVEY.md section 8c): std / Eigen headers are included first so only vacancy's own
// classes become structs.
#include <algorithm>
#include <array>
#include <cassert>
#include <chrono>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <fstream>
#include <functional>
#include <iomanip>
#include <limits>
#include <map>
#include <memory>
#include <random>
#include <sstream>
#include <string>
#include <utility>
#include <vector>

#include "Eigen/Geometry"

#define class struct
#include "vacancy/voxel_carver.h"
#undef class

#ifdef _OPENMP
#include <omp.h>
#endif

namespace {

struct RefCarver {
  vacancy::VoxelCarver carver;
  vacancy::Mesh mesh;
  bool ok{false};
};

Eigen::Affine3d AffineFromRowMajor(const double* m16) {
  Eigen::Affine3d a = Eigen::Affine3d::Identity();
  for (int r = 0; r < 3; r++) {
    for (int c = 0; c < 4; c++) {
      a.matrix()(r, c) = m16[r * 4 + c];
    }
  }
  return a;
}

// model 0 = pinhole, 1 = ortho (the only two concrete cameras: camera.h:66,116)
std::unique_ptr<vacancy::Camera> MakeCamera(int model, int width, int height,
                                            const double* c2w16, float fx,
                                            float fy, float cx, float cy) {
  Eigen::Affine3d c2w = AffineFromRowMajor(c2w16);
  if (model == 0) {
    return std::unique_ptr<vacancy::Camera>(new vacancy::PinholeCamera(
        width, height, c2w, Eigen::Vector2f(cx, cy), Eigen::Vector2f(fx, fy)));
  }
  return std::unique_ptr<vacancy::Camera>(
      new vacancy::OrthoCamera(width, height, c2w));
}

void FillMask(const uint8_t* src, int w, int h, vacancy::Image1b* img) {
  img->Init(w, h, 0);
  std::memcpy(img->data_ptr()->data(), src, static_cast<size_t>(w) * h);
}

}  // namespace

extern "C" {

int ref_omp_threads() {
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}

// torch.distributed.run exports OMP_NUM_THREADS=1 to its workers: the timed reference arm sets the
// thread count explicitly (n < 0: every core of the machine, 0: only report).
int ref_set_omp_threads(int n) {
#ifdef _OPENMP
  if (n < 0) n = omp_get_num_procs();
  if (n > 0) omp_set_num_threads(n);
  return omp_get_max_threads();
#else
  (void)n;
  return 1;
#endif
}

void ref_set_quiet(int quiet) {
  vacancy::set_log_level(quiet ? vacancy::LogLevel::kNone
                               : vacancy::LogLevel::kVerbose);
}

float ref_invalid_sdf() { return vacancy::InvalidSdf::kVal; }

// voxel_carver.cc:102-167
void ref_distance_transform_l1(const uint8_t* mask, int w, int h, int rx0,
                               int ry0, int rx1, int ry1, float* out) {
  vacancy::Image1b m;
  FillMask(mask, w, h, &m);
  vacancy::Image1f d;
  vacancy::DistanceTransformL1(m, Eigen::Vector2i(rx0, ry0),
                               Eigen::Vector2i(rx1, ry1), &d);
  std::memcpy(out, d.data().data(), sizeof(float) * w * h);
}

// voxel_carver.cc:169-237
void ref_make_sdf(const uint8_t* mask, int w, int h, int rx0, int ry0, int rx1,
                  int ry1, int normalize, int truncate, float band,
                  float* out) {
  vacancy::Image1b m;
  FillMask(mask, w, h, &m);
  vacancy::Image1f d;
  vacancy::MakeSignedDistanceField(m, Eigen::Vector2i(rx0, ry0),
                                   Eigen::Vector2i(rx1, ry1), &d,
                                   normalize != 0, truncate != 0, band);
  std::memcpy(out, d.data().data(), sizeof(float) * w * h);
}

// voxel_carver.cc:239-267
void ref_signed_distance_to_color(const float* sdf, int w, int h, float min_negative_d,
                                  float max_positive_d, uint8_t* rgb) {
  vacancy::Image1f img;
  img.Init(w, h, 0.0f);
  std::memcpy(img.data_ptr()->data(), sdf, sizeof(float) * w * h);
  vacancy::Image3b vis;
  vacancy::SignedDistance2Color(img, &vis, min_negative_d, max_positive_d);
  std::memcpy(rgb, vis.data().data(), static_cast<size_t>(w) * h * 3);
}

// Times `reps` calls of MakeSignedDistanceField on an already-built Image1b; seconds.
double ref_time_make_sdf(const uint8_t* mask, int w, int h, int normalize,
                         int truncate, float band, int reps) {
  vacancy::Image1b m;
  FillMask(mask, w, h, &m);
  vacancy::Image1f d;
  auto t0 = std::chrono::steady_clock::now();
  for (int i = 0; i < reps; i++) {
    vacancy::MakeSignedDistanceField(m, Eigen::Vector2i(0, 0),
                                     Eigen::Vector2i(w - 1, h - 1), &d,
                                     normalize != 0, truncate != 0, band);
  }
  auto t1 = std::chrono::steady_clock::now();
  return std::chrono::duration<double>(t1 - t0).count();
}

// camera.cc:39-42 (set_c2w -> inverse in double) + voxel_carver.cc:438 (cast<float>)
void ref_w2c_from_c2w(const double* c2w16, float* w2c12) {
  vacancy::PinholeCamera cam(4, 4, AffineFromRowMajor(c2w16),
                             Eigen::Vector2f(1, 1), Eigen::Vector2f(1, 1));
  Eigen::Affine3f w2c = cam.w2c().cast<float>();
  for (int r = 0; r < 3; r++) {
    for (int c = 0; c < 4; c++) {
      w2c12[r * 4 + c] = w2c.matrix()(r, c);
    }
  }
}

// examples.cc:40-52: pose = Translation3d * Quaterniond
void ref_tum_pose_to_c2w(const double* t3, const double* qxyzw, double* c2w16) {
  Eigen::Translation3d t(t3[0], t3[1], t3[2]);
  Eigen::Quaterniond q;
  q.x() = qxyzw[0];
  q.y() = qxyzw[1];
  q.z() = qxyzw[2];
  q.w() = qxyzw[3];
  Eigen::Affine3d a = t * q;
  for (int r = 0; r < 4; r++) {
    for (int c = 0; c < 4; c++) {
      c2w16[r * 4 + c] = a.matrix()(r, c);
    }
  }
}

// camera.cc ray_c: the synthetic-silhouette generator of SURVEY.md 8(d) uses it
void ref_pinhole_ray_c(int width, int height, float fx, float fy, float cx,
                       float cy, float x, float y, float* dir3) {
  vacancy::PinholeCamera cam(width, height, Eigen::Affine3d::Identity(),
                             Eigen::Vector2f(cx, cy), Eigen::Vector2f(fx, fy));
  Eigen::Vector3f d;
  cam.ray_c(x, y, &d);
  dir3[0] = d.x();
  dir3[1] = d.y();
  dir3[2] = d.z();
}

// image.h:77-101 (stb). Returns 0 on success. out may be NULL to query the size.
int ref_load_png_gray(const char* path, uint8_t* out, int* w, int* h) {
  vacancy::Image1b img;
  if (!img.Load(path)) {
    return 1;
  }
  *w = img.width();
  *h = img.height();
  if (out != nullptr) {
    std::memcpy(out, img.data().data(),
                static_cast<size_t>(img.width()) * img.height());
  }
  return 0;
}

// voxel_carver.h:43-60 options, field for field; returns NULL when Init() fails
void* ref_carver_create(const float* bb_min, const float* bb_max,
                        float resolution, int sdf_minmax_normalize,
                        int voxel_update, int sdf_interp, int update_outside,
                        int voxel_max_update_num, float voxel_update_weight,
                        int use_truncation, float truncation_band) {
  vacancy::VoxelCarverOption opt;
  opt.bb_min = Eigen::Vector3f(bb_min[0], bb_min[1], bb_min[2]);
  opt.bb_max = Eigen::Vector3f(bb_max[0], bb_max[1], bb_max[2]);
  opt.resolution = resolution;
  opt.sdf_minmax_normalize = sdf_minmax_normalize != 0;
  opt.update_option.voxel_update =
      static_cast<vacancy::VoxelUpdate>(voxel_update);
  opt.update_option.sdf_interp =
      static_cast<vacancy::SdfInterpolation>(sdf_interp);
  opt.update_option.update_outside =
      static_cast<vacancy::UpdateOutsideImage>(update_outside);
  opt.update_option.voxel_max_update_num = voxel_max_update_num;
  opt.update_option.voxel_update_weight = voxel_update_weight;
  opt.update_option.use_truncation = use_truncation != 0;
  opt.update_option.truncation_band = truncation_band;
  RefCarver* rc = new RefCarver;
  rc->carver.set_option(opt);
  rc->ok = rc->carver.Init();
  if (!rc->ok) {
    delete rc;
    return nullptr;
  }
  return rc;
}

void ref_carver_destroy(void* h) { delete static_cast<RefCarver*>(h); }

void ref_carver_dims(void* h, int* dims3) {
  RefCarver* rc = static_cast<RefCarver*>(h);
  const Eigen::Vector3i& n = rc->carver.voxel_grid_->voxel_num();
  dims3[0] = n.x();
  dims3[1] = n.y();
  dims3[2] = n.z();
}

// Carve(camera, silhouette, roi_min, roi_max, sdf*) voxel_carver.cc:394-413
int ref_carver_carve_mask(void* h, int model, int width, int height,
                          const double* c2w16, float fx, float fy, float cx,
                          float cy, const uint8_t* mask, int mw, int mh,
                          int rx0, int ry0, int rx1, int ry1, float* sdf_out) {
  RefCarver* rc = static_cast<RefCarver*>(h);
  auto cam = MakeCamera(model, width, height, c2w16, fx, fy, cx, cy);
  vacancy::Image1b m;
  FillMask(mask, mw, mh, &m);
  vacancy::Image1f sdf;
  bool ret = rc->carver.Carve(*cam, m, Eigen::Vector2i(rx0, ry0),
                              Eigen::Vector2i(rx1, ry1), &sdf);
  if (sdf_out != nullptr) {
    std::memcpy(sdf_out, sdf.data().data(), sizeof(float) * mw * mh);
  }
  return ret ? 1 : 0;
}

// Carve(camera, roi_min, roi_max, sdf) voxel_carver.cc:415-496. Returns seconds spent in
// the call in *seconds when non-NULL (image construction excluded).
int ref_carver_carve_sdf(void* h, int model, int width, int height,
                         const double* c2w16, float fx, float fy, float cx,
                         float cy, const float* sdf, int sw, int sh, int rx0,
                         int ry0, int rx1, int ry1, double* seconds) {
  RefCarver* rc = static_cast<RefCarver*>(h);
  auto cam = MakeCamera(model, width, height, c2w16, fx, fy, cx, cy);
  vacancy::Image1f img;
  img.Init(sw, sh, 0.0f);
  std::memcpy(img.data_ptr()->data(), sdf, sizeof(float) * sw * sh);
  auto t0 = std::chrono::steady_clock::now();
  bool ret = rc->carver.Carve(*cam, Eigen::Vector2i(rx0, ry0),
                              Eigen::Vector2i(rx1, ry1), img);
  auto t1 = std::chrono::steady_clock::now();
  if (seconds != nullptr) {
    *seconds = std::chrono::duration<double>(t1 - t0).count();
  }
  return ret ? 1 : 0;
}

// SoA dump of the private AoS grid (voxel_carver.h:62-72)
void ref_carver_download(void* h, float* sdf, int* update_num) {
  RefCarver* rc = static_cast<RefCarver*>(h);
  const std::vector<vacancy::Voxel>& v = rc->carver.voxel_grid_->voxels_;
  for (size_t i = 0; i < v.size(); i++) {
    if (sdf != nullptr) sdf[i] = v[i].sdf;
    if (update_num != nullptr) update_num[i] = v[i].update_num;
  }
}

// Overwrites the grid state (tests feed hand-made grids to the extractors).
void ref_carver_upload(void* h, const float* sdf, const int* update_num) {
  RefCarver* rc = static_cast<RefCarver*>(h);
  std::vector<vacancy::Voxel>& v = rc->carver.voxel_grid_->voxels_;
  for (size_t i = 0; i < v.size(); i++) {
    v[i].sdf = sdf[i];
    v[i].update_num = update_num[i];
  }
}

// voxel centres along each axis (voxel_carver.cc:315-326)
void ref_carver_positions(void* h, float* px, float* py, float* pz) {
  RefCarver* rc = static_cast<RefCarver*>(h);
  vacancy::VoxelGrid* g = rc->carver.voxel_grid_.get();
  const Eigen::Vector3i& n = g->voxel_num();
  for (int x = 0; x < n.x(); x++) px[x] = g->get(x, 0, 0).pos.x();
  for (int y = 0; y < n.y(); y++) py[y] = g->get(0, y, 0).pos.y();
  for (int z = 0; z < n.z(); z++) pz[z] = g->get(0, 0, z).pos.z();
}

// ExtractIsoSurface voxel_carver.cc:540-543 -> marching_cubes.cc:63-228
void ref_carver_extract_iso(void* h, double iso_level, int linear_interp,
                            long long* n_vertices, long long* n_faces,
                            double* seconds) {
  RefCarver* rc = static_cast<RefCarver*>(h);
  auto t0 = std::chrono::steady_clock::now();
  rc->carver.ExtractIsoSurface(&rc->mesh, iso_level, linear_interp != 0);
  auto t1 = std::chrono::steady_clock::now();
  *n_vertices = static_cast<long long>(rc->mesh.vertices().size());
  *n_faces = static_cast<long long>(rc->mesh.vertex_indices().size());
  if (seconds != nullptr) {
    *seconds = std::chrono::duration<double>(t1 - t0).count();
  }
}

// ExtractVoxel voxel_carver.cc:530-538 -> extract_voxel.cc:258-317
void ref_carver_extract_voxel(void* h, int inside_empty, long long* n_vertices,
                              long long* n_faces, double* seconds) {
  RefCarver* rc = static_cast<RefCarver*>(h);
  auto t0 = std::chrono::steady_clock::now();
  rc->carver.ExtractVoxel(&rc->mesh, inside_empty != 0);
  auto t1 = std::chrono::steady_clock::now();
  *n_vertices = static_cast<long long>(rc->mesh.vertices().size());
  *n_faces = static_cast<long long>(rc->mesh.vertex_indices().size());
  if (seconds != nullptr) {
    *seconds = std::chrono::duration<double>(t1 - t0).count();
  }
}

// Copies the mesh left by the last extract call (float xyz per vertex, int32 triple per face)
void ref_carver_mesh_copy(void* h, float* vertices, int* faces) {
  RefCarver* rc = static_cast<RefCarver*>(h);
  const auto& v = rc->mesh.vertices();
  const auto& f = rc->mesh.vertex_indices();
  for (size_t i = 0; i < v.size(); i++) {
    vertices[3 * i + 0] = v[i].x();
    vertices[3 * i + 1] = v[i].y();
    vertices[3 * i + 2] = v[i].z();
  }
  for (size_t i = 0; i < f.size(); i++) {
    faces[3 * i + 0] = f[i][0];
    faces[3 * i + 1] = f[i][1];
    faces[3 * i + 2] = f[i][2];
  }
}

// mesh.cc:583-631 ASCII PLY of the last mesh; returns 0 on success
int ref_carver_mesh_write_ply(void* h, const char* path) {
  RefCarver* rc = static_cast<RefCarver*>(h);
  return rc->mesh.WritePly(path) ? 0 : 1;
}

// mesh.cc:583-631 ASCII PLY of an arbitrary mesh given as arrays; returns 0 on success
int ref_write_ply(const char* path, const float* vertices, long long n_vertices, const int* faces,
                  long long n_faces) {
  std::vector<Eigen::Vector3f> v(static_cast<size_t>(n_vertices));
  std::vector<Eigen::Vector3i> f(static_cast<size_t>(n_faces));
  for (long long i = 0; i < n_vertices; i++)
    v[i] = Eigen::Vector3f(vertices[3 * i], vertices[3 * i + 1], vertices[3 * i + 2]);
  for (long long i = 0; i < n_faces; i++)
    f[i] = Eigen::Vector3i(faces[3 * i], faces[3 * i + 1], faces[3 * i + 2]);
  vacancy::Mesh mesh;
  mesh.set_vertices(v);
  mesh.set_vertex_indices(f);
  return mesh.WritePly(path) ? 0 : 1;
}

// Bourke tables as the reference ships them (marching_cubes_lut.cc:15-298)
void ref_mc_tables(int* edge_table256, int* tri_table256x16);

}  // extern "C"

#include "vacancy/marching_cubes_lut.h"

extern "C" void ref_mc_tables(int* edge_table256, int* tri_table256x16) {
  for (int i = 0; i < 256; i++) {
    edge_table256[i] = vacancy::marching_cubes_lut::kEdgeTable[i];
    for (int j = 0; j < 16; j++) {
      tri_table256x16[i * 16 + j] =
          vacancy::marching_cubes_lut::kTriTable[i][j];
    }
  }
}
