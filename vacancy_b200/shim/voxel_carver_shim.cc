// Drop-in replacement for the reference's src/vacancy/voxel_carver.cc, marching_cubes*.cc and
// extract_voxel.cc: it implements every symbol that include/vacancy/voxel_carver.h declares
// (voxel_carver.h:39-128), against the reference's UNCHANGED headers, on top of the C ABI of
// libvacancy_b200.so (include/vacancy_b200.h).  Host code written for the reference - examples.cc
// for one - compiles and links against it without a source change.
//
// What runs where: DistanceTransformL1 / MakeSignedDistanceField, Carve, ExtractVoxel and
// ExtractIsoSurface run on the GPU through the C ABI, with no CPU fallback - if the library
// reports an error the call logs it with LOGE and returns false (or leaves the mesh empty), the
// reference's own error convention.  Only the small VoxelGrid value class stays
// ordinary host code.
//
// The reference class has no member that could hold a device handle (voxel_carver.h:95-97), and the
// header must not change, so carvers are tracked in a side table keyed by `this`; the AoS
// `voxel_grid_` member is simply never allocated (the grid lives on the device as SoA).  Because
// the key is the object's address, a VoxelCarver must stay where it was when Init() ran: a copy or
// a moved-to object has no device grid (Carve returns false on it, as before Init), and the
// original keeps its handle until it is destroyed.
//
// Placement comes from the environment, so that host code needs no change to use more GPUs:
//   VACANCY_B200_DEVICE=<n>        the one GPU to run on (default 0);
//   VACANCY_B200_DEVICES=0,1,2,3   carve ONE grid on several GPUs (z slabs, vcb_sharded_*): same
//                                  calls, same results, the mesh arrives whole; the engine places
//                                  the slab boundaries itself (VACANCY_B200_AUTO_BALANCE=0: equal).
#include "vacancy/voxel_carver.h"

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <limits>
#include <mutex>
#include <unordered_map>
#include <vector>

#include "vacancy_b200.h"
#include "vacancy_b200_batch.h"

namespace {

// One grid: either a single carver or a z-slab engine over several GPUs; the calls below dispatch.
struct Handle {
  vcb_carver* one = nullptr;
  vcb_sharded* many = nullptr;
  const char* error() const { return many ? vcb_sharded_last_error(many) : vcb_last_error(one); }
  int carve_masks(int32_t n, const vcb_view* v, const uint8_t* const* m, float* const* out) const {
    return many ? vcb_sharded_carve_masks(many, n, v, m, out) : vcb_carve_masks(one, n, v, m, out);
  }
  int carve_sdfs(int32_t n, const vcb_view* v, const float* const* img) const {
    return many ? vcb_sharded_carve_sdfs(many, n, v, img) : vcb_carve_sdfs(one, n, v, img);
  }
  int iso_count(double iso, int32_t interp, int64_t* nv, int64_t* nf) const {
    return many ? vcb_sharded_extract_iso_count(many, iso, interp, nv, nf)
                : vcb_extract_iso_count(one, iso, interp, nv, nf);
  }
  int iso_fill(float* v, int32_t* f) const {
    return many ? vcb_sharded_extract_iso_fill(many, v, f) : vcb_extract_iso_fill(one, v, f);
  }
  int voxel_count(int32_t inside_empty, int64_t* n) const {
    return many ? vcb_sharded_extract_voxel_count(many, inside_empty, n)
                : vcb_extract_voxel_count(one, inside_empty, n);
  }
  int voxel_fill(float* v, int32_t* f) const {
    return many ? vcb_sharded_extract_voxel_fill(many, v, f) : vcb_extract_voxel_fill(one, v, f);
  }
  // the carvers that hold the parts of the last extracted mesh, in slab order
  std::vector<vcb_carver*> parts() const {
    std::vector<vcb_carver*> out;
    if (!many) {
      out.push_back(one);
      return out;
    }
    int32_t n_local = 0;
    vcb_sharded_info(many, nullptr, &n_local, nullptr);
    for (int32_t i = 0; i < n_local; i++) {
      vcb_carver* c = nullptr;
      vcb_sharded_local(many, i, &c, nullptr);
      out.push_back(c);
    }
    return out;
  }
  void destroy() {
    if (one) vcb_destroy(one);
    if (many) vcb_sharded_destroy(many);
    one = nullptr;
    many = nullptr;
  }
};

std::mutex g_mutex;
std::unordered_map<const vacancy::VoxelCarver*, Handle> g_carvers;

const Handle* find_carver(const vacancy::VoxelCarver* self) {
  std::lock_guard<std::mutex> lock(g_mutex);
  auto it = g_carvers.find(self);
  return it == g_carvers.end() ? nullptr : &it->second;  // node addresses are stable
}

void drop_carver(const vacancy::VoxelCarver* self) {
  Handle h;
  {
    std::lock_guard<std::mutex> lock(g_mutex);
    auto it = g_carvers.find(self);
    if (it != g_carvers.end()) {
      h = it->second;
      g_carvers.erase(it);
    }
  }
  h.destroy();
}

int device_ordinal() {
  const char* e = std::getenv("VACANCY_B200_DEVICE");
  return e != nullptr ? std::atoi(e) : 0;
}

// VACANCY_B200_DEVICES=0,1,... -> device list (empty: single-GPU mode)
std::vector<int32_t> device_list() {
  std::vector<int32_t> out;
  const char* e = std::getenv("VACANCY_B200_DEVICES");
  if (e == nullptr) return out;
  for (const char* p = e; *p != 0;) {
    char* end = nullptr;
    const long v = std::strtol(p, &end, 10);
    if (end == p) break;
    out.push_back(static_cast<int32_t>(v));
    p = (*end == ',') ? end + 1 : end;
    if (*end != ',' && *end != 0) break;
  }
  return out;
}

// Camera -> vcb_view. The fusion kernel restates exactly two projection models (camera.cc:131-137
// and :201-205); any other subclass of the abstract Camera cannot run on the device.
bool make_view(const vacancy::Camera& camera, int width, int height,
               const Eigen::Vector2i& roi_min, const Eigen::Vector2i& roi_max, vcb_view* v) {
  std::memset(v, 0, sizeof(*v));
  const Eigen::Affine3f w2c = camera.w2c().cast<float>();  // voxel_carver.cc:438
  for (int r = 0; r < 3; r++) {
    for (int c = 0; c < 4; c++) v->w2c[r * 4 + c] = w2c.matrix()(r, c);
  }
  if (const auto* pin = dynamic_cast<const vacancy::PinholeCamera*>(&camera)) {
    v->model = VCB_CAMERA_PINHOLE;
    v->fx = pin->focal_length()[0];
    v->fy = pin->focal_length()[1];
    v->cx = pin->principal_point()[0];
    v->cy = pin->principal_point()[1];
  } else if (dynamic_cast<const vacancy::OrthoCamera*>(&camera) != nullptr) {
    v->model = VCB_CAMERA_ORTHO;
  } else {
    vacancy::LOGE("VoxelCarver::Carve unsupported Camera subclass (pinhole and ortho only)\n");
    return false;
  }
  v->width = width;
  v->height = height;
  v->roi_min[0] = roi_min.x();
  v->roi_min[1] = roi_min.y();
  v->roi_max[0] = roi_max.x();
  v->roi_max[1] = roi_max.y();
  return true;
}

bool report(const Handle* h, int rc, const char* what) {
  if (rc == VCB_OK) return true;
  vacancy::LOGE("%s: %s\n", what, h != nullptr ? h->error() : vcb_last_error(nullptr));
  return false;
}

}  // namespace

namespace vacancy {

const float InvalidSdf::kVal = std::numeric_limits<float>::lowest();

// ---- free functions, voxel_carver.h:120-128 -----------------------------------------------------
void DistanceTransformL1(const Image1b& mask, const Eigen::Vector2i& roi_min,
                         const Eigen::Vector2i& roi_max, Image1f* dist) {
  dist->Init(mask.width(), mask.height(), 0.0f);
  const int32_t rmin[2] = {roi_min.x(), roi_min.y()}, rmax[2] = {roi_max.x(), roi_max.y()};
  report(nullptr,
         vcb_distance_transform_l1(device_ordinal(), mask.data().data(), mask.width(),
                                   mask.height(), rmin, rmax, dist->data_ptr()->data()),
         "DistanceTransformL1");
}

void MakeSignedDistanceField(const Image1b& mask, const Eigen::Vector2i& roi_min,
                             const Eigen::Vector2i& roi_max, Image1f* dist,
                             bool minmax_normalize, bool use_truncation, float truncation_band) {
  dist->Init(mask.width(), mask.height(), 0.0f);
  const int32_t rmin[2] = {roi_min.x(), roi_min.y()}, rmax[2] = {roi_max.x(), roi_max.y()};
  report(nullptr,
         vcb_make_signed_distance_field(device_ordinal(), mask.data().data(), mask.width(),
                                        mask.height(), rmin, rmax, minmax_normalize ? 1 : 0,
                                        use_truncation ? 1 : 0, truncation_band,
                                        dist->data_ptr()->data()),
         "MakeSignedDistanceField");
}

// Debugging colour map (examples.cc:131-133), voxel_carver.cc:239-267
void SignedDistance2Color(const Image1f& sdf, Image3b* vis_sdf, float min_negative_d,
                          float max_positive_d) {
  assert(min_negative_d < 0 && 0 < max_positive_d && vis_sdf != nullptr);
  vis_sdf->Init(sdf.width(), sdf.height());
  report(nullptr,
         vcb_signed_distance_to_color(device_ordinal(), sdf.data().data(), sdf.width(),
                                      sdf.height(), min_negative_d, max_positive_d,
                                      vis_sdf->data_ptr()->data()),
         "SignedDistance2Color");
}

// ---- Voxel / VoxelGrid: plain host value types kept for source compatibility ----------------------
Voxel::Voxel() {}
Voxel::~Voxel() {}
VoxelGrid::VoxelGrid() {}
VoxelGrid::~VoxelGrid() {}

bool VoxelGrid::Init(const Eigen::Vector3f& bb_max, const Eigen::Vector3f& bb_min,
                     float resolution) {
  if (resolution < std::numeric_limits<float>::min()) {
    LOGE("resolution must be positive %f\n", resolution);
    return false;
  }
  if ((bb_max.array() <= bb_min.array()).any()) {
    LOGE("input bounding box is invalid\n");
    return false;
  }
  bb_max_ = bb_max;
  bb_min_ = bb_min;
  resolution_ = resolution;
  const Eigen::Vector3f diff = bb_max_ - bb_min_;
  for (int a = 0; a < 3; a++) voxel_num_[a] = static_cast<int>(diff[a] / resolution_);
  xy_slice_num_ = voxel_num_[0] * voxel_num_[1];
  voxels_.assign(static_cast<size_t>(xy_slice_num_) * voxel_num_[2], Voxel());
  const float half = resolution_ * 0.5f;
  int id = 0;
  for (int z = 0; z < voxel_num_[2]; z++) {
    for (int y = 0; y < voxel_num_[1]; y++) {
      for (int x = 0; x < voxel_num_[0]; x++, id++) {
        Voxel& v = voxels_[id];
        v.index = Eigen::Vector3i(x, y, z);
        v.id = id;
        const int k[3] = {x, y, z};
        for (int a = 0; a < 3; a++) {
          v.pos[a] = diff[a] * (static_cast<float>(k[a]) / static_cast<float>(voxel_num_[a])) +
                     bb_min_[a] + half;
        }
        v.sdf = InvalidSdf::kVal;
      }
    }
  }
  return true;
}

const Eigen::Vector3i& VoxelGrid::voxel_num() const { return voxel_num_; }
const Voxel& VoxelGrid::get(int x, int y, int z) const {
  return voxels_[z * xy_slice_num_ + (y * voxel_num_.x() + x)];
}
Voxel* VoxelGrid::get_ptr(int x, int y, int z) {
  return &voxels_[z * xy_slice_num_ + (y * voxel_num_.x() + x)];
}
float VoxelGrid::resolution() const { return resolution_; }
void VoxelGrid::ResetOnSurface() {
  for (Voxel& v : voxels_) v.on_surface = false;
}
bool VoxelGrid::initialized() const { return !voxels_.empty(); }

// ---- VoxelCarver, voxel_carver.h:95-118 ----------------------------------------------------------
VoxelCarver::VoxelCarver() {}
VoxelCarver::~VoxelCarver() { drop_carver(this); }
VoxelCarver::VoxelCarver(VoxelCarverOption option) { set_option(option); }
void VoxelCarver::set_option(VoxelCarverOption option) { option_ = option; }

bool VoxelCarver::Init() {  // voxel_carver.cc:375-392; calling it again re-creates the grid
  drop_carver(this);
  if (vcb_abi_version() != VCB_ABI_VERSION) {
    LOGE("VoxelCarver::Init libvacancy_b200.so has ABI %d, this shim was built for %d\n",
         vcb_abi_version(), VCB_ABI_VERSION);
    return false;
  }
  vcb_options o;
  vcb_options_default(&o);
  for (int a = 0; a < 3; a++) {
    o.bb_max[a] = option_.bb_max[a];
    o.bb_min[a] = option_.bb_min[a];
  }
  o.resolution = option_.resolution;
  o.sdf_minmax_normalize = option_.sdf_minmax_normalize ? 1 : 0;
  const VoxelUpdateOption& u = option_.update_option;
  o.voxel_update = static_cast<int32_t>(u.voxel_update);
  o.sdf_interp = static_cast<int32_t>(u.sdf_interp);
  o.update_outside = static_cast<int32_t>(u.update_outside);
  o.voxel_max_update_num = u.voxel_max_update_num;
  o.voxel_update_weight = u.voxel_update_weight;
  o.use_truncation = u.use_truncation ? 1 : 0;
  o.truncation_band = u.truncation_band;
  o.device = device_ordinal();
  Handle h;
  const std::vector<int32_t> devices = device_list();
  if (devices.size() > 1) {
    if (!report(nullptr,
                vcb_sharded_create(&o, devices.data(), static_cast<int32_t>(devices.size()), nullptr,
                                   &h.many),
                "VoxelCarver::Init"))
      return false;
    // host code written for the reference cannot re-balance the slabs: let the engine place the
    // boundaries itself when the first views arrive (VACANCY_B200_AUTO_BALANCE=0 keeps them equal)
    const char* ab = std::getenv("VACANCY_B200_AUTO_BALANCE");
    vcb_sharded_set_auto_balance(h.many, (ab != nullptr && ab[0] == '0') ? 0 : 1);
  } else {
    if (devices.size() == 1) o.device = devices[0];
    if (!report(nullptr, vcb_create(&o, &h.one), "VoxelCarver::Init")) return false;
  }
  std::lock_guard<std::mutex> lock(g_mutex);
  g_carvers[this] = h;
  return true;
}

bool VoxelCarver::Carve(const Camera& camera, const Image1b& silhouette,
                        const Eigen::Vector2i& roi_min, const Eigen::Vector2i& roi_max,
                        Image1f* sdf) {
  const Handle* h = find_carver(this);
  if (h == nullptr) {  // the reference dereferences a null grid here (voxel_carver.cc:397)
    LOGE("VoxelCarver::Carve voxel grid has not been initialized\n");
    return false;
  }
  vcb_view view;
  if (!make_view(camera, silhouette.width(), silhouette.height(), roi_min, roi_max, &view)) {
    return false;
  }
  const uint8_t* mask = silhouette.data().data();
  float* out = nullptr;
  if (sdf != nullptr) {  // filled like voxel_carver.cc:104 / :405
    sdf->Init(silhouette.width(), silhouette.height(), 0.0f);
    out = sdf->data_ptr()->data();
  }
  return report(h, h->carve_masks(1, &view, &mask, sdf != nullptr ? &out : nullptr),
                "VoxelCarver::Carve");
}

bool VoxelCarver::Carve(const Camera& camera, const Eigen::Vector2i& roi_min,
                        const Eigen::Vector2i& roi_max, const Image1f& sdf) {
  const Handle* h = find_carver(this);
  if (h == nullptr) {
    LOGE("VoxelCarver::Carve voxel grid has not been initialized\n");
    return false;
  }
  vcb_view view;
  if (!make_view(camera, sdf.width(), sdf.height(), roi_min, roi_max, &view)) return false;
  const float* img = sdf.data().data();
  return report(h, h->carve_sdfs(1, &view, &img), "VoxelCarver::Carve");
}

bool VoxelCarver::Carve(const Camera& camera, const Image1b& silhouette) {
  return Carve(camera, silhouette, Eigen::Vector2i(0, 0),
               Eigen::Vector2i(silhouette.width() - 1, silhouette.height() - 1), nullptr);
}

bool VoxelCarver::Carve(const Camera& camera, const Image1b& silhouette, Image1f* sdf) {
  return Carve(camera, silhouette, Eigen::Vector2i(0, 0),
               Eigen::Vector2i(silhouette.width() - 1, silhouette.height() - 1), sdf);
}

bool VoxelCarver::Carve(const Camera& camera, const Image1f& sdf) {
  return Carve(camera, Eigen::Vector2i(0, 0), Eigen::Vector2i(sdf.width() - 1, sdf.height() - 1),
               sdf);
}

bool VoxelCarver::Carve(const std::vector<Camera>& cameras,
                        const std::vector<Image1b>& silhouettes) {
  assert(cameras.size() == silhouettes.size());
  for (size_t i = 0; i < cameras.size(); i++) {  // voxel_carver.cc:520-525
    if (!Carve(cameras[i], silhouettes[i])) return false;
  }
  return true;
}

namespace {
// Mesh setters deep-copy (mesh.cc:241-271); Vector3f / Vector3i are 3 packed scalars
void fill_mesh(Mesh* mesh, const std::vector<float>& v, const std::vector<int32_t>& f) {
  std::vector<Eigen::Vector3f> vertices(v.size() / 3);
  std::vector<Eigen::Vector3i> faces(f.size() / 3);
  static_assert(sizeof(Eigen::Vector3f) == 12 && sizeof(Eigen::Vector3i) == 12, "packed");
  if (!v.empty()) std::memcpy(vertices.data(), v.data(), v.size() * sizeof(float));
  if (!f.empty()) std::memcpy(faces.data(), f.data(), f.size() * sizeof(int32_t));
  mesh->set_vertices(vertices);
  mesh->set_vertex_indices(faces);
}
}  // namespace

void VoxelCarver::ExtractVoxel(Mesh* mesh, bool inside_empty) {  // voxel_carver.cc:530-538
  mesh->Clear();
  const Handle* h = find_carver(this);
  if (h == nullptr) return;
  int64_t n = 0;
  if (!report(h, h->voxel_count(inside_empty ? 1 : 0, &n), "VoxelCarver::ExtractVoxel")) {
    return;
  }
  std::vector<float> v(static_cast<size_t>(n) * 72);
  std::vector<int32_t> f(static_cast<size_t>(n) * 36);
  if (n > 0 && !report(h, h->voxel_fill(v.data(), f.data()), "VoxelCarver::ExtractVoxel")) {
    return;
  }
  fill_mesh(mesh, v, f);
}

void VoxelCarver::ExtractIsoSurface(Mesh* mesh, double iso_level, bool linear_interp) {
  mesh->Clear();  // marching_cubes.cc:68
  const Handle* h = find_carver(this);
  if (h == nullptr) return;
  int64_t nv = 0, nf = 0;
  if (!report(h, h->iso_count(iso_level, linear_interp ? 1 : 0, &nv, &nf),
              "VoxelCarver::ExtractIsoSurface")) {
    return;
  }
  std::vector<float> v(static_cast<size_t>(nv) * 3);
  std::vector<int32_t> f(static_cast<size_t>(nf) * 3);
  if (!report(h, h->iso_fill(v.data(), f.data()), "VoxelCarver::ExtractIsoSurface")) {
    return;
  }
  fill_mesh(mesh, v, f);
}

}  // namespace vacancy

namespace vacancy_b200 {

bool CarveBatch(vacancy::VoxelCarver* carver, const std::vector<const vacancy::Camera*>& cameras,
                const std::vector<const vacancy::Image1b*>& silhouettes) {
  if (carver == nullptr || cameras.size() != silhouettes.size()) {
    vacancy::LOGE("CarveBatch: need a carver and as many cameras as silhouettes\n");
    return false;
  }
  const Handle* h = find_carver(carver);
  if (h == nullptr) {
    vacancy::LOGE("VoxelCarver::Carve voxel grid has not been initialized\n");
    return false;
  }
  std::vector<vcb_view> views(cameras.size());
  std::vector<const uint8_t*> masks(cameras.size());
  for (size_t i = 0; i < cameras.size(); i++) {
    const vacancy::Image1b& m = *silhouettes[i];
    if (!make_view(*cameras[i], m.width(), m.height(), Eigen::Vector2i(0, 0),
                   Eigen::Vector2i(m.width() - 1, m.height() - 1), &views[i])) {
      return false;
    }
    masks[i] = m.data().data();
  }
  return report(h, h->carve_masks(static_cast<int32_t>(views.size()), views.data(), masks.data(),
                                  nullptr),
                "CarveBatch");
}

namespace {
bool write_file(const std::string& path, const std::vector<char>& text) {
  FILE* f = fopen(path.c_str(), "wb");
  if (f == nullptr) {
    vacancy::LOGE("couldn't open ply: %s\n", path.c_str());  // mesh.cc:586-589
    return false;
  }
  const bool ok = text.empty() || fwrite(text.data(), 1, text.size(), f) == text.size();
  return fclose(f) == 0 && ok;
}
}  // namespace

bool ExtractIsoSurfaceToPly(vacancy::VoxelCarver* carver, const std::string& ply_path,
                            double iso_level, bool linear_interp) {
  const Handle* h = carver ? find_carver(carver) : nullptr;
  if (h == nullptr) {
    vacancy::LOGE("ExtractIsoSurfaceToPly: voxel grid has not been initialized\n");
    return false;
  }
  int64_t nv = 0, nf = 0;
  if (!report(h, h->iso_count(iso_level, linear_interp ? 1 : 0, &nv, &nf), "ExtractIsoSurfaceToPly"))
    return false;
  // the file is the header for the global counts, every slab's vertex lines in slab order, then
  // every slab's face lines (vertex ids are already global)
  const std::vector<vcb_carver*> parts = h->parts();
  std::vector<size_t> vb(parts.size()), fb(parts.size());
  size_t hb = 0, total_v = 0, total_f = 0;
  vcb_ply_header(nv, nf, nullptr, 0, &hb);
  for (size_t i = 0; i < parts.size(); i++) {
    if (vcb_ply_body_size(parts[i], &vb[i], &fb[i]) != VCB_OK) {
      vacancy::LOGE("ExtractIsoSurfaceToPly: %s\n", vcb_last_error(parts[i]));
      return false;
    }
    total_v += vb[i];
    total_f += fb[i];
  }
  std::vector<char> text(hb + total_v + total_f);
  vcb_ply_header(nv, nf, text.data(), text.size(), nullptr);
  size_t ov = hb, of = hb + total_v;
  for (size_t i = 0; i < parts.size(); i++) {
    if (vcb_ply_body_fill(parts[i], text.data() + ov, text.data() + of) != VCB_OK) {
      vacancy::LOGE("ExtractIsoSurfaceToPly: %s\n", vcb_last_error(parts[i]));
      return false;
    }
    ov += vb[i];
    of += fb[i];
  }
  return write_file(ply_path, text);
}

bool WritePly(const vacancy::Mesh& mesh, const std::string& ply_path, int device) {
  if (!mesh.vertex_colors().empty()) {  // colours are not covered: the reference's own writer
    return mesh.WritePly(ply_path);
  }
  const auto& v = mesh.vertices();
  const auto& f = mesh.vertex_indices();
  static_assert(sizeof(Eigen::Vector3f) == 12 && sizeof(Eigen::Vector3i) == 12, "packed triples");
  const float* vp = v.empty() ? nullptr : v[0].data();
  const int32_t* fp = f.empty() ? nullptr : f[0].data();
  size_t n = 0;
  if (vcb_ply_format(device, vp, static_cast<int64_t>(v.size()), fp, static_cast<int64_t>(f.size()),
                     nullptr, 0, &n) != VCB_OK) {
    vacancy::LOGE("WritePly: %s\n", vcb_last_error(nullptr));
    return false;
  }
  std::vector<char> text(n);
  if (vcb_ply_format(device, vp, static_cast<int64_t>(v.size()), fp, static_cast<int64_t>(f.size()),
                     text.data(), text.size(), &n) != VCB_OK) {
    vacancy::LOGE("WritePly: %s\n", vcb_last_error(nullptr));
    return false;
  }
  return write_file(ply_path, text);
}

}  // namespace vacancy_b200
