// Grid checkpoints: the SoA state of a carver's z slab to and from a raw
// file.  The reference never serialises its grid (voxel_carver.h:97 keeps it private; SURVEY.md
// section 5.4) - this is an addition for long carving sessions: save after N views, restore later
// (into the same or a different slab partition) and go on carving with bit-identical results.
//
// File layout (little endian):
//   char     magic[8]  = "VCBGRID1"
//   int32    nx, ny, nz            global grid
//   int32    z_begin, z_end        planes stored in this file
//   int32    implied               1: update_num < 1 exactly where sdf == kVal (a carved grid)
//   int64    max_count             upper bound of update_num
//   vcb_options options            the options of the carver that wrote the file
//   float    sdf[(z_end - z_begin) * ny * nx]
//   int32    update_num[(z_end - z_begin) * ny * nx]
// A file may be restored into any carver of the same grid and carving options whose slab overlaps
// the file's planes; the overlap is copied.
#include <stdio.h>
#include <string.h>

#include <algorithm>

#include "carver_internal.cuh"

using namespace vcb;

namespace {

struct GridFileHeader {
  char magic[8];
  int32_t nx, ny, nz;
  int32_t z_begin, z_end;
  int32_t implied;
  int64_t max_count;
  vcb_options options;
};

constexpr size_t kIoChunk = (size_t)64 << 20;  // bytes per pinned staging round

// the options that decide what the numbers in the grid mean (placement fields do not)
bool same_carving(const vcb_options& a, const vcb_options& b) {
  return memcmp(a.bb_max, b.bb_max, sizeof(a.bb_max)) == 0 &&
         memcmp(a.bb_min, b.bb_min, sizeof(a.bb_min)) == 0 && a.resolution == b.resolution &&
         a.voxel_update == b.voxel_update && a.voxel_max_update_num == b.voxel_max_update_num &&
         a.voxel_update_weight == b.voxel_update_weight && a.use_truncation == b.use_truncation &&
         a.truncation_band == b.truncation_band;
}

struct Pinned {
  void* p = nullptr;
  ~Pinned() {
    if (p) cudaFreeHost(p);
  }
};

}  // namespace

extern "C" {

int vcb_save_grid(vcb_carver* c, const char* path) {
  if (!c || !path) return fail(c, VCB_ERR_INVALID, "null argument");
  int rc = vcb_flush(c);
  if (rc != VCB_OK) return rc;
  FILE* f = fopen(path, "wb");
  if (!f) return fail(c, VCB_ERR_INVALID, "cannot open %s for writing", path);
  GridFileHeader h;
  memset(&h, 0, sizeof(h));
  memcpy(h.magic, "VCBGRID1", 8);
  h.nx = c->nx;
  h.ny = c->ny;
  h.nz = c->nz;
  h.z_begin = c->z_begin;
  h.z_end = c->z_end;
  h.implied = c->un_implied ? 1 : 0;
  h.max_count = c->max_count;
  h.options = c->opt;
  bool ok = fwrite(&h, sizeof(h), 1, f) == 1;
  Pinned buf;
  cudaError_t e = cudaMallocHost(&buf.p, kIoChunk);
  const size_t bytes = (size_t)c->nx * c->ny * (size_t)(c->z_end - c->z_begin) * 4;
  const char* src[2] = {reinterpret_cast<const char*>(c->sdf), reinterpret_cast<const char*>(c->un)};
  for (int a = 0; a < 2 && ok && e == cudaSuccess; a++) {
    for (size_t off = 0; off < bytes && ok && e == cudaSuccess; off += kIoChunk) {
      const size_t n = std::min(kIoChunk, bytes - off);
      e = cudaMemcpyAsync(buf.p, src[a] + off, n, cudaMemcpyDeviceToHost, c->stream);
      if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
      if (e == cudaSuccess) ok = fwrite(buf.p, 1, n, f) == n;
    }
  }
  ok = (fclose(f) == 0) && ok;
  if (e != cudaSuccess) return fail(c, VCB_ERR_CUDA, "vcb_save_grid: %s", cudaGetErrorString(e));
  if (!ok) return fail(c, VCB_ERR_INVALID, "vcb_save_grid: short write to %s", path);
  return VCB_OK;
}

int vcb_load_grid(vcb_carver* c, const char* path, int32_t* planes_loaded) {
  if (!c || !path) return fail(c, VCB_ERR_INVALID, "null argument");
  if (planes_loaded) *planes_loaded = 0;
  int rc = vcb_flush(c);
  if (rc != VCB_OK) return rc;
  FILE* f = fopen(path, "rb");
  if (!f) return fail(c, VCB_ERR_INVALID, "cannot open %s", path);
  GridFileHeader h;
  if (fread(&h, sizeof(h), 1, f) != 1 || memcmp(h.magic, "VCBGRID1", 8) != 0) {
    fclose(f);
    return fail(c, VCB_ERR_INVALID, "%s is not a grid checkpoint", path);
  }
  if (h.nx != c->nx || h.ny != c->ny || h.nz != c->nz || !same_carving(h.options, c->opt)) {
    fclose(f);
    return fail(c, VCB_ERR_INVALID,
                "%s was written for another grid or other carving options (%d x %d x %d)", path,
                h.nx, h.ny, h.nz);
  }
  const int z0 = std::max(h.z_begin, c->z_begin), z1 = std::min(h.z_end, c->z_end);
  if (z0 >= z1) {  // nothing of this file belongs to this slab
    fclose(f);
    return VCB_OK;
  }
  Pinned buf;
  cudaError_t e = cudaMallocHost(&buf.p, kIoChunk);
  const size_t plane = (size_t)c->nx * c->ny * 4;
  const size_t file_array = plane * (size_t)(h.z_end - h.z_begin);
  const size_t bytes = plane * (size_t)(z1 - z0);
  char* dst[2] = {reinterpret_cast<char*>(c->sdf), reinterpret_cast<char*>(c->un)};
  bool ok = true;
  for (int a = 0; a < 2 && ok && e == cudaSuccess; a++) {
    const size_t file_off = sizeof(h) + (size_t)a * file_array + plane * (size_t)(z0 - h.z_begin);
    ok = fseeko(f, (off_t)file_off, SEEK_SET) == 0;
    char* d = dst[a] + plane * (size_t)(z0 - c->z_begin);
    for (size_t off = 0; off < bytes && ok && e == cudaSuccess; off += kIoChunk) {
      const size_t n = std::min(kIoChunk, bytes - off);
      ok = fread(buf.p, 1, n, f) == n;
      if (ok) e = cudaMemcpyAsync(d + off, buf.p, n, cudaMemcpyHostToDevice, c->stream);
      if (ok && e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
    }
  }
  fclose(f);
  if (e != cudaSuccess) return fail(c, VCB_ERR_CUDA, "vcb_load_grid: %s", cudaGetErrorString(e));
  if (!ok) return fail(c, VCB_ERR_INVALID, "vcb_load_grid: %s is truncated", path);
  if (h.max_count > c->max_count) c->max_count = h.max_count;
  if (!h.implied) c->un_implied = false;
  c->iso.n_vertices = c->iso.n_faces = -1;
  c->iso.phase = 0;
  c->vox.n_voxels = -1;
  if (planes_loaded) *planes_loaded = z1 - z0;
  return VCB_OK;
}

}  // extern "C"
