// Internal declarations shared by the CUDA translation units of libvacancy_b200.so.
// sm_100a only; compiled with -fmad=false so that no float/double multiply-add is ever contracted:
// the reference is built for baseline x86-64 (one rounding per operation) and the fused grid must
// reproduce it bit for bit.  Where an FMA is wanted it is written explicitly (__fmaf_rn).
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include <string>
#include <vector>

#include "../../include/vacancy_b200.h"

namespace vcb {

constexpr float kInvalidSdf = -3.402823466e+38f;  // numeric_limits<float>::lowest(), voxel_carver.cc:100
constexpr int kMaxViewsPerLaunch = 64;            // views fused per kernel launch
constexpr int kSdfSubBatch = 8;                   // views per 2D-SDF kernel group (scratch stays L2-sized)
constexpr int kSegVoxels = 32;                    // one extraction segment = 32 voxels along x
constexpr int16_t kRowInf = 32767;                // "no opposite pixel" marker in the int16 passes
constexpr int kMaxImageDim = 16384;               // keeps L1 distances < kRowInf

// ---- per-view parameters of the fusion kernel (lives in the kernel parameter / constant bank) ----
struct FuseView {
  float m[12];         // w2c row-major 3x4
  float fx, fy, cx, cy;
  float roi_min_x, roi_min_y, roi_max_x, roi_max_y;  // ROI as float (compared against float coords)
  int roi_min_xi, roi_min_yi, roi_max_xi, roi_max_yi;
  int pitch;           // image width in pixels
  int model;           // VCB_CAMERA_*
  int fast;            // set by launch_fusion: pinhole with intrinsics in the shared-reciprocal range
  const float* image;  // device SDF image
  const float* max_sdf;  // device scalar: max over the whole image (only read for VCB_OUTSIDE_MAX)
  // 8x8-pixel block flags of the image, one bit per block, tile_rows x tile_words words each:
  // [all 64 texels == 1.0f] followed by [all 64 texels == kVal]; NULL = no flags
  const uint32_t* tiles;
  int tile_words, tile_rows;
};

constexpr int kTilePx = 8;  // edge of a flagged pixel block
inline int tile_words_for(int width) { return ((width + kTilePx - 1) / kTilePx + 31) / 32; }
inline int tile_rows_for(int height) { return (height + kTilePx - 1) / kTilePx; }
inline size_t tile_flag_bytes(int width, int height) {
  return (size_t)2 * tile_words_for(width) * tile_rows_for(height) * sizeof(uint32_t);
}
struct TileBatch {
  const float* image[kSdfSubBatch];
  uint32_t* flags[kSdfSubBatch];
  int width[kSdfSubBatch], height[kSdfSubBatch];
};
void launch_tile_flags(const TileBatch& b, int n, cudaStream_t stream, int64_t* launches);

struct FuseParams {
  float* sdf;            // slab state, plane z_begin first
  int32_t* update_num;
  const float* pos_x;    // global position tables (device)
  const float* pos_y;
  const float* pos_z;
  const float2* wtab;    // weighted-average table: {w*float(n), 1/(w*float(n+1))}, n in [0, wtab_n)
  int wtab_n;
  int wtab_fill;         // entries a voxel can index during this launch (<= wtab_n)
  int nx, ny;
  int z_begin, z_end;    // owned planes
  int n_views;
  int max_update_num;
  int cap_possible;      // host bound: update_num may exceed max_update_num during this launch
  int use_truncation;
  int update_outside;
  float weight;
  float neg_zero;        // -0.0f, set by launch_fusion (an addend the compiler cannot fold, see fusion.cu)
  // per-launch tables: `tables` is a device block of fusion_table_bytes() the caller provides,
  // launch_fusion fills it (prep kernels) and sets the three pointers into it
  unsigned char* tables;
  const float4* view_tab;
  const float4* mz_tab;
  const unsigned char* tile_class;  // [fusion block][view], nullptr = every tile is "mixed"
  // running totals over all launches (device, 8 counters; read and cleared by
  // vcb_fusion_class_counts): [0..4] (tile, view) pairs per TileClass - what the classification
  // skipped; [5], [6] view iterations the warps executed on the all-ones / the sampling paths
  unsigned long long* class_counts;
  int count_iterations;  // fill [5], [6] (vcb_set_statistics; costs about 1 % of the kernel's time)
  FuseView views[kMaxViewsPerLaunch];
};
size_t fusion_table_bytes(int nx, int ny, int nz_slab);
int fusion_tile_planes();  // z planes per fusion tile: a slab pays for whole tiles
void launch_layer_costs(FuseParams& p, int interp, unsigned long long* layer_cost,
                        cudaStream_t stream);

// ---- per-view descriptor of the 2D SDF kernels (device array) ----
struct SdfView {
  const uint8_t* mask;   // device, width x height
  float* out;            // device float image
  unsigned char* scratch;  // sdf_scratch_bytes(width, height) bytes, 256-byte aligned
  int* absmax;           // scratch: [0] max |D| over the ROI (INT_MAX when some pixel has no seed)
  float* max_out;        // device scalar: max over the final float image
  uint32_t* tile_flags;  // optional: 8x8-block flags of the final image (see FuseView::tiles)
  int width, height;
  int rx0, ry0, rx1, ry1;
};

// kSdfSigned: MakeSignedDistanceField (L1, the reference); kSdfDistanceL1: DistanceTransformL1;
// kSdfSignedL2: the same signed field with the exact Euclidean metric (no reference counterpart)
enum SdfMode { kSdfSigned = 0, kSdfDistanceL1 = 1, kSdfSignedL2 = 2 };

// per-view scratch of the 2D SDF kernels (the Euclidean mode needs three more int32 planes)
size_t sdf_scratch_bytes(int width, int height, int mode = kSdfSigned);

// up to kSdfSubBatch view descriptors, passed to the kernels by value (kernel parameter bank)
struct SdfBatch {
  SdfView v[kSdfSubBatch];
};

void launch_sdf2d(const SdfView* views_host, int n_views, int mode, int normalize, int truncate,
                  float band, cudaStream_t stream, int64_t* launches);
void launch_sdf_color(const float* sdf, int64_t n, float min_neg, float max_pos, uint8_t* rgb,
                      cudaStream_t stream);
void launch_image_max(const float* image, int64_t n, float* max_out, cudaStream_t stream,
                      int64_t* launches);
void launch_fusion(FuseParams& p, int interp, int update, cudaStream_t stream,
                   int64_t* launches);

// ---- extraction ----
struct GridDesc {
  const float* sdf;          // slab state (plane z_begin first)
  const int32_t* update_num;
  const float* pos_x;
  const float* pos_y;
  const float* pos_z;
  int nx, ny, nz;            // GLOBAL dims
  int z_begin, z_end;        // owned planes
  int halo;                  // number of valid sdf planes stored below z_begin (0 or 2)
  float resolution;
};

// Entry points that are not tied to a carver (the free functions) switch to the device they were
// asked to run on and put the caller's current device back when they return.
struct DeviceGuard {
  int prev = -1;
  DeviceGuard() { if (cudaGetDevice(&prev) != cudaSuccess) prev = -1; }
  ~DeviceGuard() { if (prev >= 0) cudaSetDevice(prev); }
};

struct Scratch {  // grow-only device buffer
  void* ptr = nullptr;
  size_t bytes = 0;
  cudaError_t reserve(size_t n);
  void release();
};

struct PinnedBuffer {  // grow-only page-locked host buffer (ingest.cu)
  void* ptr = nullptr;
  size_t bytes = 0;
  cudaError_t reserve(size_t n);
  void release();
};

// exclusive prefix sum of n uint64 values (in -> out, may alias); *total_dev receives the sum
cudaError_t scan_exclusive_u64(const uint64_t* in, uint64_t* out, int64_t n, uint64_t* total_dev,
                               Scratch* tmp, cudaStream_t stream, int64_t* launches);

struct VoxelExtract {  // state kept between _count and _fill
  Scratch bits, segbase, tmp, stage;
  int64_t n_voxels = -1;
  int inside_empty = 0;
};
cudaError_t voxel_extract_count(const GridDesc& g, int inside_empty, bool un_implied,
                                VoxelExtract* st, cudaStream_t stream, int64_t* launches);
cudaError_t voxel_extract_fill(const GridDesc& g, VoxelExtract* st, float* vertices_host,
                               int32_t* faces_host, cudaStream_t stream, int64_t* launches);
cudaError_t voxel_extract_flags(const GridDesc& g, VoxelExtract* st, uint8_t* flags_host,
                                cudaStream_t stream, int64_t* launches);

// ---- ASCII PLY text of a device-resident mesh (ply.cu) ----
struct PlyText {
  Scratch lens, tmp, text;  // per-line offsets, scan scratch, the characters (vertex lines, then faces)
  uint64_t vertex_bytes = 0, face_bytes = 0;
  int64_t nv = 0, nf = 0;
  bool sized = false;
};
cudaError_t ply_measure(const float* verts_dev, int64_t nv, const int32_t* faces_dev, int64_t nf,
                        PlyText* st, cudaStream_t stream, int64_t* launches);
cudaError_t ply_emit(const float* verts_dev, const int32_t* faces_dev, PlyText* st,
                     cudaStream_t stream, int64_t* launches);

struct IsoExtract {
  Scratch ibits, vbits, cbits, abits, segcnt, segbase, active, tmp, hkeys, hvals, verts, faces,
      counters, cinfo, segwid, vb, slot_cube, slot_pre, cube_slot;
  int64_t n_vertices = -1, n_faces = -1;
  int phase = 0;          // 0 nothing, 1 counted, 2 vertices emitted, 3 faces emitted
  unsigned n_active = 0;  // active segments (slots)
  unsigned n_cubes = 0;   // active cubes
  uint64_t hmask = 0;
  double iso_level = 0.0;
  int linear_interp = 1;
};
cudaError_t iso_count(const GridDesc& g, bool un_implied, double iso_level, int linear_interp,
                      IsoExtract* st, cudaStream_t stream, int64_t* launches);
cudaError_t iso_emit_vertices(const GridDesc& g, int64_t vertex_offset, IsoExtract* st,
                              cudaStream_t stream, int64_t* launches);
cudaError_t iso_export_top_ids(const GridDesc& g, IsoExtract* st, int32_t* out_dev,
                               cudaStream_t stream, int64_t* launches);
cudaError_t iso_import_bottom_ids(const GridDesc& g, IsoExtract* st, const int32_t* in_dev,
                                  cudaStream_t stream, int64_t* launches);
cudaError_t iso_emit_faces(const GridDesc& g, IsoExtract* st, cudaStream_t stream,
                           int64_t* launches);

// text returned by vcb_last_error(NULL) on this thread (for entry points outside carver.cu)
void set_global_error(const char* text);

void init_mc_tables();  // uploads the constant-memory tables once per device

}  // namespace vcb
