// Internals of the carver handle shared by carver.cu (the C ABI of one carver) and sharded.cu (the
// z-slab engine that drives several carvers): the handle's layout and the few helpers both need.
#pragma once

#include "common.cuh"

namespace vcb {
struct TimedSpan {
  cudaEvent_t a, b;
  int stage;
};

}  // namespace vcb

constexpr int kMaxStageSets = vcb::kMaxViewsPerLaunch / vcb::kSdfSubBatch + 2;  // one launch ahead + 2
constexpr size_t kStageBudgetBytes = (size_t)512 << 20;                // never more than this

struct vcb_carver {
  vcb_options opt;
  int nx = 0, ny = 0, nz = 0;
  int z_begin = 0, z_end = 0;
  int halo = 0;
  cudaStream_t stream = nullptr;
  std::string err;

  float* sdf_alloc = nullptr;  // halo planes first
  float* sdf = nullptr;        // plane z_begin
  int32_t* un = nullptr;
  float *pos_x = nullptr, *pos_y = nullptr, *pos_z = nullptr;
  std::vector<float> h_pos[3];
  float2* wtab = nullptr;
  int wtab_n = 0;
  int64_t max_count = 0;  // upper bound of update_num over the grid (views fused / uploaded)
  bool un_implied = true;  // update_num < 1  =>  sdf == kVal (true unless a grid was uploaded)
  // a caller's own SDF images were fused without truncation: a touched voxel may then hold kVal
  // itself (voxel_carver.cc:482-486 assigns whatever was sampled), so ExtractVoxel reads update_num
  bool kval_touched_possible = false;

  // queued views
  vcb::FuseParams fp;
  int pending = 0;
  vcb::Scratch pool;          // vcb::kMaxViewsPerLaunch image slots
  size_t slot_floats = 0;
  vcb::Scratch max_sdf;       // one float per slot
  vcb::Scratch fuse_tabs;     // per-launch tables of the fusion kernel (fusion_table_bytes)
  vcb::Scratch cls_counts;    // 5 device counters: (tile, view) pairs per tile class (fusion.cu)
  uint64_t unclassified_pairs = 0;  // pairs of launches that ran without tile classification
  vcb::PlyText ply;           // ASCII PLY text of the last extracted mesh (vcb_ply_body_*)
  vcb::Scratch tiles;         // 8x8-block flags per slot (tile_flag_bytes), see fusion.cu
  size_t tile_slot_bytes = 0;
  bool use_tiles = true;
  // 2D SDF scratch
  vcb::Scratch mask_stage, sdf_scratch, absmax;
  // host masks are uploaded on a separate stream into a 2-deep ring of staging sets, so that the
  // copy of one sub-batch overlaps the kernels of the previous ones
  cudaStream_t copy_stream = nullptr;
  // host masks travel through a ring of staging sets (vcb::kSdfSubBatch masks each) on their own copy
  // stream; with enough sets the copies of the next 64 views run under the current fusion launch
  cudaEvent_t ev_copied[kMaxStageSets] = {}, ev_consumed[kMaxStageSets] = {};
  bool consumed_valid[kMaxStageSets] = {};
  int stage_sets = 2;
  unsigned stage_round = 0;
  size_t sdf_pix = 0;      // mask staging bytes per view
  size_t sdf_scr = 0;      // 2D-SDF scratch bytes per view

  vcb::VoxelExtract vox;
  vcb::IsoExtract iso;
  vcb::PinnedBuffer ingest_pin;  // decoded mask files on their way to the GPU (vcb_carve_mask_files)

  bool timing = true;
  std::vector<vcb::TimedSpan> spans;
  float stage_ms[4] = {0, 0, 0, 0};
  int64_t stage_launches[4] = {0, 0, 0, 0};
};


namespace vcb {

// records the error text on the carver (and for vcb_last_error(NULL) on this thread); returns code
int fail(vcb_carver* c, int code, const char* fmt, ...);

#define CU(c, call)                                                                          \
  do {                                                                                       \
    cudaError_t e_ = (call);                                                                 \
    if (e_ != cudaSuccess)                                                                   \
      return vcb::fail(c, VCB_ERR_CUDA, "%s: %s (%s:%d)", #call, cudaGetErrorString(e_),      \
                       __FILE__, __LINE__);                                                  \
  } while (0)

// makes `device` current and uploads the constant tables on first use
int use_device(vcb_carver* c, int device);
int validate(const vcb_options* o, int dims[3]);  // voxel_carver.cc:375-389 and :276-301
int check_view(vcb_carver* c, const vcb_view& v);
int flush_pending(vcb_carver* c);                 // launches the fusion of the queued views
// mask files -> pixels on host threads (ingest.cu); wanted == nullptr decodes all n
int decode_mask_batch(const char* const* paths, const vcb_view* views, int n, const bool* wanted,
                      uint8_t* out, size_t stride, int n_threads, std::string* err);

// Fuses n <= kMaxViewsPerLaunch views whose float SDF images, 8x8-block flags (may be NULL) and
// image maxima are already on this carver's device, in one launch on the carver's stream (earlier
// queued views are flushed first).  Used by the z-slab engine, which prepares the images itself.
int fuse_prepared(vcb_carver* c, int n, const vcb_view* views, const float* const* images,
                  const uint32_t* const* tiles, const float* const* max_sdf);

// cost per z tile layer (fusion_tile_planes() planes each) of the whole grid for prepared views
int layer_costs(vcb_carver* c, int n, const vcb_view* views, const float* const* images,
                const uint32_t* const* tiles, const float* const* max_sdf,
                std::vector<uint64_t>* cost);

// CUDA-event span around kernels of one stage (0 2D SDF, 1 fusion, 2 ExtractVoxel, 3 marching
// cubes) on `stream` (default: the carver's); summed by vcb_stage_times
struct Span {
  vcb_carver* c;
  int stage;
  cudaStream_t stream;
  cudaEvent_t a = nullptr, b = nullptr;
  Span(vcb_carver* c_, int stage_) : Span(c_, stage_, c_->stream) {}
  Span(vcb_carver* c_, int stage_, cudaStream_t s_) : c(c_), stage(stage_), stream(s_) {
    if (c->timing) {
      cudaEventCreate(&a);
      cudaEventCreate(&b);
      cudaEventRecord(a, stream);
    }
  }
  ~Span() {
    if (c->timing) {
      cudaEventRecord(b, stream);
      c->spans.push_back({a, b, stage});
    }
  }
};

}  // namespace vcb
