// The carver handle and the C ABI (include/vacancy_b200.h): device state, view queue, stream
// plumbing.  Mirrors vacancy::VoxelCarver (voxel_carver.h:95-118, voxel_carver.cc:367-543) for the
// carving hot path; the kernels are in sdf2d.cu, fusion.cu, extract_voxel.cu, marching_cubes.cu.
//
// Device layout (all SoA, x fastest):
//   sdf[]        float  per owned voxel (+ up to 2 halo planes below the slab for marching cubes)
//   update_num[] int32  per owned voxel
//   pos_x/y/z[]  float  voxel-centre tables of the GLOBAL grid (voxel_carver.cc:315-326)
//   image pool   kMaxViewsPerLaunch float SDF images waiting to be fused
//   SDF scratch  kSdfSubBatch x (mask copy + 2 int16 images)
#include <float.h>
#include <math.h>
#include <stdarg.h>
#include <stdio.h>
#include <string.h>

#include <mutex>

#include "carver_internal.cuh"

using namespace vcb;

namespace {

thread_local std::string g_last_error = "";

// VoxelGrid::Init's state: sdf = kVal, update_num = 0 (voxel_carver.cc:342, voxel_carver.h:67-68)
__global__ void __launch_bounds__(256) grid_reset_kernel(float4* __restrict__ sdf4,
                                                         int4* __restrict__ un4, size_t n4_sdf,
                                                         size_t n4_un, float* __restrict__ sdf_tail,
                                                         int* __restrict__ un_tail, int tail_sdf,
                                                         int tail_un) {
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  const size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  const float4 k4 = make_float4(kInvalidSdf, kInvalidSdf, kInvalidSdf, kInvalidSdf);
  for (size_t i = t; i < n4_sdf; i += stride) sdf4[i] = k4;
  for (size_t i = t; i < n4_un; i += stride) un4[i] = make_int4(0, 0, 0, 0);
  if (t < (size_t)tail_sdf) sdf_tail[t] = kInvalidSdf;
  if (t < (size_t)tail_un) un_tail[t] = 0;
}

}  // namespace

namespace vcb {  // helpers shared with sharded.cu (declared in carver_internal.cuh)

int fail(vcb_carver* c, int code, const char* fmt, ...) {
  char buf[512];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof(buf), fmt, ap);
  va_end(ap);
  if (c) {
    c->err = buf;
  }
  g_last_error = buf;
  return code;
}

std::mutex g_init_mutex;
bool g_tables_ready[64] = {false};

int use_device(vcb_carver* c, int device) {
  CU(c, cudaSetDevice(device));
  std::lock_guard<std::mutex> lock(g_init_mutex);
  if (device >= 0 && device < 64 && !g_tables_ready[device]) {
    init_mc_tables();
    CU(c, cudaGetLastError());
    g_tables_ready[device] = true;
  }
  return VCB_OK;
}

void resolve_spans(vcb_carver* c) {
  for (auto& s : c->spans) {
    float ms = 0;
    cudaEventSynchronize(s.b);
    if (cudaEventElapsedTime(&ms, s.a, s.b) == cudaSuccess) c->stage_ms[s.stage] += ms;
    cudaEventDestroy(s.a);
    cudaEventDestroy(s.b);
  }
  c->spans.clear();
}

// voxel_carver.cc:375-389 and :276-301
int validate(const vcb_options* o, int dims[3]) {
  if (o->voxel_max_update_num < 1) return fail(nullptr, VCB_ERR_INVALID, "voxel_max_update_num must be positive");
  if (!(o->voxel_update_weight >= FLT_MIN)) return fail(nullptr, VCB_ERR_INVALID, "voxel_update_weight must be positive");
  if (!(o->truncation_band >= FLT_MIN)) return fail(nullptr, VCB_ERR_INVALID, "truncation_band must be positive");
  if (!(o->resolution >= FLT_MIN)) return fail(nullptr, VCB_ERR_INVALID, "resolution must be positive %f", o->resolution);
  if (o->bb_max[0] <= o->bb_min[0] || o->bb_max[1] <= o->bb_min[1] || o->bb_max[2] <= o->bb_min[2])
    return fail(nullptr, VCB_ERR_INVALID, "input bounding box is invalid");
  if (o->voxel_update != VCB_UPDATE_MAX && o->voxel_update != VCB_UPDATE_WEIGHTED_AVERAGE)
    return fail(nullptr, VCB_ERR_INVALID, "unknown voxel_update %d", o->voxel_update);
  if (o->sdf_interp != VCB_INTERP_NN && o->sdf_interp != VCB_INTERP_BILINEAR)
    return fail(nullptr, VCB_ERR_INVALID, "unknown sdf_interp %d", o->sdf_interp);
  if (o->update_outside != VCB_OUTSIDE_NONE && o->update_outside != VCB_OUTSIDE_MAX)
    return fail(nullptr, VCB_ERR_INVALID, "unknown update_outside %d", o->update_outside);
  if (o->sdf_metric != VCB_METRIC_L1 && o->sdf_metric != VCB_METRIC_L2)
    return fail(nullptr, VCB_ERR_INVALID, "unknown sdf_metric %d", o->sdf_metric);
  for (int i = 0; i < 3; i++) {
    const float diff = o->bb_max[i] - o->bb_min[i];
    const float q = diff / o->resolution;  // voxel_carver.cc:294-296 (truncating cast)
    if (!(q < 2147483648.0f)) return fail(nullptr, VCB_ERR_INVALID, "too many voxels");
    dims[i] = (int)q;
    if (dims[i] < 1) return fail(nullptr, VCB_ERR_INVALID, "bounding box smaller than one voxel");
  }
  return VCB_OK;
}

int check_view(vcb_carver* c, const vcb_view& v) {
  if (v.width < 1 || v.height < 1 || v.width > kMaxImageDim || v.height > kMaxImageDim)
    return fail(c, VCB_ERR_INVALID, "image size %dx%d out of range (max %d)", v.width, v.height,
                kMaxImageDim);
  if (v.roi_min[0] < 0 || v.roi_min[1] < 0 || v.roi_max[0] >= v.width || v.roi_max[1] >= v.height ||
      v.roi_min[0] > v.roi_max[0] || v.roi_min[1] > v.roi_max[1])
    return fail(c, VCB_ERR_INVALID, "roi [%d,%d]-[%d,%d] outside the %dx%d image", v.roi_min[0],
                v.roi_min[1], v.roi_max[0], v.roi_max[1], v.width, v.height);
  if (v.model != VCB_CAMERA_PINHOLE && v.model != VCB_CAMERA_ORTHO)
    return fail(c, VCB_ERR_UNSUPPORTED, "camera model %d: only pinhole and ortho are supported", v.model);
  return VCB_OK;
}

void fill_fuse_view(FuseView* f, const vcb_view& v, const float* image, const float* max_sdf) {
  memcpy(f->m, v.w2c, sizeof(f->m));
  f->fx = v.fx;
  f->fy = v.fy;
  f->cx = v.cx;
  f->cy = v.cy;
  f->roi_min_x = (float)v.roi_min[0];
  f->roi_min_y = (float)v.roi_min[1];
  f->roi_max_x = (float)v.roi_max[0];
  f->roi_max_y = (float)v.roi_max[1];
  f->roi_min_xi = v.roi_min[0];
  f->roi_min_yi = v.roi_min[1];
  f->roi_max_xi = v.roi_max[0];
  f->roi_max_yi = v.roi_max[1];
  f->pitch = v.width;
  f->model = v.model;
  f->image = image;
  f->max_sdf = max_sdf;
  f->tiles = nullptr;
  f->tile_words = f->tile_rows = 0;
}

// flag storage for every slot of the view queue, sized for the largest image seen so far
int ensure_tiles(vcb_carver* c, int width, int height) {
  const size_t need = (tile_flag_bytes(width, height) + 255) & ~(size_t)255;
  if (need <= c->tile_slot_bytes && c->tiles.ptr) return VCB_OK;
  int rc = flush_pending(c);
  if (rc != VCB_OK) return rc;
  CU(c, cudaStreamSynchronize(c->stream));
  CU(c, c->tiles.reserve(need * kMaxViewsPerLaunch));
  c->tile_slot_bytes = need;
  return VCB_OK;
}

// flag storage of one queue slot (NULL when tile classification is off)
uint32_t* tile_slot(vcb_carver* c, int slot) {
  if (!c->use_tiles || c->opt.sdf_interp != VCB_INTERP_BILINEAR || !c->tiles.ptr) return nullptr;
  return reinterpret_cast<uint32_t*>(static_cast<unsigned char*>(c->tiles.ptr) +
                                     (size_t)slot * c->tile_slot_bytes);
}

// queues the block-flag kernel for `n` freshly written images and points their views at the flags
void attach_tiles(vcb_carver* c, const int* slots, const float* const* images, const vcb_view* views,
                  int n) {
  if (!c->use_tiles || c->opt.sdf_interp != VCB_INTERP_BILINEAR) return;
  TileBatch b;
  for (int i = 0; i < n; i++) {
    uint32_t* flags = reinterpret_cast<uint32_t*>(static_cast<unsigned char*>(c->tiles.ptr) +
                                                  (size_t)slots[i] * c->tile_slot_bytes);
    b.image[i] = images[i];
    b.flags[i] = flags;
    b.width[i] = views[i].width;
    b.height[i] = views[i].height;
    FuseView& f = c->fp.views[slots[i]];
    f.tiles = flags;
    f.tile_words = tile_words_for(views[i].width);
    f.tile_rows = tile_rows_for(views[i].height);
  }
  Span span(c, 0);
  launch_tile_flags(b, n, c->stream, &c->stage_launches[0]);
}

GridDesc grid_desc(const vcb_carver* c) {
  GridDesc g;
  g.sdf = c->sdf;
  g.update_num = c->un;
  g.pos_x = c->pos_x;
  g.pos_y = c->pos_y;
  g.pos_z = c->pos_z;
  g.nx = c->nx;
  g.ny = c->ny;
  g.nz = c->nz;
  g.z_begin = c->z_begin;
  g.z_end = c->z_end;
  g.halo = c->halo;
  g.resolution = c->opt.resolution;
  return g;
}

int64_t slab_voxels(const vcb_carver* c) {
  return (int64_t)c->nx * c->ny * (c->z_end - c->z_begin);
}

void launch_grid_reset(vcb_carver* c) {
  const size_t plane = (size_t)c->nx * c->ny;
  const size_t n_sdf = plane * (size_t)(c->z_end - c->z_begin + c->halo);
  const size_t n_un = plane * (size_t)(c->z_end - c->z_begin);
  // cudaMalloc pointers are 256-byte aligned, so the float4 / int4 views start aligned
  grid_reset_kernel<<<148 * 8, 256, 0, c->stream>>>(
      reinterpret_cast<float4*>(c->sdf_alloc), reinterpret_cast<int4*>(c->un), n_sdf / 4, n_un / 4,
      c->sdf_alloc + (n_sdf / 4) * 4, c->un + (n_un / 4) * 4, (int)(n_sdf % 4), (int)(n_un % 4));
}

// weighted-average factors by update count, voxel_carver.cc:91-93: {w*float(n), 1/(w*float(n+1))}.
// The fusion kernel indexes the table with update_num, so it must reach every count a voxel that
// still takes updates can hold: min(voxel_max_update_num, views applied so far) + 1.
int ensure_wtab(vcb_carver* c, int64_t needed) {
  if (needed <= c->wtab_n) return VCB_OK;
  int64_t n = 512;
  while (n < needed) n *= 2;
  std::vector<float2> t((size_t)n);
  const float w = c->opt.voxel_update_weight;
  for (int64_t k = 0; k < n; k++) {
    volatile float wn = w * (float)k;
    volatile float den = w * (float)(k + 1);
    volatile float inv = 1.0f / den;
    t[(size_t)k].x = wn;
    t[(size_t)k].y = inv;
  }
  CU(c, cudaStreamSynchronize(c->stream));
  if (c->wtab) CU(c, cudaFree(c->wtab));
  c->wtab = nullptr;
  c->wtab_n = 0;
  CU(c, cudaMalloc(&c->wtab, (size_t)n * sizeof(float2)));
  CU(c, cudaMemcpy(c->wtab, t.data(), (size_t)n * sizeof(float2), cudaMemcpyHostToDevice));
  c->wtab_n = (int)n;
  c->fp.wtab = c->wtab;
  c->fp.wtab_n = c->wtab_n;
  return VCB_OK;
}

// launches the fusion kernel over the queued views
int flush_pending(vcb_carver* c) {
  if (c->pending == 0) return VCB_OK;
  c->max_count += c->pending;
  {
    const int64_t cap = (int64_t)c->opt.voxel_max_update_num + 1;
    const int rc = ensure_wtab(c, (c->max_count < cap ? c->max_count : cap) + 2);
    if (rc != VCB_OK) return rc;
  }
  {
    const int64_t cap = (int64_t)c->opt.voxel_max_update_num + 1;
    const int64_t fill = (c->max_count < cap ? c->max_count : cap) + 2;
    c->fp.wtab_fill = (int)(fill < c->wtab_n ? fill : c->wtab_n);
    CU(c, c->fuse_tabs.reserve(fusion_table_bytes(c->nx, c->ny, c->z_end - c->z_begin)));
    c->fp.tables = static_cast<unsigned char*>(c->fuse_tabs.ptr);
  }
  c->fp.n_views = c->pending;
  // the gate tests update_num BEFORE a view is applied: the largest count it can meet in this
  // launch is (counts so far) + (views of the launch) - 1 = max_count - 1
  c->fp.cap_possible = c->max_count - 1 > (int64_t)c->opt.voxel_max_update_num ? 1 : 0;
  {
    Span span(c, 1);
    launch_fusion(c->fp, c->opt.sdf_interp, c->opt.voxel_update, c->stream, &c->stage_launches[1]);
  }
  if (c->fp.tile_class == nullptr) {  // no classification: every (tile, view) pair took the general path
    const uint64_t tiles = (uint64_t)((c->nx + 31) / 32) * ((c->ny + 7) / 8) *
                           ((c->z_end - c->z_begin + 7) / 8);
    c->unclassified_pairs += tiles * (uint64_t)c->pending;
  }
  c->pending = 0;
  CU(c, cudaGetLastError());
  return VCB_OK;
}

int fuse_prepared(vcb_carver* c, int n, const vcb_view* views, const float* const* images,
                  const uint32_t* const* tiles, const float* const* max_sdf) {
  if (n < 0 || n > kMaxViewsPerLaunch) return fail(c, VCB_ERR_INVALID, "fuse_prepared: %d views", n);
  int rc = flush_pending(c);
  if (rc != VCB_OK || n == 0) return rc;
  const bool flags = c->use_tiles && c->opt.sdf_interp == VCB_INTERP_BILINEAR;
  for (int i = 0; i < n; i++) {
    FuseView& f = c->fp.views[i];
    fill_fuse_view(&f, views[i], images[i], max_sdf[i]);
    if (flags && tiles && tiles[i]) {
      f.tiles = tiles[i];
      f.tile_words = tile_words_for(views[i].width);
      f.tile_rows = tile_rows_for(views[i].height);
    }
  }
  c->pending = n;
  return flush_pending(c);
}

// Cost estimate per z tile layer of the WHOLE grid for the given prepared views (on this carver's
// device and stream; nothing is fused).  Used by the z-slab engine to place the slab boundaries.
int layer_costs(vcb_carver* c, int n, const vcb_view* views, const float* const* images,
                const uint32_t* const* tiles, const float* const* max_sdf,
                std::vector<uint64_t>* cost) {
  if (n < 1 || n > kMaxViewsPerLaunch) return fail(c, VCB_ERR_INVALID, "layer_costs: %d views", n);
  int rc = use_device(c, c->opt.device);
  if (rc != VCB_OK) return rc;
  FuseParams p = c->fp;
  p.z_begin = 0;
  p.z_end = c->nz;
  p.n_views = n;
  const bool flags = c->use_tiles && c->opt.sdf_interp == VCB_INTERP_BILINEAR;
  for (int i = 0; i < n; i++) {
    FuseView& f = p.views[i];
    fill_fuse_view(&f, views[i], images[i], max_sdf[i]);
    if (flags && tiles && tiles[i]) {
      f.tiles = tiles[i];
      f.tile_words = tile_words_for(views[i].width);
      f.tile_rows = tile_rows_for(views[i].height);
    }
  }
  const int layers = (c->nz + fusion_tile_planes() - 1) / fusion_tile_planes();
  Scratch tabs, dev_cost;
  CU(c, tabs.reserve(fusion_table_bytes(c->nx, c->ny, c->nz)));
  cudaError_t e = dev_cost.reserve((size_t)layers * sizeof(unsigned long long));
  if (e == cudaSuccess) e = cudaMemsetAsync(dev_cost.ptr, 0, (size_t)layers * sizeof(unsigned long long), c->stream);
  if (e == cudaSuccess) {
    p.tables = static_cast<unsigned char*>(tabs.ptr);
    launch_layer_costs(p, c->opt.sdf_interp, static_cast<unsigned long long*>(dev_cost.ptr), c->stream);
    e = cudaGetLastError();
  }
  cost->assign(layers, 0);
  if (e == cudaSuccess)
    e = cudaMemcpyAsync(cost->data(), dev_cost.ptr, (size_t)layers * sizeof(uint64_t),
                        cudaMemcpyDeviceToHost, c->stream);
  if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
  tabs.release();
  dev_cost.release();
  if (e != cudaSuccess) return fail(c, VCB_ERR_CUDA, "layer_costs: %s", cudaGetErrorString(e));
  return VCB_OK;
}

int ensure_pool(vcb_carver* c, size_t floats_per_image) {
  if (floats_per_image <= c->slot_floats && c->pool.ptr) return VCB_OK;
  int rc = flush_pending(c);
  if (rc != VCB_OK) return rc;
  CU(c, cudaStreamSynchronize(c->stream));
  CU(c, c->pool.reserve(floats_per_image * sizeof(float) * kMaxViewsPerLaunch));
  c->slot_floats = floats_per_image;
  return VCB_OK;
}

int ensure_sdf_scratch(vcb_carver* c, size_t pix, size_t scr) {
  if (pix <= c->sdf_pix && scr <= c->sdf_scr && c->sdf_scratch.ptr) return VCB_OK;
  CU(c, cudaStreamSynchronize(c->stream));
  pix = pix > c->sdf_pix ? pix : c->sdf_pix;
  scr = scr > c->sdf_scr ? scr : c->sdf_scr;
  pix = (pix + 255) & ~(size_t)255;
  if (c->copy_stream) CU(c, cudaStreamSynchronize(c->copy_stream));
  {
    size_t sets = kStageBudgetBytes / (pix * kSdfSubBatch);
    sets = sets < 2 ? 2 : (sets > (size_t)kMaxStageSets ? (size_t)kMaxStageSets : sets);
    c->stage_sets = (int)sets;
  }
  CU(c, c->mask_stage.reserve(pix * kSdfSubBatch * c->stage_sets));
  CU(c, c->sdf_scratch.reserve(scr * kSdfSubBatch));
  CU(c, c->absmax.reserve(sizeof(int) * kSdfSubBatch));
  for (int s2 = 0; s2 < kMaxStageSets; s2++) c->consumed_valid[s2] = false;
  c->stage_round = 0;
  c->sdf_pix = pix;
  c->sdf_scr = scr;
  return VCB_OK;
}

// shared body of vcb_carve_masks / vcb_carve_masks_device
int carve_masks_impl(vcb_carver* c, int32_t n_views, const vcb_view* views,
                     const uint8_t* const* masks, float* const* sdf_out, bool device_io) {
  if (!c) return fail(nullptr, VCB_ERR_INVALID, "null carver");
  if (n_views < 0 || (n_views > 0 && (!views || !masks))) return fail(c, VCB_ERR_INVALID, "null argument");
  int rc = use_device(c, c->opt.device);
  if (rc != VCB_OK) return rc;
  size_t max_pix = 0, max_scr = 0;
  for (int i = 0; i < n_views; i++) {
    if ((rc = check_view(c, views[i])) != VCB_OK) return rc;
    if (!masks[i]) return fail(c, VCB_ERR_INVALID, "mask %d is null", i);
    max_pix = max(max_pix, (size_t)views[i].width * views[i].height);
    max_scr = max(max_scr, sdf_scratch_bytes(views[i].width, views[i].height,
                                             c->opt.sdf_metric == VCB_METRIC_L2 ? kSdfSignedL2
                                                                                : kSdfSigned));
  }
  if (n_views == 0) return VCB_OK;
  if ((rc = ensure_pool(c, max_pix)) != VCB_OK) return rc;
  if ((rc = ensure_sdf_scratch(c, max_pix, max_scr)) != VCB_OK) return rc;
  {
    int mw = 0, mh = 0;
    for (int i = 0; i < n_views; i++) {
      mw = max(mw, views[i].width);
      mh = max(mh, views[i].height);
    }
    if ((rc = ensure_tiles(c, mw, mh)) != VCB_OK) return rc;
  }
  float* pool = static_cast<float*>(c->pool.ptr);
  float* max_sdf = static_cast<float*>(c->max_sdf.ptr);
  if (!device_io && c->copy_stream == nullptr) {
    CU(c, cudaStreamCreateWithFlags(&c->copy_stream, cudaStreamNonBlocking));
    for (int s = 0; s < kMaxStageSets; s++) {
      CU(c, cudaEventCreateWithFlags(&c->ev_copied[s], cudaEventDisableTiming));
      CU(c, cudaEventCreateWithFlags(&c->ev_consumed[s], cudaEventDisableTiming));
    }
  }
  for (int first = 0; first < n_views;) {
    int nb = min(kSdfSubBatch, n_views - first);
    nb = min(nb, kMaxViewsPerLaunch - c->pending);
    SdfView hv[kSdfSubBatch];
    const int set = (int)(c->stage_round % (unsigned)c->stage_sets);
    if (!device_io && c->consumed_valid[set]) {
      // the staging set may only be overwritten once the kernels that read it have finished
      CU(c, cudaStreamWaitEvent(c->copy_stream, c->ev_consumed[set], 0));
    }
    for (int i = 0; i < nb; i++) {
      const vcb_view& v = views[first + i];
      const size_t pix = (size_t)v.width * v.height;
      const int slot = c->pending + i;
      if (device_io) {
        hv[i].mask = masks[first + i];
      } else {
        uint8_t* dst = static_cast<uint8_t*>(c->mask_stage.ptr) +
                       (size_t)(set * kSdfSubBatch + i) * c->sdf_pix;
        CU(c, cudaMemcpyAsync(dst, masks[first + i], pix, cudaMemcpyHostToDevice, c->copy_stream));
        hv[i].mask = dst;
      }
      hv[i].out = pool + (size_t)slot * c->slot_floats;
      hv[i].scratch = static_cast<unsigned char*>(c->sdf_scratch.ptr) + (size_t)i * c->sdf_scr;
      hv[i].absmax = static_cast<int*>(c->absmax.ptr) + i;
      hv[i].max_out = max_sdf + slot;
      hv[i].tile_flags = tile_slot(c, slot);  // written by the finalize pass
      hv[i].width = v.width;
      hv[i].height = v.height;
      hv[i].rx0 = v.roi_min[0];
      hv[i].ry0 = v.roi_min[1];
      hv[i].rx1 = v.roi_max[0];
      hv[i].ry1 = v.roi_max[1];
      fill_fuse_view(&c->fp.views[slot], v, hv[i].out, hv[i].max_out);
    }
    if (!device_io) {
      CU(c, cudaEventRecord(c->ev_copied[set], c->copy_stream));
      CU(c, cudaStreamWaitEvent(c->stream, c->ev_copied[set], 0));
    }
    {
      Span span(c, 0);
      launch_sdf2d(hv, nb, c->opt.sdf_metric == VCB_METRIC_L2 ? kSdfSignedL2 : kSdfSigned,
                   c->opt.sdf_minmax_normalize, c->opt.use_truncation,
                   c->opt.truncation_band, c->stream, &c->stage_launches[0]);
    }
    CU(c, cudaGetLastError());
    for (int i = 0; i < nb; i++) {
      FuseView& f = c->fp.views[c->pending + i];
      f.tiles = hv[i].tile_flags;
      f.tile_words = tile_words_for(hv[i].width);
      f.tile_rows = tile_rows_for(hv[i].height);
    }
    if (!device_io) {
      CU(c, cudaEventRecord(c->ev_consumed[set], c->stream));
      c->consumed_valid[set] = true;
      c->stage_round++;
    }
    if (sdf_out) {
      for (int i = 0; i < nb; i++) {
        if (!sdf_out[first + i]) continue;
        const size_t bytes = (size_t)hv[i].width * hv[i].height * sizeof(float);
        CU(c, cudaMemcpyAsync(sdf_out[first + i], hv[i].out, bytes,
                              device_io ? cudaMemcpyDeviceToDevice : cudaMemcpyDeviceToHost,
                              c->stream));
      }
    }
    c->pending += nb;
    first += nb;
    if (c->pending == kMaxViewsPerLaunch) {
      if ((rc = flush_pending(c)) != VCB_OK) return rc;
    }
  }
  if (!device_io) {
    // host buffers are borrowed for the duration of the call only (voxel_carver.cc:394-413): the
    // masks have all been read once the copy stream is idle; the kernels may still be running
    CU(c, cudaStreamSynchronize(c->copy_stream));
    if (sdf_out) CU(c, cudaStreamSynchronize(c->stream));
  }
  return VCB_OK;
}

void set_global_error(const char* text) { g_last_error = text; }
}  // namespace vcb

extern "C" {

const char* vcb_last_error(const vcb_carver* c) { return c ? c->err.c_str() : g_last_error.c_str(); }

int vcb_abi_version(void) { return VCB_ABI_VERSION; }

float vcb_invalid_sdf(void) { return kInvalidSdf; }

void vcb_options_default(vcb_options* o) {
  memset(o, 0, sizeof(*o));
  o->resolution = 0.1f;
  o->sdf_minmax_normalize = 1;
  o->voxel_update = VCB_UPDATE_MAX;
  o->sdf_interp = VCB_INTERP_BILINEAR;
  o->update_outside = VCB_OUTSIDE_NONE;
  o->voxel_max_update_num = 255;
  o->voxel_update_weight = 1.0f;
  o->use_truncation = 0;
  o->truncation_band = 0.1f;
}

int vcb_create(const vcb_options* opt, vcb_carver** out) {
  if (!opt || !out) return fail(nullptr, VCB_ERR_INVALID, "null argument");
  *out = nullptr;
  int dims[3];
  int rc = validate(opt, dims);
  if (rc != VCB_OK) return rc;
  int z0 = opt->z_begin, z1 = opt->z_end;
  if (z0 == 0 && z1 == 0) z1 = dims[2];
  if (z0 < 0 || z1 > dims[2] || z0 >= z1)
    return fail(nullptr, VCB_ERR_INVALID, "z slab [%d,%d) outside the grid of %d planes", z0, z1, dims[2]);
  int n_dev = 0;
  cudaError_t e = cudaGetDeviceCount(&n_dev);
  if (e != cudaSuccess || n_dev < 1)
    return fail(nullptr, VCB_ERR_CUDA, "no CUDA device: %s (this library has no CPU fallback)",
                cudaGetErrorString(e));
  if (opt->device < 0 || opt->device >= n_dev)
    return fail(nullptr, VCB_ERR_INVALID, "device %d not in [0,%d)", opt->device, n_dev);

  vcb_carver* c = new vcb_carver;
  c->opt = *opt;
  c->nx = dims[0];
  c->ny = dims[1];
  c->nz = dims[2];
  c->z_begin = z0;
  c->z_end = z1;
  c->halo = z0 < 2 ? z0 : 2;
  {
    const char* e = getenv("VACANCY_B200_NO_TILES");  // debugging switch: no tile classification
    c->use_tiles = !(e && e[0] == '1');
  }
#define CREATE_TRY(call)                                                                     \
  do {                                                                                       \
    cudaError_t e_ = (call);                                                                 \
    if (e_ != cudaSuccess) {                                                                 \
      fail(nullptr, VCB_ERR_CUDA, "%s: %s", #call, cudaGetErrorString(e_));                   \
      vcb_destroy(c);                                                                        \
      return VCB_ERR_CUDA;                                                                   \
    }                                                                                        \
  } while (0)
  if (use_device(c, opt->device) != VCB_OK) {
    g_last_error = c->err;
    delete c;
    return VCB_ERR_CUDA;
  }
  const size_t plane = (size_t)c->nx * c->ny;
  const size_t n = plane * (size_t)(z1 - z0);
  CREATE_TRY(cudaMalloc(&c->sdf_alloc, (n + plane * c->halo) * sizeof(float)));
  c->sdf = c->sdf_alloc + plane * c->halo;
  CREATE_TRY(cudaMalloc(&c->un, n * sizeof(int32_t)));
  // voxel centres: diff * (float(k) / float(n)) + bb_min + resolution * 0.5f, voxel_carver.cc:309-326
  const float offset = opt->resolution * 0.5f;
  float** dpos[3] = {&c->pos_x, &c->pos_y, &c->pos_z};
  for (int a = 0; a < 3; a++) {
    const float diff = opt->bb_max[a] - opt->bb_min[a];
    c->h_pos[a].resize(dims[a]);
    for (int k = 0; k < dims[a]; k++) {
      volatile float frac = (float)k / (float)dims[a];
      volatile float t = diff * frac;
      volatile float t2 = t + opt->bb_min[a];
      c->h_pos[a][k] = t2 + offset;
    }
    CREATE_TRY(cudaMalloc(dpos[a], dims[a] * sizeof(float)));
    CREATE_TRY(cudaMemcpy(*dpos[a], c->h_pos[a].data(), dims[a] * sizeof(float), cudaMemcpyHostToDevice));
  }
  CREATE_TRY(c->max_sdf.reserve(sizeof(float) * kMaxViewsPerLaunch));
  CREATE_TRY(c->cls_counts.reserve(8 * sizeof(unsigned long long)));
  CREATE_TRY(cudaMemset(c->cls_counts.ptr, 0, 8 * sizeof(unsigned long long)));
  // sdf = kVal, update_num = 0 (voxel_carver.cc:342, voxel_carver.h:67-68)
  launch_grid_reset(c);
  CREATE_TRY(cudaGetLastError());
  CREATE_TRY(cudaStreamSynchronize(c->stream));
  memset(&c->fp, 0, sizeof(c->fp));
  c->fp.sdf = c->sdf;
  c->fp.update_num = c->un;
  c->fp.pos_x = c->pos_x;
  c->fp.pos_y = c->pos_y;
  c->fp.pos_z = c->pos_z;
  c->fp.nx = c->nx;
  c->fp.ny = c->ny;
  c->fp.z_begin = z0;
  c->fp.z_end = z1;
  c->fp.max_update_num = opt->voxel_max_update_num;
  c->fp.use_truncation = opt->use_truncation;
  c->fp.update_outside = opt->update_outside;
  c->fp.weight = opt->voxel_update_weight;
  c->fp.class_counts = static_cast<unsigned long long*>(c->cls_counts.ptr);
#undef CREATE_TRY
  {
    const int64_t cap = (int64_t)opt->voxel_max_update_num + 2;
    if (ensure_wtab(c, cap < 257 ? cap : 257) != VCB_OK) {
      g_last_error = c->err;
      vcb_destroy(c);
      return VCB_ERR_CUDA;
    }
  }
  *out = c;
  return VCB_OK;
}

void vcb_destroy(vcb_carver* c) {
  if (!c) return;
  cudaSetDevice(c->opt.device);
  cudaStreamSynchronize(c->stream);
  resolve_spans(c);
  if (c->copy_stream) {
    cudaStreamSynchronize(c->copy_stream);
    cudaStreamDestroy(c->copy_stream);
    for (int s = 0; s < kMaxStageSets; s++) {
      cudaEventDestroy(c->ev_copied[s]);
      cudaEventDestroy(c->ev_consumed[s]);
    }
  }
  cudaFree(c->sdf_alloc);
  cudaFree(c->un);
  cudaFree(c->pos_x);
  cudaFree(c->pos_y);
  cudaFree(c->pos_z);
  cudaFree(c->wtab);
  Scratch* all[] = {&c->cls_counts, &c->pool, &c->max_sdf, &c->tiles, &c->fuse_tabs, &c->ply.lens, &c->ply.tmp, &c->ply.text, &c->mask_stage, &c->sdf_scratch, &c->absmax, &c->vox.bits, &c->vox.segbase, &c->vox.tmp, &c->vox.stage,
                    &c->iso.ibits, &c->iso.vbits, &c->iso.cbits, &c->iso.abits, &c->iso.segcnt,
                    &c->iso.segbase, &c->iso.active, &c->iso.tmp, &c->iso.hkeys, &c->iso.hvals,
                    &c->iso.verts, &c->iso.faces, &c->iso.counters, &c->iso.cinfo,
                    &c->iso.segwid, &c->iso.vb, &c->iso.slot_cube, &c->iso.slot_pre,
                    &c->iso.cube_slot};
  for (Scratch* s : all) s->release();
  c->ingest_pin.release();
  delete c;
}

int vcb_reset(vcb_carver* c) {
  if (!c) return fail(nullptr, VCB_ERR_INVALID, "null carver");
  int rc = use_device(c, c->opt.device);
  if (rc != VCB_OK) return rc;
  c->pending = 0;  // views queued but not yet fused belong to the grid being discarded
  c->max_count = 0;
  launch_grid_reset(c);
  c->stage_launches[1] += 1;
  c->un_implied = true;
  c->kval_touched_possible = false;
  c->iso.n_vertices = c->iso.n_faces = -1;
  c->vox.n_voxels = -1;
  CU(c, cudaGetLastError());
  return VCB_OK;
}

int vcb_grid_dims(const vcb_carver* c, int32_t dims3[3]) {
  if (!c || !dims3) return fail(nullptr, VCB_ERR_INVALID, "null argument");
  dims3[0] = c->nx;
  dims3[1] = c->ny;
  dims3[2] = c->nz;
  return VCB_OK;
}

int vcb_slab_range(const vcb_carver* c, int32_t* z_begin, int32_t* z_end) {
  if (!c) return fail(nullptr, VCB_ERR_INVALID, "null argument");
  if (z_begin) *z_begin = c->z_begin;
  if (z_end) *z_end = c->z_end;
  return VCB_OK;
}

int vcb_grid_positions(const vcb_carver* c, float* px, float* py, float* pz) {
  if (!c) return fail(nullptr, VCB_ERR_INVALID, "null argument");
  if (px) memcpy(px, c->h_pos[0].data(), c->nx * sizeof(float));
  if (py) memcpy(py, c->h_pos[1].data(), c->ny * sizeof(float));
  if (pz) memcpy(pz, c->h_pos[2].data(), c->nz * sizeof(float));
  return VCB_OK;
}

int vcb_set_stream(vcb_carver* c, void* cuda_stream) {
  if (!c) return fail(nullptr, VCB_ERR_INVALID, "null carver");
  int rc = use_device(c, c->opt.device);
  if (rc != VCB_OK) return rc;
  if ((rc = flush_pending(c)) != VCB_OK) return rc;
  CU(c, cudaStreamSynchronize(c->stream));
  c->stream = static_cast<cudaStream_t>(cuda_stream);
  return VCB_OK;
}

int vcb_flush(vcb_carver* c) {
  if (!c) return fail(nullptr, VCB_ERR_INVALID, "null carver");
  int rc = use_device(c, c->opt.device);
  if (rc != VCB_OK) return rc;
  return flush_pending(c);
}

int vcb_synchronize(vcb_carver* c) {
  int rc = vcb_flush(c);
  if (rc != VCB_OK) return rc;
  CU(c, cudaStreamSynchronize(c->stream));
  return VCB_OK;
}

int vcb_carve_masks(vcb_carver* c, int32_t n_views, const vcb_view* views,
                    const uint8_t* const* masks, float* const* sdf_out) {
  return carve_masks_impl(c, n_views, views, masks, sdf_out, false);
}

int vcb_carve_masks_device(vcb_carver* c, int32_t n_views, const vcb_view* views,
                           const uint8_t* const* masks_dev, float* const* sdf_out_dev) {
  return carve_masks_impl(c, n_views, views, masks_dev, sdf_out_dev, true);
}

int vcb_carve_sdfs(vcb_carver* c, int32_t n_views, const vcb_view* views,
                   const float* const* sdfs) {
  if (!c) return fail(nullptr, VCB_ERR_INVALID, "null carver");
  if (n_views < 0 || (n_views > 0 && (!views || !sdfs))) return fail(c, VCB_ERR_INVALID, "null argument");
  int rc = use_device(c, c->opt.device);
  if (rc != VCB_OK) return rc;
  size_t max_pix = 0;
  for (int i = 0; i < n_views; i++) {
    if ((rc = check_view(c, views[i])) != VCB_OK) return rc;
    if (!sdfs[i]) return fail(c, VCB_ERR_INVALID, "sdf %d is null", i);
    max_pix = max(max_pix, (size_t)views[i].width * views[i].height);
  }
  if (n_views == 0) return VCB_OK;
  if (!c->opt.use_truncation) c->kval_touched_possible = true;
  if ((rc = ensure_pool(c, max_pix)) != VCB_OK) return rc;
  {
    int mw = 0, mh = 0;
    for (int i = 0; i < n_views; i++) {
      mw = max(mw, views[i].width);
      mh = max(mh, views[i].height);
    }
    if ((rc = ensure_tiles(c, mw, mh)) != VCB_OK) return rc;
  }
  float* pool = static_cast<float*>(c->pool.ptr);
  float* max_sdf = static_cast<float*>(c->max_sdf.ptr);
  for (int i = 0; i < n_views; i++) {
    const vcb_view& v = views[i];
    const size_t pix = (size_t)v.width * v.height;
    const int slot = c->pending;
    float* dst = pool + (size_t)slot * c->slot_floats;
    CU(c, cudaMemcpyAsync(dst, sdfs[i], pix * sizeof(float), cudaMemcpyHostToDevice, c->stream));
    if (c->opt.update_outside == VCB_OUTSIDE_MAX) {  // voxel_carver.cc:436
      Span span(c, 0);
      launch_image_max(dst, (int64_t)pix, max_sdf + slot, c->stream, &c->stage_launches[0]);
    }
    fill_fuse_view(&c->fp.views[slot], v, dst, max_sdf + slot);
    {
      const float* img = dst;
      attach_tiles(c, &slot, &img, &v, 1);
    }
    c->pending++;
    if (c->pending == kMaxViewsPerLaunch) {
      if ((rc = flush_pending(c)) != VCB_OK) return rc;
    }
  }
  CU(c, cudaStreamSynchronize(c->stream));
  return VCB_OK;
}

int vcb_carve_sdfs_device(vcb_carver* c, int32_t n_views, const vcb_view* views,
                          const float* const* sdfs_dev) {
  if (!c) return fail(nullptr, VCB_ERR_INVALID, "null carver");
  if (n_views < 0 || (n_views > 0 && (!views || !sdfs_dev))) return fail(c, VCB_ERR_INVALID, "null argument");
  int rc = use_device(c, c->opt.device);
  if (rc != VCB_OK) return rc;
  for (int i = 0; i < n_views; i++) {
    if ((rc = check_view(c, views[i])) != VCB_OK) return rc;
    if (!sdfs_dev[i]) return fail(c, VCB_ERR_INVALID, "sdf %d is null", i);
  }
  if ((rc = flush_pending(c)) != VCB_OK) return rc;
  if (n_views > 0 && !c->opt.use_truncation) c->kval_touched_possible = true;
  {
    int mw = 0, mh = 0;
    for (int i = 0; i < n_views; i++) {
      mw = max(mw, views[i].width);
      mh = max(mh, views[i].height);
    }
    if (n_views > 0 && (rc = ensure_tiles(c, mw, mh)) != VCB_OK) return rc;
  }
  float* max_sdf = static_cast<float*>(c->max_sdf.ptr);
  for (int i = 0; i < n_views; i++) {
    const vcb_view& v = views[i];
    const int slot = c->pending;
    if (c->opt.update_outside == VCB_OUTSIDE_MAX) {
      Span span(c, 0);
      launch_image_max(sdfs_dev[i], (int64_t)v.width * v.height, max_sdf + slot, c->stream,
                       &c->stage_launches[0]);
    }
    fill_fuse_view(&c->fp.views[slot], v, sdfs_dev[i], max_sdf + slot);
    attach_tiles(c, &slot, &sdfs_dev[i], &v, 1);
    c->pending++;
    if (c->pending == kMaxViewsPerLaunch || i == n_views - 1) {
      if ((rc = flush_pending(c)) != VCB_OK) return rc;
    }
  }
  return VCB_OK;
}

static int sdf_free_function(int32_t device, const uint8_t* mask, int32_t width, int32_t height,
                             const int32_t roi_min[2], const int32_t roi_max[2], int mode,
                             int normalize, int truncate, float band, float* dist) {
  if (!mask || !dist || !roi_min || !roi_max) return fail(nullptr, VCB_ERR_INVALID, "null argument");
  vcb_view v;
  memset(&v, 0, sizeof(v));
  v.width = width;
  v.height = height;
  v.roi_min[0] = roi_min[0];
  v.roi_min[1] = roi_min[1];
  v.roi_max[0] = roi_max[0];
  v.roi_max[1] = roi_max[1];
  int rc = check_view(nullptr, v);
  if (rc != VCB_OK) return rc;
  DeviceGuard restore_callers_device;
  if ((rc = use_device(nullptr, device)) != VCB_OK) return rc;
  const size_t pix = (size_t)width * height;
  Scratch buf;
  // layout: float out | kernel scratch | mask | SdfView | int absmax | float max
  const size_t off_scr = (pix * sizeof(float) + 255) & ~(size_t)255;
  const size_t off_mask = off_scr + sdf_scratch_bytes(width, height, mode);
  const size_t off_view = (off_mask + pix + 255) / 256 * 256;
  const size_t off_misc = off_view + 256;
  CU(nullptr, buf.reserve(off_misc + 256));
  char* base = static_cast<char*>(buf.ptr);
  SdfView hv;
  hv.mask = reinterpret_cast<uint8_t*>(base + off_mask);
  hv.out = reinterpret_cast<float*>(base);
  hv.scratch = reinterpret_cast<unsigned char*>(base + off_scr);
  hv.tile_flags = nullptr;
  hv.absmax = reinterpret_cast<int*>(base + off_misc);
  hv.max_out = reinterpret_cast<float*>(base + off_misc + 64);
  hv.width = width;
  hv.height = height;
  hv.rx0 = roi_min[0];
  hv.ry0 = roi_min[1];
  hv.rx1 = roi_max[0];
  hv.ry1 = roi_max[1];
  cudaError_t e = cudaMemcpy(base + off_mask, mask, pix, cudaMemcpyHostToDevice);
  if (e == cudaSuccess) {
    launch_sdf2d(&hv, 1, mode, normalize, truncate, band, nullptr, nullptr);
    e = cudaGetLastError();
  }
  if (e == cudaSuccess) e = cudaMemcpy(dist, base, pix * sizeof(float), cudaMemcpyDeviceToHost);
  buf.release();
  if (e != cudaSuccess) return fail(nullptr, VCB_ERR_CUDA, "2D SDF: %s", cudaGetErrorString(e));
  return VCB_OK;
}

int vcb_signed_distance_to_color(int32_t device, const float* sdf, int32_t width, int32_t height,
                                 float min_negative_d, float max_positive_d, uint8_t* rgb) {
  if (!sdf || !rgb) return fail(nullptr, VCB_ERR_INVALID, "null argument");
  if (width < 1 || height < 1) return fail(nullptr, VCB_ERR_INVALID, "empty image");
  if (!(min_negative_d < 0.0f) || !(0.0f < max_positive_d))  // the reference asserts these
    return fail(nullptr, VCB_ERR_INVALID, "need min_negative_d < 0 < max_positive_d");
  DeviceGuard restore_callers_device;
  int rc = use_device(nullptr, device);
  if (rc != VCB_OK) return rc;
  const size_t n = (size_t)width * height;
  Scratch buf;
  const size_t off_rgb = (n * sizeof(float) + 255) & ~(size_t)255;
  CU(nullptr, buf.reserve(off_rgb + n * 3));
  char* base = static_cast<char*>(buf.ptr);
  cudaError_t e = cudaMemcpy(base, sdf, n * sizeof(float), cudaMemcpyHostToDevice);
  if (e == cudaSuccess) {
    launch_sdf_color(reinterpret_cast<const float*>(base), (int64_t)n, min_negative_d,
                     max_positive_d, reinterpret_cast<uint8_t*>(base + off_rgb), nullptr);
    e = cudaGetLastError();
  }
  if (e == cudaSuccess) e = cudaMemcpy(rgb, base + off_rgb, n * 3, cudaMemcpyDeviceToHost);
  buf.release();
  if (e != cudaSuccess) return fail(nullptr, VCB_ERR_CUDA, "SignedDistance2Color: %s", cudaGetErrorString(e));
  return VCB_OK;
}

int vcb_distance_transform_l1(int32_t device, const uint8_t* mask, int32_t width, int32_t height,
                              const int32_t roi_min[2], const int32_t roi_max[2], float* dist) {
  return sdf_free_function(device, mask, width, height, roi_min, roi_max, kSdfDistanceL1, 0, 0,
                           0.1f, dist);
}

int vcb_make_signed_distance_field(int32_t device, const uint8_t* mask, int32_t width,
                                   int32_t height, const int32_t roi_min[2],
                                   const int32_t roi_max[2], int32_t minmax_normalize,
                                   int32_t use_truncation, float truncation_band, float* dist) {
  return sdf_free_function(device, mask, width, height, roi_min, roi_max, kSdfSigned,
                           minmax_normalize, use_truncation, truncation_band, dist);
}

int vcb_make_signed_distance_field_l2(int32_t device, const uint8_t* mask, int32_t width,
                                      int32_t height, const int32_t roi_min[2],
                                      const int32_t roi_max[2], int32_t minmax_normalize,
                                      int32_t use_truncation, float truncation_band, float* dist) {
  return sdf_free_function(device, mask, width, height, roi_min, roi_max, kSdfSignedL2,
                           minmax_normalize, use_truncation, truncation_band, dist);
}

int vcb_extract_iso_begin(vcb_carver* c, double iso_level, int32_t linear_interp,
                          int64_t* n_vertices, int64_t* n_faces) {
  if (!c) return fail(nullptr, VCB_ERR_INVALID, "null carver");
  int rc = vcb_flush(c);
  if (rc != VCB_OK) return rc;
  if (!c->un_implied && c->halo != 0)
    return fail(c, VCB_ERR_UNSUPPORTED,
                "marching cubes on a z slab needs the uploaded grid to keep update_num < 1 <=> sdf == "
                "kVal (update_num has no halo planes); upload sdf and update_num of a carved grid "
                "together");
  {
    Span span(c, 3);
    cudaError_t e = iso_count(grid_desc(c), c->un_implied, iso_level, linear_interp, &c->iso,
                              c->stream, &c->stage_launches[3]);
    if (e != cudaSuccess) return fail(c, VCB_ERR_CUDA, "marching cubes: %s", cudaGetErrorString(e));
  }
  if (c->iso.n_vertices > INT32_MAX || c->iso.n_faces > INT32_MAX)
    return fail(c, VCB_ERR_UNSUPPORTED, "mesh exceeds int32 indices (mesh.cc:241-271)");
  if (n_vertices) *n_vertices = c->iso.n_vertices;
  if (n_faces) *n_faces = c->iso.n_faces;
  return VCB_OK;
}

int vcb_extract_iso_vertices(vcb_carver* c, int64_t vertex_offset) {
  if (!c) return fail(nullptr, VCB_ERR_INVALID, "null carver");
  if (c->iso.phase != 1) return fail(c, VCB_ERR_STATE, "vcb_extract_iso_begin has not run");
  if (vertex_offset < 0 || vertex_offset + c->iso.n_vertices > INT32_MAX)
    return fail(c, VCB_ERR_UNSUPPORTED, "vertex ids exceed int32 (mesh.cc:241-271)");
  int rc = use_device(c, c->opt.device);
  if (rc != VCB_OK) return rc;
  Span span(c, 3);
  cudaError_t e = iso_emit_vertices(grid_desc(c), vertex_offset, &c->iso, c->stream,
                                    &c->stage_launches[3]);
  if (e != cudaSuccess) return fail(c, VCB_ERR_CUDA, "marching cubes vertices: %s", cudaGetErrorString(e));
  return VCB_OK;
}

int vcb_extract_iso_export_top(vcb_carver* c, int32_t* ids_dev) {
  if (!c || !ids_dev) return fail(c, VCB_ERR_INVALID, "null argument");
  if (c->iso.phase < 2) return fail(c, VCB_ERR_STATE, "vcb_extract_iso_vertices has not run");
  int rc = use_device(c, c->opt.device);
  if (rc != VCB_OK) return rc;
  Span span(c, 3);
  cudaError_t e = iso_export_top_ids(grid_desc(c), &c->iso, ids_dev, c->stream,
                                     &c->stage_launches[3]);
  if (e != cudaSuccess) return fail(c, VCB_ERR_CUDA, "export ids: %s", cudaGetErrorString(e));
  return VCB_OK;
}

int vcb_extract_iso_import_bottom(vcb_carver* c, const int32_t* ids_dev) {
  if (!c || !ids_dev) return fail(c, VCB_ERR_INVALID, "null argument");
  if (c->iso.phase != 2) return fail(c, VCB_ERR_STATE, "vcb_extract_iso_vertices has not run");
  if (c->z_begin < 1) return fail(c, VCB_ERR_INVALID, "the first slab has no slab below it");
  int rc = use_device(c, c->opt.device);
  if (rc != VCB_OK) return rc;
  Span span(c, 3);
  cudaError_t e = iso_import_bottom_ids(grid_desc(c), &c->iso, ids_dev, c->stream,
                                        &c->stage_launches[3]);
  if (e != cudaSuccess) return fail(c, VCB_ERR_CUDA, "import ids: %s", cudaGetErrorString(e));
  return VCB_OK;
}

int vcb_extract_iso_faces(vcb_carver* c) {
  if (!c) return fail(nullptr, VCB_ERR_INVALID, "null carver");
  if (c->iso.phase != 2) return fail(c, VCB_ERR_STATE, "vcb_extract_iso_vertices has not run");
  int rc = use_device(c, c->opt.device);
  if (rc != VCB_OK) return rc;
  Span span(c, 3);
  cudaError_t e = iso_emit_faces(grid_desc(c), &c->iso, c->stream, &c->stage_launches[3]);
  if (e != cudaSuccess) return fail(c, VCB_ERR_CUDA, "marching cubes faces: %s", cudaGetErrorString(e));
  return VCB_OK;
}

int vcb_extract_iso_count(vcb_carver* c, double iso_level, int32_t linear_interp,
                          int64_t* n_vertices, int64_t* n_faces) {
  int rc = vcb_extract_iso_begin(c, iso_level, linear_interp, n_vertices, n_faces);
  if (rc != VCB_OK) return rc;
  if ((rc = vcb_extract_iso_vertices(c, 0)) != VCB_OK) return rc;
  return vcb_extract_iso_faces(c);
}

int vcb_halo_device(vcb_carver* c, float** halo_dev, int32_t* n_planes) {
  if (!c) return fail(nullptr, VCB_ERR_INVALID, "null carver");
  if (halo_dev) *halo_dev = c->sdf_alloc;
  if (n_planes) *n_planes = c->halo;
  return VCB_OK;
}

int vcb_extract_iso_fill(vcb_carver* c, float* vertices, int32_t* faces) {
  if (!c) return fail(nullptr, VCB_ERR_INVALID, "null carver");
  if (c->iso.phase != 3) return fail(c, VCB_ERR_STATE, "vcb_extract_iso_count has not run");
  int rc = use_device(c, c->opt.device);
  if (rc != VCB_OK) return rc;
  if (vertices && c->iso.n_vertices > 0)
    CU(c, cudaMemcpyAsync(vertices, c->iso.verts.ptr, (size_t)c->iso.n_vertices * 12,
                          cudaMemcpyDeviceToHost, c->stream));
  if (faces && c->iso.n_faces > 0)
    CU(c, cudaMemcpyAsync(faces, c->iso.faces.ptr, (size_t)c->iso.n_faces * 12,
                          cudaMemcpyDeviceToHost, c->stream));
  CU(c, cudaStreamSynchronize(c->stream));
  return VCB_OK;
}

int vcb_extract_iso_device(vcb_carver* c, const float** vertices_dev, const int32_t** faces_dev) {
  if (!c) return fail(nullptr, VCB_ERR_INVALID, "null carver");
  if (c->iso.phase != 3) return fail(c, VCB_ERR_STATE, "vcb_extract_iso_count has not run");
  if (vertices_dev) *vertices_dev = static_cast<const float*>(c->iso.verts.ptr);
  if (faces_dev) *faces_dev = static_cast<const int32_t*>(c->iso.faces.ptr);
  return VCB_OK;
}

int vcb_ply_body_size(vcb_carver* c, size_t* vertex_bytes, size_t* face_bytes) {
  if (!c) return fail(nullptr, VCB_ERR_INVALID, "null carver");
  if (c->iso.phase != 3) return fail(c, VCB_ERR_STATE, "no extracted mesh: run vcb_extract_iso_count first");
  int rc = use_device(c, c->opt.device);
  if (rc != VCB_OK) return rc;
  {
    Span span(c, 3);
    CU(c, ply_measure(static_cast<const float*>(c->iso.verts.ptr), c->iso.n_vertices,
                      static_cast<const int32_t*>(c->iso.faces.ptr), c->iso.n_faces, &c->ply,
                      c->stream, &c->stage_launches[3]));
  }
  if (vertex_bytes) *vertex_bytes = (size_t)c->ply.vertex_bytes;
  if (face_bytes) *face_bytes = (size_t)c->ply.face_bytes;
  return VCB_OK;
}

int vcb_ply_body_fill(vcb_carver* c, char* vertex_text, char* face_text) {
  if (!c) return fail(nullptr, VCB_ERR_INVALID, "null carver");
  if (c->iso.phase != 3 || !c->ply.sized || c->ply.nv != c->iso.n_vertices ||
      c->ply.nf != c->iso.n_faces)
    return fail(c, VCB_ERR_STATE, "vcb_ply_body_size has not run for this mesh");
  int rc = use_device(c, c->opt.device);
  if (rc != VCB_OK) return rc;
  {
    Span span(c, 3);
    CU(c, ply_emit(static_cast<const float*>(c->iso.verts.ptr),
                   static_cast<const int32_t*>(c->iso.faces.ptr), &c->ply, c->stream,
                   &c->stage_launches[3]));
  }
  const char* text = static_cast<const char*>(c->ply.text.ptr);
  if (vertex_text && c->ply.vertex_bytes > 0)
    CU(c, cudaMemcpyAsync(vertex_text, text, (size_t)c->ply.vertex_bytes, cudaMemcpyDeviceToHost,
                          c->stream));
  if (face_text && c->ply.face_bytes > 0)
    CU(c, cudaMemcpyAsync(face_text, text + c->ply.vertex_bytes, (size_t)c->ply.face_bytes,
                          cudaMemcpyDeviceToHost, c->stream));
  CU(c, cudaStreamSynchronize(c->stream));
  return VCB_OK;
}

int vcb_extract_voxel_count(vcb_carver* c, int32_t inside_empty, int64_t* n_voxels) {
  if (!c) return fail(nullptr, VCB_ERR_INVALID, "null carver");
  int rc = vcb_flush(c);
  if (rc != VCB_OK) return rc;
  if (inside_empty && c->z_begin > 0 && (!c->un_implied || c->kval_touched_possible))
    return fail(c, VCB_ERR_UNSUPPORTED,
                "inside_empty on a z slab needs the uploaded grid to keep update_num < 1 <=> sdf == kVal "
                "(update_num has no halo plane)");
  {
    Span span(c, 2);
    cudaError_t e = voxel_extract_count(grid_desc(c), inside_empty,
                                        c->un_implied && !c->kval_touched_possible, &c->vox, c->stream,
                                        &c->stage_launches[2]);
    if (e != cudaSuccess) return fail(c, VCB_ERR_CUDA, "ExtractVoxel: %s", cudaGetErrorString(e));
  }
  if (n_voxels) *n_voxels = c->vox.n_voxels;
  return VCB_OK;
}

int vcb_extract_voxel_fill(vcb_carver* c, float* vertices, int32_t* faces) {
  if (!c || !vertices || !faces) return fail(c, VCB_ERR_INVALID, "null argument");
  if (c->vox.n_voxels < 0) return fail(c, VCB_ERR_STATE, "vcb_extract_voxel_count has not run");
  if (c->vox.n_voxels * 24 > INT32_MAX)
    return fail(c, VCB_ERR_UNSUPPORTED, "voxel mesh exceeds int32 indices (mesh.cc:241-271)");
  int rc = use_device(c, c->opt.device);
  if (rc != VCB_OK) return rc;
  Span span(c, 2);
  cudaError_t e = voxel_extract_fill(grid_desc(c), &c->vox, vertices, faces, c->stream,
                                     &c->stage_launches[2]);
  if (e != cudaSuccess) return fail(c, VCB_ERR_CUDA, "ExtractVoxel fill: %s", cudaGetErrorString(e));
  return VCB_OK;
}

int vcb_extract_voxel_occupancy(vcb_carver* c, uint8_t* flags) {
  if (!c || !flags) return fail(c, VCB_ERR_INVALID, "null argument");
  if (c->vox.n_voxels < 0) return fail(c, VCB_ERR_STATE, "vcb_extract_voxel_count has not run");
  int rc = use_device(c, c->opt.device);
  if (rc != VCB_OK) return rc;
  cudaError_t e = voxel_extract_flags(grid_desc(c), &c->vox, flags, c->stream,
                                      &c->stage_launches[2]);
  if (e != cudaSuccess) return fail(c, VCB_ERR_CUDA, "occupancy: %s", cudaGetErrorString(e));
  return VCB_OK;
}

int vcb_download_grid(vcb_carver* c, float* sdf, int32_t* update_num) {
  if (!c) return fail(nullptr, VCB_ERR_INVALID, "null carver");
  int rc = vcb_flush(c);
  if (rc != VCB_OK) return rc;
  const size_t n = (size_t)slab_voxels(c);
  if (sdf) CU(c, cudaMemcpyAsync(sdf, c->sdf, n * sizeof(float), cudaMemcpyDeviceToHost, c->stream));
  if (update_num)
    CU(c, cudaMemcpyAsync(update_num, c->un, n * sizeof(int32_t), cudaMemcpyDeviceToHost, c->stream));
  CU(c, cudaStreamSynchronize(c->stream));
  return VCB_OK;
}

int vcb_upload_grid(vcb_carver* c, const float* sdf, const int32_t* update_num) {
  if (!c) return fail(nullptr, VCB_ERR_INVALID, "null carver");
  int rc = vcb_flush(c);
  if (rc != VCB_OK) return rc;
  const size_t n = (size_t)slab_voxels(c);
  if (sdf) CU(c, cudaMemcpyAsync(c->sdf, sdf, n * sizeof(float), cudaMemcpyHostToDevice, c->stream));
  if (update_num)
    CU(c, cudaMemcpyAsync(c->un, update_num, n * sizeof(int32_t), cudaMemcpyHostToDevice, c->stream));
  CU(c, cudaStreamSynchronize(c->stream));
  if (update_num) {
    for (size_t i = 0; i < n; i++) {
      if (update_num[i] > c->max_count) c->max_count = update_num[i];
    }
  }
  // A grid that a carver produced keeps "update_num < 1 <=> sdf == kVal" (a restored checkpoint
  // does); only then may marching cubes go on reading sdf alone. Anything else is arbitrary state:
  // marching cubes must read update_num itself (and, lacking halo planes of it, cannot run on a
  // slab above the first - see vcb_extract_iso_begin).
  bool consistent = sdf != nullptr && update_num != nullptr;
  for (size_t i = 0; consistent && i < n; i++) {
    consistent = (update_num[i] < 1) == (sdf[i] == kInvalidSdf);
  }
  if (!consistent) c->un_implied = false;
  return VCB_OK;
}

int vcb_grid_device(vcb_carver* c, float** sdf_dev, int32_t** update_num_dev) {
  if (!c) return fail(nullptr, VCB_ERR_INVALID, "null carver");
  if (sdf_dev) *sdf_dev = c->sdf;
  if (update_num_dev) *update_num_dev = c->un;
  return VCB_OK;
}

int vcb_set_statistics(vcb_carver* c, int32_t enabled) {
  if (!c) return fail(nullptr, VCB_ERR_INVALID, "null carver");
  c->fp.count_iterations = enabled ? 1 : 0;
  return VCB_OK;
}

int vcb_fusion_class_counts(vcb_carver* c, uint64_t pairs[5], int64_t* voxels_per_tile,
                            uint64_t warp_iterations[2]) {
  if (!c || !pairs) return fail(c, VCB_ERR_INVALID, "null argument");
  int rc = use_device(c, c->opt.device);
  if (rc != VCB_OK) return rc;
  unsigned long long h[8];
  CU(c, cudaMemcpyAsync(h, c->cls_counts.ptr, sizeof(h), cudaMemcpyDeviceToHost, c->stream));
  CU(c, cudaMemsetAsync(c->cls_counts.ptr, 0, sizeof(h), c->stream));
  CU(c, cudaStreamSynchronize(c->stream));
  for (int k = 0; k < 5; k++) pairs[k] = h[k];
  pairs[0] += c->unclassified_pairs;
  c->unclassified_pairs = 0;
  if (voxels_per_tile) *voxels_per_tile = 32 * 8 * 8;
  if (warp_iterations) {
    warp_iterations[0] = h[5];
    warp_iterations[1] = h[6];
  }
  return VCB_OK;
}

int vcb_stage_times(vcb_carver* c, float ms[4], int64_t launches[4]) {
  if (!c) return fail(nullptr, VCB_ERR_INVALID, "null carver");
  int rc = use_device(c, c->opt.device);
  if (rc != VCB_OK) return rc;
  CU(c, cudaStreamSynchronize(c->stream));
  resolve_spans(c);
  for (int i = 0; i < 4; i++) {
    if (ms) ms[i] = c->stage_ms[i];
    if (launches) launches[i] = c->stage_launches[i];
    c->stage_ms[i] = 0;
    c->stage_launches[i] = 0;
  }
  return VCB_OK;
}

}  // extern "C"
