// K0 - per-view 2D silhouette signed distance field on the GPU.
//
// Replaces DistanceTransformL1 (voxel_carver.cc:102-167) and MakeSignedDistanceField
// (voxel_carver.cc:169-237).  The reference runs a two-sweep 4-neighbour chamfer twice (mask and
// inverted mask); that transform is the exact city-block (L1) distance to the nearest pixel of the
// opposite class inside the ROI, and L1 is separable:
//     D(x, y) = min over y' of ( |y - y'| + R(x, y') ),   R = distance along the row.
// Distances are integers <= width + height - 2, so the passes run in int16/int32 and the float
// image produced at the end is bit-identical to the reference's.
//
//   sdf_row_kernel : one warp per ROI row; 32-pixel chunks, ballot + clz/ffs give the nearest
//                    opposite-class pixel to the left (forward sweep) and right (backward sweep).
//                    Output: signed row distance (fg negative, bg positive), kRowInf = none.
//   sdf_col_kernel : one block per 32 columns x 32 row segments; prefix-min of (g(y') - y') and
//                    suffix-min of (g(y') + y') per segment, carries exchanged through shared memory,
//                    for both polarities at once.  Output: signed L1 distance + max |D| (atomic).
//   sdf_finalize_kernel : int -> float, whole-image min/max normalisation (:205-222), truncation
//                    (:225-236), zeros outside the ROI (:104), and the image maximum that
//                    Carve needs for UpdateOutsideImage::kMax (:436).
// Algorithmic bytes: 1 B (mask) + 4 B (float SDF) per pixel; the int16 scratch adds 8 B/pixel of
// L2-resident traffic (views are processed kSdfSubBatch at a time to keep it there).
#include <float.h>
#include <stdio.h>

#include <mutex>

#include "common.cuh"

namespace vcb {

namespace {

__device__ __forceinline__ float atomic_max_float(float* addr, float value) {
  return value >= 0.0f
             ? __int_as_float(atomicMax(reinterpret_cast<int*>(addr), __float_as_int(value)))
             : __uint_as_float(
                   atomicMin(reinterpret_cast<unsigned*>(addr), __float_as_uint(value)));
}

constexpr int kRowWarps = 8;

__global__ void __launch_bounds__(kRowWarps * 32) sdf_row_kernel(const SdfView* __restrict__ views) {
  const SdfView v = views[blockIdx.y];
  const int lane = threadIdx.x & 31;
  const int warp = blockIdx.x * kRowWarps + (threadIdx.x >> 5);
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    *v.absmax = 0;
    *v.max_out = -FLT_MAX;
  }
  const int y = v.ry0 + warp;
  if (y > v.ry1) return;
  const uint8_t* __restrict__ m = v.mask + (size_t)y * v.width;
  int16_t* __restrict__ r = v.rowd + (size_t)y * v.width;
  constexpr int kFar = 1 << 20;
  const unsigned lt_mask = (1u << lane) - 1u;
  const unsigned gt_mask = ~((2u << lane) - 1u);

  int last_f = -kFar, last_b = -kFar;
  const int n_chunks = (v.rx1 - v.rx0) / 32 + 1;
  for (int c = 0; c < n_chunks; c++) {
    const int x0 = v.rx0 + c * 32;
    const int x = x0 + lane;
    const bool in = x <= v.rx1;
    const bool fg = in && (m[x] == 255);  // only 255 is foreground, voxel_carver.cc:109
    const unsigned fgb = __ballot_sync(0xffffffffu, fg);
    const unsigned bgb = __ballot_sync(0xffffffffu, in) & ~fgb;
    const unsigned left = (fg ? bgb : fgb) & lt_mask;
    const int xl = left ? (x0 + 31 - __clz(left)) : (fg ? last_b : last_f);
    const int dl = x - xl;
    if (in) r[x] = (int16_t)(dl > 32766 ? kRowInf : dl);
    if (fgb) last_f = x0 + 31 - __clz(fgb);
    if (bgb) last_b = x0 + 31 - __clz(bgb);
  }
  int next_f = 2 * kFar, next_b = 2 * kFar;
  for (int c = n_chunks - 1; c >= 0; c--) {
    const int x0 = v.rx0 + c * 32;
    const int x = x0 + lane;
    const bool in = x <= v.rx1;
    const bool fg = in && (m[x] == 255);
    const unsigned fgb = __ballot_sync(0xffffffffu, fg);
    const unsigned bgb = __ballot_sync(0xffffffffu, in) & ~fgb;
    const unsigned right = (fg ? bgb : fgb) & gt_mask;
    const int xr = right ? (x0 + __ffs(right) - 1) : (fg ? next_b : next_f);
    if (in) {
      int d = min((int)r[x], xr - x);
      if (d > 32766) d = kRowInf;
      r[x] = (int16_t)(fg ? -d : d);
    }
    if (fgb) next_f = x0 + __ffs(fgb) - 1;
    if (bgb) next_b = x0 + __ffs(bgb) - 1;
  }
}

constexpr int kColInf = 1 << 28;

__global__ void __launch_bounds__(1024) sdf_col_kernel(const SdfView* __restrict__ views) {
  const SdfView v = views[blockIdx.y];
  const int cx = threadIdx.x, seg = threadIdx.y;
  const int x = v.rx0 + blockIdx.x * 32 + cx;
  const bool col_ok = x <= v.rx1;
  const int roi_h = v.ry1 - v.ry0 + 1;
  const int seglen = (roi_h + 31) / 32;
  const int ys = v.ry0 + seg * seglen;
  const int ye = min(ys + seglen, v.ry1 + 1);
  const int W = v.width;

  __shared__ int s_pa[32][33], s_qa[32][33], s_pb[32][33], s_qb[32][33];

  // phase 1: per-segment prefix / suffix seeds for both polarities
  int pa = kColInf, qa = kColInf, pb = kColInf, qb = kColInf;
  if (col_ok) {
    for (int y = ys; y < ye; y++) {
      const int r = v.rowd[(size_t)y * W + x];
      const int ga = r < 0 ? (r == -kRowInf ? kColInf : -r) : 0;  // distance-to-background seeds
      const int gb = r > 0 ? (r == kRowInf ? kColInf : r) : 0;    // distance-to-foreground seeds
      pa = min(pa, ga - y);
      qa = min(qa, ga + y);
      pb = min(pb, gb - y);
      qb = min(qb, gb + y);
    }
  }
  s_pa[seg][cx] = pa;
  s_qa[seg][cx] = qa;
  s_pb[seg][cx] = pb;
  s_qb[seg][cx] = qb;
  __syncthreads();
  // phase 2: carries from the other segments of this column
  int cpa = kColInf, cpb = kColInf, cqa = kColInf, cqb = kColInf;
  for (int s = 0; s < seg; s++) {
    cpa = min(cpa, s_pa[s][cx]);
    cpb = min(cpb, s_pb[s][cx]);
  }
  for (int s = seg + 1; s < 32; s++) {
    cqa = min(cqa, s_qa[s][cx]);
    cqb = min(cqb, s_qb[s][cx]);
  }
  // phase 3: forward then backward over the segment
  int local_max = 0;
  if (col_ok) {
    int ra = cpa, rb = cpb;
    for (int y = ys; y < ye; y++) {
      const size_t i = (size_t)y * W + x;
      const int r = v.rowd[i];
      const int ga = r < 0 ? (r == -kRowInf ? kColInf : -r) : 0;
      const int gb = r > 0 ? (r == kRowInf ? kColInf : r) : 0;
      ra = min(ra, ga - y);
      rb = min(rb, gb - y);
      const int f = (r < 0 ? ra : rb) + y;
      v.cold[i] = (int16_t)(f > 32766 ? kRowInf : f);
    }
    ra = cqa;
    rb = cqb;
    for (int y = ye - 1; y >= ys; y--) {
      const size_t i = (size_t)y * W + x;
      const int r = v.rowd[i];
      const int ga = r < 0 ? (r == -kRowInf ? kColInf : -r) : 0;
      const int gb = r > 0 ? (r == kRowInf ? kColInf : r) : 0;
      ra = min(ra, ga + y);
      rb = min(rb, gb + y);
      int d = min((int)v.cold[i], (r < 0 ? ra : rb) - y);
      if (d > 32766) {
        d = kRowInf;
        local_max = INT_MAX;
      } else {
        local_max = max(local_max, d);
      }
      v.cold[i] = (int16_t)(r < 0 ? -d : d);
    }
  }
  local_max = __reduce_max_sync(0xffffffffu, local_max);
  if (cx == 0 && local_max > 0) atomicMax(v.absmax, local_max);
}

__global__ void __launch_bounds__(256) sdf_finalize_kernel(const SdfView* __restrict__ views,
                                                           int mode, int normalize, int truncate,
                                                           float band) {
  const SdfView v = views[blockIdx.y];
  const int64_t n = (int64_t)v.width * v.height;
  float norm = 1.0f;
  bool do_norm = false;
  if (normalize && mode == kSdfSigned) {
    const int am_i = *v.absmax;
    const float am = am_i == INT_MAX ? FLT_MAX : (float)am_i;
    if (am > FLT_MIN) {  // voxel_carver.cc:213
      do_norm = true;
      norm = __fdiv_rn(1.0f, am);
    }
  }
  float tmax = -FLT_MAX;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n;
       i += (int64_t)gridDim.x * blockDim.x) {
    const int x = (int)(i % v.width), y = (int)(i / v.width);
    float f = 0.0f;  // outside the ROI, voxel_carver.cc:104
    if (x >= v.rx0 && x <= v.rx1 && y >= v.ry0 && y <= v.ry1) {
      const int d = v.cold[i];
      if (mode == kSdfSigned) {
        f = d == kRowInf ? FLT_MAX : (d == -kRowInf ? -FLT_MAX : (float)d);
        if (do_norm) f = __fmul_rn(f, norm);
        if (truncate) {  // voxel_carver.cc:225-236
          if (-band >= f) {
            f = kInvalidSdf;
          } else {
            const float q = __fdiv_rn(f, band);
            f = q < 1.0f ? q : 1.0f;
          }
        }
      } else {  // DistanceTransformL1 of the mask itself: foreground distance, others 0
        f = d < 0 ? (d == -kRowInf ? FLT_MAX : (float)(-d)) : 0.0f;
      }
    }
    v.out[i] = f;
    tmax = fmaxf(tmax, f);
  }
  for (int o = 16; o > 0; o >>= 1) tmax = fmaxf(tmax, __shfl_xor_sync(0xffffffffu, tmax, o));
  if ((threadIdx.x & 31) == 0) atomic_max_float(v.max_out, tmax);
}

__global__ void __launch_bounds__(256) image_max_kernel(const float* __restrict__ img, int64_t n,
                                                        float* __restrict__ out) {
  float tmax = -FLT_MAX;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n;
       i += (int64_t)gridDim.x * blockDim.x) {
    const float f = img[i];
    tmax = (tmax < f) ? f : tmax;  // std::max_element order semantics (NaN never wins)
  }
  for (int o = 16; o > 0; o >>= 1) {
    const float t = __shfl_xor_sync(0xffffffffu, tmax, o);
    tmax = (tmax < t) ? t : tmax;
  }
  if ((threadIdx.x & 31) == 0) atomic_max_float(out, tmax);
}

__global__ void set_float_kernel(float* p, float v) { *p = v; }

}  // namespace

void launch_sdf2d(const SdfView* views_dev, const SdfView* views_host, int n_views, int mode,
                  int normalize, int truncate, float band, cudaStream_t stream,
                  int64_t* launches) {
  int max_rows = 0, max_cols = 0;
  int64_t max_pix = 0;
  for (int i = 0; i < n_views; i++) {
    const SdfView& v = views_host[i];
    max_rows = max(max_rows, v.ry1 - v.ry0 + 1);
    max_cols = max(max_cols, v.rx1 - v.rx0 + 1);
    max_pix = max(max_pix, (int64_t)v.width * v.height);
  }
  dim3 grow((max_rows + kRowWarps - 1) / kRowWarps, n_views);
  sdf_row_kernel<<<grow, kRowWarps * 32, 0, stream>>>(views_dev);
  dim3 gcol((max_cols + 31) / 32, n_views);
  sdf_col_kernel<<<gcol, dim3(32, 32), 0, stream>>>(views_dev);
  int64_t fblocks = (max_pix + 256 * 8 - 1) / (256 * 8);
  if (fblocks < 1) fblocks = 1;
  if (fblocks > 4096) fblocks = 4096;
  dim3 gfin((unsigned)fblocks, n_views);
  sdf_finalize_kernel<<<gfin, 256, 0, stream>>>(views_dev, mode, normalize, truncate, band);
  if (launches) *launches += 3;
}

void launch_image_max(const float* image, int64_t n, float* max_out, cudaStream_t stream,
                      int64_t* launches) {
  set_float_kernel<<<1, 1, 0, stream>>>(max_out, -FLT_MAX);
  int64_t blocks = (n + 256 * 8 - 1) / (256 * 8);
  if (blocks < 1) blocks = 1;
  if (blocks > 2048) blocks = 2048;
  image_max_kernel<<<(unsigned)blocks, 256, 0, stream>>>(image, n, max_out);
  if (launches) *launches += 2;
}

}  // namespace vcb

// ------------------------------------------------------------------------------------------------
// vcb_make_signed_distance_field_noalloc: the 2D SDF of one mask without a single allocation.
//
// The three kernels above need 4 bytes of scratch per pixel (two int16 images) plus a small
// descriptor block.  Here that scratch is this fixed array, part of the library image: every GPU
// that loads libvacancy_b200.so carries it (16 MiB + 256 B) until the process exits, and images of
// more than kNoAllocMaxPixels pixels are refused.  All calls share the one array.
namespace vcb {
namespace {

constexpr int64_t kNoAllocMaxPixels = 2048LL * 2048LL;
constexpr size_t kNoAllocHeader = 256;  // SdfView descriptor, abs-max and image-max cells
constexpr size_t kNoAllocBytes = kNoAllocHeader + 2 * sizeof(int16_t) * kNoAllocMaxPixels;

static __device__ __align__(256) unsigned char g_noalloc_scratch[kNoAllocBytes];

// In host code `g_noalloc_scratch` is only a handle for the runtime, not a device address: the
// real address is asked for once per device (with that device current) and cached here.
std::mutex g_noalloc_mutex;
unsigned char* g_noalloc_ptr[64] = {nullptr};

__global__ void sdf_noalloc_set_view_kernel(unsigned char* scratch, SdfView v) {
  *reinterpret_cast<SdfView*>(scratch) = v;
}

}  // namespace
}  // namespace vcb

extern "C" int vcb_make_signed_distance_field_noalloc(
    int32_t device, const uint8_t* mask_dev, int32_t width, int32_t height,
    const int32_t roi_min[2], const int32_t roi_max[2], int32_t minmax_normalize,
    int32_t use_truncation, float truncation_band, float* dist_dev, void* cuda_stream) {
  using namespace vcb;
  char msg[256];
  if (!mask_dev || !dist_dev || !roi_min || !roi_max) {
    set_global_error("vcb_make_signed_distance_field_noalloc: null argument");
    return VCB_ERR_INVALID;
  }
  if (width < 1 || height < 1 || width > kMaxImageDim || height > kMaxImageDim ||
      (int64_t)width * height > kNoAllocMaxPixels) {
    snprintf(msg, sizeof(msg),
             "vcb_make_signed_distance_field_noalloc: image %d x %d = %lld pixels exceeds the fixed "
             "scratch of %lld pixels; use vcb_make_signed_distance_field for larger images",
             width, height, (long long)width * height, (long long)kNoAllocMaxPixels);
    set_global_error(msg);
    return VCB_ERR_INVALID;
  }
  if (roi_min[0] < 0 || roi_min[1] < 0 || roi_max[0] >= width || roi_max[1] >= height ||
      roi_min[0] > roi_max[0] || roi_min[1] > roi_max[1]) {
    set_global_error("vcb_make_signed_distance_field_noalloc: roi outside the image");
    return VCB_ERR_INVALID;
  }
  if (device < 0 || device >= 64) {
    set_global_error("vcb_make_signed_distance_field_noalloc: device ordinal out of range");
    return VCB_ERR_INVALID;
  }
  cudaError_t e = cudaSetDevice(device);
  unsigned char* scratch = nullptr;
  if (e == cudaSuccess) {
    std::lock_guard<std::mutex> lock(g_noalloc_mutex);
    if (g_noalloc_ptr[device] == nullptr) {  // first call on this device: resolve the symbol
      void* p = nullptr;
      e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
      if (e == cudaSuccess) g_noalloc_ptr[device] = static_cast<unsigned char*>(p);
    }
    scratch = g_noalloc_ptr[device];
  }
  if (e != cudaSuccess || scratch == nullptr) {
    snprintf(msg, sizeof(msg), "vcb_make_signed_distance_field_noalloc: %s", cudaGetErrorString(e));
    set_global_error(msg);
    return VCB_ERR_CUDA;
  }
  cudaStream_t stream = static_cast<cudaStream_t>(cuda_stream);
  SdfView v;
  v.mask = mask_dev;
  v.out = dist_dev;
  v.rowd = reinterpret_cast<int16_t*>(scratch + kNoAllocHeader);
  v.cold = v.rowd + kNoAllocMaxPixels;
  v.absmax = reinterpret_cast<int*>(scratch + 128);      // both cells are re-initialised by
  v.max_out = reinterpret_cast<float*>(scratch + 192);   // sdf_row_kernel at the start of a call
  v.width = width;
  v.height = height;
  v.rx0 = roi_min[0];
  v.ry0 = roi_min[1];
  v.rx1 = roi_max[0];
  v.ry1 = roi_max[1];
  sdf_noalloc_set_view_kernel<<<1, 1, 0, stream>>>(scratch, v);
  launch_sdf2d(reinterpret_cast<const SdfView*>(scratch), &v, 1, kSdfSigned, minmax_normalize,
               use_truncation, truncation_band, stream, nullptr);
  e = cudaGetLastError();
  if (e != cudaSuccess) {
    snprintf(msg, sizeof(msg), "vcb_make_signed_distance_field_noalloc: %s", cudaGetErrorString(e));
    set_global_error(msg);
    return VCB_ERR_CUDA;
  }
  return VCB_OK;
}
