// K2 - ExtractVoxel on the GPU, plus the device-wide scan shared with marching cubes.
//
// Replaces vacancy::ExtractVoxel (extract_voxel.cc:258-317) and UpdateOnSurface (:15-79).
//   voxel_flag_kernel : one warp per 32-voxel segment along x; the emit predicate of :279-289
//                       (or the on_surface rule of :26-77) -> ballot word + popcount per segment.
//   scan_exclusive_u64: 3-kernel exclusive prefix sum over the per-segment counts; segments are
//                       in raster order (x fastest), so the prefix IS the reference's emission order.
//   voxel_emit_kernel : every flagged voxel becomes MakeCube(resolution) (mesh.cc:728-780) moved to
//                       the voxel centre: 24 vertices + 12 faces, written cooperatively by the warp.
// HBM-bound: 8 B/voxel read (sdf + update_num) + 432 B per emitted voxel written.
#include <float.h>

#include "common.cuh"

namespace vcb {

cudaError_t Scratch::reserve(size_t n) {
  if (n <= bytes) return cudaSuccess;
  if (ptr) cudaFree(ptr);
  ptr = nullptr;
  bytes = 0;
  cudaError_t e = cudaMalloc(&ptr, n);
  if (e == cudaSuccess) bytes = n;
  return e;
}

void Scratch::release() {
  if (ptr) cudaFree(ptr);
  ptr = nullptr;
  bytes = 0;
}

// ------------------------------------------------------------------------------------------------
namespace {

constexpr int kScanThreads = 256;
constexpr int kScanItems = 16;
constexpr int kScanTile = kScanThreads * kScanItems;

__device__ __forceinline__ uint64_t block_exclusive_scan(uint64_t v, uint64_t* total) {
  __shared__ uint64_t warp_sums[kScanThreads / 32];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  uint64_t inc = v;
  for (int o = 1; o < 32; o <<= 1) {
    const uint64_t t = __shfl_up_sync(0xffffffffu, inc, o);
    if (lane >= o) inc += t;
  }
  if (lane == 31) warp_sums[warp] = inc;
  __syncthreads();
  uint64_t warp_off = 0, all = 0;
  for (int w = 0; w < kScanThreads / 32; w++) {
    const uint64_t t = warp_sums[w];
    if (w < warp) warp_off += t;
    all += t;
  }
  __syncthreads();
  *total = all;
  return warp_off + inc - v;
}

__global__ void __launch_bounds__(kScanThreads) scan_reduce_kernel(const uint64_t* __restrict__ in,
                                                                   int64_t n,
                                                                   uint64_t* __restrict__ sums) {
  const int64_t base = (int64_t)blockIdx.x * kScanTile + (int64_t)threadIdx.x * kScanItems;
  uint64_t s = 0;
#pragma unroll
  for (int i = 0; i < kScanItems; i++) {
    if (base + i < n) s += in[base + i];
  }
  uint64_t total;
  block_exclusive_scan(s, &total);
  if (threadIdx.x == 0) sums[blockIdx.x] = total;
}

__global__ void __launch_bounds__(kScanThreads) scan_sums_kernel(uint64_t* __restrict__ sums,
                                                                 int64_t n_blocks,
                                                                 uint64_t* __restrict__ total_out) {
  uint64_t carry = 0;
  for (int64_t base = 0; base < n_blocks; base += kScanThreads) {
    const int64_t i = base + threadIdx.x;
    const uint64_t v = i < n_blocks ? sums[i] : 0;
    uint64_t total;
    const uint64_t ex = block_exclusive_scan(v, &total);
    if (i < n_blocks) sums[i] = carry + ex;
    carry += total;
  }
  if (threadIdx.x == 0) *total_out = carry;
}

__global__ void __launch_bounds__(kScanThreads) scan_apply_kernel(const uint64_t* __restrict__ in,
                                                                  uint64_t* __restrict__ out,
                                                                  int64_t n,
                                                                  const uint64_t* __restrict__ sums) {
  const int64_t base = (int64_t)blockIdx.x * kScanTile + (int64_t)threadIdx.x * kScanItems;
  uint64_t v[kScanItems];
  uint64_t s = 0;
#pragma unroll
  for (int i = 0; i < kScanItems; i++) {
    v[i] = (base + i < n) ? in[base + i] : 0;
    s += v[i];
  }
  uint64_t total;
  uint64_t run = block_exclusive_scan(s, &total) + sums[blockIdx.x];
#pragma unroll
  for (int i = 0; i < kScanItems; i++) {
    if (base + i < n) out[base + i] = run;
    run += v[i];
  }
}

}  // namespace

cudaError_t scan_exclusive_u64(const uint64_t* in, uint64_t* out, int64_t n, uint64_t* total_dev,
                               Scratch* tmp, cudaStream_t stream, int64_t* launches) {
  const int64_t n_blocks = (n + kScanTile - 1) / kScanTile;
  cudaError_t e = tmp->reserve((size_t)(n_blocks > 0 ? n_blocks : 1) * sizeof(uint64_t));
  if (e != cudaSuccess) return e;
  uint64_t* sums = static_cast<uint64_t*>(tmp->ptr);
  if (n_blocks > 0) {
    scan_reduce_kernel<<<(unsigned)n_blocks, kScanThreads, 0, stream>>>(in, n, sums);
  }
  scan_sums_kernel<<<1, kScanThreads, 0, stream>>>(sums, n_blocks, total_dev);
  if (n_blocks > 0) {
    scan_apply_kernel<<<(unsigned)n_blocks, kScanThreads, 0, stream>>>(in, out, n, sums);
  }
  if (launches) *launches += 3;
  return cudaGetLastError();
}

// ------------------------------------------------------------------------------------------------
namespace {

// MakeCube corner signs and faces, mesh.cc:740-780 (vertex i = (sx, sy, sz) * resolution/2)
__constant__ signed char c_cube_sign[24][3] = {
    {-1, 1, -1},  {1, 1, -1},  {1, 1, 1},   {-1, 1, 1},   {-1, -1, -1}, {1, -1, -1},
    {1, -1, 1},   {-1, -1, 1}, {1, 1, -1},  {1, 1, 1},    {1, -1, 1},   {1, -1, -1},
    {-1, 1, -1},  {-1, 1, 1},  {-1, -1, 1}, {-1, -1, -1}, {-1, 1, -1},  {1, 1, -1},
    {1, -1, -1},  {-1, -1, -1}, {-1, 1, 1}, {1, 1, 1},    {1, -1, 1},   {-1, -1, 1}};
__constant__ unsigned char c_cube_face[36] = {0,  2,  1,  0,  3,  2,  4,  5,  6,  4,  6,  7,
                                              8,  9,  10, 8,  10, 11, 12, 14, 13, 12, 15, 14,
                                              16, 17, 18, 16, 18, 19, 20, 22, 21, 20, 23, 22};

struct SegGeom {
  int segs_per_row;
  int64_t n_segs;  // segs_per_row * ny * (z_end - z_begin)
};

__host__ __device__ inline SegGeom seg_geom(const GridDesc& g) {
  SegGeom s;
  s.segs_per_row = (g.nx + kSegVoxels - 1) / kSegVoxels;
  s.n_segs = (int64_t)s.segs_per_row * g.ny * (g.z_end - g.z_begin);
  return s;
}

__global__ void __launch_bounds__(256) voxel_flag_kernel(const GridDesc g, int inside_empty,
                                                         uint32_t* __restrict__ bits,
                                                         uint64_t* __restrict__ segcnt) {
  const SegGeom sg = seg_geom(g);
  const int lane = threadIdx.x & 31;
  const int64_t seg = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (seg >= sg.n_segs) return;
  const int sx = (int)(seg % sg.segs_per_row);
  const int64_t row = seg / sg.segs_per_row;
  const int y = (int)(row % g.ny);
  const int zl = (int)(row / g.ny);  // local plane
  const int x = sx * kSegVoxels + lane;
  bool emit = false;
  if (x < g.nx) {
    const size_t plane = (size_t)g.nx * g.ny;
    const size_t id = (size_t)zl * plane + (size_t)y * g.nx + x;
    const float s = g.sdf[id];
    const int n = g.update_num[id];
    if (!inside_empty) {
      emit = !(s > 0.0f || n < 1);  // extract_voxel.cc:285
    } else if (n >= 1) {            // extract_voxel.cc:26-77
      const bool tiny = fabsf(s) < FLT_MIN;
      // previous voxel along x, y, z; the z neighbour of the first owned plane is the halo plane
      if (x >= 1 && g.update_num[id - 1] >= 1) {
        emit |= (__fmul_rn(s, g.sdf[id - 1]) < 0.0f) || tiny;
      }
      if (y >= 1 && g.update_num[id - g.nx] >= 1) {
        emit |= (__fmul_rn(s, g.sdf[id - g.nx]) < 0.0f) || tiny;
      }
      if (zl >= 1 && g.update_num[id - plane] >= 1) {
        emit |= (__fmul_rn(s, g.sdf[id - plane]) < 0.0f) || tiny;
      }
    }
  }
  const unsigned word = __ballot_sync(0xffffffffu, emit);
  if (lane == 0) {
    bits[seg] = word;
    segcnt[seg] = (uint64_t)__popc(word);
  }
}

__global__ void __launch_bounds__(256) voxel_emit_kernel(const GridDesc g,
                                                         const uint32_t* __restrict__ bits,
                                                         const uint64_t* __restrict__ segbase,
                                                         int64_t first, int64_t count,
                                                         float* __restrict__ vertices,
                                                         int32_t* __restrict__ faces) {
  const SegGeom sg = seg_geom(g);
  const int lane = threadIdx.x & 31;
  const int64_t seg = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (seg >= sg.n_segs) return;
  unsigned word = bits[seg];
  if (word == 0) return;
  int64_t k = (int64_t)segbase[seg];  // global cube number of the first flagged voxel
  if (k >= first + count || k + __popc(word) <= first) return;
  const int sx = (int)(seg % sg.segs_per_row);
  const int64_t row = seg / sg.segs_per_row;
  const int y = (int)(row % g.ny);
  const int z = (int)(row / g.ny) + g.z_begin;
  const float h = g.resolution / 2;  // mesh.cc:736-738
  const float py = g.pos_y[y], pz = g.pos_z[z];
  while (word) {
    const int b = __ffs(word) - 1;
    word &= word - 1;
    if (k >= first && k < first + count) {
      const float px = g.pos_x[sx * kSegVoxels + b];
      float* v = vertices + (k - first) * 72;
      int32_t* f = faces + (k - first) * 36;
      for (int i = lane; i < 72; i += 32) {
        const int vi = i / 3, a = i - vi * 3;
        const float c = c_cube_sign[vi][a] < 0 ? -h : h;
        v[i] = __fadd_rn(c, a == 0 ? px : (a == 1 ? py : pz));  // Translate: v + t, mesh.cc:161-164
      }
      for (int i = lane; i < 36; i += 32) f[i] = (int32_t)(c_cube_face[i] + 24 * k);
    }
    k++;
  }
}

__global__ void __launch_bounds__(256) voxel_unpack_kernel(const GridDesc g,
                                                           const uint32_t* __restrict__ bits,
                                                           uint8_t* __restrict__ flags) {
  const SegGeom sg = seg_geom(g);
  const int lane = threadIdx.x & 31;
  const int64_t seg = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (seg >= sg.n_segs) return;
  const int sx = (int)(seg % sg.segs_per_row);
  const int64_t row = seg / sg.segs_per_row;
  const int x = sx * kSegVoxels + lane;
  if (x < g.nx) flags[row * g.nx + x] = (bits[seg] >> lane) & 1u;
}

}  // namespace

cudaError_t voxel_extract_count(const GridDesc& g, int inside_empty, VoxelExtract* st,
                                cudaStream_t stream, int64_t* launches) {
  const SegGeom sg = seg_geom(g);
  st->n_voxels = -1;
  st->inside_empty = inside_empty;
  cudaError_t e;
  if ((e = st->bits.reserve((size_t)sg.n_segs * sizeof(uint32_t))) != cudaSuccess) return e;
  // segbase holds n_segs prefixes followed by the total
  if ((e = st->segbase.reserve((size_t)(sg.n_segs + 1) * sizeof(uint64_t))) != cudaSuccess) return e;
  uint32_t* bits = static_cast<uint32_t*>(st->bits.ptr);
  uint64_t* segbase = static_cast<uint64_t*>(st->segbase.ptr);
  GridDesc gd = g;
  if (inside_empty && g.halo == 0 && g.z_begin > 0) return cudaErrorInvalidValue;
  const int64_t threads = sg.n_segs * 32;
  const unsigned blocks = (unsigned)((threads + 255) / 256);
  voxel_flag_kernel<<<blocks, 256, 0, stream>>>(gd, inside_empty, bits, segbase);
  if (launches) *launches += 1;
  e = scan_exclusive_u64(segbase, segbase, sg.n_segs, segbase + sg.n_segs, &st->tmp, stream,
                         launches);
  if (e != cudaSuccess) return e;
  uint64_t total = 0;
  e = cudaMemcpyAsync(&total, segbase + sg.n_segs, sizeof(total), cudaMemcpyDeviceToHost, stream);
  if (e != cudaSuccess) return e;
  if ((e = cudaStreamSynchronize(stream)) != cudaSuccess) return e;
  st->n_voxels = (int64_t)total;
  return cudaSuccess;
}

cudaError_t voxel_extract_fill(const GridDesc& g, VoxelExtract* st, float* vertices_host,
                               int32_t* faces_host, cudaStream_t stream, int64_t* launches) {
  const SegGeom sg = seg_geom(g);
  const int64_t chunk = 1 << 18;  // cubes per staging round (113 MB of staging)
  cudaError_t e = st->stage.reserve((size_t)chunk * (72 * sizeof(float) + 36 * sizeof(int32_t)));
  if (e != cudaSuccess) return e;
  float* dv = static_cast<float*>(st->stage.ptr);
  int32_t* df = reinterpret_cast<int32_t*>(dv + chunk * 72);
  const int64_t threads = sg.n_segs * 32;
  const unsigned blocks = (unsigned)((threads + 255) / 256);
  for (int64_t first = 0; first < st->n_voxels; first += chunk) {
    const int64_t count = st->n_voxels - first < chunk ? st->n_voxels - first : chunk;
    voxel_emit_kernel<<<blocks, 256, 0, stream>>>(g, static_cast<uint32_t*>(st->bits.ptr),
                                                  static_cast<uint64_t*>(st->segbase.ptr), first,
                                                  count, dv, df);
    if (launches) *launches += 1;
    e = cudaMemcpyAsync(vertices_host + first * 72, dv, (size_t)count * 72 * sizeof(float),
                        cudaMemcpyDeviceToHost, stream);
    if (e != cudaSuccess) return e;
    e = cudaMemcpyAsync(faces_host + first * 36, df, (size_t)count * 36 * sizeof(int32_t),
                        cudaMemcpyDeviceToHost, stream);
    if (e != cudaSuccess) return e;
    if ((e = cudaStreamSynchronize(stream)) != cudaSuccess) return e;
  }
  return cudaGetLastError();
}

cudaError_t voxel_extract_flags(const GridDesc& g, VoxelExtract* st, uint8_t* flags_host,
                                cudaStream_t stream, int64_t* launches) {
  const SegGeom sg = seg_geom(g);
  const size_t n = (size_t)g.nx * g.ny * (g.z_end - g.z_begin);
  cudaError_t e = st->stage.reserve(n);
  if (e != cudaSuccess) return e;
  const int64_t threads = sg.n_segs * 32;
  const unsigned blocks = (unsigned)((threads + 255) / 256);
  voxel_unpack_kernel<<<blocks, 256, 0, stream>>>(g, static_cast<uint32_t*>(st->bits.ptr),
                                                  static_cast<uint8_t*>(st->stage.ptr));
  if (launches) *launches += 1;
  e = cudaMemcpyAsync(flags_host, st->stage.ptr, n, cudaMemcpyDeviceToHost, stream);
  if (e != cudaSuccess) return e;
  return cudaStreamSynchronize(stream);
}

}  // namespace vcb
