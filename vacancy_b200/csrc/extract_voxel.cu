// K2 - ExtractVoxel on the GPU, plus the device-wide scan shared with marching cubes.
//
// Replaces vacancy::ExtractVoxel (extract_voxel.cc:258-317) and UpdateOnSurface (:15-79).
//   voxel_occupied_kernel / voxel_flag_kernel : block per 4 grid rows, 16 coalesced loads in flight
//                       per thread; the emit predicate of :279-289 (lean kernel for carved grids,
//                       where "updated" is sdf != kVal) or the on_surface rule of :26-77 -> one
//                       ballot word per 32 voxels + one count per row.
//   scan_exclusive_u64: 3-kernel exclusive prefix sum over the per-ROW counts; rows are in raster
//                       order and a warp sums the popcounts before its segment, so numbering IS the
//                       reference's emission order.
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

// One block per kFlagRows grid rows; a warp iteration covers 128 voxels of each row with four
// coalesced loads of sdf and of update_num per row, evaluates the emit predicate, and turns every
// 32 voxels into one bitmap word with a ballot; each row's count is the sum of its popcounts.
constexpr int kFlagRows = 4;

// kImplied: the grid was carved (never uploaded), so update_num >= 1 exactly where sdf != kVal and
// update_num need not be read at all - half the traffic of the pass.
template <bool kImplied>
__device__ __forceinline__ bool updated(const GridDesc& g, size_t id, float s) {
  return kImplied ? (s != kInvalidSdf) : (g.update_num[id] >= 1);
}

template <bool kImplied>
__device__ __forceinline__ bool emit_predicate(const GridDesc& g, int inside_empty, float s, int n,
                                               size_t id, int x, int y, int zl) {
  if (!inside_empty) return !(s > 0.0f || n < 1);  // extract_voxel.cc:285
  if (n < 1) return false;                         // extract_voxel.cc:26-77
  const size_t plane = (size_t)g.nx * g.ny;
  const bool tiny = fabsf(s) < FLT_MIN;
  bool emit = false;
  if (x >= 1) {
    const float o = g.sdf[id - 1];
    if (updated<kImplied>(g, id - 1, o)) emit |= (__fmul_rn(s, o) < 0.0f) || tiny;
  }
  if (y >= 1) {
    const float o = g.sdf[id - g.nx];
    if (updated<kImplied>(g, id - g.nx, o)) emit |= (__fmul_rn(s, o) < 0.0f) || tiny;
  }
  if (zl >= 1) {
    const float o = g.sdf[id - plane];
    if (updated<kImplied>(g, id - plane, o)) emit |= (__fmul_rn(s, o) < 0.0f) || tiny;
  } else if (g.z_begin > 0 && g.halo >= 1) {
    // first plane of a z slab: the voxel below lives in the halo plane stored right under the slab
    // (filled from the lower slab before this kernel runs); update_num has no halo, but a grid that
    // was carved keeps update_num >= 1 <=> sdf != kVal (the caller checks that this holds)
    const float below = (g.sdf - plane)[id];
    if (below != kInvalidSdf) emit |= (__fmul_rn(s, below) < 0.0f) || tiny;
  }
  return emit;
}

template <bool kImplied, bool kSurface>
__global__ void __launch_bounds__(256) voxel_flag_kernel(const GridDesc g,
                                                         uint32_t* __restrict__ bits,
                                                         uint64_t* __restrict__ rowcnt) {
  const SegGeom sg = seg_geom(g);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, warps = blockDim.x >> 5;
  const int64_t n_rows = (int64_t)g.ny * (g.z_end - g.z_begin);
  const int64_t row0 = (int64_t)blockIdx.x * kFlagRows;
  const int groups = (sg.segs_per_row + 3) / 4;
  __shared__ unsigned s_cnt[kFlagRows];
  if (threadIdx.x < kFlagRows) s_cnt[threadIdx.x] = 0;
  __syncthreads();
  unsigned cnt[kFlagRows];
  int ry[kFlagRows], rz[kFlagRows];  // (y, slab-local z) of the block's rows, hoisted out of the loop
#pragma unroll
  for (int r = 0; r < kFlagRows; r++) {
    cnt[r] = 0;
    ry[r] = (int)((row0 + r) % g.ny);
    rz[r] = (int)((row0 + r) / g.ny);
  }
  for (int c = warp; c < groups; c += warps) {
    const int x = c * 128 + lane;
    float s[kFlagRows][4];
    int n[kFlagRows][4];
#pragma unroll
    for (int r = 0; r < kFlagRows; r++) {
      const bool row_ok = row0 + r < n_rows;
#pragma unroll
      for (int q = 0; q < 4; q++) {
        const bool ok = row_ok && x + 32 * q < g.nx;
        const size_t id = (size_t)(row0 + r) * g.nx + x + 32 * q;
        s[r][q] = ok ? g.sdf[id] : 1.0f;
        n[r][q] = !ok ? 0 : (kImplied ? (s[r][q] != kInvalidSdf ? 1 : 0) : g.update_num[id]);
      }
    }
#pragma unroll
    for (int r = 0; r < kFlagRows; r++) {
      const int64_t row = row0 + r;
      const int y = ry[r], zl = rz[r];
      unsigned word = 0;
#pragma unroll
      for (int q = 0; q < 4; q++) {
        const int xx = x + 32 * q;
        const bool ok = row < n_rows && xx < g.nx;
        const bool emit = ok && emit_predicate<kImplied>(g, kSurface ? 1 : 0, s[r][q], n[r][q],
                                               (size_t)row * g.nx + xx, xx, y, zl);
        const unsigned w_q = __ballot_sync(0xffffffffu, emit);
        if (lane == q) word = w_q;
        cnt[r] += lane == 0 ? __popc(w_q) : 0;
      }
      const int w = c * 4 + lane;
      if (lane < 4 && row < n_rows && w < sg.segs_per_row) bits[row * sg.segs_per_row + w] = word;
    }
  }
#pragma unroll
  for (int r = 0; r < kFlagRows; r++) {
    if (lane == 0 && cnt[r]) atomicAdd(&s_cnt[r], cnt[r]);
  }
  __syncthreads();
  if (threadIdx.x < kFlagRows && row0 + threadIdx.x < n_rows) {
    rowcnt[row0 + threadIdx.x] = s_cnt[threadIdx.x];
  }
}

// The common case - a carved grid, default rule (emit iff the voxel was updated and sdf <= 0,
// extract_voxel.cc:285; updated <=> sdf != kVal) - as a lean dense pass of its own: per warp iteration
// four coalesced 128-byte loads per row for four rows (16 loads in flight per thread), one ballot per
// load, lanes 0-3 store the four bitmap words of a row together.
__global__ void __launch_bounds__(256)
    voxel_occupied_kernel(const float* __restrict__ sdf, int nx, int spr, int64_t n_rows,
                          uint32_t* __restrict__ bits, uint64_t* __restrict__ rowcnt) {
  const int lane = threadIdx.x & 31;
  const int warps = blockDim.x >> 5;
  const int64_t row0 = (int64_t)blockIdx.x * kFlagRows;
  const int rows = (int)min((int64_t)kFlagRows, n_rows - row0);
  const float* __restrict__ src = sdf + row0 * nx;
  uint32_t* __restrict__ dst = bits + row0 * spr;
  __shared__ unsigned s_cnt[kFlagRows];
  if (threadIdx.x < kFlagRows) s_cnt[threadIdx.x] = 0;
  __syncthreads();
  const int groups = (spr + 3) / 4;
  for (int c = threadIdx.x >> 5; c < groups; c += warps) {
    const int x = c * 128 + lane;
    float s[kFlagRows][4];
#pragma unroll
    for (int r = 0; r < kFlagRows; r++) {
#pragma unroll
      for (int q = 0; q < 4; q++) {
        s[r][q] = (r < rows && x + 32 * q < nx) ? __ldg(src + (size_t)r * nx + x + 32 * q) : kInvalidSdf;
      }
    }
#pragma unroll
    for (int r = 0; r < kFlagRows; r++) {
      unsigned word = 0, n = 0;
#pragma unroll
      for (int q = 0; q < 4; q++) {
        const unsigned w_q = __ballot_sync(0xffffffffu, s[r][q] != kInvalidSdf && !(s[r][q] > 0.0f));
        if (lane == q) word = w_q;
        n += __popc(w_q);
      }
      const int w = c * 4 + lane;
      if (lane < 4 && r < rows && w < spr) dst[(size_t)r * spr + w] = word;
      if (lane == 0 && n) atomicAdd(&s_cnt[r], n);
    }
  }
  __syncthreads();
  if (threadIdx.x < rows) rowcnt[row0 + threadIdx.x] = s_cnt[threadIdx.x];
}

// number of flagged voxels of the row before segment `sx` (sum of popcounts, by one warp)
__device__ __forceinline__ unsigned row_prefix(const uint32_t* __restrict__ row_bits, int sx,
                                               int lane) {
  unsigned c = 0;
  for (int s0 = 0; s0 < sx; s0 += 32) {
    const int i = s0 + lane;
    c += i < sx ? __popc(row_bits[i]) : 0;
  }
  return __reduce_add_sync(0xffffffffu, c);
}

__global__ void __launch_bounds__(256) voxel_emit_kernel(const GridDesc g,
                                                         const uint32_t* __restrict__ bits,
                                                         const uint64_t* __restrict__ rowbase,
                                                         int64_t first, int64_t count,
                                                         float* __restrict__ vertices,
                                                         int32_t* __restrict__ faces) {
  const SegGeom sg = seg_geom(g);
  const int lane = threadIdx.x & 31;
  const int64_t seg = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (seg >= sg.n_segs) return;
  unsigned word = bits[seg];
  if (word == 0) return;
  const int sx = (int)(seg % sg.segs_per_row);
  const int64_t row = seg / sg.segs_per_row;
  // global cube number of the segment's first flagged voxel
  int64_t k = (int64_t)rowbase[row] + row_prefix(bits + row * sg.segs_per_row, sx, lane);
  if (k >= first + count || k + __popc(word) <= first) return;
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

cudaError_t voxel_extract_count(const GridDesc& g, int inside_empty, bool un_implied,
                                VoxelExtract* st, cudaStream_t stream, int64_t* launches) {
  const SegGeom sg = seg_geom(g);
  st->n_voxels = -1;
  st->inside_empty = inside_empty;
  if (inside_empty && g.z_begin > 0 && g.halo < 1) return cudaErrorNotSupported;  // needs the plane below
  const int64_t n_rows = (int64_t)g.ny * (g.z_end - g.z_begin);
  cudaError_t e;
  if ((e = st->bits.reserve((size_t)sg.n_segs * sizeof(uint32_t))) != cudaSuccess) return e;
  // per grid row: [0, n_rows) counts, [n_rows, 2 n_rows) exclusive prefix, then the total
  if ((e = st->segbase.reserve((size_t)(2 * n_rows + 1) * sizeof(uint64_t))) != cudaSuccess) return e;
  uint32_t* bits = static_cast<uint32_t*>(st->bits.ptr);
  uint64_t* rowcnt = static_cast<uint64_t*>(st->segbase.ptr);
  uint64_t* rowbase = rowcnt + n_rows;
  const unsigned flag_blocks = (unsigned)((n_rows + kFlagRows - 1) / kFlagRows);
  if (un_implied && inside_empty) {
    voxel_flag_kernel<true, true><<<flag_blocks, 256, 0, stream>>>(g, bits, rowcnt);
  } else if (un_implied) {
    voxel_occupied_kernel<<<flag_blocks, 256, 0, stream>>>(g.sdf, g.nx, sg.segs_per_row, n_rows, bits,
                                                           rowcnt);
  } else if (inside_empty) {
    voxel_flag_kernel<false, true><<<flag_blocks, 256, 0, stream>>>(g, bits, rowcnt);
  } else {
    voxel_flag_kernel<false, false><<<flag_blocks, 256, 0, stream>>>(g, bits, rowcnt);
  }
  if (launches) *launches += 1;
  e = scan_exclusive_u64(rowcnt, rowbase, n_rows, rowbase + n_rows, &st->tmp, stream, launches);
  if (e != cudaSuccess) return e;
  uint64_t total = 0;
  e = cudaMemcpyAsync(&total, rowbase + n_rows, sizeof(total), cudaMemcpyDeviceToHost, stream);
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
    voxel_emit_kernel<<<blocks, 256, 0, stream>>>(
        g, static_cast<uint32_t*>(st->bits.ptr),
        static_cast<uint64_t*>(st->segbase.ptr) + (int64_t)g.ny * (g.z_end - g.z_begin), first,
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
