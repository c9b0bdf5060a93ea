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
//   sdf_bits_kernel: the foreground bit-plane (1 bit per pixel, ROI applied), 32 pixels per thread.
//   sdf_row_kernel : one warp per ROI row, one lane per 32-pixel word: warp max/min scans carry the
//                    nearest foreground / background pixel across words, clz/ffs resolve it inside a
//                    word.  Output: signed row distance (fg negative, bg positive), kRowInf = none.
//   sdf_col_pack_kernel : the column pass of every image with width + height <= 16000 (all the
//                    benchmark configurations): two columns per thread on the packed 16-bit SIMD
//                    integer instructions (VIADDMNMX.S16x2), 16 row segments per block; a forward
//                    sweep per segment without carries, segment summaries through shared memory, a
//                    backward sweep with both carries; both polarities at once.  Output: signed L1
//                    distance (int16) + max |D| (atomic).
//   sdf_col_smem_kernel / sdf_col_kernel : the same pass for larger images (strip staged in shared
//                    memory / three sweeps over global memory), int32 arithmetic.
//   sdf_lut_kernel : the float value of a pixel depends on its integer distance only, so int ->
//                    float, min/max normalisation (:205-222) and truncation (:225-236) run once per
//                    possible distance into a per-view table.
//   sdf_finalize_kernel : table gather per pixel, zeros outside the ROI (:104), the image maximum
//                    that Carve needs for UpdateOutsideImage::kMax (:436) and the 8x8-block flags
//                    the fusion kernel classifies voxel tiles with.
// Algorithmic bytes: 1 B (mask) + 4 B (float SDF) per pixel; the int16 scratch adds L2-resident
// traffic (views are processed kSdfSubBatch at a time to keep it there).
#include <float.h>
#include <stdio.h>

#include <mutex>

#include "common.cuh"

namespace vcb {

// Per-view scratch: two int16 images with the row pitch rounded up to 32 pixels (so that every
// scratch access is aligned whatever the image width) and one bit per pixel.
__host__ __device__ inline int sdf_pitch(int width) { return (width + 31) & ~31; }
__host__ __device__ inline size_t sdf_plane_bytes(int width, int height) {
  return ((size_t)sdf_pitch(width) * height * sizeof(int16_t) + 255) & ~(size_t)255;
}
__host__ __device__ inline size_t sdf_bits_bytes(int width, int height) {
  return ((size_t)(sdf_pitch(width) / 32) * height * sizeof(uint32_t) + 255) & ~(size_t)255;
}
// value table of the finalize pass (L1 modes): one float per possible signed distance, see
// sdf_lut_kernel; index = clamp(d, -n, n) + n with n = width + height
__host__ __device__ inline int sdf_lut_entries(int width, int height) { return 2 * (width + height) + 1; }
__host__ __device__ inline size_t sdf_lut_bytes(int width, int height) {
  return ((size_t)sdf_lut_entries(width, height) * sizeof(float) + 255) & ~(size_t)255;
}
size_t sdf_scratch_bytes(int width, int height, int mode) {
  const size_t base = 2 * sdf_plane_bytes(width, height) + sdf_bits_bytes(width, height) +
                      sdf_lut_bytes(width, height);
  // Euclidean mode: parabola stack (two int32 planes) + squared distances (one int32 plane)
  return mode == kSdfSignedL2 ? base + 6 * sdf_plane_bytes(width, height) : base;
}

namespace {

__device__ __forceinline__ int16_t* view_rowd(const SdfView& v) {
  return reinterpret_cast<int16_t*>(v.scratch);
}
__device__ __forceinline__ int16_t* view_cold(const SdfView& v) {
  return reinterpret_cast<int16_t*>(v.scratch + sdf_plane_bytes(v.width, v.height));
}
__device__ __forceinline__ uint32_t* view_bits(const SdfView& v) {
  return reinterpret_cast<uint32_t*>(v.scratch + 2 * sdf_plane_bytes(v.width, v.height));
}
__device__ __forceinline__ float* view_lut(const SdfView& v) {
  return reinterpret_cast<float*>(v.scratch + 2 * sdf_plane_bytes(v.width, v.height) +
                                  sdf_bits_bytes(v.width, v.height));
}
// int32 planes of the Euclidean mode: 0 = stack of parabolas, 1 = their start rows, 2 = result
__device__ __forceinline__ int32_t* view_l2_plane(const SdfView& v, int k) {
  return reinterpret_cast<int32_t*>(v.scratch + 2 * sdf_plane_bytes(v.width, v.height) +
                                    sdf_bits_bytes(v.width, v.height) +
                                    sdf_lut_bytes(v.width, v.height) +
                                    (size_t)k * 2 * sdf_plane_bytes(v.width, v.height));
}

__device__ __forceinline__ float atomic_max_float(float* addr, float value) {
  return value >= 0.0f
             ? __int_as_float(atomicMax(reinterpret_cast<int*>(addr), __float_as_int(value)))
             : __uint_as_float(
                   atomicMin(reinterpret_cast<unsigned*>(addr), __float_as_uint(value)));
}

// bits [lo, hi] (inclusive) of a 32-bit word whose bit 0 is pixel x0; empty when out of range
__device__ __forceinline__ unsigned roi_word_mask(int x0, int rx0, int rx1) {
  const int lo = max(rx0 - x0, 0), hi = min(rx1 - x0, 31);
  if (lo > 31 || hi < 0 || lo > hi) return 0u;
  return (0xffffffffu >> (31 - hi)) & (0xffffffffu << lo);
}

// ---- foreground bits of 32 pixels of row y starting at x0 = 32 * w, ROI applied -----------------
// Only the value 255 is foreground (voxel_carver.cc:109).
__device__ __forceinline__ unsigned mask_word(const SdfView& v, int y, int w) {
  const int x0 = w * 32;
  const uint8_t* __restrict__ m = v.mask + (size_t)y * v.width + x0;
  unsigned f = 0;
  if ((v.width & 3) == 0 && (reinterpret_cast<size_t>(v.mask) & 3) == 0 && x0 + 32 <= v.width) {
    // byte == 0xFF  <=>  bit 7 set and the low 7 bits + 1 carry into bit 7; gather the four flags
    const uint32_t* __restrict__ m4 = reinterpret_cast<const uint32_t*>(m);
#pragma unroll
    for (int i = 0; i < 8; i++) {
      const unsigned q = __ldg(m4 + i);
      const unsigned t = ((q & 0x7f7f7f7fu) + 0x01010101u) & q & 0x80808080u;
      f |= ((((t >> 7) * 0x00204081u) >> 21) & 0xfu) << (4 * i);
    }
  } else {
    for (int i = 0; i < 32; i++) {
      if (x0 + i < v.width && m[i] == 255) f |= 1u << i;
    }
  }
  return f & roi_word_mask(x0, v.rx0, v.rx1);
}

// ---- pass 1: signed distance along the row. One warp per ROI row, one lane per 32-pixel word. ---
// The lane builds its word of the foreground bit-plane straight from the mask bytes (the words of
// the row stay in shared memory for the second sweep; no bit-plane goes through global memory).
// For every ROI pixel: distance to the nearest pixel of the OTHER class in its row (inside the ROI):
// negative on foreground, positive on background, +-kRowInf when the row has none.
constexpr int kRowWarps = 8;
constexpr int kMaxWordsPerRow = kMaxImageDim / 32;

__global__ void __launch_bounds__(kRowWarps * 32) sdf_row_kernel(const __grid_constant__ SdfBatch B) {
  const SdfView& v = B.v[blockIdx.y];
  const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
  if (blockIdx.x == 0 && threadIdx.x == 0) {  // cells the later passes accumulate into
    *v.absmax = 0;
    *v.max_out = -FLT_MAX;
  }
  const int y = v.ry0 + blockIdx.x * kRowWarps + wib;
  if (y > v.ry1) return;
  const int pitch = sdf_pitch(v.width);
  int16_t* __restrict__ out = view_rowd(v) + (size_t)y * pitch;
  const int w0 = v.rx0 >> 5, w1 = v.rx1 >> 5;
  constexpr int kFar = 1 << 20;
  // position of the last foreground / background pixel strictly before each word
  __shared__ int s_last_f[kRowWarps][kMaxWordsPerRow], s_last_b[kRowWarps][kMaxWordsPerRow];
  __shared__ unsigned s_bits[kRowWarps][kMaxWordsPerRow];

  int carry_f = -kFar, carry_b = -kFar;
  for (int wb = w0; wb <= w1; wb += 32) {
    const int w = wb + lane;
    const bool ok = w <= w1;
    const unsigned vm = ok ? roi_word_mask(w * 32, v.rx0, v.rx1) : 0u;
    const unsigned fw = ok ? mask_word(v, y, w) : 0u, bw = ~fw & vm;
    if (ok) s_bits[wib][w - w0] = fw;
    int hf = fw ? w * 32 + 31 - __clz(fw) : -kFar;
    int hb = bw ? w * 32 + 31 - __clz(bw) : -kFar;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {  // inclusive max-scan over the lanes
      const int tf = __shfl_up_sync(0xffffffffu, hf, o), tb = __shfl_up_sync(0xffffffffu, hb, o);
      if (lane >= o) {
        hf = max(hf, tf);
        hb = max(hb, tb);
      }
    }
    int ef = __shfl_up_sync(0xffffffffu, hf, 1), eb = __shfl_up_sync(0xffffffffu, hb, 1);
    if (lane == 0) {
      ef = -kFar;
      eb = -kFar;
    }
    if (ok) {
      s_last_f[wib][w - w0] = max(carry_f, ef);
      s_last_b[wib][w - w0] = max(carry_b, eb);
    }
    carry_f = max(carry_f, __shfl_sync(0xffffffffu, hf, 31));
    carry_b = max(carry_b, __shfl_sync(0xffffffffu, hb, 31));
  }
  __syncwarp();
  carry_f = 2 * kFar;  // now: first foreground / background pixel strictly after the word
  carry_b = 2 * kFar;
  const int n_chunks = (w1 - w0) / 32 + 1;
  for (int c = n_chunks - 1; c >= 0; c--) {
    const int w = w0 + c * 32 + lane;
    const bool ok = w <= w1;
    const unsigned vm = ok ? roi_word_mask(w * 32, v.rx0, v.rx1) : 0u;
    const unsigned fw = ok ? s_bits[wib][w - w0] : 0u, bw = ~fw & vm;
    int lf = fw ? w * 32 + __ffs(fw) - 1 : 2 * kFar;
    int lb = bw ? w * 32 + __ffs(bw) - 1 : 2 * kFar;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {  // inclusive min-scan from the high lanes down
      const int tf = __shfl_down_sync(0xffffffffu, lf, o), tb = __shfl_down_sync(0xffffffffu, lb, o);
      if (lane + o < 32) {
        lf = min(lf, tf);
        lb = min(lb, tb);
      }
    }
    int nf = __shfl_down_sync(0xffffffffu, lf, 1), nb = __shfl_down_sync(0xffffffffu, lb, 1);
    if (lane == 31) {
      nf = 2 * kFar;
      nb = 2 * kFar;
    }
    nf = min(nf, carry_f);
    nb = min(nb, carry_b);
    carry_f = min(carry_f, __shfl_sync(0xffffffffu, lf, 0));
    carry_b = min(carry_b, __shfl_sync(0xffffffffu, lb, 0));
    if (ok) {
      const int pf = s_last_f[wib][w - w0], pb = s_last_b[wib][w - w0];
      const int xb = w * 32;
      uint32_t packed[16];
#pragma unroll
      for (int i = 0; i < 32; i++) {
        const bool fg = (fw >> i) & 1u;
        const unsigned opp = fg ? bw : fw;  // the other class
        const unsigned lo = opp & ((2u << i) - 1u), hi = opp >> i;
        const int x = xb + i;
        const int dl = lo ? i - (31 - __clz(lo)) : x - (fg ? pb : pf);
        const int dr = hi ? __ffs(hi) - 1 : (fg ? nb : nf) - x;
        int d = min(min(dl, dr), (int)kRowInf);
        d = ((vm >> i) & 1u) ? (fg ? -d : d) : 0;
        if (i & 1) {
          packed[i >> 1] |= (uint32_t)(uint16_t)d << 16;
        } else {
          packed[i >> 1] = (uint32_t)(uint16_t)d;
        }
      }
      uint4* dst = reinterpret_cast<uint4*>(out + xb);
#pragma unroll
      for (int q = 0; q < 4; q++) {
        dst[q] = make_uint4(packed[4 * q], packed[4 * q + 1], packed[4 * q + 2], packed[4 * q + 3]);
      }
    }
  }
}

// ---- pass 2: combine along the column. Block = 32 columns x kColSegs row segments. --------------
// D(x, y) = min over y' of |y - y'| + g(x, y'), g = 0 on pixels of the other class, the row distance
// elsewhere: prefix-min of (g - y') and suffix-min of (g + y') per segment, carries through shared
// memory, for the distance-to-background (a) and distance-to-foreground (b) transforms at once.
constexpr int kColInf = 1 << 28;
constexpr int kColSegs = 32;

__global__ void __launch_bounds__(32 * kColSegs) sdf_col_kernel(const __grid_constant__ SdfBatch B) {
  const SdfView& v = B.v[blockIdx.y];
  const int cx = threadIdx.x, seg = threadIdx.y;
  const int x = v.rx0 + blockIdx.x * 32 + cx;
  const bool col_ok = x <= v.rx1;
  const int roi_h = v.ry1 - v.ry0 + 1;
  const int seglen = (roi_h + kColSegs - 1) / kColSegs;
  const int ys = v.ry0 + seg * seglen;
  const int ye = min(ys + seglen, v.ry1 + 1);
  const int pitch = sdf_pitch(v.width);
  const int16_t* __restrict__ rowd = view_rowd(v);
  int16_t* __restrict__ cold = view_cold(v);

  __shared__ int s_pa[kColSegs][33], s_qa[kColSegs][33], s_pb[kColSegs][33], s_qb[kColSegs][33];

  int pa = kColInf, qa = kColInf, pb = kColInf, qb = kColInf;
  if (col_ok) {
#pragma unroll 8
    for (int y = ys; y < ye; y++) {
      const int r = __ldg(rowd + (size_t)y * pitch + x);
      const int ga = r < 0 ? (r == -kRowInf ? kColInf : -r) : 0;
      const int gb = r > 0 ? (r == kRowInf ? kColInf : r) : 0;
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
  int cpa = kColInf, cpb = kColInf, cqa = kColInf, cqb = kColInf;
  for (int s = 0; s < seg; s++) {
    cpa = min(cpa, s_pa[s][cx]);
    cpb = min(cpb, s_pb[s][cx]);
  }
  for (int s = seg + 1; s < kColSegs; s++) {
    cqa = min(cqa, s_qa[s][cx]);
    cqb = min(cqb, s_qb[s][cx]);
  }
  int local_max = 0;
  if (col_ok) {
    int ra = cpa, rb = cpb;
#pragma unroll 8
    for (int y = ys; y < ye; y++) {
      const size_t i = (size_t)y * pitch + x;
      const int r = __ldg(rowd + i);
      const int ga = r < 0 ? (r == -kRowInf ? kColInf : -r) : 0;
      const int gb = r > 0 ? (r == kRowInf ? kColInf : r) : 0;
      ra = min(ra, ga - y);
      rb = min(rb, gb - y);
      const int f = (r < 0 ? ra : rb) + y;
      cold[i] = (int16_t)(f > 32766 ? kRowInf : f);
    }
    ra = cqa;
    rb = cqb;
#pragma unroll 8
    for (int y = ye - 1; y >= ys; y--) {
      const size_t i = (size_t)y * pitch + x;
      const int r = __ldg(rowd + i);
      const int ga = r < 0 ? (r == -kRowInf ? kColInf : -r) : 0;
      const int gb = r > 0 ? (r == kRowInf ? kColInf : r) : 0;
      ra = min(ra, ga + y);
      rb = min(rb, gb + y);
      int d = min((int)cold[i], (r < 0 ? ra : rb) - y);
      if (d > 32766) {
        d = kRowInf;
        local_max = INT_MAX;
      } else {
        local_max = max(local_max, d);
      }
      cold[i] = (int16_t)(r < 0 ? -d : d);
    }
  }
  local_max = __reduce_max_sync(0xffffffffu, local_max);
  if (cx == 0 && local_max > 0) atomicMax(v.absmax, local_max);
}

// Same pass with the block's 32-column strip staged in shared memory (up to 32 x 3168 int16).
// Seeds need no special "infinite" value here: a row distance of kRowInf (32767) exceeds every
// finite L1 distance (<= width + height - 2 <= 32766), so a result >= 32767 simply means "no seed".
// Recurrences: forward f(y) = min(f(y-1) + 1, g(y)), backward b(y) = min(b(y+1) + 1, g(y)), each one
// fused add-min per polarity; carries between the 32 row segments go through shared memory.
constexpr int kColMaxSegLen = 99;  // 32 cols * 32 segs * 99 rows * 2 B = 198 KB (+17 KB of carries)
constexpr int kColFar = 1 << 20;

__global__ void __launch_bounds__(1024) sdf_col_smem_kernel(const __grid_constant__ SdfBatch B) {
  const SdfView& v = B.v[blockIdx.y];
  const int cx = threadIdx.x, seg = threadIdx.y;
  const int x = v.rx0 + blockIdx.x * 32 + cx;
  if (v.rx0 + blockIdx.x * 32 > v.rx1) return;  // whole block outside this view's ROI
  const bool col_ok = x <= v.rx1;
  const int roi_h = v.ry1 - v.ry0 + 1;
  const int seglen = (roi_h + 31) / 32;
  const int ys = v.ry0 + seg * seglen;
  const int rows = max(0, min(seglen, v.ry1 + 1 - ys));
  const int pitch = sdf_pitch(v.width);
  const int16_t* __restrict__ rowd = view_rowd(v) + (size_t)ys * pitch + x;
  int16_t* __restrict__ cold = view_cold(v) + (size_t)ys * pitch + x;

  extern __shared__ int16_t s_tile[];                       // [seg][row][cx]
  // per (segment, column): forward value at the segment's last row / backward value at its first
  // row, each computed from the segment's own seeds only, for both polarities
  __shared__ int s_fa[32][33], s_fb[32][33], s_ba[32][33], s_bb[32][33];
  int16_t* mine = s_tile + (size_t)seg * seglen * 32 + cx;  // stride 32 between rows

  int fa = kColFar, fb = kColFar, ba = kColFar, bb = kColFar;
  if (col_ok) {
#pragma unroll 8
    for (int i = 0; i < rows; i++) {
      const int r = __ldg(rowd + (size_t)i * pitch);
      mine[i * 32] = (int16_t)r;
      const int ga = max(-r, 0), gb = max(r, 0);  // seeds: 0 on the other class, row distance else
      fa = min(fa + 1, ga);
      fb = min(fb + 1, gb);
      ba = min(ba, ga + i);
      bb = min(bb, gb + i);
    }
  }
  s_fa[seg][cx] = fa;
  s_fb[seg][cx] = fb;
  s_ba[seg][cx] = ba;
  s_bb[seg][cx] = bb;
  __syncthreads();
  // carry-in: value of the forward recurrence just above the segment / backward just below it
  int cfa = kColFar, cfb = kColFar, cba = kColFar, cbb = kColFar;
  for (int s = 0; s < seg; s++) {  // rows between segment s's last row and ys - 1
    const int gap = (seg - 1 - s) * seglen;
    cfa = min(cfa, s_fa[s][cx] + gap);
    cfb = min(cfb, s_fb[s][cx] + gap);
  }
  for (int s = seg + 1; s < 32; s++) {  // rows between this segment's last row + 1 and s's first
    const int gap = (s - seg - 1) * seglen + (seglen - rows);
    cba = min(cba, s_ba[s][cx] + gap);
    cbb = min(cbb, s_bb[s][cx] + gap);
  }
  int local_max = 0;
  if (col_ok) {
    // backward sweep: b replaces the seed in the strip, the sign keeps the pixel's class
    ba = cba;
    bb = cbb;
    for (int i = rows - 1; i >= 0; i--) {
      const int r = mine[i * 32];
      ba = min(ba + 1, max(-r, 0));
      bb = min(bb + 1, max(r, 0));
      const int b = min(r < 0 ? ba : bb, (int)kRowInf);
      mine[i * 32] = (int16_t)(r < 0 ? -b : b);
    }
    // forward sweep: seeds again from global memory (L2-hot), min with b, emit
    fa = cfa;
    fb = cfb;
#pragma unroll 8
    for (int i = 0; i < rows; i++) {
      const int r = __ldg(rowd + (size_t)i * pitch);
      const int b = mine[i * 32];
      fa = min(fa + 1, max(-r, 0));
      fb = min(fb + 1, max(r, 0));
      const int d = min(min(r < 0 ? fa : fb, abs(b)), (int)kRowInf);
      local_max = max(local_max, d);
      cold[(size_t)i * pitch] = (int16_t)(r < 0 ? -d : d);
    }
  }
  if (local_max >= kRowInf) local_max = INT_MAX;
  local_max = __reduce_max_sync(0xffffffffu, local_max);
  if (cx == 0 && local_max > 0) atomicMax(v.absmax, local_max);
}

// ---- pass 2, packed: two columns per thread on the 16-bit SIMD integer instructions ------------
// (VIADDMNMX.S16x2 / VIMNMX.S16x2 / VIADD.16x2 of sm_100a).  Used when every distance of the batch
// stays far below 2^15 (launch_sdf2d checks), with kPackInf standing for "no seed yet".
// Two sweeps over the data instead of three, and no strip in shared memory:
//   forward  (per segment, no carry): f(y) = min(f(y-1) + 1, g(y)); the signed result fb - fa (one of
//            the two polarities is 0 on every pixel) goes to the output plane, together with the
//            segment summaries F = f(last row) and Bl = min_i (f(i) + i);
//   carries: forward carry cf from the F of the segments above; the true forward value is
//            min(f(i), cf + i + 1), so the segment's backward summary is B = min(Bl, cf + 1);
//   backward (with both carries): d(y) = min(d(y+1) + 1, min(f(y), cf + y + 1)), written over f.
// A column with no pixel of the other class anywhere in the ROI ends with values >= kPackInf; the
// finalize pass clamps every |d| > width + height to "none", and the maximum is reported as INT_MAX.
constexpr int kPackSegs = 16;
constexpr int kPackInf = 0x3FFF;
constexpr unsigned kPackInf2 = 0x3FFF3FFFu, kPackOne2 = 0x00010001u, kPackMinusOne2 = 0xFFFFFFFFu;

__device__ __forceinline__ unsigned pack_dup(int v) { return (unsigned)(v & 0xFFFF) * 0x00010001u; }

__global__ void __launch_bounds__(32 * kPackSegs, 2)
    sdf_col_pack_kernel(const __grid_constant__ SdfBatch B) {
  const SdfView& v = B.v[blockIdx.y];
  const int lane = threadIdx.x, seg = threadIdx.y;
  const int xb = (v.rx0 & ~1) + blockIdx.x * 64;
  if (xb > v.rx1) return;  // whole block outside this view's ROI
  const int x = xb + 2 * lane;  // columns x, x + 1
  const bool pair_ok = x <= v.rx1;
  const unsigned valid = ((x >= v.rx0 && x <= v.rx1) ? 0x0000FFFFu : 0u) |
                         ((x + 1 >= v.rx0 && x + 1 <= v.rx1) ? 0xFFFF0000u : 0u);
  const int roi_h = v.ry1 - v.ry0 + 1;
  const int seglen = (roi_h + kPackSegs - 1) / kPackSegs;
  const int ys = v.ry0 + seg * seglen;
  const int rows = max(0, min(seglen, v.ry1 + 1 - ys));
  const int pitch2 = sdf_pitch(v.width) / 2;  // row pitch in 32-bit words
  const unsigned* __restrict__ rowd =
      reinterpret_cast<const unsigned*>(view_rowd(v)) + (size_t)ys * pitch2 + (x >> 1);
  unsigned* __restrict__ cold =
      reinterpret_cast<unsigned*>(view_cold(v)) + (size_t)ys * pitch2 + (x >> 1);

  __shared__ unsigned s_fa[kPackSegs][33], s_fb[kPackSegs][33], s_ba[kPackSegs][33],
      s_bb[kPackSegs][33];

  unsigned fa = kPackInf2, fb = kPackInf2, bla = kPackInf2, blb = kPackInf2;
  if (pair_ok) {
    unsigned iv = 0;  // (i, i)
#pragma unroll 8
    for (int i = 0; i < rows; i++) {
      const unsigned r = __ldg(rowd + (size_t)i * pitch2);
      const unsigned ga = __vmaxs2(__vneg2(r), 0u);  // seeds: 0 on the other class, row distance else
      const unsigned gb = __vmaxs2(r, 0u);
      fa = __viaddmin_s16x2(fa, kPackOne2, ga);
      fb = __viaddmin_s16x2(fb, kPackOne2, gb);
      bla = __viaddmin_s16x2(fa, iv, bla);
      blb = __viaddmin_s16x2(fb, iv, blb);
      iv = __vadd2(iv, kPackOne2);
      cold[(size_t)i * pitch2] = __vsub2(fb, fa);
    }
  }
  s_fa[seg][lane] = fa;
  s_fb[seg][lane] = fb;
  __syncthreads();
  // forward carry: the recurrence just above this segment, from the segments above (all full)
  unsigned cfa = kPackInf2, cfb = kPackInf2;
  for (int s = 0; s < seg; s++) {
    const unsigned gap = pack_dup((seg - 1 - s) * seglen);
    cfa = __vmins2(cfa, __vadd2(s_fa[s][lane], gap));
    cfb = __vmins2(cfb, __vadd2(s_fb[s][lane], gap));
  }
  // backward summary of this segment with the carry applied (an empty segment has none)
  s_ba[seg][lane] = rows > 0 ? __vmins2(bla, __vadd2(cfa, kPackOne2)) : kPackInf2;
  s_bb[seg][lane] = rows > 0 ? __vmins2(blb, __vadd2(cfb, kPackOne2)) : kPackInf2;
  __syncthreads();
  unsigned da = kPackInf2, db = kPackInf2;
  for (int s = seg + 1; s < kPackSegs; s++) {  // a segment below exists only if this one is full
    const unsigned gap = pack_dup((s - seg - 1) * seglen);
    da = __vmins2(da, __vadd2(s_ba[s][lane], gap));
    db = __vmins2(db, __vadd2(s_bb[s][lane], gap));
  }
  unsigned mx = 0;
  if (pair_ok) {
    unsigned ca = __vadd2(cfa, pack_dup(rows)), cb = __vadd2(cfb, pack_dup(rows));  // cf + i + 1
#pragma unroll 8
    for (int i = rows - 1; i >= 0; i--) {
      const unsigned f = cold[(size_t)i * pitch2];  // this thread's own forward result
      const unsigned fat = __vmins2(__vmaxs2(__vneg2(f), 0u), ca);
      const unsigned fbt = __vmins2(__vmaxs2(f, 0u), cb);
      da = __viaddmin_s16x2(da, kPackOne2, fat);
      db = __viaddmin_s16x2(db, kPackOne2, fbt);
      ca = __vadd2(ca, kPackMinusOne2);
      cb = __vadd2(cb, kPackMinusOne2);
      mx = __vmaxs2(mx, __vmaxs2(da, db));
      cold[(size_t)i * pitch2] = __vsub2(db, da);
    }
  }
  mx &= valid;
  int local_max = max((int)(mx & 0xFFFFu), (int)(mx >> 16));
  if (local_max >= kPackInf) local_max = INT_MAX;
  local_max = __reduce_max_sync(0xffffffffu, local_max);
  if (lane == 0 && local_max > 0) atomicMax(v.absmax, local_max);
}

// ---- pass 2, packed, forward results kept on chip ----------------------------------------------
// The same two sweeps as sdf_col_pack_kernel, but the block's strip (2 * kLanes columns x all ROI
// rows) of forward results stays in shared memory between the sweeps instead of going out to the
// int16 plane and back: the row distances are read once and the final distances written once.
// kLanes = 32: 64 columns x 16 row segments (ROI heights up to ~1500 rows); kLanes = 16: 32 columns
// x 32 segments (up to ~3000 rows; a warp then covers two segments of 16 lanes, 64-byte accesses).
template <int kLanes>
__global__ void __launch_bounds__(512, 1)
    sdf_col_pack_smem_kernel(const __grid_constant__ SdfBatch B, int seglen_max) {
  constexpr int kSegs = 512 / kLanes;
  const SdfView& v = B.v[blockIdx.y];
  const int lane = threadIdx.x, seg = threadIdx.y;
  const int xb = (v.rx0 & ~1) + blockIdx.x * (2 * kLanes);
  if (xb > v.rx1) return;  // whole block outside this view's ROI
  const int x = xb + 2 * lane;  // columns x, x + 1
  const bool pair_ok = x <= v.rx1;
  const unsigned valid = ((x >= v.rx0 && x <= v.rx1) ? 0x0000FFFFu : 0u) |
                         ((x + 1 >= v.rx0 && x + 1 <= v.rx1) ? 0xFFFF0000u : 0u);
  const int roi_h = v.ry1 - v.ry0 + 1;
  const int seglen = (roi_h + kSegs - 1) / kSegs;
  const int ys = v.ry0 + seg * seglen;
  const int rows = max(0, min(seglen, v.ry1 + 1 - ys));
  const int pitch2 = sdf_pitch(v.width) / 2;  // row pitch in 32-bit words
  const unsigned* __restrict__ rowd =
      reinterpret_cast<const unsigned*>(view_rowd(v)) + (size_t)ys * pitch2 + (x >> 1);
  unsigned* __restrict__ cold =
      reinterpret_cast<unsigned*>(view_cold(v)) + (size_t)ys * pitch2 + (x >> 1);

  extern __shared__ unsigned s_strip[];  // [segment][row][lane]; segments padded apart by 16 banks
  unsigned* mine = s_strip + (size_t)seg * (seglen_max * kLanes + 16) + lane;
  __shared__ unsigned s_fa[kSegs][kLanes + 1], s_fb[kSegs][kLanes + 1], s_ba[kSegs][kLanes + 1],
      s_bb[kSegs][kLanes + 1];

  unsigned fa = kPackInf2, fb = kPackInf2, bla = kPackInf2, blb = kPackInf2;
  if (pair_ok) {
    unsigned iv = 0;  // (i, i)
#pragma unroll 8
    for (int i = 0; i < rows; i++) {
      const unsigned r = __ldg(rowd + (size_t)i * pitch2);
      const unsigned ga = __vmaxs2(__vneg2(r), 0u);  // seeds: 0 on the other class, row distance else
      const unsigned gb = __vmaxs2(r, 0u);
      fa = __viaddmin_s16x2(fa, kPackOne2, ga);
      fb = __viaddmin_s16x2(fb, kPackOne2, gb);
      bla = __viaddmin_s16x2(fa, iv, bla);
      blb = __viaddmin_s16x2(fb, iv, blb);
      iv = __vadd2(iv, kPackOne2);
      mine[i * kLanes] = __vsub2(fb, fa);
    }
  }
  s_fa[seg][lane] = fa;
  s_fb[seg][lane] = fb;
  __syncthreads();
  unsigned cfa = kPackInf2, cfb = kPackInf2;
  for (int s = 0; s < seg; s++) {
    const unsigned gap = pack_dup((seg - 1 - s) * seglen);
    cfa = __vmins2(cfa, __vadd2(s_fa[s][lane], gap));
    cfb = __vmins2(cfb, __vadd2(s_fb[s][lane], gap));
  }
  s_ba[seg][lane] = rows > 0 ? __vmins2(bla, __vadd2(cfa, kPackOne2)) : kPackInf2;
  s_bb[seg][lane] = rows > 0 ? __vmins2(blb, __vadd2(cfb, kPackOne2)) : kPackInf2;
  __syncthreads();
  unsigned da = kPackInf2, db = kPackInf2;
  for (int s = seg + 1; s < kSegs; s++) {
    const unsigned gap = pack_dup((s - seg - 1) * seglen);
    da = __vmins2(da, __vadd2(s_ba[s][lane], gap));
    db = __vmins2(db, __vadd2(s_bb[s][lane], gap));
  }
  unsigned mx = 0;
  if (pair_ok) {
    unsigned ca = __vadd2(cfa, pack_dup(rows)), cb = __vadd2(cfb, pack_dup(rows));  // cf + i + 1
#pragma unroll 8
    for (int i = rows - 1; i >= 0; i--) {
      const unsigned f = mine[i * kLanes];  // this thread's own forward result
      const unsigned fat = __vmins2(__vmaxs2(__vneg2(f), 0u), ca);
      const unsigned fbt = __vmins2(__vmaxs2(f, 0u), cb);
      da = __viaddmin_s16x2(da, kPackOne2, fat);
      db = __viaddmin_s16x2(db, kPackOne2, fbt);
      ca = __vadd2(ca, kPackMinusOne2);
      cb = __vadd2(cb, kPackMinusOne2);
      mx = __vmaxs2(mx, __vmaxs2(da, db));
      cold[(size_t)i * pitch2] = __vsub2(db, da);
    }
  }
  mx &= valid;
  int local_max = max((int)(mx & 0xFFFFu), (int)(mx >> 16));
  if (local_max >= kPackInf) local_max = INT_MAX;
  local_max = __reduce_max_sync(0xffffffffu, local_max);
  if ((threadIdx.y * kLanes + threadIdx.x) % 32 == 0 && local_max > 0) atomicMax(v.absmax, local_max);
}

// ---- pass 2 (Euclidean mode): exact squared distance along the column --------------------------
// No counterpart in the reference (its transform is the L1 one above); offered because hosts may
// want the true Euclidean silhouette distance.  D2(x, y) = min over y' of (y - y')^2 + g(x, y')^2
// with g the row distance (0 on pixels of the other class): the lower envelope of parabolas, built
// with integer arithmetic only (Meijster, Roerdink, Hesselink 2000, phase 2), one thread per column,
// once per polarity.  The stack lives in global memory (coalesced across the columns of a warp).
__global__ void __launch_bounds__(128) sdf_col_l2_kernel(const __grid_constant__ SdfBatch B) {
  const SdfView& v = B.v[blockIdx.y];
  const int x = v.rx0 + blockIdx.x * blockDim.x + threadIdx.x;
  int local_max = 0;
  if (x <= v.rx1) {
    const int pitch = sdf_pitch(v.width);
    const int m = v.ry1 - v.ry0 + 1;
    const int16_t* __restrict__ rowd = view_rowd(v) + (size_t)v.ry0 * pitch + x;
    int32_t* __restrict__ st = view_l2_plane(v, 0) + x;  // (g << 16) | row of each parabola
    int32_t* __restrict__ tt = view_l2_plane(v, 1) + x;  // first row where it is the lowest
    int32_t* __restrict__ out = view_l2_plane(v, 2) + (size_t)v.ry0 * pitch + x;
    for (int pol = 0; pol < 2; pol++) {  // 0: distance to background (for foreground pixels)
      int q = -1;
      for (int u = 0; u < m; u++) {
        const int r = __ldg(rowd + (size_t)u * pitch);
        const int gu = pol == 0 ? max(-r, 0) : max(r, 0);
        if (gu >= kRowInf) continue;  // no pixel of the other class in this row
        while (q >= 0) {
          const int e = st[(size_t)q * pitch], i = e & 0xffff, gi = e >> 16, tq = tt[(size_t)q * pitch];
          if ((tq - i) * (tq - i) + gi * gi > (tq - u) * (tq - u) + gu * gu) {
            q--;
          } else {
            break;
          }
        }
        if (q < 0) {
          q = 0;
          st[0] = (gu << 16) | u;
          tt[0] = 0;
        } else {
          const int e = st[(size_t)q * pitch], i = e & 0xffff, gi = e >> 16;
          const long long num = (long long)u * u - (long long)i * i + (long long)gu * gu -
                                (long long)gi * gi;
          const int w = 1 + (int)(num / (2 * (u - i)));
          if (w < m) {
            q++;
            st[(size_t)q * pitch] = (gu << 16) | u;
            tt[(size_t)q * pitch] = w;
          }
        }
      }
      for (int u = m - 1; u >= 0; u--) {
        const int r = __ldg(rowd + (size_t)u * pitch);
        int d2 = INT_MAX;  // no seed in the whole column range: "infinitely far"
        if (q >= 0) {
          const int e = st[(size_t)q * pitch], i = e & 0xffff, gi = e >> 16;
          d2 = (u - i) * (u - i) + gi * gi;
          if (u == tt[(size_t)q * pitch]) q--;
        }
        if ((pol == 0) == (r < 0)) {  // this pixel belongs to the polarity being resolved
          out[(size_t)u * pitch] = r < 0 ? -d2 : d2;
          local_max = max(local_max, d2);
        }
      }
    }
  }
  local_max = __reduce_max_sync(0xffffffffu, local_max);
  if ((threadIdx.x & 31) == 0 && local_max > 0) atomicMax(v.absmax, local_max);
}

// ---- pass 3: int -> float, normalisation, truncation, image maximum, 8x8-block flags. -----------
// One block per band of 8 image rows, one thread per 8x8-pixel block of the band: 16-byte loads of
// the int16 distances, 16-byte stores of the floats (when the width allows), and - because a thread
// owns a whole 8x8 block - the "all texels 1.0f" / "all texels kVal" flags that the fusion kernel
// uses to classify voxel tiles come for free (one ballot per 32 blocks).
__device__ __forceinline__ float sdf_value(int d, int mode, bool do_norm, float norm, int truncate,
                                           float band) {
  float f;
  if (mode == kSdfSigned || mode == kSdfSignedL2) {
    if (mode == kSdfSigned) {
      f = d == kRowInf ? FLT_MAX : (d == -kRowInf ? -FLT_MAX : (float)d);
    } else {  // d = signed squared distance; the distance is the double square root rounded to float
      const int a = d < 0 ? -d : d;
      const float e = a == INT_MAX ? FLT_MAX : (float)__dsqrt_rn((double)a);
      f = d < 0 ? -e : e;
    }
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
  return f;
}

// normalisation factor of a view, voxel_carver.cc:205-222 (do_norm false when it does not apply)
__device__ __forceinline__ float sdf_norm_of(const SdfView& v, int mode, int normalize,
                                             bool* do_norm) {
  *do_norm = false;
  if (normalize && (mode == kSdfSigned || mode == kSdfSignedL2)) {
    const int am_i = *v.absmax;
    const float am = am_i == INT_MAX
                         ? FLT_MAX
                         : (mode == kSdfSignedL2 ? (float)__dsqrt_rn((double)am_i) : (float)am_i);
    if (am > FLT_MIN) {  // voxel_carver.cc:213
      *do_norm = true;
      return __fdiv_rn(1.0f, am);
    }
  }
  return 1.0f;
}

// L1 modes: the final value of a pixel depends on its integer distance only, so the conversion,
// normalisation and truncation run once per possible distance instead of once per pixel.
// Entry clamp(d, -n, n) + n, n = width + height (no finite distance reaches n); the two ends stand
// for "no pixel of the other class" (+-kRowInf in the exact passes, >= kPackInf in the packed one).
__global__ void __launch_bounds__(256) sdf_lut_kernel(const __grid_constant__ SdfBatch B, int mode,
                                                      int normalize, int truncate, float band) {
  const SdfView& v = B.v[blockIdx.y];
  const int n = v.width + v.height;
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i > 2 * n) return;
  bool do_norm;
  const float norm = sdf_norm_of(v, mode, normalize, &do_norm);
  const int d = i == 0 ? -kRowInf : (i == 2 * n ? (int)kRowInf : i - n);
  view_lut(v)[i] = sdf_value(d, mode, do_norm, norm, truncate, band);
}

__global__ void __launch_bounds__(256) sdf_finalize_kernel(const __grid_constant__ SdfBatch B,
                                                           int mode, int normalize, int truncate,
                                                           float band) {
  const SdfView& v = B.v[blockIdx.y];
  const int y0 = blockIdx.x * kTilePx;
  if (y0 >= v.height) return;
  bool do_norm = false;
  const float norm = mode == kSdfSignedL2 ? sdf_norm_of(v, mode, normalize, &do_norm) : 1.0f;
  const float* __restrict__ lut = view_lut(v);  // L1 modes, see sdf_lut_kernel
  const int lut_n = v.width + v.height;
  const int pitch = sdf_pitch(v.width);
  const int16_t* __restrict__ cold = view_cold(v);
  const int32_t* __restrict__ d2p = mode == kSdfSignedL2 ? view_l2_plane(v, 2) : nullptr;
  const int nbx = (v.width + kTilePx - 1) / kTilePx;
  const int words = (nbx + 31) / 32;
  const bool vec_out = (v.width & 3) == 0;
  const int lane = threadIdx.x & 31;
  float tmax = -FLT_MAX;
  // the loop bound is rounded up to whole warps so that every lane reaches the ballots
  for (int bx0 = 0; bx0 < nbx; bx0 += blockDim.x) {
    const int bx = bx0 + threadIdx.x;
    const int x0 = bx * kTilePx;
    bool ones = bx < nbx, kval = bx < nbx;
    if (bx < nbx) {
      for (int j = 0; j < kTilePx; j++) {
        const int y = y0 + j;
        if (y >= v.height) break;
        const bool row_in = y >= v.ry0 && y <= v.ry1;
        float f[kTilePx];
        // the int16 scratch rows are padded to 32 pixels: the 16-byte load is always in bounds
        int dd[kTilePx];
        if (mode == kSdfSignedL2) {
          const int4* __restrict__ p = reinterpret_cast<const int4*>(d2p + (size_t)y * pitch + x0);
          const int4 lo = row_in ? __ldg(p) : make_int4(0, 0, 0, 0);
          const int4 hi = row_in ? __ldg(p + 1) : make_int4(0, 0, 0, 0);
          dd[0] = lo.x; dd[1] = lo.y; dd[2] = lo.z; dd[3] = lo.w;
          dd[4] = hi.x; dd[5] = hi.y; dd[6] = hi.z; dd[7] = hi.w;
        } else {
          const uint4 raw = row_in
                                ? __ldg(reinterpret_cast<const uint4*>(cold + (size_t)y * pitch + x0))
                                : make_uint4(0, 0, 0, 0);
          const unsigned rw[4] = {raw.x, raw.y, raw.z, raw.w};
#pragma unroll
          for (int i = 0; i < kTilePx; i++) {
            dd[i] = (int)(int16_t)((rw[i >> 1] >> (16 * (i & 1))) & 0xffffu);
          }
        }
#pragma unroll
        for (int i = 0; i < kTilePx; i++) {
          const int x = x0 + i;
          const int d = dd[i];
          // zeros outside the ROI, voxel_carver.cc:104
          const bool in_roi = row_in && x >= v.rx0 && x <= v.rx1;
          if (mode == kSdfSignedL2) {
            f[i] = in_roi ? sdf_value(d, mode, do_norm, norm, truncate, band) : 0.0f;
          } else {
            f[i] = in_roi ? __ldg(lut + (min(max(d, -lut_n), lut_n) + lut_n)) : 0.0f;
          }
          if (x < v.width) {
            tmax = fmaxf(tmax, f[i]);
            ones = ones && (f[i] == 1.0f);
            kval = kval && (f[i] == kInvalidSdf);
          }
        }
        float* __restrict__ o = v.out + (size_t)y * v.width + x0;
        if (vec_out && x0 + kTilePx <= v.width) {
          reinterpret_cast<float4*>(o)[0] = make_float4(f[0], f[1], f[2], f[3]);
          reinterpret_cast<float4*>(o)[1] = make_float4(f[4], f[5], f[6], f[7]);
        } else {
#pragma unroll
          for (int i = 0; i < kTilePx; i++) {
            if (x0 + i < v.width) o[i] = f[i];
          }
        }
      }
    }
    const unsigned ow = __ballot_sync(0xffffffffu, ones), kw = __ballot_sync(0xffffffffu, kval);
    if (v.tile_flags != nullptr && lane == 0 && bx < nbx) {
      const int nby = (v.height + kTilePx - 1) / kTilePx;
      v.tile_flags[blockIdx.x * words + (bx >> 5)] = ow;
      v.tile_flags[(nby + blockIdx.x) * words + (bx >> 5)] = kw;
    }
  }
  for (int o = 16; o > 0; o >>= 1) tmax = fmaxf(tmax, __shfl_xor_sync(0xffffffffu, tmax, o));
  if (lane == 0) atomic_max_float(v.max_out, tmax);
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

// voxel_carver.cc:246-265, one pixel per thread; the uint8 conversions truncate like static_cast
__global__ void __launch_bounds__(256) sdf_color_kernel(const float* __restrict__ sdf, int64_t n,
                                                        float min_neg, float max_pos,
                                                        uint8_t* __restrict__ rgb) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float d = sdf[i];
  uint8_t r, g, b;
  if (d > 0.0f) {
    float t = __fdiv_rn(__fsub_rn(max_pos, d), max_pos);
    t = t < 0.0f ? 0.0f : t;  // std::max(t, 0.0f)
    t = 1.0f < t ? 1.0f : t;  // std::min(., 1.0f)
    const uint8_t c = (uint8_t)__fmul_rn(255.0f, t);
    r = 255;
    g = c;
    b = c;
  } else {
    float t = __fdiv_rn(__fsub_rn(d, min_neg), -min_neg);
    t = t < 0.0f ? 0.0f : t;
    t = 1.0f < t ? 1.0f : t;
    const uint8_t c = (uint8_t)__fmul_rn(255.0f, t);
    r = c;
    g = c;
    b = 255;
  }
  rgb[3 * i + 0] = r;
  rgb[3 * i + 1] = g;
  rgb[3 * i + 2] = b;
}

}  // namespace

void launch_sdf2d(const SdfView* views_host, int n_views, int mode, int normalize, int truncate,
                  float band, cudaStream_t stream, int64_t* launches) {
  int max_rows = 0, max_cols = 0, max_h = 0, max_words = 0, max_dist_bound = 0, max_lut = 0;
  SdfBatch views_dev;
  for (int i = 0; i < n_views; i++) {
    const SdfView& v = views_host[i];
    views_dev.v[i] = v;
    max_rows = max(max_rows, v.ry1 - v.ry0 + 1);
    max_cols = max(max_cols, v.rx1 - v.rx0 + 1);
    max_h = max(max_h, v.height);
    max_words = max(max_words, (v.ry1 - v.ry0 + 1) * (sdf_pitch(v.width) / 32));
    max_dist_bound = max(max_dist_bound, v.width + v.height);
    max_lut = max(max_lut, sdf_lut_entries(v.width, v.height));
  }
  sdf_row_kernel<<<dim3((max_rows + kRowWarps - 1) / kRowWarps, n_views), kRowWarps * 32, 0,
                   stream>>>(views_dev);
  const int max_seglen = (max_rows + 31) / 32;
  if (mode == kSdfSignedL2) {
    sdf_col_l2_kernel<<<dim3((max_cols + 127) / 128, n_views), 128, 0, stream>>>(views_dev);
  } else if (max_dist_bound <= 16000) {
    // forward results on chip when the strip fits one block's shared memory (64 columns, else 32)
    constexpr size_t kStripBudget = 200 * 1024;
    const int seglen16 = (max_rows + 15) / 16, seglen32 = (max_rows + 31) / 32;
    const size_t strip64 = (size_t)16 * (seglen16 * 32 + 16) * sizeof(unsigned);
    const size_t strip32 = (size_t)32 * (seglen32 * 16 + 16) * sizeof(unsigned);
    static bool strip_attr[64] = {false};
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev >= 0 && dev < 64 && !strip_attr[dev]) {
      cudaFuncSetAttribute(sdf_col_pack_smem_kernel<32>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                           (int)kStripBudget);
      cudaFuncSetAttribute(sdf_col_pack_smem_kernel<16>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                           (int)kStripBudget);
      strip_attr[dev] = true;
    }
    if (strip64 <= kStripBudget) {
      sdf_col_pack_smem_kernel<32><<<dim3((max_cols + 1 + 63) / 64, n_views), dim3(32, 16), strip64,
                                     stream>>>(views_dev, seglen16);
    } else if (strip32 <= kStripBudget) {
      sdf_col_pack_smem_kernel<16><<<dim3((max_cols + 1 + 31) / 32, n_views), dim3(16, 32), strip32,
                                     stream>>>(views_dev, seglen32);
    } else {
      sdf_col_pack_kernel<<<dim3((max_cols + 1 + 63) / 64, n_views), dim3(32, kPackSegs), 0,
                            stream>>>(views_dev);
    }
  } else if (max_seglen <= kColMaxSegLen) {
    const size_t smem = (size_t)32 * 32 * max_seglen * sizeof(int16_t);
    static bool attr_set[64] = {false};
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev >= 0 && dev < 64 && !attr_set[dev]) {
      cudaFuncSetAttribute(sdf_col_smem_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                           32 * 32 * kColMaxSegLen * (int)sizeof(int16_t));
      attr_set[dev] = true;
    }
    sdf_col_smem_kernel<<<dim3((max_cols + 31) / 32, n_views), dim3(32, 32), smem, stream>>>(views_dev);
  } else {
    sdf_col_kernel<<<dim3((max_cols + 31) / 32, n_views), dim3(32, kColSegs), 0, stream>>>(views_dev);
  }
  if (mode != kSdfSignedL2) {
    sdf_lut_kernel<<<dim3((max_lut + 255) / 256, n_views), 256, 0, stream>>>(views_dev, mode,
                                                                              normalize, truncate, band);
    if (launches) *launches += 1;
  }
  sdf_finalize_kernel<<<dim3((max_h + kTilePx - 1) / kTilePx, n_views), 256, 0, stream>>>(
      views_dev, mode, normalize, truncate, band);
  if (launches) *launches += 3;
}

void launch_sdf_color(const float* sdf, int64_t n, float min_neg, float max_pos, uint8_t* rgb,
                      cudaStream_t stream) {
  sdf_color_kernel<<<(unsigned)((n + 255) / 256), 256, 0, stream>>>(sdf, n, min_neg, max_pos, rgb);
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
// The kernels above need 4 bytes of scratch per pixel (two int16 images), one bit per pixel (the
// foreground bit-plane), a value table (one float per possible distance) and two cells (the
// largest |distance| and the image maximum).  Here all of it is this fixed array, part of the library image: every GPU
// that loads libvacancy_b200.so carries it (16.75 MiB) until the process exits, and images of more
// than kNoAllocMaxPixels pixels (width rounded up to a multiple of 32) are refused.  All calls
// share the one array.
namespace vcb {
namespace {

constexpr int64_t kNoAllocMaxPixels = 2048LL * 2048LL;  // counted with the width rounded up to 32
constexpr size_t kNoAllocHeader = 256;  // abs-max and image-max cells
// two int16 images + one bit per pixel + the value table (at most 2 * (16384 + 16384) + 1 floats)
// + their 256-byte alignment slack
constexpr size_t kNoAllocBytes = kNoAllocHeader + 2 * sizeof(int16_t) * kNoAllocMaxPixels +
                                 kNoAllocMaxPixels / 8 + sizeof(float) * (4 * kMaxImageDim + 1) +
                                 4 * 256;

static __device__ __align__(256) unsigned char g_noalloc_scratch[kNoAllocBytes];

// In host code `g_noalloc_scratch` is only a handle for the runtime, not a device address: the
// real address is asked for once per device (with that device current) and cached here.
std::mutex g_noalloc_mutex;
unsigned char* g_noalloc_ptr[64] = {nullptr};

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
      (int64_t)sdf_pitch(width) * height > kNoAllocMaxPixels) {
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
  DeviceGuard restore_callers_device;  // the caller's current device is put back on return
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
  v.scratch = scratch + kNoAllocHeader;
  v.tile_flags = nullptr;
  v.absmax = reinterpret_cast<int*>(scratch + 128);      // both cells are re-initialised by
  v.max_out = reinterpret_cast<float*>(scratch + 192);   // sdf_bits_kernel at the start of a call
  v.width = width;
  v.height = height;
  v.rx0 = roi_min[0];
  v.ry0 = roi_min[1];
  v.rx1 = roi_max[0];
  v.ry1 = roi_max[1];
  launch_sdf2d(&v, 1, kSdfSigned, minmax_normalize, use_truncation, truncation_band, stream,
               nullptr);
  e = cudaGetLastError();
  if (e != cudaSuccess) {
    snprintf(msg, sizeof(msg), "vcb_make_signed_distance_field_noalloc: %s", cudaGetErrorString(e));
    set_global_error(msg);
    return VCB_ERR_CUDA;
  }
  return VCB_OK;
}
