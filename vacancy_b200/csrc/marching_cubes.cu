// K3 - marching cubes on the GPU with the reference's exact vertex numbering and face order.
//
// Replaces vacancy::MarchingCubes (marching_cubes.cc:63-228) and VertexInterp (:25-57).
// The reference walks cubes in raster order (z, y, x; x fastest), each cube indexed by its MAX
// corner voxel, and welds vertices through a std::map keyed by the edge's two voxel ids: a vertex is
// numbered the first time any cube references its edge.  That sequential rule is restated
// order-independently:
//   * cube c is "non-skipped" iff update_num(c) >= 1 and none of its 8 corners holds kVal (:88-112);
//   * the OWNER of a grid edge is the first non-skipped cube, in raster order, among the <= 4 cubes
//     sharing it; every non-skipped cube that contains a crossed edge references it;
//   * inside the owner the vertex takes the rank of its first reference in the triangle-table walk
//     (i ascending, j = 0..2 reading tri[i + 2 - j], :199-218) among the edges the cube owns, and
//     its position is VertexInterp evaluated with the owner's local edge direction;
//   * vertex id = (exclusive prefix sum over raster order of owned-edge counts) + rank;
//     face id   = (exclusive prefix sum of triangle counts) + position in the cube's list.
//
// Passes (S = segment of 32 cubes along x):
//   mc_bits_kernel   dense, HBM-bound: reads each sdf value once (4 B/voxel), writes 2 bits/voxel:
//                    I = (double)sdf < iso  (:121-128)   V = sdf != kVal (:103-112)
//   mc_cube_kernel   bitmap-only (1/16 of the data): per S, word-parallel "all 8 corners valid" (C)
//                    and "8 inside-bits not all equal" -> active-cube word A; S with A != 0 get a slot
//                    in the compact lists, every owned S its entry in the dense S -> slot map.
//   mc_list_kernel   thread per slot: the slot's cubes get consecutive places in the cube list.
//   mc_count_kernel  thread per ACTIVE CUBE: cube index, edge ownership -> counts per S and per row.
//   scan             exclusive prefix over the rows in raster order (extract_voxel.cu), then
//   mc_rowprefix_kernel  warp per row: ids of each slot's first vertex / face.
//   mc_vertex_kernel thread per active cube: owned vertices -> positions (float64 interpolation,
//                    :28-46), written at their final index; ids of the vertices the cube is the home
//                    of stay in cube_ids (the rare others go into a hash table).
//   mc_face_kernel   thread per active cube: triangles in table order, ids read from the home cubes.
// Algorithmic bytes: 4 B/voxel + 12 B/vertex + 12 B/face.
#include <float.h>
#include <math.h>

#include "common.cuh"
#include "mc_tables_packed.h"

namespace vcb {

namespace {

__device__ uint64_t d_tri[256];       // packed triangle table, nibble i = edge of tri[i], 0xF = end
__device__ uint16_t d_edge_mask[256]; // Bourke edgeTable, derived: edge flagged iff corners differ
__device__ uint8_t d_ntri[256];
// distinct edges of a configuration in the order the triangle walk first references them (i
// ascending, j = 0..2 reading tri[i + 2 - j], marching_cubes.cc:199-218), one nibble each, 0xF ends:
// the k-th OWNED edge of this list is the cube's k-th vertex
__device__ uint64_t d_order[256];

// corner k -> offset from the cube's max corner (marching_cubes.cc:93-101)
__constant__ signed char c_corner[8][3] = {{-1, -1, -1}, {0, -1, -1}, {0, 0, -1}, {-1, 0, -1},
                                           {-1, -1, 0},  {0, -1, 0},  {0, 0, 0},  {-1, 0, 0}};
// edge e -> (interpolate from corner, to corner) (marching_cubes.cc:138-197)
__constant__ unsigned char c_edge_corner[12][2] = {{0, 1}, {1, 2}, {2, 3}, {3, 0}, {4, 5}, {5, 6},
                                                   {6, 7}, {7, 4}, {0, 4}, {1, 5}, {2, 6}, {3, 7}};

struct McGeom {
  int nx, ny;
  int spr;        // segments (words) per row
  int zb;         // first stored plane (z_begin - halo)
  int z_begin, z_end;
  int cz0;        // first owned cube layer = max(1, z_begin)
  int64_t n_words;   // spr * ny * stored planes
  int64_t n_segs;    // spr * ny * owned planes
};

__host__ __device__ inline McGeom mc_geom(const GridDesc& g) {
  McGeom m;
  m.nx = g.nx;
  m.ny = g.ny;
  m.spr = (g.nx + 31) / 32;
  m.zb = g.z_begin - g.halo;
  m.z_begin = g.z_begin;
  m.z_end = g.z_end;
  m.cz0 = g.z_begin > 1 ? g.z_begin : 1;
  m.n_words = (int64_t)m.spr * g.ny * (g.z_end - m.zb);
  m.n_segs = (int64_t)m.spr * g.ny * (g.z_end - g.z_begin);
  return m;
}

// word index in the stored-plane bitmaps / owned-plane segment arrays
__device__ __forceinline__ int64_t word_index(const McGeom& m, int sx, int y, int z) {
  return ((int64_t)(z - m.zb) * m.ny + y) * m.spr + sx;
}
__device__ __forceinline__ int64_t seg_index(const McGeom& m, int sx, int y, int z) {
  return ((int64_t)(z - m.z_begin) * m.ny + y) * m.spr + sx;
}
__device__ __forceinline__ unsigned get_bit(const uint32_t* __restrict__ bits, const McGeom& m,
                                            int x, int y, int z) {
  return (bits[word_index(m, x >> 5, y, z)] >> (x & 31)) & 1u;
}

// ---- dense pass ---------------------------------------------------------------------------------
// One block per kBitsRows grid rows; a warp iteration covers 128 voxels of each row with four
// coalesced 128-byte loads per row (16 independent loads in flight per thread), turns every 32
// voxels into one word of each bitmap with a ballot, and lanes 0-3 store the four words together.
// ~0.25 instructions per voxel: the pass is bound by HBM, not by issue.
constexpr int kBitsRows = 4;

__global__ void __launch_bounds__(256) mc_bits_kernel(const float* __restrict__ sdf_stored,
                                                      int nx, int spr, int64_t n_rows, float thr,
                                                      uint32_t* __restrict__ ibits,
                                                      uint32_t* __restrict__ vbits) {
  const int lane = threadIdx.x & 31;
  const int warps = blockDim.x >> 5;
  const int64_t row0 = (int64_t)blockIdx.x * kBitsRows;
  const int rows = (int)min((int64_t)kBitsRows, n_rows - row0);
  const float* __restrict__ src = sdf_stored + row0 * nx;
  uint32_t* __restrict__ idst = ibits + row0 * spr;
  uint32_t* __restrict__ vdst = vbits + row0 * spr;
  const int groups = (spr + 3) / 4;
  for (int c = threadIdx.x >> 5; c < groups; c += warps) {
    const int x = c * 128 + lane;
    float s[kBitsRows][4];
#pragma unroll
    for (int r = 0; r < kBitsRows; r++) {
#pragma unroll
      for (int q = 0; q < 4; q++) {
        s[r][q] = (r < rows && x + 32 * q < nx) ? __ldg(src + r * nx + x + 32 * q) : kInvalidSdf;
      }
    }
#pragma unroll
    for (int r = 0; r < kBitsRows; r++) {
      unsigned iw = 0, vw = 0;
#pragma unroll
      for (int q = 0; q < 4; q++) {
        const bool valid = s[r][q] != kInvalidSdf;  // positions beyond nx hold kVal
        const unsigned i_q = __ballot_sync(0xffffffffu, valid && (s[r][q] < thr));
        const unsigned v_q = __ballot_sync(0xffffffffu, valid);
        if (lane == q) {
          iw = i_q;
          vw = v_q;
        }
      }
      const int w = c * 4 + lane;
      if (lane < 4 && r < rows && w < spr) {
        idst[r * spr + w] = iw;
        vdst[r * spr + w] = vw;
      }
    }
  }
}

__global__ void __launch_bounds__(256) mc_ubits_kernel(const int32_t* __restrict__ un, int nx,
                                                       int spr, int64_t n_rows,
                                                       uint32_t* __restrict__ ubits) {
  const int lane = threadIdx.x & 31;
  const int warps = blockDim.x >> 5;
  const int64_t row = blockIdx.x;
  for (int sx = threadIdx.x >> 5; sx < spr; sx += warps) {
    const int x = sx * 32 + lane;
    const bool u = (x < nx) && (un[row * nx + x] >= 1);
    const unsigned w = __ballot_sync(0xffffffffu, u);
    if (lane == 0) ubits[row * spr + sx] = w;
  }
}

constexpr unsigned kNoSlot = 0xffffffffu;  // segwid entry of a segment without active cubes

// ---- bitmap pass --------------------------------------------------------------------------------
// pair word: bit i = b[x] & b[x-1]   /  any word: bit i = b[x] | b[x-1]
__device__ __forceinline__ unsigned shift_in(unsigned w, unsigned prev) {
  return (w << 1) | (prev >> 31);
}

// Per stored word: the cube-valid bits; per OWNED segment with at least one active cube (valid, corners
// not all on one side) a slot in the compact lists - its segment index and its word of active cubes -
// plus the dense map segment -> slot that the emit passes use to find a neighbour's data.
__global__ void __launch_bounds__(256)
    mc_cube_kernel(const McGeom m, const uint32_t* __restrict__ ibits,
                   const uint32_t* __restrict__ vbits, const uint32_t* __restrict__ ubits,
                   uint32_t* __restrict__ cbits, uint32_t* __restrict__ slot_seg,
                   uint32_t* __restrict__ slot_word, unsigned* __restrict__ counters,
                   uint32_t* __restrict__ segwid) {
  const int64_t w = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const bool in_range = w < m.n_words;
  const int sx = (int)(w % m.spr);
  const int64_t row = w / m.spr;
  const int y = (int)(row % m.ny);
  const int z = (int)(row / m.ny) + m.zb;
  unsigned c = 0, a = 0;
  if (in_range && y >= 1 && z >= 1 && z > m.zb) {
    unsigned all_v = 0xffffffffu, all_i = 0xffffffffu, any_i = 0;
#pragma unroll
    for (int r = 0; r < 4; r++) {
      const int64_t wi = word_index(m, sx, y - (r & 1), z - (r >> 1));
      const unsigned v = vbits[wi], i = ibits[wi];
      const unsigned vp = sx > 0 ? vbits[wi - 1] : 0u, ip = sx > 0 ? ibits[wi - 1] : 0u;
      all_v &= v & shift_in(v, vp);
      all_i &= i & shift_in(i, ip);
      any_i |= i | shift_in(i, ip);
    }
    c = all_v;  // x == 0 is excluded automatically: bit -1 shifts in as 0
    if (ubits) c &= ubits[word_index(m, sx, y, z)];  // marching_cubes.cc:88-90
    a = c & any_i & ~all_i;
  }
  if (in_range) cbits[w] = c;
  // Slots are handed out per BLOCK: one 64-bit atomic (slots in the low half, active cubes in the
  // high half) for all its active segments - one atomic per segment or per warp serialises on the
  // counter's address (measured: 0.65 ms of this pass at 1024^3).
  const bool owned = in_range && z >= m.z_begin;
  if (!owned) a = 0;
  const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const unsigned act = __ballot_sync(0xffffffffu, a != 0);
  const unsigned cubes = __reduce_add_sync(0xffffffffu, (unsigned)__popc(a));
  __shared__ unsigned s_slots[8], s_cubes[8];
  __shared__ unsigned s_base;
  if (lane == 0) {
    s_slots[warp] = __popc(act);
    s_cubes[warp] = cubes;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    unsigned ns = 0, nc = 0;
    for (int i = 0; i < 8; i++) {
      ns += s_slots[i];
      nc += s_cubes[i];
    }
    s_base = ns ? (unsigned)atomicAdd(reinterpret_cast<unsigned long long*>(counters),
                                      ((unsigned long long)nc << 32) | ns)
                : 0u;
  }
  __syncthreads();
  unsigned base = s_base;
  for (unsigned i = 0; i < warp; i++) base += s_slots[i];
  if (owned) {
    const int64_t seg = seg_index(m, sx, y, z);
    unsigned slot = kNoSlot;
    if (a) {
      slot = base + __popc(act & ((1u << lane) - 1u));
      slot_seg[slot] = (uint32_t)seg;
      slot_word[slot] = a;
    }
    segwid[seg] = slot;
  }
}

// ---- per-cube analysis --------------------------------------------------------------------------

struct CubeInfo {
  int ci;          // cube index, marching_cubes.cc:121-128
  unsigned owned;  // crossed edges whose vertex this cube creates
};

// Configuration and owned edges of the cube of lane `lane` in segment (sx, y, z).  Every lane of
// the warp reads the same few bitmap words (broadcast loads): the rows (y-1, y) x (z-1, z) of the
// inside bits for the 8 corners, and the rows of the cube-valid bits that hold the <= 9 earlier
// cubes sharing an edge with this one; bit `lane` of a word shifted by one position is the
// neighbour at x - 1 / x + 1.
__device__ __forceinline__ unsigned shift_out(unsigned w, unsigned next) {
  return (w >> 1) | (next << 31);
}

__device__ __forceinline__ CubeInfo cube_info(const McGeom& m, const uint32_t* __restrict__ ibits,
                                              const uint32_t* __restrict__ cbits, int sx, int y,
                                              int z, int lane) {
  const int64_t w11 = word_index(m, sx, y, z);            // row (y, z)
  const int64_t w01 = w11 - m.spr;                        // row (y - 1, z)
  const int64_t w10 = w11 - (int64_t)m.spr * m.ny;        // row (y, z - 1)
  const int64_t w00 = w10 - m.spr;                        // row (y - 1, z - 1)
  const bool has_prev = sx > 0, has_next = sx + 1 < m.spr;
  // inside bits: corner k of the cube at bit `lane`
  const unsigned i00 = ibits[w00], i01 = ibits[w01], i10 = ibits[w10], i11 = ibits[w11];
  const unsigned p00 = has_prev ? ibits[w00 - 1] : 0u, p01 = has_prev ? ibits[w01 - 1] : 0u;
  const unsigned p10 = has_prev ? ibits[w10 - 1] : 0u, p11 = has_prev ? ibits[w11 - 1] : 0u;
  const unsigned corner[8] = {shift_in(i00, p00), i00, i10, shift_in(i10, p10),
                              shift_in(i01, p01), i01, i11, shift_in(i11, p11)};
  int ci = 0;
#pragma unroll
  for (int k = 0; k < 8; k++) ci |= (int)((corner[k] >> lane) & 1u) << k;
  CubeInfo info;
  info.ci = ci;
  // earlier cubes (raster order) that share an edge with this one; rows that do not exist and
  // cubes outside the grid read as 0 (mc_cube_kernel leaves their bits clear)
  const unsigned c00 = cbits[w00], c10 = cbits[w10], c01 = cbits[w01], c11 = cbits[w11];
  const unsigned c20 = y + 1 < m.ny ? cbits[w10 + m.spr] : 0u;  // row (y + 1, z - 1)
  const unsigned c10p = has_prev ? cbits[w10 - 1] : 0u, c10n = has_next ? cbits[w10 + 1] : 0u;
  const unsigned c01p = has_prev ? cbits[w01 - 1] : 0u, c01n = has_next ? cbits[w01 + 1] : 0u;
  const unsigned c11p = has_prev ? cbits[w11 - 1] : 0u;
  const unsigned n1 = (c00 >> lane) & 1u, n2 = (c10 >> lane) & 1u, n3 = (c20 >> lane) & 1u;
  const unsigned n4 = (shift_in(c10, c10p) >> lane) & 1u, n5 = (shift_out(c10, c10n) >> lane) & 1u;
  const unsigned n6 = (c01 >> lane) & 1u, n7 = (shift_in(c11, c11p) >> lane) & 1u;
  const unsigned n8 = (shift_in(c01, c01p) >> lane) & 1u, n9 = (shift_out(c01, c01n) >> lane) & 1u;
  unsigned taken = 0;  // edges already created by an earlier cube
  taken |= (n1 | n2 | n6) << 0;
  taken |= (n2 | n5) << 1;
  taken |= (n2 | n3) << 2;
  taken |= (n4 | n2 | n7) << 3;
  taken |= (n6) << 4;
  taken |= (n7) << 7;
  taken |= (n8 | n6 | n7) << 8;
  taken |= (n6 | n9) << 9;
  taken |= (n7) << 11;
  info.owned = d_edge_mask[ci] & ~taken;
  return info;
}

// canonical edge key: (lower voxel id, axis)
__device__ __forceinline__ uint64_t edge_key(const McGeom& m, int x, int y, int z, int e) {
  const int a = c_edge_corner[e][0], b = c_edge_corner[e][1];
  const int xa = x + c_corner[a][0], ya = y + c_corner[a][1], za = z + c_corner[a][2];
  const int xb = x + c_corner[b][0], yb = y + c_corner[b][1], zb = z + c_corner[b][2];
  const int lx = min(xa, xb), ly = min(ya, yb), lz = min(za, zb);
  const int axis = (xa != xb) ? 0 : ((ya != yb) ? 1 : 2);
  const uint64_t id = ((uint64_t)lz * m.ny + ly) * m.nx + lx;
  return (id << 2) | (uint64_t)axis;
}

constexpr uint64_t kEmptyKey = 0xffffffffffffffffull;

__device__ __forceinline__ uint64_t hash_slot(uint64_t key, uint64_t mask) {
  key *= 0x9E3779B97F4A7C15ull;
  return (key >> 20) & mask;
}

__device__ __forceinline__ void hash_insert(uint64_t* keys, int32_t* vals, uint64_t mask,
                                            uint64_t key, int32_t val) {
  uint64_t slot = hash_slot(key, mask);
  while (true) {
    const unsigned long long prev =
        atomicCAS(reinterpret_cast<unsigned long long*>(keys + slot), kEmptyKey, key);
    if (prev == kEmptyKey || prev == key) {
      vals[slot] = val;
      return;
    }
    slot = (slot + 1) & mask;
  }
}

__device__ __forceinline__ int32_t hash_find(const uint64_t* __restrict__ keys,
                                             const int32_t* __restrict__ vals, uint64_t mask,
                                             uint64_t key) {
  uint64_t slot = hash_slot(key, mask);
  while (true) {
    const uint64_t k = keys[slot];
    if (k == key) return vals[slot];
    if (k == kEmptyKey) return -1;
    slot = (slot + 1) & mask;
  }
}

// ---- compact lists of the sparse passes ---------------------------------------------------------
// slot  = an owned segment (32 cubes along x) with at least one active cube, in no particular order:
//         slot_seg[slot], slot_word[slot] (its active cubes), slot_cube[slot] (index of its first cube
//         in the cube list), slot_cnt[slot] ((vertices << 16) | faces of the segment), slot_pre[slot]
//         (ids of the segment's first vertex / face);
// cube  = one active cube; the cubes of a slot are consecutive, in x order: cube_slot[c], cube_info[c]
//         (configuration | owned edges << 8), cube_ids[c][3] (below).
// The sparse passes run ONE THREAD PER ACTIVE CUBE: on a surface that crosses a segment once or
// twice a warp per segment would leave 30 lanes idle.
//
// Where a vertex id is kept.  Among the <= 4 cubes sharing a grid edge the first in raster order
// is the one whose max corner is the edge's upper endpoint - the edge's HOME cube, which sees it as
// e6 (x-edge), e5 (y-edge) or e10 (z-edge).  When the home cube is non-skipped it owns the vertex;
// the vertex pass leaves, for every active cube, the ids of the (up to three) vertices it is the
// home of: cube_ids[c][axis], -1 where it owns none.  Only vertices whose home cube is skipped or
// does not exist (next to kVal / un-updated voxels, on the grid's low faces, or below this z slab)
// go through the hash table - in the interior of a carved grid none.
struct McLists {
  const uint32_t* segwid;     // owned segment -> slot (kNoSlot: no active cube)
  const uint32_t* slot_seg;
  const uint32_t* slot_word;
  const uint32_t* slot_cube;
  const uint32_t* cube_slot;
  const uint32_t* cube_info;
  const int32_t* cube_ids;
  const uint64_t* hkeys;
  const int32_t* hvals;
  uint64_t hmask;
};

// id of the vertex on the edge with upper endpoint (hx, hy, hz) along `axis` (-1: none)
__device__ __forceinline__ int32_t vertex_id_of_edge(const McGeom& m, const McLists& t, int hx,
                                                     int hy, int hz, int axis) {
  if (hx >= 1 && hy >= 1 && hz >= m.cz0) {  // a cube of this slab
    const unsigned slot = t.segwid[seg_index(m, hx >> 5, hy, hz)];
    if (slot != kNoSlot) {
      const unsigned word = t.slot_word[slot], bit = 1u << (hx & 31);
      if (word & bit) {
        const unsigned c = t.slot_cube[slot] + __popc(word & (bit - 1u));
        const int32_t id = t.cube_ids[(size_t)c * 3 + axis];
        if (id >= 0) return id;
      }
    }
  }
  const uint64_t id = ((uint64_t)(hz - (axis == 2)) * m.ny + (hy - (axis == 1))) * m.nx +
                      (hx - (axis == 0));
  return hash_find(t.hkeys, t.hvals, t.hmask, (id << 2) | (uint64_t)axis);
}

// the active cube c: its slot, its position in the slot, its coordinates
struct CubeAt {
  unsigned slot, k;
  int x, y, z;
  int64_t row;  // owned row (y, z) index
};
__device__ __forceinline__ CubeAt cube_at(const McGeom& m, const McLists& t, unsigned c) {
  CubeAt q;
  q.slot = t.cube_slot[c];
  q.k = c - t.slot_cube[q.slot];
  const int64_t seg = t.slot_seg[q.slot];
  const int lane = (int)__fns(t.slot_word[q.slot], 0, (int)q.k + 1);  // the k-th active cube
  q.row = seg / m.spr;
  q.x = (int)(seg - q.row * m.spr) * 32 + lane;
  q.y = (int)(q.row % m.ny);
  q.z = (int)(q.row / m.ny) + m.z_begin;
  return q;
}

// One thread per slot: where the slot's cubes sit in the cube list.
__global__ void __launch_bounds__(256)
    mc_list_kernel(unsigned n_slots, const uint32_t* __restrict__ slot_word,
                   unsigned* __restrict__ counters, uint32_t* __restrict__ slot_cube,
                   uint32_t* __restrict__ slot_cnt, uint32_t* __restrict__ cube_slot) {
  const unsigned slot = blockIdx.x * blockDim.x + threadIdx.x;
  if (slot >= n_slots) return;
  const int n = __popc(slot_word[slot]);
  const unsigned base = atomicAdd(counters + 2, (unsigned)n);
  slot_cube[slot] = base;
  slot_cnt[slot] = 0;
  for (int k = 0; k < n; k++) cube_slot[base + k] = slot;
}

// One thread per active cube: configuration and owned edges; counts per segment and per row.
__global__ void __launch_bounds__(256)
    mc_count_kernel(const McGeom m, const McLists t, unsigned n_cubes,
                    const uint32_t* __restrict__ ibits, const uint32_t* __restrict__ cbits,
                    uint32_t* __restrict__ cube_info_out, uint32_t* __restrict__ slot_cnt,
                    unsigned long long* __restrict__ rowcnt) {
  const unsigned c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= n_cubes) return;
  const CubeAt q = cube_at(m, t, c);
  const CubeInfo info = cube_info(m, ibits, cbits, q.x >> 5, q.y, q.z, q.x & 31);
  const unsigned nv = __popc(info.owned), nt = d_ntri[info.ci];
  cube_info_out[c] = (unsigned)info.ci | (info.owned << 8);
  // per segment: (vertices << 16) | faces (<= 384 and <= 160); per row of segments: 64-bit sums
  atomicAdd(slot_cnt + q.slot, (nv << 16) | nt);
  atomicAdd(rowcnt + q.row, ((unsigned long long)nv << 32) | nt);
}

// One warp per owned row with any active cube: ids of each segment's first vertex / face = the row's
// exclusive prefix (scan over rows) + the counts of the segments before it in the row.
__global__ void __launch_bounds__(256)
    mc_rowprefix_kernel(const McGeom m, int64_t n_rows, const uint32_t* __restrict__ segwid,
                        const uint32_t* __restrict__ slot_cnt,
                        const unsigned long long* __restrict__ rowcnt,
                        const uint64_t* __restrict__ rowbase, uint2* __restrict__ slot_pre) {
  const int lane = threadIdx.x & 31;
  const int64_t row = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (row >= n_rows || rowcnt[row] == 0) return;
  const uint64_t rb = rowbase[row];
  unsigned run_v = (unsigned)(rb >> 32), run_f = (unsigned)(rb & 0xffffffffull);
  for (int s0 = 0; s0 < m.spr; s0 += 32) {
    const int sx = s0 + lane;
    const unsigned slot = sx < m.spr ? segwid[row * m.spr + sx] : kNoSlot;
    const unsigned cnt = slot != kNoSlot ? slot_cnt[slot] : 0u;
    unsigned iv = cnt >> 16, it = cnt & 0xffffu;
    const unsigned nv = iv, nt = it;
    for (int o = 1; o < 32; o <<= 1) {
      const unsigned a = __shfl_up_sync(0xffffffffu, iv, o), b = __shfl_up_sync(0xffffffffu, it, o);
      if (lane >= o) {
        iv += a;
        it += b;
      }
    }
    if (slot != kNoSlot) slot_pre[slot] = make_uint2(run_v + iv - nv, run_f + it - nt);
    run_v += __shfl_sync(0xffffffffu, iv, 31);
    run_f += __shfl_sync(0xffffffffu, it, 31);
  }
}

// marching_cubes.cc:25-47, float64, one rounding per operation
__device__ __forceinline__ void vertex_interp(double iso, const float p1[3], const float p2[3],
                                              double v1, double v2, float out[3]) {
  if (fabs(iso - v1) < 0.00001) {
    out[0] = p1[0]; out[1] = p1[1]; out[2] = p1[2];
    return;
  }
  if (fabs(iso - v2) < 0.00001) {
    out[0] = p2[0]; out[1] = p2[1]; out[2] = p2[2];
    return;
  }
  if (fabs(v1 - v2) < 0.00001) {
    out[0] = p1[0]; out[1] = p1[1]; out[2] = p1[2];
    return;
  }
  const double mu = __ddiv_rn(__dsub_rn(iso, v1), __dsub_rn(v2, v1));
#pragma unroll
  for (int a = 0; a < 3; a++) {
    const double d = __dsub_rn((double)p2[a], (double)p1[a]);
    out[a] = (float)__dadd_rn((double)p1[a], __dmul_rn(mu, d));
  }
}

// One thread per active cube: the vertices it owns, at their final index, in the order the triangle
// walk first references their edges.
__global__ void __launch_bounds__(128)
    mc_vertex_kernel(const McGeom m, const GridDesc g, const McLists t, unsigned n_cubes,
                     const uint2* __restrict__ slot_pre, double iso, int linear_interp,
                     int64_t vertex_offset, float* __restrict__ vertices,
                     int32_t* __restrict__ cube_ids, uint64_t* hkeys, int32_t* hvals,
                     uint64_t hmask) {
  const unsigned c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= n_cubes) return;
  const CubeAt q = cube_at(m, t, c);
  const unsigned packed = t.cube_info[c];
  const int ci = (int)(packed & 0xffu);
  const unsigned owned = (packed >> 8) & 0xfffu;
  // first vertex of this cube: the segment's first + what the segment's earlier cubes own
  int64_t vid = slot_pre[q.slot].x;
  for (unsigned k = 0; k < q.k; k++) vid += __popc((t.cube_info[c - q.k + k] >> 8) & 0xfffu);
  // ids of the vertices on the three edges that end in this cube's max corner (e6 along x, e5 along
  // y, e10 along z) - the edges this cube is the HOME of; -1 where it owns none (see McLists)
  int32_t home_id[3] = {-1, -1, -1};
  const int x = q.x, y = q.y, z = q.z;
  const size_t plane = (size_t)m.nx * m.ny;
  const uint64_t ord = owned ? d_order[ci] : ~0ull;
  for (int k = 0; k < 12; k++) {
    const int e = (int)((ord >> (4 * k)) & 0xF);
    if (e == 0xF) break;
    if (!((owned >> e) & 1u)) continue;
    // the two end points of the edge, loaded on demand (most cubes own one or two vertices)
    const int a = c_edge_corner[e][0], b = c_edge_corner[e][1];
    const int ax = x + c_corner[a][0], ay = y + c_corner[a][1], az = z + c_corner[a][2];
    const int bx = x + c_corner[b][0], by = y + c_corner[b][1], bz = z + c_corner[b][2];
    const float p1[3] = {g.pos_x[ax], g.pos_y[ay], g.pos_z[az]};
    float p[3];
    if (linear_interp) {
      const float p2[3] = {g.pos_x[bx], g.pos_y[by], g.pos_z[bz]};
      // g.sdf points at plane z_begin; halo planes sit at negative offsets
      const float v1 = g.sdf[((ptrdiff_t)az - g.z_begin) * (ptrdiff_t)plane + (ptrdiff_t)ay * m.nx + ax];
      const float v2 = g.sdf[((ptrdiff_t)bz - g.z_begin) * (ptrdiff_t)plane + (ptrdiff_t)by * m.nx + bx];
      vertex_interp(iso, p1, p2, (double)v1, (double)v2, p);
    } else {  // marching_cubes.cc:55
      p[0] = p1[0]; p[1] = p1[1]; p[2] = p1[2];
    }
    vertices[vid * 3 + 0] = p[0];
    vertices[vid * 3 + 1] = p[1];
    vertices[vid * 3 + 2] = p[2];
    const int32_t gid = (int32_t)(vid + vertex_offset);
    if (e == 6) {
      home_id[0] = gid;
    } else if (e == 5) {
      home_id[1] = gid;
    } else if (e == 10) {
      home_id[2] = gid;
    } else {  // owned although another cube is its home: that cube is skipped or does not exist
      hash_insert(hkeys, hvals, hmask, edge_key(m, x, y, z, e), gid);
    }
    vid++;
  }
  int32_t* __restrict__ out = cube_ids + (size_t)c * 3;
  out[0] = home_id[0];
  out[1] = home_id[1];
  out[2] = home_id[2];
}

// One thread per active cube: its triangles in table order, vertex ids from the home cubes.
__global__ void __launch_bounds__(128)
    mc_face_kernel(const McGeom m, const McLists t, unsigned n_cubes,
                   const uint2* __restrict__ slot_pre, int32_t* __restrict__ faces) {
  const unsigned c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= n_cubes) return;
  const CubeAt q = cube_at(m, t, c);
  const int ci = (int)(t.cube_info[c] & 0xffu);
  const unsigned nt = d_ntri[ci];
  if (nt == 0) return;
  int64_t fid = slot_pre[q.slot].y;
  for (unsigned k = 0; k < q.k; k++) fid += d_ntri[t.cube_info[c - q.k + k] & 0xffu];
  // each crossed edge is looked up once, however many triangles share it
  int32_t vid[12];
  unsigned em = d_edge_mask[ci];
  while (em) {
    const int e = __ffs(em) - 1;
    em &= em - 1;
    const int a = c_edge_corner[e][0], b = c_edge_corner[e][1];
    const int hx = q.x + max(c_corner[a][0], c_corner[b][0]);
    const int hy = q.y + max(c_corner[a][1], c_corner[b][1]);
    const int hz = q.z + max(c_corner[a][2], c_corner[b][2]);
    const int axis = c_corner[a][0] != c_corner[b][0] ? 0 : (c_corner[a][1] != c_corner[b][1] ? 1 : 2);
    vid[e] = vertex_id_of_edge(m, t, hx, hy, hz, axis);
  }
  const uint64_t tri = d_tri[ci];
  for (unsigned k = 0; k < nt; k++) {
#pragma unroll
    for (int j = 0; j < 3; j++) {  // face[j] = vertex of edge tri[i + (2 - j)], :202-204
      const int e = (int)((tri >> (4 * (3 * k + 2 - j))) & 0xF);
      faces[fid * 3 + j] = vid[e];
    }
    fid++;
  }
}

const uint64_t h_tri[256] = VCB_MC_TRI_PACKED_INIT;
const unsigned char h_edge_corner[12][2] = {{0, 1}, {1, 2}, {2, 3}, {3, 0}, {4, 5}, {5, 6},
                                            {6, 7}, {7, 4}, {0, 4}, {1, 5}, {2, 6}, {3, 7}};

}  // namespace

void init_mc_tables() {
  uint16_t em[256];
  uint8_t nt[256];
  for (int c = 0; c < 256; c++) {
    int mask = 0;
    for (int e = 0; e < 12; e++) {
      const int a = (c >> h_edge_corner[e][0]) & 1, b = (c >> h_edge_corner[e][1]) & 1;
      if (a != b) mask |= 1 << e;
    }
    em[c] = (uint16_t)mask;
    int n = 0;
    while (n < 5 && ((h_tri[c] >> (12 * n)) & 0xF) != 0xF) n++;
    nt[c] = (uint8_t)n;
  }
  uint64_t order[256];
  for (int c = 0; c < 256; c++) {
    uint64_t o = 0;
    unsigned seen = 0;
    int k = 0;
    for (int i = 0; i < 15 && ((h_tri[c] >> (4 * i)) & 0xF) != 0xF; i += 3) {
      for (int j = 0; j < 3; j++) {
        const unsigned e = (unsigned)((h_tri[c] >> (4 * (i + 2 - j))) & 0xF);
        if ((seen >> e) & 1u) continue;
        seen |= 1u << e;
        o |= (uint64_t)e << (4 * k++);
      }
    }
    for (; k < 16; k++) o |= (uint64_t)0xF << (4 * k);
    order[c] = o;
  }
  cudaMemcpyToSymbol(d_order, order, sizeof(order));
  cudaMemcpyToSymbol(d_tri, h_tri, sizeof(h_tri));
  cudaMemcpyToSymbol(d_edge_mask, em, sizeof(em));
  cudaMemcpyToSymbol(d_ntri, nt, sizeof(nt));
}

// smallest float t with (double)s < iso  <=>  s < t for every float s
static float iso_threshold(double iso) {
  if (iso != iso) return NAN;
  float t = (float)iso;                         // round to nearest
  if ((double)t < iso) t = nextafterf(t, INFINITY);  // round up
  return t;
}

namespace {

// x- and y-edges lying in one voxel plane: id of the vertex on each (or -1). Used to hand the
// vertices on a slab's top plane to the slab above it.
__global__ void __launch_bounds__(256)
    mc_export_plane_kernel(const McGeom m, int z, const uint32_t* __restrict__ ibits,
                           const McLists t, int32_t* __restrict__ out) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t n = (int64_t)m.nx * m.ny;
  if (i >= 2 * n) return;
  const int axis = (int)(i / n);
  const int lx = (int)((i % n) % m.nx), ly = (int)((i % n) / m.nx);  // lower endpoint of the edge
  const int hx = lx + (axis == 0), hy = ly + (axis == 1);
  int32_t id = -1;
  // a vertex can only sit on an edge whose endpoints lie on different sides of the iso level
  if (hx < m.nx && hy < m.ny && get_bit(ibits, m, lx, ly, z) != get_bit(ibits, m, hx, hy, z)) {
    id = vertex_id_of_edge(m, t, hx, hy, z, axis);
  }
  out[i] = id;
}

__global__ void __launch_bounds__(256)
    mc_import_plane_kernel(int nx, int ny, int z, const int32_t* __restrict__ in, uint64_t* hkeys,
                           int32_t* hvals, uint64_t hmask) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t n = (int64_t)nx * ny;
  if (i >= 2 * n) return;
  const int32_t v = in[i];
  if (v < 0) return;
  const int axis = (int)(i / n);
  const uint64_t id = (uint64_t)z * n + (uint64_t)(i % n);
  hash_insert(hkeys, hvals, hmask, (id << 2) | (uint64_t)axis, v);
}

}  // namespace

#define VCB_TRY(x) \
  if ((e = (x)) != cudaSuccess) return e

// Scratch of the sparse passes (IsoExtract): abits = slot_word, active = slot_seg, segcnt = slot_cnt,
// slot_cube / slot_pre / cube_slot as named, cinfo = cube_info, vb = cube_ids.
static McLists mc_lists(const IsoExtract* st) {
  McLists t;
  t.segwid = static_cast<const uint32_t*>(st->segwid.ptr);
  t.slot_seg = static_cast<const uint32_t*>(st->active.ptr);
  t.slot_word = static_cast<const uint32_t*>(st->abits.ptr);
  t.slot_cube = static_cast<const uint32_t*>(st->slot_cube.ptr);
  t.cube_slot = static_cast<const uint32_t*>(st->cube_slot.ptr);
  t.cube_info = static_cast<const uint32_t*>(st->cinfo.ptr);
  t.cube_ids = static_cast<const int32_t*>(st->vb.ptr);
  t.hkeys = static_cast<const uint64_t*>(st->hkeys.ptr);
  t.hvals = static_cast<const int32_t*>(st->hvals.ptr);
  t.hmask = st->hmask;
  return t;
}

// Phase 1: classification, counts and the raster-order prefix sums. Leaves n_vertices / n_faces.
cudaError_t iso_count(const GridDesc& g, bool un_implied, double iso_level, int linear_interp,
                      IsoExtract* st, cudaStream_t stream, int64_t* launches) {
  const McGeom m = mc_geom(g);
  st->n_vertices = st->n_faces = -1;
  st->phase = 0;
  st->iso_level = iso_level;
  st->linear_interp = linear_interp;
  if (m.n_segs >= (int64_t)0xffffffffll) return cudaErrorInvalidValue;
  cudaError_t e;
  const size_t wbytes = (size_t)m.n_words * sizeof(uint32_t);
  VCB_TRY(st->ibits.reserve(wbytes));
  VCB_TRY(st->vbits.reserve(wbytes));
  VCB_TRY(st->cbits.reserve(wbytes));
  // the slot lists are reserved for the worst case (every segment active) and touched sparsely
  VCB_TRY(st->abits.reserve((size_t)m.n_segs * sizeof(uint32_t)));
  VCB_TRY(st->active.reserve((size_t)m.n_segs * sizeof(uint32_t)));
  VCB_TRY(st->segwid.reserve((size_t)m.n_segs * sizeof(uint32_t)));
  const int64_t n_rows_owned = (int64_t)g.ny * (g.z_end - g.z_begin);
  // per row of segments: [0, n_rows) counts, then [n_rows + 1 .. ) exclusive prefix + total
  VCB_TRY(st->segbase.reserve((size_t)(2 * n_rows_owned + 2) * sizeof(uint64_t)));
  VCB_TRY(st->counters.reserve(256));
  uint32_t* ibits = static_cast<uint32_t*>(st->ibits.ptr);
  uint32_t* vbits = static_cast<uint32_t*>(st->vbits.ptr);
  uint32_t* cbits = static_cast<uint32_t*>(st->cbits.ptr);
  uint64_t* rowcnt = static_cast<uint64_t*>(st->segbase.ptr);
  uint64_t* rowbase = rowcnt + n_rows_owned;
  unsigned* counters = static_cast<unsigned*>(st->counters.ptr);  // slots, cubes, list cursor
  uint32_t* ubits = nullptr;

  const size_t plane = (size_t)g.nx * g.ny;
  const float* sdf_stored = g.sdf - (ptrdiff_t)g.halo * (ptrdiff_t)plane;
  const int64_t n_rows = (int64_t)g.ny * (g.z_end - m.zb);
  VCB_TRY(cudaMemsetAsync(counters, 0, 4 * sizeof(unsigned), stream));
  mc_bits_kernel<<<(unsigned)((n_rows + kBitsRows - 1) / kBitsRows), 256, 0, stream>>>(
      sdf_stored, g.nx, m.spr, n_rows, iso_threshold(iso_level), ibits, vbits);
  if (launches) *launches += 1;
  if (!un_implied) {
    if (g.halo != 0) return cudaErrorInvalidValue;  // update_num has no halo planes
    VCB_TRY(st->tmp.reserve(wbytes));
    ubits = static_cast<uint32_t*>(st->tmp.ptr);
    mc_ubits_kernel<<<(unsigned)n_rows, 256, 0, stream>>>(g.update_num, g.nx, m.spr, n_rows, ubits);
    if (launches) *launches += 1;
  }
  mc_cube_kernel<<<(unsigned)((m.n_words + 255) / 256), 256, 0, stream>>>(
      m, ibits, vbits, ubits, cbits, static_cast<uint32_t*>(st->active.ptr),
      static_cast<uint32_t*>(st->abits.ptr), counters, static_cast<uint32_t*>(st->segwid.ptr));
  VCB_TRY(cudaMemsetAsync(rowcnt, 0, (size_t)n_rows_owned * sizeof(uint64_t), stream));
  if (launches) *launches += 1;
  unsigned h_counts[2] = {0, 0};
  VCB_TRY(cudaMemcpyAsync(h_counts, counters, sizeof(h_counts), cudaMemcpyDeviceToHost, stream));
  VCB_TRY(cudaStreamSynchronize(stream));
  st->n_active = h_counts[0];
  st->n_cubes = h_counts[1];
  if (st->n_cubes > 0) {
    const unsigned ns = st->n_active, nc = st->n_cubes;
    VCB_TRY(st->slot_cube.reserve((size_t)ns * sizeof(uint32_t)));
    VCB_TRY(st->segcnt.reserve((size_t)ns * sizeof(uint32_t)));
    VCB_TRY(st->slot_pre.reserve((size_t)ns * sizeof(uint2)));
    VCB_TRY(st->cube_slot.reserve((size_t)nc * sizeof(uint32_t)));
    VCB_TRY(st->cinfo.reserve((size_t)nc * sizeof(uint32_t)));
    VCB_TRY(st->vb.reserve((size_t)nc * 3 * sizeof(int32_t)));
    mc_list_kernel<<<(ns + 255) / 256, 256, 0, stream>>>(
        ns, static_cast<uint32_t*>(st->abits.ptr), counters, static_cast<uint32_t*>(st->slot_cube.ptr),
        static_cast<uint32_t*>(st->segcnt.ptr), static_cast<uint32_t*>(st->cube_slot.ptr));
    mc_count_kernel<<<(nc + 255) / 256, 256, 0, stream>>>(
        m, mc_lists(st), nc, ibits, cbits, static_cast<uint32_t*>(st->cinfo.ptr),
        static_cast<uint32_t*>(st->segcnt.ptr), reinterpret_cast<unsigned long long*>(rowcnt));
    if (launches) *launches += 2;
  }
  // the ubits scratch (st->tmp) is dead from here on; the scan reuses st->tmp
  VCB_TRY(scan_exclusive_u64(rowcnt, rowbase, n_rows_owned, rowbase + n_rows_owned, &st->tmp,
                             stream, launches));
  if (st->n_cubes > 0) {
    mc_rowprefix_kernel<<<(unsigned)((n_rows_owned * 32 + 255) / 256), 256, 0, stream>>>(
        m, n_rows_owned, static_cast<uint32_t*>(st->segwid.ptr), static_cast<uint32_t*>(st->segcnt.ptr),
        reinterpret_cast<unsigned long long*>(rowcnt), rowbase, static_cast<uint2*>(st->slot_pre.ptr));
    if (launches) *launches += 1;
  }
  uint64_t total = 0;
  VCB_TRY(cudaMemcpyAsync(&total, rowbase + n_rows_owned, sizeof(total), cudaMemcpyDeviceToHost,
                          stream));
  VCB_TRY(cudaStreamSynchronize(stream));
  st->n_vertices = (int64_t)(total >> 32);
  st->n_faces = (int64_t)(total & 0xffffffffull);
  st->phase = 1;
  return cudaSuccess;
}

// Phase 2: vertices at their final local index; the ids kept for the face pass are global
// (local index + vertex_offset).
cudaError_t iso_emit_vertices(const GridDesc& g, int64_t vertex_offset, IsoExtract* st,
                              cudaStream_t stream, int64_t* launches) {
  if (st->phase != 1) return cudaErrorInvalidValue;
  const McGeom m = mc_geom(g);
  cudaError_t e;
  const int64_t nv = st->n_vertices, nf = st->n_faces;
  // slots for the vertices handed over by the slab below (at most one per in-plane edge)
  const uint64_t imported = g.z_begin > 0 ? 2ull * g.nx * g.ny : 0ull;
  uint64_t cap = 1024;
  while (cap < ((uint64_t)nv + imported) * 2) cap <<= 1;
  st->hmask = cap - 1;
  VCB_TRY(st->hkeys.reserve(cap * sizeof(uint64_t)));
  VCB_TRY(st->hvals.reserve(cap * sizeof(int32_t)));
  VCB_TRY(st->verts.reserve((size_t)(nv > 0 ? nv : 1) * 3 * sizeof(float)));
  VCB_TRY(st->faces.reserve((size_t)(nf > 0 ? nf : 1) * 3 * sizeof(int32_t)));
  VCB_TRY(cudaMemsetAsync(st->hkeys.ptr, 0xff, cap * sizeof(uint64_t), stream));
  if (st->n_cubes > 0) {
    mc_vertex_kernel<<<(st->n_cubes + 127) / 128, 128, 0, stream>>>(
        m, g, mc_lists(st), st->n_cubes, static_cast<uint2*>(st->slot_pre.ptr), st->iso_level,
        st->linear_interp, vertex_offset, static_cast<float*>(st->verts.ptr),
        static_cast<int32_t*>(st->vb.ptr), static_cast<uint64_t*>(st->hkeys.ptr),
        static_cast<int32_t*>(st->hvals.ptr), st->hmask);
    if (launches) *launches += 1;
  }
  st->phase = 2;
  return cudaGetLastError();
}

cudaError_t iso_export_top_ids(const GridDesc& g, IsoExtract* st, int32_t* out_dev,
                               cudaStream_t stream, int64_t* launches) {
  if (st->phase < 2) return cudaErrorInvalidValue;
  const int64_t n2 = 2ll * g.nx * g.ny;
  mc_export_plane_kernel<<<(unsigned)((n2 + 255) / 256), 256, 0, stream>>>(
      mc_geom(g), g.z_end - 1, static_cast<uint32_t*>(st->ibits.ptr), mc_lists(st), out_dev);
  if (launches) *launches += 1;
  return cudaGetLastError();
}

cudaError_t iso_import_bottom_ids(const GridDesc& g, IsoExtract* st, const int32_t* in_dev,
                                  cudaStream_t stream, int64_t* launches) {
  if (st->phase != 2 || g.z_begin < 1) return cudaErrorInvalidValue;
  const int64_t n2 = 2ll * g.nx * g.ny;
  mc_import_plane_kernel<<<(unsigned)((n2 + 255) / 256), 256, 0, stream>>>(
      g.nx, g.ny, g.z_begin - 1, in_dev, static_cast<uint64_t*>(st->hkeys.ptr),
      static_cast<int32_t*>(st->hvals.ptr), st->hmask);
  if (launches) *launches += 1;
  return cudaGetLastError();
}

// Phase 3: triangles in table order with ids looked up by edge key.
cudaError_t iso_emit_faces(const GridDesc& g, IsoExtract* st, cudaStream_t stream,
                           int64_t* launches) {
  if (st->phase != 2) return cudaErrorInvalidValue;
  const McGeom m = mc_geom(g);
  if (st->n_cubes > 0 && st->n_faces > 0) {
    mc_face_kernel<<<(st->n_cubes + 127) / 128, 128, 0, stream>>>(
        m, mc_lists(st), st->n_cubes, static_cast<uint2*>(st->slot_pre.ptr),
        static_cast<int32_t*>(st->faces.ptr));
    if (launches) *launches += 1;
  }
  st->phase = 3;
  return cudaGetLastError();
}

#undef VCB_TRY

}  // namespace vcb
