// The z-slab engine: ONE voxel grid carved by several GPUs behind the C ABI (vcb_sharded_*,
// include/vacancy_b200.h).  It stands where a single vacancy::VoxelCarver stands
// (voxel_carver.h:95-118): the caller carves views and extracts ONE mesh; the slabs together
// produce exactly the grid, and exactly the marching-cubes mesh (vertex numbering and face order of
// marching_cubes.cc:199-218 over the whole grid), that one carver would.
//
// Two ways to span the GPUs, same engine:
//   * single process, a list of devices (vcb_sharded_create): exchanges are peer copies between the
//     devices (copy engines over NVLink, no SM time), ordered with events.  A device may appear more
//     than once ("virtual slabs"), which is how the single-GPU tests drive this code.
//   * one process per GPU (vcb_sharded_create_rank): the small exchanges (halo planes, counts, id
//     planes, mesh parts) are NCCL collectives / send-recv on the communicator built from a unique
//     id the launcher distributes; the bulk - the SDF images - is pulled by the copy engines from
//     the other ranks' pools, mapped once with CUDA IPC, with a tiny NCCL all-gather as the signal
//     (see "the pools of the other ranks" below).  NCCL is resolved with dlopen("libnccl.so.2")
//     when the first communicator is made - a process that already loaded NCCL (torch) shares
//     it, a single-GPU host never needs it.
//
// Carving, per chunk of <= 64 views (one fusion launch):
//   1. view i of the chunk belongs to rank i mod N: only that rank receives the mask (its own PCIe
//      link) and runs the 2D SDF kernels (voxel_carver.cc:169-237) for it, on a side stream;
//   2. the float SDF images, their 8x8-block flags and their maxima are gathered into every
//      rank's pool (rank r's results occupy slots [r*J, (r+1)*J), J = ceil(views / N));
//   3. every rank fuses all views of the chunk into its slab (voxel_carver.cc:415-496), in call
//      order, on its main stream.
// The pools are double-buffered: steps 1-2 of chunk k+1 run under step 3 of chunk k.
//
// Marching cubes follows the phase protocol of the C ABI (vcb_extract_iso_begin ...): halo planes
// up, counts all-gathered, vertex-id planes up; the mesh is then copied into the caller's buffers
// (single process: every device writes its part; one process per GPU: parts are sent to rank 0).
#include <dlfcn.h>
#include <nccl.h>
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <mutex>
#include <thread>

#include "carver_internal.cuh"

using namespace vcb;

namespace {

// ---- NCCL through dlopen ------------------------------------------------------------------------
struct NcclApi {
  void* handle = nullptr;
  ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
  ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
  ncclResult_t (*CommInitRankConfig)(ncclComm_t*, int, ncclUniqueId, int, ncclConfig_t*) = nullptr;
  ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
  ncclResult_t (*AllGather)(const void*, void*, size_t, ncclDataType_t, ncclComm_t,
                            cudaStream_t) = nullptr;
  ncclResult_t (*Send)(const void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*Recv)(void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*GroupStart)() = nullptr;
  ncclResult_t (*GroupEnd)() = nullptr;
  const char* (*GetErrorString)(ncclResult_t) = nullptr;
  std::string error;
};

NcclApi g_nccl;
std::mutex g_nccl_mutex;

bool nccl_load() {
  std::lock_guard<std::mutex> lock(g_nccl_mutex);
  if (g_nccl.handle) return true;
  // a copy the process already holds (torch ships its own) wins over the system one
  void* h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL | RTLD_NOLOAD);
  if (!h) h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
  if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
  if (!h) {
    g_nccl.error = std::string("cannot load libnccl.so.2: ") + dlerror();
    return false;
  }
#define VCB_SYM(field, name)                                              \
  g_nccl.field = reinterpret_cast<decltype(g_nccl.field)>(dlsym(h, name)); \
  if (!g_nccl.field) {                                                    \
    g_nccl.error = std::string("libnccl.so.2 lacks ") + name;             \
    return false;                                                         \
  }
  VCB_SYM(GetUniqueId, "ncclGetUniqueId")
  VCB_SYM(CommInitRank, "ncclCommInitRank")
  VCB_SYM(CommDestroy, "ncclCommDestroy")
  VCB_SYM(AllGather, "ncclAllGather")
  VCB_SYM(Send, "ncclSend")
  VCB_SYM(Recv, "ncclRecv")
  VCB_SYM(GroupStart, "ncclGroupStart")
  VCB_SYM(GroupEnd, "ncclGroupEnd")
  VCB_SYM(GetErrorString, "ncclGetErrorString")
#undef VCB_SYM
  g_nccl.CommInitRankConfig =
      reinterpret_cast<decltype(g_nccl.CommInitRankConfig)>(dlsym(h, "ncclCommInitRankConfig"));
  g_nccl.handle = h;
  return true;
}

// ---- the engine's state -------------------------------------------------------------------------
struct ShardRank {  // one slab held by this process
  int rank = 0, device = 0;
  vcb_carver* car = nullptr;
  cudaStream_t main = nullptr, side = nullptr, copy = nullptr;
  ncclComm_t comm = nullptr;
  // double-buffered pools of one chunk: float images, block flags, image maxima; mask staging
  Scratch img[2], tiles[2], maxs[2], stage[2];
  Scratch sdf_scr, absmax;            // 2D SDF scratch for kSdfSubBatch views
  Scratch ids_out, ids_in, cnt, gather;
  // one process per GPU: the other ranks' pools mapped into this process (CUDA IPC), [array][half]
  // [rank]; array 0 images, 1 block flags, 2 maxima.  Empty when the gathers go through NCCL.
  std::vector<void*> peer[3][2];
  Scratch ipc_buf, signal;
  // pulls from different ranks go out on several streams so that several copy engines work
  static constexpr int kPullStreams = 4;
  cudaStream_t pull[kPullStreams] = {};
  cudaEvent_t ev_signal = nullptr, ev_pulled[kPullStreams] = {};
  cudaEvent_t ev_staged[2] = {}, ev_sdf[2] = {}, ev_gathered[2] = {}, ev_fused[2] = {}, ev_misc = nullptr;
  bool used[2] = {false, false};
  int64_t nv = 0, nf = 0, voff = 0, foff = 0;  // this slab's part of the last extracted mesh
  int64_t nvox = 0, vox_off = 0;               // ... of the last ExtractVoxel
};

}  // namespace

struct vcb_sharded {
  vcb_options opt;
  int world = 1;
  int nx = 0, ny = 0, nz = 0;
  bool multi_process = false;
  std::vector<int> bounds;       // world + 1 plane indices
  std::vector<ShardRank> local;  // the slabs of this process, in slab order
  std::string err;
  // the chunk being collected
  int chunk_cap = kMaxViewsPerLaunch;
  // After Init() / reset the first chunk may be kept small (two views per GPU, see queue_view):
  // nothing can overlap its transforms and its gather, so the fusion starts sooner.
  bool ramp = true;
  // vcb_sharded_set_auto_balance: before the first chunk is fused the slab boundaries are placed
  // from a cost estimate of that chunk's tile classes (the grid holds nothing yet, so moving the
  // boundaries costs a re-allocation, not a migration)
  bool balance_pending = false;
  bool probing = false;  // the chunk being collected is a sample for the estimate, not to be fused
  int queued = 0, half = 0;
  bool q_device = false;
  vcb_view qviews[kMaxViewsPerLaunch];
  const uint8_t* qmask[kMaxViewsPerLaunch];
  size_t slot_img = 0, slot_tiles = 0, stage_pix = 0, scr_bytes = 0;
  // last extraction
  std::vector<int64_t> counts;   // (vertices, faces) per rank
  int64_t total_v = -1, total_f = -1;
  int iso_phase = 0;             // 0 nothing, 1 mesh on the devices
  int64_t total_vox = -1;
  std::vector<int64_t> vox_counts;
  PinnedBuffer ingest_pin;       // decoded mask files (vcb_sharded_carve_mask_files)
  bool ce_gather = true;         // one process per GPU: gather with copy engines over IPC mappings
};

namespace {

int sfail(vcb_sharded* s, int code, const char* fmt, ...) {
  char buf[600];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof(buf), fmt, ap);
  va_end(ap);
  if (s) s->err = buf;
  set_global_error(buf);
  return code;
}

#define SCU(s, call)                                                                            \
  do {                                                                                          \
    cudaError_t e_ = (call);                                                                    \
    if (e_ != cudaSuccess)                                                                      \
      return sfail(s, VCB_ERR_CUDA, "%s: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__,   \
                   __LINE__);                                                                   \
  } while (0)
#define SNCCL(s, call)                                                                          \
  do {                                                                                          \
    ncclResult_t r_ = (call);                                                                   \
    if (r_ != ncclSuccess)                                                                      \
      return sfail(s, VCB_ERR_CUDA, "%s: %s (%s:%d)", #call, g_nccl.GetErrorString(r_),          \
                   __FILE__, __LINE__);                                                         \
  } while (0)
// a failed call on one slab's carver: carry its text up
#define SCAR(s, L, call)                                                                        \
  do {                                                                                          \
    int rc_ = (call);                                                                           \
    if (rc_ != VCB_OK) return sfail(s, rc_, "slab %d: %s", (L).rank, vcb_last_error((L).car));   \
  } while (0)

size_t align256(size_t n) { return (n + 255) & ~(size_t)255; }

bool env_flag(const char* name, bool fallback) {
  const char* e = getenv(name);
  return e ? (e[0] == '1') : fallback;
}

ShardRank* local_of(vcb_sharded* s, int rank) {
  for (auto& L : s->local)
    if (L.rank == rank) return &L;
  return nullptr;
}

// Runs f(slab) for every slab of this process; with several slabs each gets its own host thread,
// because the extraction phases block on their device (they read counts back).
template <class F>
int for_each_local(vcb_sharded* s, F f) {
  if (s->local.size() == 1) return f(s->local[0]);
  std::vector<int> rc(s->local.size(), VCB_OK);
  std::vector<std::string> text(s->local.size());
  std::vector<std::thread> th;
  for (size_t i = 0; i < s->local.size(); i++) {
    th.emplace_back([&, i] {
      rc[i] = f(s->local[i]);
      if (rc[i] != VCB_OK) text[i] = vcb_last_error(s->local[i].car);
    });
  }
  for (auto& t : th) t.join();
  for (size_t i = 0; i < rc.size(); i++) {
    if (rc[i] != VCB_OK) return sfail(s, rc[i], "slab %d: %s", s->local[i].rank, text[i].c_str());
  }
  return VCB_OK;
}

void default_bounds(int nz, int world, std::vector<int>* b) {
  b->assign(world + 1, 0);
  const int base = nz / world, extra = nz % world;
  for (int r = 0; r < world; r++) (*b)[r + 1] = (*b)[r] + base + (r < extra ? 1 : 0);
}

int check_bounds(vcb_sharded* s, const std::vector<int>& b) {
  if ((int)b.size() != s->world + 1 || b[0] != 0 || b[s->world] != s->nz)
    return sfail(s, VCB_ERR_INVALID, "slab bounds must run from 0 to %d", s->nz);
  for (int r = 0; r < s->world; r++) {
    // the slab above keeps two halo planes of this one (marching cubes)
    if (b[r + 1] - b[r] < (s->world > 1 ? 2 : 1))
      return sfail(s, VCB_ERR_INVALID, "slab %d has %d planes; every slab needs at least 2", r,
                   b[r + 1] - b[r]);
  }
  return VCB_OK;
}

void destroy_carvers(vcb_sharded* s) {
  for (auto& L : s->local) {
    if (L.car) {
      cudaSetDevice(L.device);
      cudaStreamSynchronize(L.side);
      cudaStreamSynchronize(L.copy);
      vcb_destroy(L.car);  // synchronises its stream
      L.car = nullptr;
    }
    L.used[0] = L.used[1] = false;
  }
}

int create_carvers(vcb_sharded* s) {
  for (auto& L : s->local) {
    vcb_options o = s->opt;
    o.device = L.device;
    o.z_begin = s->bounds[L.rank];
    o.z_end = s->bounds[L.rank + 1];
    if (s->world == 1) o.z_begin = o.z_end = 0;
    int rc = vcb_create(&o, &L.car);
    if (rc != VCB_OK) return sfail(s, rc, "slab %d: %s", L.rank, vcb_last_error(nullptr));
    SCAR(s, L, vcb_set_stream(L.car, L.main));
  }
  return VCB_OK;
}

int init_common(vcb_sharded* s) {
  int dims[3];
  int rc = validate(&s->opt, dims);
  if (rc != VCB_OK) return sfail(s, rc, "%s", vcb_last_error(nullptr));
  s->nx = dims[0];
  s->ny = dims[1];
  s->nz = dims[2];
  s->chunk_cap = kMaxViewsPerLaunch / s->world * s->world;
  if (s->chunk_cap < 1) return sfail(s, VCB_ERR_INVALID, "more than %d slabs", kMaxViewsPerLaunch);
  for (auto& L : s->local) {
    SCU(s, cudaSetDevice(L.device));
    int lo = 0, hi = 0;
    SCU(s, cudaDeviceGetStreamPriorityRange(&lo, &hi));
    SCU(s, cudaStreamCreateWithPriority(&L.main, cudaStreamNonBlocking, lo));
    // the 2D SDF kernels and the gathers are small: they go ahead of the fusion blocks still queued.
    // With one slab there is nothing to gather, and running the transforms next to the fusion kernel
    // only makes both slower (measured: 416.7 vs 409 ms per cfg4 pass): one stream then.
    if (s->world > 1 || env_flag("VACANCY_B200_OVERLAP_SDF", false)) {
      SCU(s, cudaStreamCreateWithPriority(&L.side, cudaStreamNonBlocking, hi));
    } else {
      L.side = L.main;
    }
    SCU(s, cudaStreamCreateWithPriority(&L.copy, cudaStreamNonBlocking, hi));
    for (int h = 0; h < 2; h++) {
      SCU(s, cudaEventCreateWithFlags(&L.ev_staged[h], cudaEventDisableTiming));
      SCU(s, cudaEventCreateWithFlags(&L.ev_sdf[h], cudaEventDisableTiming));
      SCU(s, cudaEventCreateWithFlags(&L.ev_gathered[h], cudaEventDisableTiming));
      SCU(s, cudaEventCreateWithFlags(&L.ev_fused[h], cudaEventDisableTiming));
    }
    SCU(s, cudaEventCreateWithFlags(&L.ev_misc, cudaEventDisableTiming));
    SCU(s, cudaEventCreateWithFlags(&L.ev_signal, cudaEventDisableTiming));
    for (int p = 0; p < ShardRank::kPullStreams; p++) {
      SCU(s, cudaStreamCreateWithPriority(&L.pull[p], cudaStreamNonBlocking, hi));
      SCU(s, cudaEventCreateWithFlags(&L.ev_pulled[p], cudaEventDisableTiming));
    }
  }
  return VCB_OK;
}

// buffers sized for images of w x h (grow only); flushes and waits before reallocating
int flush_chunk(vcb_sharded* s);

// ---- one process per GPU: the pools of the other ranks, mapped with CUDA IPC -----------------------
// NCCL's all-gather kernels take SMs away from the fusion kernel they run beside (measured on 8
// GPUs: 52-55.7 ms of fusion per rank against 49.5-52.9 ms with NCCL held to 2 blocks, which then
// cannot move the data in time).  So the bulk of the gather is done by the copy engines: every
// rank maps the other ranks' pools once (cudaIpcGetMemHandle / cudaIpcOpenMemHandle, the handles
// travel through one small NCCL all-gather) and pulls the slots it needs with plain device-to-
// device copies over NVLink.  NCCL stays the signalling plane: a 4-byte all-gather per chunk tells
// every rank that all ranks have finished writing their slots (and reading the previous chunk).
int close_peer_maps(vcb_sharded* s) {
  for (auto& L : s->local) {
    SCU(s, cudaSetDevice(L.device));
    for (int a = 0; a < 3; a++) {
      for (int hf = 0; hf < 2; hf++) {
        for (size_t r = 0; r < L.peer[a][hf].size(); r++) {
          if (L.peer[a][hf][r] && (int)r != L.rank) cudaIpcCloseMemHandle(L.peer[a][hf][r]);
        }
        L.peer[a][hf].clear();
      }
    }
  }
  return VCB_OK;
}

// every rank reaches this point: a 4-byte all-gather on `stream`, then wait for it on the host
int nccl_barrier(vcb_sharded* s, ShardRank& L, cudaStream_t stream, bool wait) {
  SCU(s, L.signal.reserve(sizeof(int) * (size_t)s->world));
  int* d = static_cast<int*>(L.signal.ptr);
  SNCCL(s, g_nccl.AllGather(d + L.rank, d, sizeof(int), ncclUint8, L.comm, stream));
  if (wait) SCU(s, cudaStreamSynchronize(stream));
  return VCB_OK;
}

int open_peer_maps(vcb_sharded* s) {
  ShardRank& L = s->local[0];
  SCU(s, cudaSetDevice(L.device));
  constexpr size_t kPer = 6 * sizeof(cudaIpcMemHandle_t);
  SCU(s, L.ipc_buf.reserve(kPer * (size_t)s->world));
  cudaIpcMemHandle_t mine[6];
  Scratch* arrays[3] = {L.img, L.tiles, L.maxs};
  for (int a = 0; a < 3; a++)
    for (int hf = 0; hf < 2; hf++) SCU(s, cudaIpcGetMemHandle(&mine[a * 2 + hf], arrays[a][hf].ptr));
  unsigned char* d = static_cast<unsigned char*>(L.ipc_buf.ptr);
  SCU(s, cudaMemcpyAsync(d + kPer * L.rank, mine, kPer, cudaMemcpyHostToDevice, L.side));
  SNCCL(s, g_nccl.AllGather(d + kPer * L.rank, d, kPer, ncclUint8, L.comm, L.side));
  std::vector<cudaIpcMemHandle_t> all(6 * (size_t)s->world);
  SCU(s, cudaMemcpyAsync(all.data(), d, kPer * s->world, cudaMemcpyDeviceToHost, L.side));
  SCU(s, cudaStreamSynchronize(L.side));
  for (int a = 0; a < 3; a++) {
    for (int hf = 0; hf < 2; hf++) {
      L.peer[a][hf].assign(s->world, nullptr);
      for (int r = 0; r < s->world; r++) {
        if (r == L.rank) {
          L.peer[a][hf][r] = arrays[a][hf].ptr;
          continue;
        }
        SCU(s, cudaIpcOpenMemHandle(&L.peer[a][hf][r], all[(size_t)r * 6 + a * 2 + hf],
                                    cudaIpcMemLazyEnablePeerAccess));
      }
    }
  }
  return VCB_OK;
}

int ensure_buffers(vcb_sharded* s, int w, int h) {
  const int mode = s->opt.sdf_metric == VCB_METRIC_L2 ? kSdfSignedL2 : kSdfSigned;
  const size_t img = align256((size_t)w * h * sizeof(float));
  const size_t tl = align256(tile_flag_bytes(w, h));
  const size_t pix = align256((size_t)w * h);
  const size_t scr = align256(sdf_scratch_bytes(w, h, mode));
  if (img <= s->slot_img && tl <= s->slot_tiles && pix <= s->stage_pix && scr <= s->scr_bytes)
    return VCB_OK;
  int rc = flush_chunk(s);
  if (rc != VCB_OK) return rc;
  s->slot_img = img > s->slot_img ? img : s->slot_img;
  s->slot_tiles = tl > s->slot_tiles ? tl : s->slot_tiles;
  s->stage_pix = pix > s->stage_pix ? pix : s->stage_pix;
  s->scr_bytes = scr > s->scr_bytes ? scr : s->scr_bytes;
  // every process of a multi-process run sees the same views, so all ranks resize alike
  for (auto& L : s->local) {
    SCU(s, cudaSetDevice(L.device));
    SCU(s, cudaStreamSynchronize(L.main));
    SCU(s, cudaStreamSynchronize(L.side));
    SCU(s, cudaStreamSynchronize(L.copy));
  }
  const bool ipc = s->multi_process && s->ce_gather;
  if (ipc) {
    // nobody may free a pool that another rank still has mapped: unmap, meet, then re-allocate
    if ((rc = close_peer_maps(s)) != VCB_OK) return rc;
    if ((rc = nccl_barrier(s, s->local[0], s->local[0].side, true)) != VCB_OK) return rc;
  }
  for (auto& L : s->local) {
    SCU(s, cudaSetDevice(L.device));
    for (int hf = 0; hf < 2; hf++) {
      SCU(s, L.img[hf].reserve(s->slot_img * s->chunk_cap));
      SCU(s, L.tiles[hf].reserve(s->slot_tiles * s->chunk_cap));
      SCU(s, L.maxs[hf].reserve(align256(sizeof(float) * s->chunk_cap)));
      SCU(s, L.stage[hf].reserve(s->stage_pix * (s->chunk_cap / s->world)));
      L.used[hf] = false;
    }
    SCU(s, L.sdf_scr.reserve(s->scr_bytes * kSdfSubBatch));
    SCU(s, L.absmax.reserve(sizeof(int) * kSdfSubBatch));
  }
  if (ipc && (rc = open_peer_maps(s)) != VCB_OK) return rc;
  return VCB_OK;
}

// Slab boundaries from the tile classes of the chunk that has just been gathered (half `h`, `cnt`
// views, J slots per rank).  Every process computes the same integer costs from the same images, so
// all ranks arrive at the same boundaries without talking to each other.  The grid must still be
// fresh: the carvers are simply re-created on the new plane ranges.
int place_boundaries(vcb_sharded* s, int cnt, int J, int h) {
  const int N = s->world, q = fusion_tile_planes();
  const int layers = (s->nz + q - 1) / q;
  if (layers < 2 * N) return VCB_OK;  // too thin to place anything
  ShardRank& L = s->local[0];
  SCU(s, cudaSetDevice(L.device));
  SCU(s, cudaStreamSynchronize(L.side));  // this chunk's images, flags and maxima are in the pool
  const float* images[kMaxViewsPerLaunch];
  const uint32_t* tiles[kMaxViewsPerLaunch];
  const float* maxs[kMaxViewsPerLaunch];
  for (int idx = 0; idx < cnt; idx++) {
    const int slot = (idx % N) * J + idx / N;
    images[idx] = reinterpret_cast<const float*>(static_cast<unsigned char*>(L.img[h].ptr) +
                                                 (size_t)slot * s->slot_img);
    tiles[idx] = reinterpret_cast<const uint32_t*>(static_cast<unsigned char*>(L.tiles[h].ptr) +
                                                   (size_t)slot * s->slot_tiles);
    maxs[idx] = static_cast<const float*>(L.maxs[h].ptr) + slot;
  }
  std::vector<uint64_t> cost;
  SCAR(s, L, layer_costs(L.car, cnt, s->qviews, images, tiles, maxs, &cost));
  uint64_t total = 0;
  for (uint64_t c : cost) total += c;
  if (env_flag("VACANCY_B200_DEBUG_BALANCE", false)) {
    fprintf(stderr, "vacancy_b200: cost per layer of %d planes (rank %d):", q, L.rank);
    for (uint64_t c : cost) fprintf(stderr, " %llu", (unsigned long long)c);
    fprintf(stderr, "\n");
  }
  if (total == 0) return VCB_OK;
  std::vector<int> b(N + 1, 0);
  if (vcb_balance_bounds(cost.data(), layers, q, s->nz, N, b.data()) != VCB_OK) return VCB_OK;
  if (b == s->bounds) return VCB_OK;
  int rc = check_bounds(s, b);
  if (rc != VCB_OK) return rc;
  destroy_carvers(s);
  s->bounds = b;
  return create_carvers(s);
}

// steps 1-3 of the header comment for the views collected so far
int flush_chunk(vcb_sharded* s) {
  const int cnt = s->queued;
  if (cnt == 0) return VCB_OK;
  const int N = s->world, h = s->half;
  const int J = (cnt + N - 1) / N;  // slots per rank in this chunk
  const int mode = s->opt.sdf_metric == VCB_METRIC_L2 ? kSdfSignedL2 : kSdfSigned;
  // ---- 1. each slab transforms the views it owns, on its side stream ----
  for (auto& L : s->local) {
    SCU(s, cudaSetDevice(L.device));
    if (L.used[h]) {
      // this half is free once the fusion launch that read it has finished - and, between the
      // devices of one process, once every device has copied this device's part out of it
      SCU(s, cudaStreamWaitEvent(L.side, L.ev_fused[h], 0));
      if (!s->multi_process)
        for (auto& O : s->local) SCU(s, cudaStreamWaitEvent(L.side, O.ev_gathered[h], 0));
    }
    if (!s->q_device) {
      SCU(s, cudaEventRecord(L.ev_staged[h], L.copy));
      SCU(s, cudaStreamWaitEvent(L.side, L.ev_staged[h], 0));
    }
    int owned = 0;
    for (int j = 0; j * N + L.rank < cnt; j++) owned++;
    for (int first = 0; first < owned; first += kSdfSubBatch) {
      const int nb = owned - first < kSdfSubBatch ? owned - first : kSdfSubBatch;
      SdfView hv[kSdfSubBatch];
      for (int i = 0; i < nb; i++) {
        const int j = first + i, idx = j * N + L.rank, slot = L.rank * J + j;
        const vcb_view& v = s->qviews[idx];
        hv[i].mask = s->q_device ? s->qmask[idx]
                                 : static_cast<const uint8_t*>(L.stage[h].ptr) + (size_t)j * s->stage_pix;
        hv[i].out = reinterpret_cast<float*>(static_cast<unsigned char*>(L.img[h].ptr) +
                                             (size_t)slot * s->slot_img);
        hv[i].scratch = static_cast<unsigned char*>(L.sdf_scr.ptr) + (size_t)i * s->scr_bytes;
        hv[i].absmax = static_cast<int*>(L.absmax.ptr) + i;
        hv[i].max_out = static_cast<float*>(L.maxs[h].ptr) + slot;
        hv[i].tile_flags = reinterpret_cast<uint32_t*>(static_cast<unsigned char*>(L.tiles[h].ptr) +
                                                       (size_t)slot * s->slot_tiles);
        hv[i].width = v.width;
        hv[i].height = v.height;
        hv[i].rx0 = v.roi_min[0];
        hv[i].ry0 = v.roi_min[1];
        hv[i].rx1 = v.roi_max[0];
        hv[i].ry1 = v.roi_max[1];
      }
      Span span(L.car, 0, L.side);
      launch_sdf2d(hv, nb, mode, s->opt.sdf_minmax_normalize, s->opt.use_truncation,
                   s->opt.truncation_band, L.side, &L.car->stage_launches[0]);
    }
    SCU(s, cudaGetLastError());
    SCU(s, cudaEventRecord(L.ev_sdf[h], L.side));
  }
  // ---- 2. all-gather of the three pools: rank r's slots are [r*J, (r+1)*J) ----
  if (N > 1) {
    if (s->multi_process && s->ce_gather) {
      // signal: every rank has written its slots of this chunk (and has finished pulling the
      // previous chunk, which sits before this call on its side stream); then pull the other
      // ranks' slots with the copy engines
      ShardRank& L = s->local[0];
      int rc = nccl_barrier(s, L, L.side, false);
      if (rc != VCB_OK) return rc;
      const size_t slot[3] = {s->slot_img, s->slot_tiles, sizeof(float)};
      Scratch* arrays[3] = {L.img, L.tiles, L.maxs};
      SCU(s, cudaEventRecord(L.ev_signal, L.side));
      for (int p = 0; p < ShardRank::kPullStreams; p++) SCU(s, cudaStreamWaitEvent(L.pull[p], L.ev_signal, 0));
      for (int k = 1; k < N; k++) {
        const int r = (L.rank + k) % N;  // every rank starts with another peer
        int owned = 0;
        for (int j = 0; j * N + r < cnt; j++) owned++;
        if (owned == 0) continue;
        cudaStream_t ps = L.pull[k % ShardRank::kPullStreams];
        for (int a = 0; a < 3; a++) {
          const size_t off = (size_t)r * J * slot[a];
          SCU(s, cudaMemcpyAsync(static_cast<unsigned char*>(arrays[a][h].ptr) + off,
                                 static_cast<const unsigned char*>(L.peer[a][h][r]) + off,
                                 (size_t)owned * slot[a], cudaMemcpyDeviceToDevice, ps));
        }
      }
      for (int p = 0; p < ShardRank::kPullStreams; p++) {
        SCU(s, cudaEventRecord(L.ev_pulled[p], L.pull[p]));
        SCU(s, cudaStreamWaitEvent(L.side, L.ev_pulled[p], 0));
      }
    } else if (s->multi_process) {
      ShardRank& L = s->local[0];
      SNCCL(s, g_nccl.GroupStart());
      SNCCL(s, g_nccl.AllGather(static_cast<unsigned char*>(L.img[h].ptr) + (size_t)L.rank * J * s->slot_img,
                                L.img[h].ptr, (size_t)J * s->slot_img, ncclUint8, L.comm, L.side));
      SNCCL(s, g_nccl.AllGather(static_cast<unsigned char*>(L.tiles[h].ptr) + (size_t)L.rank * J * s->slot_tiles,
                                L.tiles[h].ptr, (size_t)J * s->slot_tiles, ncclUint8, L.comm, L.side));
      SNCCL(s, g_nccl.AllGather(static_cast<float*>(L.maxs[h].ptr) + (size_t)L.rank * J, L.maxs[h].ptr,
                                (size_t)J * sizeof(float), ncclUint8, L.comm, L.side));
      SNCCL(s, g_nccl.GroupEnd());
    } else {
      for (auto& D : s->local) {
        SCU(s, cudaSetDevice(D.device));
        for (auto& S : s->local) {
          if (&S == &D) continue;
          int owned = 0;
          for (int j = 0; j * N + S.rank < cnt; j++) owned++;
          if (owned == 0) continue;
          SCU(s, cudaStreamWaitEvent(D.side, S.ev_sdf[h], 0));
          const size_t o_img = (size_t)S.rank * J * s->slot_img, o_tl = (size_t)S.rank * J * s->slot_tiles,
                       o_mx = (size_t)S.rank * J * sizeof(float);
          SCU(s, cudaMemcpyPeerAsync(static_cast<unsigned char*>(D.img[h].ptr) + o_img, D.device,
                                     static_cast<unsigned char*>(S.img[h].ptr) + o_img, S.device,
                                     (size_t)owned * s->slot_img, D.side));
          SCU(s, cudaMemcpyPeerAsync(static_cast<unsigned char*>(D.tiles[h].ptr) + o_tl, D.device,
                                     static_cast<unsigned char*>(S.tiles[h].ptr) + o_tl, S.device,
                                     (size_t)owned * s->slot_tiles, D.side));
          SCU(s, cudaMemcpyPeerAsync(static_cast<unsigned char*>(D.maxs[h].ptr) + o_mx, D.device,
                                     static_cast<unsigned char*>(S.maxs[h].ptr) + o_mx, S.device,
                                     (size_t)owned * sizeof(float), D.side));
        }
      }
    }
  }
  // ---- 2b. auto-balance: place the slab boundaries from this chunk's tile classes ----
  if (s->balance_pending && N > 1) {
    s->balance_pending = false;
    if (s->probing) {
      // a sample of the views of a long call: estimate, then forget the chunk (its views come
      // again, in call order); nothing may still be reading the pools when they are reused
      for (auto& L : s->local) {
        SCU(s, cudaSetDevice(L.device));
        SCU(s, cudaStreamSynchronize(L.side));
      }
      s->queued = 0;
      return place_boundaries(s, cnt, J, h);
    }
    int rc = place_boundaries(s, cnt, J, h);
    if (rc != VCB_OK) return rc;
  }
  // ---- 3. every slab fuses the whole chunk, in call order ----
  for (auto& L : s->local) {
    SCU(s, cudaSetDevice(L.device));
    SCU(s, cudaEventRecord(L.ev_gathered[h], L.side));
    SCU(s, cudaStreamWaitEvent(L.main, L.ev_gathered[h], 0));
    const float* images[kMaxViewsPerLaunch];
    const uint32_t* tiles[kMaxViewsPerLaunch];
    const float* maxs[kMaxViewsPerLaunch];
    for (int idx = 0; idx < cnt; idx++) {
      const int slot = (idx % N) * J + idx / N;
      images[idx] = reinterpret_cast<const float*>(static_cast<unsigned char*>(L.img[h].ptr) +
                                                   (size_t)slot * s->slot_img);
      tiles[idx] = reinterpret_cast<const uint32_t*>(static_cast<unsigned char*>(L.tiles[h].ptr) +
                                                     (size_t)slot * s->slot_tiles);
      maxs[idx] = static_cast<const float*>(L.maxs[h].ptr) + slot;
    }
    SCAR(s, L, fuse_prepared(L.car, cnt, s->qviews, images, tiles, maxs));
    SCU(s, cudaEventRecord(L.ev_fused[h], L.main));
    L.used[h] = true;
  }
  s->half ^= 1;
  s->queued = 0;
  s->ramp = false;
  return VCB_OK;
}

int queue_view(vcb_sharded* s, const vcb_view& v, const uint8_t* mask, bool device_mask) {
  int rc = check_view(nullptr, v);
  if (rc != VCB_OK) return sfail(s, rc, "%s", vcb_last_error(nullptr));
  if (!mask) return sfail(s, VCB_ERR_INVALID, "null mask");
  if ((rc = ensure_buffers(s, v.width, v.height)) != VCB_OK) return rc;
  if (s->queued > 0 && s->q_device != device_mask && (rc = flush_chunk(s)) != VCB_OK) return rc;
  s->q_device = device_mask;
  const int idx = s->queued, h = s->half;
  ShardRank* L = local_of(s, idx % s->world);
  if (L && !device_mask) {
    const int j = idx / s->world;
    SCU(s, cudaSetDevice(L->device));
    // the staging half may be overwritten once the 2D SDF kernels that read it have run
    if (j == 0 && L->used[h]) SCU(s, cudaStreamWaitEvent(L->copy, L->ev_sdf[h], 0));
    SCU(s, cudaMemcpyAsync(static_cast<unsigned char*>(L->stage[h].ptr) + (size_t)j * s->stage_pix, mask,
                           (size_t)v.width * v.height, cudaMemcpyHostToDevice, L->copy));
  }
  s->qviews[idx] = v;
  s->qmask[idx] = mask;
  s->queued++;
  int cap = s->chunk_cap;
  if (s->ramp && !s->balance_pending && s->world > 1 && 2 * s->world < s->chunk_cap) {
    // a small first chunk costs one more launch (one more pass over the slab's state) and saves most
    // of the first gather, which nothing can hide: worth it when the slabs are thin (many GPUs)
    // (every rank must take the same decision: the average slab, not this rank's)
    const double state_s = 16.0 * s->nx * s->ny * ((double)s->nz / s->world) / 6.5e12;
    const double gather_s = (double)(s->chunk_cap - 2 * s->world) * s->slot_img * (s->world - 1) /
                            s->world / 7.0e11;
    if (state_s < 0.5 * gather_s) cap = 2 * s->world;
  }
  if (s->queued >= cap) return flush_chunk(s);
  return VCB_OK;
}

// ---- auto-balance from a sample of a long call's views ---------------------------------------------
bool wants_probe(const vcb_sharded* s, int n_views) {
  return s->balance_pending && s->world > 1 && n_views > s->chunk_cap && s->queued == 0;
}
// k-th of the chunk_cap sample views of a call with n_views views (evenly spread)
int probe_index(const vcb_sharded* s, int k, int n_views) {
  return (int)((int64_t)k * n_views / s->chunk_cap);
}
// Transforms and gathers exactly one chunk of sample views (sample k is owned by rank k mod world,
// only owners read their mask), places the slab boundaries from its tile classes and forgets it.
int run_probe(vcb_sharded* s, const vcb_view* sample_views, const uint8_t* const* sample_masks,
              bool device_io) {
  s->probing = true;
  int rc = VCB_OK;
  for (int k = 0; k < s->chunk_cap && rc == VCB_OK; k++) {
    rc = queue_view(s, sample_views[k], sample_masks[k], device_io);  // the last one flushes
  }
  s->probing = false;
  s->queued = 0;
  if (rc != VCB_OK) return rc;
  if (!device_io) {  // the sample's host masks are borrowed until the copies are done
    for (auto& L : s->local) {
      SCU(s, cudaSetDevice(L.device));
      SCU(s, cudaStreamSynchronize(L.copy));
    }
  }
  return VCB_OK;
}

int carve_impl(vcb_sharded* s, int32_t n_views, const vcb_view* views, const uint8_t* const* masks,
               float* const* sdf_out, bool device_io) {
  if (!s) return sfail(nullptr, VCB_ERR_INVALID, "null handle");
  if (n_views < 0 || (n_views > 0 && (!views || !masks))) return sfail(s, VCB_ERR_INVALID, "null argument");
  bool want_sdf = false;
  for (int i = 0; sdf_out && i < n_views; i++) want_sdf = want_sdf || sdf_out[i] != nullptr;
  if (wants_probe(s, n_views) && !want_sdf) {
    // Auto-balance with many views in one call: the first chunk alone may see the grid from one
    // side only (a turntable's first 64 views), so the estimate uses views spread over the call.
    std::vector<vcb_view> sv(s->chunk_cap);
    std::vector<const uint8_t*> sm(s->chunk_cap);
    for (int k = 0; k < s->chunk_cap; k++) {
      sv[k] = views[probe_index(s, k, n_views)];
      sm[k] = masks[probe_index(s, k, n_views)];
    }
    int rc = run_probe(s, sv.data(), sm.data(), device_io);
    if (rc != VCB_OK) return rc;
  }
  for (int i = 0; i < n_views; i++) {
    if (want_sdf && sdf_out[i]) {
      // the caller wants this view's 2D SDF back (the reference's `Image1f* sdf`): the view gets a
      // chunk of its own and the image is read from the first slab's pool after the gather
      int rc = flush_chunk(s);
      if (rc != VCB_OK) return rc;
      const int h = s->half;
      if ((rc = queue_view(s, views[i], masks[i], device_io)) != VCB_OK) return rc;
      if ((rc = flush_chunk(s)) != VCB_OK) return rc;
      ShardRank& L = s->local[0];
      SCU(s, cudaSetDevice(L.device));
      SCU(s, cudaMemcpyAsync(sdf_out[i], L.img[h].ptr, (size_t)views[i].width * views[i].height * sizeof(float),
                             device_io ? cudaMemcpyDeviceToDevice : cudaMemcpyDeviceToHost, L.main));
      if (!device_io) SCU(s, cudaStreamSynchronize(L.main));
    } else {
      int rc = queue_view(s, views[i], masks[i], device_io);
      if (rc != VCB_OK) return rc;
    }
  }
  if (!device_io) {
    // host masks are borrowed for the duration of the call (voxel_carver.cc:394-413)
    for (auto& L : s->local) {
      SCU(s, cudaSetDevice(L.device));
      SCU(s, cudaStreamSynchronize(L.copy));
    }
  } else if (s->queued > 0) {
    // device masks must only stay valid until the streams have passed this call: nothing may be
    // left in the queue that still points at them
    return flush_chunk(s);
  }
  return VCB_OK;
}

// rank r's buffer -> rank r + 1 (`bytes` each), on the main streams; src / dst are per local slab
int shift_up(vcb_sharded* s, const std::vector<const void*>& src, const std::vector<void*>& dst, size_t bytes) {
  if (s->world == 1) return VCB_OK;
  if (s->multi_process) {
    ShardRank& L = s->local[0];
    SNCCL(s, g_nccl.GroupStart());
    if (L.rank + 1 < s->world) SNCCL(s, g_nccl.Send(src[0], bytes, ncclUint8, L.rank + 1, L.comm, L.main));
    if (L.rank > 0) SNCCL(s, g_nccl.Recv(dst[0], bytes, ncclUint8, L.rank - 1, L.comm, L.main));
    SNCCL(s, g_nccl.GroupEnd());
    return VCB_OK;
  }
  for (size_t i = 0; i + 1 < s->local.size(); i++) {
    ShardRank& lo = s->local[i];
    ShardRank& hi = s->local[i + 1];
    SCU(s, cudaSetDevice(lo.device));
    SCU(s, cudaEventRecord(lo.ev_misc, lo.main));
    SCU(s, cudaSetDevice(hi.device));
    SCU(s, cudaStreamWaitEvent(hi.main, lo.ev_misc, 0));
    SCU(s, cudaMemcpyPeerAsync(dst[i + 1], hi.device, src[i], lo.device, bytes, hi.main));
  }
  return VCB_OK;
}

// the two sdf planes below every slab but the first (marching cubes, inside_empty ExtractVoxel)
int exchange_halo(vcb_sharded* s) {
  if (s->world == 1) return VCB_OK;
  const size_t plane = (size_t)s->nx * s->ny;
  std::vector<const void*> src;
  std::vector<void*> dst;
  for (auto& L : s->local) {
    float* sdf = nullptr;
    SCAR(s, L, vcb_grid_device(L.car, &sdf, nullptr));
    const int planes = s->bounds[L.rank + 1] - s->bounds[L.rank];
    src.push_back(sdf + (size_t)(planes - 2) * plane);
    float* halo = nullptr;
    int32_t np = 0;
    SCAR(s, L, vcb_halo_device(L.car, &halo, &np));
    dst.push_back(halo);
  }
  return shift_up(s, src, dst, 2 * plane * sizeof(float));
}

// a pair of counts per slab (mine[2 * i], mine[2 * i + 1] for the i-th local slab) -> the pairs of
// every rank, out[2 * world]
int gather_counts(vcb_sharded* s, const std::vector<int64_t>& local_pairs, std::vector<int64_t>* out) {
  out->assign(2 * (size_t)s->world, 0);
  if (!s->multi_process) {
    for (size_t i = 0; i < s->local.size(); i++) {
      (*out)[2 * s->local[i].rank] = local_pairs[2 * i];
      (*out)[2 * s->local[i].rank + 1] = local_pairs[2 * i + 1];
    }
    return VCB_OK;
  }
  ShardRank& L = s->local[0];
  SCU(s, cudaSetDevice(L.device));
  SCU(s, L.cnt.reserve(sizeof(int64_t) * 2 * s->world));
  int64_t* d = static_cast<int64_t*>(L.cnt.ptr);
  const int64_t mine[2] = {local_pairs[0], local_pairs[1]};
  SCU(s, cudaMemcpyAsync(d + 2 * L.rank, mine, sizeof(mine), cudaMemcpyHostToDevice, L.main));
  SNCCL(s, g_nccl.AllGather(d + 2 * L.rank, d, sizeof(mine), ncclUint8, L.comm, L.main));
  SCU(s, cudaMemcpyAsync(out->data(), d, sizeof(int64_t) * 2 * s->world, cudaMemcpyDeviceToHost, L.main));
  SCU(s, cudaStreamSynchronize(L.main));
  return VCB_OK;
}

void release_rank(ShardRank& L) {
  cudaSetDevice(L.device);
  Scratch* all[] = {&L.img[0], &L.img[1], &L.tiles[0], &L.tiles[1], &L.maxs[0], &L.maxs[1],
                    &L.stage[0], &L.stage[1], &L.sdf_scr, &L.absmax, &L.ids_out, &L.ids_in,
                    &L.cnt, &L.gather, &L.ipc_buf, &L.signal};
  for (Scratch* b : all) b->release();
  for (int h = 0; h < 2; h++) {
    if (L.ev_staged[h]) cudaEventDestroy(L.ev_staged[h]);
    if (L.ev_sdf[h]) cudaEventDestroy(L.ev_sdf[h]);
    if (L.ev_gathered[h]) cudaEventDestroy(L.ev_gathered[h]);
    if (L.ev_fused[h]) cudaEventDestroy(L.ev_fused[h]);
  }
  if (L.ev_misc) cudaEventDestroy(L.ev_misc);
  if (L.ev_signal) cudaEventDestroy(L.ev_signal);
  for (int p = 0; p < ShardRank::kPullStreams; p++) {
    if (L.ev_pulled[p]) cudaEventDestroy(L.ev_pulled[p]);
    if (L.pull[p]) cudaStreamDestroy(L.pull[p]);
  }
  if (L.comm) g_nccl.CommDestroy(L.comm);
  if (L.main) cudaStreamDestroy(L.main);
  if (L.side && L.side != L.main) cudaStreamDestroy(L.side);
  if (L.copy) cudaStreamDestroy(L.copy);
}

int finish_create(vcb_sharded* s, const int32_t* z_bounds, vcb_sharded** out) {
  int rc = init_common(s);
  if (rc == VCB_OK) {
    if (z_bounds) {
      s->bounds.assign(z_bounds, z_bounds + s->world + 1);
    } else {
      default_bounds(s->nz, s->world, &s->bounds);
    }
    rc = check_bounds(s, s->bounds);
  }
  if (rc == VCB_OK) rc = create_carvers(s);
  if (rc != VCB_OK) {
    set_global_error(s->err.c_str());
    vcb_sharded_destroy(s);
    return rc;
  }
  *out = s;
  return VCB_OK;
}

}  // namespace

extern "C" {

const char* vcb_sharded_last_error(const vcb_sharded* s) { return s ? s->err.c_str() : vcb_last_error(nullptr); }

int vcb_sharded_create(const vcb_options* opt, const int32_t* devices, int32_t n_devices,
                       const int32_t* z_bounds, vcb_sharded** out) {
  if (!opt || !out || !devices || n_devices < 1) return sfail(nullptr, VCB_ERR_INVALID, "null argument");
  *out = nullptr;
  int n_dev = 0;
  cudaError_t e = cudaGetDeviceCount(&n_dev);
  if (e != cudaSuccess || n_dev < 1)
    return sfail(nullptr, VCB_ERR_CUDA, "no CUDA device: %s (this library has no CPU fallback)",
                 cudaGetErrorString(e));
  for (int i = 0; i < n_devices; i++) {
    if (devices[i] < 0 || devices[i] >= n_dev)
      return sfail(nullptr, VCB_ERR_INVALID, "device %d not in [0,%d)", devices[i], n_dev);
  }
  DeviceGuard restore_callers_device;
  vcb_sharded* s = new vcb_sharded;
  s->opt = *opt;
  s->world = n_devices;
  s->multi_process = false;
  s->local.resize(n_devices);
  for (int i = 0; i < n_devices; i++) {
    s->local[i].rank = i;
    s->local[i].device = devices[i];
  }
  // direct device-to-device copies over NVLink where the pair allows it (else staged by the driver)
  for (int i = 0; i < n_devices; i++) {
    cudaSetDevice(devices[i]);
    for (int j = 0; j < n_devices; j++) {
      int can = 0;
      if (devices[i] != devices[j] &&
          cudaDeviceCanAccessPeer(&can, devices[i], devices[j]) == cudaSuccess && can) {
        cudaDeviceEnablePeerAccess(devices[j], 0);
      }
    }
  }
  cudaGetLastError();  // "already enabled" is not an error
  return finish_create(s, z_bounds, out);
}

int vcb_sharded_unique_id(void* id128) {
  if (!id128) return sfail(nullptr, VCB_ERR_INVALID, "null argument");
  if (!nccl_load()) return sfail(nullptr, VCB_ERR_UNSUPPORTED, "%s", g_nccl.error.c_str());
  static_assert(sizeof(ncclUniqueId) == VCB_UNIQUE_ID_BYTES, "unique id size");
  ncclUniqueId id;
  ncclResult_t r = g_nccl.GetUniqueId(&id);
  if (r != ncclSuccess) return sfail(nullptr, VCB_ERR_CUDA, "ncclGetUniqueId: %s", g_nccl.GetErrorString(r));
  memcpy(id128, &id, sizeof(id));
  return VCB_OK;
}

int vcb_sharded_create_rank(const vcb_options* opt, int32_t device, int32_t rank, int32_t world,
                            const void* id128, const int32_t* z_bounds, vcb_sharded** out) {
  if (!opt || !out || world < 1 || rank < 0 || rank >= world || (world > 1 && !id128))
    return sfail(nullptr, VCB_ERR_INVALID, "bad argument");
  *out = nullptr;
  int n_dev = 0;
  cudaError_t e = cudaGetDeviceCount(&n_dev);
  if (e != cudaSuccess || n_dev < 1)
    return sfail(nullptr, VCB_ERR_CUDA, "no CUDA device: %s (this library has no CPU fallback)",
                 cudaGetErrorString(e));
  if (device < 0 || device >= n_dev) return sfail(nullptr, VCB_ERR_INVALID, "device %d not in [0,%d)", device, n_dev);
  DeviceGuard restore_callers_device;
  vcb_sharded* s = new vcb_sharded;
  s->opt = *opt;
  s->world = world;
  s->multi_process = world > 1;
  s->ce_gather = env_flag("VACANCY_B200_CE_GATHER", true);
  s->local.resize(1);
  s->local[0].rank = rank;
  s->local[0].device = device;
  if (world > 1) {
    if (!nccl_load()) {
      delete s;
      return sfail(nullptr, VCB_ERR_UNSUPPORTED, "%s", g_nccl.error.c_str());
    }
    ncclUniqueId id;
    memcpy(&id, id128, sizeof(id));
    cudaSetDevice(device);
    // VACANCY_B200_NCCL_CTAS caps the blocks NCCL may use (experiments; with the copy-engine gather
    // NCCL only carries signals and small exchanges).
    ncclResult_t r;
    const char* ctas = getenv("VACANCY_B200_NCCL_CTAS");
    const int max_ctas = ctas ? atoi(ctas) : 0;  // 0: NCCL's own choice
    if (g_nccl.CommInitRankConfig && max_ctas > 0) {
      ncclConfig_t cfg = NCCL_CONFIG_INITIALIZER;
      cfg.minCTAs = 1;
      cfg.maxCTAs = max_ctas;
      r = g_nccl.CommInitRankConfig(&s->local[0].comm, world, id, rank, &cfg);
    } else {
      r = g_nccl.CommInitRank(&s->local[0].comm, world, id, rank);
    }
    if (r != ncclSuccess) {
      delete s;
      return sfail(nullptr, VCB_ERR_CUDA, "ncclCommInitRank: %s", g_nccl.GetErrorString(r));
    }
  }
  return finish_create(s, z_bounds, out);
}

void vcb_sharded_destroy(vcb_sharded* s) {
  if (!s) return;
  DeviceGuard restore_callers_device;
  destroy_carvers(s);
  if (s->multi_process && s->ce_gather && !s->local.empty() && !s->local[0].peer[0][0].empty()) {
    // the ranks unmap each other's pools, meet, and only then free them (every rank destroys its
    // handle at the same point of the program, like every other call on it)
    close_peer_maps(s);
    nccl_barrier(s, s->local[0], s->local[0].side, true);
  }
  for (auto& L : s->local) release_rank(L);
  s->ingest_pin.release();
  delete s;
}

int vcb_sharded_info(const vcb_sharded* s, int32_t* world, int32_t* n_local, int32_t dims3[3]) {
  if (!s) return sfail(nullptr, VCB_ERR_INVALID, "null handle");
  if (world) *world = s->world;
  if (n_local) *n_local = (int32_t)s->local.size();
  if (dims3) {
    dims3[0] = s->nx;
    dims3[1] = s->ny;
    dims3[2] = s->nz;
  }
  return VCB_OK;
}

int vcb_sharded_bounds(const vcb_sharded* s, int32_t* z_bounds) {
  if (!s || !z_bounds) return sfail(nullptr, VCB_ERR_INVALID, "null argument");
  for (int r = 0; r <= s->world; r++) z_bounds[r] = s->bounds[r];
  return VCB_OK;
}

int vcb_sharded_local(vcb_sharded* s, int32_t i, vcb_carver** carver, int32_t* rank) {
  if (!s || i < 0 || i >= (int)s->local.size()) return sfail(s, VCB_ERR_INVALID, "no such local slab");
  if (carver) *carver = s->local[i].car;
  if (rank) *rank = s->local[i].rank;
  return VCB_OK;
}

int vcb_balance_bounds(const uint64_t* layer_cost, int32_t n_layers, int32_t planes_per_layer,
                       int32_t nz, int32_t world, int32_t* z_bounds) {
  if (!layer_cost || !z_bounds || world < 1 || planes_per_layer < 1 || n_layers < world ||
      (int64_t)n_layers * planes_per_layer < nz || (int64_t)(n_layers - 1) * planes_per_layer >= nz)
    return sfail(nullptr, VCB_ERR_INVALID, "vcb_balance_bounds: bad argument");
  uint64_t total = 0;
  for (int i = 0; i < n_layers; i++) total += layer_cost[i];
  z_bounds[0] = 0;
  z_bounds[world] = nz;
  uint64_t run = 0;
  int layer = 0;
  for (int r = 1; r < world; r++) {
    // first layer boundary at which the running cost reaches r / world of the total (or the one
    // before it, whichever is closer), leaving at least one layer to every slab still to come
    const uint64_t target = total / world * r;
    while (layer < n_layers && run + layer_cost[layer] <= target) run += layer_cost[layer++];
    int k = layer;
    if (k < n_layers && target - run > run + layer_cost[k] - target) k++;
    const int lo = z_bounds[r - 1] / planes_per_layer + 1, hi = n_layers - (world - r);
    k = k < lo ? lo : (k > hi ? hi : k);
    z_bounds[r] = k * planes_per_layer;
    while (layer < k) run += layer_cost[layer++];
  }
  return VCB_OK;
}

int vcb_sharded_set_auto_balance(vcb_sharded* s, int32_t enabled) {
  if (!s) return sfail(nullptr, VCB_ERR_INVALID, "null handle");
  s->balance_pending = enabled != 0;
  return VCB_OK;
}

int vcb_sharded_stream(vcb_sharded* s, int32_t i, void** main_stream) {
  if (!s || !main_stream || i < 0 || i >= (int)s->local.size())
    return sfail(s, VCB_ERR_INVALID, "no such local slab");
  *main_stream = s->local[i].main;
  return VCB_OK;
}

int vcb_sharded_repartition(vcb_sharded* s, const int32_t* z_bounds) {
  if (!s) return sfail(nullptr, VCB_ERR_INVALID, "null handle");
  DeviceGuard restore_callers_device;
  std::vector<int> b;
  if (z_bounds) {
    b.assign(z_bounds, z_bounds + s->world + 1);
  } else {
    default_bounds(s->nz, s->world, &b);
  }
  int rc = check_bounds(s, b);
  if (rc != VCB_OK) return rc;
  s->queued = 0;
  s->iso_phase = 0;
  s->total_vox = -1;
  if (b == s->bounds) return vcb_sharded_reset(s);
  destroy_carvers(s);
  s->bounds = b;
  return create_carvers(s);
}

int vcb_sharded_reset(vcb_sharded* s) {
  if (!s) return sfail(nullptr, VCB_ERR_INVALID, "null handle");
  DeviceGuard restore_callers_device;
  s->queued = 0;  // views not yet fused belong to the grid being discarded
  s->ramp = true;
  s->iso_phase = 0;
  s->total_vox = -1;
  for (auto& L : s->local) SCAR(s, L, vcb_reset(L.car));
  return VCB_OK;
}

int vcb_sharded_carve_masks(vcb_sharded* s, int32_t n_views, const vcb_view* views,
                            const uint8_t* const* masks, float* const* sdf_out) {
  DeviceGuard restore_callers_device;
  return carve_impl(s, n_views, views, masks, sdf_out, false);
}

int vcb_sharded_carve_masks_device(vcb_sharded* s, int32_t n_views, const vcb_view* views,
                                   const uint8_t* const* masks_dev, float* const* sdf_out_dev) {
  DeviceGuard restore_callers_device;
  return carve_impl(s, n_views, views, masks_dev, sdf_out_dev, true);
}

int vcb_sharded_carve_sdfs(vcb_sharded* s, int32_t n_views, const vcb_view* views,
                           const float* const* sdfs) {
  if (!s) return sfail(nullptr, VCB_ERR_INVALID, "null handle");
  DeviceGuard restore_callers_device;
  // there is no transform to share out: every slab takes the images as they are, after the views
  // carved before them (the slab's own queue keeps the call order)
  int rc = flush_chunk(s);
  if (rc != VCB_OK) return rc;
  s->balance_pending = false;  // the grid is no longer empty: the boundaries stay where they are
  for (auto& L : s->local) SCAR(s, L, vcb_carve_sdfs(L.car, n_views, views, sdfs));
  return VCB_OK;
}

int vcb_sharded_carve_mask_files(vcb_sharded* s, int32_t n_views, const vcb_view* views,
                                 const char* const* paths, int32_t n_threads) {
  if (!s) return sfail(nullptr, VCB_ERR_INVALID, "null handle");
  if (n_views < 0 || (n_views > 0 && (!views || !paths))) return sfail(s, VCB_ERR_INVALID, "null argument");
  DeviceGuard restore_callers_device;
  size_t stride = 0;
  for (int i = 0; i < n_views; i++) {
    int rc = check_view(nullptr, views[i]);
    if (rc != VCB_OK) return sfail(s, rc, "%s", vcb_last_error(nullptr));
    if (!paths[i]) return sfail(s, VCB_ERR_INVALID, "path %d is null", i);
    if ((rc = ensure_buffers(s, views[i].width, views[i].height)) != VCB_OK) return rc;
    stride = std::max(stride, align256((size_t)views[i].width * views[i].height));
  }
  // One chunk of files at a time, starting on a chunk boundary, so that view i of the batch is
  // owned by rank i mod world: this process decodes only the files whose masks it will upload.
  const int batch = s->chunk_cap;
  SCU(s, s->ingest_pin.reserve(stride * (size_t)std::min(batch, std::max(n_views, 1))));
  uint8_t* pin = static_cast<uint8_t*>(s->ingest_pin.ptr);
  if (wants_probe(s, n_views)) {
    // auto-balance: decode the sample files (each process those it owns in the sample chunk) first
    std::vector<vcb_view> sv(batch);
    std::vector<const char*> sp(batch);
    std::vector<char> mine(batch);
    std::vector<const uint8_t*> sm(batch);
    for (int k = 0; k < batch; k++) {
      sv[k] = views[probe_index(s, k, n_views)];
      sp[k] = paths[probe_index(s, k, n_views)];
      mine[k] = local_of(s, k % s->world) != nullptr;
      sm[k] = pin + (size_t)k * stride;
    }
    std::string err;
    int rc = decode_mask_batch(sp.data(), sv.data(), batch, reinterpret_cast<const bool*>(mine.data()),
                               pin, stride, n_threads, &err);
    if (rc != VCB_OK) return sfail(s, rc, "%s", err.c_str());
    if ((rc = run_probe(s, sv.data(), sm.data(), false)) != VCB_OK) return rc;
  }
  for (int first = 0; first < n_views; first += batch) {
    const int nb = std::min(batch, n_views - first);
    int rc = flush_chunk(s);
    if (rc != VCB_OK) return rc;
    std::vector<char> mine(nb);
    for (int i = 0; i < nb; i++) mine[i] = local_of(s, i % s->world) != nullptr;
    std::string err;
    rc = decode_mask_batch(paths + first, views + first, nb, reinterpret_cast<const bool*>(mine.data()),
                           pin, stride, n_threads, &err);
    if (rc != VCB_OK) return sfail(s, rc, "%s", err.c_str());
    std::vector<const uint8_t*> masks(nb);
    for (int i = 0; i < nb; i++) masks[i] = pin + (size_t)i * stride;  // read only where mine[i]
    if ((rc = carve_impl(s, nb, views + first, masks.data(), nullptr, false)) != VCB_OK) return rc;
  }
  return VCB_OK;
}

int vcb_sharded_flush(vcb_sharded* s) {
  if (!s) return sfail(nullptr, VCB_ERR_INVALID, "null handle");
  DeviceGuard restore_callers_device;
  return flush_chunk(s);
}

int vcb_sharded_synchronize(vcb_sharded* s) {
  if (!s) return sfail(nullptr, VCB_ERR_INVALID, "null handle");
  DeviceGuard restore_callers_device;
  int rc = flush_chunk(s);
  if (rc != VCB_OK) return rc;
  for (auto& L : s->local) {
    SCU(s, cudaSetDevice(L.device));
    SCU(s, cudaStreamSynchronize(L.copy));
    SCU(s, cudaStreamSynchronize(L.side));
    SCU(s, cudaStreamSynchronize(L.main));
  }
  return VCB_OK;
}

int vcb_sharded_extract_iso_count(vcb_sharded* s, double iso_level, int32_t linear_interp,
                                  int64_t* n_vertices, int64_t* n_faces) {
  if (!s) return sfail(nullptr, VCB_ERR_INVALID, "null handle");
  DeviceGuard restore_callers_device;
  s->iso_phase = 0;
  int rc = flush_chunk(s);
  if (rc != VCB_OK) return rc;
  if ((rc = exchange_halo(s)) != VCB_OK) return rc;
  rc = for_each_local(s, [&](ShardRank& L) {
    return vcb_extract_iso_begin(L.car, iso_level, linear_interp, &L.nv, &L.nf);
  });
  if (rc != VCB_OK) return rc;
  {
    std::vector<int64_t> mine;
    for (auto& L : s->local) {
      mine.push_back(L.nv);
      mine.push_back(L.nf);
    }
    if ((rc = gather_counts(s, mine, &s->counts)) != VCB_OK) return rc;
  }
  int64_t tv = 0, tf = 0;
  for (int r = 0; r < s->world; r++) {
    if (ShardRank* L = local_of(s, r)) {
      L->voff = tv;
      L->foff = tf;
    }
    tv += s->counts[2 * r];
    tf += s->counts[2 * r + 1];
  }
  if (tv > INT32_MAX || tf > INT32_MAX)
    return sfail(s, VCB_ERR_UNSUPPORTED, "mesh exceeds int32 indices (mesh.cc:241-271)");
  const size_t id_bytes = (size_t)2 * s->nx * s->ny * sizeof(int32_t);
  rc = for_each_local(s, [&](ShardRank& L) {
    int r2 = vcb_extract_iso_vertices(L.car, L.voff);
    if (r2 != VCB_OK || s->world == 1) return r2;
    cudaSetDevice(L.device);
    if (L.ids_out.reserve(id_bytes) != cudaSuccess || L.ids_in.reserve(id_bytes) != cudaSuccess)
      return fail(L.car, VCB_ERR_CUDA, "out of device memory for the vertex-id planes");
    if (L.rank + 1 < s->world) r2 = vcb_extract_iso_export_top(L.car, static_cast<int32_t*>(L.ids_out.ptr));
    return r2;
  });
  if (rc != VCB_OK) return rc;
  if (s->world > 1) {
    std::vector<const void*> src;
    std::vector<void*> dst;
    for (auto& L : s->local) {
      src.push_back(L.ids_out.ptr);
      dst.push_back(L.ids_in.ptr);
    }
    if ((rc = shift_up(s, src, dst, id_bytes)) != VCB_OK) return rc;
  }
  rc = for_each_local(s, [&](ShardRank& L) {
    int r2 = VCB_OK;
    if (L.rank > 0) r2 = vcb_extract_iso_import_bottom(L.car, static_cast<int32_t*>(L.ids_in.ptr));
    return r2 != VCB_OK ? r2 : vcb_extract_iso_faces(L.car);
  });
  if (rc != VCB_OK) return rc;
  s->total_v = tv;
  s->total_f = tf;
  s->iso_phase = 1;
  if (n_vertices) *n_vertices = tv;
  if (n_faces) *n_faces = tf;
  return VCB_OK;
}

int vcb_sharded_extract_iso_fill(vcb_sharded* s, float* vertices, int32_t* faces) {
  if (!s) return sfail(nullptr, VCB_ERR_INVALID, "null handle");
  if (s->iso_phase != 1) return sfail(s, VCB_ERR_STATE, "vcb_sharded_extract_iso_count has not run");
  DeviceGuard restore_callers_device;
  if (!s->multi_process) {
    // every device writes its part of the caller's arrays over its own PCIe link
    for (auto& L : s->local) {
      SCU(s, cudaSetDevice(L.device));
      const float* dv = nullptr;
      const int32_t* df = nullptr;
      SCAR(s, L, vcb_extract_iso_device(L.car, &dv, &df));
      if (vertices && L.nv > 0)
        SCU(s, cudaMemcpyAsync(vertices + 3 * L.voff, dv, (size_t)L.nv * 12, cudaMemcpyDeviceToHost, L.main));
      if (faces && L.nf > 0)
        SCU(s, cudaMemcpyAsync(faces + 3 * L.foff, df, (size_t)L.nf * 12, cudaMemcpyDeviceToHost, L.main));
    }
    for (auto& L : s->local) {
      SCU(s, cudaSetDevice(L.device));
      SCU(s, cudaStreamSynchronize(L.main));
    }
    return VCB_OK;
  }
  // one process per GPU: the parts travel to rank 0's GPU over NVLink, then to its host
  ShardRank& L = s->local[0];
  SCU(s, cudaSetDevice(L.device));
  const float* dv = nullptr;
  const int32_t* df = nullptr;
  SCAR(s, L, vcb_extract_iso_device(L.car, &dv, &df));
  if (L.rank != 0) {
    SNCCL(s, g_nccl.GroupStart());
    if (L.nv > 0) SNCCL(s, g_nccl.Send(dv, (size_t)L.nv * 12, ncclUint8, 0, L.comm, L.main));
    if (L.nf > 0) SNCCL(s, g_nccl.Send(df, (size_t)L.nf * 12, ncclUint8, 0, L.comm, L.main));
    SNCCL(s, g_nccl.GroupEnd());
    SCU(s, cudaStreamSynchronize(L.main));
    return VCB_OK;
  }
  const size_t others = (size_t)(s->total_v - L.nv + s->total_f - L.nf) * 12;
  SCU(s, L.gather.reserve(others > 0 ? others : 1));
  unsigned char* g = static_cast<unsigned char*>(L.gather.ptr);
  const size_t g_faces = (size_t)(s->total_v - L.nv) * 12;  // faces follow the vertices
  SNCCL(s, g_nccl.GroupStart());
  size_t ov = 0, of = 0;
  for (int r = 1; r < s->world; r++) {
    const size_t bv = (size_t)s->counts[2 * r] * 12, bf = (size_t)s->counts[2 * r + 1] * 12;
    if (bv) SNCCL(s, g_nccl.Recv(g + ov, bv, ncclUint8, r, L.comm, L.main));
    if (bf) SNCCL(s, g_nccl.Recv(g + g_faces + of, bf, ncclUint8, r, L.comm, L.main));
    ov += bv;
    of += bf;
  }
  SNCCL(s, g_nccl.GroupEnd());
  if (vertices) {
    if (L.nv > 0) SCU(s, cudaMemcpyAsync(vertices, dv, (size_t)L.nv * 12, cudaMemcpyDeviceToHost, L.main));
    if (ov > 0) SCU(s, cudaMemcpyAsync(vertices + 3 * L.nv, g, ov, cudaMemcpyDeviceToHost, L.main));
  }
  if (faces) {
    if (L.nf > 0) SCU(s, cudaMemcpyAsync(faces, df, (size_t)L.nf * 12, cudaMemcpyDeviceToHost, L.main));
    if (of > 0) SCU(s, cudaMemcpyAsync(faces + 3 * L.nf, g + g_faces, of, cudaMemcpyDeviceToHost, L.main));
  }
  SCU(s, cudaStreamSynchronize(L.main));
  return VCB_OK;
}

int vcb_sharded_local_mesh(vcb_sharded* s, int32_t i, int64_t* n_vertices, int64_t* n_faces,
                           int64_t* vertex_offset, int64_t* face_offset) {
  if (!s || i < 0 || i >= (int)s->local.size()) return sfail(s, VCB_ERR_INVALID, "no such local slab");
  if (s->iso_phase != 1) return sfail(s, VCB_ERR_STATE, "vcb_sharded_extract_iso_count has not run");
  const ShardRank& L = s->local[i];
  if (n_vertices) *n_vertices = L.nv;
  if (n_faces) *n_faces = L.nf;
  if (vertex_offset) *vertex_offset = L.voff;
  if (face_offset) *face_offset = L.foff;
  return VCB_OK;
}

int vcb_sharded_extract_voxel_count(vcb_sharded* s, int32_t inside_empty, int64_t* n_voxels) {
  if (!s) return sfail(nullptr, VCB_ERR_INVALID, "null handle");
  DeviceGuard restore_callers_device;
  s->total_vox = -1;
  int rc = flush_chunk(s);
  if (rc != VCB_OK) return rc;
  // the surface test looks one plane down (extract_voxel.cc:15-79): the halo serves it
  if (inside_empty && (rc = exchange_halo(s)) != VCB_OK) return rc;
  rc = for_each_local(s, [&](ShardRank& L) {
    return vcb_extract_voxel_count(L.car, inside_empty, &L.nvox);
  });
  if (rc != VCB_OK) return rc;
  std::vector<int64_t> mine;  // cube counts of every rank (the same exchange as the mesh counts)
  for (auto& L : s->local) {
    mine.push_back(L.nvox);
    mine.push_back(0);
  }
  if ((rc = gather_counts(s, mine, &s->vox_counts)) != VCB_OK) return rc;
  int64_t total = 0;
  for (int r = 0; r < s->world; r++) {
    if (ShardRank* L = local_of(s, r)) L->vox_off = total;
    total += s->vox_counts[2 * r];
  }
  s->total_vox = total;
  if (n_voxels) *n_voxels = total;
  return VCB_OK;
}

int vcb_sharded_extract_voxel_fill(vcb_sharded* s, float* vertices, int32_t* faces) {
  if (!s || !vertices || !faces) return sfail(s, VCB_ERR_INVALID, "null argument");
  if (s->total_vox < 0) return sfail(s, VCB_ERR_STATE, "vcb_sharded_extract_voxel_count has not run");
  if (s->total_vox * 24 > INT32_MAX)
    return sfail(s, VCB_ERR_UNSUPPORTED, "voxel mesh exceeds int32 indices (mesh.cc:241-271)");
  if (s->multi_process)
    return sfail(s, VCB_ERR_UNSUPPORTED,
                 "with one process per GPU each rank fills its own cubes: use vcb_sharded_local + "
                 "vcb_extract_voxel_fill and add 24 * (cubes of the lower slabs) to the face indices");
  DeviceGuard restore_callers_device;
  int rc = for_each_local(s, [&](ShardRank& L) {
    if (L.nvox == 0) return (int)VCB_OK;
    float* v = vertices + L.vox_off * 72;
    int32_t* f = faces + L.vox_off * 36;
    int r2 = vcb_extract_voxel_fill(L.car, v, f);
    if (r2 != VCB_OK) return r2;
    // each slab numbers its cubes from 0 (extract_voxel.cc:289-311 counts over the whole grid)
    const int32_t shift = (int32_t)(L.vox_off * 24);
    if (shift != 0)
      for (int64_t k = 0; k < L.nvox * 36; k++) f[k] += shift;
    return (int)VCB_OK;
  });
  return rc;
}

int vcb_sharded_save_grid(vcb_sharded* s, const char* prefix) {
  if (!s || !prefix) return sfail(s, VCB_ERR_INVALID, "null argument");
  DeviceGuard restore_callers_device;
  int rc = flush_chunk(s);
  if (rc != VCB_OK) return rc;
  return for_each_local(s, [&](ShardRank& L) {
    char path[4096];
    snprintf(path, sizeof(path), "%s.z%05d-%05d.vcbgrid", prefix, s->bounds[L.rank],
             s->bounds[L.rank + 1]);
    return vcb_save_grid(L.car, path);
  });
}

int vcb_sharded_load_grid(vcb_sharded* s, const char* const* paths, int32_t n_paths) {
  if (!s || !paths || n_paths < 1) return sfail(s, VCB_ERR_INVALID, "null argument");
  DeviceGuard restore_callers_device;
  int rc = flush_chunk(s);
  if (rc != VCB_OK) return rc;
  s->balance_pending = false;  // a restored grid must not be re-partitioned away
  s->iso_phase = 0;
  s->total_vox = -1;
  return for_each_local(s, [&](ShardRank& L) {
    int covered = 0;
    for (int32_t i = 0; i < n_paths; i++) {
      int32_t got = 0;
      const int r2 = vcb_load_grid(L.car, paths[i], &got);
      if (r2 != VCB_OK) return r2;
      covered += got;
    }
    const int planes = s->bounds[L.rank + 1] - s->bounds[L.rank];
    if (covered != planes)
      return fail(L.car, VCB_ERR_INVALID, "the checkpoint files cover %d of the slab's %d planes",
                  covered, planes);
    return (int)VCB_OK;
  });
}

int vcb_sharded_stage_times(vcb_sharded* s, int32_t i, float ms[4], int64_t launches[4]) {
  if (!s || i < 0 || i >= (int)s->local.size()) return sfail(s, VCB_ERR_INVALID, "no such local slab");
  DeviceGuard restore_callers_device;
  ShardRank& L = s->local[i];
  SCU(s, cudaSetDevice(L.device));
  SCU(s, cudaStreamSynchronize(L.side));
  SCAR(s, L, vcb_stage_times(L.car, ms, launches));
  return VCB_OK;
}

}  // extern "C"
