"""z-slab sharding of one voxel grid across GPUs: thin Python mirror of the
vcb_sharded_* part of the C ABI (csrc/sharded.cu), where the whole engine lives - the split of the
2D SDF work across the GPUs, the all-gather of the SDF images, the halo / count / vertex-id
exchanges of marching cubes and the gather of the mesh.

The reference is single-process; what is mirrored is its result: the slabs together produce exactly
the grid, and exactly the marching-cubes mesh (same vertex numbering, same face order), that one
carver over the whole grid produces.

  `LocalSlabs`     ONE process drives all slabs (vcb_sharded_create; peer copies between the
                   devices).  With the same device listed N times the slabs are "virtual": the tests
                   use that to prove the slab algorithm on a single GPU.
  `ShardedCarver`  one process per GPU under torch.distributed (vcb_sharded_create_rank; NCCL inside
                   the library).  torch only carries the NCCL unique id from rank 0 to the other
                   ranks and the slab costs for re-balancing - no tensor of the data path is a torch
                   tensor.
"""
import ctypes as C

import numpy as np

from . import capi
from .carver import VoxelCarver, make_view


def slab_ranges(nz, world):
    """Contiguous, balanced z ranges; every slab gets at least 2 planes (the halo depth)."""
    if world < 1 or nz < 2 * world:
        raise ValueError("cannot split %d planes into %d slabs of >= 2 planes" % (nz, world))
    base, extra = divmod(nz, world)
    out, z = [], 0
    for r in range(world):
        n = base + (1 if r < extra else 0)
        out.append((z, z + n))
        z += n
    return out


FUSION_TILE_PLANES = 8  # z planes per fusion tile (kZT in csrc/fusion.cu): a slab pays for whole tiles


def rebalanced_ranges(ranges, cost, min_planes=2, quantum=1):
    """New contiguous z ranges with (approximately) equal cost, given the measured cost of each
    current slab (e.g. its fusion time); the cost per plane is taken as constant inside a slab.
    With quantum > 1 the boundaries are multiples of `quantum` planes and a slab's cost is spread
    over the whole tiles it pays for (the fusion kernel works on tiles of FUSION_TILE_PLANES planes
    counted from the slab's first plane, so planes short of a whole tile are wasted work).
    Deterministic, so every rank computes the same partition from the gathered costs."""
    world = len(ranges)
    nz = ranges[-1][1]
    q = max(1, int(quantum))
    if nz < world * max(min_planes, q):
        q = 1  # too few planes to hand out whole tiles
    density = np.empty(nz, np.float64)
    for (a, b), c in zip(ranges, cost):
        paid = -(-(b - a) // q) * q
        density[a:b] = max(float(c), 1e-9) / paid
    cum = np.concatenate([[0.0], np.cumsum(density)])
    step = max(min_planes, q)
    bounds = [0]
    for r in range(1, world):
        target = cum[-1] * r / world
        z = int(np.searchsorted(cum, target))
        if q > 1:
            z = int(round(z / q)) * q
        z = max(z, bounds[-1] + step)
        z = min(z, (nz - step * (world - r)) // q * q if q > 1 else nz - step * (world - r))
        bounds.append(z)
    bounds.append(nz)
    return [(bounds[i], bounds[i + 1]) for i in range(world)]


def exclusive_offsets(counts):
    """counts[r] = (vertices, faces) of slab r -> per-slab (vertex_offset, face_offset), totals."""
    counts = np.asarray(counts, np.int64).reshape(-1, 2)
    off = np.zeros_like(counts)
    off[1:] = np.cumsum(counts, axis=0)[:-1]
    return off, counts.sum(axis=0)


def gather_offsets(nv, nf, world, group=None, device="cpu"):
    """all_gather of every slab's (vertices, faces) -> (per-slab offsets, totals)."""
    import torch
    import torch.distributed as dist
    mine = torch.tensor([nv, nf], dtype=torch.int64, device=device)
    allc = [torch.empty_like(mine) for _ in range(world)]
    dist.all_gather(allc, mine, group=group)
    return exclusive_offsets(torch.stack(allc).cpu().numpy())


def agreed_ranges(ranges, my_cost, world, group=None, device="cpu"):
    """Every rank contributes the measured cost of its slab; all ranks get the same new partition
    (whole fusion tiles) from the gathered costs."""
    import torch
    import torch.distributed as dist
    mine = torch.tensor([float(my_cost)], dtype=torch.float64, device=device)
    allc = [torch.empty_like(mine) for _ in range(world)]
    dist.all_gather(allc, mine, group=group)
    return rebalanced_ranges(ranges, [float(t.item()) for t in allc], quantum=FUSION_TILE_PLANES)


def assemble_ply(totals, parts):
    """totals = global (vertices, faces); parts = per-slab (vertex_text, face_text) in slab order."""
    from . import capi as _capi
    lib = _capi.load_library()
    n = C.c_size_t()
    lib.vcb_ply_header(int(totals[0]), int(totals[1]), None, 0, C.byref(n))
    head = C.create_string_buffer(n.value)
    lib.vcb_ply_header(int(totals[0]), int(totals[1]), head, n.value, None)
    return head.raw + b"".join(p[0] for p in parts) + b"".join(p[1] for p in parts)


def _options_struct(option):
    from .carver import _fill_options
    lib = capi.load_library()
    o = capi.Options()
    lib.vcb_options_default(C.byref(o))
    _fill_options(o, option, 0, None)
    return o


def _bounds_array(ranges):
    b = [ranges[0][0]] + [r[1] for r in ranges]
    return (C.c_int32 * len(b))(*b)


class _Engine:
    """Owns a vcb_sharded handle; what LocalSlabs and ShardedCarver share."""

    def __init__(self):
        self._lib = capi.load_library()
        self._h = None
        self._pin = None

    def _check(self, rc):
        capi.check_sharded(self._lib, self._h, rc)

    def _after_create(self, option):
        self._option = option
        w, nl = C.c_int32(), C.c_int32()
        d = (C.c_int32 * 3)()
        self._check(self._lib.vcb_sharded_info(self._h, C.byref(w), C.byref(nl), d))
        self.world, self.n_local, self.dims = w.value, nl.value, (d[0], d[1], d[2])
        self._refresh()

    def _refresh(self):
        b = (C.c_int32 * (self.world + 1))()
        self._check(self._lib.vcb_sharded_bounds(self._h, b))
        self.ranges = [(b[i], b[i + 1]) for i in range(self.world)]
        self.slabs, self.local_ranks = [], []
        for i in range(self.n_local):
            h, r = C.c_void_p(), C.c_int32()
            self._check(self._lib.vcb_sharded_local(self._h, i, C.byref(h), C.byref(r)))
            self.slabs.append(VoxelCarver.borrowed(h.value, self._option))
            self.local_ranks.append(r.value)

    def set_auto_balance(self, enabled=True):
        """Let the engine place the slab boundaries itself from the first chunk of views of a fresh
        grid (vcb_sharded_set_auto_balance); call before the first carve."""
        self._check(self._lib.vcb_sharded_set_auto_balance(self._h, int(bool(enabled))))
        self._auto_balance = bool(enabled)

    def _refresh_if_moved(self):
        # with auto-balance the first carve call may re-create the slabs on other boundaries
        if getattr(self, "_auto_balance", False):
            b = (C.c_int32 * (self.world + 1))()
            self._check(self._lib.vcb_sharded_bounds(self._h, b))
            if [(b[i], b[i + 1]) for i in range(self.world)] != self.ranges:
                self.slabs = []
                self._refresh()
            self._auto_balance = False  # placed once

    def close(self):
        if getattr(self, "_h", None):
            self.slabs = []
            self._lib.vcb_sharded_destroy(self._h)
            self._h = None

    def __del__(self):
        self.close()

    # ---- carving ----------------------------------------------------------------------------------
    def reset(self):
        self._check(self._lib.vcb_sharded_reset(self._h))

    def make_views(self, cameras, width, height):
        return [make_view(c, width, height) for c in cameras]

    def carve_views(self, views, masks, device=False):
        """masks: numpy uint8 arrays (HOST) or integers (DEVICE pointers readable from the owning
        GPU); views fused in call order, 64 per launch."""
        n = len(views)
        if n == 0:
            return
        # hosts that carve the same views again and again hand in ready-made ctypes arrays
        # (pack_views / pack_masks) and skip the per-call marshalling
        arr = views if isinstance(views, C.Array) else (capi.View * n)(*views)
        ptrs = masks if isinstance(masks, C.Array) else self.pack_masks(masks)
        fn = (self._lib.vcb_sharded_carve_masks_device if device
              else self._lib.vcb_sharded_carve_masks)
        self._check(fn(self._h, n, arr, ptrs, None))
        self._refresh_if_moved()

    def carve_files(self, cameras, paths, size, n_threads=0):
        """Silhouette FILES of `size` = (width, height) (vcb_sharded_carve_mask_files): every
        process decodes only the files whose masks it uploads."""
        n = len(paths)
        views = (capi.View * n)(*[make_view(c, size[0], size[1]) for c in cameras])
        arr = (C.c_char_p * n)(*[str(p).encode() for p in paths])
        self._check(self._lib.vcb_sharded_carve_mask_files(self._h, n, views, arr, int(n_threads)))
        self._refresh_if_moved()
        return True

    @staticmethod
    def pack_views(views):
        return (capi.View * len(views))(*views)

    @staticmethod
    def pack_masks(masks):
        ptrs = (C.c_void_p * len(masks))()
        for i, m in enumerate(masks):
            ptrs[i] = m if isinstance(m, int) else m.ctypes.data
        return ptrs

    def carve_batch(self, cameras, images):
        imgs = [np.ascontiguousarray(i, np.uint8) for i in images]
        views = [make_view(c, m.shape[1], m.shape[0]) for c, m in zip(cameras, imgs)]
        self.carve_views(views, imgs)
        return True

    def flush(self):
        self._check(self._lib.vcb_sharded_flush(self._h))

    def synchronize(self):
        self._check(self._lib.vcb_sharded_synchronize(self._h))

    # ---- extraction -------------------------------------------------------------------------------
    def extract_iso_counts(self, iso=0.0, interp=True):
        """Runs marching cubes over the whole grid, every slab's part of the mesh staying on its
        GPU; returns the GLOBAL (vertices, faces) counts (the same on every rank)."""
        nv, nf = C.c_int64(), C.c_int64()
        self._check(self._lib.vcb_sharded_extract_iso_count(self._h, float(iso), int(interp),
                                                            C.byref(nv), C.byref(nf)))
        self._totals = (nv.value, nf.value)
        return self._totals

    def _pinned(self, nbytes):
        import torch
        if self._pin is None or self._pin.numel() < nbytes:
            self._pin = torch.empty(max(nbytes, 1) * 5 // 4, dtype=torch.uint8, pin_memory=True)
        return self._pin.numpy()

    def mesh_host(self, receive=True):
        """The WHOLE mesh of the last extract_iso_counts on the host: (vertices [V,3] float32, faces
        [F,3] int32), views of a pinned buffer that is reused by the next call.  With one process
        per GPU only rank 0 receives it (receive=False on the others, which just send their
        part and get None)."""
        nv, nf = self._totals
        if not receive:
            self._check(self._lib.vcb_sharded_extract_iso_fill(self._h, None, None))
            return None
        raw = self._pinned((nv + nf) * 12)
        v = raw[:nv * 12].view(np.float32).reshape(nv, 3)
        f = raw[nv * 12:(nv + nf) * 12].view(np.int32).reshape(nf, 3)
        self._check(self._lib.vcb_sharded_extract_iso_fill(
            self._h, v.ctypes.data_as(C.POINTER(C.c_float)), f.ctypes.data_as(C.POINTER(C.c_int32))))
        return v, f

    def local_mesh_counts(self, i=0):
        a = [C.c_int64() for _ in range(4)]
        self._check(self._lib.vcb_sharded_local_mesh(self._h, i, *[C.byref(x) for x in a]))
        return tuple(x.value for x in a)  # vertices, faces, vertex offset, face offset

    def extract_voxel_count(self, inside_empty=False):
        n = C.c_int64()
        self._check(self._lib.vcb_sharded_extract_voxel_count(self._h, int(inside_empty), C.byref(n)))
        return n.value

    def save_grid(self, prefix):
        """One checkpoint file per local slab: <prefix>.z<begin>-<end>.vcbgrid; returns the paths."""
        self._check(self._lib.vcb_sharded_save_grid(self._h, str(prefix).encode()))
        return ["%s.z%05d-%05d.vcbgrid" % (prefix, *self.ranges[r]) for r in self.local_ranks]

    def load_grid(self, paths):
        """Restores every local slab from checkpoint files that together cover it (any partition)."""
        arr = (C.c_char_p * len(paths))(*[str(p).encode() for p in paths])
        self._check(self._lib.vcb_sharded_load_grid(self._h, arr, len(paths)))

    def main_stream(self, i=0):
        """cudaStream_t (as an integer) that orders slab i's reset, fusion and extraction."""
        p = C.c_void_p()
        self._check(self._lib.vcb_sharded_stream(self._h, i, C.byref(p)))
        return p.value or 0

    def stage_times(self, i=0):
        ms = (C.c_float * 4)()
        n = (C.c_int64 * 4)()
        self._check(self._lib.vcb_sharded_stage_times(self._h, i, ms, n))
        names = ("sdf2d", "fusion", "extract_voxel", "marching_cubes")
        return {k: (ms[j], n[j]) for j, k in enumerate(names)}


class LocalSlabs(_Engine):
    """N slabs of one grid driven by ONE process (vcb_sharded_create).  `devices` lists the CUDA
    device of every slab; the default puts all `n_slabs` on one GPU ("virtual slabs"), which runs
    the multi-GPU algorithm - 2D SDF split, gathers, halo and id exchanges - on a single device."""

    def __init__(self, option, n_slabs, device=0, devices=None, ranges=None):
        super().__init__()
        devices = list(devices) if devices is not None else [device] * n_slabs
        o = _options_struct(option)
        h = C.c_void_p()
        dev = (C.c_int32 * len(devices))(*devices)
        rc = self._lib.vcb_sharded_create(C.byref(o), dev, len(devices),
                                          _bounds_array(ranges) if ranges else None, C.byref(h))
        if rc != 0:
            raise capi.VcbError(rc, self._lib.vcb_last_error(None).decode())
        self._h = h
        self._after_create(option)

    @property
    def car(self):
        return self.slabs[0]

    @property
    def slab_voxels(self):
        return self.slabs[0].slab_voxels

    def rebalance(self):
        """Re-partitions the grid so that every slab carries about the same fusion time, measured
        since the last call of stage_times (the grid state is discarded, like Init())."""
        cost = [self.stage_times(i)["fusion"][0] for i in range(self.n_local)]
        new = rebalanced_ranges(self.ranges, cost, quantum=FUSION_TILE_PLANES)
        if new != self.ranges:
            self.slabs = []
            self._check(self._lib.vcb_sharded_repartition(self._h, _bounds_array(new)))
            self._refresh()
        return self.ranges

    def download_grid(self):
        self.flush()
        parts = [s.download_grid() for s in self.slabs]
        return np.concatenate([p[0] for p in parts]), np.concatenate([p[1] for p in parts])

    def extract_iso(self, iso=0.0, interp=True):
        self.extract_iso_counts(iso, interp)
        v, f = self.mesh_host()
        return v.copy(), f.copy()

    def ply_text(self):
        """ASCII PLY of the mesh of the last extract_iso, assembled from the slabs' text blocks:
        header for the global counts, every slab's vertex lines, then every slab's face lines."""
        return assemble_ply(self._totals, [s.ply_text(header=False, split=True) for s in self.slabs])

    def occupancy(self, inside_empty=False):
        total = self.extract_voxel_count(inside_empty)
        flags = []
        for s in self.slabs:
            buf = np.empty(s.slab_voxels, np.uint8)
            s._check(s._lib.vcb_extract_voxel_occupancy(
                s._h, buf.ctypes.data_as(C.POINTER(C.c_uint8))))
            flags.append(buf)
        return np.concatenate(flags), total

    def extract_voxel(self, inside_empty=False):
        n = self.extract_voxel_count(inside_empty)
        v = np.empty((n * 24, 3), np.float32)
        f = np.empty((n * 12, 3), np.int32)
        if n:
            self._check(self._lib.vcb_sharded_extract_voxel_fill(
                self._h, v.ctypes.data_as(C.POINTER(C.c_float)),
                f.ctypes.data_as(C.POINTER(C.c_int32))))
        return v, f


def broadcast_unique_id(rank, world, group=None, device=None):
    """The 128-byte NCCL id of vcb_sharded_unique_id, made on rank 0 and broadcast through
    torch.distributed (on `device` for an NCCL process group, on the CPU for gloo)."""
    import torch
    import torch.distributed as dist
    lib = capi.load_library()
    buf = (C.c_uint8 * capi.UNIQUE_ID_BYTES)()
    if rank == 0:
        capi.check(lib, None, lib.vcb_sharded_unique_id(buf))
    t = torch.tensor(list(buf), dtype=torch.uint8, device=device if device is not None else "cpu")
    if world > 1:
        dist.broadcast(t, src=0, group=group)
    return bytes(t.cpu().tolist())


class ShardedCarver(_Engine):
    """This rank's slab of the grid, one process per GPU (world == 1: the whole grid)."""

    def __init__(self, option, rank, world, device, stream=None, group=None, ranges=None):
        super().__init__()
        import torch
        self.rank, self.group, self.torch = rank, group, torch
        self.device_index = device
        o = _options_struct(option)
        uid = None
        if world > 1:
            on = torch.device("cuda", device) if torch.distributed.get_backend(group) == "nccl" else None
            uid = broadcast_unique_id(rank, world, group, on)
        h = C.c_void_p()
        rc = self._lib.vcb_sharded_create_rank(C.byref(o), device, rank, world, uid,
                                               _bounds_array(ranges) if ranges else None, C.byref(h))
        if rc != 0:
            raise capi.VcbError(rc, self._lib.vcb_last_error(None).decode())
        self._h = h
        self._after_create(option)

    @property
    def car(self):
        return self.slabs[0]

    @property
    def slab_voxels(self):
        return self.slabs[0].slab_voxels

    def rebalance(self, my_cost):
        """Re-partitions the grid so that every slab carries about the same measured cost (call
        it with this rank's fusion time of a representative pass). The grid state is discarded,
        like Init(). Returns the new ranges."""
        if self.world == 1:
            return self.ranges
        torch = self.torch
        on = torch.device("cuda", self.device_index) if torch.distributed.get_backend(self.group) == "nccl" else "cpu"
        new = agreed_ranges(self.ranges, my_cost, self.world, self.group, on)
        if new != self.ranges:
            self.slabs = []
            self._check(self._lib.vcb_sharded_repartition(self._h, _bounds_array(new)))
            self._refresh()
        return self.ranges

    def carve_device_masks(self, views, dev_masks):
        """dev_masks: torch uint8 [V, H, W] on this rank's GPU (every rank holds all masks; only a
        view's owner reads its mask)."""
        stride = dev_masks.stride(0)
        base = dev_masks.data_ptr()
        self.carve_views(views, [base + i * stride for i in range(dev_masks.shape[0])], device=True)

    def carve_host_masks(self, cameras, host_masks):
        """Masks in HOST memory (pinned for asynchronous copies): each rank uploads only the masks of
        the views it owns, over its own PCIe link."""
        h, w = host_masks[0].shape
        self.carve_views(self.make_views(cameras, w, h), host_masks)

    def extract_iso_host(self, iso=0.0, interp=True):
        """Extraction + the WHOLE mesh on rank 0's host (None on the other ranks)."""
        self.extract_iso_counts(iso, interp)
        return self.mesh_host(receive=self.rank == 0)

    def ply_parts(self):
        """(vertex_text, face_text) of THIS rank's part of the last extracted mesh (global ids)."""
        return self.car.ply_text(header=False, split=True)

    def ply_text_on_rank0(self, totals):
        """Gathers every rank's text blocks on rank 0 and returns the whole file there (None on
        the other ranks); totals = the global counts extract_iso_counts returned."""
        mine = self.ply_parts()
        if self.world == 1:
            return assemble_ply(totals, [mine])
        parts = [None] * self.world if self.rank == 0 else None
        self.torch.distributed.gather_object(mine, parts, dst=0, group=self.group)
        return assemble_ply(totals, parts) if self.rank == 0 else None
