"""z-slab sharding of one voxel grid across GPUs.

The reference is single-process; what is mirrored here is its result: the slabs together must
produce exactly the grid, and exactly the marching-cubes mesh (same vertex numbering, same face
order), that one carver over the whole grid produces.

* fusion needs no exchange: every slab fuses every view into its own planes;
* marching cubes needs (1) the two sdf planes below each slab (halo), (2) the per-slab vertex / face
  counts to turn local numbers into global ones, (3) the ids of the vertices lying on the slab's
  top plane, handed to the slab above.

Two drivers share the same C-ABI phases:
  `LocalSlabs`     all slabs in ONE process on one GPU (device-to-device copies) - used by the tests
                   to prove the slab algorithm equals the unsharded result on a single GPU;
  `ShardedCarver`  one process per GPU under torch.distributed (NCCL over NVLink; gloo for the
                   CPU tests of the host logic): send/recv of halo planes and id planes,
                   all_gather of the counts.
torch is plumbing only (device buffers for NCCL, streams, the process group).
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


def assemble_ply(totals, parts):
    """totals = global (vertices, faces); parts = per-slab (vertex_text, face_text) in slab order."""
    from . import capi as _capi
    lib = _capi.load_library()
    n = C.c_size_t()
    lib.vcb_ply_header(int(totals[0]), int(totals[1]), None, 0, C.byref(n))
    head = C.create_string_buffer(n.value)
    lib.vcb_ply_header(int(totals[0]), int(totals[1]), head, n.value, None)
    return head.raw + b"".join(p[0] for p in parts) + b"".join(p[1] for p in parts)


class _DevArray:
    """Exposes raw device memory of the carver to torch through __cuda_array_interface__."""

    def __init__(self, ptr, shape, typestr):
        self.__cuda_array_interface__ = {"shape": tuple(shape), "typestr": typestr,
                                         "data": (int(ptr), False), "version": 2}


def device_tensor(ptr, shape, typestr, device):
    import torch
    return torch.as_tensor(_DevArray(ptr, shape, typestr), device=device)


class _Slab:
    """One carver + torch views of its state / halo, and the phase calls."""

    def __init__(self, option, z_range, device, stream=None):
        import torch
        self.car = VoxelCarver(option, device=device, z_range=z_range)
        if not self.car.Init():
            raise capi.VcbError(-1, self.car.last_error)
        self.device = torch.device("cuda", device)
        self.lib, self.h = self.car._lib, self.car._h
        if stream is not None:
            self.car.set_stream(stream.cuda_stream)
        nx, ny, nz = self.car.dims
        self.nx, self.ny = nx, ny
        self.z0, self.z1 = self.car.z_range
        sdf_ptr, _ = self.car.grid_device()
        self.sdf = device_tensor(sdf_ptr, (self.z1 - self.z0, ny, nx), "<f4", self.device)
        hp, hn = C.c_void_p(), C.c_int32()
        capi.check(self.lib, self.h, self.lib.vcb_halo_device(self.h, C.byref(hp), C.byref(hn)))
        self.halo_planes = hn.value
        self.halo = (device_tensor(hp.value, (hn.value, ny, nx), "<f4", self.device)
                     if hn.value else None)
        self.ids_out = torch.empty((2, ny, nx), dtype=torch.int32, device=self.device)
        self.ids_in = torch.empty((2, ny, nx), dtype=torch.int32, device=self.device)

    def top_planes(self, n):
        return self.sdf[self.sdf.shape[0] - n:]

    def begin(self, iso, interp):
        nv, nf = C.c_int64(), C.c_int64()
        capi.check(self.lib, self.h, self.lib.vcb_extract_iso_begin(
            self.h, float(iso), int(interp), C.byref(nv), C.byref(nf)))
        return nv.value, nf.value

    def vertices(self, offset):
        capi.check(self.lib, self.h, self.lib.vcb_extract_iso_vertices(self.h, int(offset)))

    def export_top(self):
        capi.check(self.lib, self.h, self.lib.vcb_extract_iso_export_top(
            self.h, C.c_void_p(self.ids_out.data_ptr())))
        return self.ids_out

    def import_bottom(self, ids):
        capi.check(self.lib, self.h, self.lib.vcb_extract_iso_import_bottom(
            self.h, C.c_void_p(ids.data_ptr())))

    def faces(self):
        capi.check(self.lib, self.h, self.lib.vcb_extract_iso_faces(self.h))

    def mesh_host(self, nv, nf):
        """This slab's mesh on the host.  The arrays are views of a pinned buffer the slab keeps and
        reuses (pageable destinations make the copy several times slower): valid until the next
        call, copy them to keep them."""
        import torch
        need = (nv + nf) * 12
        if getattr(self, "_mesh_pin", None) is None or self._mesh_pin.numel() < need:
            self._mesh_pin = torch.empty(max(need, 1) * 5 // 4, dtype=torch.uint8, pin_memory=True)
        raw = self._mesh_pin.numpy()
        v = raw[:nv * 12].view(np.float32).reshape(nv, 3)
        f = raw[nv * 12:need].view(np.int32).reshape(nf, 3)
        capi.check(self.lib, self.h, self.lib.vcb_extract_iso_fill(
            self.h, v.ctypes.data_as(C.POINTER(C.c_float)), f.ctypes.data_as(C.POINTER(C.c_int32))))
        return v, f


class LocalSlabs:
    """N slabs of one grid in one process on one GPU: the multi-GPU algorithm with copies in place
    of NCCL.  `extract_iso` returns the concatenated mesh, which must equal the unsharded one."""

    def __init__(self, option, n_slabs, device=0):
        probe = VoxelCarver(option, device=device)
        if not probe.Init():
            raise capi.VcbError(-1, probe.last_error)
        self.dims = probe.dims
        probe.close()
        self.slabs = [_Slab(option, zr, device) for zr in slab_ranges(self.dims[2], n_slabs)]

    def carve_batch(self, cameras, images):
        for s in self.slabs:
            assert s.car.carve_batch(cameras, images)

    def download_grid(self):
        parts = [s.car.download_grid() for s in self.slabs]
        return np.concatenate([p[0] for p in parts]), np.concatenate([p[1] for p in parts])

    def extract_iso(self, iso=0.0, interp=True):
        import torch
        for lo, hi in zip(self.slabs[:-1], self.slabs[1:]):
            lo.car.flush()
            hi.car.flush()
            torch.cuda.synchronize()
            hi.halo.copy_(lo.top_planes(hi.halo_planes))
        torch.cuda.synchronize()
        counts = [s.begin(iso, interp) for s in self.slabs]
        off, total = exclusive_offsets(counts)
        for s, o in zip(self.slabs, off):
            s.vertices(o[0])
        torch.cuda.synchronize()
        for lo, hi in zip(self.slabs[:-1], self.slabs[1:]):
            ids = lo.export_top()
            torch.cuda.synchronize()
            hi.import_bottom(ids)
        for s in self.slabs:
            s.faces()
        parts = [s.mesh_host(*c) for s, c in zip(self.slabs, counts)]
        self._totals = (int(total[0]), int(total[1]))
        return np.concatenate([p[0] for p in parts]), np.concatenate([p[1] for p in parts])

    def ply_text(self):
        """ASCII PLY of the mesh of the last extract_iso, assembled from the slabs' text blocks:
        header for the global counts, every slab's vertex lines, then every slab's face lines."""
        return assemble_ply(self._totals, [s.car.ply_text(header=False, split=True)
                                           for s in self.slabs])

    def occupancy(self, inside_empty=False):
        parts = [s.car.occupancy(inside_empty) for s in self.slabs]
        return np.concatenate([p[0] for p in parts]), sum(p[1] for p in parts)


class ShardedCarver:
    """This rank's slab of the grid under torch.distributed (world == 1: the whole grid)."""

    def __init__(self, option, rank, world, device, stream=None, group=None):
        import torch
        self.rank, self.world, self.group = rank, world, group
        self.torch = torch
        # number of z planes as VoxelGrid::Init computes it (float32, voxel_carver.cc:294-296)
        nz = int((np.float32(option.bb_max[2]) - np.float32(option.bb_min[2])) /
                 np.float32(option.resolution))
        self.option, self.device_index = option, device
        self.ranges = slab_ranges(nz, world) if world > 1 else [(0, nz)]
        self.stream = stream if stream is not None else torch.cuda.current_stream()
        self._build()

    def _build(self):
        self.z_range = self.ranges[self.rank] if self.world > 1 else None
        self.slab = _Slab(self.option, self.z_range, self.device_index, self.stream)
        self.car = self.slab.car
        self.dims = self.car.dims
        self.slab_voxels = self.car.slab_voxels
        self._counts = None

    def rebalance(self, my_cost):
        """Re-partitions the grid so that every slab carries about the same measured cost (call
        it with this rank's fusion time of a representative pass). The grid state is discarded,
        like Init(). Returns the new ranges."""
        if self.world == 1:
            return self.ranges
        torch = self.torch
        mine = torch.tensor([float(my_cost)], dtype=torch.float64, device=self.slab.device)
        allc = [torch.empty_like(mine) for _ in range(self.world)]
        torch.distributed.all_gather(allc, mine, group=self.group)
        cost = [float(t.item()) for t in allc]
        new = rebalanced_ranges(self.ranges, cost, quantum=FUSION_TILE_PLANES)
        if new != self.ranges:
            self.ranges = new
            self.car.synchronize()
            self.slab.car.close()
            self.slab = None
            self._build()
        return self.ranges

    # ---- carving: no exchange -----------------------------------------------------------------
    def reset(self):
        self.car.reset()

    def make_views(self, cameras, width, height):
        return [make_view(c, width, height) for c in cameras]

    def carve_device_masks(self, views, dev_masks):
        """dev_masks: torch uint8 [V, H, W] on this rank's GPU."""
        stride = dev_masks.stride(0)
        base = dev_masks.data_ptr()
        self.car.carve_device(views, [base + i * stride for i in range(dev_masks.shape[0])],
                              masks=True)

    def carve_host_masks(self, cameras, host_masks, chunk=64):
        """Masks in HOST memory (pinned for asynchronous copies). One GPU: straight through the host
        entry point. Several GPUs: every rank uploads only its 1/world share over its own PCIe link
        and the shares are all-gathered over NVLink (NCCL), so no rank moves the whole set from the
        host; then each rank fuses every view into its slab.  The upload + all-gather runs on a side
        stream in chunks of `chunk` views (one fusion launch), so the masks of the next launch travel
        while the current one is being fused."""
        if self.world == 1:
            assert self.car.carve_batch(cameras, host_masks)
            return
        torch = self.torch
        n = len(host_masks)
        h, w = host_masks[0].shape
        plan, total = [], 0  # (first view, views, views per rank, offset in the device buffer)
        for first in range(0, n, chunk):
            cnt = min(chunk, n - first)
            per = (cnt + self.world - 1) // self.world
            plan.append((first, cnt, per, total))
            total += per * self.world
        buf = getattr(self, "_mask_buf", None)
        if buf is None or tuple(buf.shape) != (total, h, w):
            buf = self._mask_buf = torch.empty((total, h, w), dtype=torch.uint8,
                                               device=self.slab.device)
        side = getattr(self, "_side_stream", None)
        if side is None:
            side = self._side_stream = torch.cuda.Stream(device=self.slab.device)
        side.wait_stream(self.stream)  # kernels of the previous pass may still read the buffer
        events = []
        with torch.cuda.stream(side):
            for first, cnt, per, off in plan:
                lo, hi = min(cnt, self.rank * per), min(cnt, (self.rank + 1) * per)
                for i in range(lo, hi):
                    buf[off + i].copy_(torch.from_numpy(host_masks[first + i]), non_blocking=True)
                mine = buf[off + self.rank * per:off + (self.rank + 1) * per]
                torch.distributed.all_gather_into_tensor(
                    buf[off:off + per * self.world].view(self.world, -1), mine.view(-1),
                    group=self.group)
                ev = torch.cuda.Event()
                ev.record(side)
                events.append(ev)
        views = self.make_views(cameras, w, h)
        for (first, cnt, per, off), ev in zip(plan, events):
            self.stream.wait_event(ev)
            self.carve_device_masks(views[first:first + cnt], buf[off:off + cnt])

    # ---- marching cubes ---------------------------------------------------------------------------
    def _sendrecv(self, send, recv):
        """send -> rank + 1, recv <- rank - 1 (either may be None), ordered on the carver's stream."""
        dist = self.torch.distributed
        ops = []
        if send is not None and self.rank + 1 < self.world:
            ops.append(dist.P2POp(dist.isend, send, self.rank + 1, self.group))
        if recv is not None and self.rank > 0:
            ops.append(dist.P2POp(dist.irecv, recv, self.rank - 1, self.group))
        if ops:
            for w in dist.batch_isend_irecv(ops):
                w.wait()

    def extract_iso_counts(self, iso=0.0, interp=True):
        """Runs the extraction, leaving this rank's part of the mesh on its GPU; returns the GLOBAL
        (vertices, faces) counts."""
        torch = self.torch
        s = self.slab
        if self.world == 1:
            nv, nf = s.begin(iso, interp)
            s.vertices(0)
            s.faces()
            self._counts = (nv, nf)
            self._offsets = (0, 0)
            return nv, nf
        with torch.cuda.stream(self.stream):
            self.car.flush()
            # the slab above keeps min(2, z_begin) halo planes; every slab has >= 2 planes
            self._sendrecv(s.top_planes(2).contiguous(), s.halo)
            nv, nf = s.begin(iso, interp)
            off, total = gather_offsets(nv, nf, self.world, self.group, s.device)
            s.vertices(off[self.rank][0])
            self._sendrecv(s.export_top(), s.ids_in)
            if self.rank > 0:
                s.import_bottom(s.ids_in)
            s.faces()
        self._counts = (nv, nf)
        self._offsets = tuple(int(x) for x in off[self.rank])
        return int(total[0]), int(total[1])

    def extract_iso_host(self, iso=0.0, interp=True):
        """Extraction + copy of THIS rank's part of the mesh to the host (vertex ids are global)."""
        self.extract_iso_counts(iso, interp)
        return self.slab.mesh_host(*self._counts)

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

    def stage_times(self):
        return self.car.stage_times()
