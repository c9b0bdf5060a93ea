#!/usr/bin/env python
"""Benchmark of the carving hot path (BASELINE.json metric: Gvoxel-view updates/s for carving plus
marching-cubes ms at 1024^3).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload cfg4]

One "step" = the whole hot path on one batch of synthetic input: reset the grid (Init), build the
2D SDF of every view, fuse all views, extract the iso surface.  Prints ONE JSON line (rank 0).
  value      Nvox * V / step time, inputs (masks) already resident in HBM, CUDA-event timed
  e2e        same step through the public host API (VoxelCarver.Carve on pinned HOST masks, mesh
             copied back to the host), host<->device copies inside the timed region
  roofline   fusion kernel: algorithmic bytes per launch / measured launch time vs measured HBM peak
  cpu_baseline   the unmodified reference (oracle/_ref) timed on this box's host cores on a bounded
             sample of the same workload (N=1 only)
`--impl reference` times only the reference on the host cores (rank 0; other ranks exit).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

from vacancy_b200 import synthetic as S  # noqa: E402


def load_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            p = json.load(f)
        return float(p["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)", p
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)", {}


class ClockSampler:
    """nvidia-smi clocks / throttle reasons while the timed region runs."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.proc, self.lines = index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                 "--format=csv,noheader,nounits", "-lms", "100"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        self.t.join(timeout=2)
        sm, mx, pw, reasons = [], [], [], set()
        names = ("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap")
        for ln in self.lines:
            p = [x.strip() for x in ln.split(",")]
            if len(p) < 7:
                continue
            try:
                sm.append(float(p[0]))
                mx.append(float(p[1]))
                pw.append(float(p[2]))
            except ValueError:
                continue
            for n, v in zip(names, p[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        return {"sm_mhz": float(np.median(sm)), "sm_max_mhz": float(max(mx)),
                "power_w_max": float(max(pw)), "samples": len(sm), "reasons": sorted(reasons)}


def workload(name):
    cfg = dict(S.CONFIGS[name])
    cfg["name"] = name
    return cfg


def describe(cfg):
    return ("%s: %d^3 grid, %d views %dx%d, %s, marching cubes" %
            (cfg["name"], cfg["grid"], cfg["views"], cfg["image"], cfg["image"],
             "weighted-average TSDF (truncation 0.1, max_update_num 255)" if cfg["weighted"]
             else "kMax update, no truncation"))


# ---------------------------------------------------------------------------------------------------
def reference_sample(cfg, n_views, planes):
    """Times the UNMODIFIED reference (oracle/_ref) on a bounded sample of the workload: a slab of
    `planes` z planes of the grid (same voxel size, same bbox in x and y) and `n_views` of its
    views, each stage timed separately (MakeSignedDistanceField, Carve(camera, sdf),
    ExtractIsoSurface) like SURVEY.md section 8(d); extrapolated per unit to the whole workload."""
    from oracle import bindings as B
    try:
        ref, kind = B.RefLib(), "reference"
    except (FileNotFoundError, OSError):
        # the compiled reference did not travel: time the plain-C restatement instead
        B.build(ref=False, port=True)
        ref, kind = B.PortLib(), "port"
    n, V, size = cfg["grid"], cfg["views"], cfg["image"]
    planes = min(planes, n)
    half = planes / n
    res = float(np.float32(2.0) / np.float32(n))
    opt = B.make_options((-1.0, -1.0, -half), (1.0, 1.0, half), res, update=int(cfg["weighted"]),
                         truncation=cfg["weighted"], band=0.1)
    t0 = time.perf_counter()
    car = ref.carver(opt)
    t_init = time.perf_counter() - t0
    assert car.dims == (n, n, planes), car.dims
    cams = S.circle_cameras(V, size)
    pick = [int(round(i * V / n_views)) % V for i in range(n_views)]
    # torch.distributed.run hands its workers OMP_NUM_THREADS=1: ask for every core explicitly
    threads = ref.set_threads(-1)
    t_sdf = t_fuse = 0.0
    for vi in pick:
        mask = S.silhouette(cams[vi])
        oc = B.Camera(size, size, cams[vi].c2w(), *cams[vi]._intrinsics())
        if kind == "reference":
            t_sdf += ref.time_make_sdf(mask, True, cfg["weighted"], 0.1, 1)
            sdf = ref.make_sdf(mask, None, True, cfg["weighted"], 0.1)
            t_fuse += car.carve_sdf(oc, sdf)
        else:
            t1 = time.perf_counter()
            sdf = ref.make_sdf(mask, None, True, cfg["weighted"], 0.1)
            t2 = time.perf_counter()
            car.carve_sdf(oc, sdf)
            t_sdf += t2 - t1
            t_fuse += time.perf_counter() - t2
    tm = []
    t1 = time.perf_counter()
    v, f = car.extract_iso(0.0, True, timing=tm)
    t_mc = tm[0] if tm else time.perf_counter() - t1
    car.close()
    scale = n / planes
    t_full = V * (t_sdf / n_views + t_fuse / n_views * scale) + t_mc * scale
    nvox = float(n) ** 3
    return dict(
        value=nvox * V / t_full / 1e9, t_sample=t_sdf + t_fuse + t_mc, cores=threads, kind=kind,
        sample=("%d of %d views on a %dx%dx%d z-slab of the %d^3 grid; per-view SDF %.3f s (1 thread), "
                "fusion %.3f s/view on the slab (%d OpenMP threads), marching cubes %.3f s on the slab "
                "(1 thread); extrapolated per unit to the full workload (%.0f s); grid Init %.1f s not "
                "counted" % (n_views, V, n, n, planes, n, t_sdf / n_views, t_fuse / n_views,
                             threads, t_mc, t_full, t_init)),
        stages=dict(sdf2d_s_per_view=t_sdf / n_views, fusion_s_per_view_slab=t_fuse / n_views,
                    mc_s_slab=t_mc, fusion_gvv_per_s=nvox / scale * n_views / t_fuse / 1e9,
                    mc_gvoxel_per_s=nvox / scale / t_mc / 1e9, mesh=[len(v), len(f)]))


def run_reference_arm(args, cfg):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    total = args.steps + args.warmup
    n_views = 8 if total <= 12 else max(4, int(96 / total))  # 8 sample views
    planes = 32
    vals, times = [], []
    last = None
    for i in range(total):
        r = reference_sample(cfg, n_views, planes)
        if i >= args.warmup:
            vals.append(r["value"])
            times.append(r["t_sample"])
            last = r
    value = float(np.mean(vals))
    line = {
        "impl": "reference", "metric": "carve_gvoxel_view_updates_per_s", "value": value,
        "unit": "Gvoxel-view/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": float(np.mean(times)) * 1e3, "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": describe(cfg), "host": "CPU only (reference has no GPU path)"},
        "cpu_baseline": {"value": value, "unit": "Gvoxel-view/s", "cores": last["cores"],
                         "kind": last["kind"], "sample": last["sample"], "stages": last["stages"]},
        "e2e": {"value": value, "unit": "Gvoxel-view/s", "h2d_bytes_per_step": 0,
                "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))


# ---------------------------------------------------------------------------------------------------
def run_ours(args, cfg):
    import torch
    import torch.distributed as dist

    from vacancy_b200 import sharded

    procs = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device - this library has no CPU fallback")
    torch.cuda.set_device(local)
    if procs > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    # `world` = number of z slabs = GPUs.  Launched by torch.distributed.run: one process per GPU
    # (vcb_sharded_create_rank).  Launched plainly with --gpus N > 1: ONE process drives the N GPUs
    # (vcb_sharded_create - what a C++ host gets through VACANCY_B200_DEVICES).
    single_process = procs == 1 and args.gpus > 1
    world = args.gpus if single_process else procs
    if single_process and torch.cuda.device_count() < world:
        raise SystemExit("bench.py: --gpus %d but %d CUDA devices" % (world, torch.cuda.device_count()))
    n, V, size = cfg["grid"], cfg["views"], cfg["image"]
    hbm_peak, peak_src, _ = load_peaks()

    if single_process:
        car = sharded.LocalSlabs(S.grid_option(n, weighted=cfg["weighted"]), world,
                                 devices=list(range(world)))
    else:
        car = sharded.ShardedCarver(S.grid_option(n, weighted=cfg["weighted"]), rank, world, local)
    # the engine orders a slab's reset / fusion / extraction on its own main stream: time on it
    stream = torch.cuda.ExternalStream(car.main_stream(0), device=torch.device("cuda", local))
    cams = S.circle_cameras(V, size)
    # the sphere at the origin looks the same from every camera on the circle: rasterise once
    mask0 = S.silhouette(cams[0])
    host_masks = torch.empty((V, size, size), dtype=torch.uint8, pin_memory=True)
    host_masks[:] = torch.from_numpy(mask0)
    dev_masks = host_masks.to("cuda", non_blocking=False)
    views = car.make_views(cams, size, size)
    torch.cuda.synchronize()

    def barrier():
        if procs > 1:
            dist.barrier()
        for d in (range(world) if single_process else [local]):
            torch.cuda.synchronize(d)

    packed_dev = car.pack_masks([dev_masks.data_ptr() + i * dev_masks.stride(0) for i in range(V)])
    packed_views = car.pack_views(views)

    def step_resident():
        car.reset()
        car.carve_views(packed_views, packed_dev, device=True)
        return car.extract_iso_counts(0.0, True)

    host_np = [host_masks[i].numpy() for i in range(V)]
    # what a host hands over per pass: the view structs and the pointers to its pinned masks
    packed_masks = car.pack_masks(host_np)
    e2e_mesh = {}

    def step_e2e():
        # the call a host makes: pinned HOST masks in, the WHOLE mesh on one host out (rank 0's;
        # the other ranks send their part over NVLink inside vcb_sharded_extract_iso_fill)
        car.reset()
        car.carve_views(packed_views, packed_masks)
        total = car.extract_iso_counts(0.0, True)
        mesh = car.mesh_host(receive=rank == 0)
        if mesh is not None:
            e2e_mesh["v"], e2e_mesh["f"] = mesh
        return total[0], total[1], (total[0] + total[1]) * 12 if rank == 0 else 0

    def timed(fn, steps):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        out = None
        for _ in range(steps):
            out = fn()
        e1.record(stream)
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], device="cuda")
        if procs > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item()), out

    if args.auto_balance and world > 1:
        car.set_auto_balance(True)
        args.no_rebalance = True
    for _ in range(args.warmup):
        step_resident()
    if world > 1 and not args.no_rebalance:
        # the slabs do not cost the same (edge slabs see the cameras differently): re-partition once
        # from the measured fusion time of the warm-up passes, then warm up again
        for _ in range(2):  # second round refines the first estimate
            if single_process:
                ranges = car.rebalance()
            else:
                ranges = car.rebalance(car.stage_times()["fusion"][0])
            for _ in range(args.warmup):
                step_resident()
    else:
        ranges = car.ranges
    for i in range(car.n_local):
        car.stage_times(i)  # clear the per-stage counters
    car.car.fusion_class_counts()  # ... and the tile-class counters
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    ms_total, counts = timed(step_resident, args.steps)
    clocks = sampler.stop() if rank == 0 else None
    stages = car.stage_times()
    other_stages = [car.stage_times(i) for i in range(1, car.n_local)]
    classes = car.car.fusion_class_counts()  # (tile, view) pairs of this rank's launches, by class
    # one more pass, untimed, with the kernel counting the view iterations its warps execute (the
    # counting costs ~1 % of the kernel's time, so it stays out of the timed region)
    car.car.set_statistics(True)
    step_resident()
    iters = car.car.fusion_class_counts()["warp_iterations"]
    car.car.set_statistics(False)
    for i in range(car.n_local):
        car.stage_times(i)
    ms_step = ms_total / args.steps
    per_rank = None
    if single_process:
        alls = [stages] + other_stages
        per_rank = {k + "_ms": [round(st[name][0] / args.steps, 2) for st in alls]
                    for k, name in (("fusion", "fusion"), ("sdf2d", "sdf2d"),
                                    ("marching_cubes", "marching_cubes"))}
    elif world > 1:  # how evenly the slabs split the work (rank order = slab order, bottom to top)
        mine = torch.tensor([stages["fusion"][0] / args.steps, stages["sdf2d"][0] / args.steps,
                             stages["marching_cubes"][0] / args.steps], device="cuda")
        allr = [torch.empty_like(mine) for _ in range(world)]
        dist.all_gather(allr, mine)
        per_rank = {"fusion_ms": [round(float(t[0]), 2) for t in allr],
                    "sdf2d_ms": [round(float(t[1]), 2) for t in allr],
                    "marching_cubes_ms": [round(float(t[2]), 2) for t in allr]}

    for _ in range(min(args.warmup, 2)):
        step_e2e()
    e2e_steps = max(1, min(args.steps, 3))
    ms_e2e_total, e2e_out = timed(step_e2e, e2e_steps)
    ms_e2e = ms_e2e_total / e2e_steps
    # the host path must produce the mesh of the resident path (same global counts on every rank)
    assert (int(e2e_out[0]), int(e2e_out[1])) == (int(counts[0]), int(counts[1])), (e2e_out, counts)
    # ... and, by CONTENT, the mesh a single GPU produces: sha256 of the vertex and face buffers
    # gathered on rank 0 against the value committed from a 1-GPU run (tests/golden/bench_mesh_sha.json)
    mesh_sha = mesh_sha_ok = None
    if rank == 0:
        import hashlib
        hsh = hashlib.sha256()
        hsh.update(np.ascontiguousarray(e2e_mesh["v"]).tobytes())
        hsh.update(np.ascontiguousarray(e2e_mesh["f"]).tobytes())
        mesh_sha = hsh.hexdigest()
        gpath = os.path.join(ROOT, "tests", "golden", "bench_mesh_sha.json")
        if os.path.exists(gpath):
            with open(gpath) as f:
                want = json.load(f).get(cfg["name"])
            if want:
                mesh_sha_ok = want["sha256"] == mesh_sha
                assert mesh_sha_ok, "mesh differs from the 1-GPU golden: %s != %s" % (mesh_sha, want["sha256"])

    nvox = float(n) ** 3
    updates = nvox * V
    # fusion roofline: algorithmic bytes of one launch = 16 B per owned voxel + 4 B per SDF pixel of
    # the views in that launch; duration from CUDA events around each launch
    # the library fuses at most 64 queued views per fuse_kernel launch (kMaxViewsPerLaunch); the two
    # small per-launch table / tile-class kernels that precede it are inside the same time span
    fus_ms = stages["fusion"][0]
    fus_launches = -(-V // 64) * args.steps
    views_per_launch = V / (fus_launches / args.steps)
    bytes_per_launch = 16.0 * car.slab_voxels + 4.0 * size * size * views_per_launch
    fus_ms_per_launch = fus_ms / fus_launches
    achieved = bytes_per_launch / (fus_ms_per_launch * 1e-3) / 1e9
    launches_per_step = sum(v[1] for v in stages.values()) / args.steps
    traffic = None
    issue = None
    tpath = os.path.join(ROOT, "profiles", "fuse_kernel_traffic.json")
    if world == 1 and os.path.exists(tpath):
        with open(tpath) as f:
            t = json.load(f).get(cfg["name"])
        if t:  # ncu capture of a full 64-view launch; scale to this run's average views per launch
            traffic = (t["dram_bytes_read"] + t["dram_bytes_write"]) * views_per_launch / \
                t["views_in_launch"]
            issue = {k: t[k] for k in ("thread_instructions_per_voxel_view", "issue_active",
                                       "fma_pipe_active", "source") if k in t}

    # ---- what binds fuse_kernel (batched fusion is bound by FP32 issue, not HBM) ----
    # The classification kernel counts the (tile, view) pairs per class; a pair is 2048 voxel-views.
    # Work is counted over EXECUTED pairs only: the sampling classes at the reference's operation count
    # per voxel-view (54 weighted / 48 kMax: w2c 18, z test 1, project 6, ROI test 4, bilinear 17,
    # truncation test 1, update 7 / 1), the load-free "all ones" class without the four texel products,
    # the ROI test and the truncation test (44 / 38); skipped pairs count nothing.  Peak = the packed
    # FP32 issue rate measured on a B200 (FFMA2 / FADD2 microbenchmark, 36.95 T lane-ops/s at 1965 MHz,
    # profiles/r01_f32x2_issue_microbench.txt), scaled to the SM clock sampled during this run.
    vpt = classes["voxels_per_tile"]
    ops_s = 54.0 if cfg["weighted"] else 48.0
    ops_1 = ops_s - 10.0
    pairs_s = classes["general"] + classes["inside"]
    pairs_1 = classes["all_ones"]
    pairs_skip = classes["all_kval"] + classes["outside"]
    pairs_all = max(1, pairs_s + pairs_1 + pairs_skip)
    executed_vv = float(pairs_s + pairs_1) * vpt / args.steps  # this rank, per step
    lane_ops = (pairs_s * ops_s + pairs_1 * ops_1) * vpt / args.steps
    sm_mhz = (clocks or {}).get("sm_mhz") or 1965.0
    fp32_peak = 36.95 * sm_mhz / 1965.0
    fp32_achieved = lane_ops / (fus_ms / args.steps * 1e-3) / 1e12
    hbm = {"kernel": "fuse_kernel", "achieved": achieved, "peak": hbm_peak, "unit": "GB/s",
           "frac": achieved / hbm_peak, "traffic": traffic, "peak_source": peak_src,
           "traffic_source": "ncu --set full capture of a 64-view launch, scaled to the average "
                             "views per launch (profiles/fuse_kernel_traffic.json)" if traffic else None,
           "bytes_per_launch": bytes_per_launch, "ms_per_launch": fus_ms_per_launch,
           "note": "algorithmic bytes (16 B per voxel + 4 B per SDF pixel per launch): HBM does not "
                   "bind a 64-view launch"}
    # The kernel does not execute the reference's operation list literally (products hoisted out of
    # the z column, shared reciprocal, ...): what its FP32 pipe really issues per voxel-view is the
    # static SASS count of each class's loop (29 packed FP32 instructions per
    # voxel pair on the all-ones path, 38 on the sampling paths, 3 fewer with the kMax update; one
    # packed instruction = 2 lane-operations, one per voxel of the pair).  That fraction is the one
    # ncu's FMA-pipe utilisation measures, and it is `frac`; the algorithmic one is kept beside it.
    sass_s = 38.0 if cfg["weighted"] else 35.0
    sass_1 = sass_s - 9.0
    # warps leave the view loop once the update cap has retired all their voxels: count the view
    # iterations they really executed (256 voxel-views of instructions each), not the pairs
    pipe_achieved = (iters["sampling"] * sass_s + iters["all_ones"] * sass_1) * 256.0 / \
        (fus_ms / args.steps * 1e-3) / 1e12
    roofline = {"kernel": "fuse_kernel", "bound": "fp32_issue", "achieved": pipe_achieved,
                "peak": fp32_peak, "unit": "T lane-ops/s", "frac": pipe_achieved / fp32_peak,
                "achieved_is": "FP32 lane-operations the kernel issues (static SASS count per path x "
                               "view iterations the warps executed, counted in the kernel) / fusion time",
                "issued_voxel_views_per_step": (iters["sampling"] + iters["all_ones"]) * 256.0,
                "algorithmic": {"achieved": fp32_achieved, "frac": fp32_achieved / fp32_peak,
                                "ops_per_voxel_view": {"sampling": ops_s, "all_ones": ops_1}},
                "traffic": traffic,
                "peak_source": "measured packed-FP32 issue rate of a B200 (profiles/"
                               "r01_f32x2_issue_microbench.txt: 36.95 T lane-ops/s at 1965 MHz) x "
                               "sampled SM clock / 1965",
                "work": "reference operations per voxel-view over EXECUTED voxel-views only (%g "
                        "sampling, %g load-free all-ones class, 0 for skipped tiles)" % (ops_s, ops_1),
                "executed_voxel_views_per_step": executed_vv,
                "executed_share": (pairs_s + pairs_1) / pairs_all,
                "hbm_frac": achieved / hbm_peak, "ms_per_launch": fus_ms_per_launch,
                "rank": rank}
    tile_classes = {k: classes[k] / pairs_all for k in ("general", "inside", "all_ones", "all_kval",
                                                         "outside")}

    line = {
        "metric": "carve_gvoxel_view_updates_per_s", "value": updates / (ms_step * 1e-3) / 1e9,
        "unit": "Gvoxel-view/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms_step, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "config": {"workload": describe(cfg), "grid": [n, n, n], "views": V, "image": [size, size],
                   "parallelism": "z-slab x%d, %s" % (world, "one process driving all GPUs"
                                                       if single_process else "one process per GPU"),
                   "slabs": [list(r) for r in ranges],
                   "l2": "inputs larger than L2 (masks %.2f GB, grid state %.2f GB per step)"
                         % (V * size * size / 1e9, 8 * nvox / 1e9),
                   "step": "reset + 2D SDF of every view + fusion (64 views/launch) + marching cubes"},
        "clocks": clocks,
        "e2e": {"value": updates / (ms_e2e * 1e-3) / 1e9, "unit": "Gvoxel-view/s",
                "ms_per_step": ms_e2e, "h2d_bytes_per_step": int(V * size * size),
                "d2h_bytes_per_step": int(e2e_out[2]), "steps": e2e_steps,
                "api": "vcb_sharded_carve_masks(host pinned masks; each rank uploads the views it "
                       "owns) + vcb_sharded_extract_iso_count/_fill -> the whole mesh on rank 0's host",
                "mesh_sha256": mesh_sha, "mesh_equals_1gpu_golden": mesh_sha_ok},
        "gpu_launches": int(round(launches_per_step * args.steps)),
        "roofline": roofline,
        "stages_ms_per_step": {k: v[0] / args.steps for k, v in stages.items()},
        "mc": {"ms": stages["marching_cubes"][0] / args.steps,
               "gvoxel_per_s": nvox / world / (stages["marching_cubes"][0] / args.steps * 1e-3) / 1e9,
               "vertices": int(counts[0]), "faces": int(counts[1]),
               # algorithmic bytes: 4 B per voxel + 12 B per vertex + 12 B per face
               "algorithmic_gbs": (4.0 * nvox + 12.0 * (counts[0] + counts[1])) / world /
               (stages["marching_cubes"][0] / args.steps * 1e-3) / 1e9,
               "frac_of_hbm_peak": (4.0 * nvox + 12.0 * (counts[0] + counts[1])) / world /
               (stages["marching_cubes"][0] / args.steps * 1e-3) / 1e9 / hbm_peak},
        "fusion_gvv_per_s": updates / world / (fus_ms / args.steps * 1e-3) / 1e9,
        "per_rank": per_rank,
    }
    line["hbm"] = hbm
    line["tile_classes"] = tile_classes
    line["issue"] = issue
    if world == 1 and not args.no_cpu_baseline:
        try:
            r = reference_sample(cfg, 4, 32)
            line["cpu_baseline"] = {"value": r["value"], "unit": "Gvoxel-view/s", "cores": r["cores"],
                                    "kind": r["kind"], "sample": r["sample"],
                                    "stages": r["stages"]}
        except Exception as e:  # the checker being absent must not hide the GPU numbers
            line["cpu_baseline"] = {"value": None, "unit": "Gvoxel-view/s", "cores": 0,
                                    "kind": "reference", "sample": "unavailable: %r" % (e,)}
    if rank == 0:
        print(json.dumps(line))
    car.close()  # every rank at the same point: the ranks unmap each other's buffers before freeing
    if procs > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="cfg4", choices=sorted(S.CONFIGS))
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-rebalance", action="store_true",
                    help="keep the equal-thickness z slabs (multi-GPU)")
    ap.add_argument("--auto-balance", action="store_true",
                    help="no measured re-partitioning: the engine places the slab boundaries itself "
                         "from the tile classes of the first chunk of views (multi-GPU)")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    cfg = workload(args.workload)
    if args.impl == "reference":
        run_reference_arm(args, cfg)
    else:
        run_ours(args, cfg)


if __name__ == "__main__":
    main()
