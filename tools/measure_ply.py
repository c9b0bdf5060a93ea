"""Dev tool: times the PLY text path (F2) on a synthetic config's iso-surface; one JSON line.
GPU: vcb_ply_body_size + vcb_ply_body_fill (formatting on the device, text copied to pinned host
memory).  CPU: the reference's Mesh::WritePly (oracle/_ref) on a bounded sample of the same mesh,
extrapolated per line."""
import ctypes as C
import json
import os
import sys
import tempfile
import time

import numpy as np
import torch

sys.path.insert(0, ".")
from oracle import bindings as B  # noqa: E402
from vacancy_b200 import sharded, synthetic as S  # noqa: E402

wl = sys.argv[1] if len(sys.argv) > 1 else "cfg4"
cfg = S.CONFIGS[wl]
car = sharded.ShardedCarver(S.grid_option(cfg["grid"], weighted=cfg["weighted"]), 0, 1, 0)
cams = S.circle_cameras(cfg["views"], cfg["image"])
masks = torch.stack([torch.from_numpy(S.silhouette(c)) for c in cams[::max(1, len(cams) // 36)]]).cuda()
views = car.make_views(cams[::max(1, len(cams) // 36)], cfg["image"], cfg["image"])
car.carve_device_masks(views, masks)
nv, nf = car.extract_iso_counts(0.0, True)
lib, h = car.car._lib, car.car._h
vb, fb = C.c_size_t(), C.c_size_t()


def gpu_pass(pin):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    assert lib.vcb_ply_body_size(h, C.byref(vb), C.byref(fb)) == 0
    t1 = time.perf_counter()
    assert lib.vcb_ply_body_fill(h, C.c_void_p(pin.data_ptr()), C.c_void_p(pin.data_ptr() + vb.value)) == 0
    t2 = time.perf_counter()
    return (t1 - t0) * 1e3, (t2 - t1) * 1e3


assert lib.vcb_ply_body_size(h, C.byref(vb), C.byref(fb)) == 0
pin = torch.empty(vb.value + fb.value, dtype=torch.uint8, pin_memory=True)
gpu_pass(pin)
size_ms, fill_ms = gpu_pass(pin)
out = {"workload": wl, "vertices": nv, "faces": nf, "text_bytes": vb.value + fb.value,
       "gpu_size_ms": size_ms, "gpu_fill_ms": fill_ms,
       "gpu_lines_per_s": (nv + nf) / ((size_ms + fill_ms) * 1e-3)}
try:
    ref = B.RefLib()
    v, f = car.mesh_host()
    n = min(nv, 300000)
    with tempfile.TemporaryDirectory() as d:
        t0 = time.perf_counter()
        ref.write_ply(os.path.join(d, "s.ply"), v[:n], f[:n] % n)
        cpu = time.perf_counter() - t0
    text = bytes(pin.numpy()[:64])
    out.update({"cpu_reference_sample_lines": 2 * n, "cpu_reference_sample_s": cpu,
                "cpu_reference_lines_per_s": 2 * n / cpu,
                "cpu_reference_extrapolated_s": cpu / (2 * n) * (nv + nf),
                "first_line": text.split(b"\n")[0].decode()})
except (OSError, FileNotFoundError) as e:
    out["cpu_reference"] = "unavailable: %r" % (e,)
print(json.dumps(out))
