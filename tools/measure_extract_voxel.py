"""Dev tool: times K2 (ExtractVoxel) on the synthetic configs; prints one JSON line per config."""
import json
import sys
import time

import torch

sys.path.insert(0, ".")
from vacancy_b200 import sharded, synthetic as S  # noqa: E402

for wl in sys.argv[1:] or ["cfg2", "cfg4"]:
    cfg = S.CONFIGS[wl]
    car = sharded.ShardedCarver(S.grid_option(cfg["grid"], weighted=cfg["weighted"]), 0, 1, 0)
    cams = S.circle_cameras(cfg["views"], cfg["image"])
    m = torch.from_numpy(S.silhouette(cams[0])).cuda()
    masks = m.unsqueeze(0).expand(cfg["views"], -1, -1).contiguous()
    car.carve_device_masks(car.make_views(cams, cfg["image"], cfg["image"]), masks)
    car.car.synchronize()
    out = {"workload": wl}
    for inside_empty in (False, True):
        car.car.ExtractVoxel(inside_empty, copy=False)  # warm-up
        car.car.stage_times()
        t0 = time.perf_counter()
        n = car.car.ExtractVoxel(inside_empty, copy=False)
        torch.cuda.synchronize()
        wall = (time.perf_counter() - t0) * 1e3
        ms = car.car.stage_times()["extract_voxel"][0]
        nvox = cfg["grid"] ** 3
        out["inside_empty" if inside_empty else "occupied"] = {
            "cubes": n, "count_ms": ms, "wall_ms": wall, "gvoxel_per_s": nvox / ms / 1e6,
            # a carved grid is classified from sdf alone (4 B per voxel; update_num is implied)
            "bytes_per_voxel": 4, "algorithmic_gbs": 4.0 * nvox / ms / 1e6}
    if wl == "cfg2":
        t0 = time.perf_counter()
        v, f = car.car.ExtractVoxel(False)
        out["fill_to_host_ms"] = (time.perf_counter() - t0) * 1e3
        out["fill_bytes"] = int(v.nbytes + f.nbytes)
    print(json.dumps(out))
