"""Dev tool: one full pass of a synthetic config (reset, 2D SDF + fusion of every view, marching
cubes, ExtractVoxel count) for profiling single kernels with ncu, e.g.
  ncu --set full --clock-control none --import-source on -k regex:^mc_vertex_kernel$ -c 1 \
      -o gpurun_out/prof python tools/profile_stages.py cfg4
Prints one JSON line with the stage times of an unprofiled pass (meaningless under ncu)."""
import json
import sys

import torch

sys.path.insert(0, ".")
from vacancy_b200 import sharded, synthetic as S  # noqa: E402

wl = sys.argv[1] if len(sys.argv) > 1 else "cfg4"
passes = int(sys.argv[2]) if len(sys.argv) > 2 else 1
cfg = S.CONFIGS[wl]
car = sharded.ShardedCarver(S.grid_option(cfg["grid"], weighted=cfg["weighted"]), 0, 1, 0)
cams = S.circle_cameras(cfg["views"], cfg["image"])
m = torch.from_numpy(S.silhouette(cams[0])).cuda()
masks = m.unsqueeze(0).expand(cfg["views"], -1, -1).contiguous()
views = car.make_views(cams, cfg["image"], cfg["image"])
out = {"workload": wl}
for _ in range(passes):
    car.reset()
    car.carve_device_masks(views, masks)
    out["mesh"] = car.extract_iso_counts(0.0, True)
    out["occupied"] = car.extract_voxel_count(False)
    out["surface"] = car.extract_voxel_count(True)
    out["stages_ms"] = {k: v[0] for k, v in car.stage_times().items()}
print(json.dumps(out))
