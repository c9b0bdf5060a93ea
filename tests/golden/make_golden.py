"""Generates the committed golden fixtures FROM THE REFERENCE ITSELF (dev container only).

Needs /root/reference (read-only) and oracle/_ref/libvacancy_ref.so (`make -C oracle ref`), i.e.
the unmodified reference compiled by oracle/Makefile.  Writes, next to this file:
  bunny_inputs.npz   the bundled config-1 inputs: 6 masks (data/mask_0000{0..5}.png decoded by the
                     reference's stb loader), the TUM poses as c2w (examples.cc:40-52) and the
                     float32 w2c the reference's Camera derives from them (camera.cc:39-42)
  bunny_golden.json  per-view statistics of the reference run of examples.cc's scenario
                     (SURVEY.md section 4 table), kMax and TSDF modes, several iso levels
  bunny_final.npz    final grid (sdf, update_num), final iso-0 mesh, per-view SDF of view 0
  synth_golden.json  mesh / occupancy counts for small synthetic scenes (vacancy_b200.synthetic)
Run:  python tests/golden/make_golden.py
"""
import hashlib
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

from oracle import bindings as B  # noqa: E402
from vacancy_b200 import synthetic as S  # noqa: E402

REF_DATA = "/root/reference/data"

# examples.cc:87-101,108-115
BB_MIN = [-250.0 - 20.0, -344.586151 - 20.0, -129.982697 - 20.0]
BB_MAX = [250.0 + 20.0, 150.542343 + 20.0, 257.329224 + 20.0]
RESOLUTION = 10.0
INTRINSICS = dict(fx=258.65, fy=258.25, cx=159.3, cy=127.65, width=320, height=240)


def bunny_bbox():
    # the example adds the offset in float32 (examples.cc:91-99)
    mn = np.array([-250.0, -344.586151, -129.982697], np.float32) - np.float32(20.0)
    mx = np.array([250.0, 150.542343, 257.329224], np.float32) + np.float32(20.0)
    return mn, mx


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()[:16]


def grid_stats(sdf, un):
    upd = un >= 1
    return dict(updated=int(upd.sum()), occupied=int(((sdf <= 0) & upd).sum()),
                max_update_num=int(un.max()), sum_sdf=float(sdf[upd].astype(np.float64).sum()),
                sdf_sha=sha(sdf), un_sha=sha(un))


def main():
    ref = B.RefLib()
    masks = np.stack([ref.load_png_gray(os.path.join(REF_DATA, "mask_%05d.png" % i))
                      for i in range(6)])
    poses = []
    with open(os.path.join(REF_DATA, "tumpose.txt")) as f:
        for line in f:
            p = line.split(" ")
            assert len(p) == 8
            vals = [float(x) for x in p[1:]]
            poses.append(ref.tum_pose_to_c2w(vals[:3], vals[3:]))
    c2w = np.stack(poses)
    w2c = np.stack([ref.w2c_from_c2w(m) for m in c2w])
    np.savez_compressed(os.path.join(HERE, "bunny_inputs.npz"), masks=masks, c2w=c2w, w2c=w2c)

    mn, mx = bunny_bbox()
    golden = {"dims": None, "modes": {}}
    final = {}
    for mode, kw in (("max", dict()), ("tsdf", dict(update=1, truncation=True, band=0.1))):
        opt = B.make_options(mn, mx, RESOLUTION, **kw)
        car = ref.carver(opt)
        golden["dims"] = list(car.dims)
        rows = []
        for i in range(6):
            cam = B.Camera(320, 240, c2w[i], INTRINSICS["fx"], INTRINSICS["fy"], INTRINSICS["cx"],
                           INTRINSICS["cy"])
            sdf_img = car.carve_mask(cam, masks[i])
            sdf, un = car.download()
            row = grid_stats(sdf, un)
            row["sdf2d"] = dict(min=float(sdf_img.min()), max=float(sdf_img.max()),
                                sum=float(sdf_img.astype(np.float64).sum()), sha=sha(sdf_img))
            v, f = car.extract_iso(0.0, True)
            row["mc"] = [len(v), len(f)]
            row["mc_sha"] = [sha(v), sha(f)]
            nv, nf = car.extract_voxel(False, copy=False)
            row["voxel_cubes"] = nv // 24
            nv, nf = car.extract_voxel(True, copy=False)
            row["surface_cubes"] = nv // 24
            rows.append(row)
            if i == 0 and mode == "max":
                final["sdf2d_view0"] = sdf_img
        extra = {}
        for iso, interp in ((0.0, False), (0.05, True), (-0.02, True)):
            v, f = car.extract_iso(iso, interp)
            extra["iso%g_interp%d" % (iso, interp)] = [len(v), len(f), sha(v), sha(f)]
        golden["modes"][mode] = dict(views=rows, extra=extra)
        sdf, un = car.download()
        v, f = car.extract_iso(0.0, True)
        final[mode + "_sdf"], final[mode + "_un"] = sdf, un
        final[mode + "_verts"], final[mode + "_faces"] = v, f
        car.close()
    with open(os.path.join(HERE, "bunny_golden.json"), "w") as f:
        json.dump(golden, f, indent=1)
    np.savez_compressed(os.path.join(HERE, "bunny_final.npz"), **final)

    # small synthetic scenes: counts only
    synth = {}
    for name, n, views, size, weighted, shape in (
            ("sphere32_max", 32, 6, 128, False, S.SPHERE),
            ("dumbbell48_tsdf", 48, 10, 160, True, S.DUMBBELL),
            ("dumbbell64_max", 64, 12, 256, False, S.DUMBBELL)):
        o = S.grid_option(n, weighted=weighted)
        opt = B.make_options(o.bb_min, o.bb_max, o.resolution, update=int(weighted),
                             truncation=weighted)
        car = ref.carver(opt)
        for cam in S.circle_cameras(views, size, tilt=0.4):
            rc = B.Camera(size, size, cam.c2w(), *cam._intrinsics())
            car.carve_mask(rc, S.silhouette(cam, shape))
        sdf, un = car.download()
        v, f = car.extract_iso(0.0, True)
        row = grid_stats(sdf, un)
        row["mc"] = [len(v), len(f)]
        row["mc_sha"] = [sha(v), sha(f)]
        row["voxel_cubes"] = car.extract_voxel(False, copy=False)[0] // 24
        row["surface_cubes"] = car.extract_voxel(True, copy=False)[0] // 24
        row["dims"] = list(car.dims)
        synth[name] = row
        car.close()
    with open(os.path.join(HERE, "synth_golden.json"), "w") as f:
        json.dump(synth, f, indent=1)
    print(json.dumps({m: [(r["updated"], r["occupied"], r["mc"], r["surface_cubes"])
                          for r in golden["modes"][m]["views"]] for m in golden["modes"]}))
    print(json.dumps({k: (v["mc"], v["voxel_cubes"]) for k, v in synth.items()}))


if __name__ == "__main__":
    main()
