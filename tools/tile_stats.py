import ctypes, os, sys, json
sys.path.insert(0, '/root/repo')
os.environ['VACANCY_B200_LIB'] = '/root/repo/build_variants/lib_stats.so'
import numpy as np, torch
from vacancy_b200 import synthetic as S, capi
from vacancy_b200 import sharded
wl = sys.argv[1]
cfg = S.CONFIGS[wl]
car = sharded.ShardedCarver(S.grid_option(cfg['grid'], weighted=cfg['weighted']), 0, 1, 0)
cams = S.circle_cameras(cfg['views'], cfg['image'])
m = torch.from_numpy(S.silhouette(cams[0])).cuda()
masks = m.unsqueeze(0).expand(cfg['views'], -1, -1).contiguous()
views = car.make_views(cams, cfg['image'], cfg['image'])
car.reset(); car.carve_device_masks(views, masks); car.car.synchronize()
out = (ctypes.c_ulonglong * 4)()
capi.load_library().vcb_debug_tile_stats(out)
tot = sum(out)
print(wl, [round(x / tot, 4) for x in out], tot)
