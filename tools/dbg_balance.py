import sys; sys.path.insert(0,'.')
import numpy as np
from vacancy_b200 import sharded, synthetic as S
o = S.grid_option(256, weighted=True)
cams = S.circle_cameras(70, 512, tilt=0.2)
masks = [S.silhouette(c, ((0.1, 0.0, -0.55, 0.3),)) for c in cams]
eng = sharded.LocalSlabs(o, 4)
eng.set_auto_balance(True)
eng.carve_batch(cams, masks)
print(eng.ranges)
