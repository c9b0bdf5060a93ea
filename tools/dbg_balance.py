import sys; sys.path.insert(0,'.')
import numpy as np
from vacancy_b200 import sharded, synthetic as S
o = S.grid_option(256, weighted=True)
from vacancy_b200 import PinholeCamera
cams = [PinholeCamera(512, 512, c.c2w(), c.principal_point(), (3.5 * 512, 3.5 * 512))
        for c in S.circle_cameras(70, 512, tilt=0.2)]
masks = [S.silhouette(c, S.DUMBBELL) for c in cams]
eng = sharded.LocalSlabs(o, 4)
eng.set_auto_balance(True)
eng.carve_batch(cams, masks)
print(eng.ranges)
