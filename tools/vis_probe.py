"""Visibility throughput probe (not part of the test-suite): S4-like scene, many viewpoints."""
import sys
import time

import numpy as np

sys.path.insert(0, ".")
from sspp_b200 import scenes
from sspp_b200.engine import Engine

nv = int(sys.argv[1]) if len(sys.argv) > 1 else 65536
npts = int(sys.argv[2]) if len(sys.argv) > 2 else 4_000_000
km, dm, occ, known = scenes.synthetic_map(n_vox=512, res=0.3, seed=42)
pts = scenes.surface_points(occ, known, km, dm, 0.3, npts, seed=45)
poses = scenes.viewpoints_near(pts, nv, standoff=3.0, seed=46)
eng = Engine(0)
eng.map_create(0.3, km, dm, True)
eng.map_upload_dense(occ, known)
eng.set_model_points(pts)
for rep in range(3):
    t0 = time.perf_counter()
    st = eng.visibility(poses, 58.0, 45.0, 0.7, 6.0, fetch=False)
    eng.synchronize()
    dt = time.perf_counter() - t0
    print(f"nv={nv} np={npts} rays={st.rays} visible={st.visible_pairs} t={dt*1e3:.2f} ms  {st.rays/dt/1e9:.3f} Grays/s")
