"""Small end-to-end pass over every entry point, for compute-sanitizer runs (not part of the test-suite)."""
import sys

import numpy as np

sys.path.insert(0, ".")
from sspp_b200 import scenes
from sspp_b200.engine import Engine

res = 0.3
km, dm, occ, known = scenes.synthetic_map(n_vox=64, res=res, seed=3, n_boxes=8, n_spheres=6, region=14.0, unknown_shell=0.15)
eng = Engine(0)
eng.map_create(res, km, dm, True)
eng.map_upload_dense(occ, known)
eng.map_set_box([0, 0, 0], [1.0, 1.0, 1.0], False)
eng.map_upload_boxes([[1.05, 1.05, 1.05]], [0.3], [1])
nodes = np.concatenate([scenes.lattice_nodes(9, 1.0), scenes.uniform_nodes(700, 12.0, 1), scenes.lattice_nodes(3, 1.0)])
eng.set_nodes(nodes)
for kw in (dict(), dict(use_pyramid=False), dict(count_steps=True), dict(rows=(100, 900)), dict(use_map=False), dict(position_mode=1),
           dict(box=(0.9, 0.9, 0.6)), dict(radius=3.2)):
    a = dict(radius=1.6, box=(0.5, 0.5, 0.3))
    a.update(kw)
    st = eng.connect(**a)
    rp, col = eng.connect_fetch(st)
    print(kw, st.edges, st.tested_pairs, st.lattice_pairs)
rng = np.random.RandomState(0)
S = rng.uniform(-9, 9, (5000, 3))
E = S + rng.normal(size=S.shape)
print(np.bincount(eng.check_segments(S, E)), np.bincount(eng.check_segments_noalloc(S, E)), np.bincount(eng.validity(S)))
O = S.astype(np.float32)
D = rng.normal(size=S.shape).astype(np.float32)
print(eng.cast_rays(O, D, True, rng.uniform(1, 8, len(S)))[0].sum())
pts = scenes.surface_points(occ, known, km, dm, res, 20000, seed=4)
eng.set_model_points(pts)
st, vp, idx = eng.visibility(scenes.viewpoints_near(pts, 50, 3.0, 5), 58.0, 45.0, 0.7, 6.0)
print(st.rays, st.visible_pairs)
o2, k2 = eng.map_download_dense()
eng.close()
print("done")
