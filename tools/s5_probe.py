"""Config 5 scale probe: 1024^3 voxels @0.15 m, 216^3 lattice @0.7 m (10.08M nodes), R=1.05 -> 100 rays/pair."""
import sys
import time

import numpy as np

sys.path.insert(0, ".")
from sspp_b200 import scenes
from sspp_b200.engine import Engine

t0 = time.time()
km, dm, occ, known = scenes.synthetic_map(n_vox=1024, res=0.15, seed=42)
print("map gen %.1f s, fill %.3f" % (time.time() - t0, np.unpackbits(occ.view(np.uint8)).mean()))
nodes = scenes.lattice_nodes(216, 0.7)
print("nodes", len(nodes))
eng = Engine(0)
eng.map_create(0.15, km, dm, True)
eng.map_upload_dense(occ, known)
eng.set_nodes(nodes)
for rep in range(3):
    t0 = time.perf_counter()
    st = eng.connect(radius=1.05, box=(0.5, 0.5, 0.3))
    dt = time.perf_counter() - t0
    print("connect %.2f ms: tested %d edges %d rays/pair %d pyramid_rows %d box_pairs %d lattice_pairs %d -> %.2f Gedges/s  phases %s"
          % (dt * 1e3, st.tested_pairs, st.edges, st.rays_per_pair, st.pyramid_rows, st.box_pairs, st.lattice_pairs,
             st.tested_pairs / dt / 1e9, np.round(eng.connect_timings()[:6], 2)))
sa = eng.connect(radius=1.05, box=(0.5, 0.5, 0.3), use_pyramid=False, rows=(5_000_000, 5_200_000))
ra, ca = eng.connect_fetch(sa)
sb = eng.connect(radius=1.05, box=(0.5, 0.5, 0.3), use_pyramid=True, rows=(5_000_000, 5_200_000))
rb, cb = eng.connect_fetch(sb)
print("shortcuts leave the CSR of rows 5.0M-5.2M unchanged:", bool(np.array_equal(ra, rb) and np.array_equal(ca, cb)), sa.edges)
