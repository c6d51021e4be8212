"""S4b probe: 10^6 nodes uniform in the 100 m cube of the 512^3 map, R = 1.5."""
import sys
import time

import numpy as np

sys.path.insert(0, ".")
from sspp_b200 import scenes
from sspp_b200.engine import Engine

km, dm, occ, known = scenes.synthetic_map(n_vox=512, res=0.3, seed=42)
nodes = scenes.uniform_nodes(1_000_000, extent=100.0, seed=43)
eng = Engine(0)
eng.map_create(0.3, km, dm, True)
eng.map_upload_dense(occ, known)
eng.set_nodes(nodes)
for rep in range(4):
    t0 = time.perf_counter()
    st = eng.connect(radius=1.5, box=(0.5, 0.5, 0.3))
    dt = time.perf_counter() - t0
    print("connect %.2f ms: tested %d edges %d pyramid_rows %d box_pairs %d lattice_pairs %d -> %.2f Gedges/s  phases %s"
          % (dt * 1e3, st.tested_pairs, st.edges, st.pyramid_rows, st.box_pairs, st.lattice_pairs, st.tested_pairs / dt / 1e9,
             np.round(eng.connect_timings()[:6], 3)))
