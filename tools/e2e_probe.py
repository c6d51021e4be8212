"""e2e probe: sspp_gpu_connect_nodes_host on S4a with pinned host buffers, wall clock per call."""
import os, sys, time
import numpy as np
sys.path.insert(0, ".")
import torch
from sspp_b200 import scenes
from sspp_b200.engine import Engine
km, dm, occ, known = scenes.synthetic_map(n_vox=512, res=0.3, seed=42)
nodes = scenes.lattice_nodes(100, 1.0)
n = len(nodes)
eng = Engine(0)
eng.map_create(0.3, km, dm, True)
eng.map_upload_dense(occ, known)
pin = torch.from_numpy(nodes).pin_memory().numpy()
rp = torch.empty(n + 1, dtype=torch.int64).pin_memory().numpy()
col = torch.empty(18_000_000, dtype=torch.int32).pin_memory().numpy()
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
for _ in range(3):
    eng.connect_nodes_host(pin, rp, col, radius=1.5, box=(0.5, 0.5, 0.3))
t = []
for _ in range(10):
    flush.zero_(); torch.cuda.synchronize()
    t0 = time.perf_counter()
    st = eng.connect_nodes_host(pin, rp, col, radius=1.5, box=(0.5, 0.5, 0.3))
    t.append(time.perf_counter() - t0)
print("chunks=%s shape=%s e2e %.3f ms (min %.3f) edges %d phases %s" % (os.environ.get("SSPP_HOST_CHUNKS", "4"), os.environ.get("SSPP_HOST_CHUNK_SHAPE", "falling"),
      np.mean(t) * 1e3, np.min(t) * 1e3, st.edges, np.round(eng.connect_timings()[:6], 3)))
