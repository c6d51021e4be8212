#!/usr/bin/env python
"""The bench scene's kernels, once, for `ncu --set full` (run under gpurun, one GPU):

    python tools/profile_target.py && \\
    ncu --set full --import-source on --clock-control none --profile-from-start off \\
        --metrics lts__t_bytes.sum,lts__t_sectors_op_read.sum,lts__t_sectors_op_write.sum,l1tex__t_bytes.sum \\
        -o gpurun_out/r02_prof python tools/profile_target.py

Same scene and calls as bench.py (S4a connect pass; the 65 536 x 4e6 visibility pass), warmed up, with
the profiler switched on (cudaProfilerStart) around ONE connect pass and ONE visibility pass only, so
the capture holds one warm launch of every kernel of the step.  tools/ncu_summary.py turns the
report into the per-kernel text files under profiles/.
"""
import os
import sys
import time

import numpy as np

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)
import bench  # noqa: E402  (scene constants and helpers)
import torch  # noqa: E402
from sspp_b200 import scenes  # noqa: E402
from sspp_b200.engine import Engine  # noqa: E402


def main():
    what = sys.argv[1] if len(sys.argv) > 1 else "all"
    key_min, dims, occ, known = scenes.synthetic_map(n_vox=bench.N_VOX, res=bench.RES, seed=42)
    nodes = scenes.lattice_nodes(bench.LATTICE, 1.0)
    eng = Engine(0)
    eng.map_create(bench.RES, key_min, dims, True)
    eng.map_upload_dense(occ, known)
    eng.set_nodes(nodes)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    rt = torch.cuda.cudart()
    for _ in range(3):
        st = eng.connect(radius=bench.RADIUS, box=bench.BOX)
    if what in ("all", "vis"):
        pts, poses = bench.visibility_scene(occ, known, key_min, dims)
        eng.set_model_points(pts)
        fr = (bench.VIS["hfov"], bench.VIS["vfov"], bench.VIS["near"], bench.VIS["far"])
        for _ in range(3):
            vs = eng.visibility(poses, *fr, fetch=False)
    flush.zero_()
    torch.cuda.synchronize()
    rt.cudaProfilerStart()
    if what in ("all", "connect"):
        st = eng.connect(radius=bench.RADIUS, box=bench.BOX)
    if what in ("all", "vis"):
        vs = eng.visibility(poses, *fr, fetch=False)
    eng.synchronize()
    rt.cudaProfilerStop()
    t = eng.connect_timings()
    print("connect: tested %d edges %d, phases ms total %.3f binning %.3f count %.3f edge %.3f csr %.3f k_edges %.3f"
          % (st.tested_pairs, st.edges, t[0], t[1], t[2], t[3], t[4], t[5]))
    if what in ("all", "vis"):
        v = eng.visibility_timings()
        print("visibility: rays %d visible %d, phases ms total %.3f binning %.3f gather %.3f rays %.3f compact+sort %.3f"
              % (vs.rays, vs.visible_pairs, v[0], v[1], v[2], v[3], v[4]))
    eng.close()


if __name__ == "__main__":
    main()
