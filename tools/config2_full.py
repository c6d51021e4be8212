#!/usr/bin/env python
"""BASELINE configs[1] at full size (14 765 way-points / camera poses of the reference's fixtures): times the
engine's connectNodes (without and with the aircraft stand-in map) and its visibility pass, and the reference's
own compiled connectNodes (oracle/_ref, one thread) on the same poses.  Prints one JSON object."""
import json
import os
import sys
import time

import numpy as np

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)
sys.path.insert(0, os.path.join(REPO, "tests"))
from sspp_b200 import scenes  # noqa: E402
from sspp_b200.engine import Engine  # noqa: E402


def main():
    g = np.load(os.path.join(REPO, "tests", "golden", "config2_full.npz"))
    poses = g["poses"]
    nodes = np.ascontiguousarray(poses[:, :3])
    cam = np.ascontiguousarray(g["cam_rows"][g["ids"]])
    res = 0.25
    km, dm, occ, known = scenes.aircraft_model(res)
    pts = scenes.surface_points(occ, known, km, dm, res, 1_000_000, seed=1234)
    eng = Engine(0)
    eng.map_create(res, km, dm, True)
    eng.map_upload_dense(occ, known)
    eng.set_nodes(nodes)
    out = {"config": "BASELINE configs[1] at full size: 14765 poses of SearchSpaceUAV/Cam_1.5m_1to4_NEW, R 2.5, box (0.5,0.5,0.3), "
                     "aircraft stand-in voxelised at 0.25 m (48 rays per pair), 10^6 model points, frustum 58x45 deg 0.7-6 m"}
    for name, kw in (("connect_no_map", dict(use_map=False)), ("connect_with_map", dict(box=(0.5, 0.5, 0.3)))):
        for _ in range(3):
            st = eng.connect(radius=2.5, **kw)
        ms = []
        for _ in range(10):
            st = eng.connect(radius=2.5, **kw)
            ms.append(float(eng.connect_timings()[0]))
        out[name] = {"device_ms": float(np.mean(ms)), "tested_pairs": int(st.tested_pairs), "in_radius_pairs": int(st.in_radius_pairs),
                     "edges": int(st.edges), "edges_per_s": st.tested_pairs / (np.mean(ms) * 1e-3)}
    rp = np.zeros(len(nodes) + 1, dtype=np.int64)
    col = np.zeros(int(st.in_radius_pairs) + 64, dtype=np.int32)
    t = []
    for _ in range(6):
        t0 = time.perf_counter()
        eng.connect_nodes_host(nodes, rp, col, radius=2.5, box=(0.5, 0.5, 0.3))
        t.append(time.perf_counter() - t0)
    out["connect_with_map"]["e2e_ms"] = float(np.mean(t[1:]) * 1e3)
    eng.set_model_points(pts)
    for _ in range(3):
        vs = eng.visibility(cam, 58.0, 45.0, 0.7, 6.0, fetch=False)
    ms = []
    for _ in range(10):
        vs = eng.visibility(cam, 58.0, 45.0, 0.7, 6.0, fetch=False)
        ms.append(float(eng.visibility_timings()[0]))
    out["visibility"] = {"device_ms": float(np.mean(ms)), "rays": int(vs.rays), "visible_pairs": int(vs.visible_pairs),
                         "rays_per_s": vs.rays / (np.mean(ms) * 1e-3)}
    eng.close()
    import oracle_lib
    if os.path.exists(oracle_lib.LIB_REF):
        O = oracle_lib.Oracle(oracle_lib.LIB_REF)
        w = O.world(res, True)
        w.build_dense(km, dm, occ, known)
        for name, world in (("reference_connect_no_map", None), ("reference_connect_with_map", w)):
            p = O.planner(2.5, world)
            p.insert_nodes(poses[::-1])  # head insertion: reversed so that the list order equals `poses`
            t0 = time.perf_counter()
            E = p.connect()
            dt = time.perf_counter() - t0
            out[name] = {"seconds": dt, "connections": int(E), "threads": 1,
                         "what": "the reference's own compiled PathPlanner::connectNodes (src/pathplanner.cpp:527-567), O(N^2) scan"}
            p.close()
        w.close()
    print(json.dumps(out))


if __name__ == "__main__":
    main()
