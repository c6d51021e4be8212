"""Config 3: 100 consecutive local replans on a small octomap, GPU vs the reference's CPU path.
octomap res 0.3, free box 15x15x10 at the origin, occupied sphere r=1 at the origin, grid 10x10x0 @1.0 m
(121 nodes) around a jittered grid_start, R 1.5, box (0.5,0.5,0.3).  Fresh planner per call."""
import os
import sys
import time

import numpy as np

sys.path.insert(0, ".")
sys.path.insert(0, "tests")
import host_lib
import oracle_lib

rng = np.random.RandomState(11)
starts = [(-2.12 + rng.uniform(-0.5, 0.5), 1.2 + rng.uniform(-0.5, 0.5), 0.2) for _ in range(100)]
g = np.load("tests/golden/reactive_sphere.npz")
keys = g["sphere_keys"].astype(np.float64)


def gpu_run():
    p = host_lib.HostPlanner(1.5, device=0)
    p.map_create(0.3, [32768 - 64] * 3, [128] * 3, False)
    p.map_set_box([0, 0, 0], [15, 15, 10], False)
    p.map_insert_boxes((keys - 32768 + 0.5) * 0.3, np.full(len(keys), 0.3), np.ones(len(keys), dtype=np.uint8))
    p.close()
    lat, paths = [], []
    for s in starts:
        t0 = time.perf_counter()
        p = host_lib.HostPlanner(1.5, device=0)          # fresh planner (and context) per request
        p.map_create(0.3, [32768 - 64] * 3, [128] * 3, False)
        p.map_set_box([0, 0, 0], [15, 15, 10], False)
        p.map_insert_boxes((keys - 32768 + 0.5) * 0.3, np.full(len(keys), 0.3), np.ones(len(keys), dtype=np.uint8))
        t1 = time.perf_counter()
        p.set_goal([2, 2, 0, 0, 0, 0, 0], 0.5)
        p.generate_grid(list(s), [10, 10, 0], 1.0, False, 90.0)
        p.connect()
        t2 = time.perf_counter()
        path, created, expanded, _ = p.astar([s[0], s[1], s[2], 0, 0, 0, 0])
        t3 = time.perf_counter()
        p.close()
        lat.append((t1 - t0, t2 - t1, t3 - t2))
        paths.append(path)
    return np.array(lat), paths


def cpu_run():
    have_ref = os.path.exists(oracle_lib.LIB_REF)
    O = oracle_lib.Oracle(oracle_lib.LIB_REF if have_ref else oracle_lib.LIB_ORACLE)
    w = O.world(0.3, False)
    w.set_box([0, 0, 0], [15, 15, 10], False)
    w.set_keys(g["sphere_keys"], True)
    lat, paths = [], []
    for s in starts:
        t1 = time.perf_counter()
        p = O.planner(1.5, w)
        p.set_goal([2, 2, 0, 0, 0, 0, 0], 0.5)
        p.generate_grid(list(s), [10, 10, 0], 1.0, False, 90.0)
        p.connect()
        t2 = time.perf_counter()
        path, created, expanded = p.astar([s[0], s[1], s[2], 0, 0, 0, 0])
        t3 = time.perf_counter()
        p.close()
        lat.append((0.0, t2 - t1, t3 - t2))
        paths.append(path)
    return np.array(lat), paths, have_ref


lg, pg = gpu_run()
lc, pc, have_ref = cpu_run()
same = all(np.array_equal(a, b) for a, b in zip(pg, pc))
print("paths identical to the %s on all 100 replans: %s" % ("reference's compiled sources" if have_ref else "oracle", same))
print("GPU  per replan: map ingest %.3f ms, grid+connect %.3f ms (median), A* %.3f ms" % tuple(np.median(lg, axis=0) * 1e3))
print("CPU  per replan: grid+connect %.3f ms (median), A* %.3f ms" % tuple(np.median(lc, axis=0)[1:] * 1e3))
