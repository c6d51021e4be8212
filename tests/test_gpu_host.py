"""The drop-in flow end to end on the GPU: host C++ mirror (PathPlanner / DistanceHeuristic / GpuWorld)
-> C ABI -> CUDA, checked against the vectors the reference's compiled sources produced."""
import os

import numpy as np
import pytest

import host_lib

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def test_distance_heuristic_test_scenario(eng):
    """src/distance_heuristic_test.cpp:40-117 through the mirror: grid -> connectNodes (GPU) -> A*"""
    g = np.load(os.path.join(GOLD, "dht_known_answer.npz"))
    p = host_lib.HostPlanner(1.5, device=0)
    p.set_goal([4, 4, 2, 0, 0, 0, 0], 1.0)
    p.generate_grid([0, 0, 0], [10, 10, 2], 1.0, True, 90.0)
    edges, tested = p.connect()
    assert (edges, tested) == (80580, 76224)
    rp, col = p.csr()
    assert np.array_equal(rp, g["row_ptr"]) and np.array_equal(col, g["col_idx"])
    path, created, expanded, _ = p.astar([0, 0, 0, 0, 0, 0, 0])
    assert np.array_equal(path, g["path"]) and (created, expanded) == (9480, 153)
    p.close()


def test_reactive_planner_scenario(eng):
    """src/reactive_planner_server.cpp:40-42,58-131: setFree box, occupied sphere, grid, connect, A*;
    a second request re-uses the planner, so nodes are kept and edges double up, as in the reference"""
    g = np.load(os.path.join(GOLD, "reactive_sphere.npz"))
    p = host_lib.HostPlanner(1.5, device=0)
    p.map_create(0.3, [32768 - 64] * 3, [128] * 3, False)
    p.map_set_box([0, 0, 0], [15, 15, 10], False)
    k = g["sphere_keys"].astype(np.float64)
    p.map_insert_boxes((k - 32768 + 0.5) * 0.3, np.full(len(k), 0.3), np.ones(len(k), dtype=np.uint8))
    p.set_goal([2, 2, 0, 0, 0, 0, 0], 0.5)
    p.generate_grid([-2.12, 1.2, 0.2], [10, 10, 0], 1.0, False, 90.0)
    edges, tested = p.connect()
    assert (edges, tested) == (777, 840)
    rp, col = p.csr()
    assert np.array_equal(rp, g["row_ptr"]) and np.array_equal(col, g["col_idx"])
    path, created, expanded, _ = p.astar([-2.12, 1.2, 0.2, 0, 0, 0, 0])
    assert np.array_equal(path, g["path"]) and (created, expanded) == (int(g["created"]), int(g["expanded"]))
    # second request on the same planner (reference: nodes accumulate, children duplicate)
    p.generate_grid([-2.12, 1.2, 0.2], [10, 10, 0], 1.0, False, 90.0)
    assert len(p.nodes()[0]) == 121
    p.connect()
    rp2, col2 = p.csr()
    assert np.array_equal(rp2, 2 * rp)
    p.close()


def test_validity_filtering_in_grid_generation(eng, orc):
    """samplesFiltering with the validity extension: the batched filter keeps exactly the samples the
    oracle's getCellStatusBoundingBox calls free, in generation order"""
    from sspp_b200 import scenes
    res = 0.3
    km, dm, occ, known = scenes.synthetic_map(n_vox=64, res=res, seed=3, n_boxes=8, n_spheres=6, region=14.0)
    p = host_lib.HostPlanner(1.5, device=0)
    p.map_create(res, km, dm, True)
    p.map_upload_dense(occ, known)
    p.set_filter_by_validity(True)
    p.generate_grid([0, 0, 0], [12, 12, 2], 1.0, True, 180.0, filtering=True)
    kept, _ = p.nodes()
    q = host_lib.HostPlanner(1.5)
    q.generate_grid([0, 0, 0], [12, 12, 2], 1.0, True, 180.0)
    allp, _ = q.nodes()
    w = orc.world(res, True)
    w.build_dense(km, dm, occ, known)
    free = w.cell_status_bbox(allp[:, :3], (0.5, 0.5, 0.3)) == 0
    assert 0 < free.sum() < len(allp)
    assert np.array_equal(kept, allp[free])
    p.close()
    q.close()
    w.close()
