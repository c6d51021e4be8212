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


def test_coverage_heuristic_plans_a_view_path(eng, orc):
    """CoverageHeuristic (authored; the reference snapshot has no source for it) over the engine's visibility
    pass: the visible set of every way-point on the path equals the oracle's for its sensor pose; the
    coverage along the path is the size of the running union; the search stops within tolerance of the target"""
    from sspp_b200 import scenes
    res = 0.25
    km, dm, occ, known = scenes.synthetic_map(n_vox=96, res=res, seed=4, n_boxes=10, n_spheres=8, region=16.0)
    pts = scenes.surface_points(occ, known, km, dm, res, 30000, seed=45)
    cam = dict(mount_xyz=[0.1, 0.0, -0.05], mount_rpy=[0.0, 0.3, 0.0], hfov=58.0, vfov=45.0, near=0.7, far=6.0)
    p = host_lib.HostPlanner(1.6, device=0, coverage=cam)
    p.map_create(res, km, dm, True)
    p.map_upload_dense(occ, known)
    p.set_model_points(pts)
    p.set_filter_by_validity(True)
    p.generate_grid([0, 0, 0], [14, 14, 4], 1.5, True, 90.0, filtering=True)
    poses, _ = p.nodes()
    assert len(poses) > 100
    edges, tested = p.connect()
    assert 0 < edges
    p.set_coverage_goal(12.0, 0.5)
    path, created, expanded, order = p.astar([-6, -6, 0, 0, 0, 0, 1])
    assert len(path) >= 2
    cov = path[:, 8]
    assert np.all(np.diff(cov) >= 0) and cov[-1] >= 12.0 - 0.5
    rays, calls = p.coverage_counters()
    assert rays > 0 and calls <= expanded + 1   # one batched visibility call per expansion at most
    # visible sets along the path against the oracle, and coverage = |union| / |model|
    w = orc.world(res, True)
    w.build_dense(km, dm, occ, known)
    sens = p.sensor_poses()
    union = set()
    for k in range(len(path)):
        i = int(np.nonzero(np.all(np.abs(poses - path[k, :7]) < 1e-12, axis=1))[0][0])
        got = p.visible_from(i)
        _, want, _, _ = w.visibility(sens[i:i + 1], 58.0, 45.0, 0.7, 6.0, pts)
        assert np.array_equal(got, want)
        union |= set(got.tolist())
        assert abs(cov[k] - 100.0 * len(union) / len(pts)) < 1e-9
    p.close()
    w.close()


def test_casarea_scenario_through_the_mirror(eng):
    """config 1 on the casarea map: grid -> connectNodes (GPU, 847 rays per pair) -> A*, against the path the
    reference's compiled sources found on the same map"""
    from sspp_b200 import scenes
    g = np.load(os.path.join(GOLD, "casarea_s1m.npz"))
    occ2 = np.unpackbits(g["occ2d"], axis=1)[:, :640].astype(bool)
    known2 = np.unpackbits(g["known2d"], axis=1)[:, :640].astype(bool)
    nz = 40
    p = host_lib.HostPlanner(1.5, device=0)
    p.map_create(0.05, [32768 - 320, 32768 - 240, 32768 - 20], [640, 480, nz], True)
    p.map_upload_dense(scenes.pack_bits(np.repeat(occ2[None], nz, axis=0)), scenes.pack_bits(np.repeat(known2[None], nz, axis=0)))
    p.set_goal([5, 5, 0.5, 0, 0, 0, 0], 0.6)
    p.generate_grid([0, 0, 0], [10, 10, 1], 1.0, False, 90.0)
    assert np.array_equal(p.nodes()[0], g["poses"])
    edges, tested = p.connect()
    assert (edges, tested) == (int(g["connections"]), int(g["collision_calls"]))
    path, created, expanded, _ = p.astar([-2, -5, 0.5, 0, 0, 0, 0])
    assert len(path) > 5 and np.array_equal(path, g["path"])
    assert (created, expanded) == (int(g["created"]), int(g["expanded"]))
    p.close()
