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


def test_coverage_cache_follows_map_and_node_changes(eng, orc):
    """The coverage heuristic caches a way-point's visible set.  The cache has to follow the world: after the
    map changes between two queries (the reactive path: setOccupied / setFree / a new cloud) the answer is the
    oracle's for the NEW map, and a way-point created after another was removed (its memory may be reused)
    gets its own visible set, not the removed one's."""
    from sspp_b200 import scenes
    res = 0.25
    km, dm, occ, known = scenes.synthetic_map(n_vox=96, res=res, seed=4, n_boxes=10, n_spheres=8, region=16.0)
    pts = scenes.surface_points(occ, known, km, dm, res, 30000, seed=45)
    cam = dict(mount_xyz=[0.1, 0.0, -0.05], mount_rpy=[0.0, 0.3, 0.0], hfov=58.0, vfov=45.0, near=0.7, far=6.0)
    p = host_lib.HostPlanner(1.6, device=0, coverage=cam)
    p.map_create(res, km, dm, True)
    p.map_upload_dense(occ, known)
    p.set_model_points(pts)
    p.generate_grid([0, 0, 0], [14, 14, 4], 1.5, True, 90.0)
    poses, _ = p.nodes()
    sens = p.sensor_poses()
    w = orc.world(res, True)
    w.build_dense(km, dm, occ, known)
    # a way-point that sees something
    seen = [(i, p.visible_from(i)) for i in range(0, len(poses), 7)]
    i, before = max(seen, key=lambda t: len(t[1]))
    assert len(before) > 20
    _, want, _, _ = w.visibility(sens[i:i + 1], 58.0, 45.0, 0.7, 6.0, pts)
    assert np.array_equal(before, want)
    # a wall right in front of that sensor, on the engine and on the oracle alike
    q = sens[i, 3:7]
    fwd = np.array([1 - 2 * (q[1] ** 2 + q[2] ** 2), 2 * (q[0] * q[1] + q[3] * q[2]), 2 * (q[0] * q[2] - q[3] * q[1])])
    wall = sens[i, :3] + 0.9 * fwd
    p.map_set_box(wall, [1.6, 1.6, 1.6], True)
    w.set_box(wall, [1.6, 1.6, 1.6], True)
    after = p.visible_from(i)
    _, want2, _, _ = w.visibility(sens[i:i + 1], 58.0, 45.0, 0.7, 6.0, pts)
    assert np.array_equal(after, want2)
    assert len(after) < len(before)  # the stale answer would have been `before`
    # remove that way-point and insert a new one elsewhere: whatever address it gets, it is a new way-point
    assert p.remove_node(poses[i])
    far = poses[i].copy()
    far[:3] += [0.37, -0.21, 0.11]
    p.insert_nodes(far[None, :])
    poses2, _ = p.nodes()
    j = int(np.nonzero(np.all(np.abs(poses2 - far) < 1e-12, axis=1))[0][0])
    sens2 = p.sensor_poses()
    _, want3, _, _ = w.visibility(sens2[j:j + 1], 58.0, 45.0, 0.7, 6.0, pts)
    assert np.array_equal(p.visible_from(j), want3)
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


def test_connect_to_nn_with_map_matches_reference(eng, ref):
    """connectToNN with a map behind the heuristic: the reference tests every yaw pair through
    getLineStatusBoundingBox; the mirror tests each segment once on the GPU — same links"""
    from sspp_b200 import scenes
    res = 0.3
    km, dm, occ, known = scenes.synthetic_map(n_vox=64, res=res, seed=13, n_boxes=14, n_spheres=8, region=16.0)
    w = ref.world(res, True)
    w.build_dense(km, dm, occ, known)
    r = ref.planner(1.5, w)
    p = host_lib.HostPlanner(1.5, device=0)
    p.map_create(res, km, dm, True)
    p.map_upload_dense(occ, known)
    for pl in (r, p):
        pl.generate_grid([0, 0, 0], [8, 8, 1], 1.0, True, 90.0)
        pl.generate_grid([0.5, 0.5, 0.5], [6, 6, 0], 0.5, True, 90.0)
    rng = np.random.RandomState(7)
    hull1 = np.stack([rng.randint(-4, 5, 80), rng.randint(-4, 5, 80), rng.choice([-0.5, 0.5], 80)], axis=1).astype(np.float32)
    hull2 = np.stack([rng.randint(-5, 8, 120) * 0.5, rng.randint(-5, 8, 120) * 0.5, np.full(120, 0.5)], axis=1).astype(np.float32)
    r.connect_to_nn(hull1, hull2, 90.0)
    p.connect_to_nn(hull1, hull2, 90.0)
    (rp_r, col_r), (rp_h, col_h) = r.csr(), p.csr()
    assert 0 < len(col_r)
    assert np.array_equal(rp_h, rp_r) and np.array_equal(col_h, col_r)
    r.close()
    p.close()
    w.close()


def test_clustering_is_connected_components(eng):
    """clusteringPointCloud of the coverage heuristic = connected components of the radius graph (built by the
    engine's radius search), small components dropped — checked against scipy on the same points"""
    from scipy.sparse import coo_matrix
    from scipy.sparse.csgraph import connected_components
    from scipy.spatial import cKDTree
    cam = dict(mount_xyz=[0, 0, 0], mount_rpy=[0, 0, 0], hfov=58.0, vfov=45.0, near=0.7, far=6.0)
    p = host_lib.HostPlanner(1.5, device=0, coverage=cam)
    rng = np.random.RandomState(1)
    blobs = [rng.normal(c, 0.4, (n, 3)) for c, n in (((0, 0, 0), 400), ((5, 0, 0), 300), ((0, 6, 1), 200), ((9, 9, 9), 4))]
    pts = np.concatenate(blobs + [rng.uniform(-12, 12, (30, 3))]).astype(np.float32)
    tol = 0.5
    p.set_clustering(tol, 10, 3.0)
    lab = p.cluster_labels(pts)
    d = pts.astype(np.float64)
    pairs = cKDTree(d).query_pairs(tol, output_type="ndarray")
    # the engine's predicate is sqrt(d2) <= tol in double; the KD-tree uses the same closed ball
    A = coo_matrix((np.ones(len(pairs)), (pairs[:, 0], pairs[:, 1])), shape=(len(pts), len(pts)))
    ncomp, comp = connected_components(A, directed=False)
    sizes = np.bincount(comp)
    for c in range(ncomp):
        members = np.nonzero(comp == c)[0]
        labs = np.unique(lab[members])
        assert len(labs) == 1
        assert (labs[0] >= 0) == (sizes[c] >= 10)
    big = [c for c in range(ncomp) if sizes[c] >= 10]
    assert len(np.unique(lab[lab >= 0])) == len(big) >= 3
    p.close()


def test_dynamic_sampling_loop(eng):
    """PathPlanner::dynamicNodesGenerationAndConnection with the coverage heuristic: coarse samples first, then
    finer samples around what is still unseen, stages linked by connectToNN; coverage can only grow and the
    graph stays connected enough to plan on"""
    from sspp_b200 import scenes
    res = 0.25
    km, dm, occ, known = scenes.synthetic_map(n_vox=96, res=res, seed=4, n_boxes=6, n_spheres=4, region=10.0)
    pts = scenes.surface_points(occ, known, km, dm, res, 20000, seed=45)
    cam = dict(mount_xyz=[0.0, 0.0, 0.0], mount_rpy=[0.0, 0.0, 0.0], hfov=90.0, vfov=70.0, near=0.3, far=7.0)
    coarse = host_lib.HostPlanner(1.6, device=0, coverage=cam)
    coarse.map_create(res, km, dm, True)
    coarse.map_upload_dense(occ, known)
    coarse.set_model_points(pts)
    coarse.set_clustering(0.6, 10, 2.0)
    n0, e0, cov0 = coarse.dynamic_sampling([0, 0, 0], [16, 16, 6], 4.0, 5.0)   # decrement > start: the coarse stage only
    fine = host_lib.HostPlanner(1.6, device=0, coverage=cam)
    fine.map_create(res, km, dm, True)
    fine.map_upload_dense(occ, known)
    fine.set_model_points(pts)
    fine.set_clustering(0.6, 10, 2.0)
    n1, e1, cov1 = fine.dynamic_sampling([0, 0, 0], [16, 16, 6], 4.0, 1.5)     # stages at 4.0, 2.5, 1.0
    assert n0 > 0 and cov0 > 0
    assert n1 > n0 and e1 > e0 and cov1 > cov0          # refinement adds way-points, links and coverage
    assert cov1 <= len(pts)
    rp, col = fine.csr()
    assert len(col) == e1 and col.max() < n1
    coarse.close()
    fine.close()


def test_dynamic_sampling_loop_equals_the_references_own_loop(eng, ref):
    """f3 pinned: the REFERENCE'S OWN PathPlanner::dynamicNodesGenerationAndConnection (src/pathplanner.cpp:765-891,
    compiled unmodified into oracle/_ref) is driven through a harness heuristic that implements the authored coverage
    semantics on the oracle (oracle/ref_harness.cpp: HarnessCoverage), and the mirror's GPU-backed loop must build
    the same search space: the same way-points in the same list order, the same adjacency, the same connection count
    and the same covered model points — over three stages (3.0 m, 1.9 m, 0.8 m samples; the fourth resolution is
    negative and ends the loop).  The scene is small because the reference checks for duplicate connections in
    O(E^2) after every stage (checkSearchSpaceDuplications, :568-610): 645 way-points / 38 864 connections take it 11 s."""
    from sspp_b200 import scenes
    res = 0.25
    km, dm, occ, known = scenes.synthetic_map(n_vox=64, res=res, seed=4, n_boxes=4, n_spheres=3, region=7.0)
    allp = scenes.surface_points(occ, known, km, dm, res, 20000, seed=45)
    c = allp[np.argmin(np.linalg.norm(allp, axis=1))]
    pts = np.ascontiguousarray(allp[np.linalg.norm(allp - c, axis=1) < 3.0][:1500])  # the part of the model to inspect
    cam = dict(mount_xyz=[0.0, 0.0, 0.0], mount_rpy=[0.0, 0.0, 0.0], hfov=90.0, vfov=70.0, near=0.3, far=5.0)
    w = ref.world(res, True)
    w.build_dense(km, dm, occ, known)
    r = ref.coverage_planner(1.6, w, cam)
    r.set_model_points(pts)
    r.set_clustering(0.6, 20, 0.8)
    p = host_lib.HostPlanner(1.6, device=0, coverage=cam)
    p.map_create(res, km, dm, True)
    p.map_upload_dense(occ, known)
    p.set_model_points(pts)
    p.set_clustering(0.6, 20, 0.8)
    args = ([0, 0, 0], [10, 10, 4], 3.0, 1.1)
    n_r, e_r, cov_r = r.dynamic_sampling(*args)
    n_h, e_h, cov_h = p.dynamic_sampling(*args)
    print("dynamic sampling: reference %d way-points, %d connections, %d of %d model points covered; mirror %d / %d / %d"
          % (n_r, e_r, cov_r, len(pts), n_h, e_h, cov_h))
    assert n_r > 300 and e_r > 10000 and 0 < cov_r <= len(pts)
    assert (n_h, e_h, cov_h) == (n_r, e_r, cov_r)
    poses_r, ids_r = r.nodes()
    poses_h, ids_h = p.nodes()
    assert np.array_equal(poses_h, poses_r) and np.array_equal(ids_h, ids_r)
    (rp_r, col_r), (rp_h, col_h) = r.csr(), p.csr()
    assert np.array_equal(rp_h, rp_r) and np.array_equal(col_h, col_r)
    r.close()
    p.close()
    w.close()
