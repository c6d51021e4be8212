"""The host C++ mirror of the reference's planner classes (sspp_b200/host) against the reference's
own compiled sources (oracle/_ref) and the committed golden vectors — CPU only: adjacency is
injected, so these tests exercise sampling, the de-duplicating node store, and the A* consumer."""
import os

import numpy as np
import pytest

import host_lib

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def test_regular_grid_matches_golden(built):
    """src/distance_heuristic_test.cpp:52-94: 10x10x2 grid at 1.0 with four yaws -> 1452 nodes, list order"""
    g = np.load(os.path.join(GOLD, "dht_known_answer.npz"))
    p = host_lib.HostPlanner(1.5)
    p.generate_grid([0, 0, 0], [10, 10, 2], 1.0, True, 90.0)
    poses, ids = p.nodes()
    assert np.array_equal(poses, g["poses"]) and np.array_equal(ids, g["ids"])
    p.close()


@pytest.mark.parametrize("res,size,orient", [(0.3, (3.0, 2.0, 0.9), None), (0.7, (5.0, 5.0, 0.0), 45.0), (1.0, (10, 10, 0), None),
                                             (0.1, (1.0, 0.5, 0.2), 120.0)])
def test_regular_grid_matches_reference_live(ref, res, size, orient):
    """non-representable float resolutions: the sample count follows the accumulated rounding of `double += float`"""
    start = [-2.12, 1.2, 0.2]
    r = ref.planner(1.5)
    r.generate_grid(start, size, res, orient is not None, orient or 360.0)
    pr, ir = r.nodes()
    p = host_lib.HostPlanner(1.5)
    p.generate_grid(start, size, res, orient is not None, orient or 360.0)
    ph, ih = p.nodes()
    assert np.array_equal(ph, pr) and np.array_equal(ih, ir)
    r.close()
    p.close()


def test_node_store_dedup_and_removal(ref):
    """insertNode refuses a pose within 1e-12 in all seven components (and the id still advances);
    removeNode unlinks; list order is newest first — same as the reference's list"""
    rng = np.random.RandomState(0)
    base = np.concatenate([rng.uniform(-5, 5, (300, 3)), np.tile([0, 0, 0, 1.0], (300, 1))], axis=1)
    dup = base[::7].copy()
    near = base[::11].copy()
    near[:, 0] += 5e-13                      # inside the tolerance: a duplicate
    far = base[::13].copy()
    far[:, 1] += 5e-12                       # outside: a new node
    yaw = base[::5].copy()
    yaw[:, 5], yaw[:, 6] = np.sin(0.4), np.cos(0.4)   # same position, other orientation: a new node
    poses = np.concatenate([base, dup, near, far, yaw])
    r = ref.planner(1.0)
    r.insert_nodes(poses)
    p = host_lib.HostPlanner(1.0)
    p.insert_nodes(poses)
    (pr, ir), (ph, ih) = r.nodes(), p.nodes()
    assert len(ph) == 300 + len(far) + len(yaw)
    assert np.array_equal(ph, pr) and np.array_equal(ih, ir)
    assert p.remove_node(base[10]) and not p.remove_node(base[10])
    assert len(p.nodes()[0]) == len(ph) - 1
    r.close()
    p.close()


def test_astar_known_answer(built):
    """A* over the golden adjacency of the distance_heuristic_test scenario: 153 expanded, 9480 created,
    path cost 6.656854249492381 — the numbers the reference's compiled sources produce"""
    g = np.load(os.path.join(GOLD, "dht_known_answer.npz"))
    p = host_lib.HostPlanner(1.5)
    p.generate_grid([0, 0, 0], [10, 10, 2], 1.0, True, 90.0)
    p.set_adjacency(g["row_ptr"], g["col_idx"])
    edges, _ = p.connect()
    assert edges == 80580
    rp, col = p.csr()
    assert np.array_equal(rp, g["row_ptr"]) and np.array_equal(col, g["col_idx"])
    p.set_goal([4, 4, 2, 0, 0, 0, 0], 1.0)
    path, created, expanded, order = p.astar([0, 0, 0, 0, 0, 0, 0])
    assert (created, expanded) == (int(g["created"]), int(g["expanded"])) == (9480, 153)
    assert np.array_equal(path, g["path"])
    assert path[-1, 7] == 6.656854249492381
    p.close()


def test_astar_reactive_golden(built):
    g = np.load(os.path.join(GOLD, "reactive_sphere.npz"))
    p = host_lib.HostPlanner(1.5)
    p.generate_grid([-2.12, 1.2, 0.2], [10, 10, 0], 1.0, False, 90.0)
    assert np.array_equal(p.nodes()[0], g["poses"])
    p.set_adjacency(g["row_ptr"], g["col_idx"])
    p.connect()
    p.set_goal([2, 2, 0, 0, 0, 0, 0], 0.5)
    path, created, expanded, _ = p.astar([-2.12, 1.2, 0.2, 0, 0, 0, 0])
    assert np.array_equal(path, g["path"]) and (created, expanded) == (int(g["created"]), int(g["expanded"]))
    p.close()


def test_astar_matches_reference_on_obstacle_graphs(ref):
    """random obstacle scenes: the reference plans over its own connectNodes graph; the mirror plans over
    the same adjacency; path, g/h/f of every path node, nodes created and expanded must be identical —
    including goals that cannot be reached (both give up the same way)"""
    from sspp_b200 import scenes
    res = 0.3
    km, dm, occ, known = scenes.synthetic_map(n_vox=64, res=res, seed=13, n_boxes=14, n_spheres=8, region=16.0)
    w = ref.world(res, True)
    w.build_dense(km, dm, occ, known)
    r = ref.planner(1.5, w)
    r.generate_grid([0, 0, 0], [14, 14, 2], 1.0, False, 360.0)
    poses, _ = r.nodes()
    r.connect()
    rp, col = r.csr()
    assert 0 < len(col) < 18 * len(poses)
    p = host_lib.HostPlanner(1.5)
    p.generate_grid([0, 0, 0], [14, 14, 2], 1.0, False, 360.0)
    p.set_adjacency(rp, col)
    p.connect()
    rng = np.random.RandomState(5)
    found = 0
    for trial in range(12):
        a, b = poses[rng.randint(len(poses))], poses[rng.randint(len(poses))]
        goal = [b[0] + 0.2, b[1] - 0.1, b[2], 0, 0, 0, 0]
        start = [a[0] + 0.3, a[1] + 0.3, a[2], 0, 0, 0, 0]
        r.set_goal(goal, 0.6)
        p.set_goal(goal, 0.6)
        path_r, created_r, expanded_r = r.astar(start)
        path_h, created_h, expanded_h, _ = p.astar(start)
        assert np.array_equal(path_h, path_r), trial
        assert created_h == created_r and (expanded_r < 0 or expanded_h == expanded_r), trial
        found += len(path_r) > 0
    assert found >= 4
    r.close()
    p.close()
    w.close()


def test_connect_without_engine_is_an_error(built):
    """no CPU loop behind connectNodes(): without an adjacency source the call fails loudly"""
    p = host_lib.HostPlanner(1.5)
    p.generate_grid([0, 0, 0], [2, 2, 0], 1.0)
    with pytest.raises(host_lib.HostError):
        p.connect()
    p.close()


def test_sensor_pose_composition(built):
    """Sensors::robot2sensorTransformation = T_robot * T_mount * Rx(+90deg): identity robot + identity mount
    gives the pure Rx(+90deg) quaternion; a yawed, translated robot moves the mount point accordingly"""
    out = host_lib.robot2sensor([0, 0, 0], [0, 0, 0], [[0, 0, 0, 0, 0, 0, 1.0]])
    assert np.allclose(out[0], [0, 0, 0, np.sin(np.pi / 4), 0, 0, np.cos(np.pi / 4)], atol=1e-15)
    yaw = np.pi / 2
    out = host_lib.robot2sensor([1.0, 0, 0.5], [0, 0, 0], [[2, 3, 4, 0, 0, np.sin(yaw / 2), np.cos(yaw / 2)]])
    assert np.allclose(out[0, :3], [2, 4, 4.5], atol=1e-12)


def test_file_formats_match_reference(ref, tmp_path):
    """loadRegularGrid (three pose files in lock-step), saveSearchSpace (14 columns, 3 decimals) and
    loadSearchSpace — including the reference's habit of pushing the last edge twice when the file ends
    with a newline (src/pathplanner.cpp:319-356) — against the reference's own code"""
    g = np.load(os.path.join(GOLD, "uav_fixture_R25.npz"))
    poses = g["poses"][::-1][:400]                      # file order = insertion order = reverse list order
    cams = [poses.copy(), poses.copy()]
    cams[0][:, 2] -= 0.055
    cams[1][:, 0] += 0.1
    files = []
    for k, arr in enumerate([poses] + cams):
        fn = str(tmp_path / f"ss{k}.txt")
        np.savetxt(fn, arr, fmt="%.6f")
        files.append(fn)
    r = ref.planner(2.5)
    r.load_regular_grid(*files)
    p = host_lib.HostPlanner(2.5)
    p.load_regular_grid(*files)
    (pr, ir), (ph, ih) = r.nodes(), p.nodes()
    assert len(ph) > 50 and np.array_equal(ph, pr) and np.array_equal(ih, ir)
    # same graph on both sides, then save with each implementation: identical bytes
    r.connect()
    rp, col = r.csr()
    p.set_adjacency(rp, col)
    p.connect()
    f_ref, f_host = str(tmp_path / "ref_edges.txt"), str(tmp_path / "host_edges.txt")
    r.save_search_space(f_ref)
    p.save_search_space(f_host)
    assert open(f_ref, "rb").read() == open(f_host, "rb").read() and os.path.getsize(f_ref) > 1000
    # load the edge list back with each implementation: same nodes, same adjacency (3-decimal poses,
    # last edge doubled)
    r2 = ref.planner(2.5)
    r2.load_search_space(f_ref)
    p2 = host_lib.HostPlanner(2.5)
    p2.load_search_space(f_ref)
    (pr2, _), (ph2, _) = r2.nodes(), p2.nodes()
    assert np.array_equal(ph2, pr2)
    rp_r, col_r = r2.csr()
    rp_h, col_h = p2.csr()
    assert np.array_equal(rp_h, rp_r) and np.array_equal(col_h, col_r)
    assert len(col_r) == len(col) + 1   # the doubled last edge
    for x in (r, p, r2, p2):
        x.close()


def test_binary_graph_file_round_trip(built, tmp_path):
    """the lossless binary dump brings back list order, ids, full-precision poses and the children order;
    A* over the reloaded graph walks the same path"""
    g = np.load(os.path.join(GOLD, "dht_known_answer.npz"))
    p = host_lib.HostPlanner(1.5)
    p.generate_grid([0, 0, 0], [10, 10, 2], 1.0, True, 90.0)
    p.set_adjacency(g["row_ptr"], g["col_idx"])
    p.connect()
    fn = str(tmp_path / "graph.ssppcsr")
    p.save_binary(fn)
    q = host_lib.HostPlanner(1.5)
    q.load_binary(fn)
    (pp, ip), (pq, iq) = p.nodes(), q.nodes()
    assert np.array_equal(pp, pq) and np.array_equal(ip, iq)
    rp, col = q.csr()
    assert np.array_equal(rp, g["row_ptr"]) and np.array_equal(col, g["col_idx"])
    q.set_goal([4, 4, 2, 0, 0, 0, 0], 1.0)
    path, created, expanded, _ = q.astar([0, 0, 0, 0, 0, 0, 0])
    assert np.array_equal(path, g["path"]) and (created, expanded) == (9480, 153)
    with open(fn, "r+b") as f:
        f.write(b"XXXX")
    with pytest.raises(host_lib.HostError):
        host_lib.HostPlanner(1.5).load_binary(fn)
    p.close()
    q.close()


def test_connect_to_nn_matches_reference(ref):
    """PathPlanner::connectToNN (src/pathplanner.cpp:658-763): nearest neighbour per hull point, all yaw
    pairs, links in both directions, nodes looked up by pose — same adjacency as the reference's code"""
    r = ref.planner(1.5)
    p = host_lib.HostPlanner(1.5)
    for pl in (r, p):
        pl.generate_grid([0, 0, 0], [6, 6, 1], 1.0, True, 90.0)        # integer / half coordinates: exact as float
        pl.generate_grid([0.5, 0.5, 0.5], [4, 4, 0], 0.5, True, 90.0)
    n = len(p.nodes()[0])
    p.set_adjacency(np.zeros(n + 1, dtype=np.int64), np.zeros(0, dtype=np.int32))
    rng = np.random.RandomState(3)
    hull1 = np.stack([rng.randint(-3, 4, 40), rng.randint(-3, 4, 40), rng.choice([-0.5, 0.5], 40)], axis=1).astype(np.float32)
    hull2 = np.stack([rng.randint(-3, 6, 60) * 0.5, rng.randint(-3, 6, 60) * 0.5, np.full(60, 0.5)], axis=1).astype(np.float32)
    hull2[:5] = hull1[:5]            # same position: linked across yaws without a test
    hull2[5] = [9.25, 9.25, 9.25]    # not a node: no link
    r.connect_to_nn(hull1, hull2, 90.0)
    p.connect_to_nn(hull1, hull2, 90.0)
    (rp_r, col_r), (rp_h, col_h) = r.csr(), p.csr()
    assert len(col_r) > 100
    assert np.array_equal(rp_h, rp_r) and np.array_equal(col_h, col_r)
    # (the reference's two-argument-less constructor leaves MAXNODES uninitialised, src/pathplanner.cpp:40-50, so its
    # value after `+= numConnections` is not comparable; the mirror starts from 0)
    assert p.maxnodes() == len(col_h)
    r.close()
    p.close()
