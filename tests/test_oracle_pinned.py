"""The oracle pinned: (1) the reference's own planner sources (oracle/_ref) reproduce the known
answers of src/distance_heuristic_test.cpp and the committed golden vectors; (2) the restated
connectNodes loop (oracle_capi.cpp orc_connect) equals the reference's loop on the same inputs."""
import os

import numpy as np
import pytest

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def col_hash(rp, col):
    rows = np.repeat(np.arange(len(rp) - 1, dtype=np.uint64), np.diff(rp))
    return np.bitwise_xor.reduce((col.astype(np.uint64) + np.uint64(1)) * np.uint64(0x9E3779B97F4A7C15) ^
                                 rows * np.uint64(0xC2B2AE3D27D4EB4F))


def test_reference_known_answer_distance_heuristic_test(ref):
    """1452 nodes, 80 580 connections, degree 27/75, 153 expanded / 9 480 created,
    6-node path of cost 6.656854249492381 — re-derived from the reference's compiled sources."""
    p = ref.planner(1.5)
    p.set_goal([4, 4, 2, 0, 0, 0, 0], 1.0)
    p.generate_grid([0, 0, 0], [10, 10, 2], 1.0, True, 90.0)
    poses, ids = p.nodes()
    assert len(poses) == 1452
    assert p.connect() == 80580
    rp, col = p.csr()
    deg = np.diff(rp)
    assert (deg.min(), deg.max()) == (27, 75) and abs(deg.mean() - 55.49586776859504) < 1e-12
    path, created, expanded = p.astar([0, 0, 0, 0, 0, 0, 0])
    assert (created, expanded) == (9480, 153)
    assert path[:, :3].tolist() == [[0, 0, 0], [1, 1, 0], [2, 2, 0], [3, 2, 1], [3, 3, 1], [4, 4, 1]]
    assert path[-1, 7] == 6.656854249492381
    g = np.load(os.path.join(GOLD, "dht_known_answer.npz"))
    assert np.array_equal(poses, g["poses"]) and np.array_equal(rp, g["row_ptr"]) and np.array_equal(col, g["col_idx"])
    assert np.array_equal(path, g["path"])
    p.close()


def test_restated_connect_matches_reference_no_map(orc):
    g = np.load(os.path.join(GOLD, "dht_known_answer.npz"))
    rp, col, cnt = orc.connect(None, g["poses"][:, :3], 1.5)
    assert np.array_equal(rp, g["row_ptr"]) and np.array_equal(col, g["col_idx"])
    # 76 224 pairs go through the heuristic call, 4 356 are same-position / other-yaw
    assert (cnt[0], cnt[1]) == (76224, 4356)


def test_restated_connect_matches_reference_with_map(orc):
    """reactive scene: octomap res 0.3, free box 15x15x10, occupied sphere r=1 (golden from oracle/_ref)."""
    g = np.load(os.path.join(GOLD, "reactive_sphere.npz"))
    w = orc.world(0.3, False)
    w.set_box([0, 0, 0], [15, 15, 10], False)
    w.set_keys(g["sphere_keys"], True)
    rp, col, cnt = orc.connect(w, g["poses"][:, :3], 1.5)
    assert np.array_equal(rp, g["row_ptr"]) and np.array_equal(col, g["col_idx"])
    assert cnt[0] == int(g["collision_calls"]) == 840 and int(g["connections"]) == 777
    w.close()


def test_restated_connect_matches_reference_uav_fixture(orc):
    g = np.load(os.path.join(GOLD, "uav_fixture_R25.npz"))
    rp, col, cnt = orc.connect(None, g["poses"][:, :3], 2.5)
    assert np.array_equal(rp, g["row_ptr"])
    assert col_hash(rp, col) == g["col_hash"] and np.array_equal(col[:4096], g["col_head"])
    assert len(col) == int(g["connections"])


def test_reference_live_equals_golden_reactive(ref):
    """when oracle/_ref is present, regenerate the reactive scene and compare with the committed vectors"""
    g = np.load(os.path.join(GOLD, "reactive_sphere.npz"))
    w = ref.world(0.3, False)
    w.set_box([0, 0, 0], [15, 15, 10], False)
    w.set_keys(g["sphere_keys"], True)
    p = ref.planner(1.5, w)
    p.set_goal([2, 2, 0, 0, 0, 0, 0], 0.5)
    p.generate_grid([-2.12, 1.2, 0.2], [10, 10, 0], 1.0, False, 90.0)
    poses, _ = p.nodes()
    assert np.array_equal(poses, g["poses"])
    assert p.connect() == int(g["connections"])
    rp, col = p.csr()
    assert np.array_equal(col, g["col_idx"])
    path, created, expanded = p.astar([-2.12, 1.2, 0.2, 0, 0, 0, 0])
    assert np.array_equal(path, g["path"]) and created == int(g["created"]) and expanded == int(g["expanded"])
    p.close()
    w.close()
