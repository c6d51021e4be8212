"""BASELINE configs[1] (coverage_heuristic_test: the SearchSpaceUAV / SearchSpaceCam_1.5m_1to4_NEW fixtures) at its
FULL size — all 14 765 row-aligned way-points and camera poses, not a sample:
  * connectNodes without a map against the graph the reference's OWN compiled connectNodes built on the same
    14 765 poses (tests/golden/config2_full.npz, made by tests/golden/make_golden.py --full): 868 286 connections;
  * connectNodes against the voxelised aircraft stand-in (48 rays per pair at 0.25 m): every ordered in-radius pair
    of the whole set against the oracle's getLineStatusBoundingBox;
  * visibility from all 14 765 camera poses; a 300-pose sample of the index sets against the oracle.
tools/config2_full.py times the same three passes."""
import os

import numpy as np
import pytest

from sspp_b200 import scenes

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
BOX = (0.5, 0.5, 0.3)
RES = 0.25


def _hash(rp, col):
    return np.bitwise_xor.reduce((col.astype(np.uint64) + 1) * np.uint64(0x9E3779B97F4A7C15) ^
                                 np.repeat(np.arange(len(rp) - 1, dtype=np.uint64), np.diff(rp)) * np.uint64(0xC2B2AE3D27D4EB4F))


def test_full_fixture_graph_equals_the_references(eng):
    g = np.load(os.path.join(GOLD, "config2_full.npz"))
    assert len(g["poses"]) == 14765
    eng.set_nodes(g["poses"][:, :3])
    st = eng.connect(radius=2.5, use_map=False)
    rp, col = eng.connect_fetch(st)
    assert st.edges == int(g["connections"]) == 868286
    assert np.array_equal(rp, g["row_ptr"]) and np.array_equal(col[:4096], g["col_head"])
    assert _hash(rp, col) == g["col_hash"]


def test_full_fixture_with_the_aircraft_map_against_the_oracle(eng, orc):
    g = np.load(os.path.join(GOLD, "config2_full.npz"))
    nodes = np.ascontiguousarray(g["poses"][:, :3])
    n = len(nodes)
    km, dm, occ, known = scenes.aircraft_model(RES)
    eng.map_create(RES, km, dm, True)
    eng.map_upload_dense(occ, known)
    eng.set_nodes(nodes)
    st = eng.connect(radius=2.5, box=BOX)
    rp, col = eng.connect_fetch(st)
    assert st.rays_per_pair == 48 and 0 < st.edges < st.in_radius_pairs
    w = orc.world(RES, True)
    w.build_dense(km, dm, occ, known)
    rows, i, j, same = scenes.window_pairs(nodes, nodes.min(0) - 1, nodes.max(0) + 1, 2.5)
    assert len(rows) == n and len(i) == st.in_radius_pairs and int((~same).sum()) == st.tested_pairs
    status = w.line_status_bbox_mt(nodes[i], nodes[j], BOX, max(1, min(os.cpu_count() or 1, 32)))
    w.close()
    want = same | (status == 0)
    order = np.lexsort((j, i))
    want_keys = (i[order].astype(np.int64) * n + j[order])[want[order]]
    got_keys = np.repeat(np.arange(n, dtype=np.int64), np.diff(rp)) * n + col
    bad = np.setxor1d(got_keys, want_keys)
    assert len(bad) == 0, "pairs that differ from the oracle: %s" % [(int(k // n), int(k % n)) for k in bad[:8]]
    assert np.array_equal(got_keys, want_keys)  # and in list order
    print("config 2 at full size with the map: %d pairs checked, %d edges, %d rejected" % (len(i), len(want_keys), int((~want).sum())))


def test_full_fixture_visibility_sample_against_the_oracle(eng, orc):
    g = np.load(os.path.join(GOLD, "config2_full.npz"))
    cam = np.ascontiguousarray(g["cam_rows"][g["ids"]])  # camera pose of every way-point, list order
    km, dm, occ, known = scenes.aircraft_model(RES)
    eng.map_create(RES, km, dm, True)
    eng.map_upload_dense(occ, known)
    pts = scenes.surface_points(occ, known, km, dm, RES, 1_000_000, seed=1234)
    eng.set_model_points(pts)
    st, vp, idx = eng.visibility(cam, 58.0, 45.0, 0.7, 6.0)
    assert st.rays > 0 and 0 < st.visible_pairs < st.rays
    sel = np.sort(np.random.RandomState(5).choice(len(cam), 300, replace=False))
    seen = 0
    for v in sel:
        c = cam[v, :3]
        near = np.nonzero(np.all(np.abs(pts - c.astype(np.float32)) < 7.6, axis=1))[0]
        kmin, dims, so, sk = scenes.map_window(occ, known, km, c - 7.8, c + 7.8, RES)
        w = orc.world(RES, True)
        w.build_dense(kmin, dims, so, sk)
        _, idx_o, _, _ = w.visibility(cam[v:v + 1], 58.0, 45.0, 0.7, 6.0, pts[near])
        w.close()
        assert np.array_equal(idx[vp[v]:vp[v + 1]], near[idx_o]), "camera pose %d" % v
        seen += len(idx_o)
    assert seen > 0
    print("config 2 at full size: %d camera poses, %d rays, %d visible pairs; %d poses checked against the oracle (%d visible)"
          % (len(cam), st.rays, st.visible_pairs, len(sel), seen))
