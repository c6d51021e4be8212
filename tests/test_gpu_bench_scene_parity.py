"""Oracle parity ON the benchmarked scenes (S4a, S4b, S5 of SURVEY.md §8 d; BASELINE configs[3] and [4]).

At 10^6 - 10^7 nodes the oracle's O(N^2) connectNodes cannot run, so the full-size tests elsewhere compare
the engine with itself.  Here the GPU builds the WHOLE graph of the real scene and the oracle is put to
sampled windows of it: for every source row inside a window, every ordered in-radius pair (i, j) of that
row is handed to the oracle's getLineStatusBoundingBox (reference predicate: src/pathplanner.cpp:541-549,
src/distance_heuristic.cpp:104-111) on an octree built over the voxels the row's rays can reach, and the
row of the GPU's full-scene CSR must be exactly {j : same position or status == kFree}, in list order.
The windows sit on obstacle surfaces (where the exact walks happen) and at random places; the rows are
tallied by the way the engine answered them (sspp_gpu_connect_row_classes: coarse pyramid, row key box,
pair by pair) and every class must be present.  A disagreeing pair is enumerated ray by ray.

The visibility half: viewpoints drawn from the whole 65 536-viewpoint set of the bench scene (not its
first 24), index sets against the oracle's brute-force frustum test + castRay.
"""
import os

import numpy as np
import pytest

from sspp_b200 import scenes

pytestmark = pytest.mark.gpu
BOX = (0.5, 0.5, 0.3)
THREADS = max(1, min(os.cpu_count() or 1, 32))


def _window_cores(rng, occ, known, km, dm, res, extent, side, n_surface, n_random):
    """Window cores (lo, hi): cubes of `side` metres centred on obstacle-surface points and at random places
    of the node cube."""
    cores = []
    surf = scenes.surface_points(occ, known, km, dm, res, 4 * n_surface, seed=int(rng.randint(1 << 30))).astype(np.float64)
    surf = surf[np.all(np.abs(surf) < extent / 2 - side, axis=1)][:n_surface]
    for c in surf:
        cores.append((c - side / 2, c + side / 2))
    for _ in range(n_random):
        c = rng.uniform(-extent / 2 + side, extent / 2 - side, 3)
        cores.append((c - side / 2, c + side / 2))
    return cores


def _enumerate_rays(eng, orc, w, res, start, end):
    """Ray-by-ray statuses of one pair on both sides, for the failure message."""
    offs = orc.ray_lattice(BOX, res)
    out = []
    for ox in offs[0]:
        for oy in offs[1]:
            for oz in offs[2]:
                o = np.array([ox, oy, oz])
                s, e = (start + o)[None, :], (end + o)[None, :]
                g = int(eng.check_segments(s, e, (0.0, 0.0, 0.0))[0])
                c = int(w.line_status(s, e)[0])
                if g != c:
                    out.append((tuple(o), g, c))
    return out


def _check_scene(eng, orc, res, km, dm, occ, known, nodes, radius, cores, min_pairs, want_classes=(1, 3)):
    n = len(nodes)
    eng.set_nodes(nodes)
    st = eng.connect(radius=radius, box=BOX)
    rp, col = eng.connect_fetch(st)
    cls = eng.connect_row_classes(st)
    tally = {"pairs": 0, "edges": 0, "rejected": 0, "rows_by_class": {0: 0, 1: 0, 2: 0, 3: 0}, "pairs_by_class": {0: 0, 1: 0, 2: 0, 3: 0},
             "rejected_in_pairwise_rows": 0}
    reach = radius + max(BOX)
    for lo, hi in cores:
        rows, i, j, same = scenes.window_pairs(nodes, lo, hi, radius)
        if len(rows) == 0:
            continue
        kmin, dims, so, sk = scenes.map_window(occ, known, km, lo - reach, hi + reach, res)
        w = orc.world(res, True)
        w.build_dense(kmin, dims, so, sk)
        status = w.line_status_bbox_mt(nodes[i], nodes[j], BOX, THREADS)
        want_edge = same | (status == 0)
        order = np.lexsort((j, i))  # rows ascending, columns in list order (both loops of connectNodes walk the list)
        i, j, want_edge, status = i[order], j[order], want_edge[order], status[order]
        want_keys = (i.astype(np.int64) * n + j)[want_edge]
        seg = [col[rp[r]:rp[r + 1]] for r in np.sort(rows)]
        got_keys = np.concatenate([np.full(len(s), r, dtype=np.int64) * n + s for r, s in zip(np.sort(rows), seg)]) if seg else np.zeros(0, np.int64)
        if not np.array_equal(got_keys, want_keys):
            bad = np.setxor1d(got_keys, want_keys)[:5]
            msg = []
            for k in bad:
                a, b = int(k // n), int(k % n)
                msg.append("pair (%d -> %d): GPU edge %s, oracle status %s, rays that differ (offset, gpu, oracle): %s"
                           % (a, b, bool(np.isin(k, got_keys)), w.line_status_bbox(nodes[a:a + 1], nodes[b:b + 1], BOX)[0],
                              _enumerate_rays(eng, orc, w, res, nodes[a], nodes[b])))
            w.close()
            raise AssertionError("GPU CSR differs from the oracle in window %s..%s:\n%s" % (np.round(lo, 2), np.round(hi, 2), "\n".join(msg)))
        w.close()
        tally["pairs"] += len(i)
        tally["edges"] += int(want_edge.sum())
        tally["rejected"] += int((~want_edge).sum())
        for c in range(4):
            tally["rows_by_class"][c] += int((cls[rows] == c).sum())
            tally["pairs_by_class"][c] += int((cls[i] == c).sum())
        tally["rejected_in_pairwise_rows"] += int(((cls[i] == 3) & ~want_edge).sum())
    print("oracle parity on the full scene:", tally, "| whole graph: tested %d edges %d pyramid_rows %d box_pairs %d lattice_pairs %d"
          % (st.tested_pairs, st.edges, st.pyramid_rows, st.box_pairs, st.lattice_pairs))
    assert tally["pairs"] >= min_pairs
    for c in want_classes:
        assert tally["pairs_by_class"][c] >= min_pairs // 20, "too few sampled pairs in rows of class %d" % c
    assert tally["rejected_in_pairwise_rows"] >= min_pairs // 50 and tally["edges"] > tally["rejected"] > 0
    return st, tally


def test_s4a_full_scene_against_oracle_windows(eng, orc):
    """The configuration bench.py times: 512^3 voxels @0.3 m, 10^6 lattice nodes, R 1.5, 27 rays per pair."""
    res = 0.3
    km, dm, occ, known = scenes.synthetic_map(n_vox=512, res=res, seed=42)
    eng.map_create(res, km, dm, True)
    eng.map_upload_dense(occ, known)
    nodes = scenes.lattice_nodes(100, 1.0)
    rng = np.random.RandomState(101)
    cores = _window_cores(rng, occ, known, km, dm, res, 100.0, 7.0, 10, 4)
    st, _ = _check_scene(eng, orc, res, km, dm, occ, known, nodes, 1.5, cores, 60000)
    assert st.tested_pairs == 17701200


def test_s4b_full_scene_against_oracle_windows(eng, orc):
    """S4b: the same map, 10^6 nodes uniform in the 100 m cube (seed 43)."""
    res = 0.3
    km, dm, occ, known = scenes.synthetic_map(n_vox=512, res=res, seed=42)
    eng.map_create(res, km, dm, True)
    eng.map_upload_dense(occ, known)
    nodes = scenes.uniform_nodes(1_000_000, extent=100.0, seed=43)
    rng = np.random.RandomState(102)
    cores = _window_cores(rng, occ, known, km, dm, res, 100.0, 7.0, 12, 4)
    _check_scene(eng, orc, res, km, dm, occ, known, nodes, 1.5, cores, 60000)


def test_s5_full_scene_against_oracle_windows(eng, orc):
    """S5 (BASELINE configs[4]): 1024^3 voxels @0.15 m, 100 rays per pair; the 216^3 lattice @0.7 m
    (10 077 696 nodes, R 1.05) and 10^7 uniform nodes (seed 44; ~48 pairs per row: rows longer than a warp)."""
    res = 0.15
    km, dm, occ, known = scenes.synthetic_map(n_vox=1024, res=res, seed=42)
    eng.map_create(res, km, dm, True)
    eng.map_upload_dense(occ, known)
    rng = np.random.RandomState(103)
    nodes = scenes.lattice_nodes(216, 0.7)
    cores = _window_cores(rng, occ, known, km, dm, res, 150.0, 4.2, 8, 3)
    st, _ = _check_scene(eng, orc, res, km, dm, occ, known, nodes, 1.05, cores, 20000)
    assert st.rays_per_pair == 100
    del nodes
    nodes = scenes.uniform_nodes(10_000_000, extent=100.0, seed=44)
    cores = _window_cores(rng, occ, known, km, dm, res, 100.0, 2.6, 6, 2)
    _check_scene(eng, orc, res, km, dm, occ, known, nodes, 1.05, cores, 20000)


def test_bench_visibility_scene_against_oracle_sample(eng, orc):
    """The visibility scene of bench.py (65 536 viewpoints x 4*10^6 surface points of the S4a map): 96 viewpoints
    drawn from the whole set, index sets equal to the oracle's; thousands of occlusion rays, occluded and not."""
    res = 0.3
    km, dm, occ, known = scenes.synthetic_map(n_vox=512, res=res, seed=42)
    eng.map_create(res, km, dm, True)
    eng.map_upload_dense(occ, known)
    pts = scenes.surface_points(occ, known, km, dm, res, 4_000_000, seed=45)
    poses = scenes.viewpoints_near(pts, 65536, standoff=3.0, seed=46)
    eng.set_model_points(pts)
    st, vp, idx = eng.visibility(poses, 58.0, 45.0, 0.7, 6.0)
    assert st.rays > 5e7 and 0 < st.visible_pairs < st.rays
    sel = np.sort(np.random.RandomState(104).choice(np.arange(24, len(poses)), 96, replace=False))
    rays_checked = visible_checked = 0
    for v in sel:
        # an octree over what this viewpoint can see: the far corners of the frustum lie
        # 6 * sqrt(1 + tan^2(29 deg) + tan^2(22.5 deg)) = 7.3 m from the sensor
        c = poses[v, :3]
        kmin, dims, so, sk = scenes.map_window(occ, known, km, c - 7.8, c + 7.8, res)
        w = orc.world(res, True)
        w.build_dense(kmin, dims, so, sk)
        near = np.nonzero(np.all(np.abs(pts - c.astype(np.float32)) < 7.6, axis=1))[0]
        vp_o, idx_o, rays_o, _ = w.visibility(poses[v:v + 1], 58.0, 45.0, 0.7, 6.0, pts[near])
        want = near[idx_o]
        got = idx[vp[v]:vp[v + 1]]
        assert np.array_equal(got, want), "viewpoint %d: GPU sees %d points, oracle %d; differing indices %s" % (
            v, len(got), len(want), np.setxor1d(got, want)[:10])
        rays_checked += int(rays_o)
        visible_checked += len(want)
        w.close()
    print("visibility parity on the bench scene: %d viewpoints, %d occlusion rays, %d visible" % (len(sel), rays_checked, visible_checked))
    assert rays_checked >= 2000 and 0 < visible_checked < rays_checked
