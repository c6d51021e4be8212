"""Coverage visibility (spec v1): frustum culling + occlusion castRay, CUDA vs oracle, through the C ABI."""
import numpy as np
import pytest

from sspp_b200 import scenes

pytestmark = pytest.mark.gpu


def scene(eng, orc, res=0.25, n_vox=96, seed=4):
    km, dm, occ, known = scenes.synthetic_map(n_vox=n_vox, res=res, seed=seed, n_boxes=10, n_spheres=8,
                                              region=n_vox * res * 0.7)
    eng.map_create(res, km, dm, True)
    eng.map_upload_dense(occ, known)
    w = orc.world(res, True)
    w.build_dense(km, dm, occ, known)
    pts = scenes.surface_points(occ, known, km, dm, res, 20000, seed=45)
    return w, pts


def test_visibility_matches_oracle(eng, orc):
    w, pts = scene(eng, orc)
    poses = scenes.viewpoints_near(pts, 60, standoff=3.0, seed=46)
    eng.set_model_points(pts)
    st, vp, idx = eng.visibility(poses, 58.0, 45.0, 0.7, 6.0)
    vp_o, idx_o, rays_o, _ = w.visibility(poses, 58.0, 45.0, 0.7, 6.0, pts)
    assert st.rays == rays_o and rays_o > 0
    assert np.array_equal(vp, vp_o) and np.array_equal(idx, idx_o)
    assert 0 < st.visible_pairs < st.rays  # some in-frustum points are occluded
    w.close()


def test_visibility_degenerate_inputs(eng, orc):
    w, pts = scene(eng, orc)
    eng.set_model_points(pts)
    # no viewpoints / no points
    st, vp, idx = eng.visibility(np.zeros((0, 7)), 58.0, 45.0, 0.7, 6.0)
    assert st.rays == 0 and vp.tolist() == [0]
    poses = scenes.viewpoints_near(pts, 8, standoff=2.0, seed=1)
    eng.set_model_points(np.zeros((0, 3), dtype=np.float32))
    st, vp, idx = eng.visibility(poses, 58.0, 45.0, 0.7, 6.0)
    assert st.visible_pairs == 0 and len(vp) == 9
    # viewpoint sitting exactly on a model point; wide and narrow frusta; many points per viewpoint
    eng.set_model_points(pts)
    poses[0, :3] = pts[5]
    for hf, vf, nr, fr in ((58.0, 45.0, 0.7, 6.0), (120.0, 100.0, 0.0, 12.0), (5.0, 5.0, 1.0, 4.0)):
        st, vp, idx = eng.visibility(poses, hf, vf, nr, fr)
        vp_o, idx_o, rays_o, _ = w.visibility(poses, hf, vf, nr, fr, pts)
        assert st.rays == rays_o and np.array_equal(vp, vp_o) and np.array_equal(idx, idx_o)
    w.close()


@pytest.mark.parametrize("n_pts,at_least,at_most", [(40000, 8193, 10**9), (6000, 513, 8192), (500, 2, 512)])
def test_visibility_large_segments_sorted(eng, orc, n_pts, at_least, at_most):
    """the three paths of the per-viewpoint sort: more than 8192 visible points for one viewpoint (bitonic
    network in global memory), 513..8192 (in shared memory), up to 512 (rank sort)"""
    res = 0.25
    km, dm = [32768 - 48] * 3, [96] * 3
    occ = np.zeros((96, 96, 96), dtype=bool)
    occ[:, :, 70:72] = True  # a wall at x ~ 5.5 m
    known = np.ones_like(occ)
    ob, kb = scenes.pack_bits(occ), scenes.pack_bits(known)
    eng.map_create(res, km, dm, True)
    eng.map_upload_dense(ob, kb)
    w = orc.world(res, True)
    w.build_dense(km, dm, ob, kb)
    rng = np.random.RandomState(0)
    pts = np.stack([np.full(n_pts, 5.6), rng.uniform(-3, 3, n_pts), rng.uniform(-3, 3, n_pts)], axis=1).astype(np.float32)
    pose = np.array([[0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 1.0]])  # looking along +x
    eng.set_model_points(pts)
    st, vp, idx = eng.visibility(pose, 90.0, 90.0, 0.5, 9.0)
    vp_o, idx_o, rays_o, _ = w.visibility(pose, 90.0, 90.0, 0.5, 9.0, pts)
    assert at_least <= st.visible_pairs <= at_most
    assert np.array_equal(vp, vp_o) and np.array_equal(idx, idx_o)
    w.close()


def test_visibility_with_reference_camera_poses(eng, orc):
    """sensor poses taken from the reference's own viewpoint fixture (resources/SearchSpaceCam_1.5m_1to4_NEW.txt,
    400 rows spread over the file: eight yaws, tilted cameras, heights -1.5..18 m) against a synthetic
    aircraft-like model (ellipsoid fuselage + wing slab) voxelised at 0.25 m — SURVEY.md S2"""
    import os
    g = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "cam_fixture_poses.npz"))
    poses = g["poses"]
    res, n = 0.25, 320
    half = n // 2
    c = (np.arange(n) - half + 0.5) * res
    zc = (np.arange(96) - 8 + 0.5) * res          # z from -2 m to 22 m
    Z, Y, X = np.meshgrid(zc, c, c, indexing="ij")
    fus = (X / 30.0) ** 2 + (Y / 3.0) ** 2 + ((Z - 5.0) / 3.0) ** 2 <= 1.0
    wing = (np.abs(X) < 4.0) & (np.abs(Y) < 30.0) & (np.abs(Z - 4.5) < 0.6)
    occ = fus | wing
    known = np.ones_like(occ)
    ob, kb = scenes.pack_bits(occ), scenes.pack_bits(known)
    km, dm = [32768 - half, 32768 - half, 32768 - 8], [n, n, 96]
    eng.map_create(res, km, dm, True)
    eng.map_upload_dense(ob, kb)
    w = orc.world(res, True)
    w.build_dense(km, dm, ob, kb)
    pts = scenes.surface_points(ob, kb, np.array(km), np.array(dm), res, 60000, seed=1234)
    eng.set_model_points(pts)
    st, vp, idx = eng.visibility(poses, 58.0, 45.0, 0.7, 6.0)
    sel = np.nonzero(np.diff(vp) > 0)[0][:48]                     # viewpoints that see something, plus a few that do not
    sel = np.concatenate([sel, np.nonzero(np.diff(vp) == 0)[0][:8]])
    assert len(sel) > 20 and st.visible_pairs > 0
    vp_o, idx_o, rays_o, _ = w.visibility(poses[sel], 58.0, 45.0, 0.7, 6.0, pts)
    for k, v in enumerate(sel):
        assert np.array_equal(idx[vp[v]:vp[v + 1]], idx_o[vp_o[k]:vp_o[k + 1]]), v
    w.close()


def test_interleaved_calls_do_not_share_state(eng, orc):
    """set_nodes -> visibility -> connect: the visibility pass must not disturb the node store's queued
    bounding box (regression test for a shared scratch buffer)"""
    w, pts = scene(eng, orc)
    nodes = scenes.uniform_nodes(1500, extent=14.0, seed=3)
    eng.set_nodes(nodes)
    eng.set_model_points(pts + 40.0)          # far from the nodes: a very different bounding box
    poses = scenes.viewpoints_near(pts, 16, standoff=3.0, seed=2)
    eng.visibility(poses, 58.0, 45.0, 0.7, 6.0)
    st = eng.connect(radius=1.6)
    rp, col = eng.connect_fetch(st)
    rp_o, col_o, _ = orc.connect(w, nodes, 1.6)
    assert np.array_equal(rp, rp_o) and np.array_equal(col, col_o)
    w.close()
