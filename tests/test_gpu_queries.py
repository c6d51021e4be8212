"""Batched segment / validity / castRay queries and map ingest, CUDA vs oracle, through the C ABI."""
import numpy as np
import pytest

from sspp_b200 import scenes

pytestmark = pytest.mark.gpu
BOX = (0.5, 0.5, 0.3)


def scene(eng, orc, res, treat, n_vox=64, seed=3, shell=0.2):
    km, dm, occ, known = scenes.synthetic_map(n_vox=n_vox, res=res, seed=seed, n_boxes=8, n_spheres=6,
                                              region=n_vox * res * 0.75, unknown_shell=shell)
    eng.map_create(res, km, dm, treat)
    eng.map_upload_dense(occ, known)
    w = orc.world(res, treat)
    w.build_dense(km, dm, occ, known)
    return w, (km, dm, occ, known)


def segments(rng, n, extent, scale):
    S = rng.uniform(-extent, extent, (n, 3))
    E = S + rng.normal(scale=scale, size=S.shape)
    k = n // 5
    S[:k] = np.round(S[:k] / 0.15) * 0.15
    E[:k] = np.round(E[:k] / 0.15) * 0.15
    E[k:k + k // 2, 1:] = S[k:k + k // 2, 1:]
    E[k + k // 2:2 * k] = S[k + k // 2:2 * k]
    return S, E


@pytest.mark.parametrize("res,treat", [(0.3, True), (0.3, False), (0.25, True), (0.15, False), (0.1, True), (0.05, True),
                                       (0.5, True), (1.0, False)])
def test_check_segments(eng, orc, res, treat):
    """10^5-ish random + adversarial segments per resolution: status of every segment equals
    getLineStatusBoundingBox of the oracle; disagreeing rays are enumerated (there must be none)."""
    w, _ = scene(eng, orc, res, treat)
    rng = np.random.RandomState(int(res * 100))
    n = 100000 if res >= 0.25 else (30000 if res >= 0.1 else 4000)
    S, E = segments(rng, n, 64 * res * 0.55, 1.2)
    got = eng.check_segments(S, E, BOX)
    want = w.line_status_bbox(S, E, BOX)
    bad = np.nonzero(got != want)[0]
    assert len(bad) == 0, f"{len(bad)} segments disagree, first: {S[bad[:3]]} -> {E[bad[:3]]}"
    assert len(np.unique(want)) >= 2
    w.close()


def test_check_segments_single_ray_box_zero(eng, orc):
    w, _ = scene(eng, orc, 0.3, True)
    rng = np.random.RandomState(5)
    S, E = segments(rng, 50000, 9.0, 2.0)
    assert np.array_equal(eng.check_segments(S, E, (0, 0, 0)), w.line_status(S, E))
    w.close()


@pytest.mark.parametrize("treat", [True, False])
def test_validity(eng, orc, treat):
    w, _ = scene(eng, orc, 0.3, treat)
    rng = np.random.RandomState(7)
    P = rng.uniform(-10.5, 10.5, (60000, 3))
    P[:5000] = np.round(P[:5000] / 0.3) * 0.3
    for box in (BOX, (0.0, 0.0, 0.0), (1.0, 0.7, 0.4)):
        got = eng.validity(P, box)
        want = w.cell_status_bbox(P, box)
        assert np.array_equal(got, want)
    assert set(np.unique(want)) == {0, 1, 2}
    w.close()


@pytest.mark.parametrize("ignore_unknown", [True, False])
def test_cast_rays(eng, orc, ignore_unknown):
    w, _ = scene(eng, orc, 0.25, True, shell=0.15)
    rng = np.random.RandomState(3)
    n = 100000
    O = rng.uniform(-7, 7, (n, 3)).astype(np.float32)
    D = rng.normal(size=(n, 3)).astype(np.float32)
    D[:500, 0] = 0
    D[500:700] = 0
    D[700:1000, 1:] = 0
    mr = rng.uniform(0.5, 8.0, n)
    mr[::3] = -1.0
    hit, keys = eng.cast_rays(O, D, ignore_unknown, mr)
    hit_o, keys_o = w.cast_rays(O, D, ignore_unknown, mr)
    assert np.array_equal(hit, hit_o) and np.array_equal(keys, keys_o)
    hit2, _ = eng.cast_rays(O, D, ignore_unknown, None)
    assert np.array_equal(hit2, w.cast_rays(O, D, ignore_unknown, None)[0])
    w.close()


def test_map_dense_roundtrip_and_padding(eng):
    """odd window sizes: padding voxels are never observed; download returns what was uploaded"""
    rng = np.random.RandomState(1)
    dims = (70, 37, 19)
    vol_k = rng.rand(dims[2], dims[1], dims[0]) < 0.7
    vol_o = (rng.rand(dims[2], dims[1], dims[0]) < 0.2) & vol_k
    occ, known = scenes.pack_bits(vol_o), scenes.pack_bits(vol_k)
    eng.map_create(0.2, [32768 - 35, 32768 - 18, 32768 - 9], dims, True)
    eng.map_upload_dense(occ, known)
    o2, k2 = eng.map_download_dense()
    assert np.array_equal(o2, occ) and np.array_equal(k2, known)


def test_map_upload_boxes_equals_octree_leaves(eng, orc):
    """map ingest from (centre, size) leaf lists — pruned leaves of several sizes — reproduces the tree"""
    res = 0.3
    km, dm, occ, known = scenes.synthetic_map(n_vox=64, res=res, seed=8, n_boxes=6, n_spheres=5, region=14.0, unknown_shell=0.2)
    w = orc.world(res, True)
    w.build_dense(km, dm, occ, known, prune=True)
    co, so = w.export_boxes(True)
    cf, sf = w.export_boxes(False)
    assert len(np.unique(sf)) > 1  # pruned free space comes as big leaves
    eng.map_create(res, km, dm, True)
    eng.map_upload_boxes(np.concatenate([co, cf]), np.concatenate([so, sf]),
                         np.concatenate([np.ones(len(so), np.uint8), np.zeros(len(sf), np.uint8)]))
    o2, k2 = eng.map_download_dense()
    assert np.array_equal(o2, occ) and np.array_equal(k2, known)
    w.close()


def test_map_set_box_equals_oracle(eng, orc):
    res = 0.3
    km, dm = [32768 - 40] * 3, [80] * 3
    eng.map_create(res, km, dm, True)
    w = orc.world(res, True)
    ops = [([0, 0, 0], [15, 15, 10], False), ([1.0, -2.0, 0.5], [2.0, 1.0, 3.0], True), ([1.2, -2.2, 0.7], [0.5, 0.5, 0.5], False),
           ([-3.3, 3.3, -1.1], [0.05, 0.05, 0.05], True), ([11.9, 0, 0], [1.0, 1.0, 1.0], True)]
    for pos, box, occd in ops:
        eng.map_set_box(pos, box, occd)
        w.set_box(pos, box, occd)
    o2, k2 = eng.map_download_dense()
    oo, ko = w.export_dense(km, dm)
    assert np.array_equal(o2, oo) and np.array_equal(k2, ko)
    w.close()


def test_treat_unknown_toggle(eng, orc):
    w, (km, dm, occ, known) = scene(eng, orc, 0.3, True, shell=0.3)
    rng = np.random.RandomState(2)
    S, E = segments(rng, 20000, 9.5, 1.5)
    a = eng.check_segments(S, E, BOX)
    eng.map_set_treat_unknown(False)
    b = eng.check_segments(S, E, BOX)
    w2 = orc.world(0.3, False)
    w2.build_dense(km, dm, occ, known)
    assert np.array_equal(a, w.line_status_bbox(S, E, BOX)) and np.array_equal(b, w2.line_status_bbox(S, E, BOX))
    assert (a == 2).any() and not (b == 2).any()
    w.close()
    w2.close()
