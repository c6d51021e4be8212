"""Known-answer tests of the restated OctoMap / octomap_world arithmetic (oracle/octomap_restated.hpp).
The reference pins nothing here, so these are hand-derived from the published
algorithm: key mapping, DDA start/end conventions, the ray lattice, tree == dense equivalence."""
import numpy as np
import pytest

from sspp_b200 import scenes

BOX = (0.5, 0.5, 0.3)


def test_ray_lattice_counts(orc):
    # SURVEY.md §8 a6 probe values: rays per edge for the hard-coded robot box
    want = {0.3: 27, 0.25: 48, 0.2: 48, 0.1: 245, 0.05: 847, 0.5: 18, 1.0: 8, 0.15: 100}
    for res, r in want.items():
        lat = orc.ray_lattice(BOX, res)
        assert int(np.prod([len(a) for a in lat])) == r, res
    z = orc.ray_lattice(BOX, 0.15)[2]
    # accumulated rounding of `double +=`: four offsets, not five, and not the exact decimals
    assert len(z) == 4 and z[0] == -0.15 and z[2] == 0.04999999999999999 and z[3] == 0.14999999999999997
    x = orc.ray_lattice(BOX, 0.25)[0]
    assert len(x) == 4


def test_ray_keys_conventions(orc):
    w = orc.world(0.5, True)
    k = lambda c: int(np.floor(c / 0.5)) + 32768
    # same voxel: empty ray
    n, keys = w.ray_keys([0.1, 0.1, 0.1], [0.4, 0.3, 0.2])
    assert n == 0
    # axis-aligned +x over 4 voxels: start included, end excluded
    n, keys = w.ray_keys([0.25, 0.25, 0.25], [2.25, 0.25, 0.25])
    assert n == 4 and keys[:, 0].tolist() == [k(0.25) + i for i in range(4)] and set(keys[:, 1]) == {k(0.25)}
    # reversed direction visits the mirror set (minus the other end)
    n2, keys2 = w.ray_keys([2.25, 0.25, 0.25], [0.25, 0.25, 0.25])
    assert n2 == 4 and keys2[:, 0].tolist() == [k(2.25) - i for i in range(4)]
    # exact diagonal in x/y: ties go to the later axis (y before x when tMax equal; z wins over both)
    n3, keys3 = w.ray_keys([0.25, 0.25, 0.25], [1.25, 1.25, 0.25])
    assert keys3[0].tolist() == [k(0.25)] * 3
    assert keys3[1].tolist() == [k(0.25), k(0.25) + 1, k(0.25)]  # tMax0 == tMax1 -> not (t0<t1) -> dim 1
    # out of key range: reported, and getLineStatus then treats the ray as empty (free)
    n4, _ = w.ray_keys([0.0, 0.0, 0.0], [1e6, 0.0, 0.0])
    assert n4 == -1
    assert w.line_status([[0, 0, 0]], [[1e6, 0, 0]])[0] == 0
    w.close()


def test_line_status_unknown_and_occupied(orc):
    for treat in (True, False):
        w = orc.world(0.5, treat)
        # nothing observed: unknown blocks only when it counts as occupied
        assert w.line_status([[0.25, 0.25, 0.25]], [[3.25, 0.25, 0.25]])[0] == (2 if treat else 0)
        w.set_box([1.75, 0.25, 0.25], [3.4, 0.4, 0.4], False)  # free corridor
        assert w.line_status([[0.25, 0.25, 0.25]], [[3.25, 0.25, 0.25]])[0] == 0
        w.set_keys([[32768 + 3, 32768, 32768]], True)  # one occupied voxel on the way
        assert w.line_status([[0.25, 0.25, 0.25]], [[3.25, 0.25, 0.25]])[0] == 1
        # the END voxel is never tested: a segment ending inside the obstacle is free
        assert w.line_status([[0.25, 0.25, 0.25]], [[1.75, 0.25, 0.25]])[0] == 0
        # ... while one starting inside it is occupied (edges are not symmetric)
        assert w.line_status([[1.75, 0.25, 0.25]], [[0.25, 0.25, 0.25]])[0] == 1
        w.close()


def test_set_free_box_and_export(orc):
    w = orc.world(0.3, True)
    w.set_box([0, 0, 0], [1.5, 1.5, 0.9], False)
    c, s = w.export_boxes(False)
    # lattice from -0.751 step 0.3 while <= 0.751: 6 points -> keys -3..2; z: -0.451.. -> 4 points -> keys -2..1
    assert len(s) == 6 * 6 * 4 and np.allclose(s, 0.3)
    assert len(w.export_boxes(True)[1]) == 0
    w.set_box([0.3, 0.3, 0.3], [0.4, 0.4, 0.4], True)  # lattice points 0.099, 0.399 per axis -> keys 0..1
    co, so = w.export_boxes(True)
    assert len(so) == 8
    w.prune()
    co2, so2 = w.export_boxes(True)
    # keys (0..1)^3 share one parent: the 2x2x2 block collapses to a single 0.6 m leaf
    assert len(so2) == 1 and so2[0] == 0.6 and np.allclose(co2[0], 0.3)
    w.close()


@pytest.mark.parametrize("prune", [True, False])
def test_tree_equals_dense(orc, prune):
    res = 0.3
    km, dm, occ, known = scenes.synthetic_map(n_vox=48, res=res, seed=5, n_boxes=5, n_spheres=3, region=10.0, unknown_shell=0.25)
    w = orc.world(res, True)
    w.build_dense(km, dm, occ, known, prune=prune)
    o2, k2 = w.export_dense(km, dm)
    assert np.array_equal(o2, occ) and np.array_equal(k2, known)
    # pruning must not change any query
    rng = np.random.RandomState(1)
    S = rng.uniform(-6, 6, (3000, 3))
    E = S + rng.normal(scale=0.8, size=S.shape)
    a = w.line_status_bbox(S, E, BOX)
    w2 = orc.world(res, True)
    w2.build_dense(km, dm, occ, known, prune=not prune)
    assert np.array_equal(a, w2.line_status_bbox(S, E, BOX))
    assert np.array_equal(w.cell_status_bbox(S, BOX), w2.cell_status_bbox(S, BOX))
    w.close()
    w2.close()


def test_cast_ray_conventions(orc):
    w = orc.world(0.5, True)
    w.set_box([2.0, 0.25, 0.25], [4.4, 0.4, 0.4], False)
    w.set_keys([[32768 + 5, 32768, 32768]], True)
    hit, keys = w.cast_rays([[0.25, 0.25, 0.25]], [[1, 0, 0]], True)
    assert hit[0] == 1 and keys[0].tolist() == [32768 + 5, 32768, 32768]
    # range shorter than the obstacle distance: no hit
    hit, _ = w.cast_rays([[0.25, 0.25, 0.25]], [[1, 0, 0]], True, [1.0])
    assert hit[0] == 0
    # origin inside an occupied voxel: immediate hit
    hit, keys = w.cast_rays([[2.75, 0.25, 0.25]], [[1, 0, 0]], True)
    assert hit[0] == 1 and keys[0, 0] == 32768 + 5
    # unknown ahead and ignoreUnknown false: ray gives up
    hit, _ = w.cast_rays([[0.25, 0.25, 0.25]], [[0, 1, 0]], False)
    assert hit[0] == 0
    # zero direction: refused
    hit, _ = w.cast_rays([[0.25, 0.25, 0.25]], [[0, 0, 0]], True)
    assert hit[0] == 0
    w.close()
