"""The engine's own voxel-walk header (sspp_b200/csrc/dda_core.cuh) compiled for the CPU and checked
bit-for-bit against the oracle: statuses AND the number of voxel lookups, random and adversarial
segments, several resolutions.  (The same header is what the kernels run; the GPU tests repeat this
through the C ABI.)"""
import numpy as np
import pytest

import oracle_lib
from sspp_b200 import scenes

BOX = (0.5, 0.5, 0.3)


@pytest.fixture(scope="module")
def probe(built):
    return oracle_lib.Probe()


def segments(rng, n, extent, scale):
    S = rng.uniform(-extent, extent, (n, 3))
    E = S + rng.normal(scale=scale, size=S.shape)
    # adversarial share: snap coordinates to voxel boundaries / centres, axis-aligned and zero-length segments
    k = n // 5
    S[:k] = np.round(S[:k] / 0.15) * 0.15
    E[:k] = np.round(E[:k] / 0.15) * 0.15
    E[k:k + k // 2, 1:] = S[k:k + k // 2, 1:]
    E[k + k // 2:2 * k] = S[k + k // 2:2 * k]
    return S, E


@pytest.mark.parametrize("res,treat", [(0.3, True), (0.3, False), (0.25, True), (0.15, True), (0.1, False), (0.05, True)])
def test_probe_line_status_bbox_equals_oracle(orc, probe, res, treat):
    nv = 64
    km, dm, occ, known = scenes.synthetic_map(n_vox=nv, res=res, seed=11, n_boxes=8, n_spheres=6,
                                              region=nv * res * 0.8, unknown_shell=0.2)
    w = orc.world(res, treat)
    w.build_dense(km, dm, occ, known)
    rng = np.random.RandomState(int(res * 1000) + treat)
    n = 6000 if res >= 0.1 else 1500
    S, E = segments(rng, n, nv * res * 0.55, 1.2)
    so, steps, rays = w.line_status_bbox(S, E, BOX, True)
    sp, steps_p = probe.line_status_bbox(res, km, dm, occ, known, treat, S, E, BOX)
    assert np.array_equal(so, sp)
    assert steps == steps_p
    assert len(np.unique(so)) >= 2  # the scene exercises more than one outcome
    w.close()


@pytest.mark.parametrize("res", [0.3, 0.25, 0.1])
def test_probe_tie_rule_on_exact_diagonals(orc, probe, res):
    """Segments from voxel centres (and from points a few float ulps off them) along exact diagonals: tMax
    ties at every step, so the walk is decided by the tie rule alone; and segments ending exactly on
    voxel faces.  Status and lookup count must match the oracle ray by ray."""
    nv = 64
    km, dm, occ, known = scenes.synthetic_map(n_vox=nv, res=res, seed=21, n_boxes=10, n_spheres=8,
                                              region=nv * res * 0.8, unknown_shell=0.15)
    w = orc.world(res, True)
    w.build_dense(km, dm, occ, known)
    rng = np.random.RandomState(7)
    n = 3000
    k0 = rng.randint(-20, 20, (n, 3))
    S = (k0 + 0.5) * res
    dirs = np.array([[1, 1, 0], [1, 0, 1], [0, 1, 1], [1, 1, 1], [-1, 1, 0], [1, -1, 1], [-1, -1, -1], [1, -1, 0], [2, 1, 0], [1, 2, 2]])
    d = dirs[rng.randint(len(dirs), size=n)]
    steps = rng.randint(1, 14, (n, 1))
    E = S + d * steps * res
    third = n // 3
    # second third: nudge the start by a few float32 ulps; last third: end exactly on a voxel face
    S32 = S.astype(np.float32)
    S[third:2 * third] = np.nextafter(S32[third:2 * third], np.float32(1e9) * np.sign(rng.normal(size=(third, 3))).astype(np.float32)).astype(np.float64)
    E[2 * third:] = np.round(E[2 * third:] / res) * res
    zero = (0.0, 0.0, 0.0)  # a box of size 0: the lattice is the single centre ray
    so, st_o, rays_o = w.line_status_bbox(S, E, zero, True)
    sp, st_p = probe.line_status_bbox(res, km, dm, occ, known, True, S, E, zero)
    assert rays_o == n or rays_o <= n
    assert np.array_equal(so, sp) and st_o == st_p
    assert len(np.unique(so)) >= 2
    so2, st2, _ = w.line_status_bbox(S, E, BOX, True)
    sp2, st2p = probe.line_status_bbox(res, km, dm, occ, known, True, S, E, BOX)
    assert np.array_equal(so2, sp2) and st2 == st2p
    w.close()


def test_probe_lattice_equals_oracle(orc, probe):
    for res in (0.3, 0.25, 0.2, 0.15, 0.1, 0.05, 0.5, 1.0, 0.07):
        for size in BOX + (0.0, 1.0):
            a, n = probe.ray_lattice(size, res)
            b = orc.ray_lattice((size, size, size), res)[0]
            assert n == len(b) and np.array_equal(a, b)


def test_probe_cast_ray_range_boundary(orc, probe):
    """max_range set right at the distance of the voxel a ray would hit (a hair below, exactly, a hair
    above, and a voxel either side): the range test decides, and the engine's version of it — which
    skips the bookkeeping while the range provably cannot be exceeded — must decide as the oracle does."""
    res = 0.25
    km, dm, occ, known = scenes.synthetic_map(n_vox=64, res=res, seed=2, n_boxes=8, n_spheres=6, region=12.0)
    w = orc.world(res, True)
    w.build_dense(km, dm, occ, known)
    rng = np.random.RandomState(5)
    n = 6000
    O = rng.uniform(-6, 6, (n, 3)).astype(np.float32)
    D = rng.normal(size=(n, 3)).astype(np.float32)
    hit, key = w.cast_rays(O, D, True, np.full(n, -1.0))
    sel = hit.astype(bool)
    assert sel.sum() > 1000
    O, D, key = O[sel], D[sel], key[sel]
    centre = ((key.astype(np.float64) - 32768 + 0.5) * res).astype(np.float32)
    dist = np.linalg.norm(centre.astype(np.float64) - O.astype(np.float64), axis=1)
    total = 0
    for f in (1 - 1e-6, 1.0, 1 + 1e-6, 0.93, 1.07):
        mr = dist * f
        mr[mr <= 0] = 1e-3
        ho, ko = w.cast_rays(O, D, True, mr)
        hp, kp = probe.cast_rays(res, km, dm, occ, known, O, D, True, mr)
        assert np.array_equal(ho, hp) and np.array_equal(ko, kp)
        total += int(ho.sum())
    assert 0 < total < 5 * len(O)
    w.close()


@pytest.mark.parametrize("ignore_unknown", [True, False])
def test_probe_cast_ray_equals_oracle(orc, probe, ignore_unknown):
    res = 0.25
    km, dm, occ, known = scenes.synthetic_map(n_vox=64, res=res, seed=2, n_boxes=8, n_spheres=6, region=12.0, unknown_shell=0.15)
    w = orc.world(res, True)
    w.build_dense(km, dm, occ, known)
    rng = np.random.RandomState(3)
    n = 8000
    O = rng.uniform(-6, 6, (n, 3)).astype(np.float32)
    D = rng.normal(size=(n, 3)).astype(np.float32)
    D[:200, 0] = 0
    D[200:300] = 0
    D[300:400, 1:] = 0
    mr = rng.uniform(0.5, 8.0, n)
    mr[::3] = -1.0
    ho, ko = w.cast_rays(O, D, ignore_unknown, mr)
    hp, kp = probe.cast_rays(res, km, dm, occ, known, O, D, ignore_unknown, mr)
    assert np.array_equal(ho, hp) and np.array_equal(ko, kp)
    assert 0 < ho.sum() < n
    w.close()
