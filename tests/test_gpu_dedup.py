"""Batched SearchSpace::insertNode on the GPU (sspp_gpu_dedup_poses) against the oracle's sequential
restatement (itself pinned to the reference's SearchSpace in tests/test_dedup_oracle.py)."""
import time

import numpy as np
import pytest

from dedup_cases import near_duplicate_batch
from sspp_b200.engine import SsppError

pytestmark = pytest.mark.gpu


def check(eng, orc, P):
    keep, dup, stored = eng.dedup_poses(P)
    keep_o, dup_o = orc.dedup_poses(P)
    assert np.array_equal(keep, keep_o), "keep mask differs"
    assert np.array_equal(dup, dup_o), "returned nodes differ"
    assert stored == int(keep_o.sum())
    keep2, none, stored2 = eng.dedup_poses(P, want_dup_of=False)
    assert none is None and np.array_equal(keep2, keep_o) and stored2 == stored
    return keep, dup


@pytest.mark.parametrize("seed", [0, 1, 2])
def test_near_duplicates_match_sequential_scan(eng, orc, seed):
    keep, _ = check(eng, orc, near_duplicate_batch(20000, seed))
    assert 0 < keep.sum() < 20000


def test_chains_orientation_and_latest_match(eng, orc):
    P = np.zeros((4, 7))
    P[:, 6] = 1.0
    P[1, 0], P[2, 0], P[3, 0] = 0.8e-12, 1.6e-12, 2.0e-12
    keep, dup = check(eng, orc, P)
    assert keep.tolist() == [1, 0, 1, 0] and dup.tolist() == [-1, 0, -1, 2]
    Q = np.zeros((3, 7))
    Q[1, 1], Q[2, 1] = 1.5e-12, 0.75e-12
    keep, dup = check(eng, orc, Q)
    assert keep.tolist() == [1, 1, 0] and dup.tolist() == [-1, -1, 1]
    R = np.zeros((3, 7))
    R[:, 6] = 1.0
    R[1, 5] = 0.5
    keep, dup = check(eng, orc, R)
    assert keep.tolist() == [1, 1, 0]


def test_long_chain_needs_many_rounds(eng, orc):
    """300 poses 0.8e-12 apart: every verdict depends on the one before it (stored, refused, stored, ...)"""
    P = np.zeros((300, 7))
    P[:, 2] = np.arange(300) * 0.8e-12
    keep, _ = check(eng, orc, P)
    assert keep.tolist() == [1, 0] * 150
    rng = np.random.RandomState(4)
    check(eng, orc, P[rng.permutation(300)])


def test_grid_with_yaw_samples_inserted_twice(eng, orc):
    xs, ys, zs = np.meshgrid(np.arange(20) * 0.5, np.arange(20) * 0.5, np.arange(5) * 0.5, indexing="ij")
    pos = np.stack([xs.ravel(), ys.ravel(), zs.ravel()], axis=1)
    yaws = np.arange(8) * (45.0 * np.pi / 180.0)
    P = np.zeros((len(pos) * 8, 7))
    P[:, :3] = np.repeat(pos, 8, axis=0)
    P[:, 5] = np.tile(np.sin(yaws / 2), len(pos))
    P[:, 6] = np.tile(np.cos(yaws / 2), len(pos))
    n = len(P)
    keep, dup = check(eng, orc, np.concatenate([P, P]))
    assert keep[:n].all() and not keep[n:].any()
    assert np.array_equal(dup[n:], np.arange(n))


def test_degenerate_extents(eng, orc):
    rng = np.random.RandomState(2)
    flat = np.zeros((500, 7))  # a 2-D grid: z extent 0
    flat[:, 0] = rng.randint(0, 10, 500)
    flat[:, 1] = rng.randint(0, 10, 500)
    flat[:, 6] = 1.0
    check(eng, orc, flat)
    spot = np.zeros((400, 7))  # one position, many orientations: a single cell holds everything
    spot[:, 3:] = rng.randint(0, 6, (400, 4)) * 0.1
    check(eng, orc, spot)
    check(eng, orc, np.array([[1.0, 2, 3, 0, 0, 0, 1]]))
    keep, dup, stored = eng.dedup_poses(np.zeros((0, 7)))
    assert len(keep) == 0 and stored == 0


def test_errors(eng):
    P = np.zeros((4, 7))
    with pytest.raises(SsppError):
        eng.dedup_poses(P, tolerance=0.0)
    P[2, 1] = np.nan
    with pytest.raises(SsppError):
        eng.dedup_poses(P)


def test_full_size_lattice_inserted_twice(eng):
    """10^6 lattice poses, then the same poses again in another order: the first million is stored, every
    pose of the second million is refused and returns its first-pass twin (size-independent property)."""
    from sspp_b200 import scenes
    xyz = scenes.lattice_nodes(100, 1.0)
    n = len(xyz)
    P = np.zeros((2 * n, 7))
    perm = np.random.RandomState(0).permutation(n)
    P[:n, :3] = xyz
    P[n:, :3] = xyz[perm]
    P[:, 6] = 1.0
    eng.dedup_poses(P[:1000])
    t0 = time.perf_counter()
    keep, dup, stored = eng.dedup_poses(P)
    dt = time.perf_counter() - t0
    assert stored == n and keep[:n].all() and not keep[n:].any()
    assert np.array_equal(dup[n:], perm) and (dup[:n] == -1).all()
    print("dedup of %d poses: %.1f ms (host buffers in and out)" % (2 * n, dt * 1e3))
