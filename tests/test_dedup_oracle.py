"""The oracle's batched insertNode (orc_dedup_poses) pinned against the reference's own SearchSpace
(oracle/_ref: src/searchspace.cpp compiled unmodified), plus hand-made chains with known verdicts."""
import numpy as np

from dedup_cases import near_duplicate_batch


def test_chain_verdicts(orc):
    # a ~ b ~ c but a !~ c: b is refused (matches a), c is stored (its only match, b, was never stored)
    P = np.zeros((4, 7))
    P[:, 6] = 1.0
    P[1, 0] = 0.8e-12
    P[2, 0] = 1.6e-12
    P[3, 0] = 2.0e-12  # matches c only (|2.0-1.6| < 1, |2.0-0.8| > 1): refused, returns c
    keep, dup = orc.dedup_poses(P)
    assert keep.tolist() == [1, 0, 1, 0]
    assert dup.tolist() == [-1, 0, -1, 2]


def test_orientation_counts(orc):
    P = np.zeros((3, 7))
    P[:, 6] = 1.0
    P[1, 5] = 0.5  # same position, other yaw: stored
    keep, dup = orc.dedup_poses(P)
    assert keep.tolist() == [1, 1, 0] and dup.tolist() == [-1, -1, 0]


def test_latest_stored_match_is_returned(orc):
    # poses 0 and 1 are 1.5e-12 apart (both stored); pose 2 sits between them and matches both:
    # nodeExists walks from the head, so it returns pose 1
    P = np.zeros((3, 7))
    P[1, 1] = 1.5e-12
    P[2, 1] = 0.75e-12
    keep, dup = orc.dedup_poses(P)
    assert keep.tolist() == [1, 1, 0] and dup.tolist() == [-1, -1, 1]


def test_pinned_against_reference(orc, ref):
    for seed in range(3):
        P = near_duplicate_batch(3000, seed)
        keep, dup = orc.dedup_poses(P)
        p = ref.planner(1.5)
        ids = p.insert_nodes_ids(P)
        poses_ref, ids_ref = p.nodes()
        p.close()
        exp_keep = (ids == np.arange(len(P))).astype(np.uint8)
        assert np.array_equal(keep, exp_keep)
        assert np.array_equal(dup, np.where(exp_keep == 1, -1, ids))
        # the list the reference ends up with: stored poses, latest first
        assert np.array_equal(ids_ref, np.nonzero(keep)[0][::-1])
        assert np.array_equal(poses_ref, P[keep == 1][::-1])
        assert 0 < keep.sum() < len(P)
