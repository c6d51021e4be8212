"""sspp_gpu_check_segments_noalloc: the segment check for callers that cannot allocate device memory.

Some users of this library call it from places where cudaMalloc is off the table: a replanning
callback that runs after another component has taken the rest of the GPU's memory, or a control loop
that cannot tolerate the device-wide stall an allocation causes.  For them the library offers one
entry point whose working memory is a fixed array compiled into the library.  This test calls it
through the C ABI and requires its answers to equal, byte for byte, those of the regular
sspp_gpu_check_segments on the same segments — on two calls in a row with different inputs, so that
whatever the first call left in the shared array cannot leak into the second.  The same batches are
also put to the CPU oracle directly, and — the point of the entry point — the amount of free device
memory is read before and after warmed calls and must not move.  It also checks that a batch beyond
the stated limit is refused rather than truncated.

The file name sorts last on purpose: if this test fails under `-x`, every other test has already run.
"""
import numpy as np
import pytest

from sspp_b200 import scenes
from sspp_b200.engine import SsppError

pytestmark = pytest.mark.gpu
MAX_SEGMENTS = 16384  # SSPP_NOALLOC_MAX_SEGMENTS in include/sspp_gpu.h


def _segments(rng, n, extent, scale):
    S = rng.uniform(-extent, extent, (n, 3))
    E = S + rng.normal(scale=scale, size=S.shape)
    E[: n // 10] = S[: n // 10]  # some zero-length segments
    return S, E


def test_noalloc_segment_check_equals_regular_entry_point(eng, orc):
    import torch
    res = 0.3
    km, dm, occ, known = scenes.synthetic_map(n_vox=64, res=res, seed=3, n_boxes=8, n_spheres=6, region=14.0,
                                              unknown_shell=0.2)
    eng.map_create(res, km, dm, True)
    eng.map_upload_dense(occ, known)
    w = orc.world(res, True)
    w.build_dense(km, dm, occ, known)
    rng = np.random.RandomState(17)
    # first call: a full-size batch with the robot box of the reference (27 rays per segment)
    S1, E1 = _segments(rng, MAX_SEGMENTS, 9.5, 1.2)
    want1 = eng.check_segments(S1, E1, (0.5, 0.5, 0.3))
    got1 = eng.check_segments_noalloc(S1, E1, (0.5, 0.5, 0.3))
    assert np.array_equal(got1, want1)  # integer status bytes: exact equality is the tolerance
    assert len(np.unique(want1)) == 3   # free, occupied and unknown all occur
    # second call right after: fewer segments, another box (another ray table), other coordinates
    S2, E2 = _segments(rng, 5000, 9.5, 2.5)
    want2 = eng.check_segments(S2, E2, (0.9, 0.2, 0.0))
    got2 = eng.check_segments_noalloc(S2, E2, (0.9, 0.2, 0.0))
    assert np.array_equal(got2, want2)
    # and the first batch again gives the first answer again
    assert np.array_equal(eng.check_segments_noalloc(S1, E1, (0.5, 0.5, 0.3)), want1)
    # the oracle's getLineStatusBoundingBox on the same batches, compared with the new entry point directly
    assert np.array_equal(got1, w.line_status_bbox(S1, E1, (0.5, 0.5, 0.3)))
    assert np.array_equal(got2, w.line_status_bbox(S2, E2, (0.9, 0.2, 0.0)))
    w.close()
    # what the entry point is for: once warmed up (address looked up, kernel loaded) a call takes no device
    # memory — the free amount the driver reports is the same before and after, whatever the batch size
    eng.synchronize()
    free0, _ = torch.cuda.mem_get_info(0)
    for S, E, box in ((S1, E1, (0.5, 0.5, 0.3)), (S2, E2, (0.9, 0.2, 0.0)), (S1[:77], E1[:77], (0.5, 0.5, 0.3))):
        eng.check_segments_noalloc(S, E, box)
        free1, _ = torch.cuda.mem_get_info(0)
        assert free1 == free0, "check_segments_noalloc changed the free device memory by %d bytes" % (free0 - free1)


def test_noalloc_rejects_oversized_batch(eng):
    eng.map_create(0.3, [32768 - 32] * 3, [64] * 3, True)
    S = np.zeros((MAX_SEGMENTS + 1, 3))
    with pytest.raises(SsppError) as e:
        eng.check_segments_noalloc(S, S + 1.0)
    assert e.value.code == 4 and "16384" in str(e.value)
    # an empty batch is fine and touches nothing
    assert len(eng.check_segments_noalloc(np.zeros((0, 3)), np.zeros((0, 3)))) == 0
