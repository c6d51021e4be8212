"""Incremental map update: point-cloud insertion (OctomapWorld::insertPointcloudIntoMapImpl restated in the
oracle) — bit planes AND float log-odds of every voxel must equal the oracle's octree after every cloud."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

KM, DM = [32768 - 48] * 3, [96] * 3


def cloud(rng, n, origin, rmin, rmax, dup=0.2):
    d = rng.normal(size=(n, 3))
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    pts = np.asarray(origin) + d * rng.uniform(rmin, rmax, (n, 1))
    k = int(n * dup)
    pts[:k] = pts[n - k:] + rng.uniform(-0.05, 0.05, (k, 3))  # several points per voxel, in both orders
    rng.shuffle(pts)
    return pts.astype(np.float32)


def compare(eng, w):
    o, k = eng.map_download_dense()
    oo, ko = w.export_dense(KM, DM)
    assert np.array_equal(k, ko), "known plane differs"
    assert np.array_equal(o, oo), "occupied plane differs"
    L = eng.map_download_logodds()
    Lo = w.export_logodds(KM, DM)
    assert np.array_equal(np.isnan(L), np.isnan(Lo))
    assert np.array_equal(L[~np.isnan(L)], Lo[~np.isnan(Lo)]), "log-odds differ"
    return o, k


@pytest.mark.parametrize("res,max_range", [(0.3, 5.0), (0.3, -1.0), (0.2, 3.0)])
def test_pointcloud_insertion_matches_oracle(eng, orc, res, max_range):
    eng.map_create(res, KM, DM, False)
    w = orc.world(res, False)
    rng = np.random.RandomState(int(res * 100) + int(max_range))
    origins = [(0.1, 0.2, 0.3), (0.1, 0.2, 0.3), (1.5, -1.0, 0.4), (-2.0, 0.5, -0.7), (0.1, 0.2, 0.3), (0.1, 0.2, 0.3)]
    occ_seen = 0
    for i, org in enumerate(origins):
        pts = cloud(rng, 6000, org, 1.0, 7.0 if max_range > 0 else 5.0)
        eng.map_insert_pointcloud(pts, org, max_range)
        w.insert_pointcloud(pts, org, max_range)
        o, k = compare(eng, w)
        occ_seen = max(occ_seen, int(np.unpackbits(o.view(np.uint8)).sum()))
    assert occ_seen > 0  # repeated hits push cells over the occupancy threshold
    w.close()


def test_pointcloud_mixed_with_direct_edits(eng, orc):
    """setFree / leaf uploads between clouds: the log-odds grid follows the planes (edited cells at the clamping values)"""
    res = 0.3
    eng.map_create(res, KM, DM, True)
    w = orc.world(res, True)
    rng = np.random.RandomState(9)
    eng.map_set_box([0, 0, 0], [4, 4, 3], False)
    w.set_box([0, 0, 0], [4, 4, 3], False)
    for i in range(3):
        org = (0.2 * i, -0.1, 0.3)
        pts = cloud(rng, 4000, org, 0.8, 4.0)
        eng.map_insert_pointcloud(pts, org, 6.0)
        w.insert_pointcloud(pts, org, 6.0)
        compare(eng, w)
        eng.map_set_box([1.0, 1.0, 0.0], [0.9, 0.9, 0.9], i % 2 == 0)
        w.set_box([1.0, 1.0, 0.0], [0.9, 0.9, 0.9], i % 2 == 0)
        compare(eng, w)
    # the updated map answers queries like the oracle's
    S = rng.uniform(-5, 5, (20000, 3))
    E = S + rng.normal(size=S.shape)
    assert np.array_equal(eng.check_segments(S, E), w.line_status_bbox(S, E, (0.5, 0.5, 0.3)))
    w.close()


def test_pointcloud_edge_cases(eng, orc):
    res = 0.3
    eng.map_create(res, KM, DM, False)
    w = orc.world(res, False)
    org = (0.0, 0.0, 0.0)
    pts = np.array([[0.05, 0.05, 0.05],          # same voxel as the sensor: no free cells, the voxel gets a hit
                    [1.0, 0.0, 0.0], [1.01, 0.0, 0.0], [1.02, 0.01, 0.0],   # three points in one voxel
                    [50.0, 0.0, 0.0],            # far beyond the range and the window
                    [0.0, 0.0, 0.0]], dtype=np.float32)   # zero-length ray
    for _ in range(3):
        eng.map_insert_pointcloud(pts, org, 4.0)
        w.insert_pointcloud(pts, org, 4.0)
        compare(eng, w)
    eng.map_insert_pointcloud(np.zeros((0, 3), dtype=np.float32), org, 4.0)
    compare(eng, w)
    w.close()
