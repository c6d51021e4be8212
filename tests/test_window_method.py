"""The window method of tests/test_gpu_bench_scene_parity.py (oracle octree over a cut-out of the map, the
in-radius pairs of the rows inside it found with a k-d tree) against the oracle's own whole-scene O(N^2)
connectNodes on a scene small enough for it: the sampled rows must come out identical.  CPU only — this
pins the test's own machinery (scenes.map_window / scenes.window_pairs) before it judges the GPU."""
import numpy as np

from sspp_b200 import scenes

BOX = (0.5, 0.5, 0.3)


def test_window_rows_equal_whole_scene_oracle(orc):
    res = 0.3
    km, dm, occ, known = scenes.synthetic_map(n_vox=96, res=res, seed=5, n_boxes=14, n_spheres=10, region=22.0, unknown_shell=0.1)
    nodes = scenes.uniform_nodes(5000, extent=24.0, seed=3)
    nodes[100:102] = [[-1.0, 0.5, -2.0], [-1.0, 0.5, -2.0]]  # a same-position pair inside the first window
    w = orc.world(res, True)
    w.build_dense(km, dm, occ, known)
    rp, col, _ = orc.connect(w, nodes, 1.5, BOX)
    w.close()
    n, n_same = len(nodes), 0
    for lo in ([-4.0, -3.0, -5.0], [3.5, 2.0, -1.0]):
        lo = np.array(lo)
        hi = lo + 7.0
        rows, i, j, same = scenes.window_pairs(nodes, lo, hi, 1.5)
        kmin, dims, so, sk = scenes.map_window(occ, known, km, lo - 2.0, hi + 2.0, res)
        w2 = orc.world(res, True)
        w2.build_dense(kmin, dims, so, sk)
        n_same += int(same.sum())
        want = same | (w2.line_status_bbox_mt(nodes[i], nodes[j], BOX, 4) == 0)
        w2.close()
        order = np.lexsort((j, i))
        i, j, want = i[order], j[order], want[order]
        got = np.concatenate([np.full(rp[r + 1] - rp[r], r, dtype=np.int64) * n + col[rp[r]:rp[r + 1]] for r in np.sort(rows)])
        assert len(rows) > 50 and np.array_equal(got, (i.astype(np.int64) * n + j)[want])
    assert n_same == 2
