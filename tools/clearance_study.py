"""Numpy + scipy study, host only, not part of the test-suite (S4a): how many of the pairs that reach the pair stage of k_edges a Chebyshev clearance field would accept,
against what the key-box test accepts.  Key boxes as in k_edges (extreme rays, one voxel of overshoot)."""
import sys, time
import numpy as np
sys.path.insert(0, '/root/repo')
from sspp_b200 import scenes
from scipy import ndimage
res = 0.3
km, dm, occ, known = scenes.synthetic_map(n_vox=512, res=res, seed=42)
print(km, dm, occ.shape, occ.dtype)
t0 = time.time()
blocked = scenes.unpack_bits(occ, 512) | ~scenes.unpack_bits(known, 512)   # (z, y, x)
print("blocked fraction %.4f" % blocked.mean(), time.time() - t0)
cdt = ndimage.distance_transform_cdt(~blocked, metric='chessboard').astype(np.int16)
print("cdt done", time.time() - t0, cdt.max())
I = np.zeros((513, 513, 513), dtype=np.int32)
I[1:, 1:, 1:] = blocked.cumsum(0, dtype=np.int32).cumsum(1, dtype=np.int32).cumsum(2, dtype=np.int32)
print("integral done", time.time() - t0)
def box_count(lo, hi):  # lo, hi: (n,3) voxel indices (x,y,z), inclusive, clipped to the window
    lo = np.clip(lo, 0, 511); hi = np.clip(hi, 0, 511)
    x0, y0, z0 = lo[:, 0], lo[:, 1], lo[:, 2]; x1, y1, z1 = hi[:, 0] + 1, hi[:, 1] + 1, hi[:, 2] + 1
    return (I[z1, y1, x1] - I[z0, y1, x1] - I[z1, y0, x1] - I[z1, y1, x0] + I[z0, y0, x1] + I[z0, y1, x0] + I[z1, y0, x0] - I[z0, y0, x0])
nodes = scenes.lattice_nodes(100, 1.0)
half = np.array([0.25, 0.25, 0.15])
key = lambda p: np.floor(p / res).astype(np.int64) + 32768 - km[None, :]
# rows whose reach is free (pyramid stage, approximated at voxel instead of 4^3 granularity)
reach = np.array([1.5, 1.5, 1.5]) + half
rl, rh = key(nodes - reach) - 2, key(nodes + reach) + 2
row_free = box_count(rl, rh) == 0
print("rows with a free reach: %d of %d (engine: 529306 at 4^3 granularity)" % (row_free.sum(), len(nodes)))
offs = [(dx, dy, dz) for dx in (-1, 0, 1) for dy in (-1, 0, 1) for dz in (-1, 0, 1) if 0 < dx * dx + dy * dy + dz * dz <= 2]
tot = box_free = clear_ok = 0
work = ~row_free
P = nodes[work]
for d in offs:
    Q = P + np.array(d, dtype=np.float64)
    inside = np.all(np.abs(Q) <= 49.5 + 1e-9, axis=1)
    a, b = P[inside], Q[inside]
    lo = key(np.minimum(a, b) - half); hi = key(np.maximum(a, b) + half)
    moves = np.array(d) != 0
    lo = lo - moves[None, :]; hi = hi + moves[None, :]   # one voxel of overshoot on the axes the pair moves along
    n = len(a); tot += n
    box_free += int((box_count(lo, hi) == 0).sum())
    c = (lo + hi) // 2
    c = np.clip(c, 0, 511)
    rad = ((hi - lo + 1) // 2).max(axis=1)      # Chebyshev radius that covers the box from its centre voxel
    clear_ok += int((cdt[c[:, 2], c[:, 1], c[:, 0]] > rad).sum())
print("pairs of work rows %d; key box free %d (%.1f %%); clearance(centre) > half-extent %d (%.1f %%)" % (tot, box_free, 100 * box_free / tot, clear_ok, 100 * clear_ok / tot))

# how many rows a pyramid level of g x g x g blocks clears (reach grown to whole blocks): g = 4 is what k_count reads today
for g in (4, 2):
    cb = blocked.reshape(512 // g, g, 512 // g, g, 512 // g, g).any(axis=(1, 3, 5))
    n = 512 // g
    J = np.zeros((n + 1,) * 3, dtype=np.int32)
    J[1:, 1:, 1:] = cb.cumsum(0, dtype=np.int32).cumsum(1, dtype=np.int32).cumsum(2, dtype=np.int32)
    lo = np.clip(rl, 0, 511) // g
    hi = np.clip(rh, 0, 511) // g
    x0, y0, z0 = lo[:, 0], lo[:, 1], lo[:, 2]
    x1, y1, z1 = hi[:, 0] + 1, hi[:, 1] + 1, hi[:, 2] + 1
    cnt = J[z1, y1, x1] - J[z0, y1, x1] - J[z1, y0, x1] - J[z1, y1, x0] + J[z0, y0, x1] + J[z0, y1, x0] + J[z1, y0, x0] - J[z0, y0, x0]
    print("rows cleared by a %d^3 pyramid level: %d" % (g, int((cnt == 0).sum())))
