"""Synthetic scenes of SURVEY.md §8(d) (numpy only; shared by bench.py and tests/).

Dense maps are returned in the C-ABI's host format: uint32 words, shape (dimz, dimy, ceil(dimx/32)),
x fastest, bit = (word >> (x & 31)) & 1, over the key window [key_min, key_min + dims).
"""
import numpy as np


def pack_bits(vol):
    """bool array (z, y, x) -> uint32 (z, y, ceil(x/32))."""
    z, y, x = vol.shape
    pad = (-x) % 32
    if pad:
        vol = np.concatenate([vol, np.zeros((z, y, pad), dtype=bool)], axis=2)
    b = np.packbits(vol, axis=2, bitorder="little")
    return np.ascontiguousarray(b).view(np.uint32).reshape(z, y, -1)


def unpack_bits(words, dimx):
    z, y, _ = words.shape
    b = np.unpackbits(np.ascontiguousarray(words).view(np.uint8).reshape(z, y, -1), axis=2, bitorder="little")
    return b[:, :, :dimx].astype(bool)


def synthetic_map(n_vox=512, res=0.3, seed=42, n_boxes=400, n_spheres=200, region=100.0, unknown_shell=0.0):
    """S4/S5 map: `n_vox`^3 voxels centred on the origin (keys 32768-n/2 ... 32768+n/2-1), random
    axis-aligned boxes (edge U(1,8) m) and spheres (radius U(1,5) m) with centres uniform in the
    `region`-metre cube that holds the nodes, everything else known-free; `unknown_shell` > 0 leaves
    that fraction of the half-extent, at the outside, never observed.  A voxel is occupied when its
    centre lies inside a shape."""
    rng = np.random.RandomState(seed)
    half = n_vox // 2
    occ = np.zeros((n_vox, n_vox, n_vox), dtype=bool)  # (z, y, x)
    centres = (np.arange(n_vox) - half + 0.5) * res     # voxel centre coordinate per index

    def idx_range(lo, hi):
        i0 = int(np.searchsorted(centres, lo, side="left"))
        i1 = int(np.searchsorted(centres, hi, side="right"))
        return max(i0, 0), min(i1, n_vox)

    for _ in range(n_boxes):
        c = rng.uniform(-region / 2, region / 2, 3)
        e = rng.uniform(1.0, 8.0, 3)
        x0, x1 = idx_range(c[0] - e[0] / 2, c[0] + e[0] / 2)
        y0, y1 = idx_range(c[1] - e[1] / 2, c[1] + e[1] / 2)
        z0, z1 = idx_range(c[2] - e[2] / 2, c[2] + e[2] / 2)
        occ[z0:z1, y0:y1, x0:x1] = True
    for _ in range(n_spheres):
        c = rng.uniform(-region / 2, region / 2, 3)
        r = rng.uniform(1.0, 5.0)
        x0, x1 = idx_range(c[0] - r, c[0] + r)
        y0, y1 = idx_range(c[1] - r, c[1] + r)
        z0, z1 = idx_range(c[2] - r, c[2] + r)
        if x1 <= x0 or y1 <= y0 or z1 <= z0:
            continue
        dz = (centres[z0:z1] - c[2])[:, None, None]
        dy = (centres[y0:y1] - c[1])[None, :, None]
        dx = (centres[x0:x1] - c[0])[None, None, :]
        occ[z0:z1, y0:y1, x0:x1] |= (dx * dx + dy * dy + dz * dz) <= r * r
    known = np.ones_like(occ)
    if unknown_shell > 0:
        m = int(round(half * unknown_shell))
        if m > 0:
            known[:] = False
            known[m:-m, m:-m, m:-m] = True
            occ &= known
    key_min = np.array([32768 - half] * 3, dtype=np.int32)
    dims = np.array([n_vox] * 3, dtype=np.int32)
    return key_min, dims, pack_bits(occ), pack_bits(known)


def lattice_nodes(n_per_axis=100, spacing=1.0, centre=(0.0, 0.0, 0.0)):
    """Regular lattice in LIST order: the reference generates z outer, x middle, y inner
    (src/pathplanner.cpp:122-126) and inserts at the head (src/searchspace.cpp:98-101), so the list
    is the generation order reversed."""
    n = (n_per_axis, n_per_axis, n_per_axis) if np.isscalar(n_per_axis) else tuple(n_per_axis)
    ax = [(np.arange(n[a]) - (n[a] - 1) / 2.0) * spacing + centre[a] for a in range(3)]
    zz, xx, yy = np.meshgrid(ax[2], ax[0], ax[1], indexing="ij")  # z outer, x, y inner
    xyz = np.stack([xx.ravel(), yy.ravel(), zz.ravel()], axis=1)
    return np.ascontiguousarray(xyz[::-1])


def uniform_nodes(n, extent=100.0, seed=43):
    rng = np.random.RandomState(seed)
    return np.ascontiguousarray(rng.uniform(-extent / 2, extent / 2, (n, 3)))


def surface_points(occ_bits, known_bits, key_min, dims, res, n, seed=45):
    """Model points for visibility: uniform samples inside occupied voxels that have a free
    6-neighbour (the visible shell of the obstacle set), float32."""
    occ = unpack_bits(occ_bits, int(dims[0]))
    free = unpack_bits(known_bits, int(dims[0])) & ~occ
    shell = np.zeros_like(occ)
    shell[1:, :, :] |= free[:-1, :, :]
    shell[:-1, :, :] |= free[1:, :, :]
    shell[:, 1:, :] |= free[:, :-1, :]
    shell[:, :-1, :] |= free[:, 1:, :]
    shell[:, :, 1:] |= free[:, :, :-1]
    shell[:, :, :-1] |= free[:, :, 1:]
    shell &= occ
    zi, yi, xi = np.nonzero(shell)
    if len(zi) == 0:
        return np.zeros((0, 3), dtype=np.float32)
    rng = np.random.RandomState(seed)
    pick = rng.randint(0, len(zi), n)
    jit = rng.uniform(0.05, 0.95, (n, 3))
    k = np.stack([xi[pick] + int(key_min[0]), yi[pick] + int(key_min[1]), zi[pick] + int(key_min[2])], axis=1)
    return ((k - 32768 + jit) * res).astype(np.float32)


def viewpoints_near(points, n, standoff=3.0, seed=46):
    """Sensor poses (x,y,z,qx,qy,qz,qw) looking roughly at random model points from `standoff` m away.
    Camera convention of the visibility spec: X forward, Y up, Z right."""
    rng = np.random.RandomState(seed)
    tgt = points[rng.randint(0, len(points), n)].astype(np.float64)
    d = rng.normal(size=(n, 3))
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    pos = tgt + d * standoff
    fwd = -d
    up0 = np.where(np.abs(fwd[:, 2:3]) < 0.9, np.array([[0.0, 0.0, 1.0]]), np.array([[1.0, 0.0, 0.0]]))
    right = np.cross(fwd, up0)
    right /= np.linalg.norm(right, axis=1, keepdims=True)
    up = np.cross(right, fwd)
    R = np.stack([fwd, up, right], axis=2)  # columns: view, up, right
    q = np.zeros((n, 4))
    for i in range(n):
        m = R[i]
        tr = m[0, 0] + m[1, 1] + m[2, 2]
        if tr > 0:
            s = np.sqrt(tr + 1.0) * 2
            q[i] = [(m[2, 1] - m[1, 2]) / s, (m[0, 2] - m[2, 0]) / s, (m[1, 0] - m[0, 1]) / s, 0.25 * s]
        else:
            j = int(np.argmax([m[0, 0], m[1, 1], m[2, 2]]))
            k, l = (j + 1) % 3, (j + 2) % 3
            s = np.sqrt(m[j, j] - m[k, k] - m[l, l] + 1.0) * 2
            v = np.zeros(4)
            v[j] = 0.25 * s
            v[3] = (m[l, k] - m[k, l]) / s
            v[k] = (m[k, j] + m[j, k]) / s
            v[l] = (m[l, j] + m[j, l]) / s
            q[i] = v
    return np.ascontiguousarray(np.concatenate([pos, q], axis=1))


def map_window(occ_bits, known_bits, key_min, lo_xyz, hi_xyz, res, margin_vox=3):
    """The sub-volume of a packed map that covers the metric box [lo_xyz, hi_xyz] grown by `margin_vox`
    voxels, as (key_min, dims, occ_bits, known_bits) in the same host format — what an oracle octree
    needs to answer queries whose rays stay inside the box (the oracle cannot hold a 512^3 map)."""
    nz, ny, nxw = occ_bits.shape
    nx = nxw * 32
    lo, hi = [], []
    for a, n in zip(range(3), (nx, ny, nz)):
        k0 = int(np.floor(lo_xyz[a] / res)) + 32768 - margin_vox - int(key_min[a])
        k1 = int(np.floor(hi_xyz[a] / res)) + 32768 + margin_vox + 1 - int(key_min[a])
        lo.append(max(k0, 0))
        hi.append(min(k1, n))
    xw0, xw1 = lo[0] // 32, (hi[0] + 31) // 32
    sub_o = unpack_bits(occ_bits[lo[2]:hi[2], lo[1]:hi[1], xw0:xw1], (xw1 - xw0) * 32)[:, :, lo[0] - xw0 * 32: hi[0] - xw0 * 32]
    sub_k = unpack_bits(known_bits[lo[2]:hi[2], lo[1]:hi[1], xw0:xw1], (xw1 - xw0) * 32)[:, :, lo[0] - xw0 * 32: hi[0] - xw0 * 32]
    kmin = np.array([int(key_min[a]) + lo[a] for a in range(3)], dtype=np.int32)
    dims = np.array([hi[a] - lo[a] for a in range(3)], dtype=np.int32)
    return kmin, dims, pack_bits(sub_o), pack_bits(sub_k)


def window_pairs(nodes, core_lo, core_hi, radius):
    """Rows = nodes inside the box [core_lo, core_hi); returns (rows, i, j, same): for every row all ordered
    in-radius pairs (i = row's list index, j = the other node's), with the reference's predicate
    sqrt(dx*dx + dy*dy + dz*dz) <= radius in double, left to right (Dist, include/sspp/utils.h:356-359;
    src/pathplanner.cpp:541), i != j; `same` marks pairs at identical positions (connected without a test)."""
    from scipy.spatial import cKDTree
    core_lo, core_hi = np.asarray(core_lo, float), np.asarray(core_hi, float)
    near = np.nonzero(np.all((nodes >= core_lo - radius * 1.01) & (nodes <= core_hi + radius * 1.01), axis=1))[0]
    sub = nodes[near]
    in_core = np.all((sub >= core_lo) & (sub < core_hi), axis=1)
    rows_local = np.nonzero(in_core)[0]
    tree = cKDTree(sub)
    nb = tree.query_ball_point(sub[rows_local], radius * 1.0001)
    ii = np.repeat(rows_local, [len(x) for x in nb])
    jj = np.concatenate([np.asarray(x, dtype=np.int64) for x in nb]) if len(nb) else np.zeros(0, dtype=np.int64)
    keep = ii != jj
    ii, jj = ii[keep], jj[keep]
    d = sub[jj] - sub[ii]
    dist = np.sqrt(d[:, 0] * d[:, 0] + d[:, 1] * d[:, 1] + d[:, 2] * d[:, 2])
    keep = dist <= radius
    ii, jj = ii[keep], jj[keep]
    same = np.all(sub[ii] == sub[jj], axis=1)
    return near[rows_local], near[ii], near[jj], same


def aircraft_model(res=0.25):
    """S2 stand-in for the inspected aircraft (SURVEY.md §8 d; the reference ships the viewpoints of
    resources/SearchSpaceUAV_1.5m_1to4_NEW.txt but not the model they were sampled around): an ellipsoid fuselage along
    y, two box wings, tail plane and fin, inside x in [-30, 30], y in [-36, 36], z in [0, 15], voxelised at `res`
    over the bounding box of the fixture's way-points (x +-34.5, y +-40.5, z -3 .. 19.5).  Everything else known-free.
    -> key_min, dims, occ_bits, known_bits"""
    lo = np.array([-34.5, -40.5, -3.0])
    hi = np.array([34.5, 40.5, 19.5])
    k0 = np.floor(lo / res).astype(np.int64)
    n = (np.ceil(hi / res).astype(np.int64) - k0)
    cx = (k0[0] + np.arange(n[0]) + 0.5) * res
    cy = (k0[1] + np.arange(n[1]) + 0.5) * res
    cz = (k0[2] + np.arange(n[2]) + 0.5) * res
    Z, Y, X = np.meshgrid(cz, cy, cx, indexing="ij")
    occ = ((X / 3.0) ** 2 + (Y / 34.0) ** 2 + ((Z - 5.0) / 3.0) ** 2) <= 1.0                      # fuselage
    occ |= (np.abs(X) <= 29.0) & (np.abs(Y - 2.0) <= 4.0 - np.abs(X) * 0.08) & (np.abs(Z - 4.0) <= 0.5)  # wings, tapering
    occ |= (np.abs(X) <= 9.0) & (np.abs(Y + 30.0) <= 2.0) & (np.abs(Z - 6.0) <= 0.4)                  # tail plane
    occ |= (np.abs(X) <= 0.5) & (np.abs(Y + 30.0) <= 3.5) & (Z >= 7.0) & (Z <= 14.5)                   # fin
    known = np.ones_like(occ)
    key_min = (k0 + 32768).astype(np.int32)
    return key_min, n.astype(np.int32), pack_bits(occ), pack_bits(known)
