"""Parity of the CUDA connection pass with the oracle, through the C ABI (bit-exact CSR)."""
import os

import numpy as np
import pytest

from sspp_b200 import scenes
from sspp_b200.engine import SsppError

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
BOX = (0.5, 0.5, 0.3)


def make_map(eng, orc, res, treat, n_vox=64, seed=3, shell=0.0, region=None, n_boxes=8, n_spheres=6):
    region = region if region is not None else n_vox * res * 0.75
    km, dm, occ, known = scenes.synthetic_map(n_vox=n_vox, res=res, seed=seed, n_boxes=n_boxes, n_spheres=n_spheres,
                                              region=region, unknown_shell=shell)
    eng.map_create(res, km, dm, treat)
    eng.map_upload_dense(occ, known)
    w = orc.world(res, treat)
    w.build_dense(km, dm, occ, known)
    return w


def assert_same_graph(eng, orc, w, nodes, radius, box=BOX, mode=0, **kw):
    eng.set_nodes(nodes)
    st = eng.connect(radius=radius, box=box, use_map=w is not None, position_mode=mode, **kw)
    rp, col = eng.connect_fetch(st)
    rp_o, col_o, cnt = orc.connect(w, nodes, radius, box, mode)
    assert np.array_equal(rp, rp_o), "row_ptr differs"
    assert np.array_equal(col, col_o), "col_idx differs"
    assert st.edges == len(col_o) and st.tested_pairs == cnt[0] and st.in_radius_pairs == cnt[0] + cnt[1]
    return st, cnt


def test_known_answer_no_map(eng, orc):
    """src/distance_heuristic_test.cpp scenario (golden from the reference's own sources): 1452 nodes, 80 580 edges"""
    g = np.load(os.path.join(GOLD, "dht_known_answer.npz"))
    eng.set_nodes(g["poses"][:, :3])
    st = eng.connect(radius=1.5, use_map=False)
    rp, col = eng.connect_fetch(st)
    assert st.edges == 80580 and st.tested_pairs == 76224
    assert np.array_equal(rp, g["row_ptr"]) and np.array_equal(col, g["col_idx"])


def test_reactive_sphere_golden(eng, orc):
    """reactive scene through map_set_box (setFree) + leaf upload; golden adjacency from oracle/_ref"""
    g = np.load(os.path.join(GOLD, "reactive_sphere.npz"))
    eng.map_create(0.3, [32768 - 64] * 3, [128] * 3, False)
    eng.map_set_box([0, 0, 0], [15, 15, 10], False)
    k = g["sphere_keys"].astype(np.float64)
    eng.map_upload_boxes((k - 32768 + 0.5) * 0.3, np.full(len(k), 0.3), np.ones(len(k), dtype=np.uint8))
    eng.set_nodes(g["poses"][:, :3])
    st = eng.connect(radius=1.5, box=BOX, use_map=True)
    rp, col = eng.connect_fetch(st)
    assert st.tested_pairs == 840 and st.edges == 777
    assert np.array_equal(rp, g["row_ptr"]) and np.array_equal(col, g["col_idx"])


def test_uav_fixture_golden(eng):
    g = np.load(os.path.join(GOLD, "uav_fixture_R25.npz"))
    eng.set_nodes(g["poses"][:, :3])
    st = eng.connect(radius=2.5, use_map=False)
    rp, col = eng.connect_fetch(st)
    assert st.edges == int(g["connections"]) and np.array_equal(rp, g["row_ptr"])
    assert np.array_equal(col[:4096], g["col_head"])


@pytest.mark.parametrize("res,treat,shell", [(0.3, True, 0.0), (0.3, False, 0.2), (0.3, True, 0.2), (0.15, True, 0.0),
                                             (0.5, False, 0.0), (1.0, True, 0.0), (0.1, True, 0.1)])
def test_lattice_nodes_with_map(eng, orc, res, treat, shell):
    n_vox = 64 if res >= 0.3 else 96
    w = make_map(eng, orc, res, treat, n_vox=n_vox, shell=shell)
    ext = n_vox * res
    npa = max(4, min(14, int(ext * 0.8)))
    spacing = min(1.0, ext * 0.8 / npa)
    nodes = scenes.lattice_nodes(npa, spacing)
    st, cnt = assert_same_graph(eng, orc, w, nodes, 1.5 * spacing)
    assert 0 < st.edges < st.in_radius_pairs  # some edges collide, some do not
    w.close()


@pytest.mark.parametrize("pyramid", [True, False])
def test_random_nodes_with_map(eng, orc, pyramid):
    w = make_map(eng, orc, 0.3, True, n_vox=96, seed=9)
    nodes = scenes.uniform_nodes(3000, extent=22.0, seed=5)
    st, _ = assert_same_graph(eng, orc, w, nodes, 1.7, use_pyramid=pyramid)
    assert (st.pyramid_rows > 0) == pyramid
    w.close()


@pytest.mark.parametrize("slot", ["0", "4", "32", "64"])
def test_pair_storage_modes_give_the_same_graph(eng, orc, slot, monkeypatch):
    """Fixed slots per row (pairs written by the counting pass), scanned offsets (second neighbour pass),
    and the switch from one to the other when a row overflows its slot (slot 4: almost every row does;
    slot 32: the dense clump does) all build the same CSR."""
    monkeypatch.setenv("SSPP_CONNECT_SLOT", slot)
    w = make_map(eng, orc, 0.3, True, n_vox=96, seed=9)
    rng = np.random.RandomState(8)
    nodes = np.concatenate([scenes.uniform_nodes(2500, extent=22.0, seed=5), rng.uniform(-1.0, 1.0, (60, 3))])
    st, cnt = assert_same_graph(eng, orc, w, nodes, 1.7)
    assert st.in_radius_pairs > 10000
    st2, _ = assert_same_graph(eng, orc, w, nodes, 1.7, rows=None)  # again on warm buffers
    assert st2.edges == st.edges
    eng.set_nodes(nodes)
    full = eng.connect_fetch(eng.connect(radius=1.7, box=BOX))
    a = eng.connect_fetch(eng.connect(radius=1.7, box=BOX, rows=(0, 1000)))
    b = eng.connect_fetch(eng.connect(radius=1.7, box=BOX, rows=(1000, len(nodes))))
    assert np.array_equal(np.concatenate([a[0][:-1], b[0] + a[0][-1]]), full[0])
    assert np.array_equal(np.concatenate([a[1], b[1]]), full[1])
    w.close()


def test_duplicates_and_position_modes(eng, orc):
    """same-position nodes (other yaw) are connected without a test; mode 1 uses the 1e-12 tolerance"""
    w = make_map(eng, orc, 0.3, True, n_vox=64, seed=4)
    base = scenes.lattice_nodes(6, 1.0)
    jit = base[:50] + 1e-13   # equal under isPositionEqual, different under exact !=
    nodes = np.concatenate([base, base[::3], jit])
    for mode in (0, 1):
        st, cnt = assert_same_graph(eng, orc, w, nodes, 1.5, mode=mode)
        assert cnt[1] > 0
    w.close()


def test_large_box_many_rays(eng, orc):
    """r > 32 rays per pair (chunked lattice) and a tile too large to stage (direct lookups)"""
    w = make_map(eng, orc, 0.1, True, n_vox=128, seed=6, region=9.0)
    nodes = scenes.uniform_nodes(400, extent=8.0, seed=8)
    assert_same_graph(eng, orc, w, nodes, 1.2, box=(0.5, 0.5, 0.3))    # 245 rays
    assert_same_graph(eng, orc, w, nodes[:150], 4.0, box=(0.3, 0.2, 0.0))  # reach too wide for the smem tile
    w.close()


def test_edge_cases(eng, orc):
    w = make_map(eng, orc, 0.3, True)
    # empty node set
    eng.set_nodes(np.zeros((0, 3)))
    st = eng.connect(radius=1.5)
    rp, col = eng.connect_fetch(st)
    assert st.edges == 0 and rp.tolist() == [0] and len(col) == 0
    # a single node, and two coincident nodes
    assert_same_graph(eng, orc, w, np.array([[0.1, 0.2, 0.3]]), 1.5)
    assert_same_graph(eng, orc, w, np.array([[0.1, 0.2, 0.3], [0.1, 0.2, 0.3]]), 1.5)
    # radius 0 and negative radius
    nodes = scenes.lattice_nodes(4, 1.0)
    assert_same_graph(eng, orc, w, nodes, 0.0)
    assert_same_graph(eng, orc, w, nodes, -1.0)
    # radius exactly on the lattice spacing: `<=` keeps the axis neighbours
    st, _ = assert_same_graph(eng, orc, w, nodes, 1.0)
    # nodes outside the mapped window (unknown there)
    far = nodes + 500.0
    assert_same_graph(eng, orc, w, far, 1.5)
    # zero-size robot box: a single ray
    assert_same_graph(eng, orc, w, nodes, 1.5, box=(0.0, 0.0, 0.0))
    w.close()


def test_errors(eng):
    with pytest.raises(SsppError) as e:
        eng.connect(radius=1.5, use_map=True)          # no map yet
    assert e.value.code == 3
    eng.set_nodes(np.array([[0.0, np.nan, 0.0], [1.0, 0.0, 0.0]]))
    with pytest.raises(SsppError) as e:
        eng.connect(radius=1.5, use_map=False)
    assert e.value.code == 1 and "finite" in str(e.value)
    eng.set_nodes(np.zeros((4, 3)))
    with pytest.raises(SsppError):
        eng.connect(radius=1.5, use_map=False, rows=(2, 9))
    with pytest.raises(SsppError):
        eng.map_create(0.3, [65000] * 3, [1024] * 3, True)   # window leaves the 16-bit key space


def test_row_sharding_concatenates_to_full_csr(eng, orc):
    """multi-GPU partition: CSR blocks of contiguous row ranges concatenate to the 1-GPU CSR"""
    w = make_map(eng, orc, 0.3, True, n_vox=96, seed=9)
    nodes = scenes.uniform_nodes(2500, extent=20.0, seed=12)
    eng.set_nodes(nodes)
    st = eng.connect(radius=1.7, box=BOX)
    rp, col = eng.connect_fetch(st)
    parts_rp, parts_col = [np.zeros(1, dtype=np.int64)], []
    bounds = [0, 700, 701, 1800, 2500]
    for lo, hi in zip(bounds[:-1], bounds[1:]):
        s = eng.connect(radius=1.7, box=BOX, rows=(lo, hi))
        r, c = eng.connect_fetch(s)
        assert s.rows == hi - lo
        parts_rp.append(r[1:] + parts_rp[-1][-1])
        parts_col.append(c)
    assert np.array_equal(np.concatenate(parts_rp), rp) and np.array_equal(np.concatenate(parts_col), col)
    w.close()


def test_connect_nodes_host_and_determinism(eng, orc):
    w = make_map(eng, orc, 0.3, True, n_vox=96, seed=9)
    nodes = scenes.uniform_nodes(2000, extent=20.0, seed=2)
    rp_o, col_o, _ = orc.connect(w, nodes, 1.6, BOX)
    rp = np.zeros(len(nodes) + 1, dtype=np.int64)
    col = np.zeros(len(col_o) + 16, dtype=np.int32)
    for _ in range(3):
        st = eng.connect_nodes_host(nodes, rp, col, radius=1.6, box=BOX)
        assert np.array_equal(rp, rp_o) and np.array_equal(col[:st.edges], col_o)
    small = np.zeros(4, dtype=np.int32)
    with pytest.raises(SsppError) as e:
        eng.connect_nodes_host(nodes, rp, small, radius=1.6, box=BOX)
    assert e.value.code == 4
    w.close()


@pytest.mark.parametrize("chunks,slot", [(None, None), ("3", None), ("8", None), ("4", "0"), ("4", "4")])
def test_connect_nodes_host_chunked_delivery(eng, chunks, slot, monkeypatch):
    """Above 65 536 rows the host-buffer call runs the edge test in row chunks and sends each finished
    chunk home while the next one is computed: same CSR as connect + fetch, whatever the chunking and
    the pair storage (slot 4: the slots overflow and the chunked phase is repeated with scanned offsets)."""
    if chunks:
        monkeypatch.setenv("SSPP_HOST_CHUNKS", chunks)
    if slot:
        monkeypatch.setenv("SSPP_CONNECT_SLOT", slot)
    km, dm, occ, known = scenes.synthetic_map(n_vox=160, res=0.3, seed=12, n_boxes=30, n_spheres=20, region=40.0)
    eng.map_create(0.3, km, dm, True)
    eng.map_upload_dense(occ, known)
    nodes = scenes.lattice_nodes(44, 1.0)  # 85 184 rows
    if chunks == "3":  # an unordered node set: rows of a chunk are scattered over the cells
        nodes = np.ascontiguousarray(nodes[np.random.RandomState(3).permutation(len(nodes))] +
                                     np.random.RandomState(4).uniform(-0.2, 0.2, nodes.shape))
    eng.set_nodes(nodes)
    st0 = eng.connect(radius=1.5, box=BOX)
    rp0, col0 = eng.connect_fetch(st0)
    assert st0.pyramid_rows > 0 and st0.lattice_pairs > 0 and st0.edges < st0.in_radius_pairs
    rp = np.full(len(nodes) + 1, -7, dtype=np.int64)
    col = np.full(st0.edges + 5, -7, dtype=np.int32)
    for _ in range(2):
        st = eng.connect_nodes_host(nodes, rp, col, radius=1.5, box=BOX)
        assert st.edges == st0.edges and st.tested_pairs == st0.tested_pairs
        assert np.array_equal(rp, rp0) and np.array_equal(col[: st.edges], col0)
        assert (col[st.edges:] == -7).all()
    # without a map every in-radius pair is an edge; the chunks still line up
    st1 = eng.connect(radius=1.5, use_map=False)
    rp1, col1 = eng.connect_fetch(st1)
    colb = np.zeros(st1.edges, dtype=np.int32)
    st2 = eng.connect_nodes_host(nodes, rp, colb, radius=1.5, use_map=False)
    assert st2.edges == st1.edges and np.array_equal(rp, rp1) and np.array_equal(colb, col1)
    # a column buffer that is too small is refused with the size needed
    with pytest.raises(SsppError) as e:
        eng.connect_nodes_host(nodes, rp, np.zeros(st0.edges // 2, dtype=np.int32), radius=1.5, box=BOX)
    assert e.value.code == 4
    rp3, col3 = eng.connect_fetch(st0)  # the result stays resident: a caller can size its buffer and fetch
    assert np.array_equal(rp3, rp0) and np.array_equal(col3, col0)


def test_full_size_properties(eng):
    """BASELINE config 4 shape (512^3 voxels, 10^6 lattice nodes): size-independent properties.
    (a) with an empty (all-free) map the graph equals the map-free graph; (b) the pyramid skip does
    not change the graph; (c) interior degree is 18 without obstacles; (d) rows are ascending."""
    res = 0.3
    km, dm, occ, known = scenes.synthetic_map(n_vox=512, res=res, seed=42)
    nodes = scenes.lattice_nodes(100, 1.0)
    eng.set_nodes(nodes)
    s0 = eng.connect(radius=1.5, use_map=False)
    rp0, col0 = eng.connect_fetch(s0)
    assert s0.tested_pairs == 17701200  # candidate pairs of the 100^3 lattice at R = 1.5
    assert np.diff(rp0).max() == 18
    eng.map_create(res, km, dm, True)
    eng.map_upload_dense(np.zeros_like(occ), known)
    s1 = eng.connect(radius=1.5, box=BOX, use_map=True, use_pyramid=False)
    rp1, col1 = eng.connect_fetch(s1)
    assert np.array_equal(rp0, rp1) and np.array_equal(col0, col1)
    eng.map_upload_dense(occ, known)
    sa = eng.connect(radius=1.5, box=BOX, use_map=True, use_pyramid=True)
    rpa, cola = eng.connect_fetch(sa)
    sb = eng.connect(radius=1.5, box=BOX, use_map=True, use_pyramid=False)
    rpb, colb = eng.connect_fetch(sb)
    assert np.array_equal(rpa, rpb) and np.array_equal(cola, colb)
    assert 0 < sa.edges < s0.edges and sa.pyramid_rows > 0
    # every accepted edge is also an edge of the map-free graph, and columns ascend within a row
    rows = np.repeat(np.arange(len(nodes)), np.diff(rpa))
    d = cola[1:] - cola[:-1]
    assert np.all((d > 0) | (rows[1:] != rows[:-1]))
    key0 = rows.astype(np.int64) * len(nodes) + cola
    rows0 = np.repeat(np.arange(len(nodes)), np.diff(rp0))
    assert np.isin(key0, rows0.astype(np.int64) * len(nodes) + col0, assume_unique=True).all()


def test_full_size_random_nodes_all_paths_agree(eng, monkeypatch):
    """S4b shape (10^6 nodes uniform in the 100 m cube of the 512^3 map): every way the engine can run the
    pass gives the same graph — fixed pair slots, scanned offsets, row shards put back together, the
    chunked delivery to host buffers — and graph symmetry holds where the test is symmetric: without
    a map, (i, j) is an edge iff (j, i) is."""
    res = 0.3
    km, dm, occ, known = scenes.synthetic_map(n_vox=512, res=res, seed=42)
    nodes = scenes.uniform_nodes(1_000_000, extent=100.0, seed=43)
    n = len(nodes)
    eng.map_create(res, km, dm, True)
    eng.map_upload_dense(occ, known)
    eng.set_nodes(nodes)
    s_slot = eng.connect(radius=1.5, box=BOX)
    rp, col = eng.connect_fetch(s_slot)
    assert 0 < s_slot.edges < s_slot.in_radius_pairs and s_slot.pyramid_rows > 0 and s_slot.lattice_pairs > 0
    monkeypatch.setenv("SSPP_CONNECT_SLOT", "0")
    s_scan = eng.connect(radius=1.5, box=BOX)
    rp2, col2 = eng.connect_fetch(s_scan)
    assert np.array_equal(rp, rp2) and np.array_equal(col, col2)
    monkeypatch.delenv("SSPP_CONNECT_SLOT")
    cuts = [0, 123_457, 600_000, n]
    parts = [eng.connect_fetch(eng.connect(radius=1.5, box=BOX, rows=(cuts[g], cuts[g + 1]))) for g in range(3)]
    assert np.array_equal(np.concatenate([p[1] for p in parts]), col)
    base = np.cumsum([0] + [len(p[1]) for p in parts])
    assert np.array_equal(np.concatenate([parts[g][0][:-1] + base[g] for g in range(3)] + [[base[3]]]), rp)
    rph = np.zeros(n + 1, dtype=np.int64)
    colh = np.zeros(s_slot.edges, dtype=np.int32)
    s_host = eng.connect_nodes_host(nodes, rph, colh, radius=1.5, box=BOX)
    assert s_host.edges == s_slot.edges and np.array_equal(rph, rp) and np.array_equal(colh, col)
    s_free = eng.connect(radius=1.5, use_map=False)
    rpf, colf = eng.connect_fetch(s_free)
    rows = np.repeat(np.arange(n, dtype=np.int64), np.diff(rpf))
    fwd = np.sort(rows * n + colf)
    bwd = np.sort(colf.astype(np.int64) * n + rows)
    assert np.array_equal(fwd, bwd)


@pytest.mark.parametrize("res,box", [(0.3, BOX), (0.15, BOX), (1.0, BOX), (0.1, (0.5, 0.5, 0.3))])
def test_voxel_step_count_matches_reference_semantics(eng, orc, res, box):
    """the instrumented run counts exactly the lookups getLineStatusBoundingBox performs (rays in
    lattice order, stopping at the first non-free voxel of the first failing ray) — the Σsteps term
    of the roofline's byte model"""
    n_vox = 64 if res >= 0.3 else 96
    w = make_map(eng, orc, res, True, n_vox=n_vox, seed=21, shell=0.1)
    ext = n_vox * res
    nodes = scenes.uniform_nodes(600, extent=ext * 0.7, seed=4)
    radius = min(1.5, ext / 6)
    eng.set_nodes(nodes)
    st = eng.connect(radius=radius, box=box, count_steps=True)
    rp, col = eng.connect_fetch(st)
    rp_o, col_o, cnt = orc.connect(w, nodes, radius, box)
    assert np.array_equal(col, col_o)
    assert st.voxel_steps == cnt[2] and st.voxel_steps > 0
    assert st.tested_pairs * st.rays_per_pair >= cnt[3]
    w.close()


def test_device_block_assembly(eng, orc):
    """sspp_gpu_assemble_csr_blocks: padded per-rank blocks (as an all-gather leaves them) are packed into the
    whole-graph CSR on the device; a slot too small for its block is reported, not truncated"""
    import torch
    w = make_map(eng, orc, 0.3, True, n_vox=96, seed=9)
    nodes = scenes.uniform_nodes(2500, extent=20.0, seed=12)
    eng.set_nodes(nodes)
    st = eng.connect(radius=1.7, box=BOX)
    rp, col = eng.connect_fetch(st)
    world = 3
    bounds = [0, 833, 1666, 2500]
    stride_r = 835
    blocks = []
    for lo, hi in zip(bounds[:-1], bounds[1:]):
        s = eng.connect(radius=1.7, box=BOX, rows=(lo, hi))
        blocks.append(eng.connect_fetch(s))
    stride_e = max(len(c) for _, c in blocks) + 7
    dev = torch.device("cuda", 0)
    rp_all = torch.zeros(world * stride_r, dtype=torch.int64, device=dev)
    col_all = torch.full((world * stride_e,), -1, dtype=torch.int32, device=dev)
    for g, (r, c) in enumerate(blocks):
        rp_all[g * stride_r: g * stride_r + len(r)] = torch.from_numpy(r).to(dev)
        col_all[g * stride_e: g * stride_e + len(c)] = torch.from_numpy(c).to(dev)
    base = torch.tensor(bounds, dtype=torch.int64, device=dev)
    out_rp = torch.zeros(len(nodes) + 1, dtype=torch.int64, device=dev)
    out_col = torch.zeros(world * stride_e, dtype=torch.int32, device=dev)
    over = torch.zeros(1, dtype=torch.int64, device=dev)
    torch.cuda.synchronize()
    eng.assemble_csr_blocks(rp_all.data_ptr(), col_all.data_ptr(), world, stride_r, stride_e, base.data_ptr(), out_rp.data_ptr(),
                            out_col.data_ptr(), over.data_ptr())
    eng.synchronize()
    assert int(over.item()) == 0
    assert np.array_equal(out_rp.cpu().numpy(), rp) and np.array_equal(out_col.cpu().numpy()[:len(col)], col)
    # slots one entry too small for the largest block
    small = max(len(c) for _, c in blocks) - 1
    eng.assemble_csr_blocks(rp_all.data_ptr(), col_all.data_ptr(), world, stride_r, small, base.data_ptr(), out_rp.data_ptr(),
                            out_col.data_ptr(), over.data_ptr())
    eng.synchronize()
    assert int(over.item()) == small + 1
    w.close()


@pytest.mark.parametrize("res,extent,n", [(0.3, 8.0, 4000), (0.5, 7.0, 2500), (0.1, 6.0, 1200)])
def test_dense_rows_with_map(eng, orc, res, extent, n):
    """rows with far more than 32 pairs (batches of 32, global-memory rank sort) with a map, at 27, 18 and
    245 rays per pair; also the instrumented step count on such rows"""
    n_vox = 64 if res >= 0.3 else 128
    w = make_map(eng, orc, res, True, n_vox=n_vox, seed=31, region=extent * 1.2, n_boxes=10, n_spheres=6)
    nodes = scenes.uniform_nodes(n, extent=extent, seed=17)
    st, cnt = assert_same_graph(eng, orc, w, nodes, 1.5)
    assert st.in_radius_pairs / n > 40
    eng.set_nodes(nodes)
    sc = eng.connect(radius=1.5, box=BOX, count_steps=True)
    rp, col = eng.connect_fetch(sc)
    rp_o, col_o, _ = orc.connect(w, nodes, 1.5, BOX)
    assert np.array_equal(col, col_o) and sc.voxel_steps == cnt[2]
    w.close()


def test_casarea_map_golden(eng):
    """BASELINE config 1 with a map (S1m): resources/casarea_s.png via resources/map.yaml (res 0.05, 847 rays per
    pair), extruded over z in [-1,1); adjacency from the reference's compiled sources (tests/golden/make_golden.py)"""
    g = np.load(os.path.join(GOLD, "casarea_s1m.npz"))
    occ2 = np.unpackbits(g["occ2d"], axis=1)[:, :640].astype(bool)
    known2 = np.unpackbits(g["known2d"], axis=1)[:, :640].astype(bool)
    nz = 40
    occ = scenes.pack_bits(np.repeat(occ2[None], nz, axis=0))
    known = scenes.pack_bits(np.repeat(known2[None], nz, axis=0))
    eng.map_create(0.05, [32768 - 320, 32768 - 240, 32768 - 20], [640, 480, nz], True)
    eng.map_upload_dense(occ, known)
    eng.set_nodes(g["poses"][:, :3])
    for pyr in (True, False):
        st = eng.connect(radius=1.5, box=BOX, use_pyramid=pyr)
        rp, col = eng.connect_fetch(st)
        assert st.rays_per_pair == 847 and st.tested_pairs == int(g["collision_calls"])
        assert np.array_equal(rp, g["row_ptr"]) and np.array_equal(col, g["col_idx"])
