"""Multi-GPU CSR assembly through peer memory (sspp_gpu_connect_push_csr): the blocks the ranks push
must reproduce, byte for byte, the CSR of one connect over all rows.

One GPU is enough for the kernel's logic: several contexts on the same device play the ranks and
ordinary device tensors play the peers' buffers.  The CUDA-IPC path (one process per GPU, the real
deployment) runs when the box has two or more GPUs and is skipped otherwise."""
import os
import sys

import numpy as np
import pytest

from sspp_b200 import scenes
from sspp_b200.engine import Engine, SsppError

pytestmark = pytest.mark.gpu
BOX = (0.5, 0.5, 0.3)
HERE = os.path.dirname(os.path.abspath(__file__))


def scene(eng_list, n_vox=64, res=0.3, seed=5):
    km, dm, occ, known = scenes.synthetic_map(n_vox=n_vox, res=res, seed=seed, n_boxes=8, n_spheres=6, region=n_vox * res * 0.75)
    for e in eng_list:
        e.map_create(res, km, dm, True)
        e.map_upload_dense(occ, known)


@pytest.mark.parametrize("world,cuts", [(1, None), (2, None), (3, (0.21, 0.55)), (5, None)])
def test_push_equals_one_gpu_csr(eng, world, cuts):
    import torch
    dev = torch.device("cuda", 0)
    nodes = scenes.uniform_nodes(6000, extent=13.0, seed=9)
    n = len(nodes)
    scene([eng])
    eng.set_nodes(nodes)
    st = eng.connect(radius=1.5, box=BOX)
    rp_full, col_full = eng.connect_fetch(st)
    assert st.edges > 1000

    bounds = [0] + ([int(c * n) for c in cuts] if cuts else [(n * g) // world for g in range(1, world)]) + [n]
    ranks = [Engine(0) for _ in range(world)]
    try:
        scene(ranks)
        stats = []
        for g, e in enumerate(ranks):
            e.set_nodes(nodes)
            stats.append(e.connect(radius=1.5, box=BOX, rows=(bounds[g], bounds[g + 1])))
        counts = torch.tensor([s.edges for s in stats], dtype=torch.int64, device=dev)
        assert int(counts.sum()) == st.edges
        row_base = torch.tensor(bounds, dtype=torch.int64, device=dev)
        cap = ((st.edges + 3) // 4) * 4 + 8
        out_rp = [torch.full((n + 1,), -1, dtype=torch.int64, device=dev) for _ in range(world)]
        out_col = [torch.full((cap,), -1, dtype=torch.int32, device=dev) for _ in range(world)]
        over = torch.zeros(1, dtype=torch.int64, device=dev)
        torch.cuda.synchronize()
        for g, e in enumerate(ranks):
            e.connect_push_csr(world, g, counts.data_ptr(), row_base.data_ptr(), [t.data_ptr() for t in out_rp],
                               [t.data_ptr() for t in out_col], cap, over.data_ptr())
        for e in ranks:
            e.synchronize()
        assert int(over.item()) == 0
        for k in range(world):
            assert np.array_equal(out_rp[k].cpu().numpy(), rp_full), "row_ptr of peer %d differs" % k
            c = out_col[k].cpu().numpy()
            assert np.array_equal(c[: st.edges], col_full), "col_idx of peer %d differs" % k
            assert (c[st.edges:] == -1).all(), "push wrote past the graph's edges"

        # a graph that does not fit: nothing is written, every rank reports the total
        out_rp2 = torch.full((n + 1,), -1, dtype=torch.int64, device=dev)
        out_col2 = torch.full((cap,), -1, dtype=torch.int32, device=dev)
        ranks[0].connect_push_csr(world, 0, counts.data_ptr(), row_base.data_ptr(), [out_rp2.data_ptr()] * world,
                                  [out_col2.data_ptr()] * world, st.edges - 1, over.data_ptr())
        ranks[0].synchronize()
        assert int(over.item()) == st.edges
        assert (out_rp2 == -1).all() and (out_col2 == -1).all()
    finally:
        for e in ranks:
            e.close()


def test_push_argument_errors(eng):
    import torch
    dev = torch.device("cuda", 0)
    t = torch.zeros(8, dtype=torch.int64, device=dev)
    with pytest.raises(SsppError):  # no connect result yet
        eng.connect_push_csr(1, 0, t.data_ptr(), t.data_ptr(), [t.data_ptr()], [t.data_ptr()], 8, t.data_ptr())
    eng.set_nodes(scenes.lattice_nodes(4, 1.0))
    eng.connect(radius=1.5, use_map=False)
    with pytest.raises(SsppError):  # rank outside the world
        eng.connect_push_csr(2, 2, t.data_ptr(), t.data_ptr(), [t.data_ptr()] * 2, [t.data_ptr()] * 2, 8, t.data_ptr())
    with pytest.raises(SsppError):  # more ranks than SSPP_MAX_PEERS
        eng.connect_push_csr(17, 0, t.data_ptr(), t.data_ptr(), [t.data_ptr()] * 17, [t.data_ptr()] * 17, 8, t.data_ptr())
    with pytest.raises(SsppError):  # column buffer not 16-byte aligned
        eng.connect_push_csr(1, 0, t.data_ptr(), t.data_ptr(), [t.data_ptr()], [t.data_ptr() + 4], 8, t.data_ptr())
    with pytest.raises(SsppError):  # not a peer buffer of this context
        eng.peer_buffer_release(t.data_ptr())
    ptr, handle = eng.peer_buffer_create(4096)
    assert ptr != 0 and len(handle) == 64
    eng.peer_buffer_release(ptr)


@pytest.mark.parametrize("lattice,world", [(12, 2), (12, 3), (56, 2)])
def test_connect_to_peers_blocks_pack_to_one_gpu_csr(eng, lattice, world):
    """sspp_gpu_connect_to_peers: every rank's block lands in slot [rank] of all ranks' block buffers
    (chunk by chunk above 65 536 rows per rank: lattice 56 -> 87 808 rows each), and packing the blocks
    gives the CSR of one connect over all rows."""
    import torch
    dev = torch.device("cuda", 0)
    km, dm, occ, known = scenes.synthetic_map(n_vox=224, res=0.3, seed=12, n_boxes=60, n_spheres=40, region=56.0)
    nodes = scenes.lattice_nodes(lattice, 1.0)
    if world * len(nodes) > 200000:
        nodes = np.concatenate([nodes, nodes + np.array([0.0, 0.0, 0.37])])  # two rows per position column, still one space
    n = len(nodes)
    eng.map_create(0.3, km, dm, True)
    eng.map_upload_dense(occ, known)
    eng.set_nodes(nodes)
    st = eng.connect(radius=1.5, box=BOX)
    rp_full, col_full = eng.connect_fetch(st)
    bounds = [(n * g) // world for g in range(world + 1)]
    stride_r = max(bounds[g + 1] - bounds[g] for g in range(world)) + 1
    ranks = [Engine(0) for _ in range(world)]
    try:
        for e in ranks:
            e.map_create(0.3, km, dm, True)
            e.map_upload_dense(occ, known)
            e.set_nodes(nodes)
        stride_e = st.edges  # generous: any block fits
        rp_blk = [torch.full((world * stride_r,), -1, dtype=torch.int64, device=dev) for _ in range(world)]
        col_blk = [torch.full((world * stride_e,), -1, dtype=torch.int32, device=dev) for _ in range(world)]
        torch.cuda.synchronize()
        edges = []
        for g, e in enumerate(ranks):
            s = e.connect_to_peers(world, g, [t.data_ptr() for t in rp_blk], [t.data_ptr() for t in col_blk], stride_r, stride_e,
                                   radius=1.5, box=BOX, rows=(bounds[g], bounds[g + 1]))
            edges.append(s.edges)
        assert sum(edges) == st.edges
        row_base = torch.tensor(bounds, dtype=torch.int64, device=dev)
        for k in range(world):
            out_rp = torch.empty(n + 1, dtype=torch.int64, device=dev)
            out_col = torch.empty(world * stride_e, dtype=torch.int32, device=dev)
            over = torch.zeros(1, dtype=torch.int64, device=dev)
            ranks[k].assemble_csr_blocks(rp_blk[k].data_ptr(), col_blk[k].data_ptr(), world, stride_r, stride_e, row_base.data_ptr(),
                                         out_rp.data_ptr(), out_col.data_ptr(), over.data_ptr())
            ranks[k].synchronize()
            assert int(over.item()) == 0
            assert np.array_equal(out_rp.cpu().numpy(), rp_full), "row_ptr on rank %d differs" % k
            assert np.array_equal(out_col.cpu().numpy()[: st.edges], col_full), "col_idx on rank %d differs" % k
        # a slot that is too small: the error carries the size needed (the caller resizes and runs the pass again)
        with pytest.raises(SsppError) as ei:
            ranks[0].connect_to_peers(world, 0, [t.data_ptr() for t in rp_blk], [t.data_ptr() for t in col_blk], stride_r,
                                      edges[0] // 2, radius=1.5, box=BOX, rows=(bounds[0], bounds[1]))
        assert ei.value.code == 4 and ei.value.needed_edges == edges[0]
    finally:
        for e in ranks:
            e.close()


@pytest.mark.parametrize("world", [2, 3])
def test_visibility_blocks_push_to_one_gpu_result(eng, orc, world):
    """Sharded visibility (SURVEY 8 e: "for visibility: same pattern on (counts, indices)"): every rank computes a
    contiguous block of the viewpoints, sspp_gpu_visibility_push writes the blocks into all ranks' whole-result
    buffers, and (view_ptr, point_idx) equals one call over all viewpoints — and the oracle's answer."""
    import torch
    dev = torch.device("cuda", 0)
    res = 0.3
    km, dm, occ, known = scenes.synthetic_map(n_vox=64, res=res, seed=5, n_boxes=8, n_spheres=6, region=14.0)
    pts = scenes.surface_points(occ, known, km, dm, res, 20000, seed=45)
    poses = scenes.viewpoints_near(pts, 301, standoff=2.5, seed=46)
    fr = (58.0, 45.0, 0.7, 6.0)
    eng.map_create(res, km, dm, True)
    eng.map_upload_dense(occ, known)
    eng.set_model_points(pts)
    st, vp_full, idx_full = eng.visibility(poses, *fr)
    assert st.visible_pairs > 1000
    w = orc.world(res, True)
    w.build_dense(km, dm, occ, known)
    vp_o, idx_o, _, _ = w.visibility(poses, *fr, pts)
    w.close()
    assert np.array_equal(vp_full, vp_o) and np.array_equal(idx_full, idx_o)
    nv = len(poses)
    bounds = [(nv * g) // world for g in range(world + 1)]
    ranks = [Engine(0) for _ in range(world)]
    try:
        stats = []
        for g, e in enumerate(ranks):
            e.map_create(res, km, dm, True)
            e.map_upload_dense(occ, known)
            e.set_model_points(pts)
            stats.append(e.visibility(poses[bounds[g]:bounds[g + 1]], *fr, fetch=False))
        counts = torch.tensor([s.visible_pairs for s in stats], dtype=torch.int64, device=dev)
        assert int(counts.sum()) == st.visible_pairs
        view_base = torch.tensor(bounds, dtype=torch.int64, device=dev)
        cap = ((st.visible_pairs + 3) // 4) * 4 + 8
        out_vp = [torch.full((nv + 1,), -1, dtype=torch.int64, device=dev) for _ in range(world)]
        out_idx = [torch.full((cap,), -1, dtype=torch.int32, device=dev) for _ in range(world)]
        over = torch.zeros(1, dtype=torch.int64, device=dev)
        torch.cuda.synchronize()
        for g, e in enumerate(ranks):
            e.visibility_push(world, g, counts.data_ptr(), view_base.data_ptr(), [t.data_ptr() for t in out_vp],
                              [t.data_ptr() for t in out_idx], cap, over.data_ptr())
        for e in ranks:
            e.synchronize()
        assert int(over.item()) == 0
        for k in range(world):
            assert np.array_equal(out_vp[k].cpu().numpy(), vp_full), "view_ptr of peer %d differs" % k
            assert np.array_equal(out_idx[k].cpu().numpy()[: st.visible_pairs], idx_full), "point_idx of peer %d differs" % k
        with pytest.raises(SsppError):  # no visibility result on a fresh context
            Engine(0).visibility_push(world, 0, counts.data_ptr(), view_base.data_ptr(), [t.data_ptr() for t in out_vp],
                                      [t.data_ptr() for t in out_idx], cap, over.data_ptr())
    finally:
        for e in ranks:
            e.close()


def test_balanced_cuts_split_the_work_not_the_rows(eng):
    """sspp_gpu_balanced_cuts (SURVEY 8 e "split by prefix-sum of candidate counts"): the cut points are where the
    weighted prefix of the per-row pair counts crosses g/G of the total — recomputed here from the engine's own
    per-row data — and the blocks they define still concatenate to the one-GPU CSR."""
    res = 0.3
    km, dm, occ, known = scenes.synthetic_map(n_vox=96, res=res, seed=7, n_boxes=20, n_spheres=14, region=22.0)
    rng = np.random.RandomState(5)
    # dense cloud on one side, sparse on the other: equal row counts would be badly unbalanced
    nodes = np.ascontiguousarray(np.concatenate([rng.uniform([-11, -11, -11], [0, 11, 11], (9000, 3)),
                                                 rng.uniform([0, -11, -11], [11, 11, 11], (3000, 3))]))
    n = len(nodes)
    eng.map_create(res, km, dm, True)
    eng.map_upload_dense(occ, known)
    eng.set_nodes(nodes)
    st = eng.connect(radius=1.5, box=BOX)
    rp, col = eng.connect_fetch(st)
    cls = eng.connect_row_classes(st)
    s0 = eng.connect(radius=1.5, use_map=False)
    deg0 = np.diff(eng.connect_fetch(s0)[0])  # in-radius pairs per row
    for world, heavy in ((2, 8), (5, 8), (3, 1)):
        cuts = eng.balanced_cuts(world, heavy, radius=1.5, box=BOX)
        w = deg0 * np.where(cls == 3, heavy, 1)
        prefix = np.concatenate([[0], np.cumsum(w)])
        want = [0] + [int(np.searchsorted(prefix, (int(prefix[-1]) * g) // world, side="left")) for g in range(1, world)] + [n]
        assert list(cuts) == want
        loads = [int(prefix[cuts[g + 1]] - prefix[cuts[g]]) for g in range(world)]
        assert max(loads) <= 1.05 * (prefix[-1] / world) + w.max()
        parts = [eng.connect_fetch(eng.connect(radius=1.5, box=BOX, rows=(int(cuts[g]), int(cuts[g + 1])))) for g in range(world)]
        assert np.array_equal(np.concatenate([q[1] for q in parts]), col)
        base = np.cumsum([0] + [len(q[1]) for q in parts])
        assert np.array_equal(np.concatenate([parts[g][0][:-1] + base[g] for g in range(world)] + [[base[-1]]]), rp)
    assert list(eng.balanced_cuts(1, 8, radius=1.5, box=BOX)) == [0, n]


@pytest.mark.parametrize("kind,world", [("lattice", 2), ("lattice", 3), ("random", 3), ("big", 2)])
def test_fused_exchange_packs_to_one_gpu_csr(eng, kind, world):
    """sspp_gpu_connect_fused: the edge kernel of every rank stores its finished rows straight into the slot-format images
    of all ranks (here: several contexts on one GPU share the images), and sspp_gpu_pack_slots turns a complete image
    into the CSR of one connect over all rows, byte for byte — rows answered by the pyramid, rows without any pair and
    rows decided ray by ray alike; uneven cuts; both pyramid settings."""
    import torch
    dev = torch.device("cuda", 0)
    if kind == "big":
        km, dm, occ, known = scenes.synthetic_map(n_vox=224, res=0.3, seed=12, n_boxes=60, n_spheres=40, region=56.0)
        nodes = scenes.lattice_nodes(56, 1.0)
    else:
        km, dm, occ, known = scenes.synthetic_map(n_vox=64, res=0.3, seed=5, n_boxes=8, n_spheres=6, region=64 * 0.3 * 0.75)
        nodes = scenes.lattice_nodes(14, 1.0) if kind == "lattice" else np.concatenate(
            [scenes.uniform_nodes(1500, extent=13.0, seed=9), np.array([[40.0, 40.0, 40.0], [0.1, 0.2, 0.3], [0.1, 0.2, 0.3]])])
    n, slot = len(nodes), 32
    eng.map_create(0.3, km, dm, True)
    eng.map_upload_dense(occ, known)
    eng.set_nodes(nodes)
    bounds = [0] + sorted(int(x) for x in np.random.RandomState(3).choice(np.arange(1, n), world - 1, replace=False)) + [n]
    ranks = [Engine(0) for _ in range(world)]
    try:
        for e in ranks:
            e.map_create(0.3, km, dm, True)
            e.map_upload_dense(occ, known)
            e.set_nodes(nodes)
        for pyramid in (True, False):
            st = eng.connect(radius=1.5, box=BOX, use_pyramid=pyramid)
            rp_full, col_full = eng.connect_fetch(st)
            images = [(torch.full((n * slot,), -7, dtype=torch.int32, device=dev), torch.full((n,), -7, dtype=torch.int32, device=dev))
                      for _ in range(world)]
            torch.cuda.synchronize()
            edges = 0
            for g, e in enumerate(ranks):
                s = e.connect_fused(world, g, slot, bounds[g], [im[0].data_ptr() for im in images], [im[1].data_ptr() for im in images],
                                    radius=1.5, box=BOX, rows=(bounds[g], bounds[g + 1]), use_pyramid=pyramid)
                edges += s.edges
                assert s.rows == bounds[g + 1] - bounds[g]
            assert edges == st.edges
            for k in range(world):
                assert np.array_equal(images[k][1].cpu().numpy(), np.diff(rp_full)), "degrees in the image of rank %d" % k
                out_rp = torch.empty(n + 1, dtype=torch.int64, device=dev)
                out_col = torch.full((st.edges + 8,), -1, dtype=torch.int32, device=dev)
                over = torch.zeros(1, dtype=torch.int64, device=dev)
                ranks[k].pack_slots(images[k][0].data_ptr(), images[k][1].data_ptr(), n, slot, out_rp.data_ptr(), out_col.data_ptr(),
                                    st.edges + 8, over.data_ptr())
                ranks[k].synchronize()
                assert int(over.item()) == 0
                assert np.array_equal(out_rp.cpu().numpy(), rp_full), "row_ptr on rank %d differs" % k
                c = out_col.cpu().numpy()
                assert np.array_equal(c[: st.edges], col_full) and (c[st.edges:] == -1).all(), "col_idx on rank %d differs" % k
            # a column buffer that is too small: nothing copied, the total reported
            over = torch.zeros(1, dtype=torch.int64, device=dev)
            out_col = torch.full((st.edges,), -1, dtype=torch.int32, device=dev)
            ranks[0].pack_slots(images[0][0].data_ptr(), images[0][1].data_ptr(), n, slot, out_rp.data_ptr(), out_col.data_ptr(),
                                st.edges - 1, over.data_ptr())
            ranks[0].synchronize()
            assert int(over.item()) == st.edges and (out_col == -1).all()
        with pytest.raises(SsppError):  # the fused pass needs the map
            ranks[0].connect_fused(world, 0, slot, 0, [im[0].data_ptr() for im in images], [im[1].data_ptr() for im in images],
                                   radius=1.5, use_map=False, rows=(0, bounds[1]))
        with pytest.raises(SsppError) as ei:  # rows with more pairs than a slot: refused, the caller falls back
            ranks[0].connect_fused(world, 0, slot, 0, [im[0].data_ptr() for im in images], [im[1].data_ptr() for im in images],
                                   radius=4.0, box=BOX, rows=(0, bounds[1]))
        assert ei.value.code == 4
    finally:
        for e in ranks:
            e.close()


def _ipc_rank(rank, world, port, q):
    import torch
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", device_id=dev)
    from sspp_b200 import sharded
    e = Engine(rank)
    try:
        scene([e])
        nodes = scenes.uniform_nodes(20000, extent=13.0, seed=9)
        n = len(nodes)
        e.set_nodes(nodes)
        asm = sharded.PeerCsrAssembler(e, n, dev)
        lo, hi = sharded.row_range(n, rank, world)
        ok = True
        for radius in (1.0, 1.5, 1.0):  # the second graph outgrows the first one's column buffers
            st = e.connect(radius=radius, box=BOX, rows=(lo, hi))
            rp, col = asm.assemble(st)
            rp, col = rp.cpu().numpy(), col.cpu().numpy()
            sf = e.connect(radius=radius, box=BOX)
            rpf, colf = e.connect_fetch(sf)
            ok = ok and np.array_equal(rp, rpf) and np.array_equal(col[: sf.edges], colf)
            dist.barrier()
        asm.close()
        sasm = sharded.StreamedPeerAssembler(e, n, dev)
        for radius in (1.0, 1.5, 1.0, 1.0):  # sizing pass, a graph that outgrows the slots, both buffer sets
            st, rp, col = sasm.connect(radius=radius, box=BOX)
            rp, col = rp.cpu().numpy(), col.cpu().numpy()
            sf = e.connect(radius=radius, box=BOX)
            rpf, colf = e.connect_fetch(sf)
            ok = ok and np.array_equal(rp, rpf) and np.array_equal(col[: sf.edges], colf)
            dist.barrier()
        sasm.close()
        # the exchange fused into the edge kernel (two buffer sets: three graphs in a row), then a graph it cannot serve
        sparse = scenes.uniform_nodes(3000, extent=13.0, seed=10)  # few enough pairs per row for the 32-pair slots
        e.set_nodes(sparse)
        fasm = sharded.FusedPeerAssembler(e, len(sparse), dev)
        for radius in (0.9, 1.2, 0.9):
            st, rp, col = fasm.connect(radius=radius, box=BOX)
            sf = e.connect(radius=radius, box=BOX)
            rpf, colf = e.connect_fetch(sf)
            ok = ok and sf.edges > 0 and np.array_equal(rp.cpu().numpy(), rpf) and np.array_equal(col.cpu().numpy()[: sf.edges], colf)
            dist.barrier()
        try:
            fasm.connect(radius=4.0, box=BOX)
            ok = False
        except sharded.FusedUnavailable:
            pass
        dist.barrier()
        fasm.close()
        e.set_nodes(nodes)
        # rows cut by work (every rank computes the same cuts from the same nodes and map) ...
        cuts = e.balanced_cuts(world, 8, radius=1.5, box=BOX)
        basm = sharded.PeerCsrAssembler(e, n, dev, cuts=cuts)
        st = e.connect(radius=1.5, box=BOX, rows=basm.rows)
        rp, col = basm.assemble(st)
        sf = e.connect(radius=1.5, box=BOX)
        rpf, colf = e.connect_fetch(sf)
        ok = ok and np.array_equal(rp.cpu().numpy(), rpf) and np.array_equal(col.cpu().numpy()[: sf.edges], colf)
        dist.barrier()
        basm.close()
        # ... and the visibility pass sharded by viewpoint
        km, dm, occ, known = scenes.synthetic_map(n_vox=64, res=0.3, seed=5, n_boxes=8, n_spheres=6, region=64 * 0.3 * 0.75)
        pts = scenes.surface_points(occ, known, km, dm, 0.3, 20000, seed=45)
        poses = scenes.viewpoints_near(pts, 777, standoff=2.5, seed=46)
        e.set_model_points(pts)
        vasm = sharded.PeerCsrAssembler(e, len(poses), dev, kind="visibility")
        for _ in range(2):
            vs = e.visibility(poses[vasm.rows[0]:vasm.rows[1]], 58.0, 45.0, 0.7, 6.0, fetch=False)
            vp, idx = vasm.assemble(vs)
            dist.barrier()
        vf, vpf, idxf = e.visibility(poses, 58.0, 45.0, 0.7, 6.0)
        ok = ok and np.array_equal(vp.cpu().numpy(), vpf) and np.array_equal(idx.cpu().numpy()[: vf.visible_pairs], idxf)
        dist.barrier()
        vasm.close()
        q.put((rank, bool(ok)))
    finally:
        e.close()
        dist.destroy_process_group()


def test_peer_ipc_two_processes(built):
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs on one box")
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    world = 2
    port = 29500 + (os.getpid() % 2000)
    procs = [ctx.Process(target=_ipc_rank, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(timeout=240)
    alive = [p for p in procs if p.is_alive()]
    for p in alive:
        p.kill()
    assert not alive, "a rank hung"
    res = sorted(q.get(timeout=5) for _ in range(world))
    assert res == [(0, True), (1, True)]
