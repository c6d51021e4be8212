"""Multi-GPU path on CPU: two gloo ranks each hold the CSR block of their row range (cut from the
oracle's graph); the assembled CSR must equal the whole graph on both ranks."""
import os
import socket

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import oracle_lib
from sspp_b200 import scenes, sharded


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, rp, col, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    n = len(rp) - 1
    lo, hi = sharded.row_range(n, rank, world)
    rp_l = torch.from_numpy(rp[lo:hi + 1] - rp[lo])
    col_l = torch.from_numpy(col[rp[lo]:rp[hi]].copy())
    buffers = None
    for _ in range(2):  # second round reuses the buffers
        rp_g, col_g, buffers = sharded.assemble_csr(rp_l, col_l, buffers=buffers)
    ok = bool(torch.equal(rp_g, torch.from_numpy(rp)) and torch.equal(col_g, torch.from_numpy(col)))
    out[rank] = ok
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_csr_assembly(built):
    O = oracle_lib.Oracle()
    res = 0.3
    km, dm, occ, known = scenes.synthetic_map(n_vox=64, res=res, seed=3, n_boxes=8, n_spheres=6, region=14.0)
    w = O.world(res, True)
    w.build_dense(km, dm, occ, known)
    nodes = scenes.uniform_nodes(701, extent=12.0, seed=9)   # odd count: uneven blocks
    rp, col, _ = O.connect(w, nodes, 1.7)
    w.close()
    assert len(col) > 0
    ctx = mp.get_context("spawn")
    out = ctx.Manager().dict()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, rp, col, out)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    assert out.get(0) is True and out.get(1) is True


def _vis_worker(rank, world, port, res, km, dm, occ, known, pts, poses, vp_full, idx_full, out):
    """one rank of the sharded visibility pass: the oracle's visibility on ITS block of the viewpoints (model
    points and map replicated), then the same assembly as for the adjacency"""
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    O = oracle_lib.Oracle()
    w = O.world(res, True)
    w.build_dense(km, dm, occ, known)
    lo, hi = sharded.row_cuts(len(poses), world)[rank:rank + 2]
    vp_l, idx_l, _, _ = w.visibility(poses[lo:hi], 58.0, 45.0, 0.7, 6.0, pts)
    w.close()
    vp_g, idx_g, _ = sharded.assemble_csr(torch.from_numpy(vp_l), torch.from_numpy(idx_l.copy()))
    out[rank] = bool(torch.equal(vp_g, torch.from_numpy(vp_full)) and torch.equal(idx_g, torch.from_numpy(idx_full)))
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_visibility_assembly(built):
    """visibility shards by viewpoint (the hook runs per way-point, include/sspp/heuristic_interface.h:40): two gloo
    ranks, contiguous viewpoint blocks, (view_ptr, point_idx) assembled on both = one pass over all viewpoints"""
    O = oracle_lib.Oracle()
    res = 0.3
    km, dm, occ, known = scenes.synthetic_map(n_vox=64, res=res, seed=3, n_boxes=8, n_spheres=6, region=14.0)
    pts = scenes.surface_points(occ, known, km, dm, res, 5000, seed=45)
    poses = scenes.viewpoints_near(pts, 41, standoff=2.5, seed=46)  # odd count: uneven blocks
    w = O.world(res, True)
    w.build_dense(km, dm, occ, known)
    vp_full, idx_full, rays, _ = w.visibility(poses, 58.0, 45.0, 0.7, 6.0, pts)
    w.close()
    assert len(idx_full) > 0
    ctx = mp.get_context("spawn")
    out = ctx.Manager().dict()
    port = _free_port()
    procs = [ctx.Process(target=_vis_worker, args=(r, 2, port, res, km, dm, occ, known, pts, poses, vp_full, idx_full, out)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    assert out.get(0) is True and out.get(1) is True


def test_row_cuts_accept_balanced_cut_points():
    assert sharded.row_cuts(10, 3) == [0, 3, 6, 10]
    assert sharded.row_cuts(10, 3, np.array([0, 7, 7, 10])) == [0, 7, 7, 10]
    import pytest
    with pytest.raises(AssertionError):
        sharded.row_cuts(10, 3, [0, 8, 7, 10])


def test_row_ranges_partition():
    for n in (0, 1, 7, 1000003):
        for world in (1, 2, 3, 8):
            cuts = [sharded.row_range(n, r, world) for r in range(world)]
            assert cuts[0][0] == 0 and cuts[-1][1] == n
            assert all(cuts[i][1] == cuts[i + 1][0] for i in range(world - 1))
