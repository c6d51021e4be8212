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


def test_row_ranges_partition():
    for n in (0, 1, 7, 1000003):
        for world in (1, 2, 3, 8):
            cuts = [sharded.row_range(n, r, world) for r in range(world)]
            assert cuts[0][0] == 0 and cuts[-1][1] == n
            assert all(cuts[i][1] == cuts[i + 1][0] for i in range(world - 1))
