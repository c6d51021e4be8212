"""Row-sharded graph build over the GPUs of one box (SURVEY.md §8 e).

The path shards by source row: rank g owns the contiguous list-index block [n*g/G, n*(g+1)/G) of the
search space, with the map and the node set replicated; no collective touches the data path.  The
one exchange step is the assembly of the CSR: an all-gather of (rows, edges) per rank, then of the
padded row_ptr / col_idx blocks (NCCL has no all-gather-v), then every rank rebases the row pointers
by the running edge count.  Works on any torch.distributed backend (nccl on GPUs, gloo in the tests).

On the GPUs of one box the product path is PeerCsrAssembler: the ranks exchange CUDA IPC handles of
their whole-graph CSR buffers once, and from then on every connect ends with ONE kernel per rank
(sspp_gpu_connect_push_csr) that writes the rank's block, rebased, into all ranks' CSRs through
NVLink peer memory; NCCL only carries the 8-byte edge counts and the closing barrier.
GpuCsrAssembler (NCCL all-gather of padded slots + a packing kernel) stays as the comparison.

StreamedPeerAssembler goes one step further and overlaps the exchange with the edge test:
sspp_gpu_connect_to_peers runs the test in row chunks and a small push kernel on a second stream stores
every finished chunk into slot [rank] of all ranks' block buffers over NVLink while the next chunk's edge
kernel runs (all of it queued by the host in one go: the chunk sizes stay on the device); what is left
after the pass is one MAX all-reduce (barrier + "does every block fit") and the local packing kernel.

The same push serves the visibility pass (PeerCsrAssembler(kind="visibility")): viewpoints are split in
contiguous blocks like rows, and (view_ptr, point_idx) is assembled on every rank.  Row blocks may be cut
by work instead of by count (Engine.balanced_cuts -> the `cuts` argument of the assemblers).
"""
import torch
import torch.distributed as dist


def row_range(n, rank, world):
    """Contiguous block of list indices owned by `rank` (equal row counts)."""
    return (n * rank) // world, (n * (rank + 1)) // world


def row_cuts(n, world, cuts=None):
    """world + 1 cut points: the given ones (e.g. Engine.balanced_cuts: equal WORK per block) or equal row counts."""
    if cuts is None:
        return [row_range(n, g, world)[0] for g in range(world)] + [n]
    cuts = [int(c) for c in cuts]
    assert len(cuts) == world + 1 and cuts[0] == 0 and cuts[-1] == n and all(a <= b for a, b in zip(cuts, cuts[1:]))
    return cuts


def gather_counts(rows, edges, device):
    world = dist.get_world_size()
    mine = torch.tensor([rows, edges], dtype=torch.int64, device=device)
    allc = [torch.empty_like(mine) for _ in range(world)]
    dist.all_gather(allc, mine)
    return torch.stack(allc).cpu()


def assemble_csr(row_ptr_local, col_local, counts=None, buffers=None):
    """row_ptr_local: int64 tensor (rows+1, relative to the block); col_local: int32 tensor (edges).
    Returns (row_ptr, col_idx) of the whole graph on every rank, as tensors on the inputs' device."""
    world = dist.get_world_size()
    device = row_ptr_local.device
    rows, edges = row_ptr_local.numel() - 1, col_local.numel()
    if counts is None:
        counts = gather_counts(rows, edges, device)
    max_r, max_e = int(counts[:, 0].max()), max(int(counts[:, 1].max()), 1)
    if buffers is None or buffers[0].numel() != max_r + 1 or buffers[1].numel() != max_e:
        buffers = (torch.zeros(max_r + 1, dtype=torch.int64, device=device), torch.zeros(max_e, dtype=torch.int32, device=device),
                   torch.empty(world * (max_r + 1), dtype=torch.int64, device=device),
                   torch.empty(world * max_e, dtype=torch.int32, device=device))
    rp_s, col_s, rp_all, col_all = buffers
    rp_s[: rows + 1].copy_(row_ptr_local)
    col_s[:edges].copy_(col_local)
    dist.all_gather_into_tensor(rp_all, rp_s)
    dist.all_gather_into_tensor(col_all, col_s)
    rp_parts, col_parts, base = [torch.zeros(1, dtype=torch.int64, device=device)], [], 0
    for g in range(world):
        r, e = int(counts[g, 0]), int(counts[g, 1])
        blk = rp_all[g * (max_r + 1): g * (max_r + 1) + r + 1]
        rp_parts.append(blk[1:] + base)
        col_parts.append(col_all[g * max_e: g * max_e + e])
        base += e
    return torch.cat(rp_parts), torch.cat(col_parts), buffers


class GpuCsrAssembler:
    """Steady-state assembly for CUDA ranks: fixed-size slots all-gathered over NCCL, then packed on the
    device by the engine (sspp_gpu_assemble_csr_blocks) — no host round trip on the data path.  The
    column slot size follows a MAX all-reduce of the ranks' block sizes (8 bytes, every call), so all
    ranks grow their slots in the same call."""

    def __init__(self, eng, n_nodes, device, cuts=None):
        self.eng = eng
        self.world = dist.get_world_size()
        self.rank = dist.get_rank()
        self.device = device
        base = row_cuts(n_nodes, self.world, cuts)
        self.rows = (base[self.rank], base[self.rank + 1])
        self.row_base = torch.tensor(base, dtype=torch.int64, device=device)
        self.stride_r = max(base[g + 1] - base[g] for g in range(self.world)) + 1
        self.n = n_nodes
        self.stride_e = 0
        self.ext_stream = torch.cuda.ExternalStream(eng.stream(), device=device)
        self.send_rp = torch.zeros(self.stride_r, dtype=torch.int64, device=device)
        self.all_rp = torch.empty(self.world * self.stride_r, dtype=torch.int64, device=device)
        self.out_rp = torch.empty(n_nodes + 1, dtype=torch.int64, device=device)
        self.overflow = torch.zeros(1, dtype=torch.int64, device=device)
        self.need = torch.zeros(1, dtype=torch.int64, device=device)
        self.send_col = self.all_col = self.out_col = None

    def _size_columns(self, edges_max):
        """Local (no collective): every rank calls it with the same all-reduced figure."""
        self.stride_e = max(int(edges_max * 1.1) + 1024, 1)
        self.send_col = torch.empty(self.stride_e, dtype=torch.int32, device=self.device)
        self.all_col = torch.empty(self.world * self.stride_e, dtype=torch.int32, device=self.device)
        self.out_col = torch.empty(self.world * self.stride_e, dtype=torch.int32, device=self.device)

    def assemble(self, stats):
        """stats: the ConnectStats of this rank's sspp_gpu_connect.  Returns (row_ptr, col_idx) device
        tensors of the whole graph (col_idx is a view of the first E entries' buffer).

        Whether the column slots have to grow is decided by ALL ranks from the same number — a MAX
        all-reduce of the block sizes, every call — never from a rank's own edge count: a rank that
        resized alone would enter a different collective than its peers."""
        self.need.fill_(int(stats.edges))
        dist.all_reduce(self.need, op=dist.ReduceOp.MAX)
        need = int(self.need.item())
        if need > self.stride_e:
            self._size_columns(need)
        for attempt in range(2):
            self.eng.connect_fetch_device(self.send_rp.data_ptr(), self.send_col.data_ptr())
            dist.all_gather_into_tensor(self.all_rp, self.send_rp)
            dist.all_gather_into_tensor(self.all_col, self.send_col)
            self.ext_stream.wait_event(torch.cuda.current_stream().record_event())
            self.eng.assemble_csr_blocks(self.all_rp.data_ptr(), self.all_col.data_ptr(), self.world, self.stride_r, self.stride_e,
                                         self.row_base.data_ptr(), self.out_rp.data_ptr(), self.out_col.data_ptr(),
                                         self.overflow.data_ptr())
            self.eng.synchronize()
            over = int(self.overflow.item())
            if over == 0:
                return self.out_rp, self.out_col
            # cannot happen after the all-reduce above; kept as a guard.  Every rank packs the same gathered
            # blocks, so all of them see the same flag and resize together.
            self.overflow.zero_()
            self._size_columns(over)
        raise RuntimeError("CSR assembly: column slots still too small after resizing")


class PeerMemoryUnavailable(RuntimeError):
    """Some rank could not create or map a peer buffer (CUDA IPC not permitted, no P2P path).  Raised on
    EVERY rank together, so the callers can fall back to the NCCL assembler without a collective mismatch."""


def exchange_peer_buffer(eng, nbytes, device):
    """Create this rank's buffer and open every peer's: -> (own address, [address of rank g's buffer here]).
    Collective.  The success of every rank is agreed on before anyone returns."""
    from .engine import SsppError
    world, rank = dist.get_world_size(), dist.get_rank()
    err, ptr, handle, peers = None, 0, bytes(64), []
    try:
        ptr, handle = eng.peer_buffer_create(nbytes)
    except SsppError as e:
        err = e
    mine = torch.frombuffer(bytearray(handle), dtype=torch.uint8).to(device)
    allh = torch.empty(world * 64, dtype=torch.uint8, device=device)
    dist.all_gather_into_tensor(allh, mine)
    raw = allh.cpu().numpy().tobytes()
    if err is None:
        try:
            for g in range(world):
                peers.append(ptr if g == rank else eng.peer_buffer_open(raw[g * 64:(g + 1) * 64]))
        except SsppError as e:
            err = e
    bad = torch.tensor([0 if err is None else 1], dtype=torch.int32, device=device)
    dist.all_reduce(bad, op=dist.ReduceOp.MAX)
    if int(bad.item()):
        for g, q in enumerate(peers):
            if g != rank:
                eng.peer_buffer_release(q)
        dist.barrier()
        if ptr:
            eng.peer_buffer_release(ptr)
        raise PeerMemoryUnavailable(str(err) if err is not None else "another rank could not map the peer buffers")
    return ptr, peers


class _DeviceArray:
    """A raw device address as a torch tensor (no copy) via __cuda_array_interface__."""

    def __init__(self, ptr, n, typestr):
        self.__cuda_array_interface__ = {"shape": (int(n),), "typestr": typestr, "data": (int(ptr), False), "version": 3,
                                         "strides": None}


def device_view(ptr, n, dtype, device):
    typestr = {torch.int64: "<i8", torch.int32: "<i4"}[dtype]
    return torch.as_tensor(_DeviceArray(ptr, n, typestr), device=device)


class PeerCsrAssembler:
    """Whole-graph CSR on every rank, written by the ranks themselves over NVLink peer memory.

    Per connect: all-gather of the edge counts (8 bytes per rank; it also tells the peers that this rank
    is done reading the previous graph), sspp_gpu_connect_push_csr on the engine's stream, and a MAX
    all-reduce of the overflow word as the closing barrier (after it every rank's pushes have landed).
    The column buffers are sized on the first call and re-made, on all ranks alike, if a graph outgrows
    them (every rank sees the same total, so the decision needs no extra message)."""

    def __init__(self, eng, n_nodes, device, cuts=None, kind="connect"):
        """kind: "connect" (rows = way-points, items = adjacency) or "visibility" (rows = viewpoints, items = visible
        point indices: the same exchange on (view_ptr, point_idx), SURVEY 8 e).  cuts: world + 1 row cut points."""
        assert kind in ("connect", "visibility")
        self.kind = kind
        self.eng = eng
        self.world = dist.get_world_size()
        self.rank = dist.get_rank()
        self.device = device
        self.n = n_nodes
        base = row_cuts(n_nodes, self.world, cuts)
        self.rows = (base[self.rank], base[self.rank + 1])
        self.row_base = torch.tensor(base, dtype=torch.int64, device=device)
        self.ext_stream = torch.cuda.ExternalStream(eng.stream(), device=device)
        self.my_count = torch.zeros(1, dtype=torch.int64, device=device)
        self.counts = torch.zeros(self.world, dtype=torch.int64, device=device)
        self.overflow = torch.zeros(1, dtype=torch.int64, device=device)
        self.rp_ptr, self.rp_peers = self._exchange((n_nodes + 1) * 8)
        self.out_rp = device_view(self.rp_ptr, n_nodes + 1, torch.int64, device)
        self.col_cap = 0
        self.col_ptr, self.col_peers, self.out_col = None, None, None

    def _exchange(self, nbytes):
        return exchange_peer_buffer(self.eng, nbytes, self.device)

    def _release(self, ptr, peers):
        torch.cuda.synchronize()
        for g, q in enumerate(peers):
            if g != self.rank:
                self.eng.peer_buffer_release(q)
        dist.barrier()  # every importer has closed its mapping before the owner frees the memory
        self.eng.peer_buffer_release(ptr)

    def _size_columns(self, total):
        if self.col_ptr is not None:
            self._release(self.col_ptr, self.col_peers)
        self.col_cap = ((int(total * 1.1) + 1024 + 3) // 4) * 4
        self.col_ptr, self.col_peers = self._exchange(self.col_cap * 4)
        self.out_col = device_view(self.col_ptr, self.col_cap, torch.int32, self.device)

    def assemble(self, stats):
        """stats: the ConnectStats of this rank's sspp_gpu_connect (or the VisibilityStats of its
        sspp_gpu_visibility).  Returns (row_ptr, col_idx) device tensors of the whole graph (col_idx is the
        buffer; row_ptr[-1] entries of it are valid)."""
        cur = torch.cuda.current_stream()
        push = self.eng.connect_push_csr if self.kind == "connect" else self.eng.visibility_push
        self.my_count.fill_(int(stats.edges if self.kind == "connect" else stats.visible_pairs))
        dist.all_gather_into_tensor(self.counts, self.my_count)
        if self.col_cap == 0:
            self._size_columns(int(self.counts.sum().item()))
        for attempt in range(2):
            self.ext_stream.wait_event(cur.record_event())
            push(self.world, self.rank, self.counts.data_ptr(), self.row_base.data_ptr(), self.rp_peers, self.col_peers, self.col_cap,
                 self.overflow.data_ptr())
            cur.wait_event(self.ext_stream.record_event())
            dist.all_reduce(self.overflow, op=dist.ReduceOp.MAX)
            over = int(self.overflow.item())
            if over == 0:
                return self.out_rp, self.out_col
            self.overflow.zero_()
            self._size_columns(over)
        raise RuntimeError("CSR assembly: column buffers still too small after resizing")

    def close(self):
        if self.col_ptr is not None:
            self._release(self.col_ptr, self.col_peers)
            self.col_ptr = None
        if self.rp_ptr is not None:
            self._release(self.rp_ptr, self.rp_peers)
            self.rp_ptr = None


class StreamedPeerAssembler:
    """connect + whole-graph CSR on every rank, the exchange overlapped with the edge test.

    Every rank owns two sets of block buffers (row_ptr blocks of stride_r int64 per rank, column blocks
    of stride_e int32 per rank) that its peers can address; the sets alternate from call to call, so a
    rank may already be writing the next graph's chunks into its peers while they still pack the last
    one (the barrier of the call in between orders the reuse of a set)."""

    def __init__(self, eng, n_nodes, device, cuts=None):
        self.eng = eng
        self.world = dist.get_world_size()
        self.rank = dist.get_rank()
        self.device = device
        self.n = n_nodes
        base = row_cuts(n_nodes, self.world, cuts)
        self.rows = (base[self.rank], base[self.rank + 1])
        self.row_base = torch.tensor(base, dtype=torch.int64, device=device)
        self.stride_r = max(base[g + 1] - base[g] for g in range(self.world)) + 1
        self.ext_stream = torch.cuda.ExternalStream(eng.stream(), device=device)
        self.out_rp = torch.empty(n_nodes + 1, dtype=torch.int64, device=device)
        self.out_col = None
        self.overflow = torch.zeros(1, dtype=torch.int64, device=device)
        self.need = torch.zeros(1, dtype=torch.int64, device=device)
        self.stride_e = 0
        self.sets = []  # [(rp_ptr, rp_peers, col_ptr, col_peers)] x 2
        self.turn = 0

    def _exchange(self, nbytes):
        return exchange_peer_buffer(self.eng, nbytes, self.device)

    def _release_sets(self):
        torch.cuda.synchronize()
        for rp_ptr, rp_peers, col_ptr, col_peers in self.sets:
            for g in range(self.world):
                if g != self.rank:
                    self.eng.peer_buffer_release(rp_peers[g])
                    self.eng.peer_buffer_release(col_peers[g])
        dist.barrier()  # every importer has closed its mappings before the owners free the memory
        for rp_ptr, rp_peers, col_ptr, col_peers in self.sets:
            self.eng.peer_buffer_release(rp_ptr)
            self.eng.peer_buffer_release(col_ptr)
        self.sets = []

    def _size(self, edges_max):
        self._release_sets()
        self.stride_e = ((int(edges_max * 1.1) + 1024 + 3) // 4) * 4
        for _ in range(2):
            rp_ptr, rp_peers = self._exchange(self.world * self.stride_r * 8)
            col_ptr, col_peers = self._exchange(self.world * self.stride_e * 4)
            self.sets.append((rp_ptr, rp_peers, col_ptr, col_peers))
        self.out_col = torch.empty(self.world * self.stride_e, dtype=torch.int32, device=self.device)
        dist.barrier()

    def connect(self, **kw):
        """This rank's rows of sspp_gpu_connect + the exchange.  Returns (stats, row_ptr, col_idx): the
        stats of this rank's block and device tensors of the whole graph."""
        from .engine import SsppError
        if self.stride_e == 0:  # first call: learn the block sizes from a plain pass
            st = self.eng.connect(rows=self.rows, **kw)
            self.need.fill_(int(st.edges))
            dist.all_reduce(self.need, op=dist.ReduceOp.MAX)
            self._size(int(self.need.item()))
        for attempt in range(3):
            rp_ptr, rp_peers, col_ptr, col_peers = self.sets[self.turn % 2]
            self.turn += 1
            try:
                st = self.eng.connect_to_peers(self.world, self.rank, rp_peers, col_peers, self.stride_r, self.stride_e,
                                               rows=self.rows, **kw)
                edges = int(st.edges)
            except SsppError as e:
                if e.code != 4:
                    raise
                st, edges = None, e.needed_edges
            # barrier: every rank's copies have landed (connect_to_peers returns after its own have);
            # the same message tells all ranks whether some block did not fit its slot
            self.need.fill_(edges)
            dist.all_reduce(self.need, op=dist.ReduceOp.MAX)
            need = int(self.need.item())
            if need > self.stride_e:
                self._size(need)
                continue
            self.eng.assemble_csr_blocks(rp_ptr, col_ptr, self.world, self.stride_r, self.stride_e, self.row_base.data_ptr(),
                                         self.out_rp.data_ptr(), self.out_col.data_ptr(), self.overflow.data_ptr())
            self.eng.synchronize()
            return st, self.out_rp, self.out_col
        raise RuntimeError("CSR assembly: block slots still too small after resizing")

    def close(self):
        self._release_sets()


class FusedUnavailable(RuntimeError):
    """The fused exchange cannot serve this graph (a row outgrows the slot, or the pass runs without a map).  Raised on
    EVERY rank together, so the callers can switch to PeerCsrAssembler without a collective mismatch."""


class FusedPeerAssembler:
    """connect + whole-graph CSR on every rank with the exchange fused INTO the edge kernel (sspp_gpu_connect_fused):
    every rank owns two slot-format images of the whole graph (rows x slot int32 + rows int32 degrees) that its peers
    can address; the edge kernel's warps store each finished row into all ranks' images over NVLink while the pass is
    still running, so when the pass ends nothing is left to send.  What follows is one all-reduce (the barrier, and the
    edge total / "did every rank succeed" in the same message) and the local packing of the image into the CSR.
    The two images alternate from call to call: a rank may already be writing the next graph into its peers while they
    still pack the last one (the barrier of the call in between orders the reuse)."""

    def __init__(self, eng, n_nodes, device, cuts=None, slot=32):
        self.eng = eng
        self.world = dist.get_world_size()
        self.rank = dist.get_rank()
        self.device = device
        self.n = n_nodes
        self.slot = slot
        base = row_cuts(n_nodes, self.world, cuts)
        self.rows = (base[self.rank], base[self.rank + 1])
        self.ext_stream = torch.cuda.ExternalStream(eng.stream(), device=device)
        self.out_rp = torch.empty(n_nodes + 1, dtype=torch.int64, device=device)
        self.out_col, self.col_cap = None, 0
        self.overflow = torch.zeros(1, dtype=torch.int64, device=device)
        self.msg = torch.zeros(2, dtype=torch.int64, device=device)  # [ranks that failed, edges]
        self.sets = []
        for _ in range(2):
            sl_ptr, sl_peers = exchange_peer_buffer(eng, max(n_nodes, 1) * slot * 4, device)
            dg_ptr, dg_peers = exchange_peer_buffer(eng, max(n_nodes, 1) * 4, device)
            self.sets.append((sl_ptr, sl_peers, dg_ptr, dg_peers))
        self.turn = 0
        dist.barrier()

    def connect(self, **kw):
        """This rank's rows of sspp_gpu_connect_fused + barrier + packing.  Returns (stats, row_ptr, col_idx): the stats
        of this rank's shard and device tensors of the whole graph (col_idx is the buffer; row_ptr[-1] entries are valid)."""
        from .engine import SsppError
        for attempt in range(2):
            sl_ptr, sl_peers, dg_ptr, dg_peers = self.sets[self.turn % 2]
            self.turn += 1
            st, failed = None, 0
            try:
                st = self.eng.connect_fused(self.world, self.rank, self.slot, self.rows[0], sl_peers, dg_peers, rows=self.rows, **kw)
            except SsppError as e:
                if e.code not in (1, 4):
                    raise
                failed = 1
            # barrier: every rank's rows have landed (connect_fused returns after its own stores have retired); the same
            # message carries the edge total and whether some rank could not run the fused pass
            self.msg[0] = failed
            self.msg[1] = int(st.edges) if st is not None else 0
            dist.all_reduce(self.msg, op=dist.ReduceOp.SUM)
            nfail, total = int(self.msg[0].item()), int(self.msg[1].item())
            if nfail:
                raise FusedUnavailable("%d rank(s) could not run the fused pass (a row outgrows the %d-pair slot?)" % (nfail, self.slot))
            if total > self.col_cap:
                self.col_cap = ((int(total * 1.1) + 1024 + 3) // 4) * 4
                self.out_col = torch.empty(self.col_cap, dtype=torch.int32, device=self.device)
            self.ext_stream.wait_event(torch.cuda.current_stream().record_event())
            self.eng.pack_slots(sl_ptr, dg_ptr, self.n, self.slot, self.out_rp.data_ptr(), self.out_col.data_ptr(), self.col_cap,
                                self.overflow.data_ptr())
            self.eng.synchronize()
            return st, self.out_rp, self.out_col
        raise RuntimeError("unreachable")

    def close(self):
        torch.cuda.synchronize()
        for sl_ptr, sl_peers, dg_ptr, dg_peers in self.sets:
            for g in range(self.world):
                if g != self.rank:
                    self.eng.peer_buffer_release(sl_peers[g])
                    self.eng.peer_buffer_release(dg_peers[g])
        dist.barrier()  # every importer has closed its mappings before the owners free the memory
        for sl_ptr, sl_peers, dg_ptr, dg_peers in self.sets:
            self.eng.peer_buffer_release(sl_ptr)
            self.eng.peer_buffer_release(dg_ptr)
        self.sets = []
