"""Row-sharded graph build over the GPUs of one box (SURVEY.md §8 e).

The path shards by source row: rank g owns the contiguous list-index block [n*g/G, n*(g+1)/G) of the
search space, with the map and the node set replicated; no collective touches the data path.  The
one exchange step is the assembly of the CSR: an all-gather of (rows, edges) per rank, then of the
padded row_ptr / col_idx blocks (NCCL has no all-gather-v), then every rank rebases the row pointers
by the running edge count.  Works on any torch.distributed backend (nccl on GPUs, gloo in the tests).
"""
import torch
import torch.distributed as dist


def row_range(n, rank, world):
    """Contiguous block of list indices owned by `rank`."""
    return (n * rank) // world, (n * (rank + 1)) // world


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
    column slot size is learned on the first call (one count all-gather) and re-learned if a block
    ever outgrows it."""

    def __init__(self, eng, n_nodes, device):
        self.eng = eng
        self.world = dist.get_world_size()
        self.rank = dist.get_rank()
        self.device = device
        base = [row_range(n_nodes, g, self.world)[0] for g in range(self.world)] + [n_nodes]
        self.row_base = torch.tensor(base, dtype=torch.int64, device=device)
        self.stride_r = max(base[g + 1] - base[g] for g in range(self.world)) + 1
        self.n = n_nodes
        self.stride_e = 0
        self.ext_stream = torch.cuda.ExternalStream(eng.stream(), device=device)
        self.send_rp = torch.zeros(self.stride_r, dtype=torch.int64, device=device)
        self.all_rp = torch.empty(self.world * self.stride_r, dtype=torch.int64, device=device)
        self.out_rp = torch.empty(n_nodes + 1, dtype=torch.int64, device=device)
        self.overflow = torch.zeros(1, dtype=torch.int64, device=device)
        self.send_col = self.all_col = self.out_col = None

    def _size_columns(self, edges):
        counts = gather_counts(0, edges, self.device)
        self.stride_e = max(int(int(counts[:, 1].max()) * 1.1) + 1024, 1)
        self.send_col = torch.empty(self.stride_e, dtype=torch.int32, device=self.device)
        self.all_col = torch.empty(self.world * self.stride_e, dtype=torch.int32, device=self.device)
        self.out_col = torch.empty(self.world * self.stride_e, dtype=torch.int32, device=self.device)

    def assemble(self, stats):
        """stats: the ConnectStats of this rank's sspp_gpu_connect.  Returns (row_ptr, col_idx) device
        tensors of the whole graph (col_idx is a view of the first E entries' buffer)."""
        if self.stride_e == 0 or stats.edges > self.stride_e:
            self._size_columns(int(stats.edges))
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
            # some rank's block outgrew its slot (all ranks see the same flag): re-learn the slot size
            self.overflow.zero_()
            self._size_columns(int(stats.edges))
        raise RuntimeError("CSR assembly: column slots still too small after resizing")
