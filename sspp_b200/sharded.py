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
