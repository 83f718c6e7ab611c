"""Multi-GPU host logic: rays shard by contiguous launch-list partition, one process per GPU,
no data-path collective; the only exchange is one gather of terminal records to rank 0.

Works on any torch.distributed backend (nccl on the GPU box; gloo in the CPU tests)."""
import torch
import torch.distributed as dist

from .sweeps import partition, partition_indices  # noqa: F401  (re-exported)

DOUBLE_ROWS_2D = 4 + 2  # u_end[4], t_end, group_delay
DOUBLE_ROWS_3D = 7 + 2
INT_ROWS = 5  # status, naccept, nreject, nfe, nreflect


def pack_terminal(u_end, t_end, group_delay, status, naccept, nreject, nfe, nreflect):
    """Terminal records of one rank as two tensors: doubles [dim+2, n] and int32 [5, n]."""
    d = torch.cat([u_end, t_end[None, :], group_delay[None, :]], dim=0).contiguous()
    i = torch.stack([status, naccept, nreject, nfe, nreflect], dim=0).contiguous()
    return d, i


_ROT_CACHE = {}


def _rotated_indices(n_total, world, device):
    """Per-rank index tensors of the "rotated" layout on `device` (cached: the gather runs every step)."""
    if device is None:
        device = torch.device("cpu")
    key = (n_total, world, str(device))
    if key not in _ROT_CACHE:
        _ROT_CACHE[key] = [torch.from_numpy(partition_indices(n_total, r, world, "rotated")).to(device) for r in range(world)]
    return _ROT_CACHE[key]


def gather_terminal(d, i, n_total, rank, world, dst=0, group=None, layout="contiguous"):
    """Gather per-rank records (partitions of n_total rays, see sweeps.partition_indices) to rank `dst`, in launch-list
    order.  Returns (doubles [rows, n_total], ints [5, n_total]) on dst, (None, None) elsewhere."""
    if world == 1:
        return d, i
    idx = None
    if layout == "strided":
        sizes = [len(range(r, n_total, world)) for r in range(world)]
    elif layout == "rotated":
        idx = _rotated_indices(n_total, world, d.device if rank == dst else None)
        sizes = [int(partition_indices(n_total, r, world, "rotated").size) for r in range(world)] if idx is None else [int(t.numel()) for t in idx]
    else:
        sizes = [partition(n_total, r, world)[1] - partition(n_total, r, world)[0] for r in range(world)]
    nmax = max(sizes)

    def padded(x):
        if x.shape[1] == nmax:
            return x
        out = x.new_zeros((x.shape[0], nmax))
        out[:, : x.shape[1]] = x
        return out

    dp, ip = padded(d), padded(i)
    if rank == dst:
        dl = [torch.empty_like(dp) for _ in range(world)]
        il = [torch.empty_like(ip) for _ in range(world)]
    else:
        dl = il = None
    dist.gather(dp, dl, dst=dst, group=group)
    dist.gather(ip, il, dst=dst, group=group)
    if rank != dst:
        return None, None
    if layout == "strided":
        D = dp.new_empty((dp.shape[0], n_total))
        I = ip.new_empty((ip.shape[0], n_total))
        for r in range(world):
            D[:, r::world] = dl[r][:, : sizes[r]]
            I[:, r::world] = il[r][:, : sizes[r]]
        return D, I
    if layout == "rotated":
        D = dp.new_empty((dp.shape[0], n_total))
        I = ip.new_empty((ip.shape[0], n_total))
        for r in range(world):
            D.index_copy_(1, idx[r], dl[r][:, : sizes[r]])
            I.index_copy_(1, idx[r], il[r][:, : sizes[r]])
        return D, I
    D = torch.cat([dl[r][:, : sizes[r]] for r in range(world)], dim=1)
    I = torch.cat([il[r][:, : sizes[r]] for r in range(world)], dim=1)
    return D, I
