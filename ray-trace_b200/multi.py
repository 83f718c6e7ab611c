"""Multi-GPU host logic for the one-process-per-GPU launch (torchrun): rays shard by launch-list
partition, no data-path collective; the only exchange is one gather of terminal records to rank 0.

(The single-process form — one Julia/Python process driving every GPU of the box through the C ABI — is
rt_solve_*_multi / api.MultiRayTracer and needs none of this.)

Works on any torch.distributed backend (nccl on the GPU box; gloo in the CPU tests).  The gather posts one receive per
rank and record type with the exact partition size (no padding, no intermediate lists), into buffers that are allocated
once per (sweep size, world, layout) and reused, and places each rank's records in launch-list order with one indexed
copy per buffer; rank 0's own records are placed while the receives are in flight."""
import torch
import torch.distributed as dist

from .sweeps import block_cyclic_indices, partition, partition_indices  # noqa: F401  (re-exported)

DOUBLE_ROWS_2D = 4 + 2  # u_end[4], t_end, group_delay
DOUBLE_ROWS_3D = 7 + 2
INT_ROWS = 5  # status, naccept, nreject, nfe, nreflect


def alloc_terminal(n, dim, device):
    """Result arrays of one rank laid out as the two gather buffers: doubles [dim+2, n] (u_end rows, t_end, group_delay)
    and int32 [5, n] (status, naccept, nreject, nfe, nreflect).  The solver writes into row views of these, so that
    pack_terminal is free.  Returns (d, i, dict of rt_result member -> tensor view)."""
    d = torch.empty((dim + 2, n), dtype=torch.float64, device=device)
    i = torch.empty((INT_ROWS, n), dtype=torch.int32, device=device)
    views = {"u_end": d[:dim], "t_end": d[dim], "group_delay": d[dim + 1], "status": i[0], "naccept": i[1],
             "nreject": i[2], "nfe": i[3], "nreflect": i[4]}
    return d, i, views


def pack_terminal(u_end, t_end, group_delay, status, naccept, nreject, nfe, nreflect):
    """Terminal records of one rank as two tensors: doubles [dim+2, n] and int32 [5, n] (copies; see alloc_terminal
    for the copy-free layout)."""
    d = torch.cat([u_end, t_end[None, :], group_delay[None, :]], dim=0).contiguous()
    i = torch.stack([status, naccept, nreject, nfe, nreflect], dim=0).contiguous()
    return d, i


def _indices(n_total, rank, world, layout):
    if layout.startswith("blockcyclic"):
        block = int(layout.split(":")[1]) if ":" in layout else 4096
        return block_cyclic_indices(n_total, block, world, rank)
    return partition_indices(n_total, rank, world, layout)


class _Plan:
    """Receive buffers, placement indices and the assembled arrays of one (n_total, world, layout, shapes, device)."""

    def __init__(self, n_total, world, dst, layout, d, i):
        self.sizes, self.place = [], []
        self.n_total, self.world = n_total, world
        self.block = int(layout.split(":")[1]) if layout.startswith("blockcyclic:") else (4096 if layout == "blockcyclic" else 0)
        self.full = n_total // (world * self.block) if self.block else 0  # complete cycles of `world` blocks
        for r in range(world):
            idx = _indices(n_total, r, world, layout)
            self.sizes.append(int(idx.size))
            if self.block:
                # the complete cycles are a strided view of the assembled array (coalesced block copies); only the
                # rays of the last, incomplete cycle are placed by index
                tail = idx[self.full * self.block:]
                self.place.append(torch.from_numpy(tail).to(d.device))
            elif layout == "contiguous":
                lo = int(idx[0]) if idx.size else 0
                self.place.append(slice(lo, lo + int(idx.size)))
            elif layout == "strided":
                self.place.append(slice(r, n_total, world))
            else:
                self.place.append(torch.from_numpy(idx).to(d.device))
        self.dbuf = [d.new_empty((d.shape[0], s)) for s in self.sizes]
        self.ibuf = [i.new_empty((i.shape[0], s)) for s in self.sizes]
        self.D = d.new_empty((d.shape[0], n_total))
        self.I = i.new_empty((i.shape[0], n_total))

    def put(self, r, d, i):
        p = self.place[r]
        if self.block:
            nb = self.full * self.block
            if nb:
                cyc = self.full * self.world * self.block
                for src, dst_ in ((d, self.D), (i, self.I)):
                    rows = src.shape[0]
                    dst_[:, :cyc].view(rows, self.full, self.world, self.block)[:, :, r, :] = \
                        src[:, :nb].view(rows, self.full, self.block)
            if p.numel():
                self.D.index_copy_(1, p, d[:, nb:])
                self.I.index_copy_(1, p, i[:, nb:])
            return
        if isinstance(p, slice):
            self.D[:, p] = d
            self.I[:, p] = i
        else:
            self.D.index_copy_(1, p, d)
            self.I.index_copy_(1, p, i)


_PLANS = {}


def gather_terminal(d, i, n_total, rank, world, dst=0, group=None, layout="contiguous", timing=None):
    """Gather per-rank records (partitions of n_total rays, see sweeps.partition_indices / block_cyclic_indices;
    layout "contiguous" | "strided" | "rotated" | "blockcyclic[:B]") to rank `dst`, in launch-list order.
    Returns (doubles [rows, n_total], ints [5, n_total]) on dst — buffers owned by this module and overwritten by the
    next call with the same shapes — and (None, None) elsewhere.  timing (a dict, CUDA tensors only): on dst it receives
    the milliseconds spent waiting for the receives and placing the records (diagnostics; synchronises the device)."""
    if world == 1:
        return d, i
    if rank != dst:
        reqs = dist.batch_isend_irecv([dist.P2POp(dist.isend, d.contiguous(), dst, group),
                                       dist.P2POp(dist.isend, i.contiguous(), dst, group)])
        for q in reqs:
            q.wait()
        return None, None
    key = (n_total, world, dst, layout, tuple(d.shape), tuple(i.shape), str(d.device))
    plan = _PLANS.get(key)
    if plan is None:
        plan = _PLANS[key] = _Plan(n_total, world, dst, layout, d, i)
    ops = []
    for r in range(world):
        if r != dst:
            ops.append(dist.P2POp(dist.irecv, plan.dbuf[r], r, group))
            ops.append(dist.P2POp(dist.irecv, plan.ibuf[r], r, group))
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(3)] if (timing is not None and d.is_cuda) else None
    if ev:
        ev[0].record()
    reqs = dist.batch_isend_irecv(ops)
    plan.put(dst, d, i)  # own records while the others are in flight
    for q in reqs:
        q.wait()
    if ev:
        ev[1].record()
    for r in range(world):
        if r != dst:
            plan.put(r, plan.dbuf[r], plan.ibuf[r])
    if ev:
        ev[2].record()
        torch.cuda.synchronize()
        timing["receive_ms"] = ev[0].elapsed_time(ev[1])
        timing["place_ms"] = ev[1].elapsed_time(ev[2])
    return plan.D, plan.I
