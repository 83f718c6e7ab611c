"""N > 1 host path on CPU: two gloo ranks each trace their contiguous partition (the oracle stands in for the GPU
solve — this test is about partitioning and the terminal-record gather, not the kernels) and rank 0's gathered
records must equal a single-partition run bit for bit (rays are independent)."""
import os
import socket
import sys

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as tmp

import oracle_lib as O

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, out_path, n_total, layout):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from raytrace_b200 import multi, sweeps
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    full = sweeps.sweep_3d(3, 7, 5)[:, :n_total]
    part = np.ascontiguousarray(full[:, multi._indices(n_total, rank, world, layout)])
    res = O.solve_batch(O.env(O.RAY3D), O.opts(reltol=1e-3, dtmax=0, dtmin=0, r_esc=10 * sweeps.RE), part, True, 1)
    T = torch.from_numpy
    d, i = multi.pack_terminal(T(res["u_end"]), T(res["t_end"]), T(res["u_end"][6].copy()), T(res["status"]),
                               T(res["naccept"]), T(res["nreject"]), T(res["nfe"]), T(res["nreflect"]))
    D, I = multi.gather_terminal(d, i, n_total, rank, world, layout=layout)
    D, I = multi.gather_terminal(d, i, n_total, rank, world, layout=layout)  # a second step reuses the plan's buffers
    if rank == 0:
        np.savez(out_path, D=D.numpy(), I=I.numpy())
    else:
        assert D is None and I is None
    dist.barrier()
    dist.destroy_process_group()


import pytest


@pytest.mark.parametrize("layout", ["contiguous", "strided", "rotated", "blockcyclic:8", "blockcyclic:64", "rootlight:3"])
def test_two_rank_partition_and_gather(tmp_path, layout):
    from raytrace_b200 import sweeps
    n_total = 101  # odd: ragged partitions 50 + 51
    out = str(tmp_path / "gathered.npz")
    tmp.spawn(_worker, args=(2, _free_port(), out, n_total, layout), nprocs=2, join=True)
    got = np.load(out)
    full = sweeps.sweep_3d(3, 7, 5)[:, :n_total]
    ref = O.solve_batch(O.env(O.RAY3D), O.opts(reltol=1e-3, dtmax=0, dtmin=0, r_esc=10 * sweeps.RE), full, True, 2)
    assert got["D"].shape == (9, n_total) and got["I"].shape == (5, n_total)
    assert np.array_equal(got["D"][:7], ref["u_end"], equal_nan=True)
    assert np.array_equal(got["D"][7], ref["t_end"]) and np.array_equal(got["D"][8], ref["u_end"][6], equal_nan=True)
    for k, name in enumerate(("status", "naccept", "nreject", "nfe", "nreflect")):
        assert np.array_equal(got["I"][k], ref[name]), name


def test_partition_edge_cases():
    from raytrace_b200 import multi
    assert [multi.partition(3, r, 8) for r in range(8)] == [(0, 0), (0, 0), (0, 1), (1, 1), (1, 1), (1, 2), (2, 2), (2, 3)]
    assert multi.partition(0, 0, 4) == (0, 0)
    d, i = torch.zeros((9, 4), dtype=torch.float64), torch.zeros((5, 4), dtype=torch.int32)
    D, I = multi.gather_terminal(d, i, 4, 0, 1)
    assert D is d and I is i
