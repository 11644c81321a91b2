"""CPU (gloo, world_size 2 and 3): the host-side slab logic -- partition, ghost-row convention and
the halo plan the engine implements -- reproduces the global arrays after an exchange."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from opt_b200 import examples, slab


def test_partition_covers_everything():
    for L in (7, 64, 513):
        for n in (1, 2, 3, 8):
            parts = slab.partition(L, n)
            assert parts[0][0] == 0 and parts[-1][1] == L
            assert all(parts[i][1] == parts[i + 1][0] for i in range(n - 1))
            assert max(e - s for s, e in parts) - min(e - s for s, e in parts) <= 1


def test_split_problem_shapes_and_contents():
    pr = examples.poisson_synthetic(16, 21)
    for n in (2, 3):
        for r in range(n):
            loc = slab.split_problem(pr, r, n, halo=1)
            g0, g1, o_lo, o_hi = slab.local_range(21, r, n, 1)
            assert loc.dims == [16, g1 - g0]
            np.testing.assert_array_equal(loc.params[0].reshape(g1 - g0, -1), pr.params[0].reshape(21, -1)[g0:g1])
            s, e = slab.partition(21, n)[r]
            assert (o_hi - o_lo) == e - s


def _worker(rank, world, port, halo, L):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    rng = np.random.default_rng(5)
    glob = rng.standard_normal((L, 6)).astype(np.float32)          # a global field, identical on all ranks
    g0, g1, o_lo, o_hi = slab.local_range(L, rank, world, halo)
    loc = torch.from_numpy(glob[g0:g1].copy())
    loc[:o_lo] = -1.0                                               # ghost rows start out stale
    loc[o_hi:] = -1.0
    reqs = []
    for peer, (a, b), (c, d) in slab.halo_plan(g1 - g0, rank, world, halo):
        reqs.append(dist.isend(loc[a:b].clone(), peer))
        buf = torch.empty_like(loc[c:d])
        reqs.append(dist.irecv(buf, peer))
        reqs.append((c, d, buf))
    for r in reqs:
        if not isinstance(r, tuple):
            r.wait()
    for r in reqs:
        if isinstance(r, tuple):
            loc[r[0]:r[1]] = r[2]
    ok = torch.tensor([float(np.array_equal(loc.numpy(), glob[g0:g1]))])
    dist.all_reduce(ok, op=dist.ReduceOp.MIN)
    # all-reduced scalars: the sum of per-rank owned dot products equals the global dot product
    part = torch.tensor([float((glob[g0:g1][o_lo:o_hi].astype(np.float64) ** 2).sum())], dtype=torch.float64)
    dist.all_reduce(part)
    ok2 = abs(part.item() - float((glob.astype(np.float64) ** 2).sum())) < 1e-9 * part.item()
    dist.destroy_process_group()
    assert ok.item() == 1.0 and ok2


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


@pytest.mark.parametrize("world,halo", [(2, 1), (3, 2)])
def test_halo_plan_reassembles_global_field(world, halo):
    mp.spawn(_worker, args=(world, _free_port(), halo, 23), nprocs=world, join=True)
