"""Host-side helpers for the multi-GPU slab mode (include/OptB200.h, "multi-GPU domain decomposition").

A grid problem is cut along its slowest dimension (d1 for 2-D, d2 for 3-D; the reference's CPU-MT
backend splits the same dimension across threads, backend_cpu_mt.t:721-740).  Rank r owns the
rows [start_r, end_r) of the global grid and its local arrays additionally carry `halo` ghost rows
on every side that faces another rank.
"""
import ctypes as C

import numpy as np

from .examples import Problem


def partition(L, nranks):
    """Owned [start, end) of every rank: L rows split as evenly as possible, in rank order."""
    base, rem = divmod(L, nranks)
    out, s = [], 0
    for r in range(nranks):
        e = s + base + (1 if r < rem else 0)
        out.append((s, e))
        s = e
    return out


def local_range(L, rank, nranks, halo):
    """(global first row, global end row, o_lo, o_hi) of rank's local array: ghost rows included in
    the global range, o_lo/o_hi = owned range in local coordinates."""
    s, e = partition(L, nranks)[rank]
    g0 = s - (halo if rank > 0 else 0)
    g1 = e + (halo if rank < nranks - 1 else 0)
    return g0, g1, s - g0, e - g0


def halo_plan(L_local, rank, nranks, halo):
    """[(peer, send rows [a,b), recv rows [c,d))] in LOCAL coordinates -- the convention the engine's
    exchange_raw implements (opt_b200/csrc/engine.cuh)."""
    o_lo = halo if rank > 0 else 0
    o_hi = L_local - halo if rank < nranks - 1 else L_local
    plan = []
    if rank > 0:
        plan.append((rank - 1, (o_lo, o_lo + halo), (o_lo - halo, o_lo)))
    if rank < nranks - 1:
        plan.append((rank + 1, (o_hi - halo, o_hi), (o_hi, o_hi + halo)))
    return plan


def split_problem(pr, rank, nranks, halo=1):
    """Local Problem of `rank`: every image sliced to its slab (+ghost rows), dims adjusted."""
    L = pr.dims[-1]
    g0, g1, _, _ = local_range(L, rank, nranks, halo)
    params = []
    for p in pr.params:
        if isinstance(p, np.ndarray):
            flat = p.reshape(L, -1)
            params.append(np.ascontiguousarray(flat[g0:g1]).reshape((g1 - g0,) + tuple(_tail_shape(p, L))))
        else:
            params.append(p)
    dims = list(pr.dims)
    dims[-1] = g1 - g0
    return Problem(pr.energy, dims, params, list(pr.unknowns), dict(pr.solver), pr.kind, pr.double)


def _tail_shape(p, L):
    n = p.size // L
    return (n,)


def owned_rows(local_array, L_local, rank, nranks, halo):
    """The owned rows of a local array as a (rows, -1) view."""
    o_lo = halo if rank > 0 else 0
    o_hi = L_local - halo if rank < nranks - 1 else L_local
    return local_array.reshape(L_local, -1)[o_lo:o_hi]


def setup_slab(solver, rank, nranks, halo=1, group=None):
    """Collective: rank 0 creates the NCCL unique id inside the library, torch.distributed carries
    the 128 bytes to the other ranks, every rank calls OptB200_PlanSetSlab."""
    import torch
    import torch.distributed as dist
    lib = solver.lib
    lib.OptB200_GetNcclUniqueId.argtypes = [C.c_void_p]
    lib.OptB200_PlanSetSlab.restype = C.c_int
    lib.OptB200_PlanSetSlab.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p]
    buf = (C.c_ubyte * 128)()
    if nranks > 1:
        if rank == 0:
            lib.OptB200_GetNcclUniqueId(buf)
        dev = "cuda" if dist.get_backend(group) == "nccl" else "cpu"
        t = torch.tensor(list(buf), dtype=torch.uint8, device=dev)
        dist.broadcast(t, src=0, group=group)
        for i, v in enumerate(t.cpu().tolist()):
            buf[i] = v
    rc = lib.OptB200_PlanSetSlab(solver.plan, rank, nranks, halo, buf)
    if rc != 0:
        raise RuntimeError("OptB200_PlanSetSlab failed")
