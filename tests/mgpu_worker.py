"""Worker for tests/test_multigpu.py (launched by torch.distributed.run, one rank per GPU): solves a
grid problem slab-decomposed over all ranks and checks per-step costs and the gathered unknowns against
  (a) the CPU ORACLE's solve of the undecomposed problem (BASELINE gates: 1e-4 fp32 / 1e-9 fp64 on every per-step cost), and
  (b) the single-GPU CUDA solve of the undecomposed problem (tighter: same arithmetic, different reduction order),
both computed on rank 0.  The cases are thin enough that interior ranks (two neighbours) exist at world sizes 4 and 8.
OPTB200_NO_P2P=1 in the environment runs the NCCL-only data plane; MGPU_CASES=a,b restricts the case list."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from opt_b200 import examples, slab, solver as S  # noqa: E402
from gpu_util import COST_RTOL, X_RTOL, run_oracle  # noqa: E402

# (builder, cost tolerance, halo rows).  optical_flow samples I_hat at x + X(x): its halo must cover the largest
# flow vector (data-dependent stencil radius), the pairwise-difference energies need one row.
CASES = {
    "poisson": (lambda: examples.poisson_synthetic(64, 96, n_iter=2, l_iter=15), 1e-5, 1),
    # two strips (one ragged) of the marching kernel, fused with PCGStep3 in slab mode (ghost rows from the exchanged p'), LM
    "poisson_lm": (lambda: examples.poisson_synthetic(148, 90, n_iter=2, l_iter=12), 1e-5, 1),
    "image_warping": (lambda: examples.image_warping_synthetic(48, 64, n_iter=3, l_iter=10), 1e-4, 1),
    "optical_flow_lm": (lambda: examples.optical_flow_synthetic(40, 60, n_iter=3, l_iter=12, kind="LMGPU"), 1e-4, 4),
    "volumetric": (lambda: examples.volumetric_synthetic(12, 10, 16, n_iter=2, l_iter=10), 1e-5, 1),
    # rows of 32 sites: blocked solver vectors + packed kernel, ghost planes exchanged as one contiguous slice
    "volumetric_packed": (lambda: examples.volumetric_synthetic(32, 11, 20, n_iter=2, l_iter=10), 1e-5, 1),
    "volumetric_packed_lm": (lambda: examples.volumetric_synthetic(64, 6, 12, n_iter=3, l_iter=8), 1e-5, 1),
    # shape_from_shading: 5-point regulariser and B_I differences -> halo 2; global row numbers in the bounds
    # predicates and the back-projection rays.  Rows of whole quads take the marching kernels, the other case the
    # generic functor kernels.
    "sfs": (lambda: examples.sfs_synthetic(64, 44, n_iter=2, l_iter=8), 1e-4, 2),
    "sfs_lm": (lambda: examples.sfs_synthetic(45, 40, n_iter=3, l_iter=8), 1e-4, 2),
    # re-weighted L_p prior: rounding differences between the two reduction orders are amplified by the IRLS weights
    # (DESIGN.md section 4), so the comparison is made in fp64
    "intrinsic_f64": (lambda: examples.intrinsic_synthetic(40, 36, n_iter=3, l_iter=8).as_double(), 1e-7, 1),
    "volumetric_lm_f64": (lambda: examples.volumetric_synthetic(8, 6, 12, n_iter=3, l_iter=12).as_double(), 1e-9, 1),
}


def solve(pr, rank=0, world=1, halo=1):
    named = S.upload(pr)
    sol = S.OptSolver(pr.dims, pr.energy, pr.kind, double=pr.double)
    if world > 1:
        slab.setup_slab(sol, rank, world, halo)
    costs = []
    sol.solve(pr.solver, named, costs)
    torch.cuda.synchronize()
    xs = [named.values[i].cpu().numpy() for i in pr.unknowns]
    sol.close()
    return costs, xs


def main():
    dist.init_process_group("nccl")
    rank, world = dist.get_rank(), dist.get_world_size()
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", rank)))
    failures = []
    only = [c for c in os.environ.get("MGPU_CASES", "").split(",") if c]
    for name, (build, tol, halo) in CASES.items():
        if only and name not in only:
            continue
        pr = build()
        if "lm" in name:
            pr.kind = "LMGPU"
        L = pr.dims[-1]
        loc = slab.split_problem(pr, rank, world, halo)
        costs, xs = solve(loc, rank, world, halo)
        owned = [slab.owned_rows(x, loc.dims[-1], rank, world, halo) for x in xs]
        gathered = []
        for o in owned:
            lst = [None] * world
            dist.all_gather_object(lst, o)
            gathered.append(np.concatenate(lst, 0))
        if rank == 0:
            ref_costs, ref_xs = solve(pr)
            ok = len(costs) == len(ref_costs) and all(abs(a - b) <= tol * abs(b) for a, b in zip(costs, ref_costs))
            for g, rx in zip(gathered, ref_xs):
                ref = rx.reshape(L, -1)
                scale = np.abs(ref[np.isfinite(ref)]).max()
                ok = ok and np.abs(g - ref).max() <= 10 * tol * scale
            # (a) the oracle on the undecomposed problem
            o_costs, o_xs, _ = run_oracle(pr)
            ctol, xtol = COST_RTOL[pr.double], X_RTOL[pr.double]
            if name.startswith("intrinsic"):
                ctol, xtol = 1e-7, 1e-6                 # re-weighted L_p prior (tests/test_parity_gpu.py)
            ok_o = len(costs) == len(o_costs) and all(abs(a - b) <= ctol * abs(b) for a, b in zip(costs, o_costs))
            for g, i in zip(gathered, pr.unknowns):
                ref = o_xs[i].reshape(L, -1)
                scale = np.abs(ref[np.isfinite(ref)]).max()
                ok_o = ok_o and np.abs(g - ref).max() <= xtol * scale
            print("mgpu %-20s world=%d p2p=%s vs_single_gpu=%s vs_oracle=%s costs=%s oracle=%s" % (
                name, world, "off" if os.environ.get("OPTB200_NO_P2P") else "on", " OK " if ok else "MISMATCH", " OK " if ok_o else "MISMATCH", costs, o_costs), flush=True)
            if not (ok and ok_o):
                failures.append(name)
        dist.barrier()
    dist.destroy_process_group()
    if failures:
        sys.exit(1)


if __name__ == "__main__":
    main()
