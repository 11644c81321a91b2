"""Shared helpers for the -m gpu parity tests: run the same Problem through the CUDA library
(via the C-ABI, opt_b200.solver.OptSolver) and through the CPU oracle."""
import numpy as np

from opt_b200 import solver as S
from oracle_api import OracleSolver

# BASELINE.json: per-iteration and final cost within 1e-4 (fp32) / 1e-9 (fp64) relative
COST_RTOL = {False: 1e-4, True: 1e-9}
# unknowns: max abs error relative to the largest |value| of that image
X_RTOL = {False: 2e-4, True: 1e-8}
# single operator applications (J^T F, J^T J p): relative to the vector's max-norm
OP_RTOL = {False: 2e-5, True: 1e-11}


def run_gpu(pr, timing=False):
    import torch
    named = S.upload(pr)
    sol = S.OptSolver(pr.dims, pr.energy, pr.kind, double=pr.double, verbosity=0, timing=timing)
    costs = []
    sol.solve(pr.solver, named, costs)
    torch.cuda.synchronize()
    xs = {i: named.values[i].cpu().numpy() for i in pr.unknowns}
    return costs, xs, sol, named


def run_oracle(pr, threads=4):
    q = pr.clone()
    o = OracleSolver(q.dims, q.energy, q.kind, double=q.double, threads=threads)
    o.set_params(**{k: (float(v) if isinstance(v, float) else int(v)) for k, v in q.solver.items()})
    o.solve(q.params)
    xs = {i: q.params[i] for i in q.unknowns}
    return o.cost_history(), xs, o


def assert_costs_close(gpu_costs, ref_costs, double, tol=None):
    assert len(gpu_costs) == len(ref_costs), (gpu_costs, ref_costs)
    tol = COST_RTOL[double] if tol is None else tol
    for k, (a, b) in enumerate(zip(gpu_costs, ref_costs)):
        assert np.isfinite(a) and abs(a - b) <= tol * abs(b), "cost after step %d: gpu %r oracle %r" % (k, a, b)


def assert_unknowns_close(gx, rx, double, tol=None):
    tol = X_RTOL[double] if tol is None else tol
    for i in gx:
        scale = np.abs(rx[i][np.isfinite(rx[i])]).max() + 1e-30
        err = np.abs(gx[i].reshape(-1) - rx[i].reshape(-1)).max()
        assert err <= tol * scale, "unknown image %d: max err %g scale %g" % (i, err, scale)
