"""Per-kernel timing table (the reference's collectPerKernelTimingInfo output, backend_cuda.t:262-297) of one bench
workload: python tools/timing_table.py <workload> [--kind LMGPU] [--double]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402,F401
import bench  # noqa: E402
from opt_b200 import solver as S  # noqa: E402

name = sys.argv[1]
pr, w, _ = bench.make_problem(name)
if "--kind" in sys.argv:
    pr.kind = sys.argv[sys.argv.index("--kind") + 1]
if "--double" in sys.argv:
    pr = pr.as_double()
named = S.upload(pr)
keep = [v.clone() if hasattr(v, "clone") else v for v in named.values]
for timing in (False, True):        # first solve warms up, second prints the table on PlanFree
    for i in pr.unknowns:
        named.values[i].copy_(keep[i])
    sol = S.OptSolver(pr.dims, pr.energy, pr.kind, double=pr.double, verbosity=1 if timing else 0, timing=timing)
    sol.solve(pr.solver, named, [])
    torch.cuda.synchronize()
    sol.close()
