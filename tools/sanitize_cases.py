"""Small single-GPU solves that exercise every asynchronous-copy / barrier protocol of the engine, for
`compute-sanitizer --tool memcheck|racecheck python tools/sanitize_cases.py` (tools/run_r2i.sh): the cp.async rings of the
poisson and shape_from_shading marching kernels, the mbarrier ring (bulk copies + 16-byte cp.async) of the packed volumetric
kernel, fused and unfused, the last-block counters, the segmented graph scatter, the ARAP gather and the CSR kernels."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import torch  # noqa: E402
from opt_b200 import examples, solver as S  # noqa: E402

CASES = [
    ("volumetric_packed", lambda: examples.volumetric_synthetic(64, 19, 37, n_iter=2, l_iter=4), {}),
    ("volumetric_packed_lm", lambda: examples.volumetric_synthetic(32, 11, 9, n_iter=2, l_iter=4), {"kind": "LMGPU"}),
    ("volumetric_scalar", lambda: examples.volumetric_synthetic(11, 9, 7, n_iter=1, l_iter=3), {}),
    ("poisson_strips", lambda: examples.poisson_synthetic(148, 83, n_iter=1, l_iter=4), {}),
    ("sfs_quads", lambda: examples.sfs_synthetic(132, 37, n_iter=2, l_iter=3), {}),
    ("sfs_quads_lm", lambda: examples.sfs_synthetic(132, 37, n_iter=2, l_iter=3), {"kind": "LMGPU"}),
    ("optical_flow", lambda: examples.optical_flow_synthetic(53, 38, n_iter=1, l_iter=3), {}),
    ("arap", lambda: examples.arap_synthetic(13, 11, n_iter=1, l_iter=3), {}),
    ("embedded", lambda: examples.embedded_synthetic(12, 10, n_iter=1, l_iter=3), {}),
    ("cotangent", lambda: examples.cotangent_synthetic(12, 10, n_iter=1, l_iter=3), {}),
    ("poisson_materialized", lambda: examples.poisson_synthetic(67, 45, n_iter=1, l_iter=3), {"materialized": True}),
    ("arap_fused_jtj", lambda: examples.arap_synthetic(13, 11, n_iter=1, l_iter=3), {"materialized": True, "fused_jtj": True}),
]


def main():
    only = sys.argv[1:]
    for name, build, opt in CASES:
        if only and name not in only:
            continue
        pr = build()
        pr.kind = opt.get("kind", pr.kind)
        named = S.upload(pr)
        sol = S.OptSolver(pr.dims, pr.energy, pr.kind, materialized=opt.get("materialized", False), fused_jtj=opt.get("fused_jtj", False))
        cost = sol.solve(pr.solver, named)
        torch.cuda.synchronize()
        sol.close()
        print("sanitize %-22s final cost %.9g" % (name, cost), flush=True)


if __name__ == "__main__":
    main()
