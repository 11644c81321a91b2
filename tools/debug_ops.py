"""GPU debugging aid: compare r = -J^T F, the preconditioner and J^T J p of the CUDA library with the
oracle for one test case and print WHERE they differ (ASCII map for 2-D problems).
usage: python tools/debug_ops.py sfs [double]"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from opt_b200 import solver as S  # noqa: E402
from oracle_api import OracleSolver  # noqa: E402
import test_parity_gpu as tp  # noqa: E402

case = sys.argv[1]
double = len(sys.argv) > 2 and sys.argv[2] == "double"
pr = tp.make(case, double=double)
named = S.upload(pr)
sol = S.OptSolver(pr.dims, pr.energy, pr.kind, double=double)
o = OracleSolver(pr.dims, pr.energy, pr.kind, double=double, threads=4)
dt = np.float64 if double else np.float32
sol.debug_pcg_begin(named)
r_gpu, pre_gpu = sol.debug_read("r"), sol.debug_read("pre")
r_ref, pre_ref = o.eval_jtf(pr.params, dt)
rng = np.random.default_rng(7)
p = rng.standard_normal(r_ref.shape).astype(dt) * (r_ref != 0)
ap_gpu = sol.debug_apply_jtj(named, p)
ap_ref = o.apply_jtj(pr.params, p)
U = sol.unknown_channels()


def show(name, a, b):
    a = np.asarray(a).reshape(-1, U); b = np.asarray(b).reshape(-1, U)
    err = np.abs(a - b).max(1)
    scale = np.abs(b).max() + 1e-30
    bad = err > 1e-4 * scale
    print("%s: max err %.3e (scale %.3e), %d of %d sites differ" % (name, err.max(), scale, bad.sum(), len(bad)))
    if len(pr.dims) == 2 and bad.any():
        W, H = pr.dims
        m = bad.reshape(H, W)
        for y in range(H):
            print("".join("#" if m[y, x] else "." for x in range(W)))
        ys, xs = np.nonzero(m)
        for y, x in list(zip(ys, xs))[:6]:
            print("  (x=%d,y=%d) gpu %s ref %s" % (x, y, a[y * W + x], b[y * W + x]))


show("r", r_gpu, r_ref)
show("pre", pre_gpu, pre_ref)
show("Ap", ap_gpu, ap_ref)
