"""debug: packed / scalar / generic J^T J p of volumetric on one grid; prints where they differ"""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from opt_b200 import examples, solver as S
dims = tuple(int(v) for v in sys.argv[1:4]) if len(sys.argv) > 3 else (64, 19, 37)
pr = examples.volumetric_synthetic(*dims, n_iter=2, l_iter=8)
rng = np.random.default_rng(3)
N = int(np.prod(pr.dims))
p = rng.standard_normal((N, 6)).astype(np.float32)
def run(env):
    for k in ("OPTB200_VOL_SCALAR", "OPTB200_GENERIC_JTJ"):
        os.environ.pop(k, None)
    if env: os.environ[env] = "1"
    sol = S.OptSolver(pr.dims, pr.energy, pr.kind)
    return sol.debug_apply_jtj(S.upload(pr), p).reshape(dims[2], dims[1], dims[0], 6)
for rep in range(2):
    pk, sc, ge = run(None), run("OPTB200_VOL_SCALAR"), run("OPTB200_GENERIC_JTJ")
    for name, a in (("packed", pk), ("scalar", sc)):
        d = np.abs(a - ge).max(-1)
        bad = np.argwhere(d > 2e-5 * np.abs(ge).max())
        print(rep, name, "max err", d.max(), "nbad", len(bad), "z", sorted(set(bad[:, 0]))[:12], "y", sorted(set(bad[:, 1]))[:12], "x", sorted(set(bad[:, 2]))[:12])
