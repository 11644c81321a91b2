"""Prints, per test case, the per-step costs of gpu-fp32, oracle-fp32 and oracle-fp64 so the fp32
tolerance of a case can be judged against the fp32 noise floor of the cost itself."""
import sys, os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
from gpu_util import run_gpu, run_oracle
import test_parity_gpu as T

for case in sorted(T.CASES):
    for kind in ("gaussNewtonGPU", "LMGPU"):
        pr = T.make(case, kind, False)
        g32 = run_gpu(pr)[0]
        o32 = run_oracle(pr)[0]
        o64 = run_oracle(T.make(case, kind, True))[0]
        g64 = run_gpu(T.make(case, kind, True))[0]
        n = min(len(g32), len(o32), len(o64))
        rel = lambda a, b: max(abs(a[i] - b[i]) / abs(b[i]) for i in range(n))
        print("%-16s %-14s steps %d/%d/%d  gpu32-vs-or32 %.2e  gpu32-vs-or64 %.2e  or32-vs-or64 %.2e  gpu64-vs-or64 %.2e  last %.6g" %
              (case, kind, len(g32), len(o32), len(o64), rel(g32, o32), rel(g32, o64), rel(o32, o64), rel(g64, o64), o64[-1]))
