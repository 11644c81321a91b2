"""Fatal errors (include/OptB200.h, OptB200_SetErrorHandler): with a handler installed a CUDA failure is reported to it and the
failing call returns, instead of the reference's print-and-exit (backend_cuda.t:26-40), which stays the default."""
import json
import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_handler_entry_points_load_without_a_gpu(built):
    """No compute: install and remove a handler, read the (empty) last error."""
    from opt_b200 import solver as S
    seen = []
    S.set_error_handler(lambda code, msg: seen.append((code, msg)))
    S.set_error_handler(None)
    code, msg = S.last_error()
    assert (code, msg) == (0, "") and seen == []


@pytest.mark.gpu
def test_out_of_memory_plan_reports_to_the_handler_and_the_library_stays_usable(built):
    """A plan whose solver vectors cannot be allocated (400000 x 400000 pixels: 5 TB): the handler gets cudaErrorMemoryAllocation
    with the reference's message text, Opt_ProblemPlan returns NULL, nothing stays allocated, and the next solve is unaffected."""
    import torch
    from opt_b200 import examples, solver as S
    from gpu_util import assert_costs_close, run_gpu, run_oracle
    pr = examples.poisson_synthetic(67, 45, n_iter=2, l_iter=12)
    before, _, _, _ = run_gpu(pr)                      # warm: context, library, allocator pools
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info()[0]
    seen = []
    S.set_error_handler(lambda code, msg: seen.append((code, msg)))
    try:
        with pytest.raises(RuntimeError, match="Opt_ProblemPlan failed"):
            S.OptSolver([400000, 400000], pr.energy, pr.kind)
        assert len(seen) == 1
        code, msg = seen[0]
        assert code == 2 and "Cuda reported error 2" in msg and "cudaMalloc" in msg and "engine.cuh" in msg
        assert S.last_error()[0] == 2
        assert torch.cuda.mem_get_info()[0] >= free0 - (1 << 20)
        after, _, sol, _ = run_gpu(pr)
        assert sol.failed() == 0, seen
        assert after == before
        rc, _, _ = run_oracle(pr)
        assert_costs_close(after, rc, False)
    finally:
        S.set_error_handler(None)


CHILD = r"""
import json, os, sys
sys.path.insert(0, %(root)r); sys.path.insert(0, os.path.join(%(root)r, "tests"))
import torch
from opt_b200 import examples, solver as S
pr = examples.poisson_synthetic(67, 45, n_iter=2, l_iter=12)
named = S.upload(pr)
sol = S.OptSolver(pr.dims, pr.energy, pr.kind)
seen = []
if %(install)d:
    S.set_error_handler(lambda code, msg: seen.append((code, msg)))
class Wild:                                   # an image pointer into nowhere: the first kernel that reads it faults
    is_cuda = True
    def is_contiguous(self): return True
    def data_ptr(self): return 0x10
named.values[pr.unknowns[0]] = Wild()
costs = []
try:
    sol.solve(pr.solver, named, costs)
except Exception as e:                        # not expected: the library returns, it does not raise
    print(json.dumps({"raised": repr(e)})); sys.stdout.flush(); os._exit(0)
print(json.dumps({"survived": True, "handler_calls": len(seen), "code": seen[0][0] if seen else None,
                  "message": seen[0][1] if seen else None, "failed": sol.failed(),
                  "cost_is_nan": bool(sol.current_cost() != sol.current_cost())}))
sys.stdout.flush()
os._exit(0)                                   # the CUDA context is gone: skip torch's teardown
"""


def run_child(install):
    r = subprocess.run([sys.executable, "-c", CHILD % {"root": ROOT, "install": install}], capture_output=True, text=True, timeout=300)
    return r


@pytest.mark.gpu
def test_kernel_fault_mid_solve_returns_to_the_caller(built):
    """A wild image pointer makes the first kernel fault (a sticky error: the context is lost, so this runs in a child process).
    With a handler the solve returns, the handler ran exactly once, the plan reads failed and its cost NaN ..."""
    r = run_child(1)
    assert r.returncode == 0, r.stderr[-2000:]
    out = json.loads(r.stdout.strip().splitlines()[-1])
    assert out.get("survived") and out["handler_calls"] == 1 and out["failed"] == 1 and out["cost_is_nan"], out
    assert out["code"] != 0 and "Cuda reported error" in out["message"], out


@pytest.mark.gpu
def test_without_a_handler_the_reference_behaviour_is_kept(built):
    """... and without one the process prints the reference's message and exits with the CUDA error code."""
    r = run_child(0)
    assert r.returncode != 0
    assert "Cuda reported error" in r.stdout and "survived" not in r.stdout
