"""GPU (needs >= 2 devices): slab-decomposed solve == the CPU oracle's and the single-GPU solve of the undecomposed problem,
at 2, 4 and 8 ranks, over peer memory and over NCCL only.  Skipped where the box has fewer devices;
run with `gpurun --gpus 8 -- python -m pytest tests/test_multigpu.py -m gpu`."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _ngpus():
    import torch
    return torch.cuda.device_count()


@pytest.mark.parametrize("p2p", ["p2p", "nccl"])
@pytest.mark.parametrize("world", [2, 4, 8])
def test_slab_solve_matches_oracle_and_single_gpu(built, world, p2p):
    if _ngpus() < world:
        pytest.skip("needs %d GPUs" % world)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(world),
           "--master-addr", "127.0.0.1", "--master-port", str(29517 + world), os.path.join(ROOT, "tests", "mgpu_worker.py")]
    env = dict(os.environ)
    env.pop("OPTB200_NO_P2P", None)
    if p2p == "nccl":
        env["OPTB200_NO_P2P"] = "1"
        env["MGPU_CASES"] = "poisson,poisson_lm,volumetric_packed,volumetric_packed_lm,sfs,optical_flow_lm"
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=900, env=env)
    sys.stdout.write(r.stdout[-6000:])
    assert r.returncode == 0, r.stderr[-4000:]
    assert "MISMATCH" not in r.stdout and r.stdout.count(" OK ") >= 10
