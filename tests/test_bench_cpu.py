"""bench.py's CPU-side contract: the reference arm (the oracle port timed on the host cores) prints one JSON line with the
keys the driver reads, on the same `config` as the GPU arm; helpers that must not fail on a box without NVML / a GPU."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def run_bench(*args, env=None):
    e = dict(os.environ)
    e.update(env or {})
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), *args], capture_output=True, text=True, timeout=600, env=e, cwd=ROOT)
    assert r.returncode == 0, r.stderr[-2000:]
    return r.stdout.strip().splitlines()


def test_reference_arm_line(built):
    lines = run_bench("--impl", "reference", "--steps", "1", "--warmup", "1")
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["metric"] == "pcg_iterations_per_second" and d["unit"] == "PCG iters/s"
    assert d["higher_is_better"] is True and d["scaling"] == "strong" and d["n_gpus"] == 1
    assert d["value"] > 0 and d["ms_per_step"] > 0 and d["gpu_launches"] == 0
    assert d["e2e"] == {"value": d["value"], "unit": d["unit"], "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    cb = d["cpu_baseline"]
    assert cb["kind"] == "port" and cb["cores"] >= 1 and cb["value"] == d["value"] and "sample" in cb
    assert d["sampled"] is True                      # 512^3 does not fit the time budget: a bounded sample, scaled
    assert d["config"]["dims"] == [512, 512, 512] and d["config"]["nIterations"] == 2 and d["config"]["lIterations"] == 60


def test_reference_arm_under_torchrun_env_prints_on_rank_0_only(built):
    """N > 1: the driver launches the reference arm under torchrun; rank 0 alone runs it, the others exit 0 without work."""
    assert run_bench("--impl", "reference", "--gpus", "2", "--steps", "1", "--warmup", "0", env={"RANK": "1", "WORLD_SIZE": "2", "LOCAL_RANK": "1"}) == []
    lines = run_bench("--impl", "reference", "--gpus", "2", "--steps", "1", "--warmup", "0", env={"RANK": "0", "WORLD_SIZE": "2", "LOCAL_RANK": "0"})
    assert len(lines) == 1 and json.loads(lines[0])["n_gpus"] == 2


def test_same_config_for_both_arms_and_cpu_binding_is_optional():
    sys.path.insert(0, ROOT)
    import argparse
    import bench
    args = argparse.Namespace(workload="volumetric_512", size=None, double=False, kind=None, materialized=0, gpus=1)
    a = bench.config_of(args, bench.WORKLOADS["volumetric_512"], 1)
    b = bench.config_of(args, bench.WORKLOADS["volumetric_512"], 1)
    assert a == b and a["workload"].startswith("volumetric_mesh_deformation 512^3")
    assert bench.bind_near_gpu(0) is None or bench.bind_near_gpu(0) >= 1     # no NVML / no GPU here: None, never an exception
