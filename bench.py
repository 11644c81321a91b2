#!/usr/bin/env python
"""bench.py -- PCG iterations/s and solve time of the GN/LM + PCG hot path on B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload NAME]

A "step" is one complete Opt_ProblemSolve of the workload (Init, nIterations Gauss-Newton steps of
lIterations PCG iterations each, final cost).  Default workload at N=1 is BASELINE.json configs[1]:
poisson_image_editing on a synthetic 4096x4096 float4 image, Gauss-Newton + PCG, fp32
(nIterations=1, lIterations=100).  Rank 0 prints ONE JSON line.

  value      PCG iterations/s with every input already resident in HBM (device arrays handed to
             the C-ABI); the unknown image is restored by a device-to-device copy each step.
  e2e        same metric through the same C-ABI call sequence but starting from pinned HOST
             buffers: H2D of every image and D2H of the solved unknown inside the timed region, every
             step; consecutive steps are double-buffered so the copies overlap the previous solve.
  roofline   dominant kernel (PCGStep1 = J^T J p): algorithmic bytes / cudaEvent duration measured
             live by the library's per-kernel timer, against MEASURED_PEAKS.json.
  cpu_baseline  the CPU oracle (a port of the reference's backend_cpu/_mt path; the reference
             itself is Terra and cannot be built here) timed on this box's host cores.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
for _p in (ROOT, os.path.join(ROOT, "tests")):
    if _p not in sys.path:
        sys.path.insert(0, _p)

from opt_b200 import examples  # noqa: E402

# Algorithmic bytes follow SURVEY.md section 8(d): per PCG iteration s*(13U + I_JTJ + 3 I_excl) per site
# (+ `extra` bytes per vertex of index traffic for graphs); PCGStep1 alone s*(2U + I_JTJ + I_excl).
# ijtj = input channels the J^T J functor needs, iexcl = channels Exclude() reads, mask = index of the
# exclude mask in problemparams.  The unknown-vector terms are counted for ACTIVE (non-excluded) sites
# only -- excluded pixels are not unknowns and hold no data -- the exclude reads for all sites; the
# all-sites figure of the SURVEY table is reported next to it as `survey_formula`.
WORKLOADS = {
    "poisson_4096": dict(builder=examples.poisson_synthetic, args=(4096, 4096), kw=dict(n_iter=1, l_iter=100),
                         U=4, ijtj=0, iexcl=1, extra=0, mask=2,
                         desc="poisson_image_editing 4096x4096 float4, GN 1x100 PCG, fp32"),
    "image_warping_512": dict(builder=examples.image_warping_synthetic, args=(512, 512), kw=dict(n_iter=8, l_iter=400),
                              U=3, ijtj=8, iexcl=1, extra=0, mask=4,
                              desc="image_warping 512x512, GN 8x400 PCG, fp32"),
    "optical_flow_2048": dict(builder=examples.optical_flow_synthetic, args=(2048, 2048), kw=dict(n_iter=10, l_iter=50),
                              U=2, ijtj=4, iexcl=0, extra=0, mask=None,
                              desc="optical_flow 2048x2048, GN 10x50 PCG, fp32"),
    "sfs_2048": dict(builder=examples.sfs_synthetic, args=(2048, 2048), kw=dict(n_iter=10, l_iter=10, holes=False),
                     U=1, ijtj=5.5, iexcl=1, extra=0, mask=None,
                     desc="shape_from_shading 2048x2048, GN 10x10 PCG, fp32"),
    "volumetric_512": dict(builder=examples.volumetric_synthetic, args=(512, 512, 512), kw=dict(n_iter=2, l_iter=60),
                           U=6, ijtj=9, iexcl=0, extra=0, mask=None,
                           desc="volumetric_mesh_deformation 512^3, GN 2x60 PCG (the driver's 60 linear iterations per nonlinear step, main.cpp:32-36), fp32"),
    "volumetric_256": dict(builder=examples.volumetric_synthetic, args=(256, 256, 256), kw=dict(n_iter=2, l_iter=60),
                           U=6, ijtj=9, iexcl=0, extra=0, mask=None,
                           desc="volumetric_mesh_deformation 256^3, GN 2x60 PCG, fp32"),
    "arap_1m": dict(builder=examples.arap_synthetic, args=(1024, 1024), kw=dict(n_iter=2, l_iter=100),
                    U=6, ijtj=9, iexcl=0, extra=48, mask=None,
                    desc="arap_mesh_deformation 1024x1024-vertex mesh (6.28M directed edges), GN 2x100 PCG, fp32"),
}


# dram__bytes_read.sum + dram__bytes_write.sum of the dominant kernel, per launch, from the `ncu --set full`
# captures summarised under profiles/ (README there): workload -> (bytes, capture)
NCU_TRAFFIC = {}


def load_traffic():
    path = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(path):
        NCU_TRAFFIC.update(json.load(open(path)))


def load_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        return float(json.load(open(path))["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index=0):
        self.rows, self.proc, self.index = [], None, index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.time(), [c.strip() for c in line.split(",")]))

    def stop(self, t0=None, t1=None):
        """Samples that arrived inside [t0, t1] (the timed region).  nvidia-smi needs a few hundred ms to come up, so the
        sampler is started before the warm-up steps; if the timed region is shorter than one sampling period, the
        samples of the warm-up steps (same load, immediately before) are reported instead and the window says so."""
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.12)
        self.proc.terminate()
        self.thread.join(timeout=2)
        ok = [(t, r) for t, r in self.rows if len(r) >= 6 and r[0].replace(".", "").isdigit()]
        inside = [r for t, r in ok if t0 is None or (t0 <= t <= (t1 or t) + 0.12)]
        window = "timed region"
        if not inside:
            inside, window = [r for _, r in ok], "warm-up + timed region (timed region shorter than one sampling period)"
        sm = [float(r[0]) for r in inside]
        mx = [float(r[1]) for r in inside if r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for k, n in enumerate(names) if any(r[2 + k] == "Active" for r in inside)]
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm), "window": window}


SLAB = {   # workloads that run domain-decomposed at N > 1 (the others run as independent replicas)
    "poisson_4096": "weak",       # N stacked 4096-row images: per-GPU work fixed
    "volumetric_512": "strong",   # the 512^3 grid cut into N slabs of planes
    "volumetric_256": "strong",
}


def make_problem(name, size=None, rank=0, world=1):
    """Returns (local problem of this rank, workload record, slab mode or None)."""
    from opt_b200 import slab
    w = WORKLOADS[name]
    args = tuple(size) if size else w["args"]
    mode = SLAB.get(name) if world > 1 else None
    if mode == "weak":
        W, H = args
        g0, g1, _, _ = slab.local_range(H * world, rank, world, 1)
        return w["builder"](W, H * world, rows=(g0, g1), period=H, **w["kw"]), w, mode
    if mode == "strong":
        W, H, D = args
        g0, g1, _, _ = slab.local_range(D, rank, world, 1)
        return w["builder"](W, H, D, planes=(g0, g1), **w["kw"]), w, mode
    return w["builder"](*args, **w["kw"]), w, None


def sz_of(pr):
    return 8 if pr.double else 4


def pcg_iterations(pr):
    return int(pr.solver["nIterations"]) * int(pr.solver["lIterations"])


# ------------------------------------------------------------------------------ CPU oracle legs
def cpu_pcg_rate(pr, threads, budget_s=20.0):
    """PCG iterations/s of the CPU oracle on the full-size workload.  Two bounded solves with L1 and
    L2 PCG iterations are timed; the rate is (L2-L1)/(t2-t1), which removes the per-solve overhead
    (cost evaluation, PCGInit1, update) that a 100-iteration solve would amortise."""
    from oracle_api import OracleSolver

    def timed(l_iter):
        q = pr.clone()
        o = OracleSolver(q.dims, q.energy, q.kind, double=q.double, threads=threads)
        o.set_params(nIterations=1, lIterations=l_iter)
        t = time.perf_counter()
        o.solve(q.params)
        dt = time.perf_counter() - t
        o.close()
        return dt

    t1 = timed(1)
    per_iter_guess = max(t1 / 4.0, 1e-4)
    l2 = int(min(40, max(3, 1 + (budget_s - 2 * t1) / per_iter_guess)))
    t2 = timed(l2)
    rate = (l2 - 1) / max(t2 - t1, 1e-9)
    sample = "full-size workload, 2 solves with lIterations=1 and %d (%.1fs + %.1fs); rate=(%d-1)/(t2-t1)" % (l2, t1, t2, l2)
    return rate, sample, t1 + t2


def host_threads():
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


def run_reference(args, pr, w):
    """--impl reference: the reference's CPU solver path on the host cores.  The reference is
    Lua/Terra and cannot be built in this image, so this is the oracle port (kind "port")."""
    threads = host_threads()
    rates, samples = [], []
    for _ in range(args.warmup):
        pass                                            # nothing to warm on the CPU path
    t0 = time.perf_counter()
    for _ in range(max(1, args.steps)):
        rate, sample, _ = cpu_pcg_rate(pr, threads, budget_s=max(4.0, 40.0 / max(1, args.steps)))
        rates.append(rate)
        samples.append(sample)
    wall = time.perf_counter() - t0
    rate = float(np.median(rates))
    line = {
        "impl": "reference", "metric": "pcg_iterations_per_second", "value": rate, "unit": "PCG iters/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": 1000.0 * pcg_iterations(pr) / rate, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64" if pr.double else "f32", "data": "synthetic",
        "config": {"workload": w["desc"], "dims": pr.dims, "solver": pr.kind, "timing": "host wall clock", "wall_s": wall},
        "cpu_baseline": {"value": rate, "unit": "PCG iters/s", "cores": threads, "kind": "port", "sample": samples[0]},
        "e2e": {"value": rate, "unit": "PCG iters/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))


# ------------------------------------------------------------------------------ GPU arm
def run_ours(args, pr, w, rank, world, mode):
    import torch
    import torch.distributed as dist
    from opt_b200 import solver as S

    dev = torch.device("cuda", int(os.environ.get("LOCAL_RANK", 0)))
    torch.cuda.set_device(dev)
    img_idx = [i for i, p in enumerate(pr.params) if isinstance(p, np.ndarray)]
    host = {i: torch.from_numpy(pr.params[i]).pin_memory() for i in img_idx}
    devt = {i: host[i].to(dev, non_blocking=True) for i in img_idx}
    pristine = {i: devt[i].clone() for i in pr.unknowns}
    named = S.NamedParameters()
    for i, p in enumerate(pr.params):
        named.set("p%d" % i, devt[i] if i in devt else p)
    from opt_b200 import slab
    sol = S.OptSolver(pr.dims, pr.energy, pr.kind, double=pr.double, verbosity=0, timing=False)
    if mode:
        slab.setup_slab(sol, rank, world, 1)
    sol.set_solver_parameters(pr.solver)
    iters = pcg_iterations(pr)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def step_resident():
        for i in pr.unknowns:
            devt[i].copy_(pristine[i], non_blocking=True)
        sol.init(named)
        while sol.step():
            pass
        return sol.current_cost()

    out_host = {i: torch.empty_like(host[i]).pin_memory() for i in pr.unknowns}

    # End to end: every step uploads its inputs from pinned host memory and downloads the solved unknowns.  The steps
    # are pipelined the way a caller streaming problems through the C ABI would do it: two sets of device buffers,
    # uploads and downloads on their own streams (the library fences its own stream against the legacy default stream at
    # every call, torch side streams are non-blocking), so step k+1's upload and step k-1's download overlap step k's
    # solve.  All of it is inside the timed region.
    devt2 = {i: torch.empty_like(devt[i]) for i in img_idx}
    named2 = S.NamedParameters()
    for i, p in enumerate(pr.params):
        named2.set("p%d" % i, devt2[i] if i in devt2 else p)
    sets = [(devt, named), (devt2, named2)]
    up_stream, down_stream = torch.cuda.Stream(), torch.cuda.Stream()      # one per DMA direction (full-duplex link)

    def run_e2e(steps):
        main_stream = torch.cuda.current_stream()
        up = [torch.cuda.Event(), torch.cuda.Event()]
        done = [torch.cuda.Event(), torch.cuda.Event()]
        drained = [None, None]                  # download of the result that last lived in this buffer set
        up_stream.wait_stream(main_stream)
        down_stream.wait_stream(main_stream)

        def upload(k):
            d = sets[k % 2][0]
            with torch.cuda.stream(up_stream):
                if drained[k % 2] is not None:
                    up_stream.wait_event(drained[k % 2])
                for i in img_idx:
                    d[i].copy_(host[i], non_blocking=True)
                up[k % 2].record(up_stream)

        dbg = [] if os.environ.get("BENCH_E2E_DEBUG") else None
        upload(0)
        cost = None
        for k in range(steps):
            if k + 1 < steps:
                upload(k + 1)
            d, nm = sets[k % 2]
            main_stream.wait_event(up[k % 2])
            if dbg is not None:
                t0 = torch.cuda.Event(enable_timing=True); t0.record(main_stream)
            sol.init(nm)
            while sol.step():
                pass
            cost = sol.current_cost()
            done[k % 2].record(main_stream)
            if dbg is not None:
                t1 = torch.cuda.Event(enable_timing=True); t1.record(main_stream)
                dbg.append((t0, t1))
            with torch.cuda.stream(down_stream):
                down_stream.wait_event(done[k % 2])
                for i in pr.unknowns:
                    out_host[i].copy_(d[i], non_blocking=True)
                drained[k % 2] = torch.cuda.Event()
                drained[k % 2].record(down_stream)
        main_stream.wait_stream(up_stream)
        main_stream.wait_stream(down_stream)
        if dbg:
            torch.cuda.synchronize()
            sys.stderr.write("e2e solve ms per step: %s\n" % ["%.2f" % a.elapsed_time(b) for a, b in dbg])
        return cost

    def timed(fn, steps, warmup, sample_clocks):
        clk = ClockSampler(dev.index or 0) if sample_clocks else None
        if clk:
            clk.start()                      # before the warm-up: nvidia-smi takes a few hundred ms to deliver its first sample
        for _ in range(warmup):
            fn()
        barrier()
        t_begin = time.time()
        l0 = sol.launch_count()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        cost = None
        for _ in range(steps):
            cost = fn()
        e1.record()
        barrier()
        ms = e0.elapsed_time(e1)
        if world > 1:
            t = torch.tensor([ms], device=dev, dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms, cost, (clk.stop(t_begin, time.time()) if clk else None), sol.launch_count() - l0

    ms, cost, clocks, launches = timed(step_resident, args.steps, args.warmup, True)
    e2e_steps = max(1, args.steps)
    run_e2e(1)
    ms_e2e, cost_e2e, _, _ = timed(lambda: run_e2e(e2e_steps), 1, 0, False)
    h2d = sum(host[i].numel() * host[i].element_size() for i in img_idx)
    d2h = sum(out_host[i].numel() * out_host[i].element_size() for i in pr.unknowns) + 8

    # per-kernel time of the dominant kernel, measured live with the library's cudaEvent timer
    roof = None
    if rank == 0 or mode:
        tsol = S.OptSolver(pr.dims, pr.energy, pr.kind, double=pr.double, verbosity=0, timing=True)
        if mode:
            slab.setup_slab(tsol, rank, world, 1)
        tsol.set_solver_parameters(pr.solver)
        for rep in range(2):
            for i in pr.unknowns:
                devt[i].copy_(pristine[i])
            tsol.reset_timing()
            it0 = tsol.pcg_iterations()
            tsol.init(named)
            while tsol.step():
                pass
        torch.cuda.synchronize()
        kt = {k: tsol.kernel_time(k) for k in ("PCGStep1", "PCGStep3+1", "PCGStep1_Graph", "PCGStep2", "PCGStep3", "PCGInit1", "computeCost")}
        # PCG iterations one solve REALLY runs (LM leaves the inner loop early on q_tolerance, solver:2180-2200): the
        # metric counts these, not nIter x lIter.  The library counts them as the reference would (the one iteration
        # LM queues ahead of the exit test's verdict is wasted work, not a counted iteration).
        executed = tsol.pcg_iterations() - it0
        if executed > 0:
            iters = executed
        # the J^T J apply runs either as PCGStep1 or fused with the previous iteration's PCGStep3
        n1, t1 = kt["PCGStep3+1"] if kt["PCGStep3+1"][0] else kt["PCGStep1"]
        fused = kt["PCGStep3+1"][0] > 0
        peak, peak_src = load_peaks()
        N = sol.site_count()
        m = w["mask"]
        n_act = int((pr.params[m] == 0).sum()) if m is not None else N
        sz, U = (8 if pr.double else 4), w["U"]

        def alg_bytes(per_active, per_site):
            return sz * (n_act * per_active + N * per_site)

        step1_words = 2 * U + w["ijtj"] + (3 * U if fused else 0)      # fused kernel also does PCGStep3's 3U words
        iter_words = 13 * U + w["ijtj"]
        if n1:
            per_launch_ms = t1 / n1
            alg = alg_bytes(step1_words, w["iexcl"]) + w["extra"] * N
            ach = alg / (per_launch_ms * 1e-3) / 1e9
            it_alg = alg_bytes(iter_words, 3 * w["iexcl"]) + w["extra"] * N
            it_rate = iters * args.steps / (ms * 1e-3)
            it_survey = sz * N * (iter_words + 3 * w["iexcl"]) + w["extra"] * N
            roof = {"bound": "hbm", "kernel": "PCGStep3+PCGStep1 (fused)" if fused else "PCGStep1",
                    "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak,
                    "traffic": (NCU_TRAFFIC.get(args.workload + ("_f64" if pr.double else "")) or {}).get("bytes_per_launch") if world == 1 and not args.size else None,
                    "traffic_source": (NCU_TRAFFIC.get(args.workload + ("_f64" if pr.double else "")) or {}).get("capture"),
                    "frac_of_nominal_8TBs": ach / 8000.0,
                    "peak_source": peak_src, "algorithmic_bytes_per_launch": alg, "ms_per_launch": per_launch_ms,
                    "active_sites": n_act, "sites": N,
                    "kernel_ms": {k: (v[1] / v[0] if v[0] else None) for k, v in kt.items()},
                    "iteration": {"algorithmic_bytes": it_alg, "achieved": it_alg * it_rate / 1e9,
                                  "frac": it_alg * it_rate / 1e9 / peak,
                                  "survey_formula": {"algorithmic_bytes": it_survey, "frac": it_survey * it_rate / 1e9 / peak}}}
        tsol.close()

    # units: one PCG iteration over one single-GPU-sized grid.  weak slab / replicas: N such units per
    # global iteration; strong slab: the global grid IS the single-GPU grid.
    mult = 1 if mode == "strong" else world
    value = mult * iters * args.steps / (ms * 1e-3)
    if rank != 0:
        return
    cpu = None
    if not args.no_cpu and world == 1:
        threads = host_threads()
        rate, sample, _ = cpu_pcg_rate(pr, threads, budget_s=args.cpu_budget)
        cpu = {"value": rate, "unit": "PCG iters/s", "cores": threads, "kind": "port", "sample": sample}
    line = {
        "metric": "pcg_iterations_per_second", "value": value, "unit": "PCG iters/s", "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True,
        "scaling": "strong" if mode == "strong" else "weak", "vs_baseline": None,
        "dtype": "f32" if not pr.double else "f64", "data": "synthetic",
        "config": {"workload": w["desc"], "dims": pr.dims, "solver": pr.kind, "pcg_iterations_per_step": iters,
                   "l2": "inputs larger than L2 (each solver vector is %.0f MB)" % (w["U"] * sol.site_count() * sz_of(pr) / 1e6),
                   "pcg_iterations_nominal": pcg_iterations(pr), "final_cost": cost, "parallelism": "1 GPU" if world == 1 else ("%d ranks, %s" % (world, (
                       "slab decomposition along the slowest dim (%s scaling), NCCL halo exchange + all-reduced PCG scalars" % mode)
                       if mode else "independent replicas"))},
        "clocks": clocks,
        "e2e": {"value": mult * iters * e2e_steps / (ms_e2e * 1e-3),
                "unit": "PCG iters/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "ms_per_step": ms_e2e / e2e_steps, "final_cost": cost_e2e},
        "gpu_launches": launches,
        "roofline": roof,
        "cpu_baseline": cpu,
    }
    print(json.dumps(line))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="poisson_4096", choices=sorted(WORKLOADS))
    ap.add_argument("--size", type=int, nargs="+", default=None, help="override the workload's grid size")
    ap.add_argument("--double", action="store_true", help="run the workload in fp64 (doublePrecision = 1)")
    ap.add_argument("--kind", default=None, choices=["gaussNewtonGPU", "LMGPU"], help="override the workload's solver kind")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--cpu-budget", type=float, default=20.0)
    args = ap.parse_args()
    load_traffic()
    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    def variant(pr):
        if args.kind:
            pr.kind = args.kind
        return pr.as_double() if args.double else pr

    if args.impl == "reference":
        if rank == 0:
            pr, w, _ = make_problem(args.workload, args.size)
            run_reference(args, variant(pr), w)
        return
    pr, w, mode = make_problem(args.workload, args.size, rank, world)
    pr = variant(pr)
    if world > 1:
        import torch
        import torch.distributed as dist
        torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", 0)))
        dist.init_process_group("nccl")
    run_ours(args, pr, w, rank, world, mode)
    if world > 1:
        import torch.distributed as dist
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
