#!/usr/bin/env python
"""bench.py -- PCG iterations/s and solve time of the GN/LM + PCG hot path on B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload NAME]

A "step" is one complete Opt_ProblemSolve of the workload (Init, nIterations Gauss-Newton steps of
lIterations PCG iterations each, final cost).  The default workload is the configuration BASELINE.json's
north_star states its targets on and the largest single-GPU configuration: volumetric_mesh_deformation on
a synthetic 512^3 grid (configs[4]; SURVEY.md section 8d C5: GN, 60 linear iterations per nonlinear step as
examples/volumetric_mesh_deformation/src/main.cpp:32-36), STRONG-scaled: at --gpus N the same 512^3 grid is
cut into N slabs of planes.  The other BASELINE configurations are selected with --workload
(tools/bench_all.sh writes one line per configuration under profiles/).  Rank 0 prints ONE JSON line.

  value      PCG iterations/s with every input already resident in HBM (device arrays handed to
             the C-ABI); the unknown image is restored by a device-to-device copy each step.
  e2e        same metric through the same C-ABI call sequence but starting from pinned HOST
             buffers: H2D of every image and D2H of the solved unknown inside the timed region, every
             step; consecutive steps are double-buffered so the copies overlap the previous solve
             (pipeline throughput; `e2e.cold_solve_ms` is one un-overlapped upload + solve + download).
  roofline   dominant kernel (PCGStep1 = J^T J p, fused with PCGStep3 where the engine fuses them):
             algorithmic bytes / cudaEvent duration measured live by the library's per-kernel timer,
             against MEASURED_PEAKS.json.
  parity     the CUDA path against the CPU oracle inside this run: the oracle solve the cpu_baseline leg
             times is repeated on the GPU with the same iteration counts (cost and unknowns compared), and for
             workloads whose full-size oracle solve does not fit the time budget (512^3) the operators
             r = -J^T F, the preconditioner and J^T J p of the FULL-SIZE plan are read back on sampled plane
             ranges and compared with the oracle evaluated on those sub-volumes.  N > 1 (strong-scaled 3-D default): the
             same operator check on every rank's slab, including the owned planes next to the ghost planes, worst rank
             reported; each rank runs on the CPUs next to its GPU (NVML affinity) so its staging buffers are NUMA-local.
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
                         U=4, ijtj=0, iexcl=1, extra=0, mask=2, ref_sample=(1024, 1024),
                         desc="poisson_image_editing 4096x4096 float4, GN 1x100 PCG, fp32"),
    "image_warping_512": dict(builder=examples.image_warping_synthetic, args=(512, 512), kw=dict(n_iter=8, l_iter=400),
                              U=3, ijtj=8, iexcl=1, extra=0, mask=4, ref_sample=(128, 128),
                              desc="image_warping 512x512, GN 8x400 PCG, fp32"),
    "optical_flow_2048": dict(builder=examples.optical_flow_synthetic, args=(2048, 2048), kw=dict(n_iter=10, l_iter=50),
                              U=2, ijtj=4, iexcl=0, extra=0, mask=None, ref_sample=(512, 512),
                              desc="optical_flow 2048x2048, GN 10x50 PCG, fp32"),
    "sfs_2048": dict(builder=examples.sfs_synthetic, args=(2048, 2048), kw=dict(n_iter=10, l_iter=10, holes=False),
                     U=1, ijtj=5.5, iexcl=1, extra=0, mask=None, ref_sample=(512, 512),
                     desc="shape_from_shading 2048x2048, GN 10x10 PCG, fp32"),
    "volumetric_512": dict(builder=examples.volumetric_synthetic, args=(512, 512, 512), kw=dict(n_iter=2, l_iter=60),
                           U=6, ijtj=9, iexcl=0, extra=0, mask=None, cpu_sample=(512, 512, 8), ref_sample=(64, 64, 16),
                           desc="volumetric_mesh_deformation 512^3, GN 2x60 PCG (the driver's 60 linear iterations per nonlinear step, main.cpp:32-36), fp32"),
    "volumetric_256": dict(builder=examples.volumetric_synthetic, args=(256, 256, 256), kw=dict(n_iter=2, l_iter=60),
                           U=6, ijtj=9, iexcl=0, extra=0, mask=None, cpu_sample=(256, 256, 16), ref_sample=(64, 64, 16),
                           desc="volumetric_mesh_deformation 256^3, GN 2x60 PCG, fp32"),
    "arap_1m": dict(builder=examples.arap_synthetic, args=(1024, 1024), kw=dict(n_iter=2, l_iter=100),
                    U=6, ijtj=9, iexcl=0, extra=48, mask=None, ref_sample=(192, 192),
                    desc="arap_mesh_deformation 1024x1024-vertex mesh (6.28M directed edges), GN 2x100 PCG, fp32"),
    # the other three graph energies (hyper-edge terms as a segmented scatter, engine.cuh k_graph_contrib / k_apply_jtj_seg).
    # ijtj = input channels J^T J reads per vertex; extra = bytes of index arrays per vertex (GV arrays x 4 B x ~6 edges)
    "embedded_1m": dict(builder=examples.embedded_synthetic, args=(1024, 1024), kw=dict(n_iter=2, l_iter=50),
                        U=12, ijtj=15, iexcl=0, extra=48, mask=None, ref_sample=(128, 128),
                        desc="embedded_mesh_deformation 1024x1024-vertex mesh (6.28M directed edges), GN 2x50 PCG, fp32"),
    "cotangent_1m": dict(builder=examples.cotangent_synthetic, args=(1024, 1024), kw=dict(n_iter=2, l_iter=50),
                         U=3, ijtj=3, iexcl=0, extra=96, mask=None, ref_sample=(128, 128),
                         desc="cotangent_mesh_smoothing 1024x1024-vertex mesh (6.28M 4-vertex hyper-edges), GN 2x50 PCG, fp32"),
    "robust_1m": dict(builder=examples.robust_synthetic, args=(1024, 1024), kw=dict(n_iter=2, l_iter=50, kind="LMGPU"),
                      U=7, ijtj=16, iexcl=0, extra=48, mask=None, ref_sample=(128, 128),
                      desc="robust_nonrigid_alignment 1024x1024-vertex mesh (6.28M directed edges), LM 2x50 PCG, fp32"),
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
def sample_problem(name, key, size=None):
    """A bounded sample of the workload for the CPU legs: the same builder on `cpu_sample` / `ref_sample` dims (3-D:
    a slab of planes from the middle of the full volume, so positions and constraints are those of the full problem).
    Returns (problem, sites of the sample / sites of the full workload, description); no `key` entry: the full size."""
    w = WORKLOADS[name]
    full = tuple(size) if size else w["args"]
    dims = w.get(key)
    if not dims or size:
        return w["builder"](*full, **w["kw"]), 1.0, "full-size workload"
    if len(full) == 3 and dims[:2] == full[:2]:
        z0 = (full[2] - dims[2]) // 2
        pr = w["builder"](*full, planes=(z0, z0 + dims[2]), **w["kw"])
        what = "planes [%d, %d) of the %s volume" % (z0, z0 + dims[2], "x".join(map(str, full)))
    else:
        pr = w["builder"](*dims, **w["kw"])
        what = "the same builder at %s instead of %s" % ("x".join(map(str, dims)), "x".join(map(str, full)))
    frac = float(np.prod(dims)) / float(np.prod(full))
    return pr, frac, "%s (%.4g of the sites; rates scaled by that factor)" % (what, frac)


def oracle_solve(pr, threads, n_iter, l_iter):
    """One whole oracle solve; returns (seconds, cost history, solved problem copy)."""
    from oracle_api import OracleSolver
    q = pr.clone()
    o = OracleSolver(q.dims, q.energy, q.kind, double=q.double, threads=threads)
    o.set_params(nIterations=int(n_iter), lIterations=int(l_iter))
    t = time.perf_counter()
    o.solve(q.params)
    dt = time.perf_counter() - t
    costs = o.cost_history()
    o.close()
    return dt, costs, q


def cpu_pcg_rate(pr, threads, budget_s=20.0):
    """PCG iterations/s of the CPU oracle on `pr`.  Two bounded solves with 1 and L2 PCG iterations are timed; the
    rate is (L2-1)/(t2-t1), which removes the per-solve overhead (cost evaluation, PCGInit1, update) that a long
    solve would amortise.  The second solve is returned so that the GPU can be checked against it."""
    t1, _, _ = oracle_solve(pr, threads, 1, 1)
    per_iter_guess = max(t1 / 4.0, 1e-4)
    l2 = int(min(40, max(3, 1 + (budget_s - 2 * t1) / per_iter_guess)))
    t2, costs, solved = oracle_solve(pr, threads, 1, l2)
    rate = (l2 - 1) / max(t2 - t1, 1e-9)
    sample = "2 solves with lIterations=1 and %d (%.1fs + %.1fs); rate=(%d-1)/(t2-t1)" % (l2, t1, t2, l2)
    return rate, sample, l2, costs, solved


CPU_THREADS = 0          # --cpu-threads


def host_threads():
    if CPU_THREADS > 0:
        return CPU_THREADS
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


def parallelism(world, mode):
    if world == 1:
        return "1 GPU"
    if not mode:
        return "%d ranks, independent replicas" % world
    return ("%d ranks, slab decomposition along the slowest dimension (%s scaling): boundary planes of the search direction "
            "and the PCG scalars move by direct stores into peer memory over NVLink (NCCL for set-up and the "
            "once-per-nonlinear-step unknown exchange)" % (world, mode))


def config_of(args, w, world):
    """Identical for both arms (the driver compares them): the workload, not the implementation."""
    dims = [int(d) for d in (args.size or w["args"])]
    kind = args.kind or w["kw"].get("kind", "gaussNewtonGPU")
    mode = SLAB.get(args.workload) if world > 1 else None
    return {"workload": w["desc"], "dims": dims, "solver": kind, "precision": "f64" if args.double else "f32",
            "nIterations": int(w["kw"]["n_iter"]), "lIterations": int(w["kw"]["l_iter"]),
            "pcg_iterations_nominal": int(w["kw"]["n_iter"]) * int(w["kw"]["l_iter"]),
            "l2": "inputs larger than L2 (each solver vector is %.0f MB)" % (w["U"] * float(np.prod(dims)) * (8 if args.double else 4) / 1e6),
            "parallelism": parallelism(world, mode)}


def run_reference(args, w):
    """--impl reference: the reference's CPU solver path on the host cores.  The reference is Lua/Terra and cannot be
    built in this image, so this is the oracle port (kind "port").  Every step is one WHOLE solve (the workload's
    nIterations x lIterations) of a bounded sample of the workload, timed by the host clock; `value` is the sample's
    PCG iterations/s scaled by (sample sites / workload sites), i.e. the rate of full-size iterations at the measured
    per-site cost ("sampled": true)."""
    threads = host_threads()
    pr, frac, what = sample_problem(args.workload, "ref_sample", args.size)
    if args.kind:
        pr.kind = args.kind
    if args.double:
        pr = pr.as_double()
    n_iter, l_iter = int(pr.solver["nIterations"]), int(pr.solver["lIterations"])
    for _ in range(min(1, args.warmup)):
        oracle_solve(pr, threads, 1, 2)                 # page in the library, spin up the OpenMP team
    times = []
    for _ in range(max(1, args.steps)):
        dt, costs, _ = oracle_solve(pr, threads, n_iter, l_iter)
        times.append(dt)
    total = float(np.sum(times))
    iters = n_iter * l_iter
    rate_sample = iters * len(times) / total
    rate = rate_sample * frac
    line = {
        "impl": "reference", "metric": "pcg_iterations_per_second", "value": rate, "unit": "PCG iters/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": 1000.0 * total / len(times), "higher_is_better": True,
        "scaling": "strong" if SLAB.get(args.workload) == "strong" else "weak",
        "vs_baseline": None, "dtype": "f64" if pr.double else "f32", "data": "synthetic",
        "config": config_of(args, w, args.gpus), "sampled": frac != 1.0,
        "cpu_baseline": {"value": rate, "unit": "PCG iters/s", "cores": threads, "kind": "port",
                         "sample": "%s; every step is one whole GN %dx%d solve of the sample (%.2f s each), value = sample "
                                   "iterations/s x %.4g" % (what, n_iter, l_iter, total / len(times), frac),
                         "sample_rate": rate_sample, "sample_dims": [int(d) for d in pr.dims], "final_cost_of_sample": costs[-1] if costs else None},
        "e2e": {"value": rate, "unit": "PCG iters/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0, "timing": "host wall clock around each whole solve",
    }
    print(json.dumps(line))


# ------------------------------------------------------------------------------ GPU arm
def bind_near_gpu(index):
    """N > 1: run this rank on the CPUs NVML names as local to its GPU, before any pinned buffer is allocated, so that the
    end-to-end leg's staging memory is first-touched on the NUMA node the GPU's PCIe root hangs off (eight ranks staging
    through one node's memory is what limited r1's end-to-end curve).  Returns the CPU count, or None if NVML has no answer."""
    try:
        import pynvml
        pynvml.nvmlInit()
        try:                                  # CUDA's ordinal need not be NVML's: go through the PCI address
            import torch
            pp = torch.cuda.get_device_properties(index)
            h = pynvml.nvmlDeviceGetHandleByPciBusId(("%08x:%02x:%02x.0" % (pp.pci_domain_id, pp.pci_bus_id, pp.pci_device_id)).encode())
        except Exception:
            h = pynvml.nvmlDeviceGetHandleByIndex(index)
        words = pynvml.nvmlDeviceGetCpuAffinity(h, (os.cpu_count() + 63) // 64)
        cpus = {64 * w + b for w, m in enumerate(words) for b in range(64) if (int(m) >> b) & 1}
        cpus &= os.sched_getaffinity(0)
        if cpus:
            os.sched_setaffinity(0, cpus)
            return len(cpus)
    except Exception:
        pass
    return None


def run_ours(args, pr, w, rank, world, mode):
    import torch
    import torch.distributed as dist
    from opt_b200 import solver as S

    dev = torch.device("cuda", int(os.environ.get("LOCAL_RANK", 0)))
    torch.cuda.set_device(dev)
    numa = bind_near_gpu(dev.index or 0) if world > 1 else None
    img_idx = [i for i, p in enumerate(pr.params) if isinstance(p, np.ndarray)]
    host = {i: torch.from_numpy(pr.params[i]).pin_memory() for i in img_idx}
    devt = {i: host[i].to(dev, non_blocking=True) for i in img_idx}
    pristine = {i: devt[i].clone() for i in pr.unknowns}
    named = S.NamedParameters()
    for i, p in enumerate(pr.params):
        named.set("p%d" % i, devt[i] if i in devt else p)
    from opt_b200 import slab
    mat = dict(materialized=args.materialized > 0, fused_jtj=args.materialized == 2)
    sol = S.OptSolver(pr.dims, pr.energy, pr.kind, double=pr.double, verbosity=0, timing=False, **mat)
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
    ms_cold, _, _, _ = timed(lambda: run_e2e(1), 1, 0, False)          # one upload + solve + download, nothing overlapped
    ms_e2e, cost_e2e, _, _ = timed(lambda: run_e2e(e2e_steps), 1, 0, False)
    h2d = sum(host[i].numel() * host[i].element_size() for i in img_idx)
    d2h = sum(out_host[i].numel() * out_host[i].element_size() for i in pr.unknowns) + 8

    # per-kernel time of the dominant kernel, measured live with the library's cudaEvent timer
    roof = None
    if rank == 0 or mode:
        tsol = S.OptSolver(pr.dims, pr.energy, pr.kind, double=pr.double, verbosity=0, timing=True, **mat)
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
        kt = {k: tsol.kernel_time(k) for k in ("PCGStep1", "PCGStep3+1", "PCGStep1_Graph", "PCGStep2", "PCGStep3", "PCGInit1", "computeCost",
                                               "Jp", "J^T", "J^TJp", "PCGStep1_Finish", "saveJToCRS", "J_transpose", "JTJ multiply")}
        # PCG iterations one solve REALLY runs (LM leaves the inner loop early on q_tolerance, solver:2180-2200): the
        # metric counts these, not nIter x lIter.  The library counts them as the reference would (the one iteration
        # LM queues ahead of the exit test's verdict is wasted work, not a counted iteration).
        executed = tsol.pcg_iterations() - it0
        if executed > 0:
            iters = executed
        # the J^T J apply runs either as PCGStep1 or fused with the previous iteration's PCGStep3
        n1, t1 = kt["PCGStep3+1"] if kt["PCGStep3+1"][0] else kt["PCGStep1"]
        fused = kt["PCGStep3+1"][0] > 0
        two_pass = not fused and n1 > 0 and kt["PCGStep1_Graph"][0] == n1      # graph energies: edge pass + vertex pass = one J^T J apply
        if two_pass:
            t1 += kt["PCGStep1_Graph"][1]
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
            roof = {"bound": "hbm", "kernel": "PCGStep3+PCGStep1 (fused)" if fused else ("PCGStep1_Graph + PCGStep1" if two_pass else "PCGStep1"),
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
        if tsol.materialized_mode() and roof is None:
            # materialized-Jacobian mode: PCGStep1 is one (J^T J) or two (J, J^T) CSR products.  Algorithmic bytes of y = A x:
            # values + column indices of every entry, the row offsets, x read once and y written once.
            peak, peak_src = load_peaks()
            sz = 8 if pr.double else 4
            ncol = sol.site_count() * w["U"]
            fusedm = tsol.materialized_mode() == 2
            nnz = tsol.materialized_nnz(1 if fusedm else 0)
            names = ["J^TJp"] if fusedm else ["Jp", "J^T"]
            t_ms = sum(kt[k][1] / kt[k][0] for k in names if kt[k][0])
            rows = ncol if fusedm else None
            alg = (nnz * (sz + 4) + (ncol + 1) * 4 + 2 * ncol * sz) if fusedm else 2 * nnz * (sz + 4) + 3 * ncol * sz
            roof = {"bound": "hbm", "kernel": " + ".join(names) + " (CSR SpMV, materialized Jacobian)", "achieved": alg / (t_ms * 1e-3) / 1e9 if t_ms else None,
                    "peak": peak, "unit": "GB/s", "frac": alg / (t_ms * 1e-3) / 1e9 / peak if t_ms else None, "traffic": None,
                    "frac_of_nominal_8TBs": alg / (t_ms * 1e-3) / 1e9 / 8000.0 if t_ms else None, "peak_source": peak_src,
                    "algorithmic_bytes_per_launch": alg, "ms_per_launch": t_ms, "nnz": nnz,
                    "kernel_ms": {k: (v[1] / v[0] if v[0] else None) for k, v in kt.items()}}
        tsol.close()

    # units: one PCG iteration over one single-GPU-sized grid.  weak slab / replicas: N such units per
    # global iteration; strong slab: the global grid IS the single-GPU grid.
    mult = 1 if mode == "strong" else world
    value = mult * iters * args.steps / (ms * 1e-3)
    slab_parity = None
    if mode == "strong" and world > 1 and not args.no_cpu and len(pr.dims) == 3:
        # N > 1: the full-size operator check on every rank's slab, worst rank reported (the whole-solve parity of the slab
        # paths against the oracle is tests/test_multigpu.py's job; the final cost below is the N = 1 run's to the last digit)
        local = operator_parity_3d(S, sol, named, pr, devt, pristine, max(1, min(host_threads(), (os.cpu_count() or world) // world)), bottom=rank == 0, top=rank == world - 1)
        t = torch.tensor([local["max_rel_err_p0"], local["max_rel_err_pre"], local["max_rel_err_JTJp"]], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        worst = [float(v) for v in t.tolist()]
        slab_parity = {"iters": None, "problem": "every rank's slab of the full-size workload (max over %d ranks)" % world,
                       "full_size_operators": {"what": local["what"], "gate": local["gate"], "max_rel_err_p0": worst[0],
                                               "max_rel_err_pre": worst[1], "max_rel_err_JTJp": worst[2],
                                               "planes_rank0": local["planes"], "ok": bool(max(worst) <= local["gate"])},
                       "ok": bool(max(worst) <= local["gate"])}
    if rank != 0:
        return
    cpu, parity = None, slab_parity
    if not args.no_cpu and world == 1:
        threads = host_threads()
        spr, frac, what = sample_problem(args.workload, "cpu_sample", args.size)
        spr.kind = pr.kind
        if pr.double:
            spr = spr.as_double()
        rate, sample, l2, ocosts, solved = cpu_pcg_rate(spr, threads, budget_s=args.cpu_budget)
        cpu = {"value": rate * frac, "unit": "PCG iters/s", "cores": threads, "kind": "port",
               "sample": "%s; %s" % (what, sample), "sample_rate": rate, "sample_dims": [int(d) for d in spr.dims]}
        # parity at the iteration count the GPU tests use (<= 12 PCG iterations in one nonlinear step): fp32 PCG runs that differ
        # only in summation order drift apart exponentially with the iteration count, so the timing solve's 40 iterations are
        # not the right yardstick for BASELINE's 1e-4 / 1e-9 gates; a shorter oracle solve is run for the comparison
        if l2 > PARITY_MAX_ITERS:
            l2 = PARITY_MAX_ITERS
            _, ocosts, solved = oracle_solve(spr, threads, 1, l2)
        parity = solve_parity(S, spr, l2, ocosts, solved)
        parity["problem"] = what
        if frac != 1.0 and len(pr.dims) == 3:
            parity["full_size_operators"] = operator_parity_3d(S, sol, named, pr, devt, pristine, threads)
            parity["ok"] = bool(parity["ok"] and parity["full_size_operators"]["ok"])
    it_ms = ms / args.steps / iters
    kern_ms = None
    if roof:
        km = roof["kernel_ms"]
        kern_ms = sum(v for k, v in km.items() if v and k in ("PCGStep1", "PCGStep3+1", "PCGStep1_Graph", "PCGStep2", "PCGStep3")
                      and not (k == "PCGStep1" and km.get("PCGStep3+1")))
    line = {
        "metric": "pcg_iterations_per_second", "value": value, "unit": "PCG iters/s", "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True,
        "scaling": "strong" if mode == "strong" or (world == 1 and SLAB.get(args.workload) == "strong") else "weak", "vs_baseline": None,
        "dtype": "f32" if not pr.double else "f64", "data": "synthetic",
        "config": config_of(args, w, world),
        "pcg_iterations_per_step": iters, "final_cost": cost,
        "iteration_ms": {"total": it_ms, "kernels": kern_ms, "overhead": (it_ms - kern_ms) if kern_ms else None,
                         "note": "total = timed region / PCG iterations; kernels = live cudaEvent time of the per-iteration kernels "
                                 "(J^T J apply + PCGStep2 [+ PCGStep3]); overhead = launch gaps, per-nonlinear-step kernels and, at N > 1, "
                                 "the cross-GPU handshakes"},
        "clocks": clocks,
        "e2e": {"value": mult * iters * e2e_steps / (ms_e2e * 1e-3),
                "unit": "PCG iters/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "ms_per_step": ms_e2e / e2e_steps, "final_cost": cost_e2e, "cold_solve_ms": ms_cold,
                "host_cpus_bound": numa,          # N > 1: rank 0's CPU set after binding next to its GPU (staging buffers NUMA-local)
                "note": "pipeline throughput: step k+1's upload and step k-1's download overlap step k's solve; cold_solve_ms is one "
                        "un-overlapped upload + solve + download"},
        "gpu_launches": launches,
        "roofline": roof,
        "parity": parity,
        "cpu_baseline": cpu,
    }
    print(json.dumps(line))


PARITY_MAX_ITERS = 12
PARITY_GATES = {False: {"cost": 1e-4, "unknowns": 2e-4, "operator": 2e-5}, True: {"cost": 1e-9, "unknowns": 1e-8, "operator": 1e-11}}


def solve_parity(S, spr, l_iter, ocosts, solved):
    """The oracle solve `solved` (nIterations=1, lIterations=l_iter on `spr`) repeated through the C ABI on the GPU."""
    import torch
    named = S.upload(spr)
    sol = S.OptSolver(spr.dims, spr.energy, spr.kind, double=spr.double)
    gcosts = []
    sol.solve({"nIterations": 1, "lIterations": int(l_iter)}, named, gcosts)
    torch.cuda.synchronize()
    gate = PARITY_GATES[bool(spr.double)]
    ok = len(gcosts) == len(ocosts)
    rel_cost = max([abs(a - b) / max(abs(b), 1e-300) for a, b in zip(gcosts, ocosts)] or [float("inf")])
    rel_x = 0.0
    for i in spr.unknowns:
        ref = solved.params[i].reshape(-1).astype(np.float64)
        got = named.values[i].cpu().numpy().reshape(-1).astype(np.float64)
        fin = np.isfinite(ref)
        rel_x = max(rel_x, float(np.abs(got[fin] - ref[fin]).max() / (np.abs(ref[fin]).max() + 1e-30)))
    sol.close()
    ok = bool(ok and np.isfinite(rel_cost) and rel_cost <= gate["cost"] and rel_x <= gate["unknowns"])
    return {"iters": int(l_iter), "rel_err_cost": float(rel_cost), "max_rel_err_unknowns": float(rel_x),
            "costs_gpu": gcosts, "costs_oracle": list(ocosts), "gate": gate, "ok": ok}


def operator_parity_3d(S, sol, named, pr, devt, pristine, threads, planes=6, bottom=True, top=True):
    """FULL-SIZE check of a 3-D workload whose oracle solve does not fit the time budget: the full-size plan runs one PCG
    iteration from the initial unknowns; p0 = M^-1 (-J^T F), M^-1 and Ap = J^T J p0 are read back on three ranges of
    `planes` planes (bottom, middle, top of the volume) and compared with the oracle evaluated on those sub-volumes
    (r, the preconditioner and J^T J p are local: radius-1 stencils, so a plane that has both its z-neighbours inside
    the sub-volume -- or lies on the boundary of the full grid -- is exact).  In slab mode (N > 1) every rank calls this on its
    own slab (the solve inside is collective); bottom / top say whether the slab's first / last plane is the boundary of the
    full grid or a ghost plane, which - like every other plane without both neighbours in the sub-volume - is left out."""
    from oracle_api import OracleSolver
    import torch
    W, H, D = [int(d) for d in pr.dims]
    plane = W * H
    for i in pr.unknowns:
        devt[i].copy_(pristine[i])
    saved = dict(pr.solver)
    sol.set_solver_parameters({"nIterations": 1, "lIterations": 1})
    sol.init(named)
    while sol.step():
        pass
    torch.cuda.synchronize()
    sol.set_solver_parameters(saved)
    gate = PARITY_GATES[bool(pr.double)]["operator"]
    dt = np.float64 if pr.double else np.float32
    out = {"planes": [], "max_rel_err_p0": 0.0, "max_rel_err_pre": 0.0, "max_rel_err_JTJp": 0.0, "gate": gate,
           "what": "p0 = M^-1(-J^T F), M^-1 and J^T J p0 of the full-size plan vs the oracle on sub-volumes"}
    for za in sorted({0, max(0, D // 2 - planes // 2), max(0, D - planes)}):
        zb = min(D, za + planes)
        n = (zb - za) * plane
        p0 = sol.debug_read_range("p", za * plane, n)
        pre = sol.debug_read_range("pre", za * plane, n)
        ap = sol.debug_read_range("Ap", za * plane, n)
        sub = [np.ascontiguousarray(q.reshape(D * plane, -1)[za * plane:zb * plane]) if isinstance(q, np.ndarray) else q for q in pr.params]
        o = OracleSolver([W, H, zb - za], pr.energy, pr.kind, double=pr.double, threads=threads)
        r_ref, pre_ref = o.eval_jtf(sub, dt)
        ap_ref = o.apply_jtj(sub, np.ascontiguousarray(p0.reshape(-1)))
        o.close()
        U = p0.shape[1]
        r_ref, pre_ref, ap_ref = r_ref.reshape(-1, U), pre_ref.reshape(-1, U), ap_ref.reshape(-1, U)
        lo = 0 if (za == 0 and bottom) else 1                     # first / last plane of the sub-volume miss a z-neighbour unless
        hi = (zb - za) if (zb == D and top) else (zb - za - 1)     # they are the boundary of the full grid
        sl = slice(lo * plane, hi * plane)
        p0_ref = (pre_ref * r_ref).astype(dt)
        e_p = float(np.abs(p0[sl] - p0_ref[sl]).max() / np.abs(p0_ref[sl]).max())
        e_m = float(np.abs(pre[sl] - pre_ref[sl]).max() / np.abs(pre_ref[sl]).max())
        e_a = float(np.abs(ap[sl] - ap_ref[sl]).max() / np.abs(ap_ref[sl]).max())
        out["planes"].append({"range": [za + lo, za + hi], "rel_err_p0": e_p, "rel_err_pre": e_m, "rel_err_JTJp": e_a})
        out["max_rel_err_p0"] = max(out["max_rel_err_p0"], e_p)
        out["max_rel_err_pre"] = max(out["max_rel_err_pre"], e_m)
        out["max_rel_err_JTJp"] = max(out["max_rel_err_JTJp"], e_a)
    out["ok"] = bool(max(out["max_rel_err_p0"], out["max_rel_err_pre"], out["max_rel_err_JTJp"]) <= gate)
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="volumetric_512", choices=sorted(WORKLOADS))
    ap.add_argument("--size", type=int, nargs="+", default=None, help="override the workload's grid size")
    ap.add_argument("--double", action="store_true", help="run the workload in fp64 (doublePrecision = 1)")
    ap.add_argument("--kind", default=None, choices=["gaussNewtonGPU", "LMGPU"], help="override the workload's solver kind")
    ap.add_argument("--materialized", type=int, default=0, choices=[0, 1, 2],
                    help="Opt_InitializationParameters.useMaterializedJTJ (1) + useFusedJTJ (2): PCGStep1 through stored CSR matrices")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--cpu-budget", type=float, default=20.0)
    ap.add_argument("--cpu-threads", type=int, default=0, help="threads of the cpu_baseline / reference legs (0 = all host cores; 1 = the serial backend_cpu row)")
    args = ap.parse_args()
    global CPU_THREADS
    CPU_THREADS = args.cpu_threads
    load_traffic()
    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    def variant(pr):
        if args.kind:
            pr.kind = args.kind
        return pr.as_double() if args.double else pr

    if args.impl == "reference":
        if rank == 0:
            run_reference(args, WORKLOADS[args.workload])
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
