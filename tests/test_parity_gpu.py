"""GPU: the CUDA engine (through the C-ABI) against the CPU oracle on identical seeded inputs, and
against the reference's golden final costs."""
import os

import numpy as np
import pytest

from opt_b200 import examples
from gpu_util import (OP_RTOL, assert_costs_close, assert_unknowns_close, run_gpu, run_oracle)

pytestmark = pytest.mark.gpu

CASES = {
    "poisson": (examples.poisson_synthetic, (67, 45), dict(n_iter=2, l_iter=12)),
    "poisson_aligned": (examples.poisson_synthetic, (128, 64), dict(n_iter=1, l_iter=20)),
    # rows of whole quads, two strips (the second one ragged: 37 quads) and six bands of the marching J^T J kernel
    "poisson_strips": (examples.poisson_synthetic, (148, 83), dict(n_iter=2, l_iter=12)),
    "image_warping": (examples.image_warping_synthetic, (61, 47), dict(n_iter=3, l_iter=10)),
    "volumetric": (examples.volumetric_synthetic, (11, 9, 7), dict(n_iter=3, l_iter=10)),
    # several 30x4x4 (fp32) / 30x4x2 (fp64) tiles of the tiled J^T J kernel per dimension, ragged edges
    "volumetric_tiles": (examples.volumetric_synthetic, (67, 10, 9), dict(n_iter=2, l_iter=8)),
    # rows that are a multiple of 32 sites: the packed fp32 kernel fed by bulk copies (ragged tiles in x and y, two
    # z segments, 30 x 16 tiles: 3 x 2 of them)
    "volumetric_packed": (examples.volumetric_synthetic, (64, 19, 37), dict(n_iter=2, l_iter=8)),
    "optical_flow": (examples.optical_flow_synthetic, (53, 38), dict(n_iter=3, l_iter=10)),
    "arap": (examples.arap_synthetic, (13, 11), dict(n_iter=3, l_iter=10)),
    "sfs": (examples.sfs_synthetic, (45, 37), dict(n_iter=3, l_iter=10)),
    # rows of whole quads: the marching J^T J kernel (two strips of 120 written pixels, five bands of 8 rows)
    "sfs_quads": (examples.sfs_synthetic, (132, 37), dict(n_iter=3, l_iter=10)),
    "embedded": (examples.embedded_synthetic, (12, 10), dict(n_iter=3, l_iter=10)),
    "cotangent": (examples.cotangent_synthetic, (12, 10), dict(n_iter=3, l_iter=10)),
    "robust": (examples.robust_synthetic, (12, 10), dict(n_iter=3, l_iter=10)),
    "intrinsic": (examples.intrinsic_synthetic, (43, 29), dict(n_iter=3, l_iter=10)),
    # Well-conditioned variants of the two cases whose fp32 gates are widened below (the fp32 and fp64 ORACLE runs themselves
    # differ by more than the standard gate on those): the same code paths, held to the standard gates on every step.
    # pNorm = 1.6: the re-weighting exponent is -0.4 instead of -1.2; interior rings on an equilateral lattice: no triangle of
    # nearly collinear edge vectors.
    "intrinsic_p16": (examples.intrinsic_synthetic, (43, 29), dict(n_iter=4, l_iter=10, pnorm=1.6)),
    "cotangent_interior": (examples.cotangent_synthetic, (12, 10), dict(n_iter=3, l_iter=10, interior_rings=True)),
}


def make(case, kind="gaussNewtonGPU", double=False):
    b, args, kw = CASES[case]
    pr = b(*args, **kw)
    pr.kind = kind
    return pr.as_double() if double else pr


@pytest.mark.parametrize("double", [False, True])
@pytest.mark.parametrize("kind", ["gaussNewtonGPU", "LMGPU"])
@pytest.mark.parametrize("case", sorted(CASES))
def test_solve_matches_oracle(built, case, kind, double):
    pr = make(case, kind, double)
    gc, gx, sol, _ = run_gpu(pr)
    rc, rx, _ = run_oracle(pr)
    if case == "intrinsic":
        # The L_p weights pow(|r_c - r_n| + 1e-7, pNorm - 2) are refreshed from the current r before every
        # nonlinear step (lib.t:112-121) and the albedo prior drives |r_c - r_n| towards 0, so from the second
        # step on the weights amplify the rounding of r (and the last-ulp differences between the device's and
        # libm's pow) by 1.2 / |r_c - r_n|.  The first step holds BASELINE's gates (1e-4 fp32 / 1e-9 fp64); the
        # re-weighted steps are held to 2e-2 (fp32) and 1e-7 (fp64).  That is the conditioning of the energy at the driver's
        # pNorm = 0.8, not of this implementation: the ORACLE's own fp32 and fp64 runs of this case are 9.5e-3 and 1.3e-2 apart
        # after steps 2 and 3.  "intrinsic_p16" runs the same code at pNorm = 1.6 and holds the standard gates on every step.
        assert_costs_close(gc[:2], rc[:2], double)
        assert_costs_close(gc[2:], rc[2:], double, tol=1e-7 if double else 2e-2)
        assert_unknowns_close(gx, rx, double, tol=1e-6 if double else 2e-2)
        return
    assert_costs_close(gc, rc, double)
    assert_unknowns_close(gx, rx, double)
    if not double:
        # fp32 is additionally held to 1e-4 of the fp64 oracle on the same (fp32-valued) inputs
        rc64, _, _ = run_oracle(pr.as_double())
        assert_costs_close(gc, rc64, False)
    assert sol.launch_count() > 0


@pytest.mark.parametrize("double", [False, True])
@pytest.mark.parametrize("case", sorted(CASES))
def test_jtf_and_jtj_match_oracle(built, case, double):
    """One application of the operators at the initial unknowns: r = -J^T F, the preconditioner and
    J^T J p for a random p (zero at excluded sites)."""
    from opt_b200 import solver as S
    from oracle_api import OracleSolver
    pr = make(case, double=double)
    named = S.upload(pr)
    sol = S.OptSolver(pr.dims, pr.energy, pr.kind, double=double)
    o = OracleSolver(pr.dims, pr.energy, pr.kind, double=double, threads=4)
    dt = np.float64 if double else np.float32
    sol.debug_pcg_begin(named)
    r_gpu, pre_gpu = sol.debug_read("r"), sol.debug_read("pre")
    r_ref, pre_ref = o.eval_jtf(pr.params, dt)
    tol = OP_RTOL[double]
    if case == "cotangent" and not double:
        # disc = (a.a)(b.b) - (a.b)^2 of the wrapped boundary rings is a catastrophic cancellation in fp32 (nearly
        # collinear unit vectors): closed form and dual numbers round differently there; fp64 agrees to 1e-12.  The oracle's
        # own fp32 and fp64 evaluations of J^T F differ by 5.5e-5 on this mesh; "cotangent_interior" (no wrapped rings) holds 2e-5
        tol = 5e-4
    assert np.abs(r_gpu - r_ref).max() <= tol * np.abs(r_ref).max()
    assert np.abs(pre_gpu - pre_ref).max() <= tol * np.abs(pre_ref).max()
    rng = np.random.default_rng(7)
    p = rng.standard_normal(r_ref.shape).astype(dt) * (r_ref != 0)      # excluded sites carry p = 0
    ap_gpu = sol.debug_apply_jtj(named, p)
    ap_ref = o.apply_jtj(pr.params, p)
    assert np.abs(ap_gpu - ap_ref).max() <= tol * np.abs(ap_ref).max()


@pytest.mark.parametrize("case,double", [("volumetric_tiles", False), ("volumetric_tiles", True), ("volumetric_packed", False), ("volumetric_packed", True)])
def test_volumetric_tiled_jtj_equals_generic(built, case, double, monkeypatch):
    """The marching kernels (precomputed trig table; fp64 scalar, fp32 packed pairs fed by bulk copies) against the
    generic one-thread-per-site functor kernel: two different derivations of the same operator."""
    from opt_b200 import solver as S
    pr = make(case, double=double)
    named = S.upload(pr)
    dt = np.float64 if double else np.float32
    rng = np.random.default_rng(11)
    N = int(np.prod(pr.dims))
    p = rng.standard_normal((N, 6)).astype(dt)
    sol = S.OptSolver(pr.dims, pr.energy, pr.kind, double=double)
    fast = sol.debug_apply_jtj(named, p)
    monkeypatch.setenv("OPTB200_GENERIC_JTJ", "1")
    generic = sol.debug_apply_jtj(named, p)
    assert not np.array_equal(fast, generic)          # really two code paths
    assert np.abs(fast - generic).max() <= OP_RTOL[double] * np.abs(generic).max()


@pytest.mark.parametrize("double", [False, True])
@pytest.mark.parametrize("case", ["poisson_aligned", "poisson_strips"])
def test_poisson_marching_jtj_equals_tile_kernel(built, case, double, monkeypatch):
    """k_poisson_march (one channel plane per warp, rows carried in registers, cp.async ring, per-pixel bits of
    active4 instead of the mask image) against the tile kernel k_poisson_apply, GN and LM solves included."""
    from opt_b200 import solver as S
    pr = make(case, double=double)
    rng = np.random.default_rng(9)
    unknown = (pr.params[2].reshape(-1) == 0)
    p = (rng.standard_normal((int(np.prod(pr.dims)), 4)) * unknown[:, None]).astype(np.float64 if double else np.float32)
    march = S.OptSolver(pr.dims, pr.energy, pr.kind, double=double).debug_apply_jtj(S.upload(pr), p)
    monkeypatch.setenv("OPTB200_POISSON_TILED", "1")          # read when the plan is made
    tiled = S.OptSolver(pr.dims, pr.energy, pr.kind, double=double).debug_apply_jtj(S.upload(pr), p)
    assert np.abs(march).max() > 0
    assert np.abs(march - tiled).max() <= OP_RTOL[double] * np.abs(tiled).max()
    for kind in ("gaussNewtonGPU", "LMGPU"):
        q = make(case, kind, double)
        ct, xt, _, _ = run_gpu(q)
        monkeypatch.delenv("OPTB200_POISSON_TILED")
        cm, xm, _, _ = run_gpu(q)
        monkeypatch.setenv("OPTB200_POISSON_TILED", "1")
        assert_costs_close(cm, ct, double)
        assert_unknowns_close(xm, xt, double)


@pytest.mark.parametrize("double", [False, True])
@pytest.mark.parametrize("kind", ["gaussNewtonGPU", "LMGPU"])
def test_sfs_marching_jtj_equals_generic(built, kind, double, monkeypatch):
    """k_sfs_march (register pipeline over rows, shuffles across lanes) against the per-site functor kernel."""
    from opt_b200 import solver as S
    pr = make("sfs_quads", kind=kind, double=double)
    named = S.upload(pr)
    rng = np.random.default_rng(5)
    p = rng.standard_normal((int(np.prod(pr.dims)), 1)).astype(np.float64 if double else np.float32)
    sol = S.OptSolver(pr.dims, pr.energy, pr.kind, double=double)
    fast = sol.debug_apply_jtj(named, p)
    monkeypatch.setenv("OPTB200_GENERIC_JTJ", "1")
    generic = sol.debug_apply_jtj(named, p)
    assert not np.array_equal(fast, generic)
    assert np.abs(fast - generic).max() <= OP_RTOL[double] * np.abs(generic).max()


def test_volumetric_packed_equals_scalar_marching_kernel(built, monkeypatch):
    """fp32: the packed-pair kernel on blocked vectors against the scalar marching kernel on SoA vectors (the plan
    picks the variant; OPTB200_VOL_SCALAR is read when the plan is made), GN and LM solves included."""
    from opt_b200 import solver as S
    pr = make("volumetric_packed")
    rng = np.random.default_rng(3)
    p = rng.standard_normal((int(np.prod(pr.dims)), 6)).astype(np.float32)
    packed = S.OptSolver(pr.dims, pr.energy, pr.kind).debug_apply_jtj(S.upload(pr), p)
    monkeypatch.setenv("OPTB200_VOL_SCALAR", "1")
    scalar = S.OptSolver(pr.dims, pr.energy, pr.kind).debug_apply_jtj(S.upload(pr), p)
    assert not np.array_equal(packed, scalar)
    assert np.abs(packed - scalar).max() <= OP_RTOL[False] * np.abs(scalar).max()
    for kind in ("gaussNewtonGPU", "LMGPU"):
        q = make("volumetric_packed", kind)
        cs, _, _, _ = run_gpu(q)
        monkeypatch.delenv("OPTB200_VOL_SCALAR")
        cp, _, _, _ = run_gpu(q)
        monkeypatch.setenv("OPTB200_VOL_SCALAR", "1")
        assert_costs_close(cp, cs, False)


@pytest.mark.parametrize("double", [False, True])
def test_arap_gather_equals_atomic_scatter(built, double, monkeypatch):
    """Graph J^T J p: the per-vertex gather over the sorted incidence lists against the reference-style
    edge kernel with global atomics."""
    from opt_b200 import solver as S
    pr = make("arap", double=double)
    named = S.upload(pr)
    dt = np.float64 if double else np.float32
    rng = np.random.default_rng(5)
    p = rng.standard_normal((pr.dims[0], 6)).astype(dt)
    sol = S.OptSolver(pr.dims, pr.energy, pr.kind, double=double)
    gathered = sol.debug_apply_jtj(named, p)
    monkeypatch.setenv("OPTB200_GRAPH_ATOMICS", "1")
    sol2 = S.OptSolver(pr.dims, pr.energy, pr.kind, double=double)
    scattered = sol2.debug_apply_jtj(named, p)
    assert np.abs(gathered - scattered).max() <= OP_RTOL[double] * np.abs(scattered).max()


GOLDEN = [("image_warping_cat512.npz", examples.image_warping_golden),
          ("volumetric_head.npz", examples.volumetric_golden),
          ("intrinsic_ye_high2_stride12.npz", examples.intrinsic_golden),
          ("optical_flow_dogdance_stride16.npz", examples.optical_flow_golden),
          ("cotangent_head.npz", examples.cotangent_golden),
          ("embedded_raptor2k.npz", examples.embedded_golden),
          ("arap_small_armadillo.npz", examples.arap_golden),
          ("poisson_stride4.npz", examples.poisson_golden)]


@pytest.mark.parametrize("fname,builder", GOLDEN)
def test_reference_golden_final_cost(built, golden_dir, fname, builder):
    """examples/test_final_cost.py:56-66 with its own acceptance threshold (rel. error < 1e-5)."""
    d = np.load(os.path.join(golden_dir, fname))
    pr = builder(d)
    gc, _, _, _ = run_gpu(pr)
    golden = float(d["golden_cost"])
    assert abs(gc[-1] - golden) / golden < 1e-5, (gc, golden)


def test_final_cost_line_and_timing_table(built, capfd):
    """stdout contract: `final cost=` line (solver:1903) that the reference's tests regex."""
    import re
    from opt_b200 import solver as S
    pr = make("poisson")
    named = S.upload(pr)
    sol = S.OptSolver(pr.dims, pr.energy, pr.kind, verbosity=1, timing=True)
    cost = sol.solve(pr.solver, named)
    out = capfd.readouterr().out
    m = re.search("final cost=(.*)", out)
    assert m and abs(float(m.group(1)) - cost) <= 1e-6 * cost
    assert "PCGStep1" in out and "PCGStep2" in out


def test_solve_equals_init_step_loop(built):
    from opt_b200 import solver as S
    pr = make("image_warping")
    a = S.OptSolver(pr.dims, pr.energy, pr.kind)
    ca = a.solve_blocking(pr.solver, S.upload(pr))
    gc, _, _, _ = run_gpu(pr)
    assert abs(ca - gc[-1]) <= 1e-6 * abs(ca)


LM_EXIT_CASES = {   # LM solves whose q_tolerance test fires (oracle: 112 of 120, 119 of 160, 98 of 120 iterations)
    "sfs": (examples.sfs_synthetic, (45, 37), dict(n_iter=4, l_iter=30)),
    "volumetric": (examples.volumetric_synthetic, (11, 9, 7), dict(n_iter=4, l_iter=40)),
    "poisson": (examples.poisson_synthetic, (67, 45), dict(n_iter=3, l_iter=40)),
}


@pytest.mark.parametrize("double", [False, True])
@pytest.mark.parametrize("case", sorted(LM_EXIT_CASES))
def test_lm_device_exit_test(built, case, double, monkeypatch):
    """The LM inner loop with the exit test on the device (one iteration queued ahead of the verdict, its step
    neutralised by the stop flag) against the reference-style loop with a blocking fetchQ per iteration and
    against the oracle: same number of PCG iterations, same costs, same unknowns."""
    b, args, kw = LM_EXIT_CASES[case]
    pr = b(*args, **kw)
    pr.kind = "LMGPU"
    if double:
        pr = pr.as_double()
    gc, gx, sol, _ = run_gpu(pr)
    rc, rx, o = run_oracle(pr)
    nominal = kw["n_iter"] * kw["l_iter"]
    assert o.pcg_iterations() < nominal                  # the test really fires in this case
    if double:
        assert sol.pcg_iterations() == o.pcg_iterations()
    assert_costs_close(gc, rc, double)
    assert_unknowns_close(gx, rx, double)
    monkeypatch.setenv("OPTB200_LM_HOSTLOOP", "1")       # read when the plan is made
    hc, hx, hsol, _ = run_gpu(pr)
    assert hsol.pcg_iterations() == sol.pcg_iterations()
    assert_costs_close(gc, hc, double, tol=1e-6 if not double else 1e-12)
    assert_unknowns_close(gx, hx, double, tol=1e-5 if not double else 1e-11)


@pytest.mark.parametrize("name", ["shape_from_shading", "volumetric_mesh_deformation", "image_warping", "arap_mesh_deformation",
                                  "optical_flow", "poisson_image_editing", "intrinsic_image_decomposition"])
def test_reference_long_run_cost(built, golden_dir, name):
    """examples/test_example_output.py: the recorded long-run costs with the reference's own tolerances (58.968529
    +- 1e-2 after GN 60x10 on the 640x480 shape-from-shading data; 19.287743 +- 1e-6 after GN 20x60; 0.078505 +- 1e-4
    after the first of 19 outer iterations of GN 8x400; 3.929534 +- 1e-5 after the first of 10 outer iterations of ARAP
    GN 20x100; 4055.09033 +- 1e-5 after the first solve of full-size optical flow GN 1x50), and the two full-size image
    runs whose reference tolerance is one fp32 ulp, held to 2e-6 / 5e-6 (see test_oracle_golden.LONG_RUNS)."""
    from test_oracle_golden import LONG_RUNS
    fname, builder, kw, golden, tol = LONG_RUNS[name]
    pr = builder(np.load(os.path.join(golden_dir, fname)), **kw)
    gc, _, sol, _ = run_gpu(pr)
    assert abs(gc[-1] - golden) / golden < tol, (gc[-1], golden)
    assert sol.pcg_iterations() == kw["n_iter"] * kw["l_iter"]


def test_embedded_31_outer_iterations(built, golden_dir):
    """examples/test_example_output.py:63,112: embedded_mesh_deformation --oIterations 31 --nIterations 5 --lIterations 125,
    recorded 0.0944651290 (niessner/Opt CUDA).  That figure is the final cost of the LAST outer iteration (markers moved all the
    way); the oracle reaches 0.09448624 in fp32 and 0.09448626 in fp64 there (tests/test_oracle_golden.py, slow case), 2.2e-4
    from the recorded value with no sensitivity to precision or to the initial rotation - so this anchor is held to 5e-4 against
    the recorded figure, not to the reference's 1e-5, and to 1e-4 against the oracle's value."""
    import torch
    from opt_b200 import solver as S
    from test_oracle_golden import EMBEDDED_31_ORACLE, EMBEDDED_31_RECORDED
    pr = examples.embedded_golden(np.load(os.path.join(golden_dir, "embedded_raptor2k.npz")), n_iter=5, l_iter=125)
    targets = [examples.embedded_outer_constraints(pr, i, 31) for i in range(31)]
    named = S.upload(pr)
    sol = S.OptSolver(pr.dims, pr.energy, pr.kind)
    last = None
    for z in targets:
        named.values[6].copy_(torch.from_numpy(z))
        costs = []
        sol.solve(pr.solver, named, costs)
        last = costs
    torch.cuda.synchronize()
    assert sol.pcg_iterations() == 31 * 5 * 125
    assert abs(last[-1] - EMBEDDED_31_ORACLE) <= 1e-4 * EMBEDDED_31_ORACLE, last
    assert abs(last[-1] - EMBEDDED_31_RECORDED) <= 5e-4 * EMBEDDED_31_RECORDED, last


TINY = [("poisson", examples.poisson_synthetic, s) for s in [(4, 1), (1, 5), (4, 4), (8, 2), (12, 3), (5, 2)]] + \
       [("optical_flow", examples.optical_flow_synthetic, s) for s in [(4, 3), (8, 2), (4, 2)]] + \
       [("image_warping", examples.image_warping_synthetic, s) for s in [(3, 3), (4, 2), (2, 5)]] + \
       [("sfs", examples.sfs_synthetic, s) for s in [(4, 4), (8, 3), (4, 9), (5, 5)]] + \
       [("volumetric", examples.volumetric_synthetic, s) for s in [(2, 2, 2), (32, 2, 3), (3, 1, 4)]]


@pytest.mark.parametrize("name,builder,size", TINY, ids=["%s-%s" % (n, "x".join(map(str, s))) for n, _, s in TINY])
def test_tiny_grids(built, name, builder, size):
    """Degenerate sizes: single rows / columns / quads, grids smaller than one tile, strip or band of the specialised
    kernels (the marching kernels must handle H = 1, a single quad per row, bands cut by the image border)."""
    pr = builder(*size, n_iter=1, l_iter=3).as_double()
    rc, rx, _ = run_oracle(pr)
    if not np.all(np.isfinite(rc)) or rc[-1] < 1e-9 * rc[0]:
        pytest.skip("degenerate system: solved exactly / alphaDen = 0 (the reference divides by it as well)")
    gc, gx, _, _ = run_gpu(pr)
    assert_costs_close(gc, rc, True)
    assert_unknowns_close(gx, rx, True)


@pytest.mark.parametrize("kind", ["gaussNewtonGPU", "LMGPU"])
def test_arap_fused_gather_equals_unfused(built, kind, monkeypatch):
    """k_gather_jtj_fused (PCGStep3 folded into the per-vertex gather; off by default because it does not pay on B200)
    against the default gather + separate PCGStep3."""
    pr = make("arap", kind)
    cu, xu, _, _ = run_gpu(pr)
    monkeypatch.setenv("OPTB200_GATHER_FUSED", "1")           # read when the plan is made
    cf, xf, _, _ = run_gpu(pr)
    assert_costs_close(cf, cu, False)
    assert_unknowns_close(xf, xu, False)


@pytest.mark.parametrize("kind", ["gaussNewtonGPU", "LMGPU"])
def test_volumetric_fused_step3_equals_unfused(built, kind, monkeypatch):
    """The packed kernel with PCGStep3 folded in (z staged next to p, p' = z + beta p formed in registers, written once)
    against the same kernel behind a separate PCGStep3 pass (OPTB200_NO_FUSE=1), and both against the oracle."""
    pr = make("volumetric_packed", kind)
    cf, xf, sol, _ = run_gpu(pr, timing=True)
    assert sol.kernel_time("PCGStep3+1")[0] > 0 and sol.kernel_time("PCGStep3")[0] == 0
    monkeypatch.setenv("OPTB200_NO_FUSE", "1")
    cu, xu, sol2, _ = run_gpu(pr, timing=True)
    assert sol2.kernel_time("PCGStep3+1")[0] == 0 and sol2.kernel_time("PCGStep3")[0] > 0
    assert_costs_close(cf, cu, False, tol=1e-5)
    assert_unknowns_close(xf, xu, False, tol=1e-4)
    rc, rx, _ = run_oracle(pr)
    assert_costs_close(cf, rc, False)
    assert_unknowns_close(xf, rx, False)


@pytest.mark.parametrize("case,mask_index,value", [("poisson_strips", 2, 255.0), ("sfs_quads", 17, 0.0), ("image_warping", 4, 255.0)])
def test_plan_reused_with_another_exclude_mask(built, case, mask_index, value):
    """a plan solved twice with two different Exclude masks must give what a fresh plan gives (sites that become
    excluded keep no stale delta / r / p / z / Ap: the engine clears the solver vectors when the mask changes)."""
    from opt_b200 import solver as S
    import torch
    pr1 = make(case)
    pr2 = make(case)
    W, H = pr2.dims[0], pr2.dims[1]
    m = pr2.params[mask_index].reshape(H, W)
    assert (m[12:24, 30:50] != value).any()             # these pixels are unknowns in the first solve ...
    m[12:24, 30:50] = value                             # ... and excluded in the second (M != 0 / D_i <= 0)
    sol = S.OptSolver(pr1.dims, pr1.energy, pr1.kind)
    n1 = S.upload(pr1)
    sol.solve(pr1.solver, n1)
    n2 = S.upload(pr2)
    c_reused = []
    sol.solve(pr2.solver, n2, c_reused)
    torch.cuda.synchronize()
    x_reused = {i: n2.values[i].cpu().numpy() for i in pr2.unknowns}
    c_fresh, x_fresh, _, _ = run_gpu(pr2)
    # block-level double atomics make two runs agree to rounding, not bit for bit; stale vectors show up at 1e-2 and more
    assert_costs_close(c_reused, c_fresh, False, tol=1e-6)
    assert_unknowns_close(x_reused, x_fresh, False, tol=1e-5)


@pytest.mark.parametrize("double", [False, True])
@pytest.mark.parametrize("case", ["embedded", "cotangent", "robust"])
def test_graph_segmented_scatter_equals_atomics(built, case, double, monkeypatch):
    """J^T J p of the hyper-edge energies as a segmented sum over vertex-sorted (edge, slot) partials (the default: no atomicAdd
    into Ap, deterministic) against the reference's form, one global atomic per edge, slot and channel (OPTB200_GRAPH_ATOMICS=1)."""
    from opt_b200 import solver as S
    pr = make(case, double=double)
    dt = np.float64 if double else np.float32
    rng = np.random.default_rng(13)
    U = {"embedded": 12, "cotangent": 3, "robust": None}[case]
    sol = S.OptSolver(pr.dims, pr.energy, pr.kind, double=double, timing=True)
    U = sol.unknown_channels()
    p = rng.standard_normal((pr.dims[0], U)).astype(dt)
    named = S.upload(pr)
    seg = sol.debug_apply_jtj(named, p)
    seg2 = sol.debug_apply_jtj(named, p)
    assert np.array_equal(seg, seg2)                    # fixed summation order: bit-reproducible
    monkeypatch.setenv("OPTB200_GRAPH_ATOMICS", "1")
    sol_a = S.OptSolver(pr.dims, pr.energy, pr.kind, double=double)
    atom = sol_a.debug_apply_jtj(S.upload(pr), p)
    assert np.abs(seg - atom).max() <= OP_RTOL[double] * np.abs(atom).max()
    monkeypatch.delenv("OPTB200_GRAPH_ATOMICS")
    cs, xs, s2, _ = run_gpu(pr, timing=True)
    assert s2.kernel_time("PCGStep1_Graph")[0] > 0
    monkeypatch.setenv("OPTB200_GRAPH_ATOMICS", "1")
    ca, xa, _, _ = run_gpu(pr)
    assert_costs_close(cs, ca, double, tol=1e-5 if not double else 1e-10)


@pytest.mark.parametrize("l_iter", [7, 8])
@pytest.mark.parametrize("case", ["poisson_strips", "volumetric_packed", "optical_flow", "sfs_quads", "image_warping"])
def test_deferred_delta_update_is_bit_identical(built, case, l_iter, monkeypatch):
    """k_pcg_step2's deferred form (delta updated every second PCG iteration for two directions at once; odd and even loop lengths)
    performs the same two FMAs per element as the eager form: costs and unknowns must agree bit for bit wherever the reductions are
    order-independent, and to rounding of the block-partial sums otherwise."""
    b, args, kw = CASES[case]
    pr = b(*args, **dict(kw, l_iter=l_iter))
    lazy_c, lazy_x, sol, _ = run_gpu(pr, timing=True)
    assert sol.kernel_time("PCGStep3+1")[0] > 0              # the fused path, where the deferred form is active
    monkeypatch.setenv("OPTB200_NO_LAZY_DELTA", "1")
    eager_c, eager_x, _, _ = run_gpu(pr)
    assert_costs_close(lazy_c, eager_c, False, tol=1e-6)
    assert_unknowns_close(lazy_x, eager_x, False, tol=1e-5)
    rc, rx, _ = run_oracle(pr)
    assert_costs_close(lazy_c, rc, False)
    assert_unknowns_close(lazy_x, rx, False)


@pytest.mark.parametrize("double", [False, True])
def test_sfs_marching_model_cost_equals_generic(built, double, monkeypatch):
    """computeModelCost on the marching pipeline (MODE 4 of k_sfs_march: the cost pass on the linearised residuals) against the
    generic per-site kernel: the LM trust-region decisions and every per-step cost must coincide, and both match the oracle."""
    pr = make("sfs_quads", "LMGPU", double)
    pr.solver["nIterations"] = 5
    cm, xm, sol, _ = run_gpu(pr, timing=True)
    assert sol.kernel_time("computeModelCost")[0] > 0
    monkeypatch.setenv("OPTB200_GENERIC_MODEL_COST", "1")
    cg, xg, _, _ = run_gpu(pr)
    assert_costs_close(cm, cg, double, tol=1e-6 if not double else 1e-12)
    assert_unknowns_close(xm, xg, double, tol=1e-5 if not double else 1e-11)
    rc, rx, _ = run_oracle(pr)
    assert_costs_close(cm, rc, double)


@pytest.mark.parametrize("double", [False, True])
def test_sfs_diagonal_only_ctc_equals_generic(built, double, monkeypatch):
    """PCGComputeCtC through SfsFunctor::diagJTJ (the diagonal of J^T J alone: gradient images and weights, no X / B_I loads, no
    gradient) against the generic kernel that takes the diagonal out of evalJTF: same summation order, so the LM steps coincide."""
    pr = make("sfs_quads", "LMGPU", double)
    pr.solver["nIterations"] = 5
    cd, xd, _, _ = run_gpu(pr)
    monkeypatch.setenv("OPTB200_GENERIC_CTC", "1")
    cg, xg, _, _ = run_gpu(pr)
    assert_costs_close(cd, cg, double, tol=1e-6 if not double else 1e-12)
    assert_unknowns_close(xd, xg, double, tol=1e-5 if not double else 1e-11)
    rc, rx, _ = run_oracle(pr)
    assert_costs_close(cd, rc, double)


@pytest.mark.parametrize("double", [False, True])
@pytest.mark.parametrize("kind", ["gaussNewtonGPU", "LMGPU"])
def test_image_warping_rotation_table_equals_inline_sincos(built, kind, double, monkeypatch):
    """image_warping reads (sin A, cos A) of a site and its neighbours from a per-pixel table refreshed by precompute; with
    OPTB200_IW_SINCOS=1 (read at plan time) J^T F and J^T J p evaluate sincos inline as before.  Same function on the same
    operands: the solves must coincide, and both match the oracle."""
    pr = make("image_warping", kind, double)
    ct, xt, _, _ = run_gpu(pr)
    monkeypatch.setenv("OPTB200_IW_SINCOS", "1")
    ci, xi, _, _ = run_gpu(pr)
    assert_costs_close(ct, ci, double, tol=1e-6 if not double else 1e-12)
    assert_unknowns_close(xt, xi, double, tol=1e-5 if not double else 1e-11)
    rc, rx, _ = run_oracle(pr)
    assert_costs_close(ct, rc, double)
    assert_unknowns_close(xt, rx, double)
