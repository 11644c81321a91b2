"""Host-side input builders for the shipped examples.

Each builder returns a `Problem`: the energy-file name, the `dimensions` array and the flat
`problemparams` list in the positions the energy file fixes, as host
numpy arrays / scalars.  They restate what the reference's example drivers do before calling
`Opt_ProblemSolve` (cited per function); synthetic builders follow SURVEY.md section 8(d).
Nothing here computes: arrays are handed to the C-ABI library (device copies are made by
`opt_b200.solver.OptSolver`) or, in tests, to the CPU oracle.
"""
from dataclasses import dataclass, field

import numpy as np


@dataclass
class Problem:
    energy: str                  # e.g. "image_warping.t"
    dims: list                   # Opt_ProblemPlan dimensions
    params: list                 # problemparams; np.ndarray | np.float32 scalar | None
    unknowns: list               # indices into params that are unknown images
    solver: dict = field(default_factory=dict)   # solver parameters (nIterations, lIterations, ...)
    kind: str = "gaussNewtonGPU"
    double: bool = False
    opt_float_scalars: list = field(default_factory=list)   # indices of Param(..., opt_float, ...) scalars

    def clone(self):
        return Problem(self.energy, list(self.dims),
                       [p.copy() if isinstance(p, np.ndarray) else p for p in self.params],
                       list(self.unknowns), dict(self.solver), self.kind, self.double, list(self.opt_float_scalars))

    def as_double(self):
        """opt_float images become double; Param(float) scalars and index/uint8 arrays stay
        (o.t:2430)."""
        q = self.clone()
        q.double = True
        q.params = [p.astype(np.float64) if isinstance(p, np.ndarray) and p.dtype == np.float32 else p
                    for p in q.params]
        for i in q.opt_float_scalars:
            q.params[i] = np.float64(q.params[i])
        return q


# ---------------------------------------------------------------------------------------------
def image_warping(mask, constraints, alpha=1.0, n_iter=1, l_iter=1):
    """examples/image_warping/src/CombinedSolver.h:129-134 (weights), :165-197 (resetGPU),
    :199-227 (setConstraintImage); main.cpp:157-177 (border pixels pinned)."""
    H, W = mask.shape
    maskf = mask.astype(np.float32)
    ys, xs = np.mgrid[0:H, 0:W]
    ur = np.stack([xs, ys], -1).astype(np.float32)
    offset = ur.copy()
    angle = np.full((H, W), 1e-5, dtype=np.float32)
    cons = [list(c) for c in constraints]
    for y in range(H):
        for x in range(W):
            if y == 0 or x == 0 or y == H - 1 or x == W - 1:
                cons.append([x, y, x, y])
    cimg = np.full((H, W, 2), -1.0, dtype=np.float32)
    a = np.float32(alpha)
    for x, y, tx, ty in cons:
        if maskf[y, x] == 0:
            cimg[y, x, 0] = (np.float32(1) - a) * np.float32(x) + a * np.float32(tx)
            cimg[y, x, 1] = (np.float32(1) - a) * np.float32(y) + a * np.float32(ty)
    w_fit, w_reg = np.sqrt(np.float32(100.0)), np.sqrt(np.float32(0.01))
    return Problem("image_warping.t", [W, H], [offset, angle, ur, cimg, maskf, w_fit, w_reg], [0, 1],
                   {"nIterations": n_iter, "lIterations": l_iter})


def image_warping_golden(npz, outer_iterations=1, n_iter=1, l_iter=1):
    """(1,1,1): the reference's recorded final cost (test_final_cost.py:58).  (19,8,400): the FIRST outer iteration
    of the long run (test_example_output.py:64,112,133), whose constraints are moved 1/19 of the way
    (CombinedSolver.h:155: alpha = (i+1)/numIter)."""
    shape = tuple(npz["shape"])
    is_zero = np.unpackbits(npz["mask_is_zero"])[: shape[0] * shape[1]].reshape(shape).astype(bool)
    mask = np.where(is_zero, 0, 255).astype(np.uint8)
    alpha = np.float32(1.0) / np.float32(outer_iterations)
    return image_warping(mask, npz["constraints"].tolist(), alpha=alpha, n_iter=n_iter, l_iter=l_iter)


def volumetric(bbox_min, bbox_max, grid=(5, 20, 5), n_iter=1, l_iter=1):
    """examples/volumetric_mesh_deformation/src/CombinedSolver.h:47-51 (weights), :81-108
    (setConstraints), :232-252 (initializeWarpGrid), :254-262 (bounding box + EPS).
    The driver stores node (i,j,k) at i*(Dy+1)*(Dz+1) + j*(Dz+1) + k and passes dims
    {Dx+1,Dy+1,Dz+1}; Opt reads d0 fastest, so the array is reproduced byte for byte."""
    f32 = np.float32
    gx, gy, gz = grid
    eps = f32(0.000001)
    mn = bbox_min.astype(f32) - eps
    mx = bbox_max.astype(f32) + eps
    delta = (mx - mn) / np.array([gx, gy, gz], dtype=f32)
    ii, jj, kk = np.meshgrid(np.arange(gx + 1), np.arange(gy + 1), np.arange(gz + 1), indexing="ij")
    fac = np.stack([ii, jj, kk], -1).astype(f32)
    pos = (mn + fac * delta).astype(f32)                  # [i][j][k][3], k fastest
    target = np.full(pos.shape, -np.inf, dtype=f32)
    target[:, 0] = pos[:, 0]
    ang = f32(np.deg2rad(np.float32(-90.0)))              # mLib matrix3x3.h:157-166 (degrees, ccw)
    s, c = f32(np.sin(ang)), f32(np.cos(ang))
    R = np.array([[c, -s, 0], [s, c, 0], [0, 0, 1]], dtype=f32)
    mid = mn + np.array([gx / 2.0, float(gy), gz / 2.0], dtype=f32) * delta
    v = pos[:, gy]
    target[:, gy] = ((v - mid) @ R.T + mid + np.array([2.5, -2.5, 0.0], dtype=f32)).astype(f32)
    n = (gx + 1) * (gy + 1) * (gz + 1)
    offset = pos.reshape(n, 3).copy()
    angle = np.zeros((n, 3), dtype=f32)
    w_fit, w_reg = np.sqrt(f32(1.0)), np.sqrt(f32(0.05))
    return Problem("volumetric_mesh_deformation.t", [gx + 1, gy + 1, gz + 1],
                   [offset, angle, offset.copy(), target.reshape(n, 3), w_fit, w_reg], [0, 1],
                   {"nIterations": n_iter, "lIterations": l_iter})


def volumetric_golden(npz, n_iter=1, l_iter=1):
    """(1,1): test_final_cost.py:66; (20,60): the long run of test_example_output.py:73,119,140."""
    return volumetric(npz["bbox_min"], npz["bbox_max"], tuple(int(g) for g in npz["grid"]), n_iter=n_iter, l_iter=l_iter)


def sfs_golden(npz, n_iter=60, l_iter=10, kind="gaussNewtonGPU"):
    """shape_from_shading on the reference's default data set (640 x 480; SFSSolverInput.h:27-62 gives the parameter
    order, main.cpp:8 the files): the long run of test_example_output.py:72,118,139 (recorded cost 58.968529, 1e-2)."""
    H, W = (int(v) for v in npz["shape"])
    D = npz["D"].astype(np.float32)
    X = D.copy() if bool(npz["X_is_D"]) else npz["X"].astype(np.float32)
    M = np.unpackbits(npz["masks"])[: 2 * H * W].reshape(2 * H, W).astype(np.uint8)
    scal = [np.float32(v) for v in npz["scalars"]]
    return Problem("shape_from_shading.t", [W, H], scal + [X, D, npz["Im"].astype(np.float32), M[:H].copy(), M[H:].copy()], [16],
                   {"nIterations": n_iter, "lIterations": l_iter}, kind=kind)


# ---------------------------------------------------------------------------------------------
# synthetic inputs of the BASELINE.json shapes, seed 1234
def poisson_synthetic(W, H, seed=1234, n_iter=1, l_iter=100, rows=None, period=None):
    """X,T in U[0,255] (4 channels, w = 255), mask = 0 (unknown) inside a centred rectangle covering
    50 % of the area, 255 elsewhere (examples/poisson_image_editing/src/CombinedSolver.h:79-82,
    main.cpp:102-106).  Every global row y has its own RNG stream, so `rows=(y0,y1)` builds just
    that band of the H-row image (multi-GPU slabs incl. ghost rows); with `period` the mask
    rectangle repeats every `period` rows (weak scaling: one rectangle per rank)."""
    y0, y1 = rows if rows else (0, H)
    period = period or H
    X = np.empty((y1 - y0, W, 4), dtype=np.float32)
    T = np.empty((y1 - y0, W, 4), dtype=np.float32)
    for y in range(y0, y1):
        rng = np.random.default_rng([seed, y])
        X[y - y0] = rng.uniform(0, 255, (W, 4))
        T[y - y0] = rng.uniform(0, 255, (W, 4))
    X[..., 3] = 255
    T[..., 3] = 255
    rw, rh = int(round(W * np.sqrt(0.5))), int(round(period * np.sqrt(0.5)))
    x0, r0 = (W - rw) // 2, (period - rh) // 2
    yy = np.arange(y0, y1) % period
    inside_y = (yy >= r0) & (yy < r0 + rh)
    M = np.full((y1 - y0, W), 255, dtype=np.float32)
    M[np.ix_(inside_y, np.arange(x0, x0 + rw))] = 0
    return Problem("poisson_image_editing.t", [W, y1 - y0], [X, T, M], [0],
                   {"nIterations": n_iter, "lIterations": l_iter})


def image_warping_synthetic(W, H, seed=1234, n_iter=2, l_iter=10, noise=0.3, w_fit=100.0, w_reg=1.0):
    """Disc-shaped mask, 8 random markers.  The default regularisation weight is 1 rather than the
    driver's 0.01: with 0.01 the normal equations are so ill-conditioned that two fp32 PCG runs that
    differ only in rounding disagree by ~1e-3 in cost after a few iterations (measured with the
    oracle, fp32 vs fp64), which would make a 1e-4 fp32 parity gate meaningless.  The driver's own
    weights are exercised by the golden test."""
    rng = np.random.default_rng(seed)
    ys, xs = np.mgrid[0:H, 0:W]
    mask = np.where(((xs - W / 2) ** 2 + (ys - H / 2) ** 2) < (0.45 * min(W, H)) ** 2, 0, 255).astype(np.uint8)
    cons = []
    for _ in range(8):
        x, y = int(rng.integers(W // 4, 3 * W // 4)), int(rng.integers(H // 4, 3 * H // 4))
        cons.append([x, y, x + int(rng.integers(-5, 6)), y + int(rng.integers(-5, 6))])
    pr = image_warping(mask, cons, n_iter=n_iter, l_iter=l_iter)
    pr.params[1] = (0.05 * rng.standard_normal((H, W))).astype(np.float32)          # non-trivial angles
    pr.params[0] = (pr.params[0] + noise * rng.standard_normal((H, W, 2))).astype(np.float32)
    pr.params[5], pr.params[6] = np.sqrt(np.float32(w_fit)), np.sqrt(np.float32(w_reg))
    return pr


def volumetric_synthetic(W, H, D, seed=1234, n_iter=2, l_iter=10, planes=None):
    """dims {W,H,D}; lattice in the unit-aspect box; plane d1==0 fixed, plane d1==H-1 rotated -90 deg
    about z and shifted.  Offsets/angles get small random values so that
    the rotation derivatives are exercised.  Every global d2-plane has its own RNG stream, so
    `planes=(z0,z1)` builds just that slab of the D-plane volume."""
    f32 = np.float32
    z0, z1 = planes if planes else (0, D)
    nz = z1 - z0
    kk, jj, ii = np.meshgrid(np.arange(z0, z1), np.arange(H), np.arange(W), indexing="ij")   # d0 fastest
    pos = np.stack([ii, jj, kk], -1).astype(f32) / f32(max(W, H, D) - 1)
    target = np.full(pos.shape, -np.inf, dtype=f32)
    target[:, 0] = pos[:, 0]
    R = np.array([[0, 1, 0], [-1, 0, 0], [0, 0, 1]], dtype=f32)
    mid = np.array([0.5 * (W - 1), H - 1, 0.5 * (D - 1)], dtype=f32) / f32(max(W, H, D) - 1)
    target[:, H - 1] = ((pos[:, H - 1] - mid) @ R.T + mid + np.array([0.25, -0.25, 0], dtype=f32)).astype(f32)
    offset = np.empty((nz, H, W, 3), dtype=f32)
    angle = np.empty((nz, H, W, 3), dtype=f32)
    for z in range(z0, z1):
        rng = np.random.default_rng([seed, z])
        offset[z - z0] = pos[z - z0] + 0.01 * rng.standard_normal((H, W, 3), dtype=f32)
        angle[z - z0] = 0.05 * rng.standard_normal((H, W, 3), dtype=f32)
    n = W * H * nz
    w_fit, w_reg = np.sqrt(f32(1.0)), np.sqrt(f32(0.05))
    return Problem("volumetric_mesh_deformation.t", [W, H, nz],
                   [offset.reshape(n, 3), angle.reshape(n, 3), pos.reshape(n, 3).copy(), target.reshape(n, 3), w_fit, w_reg], [0, 1],
                   {"nIterations": n_iter, "lIterations": l_iter})


def optical_flow_synthetic(W, H, seed=1234, n_iter=3, l_iter=20, kind="gaussNewtonGPU", shift=(0.4, -0.2)):
    """I = sum of low-frequency sinusoids in [0,255]; I_hat = I translated by `shift` px (sub-pixel, so plain Gauss-Newton stays in the linear regime);
    dx,dy by the 3x3 kernel of examples/optical_flow/src/CombinedSolver.h:152-178."""
    rng = np.random.default_rng(seed)
    ys, xs = np.mgrid[0:H, 0:W].astype(np.float64)

    fmax_x, fmax_y = max(1.0, min(4.0, W / 24.0)), max(1.0, min(4.0, H / 24.0))   # keep periods well above the shift

    def field(sx, sy):
        r = np.random.default_rng(seed + 1)
        out = np.zeros((H, W))
        for _ in range(8):
            fx, fy, ph = r.uniform(0.5, fmax_x), r.uniform(0.5, fmax_y), r.uniform(0, 2 * np.pi)
            out += np.sin(2 * np.pi * (fx * (xs + sx) / W + fy * (ys + sy) / H) + ph)
        return out

    lo, hi = field(0, 0).min(), field(0, 0).max()
    I = ((field(0, 0) - lo) / (hi - lo) * 255).astype(np.float32)
    Ih = ((field(shift[0], shift[1]) - lo) / (hi - lo) * 255).astype(np.float32)
    # computeDU / computeDV: 3x3 differences / 8 in the interior, 0 on the 1-pixel border
    dx = np.zeros((H, W), dtype=np.float32)
    dy = np.zeros((H, W), dtype=np.float32)
    c = Ih
    dx[1:-1, 1:-1] = ((c[:-2, 2:] + c[1:-1, 2:] + c[2:, 2:]) - (c[:-2, :-2] + c[1:-1, :-2] + c[2:, :-2])) / np.float32(8)
    dy[1:-1, 1:-1] = ((c[2:, :-2] + c[2:, 1:-1] + c[2:, 2:]) - (c[:-2, :-2] + c[:-2, 1:-1] + c[:-2, 2:])) / np.float32(8)
    X = np.zeros((H, W, 2), dtype=np.float32)            # driver starts from zero flow (resetFlowVectors)
    X[1:-1, 1:-1] = (0.1 * rng.standard_normal((H - 2, W - 2, 2))).astype(np.float32)   # off-grid sample positions
    w_fit, w_reg = np.sqrt(np.float32(10.0)), np.sqrt(np.float32(0.1))
    return Problem("optical_flow.t", [W, H], [w_fit, w_reg, X, I, Ih, dx, dy], [2],
                   {"nIterations": n_iter, "lIterations": l_iter}, kind=kind)


def grid_mesh_rings(nx, ny, interior_only=False):
    """(head, tail, prev, next) of every directed edge of the height-field triangulation, neighbours of a vertex in
    counter-clockwise ring order and prev/next taken cyclically in that ring
    (cotangent_mesh_smoothing/src/CombinedSolver.h:66-120).  interior_only: only vertices with their full ring of six emit
    edges (boundary vertices stay unknowns, held by the fitting term and by their interior neighbours' rings)."""
    ring = [(1, 0), (1, 1), (0, 1), (-1, 0), (-1, -1), (0, -1)]
    head, tail, prev, nxt = [], [], [], []
    for y in range(ny):
        for x in range(nx):
            nb = [(x + dx) + nx * (y + dy) for dx, dy in ring if 0 <= x + dx < nx and 0 <= y + dy < ny]
            if interior_only and len(nb) < 6: continue
            n = len(nb)
            for i in range(n):
                head.append(x + nx * y); tail.append(nb[i]); prev.append(nb[(i + n - 1) % n]); nxt.append(nb[(i + 1) % n])
    return [np.array(a, dtype=np.int32) for a in (head, tail, prev, nxt)]


def cotangent_synthetic(nx, ny, seed=1234, n_iter=3, l_iter=10, interior_rings=False):
    """cotangent_mesh_smoothing on a noisy height-field mesh: X = A = noisy positions (driver: both the input mesh,
    CombinedSolver.h:128-150), weights sqrt(1), sqrt(0.5) (main.cpp:62-63).  The open mesh's boundary vertices have rings that
    wrap across the outside of the mesh - as the driver builds them - where the two edge vectors of a 'triangle' are nearly
    collinear and the energy's (a.a)(b.b) - (a.b)^2 cancels catastrophically in fp32 (the fp32 and fp64 ORACLE evaluations of
    J^T F differ by 5e-5 there).  interior_rings: an equilateral lattice on which only full rings emit edges - the
    well-conditioned variant of the same code path (fp32 / fp64 oracle J^T F agree to 4e-7) that holds the standard gates."""
    rng = np.random.default_rng(seed)
    f32 = np.float32
    ys, xs = np.mgrid[0:ny, 0:nx]
    z = 0.1 * np.sin(6.0 * xs / nx) * np.cos(5.0 * ys / ny)
    if interior_rings: pos = np.stack([(xs - 0.5 * ys) / max(nx, ny), ys * (np.sqrt(3) / 2) / max(nx, ny), z], -1).reshape(-1, 3)
    else: pos = np.stack([xs / max(nx, ny), ys / max(nx, ny), z], -1).reshape(-1, 3)
    # noise relative to the mesh spacing h (0.12 h and 0.024 h: 0.01 / 0.002 on the 12-vertex-wide test meshes).  An absolute 0.01
    # on a 1024-wide mesh (h = 1e-3) scrambles it into folded triangles, and an unperturbed regular grid has opposite angles of
    # exactly 90 degrees: both put om = (cot + cot) / 2 on the energy's select(om > 0, om, 1e-4), where fp32 rounding decides
    h = 1.0 / max(nx, ny)
    A = (pos + 0.12 * h * rng.standard_normal(pos.shape)).astype(f32)
    X = (A + 0.024 * h * rng.standard_normal(pos.shape)).astype(f32)
    g = grid_mesh_rings(nx, ny, interior_rings)
    ecount = np.array([len(g[0])], dtype=np.int32)
    return Problem("cotangent_mesh_smoothing.t", [nx * ny, len(g[0])],
                   [np.sqrt(f32(1.0)), np.sqrt(f32(0.5)), X, A, ecount] + g, [2],
                   {"nIterations": n_iter, "lIterations": l_iter})


def robust_synthetic(nx, ny, seed=1234, n_iter=3, l_iter=10, kind="LMGPU"):
    """robust_nonrigid_alignment on a height-field mesh: point-to-plane constraints on a tenth of the vertices
    (target = displaced position, unit normal), RobustWeights = 1 (CombinedSolver.h:343), Angle = 0, weights
    sqrt(10), sqrt(64) (:158-169); the driver uses LM."""
    rng = np.random.default_rng(seed)
    f32 = np.float32
    ys, xs = np.mgrid[0:ny, 0:nx]
    z = 0.1 * np.sin(6.0 * xs / nx) * np.cos(5.0 * ys / ny)
    ur = np.stack([xs / max(nx, ny), ys / max(nx, ny), z], -1).reshape(-1, 3).astype(f32)
    n = nx * ny
    v0, v1 = grid_mesh_edges(nx, ny)
    cons = np.full((n, 3), -np.inf, dtype=f32)
    nrm = np.zeros((n, 3), dtype=f32)
    pick = rng.choice(n, max(4, n // 10), replace=False)
    cons[pick] = ur[pick] + (0.05 * rng.standard_normal((len(pick), 3))).astype(f32)
    d = rng.standard_normal((len(pick), 3)) + np.array([0, 0, 3.0])
    nrm[pick] = (d / np.linalg.norm(d, axis=1, keepdims=True)).astype(f32)
    offset = (ur + 0.005 * rng.standard_normal((n, 3))).astype(f32)
    angle = (0.02 * rng.standard_normal((n, 3))).astype(f32)
    rw = np.ones((n,), dtype=f32)
    ecount = np.array([len(v0)], dtype=np.int32)
    return Problem("robust_nonrigid_alignment.t", [n, len(v0)],
                   [np.sqrt(f32(10.0)), np.sqrt(f32(64.0)), offset, angle, rw, ur, cons, nrm, ecount, v0, v1], [2, 3, 4],
                   {"nIterations": n_iter, "lIterations": l_iter}, kind=kind)


def embedded_synthetic(nx, ny, seed=1234, n_iter=3, l_iter=10):
    """embedded_mesh_deformation on a height-field mesh: Offset = UrShape + noise, RotMatrix = identity + noise
    (driver: identity, embedded_mesh_deformation/src/CombinedSolver.h:117-131), a few vertices constrained, weights
    sqrt(3), sqrt(12), sqrt(5) (:36-45)."""
    rng = np.random.default_rng(seed)
    f32 = np.float32
    ys, xs = np.mgrid[0:ny, 0:nx]
    z = 0.1 * np.sin(6.0 * xs / nx) * np.cos(5.0 * ys / ny)
    ur = np.stack([xs / max(nx, ny), ys / max(nx, ny), z], -1).reshape(-1, 3).astype(f32)
    n = nx * ny
    v0, v1 = grid_mesh_edges(nx, ny)
    cons = np.full((n, 3), -np.inf, dtype=f32)
    pick = rng.choice(n, max(4, n // 50), replace=False)
    cons[pick] = ur[pick] + (0.05 * rng.standard_normal((len(pick), 3))).astype(f32)
    offset = (ur + 0.01 * rng.standard_normal((n, 3))).astype(f32)
    rot = (np.tile(np.eye(3, dtype=f32).reshape(1, 9), (n, 1)) + 0.05 * rng.standard_normal((n, 9))).astype(f32)
    ecount = np.array([len(v0)], dtype=np.int32)
    return Problem("embedded_mesh_deformation.t", [n, len(v0)],
                   [np.sqrt(f32(3.0)), np.sqrt(f32(12.0)), np.sqrt(f32(5.0)), offset, rot, ur, cons, ecount, v0, v1], [3, 4],
                   {"nIterations": n_iter, "lIterations": l_iter})


# lighting coefficients / intrinsics of examples/data/shape_from_shading/default.SFSSolverParameters
# (layout: shape_from_shading/src/TerraSolverParameters.h:7-31)
SFS_L = [0.6908317804336548, 0.044598858803510666, 0.01812959648668766, -0.1773163229227066, -0.04067882522940636,
         0.14467650651931763, 0.02393525093793869, -0.24658696353435516, 0.005797004792839289]


def sfs_synthetic(W, H, seed=1234, n_iter=3, l_iter=10, kind="gaussNewtonGPU", holes=True):
    """shape_from_shading: D_i = 1 + smooth field, X = D_i + small noise, Im = smooth
    intensity in [0,1], uint8 edge masks, pinhole intrinsics scaled from the 640x480 default
    (f = 0.897 W, principal point at the image centre), weights w_p, w_s, w_g = 100, 100, 1 and lighting from
    default.SFSSolverParameters.  `holes` punches invalid-depth pixels (D_i <= 0 -> Exclude)."""
    rng = np.random.default_rng(seed)
    f32 = np.float32
    ys, xs = np.mgrid[0:H, 0:W].astype(np.float64)
    smooth = np.sin(2 * np.pi * (0.7 * xs / W + 0.2)) * np.cos(2 * np.pi * (0.6 * ys / H))
    D = (1.0 + 0.03 * smooth).astype(f32)
    if holes:
        k = max(1, (W * H) // 200)
        D.reshape(-1)[rng.choice(W * H, k, replace=False)] = 0.0
        D[H // 3:H // 3 + 2, W // 4:W // 4 + 3] = -1.0
    X = (D + 0.002 * rng.standard_normal((H, W))).astype(f32)
    Im = (0.5 + 0.3 * np.cos(2 * np.pi * (0.9 * xs / W)) * np.sin(2 * np.pi * (0.8 * ys / H + 0.1))
          + 0.01 * rng.standard_normal((H, W))).astype(f32)
    mR = (rng.random((H, W)) > 0.1).astype(np.uint8)
    mC = (rng.random((H, W)) > 0.1).astype(np.uint8)
    fx = f32(574.0528564453125 / 640.0 * W)
    scal = [f32(100.0), f32(100.0), f32(1.0), fx, fx, f32(W / 2), f32(H / 2)] + [f32(v) for v in SFS_L]
    return Problem("shape_from_shading.t", [W, H], scal + [X, D, Im, mR, mC], [16],
                   {"nIterations": n_iter, "lIterations": l_iter}, kind=kind)


def intrinsic_synthetic(W, H, seed=1234, n_iter=3, l_iter=10, kind="gaussNewtonGPU", pnorm=0.8):
    """intrinsic_image_decomposition on a synthetic picture (piecewise-constant albedo times a smooth shading
    field, 1 % noise), initialised like the driver (src/CombinedSolver.h:68-108): i = log2(rgb + 0.01),
    r = log2(chroma + 0.01), s = log2(intensity + 0.01); weights sqrt(500), sqrt(1000), sqrt(10000) and
    pNorm = 0.8 (:33-46; pNorm is an opt_float Param, intrinsic_image_decomposition.t:5).  The re-weighting
    pow(|r_c - r_n| + 1e-7, pNorm - 2) makes the driver's pNorm = 0.8 ill-conditioned in fp32 from the second nonlinear step on
    (the fp32 and fp64 ORACLE runs drift 1e-2 apart by step 3); pnorm = 1.6 is the well-conditioned variant of the same code
    path (fp32 / fp64 oracle runs agree to 1e-7 over five steps) that the parity suite holds to the standard gates."""
    rng = np.random.default_rng(seed)
    f32 = np.float32
    ys, xs = np.mgrid[0:H, 0:W].astype(np.float64)
    blocks = rng.random((H // 8 + 1, W // 8 + 1, 3))
    albedo = 0.2 + 0.7 * blocks[(ys // 8).astype(int), (xs // 8).astype(int)]
    shading = 0.6 + 0.3 * np.sin(2 * np.pi * (0.8 * xs / W + 0.1)) * np.cos(2 * np.pi * 0.7 * ys / H)
    rgb = np.clip(albedo * shading[..., None] * (1.0 + 0.01 * rng.standard_normal((H, W, 3))), 0.02, 1.0).astype(f32)
    eps = f32(0.01)
    intensity = rgb.sum(-1, dtype=f32) / f32(3.0)
    img = np.log2(rgb + eps).astype(f32)
    r = np.log2(rgb / intensity[..., None] + eps).astype(f32)
    s = np.log2(intensity + eps).astype(f32)
    w = [np.sqrt(f32(500.0)), np.sqrt(f32(1000.0)), np.sqrt(f32(10000.0))]
    return Problem("intrinsic_image_decomposition.t", [W, H], w + [f32(pnorm), r, img, s], [4, 6],
                   {"nIterations": n_iter, "lIterations": l_iter}, kind=kind, opt_float_scalars=[3])


def intrinsic_golden(npz, n_iter=1, l_iter=1):
    """The reference's golden run of intrinsic_image_decomposition (examples/test_final_cost.py:60,81: stride 12):
    src/main.cpp:28-39 (strided copy) and src/CombinedSolver.h:33-46, 68-108 (weights, log2 initialisation)."""
    f32 = np.float32
    v = npz["rgb"].astype(f32) / f32(255.0)
    intensity = (v[..., 0] + v[..., 1] + v[..., 2]) / f32(3.0)
    eps = f32(0.01)
    img = np.log2(v + eps).astype(f32)
    r = np.log2(v / intensity[..., None] + eps).astype(f32)
    s = np.log2(intensity + eps).astype(f32)
    H, W = s.shape
    w = [np.sqrt(f32(500.0)), np.sqrt(f32(1000.0)), np.sqrt(f32(10000.0))]
    return Problem("intrinsic_image_decomposition.t", [W, H], w + [f32(0.8), r, img, s], [4, 6],
                   {"nIterations": n_iter, "lIterations": l_iter}, opt_float_scalars=[3])


def _filter_gaussian(img, sigma):
    """examples/optical_flow/src/ImageHelper.h:68-114: separable, kernel radius ceil(2 sigma), renormalised at the border."""
    f32 = np.float32
    R = int(np.ceil(2.0 * sigma))
    k = np.exp(-((np.arange(R + 1, dtype=f32) ** 2) / f32(2.0 * sigma * sigma))).astype(f32)

    def one_pass(im, axis):
        n = im.shape[axis]
        out = np.zeros_like(im)
        for i in range(n):
            v = np.zeros(np.delete(im.shape, axis), dtype=f32)
            w = f32(0)
            for kk in range(-R, R + 1):
                j = i + kk
                if 0 <= j < n:
                    v = v + k[abs(kk)] * np.take(im, j, axis=axis)
                    w = f32(w + k[abs(kk)])
            v = v / w
            if axis == 1:
                out[:, i] = v
            else:
                out[i, :] = v
        return out
    return one_pass(one_pass(img.astype(f32), 1), 0)


def _prewitt8(im, horizontal):
    """computeDU / computeDV, examples/optical_flow/src/CombinedSolver.h:152-178 (zero on the border)."""
    f32 = np.float32
    r = np.zeros_like(im)
    if horizontal:
        r[1:-1, 1:-1] = (-im[:-2, :-2] - im[1:-1, :-2] - im[2:, :-2] + im[:-2, 2:] + im[1:-1, 2:] + im[2:, 2:]) / f32(8.0)
    else:
        r[1:-1, 1:-1] = (-im[:-2, :-2] - im[:-2, 1:-1] - im[:-2, 2:] + im[2:, :-2] + im[2:, 1:-1] + im[2:, 2:]) / f32(8.0)
    return r


def gray_from_rgb(rgb):
    """mLib baseImage.h:1461-1471 (convertToGrayscale of an 8-bit colour image), in fp32."""
    f32 = np.float32
    a = rgb.astype(f32)
    return ((f32(0.299) * a[..., 0] + f32(0.587) * a[..., 1] + f32(0.114) * a[..., 2]) / f32(255.0)).astype(f32)


def optical_flow_golden(npz, n_iter=1, l_iter=1, outer_iterations=1):
    """The FIRST solve of the reference's golden run of optical_flow (test_final_cost.py:62, stride 16, oIterations 1):
    the coarse level of the driver's 2-level pyramid (sigma = 5, CombinedSolver.h:25-37, solved first :47-66) with zero
    flow, fit weight 10 + (50 - 10) / (oIterations * 2) (:69-93; 30 for oIterations 1) and regularisation weight 0.1.
    The fixture holds either the strided grey images or, for the full-size long run, the 8-bit colour frames."""
    f32 = np.float32
    if "source_rgb" in npz:
        src, tar = gray_from_rgb(npz["source_rgb"]), gray_from_rgb(npz["target_rgb"])
    else:
        src, tar = npz["source"], npz["target"]
    S, T = _filter_gaussian(src, 5.0), _filter_gaussian(tar, 5.0)
    H, W = S.shape
    X = np.zeros((H, W, 2), dtype=f32)
    w_fit = f32(10.0) + (f32(50.0) - f32(10.0)) / f32(outer_iterations * 2)
    return Problem("optical_flow.t", [W, H], [np.sqrt(f32(w_fit)), np.sqrt(f32(0.1)), X, S, T, _prewitt8(T, True), _prewitt8(T, False)],
                   [2], {"nIterations": n_iter, "lIterations": l_iter})


# ---------------------------------------------------------------------------------------------
# mesh drivers: connectivity the reference derives with OpenMesh, restated on plain index arrays
def _vertex_edges(nv, faces):
    """Directed edge lists (v0, v1): for every vertex, all of its one-ring neighbours.  The graph energies sum over
    edges, so the order inside a ring is immaterial for the cost; ascending is used here
    (examples/*/src/CombinedSolver.h: the vv_iter loops that fill the edge lists)."""
    nbr = [set() for _ in range(nv)]
    for f in faces:
        k = len(f)
        for a in range(k):
            x, y = int(f[a]), int(f[(a + 1) % k])
            nbr[x].add(y)
            nbr[y].add(x)
    v0 = [a for a in range(nv) for _ in nbr[a]]
    v1 = [b for a in range(nv) for b in sorted(nbr[a])]
    return np.array(v0, dtype=np.int32), np.array(v1, dtype=np.int32)


def _vertex_rings(nv, faces):
    """One-ring of every vertex in rotation order (what a half-edge circulator walks); an open fan is walked from
    one boundary end to the other."""
    succ = [dict() for _ in range(nv)]   # at vertex a: succ[a][b] = c for every oriented face (a, b, c)
    for a, b, c in faces:
        a, b, c = int(a), int(b), int(c)
        succ[a][b] = c
        succ[b][c] = a
        succ[c][a] = b
    rings = []
    for a in range(nv):
        s = succ[a]
        if not s:
            rings.append([])
            continue
        preds = set(s.values())
        starts = [b for b in s if b not in preds]
        start = starts[0] if starts else next(iter(s))
        ring, cur = [start], start
        while cur in s and s[cur] != start and len(ring) <= len(s) + 1:
            cur = s[cur]
            ring.append(cur)
        rings.append(ring)
    return rings


def _marker_constraints(nv, npz):
    cons = np.full((nv, 3), -np.inf, dtype=np.float32)
    cons[npz["marker_vertex"]] = npz["marker_pos"].astype(np.float32)
    return cons


def grid_mesh_edges(nx, ny):
    """Directed edges of a regular height-field triangulation, both directions of every mesh
    edge, sorted by head vertex (the order examples/shared/OptGraph.h:78-89 produces)."""
    idx = np.arange(nx * ny).reshape(ny, nx)
    pairs = [(idx[:, :-1], idx[:, 1:]), (idx[:-1, :], idx[1:, :]), (idx[:-1, :-1], idx[1:, 1:])]
    a = np.concatenate([p[0].ravel() for p in pairs] + [p[1].ravel() for p in pairs])
    b = np.concatenate([p[1].ravel() for p in pairs] + [p[0].ravel() for p in pairs])
    order = np.lexsort((b, a))
    return a[order].astype(np.int32), b[order].astype(np.int32)


def arap_synthetic(nx, ny, seed=1234, n_iter=2, l_iter=10):
    """examples/arap_mesh_deformation/src/CombinedSolver.h:84-102,161-167; main.cpp:137-138."""
    rng = np.random.default_rng(seed)
    f32 = np.float32
    ys, xs = np.mgrid[0:ny, 0:nx]
    z = 0.1 * np.sin(6.0 * xs / nx) * np.cos(5.0 * ys / ny)
    ur = np.stack([xs / max(nx, ny), ys / max(nx, ny), z], -1).reshape(-1, 3).astype(f32)
    n = nx * ny
    v0, v1 = grid_mesh_edges(nx, ny)
    cons = np.full((n, 3), -np.inf, dtype=f32)
    pick = rng.choice(n, max(4, n // 1000), replace=False)
    cons[pick] = ur[pick] + (0.05 * rng.standard_normal((len(pick), 3))).astype(f32)
    offset = ur.copy()
    angle = np.full((n, 3), 0.1, dtype=f32)
    w_fit, w_reg = np.sqrt(f32(4.0)), np.sqrt(f32(1.0))
    ecount = np.array([len(v0)], dtype=np.int32)
    return Problem("arap_mesh_deformation.t", [n, len(v0)],
                   [w_fit, w_reg, offset, angle, ur, cons, ecount, v0, v1], [2, 3],
                   {"nIterations": n_iter, "lIterations": l_iter})


# ---------------------------------------------------------------------------------------------
# mesh goldens (examples/test_final_cost.py:56-66; inputs stored by tools/make_golden.py)
def cotangent_golden(npz, n_iter=1, l_iter=1):
    """cotangent_mesh_smoothing on head.ply: examples/cotangent_mesh_smoothing/src/CombinedSolver.h:66-135 (every
    directed edge (head, tail) carries the ring neighbours before and after the tail), weights 1.0 / 0.5 (:18-19)."""
    f32 = np.float32
    v = npz["vertices"].astype(f32)
    head, tail, prev, nxt = [], [], [], []
    for a, ring in enumerate(_vertex_rings(len(v), npz["faces"])):
        n = len(ring)
        for i, c in enumerate(ring):
            head.append(a)
            tail.append(c)
            prev.append(ring[(i + n - 1) % n])
            nxt.append(ring[(i + 1) % n])
    g = [np.array(x, dtype=np.int32) for x in (head, tail, prev, nxt)]
    E = len(head)
    return Problem("cotangent_mesh_smoothing.t", [len(v), E],
                   [np.sqrt(f32(1.0)), np.sqrt(f32(0.5)), v.copy(), v.copy(), np.array([E], dtype=np.int32)] + g, [2],
                   {"nIterations": n_iter, "lIterations": l_iter})


def _rotation_deg(axis, deg):
    """mLib matrix3x3.h:127-174 (rotationX/Y/Z of an angle in degrees, float arithmetic)."""
    f32 = np.float32
    a = f32(f32(deg) * f32(np.pi) / f32(180.0))
    s, c = f32(np.sin(a)), f32(np.cos(a))
    if axis == "x":
        return np.array([[1, 0, 0], [0, c, -s], [0, s, c]], dtype=f32)
    if axis == "y":
        return np.array([[c, 0, s], [0, 1, 0], [-s, 0, c]], dtype=f32)
    return np.array([[c, -s, 0], [s, c, 0], [0, 0, 1]], dtype=f32)


def embedded_golden(npz, n_iter=1, l_iter=1):
    """embedded_mesh_deformation on raptor_simplify2k: examples/embedded_mesh_deformation/src/CombinedSolver.h:34-160
    (weights 3 / 12 / 5 at :39-41, rotations start as R_y R_x R_z of 1e-3 degrees, unconstrained vertices = -inf)."""
    f32 = np.float32
    v = npz["vertices"].astype(f32)
    nv = len(v)
    v0, v1 = _vertex_edges(nv, npz["faces"])
    R = (_rotation_deg("y", 1e-3) @ _rotation_deg("x", 1e-3) @ _rotation_deg("z", 1e-3)).astype(f32)
    rots = np.tile(R.reshape(1, 9), (nv, 1)).astype(f32)
    E = len(v0)
    return Problem("embedded_mesh_deformation.t", [nv, E],
                   [np.sqrt(f32(3.0)), np.sqrt(f32(12.0)), np.sqrt(f32(5.0)), v.copy(), rots, v.copy(),
                    _marker_constraints(nv, npz), np.array([E], dtype=np.int32), v0, v1], [3, 4],
                   {"nIterations": n_iter, "lIterations": l_iter})


def embedded_outer_constraints(pr, i, n_outer):
    """The Constraints image of outer iteration i of n_outer (CombinedSolver.h:57-60, 75-93: alpha = (i + 1) / n_outer in float,
    z = (1 - alpha) pt + alpha target with pt the INITIAL position of the marker vertex - m_result is only refreshed after the
    last outer iteration).  `pr` is embedded_golden()'s problem, whose params[6] are the alpha = 1 targets and params[5] the
    initial positions; the unknowns (params[3], params[4]) carry over from one outer iteration to the next."""
    f32 = np.float32
    tgt, v = np.asarray(pr.params[6]), np.asarray(pr.params[5])
    con = np.isfinite(tgt[:, 0])
    a = f32(i + 1) / f32(n_outer)
    z = tgt.astype(f32).copy()
    z[con] = ((f32(1) - a) * v[con].astype(f32) + a * tgt[con].astype(f32)).astype(f32)
    return z.astype(tgt.dtype)


def sqrt3_subdivide(v, faces):
    """One step of Kobbelt's sqrt(3) subdivision on a closed triangle mesh, as the reference's ARAP driver applies it
    before solving (examples/arap_mesh_deformation/src/main.cpp:65-81, OpenMesh Sqrt3T with numSubdivides = 1):
    old vertices relax towards their ring average with alpha(n) = (4 - 2 cos(2 pi / n)) / 9, a new vertex is
    inserted at every face centroid (of the un-relaxed positions), and every old edge is flipped so that it joins the
    two centroids next to it.  Returns the positions (old vertices first, then one per face) and, for every
    vertex, the set of its neighbours in the subdivided mesh."""
    f32 = np.float32
    v = v.astype(f32)
    nv = len(v)
    edge_faces = {}
    for fi, (a, b, c) in enumerate(faces):
        for x, y in ((a, b), (b, c), (c, a)):
            edge_faces.setdefault((min(int(x), int(y)), max(int(x), int(y))), []).append(fi)
    if any(len(l) != 2 for l in edge_faces.values()):
        raise ValueError("sqrt3_subdivide: closed manifold meshes only")
    ring = [set() for _ in range(nv)]
    for x, y in edge_faces:
        ring[x].add(y)
        ring[y].add(x)
    relaxed = v.copy()
    for a in range(nv):
        k = len(ring[a])
        alpha = (4.0 - 2.0 * np.cos(2.0 * np.pi / k)) / 9.0
        acc = np.zeros(3, dtype=f32)
        for b in ring[a]:
            acc = acc + v[b]
        relaxed[a] = acc * f32(alpha / k) + f32(1.0 - alpha) * v[a]
    cent = np.empty((len(faces), 3), dtype=f32)
    for fi, (a, b, c) in enumerate(faces):
        cent[fi] = ((v[a] + v[b]) + v[c]) / f32(3.0)
    adj = [set() for _ in range(nv + len(faces))]
    for fi, f in enumerate(faces):
        for x in f:
            adj[nv + fi].add(int(x))
            adj[int(x)].add(nv + fi)
    for l in edge_faces.values():
        adj[nv + l[0]].add(nv + l[1])
        adj[nv + l[1]].add(nv + l[0])
    return np.concatenate([relaxed, cent], 0), adj


def arap_golden(npz, n_iter=1, l_iter=1, outer_iterations=1):
    """arap_mesh_deformation on small_armadillo after one sqrt(3) subdivision: examples/arap_mesh_deformation/src/
    main.cpp:45-81 and CombinedSolver.h:23-24 (weights 4 / 1), :84-135 (angles start at 0.1, offsets = rest shape =
    subdivided positions, marker vertices constrained, everything else -inf).  outer_iterations = 10 with (20, 100)
    is the FIRST outer iteration of the long run (test_example_output.py:59,110,131): the markers are moved
    alpha = 1/10 of the way from the current vertex position to their target (CombinedSolver.h:66-68, 84-101)."""
    f32 = np.float32
    P, adj = sqrt3_subdivide(npz["vertices"], npz["faces"])
    n = len(P)
    cons = _marker_constraints(n, npz)
    if outer_iterations != 1:
        alpha = f32(1.0) / f32(outer_iterations)
        idx = npz["marker_vertex"]
        cons[idx] = (f32(1.0) - alpha) * P[idx] + alpha * cons[idx]
    v0 = np.array([a for a in range(n) for _ in adj[a]], dtype=np.int32)
    v1 = np.array([b for a in range(n) for b in sorted(adj[a])], dtype=np.int32)
    E = len(v0)
    return Problem("arap_mesh_deformation.t", [n, E],
                   [np.sqrt(f32(4.0)), np.sqrt(f32(1.0)), P.copy(), np.full((n, 3), 1e-1, dtype=f32), P.copy(),
                    cons, np.array([E], dtype=np.int32), v0, v1], [2, 3],
                   {"nIterations": n_iter, "lIterations": l_iter})


def poisson_golden(npz, n_iter=1, l_iter=1):
    """poisson_image_editing at stride 4 (test_final_cost.py:61,82): examples/poisson_image_editing/src/
    CombinedSolver.h:70-93 (alpha forced to 255; mask 0 = unknown where the driver's mask image reads 255).  The mask
    in the fixture is the one the driver effectively derives, see tools/make_golden.py:poisson."""
    f32 = np.float32
    H, W = (int(t) for t in npz["shape"])
    unknown = np.unpackbits(npz["unknown"])[:H * W].reshape(H, W).astype(bool)
    X = np.concatenate([npz["image"].astype(f32), np.full((H, W, 1), 255, dtype=f32)], -1)
    T = np.concatenate([npz["target"].astype(f32), np.full((H, W, 1), 255, dtype=f32)], -1)
    M = np.where(unknown, 0, 255).astype(f32)
    return Problem("poisson_image_editing.t", [W, H], [X, T, M], [0], {"nIterations": n_iter, "lIterations": l_iter})
