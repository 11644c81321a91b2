"""TEST INFRASTRUCTURE ONLY: numpy / pure-Python restatement of the CPU code the reference's example drivers run
around the solver path.  Checker for the CUDA kernels of opt_b200/csrc/driver_ops.cu;
never imported by the product package.  float32 throughout, one rounding per operation, the reference's order of
operations (a plain C++ build without FMA contraction).

Parity: unpinned by reference goldens except the Gaussian + Prewitt chain, whose output feeds the optical_flow
golden cost (examples/test_final_cost.py:62) that tests/test_oracle_golden.py reproduces through
opt_b200.examples; the restatements here are independent of that module and are cross-checked against it."""
import numpy as np

f32 = np.float32


def filter_gaussian(img, sigma):
    """examples/optical_flow/src/ImageHelper.h:63-114."""
    img = np.asarray(img, dtype=f32)
    R = int(np.ceil(f32(2.0) * f32(sigma)))
    x = np.arange(R + 1, dtype=f32)
    kern = np.exp(-((x * x) / (f32(2.0) * f32(sigma) * f32(sigma)))).astype(f32)

    def along_x(im):
        H, W = im.shape
        v = np.zeros((H, W), dtype=f32)
        w = np.zeros(W, dtype=f32)
        for k in range(-R, R + 1):
            lo, hi = max(0, -k), min(W, W - k)          # pixels i with 0 <= i + k < W
            if lo >= hi:
                continue
            g = kern[abs(k)]
            v[:, lo:hi] = v[:, lo:hi] + g * im[:, lo + k:hi + k]
            w[lo:hi] = w[lo:hi] + g
        return np.where(w > 0, v / np.where(w > 0, w, f32(1)), v).astype(f32)

    return along_x(along_x(img).T.copy()).T.copy()


def prewitt(im, horizontal):
    """computeDU / computeDV, examples/optical_flow/src/CombinedSolver.h:152-178."""
    im = np.asarray(im, dtype=f32)
    r = np.zeros_like(im)
    c = slice(1, -1)
    if horizontal:
        d = -im[:-2, :-2] - im[c, :-2] - im[2:, :-2] + im[:-2, 2:] + im[c, 2:] + im[2:, 2:]
    else:
        d = -im[:-2, :-2] - im[:-2, c] - im[:-2, 2:] + im[2:, :-2] + im[2:, c] + im[2:, 2:]
    r[c, c] = d / f32(8.0)
    return r


def _point_in_triangle_lk(p0, p1, p2, sx, sy):
    """examples/image_warping/src/CombinedSolver.h:63-99 with w0 = w1 = w2 = 1."""
    X = [f32(p[0] - sx * f32(1)) for p in (p0, p1, p2)]
    Y = [f32(p[1] - sy * f32(1)) for p in (p0, p1, p2)]
    d01 = f32(f32(X[0] * Y[1]) - f32(Y[0] * X[1]))
    d12 = f32(f32(X[1] * Y[2]) - f32(Y[1] * X[2]))
    d20 = f32(f32(X[2] * Y[0]) - f32(Y[2] * X[0]))
    if d01 < 0 and d12 < 0 and d20 < 0:
        return None
    with np.errstate(divide="ignore", invalid="ignore", over="ignore"):
        inv = f32(1.0) / f32(f32(d01 + d12) + d20)
        d01, d12, d20 = f32(d01 * inv), f32(d12 * inv), f32(d20 * inv)
    if d01 >= 0 and d12 >= 0 and d20 >= 0:
        return d12, d20, d01
    return None


def rasterize_warp(warp, color, mask):
    """copyResultToCPU + rasterizeTriangle, examples/image_warping/src/CombinedSolver.h:231-308: triangles drawn one
    after the other, the last writer of a pixel wins."""
    warp, color, mask = (np.asarray(a, dtype=f32) for a in (warp, color, mask))
    H, W = mask.shape
    out = np.full((H, W, 3), 255.0, dtype=f32)
    for y in range(H - 1):
        for x in range(W - 1):
            if not (mask[y, x] == 0 and mask[y + 1, x] == 0 and mask[y, x + 1] == 0 and mask[y + 1, x + 1] == 0):
                continue
            tris = (((y, x), (y, x + 1), (y + 1, x)), ((y + 1, x), (y, x + 1), (y + 1, x + 1)))
            for tri in tris:
                p = [warp[i] for i in tri]
                c = [color[i] for i in tri]
                xs, ys = [q[0] for q in p], [q[1] for q in p]
                if not all(np.isfinite(xs + ys)):
                    continue
                x_lo, x_hi = int(np.floor(min(xs))), int(np.ceil(max(xs)))
                y_lo, y_hi = int(np.floor(min(ys))), int(np.ceil(max(ys)))
                for px in range(max(x_lo, 0), min(x_hi, W - 1) + 1):
                    for py in range(max(y_lo, 0), min(y_hi, H - 1) + 1):
                        b = _point_in_triangle_lk(p[0], p[1], p[2], f32(px), f32(py))
                        if b is not None:
                            out[py, px] = (c[0] * b[0] + c[1] * b[1]) + c[2] * b[2]
    return out


def skin_trilinear(grid, dims, voxel, rel):
    """copyResultToCPUFromFloat3, examples/volumetric_mesh_deformation/src/CombinedSolver.h:302-335 (getIndex1D :273-276)."""
    grid, rel = np.asarray(grid, dtype=f32), np.asarray(rel, dtype=f32)
    ny1, nz1 = dims[1] + 1, dims[2] + 1

    def node(dx, dy, dz):
        return grid[(voxel[:, 0] + dx) * (ny1 * nz1) + (voxel[:, 1] + dy) * nz1 + (voxel[:, 2] + dz)]
    rx, ry, rz = (rel[:, i:i + 1] for i in range(3))
    ox, oy, oz = f32(1.0) - rx, f32(1.0) - ry, f32(1.0) - rz
    px00 = ox * node(0, 0, 0) + rx * node(1, 0, 0)
    px01 = ox * node(0, 0, 1) + rx * node(1, 0, 1)
    px10 = ox * node(0, 1, 0) + rx * node(1, 1, 0)
    px11 = ox * node(0, 1, 1) + rx * node(1, 1, 1)
    pxx0 = oy * px00 + ry * px10
    pxx1 = oy * px01 + ry * px11
    return (oz * pxx0 + rz * pxx1).astype(f32)


def _round(v):
    """mLib math::round (core-util/utility.h:130-133): half away from zero."""
    v = f32(v)
    return int(np.floor(f32(v + f32(0.5)))) if v > 0 else int(np.ceil(f32(v - f32(0.5))))


def render_flow_lines(image, flow, colors, skip=5):
    """renderFlowVecotors (examples/optical_flow/src/main.cpp:7-24) with ImageHelper::drawLine / drawLineInternal
    (ImageHelper.h:52-66, 118-141); the per-line colour is an input."""
    img = np.array(image, copy=True)
    H, W = flow.shape[:2]

    def put(x, y, c):
        if 0 <= x < W and 0 <= y < H:
            img[y, x] = c
    s = 0
    for j in range(1, H - 1, skip):
        for i in range(1, W - 1, skip):
            fx, fy = flow[j, i]
            ex, ey = i + _round(fx), j + _round(fy)
            cl = lambda v, hi: min(max(v, 0), hi)       # noqa: E731
            a, b = (cl(i, W - 1), cl(j, H - 1)), (cl(ex, W - 1), cl(ey, H - 1))
            if not ex >= i:
                a, b = b, a
            c = colors[s]
            s += 1
            dx = f32(b[0] - a[0])
            if dx == 0:
                for y in range(a[1], b[1]):
                    put(a[0], y, c)
                continue
            dy = f32(b[1] - a[1])
            err, d_err = f32(0), f32(abs(dy / dx))
            direction = (b[1] > a[1]) - (b[1] < a[1])
            y = a[1]
            for x in range(a[0], b[0]):
                put(x, y, c)
                err = f32(err + d_err)
                while err >= f32(0.5):
                    put(x, y, c)
                    y += direction
                    err = f32(err - f32(1))
    return img
