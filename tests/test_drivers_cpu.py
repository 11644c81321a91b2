"""CPU: the numpy restatement of the drivers' pre/post-processing (oracle/driver_ref.py) against independent
knowledge -- the golden-pinned Gaussian/Prewitt of opt_b200.examples, analytic cases of the rasteriser, the
skinning and the line drawer -- and the exported symbols of include/OptB200Drivers.h."""
import ctypes as C
import os
import re
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import driver_ref as R  # noqa: E402

from opt_b200 import examples as E  # noqa: E402

f32 = np.float32


def test_driver_symbols_exported(built):
    txt = re.sub(r"/\*.*?\*/", "", open(os.path.join(ROOT, "include", "OptB200Drivers.h")).read(), flags=re.S)
    names = sorted(set(re.findall(r"\b(OptB200_[A-Za-z]+)\s*\(", txt)))
    assert len(names) == 7          # six driver kernels + OptB200_FilterGaussianNoAlloc
    lib = C.CDLL(built)
    for n in names:
        assert hasattr(lib, n), n


def test_gaussian_and_prewitt_match_the_golden_pinned_restatement():
    # opt_b200.examples._filter_gaussian/_prewitt8 feed the optical_flow golden (final cost 0.52119255 reproduced to 1e-9)
    rng = np.random.default_rng(0)
    im = (rng.random((37, 53), dtype=f32) * 255).astype(f32)
    for sigma in (0.7, 1.0, 5.0):
        assert np.array_equal(R.filter_gaussian(im, sigma), E._filter_gaussian(im, sigma))
    assert np.array_equal(R.prewitt(im, True), E._prewitt8(im, True))
    assert np.array_equal(R.prewitt(im, False), E._prewitt8(im, False))
    # a constant image stays constant up to rounding (renormalised at the border); Prewitt of a ramp is its slope
    c = R.filter_gaussian(np.full((9, 11), 3.0, dtype=f32), 5.0)
    assert np.abs(c - 3.0).max() < 1e-6
    ramp = np.tile(np.arange(8, dtype=f32) * 2, (6, 1))
    assert np.array_equal(R.prewitt(ramp, True)[1:-1, 1:-1], np.full((4, 6), 3 * 4 / 8.0, dtype=f32))
    assert not R.prewitt(ramp, False).any()


def test_rasteriser_identity_warp_reproduces_the_image():
    H, W = 7, 9
    rng = np.random.default_rng(1)
    color = rng.random((H, W, 3), dtype=f32)
    ys, xs = np.mgrid[0:H, 0:W]
    warp = np.stack([xs, ys], axis=-1).astype(f32)
    mask = np.zeros((H, W), dtype=f32)
    out = R.rasterize_warp(warp, color, mask)
    assert np.abs(out - color).max() < 1e-6           # every pixel is a vertex of the identity mesh
    mask[3, 4] = 1                                    # the four cells around a masked vertex are not drawn
    out = R.rasterize_warp(warp, color, mask)
    assert np.array_equal(out[3, 4], np.full(3, 255, dtype=f32))
    assert np.abs(out[0, 0] - color[0, 0]).max() < 1e-6
    # a translated mesh moves the picture; uncovered pixels stay white
    out = R.rasterize_warp(warp + f32(2), color, np.zeros((H, W), dtype=f32))
    assert np.abs(out[2:, 2:] - color[:-2, :-2]).max() < 1e-6
    assert (out[:2] == 255).all() and (out[:, :2] == 255).all()


def test_skinning_reproduces_lattice_interpolation():
    dims = (3, 4, 2)
    g = np.stack(np.meshgrid(np.arange(4), np.arange(5), np.arange(3), indexing="ij"), axis=-1).reshape(-1, 3).astype(f32)
    rng = np.random.default_rng(2)
    voxel = np.stack([rng.integers(0, d, 50) for d in dims], axis=1).astype(np.int32)
    rel = rng.random((50, 3), dtype=f32)
    out = R.skin_trilinear(g * f32(0.5) + f32(1), dims, voxel, rel)     # affine lattice: skinning is exact
    assert np.abs(out - ((voxel + rel) * 0.5 + 1)).max() < 1e-6


def test_flow_lines():
    H, W = 12, 17
    flow = np.zeros((H, W, 2), dtype=f32)
    flow[1, 1] = (3.4, 0.0)           # horizontal: pixels x = 1..3 of row 1
    flow[6, 6] = (0.0, 2.6)           # vertical downwards: rows 6..8 of column 6
    flow[6, 11] = (-2.0, -2.0)        # drawn from the end point towards the start
    n = len(range(1, W - 1, 5)) * len(range(1, H - 1, 5))
    colors = np.arange(1, n + 1, dtype=np.uint8)[:, None].repeat(4, axis=1)
    img = R.render_flow_lines(np.zeros((H, W, 4), dtype=np.uint8), flow, colors)
    assert [int(v) for v in np.flatnonzero(img[1, :, 0])] == [1, 2, 3]
    assert [int(v) for v in np.flatnonzero(img[:, 6, 0])] == [6, 7, 8]
    assert img[4, 9, 0] != 0 and img[5, 10, 0] != 0 and img[6, 11, 0] == 0
