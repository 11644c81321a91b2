"""Device time of the driver pre/post-processing kernels (include/OptB200Drivers.h) against their algorithmic
HBM traffic, with the numpy restatement (oracle/driver_ref.py) timed beside them on a bounded sample:
python tools/bench_driver_ops.py [size]  -> one JSON line per kernel."""
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import driver_ref as R  # noqa: E402
from opt_b200 import drivers as D  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
peak = 6447.5
if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")):
    peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
rng = np.random.default_rng(1234)


def dev_ms(fn, reps=20):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / reps


def cpu_ms(fn):
    t = time.perf_counter()
    fn()
    return (time.perf_counter() - t) * 1e3


def line(name, ms, alg_bytes, cpu, cpu_sample):
    print(json.dumps({"kernel": name, "size": n, "ms": ms, "algorithmic_GBps": alg_bytes / ms / 1e6, "frac_of_measured_peak": alg_bytes / ms / 1e6 / peak,
                      "cpu_numpy_ms_scaled": cpu, "cpu_sample": cpu_sample}))


im = (rng.random((n, n), dtype=np.float32) * 255).astype(np.float32)
d_im = torch.from_numpy(im).cuda()
small = im[:512, :512].copy()
scale = (n / 512.0) ** 2
for sigma in (1.0, 5.0):
    line("FilterGaussian sigma=%g" % sigma, dev_ms(lambda: D.filter_gaussian(d_im, sigma)), 4 * 4 * n * n,   # two passes: read + write each
         cpu_ms(lambda: R.filter_gaussian(small, sigma)) * scale, "512x512 numpy, scaled by area")
line("PrewittDU", dev_ms(lambda: D.prewitt(d_im, True)), 2 * 4 * n * n, cpu_ms(lambda: R.prewitt(small, True)) * scale, "512x512 numpy, scaled by area")
ys, xs = np.mgrid[0:n, 0:n]
warp = (np.stack([xs, ys], axis=-1) + 0.3 * rng.standard_normal((n, n, 2))).astype(np.float32)
color = (rng.random((n, n, 3), dtype=np.float32) * 255).astype(np.float32)
mask = np.zeros((n, n), dtype=np.float32)
dw, dc, dm = (torch.from_numpy(a).cuda() for a in (warp, color, mask))
t = cpu_ms(lambda: R.rasterize_warp(warp[:48, :48], color[:48, :48], mask[:48, :48])) * (n / 48.0) ** 2
line("RasterizeWarp", dev_ms(lambda: D.rasterize_warp(dw, dc, dm), 5), (8 + 12 + 4 + 12 + 4) * n * n, t, "48x48 pure-Python loops, scaled by area")
dims = (255, 255, 255)
grid = rng.standard_normal((256 ** 3, 3)).astype(np.float32)
nv = 4_000_000
voxel = np.stack([rng.integers(0, d, nv) for d in dims], axis=1).astype(np.int32)
rel = rng.random((nv, 3), dtype=np.float32)
dg, dv, dr = (torch.from_numpy(a).cuda() for a in (grid, voxel, rel))
line("SkinTrilinear 4M vertices / 256^3 lattice", dev_ms(lambda: D.skin_trilinear(dg, dims, dv, dr)), nv * (12 + 12 + 12 + 8 * 12),
     cpu_ms(lambda: R.skin_trilinear(grid, dims, voxel, rel)), "full size numpy")
