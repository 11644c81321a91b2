"""Host-side mirror of the per-pixel / per-vertex steps the reference's example drivers run around the solver
(include/OptB200Drivers.h), over the C-ABI library, plus the optical-flow driver's hierarchical solve
(examples/optical_flow/src/CombinedSolver.h:18-110) built from them.

PyTorch provides device memory only; every computation is a CUDA kernel of libOptB200.so.  No fallback."""
import ctypes as C

import numpy as np

from . import solver as S

_bound = False


def _lib():
    global _bound
    lib, _ = S.load_library()
    if not _bound:
        P, U, L = C.c_void_p, C.c_uint, C.c_long
        lib.OptB200_FilterGaussian.argtypes = [P, U, U, C.c_float, P, P]
        lib.OptB200_PrewittDU.argtypes = [P, P, U, U, P]
        lib.OptB200_PrewittDV.argtypes = [P, P, U, U, P]
        lib.OptB200_RasterizeWarp.argtypes = [P, P, P, U, U, P, P, P]
        lib.OptB200_SkinTrilinear.argtypes = [P, C.POINTER(C.c_int), P, P, L, P, P]
        lib.OptB200_RenderFlowLines.argtypes = [P, P, P, U, U, U, P, P]
        for n in ("FilterGaussian", "PrewittDU", "PrewittDV", "RasterizeWarp", "SkinTrilinear", "RenderFlowLines"):
            getattr(lib, "OptB200_" + n).restype = C.c_int
        _bound = True
    return lib


def _check(rc, what):
    if rc != 0:
        raise RuntimeError("%s failed with CUDA error %d" % (what, rc))


def _stream():
    import torch
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _f32(t, shape_tail=None):
    import torch
    assert t.is_cuda and t.dtype == torch.float32 and t.is_contiguous()
    return t


def filter_gaussian(image, sigma):
    """In place; `image` is an H x W float32 CUDA tensor (ImageHelper::filterGaussian)."""
    import torch
    _f32(image)
    H, W = image.shape
    scratch = torch.empty_like(image)
    _check(_lib().OptB200_FilterGaussian(image.data_ptr(), W, H, float(sigma), scratch.data_ptr(), _stream()), "OptB200_FilterGaussian")
    return image


def prewitt(image, horizontal):
    """computeDU (horizontal) / computeDV of an H x W float32 CUDA tensor."""
    import torch
    _f32(image)
    H, W = image.shape
    out = torch.empty_like(image)
    fn = _lib().OptB200_PrewittDU if horizontal else _lib().OptB200_PrewittDV
    _check(fn(image.data_ptr(), out.data_ptr(), W, H, _stream()), "OptB200_Prewitt")
    return out


def rasterize_warp(warp, color, mask):
    """warp H x W x 2, color H x W x 3, mask H x W (float32 CUDA tensors) -> H x W x 3 result image."""
    import torch
    _f32(warp), _f32(color), _f32(mask)
    H, W = mask.shape
    result = torch.empty((H, W, 3), dtype=torch.float32, device=mask.device)
    winner = torch.empty((H, W), dtype=torch.int32, device=mask.device)
    _check(_lib().OptB200_RasterizeWarp(warp.data_ptr(), color.data_ptr(), mask.data_ptr(), W, H, result.data_ptr(),
                                        winner.data_ptr(), _stream()), "OptB200_RasterizeWarp")
    return result


def skin_trilinear(grid, dims, voxel, rel):
    """grid ((dims[0]+1)*(dims[1]+1)*(dims[2]+1)) x 3 float32, voxel n x 3 int32, rel n x 3 float32 -> n x 3."""
    import torch
    _f32(grid), _f32(rel)
    assert voxel.dtype == torch.int32 and voxel.is_contiguous()
    n = voxel.shape[0]
    out = torch.empty((n, 3), dtype=torch.float32, device=grid.device)
    d = (C.c_int * 3)(*[int(v) for v in dims])
    _check(_lib().OptB200_SkinTrilinear(grid.data_ptr(), d, voxel.data_ptr(), rel.data_ptr(), n, out.data_ptr(), _stream()), "OptB200_SkinTrilinear")
    return out


def render_flow_lines(image, flow, colors, skip=5):
    """image H x W x 4 uint8 (drawn into), flow H x W x 2 float32, colors nsamples x 4 uint8."""
    import torch
    assert image.dtype == torch.uint8 and colors.dtype == torch.uint8 and image.is_contiguous() and colors.is_contiguous()
    _f32(flow)
    H, W = flow.shape[:2]
    assert colors.shape[0] == flow_sample_count(W, H, skip)
    winner = torch.empty((H, W), dtype=torch.int32, device=flow.device)
    _check(_lib().OptB200_RenderFlowLines(image.data_ptr(), flow.data_ptr(), colors.data_ptr(), W, H, skip, winner.data_ptr(), _stream()), "OptB200_RenderFlowLines")
    return image


def flow_sample_count(W, H, skip=5):
    return len(range(1, W - 1, skip)) * len(range(1, H - 1, skip))


class OpticalFlowDriver:
    """The reference's optical-flow CombinedSolver (CombinedSolver.h:18-110) on the device: a two-level Gaussian
    hierarchy (sigma 1 and 5) of source and target, Prewitt derivatives of the filtered target, and for every
    level, coarse first, `num_iter` solves with the fit weight ramped from 10 towards 50.  As in the reference,
    preSingleSolve resets the flow of EVERY level to zero, after the coarse-to-fine copy (CombinedSolver.h:52-56,
    95-97 -- the copy is therefore without effect; reproduced, not fixed)."""
    SIGMAS = (1.0, 5.0)

    def __init__(self, source, target, kind="gaussNewtonGPU", num_iter=3, n_iter=1, l_iter=50, double=False):
        import torch
        self.kind, self.num_iter, self.n_iter, self.l_iter = kind, num_iter, n_iter, l_iter
        H, W = source.shape
        self.dims = [W, H]
        self.levels = []
        for sg in self.SIGMAS:
            t = filter_gaussian(torch.as_tensor(target, dtype=torch.float32).cuda().clone(), sg)
            s = filter_gaussian(torch.as_tensor(source, dtype=torch.float32).cuda().clone(), sg)
            self.levels.append(dict(source=s, target=t, du=prewitt(t, True), dv=prewitt(t, False),
                                    flow=torch.zeros((H, W, 2), dtype=torch.float32, device="cuda")))
        self.solver = S.OptSolver(self.dims, "optical_flow.t", kind, double=double)
        self.costs = []

    def solve_all(self):
        f32 = np.float32
        weight_fit = f32(10.0)
        step = f32((f32(50.0) - weight_fit) / f32(self.num_iter * len(self.levels)))
        reg_sqrt = np.sqrt(f32(0.1))
        for li in range(len(self.levels) - 1, -1, -1):
            if li < len(self.levels) - 1:
                self.levels[li]["flow"].copy_(self.levels[li + 1]["flow"])
            for lv in self.levels:                       # preSingleSolve -> resetGPU
                lv["flow"].zero_()
            lv = self.levels[li]
            for _ in range(self.num_iter):
                weight_fit = f32(weight_fit + step)       # preNonlinearSolve
                named = S.NamedParameters()
                named.set("w_fit", np.sqrt(weight_fit)).set("w_reg", reg_sqrt).set("X", lv["flow"])
                named.set("I", lv["source"]).set("I_hat", lv["target"]).set("I_hat_dx", lv["du"]).set("I_hat_dy", lv["dv"])
                self.costs.append(self.solver.solve({"nIterations": self.n_iter, "lIterations": self.l_iter}, named))
        return self.levels[0]["flow"]
