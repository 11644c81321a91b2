"""GPU: OptB200_FilterGaussianNoAlloc (include/OptB200Drivers.h) -- the separable Gaussian with its intermediate image in a
static device array inside the library instead of a caller-owned scratch buffer, for callers that cannot allocate.  Called
through the C ABI and compared with the existing entry point OptB200_FilterGaussian on two calls in a row (the second one on
a smaller image, so the array still holds the first call's contents where the second does not write), plus the size limit.
Collected last on purpose; passes alone in a fresh process and after the other tests."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
f32 = np.float32
MAX_PIXELS = 2048 * 2048            # OPTB200_GAUSS_NOALLOC_MAX_PIXELS


def _lib():
    from opt_b200 import solver as S
    lib, _ = S.load_library()
    P, U = C.c_void_p, C.c_uint
    lib.OptB200_FilterGaussian.restype = C.c_int
    lib.OptB200_FilterGaussian.argtypes = [P, U, U, C.c_float, P, P]
    lib.OptB200_FilterGaussianNoAlloc.restype = C.c_int
    lib.OptB200_FilterGaussianNoAlloc.argtypes = [P, U, U, C.c_float, P]
    return lib


def _both(lib, im, sigma):
    """(existing entry point with a torch scratch image, new entry point) on copies of `im`; default stream."""
    import torch
    H, W = im.shape
    a = torch.from_numpy(im).cuda()
    scratch = torch.empty_like(a)
    assert lib.OptB200_FilterGaussian(a.data_ptr(), W, H, sigma, scratch.data_ptr(), None) == 0
    b = torch.from_numpy(im).cuda()
    torch.cuda.synchronize()
    before = torch.cuda.memory_stats().get("allocation.all.allocated", 0)
    assert lib.OptB200_FilterGaussianNoAlloc(b.data_ptr(), W, H, sigma, None) == 0
    torch.cuda.synchronize()
    assert torch.cuda.memory_stats().get("allocation.all.allocated", 0) == before      # nothing came from torch's allocator either
    return a.cpu().numpy(), b.cpu().numpy()


def test_no_alloc_gaussian_matches_existing_entry_point_twice(built):
    lib = _lib()
    rng = np.random.default_rng(11)
    first = (rng.random((300, 517), dtype=f32) * 255).astype(f32)
    second = (rng.random((130, 33), dtype=f32) * 255).astype(f32)          # smaller: stale contents of call 1 stay beyond it
    for im, sigma in ((first, 5.0), (second, 1.0), (first, 1.0)):
        ref, got = _both(lib, im, sigma)
        assert np.isfinite(got).all() and not np.array_equal(got, im)
        assert np.abs(got - ref).max() <= 1e-6 * 255                        # the tolerance of tests/test_drivers_gpu.py
        assert np.array_equal(got, ref)                                     # same kernels: in fact bit-identical


def test_no_alloc_gaussian_rejects_images_beyond_the_static_scratch(built):
    import torch
    lib = _lib()
    im = torch.zeros(8, device="cuda")
    assert lib.OptB200_FilterGaussianNoAlloc(im.data_ptr(), 4096, 2048, 1.0, None) != 0        # 2 x the maximum: refused, no launch
    assert lib.OptB200_FilterGaussianNoAlloc(im.data_ptr(), 0, 16, 1.0, None) != 0
    torch.cuda.synchronize()
    full = torch.rand(2048, 2048, device="cuda") * 255                                         # exactly the maximum: accepted
    ref = full.clone()
    scratch = torch.empty_like(ref)
    assert lib.OptB200_FilterGaussian(ref.data_ptr(), 2048, 2048, 1.0, scratch.data_ptr(), None) == 0
    assert lib.OptB200_FilterGaussianNoAlloc(full.data_ptr(), 2048, 2048, 1.0, None) == 0
    torch.cuda.synchronize()
    assert torch.equal(full, ref)
    assert 2048 * 2048 == MAX_PIXELS
