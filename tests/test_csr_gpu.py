"""GPU: the CSR kernels of the materialized-Jacobian mode (opt_b200/csrc/csr.cuh behind include/OptB200Sparse.h) against
the reference's own known-answer test for its matrix library, API/src/linalg_cpu_test.t:20-190 (integer-valued, so the
comparison is exact), and against scipy on random matrices; then the three Opt_InitializationParameters modes
(matrix-free, useMaterializedJTJ, + useFusedJTJ) on the energies with Jacobian-row emitters, as
examples/test_final_cost.py:91-93 runs them: same recorded (1,1,1) costs, same iterates."""
import ctypes as C
import os

import numpy as np
import pytest

from opt_b200 import examples
from gpu_util import OP_RTOL, assert_costs_close, assert_unknowns_close, run_oracle

pytestmark = pytest.mark.gpu


def _lib():
    from opt_b200 import solver as S
    lib, _ = S.load_library()
    P, I = C.c_void_p, C.c_int
    lib.OptB200_CsrTranspose.argtypes = [I, I, I, P, P, P, P, P, P]
    lib.OptB200_CsrAta.argtypes = [I, I, I, P, P, P, P, P, P, I, C.POINTER(I)]
    lib.OptB200_CsrSpmv.argtypes = [I, I, I, P, P, P, P, P]
    return lib


def _dev(a):
    import torch
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def csr_transpose(lib, nrows, ncols, rp, ci, v):
    import torch
    d = [_dev(rp.astype(np.int32)), _dev(ci.astype(np.int32)), _dev(v.astype(np.float32))]
    o = [torch.zeros(ncols + 1, dtype=torch.int32, device="cuda"), torch.zeros(len(ci), dtype=torch.int32, device="cuda"),
         torch.zeros(len(ci), dtype=torch.float32, device="cuda")]
    assert lib.OptB200_CsrTranspose(nrows, ncols, len(ci), *[t.data_ptr() for t in d + o]) == 0
    return [t.cpu().numpy() for t in o]


def csr_ata(lib, nrows, ncols, rp, ci, v, cap):
    import torch
    d = [_dev(rp.astype(np.int32)), _dev(ci.astype(np.int32)), _dev(v.astype(np.float32))]
    o = [torch.zeros(ncols + 1, dtype=torch.int32, device="cuda"), torch.zeros(cap, dtype=torch.int32, device="cuda"),
         torch.zeros(cap, dtype=torch.float32, device="cuda")]
    n = C.c_int(0)
    rc = lib.OptB200_CsrAta(nrows, ncols, len(ci), *[t.data_ptr() for t in d + o], cap, C.byref(n))
    return rc, n.value, [t.cpu().numpy() for t in o]


def csr_spmv(lib, nrows, ncols, rp, ci, v, x):
    import torch
    d = [_dev(rp.astype(np.int32)), _dev(ci.astype(np.int32)), _dev(v.astype(np.float32)), _dev(x.astype(np.float32))]
    y = torch.zeros(nrows, dtype=torch.float32, device="cuda")
    assert lib.OptB200_CsrSpmv(nrows, ncols, len(ci), *[t.data_ptr() for t in d], y.data_ptr()) == 0
    return y.cpu().numpy()


def test_reference_known_answer(built):
    """API/src/linalg_cpu_test.t:20-190: A = [[1,0,3,4],[0,9,0,2],[7,0,8,0]], x = (3,5,7,9)."""
    lib = _lib()
    rp = np.array([0, 3, 5, 7]); ci = np.array([0, 2, 3, 1, 3, 0, 2]); v = np.array([1, 3, 4, 9, 2, 7, 8.0])
    rpT, ciT, vT = csr_transpose(lib, 3, 4, rp, ci, v)
    assert rpT.tolist() == [0, 2, 3, 5, 7]                       # :95-99
    assert ciT.tolist() == [0, 2, 1, 0, 2, 0, 1]                 # :101-107
    assert vT.tolist() == [1, 7, 9, 3, 8, 4, 2]                  # :109-115
    assert csr_spmv(lib, 3, 4, rp, ci, v, np.array([3, 5, 7, 9.0])).tolist() == [60, 63, 77]     # :117-119
    rc, n, (rpA, ciA, vA) = csr_ata(lib, 3, 4, rp, ci, v, 16)
    assert rc == 0 and n == 12
    assert rpA.tolist() == [0, 3, 5, 8, 12]                      # :153-157
    assert ciA[:12].tolist() == [0, 2, 3, 1, 3, 0, 2, 3, 0, 1, 2, 3]                             # :159-170
    assert vA[:12].tolist() == [50, 59, 4, 81, 18, 59, 73, 12, 4, 18, 12, 20]                   # :172-183
    rc, n, _ = csr_ata(lib, 3, 4, rp, ci, v, 5)                  # too small: reports the size
    assert rc == -1 and n == 12


@pytest.mark.parametrize("shape,density", [((200, 150), 0.03), ((1000, 40), 0.2), ((37, 500), 0.01), ((5000, 5000), 0.001)])
def test_csr_kernels_match_scipy(built, shape, density):
    import scipy.sparse as sp
    lib = _lib()
    rng = np.random.default_rng(3)
    A = sp.random(*shape, density=density, format="csr", random_state=rng, data_rvs=lambda n: (rng.integers(1, 5, n) * rng.choice([-1, 1], n)).astype(np.float64))     # no explicit zeros
    A.sort_indices()
    rp, ci, v = A.indptr, A.indices, A.data
    rpT, ciT, vT = csr_transpose(lib, shape[0], shape[1], rp, ci, v)
    AT = A.T.tocsr(); AT.sort_indices()
    assert np.array_equal(rpT, AT.indptr) and np.array_equal(ciT, AT.indices) and np.array_equal(vT, AT.data.astype(np.float32))
    x = rng.integers(-3, 4, shape[1]).astype(np.float64)
    assert np.array_equal(csr_spmv(lib, shape[0], shape[1], rp, ci, v, x), (A @ x).astype(np.float32))
    # pattern of A^T A: structural (explicit zeros of cancelling products stay), values exact (small integers)
    S = (abs(A).T @ abs(A)).tocsr(); S.sort_indices()
    rc, n, (rpA, ciA, vA) = csr_ata(lib, shape[0], shape[1], rp, ci, v, S.nnz + 8)
    assert rc == 0 and n == S.nnz
    assert np.array_equal(rpA, S.indptr) and np.array_equal(ciA[:n], S.indices)
    G = (A.T @ A).tocsr()
    rows, cols = S.nonzero()                                     # CSR order: row-major, columns ascending
    assert np.array_equal(vA[:n], np.asarray(G[rows, cols]).reshape(-1).astype(np.float32))


# ------------------------------------------------------------------------------ the three modes of the solver
MODES = [(False, False), (True, False), (True, True)]            # examples/test_final_cost.py:91-93


def _solve(pr, materialized, fused, timing=False):
    import torch
    from opt_b200 import solver as S
    named = S.upload(pr)
    sol = S.OptSolver(pr.dims, pr.energy, pr.kind, double=pr.double, materialized=materialized, fused_jtj=fused, timing=timing)
    costs = []
    sol.solve(pr.solver, named, costs)
    torch.cuda.synchronize()
    xs = {i: named.values[i].cpu().numpy() for i in pr.unknowns}
    return costs, xs, sol


GOLDENS = {"poisson": ("poisson_stride4.npz", examples.poisson_golden, 1530364.25),
           "image_warping": ("image_warping_cat512.npz", examples.image_warping_golden, 1774.3405),
           "arap": ("arap_small_armadillo.npz", examples.arap_golden, 7183.464843)}


@pytest.mark.parametrize("materialized,fused", MODES)
@pytest.mark.parametrize("name", sorted(GOLDENS))
def test_recorded_costs_in_all_three_modes(built, golden_dir, name, materialized, fused):
    """The (1,1,1) costs of examples/test_final_cost.py:56-66, relative tolerance 1e-5 as there (:156)."""
    f, builder, golden = GOLDENS[name]
    pr = builder(np.load(os.path.join(golden_dir, f)))
    costs, _, sol = _solve(pr, materialized, fused)
    assert sol.materialized_mode() == (2 if fused else 1 if materialized else 0)
    assert abs(costs[-1] - golden) <= 1e-5 * golden, (costs, golden)


CASES = {"poisson": (examples.poisson_synthetic, (67, 45), dict(n_iter=2, l_iter=12)),
         "image_warping": (examples.image_warping_synthetic, (61, 47), dict(n_iter=3, l_iter=10)),
         "arap": (examples.arap_synthetic, (13, 11), dict(n_iter=3, l_iter=10))}


@pytest.mark.parametrize("double", [False, True])
@pytest.mark.parametrize("kind", ["gaussNewtonGPU", "LMGPU"])
@pytest.mark.parametrize("materialized,fused", MODES[1:])
@pytest.mark.parametrize("name", sorted(CASES))
def test_materialized_solve_matches_oracle(built, name, materialized, fused, kind, double):
    b, args, kw = CASES[name]
    pr = b(*args, **kw)
    pr.kind = kind
    if double:
        pr = pr.as_double()
    costs, xs, sol = _solve(pr, materialized, fused, timing=True)
    assert sol.materialized_mode() == (2 if fused else 1)
    assert sol.kernel_time("J^TJp" if fused else "J^T")[0] > 0 and sol.kernel_time("PCGStep1")[0] == 0
    assert sol.materialized_nnz(0) > 0 and (sol.materialized_nnz(1) > 0) == fused
    rc, rx, _ = run_oracle(pr)
    assert_costs_close(costs, rc, double)
    assert_unknowns_close(xs, rx, double)


@pytest.mark.parametrize("materialized,fused", MODES[1:])
@pytest.mark.parametrize("name", sorted(CASES))
def test_materialized_operator_matches_oracle(built, name, materialized, fused):
    """J^T J p through the stored matrices for a random p against the oracle's J^T J p (dual numbers): checks every
    Jacobian row the functors emit."""
    from opt_b200 import solver as S
    from oracle_api import OracleSolver
    b, args, kw = CASES[name]
    pr = b(*args, **kw).as_double()
    named = S.upload(pr)
    sol = S.OptSolver(pr.dims, pr.energy, pr.kind, double=True, materialized=materialized, fused_jtj=fused)
    o = OracleSolver(pr.dims, pr.energy, pr.kind, double=True, threads=4)
    r_ref, _ = o.eval_jtf(pr.params, np.float64)
    rng = np.random.default_rng(5)
    p = rng.standard_normal(r_ref.shape) * (r_ref != 0)
    ap = sol.debug_apply_jtj(named, p)
    ap_ref = o.apply_jtj(pr.params, p)
    assert np.abs(ap - ap_ref).max() <= OP_RTOL[True] * np.abs(ap_ref).max()
