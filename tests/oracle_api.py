"""TEST INFRASTRUCTURE: Python handle on the CPU oracle (oracle/liboptoracle.so).

Mirrors the reference's OptSolver helper (examples/shared/OptSolver.h:43-112) over host numpy
arrays.  Never imported by the product package.
"""
import ctypes as C
import os
import subprocess

import numpy as np

from opt_b200 import capi

_ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
_LIB = None


def build_oracle():
    subprocess.check_call(["make", "-s", "-C", os.path.join(_ROOT, "oracle")])


def lib():
    global _LIB
    if _LIB is None:
        path = os.path.join(_ROOT, "oracle", "liboptoracle.so")
        if not os.path.exists(path):
            build_oracle()
        _LIB = C.CDLL(path)
        _LIB.OptOracle_UnknownCount.restype = C.c_long
        _LIB.OptOracle_UnknownCount.argtypes = [C.c_void_p]
        _LIB.OptOracle_PcgIterations.restype = C.c_long
        _LIB.OptOracle_PcgIterations.argtypes = [C.c_void_p]
        _LIB.OptOracle_CostHistory.restype = C.c_int
        _LIB.OptOracle_CostHistory.argtypes = [C.c_void_p, C.POINTER(C.c_double), C.c_int]
        _LIB.OptOracle_ApplyJTJ.argtypes = [C.c_void_p, C.POINTER(C.c_void_p), C.c_void_p, C.c_void_p]
        _LIB.OptOracle_EvalJTF.argtypes = [C.c_void_p, C.POINTER(C.c_void_p), C.c_void_p, C.c_void_p]
        _LIB.OptOracle_FreeState.argtypes = [C.c_void_p]
    return _LIB


def pack_params(params, keep):
    """problemparams as void*[]: arrays by address, scalars as pointers to a C float
    (examples/shared/NamedParameters.h:35-48)."""
    arr = (C.c_void_p * len(params))()
    for i, p in enumerate(params):
        if p is None:
            arr[i] = None
        elif isinstance(p, np.ndarray):
            assert p.flags["C_CONTIGUOUS"]
            arr[i] = p.ctypes.data
        else:
            if isinstance(p, (np.float64,)):        # opt_float-typed Param in double mode
                box = C.c_double(float(p))
            elif isinstance(p, (int, np.integer)):
                box = C.c_int(int(p))
            else:
                box = C.c_float(float(p))
            keep.append(box)
            arr[i] = C.addressof(box)
    return arr


class OracleSolver:
    def __init__(self, dims, energy, kind="gaussNewtonGPU", double=False, threads=1, verbosity=0):
        self.f = capi.bind(lib(), "OptOracle_")
        ip = capi.Opt_InitializationParameters()
        ip.doublePrecision = int(double)
        ip.verbosityLevel = verbosity
        ip.backend = b"backend_cpu_mt"
        ip.numthreads = threads
        self.state = self.f["NewState"](ip)
        self.problem = self.f["ProblemDefine"](self.state, energy.encode(), kind.encode())
        assert self.problem, energy
        d = (C.c_uint * len(dims))(*dims)
        self.plan = self.f["ProblemPlan"](self.state, self.problem, d)
        assert self.plan
        self._keep = []

    def set_params(self, **kw):
        for k, v in kw.items():
            box = C.c_float(v) if isinstance(v, float) else C.c_int(v)
            self.f["SetSolverParameter"](self.state, self.plan, k.encode(), C.addressof(box))

    def init(self, params):
        self._pp = pack_params(params, self._keep)
        self.f["ProblemInit"](self.state, self.plan, self._pp)
        return self.cost()

    def step(self):
        return self.f["ProblemStep"](self.state, self.plan, self._pp)

    def solve(self, params):
        self.init(params)
        while self.step():
            pass
        return self.cost()

    def cost(self):
        return self.f["ProblemCurrentCost"](self.state, self.plan)

    def cost_history(self):
        buf = (C.c_double * 4096)()
        n = lib().OptOracle_CostHistory(self.plan, buf, 4096)
        return [buf[i] for i in range(min(n, 4096))]

    def pcg_iterations(self):
        return lib().OptOracle_PcgIterations(self.plan)

    def unknown_count(self):
        return lib().OptOracle_UnknownCount(self.plan)

    def apply_jtj(self, params, p):
        pp = pack_params(params, self._keep)
        out = np.zeros_like(p)
        lib().OptOracle_ApplyJTJ(self.plan, pp, p.ctypes.data, out.ctypes.data)
        return out

    def eval_jtf(self, params, dtype):
        pp = pack_params(params, self._keep)
        n = self.unknown_count()
        g = np.zeros(n, dtype=dtype)
        pre = np.zeros(n, dtype=dtype)
        lib().OptOracle_EvalJTF(self.plan, pp, g.ctypes.data, pre.ctypes.data)
        return g, pre

    def close(self):
        if self.plan:
            self.f["PlanFree"](self.state, self.plan)
            self.f["ProblemDelete"](self.state, self.problem)
            lib().OptOracle_FreeState(self.state)
            self.plan = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
