"""Host-side mirror of the reference's driver helpers over the C-ABI library libOptB200.so.

`OptSolver` follows examples/shared/OptSolver.h:43-112 (constructor = NewState + ProblemDefine +
ProblemPlan, `solve` = setAllSolverParameters + Init/Step loop + CurrentCost) and
`NamedParameters` follows examples/shared/NamedParameters.h:35-48 (flat void*[] in first-set order).
PyTorch is used for device memory only: images are CUDA tensors whose `data_ptr()` goes into
`problemparams`; all computation happens inside the library.  There is no fallback: if the
library is missing or no CUDA device is present, construction raises.
"""
import ctypes as C
import os

import numpy as np

from . import capi

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libOptB200.so")
_lib = None
_fn = None

FLOAT_PARAMS = {"min_relative_decrease", "min_trust_region_radius", "max_trust_region_radius", "q_tolerance",
                "function_tolerance", "trust_region_radius", "radius_decrease_factor", "min_lm_diagonal",
                "max_lm_diagonal"}
INT_PARAMS = {"residual_reset_period", "nIter", "nIterations", "lIterations"}


def load_library():
    """dlopen libOptB200.so and bind every Opt.h / OptB200.h entry point (raises if absent)."""
    global _lib, _fn
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError("libOptB200.so is not built (run `python -m opt_b200.build`); "
                               "there is no fallback implementation")
        try:                    # torch bundles its own NCCL; its symbols must be resolved before the system libnccl
            import torch        # noqa: F401  that libOptB200.so links is mapped (undefined ncclDevCommCreate otherwise)
        except ImportError:
            pass
        _lib = C.CDLL(LIB_PATH)
        _fn = capi.bind(_lib, "Opt_")
        P = C.c_void_p
        _lib.OptB200_FreeState.argtypes = [P]
        _lib.OptB200_PlanLaunchCount.restype = C.c_long
        _lib.OptB200_PlanLaunchCount.argtypes = [P]
        _lib.OptB200_PlanMaterializedMode.restype = C.c_int
        _lib.OptB200_PlanMaterializedMode.argtypes = [P]
        _lib.OptB200_PlanMaterializedNnz.restype = C.c_long
        _lib.OptB200_PlanMaterializedNnz.argtypes = [P, C.c_int]
        _lib.OptB200_PlanPcgIterations.restype = C.c_long
        _lib.OptB200_PlanPcgIterations.argtypes = [P]
        _lib.OptB200_PlanKernelTime.restype = C.c_int
        _lib.OptB200_PlanKernelTime.argtypes = [P, C.c_char_p, C.POINTER(C.c_double)]
        _lib.OptB200_PlanResetTiming.argtypes = [P]
        _lib.OptB200_PlanUnknownChannels.restype = C.c_int
        _lib.OptB200_PlanUnknownChannels.argtypes = [P]
        _lib.OptB200_PlanSiteCount.restype = C.c_long
        _lib.OptB200_PlanSiteCount.argtypes = [P]
        _lib.OptB200_DebugReadVector.restype = C.c_int
        _lib.OptB200_DebugReadVector.argtypes = [P, C.c_char_p, P]
        _lib.OptB200_DebugReadVectorRange.restype = C.c_int
        _lib.OptB200_DebugReadVectorRange.argtypes = [P, C.c_char_p, C.c_long, C.c_long, P]
        _lib.OptB200_DebugApplyJTJ.argtypes = [P, C.POINTER(P), P, P]
        _lib.OptB200_DebugPcgBegin.argtypes = [P, C.POINTER(P)]
        _lib.OptB200_SetErrorHandler.argtypes = [P, P]
        _lib.OptB200_LastError.restype = C.c_char_p
        _lib.OptB200_LastError.argtypes = [C.POINTER(C.c_int)]
        _lib.OptB200_PlanFailed.restype = C.c_int
        _lib.OptB200_PlanFailed.argtypes = [P]
    return _lib, _fn


ERROR_HANDLER = C.CFUNCTYPE(None, C.c_int, C.c_char_p, C.c_void_p)
_handler_keepalive = None


def set_error_handler(fn):
    """OptB200_SetErrorHandler: fn(code, message) is called once per fatal CUDA / NCCL error and the failing call returns
    instead of the process exiting (include/OptB200.h); None restores the reference's print-and-exit."""
    global _handler_keepalive
    lib, _ = load_library()
    if fn is None:
        lib.OptB200_SetErrorHandler(None, None)
        _handler_keepalive = None
        return
    cb = ERROR_HANDLER(lambda code, msg, user: fn(int(code), (msg or b"").decode(errors="replace")))
    lib.OptB200_SetErrorHandler(C.cast(cb, C.c_void_p), None)
    _handler_keepalive = cb


def last_error():
    lib, _ = load_library()
    code = C.c_int(0)
    msg = lib.OptB200_LastError(C.byref(code))
    return code.value, (msg or b"").decode(errors="replace")


class NamedParameters:
    """Ordered name -> value map flattened into void*[] (NamedParameters.h:35-48).
    Values: CUDA tensors (images / graph index arrays, by device address), Python/NumPy floats
    (host `float*`, re-read by the library at Init and at every Step) or None."""

    def __init__(self):
        self.names, self.values, self._boxes = [], [], {}

    def set(self, name, value):
        if name in self.names:
            self.values[self.names.index(name)] = value
        else:
            self.names.append(name)
            self.values.append(value)
        return self

    def data(self):
        arr = (C.c_void_p * len(self.values))()
        for i, v in enumerate(self.values):
            if v is None:
                arr[i] = None
            elif hasattr(v, "data_ptr"):
                assert v.is_cuda and v.is_contiguous(), self.names[i]
                arr[i] = v.data_ptr()
            else:
                box = C.c_double(float(v)) if isinstance(v, np.float64) else C.c_float(float(v))
                self._boxes[i] = box
                arr[i] = C.addressof(box)
        return arr


class OptSolver:
    def __init__(self, dims, energy_file, solver_kind="gaussNewtonGPU", double=False, verbosity=0,
                 timing=False, backend="backend_cuda", materialized=False, fused_jtj=False):
        self.lib, self.f = load_library()
        ip = capi.Opt_InitializationParameters()
        ip.doublePrecision = int(double)
        ip.verbosityLevel = verbosity
        ip.collectPerKernelTimingInfo = int(timing)
        ip.backend = backend.encode()
        ip.numthreads = 1
        ip.useMaterializedJTJ = int(materialized)
        ip.useFusedJTJ = int(fused_jtj)
        self.double = double
        self.state = self.f["NewState"](ip)
        if not self.state:
            raise RuntimeError("Opt_NewState failed (backend %r)" % backend)
        self.problem = self.f["ProblemDefine"](self.state, energy_file.encode(), solver_kind.encode())
        if not self.problem:
            raise RuntimeError("Opt_ProblemDefine rejected %r" % energy_file)
        d = (C.c_uint * len(dims))(*[int(x) for x in dims])
        self.plan = self.f["ProblemPlan"](self.state, self.problem, d)
        if not self.plan:
            raise RuntimeError("Opt_ProblemPlan failed")
        self._pp = None
        self._params = None

    # --- Opt.h calls ---------------------------------------------------------------------------
    def set_solver_parameters(self, params):
        for k, v in params.items():
            if k in FLOAT_PARAMS:
                box = C.c_float(float(v))
            elif k in INT_PARAMS:
                box = C.c_uint(int(v))          # drivers pass unsigned int* (CombinedSolverParameters.h)
            else:
                box = C.c_float(float(v))
            self.f["SetSolverParameter"](self.state, self.plan, k.encode(), C.addressof(box))

    def init(self, named):
        self._params = named
        self._pp = named.data()
        self.f["ProblemInit"](self.state, self.plan, self._pp)

    def step(self):
        return self.f["ProblemStep"](self.state, self.plan, self._pp)

    def current_cost(self):
        return self.f["ProblemCurrentCost"](self.state, self.plan)

    def solve(self, solver_params, named, costs=None):
        """OptSolver::solve with launchProfiledSolve (OptUtils.h:47-64): returns the final cost and
        optionally appends the cost after Init and after every Step to `costs`."""
        self.set_solver_parameters(solver_params)
        self.init(named)
        if costs is not None:
            costs.append(self.current_cost())
        while self.step():
            if costs is not None:
                costs.append(self.current_cost())
        return self.current_cost()

    def solve_blocking(self, solver_params, named):
        """Opt_ProblemSolve in one call."""
        self.set_solver_parameters(solver_params)
        self._params = named
        self._pp = named.data()
        self.f["ProblemSolve"](self.state, self.plan, self._pp)
        return self.current_cost()

    # --- OptB200.h extensions --------------------------------------------------------------------
    def failed(self):
        """1 after a fatal error on this plan was handed to the installed error handler."""
        return self.lib.OptB200_PlanFailed(self.plan)

    def launch_count(self):
        return self.lib.OptB200_PlanLaunchCount(self.plan)

    def materialized_mode(self):
        """0 matrix-free, 1 J and J^T stored, 2 fused J^T J (the mode in effect, not the one asked for)."""
        return self.lib.OptB200_PlanMaterializedMode(self.plan)

    def materialized_nnz(self, which=0):
        return self.lib.OptB200_PlanMaterializedNnz(self.plan, which)

    def pcg_iterations(self):
        """PCG iterations executed since the plan was created (LM may leave the inner loop early)."""
        return self.lib.OptB200_PlanPcgIterations(self.plan)

    def kernel_time(self, name):
        """(launch count, total ms) of the kernels timed under `name` (needs timing=True)."""
        ms = C.c_double(0.0)
        n = self.lib.OptB200_PlanKernelTime(self.plan, name.encode(), C.byref(ms))
        return n, ms.value

    def reset_timing(self):
        self.lib.OptB200_PlanResetTiming(self.plan)

    def unknown_channels(self):
        return self.lib.OptB200_PlanUnknownChannels(self.plan)

    def site_count(self):
        return self.lib.OptB200_PlanSiteCount(self.plan)

    def _dtype(self):
        return np.float64 if self.double else np.float32

    def debug_read(self, name):
        out = np.zeros(self.site_count() * self.unknown_channels(), dtype=self._dtype())
        rc = self.lib.OptB200_DebugReadVector(self.plan, name.encode(), out.ctypes.data)
        if rc != 0:
            raise KeyError(name)
        return out

    def debug_read_range(self, name, first_site, nsites):
        """Sites [first_site, first_site + nsites) of a solver vector as [nsites, U] (full-size parity checks)."""
        out = np.zeros((nsites, self.unknown_channels()), dtype=self._dtype())
        rc = self.lib.OptB200_DebugReadVectorRange(self.plan, name.encode(), first_site, nsites, out.ctypes.data)
        if rc != 0:
            raise KeyError(name)
        return out

    def debug_apply_jtj(self, named, p):
        p = np.ascontiguousarray(p, dtype=self._dtype())
        out = np.zeros_like(p)
        pp = named.data()
        self.lib.OptB200_DebugApplyJTJ(self.plan, pp, p.ctypes.data, out.ctypes.data)
        return out

    def debug_pcg_begin(self, named):
        pp = named.data()
        self.lib.OptB200_DebugPcgBegin(self.plan, pp)

    def close(self):
        if getattr(self, "plan", None):
            self.f["PlanFree"](self.state, self.plan)
            self.plan = None
        if getattr(self, "problem", None):
            self.f["ProblemDelete"](self.state, self.problem)
            self.problem = None
        if getattr(self, "state", None):
            self.lib.OptB200_FreeState(self.state)
            self.state = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def upload(problem, device="cuda"):
    """Problem (host numpy) -> NamedParameters of CUDA tensors / host scalars, positions preserved."""
    import torch
    named = NamedParameters()
    for i, p in enumerate(problem.params):
        if isinstance(p, np.ndarray):
            named.set("p%d" % i, torch.from_numpy(p).to(device).contiguous())
        else:
            named.set("p%d" % i, p)
    return named
