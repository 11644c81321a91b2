"""ctypes mirror of include/Opt.h (reference: API/release/include/Opt.h:10-76).

Only declarations live here: the struct layout and the argument/return types of the twelve
entry points.  `bind(lib, prefix)` attaches them to a loaded shared library; the product library
exports the bare `Opt_*` names; test infrastructure may bind a differently prefixed library.
"""
import ctypes as C


class Opt_InitializationParameters(C.Structure):
    # reference: Opt.h:10-35 / createwrapper.t:98-134 -- 44 bytes
    _fields_ = [
        ("doublePrecision", C.c_int),
        ("verbosityLevel", C.c_int),
        ("collectPerKernelTimingInfo", C.c_int),
        ("backend", C.c_char * 20),
        ("numthreads", C.c_int),
        ("useMaterializedJTJ", C.c_int),
        ("useFusedJTJ", C.c_int),
    ]


assert C.sizeof(Opt_InitializationParameters) == 44

_P = C.c_void_p
SIGNATURES = {
    "NewState": (_P, [Opt_InitializationParameters]),
    "ProblemDefine": (_P, [_P, C.c_char_p, C.c_char_p]),
    "ProblemDelete": (None, [_P, _P]),
    "ProblemPlan": (_P, [_P, _P, C.POINTER(C.c_uint)]),
    "PlanFree": (None, [_P, _P]),
    "SetSolverParameter": (None, [_P, _P, C.c_char_p, _P]),
    "ProblemSolve": (None, [_P, _P, C.POINTER(_P)]),
    "ProblemInit": (None, [_P, _P, C.POINTER(_P)]),
    "ProblemStep": (C.c_int, [_P, _P, C.POINTER(_P)]),
    "ProblemCurrentCost": (C.c_double, [_P, _P]),
}


def bind(lib, prefix="Opt_"):
    """Return {short name: ctypes function} for every Opt.h entry point of `lib`."""
    out = {}
    for short, (res, args) in SIGNATURES.items():
        fn = getattr(lib, prefix + short)   # AttributeError if the library does not export it
        fn.restype = res
        fn.argtypes = args
        out[short] = fn
    return out
