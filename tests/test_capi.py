"""CPU: the C-ABI library loads and exports every symbol include/*.h declares (no compute calls)."""
import ctypes as C
import os
import re

from opt_b200 import capi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols(header):
    txt = open(os.path.join(ROOT, "include", header)).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(Opt(?:B200)?_[A-Za-z]+)\s*\(", txt)))


def test_library_exports_every_declared_symbol(built):
    lib = C.CDLL(built)
    names = []
    for header in sorted(os.listdir(os.path.join(ROOT, "include"))):          # every include/*.h, not only the two solver headers
        names += declared_symbols(header)
    assert len(names) >= 10 + 6 and "OptB200_FilterGaussianNoAlloc" in names and "OptB200_CsrSpmv" in names
    for n in names:
        assert hasattr(lib, n), n


def test_struct_layout_matches_reference():
    assert C.sizeof(capi.Opt_InitializationParameters) == 44
    assert capi.Opt_InitializationParameters.backend.offset == 12
    assert capi.Opt_InitializationParameters.numthreads.offset == 32


def test_cpu_backends_and_unknown_energies_are_rejected(built, capfd):
    lib = C.CDLL(built)
    f = capi.bind(lib)
    ip = capi.Opt_InitializationParameters()
    ip.backend = b"backend_cpu"
    assert not f["NewState"](ip)                      # no CPU fallback in the product library
    ip.backend = b"backend_cuda"
    st = f["NewState"](ip)
    assert st
    assert not f["ProblemDefine"](st, b"my_own_energy.t", b"gaussNewtonGPU")
    assert "no built-in functor" in capfd.readouterr().err
    pr = f["ProblemDefine"](st, b"../somewhere/image_warping.t", b"LMGPU")
    assert pr
    f["ProblemDelete"](st, pr)


def test_product_does_not_reference_the_oracle():
    bad = []
    for d, _, fs in os.walk(os.path.join(ROOT, "opt_b200")):
        for fn in fs:
            if fn.endswith((".py", ".c", ".cu", ".cuh", ".h")):
                txt = open(os.path.join(d, fn)).read()
                if re.search(r"oracle_api|liboptoracle|OptOracle_|#include\s+\"[^\"]*oracle", txt):
                    bad.append(fn)
    assert bad == ["build.py"] or bad == [], bad      # build.py only *builds* the checker
