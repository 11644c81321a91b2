"""CPU: the restricted energy-file front end (opt_b200/csrc/energy_check.c).  Opt_ProblemDefine must accept the shipped
energy files, reject files that declare another binding or whose expressions were edited, and fall back to the name
when the path cannot be read.  The energy texts written here are the test's own (poisson-like), not the reference's."""
import ctypes as C
import os

import pytest

from opt_b200 import capi

REF = "/root/reference/examples"
ENERGIES = ["poisson_image_editing", "image_warping", "volumetric_mesh_deformation", "optical_flow", "arap_mesh_deformation",
            "shape_from_shading", "embedded_mesh_deformation", "cotangent_mesh_smoothing", "robust_nonrigid_alignment",
            "intrinsic_image_decomposition"]

MINE = """-- my own two-image energy
local W,H = Dim("W",0), Dim("H",1)
local X = Unknown("X", opt_float4,{W,H},0) --[[ block
comment ]] local T = Array("T", opt_float4,{W,H},1)
local M = Array( "M", opt_float, {W,H}, 2)
local w = Param("w", float, 3)
local G = Graph("G", {E}, "v0", {N}, 7,
                         "v1", {N}, 8)
UsePreconditioner(false)
Exclude(Not(eq(M(0,0),0)))
Energy(w*(X(0,0) - T(0,0)))
"""


def lib_and_state(built):
    lib = C.CDLL(built)
    lib.OptB200_EnergyDescribe.argtypes = [C.c_char_p, C.c_char_p, C.c_ulong]
    lib.OptB200_EnergyExpected.restype = C.c_char_p
    lib.OptB200_EnergyExpected.argtypes = [C.c_char_p]
    f = capi.bind(lib)
    ip = capi.Opt_InitializationParameters()
    ip.backend = b"backend_cuda"
    return lib, f, f["NewState"](ip)


def describe(lib, path):
    buf = C.create_string_buffer(4096)
    rc = lib.OptB200_EnergyDescribe(str(path).encode(), buf, 4096)
    return rc, buf.value.decode()


def test_describe_parses_declarations(built, tmp_path):
    lib, _, _ = lib_and_state(built)
    p = tmp_path / "mine.t"
    p.write_text(MINE)
    rc, d = describe(lib, p)
    assert rc == 0
    lines = d.splitlines()
    assert lines[:2] == ["dim W 0", "dim H 1"]
    assert "unknown X opt_float4 {W,H} 0" in lines and "array T opt_float4 {W,H} 1" in lines and "array M opt_float {W,H} 2" in lines
    assert "param w float - 3" in lines and "graph G {E}" in lines and "  vertex v0 {N} 7" in lines and "  vertex v1 {N} 8" in lines
    assert "preconditioner false" in lines and "exclude yes" in lines and lines[-1].startswith("fingerprint ")
    # comments and white space do not change the description; an edited expression changes only the fingerprint
    p.write_text(MINE.replace("-- my own two-image energy\n", "").replace("local w = ", "local   w   =   "))
    assert describe(lib, p)[1] == d
    p.write_text(MINE.replace("X(0,0) - T(0,0)", "X(0,0) - 2*T(0,0)"))
    d2 = describe(lib, p)[1].splitlines()
    assert d2[:-1] == lines[:-1] and d2[-1] != lines[-1]
    assert describe(lib, tmp_path / "absent.t")[0] == 1


def test_define_rejects_a_file_that_is_not_the_shipped_energy(built, tmp_path, capfd, monkeypatch):
    lib, f, st = lib_and_state(built)
    p = tmp_path / "poisson_image_editing.t"
    p.write_text(MINE)                                             # right name, other declarations
    assert not f["ProblemDefine"](st, str(p).encode(), b"gaussNewtonGPU")
    assert "declarations differ" in capfd.readouterr().err
    assert f["ProblemDefine"](st, str(tmp_path / "nowhere" / "poisson_image_editing.t").encode(), b"gaussNewtonGPU")   # unreadable: by name
    assert lib.OptB200_EnergyExpected(b"poisson_image_editing").decode().startswith("dim W 0\n")
    assert lib.OptB200_EnergyExpected(b"my_energy") is None


@pytest.mark.skipif(not os.path.isdir(REF), reason="reference not mounted")
@pytest.mark.parametrize("energy", ENERGIES)
def test_shipped_energy_files_verify(built, energy, tmp_path, capfd, monkeypatch):
    lib, f, st = lib_and_state(built)
    path = "%s/%s/%s.t" % (REF, energy, energy)
    rc, d = describe(lib, path)
    assert rc == 0 and d == lib.OptB200_EnergyExpected(energy.encode()).decode()
    pr = f["ProblemDefine"](st, path.encode(), b"LMGPU")
    assert pr
    f["ProblemDelete"](st, pr)
    # the same file with one coefficient edited: declarations match, expressions do not -> rejected unless overridden
    src = open(path, encoding="utf-8", errors="replace").read()
    edited = src.replace("Energy(", "Energy(1.5*", 1)
    assert edited != src
    q = tmp_path / (energy + ".t")
    q.write_text(edited)
    assert not f["ProblemDefine"](st, str(q).encode(), b"gaussNewtonGPU")
    assert "expressions were edited" in capfd.readouterr().err
    monkeypatch.setenv("OPTB200_ALLOW_MODIFIED_ENERGY", "1")
    pr = f["ProblemDefine"](st, str(q).encode(), b"gaussNewtonGPU")
    assert pr and "warning" in capfd.readouterr().err
    f["ProblemDelete"](st, pr)
