"""CPU: the loaders of include/OptB200Formats.h (plain C, host side) on files written by the test itself, and -- in the
build container, where the reference's data directory is mounted -- on the reference's own data files."""
import ctypes as C
import os
import re
import struct

import numpy as np
import pytest

from opt_b200 import formats as F

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_DATA = "/root/reference/examples/data"


def test_format_symbols_exported(built):
    txt = re.sub(r"/\*.*?\*/", "", open(os.path.join(ROOT, "include", "OptB200Formats.h")).read(), flags=re.S)
    names = sorted(set(re.findall(r"\b(OptB200_[A-Za-z]+)\s*\(", txt)))
    assert len(names) == 8
    lib = C.CDLL(built)
    for n in names:
        assert hasattr(lib, n), n


def test_imagedump_roundtrip_and_infinity_clamp(built, tmp_path):
    a = np.arange(12, dtype=np.float32).reshape(3, 4)
    a[1, 2], a[2, 0] = np.inf, -np.inf
    p = str(tmp_path / "a.imagedump")
    F.write_imagedump(p, a)
    raw = open(p, "rb").read()
    assert struct.unpack("<4i", raw[:16]) == (4, 3, 1, 0) and len(raw) == 16 + 48      # im.t:9-15
    assert np.array_equal(F.read_imagedump(p, clamp_infinity=False), a)
    b = F.read_imagedump(p)                                                           # SimpleBuffer.cpp:34-46
    assert b[1, 2] == np.finfo(np.float32).max and b[2, 0] == np.float32(-10000.0)
    m = (np.arange(24) % 2).astype(np.uint8).reshape(6, 4)
    F.write_imagedump(p, m)
    assert np.array_equal(F.read_imagedump(p), m)
    with pytest.raises(IOError):
        F.read_imagedump(str(tmp_path / "missing.imagedump"))
    open(p, "wb").write(raw[:20])
    with pytest.raises(IOError):
        F.read_imagedump(p)


def test_text_tables(built, tmp_path):
    p = str(tmp_path / "c.constraints")
    open(p, "w").write("2\n30 132 59 44\n229 51\t157 91\n")
    assert F.read_constraints(p).tolist() == [[30, 132, 59, 44], [229, 51, 157, 91]]
    open(p, "w").write("3\n1 2 3 4\n")
    with pytest.raises(IOError):
        F.read_constraints(p)
    q = str(tmp_path / "m.mrk")
    open(q, "w").write("2 \n15.5 -52.75 12.25 0.02 68\n0.0 50.0 -50.0 0.02 7\n")
    pos, idx = F.read_markers(q)
    assert pos.tolist() == [[15.5, -52.75, 12.25], [0.0, 50.0, -50.0]] and idx.tolist() == [68, 7]


def test_sfs_parameters(built, tmp_path):
    vals = [float(i) + 0.5 for i in range(36)]
    p = str(tmp_path / "x.SFSSolverParameters")
    open(p, "wb").write(struct.pack("<36f3I", *vals, 1, 10, 8) + b"\0\0\0\0")
    s = F.read_sfs_parameters(p)
    assert (s.weightFitting, s.weightShading, s.fx, s.uy) == (0.5, 3.5, 7.5, 10.5)
    assert list(s.lightingCoefficients) == vals[27:36] and list(s.unused) == [1, 10, 8]


def test_meshes(built, tmp_path):
    v = np.array([[0, 0, 0], [1, 0, 0.5], [0, 1, 0.25], [1, 1, -2]], dtype=np.float32)
    f = np.array([[0, 1, 2], [2, 1, 3]], dtype=np.int32)
    p = str(tmp_path / "m.off")
    F.write_off(p, v, f)
    v2, f2 = F.read_mesh(p)
    assert np.array_equal(v, v2) and np.array_equal(f, f2)
    # ascii PLY with an extra vertex property in front of x y z and a colour after the face list
    q = str(tmp_path / "a.ply")
    with open(q, "w") as o:
        o.write("ply\nformat ascii 1.0\ncomment test\nelement vertex 4\nproperty float quality\nproperty float x\nproperty float y\nproperty float z\n"
                "element face 2\nproperty list uchar int vertex_indices\nend_header\n")
        for r in v:
            o.write("9 %g %g %g\n" % tuple(r))
        for r in f:
            o.write("3 %d %d %d\n" % tuple(r))
    v2, f2 = F.read_mesh(q)
    assert np.array_equal(v, v2) and np.array_equal(f, f2)
    # binary little-endian PLY: double positions, normals in between, uint face indices
    b = str(tmp_path / "b.ply")
    with open(b, "wb") as o:
        o.write(b"ply\nformat binary_little_endian 1.0\nelement vertex 4\nproperty double x\nproperty double y\nproperty float nx\nproperty double z\n"
                b"element face 2\nproperty list uchar uint vertex_indices\nend_header\n")
        for r in v:
            o.write(struct.pack("<ddfd", r[0], r[1], 7.0, r[2]))
        for r in f:
            o.write(struct.pack("<B3I", 3, *r))
    v2, f2 = F.read_mesh(b)
    assert np.array_equal(v, v2) and np.array_equal(f, f2)
    open(b, "wb").write(b"ply\nformat binary_big_endian 1.0\nend_header\n")
    with pytest.raises(IOError):
        F.read_mesh(b)


@pytest.mark.skipif(not os.path.isdir(REF_DATA), reason="reference data directory not mounted")
def test_reference_data_files(built, golden_dir):
    """The fixtures under tests/golden were derived from these files; the C loaders must see the same content."""
    g = np.load(os.path.join(golden_dir, "sfs_default.npz"))
    D = F.read_imagedump(REF_DATA + "/shape_from_shading/default_targetDepth.imagedump")
    assert np.array_equal(D, g["D"]) and D.shape == (480, 640)
    M = F.read_imagedump(REF_DATA + "/shape_from_shading/default_maskEdgeMap.imagedump")
    assert M.dtype == np.uint8 and np.array_equal(np.packbits(M), g["masks"])
    p = F.read_sfs_parameters(REF_DATA + "/shape_from_shading/default.SFSSolverParameters")
    assert np.array_equal(np.array([p.weightFitting, p.weightRegularizer, p.weightShading, p.fx, p.fy, p.ux, p.uy] + list(p.lightingCoefficients), dtype=np.float32), g["scalars"])
    c = np.load(os.path.join(golden_dir, "image_warping_cat512.npz"))["constraints"]
    assert np.array_equal(F.read_constraints(REF_DATA + "/cat512.constraints"), c)
    v, f = F.read_mesh(REF_DATA + "/head.ply")
    h = np.load(os.path.join(golden_dir, "cotangent_head.npz"))
    assert np.array_equal(v, h["vertices"]) and np.array_equal(f, h["faces"])
    v, f = F.read_mesh(REF_DATA + "/raptor_simplify2k.off")
    r = np.load(os.path.join(golden_dir, "embedded_raptor2k.npz"))
    assert np.allclose(v, r["vertices"]) and np.array_equal(f, r["faces"])


def test_imagedump_header_that_overflows_is_rejected(tmp_path):
    """width * height * channels from three unchecked int32 fields must not wrap size_t / outgrow the file."""
    import struct
    from opt_b200 import formats as F
    p = tmp_path / "bad.imagedump"
    p.write_bytes(struct.pack("<4i", 0x7fffffff, 0x7fffffff, 8, 0) + b"\0" * 64)
    with pytest.raises(OSError):
        F.read_imagedump(str(p))
    p.write_bytes(struct.pack("<4i", 4, 4, 1, 0) + b"\0" * 60)            # one float short of 4 x 4
    with pytest.raises(OSError):
        F.read_imagedump(str(p))


def test_off_comment_at_end_of_file_without_newline(tmp_path):
    """the comment stripper must stop at the terminator."""
    from opt_b200 import formats as F
    p = tmp_path / "m.off"
    p.write_text("OFF\n3 1 0\n0 0 0\n1 0 0\n0 1 0\n3 0 1 2\n# trailing comment without newline")
    v, f = F.read_mesh(str(p))
    assert v.shape == (3, 3) and f.tolist() == [[0, 1, 2]]
