"""ctypes mirror of include/OptB200Formats.h: the reference drivers' on-disk formats as numpy arrays (host side)."""
import ctypes as C

import numpy as np

from . import solver as S


class _ImageDump(C.Structure):
    _fields_ = [("width", C.c_int), ("height", C.c_int), ("channels", C.c_int), ("datatype", C.c_int), ("data", C.c_void_p)]


class SFSParameters(C.Structure):
    _fields_ = [(n, C.c_float) for n in ("weightFitting", "weightRegularizer", "weightPrior", "weightShading", "weightShadingStart",
                                         "weightShadingIncrement", "weightBoundary", "fx", "fy", "ux", "uy")] + \
               [("deltaTransform", C.c_float * 16), ("lightingCoefficients", C.c_float * 9), ("unused", C.c_uint * 3)]


assert C.sizeof(SFSParameters) == 156
_bound = False


def _lib():
    global _bound
    lib, _ = S.load_library()
    if not _bound:
        P = C.c_void_p
        lib.OptB200_ImageDumpRead.argtypes = [C.c_char_p, C.c_int, C.POINTER(_ImageDump)]
        lib.OptB200_ImageDumpWrite.argtypes = [C.c_char_p, C.POINTER(_ImageDump)]
        lib.OptB200_SFSParametersRead.argtypes = [C.c_char_p, C.POINTER(SFSParameters)]
        lib.OptB200_ConstraintsRead.argtypes = [C.c_char_p, C.POINTER(C.c_int), C.POINTER(C.POINTER(C.c_int))]
        lib.OptB200_MarkersRead.argtypes = [C.c_char_p, C.POINTER(C.c_int), C.POINTER(C.POINTER(C.c_float)), C.POINTER(C.POINTER(C.c_int))]
        lib.OptB200_MeshRead.argtypes = [C.c_char_p, C.POINTER(C.c_int), C.POINTER(C.POINTER(C.c_float)), C.POINTER(C.c_int), C.POINTER(C.POINTER(C.c_int))]
        lib.OptB200_MeshWriteOFF.argtypes = [C.c_char_p, C.c_int, P, C.c_int, P]
        lib.OptB200_FormatFree.argtypes = [P]
        _bound = True
    return lib


def _ok(rc, what, path):
    if rc != 0:
        raise IOError("%s(%s) failed: %d" % (what, path, rc))


def _take(ptr, n, dtype):
    a = np.ctypeslib.as_array(ptr, shape=(n,)).astype(dtype, copy=True) if n else np.zeros(0, dtype=dtype)
    _lib().OptB200_FormatFree(C.cast(ptr, C.c_void_p))
    return a


def read_imagedump(path, clamp_infinity=True):
    d = _ImageDump()
    _ok(_lib().OptB200_ImageDumpRead(path.encode(), int(clamp_infinity), C.byref(d)), "OptB200_ImageDumpRead", path)
    n = d.width * d.height * d.channels
    ct = C.c_float if d.datatype == 0 else C.c_ubyte
    a = _take(C.cast(d.data, C.POINTER(ct)), n, np.float32 if d.datatype == 0 else np.uint8)
    return a.reshape((d.height, d.width) if d.channels == 1 else (d.height, d.width, d.channels))


def write_imagedump(path, a):
    a = np.ascontiguousarray(a)
    assert a.dtype in (np.float32, np.uint8)
    d = _ImageDump(a.shape[1], a.shape[0], 1 if a.ndim == 2 else a.shape[2], 0 if a.dtype == np.float32 else 1, a.ctypes.data)
    _ok(_lib().OptB200_ImageDumpWrite(path.encode(), C.byref(d)), "OptB200_ImageDumpWrite", path)


def read_sfs_parameters(path):
    p = SFSParameters()
    _ok(_lib().OptB200_SFSParametersRead(path.encode(), C.byref(p)), "OptB200_SFSParametersRead", path)
    return p


def read_constraints(path):
    n, rows = C.c_int(), C.POINTER(C.c_int)()
    _ok(_lib().OptB200_ConstraintsRead(path.encode(), C.byref(n), C.byref(rows)), "OptB200_ConstraintsRead", path)
    return _take(rows, 4 * n.value, np.int32).reshape(-1, 4)


def read_markers(path):
    n, pos, idx = C.c_int(), C.POINTER(C.c_float)(), C.POINTER(C.c_int)()
    _ok(_lib().OptB200_MarkersRead(path.encode(), C.byref(n), C.byref(pos), C.byref(idx)), "OptB200_MarkersRead", path)
    return _take(pos, 3 * n.value, np.float32).reshape(-1, 3), _take(idx, n.value, np.int32)


def read_mesh(path):
    nv, nf, v, f = C.c_int(), C.c_int(), C.POINTER(C.c_float)(), C.POINTER(C.c_int)()
    _ok(_lib().OptB200_MeshRead(path.encode(), C.byref(nv), C.byref(v), C.byref(nf), C.byref(f)), "OptB200_MeshRead", path)
    return _take(v, 3 * nv.value, np.float32).reshape(-1, 3), _take(f, 3 * nf.value, np.int32).reshape(-1, 3)


def write_off(path, vertices, faces):
    v = np.ascontiguousarray(vertices, dtype=np.float32)
    f = np.ascontiguousarray(faces, dtype=np.int32)
    _ok(_lib().OptB200_MeshWriteOFF(path.encode(), len(v), v.ctypes.data, len(f), f.ctypes.data), "OptB200_MeshWriteOFF", path)
