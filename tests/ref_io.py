"""TEST INFRASTRUCTURE: array files exchanged with the oracle/_ref harnesses (oracle/ref_src/io.h) and a runner."""
import os
import struct
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_BIN = os.path.join(ROOT, "oracle", "_ref")


def write_arrays(path, arrays):
    with open(path, "wb") as f:
        f.write(struct.pack("<i", len(arrays)))
        for a in arrays:
            a = np.ascontiguousarray(a)
            f.write(struct.pack("<i", a.ndim) + struct.pack("<%di" % a.ndim, *a.shape) + struct.pack("<i", a.itemsize) + a.tobytes())


def read_arrays(path, dtypes):
    out = []
    with open(path, "rb") as f:
        (n,) = struct.unpack("<i", f.read(4))
        for k in range(n):
            (nd,) = struct.unpack("<i", f.read(4))
            dims = struct.unpack("<%di" % nd, f.read(4 * nd))
            (elem,) = struct.unpack("<i", f.read(4))
            cnt = int(np.prod(dims))
            out.append(np.frombuffer(f.read(cnt * elem), dtype=dtypes[k]).reshape(dims).copy())
    return out


def have(name):
    return os.path.exists(os.path.join(REF_BIN, name))


def run_ref(name, args, arrays, dtypes, tmp_path):
    """Runs oracle/_ref/<name> [args] in.bin out.bin and returns the output arrays."""
    src, dst = str(tmp_path / "in.bin"), str(tmp_path / "out.bin")
    write_arrays(src, arrays)
    r = subprocess.run([os.path.join(REF_BIN, name)] + list(args) + [src, dst], capture_output=True, text=True, timeout=600, cwd=str(tmp_path))
    assert r.returncode == 0, (r.stdout[-2000:], r.stderr[-2000:])
    return read_arrays(dst, dtypes)
