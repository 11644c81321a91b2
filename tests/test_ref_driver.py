"""GPU: the reference's OWN C++ driver layer running on libOptB200.so ("next" items 3 / 8).

tests/ref_driver/Makefile compiles, from /root/reference where the sources lie and with include/Opt.h first on the
include path,
  * bin/ref_harness: a caller over the unmodified examples/shared/{OptSolver,OptUtils,NamedParameters,OptImage,OptGraph}.h
    (OptSolver::solve -> launchProfiledSolve -> Opt_ProblemInit / Opt_ProblemStep, the graph flattened to
    `edgeCount, v0, v1` by NamedParameters::data()), fed with problem dumps written here, and
  * bin/poisson_image_editing: the reference's complete, unmodified example driver (main.cpp, CombinedSolver.h, mLib).
The build happens in the container (`__graft_entry__.build()` when /root/reference exists); the binaries travel to the GPU
box with the snapshot.  `final cost=` is parsed from stdout exactly as examples/test_final_cost.py:100-103 does."""
import os
import re
import struct
import subprocess
import zlib

import numpy as np
import pytest

from opt_b200 import examples

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "tests", "ref_driver", "bin")


def _need(name):
    path = os.path.join(BIN, name)
    if not os.path.exists(path):
        pytest.skip("%s not built (tests/ref_driver/Makefile needs /root/reference; build() makes it in the container)" % name)
    return path


def final_cost(output):                      # examples/test_final_cost.py:100-103
    return float(re.search("final cost=(.*)", output).group(1))


def _s(b):
    b = b.encode()
    return struct.pack("<I", len(b)) + b


def dump_problem(path, pr, names, graph=None):
    """Problem -> the dump ref_harness reads.  `names`: parameter names in problemparams order; `graph` = (name, index of
    the edge-count entry, arity): that entry and the following `arity` index arrays become one OptGraph parameter."""
    out = [b"OPTB2REF", _s(pr.energy), _s(pr.kind), struct.pack("<I", len(pr.dims))]
    out += [struct.pack("<I", int(d)) for d in pr.dims]
    out.append(struct.pack("<II", int(pr.solver["nIterations"]), int(pr.solver["lIterations"])))
    items = []
    i = 0
    image_dims = [int(d) for d in pr.dims]
    while i < len(pr.params):
        p = pr.params[i]
        if graph and i == graph[1]:
            arity = graph[2]
            idx = [np.ascontiguousarray(pr.params[i + 1 + k], dtype=np.int32) for k in range(arity)]
            items.append(struct.pack("<I", 3) + _s(graph[0]) + struct.pack("<II", arity, len(idx[0])) + b"".join(a.tobytes() for a in idx))
            i += 1 + arity
            continue
        name = names[len(items)]
        if isinstance(p, np.ndarray):
            a = np.ascontiguousarray(p, dtype=np.float32)
            if graph:                                   # graph energies: 1-D vertex images
                dims, ch = [a.shape[0]], int(a.size // a.shape[0])
            else:
                dims = image_dims
                ch = int(a.size // int(np.prod(dims)))
            items.append(struct.pack("<I", 1) + _s(name) + struct.pack("<III", int(i in pr.unknowns), ch, len(dims))
                         + b"".join(struct.pack("<I", d) for d in dims) + a.tobytes())
        else:
            items.append(struct.pack("<I", 2) + _s(name) + struct.pack("<f", float(p)))
        i += 1
    out.append(struct.pack("<I", len(items)))
    with open(path, "wb") as f:
        f.write(b"".join(out + items))


def run_harness(tmp_path, pr, names, graph=None, extra=()):
    exe = _need("ref_harness")
    path = str(tmp_path / "problem.bin")
    dump_problem(path, pr, names, graph)
    r = subprocess.run([exe, path] + list(extra), capture_output=True, text=True, timeout=600, cwd=str(tmp_path))
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    return r.stdout


def test_poisson_golden_through_reference_optsolver(built, golden_dir, tmp_path):
    """poisson_image_editing at stride 4, (1,1,1): 1530364.25 (examples/test_final_cost.py:61), image parameters."""
    npz = np.load(os.path.join(golden_dir, "poisson_stride4.npz"))
    pr = examples.poisson_golden(npz)
    out = run_harness(tmp_path, pr, ["X", "T", "M"])
    assert abs(final_cost(out) - 1530364.25) <= 1e-5 * 1530364.25, out[-2000:]
    assert "TIMING" in out and "Per-iter times ms (nonlinear,linear):" in out      # backend_cuda.t:262-297, read by outParse.py


def test_arap_golden_through_reference_optsolver(built, golden_dir, tmp_path):
    """arap_mesh_deformation, (1,1,1): 7183.464843 (test_final_cost.py:56): the graph goes through OptGraph and
    NamedParameters::data()'s `edgeCount, v0, v1` flattening (examples/shared/NamedParameters.h:35-48)."""
    npz = np.load(os.path.join(golden_dir, "arap_small_armadillo.npz"))
    pr = examples.arap_golden(npz)
    out = run_harness(tmp_path, pr, ["w_fitSqrt", "w_regSqrt", "Offset", "Angle", "UrShape", "Constraints", "G"], graph=("G", 6, 2))
    assert abs(final_cost(out) - 7183.464843) <= 1e-5 * 7183.464843, out[-2000:]


def _png(path, rgba):
    """Minimal 8-bit RGBA PNG writer (the example driver loads its inputs with LodePNG)."""
    h, w, _ = rgba.shape
    raw = b"".join(b"\x00" + rgba[y].tobytes() for y in range(h))

    def chunk(tag, data):
        c = struct.pack(">I", len(data)) + tag + data
        return c + struct.pack(">I", zlib.crc32(tag + data) & 0xffffffff)
    with open(path, "wb") as f:
        f.write(b"\x89PNG\r\n\x1a\n" + chunk(b"IHDR", struct.pack(">IIBBBBB", w, h, 8, 6, 0, 0, 0))
                + chunk(b"IDAT", zlib.compress(raw, 6)) + chunk(b"IEND", b""))


def test_unmodified_reference_poisson_example(built, golden_dir, tmp_path):
    """The reference's complete poisson_image_editing driver (its main.cpp, CombinedSolver.h, ArgParser, mLib image
    loading, OptSolver) linked against libOptB200.so, run the way examples/test_example_output.py:118,139 runs it
    (stride 1, GN 1 x 100 on the default 448 x 320 images): final cost 139582.609375 as recorded there."""
    exe = _need("poisson_image_editing")
    npz = np.load(os.path.join(golden_dir, "poisson_full.npz"))
    H, W = (int(t) for t in npz["shape"])
    data = tmp_path / "data"
    run = tmp_path / "poisson_image_editing"
    data.mkdir()
    run.mkdir()
    alpha = np.full((H, W, 1), 255, np.uint8)
    _png(str(data / "poisson0.png"), np.concatenate([npz["image"], alpha], -1))
    _png(str(data / "poisson1.png"), np.concatenate([npz["target"], alpha], -1))
    unknown = np.unpackbits(npz["unknown"])[:H * W].reshape(H, W)
    mask = np.repeat((unknown * 255).astype(np.uint8)[..., None], 4, -1)        # the driver: mask pixel 255 -> unknown
    mask[..., 3] = 255
    _png(str(data / "poisson_mask.png"), mask)
    (run / "args.config").write_text("useOpt=true\nuseOptLM=false\nuseCeres=false\n")
    r = subprocess.run([exe, "--backend", "backend_cuda", "--nIterations", "1", "--lIterations", "100", "--stride", "1", "--noOutput", "true"],
                       capture_output=True, text=True, timeout=600, cwd=str(run))
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    golden = float(npz["golden_cost_1x100"])
    got = final_cost(r.stdout)
    assert abs(got - golden) <= 2e-6 * golden, (got, golden)


@pytest.mark.parametrize("flags", [["--materialized"], ["--materialized", "--fused"]])
def test_materialized_flags_through_reference_optsolver(built, golden_dir, tmp_path, flags):
    """OptSolver forwards CombinedSolverParameters.useMaterializedJTJ / useFusedJTJ into Opt_InitializationParameters
    (examples/shared/OptSolver.h:58-59); the recorded costs hold in those modes (examples/test_final_cost.py:91-93)."""
    npz = np.load(os.path.join(golden_dir, "arap_small_armadillo.npz"))
    out = run_harness(tmp_path, examples.arap_golden(npz), ["w_fitSqrt", "w_regSqrt", "Offset", "Angle", "UrShape", "Constraints", "G"],
                      graph=("G", 6, 2), extra=flags)
    assert abs(final_cost(out) - 7183.464843) <= 1e-5 * 7183.464843, out[-2000:]
    # the kernel table of the library (collectPerKernelTimingInfo = 1 in OptSolver.h) names the sparse products it ran
    assert ("J^TJp" in out) == ("--fused" in flags) and (" J^T " in out) != ("--fused" in flags)
