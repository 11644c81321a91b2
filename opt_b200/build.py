"""In-tree build of libOptB200.so (CUDA sm_100a engine + C host) and of the CPU oracle.

`python -m opt_b200.build` or `__graft_entry__.build()`.  nvcc cross-compiles without a GPU.
Objects are rebuilt only when a source or header is newer.
"""
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CSRC = os.path.join(ROOT, "opt_b200", "csrc")
OBJ = os.path.join(ROOT, "build", "obj")
LIB = os.path.join(ROOT, "opt_b200", "libOptB200.so")
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
HOST_CC = "/usr/bin/gcc"
HOST_CXX = "/usr/bin/g++"
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
              "-ccbin", HOST_CXX, "-Xcompiler", "-fPIC", "--expt-relaxed-constexpr"]


def _headers():
    out = []
    for d, _, fs in os.walk(CSRC):
        out += [os.path.join(d, f) for f in fs if f.endswith((".cuh", ".h", ".inc"))]
    out += [os.path.join(ROOT, "include", f) for f in os.listdir(os.path.join(ROOT, "include"))]
    return out


def _stale(target, deps):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps)


def _compile(src, verbose):
    obj = os.path.join(OBJ, os.path.basename(src) + ".o")
    if not _stale(obj, [src] + _headers()):
        return obj, ""
    if src.endswith(".cu"):
        cmd = [NVCC] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else []) + ["-c", src, "-o", obj]
    else:
        cmd = [HOST_CC, "-O2", "-std=c11", "-fPIC", "-Wall", "-c", src, "-o", obj]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("build failed: %s\n%s\n%s" % (" ".join(cmd), r.stdout, r.stderr))
    return obj, r.stderr


def build(verbose=False, oracle=True):
    os.makedirs(OBJ, exist_ok=True)
    srcs = [os.path.join(CSRC, "opt_api.c"), os.path.join(CSRC, "registry.cu"), os.path.join(CSRC, "driver_ops.cu"), os.path.join(CSRC, "csr_api.cu"), os.path.join(CSRC, "formats.c"), os.path.join(CSRC, "energy_check.c")]
    ex = os.path.join(CSRC, "examples")
    srcs += sorted(os.path.join(ex, f) for f in os.listdir(ex) if f.endswith(".cu"))
    with ThreadPoolExecutor(max_workers=min(8, len(srcs))) as pool:
        results = list(pool.map(lambda s: _compile(s, verbose), srcs))
    objs = [o for o, _ in results]
    if verbose:
        for _, log in results:
            if log:
                sys.stderr.write(log)
    if _stale(LIB, objs):
        cmd = [NVCC, "-shared", "-o", LIB] + objs + ["-ccbin", HOST_CXX, "-lcudart", "-lnccl"]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError("link failed: %s\n%s" % (" ".join(cmd), r.stderr))
    # the C drivers that show the library behind the unchanged Opt.h (drivers_c/): plain C, linked like the reference's
    bindir = os.path.join(ROOT, "opt_b200", "bin")
    os.makedirs(bindir, exist_ok=True)
    ddir = os.path.join(ROOT, "drivers_c")
    for f in sorted(os.listdir(ddir)):
        if not f.endswith(".c"):
            continue
        exe = os.path.join(bindir, f[:-2])
        if _stale(exe, [os.path.join(ddir, f), LIB] + _headers()):
            cuda = os.path.dirname(os.path.dirname(NVCC))
            cmd = [HOST_CC, "-O2", "-std=c11", "-Wall", "-I" + os.path.join(cuda, "include"), os.path.join(ddir, f), "-o", exe,
                   "-L" + os.path.dirname(LIB), "-lOptB200", "-L" + os.path.join(cuda, "lib64"), "-lcudart",
                   "-Wl,-rpath,$ORIGIN/..", "-Wl,-rpath," + os.path.join(cuda, "lib64")]
            r = subprocess.run(cmd, capture_output=True, text=True)
            if r.returncode != 0:
                raise RuntimeError("driver build failed: %s\n%s" % (" ".join(cmd), r.stderr))
    if oracle:
        subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "oracle")])
        # test infrastructure that is compiled from the reference's own sources where they lie (container only; the
        # binaries travel to the GPU box): the reference's driver layer / poisson example against this library, and the
        # reference's CPU driver code as the oracle of driver_ops.cu
        if os.path.isdir("/root/reference/examples/shared"):
            for d in ("tests/ref_driver", "oracle/ref_src"):
                if os.path.exists(os.path.join(ROOT, d, "Makefile")):
                    r = subprocess.run(["make", "-s", "-j8", "-C", os.path.join(ROOT, d)], capture_output=True, text=True)
                    if r.returncode != 0:
                        raise RuntimeError("make -C %s failed:\n%s" % (d, r.stderr[-4000:]))
    return LIB


if __name__ == "__main__":
    print(build(verbose="-v" in sys.argv))
