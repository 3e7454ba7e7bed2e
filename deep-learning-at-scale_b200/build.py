"""Build libdlas_b200.so in-tree with nvcc for sm_100a (no other architecture, no JIT).

    python deep-learning-at-scale_b200/build.py [--force] [--verbose] [--trace]

--trace builds a second library, lib/libdlas_b200_trace.so, whose tcgen05 kernel carries the SM-clock phase stamps
read by tools/tc_trace.py (-DDLAS_TC_TRACE); the product library has none of that code.

Objects go to build/, the library to lib/libdlas_b200.so (git-ignored; it travels to the GPU box
with the gpurun snapshot).  -Xptxas -v output is kept in build/ptxas_<file>.log."""
import os, subprocess, sys, shutil, hashlib
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
BUILD = os.path.join(HERE, "build")
LIBDIR = os.path.join(HERE, "lib")
LIB = os.path.join(LIBDIR, "libdlas_b200.so")
NVCC = os.environ.get("NVCC") or shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
         "-Xcompiler", "-fPIC", "-Xptxas", "-v", "--expt-relaxed-constexpr"]


def sources():
    return sorted(f for f in os.listdir(CSRC) if f.endswith(".cu"))


def _stamp(path):
    h = hashlib.sha1()
    for f in sorted(os.listdir(CSRC)):
        if f.endswith((".cuh", ".h")) or f == os.path.basename(path):
            h.update(open(os.path.join(CSRC, f), "rb").read())
    h.update(open(os.path.join(HERE, "..", "include", "dlas_b200.h"), "rb").read())
    h.update(" ".join(FLAGS).encode())
    return h.hexdigest()


def compile_one(src, force=False, verbose=False, trace=False):
    path = os.path.join(CSRC, src)
    tag = "_trace" if trace else ""
    obj = os.path.join(BUILD, src[:-3] + tag + ".o")
    stamp_file = obj + ".stamp"
    stamp = _stamp(path)
    if not force and os.path.exists(obj) and os.path.exists(stamp_file) and open(stamp_file).read() == stamp:
        return obj, False
    cmd = [NVCC] + FLAGS + (["-DDLAS_TC_TRACE"] if trace else []) + ["-c", path, "-o", obj]
    r = subprocess.run(cmd, capture_output=True, text=True)
    open(os.path.join(BUILD, "ptxas_" + src[:-3] + tag + ".log"), "w").write(r.stdout + r.stderr)
    if r.returncode != 0:
        raise RuntimeError("nvcc failed for %s:\n%s" % (src, r.stdout + r.stderr))
    if verbose:
        print(r.stderr)
    open(stamp_file, "w").write(stamp)
    return obj, True


def build(force=False, verbose=False, trace=False):
    os.makedirs(BUILD, exist_ok=True)
    os.makedirs(LIBDIR, exist_ok=True)
    lib = LIB[:-3] + "_trace.so" if trace else LIB
    with ThreadPoolExecutor(max_workers=8) as ex:
        res = list(ex.map(lambda s: compile_one(s, force, verbose, trace and s == "mlp_tc.cu"), sources()))
    objs = [o for o, _ in res]
    if any(ch for _, ch in res) or not os.path.exists(lib):
        cmd = [NVCC, "-shared", "-o", lib] + objs + ["-gencode", "arch=compute_100a,code=sm_100a", "-lcudart"]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError("link failed:\n" + r.stdout + r.stderr)
    return lib


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="--verbose" in sys.argv, trace="--trace" in sys.argv))
