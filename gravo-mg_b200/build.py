"""In-tree build of the CUDA library for sm_100a (no JIT cache: the .so travels with the repo).

    python -m gravomg_b200.build            # or: from gravomg_b200.build import build; build()

Output: gravo-mg_b200/lib/libgravomg_b200.so (git-ignored, NOT gpurun-ignored).
Flags: -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -fmad=false. FMA contraction is off for
the whole library because every floating-point result on this path either feeds a comparison that
decides an integer output or must match the reference to 1e-12.
"""
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIBDIR = os.path.join(HERE, "lib")
OBJDIR = os.path.join(HERE, "build")
LIB = os.path.join(LIBDIR, "libgravomg_b200.so")
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-fmad=false", "-std=c++17",
         "-Xcompiler", "-fPIC", "-Xcompiler", "-Wall", "--expt-extended-lambda", "-Xptxas", "-v"]


def _sources():
    return sorted(f for f in os.listdir(CSRC) if f.endswith(".cu"))


def _headers_mtime():
    hs = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".h", ".cuh"))]
    hs.append(os.path.join(os.path.dirname(HERE), "include", "gravomg_b200.h"))
    return max(os.path.getmtime(h) for h in hs)


def _compile(src, verbose):
    obj = os.path.join(OBJDIR, src[:-3] + ".o")
    spath = os.path.join(CSRC, src)
    if os.path.exists(obj) and os.path.getmtime(obj) > max(os.path.getmtime(spath), _headers_mtime()):
        return obj, ""
    r = subprocess.run([NVCC, *FLAGS, "-c", spath, "-o", obj], capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"nvcc failed on {src}:\n{r.stdout}\n{r.stderr}")
    with open(obj + ".ptxas.log", "w") as f:
        f.write(r.stderr)
    return obj, r.stderr if verbose else ""


def build(verbose=False, force=False):
    os.makedirs(LIBDIR, exist_ok=True)
    os.makedirs(OBJDIR, exist_ok=True)
    if force:
        for f in os.listdir(OBJDIR):
            os.remove(os.path.join(OBJDIR, f))
    with ThreadPoolExecutor(max_workers=8) as ex:
        results = list(ex.map(lambda s: _compile(s, verbose), _sources()))
    objs = [o for o, _ in results]
    if verbose:
        for _, log in results:
            sys.stderr.write(log)
    if (not os.path.exists(LIB)) or any(os.path.getmtime(o) > os.path.getmtime(LIB) for o in objs):
        r = subprocess.run([NVCC, "-shared", "-o", LIB, *objs, "-gencode", "arch=compute_100a,code=sm_100a",
                            "-cudart", "static"], capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
    return LIB


FACADE_DIR = os.path.join(HERE, "facade")
FACADE_TEST = os.path.join(LIBDIR, "facade_test")


def build_facade_test():
    """Compile the C++ facade (namespace GravoMG, reference signatures) and its self-check program
    against the Eigen shim. Skipped quietly if the facade sources are not there yet."""
    src = os.path.join(FACADE_DIR, "facade_test.cpp")
    if not os.path.exists(src):
        return None
    root = os.path.dirname(HERE)
    srcs = [src, os.path.join(FACADE_DIR, "gravomg_facade.cpp"), os.path.join(FACADE_DIR, "facade_extras.cpp")]
    deps = srcs + [LIB, os.path.join(FACADE_DIR, "include", "gravomg", "b200_facade.h"),
                   os.path.join(root, "include", "gravomg_b200.h")]
    if os.path.exists(FACADE_TEST) and os.path.getmtime(FACADE_TEST) > max(os.path.getmtime(d) for d in deps):
        return FACADE_TEST
    cmd = ["g++", "-std=c++20", "-O2", "-ffp-contract=off", "-I", os.path.join(FACADE_DIR, "include"),
           "-I", os.path.join(root, "oracle", "eigen_shim"), "-I", os.path.join(root, "include"),
           *srcs, "-L", LIBDIR, "-lgravomg_b200",
           "-pthread", "-Wl,-rpath,$ORIGIN", "-o", FACADE_TEST]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"facade build failed:\n{r.stdout}\n{r.stderr}")
    return FACADE_TEST


if __name__ == "__main__":
    print(build(verbose="-v" in sys.argv, force="-f" in sys.argv))
    print(build_facade_test())        # links against the library and bakes in the header's struct layouts
