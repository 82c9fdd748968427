"""Build recipe for the native libraries (in-tree, so the .so files travel with the repo).

  libnorib200.so       CUDA kernels + the C ABI of include/norib200.h      (nvcc, sm_100a)
  libnorib200_host.so  C++ host mirror of the Nori plugin surface (XML, OBJ, BVH build, flatten)
  nori_b200            command-line renderer linking both

`python optix-renderer_b200/build.py` or `build_all()` rebuilds what is out of date.
"""
import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
CSRC = os.path.join(HERE, "csrc")
HOST = os.path.join(HERE, "host")
OBJ = os.path.join(HERE, "build")
LIB_CUDA = os.path.join(HERE, "libnorib200.so")
LIB_HOST = os.path.join(HERE, "libnorib200_host.so")
CLI = os.path.join(HERE, "nori_b200")

NVCC = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
CXX = shutil.which("g++") or "g++"
ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
NVCC_FLAGS = (["-O3", "-std=c++17", "-lineinfo", "-Xcompiler", "-fPIC", "--expt-relaxed-constexpr"] + ARCH
              + os.environ.get("NB_NVCC_EXTRA", "").split())
# the traversal TU must not contract a*b+c into FMA: hit parity with the x86-64 reference
# the render kernels' shading math is judged statistically (t-tests, image RMSE): approximate
# division / square root there; nb_trace.cuh asks for IEEE operations explicitly where hits depend on them
FAST = ["-prec-div=false", "-prec-sqrt=false"]
PER_FILE = {"nb_trace.cu": ["-fmad=false"], "nb_static.cu": ["-fmad=false"], "nb_mega.cu": FAST, "nb_render.cu": FAST}
CXX_FLAGS = ["-O2", "-std=c++17", "-fPIC", "-Wall", "-ffp-contract=off", "-pthread"]


def _newer(target, deps):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps if os.path.exists(d))


def _run(cmd, verbose):
    if verbose:
        print(" ".join(cmd), flush=True)
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if r.returncode != 0:
        raise RuntimeError("build failed: %s\n%s" % (" ".join(cmd), r.stdout))
    if verbose and r.stdout.strip():
        print(r.stdout)
    return r.stdout


def _headers(*dirs):
    out = []
    for d in dirs:
        if os.path.isdir(d):
            out += [os.path.join(d, f) for f in os.listdir(d) if f.endswith((".h", ".cuh", ".hpp"))]
    return out


def build_cuda(verbose=False, extra=()):
    os.makedirs(OBJ, exist_ok=True)
    hdrs = _headers(CSRC, os.path.join(ROOT, "include"))
    objs, jobs = [], []
    for f in sorted(os.listdir(CSRC)):
        if not f.endswith(".cu"):
            continue
        src = os.path.join(CSRC, f)
        obj = os.path.join(OBJ, f[:-3] + ".o")
        objs.append(obj)
        if _newer(obj, [src] + hdrs):
            jobs.append([NVCC] + NVCC_FLAGS + PER_FILE.get(f, []) + list(extra) + ["-c", src, "-o", obj])
    if jobs:                                   # translation units are independent: compile them side by side
        from concurrent.futures import ThreadPoolExecutor
        with ThreadPoolExecutor(max_workers=min(len(jobs), os.cpu_count() or 1)) as ex:
            list(ex.map(lambda cmd: _run(cmd, verbose), jobs))
    if _newer(LIB_CUDA, objs):
        _run([NVCC] + ARCH + ["-shared", "-o", LIB_CUDA] + objs + ["-cudart", "static", "-ldl"], verbose)
    return LIB_CUDA


def build_host(verbose=False):
    if not os.path.isdir(HOST):
        return None
    os.makedirs(OBJ, exist_ok=True)
    hdrs = _headers(HOST, os.path.join(ROOT, "include"))
    objs = []
    main_obj = None
    for f in sorted(os.listdir(HOST)):
        if not f.endswith(".cpp"):
            continue
        src = os.path.join(HOST, f)
        obj = os.path.join(OBJ, "host_" + f[:-4] + ".o")
        if _newer(obj, [src] + hdrs):
            _run([CXX] + CXX_FLAGS + ["-I", os.path.join(ROOT, "include"), "-c", src, "-o", obj], verbose)
        if f == "main.cpp":
            main_obj = obj
        else:
            objs.append(obj)
    if objs and _newer(LIB_HOST, objs):
        _run([CXX, "-shared", "-o", LIB_HOST] + objs + ["-L", HERE, "-lnorib200", "-Wl,-rpath,$ORIGIN", "-pthread"], verbose)
    if main_obj and _newer(CLI, [main_obj, LIB_HOST, LIB_CUDA]):
        _run([CXX, "-o", CLI, main_obj, "-L", HERE, "-lnorib200_host", "-lnorib200", "-Wl,-rpath,$ORIGIN",
              "-pthread", "-ldl"], verbose)
    return LIB_HOST


def build_all(verbose=False):
    build_cuda(verbose)
    build_host(verbose)


if __name__ == "__main__":
    build_all(verbose="-q" not in sys.argv)
