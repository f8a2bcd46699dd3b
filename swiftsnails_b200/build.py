"""In-tree native build for swiftsnails_b200.

Two artefacts, both written under ``swiftsnails_b200/lib`` (git-ignored, shipped to the GPU box
with the working-tree snapshot):

* ``libssb200_kernels.so`` — every sm_100a CUDA kernel behind the C ABI of ``include/ssb200.h``
  (``nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo``; cross-compiles without a GPU).
* ``libssb200_core.so`` + role binaries — the C++17 host core (``csrc/core``): buffers, config,
  hash fragments, ZeroMQ transfer, CPU sparse table, master/server/worker roles.

The build is incremental: an object is recompiled only when its source (or any header next to
it) is newer.  ``python -m swiftsnails_b200.build`` builds everything.
"""
from __future__ import annotations

import glob
import os
import shutil
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB_DIR = os.path.join(ROOT, "swiftsnails_b200", "lib")
OBJ_DIR = os.path.join(ROOT, "build", "obj")
BIN_DIR = os.path.join(ROOT, "swiftsnails_b200", "bin")

NVCC = os.environ.get("NVCC") or shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"


def _pick_cxx():
    """The host compiler for the C++ core. The shared library is loaded into Python processes that already hold the
    system libstdc++ (numpy, torch), so it must link libstdc++ DYNAMICALLY: a compiler that can only find libstdc++.a
    would put a second, exported copy of the runtime inside libssb200_core.so, and iostream calls then mix the two
    (std::ofstream in ssb_cpu_table_dump crashed that way). Such a compiler is passed over."""
    cands = [os.environ.get("CXX"), shutil.which("g++"), "/usr/bin/g++"]
    seen = []
    for c in cands:
        if not c or c in seen:
            continue
        seen.append(c)
        try:
            so = subprocess.run([c, "-print-file-name=libstdc++.so"], capture_output=True, text=True).stdout.strip()
        except OSError:
            continue
        if os.path.isabs(so) and os.path.exists(so):
            return c
    return seen[0] if seen else "g++"


CXX = _pick_cxx()

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
    "-Xcompiler", "-fPIC", "--expt-relaxed-constexpr", "-Xptxas", "-v",
]
CXX_FLAGS = ["-O2", "-g", "-std=c++17", "-fPIC", "-pthread", "-Wall", "-Wno-unused-function"]


def _newest_mtime(paths):
    return max((os.path.getmtime(p) for p in paths if os.path.exists(p)), default=0.0)


def _needs(out, deps):
    return not os.path.exists(out) or os.path.getmtime(out) < _newest_mtime(deps)


def _run(cmd, log_path=None):
    proc = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if log_path:
        with open(log_path, "w") as f:
            f.write(" ".join(cmd) + "\n" + proc.stdout)
    if proc.returncode != 0:
        raise RuntimeError("build step failed: %s\n%s" % (" ".join(cmd), proc.stdout))
    return proc.stdout


def find_libzmq():
    """pyzmq bundles libzmq 4.3.5; the host core links against it (no zmq.h on this image, the
    prototypes are declared in csrc/core/zmq_api.h)."""
    env = os.environ.get("SSB200_LIBZMQ")
    if env and os.path.exists(env):
        return env
    for sp in sys.path:
        hits = glob.glob(os.path.join(sp, "pyzmq.libs", "libzmq*.so*"))
        if hits:
            return hits[0]
    for pat in ("/usr/lib/x86_64-linux-gnu/libzmq.so*", "/usr/local/lib/libzmq.so*"):
        hits = glob.glob(pat)
        if hits:
            return hits[0]
    return None


def build_kernels(verbose=False):
    os.makedirs(LIB_DIR, exist_ok=True)
    os.makedirs(OBJ_DIR, exist_ok=True)
    srcs = sorted(glob.glob(os.path.join(ROOT, "csrc", "kernels", "*.cu")))
    hdrs = (glob.glob(os.path.join(ROOT, "csrc", "kernels", "*.cuh")) + glob.glob(os.path.join(ROOT, "include", "*.h"))
            + glob.glob(os.path.join(ROOT, "include", "*.cuh")))
    objs = []
    jobs = []
    for src in srcs:
        obj = os.path.join(OBJ_DIR, os.path.basename(src)[:-3] + ".o")
        objs.append(obj)
        if _needs(obj, [src] + hdrs):
            jobs.append((src, obj))

    def compile_one(job):
        src, obj = job
        _run([NVCC] + NVCC_FLAGS + ["-c", src, "-o", obj], log_path=obj + ".log")
        return src

    with ThreadPoolExecutor(max_workers=min(8, max(1, len(jobs)))) as ex:
        for src in ex.map(compile_one, jobs):
            if verbose:
                print("nvcc", os.path.relpath(src, ROOT))
    out = os.path.join(LIB_DIR, "libssb200_kernels.so")
    if _needs(out, objs):
        _run([NVCC, "-shared", "-o", out] + objs + ["-lcudart", "-lcuda"])
        if verbose:
            print("link", os.path.relpath(out, ROOT))
    return out


def build_core(verbose=False):
    os.makedirs(LIB_DIR, exist_ok=True)
    os.makedirs(OBJ_DIR, exist_ok=True)
    os.makedirs(BIN_DIR, exist_ok=True)
    core_dir = os.path.join(ROOT, "csrc", "core")
    srcs = sorted(glob.glob(os.path.join(core_dir, "*.cc")))
    if not srcs:
        return None
    hdrs = (glob.glob(os.path.join(core_dir, "**", "*.h"), recursive=True)
            + glob.glob(os.path.join(ROOT, "csrc", "apps", "*.h"))
            + glob.glob(os.path.join(ROOT, "csrc", "tests", "*.h"))
            + glob.glob(os.path.join(ROOT, "include", "*.h")))
    libzmq = find_libzmq()
    if libzmq is None:
        raise RuntimeError("libzmq not found (pyzmq.libs); set SSB200_LIBZMQ")
    zmq_dir = os.path.dirname(libzmq)
    link_zmq = ["-L" + zmq_dir, "-l:" + os.path.basename(libzmq), "-Wl,-rpath," + zmq_dir]
    objs = []
    jobs = []
    for src in srcs:
        obj = os.path.join(OBJ_DIR, "core_" + os.path.basename(src)[:-3] + ".o")
        objs.append(obj)
        if _needs(obj, [src] + hdrs):
            jobs.append((src, obj))

    def compile_one(job):
        src, obj = job
        _run([CXX] + CXX_FLAGS + ["-I", os.path.join(ROOT, "include"), "-c", src, "-o", obj])
        return src

    with ThreadPoolExecutor(max_workers=min(8, max(1, len(jobs)))) as ex:
        for src in ex.map(compile_one, jobs):
            if verbose:
                print("g++", os.path.relpath(src, ROOT))
    out = os.path.join(LIB_DIR, "libssb200_core.so")
    if _needs(out, objs):
        _run([CXX, "-shared", "-o", out] + objs + link_zmq + ["-pthread"])
        if verbose:
            print("link", os.path.relpath(out, ROOT))
    # role binaries + C++ unit-test binary
    rpath = ["-Wl,-rpath," + LIB_DIR, "-Wl,-rpath,$ORIGIN/../lib"]
    app_srcs = sorted(glob.glob(os.path.join(ROOT, "csrc", "apps", "*.cc"))) + sorted(
        glob.glob(os.path.join(ROOT, "csrc", "tests", "*.cc")))
    app_jobs = []
    for src in app_srcs:
        exe = os.path.join(BIN_DIR, os.path.basename(src)[:-3])
        if _needs(exe, [src, out] + hdrs):
            app_jobs.append((src, exe))

    def link_one(job):
        src, exe = job
        _run([CXX] + CXX_FLAGS + ["-I", os.path.join(ROOT, "include"), "-I", core_dir, src, "-o", exe,
                                  "-L" + LIB_DIR, "-lssb200_core"] + rpath + link_zmq + ["-pthread"])
        return exe

    with ThreadPoolExecutor(max_workers=min(8, max(1, len(app_jobs)))) as ex:
        for exe in ex.map(link_one, app_jobs):
            if verbose:
                print("g++", os.path.relpath(exe, ROOT))
    return out


def build_all(verbose=False):
    k = build_kernels(verbose)
    c = build_core(verbose)
    return k, c


if __name__ == "__main__":
    k, c = build_all(verbose=True)
    print("kernels:", k)
    print("core:   ", c)
