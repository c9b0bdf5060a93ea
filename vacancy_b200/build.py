"""Builds libvacancy_b200.so in-tree with nvcc for sm_100a (no JIT cache, no torch extension).

`python -m vacancy_b200.build` or `build_library()`; the .so lands next to this file so that it
travels to the GPU box with the repository snapshot.
"""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libvacancy_b200.so")
SOURCES = ["carver.cu", "sdf2d.cu", "fusion.cu", "extract_voxel.cu", "marching_cubes.cu", "ply.cu",
           "sharded.cu", "checkpoint.cu", "ingest.cu"]
NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
    # never contract a*b+c: the fused grid must match the reference's one-rounding-per-op arithmetic
    "-fmad=false", "-prec-div=true", "-prec-sqrt=true", "-ftz=false",
    "-Xcompiler", "-fPIC,-O2,-Wall,-Wno-unused-function", "-shared",
]


LINK_LIBS = ["-ldl"]  # NCCL is resolved with dlopen at run time (csrc/sharded.cu), never linked


def _stale():
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    deps = [os.path.join(CSRC, f) for f in os.listdir(CSRC)]
    deps.append(os.path.join(HERE, "..", "include", "vacancy_b200.h"))
    return any(os.path.getmtime(d) > t for d in deps)


def build_library(force=False, verbose=False):
    if not force and not _stale():
        return LIB
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    # the sources are compiled in parallel into objects, linked into a temporary file and moved
    # into place, so that a repository snapshot taken meanwhile never sees a half-written library
    objdir = os.path.join(HERE, "build")
    os.makedirs(objdir, exist_ok=True)
    compile_flags = [f for f in NVCC_FLAGS if f != "-shared"]
    procs = []
    for src in SOURCES:
        obj = os.path.join(objdir, src.replace(".cu", ".o"))
        cmd = [nvcc] + compile_flags + (["-Xptxas", "-v"] if verbose else []) + [
            "-c", os.path.join(CSRC, src), "-o", obj]
        procs.append((src, obj, subprocess.Popen(cmd, stdout=subprocess.PIPE,
                                                 stderr=subprocess.STDOUT, text=True)))
    objs, failed, log = [], [], ""
    for src, obj, pr in procs:
        out, _ = pr.communicate()
        log += out
        objs.append(obj)
        if pr.returncode != 0:
            failed.append(src)
    if failed:
        sys.stderr.write(log)
        raise RuntimeError("nvcc failed compiling " + ", ".join(failed))
    tmp = LIB + ".tmp.%d" % os.getpid()
    r = subprocess.run([nvcc, "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-o", tmp]
                       + objs + LINK_LIBS, capture_output=True, text=True)
    r.stdout = log + r.stdout
    if r.returncode != 0:
        sys.stderr.write(r.stdout + r.stderr)
        raise RuntimeError("nvcc failed linking libvacancy_b200.so")
    os.replace(tmp, LIB)
    if verbose:
        sys.stderr.write(r.stdout + r.stderr)
    return LIB


if __name__ == "__main__":
    print(build_library(force=True, verbose="-v" in sys.argv))
