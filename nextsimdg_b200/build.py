"""Builds nextsimdg_b200/csrc/libnsdg_gpu.so with nvcc for sm_100a, in-tree.

nvcc cross-compiles without a GPU.  The library is git-ignored but travels to the GPU box with
the repo snapshot.  `python -m nextsimdg_b200.build [-v]`
"""
import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OUT = os.path.join(CSRC, "libnsdg_gpu.so")
OBJDIR = os.path.join(CSRC, "build")

ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
COMMON = ["-std=c++17", "-O3", "-lineinfo", "-Xcompiler", "-fPIC"]
# per-file extras
SOURCES = {
    "ctx.cu": [],
    "thermo.cu": [],
    # no automatic FMA contraction: every fused multiply-add of the transport arithmetic is written
    # out, so that all kernels and all unrolled loop copies round alike (see transport.cu)
    "transport.cu": ["-fmad=false"],
    "mevp.cu": [],
    "dyn.cu": [],
    "comm.cu": [],
    "noalloc.cu": [],
}


def _nvcc():
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(nvcc):
        raise RuntimeError("nvcc not found")
    return nvcc


def _stale(target, deps):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps)


def build(verbose=False, force=False):
    nvcc = _nvcc()
    os.makedirs(OBJDIR, exist_ok=True)
    headers = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".h", ".cuh"))]
    headers.append(os.path.join(HERE, "..", "include", "nsdg_gpu.h"))
    headers.append(os.path.abspath(__file__))
    objs = []
    for src, extra in SOURCES.items():
        path = os.path.join(CSRC, src)
        if not os.path.exists(path):
            raise RuntimeError(f"missing source {path}")
        obj = os.path.join(OBJDIR, src.replace(".cu", ".o"))
        objs.append(obj)
        if force or _stale(obj, [path] + headers):
            cmd = [nvcc] + ARCH + COMMON + extra + ["-Xptxas", "-v"] * int(verbose) + \
                  ["-c", path, "-o", obj]
            if verbose:
                print(" ".join(cmd), flush=True)
            r = subprocess.run(cmd, capture_output=True, text=True)
            if verbose or r.returncode:
                sys.stderr.write(r.stdout + r.stderr)
            if r.returncode:
                raise RuntimeError(f"nvcc failed on {src}")
    if force or _stale(OUT, objs):
        # NCCL is bound at run time with dlopen (comm.cu), so the library loads on a CPU-only
        # box for the symbol-export test.  libstdc++ must be the shared one: this image's g++
        # wrapper resolves -lstdc++ to the static archive, which breaks inside python.
        cmd = [nvcc] + ARCH + ["-shared", "-o", OUT] + objs + \
              ["-cudart", "static", "-Xlinker", "--no-as-needed", "-ldl", "-lpthread",
               "-Xcompiler", "-nostdlib++", "-L/usr/lib/x86_64-linux-gnu", "-l:libstdc++.so.6"]
        if verbose:
            print(" ".join(cmd), flush=True)
        r = subprocess.run(cmd, capture_output=True, text=True)
        if verbose or r.returncode:
            sys.stderr.write(r.stdout + r.stderr)
        if r.returncode:
            raise RuntimeError("link failed")
    return OUT


if __name__ == "__main__":
    print(build(verbose="-v" in sys.argv, force="-f" in sys.argv))
