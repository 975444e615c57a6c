"""Builds libbricks_b200.so (sm_100a) in-tree with nvcc.  Called by __graft_entry__.build().

The .so is git-ignored but travels to the GPU box with the repo snapshot.
"""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIBDIR = os.path.join(HERE, "lib")
# kernel experiments: BRICKS_EXP_NAME=x BRICKS_EXP_FLAGS="-DFOO" builds lib/libbricks_b200_x.so beside the
# product library (load it with BRICKS_LIB=...); the product build never sets these
EXP = os.environ.get("BRICKS_EXP_NAME", "")
LIB = os.path.join(LIBDIR, "libbricks_b200%s.so" % ("_" + EXP if EXP else ""))
SOURCES = ["context.cu", "plan.cu", "stencil_march.cu", "stencil_naive.cu", "stencil4d.cu", "exchange.cu", "step.cu", "array_copy.cu", "probe.cu",
           "host_capi.cpp"]
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
         "-Xcompiler", "-fPIC", "-Xcompiler", "-Wall", "-Xcompiler", "-Wno-unused-function",
         "--expt-relaxed-constexpr", "-ccbin", "/usr/bin/g++", "-Xcompiler", "-fopenmp"]
FLAGS += os.environ.get("BRICKS_EXP_FLAGS", "").split()


def _stale(target, deps):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps)


def build(verbose=False, force=False, ptxas_v=False):
    os.makedirs(LIBDIR, exist_ok=True)
    objdir = os.path.join(LIBDIR, "obj" + ("_" + EXP if EXP else ""))
    os.makedirs(objdir, exist_ok=True)
    headers = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cuh", ".h"))]
    for root, _, files in os.walk(os.path.join(HERE, "..", "include")):
        headers += [os.path.join(root, f) for f in files]
    objs = []
    procs = []
    for src in SOURCES:
        s = os.path.join(CSRC, src)
        o = os.path.join(objdir, os.path.splitext(src)[0] + ".o")
        objs.append(o)
        if force or _stale(o, [s] + headers):
            cmd = [NVCC] + FLAGS + (["-Xptxas", "-v"] if ptxas_v else []) + ["-c", s, "-o", o]
            if verbose:
                print(" ".join(cmd))
            procs.append((src, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT)))
    failed = False
    for src, p in procs:
        out, _ = p.communicate()
        if p.returncode != 0 or verbose or ptxas_v:
            sys.stdout.write(out.decode())
        if p.returncode != 0:
            failed = True
            print("nvcc failed on %s" % src)
    if failed:
        raise RuntimeError("libbricks_b200 build failed")
    if force or _stale(LIB, objs + [os.path.join(CSRC, "exports.map")]):
        # cudart static; NCCL resolved at load time (torch's bundled libnccl.so.2 when torch is
        # imported first, else the system one)
        cmd = [NVCC, "-shared", "-o", LIB] + objs + ["-lnccl", "-lgomp", "-Xlinker", "--version-script=" + os.path.join(CSRC, "exports.map"), "-ccbin", "/usr/bin/g++",
                                                     "-gencode", "arch=compute_100a,code=sm_100a"]
        if verbose:
            print(" ".join(cmd))
        subprocess.check_call(cmd)
    return LIB


if __name__ == "__main__":
    build(verbose="-v" in sys.argv, force="-f" in sys.argv, ptxas_v="-p" in sys.argv)
    print(LIB)
