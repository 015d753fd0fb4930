"""Builds libgomoku_b200.so in-tree with nvcc for sm_100a (the only target)."""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
SO = os.path.join(HERE, "libgomoku_b200.so")
SOURCES = ["gz_api.cu", "gz_kernels_board.cu", "gz_kernels_rollout.cu", "gz_kernels_planner.cu",
           "gz_kernels_tree.cu", "gz_kernels_selfplay.cu", "gz_kernels_dataset.cu", "gz_nn_fp32.cu", "gz_nn_bf16.cu", "gz_nn_fc.cu"]
FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
         "-Xcompiler", "-fPIC", "-fmad=true", "-cudart", "static"]
FLAGS += os.environ.get("GZ_EXTRA_NVCC_FLAGS", "").split()   # e.g. -DGZ_HEAD_TIMING: per-phase clock64 printf


def needs_build():
    if not os.path.exists(SO):
        return True
    t = os.path.getmtime(SO)
    deps = [os.path.join(CSRC, f) for f in os.listdir(CSRC)] + \
           [os.path.join(HERE, "..", "include", "gomoku_b200.h")]
    return any(os.path.getmtime(d) > t for d in deps)


def build(force=False, verbose=False, out=None, extra_flags=()):
    """out / extra_flags: an instrumented variant next to the product library (e.g. scratch/gn_timing.py builds
    libgomoku_b200_timing.so with -DGZ_GN_TIMING); the product build takes neither."""
    so = out or SO
    if out is None and not force and not needs_build():
        return SO
    objs = []
    procs = []
    bdir = os.path.join(HERE, "build" if out is None else "build_" + os.path.basename(out).split(".")[0])
    os.makedirs(bdir, exist_ok=True)
    for src in SOURCES:
        obj = os.path.join(bdir, src.replace(".cu", ".o"))
        objs.append(obj)
        cmd = ["nvcc"] + FLAGS + list(extra_flags) + (["-Xptxas", "-v"] if verbose else []) + \
              ["-c", os.path.join(CSRC, src), "-o", obj]
        procs.append((src, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT)))
    for src, p in procs:
        out = p.communicate()[0].decode()
        if verbose or p.returncode:
            sys.stderr.write(out)
        if p.returncode:
            raise RuntimeError("nvcc failed on " + src)
    subprocess.check_call(["nvcc", "-shared", "-cudart", "static", "-o", so] + objs)
    return so


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
