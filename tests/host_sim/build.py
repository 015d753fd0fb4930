"""Builds tests/host_sim/libhost_sim.so (CPU desk-check of the kernels' __host__ __device__ logic).
Test infrastructure only."""
import ctypes as C
import os
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))
_ROOT = os.path.dirname(os.path.dirname(_HERE))


def lib():
    so = os.path.join(_HERE, "libhost_sim.so")
    srcs = [os.path.join(_HERE, "host_sim.cpp")] + [
        os.path.join(_ROOT, "alphazero-gomoku_b200", "csrc", f)
        for f in ("gz_bitboard.cuh", "gz_policy.cuh", "gz_lines.cuh")]
    if not os.path.exists(so) or any(os.path.getmtime(s) > os.path.getmtime(so) for s in srcs):
        subprocess.check_call(["g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-ffp-contract=off",
                               "-Wno-unknown-pragmas", "-o", so, srcs[0]])
    L = C.CDLL(so)
    L.hs_rollout_tail.restype = C.c_double
    return L
