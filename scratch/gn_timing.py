"""Where k_gn_block's warps spend their cycles: builds an instrumented copy of the library (-DGZ_GN_TIMING, clock64
around every barrier wait of the producer / issuer / epilogue warps of block 0) and runs the planner networks once on
4096 boards.  Diagnostic, not a benchmark.  Build here (no GPU needed), run on the box:
    python scratch/gn_timing.py --build-only ; gpurun -- python scratch/gn_timing.py"""
import importlib.util
import os
import sys
ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
PKG = os.path.join(ROOT, "alphazero-gomoku_b200")
sys.path[:0] = [PKG, ROOT]
SO = os.path.join(PKG, "libgomoku_b200_timing.so")
spec = importlib.util.spec_from_file_location("gz_build", os.path.join(PKG, "build.py"))
b = importlib.util.module_from_spec(spec)
spec.loader.exec_module(b)
if "--build-only" in sys.argv or not os.path.exists(SO):
    b.build(out=SO, extra_flags=["-DGZ_GN_TIMING"])
    if "--build-only" in sys.argv:
        sys.exit(0)
import torch
import engine, workload, bench
engine.load_library(SO)
G = 4096
eng = engine.Engine(max_games=G, max_simulations=8, precision=1)
pv, gn, dq = bench.synthetic_weights()
eng.set_pv_weights(pv); eng.set_planner_weights(gn, dq)
cells, tm, mc = workload.make_positions(G, seed=2026)
boards = eng.boards_from_cells(cells, tm, mc)
for _ in range(2):
    eng.planner_nets_forward(boards)
    torch.cuda.synchronize()
    print("----", flush=True)
