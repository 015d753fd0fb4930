"""Where k_conv_umma<128,128,9>'s issuer and epilogue warps wait: an instrumented copy of the library (-DGZ_CONV_TIMING),
one policy-value forward on 4096 boards.  Diagnostic, not a benchmark.  Build here: python scratch/conv_timing.py --build-only"""
import importlib.util, os, sys
ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
PKG = os.path.join(ROOT, "alphazero-gomoku_b200")
sys.path[:0] = [PKG, ROOT]
SO = os.path.join(PKG, "libgomoku_b200_convtiming.so")
spec = importlib.util.spec_from_file_location("gz_build", os.path.join(PKG, "build.py"))
b = importlib.util.module_from_spec(spec); spec.loader.exec_module(b)
if "--build-only" in sys.argv or not os.path.exists(SO):
    b.build(out=SO, extra_flags=["-DGZ_CONV_TIMING"])
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
    eng.pv_forward(boards)
    torch.cuda.synchronize()
    print("----", flush=True)
