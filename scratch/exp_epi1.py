"""Perf experiment: k_gn_block_pairs with epilogue 1 stripped of (1) its shuffles, (2) its D_b read as well.  Results are
wrong on purpose; only the kernel time matters (is the kernel paced by tensor-memory reads?).  usage: exp_epi1.py <so>"""
import os, sys
ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
sys.path[:0] = [os.path.join(ROOT, "alphazero-gomoku_b200"), ROOT]
import torch
import engine, workload, bench
engine.load_library(sys.argv[1])
os.environ["GZ_GN_FUSED_EMBED"] = "0"
G = 4096
eng = engine.Engine(max_games=G, max_simulations=8, precision=1)
pv, gn, dq = bench.synthetic_weights()
eng.set_pv_weights(pv); eng.set_planner_weights(gn, dq)
cells, tm, mc = workload.make_positions(G, seed=2026)
boards = eng.boards_from_cells(cells, tm, mc)
for _ in range(5):
    eng.planner_nets_forward(boards)
torch.cuda.synchronize()
t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
t0.record()
for _ in range(100):
    eng.planner_nets_forward(boards)
t1.record(); torch.cuda.synchronize()
print(os.path.basename(sys.argv[1]), "planner nets forward: %.1f us" % (t0.elapsed_time(t1) * 10))
