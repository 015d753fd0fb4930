"""Per-phase cycle counts of k_head_umma (library built with GZ_EXTRA_NVCC_FLAGS=-DGZ_HEAD_TIMING)."""
import os, sys
sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", "alphazero-gomoku_b200"))
sys.path.insert(0, os.path.join(os.path.dirname(__file__), ".."))
import torch, engine, workload
from bench import synthetic_weights
G = 4096
eng = engine.Engine(max_games=G, max_simulations=8, device=0, precision=1, lanes=1, store_priors=True)
pv, gn, dq = synthetic_weights()
eng.set_pv_weights(pv); eng.set_planner_weights(gn, dq)
cells, tm, mc = workload.make_positions(G, seed=1)
boards = eng.boards_from_cells(cells, tm, mc)
for _ in range(3):
    eng.pv_forward(boards); eng.planner_nets_forward(boards)
torch.cuda.synchronize()
