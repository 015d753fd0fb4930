"""Per-phase cycle counts of k_head_umma inside a short search (library built with -DGZ_HEAD_TIMING)."""
import os, sys
sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", "alphazero-gomoku_b200"))
sys.path.insert(0, os.path.join(os.path.dirname(__file__), ".."))
import torch, engine, workload
from bench import synthetic_weights, PRESET
G = 4096
eng = engine.Engine(max_games=G, max_simulations=8, device=0, precision=1, lanes=1, store_priors=True)
pv, gn, dq = synthetic_weights()
eng.set_pv_weights(pv); eng.set_planner_weights(gn, dq)
cells, tm, mc = workload.make_positions(G, seed=1)
boards = eng.boards_from_cells(cells, tm, mc)
p = engine.make_params(2, PRESET["c_puct"], beta=PRESET["beta"], planner_steps=5, difficulty="hard", deterministic=False, seed=1, game_base=0)
eng.search(boards, tm.copy(), p)
torch.cuda.synchronize()
