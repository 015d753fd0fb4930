"""One hard-preset search of 4096 positions with S=64 simulations: the workload ncu captures a whole simulation wave of
(every kernel of the path once, the GraphNet ones five times).  Not a benchmark."""
import os, sys
sys.path[:0] = [os.path.join(os.path.dirname(__file__), "..", "alphazero-gomoku_b200"), os.path.join(os.path.dirname(__file__), "..")]
import torch
import engine, workload, bench
G, S = 4096, 64
eng = engine.Engine(max_games=G, max_simulations=S, precision=1)
pv, gn, dq = bench.synthetic_weights()
eng.set_pv_weights(pv); eng.set_planner_weights(gn, dq)
cells, tm, mc = workload.make_positions(G, seed=2026)
boards = eng.boards_from_cells(cells, tm, mc)
p = engine.make_params(S, 1.8, beta=0.2, planner_steps=5, difficulty="hard", deterministic=False, seed=1)
eng.search(boards, tm, p)
torch.cuda.synchronize()
print(eng.counters())
