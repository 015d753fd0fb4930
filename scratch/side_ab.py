"""A/B: search step time with / without the DQN side stream (GZ_NO_SIDE_STREAM=1), profiling off."""
import os, sys, time
sys.path[:0] = [os.path.join(os.path.dirname(__file__), "..", "alphazero-gomoku_b200"), os.path.join(os.path.dirname(__file__), "..")]
import torch, numpy as np
import engine, workload
import bench
G, S = 4096, 200
eng = engine.Engine(max_games=G, max_simulations=S, precision=1)
pv, gn, dq = bench.synthetic_weights()
eng.set_pv_weights(pv); eng.set_planner_weights(gn, dq)
cells, tm, mc = workload.make_positions(G, seed=2026)
boards = eng.boards_from_cells(cells, tm, mc)
p = lambda k: engine.make_params(S, 1.8, beta=0.2, planner_steps=5, difficulty="hard", deterministic=False, seed=k)
for k in range(2): eng.search(boards, tm, p(k))
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for k in range(3): eng.search(boards, tm, p(10 + k))
e1.record(); torch.cuda.synchronize()
print("side" if not os.environ.get("GZ_NO_SIDE_STREAM") else "no-side", e0.elapsed_time(e1) / 3, "ms/step")
