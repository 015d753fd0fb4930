"""Stage times of the policy-value forward alone (back to back), for grid-size experiments."""
import os, sys
sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", "alphazero-gomoku_b200"))
sys.path.insert(0, os.path.join(os.path.dirname(__file__), ".."))
import torch, engine, workload
from bench import synthetic_weights
G = 4096
eng = engine.Engine(max_games=G, max_simulations=8, device=0, precision=1, lanes=1)
pv, gn, dq = synthetic_weights()
eng.set_pv_weights(pv); eng.set_planner_weights(gn, dq)
cells, tm, mc = workload.make_positions(G, seed=1)
boards = eng.boards_from_cells(cells, tm, mc)
for _ in range(5): eng.pv_forward(boards)
torch.cuda.synchronize()
eng.set_profiling(True); eng.stage_times()
for _ in range(100): eng.pv_forward(boards)
torch.cuda.synchronize()
st = eng.stage_times()
print(os.environ.get("GZ_CONV_GRID"), {k: round(1e3 * v[0] / v[1], 1) for k, v in st.items() if v[1]})
