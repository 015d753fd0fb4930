"""Per-launch stage times of the net forwards run back to back (sustained, power capped) and with idle gaps."""
import os, sys, time
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
def run(n, gap, fn):
    for _ in range(5): fn(boards)
    torch.cuda.synchronize()
    eng.set_profiling(True); eng.stage_times()
    for _ in range(n):
        fn(boards)
        if gap: torch.cuda.synchronize(); time.sleep(gap)
    torch.cuda.synchronize()
    st = eng.stage_times(); eng.set_profiling(False)
    return {k: round(1e3 * v[0] / v[1], 1) for k, v in st.items() if v[1]}
for name, fn in (("pv", eng.pv_forward), ("planner_nets", eng.planner_nets_forward)):
    print(name, "back-to-back x400 us/stage:", run(400, 0, fn))
    print(name, "20 ms idle gaps  x40 us/stage:", run(40, 0.02, fn))
