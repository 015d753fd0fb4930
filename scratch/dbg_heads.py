import sys, time, faulthandler
faulthandler.dump_traceback_later(60, exit=True)
sys.path.insert(0, "alphazero-gomoku_b200"); sys.path.insert(0, "tests/golden")
import numpy as np, torch
import engine, workload
from weights import make_state_dict
w = {k: make_state_dict(k, 1234) for k in ("pv", "graphnet", "dqn")}
G = int(sys.argv[1]); pri = int(sys.argv[2]); lanes = int(sys.argv[3])
eng = engine.Engine(max_games=G, max_simulations=200, precision=1, store_priors=bool(pri), lanes=lanes)
eng.set_pv_weights(w["pv"]); eng.set_planner_weights(w["graphnet"], w["dqn"])
cells, tm, mc = workload.make_positions(G, seed=1)
boards = eng.boards_from_cells(cells, tm, mc)
def t(name, f):
    t0 = time.time(); r = f(); torch.cuda.synchronize(); print(name, "ok %.3fs" % (time.time() - t0), flush=True); return r
t("pv_forward", lambda: eng.pv_forward(boards))
t("planner_nets", lambda: eng.planner_nets_forward(boards))
for S, ps in ((2, 0), (2, 1), (8, 5), (200, 5)):
    p = engine.make_params(S, 1.8, beta=0.2, planner_steps=ps, difficulty="hard", seed=3)
    t("search S=%d ps=%d" % (S, ps), lambda: eng.search(boards, tm, p))
print(eng.counters())
