"""Target of the compute-sanitizer passes on record under profiles/ (memcheck, racecheck, synccheck): the smoke() shape --
16 positions, planner_steps = 2, deterministic -- through gz_search_host in both network modes, plus one policy-value
and one planner-network forward and a short self-play session.
    compute-sanitizer --tool memcheck  python scratch/sanitize_target.py --sims 24
    compute-sanitizer --tool racecheck python scratch/sanitize_target.py --sims 6"""
import argparse
import os
import sys
ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
sys.path[:0] = [os.path.join(ROOT, "alphazero-gomoku_b200"), os.path.join(ROOT, "tests", "golden")]
import engine, workload
from weights import make_state_dict
ap = argparse.ArgumentParser()
ap.add_argument("--sims", type=int, default=24)
ap.add_argument("--games", type=int, default=16)
ap.add_argument("--no-graph", action="store_true")
a = ap.parse_args()
w = {k: make_state_dict(k, 1234) for k in ("pv", "graphnet", "dqn")}
cells, tm, mc = workload.make_positions(a.games, seed=7, plies=(12, 48, 150, 190))
for prec in (1, 0):
    eng = engine.Engine(max_games=a.games, max_simulations=max(a.sims, 8), precision=prec, store_priors=True)
    if a.no_graph:
        eng.set_graph(False)
    eng.set_pv_weights(w["pv"]); eng.set_planner_weights(w["graphnet"], w["dqn"])
    p = engine.make_params(a.sims, 1.8, beta=0.2, planner_steps=2, difficulty="hard", deterministic=False, seed=5)
    cv, cval, best = eng.search_host(cells, tm, mc, tm, p)
    boards = eng.boards_from_cells(cells, tm, mc)
    eng.pv_forward(boards); eng.planner_nets_forward(boards); eng.board_ops(boards)
    eng.get_priors(0, 0)
    eng.selfplay_begin(a.games, p, 0.2, 0, 0)
    for _ in range(3):
        eng.selfplay_step()
    eng.selfplay_drain()
    c = eng.counters()
    print("precision", prec, "simulations", c["simulations"], "launches", c["kernel_launches"], flush=True)
    eng.close()
print("sanitize target done")
