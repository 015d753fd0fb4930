#!/usr/bin/env python
"""Fixtures produced by the PINNED C oracle (oracle/gomoku_oracle.c) for inputs that are too expensive to run
through the oracle on the GPU box during a test (minutes to hours of CPU): the timed configuration's stochastic
trees and large-n outcome rates.  The oracle itself is pinned bit for bit to the unmodified Python reference by
tests/test_oracle_vs_golden.py; tests/test_oracle_vs_golden.py also re-runs a few entries of these files so a
stale fixture cannot go unnoticed.

    python tests/golden/make_oracle_fixtures.py --parts config3,gaps,stoch --jobs 8

Parts (each -> tests/golden/oracle_<part>.npz):
  config3  BASELINE config 3 (hard preset as the README quotes it: S=200, c_puct 1.8, beta 0.2, planner_steps 5,
           k 16, alpha 0.75, explore 0.05), Philox randomness shared with the engine (key = SEED, counter =
           (step, simulation, game id, tag)): 132 of the 4096 workload positions (every 31st: all six stone counts),
           game ids GAME_BASE + k.  Whole trees + the near-tie diagnostics of the blended pick.
  gaps     near-tie diagnostics (min top-2 gap of the blended pick) of the DET-B reference trees in
           mcts_planner.npz, for the bf16 flip accounting.
  stoch    outcome counts (win / loss / draw / cut-off) of n = 102400 Philox playouts per position with an
           oracle-only seed, planner_steps 0 and 5, on the positions of stoch.npz: the reference sampler of the
           binomial test (SURVEY.md 8c asks for n >= 1e5 through the validated C restatement).
"""
import argparse
import multiprocessing as mp
import os
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
for p in (HERE, os.path.join(ROOT, "oracle"), os.path.join(ROOT, "alphazero-gomoku_b200")):
    sys.path.insert(0, p)

import oracle as orc  # noqa: E402
from weights import make_state_dict  # noqa: E402

# config 3 (shared with tests/test_gpu_timed_path.py)
C3 = dict(G=4096, workload_seed=3, stride=31, first=7, S=200, c_puct=1.8, beta=0.2, planner_steps=5,
          difficulty="hard", seed=0xB200C0DE, game_base=5000)
STOCH = dict(n=102400, chunk=2048, seed=0x0AC1E5EED, game_base=0)

_O = None


def _oracle():
    global _O
    if _O is None:
        _O = orc.Oracle(**{k: make_state_dict(k, 1234) for k in ("pv", "graphnet", "dqn")})
    return _O


def config3_positions():
    import workload
    cells, tm, mc = workload.make_positions(C3["G"], seed=C3["workload_seed"])
    idx = np.arange(C3["first"], C3["G"], C3["stride"])
    return idx, cells[idx], tm[idx], mc[idx]


def _tree_task(args):
    k, cells, tm, mc, sims, c_puct, beta, ps, diff, ai, seed, game = args
    O = _oracle()
    b = orc.board_from_cells(cells, tm, mc)
    t0 = time.time()
    t = O.search(b, O.params(sims, c_puct, beta, ps, ai, diff), seed=seed, game=game)
    st = t["stats"]
    return dict(k=k, n_nodes=st.n_nodes, best_move=st.best_move, parent=t["parent"], move=t["move"],
                visits=t["visits"], value=t["value"], min_gap=st.min_gap, min_gap_sim=st.min_gap_sim,
                min_gap_step=st.min_gap_step, n_argmax_picks=st.n_argmax_picks,
                n_planner_steps=st.n_planner_steps, seconds=time.time() - t0)


def _pack_trees(recs, nmax):
    out = {}
    for key, dt, fill in (("parent", np.int16, -1), ("move", np.int16, -1), ("visits", np.int32, 0),
                          ("value", np.float64, 0.0)):
        a = np.full((len(recs), nmax), fill, dt)
        for i, r in enumerate(recs):
            a[i, :r["n_nodes"]] = r[key]
        out[key] = a
    for key in ("n_nodes", "best_move", "min_gap", "min_gap_sim", "min_gap_step", "n_argmax_picks",
                "n_planner_steps"):
        out[key] = np.array([r[key] for r in recs])
    return out


def part_config3(pool, outdir):
    idx, cells, tm, mc = config3_positions()
    tasks = [(k, cells[k], int(tm[k]), int(mc[k]), C3["S"], C3["c_puct"], C3["beta"], C3["planner_steps"],
              C3["difficulty"], int(tm[k]), C3["seed"], C3["game_base"] + k) for k in range(len(idx))]
    recs = sorted(pool.imap_unordered(_tree_task, tasks), key=lambda r: r["k"])
    out = _pack_trees(recs, C3["S"] + 1)
    np.savez_compressed(os.path.join(outdir, "oracle_config3.npz"), index=idx, cells=cells, to_move=tm,
                        move_count=mc, **out)
    print("config3: %d trees, %.0f cpu-seconds, min gap %.3g" %
          (len(recs), sum(r["seconds"] for r in recs), out["min_gap"].min()))


def part_gaps(pool, outdir):
    g = np.load(os.path.join(HERE, "mcts_planner.npz"))
    tasks = [(i, g["cells"][i], int(g["to_move"][i]), int(g["move_count"][i]), int(g["sims"][i]),
              float(g["c_puct"][i]), float(g["beta"][i]), int(g["planner_steps"][i]), str(g["difficulty"][i]),
              int(g["ai_player"][i]), 0, 0) for i in range(len(g["sims"]))]
    recs = sorted(pool.imap_unordered(_tree_task, tasks), key=lambda r: r["k"])
    for i, r in enumerate(recs):   # the oracle must reproduce the reference tree before its diagnostics mean anything
        n = int(g["n_nodes"][i])
        assert r["n_nodes"] == n and np.array_equal(r["visits"], g["visits"][i][:n]) \
            and np.array_equal(r["value"], g["value"][i][:n]), i
    np.savez_compressed(os.path.join(outdir, "oracle_gaps.npz"),
                        min_gap=np.array([r["min_gap"] for r in recs]),
                        n_argmax_picks=np.array([r["n_argmax_picks"] for r in recs]))
    print("gaps:", [float("%.3g" % r["min_gap"]) for r in recs])


def _stoch_task(args):
    pos, cells, tm, mc, ps, lo, hi = args
    O = _oracle()
    if not getattr(O, '_cache_on', False):
        O.L.gzo_set_planner_cache(1 << 15)
        O._cache_on = True
    b = orc.board_from_cells(cells, tm, mc)
    p = O.params(1, 1.8, 0.2, ps, tm, "hard")
    counts = np.zeros(4, np.int64)
    steps = 0
    for i in range(lo, hi):
        rng = orc.Rng(1, STOCH["seed"] & 0xFFFFFFFF, STOCH["seed"] >> 32, STOCH["game_base"] + i, pos)
        v, fb, _ = O.simulate(b, p, rng)
        counts[{1.0: 0, -1.0: 1, 0.1: 2, 0.0: 3}[v]] += 1
        steps += fb.move_count - mc
    return pos, ps, counts, steps


def part_stoch(pool, outdir):
    g = np.load(os.path.join(HERE, "stoch.npz"))
    npos = len(g["cells"])
    tasks = []
    for ps in (0, 5):
        for i in range(npos):
            for lo in range(0, STOCH["n"], STOCH["chunk"]):
                tasks.append((i, g["cells"][i], int(g["to_move"][i]), int(g["move_count"][i]), ps, lo,
                              lo + STOCH["chunk"]))
    counts = np.zeros((2, npos, 4), np.int64)
    steps = np.zeros((2, npos), np.int64)
    t0 = time.time()
    for pos, ps, c, s in pool.imap_unordered(_stoch_task, tasks, chunksize=STOCH["n"] // STOCH["chunk"]):   # one position per worker: its memo stays warm
        counts[ps // 5, pos] += c
        steps[ps // 5, pos] += s
    np.savez_compressed(os.path.join(outdir, "oracle_stoch.npz"), cells=g["cells"], to_move=g["to_move"],
                        move_count=g["move_count"], planner_steps=np.array([0, 5]), counts=counts,
                        total_steps=steps, n=np.int64(STOCH["n"]))
    print("stoch: %.0f s" % (time.time() - t0))
    print(counts)


PARTS = dict(config3=part_config3, gaps=part_gaps, stoch=part_stoch)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--parts", default="config3,gaps,stoch")
    ap.add_argument("--jobs", type=int, default=os.cpu_count())
    ap.add_argument("--out", default=HERE)
    a = ap.parse_args()
    orc.build()
    with mp.get_context("fork").Pool(a.jobs) as pool:
        for p in a.parts.split(","):
            t0 = time.time()
            PARTS[p](pool, a.out)
            print("  [%s] %.1f s" % (p, time.time() - t0), flush=True)


if __name__ == "__main__":
    main()
