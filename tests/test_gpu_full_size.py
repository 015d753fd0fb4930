"""BASELINE.json's configurations at their full sizes on the B200:
config 2 -- medium preset, S=100, 1024 concurrent games, deterministic mode: a spread of trees against the oracle bit
for bit, and size-independent properties on all 1024;
config 3/4 -- hard preset, S=200, beta=0.2, planner_steps=5, 4096 games, Philox randomness, bf16 networks: the
statistics of a game must not depend on how the games are sharded (game ids are global), so two
half-size calls with shifted game_base must reproduce the full call bit for bit; and a rerun is identical."""
import numpy as np
import pytest

import oracle as orc

pytestmark = pytest.mark.gpu


def _root_properties(cells, cv, cval, best, S):
    empty = cells == 0
    assert not cv[~empty].any() and not cval[~empty].any()            # children only on legal cells
    n_empty = empty.sum(1)
    assert np.array_equal(cv.sum(1), np.minimum(S - 1, np.where(n_empty > 0, S - 1, 0)))   # one node per simulation
    # expansion pops the LAST unexplored move: with S-1 <= E the children are the S-1 highest empty cells
    for g in range(0, len(cells), 37):
        if n_empty[g] >= S - 1:
            want = np.flatnonzero(empty[g])[-(S - 1):]
            assert np.array_equal(np.flatnonzero(cv[g]), want)
    first_max = np.where(cv.max(1) > 0, cv.argmax(1), -1)
    # max(children, key=visits) keeps the first maximum in INSERTION order = highest cell first (reverse raster)
    for g in range(len(cells)):
        if first_max[g] >= 0:
            top = np.flatnonzero(cv[g] == cv[g].max())
            assert best[g] == top[-1]


def test_config2_medium_1024_games_deterministic(weights, oracle):
    import engine
    import workload
    G, S = 1024, 100
    eng = engine.Engine(max_games=G, max_simulations=S)
    eng.set_pv_weights(weights["pv"])
    cells, tm, mc = workload.make_positions(G, seed=2)
    p = engine.make_params(S, 1.6, beta=0.2, planner_steps=0, difficulty="medium", deterministic=True)
    cv, cval, best = eng.search_host(cells, tm, mc, tm, p)
    cv2, cval2, best2 = eng.search_host(cells, tm, mc, tm, p)
    assert np.array_equal(cv, cv2) and np.array_equal(cval, cval2) and np.array_equal(best, best2)
    _root_properties(cells, cv, cval, best, S)
    assert eng.counters()["simulations"] == 2 * G * S
    for i in range(5, G, 41):                                          # 25 trees spread over all stone counts
        b = orc.board_from_cells(cells[i], tm[i], mc[i])
        t = oracle.search(b, oracle.params(S, 1.6, 0.2, 0, int(tm[i]), "medium"))
        ch = np.flatnonzero(t["parent"] == 0)
        wv = np.zeros(225, np.int32); wx = np.zeros(225, np.float64)
        wv[t["move"][ch]] = t["visits"][ch]; wx[t["move"][ch]] = t["value"][ch]
        assert np.array_equal(cv[i], wv) and np.array_equal(cval[i], wx), i
        assert best[i] == t["stats"].best_move
    eng.close()


def test_config3_hard_4096_games_sharding_invariance(weights):
    import engine
    import workload
    G, S = 4096, 200
    eng = engine.Engine(max_games=G, max_simulations=S, precision=1)
    eng.set_pv_weights(weights["pv"])
    eng.set_planner_weights(weights["graphnet"], weights["dqn"])
    cells, tm, mc = workload.make_positions(G, seed=3)

    def run(lo, hi):
        p = engine.make_params(S, 1.8, beta=0.2, planner_steps=5, difficulty="hard", deterministic=False,
                               seed=0xB200, game_base=lo)
        return eng.search_host(cells[lo:hi], tm[lo:hi], mc[lo:hi], tm[lo:hi], p)

    cv, cval, best = run(0, G)
    c = eng.counters()
    assert c["simulations"] == G * S and c["planner_steps"] > 4 * G * S
    _root_properties(cells, cv, cval, best, S)
    assert np.all(np.abs(cval) <= cv + 1e-9)                          # every playout value lies in [-1, 1]
    assert len(np.unique(np.round(cval[cv > 0], 6))) > 2               # real random playouts: wins, losses, cut-offs
    for lo, hi in ((0, G // 2), (G // 2, G)):                          # the 2-GPU sharding of the same games
        cv_h, cval_h, best_h = run(lo, hi)
        assert np.array_equal(cv_h, cv[lo:hi]) and np.array_equal(cval_h, cval[lo:hi])
        assert np.array_equal(best_h, best[lo:hi])
    eng.close()


def test_large_batch_16384_games_matches_small_slices(weights):
    """Capacity beyond the benchmark sizes (BASELINE config 4 runs 8192 games per GPU; DESIGN.md 4 sizes 65 536): 16 384
    concurrent games put 110-111 boards on every CTA of the persistent tensor kernels (barrier phases wrap dozens of
    times, every ring slot is reused).  Slices of 128 games taken from the start, the middle and the end must reproduce
    the big call bit for bit when they are searched alone with the same global game ids -- bf16 networks, Philox
    randomness, planner-guided playouts -- and the policy-value outputs of those boards must be identical too."""
    import engine
    import workload
    G, S = 16384, 6
    eng = engine.Engine(max_games=G, max_simulations=8, precision=1)
    eng.set_pv_weights(weights["pv"])
    eng.set_planner_weights(weights["graphnet"], weights["dqn"])
    cells, tm, mc = workload.make_positions(G, seed=11)

    def run(lo, hi):
        p = engine.make_params(S, 1.8, beta=0.2, planner_steps=3, difficulty="hard", deterministic=False,
                               seed=0x16384, game_base=lo)
        return eng.search_host(cells[lo:hi], tm[lo:hi], mc[lo:hi], tm[lo:hi], p)

    cv, cval, best = run(0, G)
    assert eng.counters()["simulations"] == G * S
    boards = eng.boards_from_cells(cells, tm, mc)
    lg, pr, v = (t.cpu().numpy() for t in eng.pv_forward(boards))
    for lo in (0, 8064, G - 128):
        cv_s, cval_s, best_s = run(lo, lo + 128)
        assert np.array_equal(cv_s, cv[lo:lo + 128]) and np.array_equal(cval_s, cval[lo:lo + 128]), lo
        assert np.array_equal(best_s, best[lo:lo + 128]), lo
        lg_s, pr_s, v_s = (t.cpu().numpy() for t in eng.pv_forward(boards[lo:lo + 128].clone()))
        assert np.array_equal(lg_s, lg[lo:lo + 128]) and np.array_equal(v_s, v[lo:lo + 128]), lo
    eng.close()
