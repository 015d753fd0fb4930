"""Edge cases and error behaviour of the C ABI on the B200: ragged batch sizes (1, 127, 129, 130 boards -- the tensor
kernels work on tiles of 128), finished / full / 199-stone roots, a root with fewer empty cells than simulations,
num_simulations = 1, calls out of order, capacities exceeded.  Errors must come back as negative codes with a
message (EngineError in the wrapper), never as a crash or a silent fallback; results on the odd shapes must equal
the oracle's bit for bit."""
import ctypes as C

import numpy as np
import pytest

import oracle as orc

pytestmark = pytest.mark.gpu


def _want(oracle, cells, tm, mc, S, c_puct, plans, difficulty):
    b = orc.board_from_cells(cells, int(tm), int(mc))
    t = oracle.search(b, oracle.params(S, c_puct, 0.2, plans, int(tm), difficulty))
    ch = np.flatnonzero(t["parent"] == 0)
    wv = np.zeros(225, np.int32); wx = np.zeros(225, np.float64)
    wv[t["move"][ch]] = t["visits"][ch]; wx[t["move"][ch]] = t["value"][ch]
    return wv, wx, t["stats"].best_move


def _row(cells, r, c0, n, colour):
    for k in range(n):
        cells[r * 15 + c0 + k] = colour


@pytest.mark.parametrize("precision", [0, 1])
@pytest.mark.parametrize("G", [1, 127, 129, 130])
def test_ragged_batches_match_oracle(weights, oracle, G, precision):
    import engine
    import workload
    eng = engine.Engine(max_games=G, max_simulations=16, precision=precision)
    eng.set_pv_weights(weights["pv"])
    eng.set_planner_weights(weights["graphnet"], weights["dqn"])
    cells, tm, mc = workload.make_positions(G, seed=40 + G)
    # DET-A (no planner): the trees cannot depend on the network precision, every board is checked on both paths
    p = engine.make_params(12, 1.4, beta=0.2, planner_steps=0, difficulty="easy", deterministic=True)
    cv, cval, best = eng.search_host(cells, tm, mc, tm, p)
    for i in sorted({0, G // 2, G - 1, max(0, G - 2)}):
        wv, wx, wb = _want(oracle, cells[i], tm[i], mc[i], 12, 1.4, 0, "easy")
        assert np.array_equal(cv[i], wv) and np.array_equal(cval[i], wx) and best[i] == wb, (G, i)
    # the networks on the ragged tail: last board of the batch equals the same board evaluated alone
    boards = eng.boards_from_cells(cells, tm, mc)
    lg, pr, v = eng.pv_forward(boards)
    lg1, pr1, v1 = eng.pv_forward(boards[G - 1:G].clone())
    tol = 1e-5 if precision == 0 else 2e-2
    assert np.allclose(pr[G - 1].cpu().numpy(), pr1[0].cpu().numpy(), atol=tol)
    assert abs(float(v[G - 1]) - float(v1[0])) <= tol
    assert abs(float(pr[G - 1].sum()) - 1.0) < 1e-3
    eng.close()


@pytest.mark.parametrize("G", [1, 2, 147, 149, 300, 445])
def test_planner_networks_on_ragged_batches(weights, G):
    """k_gn_block0_embed / k_gn_block_pairs are persistent kernels of min(G, 148) CTAs whose per-CTA board count
    decides which prologue and tail paths run (one board: no pipelining at all; two: both embeds in the prologue; three
    and more: the steady state), and consecutive layers walk the boards in opposite directions.  For batch sizes that
    give 1, 2, 3 and 4 boards per CTA (and mixes of them): bf16 GraphNet / DQN outputs within 2e-2 of the fp32 engine's
    (relative to the logits' scale), and every board bit-identical to the same board evaluated in a batch of one."""
    import engine
    import workload
    cells, tm, mc = workload.make_positions(G, seed=90 + G)
    out = []
    for prec in (0, 1):
        e = engine.Engine(max_games=G, max_simulations=8, precision=prec)
        e.set_pv_weights(weights["pv"])
        e.set_planner_weights(weights["graphnet"], weights["dqn"])
        b = e.boards_from_cells(cells, tm, mc)
        gl, gp, q = (t.cpu().numpy() for t in e.planner_nets_forward(b))
        if prec == 1:
            for i in sorted({0, G // 2, G - 1}):
                gl1, gp1, q1 = (t.cpu().numpy() for t in e.planner_nets_forward(b[i:i + 1].clone()))
                assert np.array_equal(gl1[0], gl[i]) and np.array_equal(q1[0], q[i]), (G, i)
        out.append((gl, gp, q))
        e.close()
    (gl0, gp0, q0), (gl1, gp1, q1) = out
    assert np.max(np.abs(gl1 - gl0)) < 2e-2 * max(1.0, np.abs(gl0).max())
    assert np.max(np.abs(gp1 - gp0)) < 2e-2 and np.max(np.abs(q1 - q0)) < 2e-2


def test_terminal_full_and_nearly_full_roots(weights, oracle):
    import engine
    import workload
    eng = engine.Engine(max_games=8, max_simulations=32)
    eng.set_pv_weights(weights["pv"])
    eng.set_planner_weights(weights["graphnet"], weights["dqn"])
    boards = []
    won = np.zeros(225, np.uint8); _row(won, 7, 3, 5, 1); _row(won, 8, 3, 4, 2)
    boards.append((won, 2, 9))                                            # black already has five: finished root
    dense, _, _ = workload.make_positions(1, seed=9, plies=(150,))
    boards.append((dense[0], 1, 150))                                     # 75 empties
    d199, _, _ = workload.make_positions(1, seed=10, plies=(198,))
    boards.append((d199[0], 1, 198))                                      # two plies from the 200-move draw
    d3 = d199[0].copy()
    boards.append((d3, 1, 198))
    one_left = np.where(np.arange(225) == 224, 0, workload._dense_cells(np.random.default_rng(1), 225)).astype(np.uint8)
    boards.append((one_left, 1, 199))                                     # one empty cell, 199 stones
    cells = np.stack([b[0] for b in boards]); tm = np.array([b[1] for b in boards], np.uint8)
    mc = np.array([b[2] for b in boards], np.uint16)
    for S, plans in ((1, 0), (2, 0), (30, 0), (30, 2)):
        p = engine.make_params(S, 1.6, beta=0.2, planner_steps=plans, difficulty="medium", deterministic=True)
        cv, cval, best = eng.search_host(cells, tm, mc, tm, p)
        assert best[0] == -1 and not cv[0].any()                          # ai_agent.py:203-204: no children
        if S == 1:
            assert (best == -1).all() and not cv.any()                    # the first simulation never expands
        for i in range(len(boards)):
            wv, wx, wb = _want(oracle, cells[i], tm[i], mc[i], S, 1.6, plans, "medium")
            assert np.array_equal(cv[i], wv) and np.array_equal(cval[i], wx) and best[i] == wb, (S, plans, i)
    eng.close()


def test_errors_are_codes_with_messages(weights):
    import engine
    import workload
    L = engine.load_library()
    eng = engine.Engine(max_games=4, max_simulations=8)
    cells, tm, mc = workload.make_positions(6, seed=1)
    p = engine.make_params(8, 1.4, deterministic=True)
    with pytest.raises(engine.EngineError, match="gz_set_pv_weights was not called"):
        eng.search_host(cells[:2], tm[:2], mc[:2], tm[:2], p)             # eval_nodes without weights
    eng.set_pv_weights(weights["pv"])
    with pytest.raises(engine.EngineError, match="G out of range"):
        eng.search_host(cells, tm, mc, tm, p)                             # 6 boards, capacity 4
    with pytest.raises(engine.EngineError, match="num_simulations out of range"):
        eng.search_host(cells[:2], tm[:2], mc[:2], tm[:2], engine.make_params(9, 1.4))
    with pytest.raises(engine.EngineError, match="planner weights not set"):
        eng.search_host(cells[:2], tm[:2], mc[:2], tm[:2], engine.make_params(8, 1.4, planner_steps=2))
    with pytest.raises(engine.EngineError):
        eng.get_tree(0)                                                   # gz_get_tree before any gz_search
    with pytest.raises(engine.EngineError, match="gz_selfplay_begin was not called"):
        eng.selfplay_step()
    with pytest.raises(engine.EngineError, match="gz_arena_begin was not called"):
        eng.arena_stats()
    with pytest.raises(engine.EngineError, match="opponent must be"):
        eng.arena_begin(2, p, 0.0, 7, 2)
    bad = np.zeros(10, np.float32)
    assert L.gz_set_pv_weights(eng.h, bad.ctypes.data_as(C.c_void_p), C.c_int64(10)) < 0
    assert b"gz_set_pv_weights" in L.gz_last_error()
    assert L.gz_search(None, None, None, 1, None, None, None, None, None) < 0
    # the handle still works after every refused call
    cv, cval, best = eng.search_host(cells[:4], tm[:4], mc[:4], tm[:4], p)
    assert (cv.sum(1) == 7).all()
    eng.close()
