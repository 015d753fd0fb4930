"""GPU parity tests (run on the B200 box): every kernel family through the C-ABI against the golden
fixtures (produced by the unmodified Python reference) and against the oracle on seeded inputs.
Integer / index / float64-tree results must be bit-exact; network outputs within 1e-4 (fp32 mode).
"""
import numpy as np
import pytest

import oracle as orc
from conftest import golden

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def eng(weights):
    import engine
    e = engine.Engine(max_games=512, max_simulations=400, store_priors=True)
    e.set_pv_weights(weights["pv"])
    e.set_planner_weights(weights["graphnet"], weights["dqn"])
    yield e
    e.close()


def raster_bits(words):
    w = np.ascontiguousarray(words).view(np.uint32).reshape(-1, 8)
    return np.unpackbits(w.view(np.uint8).reshape(len(w), 32), axis=1, bitorder="little")[:, :225]


def status_of(g, i):
    return 0 if not g["game_over"][i] else (3 if g["winner"][i] == 0 else int(g["winner"][i]))


def test_board_import_and_ops(eng):
    g = golden("boards")
    boards = eng.boards_from_cells(g["cells"], g["to_move"], g["move_count"])
    assert np.array_equal(eng.boards_to_cells(boards).cpu().numpy(), g["cells"])
    mc, tm, st = eng.board_meta(boards)
    n = len(g["cells"])
    assert np.array_equal(mc, g["move_count"]) and np.array_equal(tm, g["to_move"])
    assert list(st) == [status_of(g, i) for i in range(n)]
    ops = {k: v.cpu().numpy() for k, v in eng.board_ops(boards).items()}
    assert np.array_equal(raster_bits(ops["legal"]), g["legal"])
    assert np.array_equal(ops["threat"], g["threat"])
    assert np.array_equal(ops["pattern"], g["pattern"].astype(np.int64))
    live = np.flatnonzero(g["game_over"] == 0)
    win = raster_bits(ops["win"].reshape(-1, 8)).reshape(n, 2, 225)
    three = raster_bits(ops["three"].reshape(-1, 8)).reshape(n, 2, 225)
    assert np.array_equal(win[live], g["win_cells"][live])
    assert np.array_equal(three[live], g["three_cells"][live])


def test_make_move_trajectories(eng):
    g = golden("boards")
    n = len(g["moves"])
    boards = eng.boards_from_cells(np.zeros((n, 225), np.uint8), np.ones(n), np.zeros(n))
    for ply in range(g["moves"].shape[1]):
        mv = g["moves"][:, ply].astype(np.int32)
        if np.all(mv < 0):
            break
        ok = eng.make_moves(boards, mv).cpu().numpy()
        assert np.array_equal(ok.astype(bool), mv >= 0)   # every recorded move was legal
    assert np.array_equal(eng.boards_to_cells(boards).cpu().numpy(), g["cells"])
    mc, tm, st = eng.board_meta(boards)
    assert np.array_equal(mc, g["move_count"]) and np.array_equal(tm, g["to_move"])
    assert list(st) == [status_of(g, i) for i in range(n)]


def test_rollout_policy_candidates(eng):
    import engine
    g = golden("tiers")
    boards = eng.boards_from_cells(g["cells"], g["to_move"], g["move_count"])
    p = engine.make_params(1, 1.8, deterministic=True)
    mv, cand = eng.rollout_policy(boards, p)
    assert np.array_equal(mv.cpu().numpy(), g["move"])
    assert np.array_equal(raster_bits(cand.cpu().numpy()), g["cand"])


@pytest.mark.parametrize("name", ["rollout", "rollout_planner"])
def test_det_rollouts_vs_reference(eng, name):
    import engine
    g = golden(name)
    boards = eng.boards_from_cells(g["start_cells"], g["to_move"], g["move_count"])
    for ps in sorted(set(g["planner_steps"])):
        sel = np.flatnonzero(g["planner_steps"] == ps)
        p = engine.make_params(1, 1.8, planner_steps=int(ps), difficulty="hard", deterministic=True)
        v, steps, final = eng.rollout(boards[sel], g["ai_player"][sel], p)
        fin = eng.boards_to_cells(final).cpu().numpy()
        bad = [int(i) for k, i in enumerate(sel) if not (
            v[k].item() == g["value"][i] and np.array_equal(fin[k], g["final_cells"][i]))]
        # the planner's blended f32 pick may flip on a near-tie; DET-A must be exact
        assert len(bad) <= (0 if ps == 0 else 1), bad
        if not bad:
            assert np.array_equal(steps.cpu().numpy(), g["steps"][sel])


def test_philox_rollouts_vs_oracle(eng, oracle):
    """Stochastic playouts: bit-exact against the oracle, which shares the Philox keying."""
    import engine
    g = golden("tiers")
    n = len(g["cells"])
    boards = eng.boards_from_cells(g["cells"], g["to_move"], g["move_count"])
    seed, base, sim = 0xC0FFEE1234, 1000, 3
    p = engine.make_params(1, 1.8, deterministic=False, seed=seed, game_base=base)
    v, steps, final = eng.rollout(boards, g["to_move"], p, sim=sim)
    fin = eng.boards_to_cells(final).cpu().numpy()
    v = v.cpu().numpy()
    for i in range(0, n, 2):
        b = orc.board_from_cells(g["cells"][i], g["to_move"][i], g["move_count"][i])
        rng = orc.Rng(1, seed & 0xFFFFFFFF, seed >> 32, base + i, sim)
        v0, fb, _ = oracle.simulate(b, oracle.params(1, 1.8, 0.2, 0, int(g["to_move"][i])), rng)
        assert v[i] == v0 and np.array_equal(fin[i], orc.cells_of(fb))


def test_stochastic_outcome_rates_vs_reference(eng):
    """Outcome frequencies of real-RNG playouts: Philox engine vs Mersenne-Twister reference,
    two-sample binomial bound with z = 3.9."""
    import engine
    g = golden("stoch")
    n_e = 512
    for i in range(len(g["cells"])):
        cells = np.repeat(g["cells"][i][None], n_e, 0)
        boards = eng.boards_from_cells(cells, np.full(n_e, g["to_move"][i]), np.full(n_e, g["move_count"][i]))
        p = engine.make_params(1, 1.8, deterministic=False, seed=77 + i)
        v, steps, _ = eng.rollout(boards, int(g["to_move"][i]), p)
        v = v.cpu().numpy()
        n_r = int(g["n"][i])
        for k, val in enumerate((1.0, -1.0, 0.1, 0.0)):
            pe, pr = np.mean(v == val), g["counts"][i][k] / n_r
            pool = (pe * n_e + pr * n_r) / (n_e + n_r)
            bound = 3.9 * np.sqrt(max(pool * (1 - pool), 1e-4) * (1 / n_e + 1 / n_r))
            assert abs(pe - pr) <= bound, (i, val, pe, pr, bound)
        ms_e, ms_r = steps.float().mean().item(), g["total_steps"][i] / n_r
        assert abs(ms_e - ms_r) <= 0.25 * max(ms_r, 4.0)


def test_networks_fp32(eng):
    g = golden("nets")
    n = len(g["cells"])
    stones = (g["cells"] != 0).sum(1)
    boards = eng.boards_from_cells(g["cells"], 1 + stones % 2, stones)
    lg, pr, v = (t.cpu().numpy() for t in eng.pv_forward(boards))
    assert np.max(np.abs(lg - g["logits"])) < 1e-4          # north_star: 1e-4 in fp32
    assert np.max(np.abs(pr - g["probs"])) < 1e-4
    assert np.max(np.abs(v - g["value"])) < 1e-4
    gl, gp, q = (t.cpu().numpy() for t in eng.planner_nets_forward(boards))
    assert np.max(np.abs(gl - g["gn_logits"])) < 1e-4
    assert np.max(np.abs(q - g["dqn_q"])) < 1e-4
    assert n == len(lg)


def test_planner_scores_topk_move(eng):
    import engine
    g = golden("planner")
    boards = eng.boards_from_cells(g["cells"], g["to_move"], g["move_count"])
    flips = 0
    for diff in ("easy", "medium", "hard"):
        sel = np.flatnonzero(g["difficulty"] == diff)
        p = engine.make_params(1, 1.8, planner_steps=5, difficulty=diff, deterministic=True)
        r = {k: v.cpu().numpy() for k, v in eng.planner_move(boards[sel], g["ai_player"][sel], p).items()}
        for k, i in enumerate(sel):
            want = g["scores"][i]
            occ = np.isnan(want)
            assert np.array_equal(r["scores2"][k][~occ], np.round(want[~occ] * 2).astype(np.int64))
            kk = int(g["k"][i])
            assert np.array_equal(r["topk"][k][:kk], g["topk"][i][:kk])
            ok = ~np.isnan(g["composed"][i][:kk])
            assert np.allclose(r["composed"][k][:kk][ok], g["composed"][i][:kk][ok], rtol=0, atol=1e-4)
            if r["move"][k] != g["move"][i]:
                srt = np.sort(g["composed"][i][:kk][ok])
                assert srt[-1] - srt[-2] < 2e-4
                flips += 1
    assert flips <= 2


def _search_group(eng, g, idx):
    import engine
    i0 = idx[0]
    p = engine.make_params(int(g["sims"][i0]), float(g["c_puct"][i0]), beta=float(g["beta"][i0]),
                           planner_steps=int(g["planner_steps"][i0]), difficulty=str(g["difficulty"][i0]),
                           deterministic=True)
    boards = eng.boards_from_cells(g["cells"][idx], g["to_move"][idx], g["move_count"][idx])
    cv, cval, best = eng.search(boards, g["ai_player"][idx], p)
    bad = []
    for k, i in enumerate(idx):
        t = eng.get_tree(k)
        n = int(g["n_nodes"][i])
        same = (len(t["parent"]) == n and np.array_equal(t["parent"], g["parent"][i][:n])
                and np.array_equal(t["move"], g["move"][i][:n])
                and np.array_equal(t["visits"], g["visits"][i][:n])
                and np.array_equal(t["value"], g["value"][i][:n])
                and best[k].item() == g["best_move"][i])
        if same:   # root-children view must agree with the tree
            ch = np.flatnonzero(g["parent"][i][:n] == 0)
            assert np.array_equal(cv[k].cpu().numpy()[g["move"][i][ch]], g["visits"][i][ch])
            assert np.array_equal(cval[k].cpu().numpy()[g["move"][i][ch]], g["value"][i][ch])
        else:
            bad.append(int(i))
    return bad


def _groups(g):
    keys = {}
    for i in range(len(g["sims"])):
        k = (int(g["sims"][i]), float(g["c_puct"][i]), float(g["beta"][i]), int(g["planner_steps"][i]),
             str(g["difficulty"][i]))
        keys.setdefault(k, []).append(i)
    return [np.array(v) for v in keys.values()]


def test_mcts_det_a_whole_trees_vs_reference(eng):
    """Every node of every tree (parent, move, visits, float64 value sum): bit-exact."""
    g = golden("mcts")
    bad = sum((_search_group(eng, g, idx) for idx in _groups(g)), [])
    assert bad == []


def test_mcts_det_b_whole_trees_vs_reference(eng):
    g = golden("mcts_planner")
    bad = sum((_search_group(eng, g, idx) for idx in _groups(g)), [])
    assert len(bad) <= 1, bad


def test_search_host_matches_device_and_oracle(eng, oracle):
    """The host-buffer entry point (behind get_move) on the config-2 shape, vs the oracle."""
    import engine
    g = golden("tiers")
    sel = np.arange(0, len(g["cells"]), 4)
    p = engine.make_params(100, 1.6, beta=0.2, difficulty="medium", deterministic=True)
    cv, cval, best = eng.search_host(g["cells"][sel], g["to_move"][sel], g["move_count"][sel],
                                     g["to_move"][sel], p)
    for k, i in enumerate(sel[:24]):
        b = orc.board_from_cells(g["cells"][i], g["to_move"][i], g["move_count"][i])
        t = oracle.search(b, oracle.params(100, 1.6, 0.2, 0, int(g["to_move"][i]), "medium"))
        ch = np.flatnonzero(t["parent"] == 0)
        want_v = np.zeros(225, np.int32); want_x = np.zeros(225, np.float64)
        want_v[t["move"][ch]] = t["visits"][ch]; want_x[t["move"][ch]] = t["value"][ch]
        assert np.array_equal(cv[k], want_v) and np.array_equal(cval[k], want_x)
        assert best[k] == t["stats"].best_move
    c = eng.counters()
    assert c["kernel_launches"] > 0 and c["simulations"] > 0
