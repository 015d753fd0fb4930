"""The reference's CLASS surface through the drop-in modules on the B200: AlphaZeroGomokuAI.get_move
(opening book + search, README presets set through the `ai.params[...] = ` idiom of training.py:240), GomokuModel.predict
on both input forms, evaluate_position, MCTSNode.prior_prob (gz_get_priors), AIFactory / BGPlannerAI with the real
Mersenne-Twister stream, `ai.model = other` / in-place weight updates re-packing the engine, time_limit, and a
checkpoint written by the reference's GomokuModel.save_model.  Expected values: tests/golden/dropin.npz, produced by
the UNMODIFIED reference classes (tests/golden/make_goldens.py --parts dropin)."""
import os
import random

import numpy as np
import pytest
import torch

from conftest import GOLDEN, golden, host_board, report
from weights import make_state_dict

pytestmark = pytest.mark.gpu


def _load(module, kind, seed=1234):
    sd = make_state_dict(kind, seed)
    module.load_state_dict({k: torch.from_numpy(np.asarray(v)) for k, v in sd.items()}, strict=True)
    module.eval()


def _ai(player, difficulty, planner_steps, **kw):
    from ai_agent import AlphaZeroGomokuAI
    kw.setdefault("deterministic", True)
    ai = AlphaZeroGomokuAI(player, difficulty, device="cuda", beta=0.2, planner_steps=planner_steps,
                           time_limit=float("inf"), **kw)
    _load(ai.model.model, "pv")
    _load(ai.bg_planner.graph_net, "graphnet")
    _load(ai.bg_planner.opp_dqn, "dqn")
    return ai


@pytest.fixture(scope="module")
def g():
    return golden("dropin")


def test_get_move_opening_book(g):
    ai = _ai(1, "easy", 0)
    for i in range(len(g["book_move"])):
        b = host_board(g["book_cells"][i], g["book_to_move"][i])
        assert b.get_move_count() == g["book_move_count"][i] < 6
        mv = ai.get_move(b)
        assert mv[0] * 15 + mv[1] == g["book_move"][i]
    # with the real RNG the candidate list handed to random.choice is the reference's (ai_agent.py:154-163)
    ai.deterministic = False
    real = random.choice
    try:
        for i in range(len(g["book_move"])):
            seen = []
            random.choice = lambda s: (seen.append(list(s)), s[-1])[1]
            mv = ai.get_move(host_board(g["book_cells"][i], g["book_to_move"][i]))
            cand = np.zeros(225, np.uint8)
            for r, c in (seen[0] if seen else []):
                cand[r * 15 + c] = 1
            assert np.array_equal(cand, g["book_cand"][i]) and bool(seen) == bool(g["book_from_choice"][i])
            if seen:
                assert mv == seen[0][-1]
    finally:
        random.choice = real


def test_get_move_search_readme_presets(g):
    """Configs 1-3 as the README quotes them, set the way training.py:240 does; one AI object per (colour, preset,
    planner_steps) serves several boards and budgets, so params are re-read and the engine re-sized between calls."""
    ais, flips, n = {}, [], len(g["search_move"])
    for i in range(n):
        key = (int(g["search_player"][i]), str(g["search_difficulty"][i]), int(g["search_planner_steps"][i]))
        if key not in ais:
            ais[key] = _ai(*key)
        ai = ais[key]
        ai.params["num_simulations"] = int(g["search_sims"][i])
        ai.params["c_puct"] = float(g["search_c_puct"][i])
        b = host_board(g["search_cells"][i], g["search_to_move"][i])
        assert b.get_move_count() == g["search_move_count"][i]
        before, played = b.board.copy(), ai.games_played
        mv = ai.get_move(b)
        assert np.array_equal(before, b.board) and ai.games_played == played + 1   # ai_agent.py:132 there
        res = ai.search_batch([b])
        same = (mv[0] * 15 + mv[1] == g["search_move"][i]
                and np.array_equal(res["child_visits"][0], g["search_child_visits"][i])
                and np.array_equal(res["child_value"][0], g["search_child_value"][i]))
        if not same:
            flips.append(i)
    report("dropin_get_move", dict(cases=n, differing=flips))
    # planner_steps = 0 is integer + float64 work: exact.  With planner steps the f32 blend may flip on a near-tie.
    assert all(g["search_planner_steps"][i] > 0 for i in flips), flips
    assert len(flips) <= 1, flips


def test_params_mutation_is_reread(g):
    ai = _ai(1, "medium", 0)
    i = int(np.argmax(g["search_move_count"] == 180))
    b = host_board(g["search_cells"][i], g["search_to_move"][i])
    for sims in (50, 8, 120):
        ai.params["num_simulations"] = sims
        cv = ai.search_batch([b])["child_visits"][0]
        assert cv.sum() == sims - 1
        assert ai._engine.counters()["last_search_simulations"] == sims


def test_time_limit_cuts_the_search(g):
    ai = _ai(1, "hard", 0)
    ai.params["num_simulations"] = 200
    i = int(np.argmax(g["search_move_count"] == 190))
    b = host_board(g["search_cells"][i], g["search_to_move"][i])
    ai.time_limit = 1e-7                        # ai_agent.py:183-189: no new simulation once the limit has passed
    mv = ai.get_move(b)
    done = ai._engine.counters()["last_search_simulations"]
    assert 1 <= done < 200 and b.is_valid_move(*mv)
    ai.time_limit = 5.0
    ai.get_move(b)
    assert ai._engine.counters()["last_search_simulations"] == 200


def test_predict_both_input_forms(g):
    from neural_network import GomokuModel
    m = GomokuModel(device="cuda")
    _load(m.model, "pv")
    for i in range(len(g["pred_cells"])):
        b = host_board(g["pred_cells"][i])
        p2, v2 = m.predict(b.get_board_state())              # (15, 15) ints
        p3, v3 = m.predict(b.get_board_tensor())             # (3, 15, 15) planes
        assert p2.shape == (225,) and p2.dtype == np.float32 and isinstance(v2, float)
        assert np.max(np.abs(p2 - g["pred_policy2d_a"][i])) < 1e-4 and abs(v2 - g["pred_value2d_a"][i]) < 1e-4
        assert np.max(np.abs(p3 - g["pred_policy3d_a"][i])) < 1e-4 and abs(v3 - g["pred_value3d_a"][i]) < 1e-4
        vm = b.get_valid_moves()
        mp = m.get_move_probabilities(b.get_board_state(), vm)
        assert np.max(np.abs(mp - g["prior_a"][i][[r * 15 + c for r, c in vm]])) < 1e-4


def test_evaluate_position_and_model_reassignment(g):
    from neural_network import GomokuModel
    boards = [host_board(g["evalpos_cells"][k], g["evalpos_to_move"][k], g["evalpos_game_over"][k],
                         g["evalpos_winner"][k]) for k in range(len(g["evalpos_cells"]))]
    other = GomokuModel(device="cuda")
    _load(other.model, "pv", 4321)
    for pi, player in enumerate((1, 2)):
        ai = _ai(player, "easy", 0)
        for k, b in enumerate(boards):
            want = g["evalpos_a"][pi][k]
            got = ai.evaluate_position(b)
            assert got == want if b.game_over else abs(got - want) < 1e-4     # +-1, and 0.0 for a draw (:456-468)
        ai.model = other                                                          # training.py:247-248
        for k, b in enumerate(boards):
            want = g["evalpos_b"][pi][k]
            got = ai.evaluate_position(b)
            assert got == want if b.game_over else abs(got - want) < 1e-4


@pytest.mark.parametrize("precision,tol", [("f32", 1e-4), ("bf16", 2e-2)])
def test_prior_prob_accessor_and_repack(g, precision, tol):
    """MCTSNode.prior_prob (ai_agent.py:564-582) through gz_get_priors, for the model the AI was built with, after
    `ai.model = other`, and after an in-place load_state_dict on the same module (weights_version must notice)."""
    from neural_network import GomokuModel
    ai = _ai(1, "easy", 0, store_priors=True, precision=precision)
    ai.params["num_simulations"] = 4
    other = GomokuModel(device="cuda")
    _load(other.model, "pv", 4321)
    worst = 0.0
    for tag, setter in (("a", lambda: None), ("b", lambda: setattr(ai, "model", other)),
                        ("a", lambda: _load(other.model, "pv", 1234))):
        setter()
        for i in range(len(g["pred_cells"])):
            b = host_board(g["pred_cells"][i])
            ai.search_batch([b])
            got = ai.root_prior_prob(b)
            want = g["prior_" + tag][i][[r * 15 + c for r, c in b.get_valid_moves()]]
            assert got.shape == want.shape and abs(got.sum() - 1.0) < 1e-5
            worst = max(worst, float(np.max(np.abs(got - want))))
            assert np.max(np.abs(got - want)) < tol, (tag, i)
    report("dropin_priors_" + precision, dict(max_abs_err=worst, tol=tol))


def test_prior_of_inner_nodes_and_errors(weights):
    """gz_get_priors on nodes below the root: node j's prior must equal the root prior of a search started from node
    j's own board; terminal nodes and handles without store_priors are errors, not zeros."""
    import engine
    g = golden("dropin")
    e = engine.Engine(max_games=4, max_simulations=16, store_priors=True)
    e.set_pv_weights(weights["pv"])
    cells = g["pred_cells"][3]
    stones = int((cells != 0).sum())
    boards = e.boards_from_cells(cells[None], [1 + stones % 2], [stones])
    p = engine.make_params(6, 1.4, deterministic=True)
    e.search(boards, [1 + stones % 2], p)
    t = e.get_tree(0)
    inner = {}
    for j in range(1, len(t["move"])):
        inner[j] = (int(t["move"][j]), e.get_priors(0, j))
    for j, (mv, (pr, v)) in inner.items():
        c2 = cells.copy()
        c2[mv] = 1 + stones % 2
        b2 = e.boards_from_cells(c2[None], [2 - stones % 2], [stones + 1])
        e.search(b2, [1], p)
        pr0, v0 = e.get_priors(0, 0)
        assert np.array_equal(pr, pr0) and v == v0 and pr[mv] == 0.0
    with pytest.raises(engine.EngineError):
        e.get_priors(0, 99)
    e2 = engine.Engine(max_games=4, max_simulations=16)
    e2.set_pv_weights(weights["pv"])
    e2.search(boards.clone(), [1], engine.make_params(4, 1.4, deterministic=True, eval_nodes=False))
    with pytest.raises(engine.EngineError):
        e2.get_priors(0, 0)
    e.close(); e2.close()


def test_reference_checkpoint_interchange(g, tmp_path):
    """A checkpoint written by the reference's GomokuModel.save_model (neural_network.py:292-306) loads into the
    drop-in and predicts the reference's outputs; what the drop-in saves has the same dictionary layout."""
    from neural_network import GomokuModel
    ref_ck = os.path.join(GOLDEN, "ref_checkpoint.pth")
    m = GomokuModel(model_path=ref_ck, device="cuda")
    b = host_board(g["pred_cells"][0])
    p, v = m.predict(b.get_board_state())
    assert np.max(np.abs(p - g["pred_policy2d_a"][0])) < 1e-4 and abs(v - g["pred_value2d_a"][0]) < 1e-4
    out = str(tmp_path / "models" / "again.pth")
    m.save_model(out)
    a, r = torch.load(out, map_location="cpu"), torch.load(ref_ck, map_location="cpu")
    assert list(a) == list(r) and a["model_type"] == r["model_type"] and a["board_size"] == r["board_size"]
    assert list(a["model_state_dict"]) == list(r["model_state_dict"])
    for k in a["model_state_dict"]:
        assert torch.equal(a["model_state_dict"][k].cpu(), r["model_state_dict"][k]), k
    m2 = GomokuModel(device="cuda")
    assert m2.load_model(out) is True and m2.load_model(str(tmp_path / "missing.pth")) is False
    p2, v2 = m2.predict(b.get_board_state())
    assert np.array_equal(p, p2) and v == v2
    assert float(g["ref_loads_dropin_checkpoint"].max()) == 0.0    # the other direction, checked at generation time


def test_ai_factory_and_bg_planner_stream(g):
    """AIFactory.create_ai (ai_agent.py:585-623) and BGPlannerAI.get_move (bg_planner.py:232-269) driven by the real
    `random` module: one random.random() per call, random.choice(top_moves) when it explores -- the same
    Mersenne-Twister stream as the reference, so seeded move sequences are the reference's."""
    from ai_agent import AIFactory, AlphaZeroGomokuAI
    from bg_planner import BGPlannerAI
    assert AIFactory.get_available_ai_types() == ["alphazero", "bg_planner"]
    assert AIFactory.get_available_difficulties() == ["easy", "medium", "hard"]
    with pytest.raises(ValueError):
        AIFactory.create_ai("minimax", 1)
    az = AIFactory.create_ai("alphazero", 2, "hard", device="cuda", planner_steps=0)
    assert type(az) is AlphaZeroGomokuAI and az.player == 2 and az.name == "AlphaZero_hard"
    assert az.params["num_simulations"] == 400 and az.params["c_puct"] == 1.8     # the code presets (ai_agent.py:84-90)
    bad = []
    for i in range(len(g["planner_moves"])):
        pl = AIFactory.create_ai("bg_planner", int(g["planner_player"][i]), str(g["planner_difficulty"][i]),
                                 device="cuda")
        assert type(pl) is BGPlannerAI
        _load(pl.graph_net, "graphnet")
        _load(pl.opp_dqn, "dqn")
        b = host_board(g["planner_cells"][i], g["planner_to_move"][i])
        random.seed(int(g["planner_seed"][i]))
        moves = [pl.get_move(b) for _ in range(g["planner_moves"].shape[1])]
        moves = [r * 15 + c for r, c in moves]
        if moves != list(g["planner_moves"][i]):
            bad.append(i)
    report("dropin_planner_stream", dict(streams=len(g["planner_moves"]), differing=bad))
    assert len(bad) <= 1, bad          # a near-tie of the f32 blend may flip one stream's argmax
