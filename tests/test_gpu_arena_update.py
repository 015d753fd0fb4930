"""SURVEY.md 8f rows 3-4 on the B200: the device-resident evaluation arena (gz_arena_*) against the reference's
evaluate_model fixtures and the oracle's restatement, move for move in deterministic mode; the policy-value update
on the GPU against the reference's train_epoch fixtures; one whole training iteration end to end."""
import numpy as np
import pytest
import torch

from conftest import golden

pytestmark = pytest.mark.gpu


def _sub(a):
    """update.npz stores tensors above 4096 elements as every 37th element of the flattened tensor."""
    return a if a.size <= 4096 else a.reshape(-1)[::37]


def _planner_sd(weights):
    return ({k: torch.as_tensor(v) for k, v in weights["graphnet"].items()},
            {k: torch.as_tensor(v) for k, v in weights["dqn"].items()})


def _model(weights):
    import neural_network
    gm = neural_network.GomokuModel(device="cuda")
    gm.model.load_state_dict({k: torch.as_tensor(t) for k, t in weights["pv"].items()}, strict=False)
    return gm


def _games_from_records(rec):
    import selfplay
    games = selfplay.split_games(rec)
    return {g: [int(p["move"]) for p in plies] for g, plies in games.items()}


def test_arena_matches_reference_evaluate_model(weights):
    """The four evaluate_model runs of tests/golden/arena.npz: every game's move list, the winners and the
    wins / losses / draws dict, bit for bit (fp32 networks, deterministic mode)."""
    import training
    z = golden("arena")
    gm = _model(weights)
    for ci, (sims, plans, ng, base, preset) in enumerate(z["cases"]):
        res, rec = training.evaluate_model(gm, gm if base else None, "cuda", games=int(ng), eval_difficulty="easy",
                                           eval_num_sim=int(sims), eval_plans=int(plans), deterministic=True,
                                           seed=1, precision="f32", keep_records=True, preset_sims=int(preset),
                                           planner_weights=_planner_sd(weights))
        got = _games_from_records(rec)
        sel = np.flatnonzero(z["case_of"] == ci)
        assert sorted(got) == list(range(len(sel)))                     # game_id = ordinal
        for o, i in enumerate(sel):
            assert got[o] == list(map(int, z["moves"][i][:z["n_moves"][i]])), "case %d game %d" % (ci, o)
        w, l, d = (int(v) for v in z["result"][ci])
        assert res == {"wins": w, "losses": l, "draws": d, "win_rate": w / max(1, w + l + d)}


@pytest.mark.parametrize("vs_random,sims_a,sims_b,plans,n", [(True, 6, 6, 1, 10), (False, 9, 4, 0, 10),
                                                             (False, 3, 5, 1, 3)])
def test_arena_matches_oracle_with_recycled_slots(weights, oracle, vs_random, sims_a, sims_b, plans, n):
    """n games in 4 (2) slots (slots are recycled, the last generation is only partly counted), asymmetric budgets.
    The planner case is kept to 3 games: the CPU oracle needs seconds per planner-guided game."""
    import ai_agent
    import oracle as orc
    import training
    ai = ai_agent.AlphaZeroGomokuAI(1, "easy", device="cuda", planner_steps=plans, deterministic=True, seed=3)
    ai.params["num_simulations"] = sims_a
    gs, ds = _planner_sd(weights)
    ai.bg_planner.graph_net.load_state_dict(gs, strict=False)
    ai.bg_planner.opp_dqn.load_state_dict(ds, strict=False)
    ai.model = _model(weights)
    arena = training.Arena(ai, n, vs_random=vs_random, sims_b=sims_b, concurrent=4 if n > 4 else 2)
    st, rec = arena.run(keep_records=True)
    want_moves, want_winner, wld = orc.arena_games(oracle, n, vs_random, sims_a, sims_b, 1.4, 0.2, plans, "easy")
    got = _games_from_records(rec)
    assert sorted(got) == list(range(n))
    for g in range(n):
        assert got[g] == want_moves[g], "game %d" % g
    assert [st["wins"], st["losses"], st["draws"]] == wld and st["games_finished"] == n


def test_stochastic_arena_beats_random_and_counts_add_up():
    """Real Philox randomness, bf16 networks: 64 games against the RandomAgent.  Every record replays through the
    host board; A (whose colour alternates with the ordinal) wins nearly always."""
    import selfplay
    import training
    from gomoku_board import GomokuBoard
    gm = __import__("neural_network").GomokuModel(device="cuda")
    res, rec = training.evaluate_model(gm, None, "cuda", games=64, eval_num_sim=24, eval_plans=1, seed=5,
                                       keep_records=True)
    assert res["wins"] + res["losses"] + res["draws"] == 64
    games = selfplay.split_games(rec)
    assert sorted(games) == list(range(64))
    a_wins = 0
    for g, plies in games.items():
        b = GomokuBoard()
        for p in plies:
            assert int(p["mover"]) == b.current_player
            assert b.make_move(int(p["move"]) // 15, int(p["move"]) % 15)
        assert b.game_over
        side_a = 1 if g % 2 == 0 else 2
        a_wins += b.winner == side_a
        a_plies = [p for p in plies if int(p["mover"]) == side_a]
        first = a_plies[0]["move"]
        assert abs(int(first) // 15 - 7) <= 1 and abs(int(first) % 15 - 7) <= 1     # A opens from the book
    assert a_wins == res["wins"] and res["wins"] >= 48


def test_update_on_gpu_matches_reference_train_epoch(weights):
    """Same batches as the CPU fixture; cuDNN / cuBLAS fp32 against the reference's CPU fp32: losses within 1e-3
    relative, and the UPDATED network as a function -- logits and value on the 128 validation samples -- within 2e-2
    (mean 2e-3) of the reference-trained one.  Weight tensors are only checked on average: Adam moves every
    weight by about lr per step whatever the gradient's size, so where the true gradient is zero (every convolution
    bias in front of a BatchNorm) rounding noise decides the direction of the step; such weights differ by several lr
    without changing the function.  BatchNorm running variances within 5 % (they see the weight differences of every layer below)."""
    import training
    z, d = golden("update"), golden("dataset")
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    x = torch.from_numpy(d["planes"][:640].astype(np.float32)).cuda()
    y = torch.from_numpy(d["move"][:640]).cuda()
    v = torch.from_numpy(d["value"][:640]).reshape(-1, 1).cuda()
    gm = _model(weights)
    opt = torch.optim.Adam(gm.model.parameters(), lr=8e-4, weight_decay=1e-5)

    def batches(lo, hi):
        return [(x[i:i + 128], y[i:i + 128], v[i:i + 128]) for i in range(lo, hi, 128)]

    tl, vl = [], []
    for ep in range(2):
        tl.append(training.train_epoch(gm, batches(0, 512), opt, "cuda", grad_clip=0.8))
        vl.append(training.validate_epoch(gm, batches(512, 640), "cuda"))
    assert np.allclose(tl, z["train_loss"], rtol=1e-3), (tl, z["train_loss"])
    assert np.allclose(vl, z["val_loss"], rtol=1e-3), (vl, z["val_loss"])
    # the updated network as a function: outputs on the validation samples against the reference-trained network
    gm.eval_mode()
    with torch.no_grad():
        lg, val = gm.model(x[512:])
    dl = np.abs(lg.cpu().numpy() - z["val_logits"]); dv = np.abs(val.cpu().numpy().reshape(-1) - z["val_value"])
    assert dl.max() <= 2e-2 and dl.mean() <= 2e-3 and dv.max() <= 2e-2, (dl.max(), dl.mean(), dv.max())
    for k, t in gm.model.state_dict().items():
        got, want = _sub(t.cpu().numpy()), z["param/" + k]
        if "num_batches_tracked" in k:
            assert np.array_equal(got, want)
        elif "running_var" in k:     # running_mean follows the (arbitrary) drift of the bias in front of it
            assert np.allclose(got, want, rtol=5e-2, atol=1e-3), k
        elif k.endswith("weight") and got.ndim > 1:   # biases in front of a BatchNorm are free to drift
            assert np.abs(got - want).mean() <= 1e-3, (k, np.abs(got - want).mean())


def test_one_training_iteration_end_to_end():
    """Self-play -> dataset (device decode, 35 % x 8 symmetries) -> two epochs -> arena: the loss is finite, the
    engine picks up the UPDATED weights (predict changes), every phase is timed."""
    import neural_network
    import training
    gm = neural_network.GomokuModel(device="cuda")
    opt = torch.optim.Adam(gm.model.parameters(), lr=8e-4, weight_decay=1e-5)
    sched = torch.optim.lr_scheduler.StepLR(opt, step_size=2, gamma=0.85)
    cells = np.zeros(225, np.uint8); cells[112] = 1; cells[96] = 2
    before, _ = gm.predict_batch(cells[None])
    out = training.run_iteration(gm, opt, sched, "cuda", games_per_iteration=32, concurrent=32, difficulty="easy",
                                 planner_steps=1, num_simulations=16, eval_games=8, seed=2)
    assert out["games"] >= 32 and out["records"] > 32 * 9
    assert out["samples"] == out["records"] + 8 * max(1, int(out["records"] * 0.35))
    assert np.isfinite(out["train_loss"]) and np.isfinite(out["val_loss"]) and 0 < out["val_loss"] < 20
    assert sum(out["eval"][k] for k in ("wins", "losses", "draws")) == 8
    assert set(out["phases"]) == {"selfplay_s", "gather_s", "dataset_s", "update_s", "eval_s"}
    after, _ = gm.predict_batch(cells[None])
    assert np.abs(after - before).max() > 1e-6


def test_graphed_update_matches_eager(weights):
    """train_epoch as replays of one captured CUDA graph (capturable Adam, DeviceLoader over device tensors) against the
    eager loop on the same samples in the same order.  Short run (2 steps): the updated networks agree as functions to
    5e-3 (mean 5e-4).  Long run (2 epochs x 9 steps incl. a ragged batch): epoch losses agree to 2e-3 -- cuDNN may pick
    other algorithms under capture and Adam amplifies rounding step by step -- and the warm-up steps needed for the
    capture leave no trace in the Adam step counters or in BatchNorm's batch counter."""
    import training
    d = golden("dataset")
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False

    class Data:
        planes = torch.from_numpy(d["planes"][:1100].astype(np.float32)).cuda()
        move = torch.from_numpy(d["move"][:1100]).cuda()
        value = torch.from_numpy(d["value"][:1100]).cuda()

    holder = training._Materialized(Data)

    def run(mode, n, epochs):
        gm = _model(weights)
        nbt0 = int(gm.model.bn.num_batches_tracked)
        opt = torch.optim.Adam(gm.model.parameters(), lr=8e-4, weight_decay=1e-5, capturable=mode == "graphed")
        loader = training.DeviceLoader(holder, list(range(n)), 128, shuffle=False)
        losses = [training.train_epoch(gm, loader, opt, "cuda", grad_clip=0.8) for _ in range(epochs)]
        gm.eval_mode()
        with torch.no_grad():
            lg, val = gm.model(Data.planes[512:768])
        steps = {int(st["step"]) for st in opt.state.values()}
        return losses, lg.cpu().numpy(), val.cpu().numpy(), steps, int(gm.model.bn.num_batches_tracked) - nbt0

    le, lge, ve, se, ne = run("eager", 256, 1)
    lg_, lgg, vg, sg, ng = run("graphed", 256, 1)
    assert se == sg == {2} and ne == ng == 2
    assert np.allclose(le, lg_, rtol=1e-4), (le, lg_)
    dl = np.abs(lge - lgg)
    assert dl.max() <= 5e-3 and dl.mean() <= 5e-4 and np.abs(ve - vg).max() <= 5e-3, (dl.max(), dl.mean())
    le, _, _, se, ne = run("eager", 1100, 2)                  # 8 full batches of 128 + a ragged one of 76, twice
    lg_, _, _, sg, ng = run("graphed", 1100, 2)
    assert se == sg == {18} and ne == ng == 18
    assert np.allclose(le, lg_, rtol=2e-3), (le, lg_)


def test_training_main_follows_the_reference_iteration_order(tmp_path):
    """training.main (training.py:361-537): per iteration self-play -> update -> iteration checkpoint -> best-model
    bookkeeping -> evaluation as `eval_games` single-game calls against the snapshot taken AFTER this iteration's
    update (iteration 1 therefore plays its own fresh snapshot, not the RandomAgent), final checkpoint at the end."""
    import os
    import training
    hist = training.main(iterations=2, games_per_iteration=4, device="cuda", models_dir=str(tmp_path), difficulty="easy",
                         num_simulations=8, planner_steps=1, epochs=1, batch_size=32, eval_games=2)
    assert len(hist) == 2
    for it, res in enumerate(hist, 1):
        assert res["games"] == 4 and res["samples"] >= 8 and np.isfinite(res["val_loss"])
        assert os.path.exists(tmp_path / ("alphazero_gomoku_iter_%d.pth" % it))
        ev = res["eval"]
        assert ev["wins"] + ev["losses"] + ev["draws"] == 2 and 0.0 <= ev["win_rate"] <= 1.0
    assert os.path.exists(tmp_path / "alphazero_gomoku_best.pth") and os.path.exists(tmp_path / "alphazero_gomoku_final.pth")
