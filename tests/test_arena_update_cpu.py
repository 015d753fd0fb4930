"""CPU checks of SURVEY.md 8f rows 3-4 against fixtures made by the reference itself (tests/golden/make_goldens.py
parts `arena` and `update`): the oracle's arena restatement reproduces training.py:221-270 evaluate_model move for
move, and the package's train_epoch / validate_epoch reproduce training.py:277-337 on the same batches."""
import os

import numpy as np
import pytest
import torch

from conftest import golden


def _sub(a):
    """update.npz stores tensors above 4096 elements as every 37th element of the flattened tensor."""
    return a if a.size <= 4096 else a.reshape(-1)[::37]


def _golden_games(z, case):
    sel = np.flatnonzero(z["case_of"] == case)
    return [list(map(int, z["moves"][i][:z["n_moves"][i]])) for i in sel], [int(z["winner"][i]) for i in sel]


def test_oracle_arena_matches_reference_evaluate_model(oracle):
    import oracle as orc
    z = golden("arena")
    for ci, (sims, plans, ng, base, preset) in enumerate(z["cases"]):
        want_moves, want_winner = _golden_games(z, ci)
        sims_a = int(preset) if base else int(sims)      # training.py:240 vs :241-246
        got_moves, got_winner, wld = orc.arena_games(oracle, int(ng), vs_random=not base, sims_a=sims_a,
                                                     sims_b=int(sims), c_puct=1.4, beta=0.2,
                                                     planner_steps=int(plans), difficulty="easy")
        assert got_moves == want_moves, "case %d" % ci
        assert got_winner == want_winner
        assert wld == list(z["result"][ci])


def test_arena_sharding_covers_every_ordinal_once():
    import training
    for games, G, world in ((10, 4, 2), (7, 8, 1), (4096, 512, 8), (5, 1, 3)):
        total = 0
        for rank in range(world):
            a = training.Arena.__new__(training.Arena)
            a.games, a.G, a.rank, a.world = games, G, rank, world
            total += a.my_games()
        assert total == games


def test_update_matches_reference_train_epoch(weights):
    """Same 640 samples, same batches (128, unshuffled), same Adam / clip / StepLR recipe, on CPU with one thread:
    the losses must agree to 1e-6 and every parameter and BN buffer to 1e-6 absolute after two epochs."""
    import neural_network
    import training
    z, d = golden("update"), golden("dataset")
    torch.set_num_threads(1)
    x = torch.from_numpy(d["planes"][:640].astype(np.float32))
    y = torch.from_numpy(d["move"][:640])
    v = torch.from_numpy(d["value"][:640]).reshape(-1, 1)
    cwd = os.getcwd()
    gm = neural_network.GomokuModel(device="cpu")
    assert os.getcwd() == cwd
    if gm.device.type != "cpu":
        gm.model.cpu(); gm.device = torch.device("cpu")
    gm.model.load_state_dict({k: torch.as_tensor(t) for k, t in weights["pv"].items()}, strict=False)
    opt = torch.optim.Adam(gm.model.parameters(), lr=8e-4, weight_decay=1e-5)
    sched = torch.optim.lr_scheduler.StepLR(opt, step_size=2, gamma=0.85)

    def batches(lo, hi):
        return [(x[i:i + 128], y[i:i + 128], v[i:i + 128]) for i in range(lo, hi, 128)]

    tl, vl = [], []
    for ep in range(2):
        tl.append(training.train_epoch(gm, batches(0, 512), opt, torch.device("cpu"), grad_clip=0.8))
        vl.append(training.validate_epoch(gm, batches(512, 640), torch.device("cpu")))
    sched.step()
    assert np.allclose(tl, z["train_loss"], rtol=0, atol=1e-6), (tl, z["train_loss"])
    assert np.allclose(vl, z["val_loss"], rtol=0, atol=1e-6), (vl, z["val_loss"])
    assert np.allclose([g["lr"] for g in opt.param_groups], z["lr"])
    gm.eval_mode()
    with torch.no_grad():
        lg, val = gm.model(x[512:])
    assert np.allclose(lg.numpy(), z["val_logits"], rtol=0, atol=1e-5)
    assert np.allclose(val.numpy().reshape(-1), z["val_value"], rtol=0, atol=1e-5)
    sd = gm.model.state_dict()
    assert len([k for k in z.files if k.startswith("param/")]) == len(sd) == 45
    for k, t in sd.items():
        assert np.allclose(_sub(t.numpy()), z["param/" + k], rtol=0, atol=1e-6), k


def test_simple_replay_from_records_matches_reference_lists():
    import training
    d = golden("dataset")
    n0 = int(d["game_n"][0])
    rep = training.SimpleReplay.from_records(d["records"][:n0])
    assert len(rep) == n0
    assert np.array_equal(np.stack(rep.states).astype(np.uint8), d["planes"][:n0])
    assert rep.move_indices == [int(m) for m in d["move"][:n0]]
    assert rep.outcomes == [int(o) for o in d["value"][:n0]]
    hand = training.SimpleReplay()
    for s, m, p in zip(rep.states, rep.move_indices, rep.players):
        hand.add(s, m, p)
    hand.finalize_with_winner(1)
    assert hand.outcomes == [1 if p == 1 else -1 for p in rep.players]
    hand2 = training.SimpleReplay(); hand2.add(rep.states[0], 0, 2); hand2.finalize_with_winner(None)
    assert hand2.outcomes == [0]


def test_device_loader_feeds_the_same_batches(weights):
    """DeviceLoader over materialized tensors (here on the CPU: the eager path of train_epoch) yields the batches a
    DataLoader(Subset(dataset, indices), 128, shuffle=False) would, ragged tail included, and a seeded shuffle is a
    permutation of the indices."""
    import random
    import neural_network
    import training
    d = golden("dataset")
    torch.set_num_threads(1)

    class Data:
        planes = torch.from_numpy(d["planes"][:300].astype(np.float32))
        move = torch.from_numpy(d["move"][:300])
        value = torch.from_numpy(d["value"][:300])

    holder = training._Materialized(Data)
    idx = list(range(299, -1, -1))
    got = list(training.DeviceLoader(holder, idx, 128, shuffle=False))
    assert [b[0].shape[0] for b in got] == [128, 128, 44] and len(training.DeviceLoader(holder, idx, 128)) == 3
    for k, (x, y, v) in enumerate(got):
        sel = torch.as_tensor(idx[128 * k:128 * (k + 1)])
        assert torch.equal(x, Data.planes[sel]) and torch.equal(y, Data.move[sel])
        assert torch.equal(v, Data.value[sel].reshape(-1, 1))
    sh = training.DeviceLoader(holder, idx, 128, shuffle=True, rng=random.Random(3))
    seen = torch.cat([y for _, y, _ in sh])
    assert sorted(seen.tolist()) == sorted(Data.move.tolist())

    def model():
        gm = neural_network.GomokuModel(device="cpu")
        if gm.device.type != "cpu":
            gm.model.cpu(); gm.device = torch.device("cpu")
        gm.model.load_state_dict({k: torch.as_tensor(t) for k, t in weights["pv"].items()}, strict=False)
        return gm, torch.optim.Adam(gm.model.parameters(), lr=8e-4, weight_decay=1e-5)

    ga, oa = model()
    gb, ob = model()
    la = training.train_epoch(ga, training.DeviceLoader(holder, idx, 128), oa, torch.device("cpu"), grad_clip=0.8)
    lb = training.train_epoch(gb, got, ob, torch.device("cpu"), grad_clip=0.8)
    assert la == lb
