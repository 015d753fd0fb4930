"""Host-side logic of the drop-in classes that needs no GPU: the opening book of get_move, the preset tables, the
factory, and the checkpoint format against a file written by the reference (tests/golden/ref_checkpoint.pth)."""
import os
import random

import numpy as np
import pytest
import torch

from conftest import GOLDEN, golden, host_board
from weights import make_state_dict


def test_opening_book_matches_reference_without_engine():
    from ai_agent import AlphaZeroGomokuAI
    g = golden("dropin")
    ai = AlphaZeroGomokuAI(1, "easy", planner_steps=0, deterministic=True)
    for i in range(len(g["book_move"])):
        mv = ai.get_move(host_board(g["book_cells"][i], g["book_to_move"][i]))
        assert mv[0] * 15 + mv[1] == g["book_move"][i]
    assert ai._engine is None                      # below six stones nothing is searched (ai_agent.py:120-121 there)
    full = host_board(np.ones(225, np.uint8))
    assert ai.get_move(full) is None               # no empty cell -> None (:117-118)


def test_presets_and_factory_surface():
    from ai_agent import AIFactory, AlphaZeroGomokuAI
    ai = AlphaZeroGomokuAI(2, "nonsense", planner_steps=-3)
    assert ai.params["num_simulations"] == 200 and ai.planner_steps == 0      # unknown -> medium; clamp (:49,92)
    want = {"easy": (100, 1.4, 0.2), "medium": (200, 1.6, 0.05), "hard": (400, 1.8, 0.01)}
    for d, (s, c, e) in want.items():
        p = AlphaZeroGomokuAI(1, d).params
        assert (p["num_simulations"], p["c_puct"], p["exploration"]) == (s, c, e)
        assert set(p) == {"device", "bg_weight", "planning_steps", "num_simulations", "c_puct", "exploration"}
    a, b = AlphaZeroGomokuAI(1, "easy"), AlphaZeroGomokuAI(1, "easy")
    a.params["num_simulations"] = 7                # per-instance dict, as in the reference
    assert b.params["num_simulations"] == 100
    with pytest.raises(ValueError):
        AIFactory.create_ai("nope", 1)


def test_checkpoint_format_equals_reference_file(tmp_path):
    from neural_network import GomokuModel
    ref = torch.load(os.path.join(GOLDEN, "ref_checkpoint.pth"), map_location="cpu")
    m = GomokuModel(device="cpu")
    sd = make_state_dict("pv", 1234)
    m.model.load_state_dict({k: torch.from_numpy(np.asarray(v)) for k, v in sd.items()}, strict=True)
    path = str(tmp_path / "models" / "x.pth")
    m.save_model(path)
    mine = torch.load(path, map_location="cpu")
    assert list(mine) == list(ref) == ["model_state_dict", "model_type", "board_size", "device"]
    assert mine["model_type"] == ref["model_type"] and mine["board_size"] == ref["board_size"] == 15
    assert list(mine["model_state_dict"]) == list(ref["model_state_dict"]) and len(ref["model_state_dict"]) == 45
    for k, t in ref["model_state_dict"].items():
        assert mine["model_state_dict"][k].dtype == t.dtype and torch.equal(mine["model_state_dict"][k], t), k
    m2 = GomokuModel(model_path=os.path.join(GOLDEN, "ref_checkpoint.pth"), device="cpu")
    for k, t in ref["model_state_dict"].items():
        assert torch.equal(m2.model.state_dict()[k], t)
    assert m2.load_model(str(tmp_path / "nope.pth")) is False
    bad = tmp_path / "bad.pth"
    bad.write_bytes(b"not a checkpoint")
    assert m2.load_model(str(bad)) is False        # every load error is a warning (neural_network.py:331-333 there)


def test_bg_planner_consumes_the_random_module_like_the_reference(monkeypatch):
    """No GPU here: the engine call is stubbed; what is checked is the host RNG protocol of get_move."""
    import bg_planner
    pl = bg_planner.BGPlannerAI(1, "easy", device="cpu")

    class FakeEngine:
        def boards_from_cells(self, *a):
            return None

        def planner_move(self, boards, player, p, **k):
            assert p.deterministic == 1
            return {"topk": torch.tensor([[30, 31, 32, 33, 34, 35, 36, 37, -1, -1, -1, -1, -1, -1, -1, -1]]),
                    "move": torch.tensor([33])}
    monkeypatch.setattr(pl, "_get_engine", lambda: FakeEngine())
    b = host_board(np.zeros(225, np.uint8))
    calls = []
    monkeypatch.setattr(random, "random", lambda: (calls.append("random"), 0.5)[1])
    monkeypatch.setattr(random, "choice", lambda s: (calls.append("choice"), s[2])[1])
    assert pl.get_move(b) == (2, 3) and calls == ["random"]            # 0.5 >= explore 0.2: the argmax
    del calls[:]
    monkeypatch.setattr(random, "random", lambda: (calls.append("random"), 0.1)[1])
    assert pl.get_move(b) == (2, 2) and calls == ["random", "choice"]  # explores: random.choice(top_moves)
