import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, "tests", "golden")
PKG = os.path.join(ROOT, "alphazero-gomoku_b200")
for p in (ROOT, PKG, GOLDEN, os.path.join(ROOT, "oracle")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def pytest_collection_modifyitems(config, items):
    try:
        import torch
        has_gpu = torch.cuda.is_available()
    except Exception:
        has_gpu = False
    if has_gpu:
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for it in items:
        if "gpu" in it.keywords:
            it.add_marker(skip)


def golden(name):
    return np.load(os.path.join(GOLDEN, name + ".npz"))


def report(name, payload):
    """Measured counts a test wants on record (flip counts, error maxima): gpurun_out/test_reports/<name>.json comes
    back from the GPU box with the run; the summaries quoted in DESIGN.md / profiles/ are copied from there."""
    import json
    d = os.path.join(ROOT, "gpurun_out", "test_reports")
    try:
        os.makedirs(d, exist_ok=True)
        with open(os.path.join(d, name + ".json"), "w") as f:
            json.dump(payload, f, indent=1, default=lambda o: o.tolist() if hasattr(o, "tolist") else str(o))
    except OSError:
        pass


def host_board(cells, to_move=None, game_over=False, winner=None):
    """A drop-in GomokuBoard holding `cells` (uint8[225]); the history order is arbitrary (only its length and the
    stones are read by the hot path)."""
    from gomoku_board import GomokuBoard
    b = GomokuBoard()
    c = np.asarray(cells).reshape(15, 15)
    b.board = c.astype(int)
    blk = [(int(r), int(q), 1) for r, q in zip(*np.nonzero(c == 1))]
    wht = [(int(r), int(q), 2) for r, q in zip(*np.nonzero(c == 2))]
    hist = []
    for i in range(max(len(blk), len(wht))):
        if i < len(blk):
            hist.append(blk[i])
        if i < len(wht):
            hist.append(wht[i])
    b.move_history = hist
    b.current_player = int(to_move) if to_move is not None else 1 + len(hist) % 2
    b.game_over = bool(game_over)
    b.winner = None if not winner else int(winner)
    return b


@pytest.fixture(scope="session")
def weights():
    from weights import make_state_dict
    return dict(pv=make_state_dict("pv", 1234), graphnet=make_state_dict("graphnet", 1234),
                dqn=make_state_dict("dqn", 1234))


@pytest.fixture(scope="session")
def oracle(weights):
    import oracle as orc
    return orc.Oracle(**weights)


_BLEND_ERR = {}


def bf16_blend_error(weights, G=4096):
    """Largest |composed_bf16 - composed_fp32| of the planner's blended candidate score (bg_planner.py:256-264) over
    the <= 16 candidates of G workload positions, both measured by the engine; the fp32 side is pinned to the
    reference within 1e-4 by test_planner_scores_topk_move.  Returns (err, fp32 result, bf16 result)."""
    if "v" not in _BLEND_ERR:
        import engine
        import workload
        cells, tm, mc = workload.make_positions(G, seed=3)
        p = engine.make_params(1, 1.8, planner_steps=5, difficulty="hard", deterministic=True)
        out = []
        for prec in (0, 1):
            e = engine.Engine(max_games=G, max_simulations=8, precision=prec)
            e.set_pv_weights(weights["pv"])
            e.set_planner_weights(weights["graphnet"], weights["dqn"])
            r = e.planner_move(e.boards_from_cells(cells, tm, mc), tm, p)
            out.append({k: v.cpu().numpy() for k, v in r.items()})
            e.close()
        valid = out[0]["topk"] >= 0
        err = float(np.max(np.abs(out[0]["composed"] - out[1]["composed"])[valid]))
        _BLEND_ERR["v"] = (err, out[0], out[1])
    return _BLEND_ERR["v"]
