"""Device-resident self-play (gz_selfplay_*) on the B200 against the reference's play_one_game /
SimpleReplay semantics (training.py:141-218, :77-97; ai_agent.py:109-166): every record must replay
through the host GomokuBoard, outcome labels must follow the winner, book plies must follow the opening
book, and in deterministic mode the searched plies must equal the oracle's search answer."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def session(weights):
    import torch
    import ai_agent
    import selfplay
    torch.manual_seed(7)
    ai = ai_agent.AlphaZeroGomokuAI(1, "easy", device="cuda", beta=0.2, planner_steps=1, deterministic=True,
                                    seed=11)
    ai.params["num_simulations"] = 12
    ai.params["c_puct"] = 1.2
    ai.model.model.load_state_dict({k: torch.as_tensor(v) for k, v in weights["pv"].items()}, strict=False)
    ai.bg_planner.graph_net.load_state_dict({k: torch.as_tensor(v) for k, v in weights["graphnet"].items()}, strict=False)
    ai.bg_planner.opp_dqn.load_state_dict({k: torch.as_tensor(v) for k, v in weights["dqn"].items()}, strict=False)
    sp = selfplay.SelfPlay(ai, 48, rank=1, world=2, seed=5)
    return ai, sp


def _replay_game(plies):
    from gomoku_board import GomokuBoard
    import selfplay
    b = GomokuBoard()
    rec = np.array(plies, dtype=selfplay.REPLAY_DTYPE)
    planes, moves, movers, outcomes = selfplay.records_to_samples(rec)
    for i in range(len(rec)):
        assert rec[i]["ply"] == i == b.get_move_count()
        assert np.array_equal(planes[i], b.get_board_tensor())       # state BEFORE the move
        assert movers[i] == b.current_player
        assert not b.game_over
        assert b.make_move(int(moves[i]) // 15, int(moves[i]) % 15)
    return b, outcomes, movers


def test_deterministic_selfplay_games_replay_and_match_oracle(session, oracle, weights):
    import oracle as orc
    ai, sp = session
    rec = sp.play_games(60, max_steps=400)
    st = sp.stats()
    assert st["games_finished"] >= 60 and st["records_dropped"] == 0
    assert st["black_wins"] + st["white_wins"] + st["draws"] == st["games_finished"]
    games = __import__("selfplay").split_games(rec)
    assert len(games) == st["games_finished"]
    assert all(48 <= (g & 0xFFFFF) < 96 for g in games)                # rank 1 of 2: slots 48..95
    assert {g >> 20 for g in games} >= {0, 1}                          # recycled slots move up by 2^20
    checked = 0
    for gid, plies in sorted(games.items()):
        b, outcomes, movers = _replay_game(plies)
        assert b.game_over and len(plies) == b.get_move_count()
        want = [0 if b.winner is None else (1 if m == b.winner else -1) for m in movers]
        assert list(outcomes) == want
        # opening book: (7,7), then centre 3x3 in raster order (deterministic mode: first element)
        assert plies[0]["move"] == 112
        assert all(abs(int(p["move"]) // 15 - 7) <= 1 and abs(int(p["move"]) % 15 - 7) <= 1 for p in plies[1:6])
        if checked < 6:                                                # searched plies vs the oracle
            for p in plies[6:9]:
                cells = np.zeros(225, np.uint8)
                planes, _, _, _ = __import__("selfplay").records_to_samples(np.array([p], dtype=rec.dtype))
                cells[planes[0, 0].reshape(-1) > 0] = 1
                cells[planes[0, 1].reshape(-1) > 0] = 2
                ob = orc.board_from_cells(cells, int(p["mover"]), int(p["ply"]))
                t = oracle.search(ob, oracle.params(12, 1.2, 0.2, 1, int(p["mover"]), "easy"))
                assert t["stats"].best_move == int(p["move"])
            checked += 1
    assert checked == 6


def test_stochastic_selfplay_statistics(weights):
    import torch
    import ai_agent
    import selfplay
    ai = ai_agent.AlphaZeroGomokuAI(1, "easy", device="cuda", planner_steps=0, seed=3, precision="bf16")
    ai.params["num_simulations"] = 8
    ai.params["exploration"] = 0.5
    sp = selfplay.SelfPlay(ai, 256, seed=9)
    rec = sp.play_games(300, max_steps=400)
    games = selfplay.split_games(rec)
    firsts = set()
    for gid, plies in games.items():
        b, outcomes, movers = _replay_game(plies)
        assert b.game_over
        firsts.add(tuple(int(p["move"]) for p in plies[1:6]))
    assert len(firsts) > 50                                            # book plies are randomised per game
    # epsilon = 0.5: about half of the searched plies are uniform random moves, so games differ
    assert len({tuple(int(p["move"]) for p in plies[6:12]) for plies in games.values()}) > 100


def test_games_total_plays_exactly_the_first_n_starts(weights):
    """gz_selfplay_begin(games_total = N): the N games with start ordinal generation * G + slot < N are played to the
    end and recorded -- the reference's `for game in range(N): play_one_game` (training.py:415-420) -- whatever their
    length; games started later only keep their slot busy (an open-ended session stopped at N finishers would keep
    whichever games ended first)."""
    import ai_agent
    import selfplay
    G, N = 64, 160
    ai = ai_agent.AlphaZeroGomokuAI(1, "easy", device="cuda", planner_steps=0, seed=21, precision="bf16")
    ai.params["num_simulations"] = 8
    sp = selfplay.SelfPlay(ai, G, seed=4, games_total=N)
    rec = sp.play_games(max_steps=2000)
    st = sp.stats()
    assert st["games_finished"] == N and st["records_dropped"] == 0
    games = selfplay.split_games(rec)
    want = {slot + (gen << 20) for gen in range(3) for slot in range(G) if gen * G + slot < N}
    assert set(games) == want and len(want) == N
    for plies in games.values():
        b, _, _ = _replay_game(plies)
        assert b.game_over
    assert st["plies_played"] == len(rec) == sum(len(p) for p in games.values())
    sp.step(3)                                              # later games are neither counted nor recorded
    assert sp.stats()["games_finished"] == N and len(sp.drain()) == 0
