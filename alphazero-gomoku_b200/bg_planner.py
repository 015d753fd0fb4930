"""BG-Planner components with the reference's surface (bg_planner.py:22-269 there).

GraphNet / OpponentDQN are torch modules only for their parameters (state-dict keys and default
initialisation match the reference, so weights interchange); their forward passes on the hot path,
the knowledge-search scoring, the stable top-k and the blended pick all run in the CUDA engine
(gz_planner_move / gz_board_ops).
"""
import random

import numpy as np
import torch
import torch.nn as nn
import torch.nn.functional as F

import engine as _engine
from gomoku_board import GomokuBoard


class GraphNet(nn.Module):
    """conv3x3(3->64)+ReLU, 4 x [conv3x3(64->64)+ReLU, conv1x1(64->64)+ReLU], conv1x1(64->2),
    flatten, Linear 450->225."""

    def __init__(self, board_size=15, in_channels=3, hidden_dim=64, num_layers=4):
        super().__init__()
        self.board_size, self.hidden_dim = board_size, hidden_dim
        self.embed = nn.Conv2d(in_channels, hidden_dim, 3, padding=1)
        self.layers = nn.ModuleList(
            nn.Conv2d(hidden_dim, hidden_dim, 3 if i % 2 == 0 else 1, padding=1 if i % 2 == 0 else 0)
            for i in range(2 * num_layers))
        cells = board_size * board_size
        self.policy_head = nn.Sequential(nn.Conv2d(hidden_dim, 2, 1), nn.Flatten(),
                                         nn.Linear(2 * cells, cells))

    def forward(self, x):
        h = F.relu(self.embed(x))
        for layer in self.layers:
            h = F.relu(layer(h))
        return self.policy_head(h)


class OpponentDQN(nn.Module):
    """Linear 675->256, ReLU, 256->256, ReLU, 256->225 on the flattened planes."""

    def __init__(self, board_size=15, in_channels=3, hidden=256):
        super().__init__()
        self.board_size = board_size
        cells = board_size * board_size
        self.net = nn.Sequential(nn.Linear(in_channels * cells, hidden), nn.ReLU(),
                                 nn.Linear(hidden, hidden), nn.ReLU(), nn.Linear(hidden, cells))

    def forward(self, planes):
        return self.net(planes.flatten(1))


def _board_arrays(board):
    return board.to_cells() if hasattr(board, "to_cells") else \
        np.ascontiguousarray(board.board, np.uint8).reshape(225)


class KnowledgeSearch:
    """Heuristic scoring (pattern score, win-now / opponent-wins-next tests, centre bias) served by
    the engine's board and planner kernels."""

    def __init__(self, board_size=15, engine=None):
        self.n = board_size
        self._engine = engine

    def _eng(self):
        if self._engine is None:
            self._engine = _engine.Engine(max_games=8, max_simulations=8)
        return self._engine

    def _device_board(self, board):
        return self._eng().boards_from_cells(_board_arrays(board)[None], [board.current_player],
                                             [board.get_move_count()])

    def _pattern_score(self, board, player):
        ops = self._eng().board_ops(self._device_board(board))
        return float(ops["pattern"][0, player - 1].item())

    def _center_bias(self, move):
        c = self.n // 2
        return max(0.0, (6 - (abs(move[0] - c) + abs(move[1] - c))) * 0.5)

    def _scores2(self, board, player):
        eng = self._eng()
        if not eng_has_planner(eng):
            raise _engine.EngineError("planner weights are not set on this engine")
        p = _engine.make_params(1, 1.0, planner_steps=1, difficulty="hard", deterministic=True)
        r = eng.planner_move(self._device_board(board), player, p)
        return r["scores2"][0].cpu().numpy(), r["topk"][0].cpu().numpy()

    def score_move(self, board, move, player):
        if not board.is_valid_move(move[0], move[1]):
            return -1e9
        s2, _ = self._scores2(board, player)
        return float(s2[move[0] * self.n + move[1]]) / 2.0

    def top_k_moves(self, board, player, k=10):
        s2, _ = self._scores2(board, player)
        moves = board.get_valid_moves()
        order = sorted(range(len(moves)), key=lambda i: -s2[moves[i][0] * self.n + moves[i][1]])
        return [moves[i] for i in order[:k]]          # sorted() is stable: ties keep raster order


def eng_has_planner(eng):
    return getattr(eng, "_planner_set", False)


class BGPlannerAI:
    """GraphNet + OpponentDQN + knowledge search; `player` is the planner's fixed colour."""

    PRESETS = {"easy": {"k": 8, "mix_alpha": 0.5, "explore": 0.2},
               "medium": {"k": 12, "mix_alpha": 0.65, "explore": 0.1},
               "hard": {"k": 16, "mix_alpha": 0.75, "explore": 0.05}}

    def __init__(self, player, difficulty="medium", device="cuda"):
        self.player = player
        self.difficulty = difficulty
        want = "cuda" if device in (None, "cpu") else device
        self.device = torch.device(want if torch.cuda.is_available() else "cpu")
        self.board_size = 15
        self.graph_net = GraphNet(self.board_size).to(self.device).eval()
        self.opp_dqn = OpponentDQN(self.board_size).to(self.device).eval()
        self.params = dict(self.PRESETS[difficulty if difficulty in self.PRESETS else "medium"])
        self._engine = None
        self._packed = None
        self.k_search = KnowledgeSearch(self.board_size)
        self._calls = 0

    def _board_to_planes(self, board_state):
        a = np.asarray(board_state)
        return np.stack([a == 1, a == 2, a == 0]).astype(np.float32)

    def _get_engine(self):
        from neural_network import weights_version
        if self._engine is None:
            self._engine = _engine.Engine(max_games=8, max_simulations=8, device=self.device.index or 0)
            self.k_search._engine = self._engine
        v = (weights_version(self.graph_net), weights_version(self.opp_dqn))
        if v != self._packed:
            self._engine.set_planner_weights(self.graph_net.state_dict(), self.opp_dqn.state_dict())
            self._engine._planner_set = True
            self._packed = v
        return self._engine

    def get_move(self, board):
        """bg_planner.py:232-269.  Knowledge-search top-k, network blend and argmax run in the engine (one
        gz_planner_move call, deterministic pick); the explore roll stays on the host with Python's `random`
        exactly where the reference has it -- one random.random() per call, always consumed, then
        random.choice(top_moves) when it explores -- so a seeded (or monkey-patched) `random` module drives this
        class the way it drives the reference's."""
        valid = board.get_valid_moves()
        if not valid:
            return None
        eng = self._get_engine()
        diff = self.difficulty if self.difficulty in self.PRESETS else "medium"
        p = _engine.make_params(1, 1.0, planner_steps=1, difficulty=diff, deterministic=True)
        p.planner_k, p.planner_alpha, p.planner_explore = \
            self.params["k"], self.params["mix_alpha"], self.params["explore"]
        boards = eng.boards_from_cells(_board_arrays(board)[None], [board.current_player],
                                       [board.get_move_count()])
        self._calls += 1
        r = eng.planner_move(boards, self.player, p)
        top = [int(m) for m in r["topk"][0].cpu().numpy()[:self.params["k"]] if m >= 0]
        if not top:                        # bg_planner.py:239-240
            return random.choice(valid)
        top_moves = [(m // self.board_size, m % self.board_size) for m in top]
        mv = int(r["move"][0].item())
        if random.random() < self.params["explore"]:
            return random.choice(top_moves)
        if mv < 0:
            return random.choice(valid)
        return (mv // self.board_size, mv % self.board_size)
