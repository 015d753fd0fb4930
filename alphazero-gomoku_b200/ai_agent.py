"""AlphaZeroGomokuAI / AIFactory with the reference's surface (ai_agent.py:23-623 there), backed by
the B200 engine.

`get_move(board)` keeps the reference's host-side dispatcher (opening book for the first six stones,
epsilon-random override, counters) and runs the search itself -- select / expand / policy-value
evaluation / planner-guided playout / backup -- as one `gz_search_host` call on the GPU.
`get_moves(boards)` is the batched form the self-play driver uses (thousands of positions per call).

Reference behaviours kept on purpose: UCB1-style score with beta * BG bonus, no
negamax, one node per simulation in reverse raster order, `params` re-read at every call, `beta` /
`planner_steps` taken from the constructor (not from `params["bg_weight"]` / `["planning_steps"]`),
`model` assignable.  The wall-clock `time_limit` is handed to the engine (gz_search_params.time_limit_s): like
the reference the search stops starting new simulations once the limit has passed (checked every four waves);
parity runs use `float("inf")`.
"""
import logging
import random
import time

import numpy as np
import torch

import engine as _engine
from bg_planner import BGPlannerAI, KnowledgeSearch
from gomoku_board import GomokuBoard
from neural_network import GomokuModel, weights_version


class AlphaZeroGomokuAI:
    DIFFICULTY_PARAMS = {
        "easy": {"device": "cpu", "bg_weight": 0.1, "planning_steps": 2, "num_simulations": 100,
                 "c_puct": 1.4, "exploration": 0.2},
        "medium": {"device": "cpu", "bg_weight": 0.2, "planning_steps": 3, "num_simulations": 200,
                   "c_puct": 1.6, "exploration": 0.05},
        "hard": {"device": "cpu", "bg_weight": 0.25, "planning_steps": 4, "num_simulations": 400,
                 "c_puct": 1.8, "exploration": 0.01},
    }

    def __init__(self, player, difficulty="medium", model_path=None, device=None, beta=0.2,
                 planner_steps=5, time_limit=5.0, time_reward_factor=0.1, deterministic=False,
                 seed=None, precision="f32", store_priors=False):
        self.player = player
        self.difficulty = difficulty
        self.name = "AlphaZero_%s" % difficulty
        want = "cuda" if device in (None, "cpu") else device
        self.device = torch.device(want if torch.cuda.is_available() else "cpu")
        self.beta = float(beta)
        self.planner_steps = int(max(0, planner_steps))
        self.time_limit = float(time_limit)
        self.time_reward_factor = float(time_reward_factor)
        self.logger = logging.getLogger(__name__)
        self.model = GomokuModel(model_path=model_path, device=str(self.device))
        self.bg_planner = BGPlannerAI(player=self.player, difficulty=self.difficulty,
                                      device=str(self.device))
        self.knowledge_search = KnowledgeSearch(board_size=15)
        self.difficulty_params = {k: dict(v) for k, v in self.DIFFICULTY_PARAMS.items()}
        self.params = self.difficulty_params.get(difficulty, self.difficulty_params["medium"])
        self.mcts_root = None
        self.mcts_cache = {}
        self.auto_save_interval = None
        self.games_played = 0
        self._last_decision_time = None
        # engine-side extensions (not in the reference): deterministic mode and a Philox seed
        # precision of the networks inside the engine: "f32" (FFMA, 1e-4 against torch) or "bf16"
        # (tcgen05 tensor cores, 2e-2): the throughput mode for batched self-play
        self.precision = precision
        # keep MCTSNode.prior_prob per node (ai_agent.py:522-523 there: computed, never read by the search)
        self.store_priors = bool(store_priors)
        self.deterministic = bool(deterministic)
        self.seed = random.getrandbits(63) if seed is None else int(seed)
        self._engine = None
        self._packed = (None, None)
        self._searches = 0

    # ---- engine plumbing
    def _get_engine(self, n_games, sims):
        e = self._engine
        if e is None or e.max_games < n_games or e.max_simulations < sims:
            if e is not None:
                e.close()
            cap_g = max(n_games, 1)
            e = _engine.Engine(max_games=cap_g, max_simulations=max(sims, 8),
                               device=self.device.index or 0,
                               precision=1 if self.precision == "bf16" else 0,
                               store_priors=self.store_priors)
            self._engine, self._packed = e, (None, None)
            self.knowledge_search._engine = e
        pv_v = weights_version(self.model.model)
        pl_v = (weights_version(self.bg_planner.graph_net), weights_version(self.bg_planner.opp_dqn))
        if self._packed[0] != pv_v:
            e.set_pv_weights(self.model.model.state_dict())
        if self._packed[1] != pl_v:
            e.set_planner_weights(self.bg_planner.graph_net.state_dict(),
                                  self.bg_planner.opp_dqn.state_dict())
            e._planner_set = True
        self._packed = (pv_v, pl_v)
        return e

    def _search_params(self, game_base=0):
        sims = int(self.params.get("num_simulations", 200))
        p = _engine.make_params(sims, float(self.params["c_puct"]), beta=self.beta,
                                planner_steps=self.planner_steps,
                                difficulty=self.difficulty if self.difficulty in
                                _engine.PLANNER_PRESETS else "medium",
                                deterministic=self.deterministic,
                                seed=(self.seed + 0x9E3779B97F4A7C15 * self._searches) & (2 ** 64 - 1),
                                game_base=game_base, time_limit=self.time_limit)
        bp = self.bg_planner.params
        p.planner_k, p.planner_alpha, p.planner_explore = bp["k"], bp["mix_alpha"], bp["explore"]
        return p

    # ---- the reference's entry points
    def get_move(self, board):
        valid_moves = board.get_valid_moves()
        if not valid_moves:
            return None
        if board.get_move_count() < 6:
            return self._opening_move(board, valid_moves)
        best_move = self._mcts_search(board, valid_moves)
        if self._roll() < self.params["exploration"]:
            return self._choice(valid_moves)
        self.games_played += 1
        if self.auto_save_interval and self.games_played % self.auto_save_interval == 0:
            self._auto_save_model()
        return best_move

    def _roll(self):
        return 1.0 if self.deterministic else random.random()

    def _choice(self, seq):
        return seq[0] if self.deterministic else random.choice(seq)

    def _opening_move(self, board, valid_moves):
        c = board.BOARD_SIZE // 2
        n = board.get_move_count()
        if n == 0 and (c, c) in valid_moves:
            return (c, c)
        if n < 6:
            for radius in (1, 2):
                ring = [m for m in valid_moves if abs(m[0] - c) <= radius and abs(m[1] - c) <= radius]
                if ring:
                    return self._choice(ring)
        return self._mcts_search(board, valid_moves)

    def _mcts_search(self, board, valid_moves):
        t0 = time.time()
        res = self.search_batch([board])
        self._last_decision_time = time.time() - t0
        best = int(res["best_move"][0])
        if best < 0:                       # root without children (finished board or S <= 1)
            return self._choice(valid_moves)
        return (best // board.BOARD_SIZE, best % board.BOARD_SIZE)

    def search_batch(self, boards, ai_players=None, game_base=0):
        """One search per board through host buffers.  Returns child visit counts, value sums and
        the best move (raster index, -1 when the root has no child) per board."""
        n = len(boards)
        cells = np.stack([b.to_cells() if hasattr(b, "to_cells") else
                          np.ascontiguousarray(b.board, np.uint8).reshape(225) for b in boards])
        to_move = np.array([b.current_player for b in boards], np.uint8)
        count = np.array([b.get_move_count() for b in boards], np.uint16)
        ai = np.full(n, self.player, np.uint8) if ai_players is None else np.asarray(ai_players, np.uint8)
        p = self._search_params(game_base)
        eng = self._get_engine(n, p.num_simulations)
        self._searches += 1
        cv, cval, best = eng.search_host(cells, to_move, count, ai, p)
        return {"child_visits": cv, "child_value": cval, "best_move": best}

    def root_prior_prob(self, board, index=0):
        """MCTSNode.prior_prob of the root of the last search (ai_agent.py:564-582 there): the network's softmax
        gathered over `board`'s valid moves in raster order and renormalised.  Needs store_priors=True."""
        pr, _ = self._engine.get_priors(index, 0)
        return np.array([pr[r * board.BOARD_SIZE + c] for r, c in board.get_valid_moves()], np.float64)

    def get_moves(self, boards):
        """Batched get_move (book, search, epsilon override) for many independent boards."""
        out = [None] * len(boards)
        todo = []
        for i, b in enumerate(boards):
            vm = b.get_valid_moves()
            if not vm:
                continue
            if b.get_move_count() < 6:
                mv = self._book_only(b, vm)
                if mv is not None:
                    out[i] = mv
                    continue
            todo.append(i)
        if todo:
            res = self.search_batch([boards[i] for i in todo])
            for k, i in enumerate(todo):
                vm = boards[i].get_valid_moves()
                best = int(res["best_move"][k])
                mv = (best // 15, best % 15) if best >= 0 else self._choice(vm)
                if self._roll() < self.params["exploration"]:
                    mv = self._choice(vm)
                else:
                    self.games_played += 1
                out[i] = mv
        return out

    def _book_only(self, board, valid_moves):
        c = board.BOARD_SIZE // 2
        if board.get_move_count() == 0 and (c, c) in valid_moves:
            return (c, c)
        for radius in (1, 2):
            ring = [m for m in valid_moves if abs(m[0] - c) <= radius and abs(m[1] - c) <= radius]
            if ring:
                return self._choice(ring)
        return None

    def _bg_score(self, board, player):
        try:
            return float(np.tanh(self.knowledge_search._pattern_score(board, player) / 10000.0))
        except Exception:
            return 0.0

    def evaluate_position(self, board):
        if board.game_over:
            if board.winner == self.player:
                return 1.0
            return -1.0 if board.winner is not None else 0.0
        _, value = self.model.predict(board.get_board_state())
        return value

    def _auto_save_model(self):
        try:
            path = self.model.auto_save_model(suffix="_%s" % self.difficulty)
            self.logger.info("Auto-saved model to %s", path)
        except Exception as exc:
            self.logger.error("Failed to auto-save model: %s", exc)


class AIFactory:
    @staticmethod
    def create_ai(ai_type, player, difficulty="medium", **kwargs):
        if ai_type == "alphazero":
            return AlphaZeroGomokuAI(player, difficulty, **kwargs)
        if ai_type == "bg_planner":
            return BGPlannerAI(player, difficulty, device=kwargs.get("device", "cuda"))
        raise ValueError("Unsupported ai_type: %s" % ai_type)

    @staticmethod
    def get_available_ai_types():
        return ["alphazero", "bg_planner"]

    @staticmethod
    def get_available_difficulties():
        return ["easy", "medium", "hard"]


__all__ = ["AlphaZeroGomokuAI", "AIFactory", "GomokuBoard", "GomokuModel", "BGPlannerAI",
           "KnowledgeSearch"]
