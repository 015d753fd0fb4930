"""ctypes binding of libgomoku_b200.so (include/gomoku_b200.h) plus a thin torch-tensor wrapper.

PyTorch is plumbing here: it owns device memory and streams; every computation on the hot path is
a kernel inside libgomoku_b200.so.  There is no CPU fallback: constructing an Engine without a
CUDA device, or without the built library, raises.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "libgomoku_b200.so")
_LIB = None

PLANNER_PRESETS = {"easy": (8, 0.5, 0.2), "medium": (12, 0.65, 0.1), "hard": (16, 0.75, 0.05)}

EXPORTS = [
    "gz_create", "gz_destroy", "gz_last_error", "gz_version", "gz_set_pv_weights",
    "gz_set_planner_weights", "gz_set_tables", "gz_pv_param_count", "gz_graphnet_param_count",
    "gz_dqn_param_count", "gz_boards_from_cells", "gz_boards_to_cells", "gz_make_moves",
    "gz_board_ops", "gz_pv_forward", "gz_planner_nets_forward", "gz_planner_move",
    "gz_rollout_policy", "gz_rollout", "gz_search", "gz_search_host", "gz_get_tree",
    "gz_get_counters", "gz_reset_counters", "gz_set_profiling", "gz_set_lanes", "gz_get_stage_times", "gz_debug_pv_layer",
    "gz_selfplay_begin", "gz_selfplay_step", "gz_selfplay_drain", "gz_selfplay_boards", "gz_selfplay_get_stats",
    "gz_dataset_build", "gz_arena_begin", "gz_arena_get_stats", "gz_get_priors", "gz_set_graph", "gz_selfplay_set_simulations", "gz_pv_forward_static",
]

STATIC_MAX_BOARDS = 16    # GZ_STATIC_MAX_BOARDS

ARENA_ENGINE, ARENA_RANDOM = 1, 2     # GZ_ARENA_ENGINE / GZ_ARENA_RANDOM

STAGES = ["tree", "pv_conv0", "pv_tower", "pv_heads", "planner_score", "gn_convs", "gn_head", "dqn",
          "planner_pick", "rollout_tail", "misc", "reserved"]


class GzConfig(C.Structure):
    _fields_ = [("device", C.c_int32), ("max_games", C.c_int32), ("max_simulations", C.c_int32),
                ("precision", C.c_int32), ("store_priors", C.c_int32), ("lanes", C.c_int32)]


class GzSearchParams(C.Structure):
    _fields_ = [("num_simulations", C.c_int32), ("planner_steps", C.c_int32),
                ("planner_k", C.c_int32), ("deterministic", C.c_int32), ("c_puct", C.c_double),
                ("beta", C.c_double), ("planner_alpha", C.c_double), ("planner_explore", C.c_double),
                ("seed", C.c_uint64), ("game_base", C.c_uint32), ("eval_nodes", C.c_int32),
                ("time_limit_s", C.c_double)]


class GzCounters(C.Structure):
    _fields_ = [(n, C.c_int64) for n in ("simulations", "nodes_created", "pv_evals", "planner_steps",
                                          "rollout_steps", "terminal_leaves", "kernel_launches",
                                          "last_search_simulations")]


class GzSelfplayStats(C.Structure):
    _fields_ = [(n, C.c_int64) for n in ("games_finished", "plies_played", "black_wins", "white_wins", "draws",
                                          "records_ready", "records_dropped", "steps")]


class GzArenaStats(C.Structure):
    _fields_ = [(n, C.c_int64) for n in ("games_finished", "wins", "losses", "draws", "plies_played", "steps")]


BOARD_BYTES = 80
# gz_replay_record (include/gomoku_b200.h): 72 bytes
REPLAY_DTYPE = np.dtype([("stones", "<u4", (2, 8)), ("move", "u1"), ("mover", "u1"), ("outcome", "i1"),
                         ("ply", "u1"), ("game_id", "<u4")])
assert REPLAY_DTYPE.itemsize == 72


class EngineError(RuntimeError):
    pass


def load_library(path=None):
    """dlopen the engine.  Raises if it was not built: the product never runs without it."""
    global _LIB
    if _LIB is None:
        path = path or _SO
        if not os.path.exists(path):
            raise EngineError("libgomoku_b200.so is missing (%s): run `python alphazero-gomoku_b200/"
                              "build.py`; there is no fallback path" % path)
        L = C.CDLL(path)
        L.gz_last_error.restype = C.c_char_p
        for f in ("gz_pv_param_count", "gz_graphnet_param_count", "gz_dqn_param_count"):
            getattr(L, f).restype = C.c_int64
        _LIB = L
    return _LIB


def flatten_state_dict(sd):
    """torch/numpy state dict -> one float32 vector in state-dict order without BN counters."""
    parts = []
    for k, v in sd.items():
        if k.endswith("num_batches_tracked"):
            continue
        a = v.detach().cpu().numpy() if hasattr(v, "detach") else np.asarray(v)
        parts.append(np.asarray(a, dtype=np.float32).reshape(-1))
    return np.ascontiguousarray(np.concatenate(parts))


def make_params(sims, c_puct, beta=0.2, planner_steps=0, difficulty="medium", deterministic=False,
                seed=0, game_base=0, eval_nodes=True, time_limit=0.0):
    k, a, e = PLANNER_PRESETS.get(difficulty, PLANNER_PRESETS["medium"])
    return GzSearchParams(int(sims), int(max(0, planner_steps)), k, int(bool(deterministic)),
                          float(c_puct), float(beta), a, e, int(seed) & (2 ** 64 - 1),
                          int(game_base), int(bool(eval_nodes)),
                          float(time_limit) if time_limit and time_limit != float("inf") else 0.0)


def _ptr(t):
    return C.c_void_p(0 if t is None else t.data_ptr())


class Engine:
    """One engine handle on one GPU; all tensors it returns live on that GPU."""

    def __init__(self, max_games, max_simulations, device=0, precision=0, store_priors=False,
                 host_tables=True, lanes=0):
        import torch
        self.torch = torch
        self.L = load_library()
        if not torch.cuda.is_available():
            raise EngineError("no CUDA device: the B200 engine has no CPU path")
        self.device = torch.device("cuda", device)
        self.max_games, self.max_simulations = int(max_games), int(max_simulations)
        cfg = GzConfig(device, self.max_games, self.max_simulations, precision, int(store_priors), int(lanes))
        self.h = C.c_void_p()
        self._ck(self.L.gz_create(C.byref(cfg), C.byref(self.h)))
        if host_tables:
            n = min(2048, self.max_simulations + 2)
            logt = np.zeros(n, np.float64)
            logt[1:] = [float(np.log(i)) for i in range(1, n)]
            tanht = np.array([float(np.tanh((i * 50.0) / 10000.0)) for i in range(4096)], np.float64)
            self._ck(self.L.gz_set_tables(self.h, logt.ctypes.data_as(C.c_void_p), n,
                                          tanht.ctypes.data_as(C.c_void_p), 4096))

    def close(self):
        if getattr(self, "h", None) is not None and self.h:
            self.L.gz_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _ck(self, rc):
        if rc != 0:
            raise EngineError("gomoku_b200 error %d: %s" % (rc, self.L.gz_last_error().decode()))

    def _stream(self):
        return C.c_void_p(self.torch.cuda.current_stream(self.device).cuda_stream)

    # ---- weights
    def set_pv_weights(self, state_dict):
        f = flatten_state_dict(state_dict)
        self._ck(self.L.gz_set_pv_weights(self.h, f.ctypes.data_as(C.c_void_p), C.c_int64(f.size)))

    def set_planner_weights(self, graphnet_sd, dqn_sd):
        g, d = flatten_state_dict(graphnet_sd), flatten_state_dict(dqn_sd)
        self._ck(self.L.gz_set_planner_weights(self.h, g.ctypes.data_as(C.c_void_p), C.c_int64(g.size),
                                               d.ctypes.data_as(C.c_void_p), C.c_int64(d.size)))

    # ---- boards
    def boards_from_cells(self, cells, to_move, move_count):
        t = self.torch
        cells = t.as_tensor(np.ascontiguousarray(cells, np.uint8)).to(self.device).reshape(-1, 225)
        G = cells.shape[0]
        tm = t.as_tensor(np.ascontiguousarray(to_move, np.uint8)).to(self.device)
        mc = t.as_tensor(np.ascontiguousarray(move_count, np.int16)).to(self.device)
        boards = t.empty((G, BOARD_BYTES), dtype=t.uint8, device=self.device)
        self._ck(self.L.gz_boards_from_cells(self.h, _ptr(cells), _ptr(tm), _ptr(mc), G, _ptr(boards),
                                             self._stream()))
        return boards

    def boards_to_cells(self, boards):
        G = boards.shape[0]
        out = self.torch.empty((G, 225), dtype=self.torch.uint8, device=self.device)
        self._ck(self.L.gz_boards_to_cells(self.h, _ptr(boards), G, _ptr(out), self._stream()))
        return out

    def board_meta(self, boards):
        """(move_count, to_move, status) numpy arrays decoded from the 80-byte records."""
        raw = boards.cpu().numpy()
        mc = raw[:, 64:66].copy().view(np.uint16).reshape(-1)
        return mc, raw[:, 66].copy(), raw[:, 67].copy()

    def make_moves(self, boards, moves):
        t = self.torch
        G = boards.shape[0]
        mv = t.as_tensor(np.ascontiguousarray(moves, np.int32)).to(self.device)
        ok = t.empty(G, dtype=t.uint8, device=self.device)
        self._ck(self.L.gz_make_moves(self.h, _ptr(boards), _ptr(mv), G, _ptr(ok), self._stream()))
        return ok

    def board_ops(self, boards):
        t = self.torch
        G = boards.shape[0]
        legal = t.empty((G, 8), dtype=t.int32, device=self.device)
        win = t.empty((G, 2, 8), dtype=t.int32, device=self.device)
        three = t.empty((G, 2, 8), dtype=t.int32, device=self.device)
        threat = t.empty((G, 2), dtype=t.int32, device=self.device)
        pattern = t.empty((G, 2), dtype=t.int64, device=self.device)
        self._ck(self.L.gz_board_ops(self.h, _ptr(boards), G, _ptr(legal), _ptr(win), _ptr(three),
                                     _ptr(threat), _ptr(pattern), self._stream()))
        return dict(legal=legal, win=win, three=three, threat=threat, pattern=pattern)

    # ---- networks
    def pv_forward(self, boards):
        t = self.torch
        G = boards.shape[0]
        logits = t.empty((G, 225), dtype=t.float32, device=self.device)
        probs = t.empty_like(logits)
        value = t.empty(G, dtype=t.float32, device=self.device)
        self._ck(self.L.gz_pv_forward(self.h, _ptr(boards), G, _ptr(logits), _ptr(probs), _ptr(value),
                                      self._stream()))
        return logits, probs, value

    def pv_forward_static(self, boards, out):
        """gz_pv_forward_static: at most STATIC_MAX_BOARDS boards, activations in the library's static device array,
        no allocation inside the call; `out` = caller-owned (logits [G,225], probs [G,225], value [G]) f32 tensors."""
        logits, probs, value = out
        self._ck(self.L.gz_pv_forward_static(self.h, _ptr(boards), int(boards.shape[0]), _ptr(logits), _ptr(probs),
                                             _ptr(value), self._stream()))
        return out

    def debug_pv_layer(self, boards, nlayers):
        G = boards.shape[0]
        out = self.torch.empty((G, 128, 225), dtype=self.torch.float32, device=self.device)
        self._ck(self.L.gz_debug_pv_layer(self.h, _ptr(boards), G, int(nlayers), _ptr(out), self._stream()))
        return out

    def planner_nets_forward(self, boards):
        t = self.torch
        G = boards.shape[0]
        lg = t.empty((G, 225), dtype=t.float32, device=self.device)
        pol = t.empty_like(lg)
        q = t.empty_like(lg)
        self._ck(self.L.gz_planner_nets_forward(self.h, _ptr(boards), G, _ptr(lg), _ptr(pol), _ptr(q),
                                                self._stream()))
        return lg, pol, q

    # ---- planner / rollout
    def _ai(self, ai_player, G):
        a = np.broadcast_to(np.asarray(ai_player, np.uint8), (G,))
        return self.torch.as_tensor(np.array(a, copy=True, order="C")).to(self.device)   # copy: fixtures are read-only

    def planner_move(self, boards, ai_player, params, sim=0, step=0):
        t = self.torch
        G = boards.shape[0]
        ai = self._ai(ai_player, G)
        scores2 = t.empty((G, 225), dtype=t.int64, device=self.device)
        topk = t.empty((G, 16), dtype=t.int32, device=self.device)
        comp = t.empty((G, 16), dtype=t.float64, device=self.device)
        move = t.empty(G, dtype=t.int32, device=self.device)
        self._ck(self.L.gz_planner_move(self.h, _ptr(boards), _ptr(ai), G, C.byref(params), sim, step,
                                        _ptr(scores2), _ptr(topk), _ptr(comp), _ptr(move),
                                        self._stream()))
        return dict(scores2=scores2, topk=topk, composed=comp, move=move)

    def rollout_policy(self, boards, params, sim=0, step=0):
        t = self.torch
        G = boards.shape[0]
        move = t.empty(G, dtype=t.int32, device=self.device)
        cand = t.empty((G, 8), dtype=t.int32, device=self.device)
        self._ck(self.L.gz_rollout_policy(self.h, _ptr(boards), G, C.byref(params), sim, step,
                                          _ptr(move), _ptr(cand), self._stream()))
        return move, cand

    def rollout(self, boards, ai_player, params, sim=0):
        t = self.torch
        G = boards.shape[0]
        ai = self._ai(ai_player, G)
        value = t.empty(G, dtype=t.float64, device=self.device)
        steps = t.empty(G, dtype=t.int32, device=self.device)
        final = t.empty((G, BOARD_BYTES), dtype=t.uint8, device=self.device)
        self._ck(self.L.gz_rollout(self.h, _ptr(boards), _ptr(ai), G, C.byref(params), sim,
                                   _ptr(value), _ptr(steps), _ptr(final), self._stream()))
        return value, steps, final

    # ---- search
    def search(self, boards, ai_player, params):
        t = self.torch
        G = boards.shape[0]
        ai = self._ai(ai_player, G)
        cv = t.empty((G, 225), dtype=t.int32, device=self.device)
        cval = t.empty((G, 225), dtype=t.float64, device=self.device)
        best = t.empty(G, dtype=t.int32, device=self.device)
        self._ck(self.L.gz_search(self.h, _ptr(boards), _ptr(ai), G, C.byref(params), _ptr(cv),
                                  _ptr(cval), _ptr(best), self._stream()))
        return cv, cval, best

    def search_host(self, cells, to_move, move_count, ai_player, params, out=None):
        """Host buffers in, host buffers out (the call behind get_move)."""
        cells = np.ascontiguousarray(cells, np.uint8).reshape(-1, 225)
        G = cells.shape[0]
        tm = np.ascontiguousarray(np.broadcast_to(np.asarray(to_move, np.uint8), (G,)))
        mc = np.ascontiguousarray(np.broadcast_to(np.asarray(move_count, np.uint16), (G,)))
        ai = np.ascontiguousarray(np.broadcast_to(np.asarray(ai_player, np.uint8), (G,)))
        if out is None:
            out = (np.empty((G, 225), np.int32), np.empty((G, 225), np.float64), np.empty(G, np.int32))
        cv, cval, best = out
        P = lambda a: a.ctypes.data_as(C.c_void_p)  # noqa: E731
        self._ck(self.L.gz_search_host(self.h, P(cells), P(tm), P(mc), P(ai), G, C.byref(params),
                                       P(cv), P(cval), P(best)))
        return cv, cval, best

    # ---- self-play
    def selfplay_begin(self, n_games, params, exploration, replay_capacity=0, games_total=0):
        """games_total = N > 0: exactly the N games with start ordinal generation * n_games + slot < N are counted and
        recorded (the reference's `for game in range(N)`); 0 = open-ended recycling."""
        self._sp_games = int(n_games)
        self._ck(self.L.gz_selfplay_begin(self.h, int(n_games), C.byref(params), C.c_double(exploration),
                                          C.c_int64(replay_capacity), C.c_int64(games_total), self._stream()))

    def arena_begin(self, n_games, params, exploration, opponent, games_total, ordinal_stride=None,
                    replay_capacity=0, sims_b=0):
        """evaluate_model's game loop (training.py:221-270) for n_games concurrent games; opponent is ARENA_ENGINE
        or ARENA_RANDOM; params.num_simulations is side A's budget, sims_b side B's (0 = the same).  Stepped /
        drained with the selfplay_* calls."""
        self._sp_games = int(n_games)
        self._ck(self.L.gz_arena_begin(self.h, int(n_games), C.byref(params), C.c_double(exploration),
                                       C.c_int32(opponent), C.c_int32(sims_b), C.c_int64(games_total),
                                       C.c_int64(ordinal_stride or n_games), C.c_int64(replay_capacity),
                                       self._stream()))

    def arena_stats(self):
        st = GzArenaStats()
        self._ck(self.L.gz_arena_get_stats(self.h, C.byref(st)))
        return {n: getattr(st, n) for n, _ in GzArenaStats._fields_}

    def selfplay_step(self):
        self._ck(self.L.gz_selfplay_step(self.h, self._stream()))

    def selfplay_set_simulations(self, n):
        self._ck(self.L.gz_selfplay_set_simulations(self.h, int(n)))

    def selfplay_stats(self):
        st = GzSelfplayStats()
        self._ck(self.L.gz_selfplay_get_stats(self.h, C.byref(st)))
        return {n: getattr(st, n) for n, _ in GzSelfplayStats._fields_}

    def selfplay_drain(self):
        """Finished-game records as a uint8 device tensor [n, 72] (view with REPLAY_DTYPE on the host)."""
        t = self.torch
        n = C.c_int64()
        st = self.selfplay_stats()
        out = t.empty((max(st["records_ready"], 1), REPLAY_DTYPE.itemsize), dtype=t.uint8, device=self.device)
        self._ck(self.L.gz_selfplay_drain(self.h, _ptr(out), C.c_int64(out.shape[0]), C.byref(n), self._stream()))
        return out[:n.value]

    def dataset_build(self, records, chosen=None, sample_ids=None, first=0, count=None, reference_exact=True):
        """gz_dataset_build: (planes f32 [count,3,15,15], move i64 [count], value f32 [count]) device tensors
        from packed replay records (uint8 device tensor [n, 72]); `chosen` = the records augmented eightfold."""
        t = self.torch
        rec = records if t.is_tensor(records) else t.as_tensor(np.ascontiguousarray(records).view(np.uint8))
        rec = rec.to(self.device).contiguous().view(t.uint8).reshape(-1, REPLAY_DTYPE.itemsize)
        n = rec.shape[0]
        if chosen is None:
            ch = t.zeros(0, dtype=t.int32)
        elif t.is_tensor(chosen):
            ch = chosen.to(t.int32)
        else:
            ch = t.as_tensor(np.asarray(chosen, np.int32))
        ch = ch.to(self.device).contiguous()
        m = ch.shape[0]
        ids = None
        if sample_ids is not None:
            ids = sample_ids.to(t.int64) if t.is_tensor(sample_ids) else t.as_tensor(np.asarray(sample_ids, np.int64))
            ids = ids.to(self.device).contiguous()
            count = ids.shape[0]
        elif count is None:
            count = n + 8 * m - first
        planes = t.empty((count, 3, 15, 15), dtype=t.float32, device=self.device)
        move = t.empty(count, dtype=t.int64, device=self.device)
        value = t.empty(count, dtype=t.float32, device=self.device)
        self._ck(self.L.gz_dataset_build(self.h, _ptr(rec) if n else None, C.c_int64(n), _ptr(ch) if m else None,
                                         C.c_int64(m), _ptr(ids) if ids is not None else None, C.c_int64(first),
                                         C.c_int64(count), int(bool(reference_exact)), _ptr(planes) if count else None,
                                         _ptr(move) if count else None, _ptr(value) if count else None,
                                         self._stream()))
        return planes, move, value

    def selfplay_boards(self):
        out = self.torch.empty((self._sp_games, BOARD_BYTES), dtype=self.torch.uint8, device=self.device)
        self._ck(self.L.gz_selfplay_boards(self.h, _ptr(out), self._stream()))
        return out

    def get_tree(self, g):
        cap = self.max_simulations + 1
        parent = np.empty(cap, np.int32); move = np.empty(cap, np.int32)
        visits = np.empty(cap, np.int32); value = np.empty(cap, np.float64)
        n = C.c_int32()
        P = lambda a: a.ctypes.data_as(C.c_void_p)  # noqa: E731
        self._ck(self.L.gz_get_tree(self.h, int(g), cap, P(parent), P(move), P(visits), P(value),
                                    C.byref(n)))
        k = n.value
        return dict(parent=parent[:k], move=move[:k], visits=visits[:k], value=value[:k])

    def get_priors(self, g, node=0):
        """(MCTSNode.prior_prob as float32[225] over raster cells (0 on occupied cells), value-head output) of a
        node of game g after a search; needs store_priors=True."""
        pr = np.empty(225, np.float32)
        v = C.c_float()
        self._ck(self.L.gz_get_priors(self.h, int(g), int(node), pr.ctypes.data_as(C.c_void_p), C.byref(v)))
        return pr, v.value

    def counters(self):
        c = GzCounters()
        self._ck(self.L.gz_get_counters(self.h, C.byref(c)))
        return {n: getattr(c, n) for n, _ in GzCounters._fields_}

    def set_profiling(self, on):
        self._ck(self.L.gz_set_profiling(self.h, int(bool(on))))

    def set_graph(self, on):
        self._ck(self.L.gz_set_graph(self.h, int(bool(on))))

    def set_lanes(self, lanes):
        self._ck(self.L.gz_set_lanes(self.h, int(lanes)))

    def stage_times(self):
        """{stage: (milliseconds, launches)} accumulated since the last call (CUDA events)."""
        ms = (C.c_double * len(STAGES))()
        ln = (C.c_int64 * len(STAGES))()
        self._ck(self.L.gz_get_stage_times(self.h, ms, ln))
        return {n: (ms[i], ln[i]) for i, n in enumerate(STAGES)}

    def reset_counters(self):
        self._ck(self.L.gz_reset_counters(self.h))
