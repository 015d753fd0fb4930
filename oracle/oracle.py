"""ctypes binding of oracle/libgomoku_oracle.so.  TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs import this
module; the product package never does (a test greps for it).
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None


class Board(C.Structure):
    _fields_ = [("cells", C.c_int8 * 225), ("to_move", C.c_int32), ("move_count", C.c_int32),
                ("game_over", C.c_int32), ("winner", C.c_int32)]


class Rng(C.Structure):
    _fields_ = [("mode", C.c_int32), ("seed_lo", C.c_uint32), ("seed_hi", C.c_uint32),
                ("game", C.c_uint32), ("sim", C.c_uint32)]


class Weights(C.Structure):
    _fields_ = [("pv", C.c_void_p), ("graphnet", C.c_void_p), ("dqn", C.c_void_p)]


class Params(C.Structure):
    _fields_ = [("num_simulations", C.c_int32), ("c_puct", C.c_double), ("beta", C.c_double),
                ("planner_steps", C.c_int32), ("planner_k", C.c_int32), ("planner_alpha", C.c_double),
                ("planner_explore", C.c_double), ("ai_player", C.c_int32), ("eval_nodes", C.c_int32),
                ("log_table", C.c_void_p), ("n_log", C.c_int32),
                ("tanh_table", C.c_void_p), ("n_tanh", C.c_int32)]


class Stats(C.Structure):
    _fields_ = [("n_nodes", C.c_int32), ("best_move", C.c_int32), ("n_pv_evals", C.c_int64),
                ("n_planner_steps", C.c_int64), ("n_rollout_steps", C.c_int64),
                ("n_terminal_leaves", C.c_int64), ("min_gap", C.c_double), ("min_gap_sim", C.c_int32),
                ("min_gap_step", C.c_int32), ("n_argmax_picks", C.c_int64)]


PLANNER_PRESETS = {"easy": (8, 0.5, 0.2), "medium": (12, 0.65, 0.1), "hard": (16, 0.75, 0.05)}


def build(force=False):
    so = os.path.join(_HERE, "libgomoku_oracle.so")
    src = [os.path.join(_HERE, f) for f in ("gomoku_oracle.c", "gomoku_oracle.h")]
    if force or not os.path.exists(so) or any(os.path.getmtime(s) > os.path.getmtime(so) for s in src):
        subprocess.check_call(["make", "-C", _HERE, "-B", "libgomoku_oracle.so"],
                              stdout=subprocess.DEVNULL)
    return so


def lib():
    global _LIB
    if _LIB is None:
        L = C.CDLL(build())
        L.gzo_pattern_score.restype = C.c_int64
        L.gzo_simulate.restype = C.c_double
        for f in ("gzo_pv_param_count", "gzo_graphnet_param_count", "gzo_dqn_param_count"):
            getattr(L, f).restype = C.c_int64
        _LIB = L
    return _LIB


def board_from_moves(moves):
    L = lib()
    b = Board()
    L.gzo_board_init(C.byref(b))
    for m in moves:
        if m < 0:
            break
        assert L.gzo_make_move(C.byref(b), int(m) // 15, int(m) % 15) == 1
    return b


def board_from_cells(cells, to_move, move_count):
    b = Board()
    arr = np.ascontiguousarray(cells, dtype=np.uint8)
    lib().gzo_board_from_cells(C.byref(b), arr.ctypes.data_as(C.c_void_p), int(to_move), int(move_count))
    return b


def cells_of(b):
    return np.frombuffer(b.cells, dtype=np.int8).astype(np.uint8).copy()


def flatten_state_dict(sd):
    """state dict (name -> ndarray) -> one float32 vector in state-dict order, BN counters dropped."""
    parts = [np.asarray(v, dtype=np.float32).reshape(-1) for k, v in sd.items()
             if not k.endswith("num_batches_tracked")]
    return np.ascontiguousarray(np.concatenate(parts))


def log_table(n):
    t = np.zeros(n, dtype=np.float64)
    t[1:] = [float(np.log(i)) for i in range(1, n)]
    return t


def tanh_table(n=4096):
    return np.array([float(np.tanh((i * 50.0) / 10000.0)) for i in range(n)], dtype=np.float64)


class Oracle:
    """Holds the flattened weights and tables and exposes the oracle's entry points."""

    def __init__(self, pv=None, graphnet=None, dqn=None, use_tables=True):
        self.L = lib()
        self._keep = []
        self.w = Weights()
        for name, sd, cnt in (("pv", pv, self.L.gzo_pv_param_count()),
                              ("graphnet", graphnet, self.L.gzo_graphnet_param_count()),
                              ("dqn", dqn, self.L.gzo_dqn_param_count())):
            if sd is not None:
                flat = flatten_state_dict(sd)
                assert flat.size == cnt, (name, flat.size, cnt)
                self._keep.append(flat)
                setattr(self.w, name, flat.ctypes.data)
        self.logt = log_table(1024) if use_tables else None
        self.tanht = tanh_table() if use_tables else None

    def params(self, sims, c_puct, beta, planner_steps, ai_player, difficulty="hard", eval_nodes=0):
        k, a, e = PLANNER_PRESETS.get(difficulty, PLANNER_PRESETS["medium"])
        p = Params(int(sims), float(c_puct), float(beta), int(planner_steps), k, a, e, int(ai_player),
                   int(eval_nodes), None, 0, None, 0)
        if self.logt is not None:
            p.log_table, p.n_log = self.logt.ctypes.data, self.logt.size
            p.tanh_table, p.n_tanh = self.tanht.ctypes.data, self.tanht.size
        return p

    # ---- rules
    def threat(self, b, player):
        return self.L.gzo_threat_level(C.byref(b), player)

    def pattern(self, b, player):
        return int(self.L.gzo_pattern_score(C.byref(b), player))

    def legal(self, b):
        m = np.zeros(225, np.uint8)
        self.L.gzo_legal_mask(C.byref(b), m.ctypes.data_as(C.c_void_p))
        return m

    # ---- rollout
    def select_move(self, b, literal=False, rng=None, step=0):
        cand = np.zeros(225, np.uint8)
        fc = C.c_int(0)
        f = self.L.gzo_select_offensive_move_literal if literal else self.L.gzo_select_offensive_move
        mv = f(C.byref(b), C.byref(rng) if rng is not None else None, step,
               cand.ctypes.data_as(C.c_void_p), C.byref(fc))
        return mv, cand, fc.value

    def simulate(self, b, p, rng=None):
        fb = Board()
        st = Stats()
        st.min_gap = float("inf")
        v = self.L.gzo_simulate(C.byref(b), C.byref(p), C.byref(self.w),
                                C.byref(rng) if rng is not None else None, C.byref(fb), C.byref(st))
        return v, fb, st

    # ---- planner
    def planner_scores(self, b, ai_player):
        s = np.zeros(225, np.float64)
        self.L.gzo_planner_scores(C.byref(b), ai_player, s.ctypes.data_as(C.c_void_p))
        return s

    def planner_move(self, b, p, rng=None, step=0):
        top = np.full(32, -1, np.int32)
        comp = np.full(32, np.nan, np.float64)
        mv = self.L.gzo_planner_move(C.byref(b), C.byref(p), C.byref(self.w),
                                     C.byref(rng) if rng is not None else None, step,
                                     top.ctypes.data_as(C.c_void_p), comp.ctypes.data_as(C.c_void_p))
        return mv, top, comp

    # ---- nets
    def pv_forward(self, cells):
        cells = np.ascontiguousarray(cells, np.uint8)
        lg = np.zeros(225, np.float32); pr = np.zeros(225, np.float32); v = C.c_float(0)
        self.L.gzo_pv_forward(C.c_void_p(self.w.pv), cells.ctypes.data_as(C.c_void_p),
                              lg.ctypes.data_as(C.c_void_p), pr.ctypes.data_as(C.c_void_p), C.byref(v))
        return lg, pr, v.value

    def graphnet_forward(self, cells):
        cells = np.ascontiguousarray(cells, np.uint8)
        lg = np.zeros(225, np.float32)
        self.L.gzo_graphnet_forward(C.c_void_p(self.w.graphnet), cells.ctypes.data_as(C.c_void_p),
                                    lg.ctypes.data_as(C.c_void_p))
        return lg

    def dqn_forward(self, cells):
        cells = np.ascontiguousarray(cells, np.uint8)
        q = np.zeros(225, np.float32)
        self.L.gzo_dqn_forward(C.c_void_p(self.w.dqn), cells.ctypes.data_as(C.c_void_p),
                               q.ctypes.data_as(C.c_void_p))
        return q

    # ---- search
    def search(self, b, p, seed=0, game=0):
        n = p.num_simulations + 2
        parent = np.full(n, -1, np.int32); move = np.full(n, -1, np.int32)
        visits = np.zeros(n, np.int32); value = np.zeros(n, np.float64)
        st = Stats()
        self.L.gzo_search(C.byref(b), C.byref(p), C.byref(self.w), C.c_uint32(seed & 0xFFFFFFFF),
                          C.c_uint32((seed >> 32) & 0xFFFFFFFF), C.c_uint32(game),
                          parent.ctypes.data_as(C.c_void_p), move.ctypes.data_as(C.c_void_p),
                          visits.ctypes.data_as(C.c_void_p), value.ctypes.data_as(C.c_void_p), n,
                          C.byref(st))
        k = st.n_nodes
        return dict(parent=parent[:k], move=move[:k], visits=visits[:k], value=value[:k], stats=st)


# ---- dataset build: CPU restatement of training.py:38-70,104-134 (numpy; test infrastructure) -------------------
def transform_planes(planes, k_rot, flip):
    """_transform_planes (training.py:44-51): np.rot90 counter-clockwise k times over axes (1, 2), then np.flip(axis=2)."""
    x = np.asarray(planes)
    for _ in range(k_rot % 4):
        x = np.rot90(x, k=1, axes=(1, 2))
    if flip:
        x = np.flip(x, axis=2)
    return np.ascontiguousarray(x)


def transform_index(idx, k_rot, flip, n=15, reference_exact=True):
    """_transform_index (training.py:53-61) as written -- (r, c) -> (c, n-1-r) per quarter turn, i.e. CLOCKWISE,
    against the planes' counter-clockwise turn -- or, with reference_exact=False, the index of the cell the
    stone really lands on under transform_planes: (r, c) -> (n-1-c, r)."""
    r, c = divmod(int(idx), n)
    for _ in range(k_rot % 4):
        r, c = (c, n - 1 - r) if reference_exact else (n - 1 - c, r)
    if flip:
        c = n - 1 - c
    return r * n + c


def augment_sample(planes, move_idx, reference_exact=True):
    """augment_sample (training.py:63-70): the 8 symmetries in the order k_rot = 0..3, flip = False, True."""
    return [(transform_planes(planes, k, f), transform_index(move_idx, k, f, reference_exact=reference_exact))
            for k in range(4) for f in (False, True)]


def dataset_samples(planes, moves, outcomes, chosen, reference_exact=True):
    """GomokuSelfPlayDataset.__init__ (training.py:104-127) for a given choice of augmented records:
    every record once, then 8 symmetries of each chosen record.  Returns (planes [N,3,15,15] f32, move [N] i64,
    value [N] f32)."""
    xs, ys, vs = [], [], []
    for i in range(len(moves)):
        xs.append(np.asarray(planes[i], np.float32)); ys.append(int(moves[i])); vs.append(float(outcomes[i]))
    for i in chosen:
        for x, y in augment_sample(planes[i], moves[i], reference_exact):
            xs.append(np.asarray(x, np.float32)); ys.append(y); vs.append(float(outcomes[i]))
    if not xs:
        return np.zeros((0, 3, 15, 15), np.float32), np.zeros(0, np.int64), np.zeros(0, np.float32)
    return np.stack(xs), np.asarray(ys, np.int64), np.asarray(vs, np.float32)


# ---- evaluation arena: CPU restatement of training.py:221-270 (deterministic mode; test infrastructure) --------------
def arena_games(orc, games, vs_random, sims_a, sims_b, c_puct, beta, planner_steps, difficulty="easy"):
    """evaluate_model's loop under the DET patches (random.choice -> seq[0], random.random -> 1.0): game g has side A
    = black when g is even; A (and B when it is an AlphaZeroGomokuAI) plays the opening book below six stones
    (ai_agent.py:138-166, first element of the candidate list), otherwise the best move of a search with
    ai_player = its own colour (the epsilon roll 1.0 never overrides); the RandomAgent plays the first empty cell.
    sims_a / sims_b: num_simulations of side A / side B.  Returns (list of move lists, winners (0 = none), [W, L, D])."""
    L = orc.L
    all_moves, winners, wld = [], [], [0, 0, 0]
    for g in range(games):
        side_a = 1 if g % 2 == 0 else 2
        b = Board()
        L.gzo_board_init(C.byref(b))
        moves = []
        while not b.game_over:
            empties = [i for i in range(225) if b.cells[i] == 0]
            if not empties:
                break
            mover_is_a = b.to_move == side_a
            if vs_random and not mover_is_a:
                mv = empties[0]
            elif b.move_count < 6:
                if b.move_count == 0:
                    mv = 112
                else:
                    mv = None
                    for radius in (1, 2):
                        ring = [i for i in empties if abs(i // 15 - 7) <= radius and abs(i % 15 - 7) <= radius]
                        if ring:
                            mv = ring[0]
                            break
                    assert mv is not None
            else:
                p = orc.params(sims_a if mover_is_a else sims_b, c_puct, beta, planner_steps, int(b.to_move), difficulty)
                t = orc.search(b, p)
                mv = t["stats"].best_move if t["stats"].best_move >= 0 else empties[0]
            assert L.gzo_make_move(C.byref(b), mv // 15, mv % 15)
            moves.append(mv)
        all_moves.append(moves)
        winners.append(int(b.winner))
        wld[2 if b.winner == 0 else (0 if b.winner == side_a else 1)] += 1
    return all_moves, winners, wld
