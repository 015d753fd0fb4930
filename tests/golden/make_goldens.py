#!/usr/bin/env python
"""Generate the golden fixtures under tests/golden/ by running the UNMODIFIED reference.

Runs only in the build container: it imports /root/reference (read-only) from a scratch cwd with no
./models directory (neural_network.py:197-212 would silently load checkpoints from there), applies
the deterministic-mode patches of SURVEY.md A.7 where a part says so, and writes small .npz files
that travel with the repo.  Nothing under tests/, bench.py or smoke() reads /root/reference at run
time; they read these fixtures.

    python tests/golden/make_goldens.py --parts all --jobs 8

Parts (each -> tests/golden/<part>.npz):
  boards    rules + masks + threat levels + pattern scores            gomoku_board.py, ai_agent.py:403-430, bg_planner.py:133-196
  tiers     candidate set handed to random.choice by the rollout policy   ai_agent.py:306-361
  rollout   DET-A playouts (planner_steps=0): value, steps, final board   ai_agent.py:251-304
  mcts      DET-A searches: the whole tree (parent, move, visits, value)  ai_agent.py:168-249,441-454,532-562
  planner   knowledge-search scores, stable top-k, blended pick (DET-B)  bg_planner.py:90-131,232-269
  nets      logits / value / softmax of the three networks (fp32, 1 thread)   neural_network.py:132-159,214-252
  mctsb     DET-B playouts and searches (planner_steps>0)
  stoch     outcome frequencies of real-RNG playouts (Mersenne Twister)
  dataset   SimpleReplay -> GomokuSelfPlayDataset with the 8-fold augmentation   training.py:38-134
  dropin    the class surface: get_move (book + search, README presets), predict, evaluate_position, prior_prob,
            BGPlannerAI through AIFactory with the real RNG stream, a checkpoint (-> dropin.npz, ref_checkpoint.pth)
"""
import argparse
import os
import sys
import tempfile
import time

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
sys.path.insert(0, "/root/reference")
_SCRATCH = tempfile.mkdtemp(prefix="gz_golden_")
os.chdir(_SCRATCH)

import random  # noqa: E402
import numpy as np  # noqa: E402
import torch  # noqa: E402

torch.set_num_threads(1)

import gomoku_board  # noqa: E402
import ai_agent  # noqa: E402
import bg_planner  # noqa: E402
import neural_network  # noqa: E402
from gomoku_board import GomokuBoard  # noqa: E402
from weights import make_state_dict  # noqa: E402

WEIGHT_SEED = 1234
_ORIG_CHOICE = random.choice
_ORIG_RANDOM = random.random


# ----------------------------------------------------------------------------- helpers
def det_patch():
    random.choice = lambda s: s[0]
    random.random = lambda: 1.0


def det_unpatch():
    random.choice = _ORIG_CHOICE
    random.random = _ORIG_RANDOM


def board_from_moves(moves):
    b = GomokuBoard()
    for m in moves:
        ok = b.make_move(int(m) // 15, int(m) % 15)
        assert ok, "illegal move in a generated sequence"
    return b


def cells_of(b):
    return b.board.astype(np.uint8).reshape(225)


def load_weights(module, kind, gain=None):
    sd = make_state_dict(kind, WEIGHT_SEED, gain)
    module.load_state_dict({k: torch.from_numpy(np.asarray(v)) for k, v in sd.items()}, strict=True)
    module.eval()


def make_ai(player, difficulty="hard", planner_steps=0, beta=0.2, sims=None, c_puct=None):
    ai = ai_agent.AlphaZeroGomokuAI(player, difficulty, device="cpu", beta=beta,
                                    planner_steps=planner_steps, time_limit=float("inf"))
    load_weights(ai.model.model, "pv")
    load_weights(ai.bg_planner.graph_net, "graphnet")
    load_weights(ai.bg_planner.opp_dqn, "dqn")
    ai.params = dict(ai.params)
    if sims is not None:
        ai.params["num_simulations"] = int(sims)
    if c_puct is not None:
        ai.params["c_puct"] = float(c_puct)
    return ai


# ----------------------------------------------------------------------------- positions
def _interleave(black, white):
    assert len(black) in (len(white), len(white) + 1)
    out = []
    for i in range(len(black)):
        out.append(black[i][0] * 15 + black[i][1])
        if i < len(white):
            out.append(white[i][0] * 15 + white[i][1])
    return out


def tactical_positions():
    """Hand-built boards: win-in-1, open three/four, edges, corners, overline, both sides armed."""
    P = []
    # black open four in the middle, white to move
    P.append(_interleave([(7, 3), (7, 4), (7, 5), (7, 6)], [(0, 0), (0, 2), (0, 4), (0, 6)]))
    # same, black to move (win-in-1 at two cells)
    P.append(_interleave([(7, 3), (7, 4), (7, 5), (7, 6)], [(0, 0), (0, 2), (0, 4)])[:7] + [0 * 15 + 8])
    # white three in a column touching the top edge, black scattered
    P.append(_interleave([(14, 0), (14, 2), (14, 4), (12, 7)], [(0, 5), (1, 5), (2, 5)]))
    # black run of four into the top-left corner along the diagonal
    P.append(_interleave([(0, 0), (1, 1), (2, 2), (3, 3)], [(0, 14), (2, 14), (4, 14), (6, 14)]))
    # anti-diagonal run of four ending at the bottom-left corner, white to move
    P.append(_interleave([(14, 0), (13, 1), (12, 2), (11, 3)], [(0, 1), (0, 3), (0, 5), (0, 7)]))
    # overline shape XXX_XX for black (filling the gap makes six), black to move
    P.append(_interleave([(5, 2), (5, 3), (5, 4), (5, 6), (5, 7)], [(9, 0), (9, 2), (9, 4), (9, 6), (9, 8)]))
    # broken three XX.X and X.XX for white, row 14 (bottom edge) and col 14 (right edge)
    P.append(_interleave([(7, 7), (0, 0), (0, 2), (2, 0), (2, 2), (4, 0)],
                         [(14, 3), (14, 4), (14, 6), (3, 14), (5, 14), (6, 14)]))
    # both players have a four: the mover wins first
    P.append(_interleave([(3, 3), (3, 4), (3, 5), (3, 6)], [(10, 3), (10, 4), (10, 5), (10, 6)]))
    P.append(_interleave([(3, 3), (3, 4), (3, 5), (3, 6), (12, 12)], [(10, 3), (10, 4), (10, 5), (10, 6)]))
    # right-edge horizontal four for black ending at col 14 (one open end only)
    P.append(_interleave([(6, 11), (6, 12), (6, 13), (6, 14)], [(8, 0), (8, 2), (8, 4), (8, 6)]))
    # live two / live three mix around the centre
    P.append(_interleave([(7, 7), (7, 8), (9, 9), (10, 10)], [(6, 6), (5, 5), (8, 3), (8, 4)]))
    # vertical white open three, black has a diagonal broken three
    P.append(_interleave([(4, 4), (5, 5), (7, 7), (11, 2)], [(6, 10), (7, 10), (8, 10)]))
    # finished games: black five, white five, black six (overline) -> game_over boards
    P.append(_interleave([(7, 3), (7, 4), (7, 5), (7, 6), (7, 7)], [(0, 0), (0, 2), (0, 4), (0, 6)]))
    P.append(_interleave([(0, 0), (0, 2), (0, 4), (0, 6), (2, 0)], [(10, 14), (11, 13), (12, 12), (13, 11), (14, 10)]))
    P.append(_interleave([(5, 2), (5, 3), (5, 4), (5, 6), (5, 7), (5, 5)], [(9, 0), (9, 2), (9, 4), (9, 6), (9, 8)]))
    return P


def dense_pattern_moves(rng, n_stones):
    """A legal alternating move sequence of n_stones on the colouring ((r+2c) mod 4 < 2 -> black),
    which has no run longer than two in any direction, in a random order.  Used for late-game
    boards (draw at 200 stones, gomoku_board.py:106)."""
    blk = [(r, c) for r in range(15) for c in range(15) if (r + 2 * c) % 4 < 2]
    wht = [(r, c) for r in range(15) for c in range(15) if (r + 2 * c) % 4 >= 2]
    rng.shuffle(blk)
    rng.shuffle(wht)
    nb, nw = (n_stones + 1) // 2, n_stones // 2
    return _interleave(blk[:nb], wht[:nw])


def random_position(rng, plies, style):
    b = GomokuBoard()
    seq = []
    while b.get_move_count() < plies:
        vm = b.get_valid_moves()
        for _try in range(50):
            if style == "cluster" and b.get_move_count() > 0:
                occ = np.argwhere(b.board != 0)
                near = [m for m in vm if np.min(np.max(np.abs(occ - np.array(m)), axis=1)) <= 2]
                pool = near or vm
            elif style == "centre":
                pool = [m for m in vm if abs(m[0] - 7) + abs(m[1] - 7) <= 6] or vm
            else:
                pool = vm
            mv = pool[int(rng.integers(len(pool)))]
            t = b.copy_board()
            t.make_move(*mv)
            if not t.game_over:
                break
        else:
            break
        b.make_move(*mv)
        seq.append(mv[0] * 15 + mv[1])
    return seq


def policy_position(seed, plies):
    """Positions reached by the reference's own rollout policy with its real RNG (runs of 3-4 appear)."""
    det_unpatch()
    random.seed(seed)
    ai = ai_agent.AlphaZeroGomokuAI(1, "easy", device="cpu", planner_steps=0)
    b = GomokuBoard()
    b.make_move(7, 7)
    seq = [7 * 15 + 7]
    while b.get_move_count() < plies:
        mv = ai._select_offensive_move(b, b.get_valid_moves())
        t = b.copy_board()
        t.make_move(*mv)
        if t.game_over:
            break
        b.make_move(*mv)
        seq.append(mv[0] * 15 + mv[1])
    return seq


def build_position_set():
    """The shared position list (move sequences).  Deterministic: PCG64 + Mersenne seeds."""
    rng = np.random.default_rng(20261104)
    pos = []
    tags = []
    for p in tactical_positions():
        pos.append(p); tags.append("tactical")
    for plies in (6, 7, 12, 13, 24, 25, 48, 49, 96, 97, 150, 151):
        for style in ("uniform", "cluster", "centre"):
            for _rep in range(3):
                pos.append(random_position(rng, plies, style)); tags.append(style)
    for n in (120, 150, 170, 171, 180, 190, 195, 196, 197, 198, 199):
        for _rep in range(3):
            pos.append(dense_pattern_moves(rng, n)); tags.append("dense")
    return pos, tags


def pad_moves(seqs, width=200):
    out = np.full((len(seqs), width), -1, dtype=np.int16)
    for i, s in enumerate(seqs):
        out[i, :len(s)] = s
    return out


# ----------------------------------------------------------------------------- part: positions
def part_positions(pool, outdir):
    pos, tags = build_position_set()
    specs = [(1000 + i, pl) for i, pl in enumerate((10, 14, 18, 22, 26, 30, 34, 38, 44, 50, 60, 70,
                                                    11, 15, 19, 23, 27, 31, 35, 39, 45, 51, 61, 71))]
    for seq in pool.starmap(policy_position, specs):
        pos.append(seq); tags.append("policy")
    np.savez_compressed(os.path.join(outdir, "positions.npz"), moves=pad_moves(pos),
                        tags=np.array(tags))
    print("positions:", len(pos))


def load_positions(outdir):
    z = np.load(os.path.join(outdir, "positions.npz"))
    return [[int(m) for m in row if m >= 0] for row in z["moves"]], list(z["tags"])


# ----------------------------------------------------------------------------- part: boards
def _board_task(seq):
    b = board_from_moves(seq)
    ai = _board_task.ai
    ks = _board_task.ks
    rec = {}
    rec["cells"] = cells_of(b)
    rec["to_move"] = b.current_player
    rec["move_count"] = b.get_move_count()
    rec["game_over"] = int(b.game_over)
    rec["winner"] = 0 if b.winner is None else int(b.winner)
    legal = np.zeros(225, np.uint8)
    for (r, c) in b.get_valid_moves():
        legal[r * 15 + c] = 1
    rec["legal"] = legal
    rec["threat"] = np.array([ai._evaluate_threat_level(b, 1), ai._evaluate_threat_level(b, 2)], np.int32)
    rec["pattern"] = np.array([ks._pattern_score(b, 1), ks._pattern_score(b, 2)], np.float64)
    # win_cells[p]: empty cells where a stone of p ends the game with winner p (forced mover),
    # three_cells[p]: empty cells where a stone of p gives threat level >= 3 for p
    win = np.zeros((2, 225), np.uint8)
    three = np.zeros((2, 225), np.uint8)
    status_after = np.zeros((2, 225), np.uint8)   # 0 live, 1/2 winner, 3 draw, 255 illegal
    status_after[:] = 255
    if not b.game_over:
        for p in (1, 2):
            for (r, c) in b.get_valid_moves():
                t = b.copy_board()
                t.current_player = p
                t.make_move(r, c)
                if t.game_over and t.winner == p:
                    win[p - 1, r * 15 + c] = 1
                if ai._evaluate_threat_level(t, p) >= 3:
                    three[p - 1, r * 15 + c] = 1
                status_after[p - 1, r * 15 + c] = (0 if not t.game_over else (3 if t.winner is None else t.winner))
    rec["win_cells"] = win
    rec["three_cells"] = three
    rec["status_after"] = status_after
    return rec


def _init_board_task():
    _board_task.ai = ai_agent.AlphaZeroGomokuAI(1, "easy", device="cpu", planner_steps=0)
    _board_task.ks = bg_planner.KnowledgeSearch(15)


def part_boards(pool, outdir):
    pos, _ = load_positions(outdir)
    # also every prefix of a few sequences, so that make_move trajectories (win / draw detection at
    # each ply) are covered, including the 200-stone draw
    rng = np.random.default_rng(7)
    extra = []
    for n in (199, 198):
        s = dense_pattern_moves(rng, n)
        b = board_from_moves(s)
        for mv in b.get_valid_moves()[:6]:
            extra.append(s + [mv[0] * 15 + mv[1]])
            if n == 198:
                base = extra[-1]
                b2 = board_from_moves(base)
                for mv2 in ([] if b2.game_over else b2.get_valid_moves()[:3]):
                    extra.append(base + [mv2[0] * 15 + mv2[1]])
    allpos = pos + extra
    recs = pool.map(_board_task, allpos, chunksize=4)
    out = {k: np.stack([np.asarray(r[k]) for r in recs]) for k in recs[0]}
    out["moves"] = pad_moves(allpos, 201)
    np.savez_compressed(os.path.join(outdir, "boards.npz"), **out)
    print("boards:", len(allpos))


# ----------------------------------------------------------------------------- part: tiers
def _tier_task(seq):
    b = board_from_moves(seq)
    if b.game_over:
        return None
    ai = _board_task.ai
    seen = []

    def rec_choice(s):
        seen.append(list(s))
        return s[0]
    random.choice = rec_choice
    mv = ai._select_offensive_move(b, b.get_valid_moves())
    det_unpatch()
    cand = np.zeros(225, np.uint8)
    if seen:
        for (r, c) in seen[0]:
            cand[r * 15 + c] = 1
    return dict(cells=cells_of(b), to_move=b.current_player, move_count=b.get_move_count(),
                move=mv[0] * 15 + mv[1], from_choice=int(bool(seen)), cand=cand)


def part_tiers(pool, outdir):
    pos, _ = load_positions(outdir)
    recs = [r for r in pool.map(_tier_task, pos, chunksize=2) if r is not None]
    out = {k: np.stack([np.asarray(r[k]) for r in recs]) for k in recs[0]}
    np.savez_compressed(os.path.join(outdir, "tiers.npz"), **out)
    print("tiers:", len(recs))


# ----------------------------------------------------------------------------- part: rollout (DET-A / DET-B)
def _rollout_task(args):
    seq, planner_steps, ai_player = args
    b = board_from_moves(seq)
    if b.game_over:
        return None
    det_patch()
    ai = make_ai(ai_player, "hard", planner_steps=planner_steps)
    final = {}
    orig = ai._get_terminal_value

    def spy(board):
        final["cells"] = cells_of(board)
        final["move_count"] = board.get_move_count()
        final["game_over"] = int(board.game_over)
        final["winner"] = 0 if board.winner is None else int(board.winner)
        return orig(board)
    ai._get_terminal_value = spy

    class N:  # the only attributes _simulate reads (ai_agent.py:258-261)
        board = b

        @staticmethod
        def is_terminal():
            return b.game_over
    v = ai._simulate(N)
    det_unpatch()
    return dict(start_cells=cells_of(b), to_move=b.current_player, move_count=b.get_move_count(),
                ai_player=ai_player, planner_steps=planner_steps, value=float(v),
                final_cells=final["cells"], final_move_count=final["move_count"],
                final_game_over=final["game_over"], final_winner=final["winner"],
                steps=final["move_count"] - b.get_move_count())


def part_rollout(pool, outdir):
    pos, tags = load_positions(outdir)
    live = [p for p in pos if not board_from_moves(p).game_over]
    sel = live[::2]
    tasks = [(p, 0, 1 + (i % 2)) for i, p in enumerate(sel)]
    recs = [r for r in pool.map(_rollout_task, tasks, chunksize=1) if r is not None]
    out = {k: np.stack([np.asarray(r[k]) for r in recs]) for k in recs[0]}
    np.savez_compressed(os.path.join(outdir, "rollout.npz"), **out)
    print("rollout:", len(recs))


# ----------------------------------------------------------------------------- part: mcts (whole tree)
def _mcts_task(args):
    seq, ai_player, sims, c_puct, beta, planner_steps, difficulty = args
    b = board_from_moves(seq)
    det_patch()
    ai = make_ai(ai_player, difficulty, planner_steps=planner_steps, beta=beta, sims=sims, c_puct=c_puct)
    nodes = []
    orig_init = ai_agent.MCTSNode.__init__

    def spy_init(self, *a, **k):
        nodes.append(self)
        orig_init(self, *a, **k)
    ai_agent.MCTSNode.__init__ = spy_init
    t0 = time.time()
    try:
        best = ai._mcts_search(b, b.get_valid_moves())
    finally:
        ai_agent.MCTSNode.__init__ = orig_init
        det_unpatch()
    idx = {id(n): i for i, n in enumerate(nodes)}
    n = len(nodes)
    parent = np.array([-1 if x.parent is None else idx[id(x.parent)] for x in nodes], np.int32)
    move = np.array([-1 if x.move is None else x.move[0] * 15 + x.move[1] for x in nodes], np.int32)
    visits = np.array([x.visits for x in nodes], np.int32)
    value = np.array([x.value for x in nodes], np.float64)
    term = np.array([int(x.board.game_over) for x in nodes], np.uint8)
    pat = np.array([ai.knowledge_search._pattern_score(x.board, ai.player) for x in nodes], np.float64)
    return dict(cells=cells_of(b), to_move=b.current_player, move_count=b.get_move_count(),
                ai_player=ai_player, sims=sims, c_puct=c_puct, beta=beta, planner_steps=planner_steps,
                difficulty=difficulty, n_nodes=n, parent=parent, move=move, visits=visits, value=value,
                terminal=term, pattern=pat, best_move=best[0] * 15 + best[1], seconds=time.time() - t0)


def _save_trees(recs, path):
    nmax = max(r["n_nodes"] for r in recs)
    out = {}
    for k in recs[0]:
        v0 = recs[0][k]
        if isinstance(v0, np.ndarray) and k in ("parent", "move", "visits", "value", "terminal", "pattern"):
            fill = -1 if k in ("parent", "move") else 0
            arr = np.full((len(recs), nmax), fill, dtype=v0.dtype)
            for i, r in enumerate(recs):
                arr[i, :r["n_nodes"]] = r[k]
            out[k] = arr
        else:
            out[k] = np.stack([np.asarray(r[k]) for r in recs])
    np.savez_compressed(path, **out)


def part_mcts(pool, outdir):
    pos, tags = load_positions(outdir)
    live = [(p, t) for p, t in zip(pos, tags) if not board_from_moves(p).game_over]
    tasks = []
    for i, (p, t) in enumerate(live):
        n = len(p)
        mover = 1 + (n % 2)
        ai_player = mover if i % 4 else 3 - mover          # mostly the natural mover, sometimes not
        if n >= 150:      # late game: E+1 < S so the UCB scan runs at the root and below
            tasks.append((p, ai_player, 200 if i % 2 else 100, 1.8 if i % 2 else 1.6, 0.2, 0, "hard"))
        elif n >= 90:
            tasks.append((p, ai_player, 160, 1.6, 0.2, 0, "medium"))
        elif n >= 40 and i % 2 == 0:
            tasks.append((p, ai_player, 40, 1.2, 0.2, 0, "easy"))
        elif i % 3 == 0:
            tasks.append((p, ai_player, 12, 1.8, 0.2, 0, "hard"))
    # a few with beta = 0 and a large beta, to separate the UCB term from the BG bonus
    dense = [p for p, t in live if t == "dense" and len(p) >= 170]
    tasks.append((dense[0], 1, 150, 1.6, 0.0, 0, "medium"))
    tasks.append((dense[1], 2, 150, 1.6, 5.0, 0, "medium"))
    tasks.append((dense[2], 1, 400, 1.8, 0.2, 0, "hard"))
    tasks.sort(key=lambda t: -t[2] * (226 - len(t[0])))
    recs = pool.map(_mcts_task, tasks, chunksize=1)
    _save_trees(recs, os.path.join(outdir, "mcts.npz"))
    print("mcts:", len(recs), "cpu-seconds", sum(r["seconds"] for r in recs))


# ----------------------------------------------------------------------------- part: planner (DET-B)
def _planner_task(args):
    seq, ai_player, difficulty = args
    b = board_from_moves(seq)
    if b.game_over:
        return None
    det_patch()
    pl = bg_planner.BGPlannerAI(ai_player, difficulty, device="cpu")
    load_weights(pl.graph_net, "graphnet")
    load_weights(pl.opp_dqn, "dqn")
    ks = pl.k_search
    scores = np.full(225, np.nan, np.float64)
    for (r, c) in b.get_valid_moves():
        scores[r * 15 + c] = ks.score_move(b, (r, c), ai_player)
    k = pl.params["k"]
    top = ks.top_k_moves(b, ai_player, k=k)
    topi = np.full(16, -1, np.int32)
    topi[:len(top)] = [r * 15 + c for r, c in top]
    planes = pl._board_to_planes(b.get_board_state())
    x = torch.from_numpy(planes).unsqueeze(0)
    with torch.no_grad():
        logits = pl.graph_net(x).squeeze(0)
        policy = torch.softmax(logits, dim=0).numpy()
        q = pl.opp_dqn(x).squeeze(0).numpy()
    a = pl.params["mix_alpha"]
    composed = np.full(16, np.nan, np.float64)
    for j, (r, c) in enumerate(top):
        composed[j] = a * float(policy[r * 15 + c]) - (1 - a) * float(q[r * 15 + c])
    mv = pl.get_move(b)
    det_unpatch()
    return dict(cells=cells_of(b), to_move=b.current_player, move_count=b.get_move_count(),
                ai_player=ai_player, difficulty=difficulty, k=k, scores=scores, topk=topi,
                gn_logits=logits.numpy().astype(np.float32), gn_policy=policy.astype(np.float32),
                dqn_q=q.astype(np.float32), composed=composed, move=mv[0] * 15 + mv[1])


def part_planner(pool, outdir):
    pos, tags = load_positions(outdir)
    tasks = []
    for i, p in enumerate(pos):
        tasks.append((p, 1 + (i % 2), ("hard", "medium", "easy")[i % 3]))
        if tags[i] == "tactical":
            tasks.append((p, 2 - (i % 2), "hard"))
    recs = [r for r in pool.map(_planner_task, tasks, chunksize=2) if r is not None]
    out = {k: np.stack([np.asarray(r[k]) for r in recs]) for k in recs[0]}
    np.savez_compressed(os.path.join(outdir, "planner.npz"), **out)
    print("planner:", len(recs))


# ----------------------------------------------------------------------------- part: nets
def part_nets(pool, outdir):
    pos, tags = load_positions(outdir)
    sel = [p for p in pos][::5][:40]
    gm = neural_network.GomokuModel(device="cpu")
    load_weights(gm.model, "pv")
    gn = bg_planner.GraphNet(15); load_weights(gn, "graphnet")
    dq = bg_planner.OpponentDQN(15); load_weights(dq, "dqn")
    rec = dict(cells=[], logits=[], value=[], probs=[], gn_logits=[], dqn_q=[], pred_value=[])
    for p in sel:
        b = board_from_moves(p)
        st = b.get_board_state()
        x = torch.from_numpy(b.get_board_tensor()).unsqueeze(0)
        with torch.no_grad():
            lg, v = gm.model(x)
            rec["gn_logits"].append(gn(x).squeeze(0).numpy())
            rec["dqn_q"].append(dq(x).squeeze(0).numpy())
        pol, val = gm.predict(st)
        rec["cells"].append(cells_of(b))
        rec["logits"].append(lg.squeeze(0).numpy())
        rec["value"].append(np.float32(v.item()))
        rec["probs"].append(pol.astype(np.float32))
        rec["pred_value"].append(np.float32(val))
    np.savez_compressed(os.path.join(outdir, "nets.npz"), **{k: np.stack(v) for k, v in rec.items()})
    print("nets:", len(sel))


# ----------------------------------------------------------------------------- part: mctsb (DET-B)
def part_mctsb(pool, outdir):
    pos, tags = load_positions(outdir)
    live = [(p, t) for p, t in zip(pos, tags) if not board_from_moves(p).game_over]
    ro = [(p, 5 if i % 2 else 2, 1 + (i % 2)) for i, (p, t) in enumerate(live[::4])]
    recs = [r for r in pool.map(_rollout_task, ro, chunksize=1) if r is not None]
    out = {k: np.stack([np.asarray(r[k]) for r in recs]) for k in recs[0]}
    np.savez_compressed(os.path.join(outdir, "rollout_planner.npz"), **out)
    print("rollout_planner:", len(recs))
    tasks = []
    late = [p for p, t in live if len(p) >= 150]
    mid = [p for p, t in live if 40 <= len(p) < 150]
    early = [p for p, t in live if len(p) < 40]
    for i, p in enumerate(late[:10]):
        tasks.append((p, 1 + (len(p) % 2), 60, 1.8, 0.2, 5, "hard"))
    for i, p in enumerate(mid[:6]):
        tasks.append((p, 1 + (len(p) % 2), 10, 1.6, 0.2, 3, "medium"))
    for i, p in enumerate(early[:6]):
        tasks.append((p, 1 + (len(p) % 2), 5, 1.8, 0.2, 5, "hard"))
    recs = pool.map(_mcts_task, tasks, chunksize=1)
    _save_trees(recs, os.path.join(outdir, "mcts_planner.npz"))
    print("mcts_planner:", len(recs), "cpu-seconds", sum(r["seconds"] for r in recs))


# ----------------------------------------------------------------------------- part: stoch
def _stoch_task(args):
    seq, seed, n = args
    b = board_from_moves(seq)
    det_unpatch()
    random.seed(seed)
    ai = make_ai(b.current_player, "hard", planner_steps=0)
    counts = np.zeros(4, np.int64)      # win, loss, draw, cut-off  (from the AI = mover's view)
    steps = 0

    class N:
        board = b

        @staticmethod
        def is_terminal():
            return False
    orig = ai._get_terminal_value
    last = {}

    def spy(board):
        last["mc"] = board.get_move_count()
        return orig(board)
    ai._get_terminal_value = spy
    for _ in range(n):
        v = ai._simulate(N)
        counts[{1.0: 0, -1.0: 1, 0.1: 2, 0.0: 3}[v]] += 1
        steps += last["mc"] - b.get_move_count()
    return dict(cells=cells_of(b), to_move=b.current_player, move_count=b.get_move_count(),
                counts=counts, total_steps=steps, n=n)


def part_stoch(pool, outdir):
    pos, tags = load_positions(outdir)
    pick = [p for p, t in zip(pos, tags) if t in ("policy", "cluster") and 12 <= len(p) <= 60][:6]
    pick += [p for p, t in zip(pos, tags) if t == "dense" and 120 <= len(p) <= 171][:2]
    tasks = []
    for i, p in enumerate(pick):
        for s in range(4):                       # 4 shards per position -> merged below
            tasks.append((p, 9000 + 10 * i + s, 40))
    recs = pool.map(_stoch_task, tasks, chunksize=1)
    merged = []
    for i in range(len(pick)):
        rs = recs[4 * i:4 * i + 4]
        m = dict(rs[0])
        m["counts"] = sum(r["counts"] for r in rs)
        m["total_steps"] = sum(r["total_steps"] for r in rs)
        m["n"] = sum(r["n"] for r in rs)
        merged.append(m)
    out = {k: np.stack([np.asarray(r[k]) for r in merged]) for k in merged[0]}
    np.savez_compressed(os.path.join(outdir, "stoch.npz"), **out)
    print("stoch:", len(merged))


# ----------------------------------------------------------------------------- part: dataset
def part_dataset(pool, outdir):
    """training.py:77-134 run as shipped: a SimpleReplay filled ply by ply from the shared position list (every
    sequence = one game: planes before the move, move index, mover), finalize_with_winner, then
    GomokuSelfPlayDataset(replay, True, 0.35) under random.seed(1234).  random.sample is wrapped only to record
    which records it chose.  Planes are saved as uint8 (they are 0/1)."""
    import training
    seqs, _tags = load_positions(outdir)
    stones, move, mover, outcome, ply, gid = [], [], [], [], [], []
    states, out_moves, out_values, chosen_all, n_all = [], [], [], [], []
    for g, seq in enumerate(seqs[:40:3] + seqs[-6:]):
        seq = seq[:60]
        board = gomoku_board.GomokuBoard()
        replay = training.SimpleReplay()
        for k, mv in enumerate(seq):
            if board.game_over:
                break
            r, c = divmod(mv, 15)
            cells = np.array(board.board, dtype=np.uint8).reshape(225)
            bits = np.zeros((2, 256), np.uint8)
            for colour in (1, 2):
                rr, cc = np.nonzero(cells.reshape(15, 15) == colour)
                bits[colour - 1, 16 * rr + cc] = 1
            stones.append(np.packbits(bits, axis=1, bitorder="little").view("<u4").reshape(2, 8))
            replay.add(board.get_board_tensor(), mv, board.current_player)
            move.append(mv); mover.append(board.current_player); ply.append(k); gid.append(g)
            assert board.make_move(r, c)
        winner = board.winner if board.game_over else (None if g % 3 else 1 + g % 2)   # unfinished: mix of labels
        replay.finalize_with_winner(winner)
        outcome.extend(replay.outcomes)
        picked = []
        real_sample = random.sample
        def recording_sample(population, k):   # noqa: E306
            out = real_sample(population, k)
            picked.append(list(out))
            return out
        random.seed(1234 + g)
        random.sample = recording_sample
        try:
            ds = training.GomokuSelfPlayDataset(replay, use_augmentation=True, augment_ratio=0.35)
        finally:
            random.sample = real_sample
        states.append(np.stack([x for x, _, _ in ds.samples]).astype(np.uint8))
        out_moves.append(np.array([y for _, y, _ in ds.samples], np.int64))
        out_values.append(np.array([v for _, _, v in ds.samples], np.float32))
        chosen_all.append(np.array(picked[0], np.int32)); n_all.append(len(replay))
    rec = np.zeros(len(move), dtype=[("stones", "<u4", (2, 8)), ("move", "u1"), ("mover", "u1"), ("outcome", "i1"),
                                     ("ply", "u1"), ("game_id", "<u4")])
    rec["stones"] = np.stack(stones); rec["move"] = move; rec["mover"] = mover; rec["outcome"] = outcome
    rec["ply"] = ply; rec["game_id"] = gid
    np.savez_compressed(os.path.join(outdir, "dataset.npz"), records=rec.view(np.uint8).reshape(len(move), 72),
                        game_n=np.array(n_all, np.int32), chosen=np.concatenate(chosen_all),
                        chosen_n=np.array([len(c) for c in chosen_all], np.int32),
                        planes=np.concatenate(states), move=np.concatenate(out_moves),
                        value=np.concatenate(out_values))
    print("dataset: %d games, %d records, %d samples" % (len(n_all), len(move), sum(len(x) for x in out_moves)))



# ----------------------------------------------------------------------------- part: arena
def part_arena(pool, outdir):
    """training.py:221-270 evaluate_model run as shipped under the DET patches (random.choice -> seq[0],
    random.random -> 1.0): the RandomAgent then plays the first empty cell in raster order.  Every AlphaZeroGomokuAI
    the function builds gets time_limit=inf and the shared synthetic planner weights (the reference draws fresh
    random planner nets per AI, bg_planner.py:210-211); GomokuBoard is wrapped only to log the moves.
    evaluate_model overrides num_simulations of the opponent only when a baseline is given (:241-246) and side A
    keeps its preset (100 for "easy"); to keep this affordable the wrapper lowers every new AI's preset to PRESET_SIMS,
    so with a baseline A searches PRESET_SIMS and B eval_num_sim simulations -- the asymmetry is what is pinned."""
    PRESET_SIMS = 5
    import training
    real_init = ai_agent.AlphaZeroGomokuAI.__init__

    def init(self, *a, **k):
        k["time_limit"] = float("inf")
        real_init(self, *a, **k)
        load_weights(self.bg_planner.graph_net, "graphnet")
        load_weights(self.bg_planner.opp_dqn, "dqn")
        self.params["num_simulations"] = PRESET_SIMS

    games = []

    class LoggedBoard(GomokuBoard):
        def __init__(self, *a, **k):
            super().__init__(*a, **k)
            self.log = []
            games.append(self)

        def make_move(self, r, c):
            ok = super().make_move(r, c)
            if ok:
                self.log.append(r * 15 + c)
            return ok

    cases = [(3, 0, 4, 0), (4, 1, 2, 0), (3, 0, 2, 1), (4, 2, 2, 1)]   # sims, planner_steps, games, baseline?
    if os.environ.get("GZ_ARENA_CASES"):   # subset (the engine-vs-engine cases cost the reference ~25-60 minutes each)
        cases = [cases[int(i)] for i in os.environ["GZ_ARENA_CASES"].split(",")]
    out = dict(cases=np.array([c + (PRESET_SIMS,) for c in cases], np.int32), moves=[], winner=[], n_moves=[], case_of=[], result=[])
    ai_agent.AlphaZeroGomokuAI.__init__ = init
    training.GomokuBoard = LoggedBoard
    det_patch()
    try:
        for ci, (sims, plans, ng, base) in enumerate(cases):
            del games[:]
            cur = neural_network.GomokuModel(device="cpu")
            load_weights(cur.model, "pv")
            t0 = time.time()
            res = training.evaluate_model(cur, cur if base else None, torch.device("cpu"), games=ng,
                                          eval_difficulty="easy", eval_num_sim=sims, eval_plans=plans)
            print("arena case", ci, res, "%.0f s" % (time.time() - t0), flush=True)
            out["result"].append([res["wins"], res["losses"], res["draws"]])
            for g in games:
                out["moves"].append(g.log); out["n_moves"].append(len(g.log))
                out["winner"].append(0 if g.winner is None else g.winner); out["case_of"].append(ci)
    finally:
        det_unpatch()
        ai_agent.AlphaZeroGomokuAI.__init__ = real_init
        training.GomokuBoard = GomokuBoard
    np.savez_compressed(os.path.join(outdir, "arena.npz"), cases=out["cases"], moves=pad_moves(out["moves"], 225),
                        n_moves=np.array(out["n_moves"], np.int32), winner=np.array(out["winner"], np.int32),
                        case_of=np.array(out["case_of"], np.int32), result=np.array(out["result"], np.int32))
    print("arena: %d games" % len(out["moves"]))


# ----------------------------------------------------------------------------- part: update
def part_update(pool, outdir):
    """training.py:277-337 train_epoch / validate_epoch run as shipped on the first 640 samples of dataset.npz
    (train: 512, validation: 128; DataLoader batch 128, shuffle=False), the policy-value net loaded with the shared
    synthetic weights, Adam(8e-4, wd 1e-5) + clip 0.8 + StepLR(2, 0.85) as training.py:387-388.  Two epochs; saves
    the losses and every parameter / buffer after the update (float32; large tensors subsampled)."""
    import contextlib
    import io
    import training
    from torch.utils.data import DataLoader, TensorDataset
    z = np.load(os.path.join(outdir, "dataset.npz"))
    x = torch.from_numpy(z["planes"][:640].astype(np.float32))
    y = torch.from_numpy(z["move"][:640])
    v = torch.from_numpy(z["value"][:640]).reshape(-1, 1)
    gm = neural_network.GomokuModel(device="cpu")
    load_weights(gm.model, "pv")
    opt = torch.optim.Adam(gm.model.parameters(), lr=8e-4, weight_decay=1e-5)
    sched = torch.optim.lr_scheduler.StepLR(opt, step_size=2, gamma=0.85)
    tr = DataLoader(TensorDataset(x[:512], y[:512], v[:512]), batch_size=128, shuffle=False)
    va = DataLoader(TensorDataset(x[512:], y[512:], v[512:]), batch_size=128, shuffle=False)
    tl, vl = [], []
    with contextlib.redirect_stdout(io.StringIO()):
        for ep in range(2):
            tl.append(training.train_epoch(gm, tr, opt, torch.device("cpu"), grad_clip=0.8, epoch_index=ep + 1, num_epochs=2))
            vl.append(training.validate_epoch(gm, va, torch.device("cpu"), epoch_index=ep + 1, num_epochs=2))
        sched.step()
    gm.eval_mode()
    with torch.no_grad():      # the updated network as a function: outputs on the validation samples
        val_logits, val_value = gm.model(x[512:])
    sd = gm.model.state_dict()
    # tensors above 4096 elements are stored as every 37th element of the flattened tensor (fixture size)
    out = {"param/" + k: (t.numpy() if t.numel() <= 4096 else t.numpy().reshape(-1)[::37]) for k, t in sd.items()}
    np.savez_compressed(os.path.join(outdir, "update.npz"), train_loss=np.array(tl, np.float64),
                        val_loss=np.array(vl, np.float64), val_logits=val_logits.numpy(),
                        val_value=val_value.numpy().reshape(-1), lr=np.array([g["lr"] for g in opt.param_groups]), **out)
    print("update: train", tl, "val", vl)


# ----------------------------------------------------------------------------- part: dropin
# The reference's CLASS surface, run unmodified: AlphaZeroGomokuAI.get_move (book and search, README
# presets through the `ai.params[...] = ` idiom of training.py:240), GomokuModel.predict on both input forms,
# evaluate_position, MCTSNode.prior_prob, BGPlannerAI.get_move with the real Mersenne-Twister stream, and a
# checkpoint written by GomokuModel.save_model.  The drop-in classes are tested against these on the GPU box.
WEIGHT_SEED_B = 4321      # a second policy-value model, for the `ai.model = other` re-pack test


def _getmove_task(args):
    seq, player, difficulty, sims, c_puct, planner_steps = args
    b = board_from_moves(seq)
    det_patch()
    ai = ai_agent.AlphaZeroGomokuAI(player, difficulty, device="cpu", beta=0.2, planner_steps=planner_steps,
                                    time_limit=float("inf"))
    load_weights(ai.model.model, "pv")
    load_weights(ai.bg_planner.graph_net, "graphnet")
    load_weights(ai.bg_planner.opp_dqn, "dqn")
    ai.params["num_simulations"] = int(sims)     # the reference's own idiom (training.py:240,246)
    ai.params["c_puct"] = float(c_puct)
    nodes = []
    orig_init = ai_agent.MCTSNode.__init__

    def spy_init(self, *a, **k):
        nodes.append(self)
        orig_init(self, *a, **k)
    ai_agent.MCTSNode.__init__ = spy_init
    t0 = time.time()
    before = cells_of(b).copy()
    try:
        mv = ai.get_move(b)
    finally:
        ai_agent.MCTSNode.__init__ = orig_init
        det_unpatch()
    assert np.array_equal(before, cells_of(b))           # get_move never mutates the caller's board
    cv = np.zeros(225, np.int32); cx = np.zeros(225, np.float64)
    for n in nodes:
        if n.parent is nodes[0]:
            cv[n.move[0] * 15 + n.move[1]] = n.visits
            cx[n.move[0] * 15 + n.move[1]] = n.value
    return dict(cells=cells_of(b), to_move=b.current_player, move_count=b.get_move_count(), player=player,
                difficulty=difficulty, sims=sims, c_puct=c_puct, planner_steps=planner_steps,
                move=mv[0] * 15 + mv[1], child_visits=cv, child_value=cx, n_nodes=len(nodes),
                games_played=ai.games_played, seconds=time.time() - t0)


def _book_records():
    """get_move below six stones (ai_agent.py:138-166): the candidate list handed to random.choice and the DET move."""
    ai = ai_agent.AlphaZeroGomokuAI(1, "easy", device="cpu", planner_steps=0, time_limit=float("inf"))
    seqs = [[]]
    b = GomokuBoard()
    det_patch()
    for _ in range(5):                                   # the reference's own DET opening line
        mv = ai.get_move(b)
        b.make_move(*mv)
        seqs.append(list(seqs[-1]) + [mv[0] * 15 + mv[1]])
    det_unpatch()
    seqs += [[0], [7 * 15 + 7, 0], [6 * 15 + 6, 8 * 15 + 8, 7 * 15 + 6], [0, 14, 210, 224, 100],
             [6 * 15 + 6, 6 * 15 + 7, 6 * 15 + 8, 7 * 15 + 6, 7 * 15 + 8]]
    recs = []
    for seq in seqs:
        b = board_from_moves(seq)
        seen = []

        def rec_choice(s):
            seen.append(list(s))
            return s[0]
        random.choice = rec_choice
        random.random = lambda: 1.0
        mv = ai.get_move(b)
        det_unpatch()
        cand = np.zeros(225, np.uint8)
        for (r, c) in (seen[0] if seen else []):
            cand[r * 15 + c] = 1
        recs.append(dict(cells=cells_of(b), to_move=b.current_player, move_count=b.get_move_count(),
                         move=mv[0] * 15 + mv[1], cand=cand, from_choice=int(bool(seen))))
    return recs


def _planner_stream_task(args):
    seq, player, difficulty, seed, calls = args
    b = board_from_moves(seq)
    det_unpatch()
    pl = ai_agent.AIFactory.create_ai("bg_planner", player, difficulty, device="cpu")
    assert type(pl) is bg_planner.BGPlannerAI
    load_weights(pl.graph_net, "graphnet")
    load_weights(pl.opp_dqn, "dqn")
    random.seed(seed)
    moves = []
    for _ in range(calls):                                # real Mersenne-Twister stream: one random.random() per
        mv = pl.get_move(b)                               # call, random.choice(top_moves) when it explores
        moves.append(mv[0] * 15 + mv[1])
    return dict(cells=cells_of(b), to_move=b.current_player, move_count=b.get_move_count(), player=player,
                difficulty=difficulty, seed=seed, moves=np.array(moves, np.int32))


def part_dropin(pool, outdir):
    import importlib.util
    import shutil
    pos, tags = load_positions(outdir)
    live = [(p, t) for p, t in zip(pos, tags) if not board_from_moves(p).game_over]
    dense = [p for p, t in live if t == "dense"]
    by_len = lambda lo, hi: [p for p in dense if lo <= len(p) <= hi]   # noqa: E731
    mid = [p for p, t in live if t in ("cluster", "centre") and 90 <= len(p) <= 151]
    # ---- get_move through the search: README presets (BASELINE configs 1-3) set through ai.params
    tasks = []
    for i, p in enumerate(by_len(170, 180)[:3]):
        tasks.append((p, 1 + len(p) % 2, "easy", 50, 1.2, 0))
    for i, p in enumerate(by_len(180, 195)[:3]):
        tasks.append((p, 1 + len(p) % 2, "medium", 100, 1.6, 0))
    for i, p in enumerate(by_len(190, 198)[:4]):
        tasks.append((p, 1 + len(p) % 2, "hard", 200, 1.8, 0))
    for i, p in enumerate(by_len(195, 198)[:3]):
        tasks.append((p, 1 + len(p) % 2, "hard", 200, 1.8, 5))     # the ctor default planner_steps
    tasks.append((by_len(190, 194)[0], 1 + len(by_len(190, 194)[0]) % 2, "easy", 50, 1.2, 2))
    tasks.append((by_len(190, 194)[1], 2 - len(by_len(190, 194)[1]) % 2, "medium", 100, 1.6, 3))   # AI colour != mover
    tasks.append((mid[0], 1 + len(mid[0]) % 2, "easy", 20, 1.2, 0))
    tasks.append((mid[-1], 1 + len(mid[-1]) % 2, "hard", 30, 1.8, 0))
    tasks.append((by_len(196, 199)[-1], 1 + len(by_len(196, 199)[-1]) % 2, "hard", 400, 1.8, 0))    # code preset S=400
    res_search = pool.map_async(_getmove_task, tasks, chunksize=1)
    # ---- BGPlannerAI through the factory, real RNG stream
    ptasks = []
    for i, p in enumerate([q for q, t in live if 6 <= len(q) <= 60][::5][:6] + by_len(150, 190)[:2]):
        ptasks.append((p, 1 + (i % 2), ("easy", "medium", "hard")[i % 3], 600 + i, 12))
    res_planner = pool.map_async(_planner_stream_task, ptasks, chunksize=1)

    out = {}
    book = _book_records()
    for k in book[0]:
        out["book_" + k] = np.stack([np.asarray(r[k]) for r in book])

    # ---- predict (both input forms), evaluate_position, prior_prob: models A (seed 1234) and B (seed 4321)
    sel = [p for p, t in live][::7][:10]
    finished = [p for p in pos if board_from_moves(p).game_over][:3]
    rng = np.random.default_rng(11)
    draw = dense_pattern_moves(rng, 200)
    assert board_from_moves(draw).game_over and board_from_moves(draw).winner is None
    ev_boards = sel[:4] + finished + [draw]
    for tag, wseed in (("a", WEIGHT_SEED), ("b", WEIGHT_SEED_B)):
        gm = neural_network.GomokuModel(device="cpu")
        sd = make_state_dict("pv", wseed)
        gm.model.load_state_dict({k: torch.from_numpy(np.asarray(v)) for k, v in sd.items()}, strict=True)
        gm.model.eval()
        pol2, val2, pol3, val3, prior = [], [], [], [], []
        for p in sel:
            b = board_from_moves(p)
            a, v = gm.predict(b.get_board_state())                 # (15,15) ints
            c, w = gm.predict(b.get_board_tensor())                # (3,15,15) planes
            pol2.append(a); val2.append(v); pol3.append(c); val3.append(w)
            node = ai_agent.MCTSNode(b, None, None, gm, {"c_puct": 1.0})
            pr = np.zeros(225, np.float64)
            for (r, c2), x in zip(b.get_valid_moves(), node.prior_prob):
                pr[r * 15 + c2] = x
            prior.append(pr)
        out["pred_policy2d_" + tag] = np.stack(pol2).astype(np.float32)
        out["pred_value2d_" + tag] = np.array(val2, np.float64)
        out["pred_policy3d_" + tag] = np.stack(pol3).astype(np.float32)
        out["pred_value3d_" + tag] = np.array(val3, np.float64)
        out["prior_" + tag] = np.stack(prior)
        ev = []
        for player in (1, 2):
            ai = ai_agent.AlphaZeroGomokuAI(player, "easy", device="cpu", planner_steps=0)
            ai.model = gm                                          # training.py:247-248 idiom
            ev.append([ai.evaluate_position(board_from_moves(p)) for p in ev_boards])
        out["evalpos_" + tag] = np.array(ev, np.float64)
        if tag == "a":
            ck = os.path.join(_SCRATCH, "ckpt", "ref_checkpoint.pth")
            gm.save_model(ck)                                      # neural_network.py:292-306
            shutil.copy(ck, os.path.join(outdir, "ref_checkpoint.pth"))
            # and the other direction: a checkpoint written by the drop-in wrapper, loaded by the reference
            pkg = os.path.join(os.path.dirname(os.path.dirname(HERE)), "alphazero-gomoku_b200")
            sys.path.insert(0, pkg)
            spec = importlib.util.spec_from_file_location("dropin_neural_network", os.path.join(pkg, "neural_network.py"))
            dnn = importlib.util.module_from_spec(spec)
            spec.loader.exec_module(dnn)
            sys.path.remove(pkg)
            dm = dnn.GomokuModel(device="cpu")
            dm.model.load_state_dict({k: torch.from_numpy(np.asarray(v)) for k, v in sd.items()}, strict=True)
            dck = os.path.join(_SCRATCH, "ckpt", "dropin_checkpoint.pth")
            dm.save_model(dck)
            back = neural_network.GomokuModel(model_path=dck, device="cpu")
            b0 = board_from_moves(sel[0])
            pa, va = back.predict(b0.get_board_state())
            out["ref_loads_dropin_checkpoint"] = np.array(
                [float(np.max(np.abs(pa - pol2[0]))), abs(va - val2[0])], np.float64)
            assert out["ref_loads_dropin_checkpoint"].max() == 0.0
    out["pred_cells"] = np.stack([cells_of(board_from_moves(p)) for p in sel])
    eb = [board_from_moves(p) for p in ev_boards]
    out["evalpos_cells"] = np.stack([cells_of(b) for b in eb])
    out["evalpos_to_move"] = np.array([b.current_player for b in eb], np.uint8)
    out["evalpos_move_count"] = np.array([b.get_move_count() for b in eb], np.int32)
    out["evalpos_game_over"] = np.array([int(b.game_over) for b in eb], np.uint8)
    out["evalpos_winner"] = np.array([0 if b.winner is None else b.winner for b in eb], np.uint8)

    recs = res_planner.get()
    for k in recs[0]:
        out["planner_" + k] = np.stack([np.asarray(r[k]) for r in recs])
    recs = res_search.get()
    for k in recs[0]:
        out["search_" + k] = np.stack([np.asarray(r[k]) for r in recs])
    np.savez_compressed(os.path.join(outdir, "dropin.npz"), **out)
    print("dropin: %d book, %d search (%.0f cpu-s), %d planner streams" %
          (len(book), len(tasks), sum(r["seconds"] for r in recs), len(ptasks)))


PARTS = dict(positions=part_positions, boards=part_boards, tiers=part_tiers, rollout=part_rollout,
             mcts=part_mcts, planner=part_planner, nets=part_nets, mctsb=part_mctsb, stoch=part_stoch,
             dataset=part_dataset, arena=part_arena, update=part_update, dropin=part_dropin)


def main():
    import multiprocessing as mp
    ap = argparse.ArgumentParser()
    ap.add_argument("--parts", default="all")
    ap.add_argument("--jobs", type=int, default=os.cpu_count())
    ap.add_argument("--out", default=HERE)
    a = ap.parse_args()
    parts = list(PARTS) if a.parts == "all" else a.parts.split(",")
    with mp.get_context("fork").Pool(a.jobs, initializer=_init_board_task) as pool:
        for p in parts:
            t0 = time.time()
            PARTS[p](pool, a.out)
            print("  [%s] %.1f s" % (p, time.time() - t0), flush=True)


if __name__ == "__main__":
    main()
