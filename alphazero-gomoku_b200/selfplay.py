"""Batched self-play on the B200 engine: the caller of the hot path that the reference writes as
`play_one_game` + `SimpleReplay` (training.py:141-218, :77-97), for thousands of games at once.

The game loop itself runs on the device (`gz_selfplay_*`, csrc/gz_kernels_selfplay.cu); this module is
the host-side mirror: starting a session from two reference-style AIs, draining packed replay records,
converting them to the reference's sample format, and all-gathering them across ranks (the only
collective of the whole pipeline, off the hot path -- SURVEY.md 8e).
"""
import numpy as np

import engine as _engine
from engine import REPLAY_DTYPE


def _raster_bits(stones):
    """[n, 8] uint32 pitch-16 bitboards -> [n, 225] uint8 (1 where a stone is)."""
    w = np.ascontiguousarray(stones, dtype="<u4")
    bits = np.unpackbits(w.view(np.uint8).reshape(len(w), 32), axis=1, bitorder="little")   # bit p = 16*r + c
    return bits.reshape(len(w), 16, 16)[:, :15, :15].reshape(len(w), 225)


def records_to_samples(records):
    """Packed records -> the reference's replay sample format (training.py:84-94):
    planes float32 [n,3,15,15] (black, white, empty -- GomokuBoard.get_board_tensor, gomoku_board.py:239-260),
    move index int64 [n], mover int64 [n], outcome int64 [n]."""
    rec = np.asarray(records).view(REPLAY_DTYPE).reshape(-1)
    black = _raster_bits(rec["stones"][:, 0])
    white = _raster_bits(rec["stones"][:, 1])
    planes = np.stack([black, white, 1 - black - white], axis=1).astype(np.float32).reshape(-1, 3, 15, 15)
    return planes, rec["move"].astype(np.int64), rec["mover"].astype(np.int64), rec["outcome"].astype(np.int64)


def split_games(records):
    """Records grouped by game id, each game in ply order (records of a game are stored contiguously)."""
    rec = np.asarray(records).view(REPLAY_DTYPE).reshape(-1)
    games = {}
    for r in rec:
        games.setdefault(int(r["game_id"]), []).append(r)
    return {g: sorted(v, key=lambda r: int(r["ply"])) for g, v in games.items()}


def game_id(rank, games_per_rank, slot, generation):
    """Global id of the generation-th game played in `slot` of `rank` (gz_replay_record.game_id): ranks own
    disjoint slot ranges, recycled games move up by 2^20, so ids never collide below 2^20 slots in total."""
    return rank * games_per_rank + slot + (generation << 20)


class SelfPlay:
    """G concurrent self-play games on one GPU.

    ai: an AlphaZeroGomokuAI (alphazero-gomoku_b200/ai_agent.py); its params / beta / planner_steps /
    model / planner nets configure every search, exactly as the two AIs sharing one model do in
    training.py:403-408.  rank / world place this GPU's games in the global id space."""

    def __init__(self, ai, n_games, rank=0, world=1, seed=0, replay_capacity=0, lanes=0, games_total=0):
        self.ai, self.n_games, self.rank, self.world = ai, int(n_games), int(rank), int(world)
        sims = int(ai.params["num_simulations"])
        self.engine = ai._get_engine(self.n_games, sims)
        p = ai._search_params(game_base=self.rank * self.n_games)
        p.seed = (int(seed) * 0x9E3779B97F4A7C15 + 0xB200) & (2 ** 64 - 1)
        self.params = p
        self.games_total = int(games_total)
        if self.games_total and not replay_capacity:     # every counted game fits: nothing is ever dropped
            replay_capacity = 200 * self.games_total
        self.engine.selfplay_begin(self.n_games, p, float(ai.params["exploration"]), replay_capacity,
                                   self.games_total)

    def step(self, n=1):
        for _ in range(n):
            self.engine.selfplay_step()

    def stats(self):
        return self.engine.selfplay_stats()

    def drain(self):
        """numpy structured array (REPLAY_DTYPE) of the games finished since the last drain."""
        t = self.engine.selfplay_drain()
        return t.cpu().numpy().reshape(-1).view(REPLAY_DTYPE).copy() if t.shape[0] else np.empty(0, REPLAY_DTYPE)

    def drain_device(self):
        return self.engine.selfplay_drain()

    def play_games(self, n_finished=None, max_steps=100000):
        """Steps until the session's games_total games (or, in an open-ended session, at least n_finished games) are
        finished; returns their records.  With games_total the recorded games are the FIRST games_total starts, all
        played to the end, like the reference's `for game in range(N): play_one_game(...)` (training.py:415-420); an
        open-ended session stopped at a count keeps whichever games finished first (short games over-represented)."""
        target = self.games_total or int(n_finished)
        out = []
        for _ in range(max_steps):
            self.step()
            st = self.stats()
            if st["records_ready"] > self.n_games * 128:
                out.append(self.drain())
            if st["games_finished"] >= target:
                break
        out.append(self.drain())
        assert self.stats()["records_dropped"] == 0, "replay buffer overflow: records were dropped"
        return np.concatenate(out)


def gather_replay(records, group=None):
    """All-gather of packed replay records across ranks (uint8 [n,72] tensor, device or host; NCCL on
    GPUs, gloo in the CPU tests).  Ranks may hold different counts: sizes are exchanged first, then the
    payload is padded to the maximum.  Returns the concatenation in rank order on every rank."""
    import torch
    import torch.distributed as dist
    t = records if isinstance(records, torch.Tensor) else torch.from_numpy(
        np.ascontiguousarray(np.asarray(records).view(np.uint8).reshape(-1, REPLAY_DTYPE.itemsize)))
    if not dist.is_initialized() or dist.get_world_size(group) == 1:
        return t
    world = dist.get_world_size(group)
    n = torch.tensor([t.shape[0]], dtype=torch.int64, device=t.device)
    sizes = [torch.zeros_like(n) for _ in range(world)]
    dist.all_gather(sizes, n, group=group)
    sizes = [int(x.item()) for x in sizes]
    m = max(sizes)
    pad = torch.zeros((m, REPLAY_DTYPE.itemsize), dtype=torch.uint8, device=t.device)
    pad[:t.shape[0]] = t
    parts = [torch.empty_like(pad) for _ in range(world)]
    dist.all_gather(parts, pad, group=group)
    return torch.cat([p[:k] for p, k in zip(parts, sizes)], dim=0)


class GomokuSelfPlayDataset:
    """GomokuSelfPlayDataset (training.py:104-134) decoded on the device from packed replay records.

    Same constructor logic as the reference: every record once, plus the 8 symmetries (augment_sample,
    training.py:63-70) of `max(1, int(n * augment_ratio))` records picked by `random.sample(range(n), k)` --
    `rng` defaults to Python's global `random`, so a seeded run picks what the reference picks.
    `reference_exact=True` keeps the reference's index transform as written (it turns the policy label the other
    way round from the planes for odd quarter turns); False labels the cell the stone lands on.
    Samples are decoded by gz_dataset_build: the whole set once (`materialize=True`: tensors .planes, .move,
    .value and the reference's __getitem__) or batch by batch with `batch(sample_ids)`."""

    def __init__(self, records, engine, use_augmentation=True, augment_ratio=0.5, rng=None, reference_exact=True,
                 materialize=True):
        import random as _random
        self.engine, self.reference_exact = engine, bool(reference_exact)
        self.records = records
        self.n = int(records.shape[0]) if hasattr(records, "shape") else len(records)
        self.chosen = []
        if use_augmentation and self.n > 0:
            aug_count = int(self.n * augment_ratio)
            self.chosen = (rng or _random).sample(list(range(self.n)), k=max(1, aug_count))
        self.planes = self.move = self.value = None
        if materialize:
            self.planes, self.move, self.value = engine.dataset_build(records, self.chosen,
                                                                      reference_exact=self.reference_exact)

    def __len__(self):
        return self.n + 8 * len(self.chosen)

    def batch(self, sample_ids):
        """(planes [B,3,15,15], move [B], value [B,1]) for the given sample indices, decoded on demand."""
        x, y, v = self.engine.dataset_build(self.records, self.chosen, sample_ids=sample_ids,
                                            reference_exact=self.reference_exact)
        return x, y, v.reshape(-1, 1)

    def __getitem__(self, idx):
        if self.planes is None:
            x, y, v = self.batch([int(idx)])
            return x[0], y[0], v[0]
        return self.planes[idx], self.move[idx], self.value[idx].reshape(1)
