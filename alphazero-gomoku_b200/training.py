"""The callers either side of the hot path, with the reference's names (training.py there): self-play data
generation, dataset build, policy-value update, evaluation arena -- SURVEY.md 8f rows 1-4.

What differs from the reference is where the loops run.  `play_one_game` / `evaluate_model` there play one game at a
time through `get_move`; here G games advance in lock-step on the device (gz_selfplay_* / gz_arena_*,
csrc/gz_kernels_selfplay.cu) and only counters and packed 72-byte replay records come back.  The update
(`train_epoch`, `validate_epoch`) is the reference's loss and optimiser recipe on stock PyTorch modules -- the
dense work of this path that matters is inference inside the search, SURVEY.md 8f row 3 -- fed by batches that
gz_dataset_build decodes on the device.  The wall-clock time-outs of the reference's loops do not exist here.
"""
import math
import os
import random

import numpy as np
import torch
import torch.nn as nn

import engine as _engine
from ai_agent import AlphaZeroGomokuAI
from gomoku_board import GomokuBoard
from neural_network import GomokuModel
from selfplay import GomokuSelfPlayDataset, SelfPlay, gather_replay, records_to_samples

__all__ = ["SimpleReplay", "GomokuSelfPlayDataset", "Arena", "play_games", "evaluate_model", "DeviceLoader",
           "GraphedTrainStep", "train_epoch", "validate_epoch", "run_iteration", "main"]


class SimpleReplay:
    """training.py:77-97.  Filled by hand like the reference's, or from packed engine records."""

    def __init__(self):
        self.states, self.move_indices, self.players, self.outcomes = [], [], [], []

    def add(self, planes, move_idx, player):
        self.states.append(np.asarray(planes, np.float32))
        self.move_indices.append(int(move_idx))
        self.players.append(int(player))

    def finalize_with_winner(self, winner):
        for p in self.players:
            self.outcomes.append(0 if winner is None else (1 if p == winner else -1))

    def __len__(self):
        return len(self.states)

    @classmethod
    def from_records(cls, records):
        planes, moves, movers, outcomes = records_to_samples(records)
        r = cls()
        r.states, r.move_indices = list(planes), [int(m) for m in moves]
        r.players, r.outcomes = [int(m) for m in movers], [int(o) for o in outcomes]
        return r


def play_games(ai, n_games, concurrent=None, rank=0, world=1, seed=0):
    """The self-play phase of an iteration (training.py:403-428: `games_per_iteration` calls of play_one_game with
    two AIs sharing one model): returns the packed records of at least n_games finished games and the statistics."""
    g = int(concurrent or min(n_games, 4096))
    sp = SelfPlay(ai, g, rank=rank, world=world, seed=seed)
    rec = sp.play_games(n_games)
    return rec, sp.stats()


class Arena:
    """evaluate_model's game loop (training.py:221-270) for `games` games, `concurrent` of them at a time.

    ai configures side A ("current"): difficulty preset, params, beta, planner_steps, planner nets.  Side B is the
    same agent on the other colour (baseline given) or the RandomAgent (baseline None).  rank / world shard the game
    ordinals: rank r plays ordinals r*G + s + k*G*world."""

    def __init__(self, ai, games, vs_random, sims_b=0, concurrent=None, rank=0, world=1, seed=0):
        self.ai, self.games = ai, int(games)
        self.G = int(concurrent or max(1, min(-(-self.games // world), 4096)))
        sims = max(int(ai.params["num_simulations"]), int(sims_b))
        self.engine = ai._get_engine(self.G, sims)
        p = ai._search_params(game_base=rank * self.G)
        p.seed = (int(seed) * 0x9E3779B97F4A7C15 + 0xA2E4A) & (2 ** 64 - 1)
        self.rank, self.world = rank, world
        self.engine.arena_begin(self.G, p, float(ai.params["exploration"]),
                                _engine.ARENA_RANDOM if vs_random else _engine.ARENA_ENGINE,
                                self.games, ordinal_stride=self.G * world, replay_capacity=self.G * 256,
                                sims_b=0 if vs_random else int(sims_b))

    def my_games(self):
        """How many of the `games` ordinals fall to this rank."""
        stride = self.G * self.world
        lo = self.rank * self.G
        return sum(max(0, min(self.G, self.games - (lo + k * stride))) for k in range(-(-self.games // stride)))

    def run(self, max_steps=100000, keep_records=False):
        want, recs = self.my_games(), []
        for _ in range(max_steps):
            st = self.engine.arena_stats()
            if st["games_finished"] >= want:
                break
            self.engine.selfplay_step()
            if keep_records and self.engine.selfplay_stats()["records_ready"] > self.G * 128:
                recs.append(self.engine.selfplay_drain().cpu().numpy())
        st = self.engine.arena_stats()
        if keep_records:
            recs.append(self.engine.selfplay_drain().cpu().numpy())
            rec = np.concatenate(recs).reshape(-1).view(_engine.REPLAY_DTYPE)
            return st, rec
        return st


def evaluate_model(current, baseline, device, games=8, eval_difficulty="easy", eval_num_sim=60, eval_plans=2,
                   deterministic=False, seed=None, precision="bf16", keep_records=False, preset_sims=None,
                   planner_weights=None):
    """training.py:221-270 with the same arguments and result dict.  Differences, all forced by batching: one
    AlphaZeroGomokuAI (one set of random planner nets) serves every game instead of a fresh pair per game, and the
    games run concurrently.  `baseline` only decides the opponent kind: the reference's two AIs differ in `.model`
    alone, which its search never reads.  Kept from the reference: with a baseline only the
    BASELINE's num_simulations becomes eval_num_sim (:241-246) -- the current model keeps its preset (100 for "easy",
    `preset_sims` overrides it for tests); against the RandomAgent the current model searches eval_num_sim (:240)."""
    ai = AlphaZeroGomokuAI(GomokuBoard.BLACK, difficulty=eval_difficulty, device=str(device),
                           planner_steps=eval_plans, deterministic=deterministic, seed=seed, precision=precision)
    if preset_sims is not None:
        ai.params["num_simulations"] = int(preset_sims)
    if baseline is None:
        ai.params["num_simulations"] = eval_num_sim
    ai.model = current
    if planner_weights is not None:
        ai.bg_planner.graph_net.load_state_dict(planner_weights[0], strict=False)
        ai.bg_planner.opp_dqn.load_state_dict(planner_weights[1], strict=False)
    arena = Arena(ai, games, vs_random=baseline is None, sims_b=eval_num_sim, seed=ai.seed)
    out = arena.run(keep_records=keep_records)
    st = out[0] if keep_records else out
    total = max(1, st["wins"] + st["losses"] + st["draws"])
    res = {"wins": st["wins"], "losses": st["losses"], "draws": st["draws"], "win_rate": st["wins"] / total}
    return (res, out[1]) if keep_records else res


# ---------------------------------------------------------------------------------------------- update (8f row 3)
class DeviceLoader:
    """DataLoader(Subset(dataset, indices), batch_size, shuffle) of training.py:446-447 without the host round trip:
    every batch is decoded from the packed records by gz_dataset_build straight into device tensors
    (x [B,3,15,15] f32, y_policy [B] i64, y_value [B,1] f32).  `shuffle` draws a fresh order per epoch from `rng`
    (Python's `random` by default, as the split in training.py:435-439 does)."""

    def __init__(self, dataset, indices, batch_size=128, shuffle=False, rng=None):
        self.dataset, self.indices, self.batch_size = dataset, list(indices), int(batch_size)
        self.shuffle, self.rng = bool(shuffle), rng or random

    def __len__(self):
        return -(-len(self.indices) // self.batch_size)

    def __iter__(self):
        idx = list(self.indices)
        if self.shuffle:
            self.rng.shuffle(idx)
        for i in range(0, len(idx), self.batch_size):
            yield self.dataset.batch(idx[i:i + self.batch_size])


def _losses(model, x, y_policy, y_value):
    policy_logits, value_pred = model.model(x)
    return nn.functional.cross_entropy(policy_logits, y_policy) + nn.functional.mse_loss(value_pred, y_value)


def train_epoch(model, loader, optimizer, device, grad_clip=1.0, epoch_index=1, num_epochs=1):
    """training.py:277-311: CE on the played move + MSE on the outcome, clip_grad_norm_, optimizer.step; returns the
    mean batch loss.  The per-batch `loss.item()` of the reference (a device sync per batch) becomes a float64
    accumulator on the device read once per epoch -- the same sum of float32 losses in float64.

    When the loader is a DeviceLoader over materialized device tensors and the optimizer was built with
    `capturable=True`, the full batches run as replays of ONE captured CUDA graph (GraphedTrainStep): at batch 128 the
    step is ~60 small kernels and launch-bound, the graph removes the launches (same kernels, same arithmetic)."""
    model.train_mode()
    if (isinstance(loader, DeviceLoader) and isinstance(loader.dataset, _Materialized)
            and optimizer.defaults.get("capturable") and torch.device(device).type == "cuda"
            and len(loader.indices) >= loader.batch_size):
        return _train_epoch_graphed(model, loader, optimizer, device, grad_clip)
    total, batches = None, 0
    for x, y_policy, y_value in loader:
        x, y_policy, y_value = x.to(device), y_policy.to(device), y_value.to(device)
        optimizer.zero_grad()
        loss = _losses(model, x, y_policy, y_value)
        loss.backward()
        if grad_clip is not None and grad_clip > 0:
            nn.utils.clip_grad_norm_(model.model.parameters(), grad_clip)
        optimizer.step()
        d = loss.detach().double()
        total = d if total is None else total + d
        batches += 1
    return float(total) / max(1, batches) if batches else 0.0


class GraphedTrainStep:
    """zero_grad + forward + CE/MSE + backward + clip_grad_norm_ + Adam.step of one batch, captured once as a CUDA graph
    and replayed per batch.  The batch is gathered INSIDE the graph from the dataset's device tensors by a static index
    buffer (`idx`), the loss is added to a float64 accumulator inside the graph, so a step costs one small copy and one
    graph launch.  The three warm-up steps PyTorch needs before capture run on a copy of the state that is restored in
    place afterwards (parameters, BatchNorm buffers, Adam moments and step counters), so the captured run starts from
    exactly the state it was given."""

    def __init__(self, model, optimizer, data, batch_size, grad_clip):
        self.model, self.optimizer, self.data, self.grad_clip = model, optimizer, data, grad_clip
        dev = data.planes.device
        self.idx = torch.zeros(batch_size, dtype=torch.long, device=dev)
        self.total = torch.zeros((), dtype=torch.float64, device=dev)
        net = model.model
        tensors = list(net.parameters()) + list(net.buffers())
        saved = [t.detach().clone() for t in tensors]
        had_state = {p: {k: (v.detach().clone() if torch.is_tensor(v) else v) for k, v in st.items()}
                     for p, st in optimizer.state.items()}
        side = torch.cuda.Stream(device=dev)
        side.wait_stream(torch.cuda.current_stream(dev))
        with torch.cuda.stream(side):
            for _ in range(3):
                self._step()
        torch.cuda.current_stream(dev).wait_stream(side)
        with torch.no_grad():                         # in place: the graph must see the same tensors
            for t, s0 in zip(tensors, saved):
                t.copy_(s0)
            for p, st in optimizer.state.items():
                for k, v in st.items():
                    if torch.is_tensor(v):
                        if p in had_state:
                            v.copy_(had_state[p][k])
                        else:
                            v.zero_()
            self.total.zero_()
        optimizer.zero_grad(set_to_none=True)
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph):
            self._step()

    def _step(self):
        d = self.data
        x = d.planes.index_select(0, self.idx)
        y = d.move.index_select(0, self.idx)
        v = d.value.index_select(0, self.idx).reshape(-1, 1)
        self.optimizer.zero_grad(set_to_none=True)
        loss = _losses(self.model, x, y, v)
        loss.backward()
        if self.grad_clip is not None and self.grad_clip > 0:
            nn.utils.clip_grad_norm_(self.model.model.parameters(), self.grad_clip)
        self.optimizer.step()
        self.total += loss.detach().double()

    def run(self, batch_indices):
        self.idx.copy_(batch_indices, non_blocking=True)
        self.graph.replay()


def _train_epoch_graphed(model, loader, optimizer, device, grad_clip):
    idx = list(loader.indices)
    if loader.shuffle:
        loader.rng.shuffle(idx)
    B = loader.batch_size
    data = loader.dataset.d
    perm = torch.as_tensor(idx, dtype=torch.long, device=data.planes.device)
    n_full = len(idx) // B
    step = GraphedTrainStep(model, optimizer, data, B, grad_clip)
    for i in range(n_full):
        step.run(perm[i * B:(i + 1) * B])
    total, batches = step.total.clone(), n_full
    if len(idx) % B:                                   # the ragged last batch runs eagerly
        tail = perm[n_full * B:]
        optimizer.zero_grad()
        loss = _losses(model, data.planes[tail], data.move[tail], data.value[tail].reshape(-1, 1))
        loss.backward()
        if grad_clip is not None and grad_clip > 0:
            nn.utils.clip_grad_norm_(model.model.parameters(), grad_clip)
        optimizer.step()
        total = total + loss.detach().double()
        batches += 1
    return float(total) / max(1, batches)


def validate_epoch(model, loader, device, epoch_index=1, num_epochs=1):
    """training.py:314-337."""
    model.eval_mode()
    total, batches = None, 0
    with torch.no_grad():
        for x, y_policy, y_value in loader:
            x, y_policy, y_value = x.to(device), y_policy.to(device), y_value.to(device)
            d = _losses(model, x, y_policy, y_value).double()
            total = d if total is None else total + d
            batches += 1
    return float(total) / max(1, batches) if batches else 0.0


# ---------------------------------------------------------------------------------------------- one iteration
def _sync_time(device):
    import time
    if torch.device(device).type == "cuda":
        torch.cuda.synchronize(device)
    return time.perf_counter()


def run_iteration(model, optimizer, scheduler, device, games_per_iteration=10, concurrent=None, difficulty="medium",
                  beta=0.2, planner_steps=5, num_simulations=None, c_puct=None, augment_ratio=0.35, train_split=0.9,
                  batch_size=128, epochs=2, grad_clip=0.8, eval_games=6, baseline=None, rank=0, world=1, seed=0,
                  precision="bf16", reference_exact=True, group=None):
    """One pass of training.py:401-508: self-play -> (all-gather) -> dataset with 8-fold augmentation -> 90/10 split ->
    `epochs` x (train_epoch, validate_epoch) -> scheduler.step -> evaluation games.  Returns the losses, the
    evaluation result and the wall time of every phase (device-synchronised).

    Multi-GPU: every rank plays games_per_iteration / world games, the packed records are all-gathered
    (the one collective), and every rank then runs the SAME update on the gathered set -- the reference's
    single-trainer semantics, no gradient exchange; rank 0's weights are broadcast at the end so ranks cannot drift."""
    import torch.distributed as dist
    phases = {}
    t0 = _sync_time(device)
    ai = AlphaZeroGomokuAI(GomokuBoard.BLACK, difficulty=difficulty, device=str(device), beta=beta,
                           planner_steps=planner_steps, seed=seed, precision=precision)
    if num_simulations is not None:
        ai.params["num_simulations"] = int(num_simulations)
    if c_puct is not None:
        ai.params["c_puct"] = float(c_puct)
    ai.model = model
    model.eval_mode()
    mine = -(-int(games_per_iteration) // world)
    sp = SelfPlay(ai, int(concurrent or min(mine, 4096)), rank=rank, world=world, seed=seed)
    # records stay on the device: drained as a uint8 tensor, gathered, decoded by gz_dataset_build
    parts = []
    while True:
        sp.step()
        st = sp.stats()
        if st["records_ready"] > sp.n_games * 128 or st["games_finished"] >= mine:
            parts.append(sp.drain_device())
        if st["games_finished"] >= mine:
            break
    records = torch.cat(parts, dim=0)
    t1 = _sync_time(device)
    phases["selfplay_s"] = t1 - t0
    records = gather_replay(records, group=group) if world > 1 else records
    t2 = _sync_time(device)
    phases["gather_s"] = t2 - t1
    rng = random.Random(1000003 * seed + 17)      # identical on every rank: same subset, same split, same order
    dataset = GomokuSelfPlayDataset(records, sp.engine, use_augmentation=True, augment_ratio=augment_ratio, rng=rng,
                                    reference_exact=reference_exact, materialize=True)
    t3 = _sync_time(device)
    phases["dataset_s"] = t3 - t2
    out = {"games": int(st["games_finished"]), "records": int(records.shape[0]), "samples": len(dataset),
           "selfplay_stats": st}
    if len(dataset) >= 8:
        idx = list(range(len(dataset)))
        rng.shuffle(idx)
        split = int(len(dataset) * train_split)
        train_loader = DeviceLoader(_Materialized(dataset), idx[:split], batch_size, shuffle=True, rng=rng)
        val_loader = DeviceLoader(_Materialized(dataset), idx[split:], batch_size, shuffle=False)
        tl, vl = [], []
        for ep in range(epochs):
            tl.append(train_epoch(model, train_loader, optimizer, device, grad_clip, ep + 1, epochs))
            vl.append(validate_epoch(model, val_loader, device, ep + 1, epochs))
        if scheduler is not None:
            scheduler.step()
        out.update(train_loss=float(np.mean(tl)), val_loss=float(np.mean(vl)))
        if world > 1 and dist.is_initialized():
            for t in list(model.model.parameters()) + list(model.model.buffers()):
                dist.broadcast(t.data, src=0, group=group)
    t4 = _sync_time(device)
    phases["update_s"] = t4 - t3
    if eval_games:
        model.eval_mode()
        out["eval"] = evaluate_model(model, baseline, device, games=eval_games, seed=seed + 1, precision=precision)
    t5 = _sync_time(device)
    phases["eval_s"] = t5 - t4
    out["phases"] = phases
    return out


class _Materialized:
    """dataset.batch() over the tensors gz_dataset_build already decoded (index_select on the device)."""

    def __init__(self, dataset):
        self.d = dataset

    def batch(self, ids):
        i = torch.as_tensor(ids, dtype=torch.long, device=self.d.planes.device)
        return self.d.planes[i], self.d.move[i], self.d.value[i].reshape(-1, 1)


def main(iterations=6, games_per_iteration=10, device=None, models_dir="models", **kw):
    """training.py:361-537: iterations of run_iteration with Adam(8e-4, wd 1e-5), StepLR(2, 0.85), per-iteration /
    best / final checkpoints and early stopping after 3 non-improving iterations."""
    device = torch.device(device or "cuda")
    os.makedirs(models_dir, exist_ok=True)
    model = GomokuModel(model_path=None, board_size=15, device=str(device))
    optimizer = torch.optim.Adam(model.model.parameters(), lr=8e-4, weight_decay=1e-5,
                                 capturable=device.type == "cuda")   # capturable: train_epoch replays a CUDA graph
    scheduler = torch.optim.lr_scheduler.StepLR(optimizer, step_size=2, gamma=0.85)
    best_val, patience, patience_count, snapshot, history = math.inf, 3, 0, None, []
    for it in range(1, iterations + 1):
        res = run_iteration(model, optimizer, scheduler, device, games_per_iteration=games_per_iteration,
                            baseline=snapshot, seed=it, **kw)
        history.append(res)
        model.save_model(os.path.join(models_dir, "alphazero_gomoku_iter_%d.pth" % it))
        if res.get("val_loss", math.inf) < best_val:
            best_val, patience_count = res["val_loss"], 0
            best_path = os.path.join(models_dir, "alphazero_gomoku_best.pth")
            model.save_model(best_path)
            snapshot = GomokuModel(model_path=best_path, board_size=15, device=str(device))
        else:
            patience_count += 1
        print("iteration %d: %s" % (it, {k: v for k, v in res.items() if k != "selfplay_stats"}))
        if patience_count >= patience:
            break
    model.save_model(os.path.join(models_dir, "alphazero_gomoku_final.pth"))
    return history


if __name__ == "__main__":
    main()
