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
from neural_network import GomokuModel, mark_weights_changed
from selfplay import GomokuSelfPlayDataset, SelfPlay, gather_replay, records_to_samples

__all__ = ["SimpleReplay", "GomokuSelfPlayDataset", "Arena", "play_games", "evaluate_model", "DeviceLoader",
           "GraphedTrainStep", "train_epoch", "validate_epoch", "run_iteration", "main", "shard_batches"]


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


def shard_batches(indices, global_batch, rank, world):
    """Data-parallel split of a sample order: global batch b is indices[b*B:(b+1)*B]; rank r trains on its r-th
    1/world of every global batch, so the mean of the ranks' gradients is the gradient of the global batch's mean loss.
    A ragged last batch is cut down to a multiple of world (at most world - 1 samples are left out).  Returns this
    rank's index list and its local batch size."""
    B = int(global_batch)
    local = B // world
    assert local * world == B, "the global batch must be a multiple of the number of ranks"
    out = []
    for lo in range(0, len(indices), B):
        chunk = indices[lo:lo + B]
        per = len(chunk) // world
        out.extend(chunk[rank * per:(rank + 1) * per])
    return out, local


def _sync_grads(model, world, group):
    """One all-reduce of the flattened fp32 gradient (2.8 MB for the policy-value net), averaged over the ranks.
    Works eagerly and inside a CUDA-graph capture (NCCL collectives are capturable)."""
    import torch.distributed as dist
    grads = [p.grad for p in model.model.parameters() if p.grad is not None]
    flat = torch.cat([g.reshape(-1) for g in grads])
    dist.all_reduce(flat, group=group)
    flat.div_(world)
    off = 0
    for g in grads:
        n = g.numel()
        g.copy_(flat[off:off + n].view_as(g))
        off += n


def _mean_over_ranks(value, count, device, world, group):
    """Sum of per-batch losses and batch count -> mean over every rank's batches (identical on all ranks)."""
    if world > 1:
        import torch.distributed as dist
        t = torch.tensor([float(value), float(count)], dtype=torch.float64, device=device)
        dist.all_reduce(t, group=group)
        value, count = t[0].item(), t[1].item()
    return float(value) / max(1.0, count) if count else 0.0


def _losses(model, x, y_policy, y_value):
    policy_logits, value_pred = model.model(x)
    return nn.functional.cross_entropy(policy_logits, y_policy) + nn.functional.mse_loss(value_pred, y_value)


def train_epoch(model, loader, optimizer, device, grad_clip=1.0, epoch_index=1, num_epochs=1, world=1, group=None):
    """training.py:277-311: CE on the played move + MSE on the outcome, clip_grad_norm_, optimizer.step; returns the
    mean batch loss.  world > 1: data-parallel -- `loader` holds this rank's share of every global batch
    (shard_batches), the gradient is averaged over the ranks before clipping (one all-reduce per step), so every rank
    applies the same update and the weights stay in sync without a broadcast; the returned loss is the mean over all
    ranks' batches.  The per-batch `loss.item()` of the reference (a device sync per batch) becomes a float64
    accumulator on the device read once per epoch -- the same sum of float32 losses in float64.

    When the loader is a DeviceLoader over materialized device tensors and the optimizer was built with
    `capturable=True`, the full batches run as replays of ONE captured CUDA graph (GraphedTrainStep): at batch 128 the
    step is ~60 small kernels and launch-bound, the graph removes the launches (same kernels, same arithmetic)."""
    model.train_mode()
    if (isinstance(loader, DeviceLoader) and isinstance(loader.dataset, _Materialized)
            and optimizer.defaults.get("capturable") and torch.device(device).type == "cuda"
            and len(loader.indices) >= loader.batch_size):
        return _train_epoch_graphed(model, loader, optimizer, device, grad_clip, world, group)
    total, batches = None, 0
    for x, y_policy, y_value in loader:
        x, y_policy, y_value = x.to(device), y_policy.to(device), y_value.to(device)
        optimizer.zero_grad()
        loss = _losses(model, x, y_policy, y_value)
        loss.backward()
        if world > 1:
            _sync_grads(model, world, group)
        if grad_clip is not None and grad_clip > 0:
            nn.utils.clip_grad_norm_(model.model.parameters(), grad_clip)
        optimizer.step()
        d = loss.detach().double()
        total = d if total is None else total + d
        batches += 1
    return _mean_over_ranks(total if total is not None else 0.0, batches, device, world, group)


class GraphedTrainStep:
    """zero_grad + forward + CE/MSE + backward + clip_grad_norm_ + Adam.step of one batch, captured once as a CUDA graph
    and replayed per batch.  The batch is gathered INSIDE the graph from the dataset's device tensors by a static index
    buffer (`idx`), the loss is added to a float64 accumulator inside the graph, so a step costs one small copy and one
    graph launch.  The three warm-up steps PyTorch needs before capture run on a copy of the state that is restored in
    place afterwards (parameters, BatchNorm buffers, Adam moments and step counters), so the captured run starts from
    exactly the state it was given."""

    def __init__(self, model, optimizer, data, batch_size, grad_clip, world=1, group=None):
        self.model, self.optimizer, self.data, self.grad_clip = model, optimizer, data, grad_clip
        self.world, self.group = world, group
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
        if self.world > 1:                             # data-parallel: the all-reduce is a node of the graph
            _sync_grads(self.model, self.world, self.group)
        if self.grad_clip is not None and self.grad_clip > 0:
            nn.utils.clip_grad_norm_(self.model.model.parameters(), self.grad_clip)
        self.optimizer.step()
        self.total += loss.detach().double()

    def run(self, batch_indices):
        self.idx.copy_(batch_indices, non_blocking=True)
        self.graph.replay()


def _train_epoch_graphed(model, loader, optimizer, device, grad_clip, world=1, group=None):
    idx = list(loader.indices)
    if loader.shuffle:
        loader.rng.shuffle(idx)
    B = loader.batch_size
    data = loader.dataset.d
    perm = torch.as_tensor(idx, dtype=torch.long, device=data.planes.device)
    n_full = len(idx) // B
    step = GraphedTrainStep(model, optimizer, data, B, grad_clip, world, group)
    for i in range(n_full):
        step.run(perm[i * B:(i + 1) * B])
    total, batches = step.total.clone(), n_full
    if len(idx) % B:                                   # the ragged last batch runs eagerly
        tail = perm[n_full * B:]
        optimizer.zero_grad()
        loss = _losses(model, data.planes[tail], data.move[tail], data.value[tail].reshape(-1, 1))
        loss.backward()
        if world > 1:
            _sync_grads(model, world, group)
        if grad_clip is not None and grad_clip > 0:
            nn.utils.clip_grad_norm_(model.model.parameters(), grad_clip)
        optimizer.step()
        total = total + loss.detach().double()
        batches += 1
    mark_weights_changed(model.model)                  # graph replays do not bump the tensors' _version
    return _mean_over_ranks(total, batches, device, world, group)


def validate_epoch(model, loader, device, epoch_index=1, num_epochs=1, world=1, group=None):
    """training.py:314-337.  world > 1: `loader` is this rank's share of the validation batches; the mean is taken
    over every rank's batches, so all ranks see the same number (and take the same best-model / early-stop decisions)."""
    model.eval_mode()
    total, batches = None, 0
    with torch.no_grad():
        for x, y_policy, y_value in loader:
            x, y_policy, y_value = x.to(device), y_policy.to(device), y_value.to(device)
            d = _losses(model, x, y_policy, y_value).double()
            total = d if total is None else total + d
            batches += 1
    return _mean_over_ranks(total if total is not None else 0.0, batches, device, world, group)


# ---------------------------------------------------------------------------------------------- one iteration
def _sync_time(device):
    import time
    if torch.device(device).type == "cuda":
        torch.cuda.synchronize(device)
    return time.perf_counter()


def run_iteration(model, optimizer, scheduler, device, games_per_iteration=10, concurrent=None, difficulty="medium",
                  beta=0.2, planner_steps=5, num_simulations=None, c_puct=None, augment_ratio=0.35, train_split=0.9,
                  batch_size=128, epochs=2, grad_clip=0.8, eval_games=6, baseline=None, rank=0, world=1, seed=0,
                  precision="bf16", reference_exact=True, group=None, update=None):
    """One pass of training.py:401-508: self-play -> (all-gather) -> dataset with 8-fold augmentation -> 90/10 split ->
    `epochs` x (train_epoch, validate_epoch) -> scheduler.step -> evaluation games.  Returns the losses, the
    evaluation result and the wall time of every phase (device-synchronised).

    Multi-GPU: every rank plays games_per_iteration / world games and the packed records are
    all-gathered (off the hot path).  update = "dp" (default when world > 1): data-parallel update -- every rank trains
    on its 1/world of each global batch of `batch_size` samples, one all-reduce of the 2.8 MB gradient per step, the
    weights stay in sync by construction, BatchNorm running statistics (per-rank batch statistics) are averaged at the
    end, losses are means over all ranks.  update = "replicated": every rank runs the SAME update on the whole
    gathered set with the global batch (the reference's single-trainer semantics, training.py:277-337,387-393), then
    rank 0 broadcasts weights, buffers and the losses it decided on."""
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
    # exactly `mine` complete games per rank, fixed by start order before any is played (training.py:415-420 plays
    # N whole games): stopping at the first `mine` finishers would keep the short games and never record a 200-ply draw
    n_slots = int(concurrent or min(mine, 4096))
    sp = SelfPlay(ai, min(n_slots, mine), rank=rank, world=world, seed=seed, games_total=mine)
    # records stay on the device: drained as a uint8 tensor, gathered, decoded by gz_dataset_build
    parts = []
    while True:
        sp.step()
        st = sp.stats()
        if st["records_ready"] > sp.n_games * 128 or st["games_finished"] >= mine:
            parts.append(sp.drain_device())
        if st["games_finished"] >= mine:
            break
    assert st["records_dropped"] == 0, "replay buffer overflow: %d records dropped" % st["records_dropped"]
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
    dp = world > 1 and dist.is_initialized() and (update or "dp") == "dp"
    out["update_mode"] = "dp" if dp else ("replicated" if world > 1 else "single")
    if len(dataset) >= 8:
        idx = list(range(len(dataset)))
        rng.shuffle(idx)
        split = int(len(dataset) * train_split)
        data = _Materialized(dataset)
        tl, vl = [], []
        if dp:
            val_idx, local_b = shard_batches(idx[split:], batch_size, rank, world)
            val_loader = DeviceLoader(data, val_idx, local_b, shuffle=False)
            train_idx = idx[:split]
            for ep in range(epochs):
                rng.shuffle(train_idx)                   # the same order on every rank (shared seed)
                mine_idx, local_b = shard_batches(train_idx, batch_size, rank, world)
                loader = DeviceLoader(data, mine_idx, local_b, shuffle=False)
                tl.append(train_epoch(model, loader, optimizer, device, grad_clip, ep + 1, epochs, world, group))
                vl.append(validate_epoch(model, val_loader, device, ep + 1, epochs, world, group))
            with torch.no_grad():                        # BatchNorm running statistics were fed rank-local batches
                for b in model.model.buffers():
                    if b.dtype.is_floating_point:
                        dist.all_reduce(b, group=group)
                        b.div_(world)
        else:
            train_loader = DeviceLoader(data, idx[:split], batch_size, shuffle=True, rng=rng)
            val_loader = DeviceLoader(data, idx[split:], batch_size, shuffle=False)
            for ep in range(epochs):
                tl.append(train_epoch(model, train_loader, optimizer, device, grad_clip, ep + 1, epochs))
                vl.append(validate_epoch(model, val_loader, device, ep + 1, epochs))
            if world > 1 and dist.is_initialized():      # rank 0 decides: weights, buffers and the losses it saw
                for t in list(model.model.parameters()) + list(model.model.buffers()):
                    dist.broadcast(t.data, src=0, group=group)
                lt = torch.tensor(tl + vl, dtype=torch.float64, device=device)
                dist.broadcast(lt, src=0, group=group)
                tl, vl = lt[:len(tl)].tolist(), lt[len(tl):].tolist()
        if scheduler is not None:
            scheduler.step()
        model._packed = None                             # graph replays update parameters without bumping _version
        out.update(train_loss=float(np.mean(tl)), val_loss=float(np.mean(vl)))
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
    eval_games = int(kw.pop("eval_games", 6))
    precision = kw.get("precision", "bf16")
    for it in range(1, iterations + 1):
        # phases one and two; the evaluation comes after the best-model bookkeeping, as in training.py:478-503
        res = run_iteration(model, optimizer, scheduler, device, games_per_iteration=games_per_iteration,
                            eval_games=0, seed=it, **kw)
        history.append(res)
        if "val_loss" not in res:        # fewer than 8 samples: the reference `continue`s -- no checkpoint, no
            continue                     # evaluation, patience untouched (training.py:431-433)
        model.save_model(os.path.join(models_dir, "alphazero_gomoku_iter_%d.pth" % it))
        if res["val_loss"] < best_val:
            best_val, patience_count = res["val_loss"], 0
            best_path = os.path.join(models_dir, "alphazero_gomoku_best.pth")
            model.save_model(best_path)
            snapshot = GomokuModel(model_path=best_path, board_size=15, device=str(device))
        else:
            patience_count += 1
        # six evaluate_model(games=1) calls: the current model is black in every one of them (training.py:489-497)
        model.eval_mode()
        ev = {"wins": 0, "losses": 0, "draws": 0}
        for eg in range(eval_games):
            st = evaluate_model(model, snapshot, device, games=1, seed=1000 * it + eg, precision=precision)
            for k in ev:
                ev[k] += st.get(k, 0)
        ev["win_rate"] = ev["wins"] / max(1, ev["wins"] + ev["losses"] + ev["draws"])
        res["eval"] = ev
        print("iteration %d: %s" % (it, {k: v for k, v in res.items() if k != "selfplay_stats"}))
        if patience_count >= patience:
            break
    model.save_model(os.path.join(models_dir, "alphazero_gomoku_final.pth"))
    return history


if __name__ == "__main__":
    main()
