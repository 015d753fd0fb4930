"""Host side of the self-play path, on CPU: the packed replay record format against the reference's
sample format (training.py:84-94, gomoku_board.py:239-260), and the replay all-gather -- the only
collective of the pipeline -- on a world-size-2 gloo group."""
import os
import sys

import numpy as np
import pytest

from conftest import PKG


def _random_game(seed):
    """A random legal game on the host board, recorded in the packed format."""
    from gomoku_board import GomokuBoard
    import selfplay
    rng = np.random.default_rng(seed)
    b = GomokuBoard()
    recs, planes = [], []
    while not b.game_over:
        r, c = b.get_valid_moves()[rng.integers(len(b.get_valid_moves()))]
        rec = np.zeros((), selfplay.REPLAY_DTYPE)
        for colour in (1, 2):
            for rr, cc in zip(*np.nonzero(b.board == colour)):
                p = 16 * int(rr) + int(cc)
                rec["stones"][colour - 1, p >> 5] |= np.uint32(1 << (p & 31))
        rec["move"], rec["mover"], rec["ply"], rec["game_id"] = r * 15 + c, b.current_player, len(recs), seed
        planes.append(b.get_board_tensor().copy())
        recs.append(rec)
        b.make_move(r, c)
    out = np.array(recs, dtype=selfplay.REPLAY_DTYPE)
    out["outcome"] = [0 if b.winner is None else (1 if m == b.winner else -1) for m in out["mover"]]
    return out, np.stack(planes), b


def test_record_format_matches_reference_samples():
    import selfplay
    assert selfplay.REPLAY_DTYPE.itemsize == 72
    rec, planes, b = _random_game(3)
    p, mv, mover, outcome = selfplay.records_to_samples(rec)
    assert p.dtype == np.float32 and p.shape == (len(rec), 3, 15, 15)
    assert np.array_equal(p, planes)                      # get_board_tensor before each move
    assert np.array_equal(mover, 1 + np.arange(len(rec)) % 2)
    assert set(np.unique(outcome)) <= {-1, 0, 1}
    games = selfplay.split_games(np.concatenate([rec, _random_game(4)[0]]))
    assert sorted(games) == [3, 4] and [int(r["ply"]) for r in games[3]] == list(range(len(rec)))


def _gather_worker(rank, world, port, q):
    sys.path.insert(0, PKG)
    import torch
    import torch.distributed as dist
    import selfplay
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    rec = np.concatenate([_random_game(10 * rank + i)[0] for i in range(rank + 1)])   # ragged: rank 1 holds more
    out = selfplay.gather_replay(rec)
    empty = selfplay.gather_replay(np.empty(0, selfplay.REPLAY_DTYPE) if rank == 0 else rec)   # one empty rank
    q.put((rank, out.numpy().tobytes(), len(rec), empty.shape[0]))
    dist.destroy_process_group()


def test_gather_replay_world_size_2_gloo():
    import torch.multiprocessing as mp
    import selfplay
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + os.getpid() % 2000
    procs = [ctx.Process(target=_gather_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in procs)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    want = np.concatenate([_random_game(0)[0], _random_game(10)[0], _random_game(11)[0]])
    for rank, blob, n_local, n_empty in res:
        got = np.frombuffer(blob, dtype=selfplay.REPLAY_DTYPE)
        assert got.tobytes() == want.tobytes()            # rank order, every rank the same
    assert res[0][3] == res[1][3] == res[1][2]            # an empty rank contributes nothing


def test_game_sharding_ids_are_global():
    """Rank r of W owns slots r*G + g; a recycled game moves up by 2^20: ids never collide across ranks."""
    import selfplay
    G, W = 8192, 8
    ids = [selfplay.game_id(r, G, g, k) for r in range(W) for g in (0, 1, G - 1) for k in range(4)]
    assert len(set(ids)) == len(ids) and max(ids) < 2 ** 32
    assert selfplay.game_id(7, G, G - 1, 0) < 2 ** 20


# ---- data-parallel update (training.run_iteration update="dp"): world-size-2 gloo, CPU ---------------------------------
def _dp_worker(rank, world, port, q):
    from conftest import GOLDEN
    sys.path[:0] = [PKG, GOLDEN]
    import torch
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    torch.set_num_threads(1)
    import training
    from neural_network import GomokuModel
    from weights import make_state_dict
    g = torch.Generator().manual_seed(5)
    n = 96
    x = (torch.rand(n, 3, 15, 15, generator=g) > 0.6).float()
    y = torch.randint(0, 225, (n,), generator=g)
    v = torch.randint(-1, 2, (n,), generator=g).float()

    class Data:
        planes, move, value = x, y, v

        def batch(self, ids):
            i = torch.as_tensor(ids, dtype=torch.long)
            return x[i], y[i], v[i].reshape(-1, 1)
    m = GomokuModel(device="cpu")
    m.model.load_state_dict({k: torch.from_numpy(np.asarray(t)) for k, t in make_state_dict("pv", 1234).items()})
    opt = torch.optim.Adam(m.model.parameters(), lr=8e-4, weight_decay=1e-5)
    order = list(range(n - 3))                                   # ragged last global batch: 93 = 2 * 32 + 29
    mine, local_b = training.shard_batches(order, 32, rank, world)
    loader = training.DeviceLoader(Data(), mine, local_b, shuffle=False)
    tl = training.train_epoch(m, loader, opt, "cpu", 0.8, 1, 1, world, None)
    vl = training.validate_epoch(m, loader, "cpu", 1, 1, world, None)
    q.put((rank, tl, vl, mine, {k: t.numpy().copy() for k, t in m.model.state_dict().items()}))
    dist.destroy_process_group()


def test_data_parallel_update_world_size_2_gloo(weights):
    """Each rank trains on its half of every global batch; after one all-reduce per step both ranks must hold the same
    weights, and they must equal a single process that averages the two half-batch gradients by hand."""
    import torch
    import torch.multiprocessing as mp
    import training
    from neural_network import GomokuModel
    port = 31500 + os.getpid() % 2000
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    ps = [ctx.Process(target=_dp_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in ps:
        p.start()
    res = sorted((q.get(timeout=300) for _ in ps), key=lambda r: r[0])
    for p in ps:
        p.join(timeout=60)
    (_, tl0, vl0, idx0, sd0), (_, tl1, vl1, idx1, sd1) = res
    assert tl0 == tl1 and vl0 == vl1                              # the mean over all ranks' batches, on every rank
    assert len(idx0) == len(idx1) == 16 + 16 + 14 and not set(idx0) & set(idx1)
    assert idx0[:16] == list(range(16)) and idx1[:16] == list(range(16, 32))
    for k in sd0:
        if "running" not in k and "num_batches" not in k:         # BatchNorm statistics are rank-local until averaged
            assert np.array_equal(sd0[k], sd1[k]), k
    # the same three steps by hand in one process
    torch.set_num_threads(1)
    g = torch.Generator().manual_seed(5)
    n = 96
    x = (torch.rand(n, 3, 15, 15, generator=g) > 0.6).float()
    y = torch.randint(0, 225, (n,), generator=g)
    v = torch.randint(-1, 2, (n,), generator=g).float()
    m = GomokuModel(device="cpu")
    m.model.load_state_dict({k: torch.from_numpy(np.asarray(t)) for k, t in weights["pv"].items()})
    twin = GomokuModel(device="cpu")                              # rank 1's BatchNorm statistics evolve separately
    twin.model.load_state_dict(m.model.state_dict())
    opt = torch.optim.Adam(m.model.parameters(), lr=8e-4, weight_decay=1e-5)
    m.train_mode(); twin.train_mode()
    for b in range(3):
        grads = []
        for net, idx in ((m, idx0), (twin, idx1)):
            lo = [0, 16, 32][b]
            ids = torch.as_tensor(idx[lo:lo + (16 if b < 2 else 14)])
            net.model.zero_grad()
            with torch.no_grad():
                for pt, ps_ in zip(twin.model.parameters(), m.model.parameters()):
                    pt.copy_(ps_)
            training._losses(net, x[ids], y[ids], v[ids].reshape(-1, 1)).backward()
            grads.append([p.grad.clone() for p in net.model.parameters()])
        for p, g0, g1 in zip(m.model.parameters(), grads[0], grads[1]):
            p.grad = (g0 + g1) / 2
        torch.nn.utils.clip_grad_norm_(m.model.parameters(), 0.8)
        opt.step()
    for (k, t), p in zip(m.model.named_parameters(), m.model.parameters()):
        assert np.allclose(sd0[k], p.detach().numpy(), rtol=0, atol=1e-6), k
