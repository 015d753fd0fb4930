"""Dataset build + symmetry augmentation (training.py:38-134).

CPU: the numpy restatement in oracle/oracle.py against fixtures produced by the unmodified reference
(tests/golden/make_goldens.py --parts dataset).  GPU: gz_dataset_build against the fixtures and the oracle."""
import random

import numpy as np
import pytest

import oracle as orc
from conftest import golden


def _games(g):
    """Per game: (records [n,72] u8, chosen, planes, move, value) slices of the fixture."""
    r0 = c0 = s0 = 0
    for n, m in zip(g["game_n"], g["chosen_n"]):
        ns = n + 8 * m
        yield (g["records"][r0:r0 + n], g["chosen"][c0:c0 + m], g["planes"][s0:s0 + ns], g["move"][s0:s0 + ns],
               g["value"][s0:s0 + ns])
        r0, c0, s0 = r0 + n, c0 + m, s0 + ns


def test_oracle_dataset_matches_reference():
    import selfplay
    g = golden("dataset")
    n_games = 0
    for rec, chosen, planes, move, value in _games(g):
        x, mv, _mover, outcome = selfplay.records_to_samples(rec)
        ox, oy, ov = orc.dataset_samples(x, mv, outcome, chosen, reference_exact=True)
        assert np.array_equal(ox, planes.astype(np.float32))
        assert np.array_equal(oy, move)
        assert np.array_equal(ov, value)
        n_games += 1
    assert n_games == len(g["game_n"]) and n_games >= 10


def test_index_transform_as_written_disagrees_with_planes_for_odd_turns():
    # the probed reference bug: put one stone, transform, look where the label points
    for idx in (0, 17, 112, 224, 3 * 15 + 11):
        planes = np.zeros((3, 15, 15), np.float32)
        planes[0].flat[idx] = 1.0
        for k in range(4):
            for flip in (False, True):
                where = int(np.argmax(orc.transform_planes(planes, k, flip)[0]))
                assert orc.transform_index(idx, k, flip, reference_exact=False) == where
                same = orc.transform_index(idx, k, flip, reference_exact=True) == where
                assert same == (k % 2 == 0 or idx == 112)      # the centre is a fixed point


@pytest.fixture(scope="module")
def engine_f32():
    import engine
    e = engine.Engine(max_games=8, max_simulations=4)
    yield e
    e.close()


def _random_records(n, seed):
    rng = np.random.default_rng(seed)
    import engine
    rec = np.zeros(n, dtype=engine.REPLAY_DTYPE)
    for i in range(n):
        k = int(rng.integers(0, 120))
        cells = rng.permutation(225)[:k + 1]
        bits = np.zeros((2, 256), np.uint8)
        for j, c in enumerate(cells[:k]):
            bits[j & 1, 16 * (c // 15) + c % 15] = 1
        rec["stones"][i] = np.packbits(bits, axis=1, bitorder="little").view("<u4").reshape(2, 8)
        rec["move"][i] = cells[k]
        rec["mover"][i] = 1 + (k & 1)
        rec["outcome"][i] = int(rng.integers(-1, 2))
        rec["ply"][i] = k
        rec["game_id"][i] = i
    return rec


@pytest.mark.gpu
def test_gpu_dataset_matches_reference_goldens(engine_f32):
    g = golden("dataset")
    for rec, chosen, planes, move, value in _games(g):
        x, y, v = engine_f32.dataset_build(rec, chosen, reference_exact=True)
        assert np.array_equal(x.cpu().numpy().reshape(-1, 3, 15, 15), planes.astype(np.float32))
        assert np.array_equal(y.cpu().numpy(), move)
        assert np.array_equal(v.cpu().numpy(), value)


@pytest.mark.gpu
@pytest.mark.parametrize("exact", [True, False])
def test_gpu_dataset_matches_oracle_and_gathers(engine_f32, exact):
    import selfplay
    rec = _random_records(300, 5)
    chosen = [7, 7, 0, 299, 150, 31]                      # repeats and both ends
    px, pm, _mover, po = selfplay.records_to_samples(rec.view(np.uint8).reshape(-1, 72))
    ox, oy, ov = orc.dataset_samples(px, pm, po, chosen, reference_exact=exact)
    x, y, v = engine_f32.dataset_build(rec.view(np.uint8).reshape(-1, 72), chosen, reference_exact=exact)
    assert np.array_equal(x.cpu().numpy(), ox) and np.array_equal(y.cpu().numpy(), oy)
    assert np.array_equal(v.cpu().numpy(), ov)
    ids = np.random.default_rng(1).integers(0, len(oy), 257)
    gx, gy, gv = engine_f32.dataset_build(rec.view(np.uint8).reshape(-1, 72), chosen, sample_ids=ids, reference_exact=exact)
    assert np.array_equal(gx.cpu().numpy(), ox[ids]) and np.array_equal(gy.cpu().numpy(), oy[ids])
    assert np.array_equal(gv.cpu().numpy(), ov[ids])
    # ranges and the empty cases
    rx, ry, _ = engine_f32.dataset_build(rec.view(np.uint8).reshape(-1, 72), chosen, first=290, count=30, reference_exact=exact)
    assert np.array_equal(rx.cpu().numpy(), ox[290:320]) and np.array_equal(ry.cpu().numpy(), oy[290:320])
    ex, ey, ev = engine_f32.dataset_build(rec.view(np.uint8).reshape(-1, 72)[:0], None)
    assert ex.shape[0] == 0 and ey.shape[0] == 0 and ev.shape[0] == 0
    nx, _, _ = engine_f32.dataset_build(rec.view(np.uint8).reshape(-1, 72), None)
    assert np.array_equal(nx.cpu().numpy(), ox[:300])


@pytest.mark.gpu
def test_gpu_dataset_class_picks_like_the_reference_and_keeps_symmetry_invariants(engine_f32):
    import selfplay
    rec = _random_records(4096, 9).view(np.uint8).reshape(-1, 72)
    random.seed(77)
    ds = selfplay.GomokuSelfPlayDataset(rec, engine_f32, use_augmentation=True, augment_ratio=0.35, reference_exact=False)
    random.seed(77)
    assert ds.chosen == random.sample(list(range(4096)), k=int(4096 * 0.35))     # training.py:118-119
    assert len(ds) == 4096 + 8 * len(ds.chosen) == ds.planes.shape[0]
    x = ds.planes.reshape(len(ds), 3, 225)
    # every sample: the three planes partition the board; every symmetry keeps the stone counts; the corrected
    # label points at an empty cell (the move is played on the position that was recorded)
    assert bool((x.sum(1) == 1).all())
    aug = x[4096:].reshape(-1, 8, 3, 225).sum(-1)
    assert bool((aug == aug[:, :1]).all())
    idx = ds.move.reshape(-1, 1)
    assert bool((x[:, 2].gather(1, idx) == 1).all())
    xb, yb, vb = ds.batch([0, 4096, len(ds) - 1])
    assert bool((xb == ds.planes[[0, 4096, len(ds) - 1]]).all()) and vb.shape == (3, 1)
    x0, y0, v0 = ds[5]
    assert x0.shape == (3, 15, 15) and y0.dim() == 0 and v0.shape == (1,)
