"""gz_pv_forward_static -- the allocation-free variant of the policy-value forward (no counterpart in the reference; added
for callers that cannot call cudaMalloc) -- against the existing gz_pv_forward
through the C-ABI, two calls in a row on different boards (the static scratch is shared and keeps the last call's
contents), plus the size limit.  Collected last on purpose; passes alone in a fresh
process and after the other tests."""
import numpy as np
import pytest

from conftest import golden

pytestmark = pytest.mark.gpu


def test_pv_forward_static_matches_pv_forward_twice_in_a_row(weights):
    import torch
    import engine
    g = golden("nets")
    e = engine.Engine(max_games=64, max_simulations=8, precision=0)
    e.set_pv_weights(weights["pv"])
    n = engine.STATIC_MAX_BOARDS
    batches = [g["cells"][:n], g["cells"][n:2 * n - 3][::-1].copy()]       # 16 boards, then 13 other boards
    out = [tuple(torch.empty(s, dtype=torch.float32, device=e.device) for s in ((len(c), 225), (len(c), 225), (len(c),)))
           for c in batches]
    boards, want = [], []
    for c in batches:
        stones = (c != 0).sum(1)
        b = e.boards_from_cells(c, 1 + stones % 2, stones)
        boards.append(b)
        want.append([t.cpu().numpy() for t in e.pv_forward(b)])             # the existing entry point
    e.pv_forward_static(boards[0], out[0])                                  # first call per device: symbol look-up
    for o in out:                                                           # (in a fresh process torch loads its own
        for t in o:                                                         # fill kernel here: before the baseline)
            t.fill_(float("nan"))
    launches0 = e.counters()["kernel_launches"]
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info()[0]
    e.pv_forward_static(boards[0], out[0])                                  # two calls in a row, nothing in between
    e.pv_forward_static(boards[1], out[1])
    torch.cuda.synchronize()
    assert torch.cuda.mem_get_info()[0] == free0                            # no device memory was taken by the calls
    assert e.counters()["kernel_launches"] == launches0 + 12
    for o, w, c in zip(out, want, batches):
        got = [t.cpu().numpy() for t in o]
        for a, b in zip(got, w):
            assert a.shape == b.shape and np.max(np.abs(a - b)) < 1e-4      # the fp32 tolerance of test_networks_fp32
    k = len(batches[0])
    assert np.max(np.abs(out[0][0].cpu().numpy() - g["logits"][:k])) < 1e-4   # and the reference's own outputs
    assert np.max(np.abs(out[0][2].cpu().numpy() - g["value"][:k])) < 1e-4
    # a precision-1 handle computes the same fp32 result through this entry point
    e1 = engine.Engine(max_games=64, max_simulations=8, precision=1)
    e1.set_pv_weights(weights["pv"])
    o1 = tuple(torch.empty_like(t) for t in out[1])
    e1.pv_forward_static(e1.boards_from_cells(batches[1], 1 + (batches[1] != 0).sum(1) % 2, (batches[1] != 0).sum(1)), o1)
    for a, b in zip(o1, want[1]):
        assert np.max(np.abs(a.cpu().numpy() - b)) < 1e-4
    e1.close()
    e.close()


def test_pv_forward_static_rejects_what_it_cannot_hold(weights):
    import torch
    import engine
    e = engine.Engine(max_games=64, max_simulations=8)
    n = engine.STATIC_MAX_BOARDS + 1
    boards = e.boards_from_cells(np.zeros((n, 225), np.uint8), np.ones(n), np.zeros(n))
    out = (torch.empty((n, 225), device=e.device), torch.empty((n, 225), device=e.device), torch.empty(n, device=e.device))
    with pytest.raises(engine.EngineError, match="gz_set_pv_weights was not called"):
        e.pv_forward_static(boards[:4], tuple(t[:4] for t in out))
    e.set_pv_weights(weights["pv"])
    with pytest.raises(engine.EngineError, match="exceeds GZ_STATIC_MAX_BOARDS"):
        e.pv_forward_static(boards, out)
    e.pv_forward_static(boards[:n - 1], tuple(t[:n - 1] for t in out))      # exactly the maximum is fine
    torch.cuda.synchronize()
    assert bool(torch.isfinite(out[0][:n - 1]).all())
    e.close()
