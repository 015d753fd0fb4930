"""Synthetic positions for benchmarks and large-size property tests:
legal mid-game boards built with numpy only -- uniform random placements for sparse boards, and a
run-free colouring ((r + 2c) mod 4 < 2 -> black; no run longer than two in any direction) for dense
ones, so that no generated board already contains a five."""
import numpy as np

PLIES = (6, 12, 24, 48, 96, 150)


def _dense_cells(rng, n_stones):
    r, c = np.divmod(np.arange(225), 15)
    black = np.flatnonzero((r + 2 * c) % 4 < 2)
    white = np.flatnonzero((r + 2 * c) % 4 >= 2)
    nb, nw = (n_stones + 1) // 2, n_stones // 2
    cells = np.zeros(225, np.uint8)
    cells[rng.permutation(black)[:nb]] = 1
    cells[rng.permutation(white)[:nw]] = 2
    return cells


def _has_five(cells):
    b = cells.reshape(15, 15)
    for colour in (1, 2):
        m = b == colour
        for dr, dc in ((0, 1), (1, 0), (1, 1), (1, -1)):
            acc = np.ones((15, 15), bool)
            for k in range(5):
                sh = np.zeros((15, 15), bool)
                r0, r1 = max(0, -k * dr), min(15, 15 - k * dr)
                c0, c1 = max(0, -k * dc), min(15, 15 - k * dc)
                sh[r0:r1, c0:c1] = m[r0 + k * dr:r1 + k * dr, c0 + k * dc:c1 + k * dc]
                acc &= sh
            if acc.any():
                return True
    return False


def make_positions(n, seed=2026, plies=PLIES):
    """-> (cells uint8[n,225], to_move uint8[n], move_count uint16[n]); board i has plies[i % len]
    stones, black to move on even counts."""
    rng = np.random.default_rng(seed)
    cells = np.zeros((n, 225), np.uint8)
    count = np.zeros(n, np.uint16)
    for i in range(n):
        k = plies[i % len(plies)]
        if k <= 60:
            for _ in range(20):
                c = np.zeros(225, np.uint8)
                pick = rng.permutation(225)[:k]
                c[pick[0::2]] = 1
                c[pick[1::2]] = 2
                if not _has_five(c):
                    break
            else:
                c = _dense_cells(rng, k)
        else:
            c = _dense_cells(rng, k)
        cells[i] = c
        count[i] = k
    to_move = (1 + count % 2).astype(np.uint8)
    return cells, to_move, count
