"""CPU desk-check: the __host__ __device__ functions the CUDA kernels are built from
(alphazero-gomoku_b200/csrc/gz_bitboard.cuh, gz_policy.cuh), compiled for the host, against the
golden fixtures and the oracle.  Bit-exact everywhere (integer / index work)."""
import ctypes as C
import os
import sys

import numpy as np

import oracle as orc
from conftest import golden

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "host_sim"))
from build import lib as hs_lib  # noqa: E402

P = lambda a: a.ctypes.data_as(C.c_void_p)  # noqa: E731


def raster_bits(words):
    return np.unpackbits(np.ascontiguousarray(words, dtype="<u4").view(np.uint8), bitorder="little")[:225]


def board_ops(L, cells, to_move, mc):
    cells = np.ascontiguousarray(cells, np.uint8)
    legal = np.zeros(8, np.uint32); win = np.zeros(16, np.uint32); three = np.zeros(16, np.uint32)
    threat = np.zeros(2, np.int32); pat = np.zeros(2, np.int64)
    st = L.hs_board_ops(P(cells), int(to_move), int(mc), P(legal), P(win), P(three), P(threat), P(pat))
    return st, raster_bits(legal), [raster_bits(win[:8]), raster_bits(win[8:])], \
        [raster_bits(three[:8]), raster_bits(three[8:])], threat, pat


def test_philox_matches_kat():
    L = hs_lib()
    ctr = np.array([0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344], np.uint32)
    key = np.array([0xa4093822, 0x299f31d0], np.uint32)
    out = np.zeros(4, np.uint32)
    L.hs_philox(P(ctr), P(key), P(out))
    assert list(out) == [0xd16cfe09, 0x94fdcceb, 0x5001e420, 0x24126ea1]


def test_board_ops_vs_golden():
    L = hs_lib()
    g = golden("boards")
    for i in range(len(g["cells"])):
        st, legal, win, three, threat, pat = board_ops(L, g["cells"][i], g["to_move"][i], g["move_count"][i])
        want_st = 0 if not g["game_over"][i] else (3 if g["winner"][i] == 0 else g["winner"][i])
        assert st == want_st
        assert np.array_equal(legal, g["legal"][i])
        assert list(threat) == list(g["threat"][i])
        assert [int(x) for x in pat] == [int(x) for x in g["pattern"][i]]
        if not g["game_over"][i]:
            for p in (0, 1):
                assert np.array_equal(win[p], g["win_cells"][i][p])
                assert np.array_equal(three[p], g["three_cells"][i][p])
        # make_move trajectory
        mv = np.ascontiguousarray(g["moves"][i][g["moves"][i] >= 0], np.int32)
        cells = np.zeros(225, np.uint8); tm = C.c_int(); mc = C.c_int()
        st2 = L.hs_make_moves(P(mv), len(mv), P(cells), C.byref(tm), C.byref(mc))
        assert st2 == want_st and np.array_equal(cells, g["cells"][i])
        assert (tm.value, mc.value) == (g["to_move"][i], g["move_count"][i])


def test_policy_vs_golden():
    L = hs_lib()
    g = golden("tiers")
    for i in range(len(g["cells"])):
        cand = np.zeros(8, np.uint32); fc = C.c_int()
        cells = np.ascontiguousarray(g["cells"][i], np.uint8)
        mv = L.hs_policy(P(cells), int(g["to_move"][i]), int(g["move_count"][i]), 1, 0, 0, 0, 0, 0,
                         P(cand), C.byref(fc))
        assert mv == g["move"][i] and fc.value == g["from_choice"][i]
        assert np.array_equal(raster_bits(cand), g["cand"][i])


def _tail(L, cells, tm, mc, ai, det, seed, game, sim, start=0):
    cells = np.ascontiguousarray(cells, np.uint8)
    fin = np.zeros(225, np.uint8); fmc = C.c_int(); fst = C.c_int(); steps = C.c_int()
    v = L.hs_rollout_tail(P(cells), int(tm), int(mc), int(ai), det, seed & 0xFFFFFFFF, seed >> 32,
                          game, sim, start, P(fin), C.byref(fmc), C.byref(fst), C.byref(steps))
    return v, fin, fmc.value, fst.value, steps.value


def test_det_rollout_vs_golden():
    L = hs_lib()
    g = golden("rollout")
    for i in range(len(g["value"])):
        v, fin, fmc, fst, steps = _tail(L, g["start_cells"][i], g["to_move"][i], g["move_count"][i],
                                        g["ai_player"][i], 1, 0, 0, 0)
        assert v == g["value"][i] and np.array_equal(fin, g["final_cells"][i])
        assert fmc == g["final_move_count"][i] and steps == g["steps"][i]


def test_philox_rollout_vs_oracle(oracle):
    """Stochastic playouts are bit-comparable because oracle and engine share the Philox keying."""
    L = hs_lib()
    g = golden("tiers")
    for i in range(0, len(g["cells"]), 3):
        b = orc.board_from_cells(g["cells"][i], g["to_move"][i], g["move_count"][i])
        for sim in range(3):
            seed = 0x1234567800000000 | (i * 7 + 1)
            rng = orc.Rng(1, seed & 0xFFFFFFFF, seed >> 32, 17 + i, sim)
            p = oracle.params(1, 1.8, 0.2, 0, int(g["to_move"][i]))
            v0, fb, _ = oracle.simulate(b, p, rng)
            v, fin, fmc, fst, steps = _tail(L, g["cells"][i], g["to_move"][i], g["move_count"][i],
                                            g["to_move"][i], 0, seed, 17 + i, sim)
            assert v == v0 and np.array_equal(fin, orc.cells_of(fb)) and fmc == fb.move_count


def test_planner_scores_vs_golden():
    L = hs_lib()
    g = golden("planner")
    for i in range(len(g["cells"])):
        s2 = np.zeros(225, np.int64)
        cells = np.ascontiguousarray(g["cells"][i], np.uint8)
        want = g["scores"][i]
        occ = np.isnan(want)
        assert np.all(want[~occ] * 2 == np.round(want[~occ] * 2))
        # per-cell form and the line-decomposed form the kernel uses (gz_lines.cuh)
        for fn in (L.hs_planner_scores2, L.hs_planner_scores2_lines):
            s2[:] = 0
            fn(P(cells), int(g["to_move"][i]), int(g["move_count"][i]), int(g["ai_player"][i]), P(s2))
            assert np.all(s2[occ] == np.iinfo(np.int64).min)
            assert np.array_equal(s2[~occ], np.round(want[~occ] * 2).astype(np.int64))
