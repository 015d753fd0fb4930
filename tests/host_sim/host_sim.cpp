// host_sim.cpp -- CPU desk-check harness for the __host__ __device__ logic in
// alphazero-gomoku_b200/csrc/*.cuh.  TEST INFRASTRUCTURE ONLY: it lets `pytest -m "not gpu"` compare
// the exact functions the CUDA kernels call with the oracle before any GPU time is spent.  It is
// not part of libgomoku_b200.so and nothing in the product imports it.
#include <cstring>

#include "../../alphazero-gomoku_b200/csrc/gz_lines.cuh"

using namespace gz;

static gz_board from_cells(const uint8_t *cells, int to_move, int move_count) {
    gz_board b;
    std::memset(&b, 0, sizeof b);
    BB bl = bb_zero(), wh = bb_zero();
    for (int i = 0; i < 225; ++i) {
        if (cells[i] == 1) bb_set(bl, cell_to_bit(i));
        if (cells[i] == 2) bb_set(wh, cell_to_bit(i));
    }
    bb_store(b.stones[0], bl);
    bb_store(b.stones[1], wh);
    b.to_move = (uint8_t)to_move;
    b.move_count = (uint16_t)move_count;
    b.status = has_five(bl) ? 1 : has_five(wh) ? 2 : (move_count >= 200 ? 3 : 0);
    return b;
}

static void to_cells(const gz_board &b, uint8_t *cells) {
    BB bl = bb_load(b.stones[0]), wh = bb_load(b.stones[1]);
    for (int i = 0; i < 225; ++i) {
        int p = cell_to_bit(i);
        cells[i] = bb_test(bl, p) ? 1 : bb_test(wh, p) ? 2 : 0;
    }
}

extern "C" {

int hs_board_ops(const uint8_t *cells, int to_move, int move_count, uint32_t *legal8,
                 uint32_t *win16, uint32_t *three16, int32_t *threat2, int64_t *pattern2) {
    gz_board b = from_cells(cells, to_move, move_count);
    BB bl = bb_load(b.stones[0]), wh = bb_load(b.stones[1]);
    BB e = bb_empty_cells(bl, wh);
    bb_to_raster(e, legal8);
    for (int c = 0; c < 2; ++c) {
        const BB &s = c ? wh : bl;
        BB w, t;
        win_three_masks(s, e, w, t);
        if (has_run3(s)) t = e;
        bb_to_raster(w, win16 + 8 * c);
        bb_to_raster(t, three16 + 8 * c);
        threat2[c] = threat_level(s);
        pattern2[c] = pattern_score(s, e);
    }
    return b.status;
}

int hs_make_moves(const int32_t *moves, int n, uint8_t *cells, int *to_move, int *move_count) {
    uint8_t z[225] = {0};
    gz_board b = from_cells(z, 1, 0);
    for (int i = 0; i < n; ++i)
        if (!board_make_move(b, moves[i])) return -1 - i;
    to_cells(b, cells);
    *to_move = b.to_move;
    *move_count = b.move_count;
    return b.status;
}

int hs_policy(const uint8_t *cells, int to_move, int move_count, int det, uint32_t seed_lo,
              uint32_t seed_hi, uint32_t game, uint32_t sim, int step, uint32_t *cand8,
              int *from_choice) {
    gz_board b = from_cells(cells, to_move, move_count);
    RolloutRng rng = {seed_lo, seed_hi, game, sim, det != 0};
    PolicyPick pk = select_offensive_move(b, rng, step);
    bb_to_raster(pk.cand, cand8);
    *from_choice = pk.from_choice;
    return bit_to_cell(pk.bit);
}

double hs_rollout_tail(const uint8_t *cells, int to_move, int move_count, int ai_player, int det,
                       uint32_t seed_lo, uint32_t seed_hi, uint32_t game, uint32_t sim,
                       int start_step, uint8_t *final_cells, int *final_move_count,
                       int *final_status, int *steps_out) {
    gz_board b = from_cells(cells, to_move, move_count);
    RolloutRng rng = {seed_lo, seed_hi, game, sim, det != 0};
    int steps = start_step, tail = 0;
    double v = rollout_tail(b, ai_player, rng, steps, tail);
    to_cells(b, final_cells);
    *final_move_count = b.move_count;
    *final_status = b.status;
    *steps_out = steps;
    return v;
}

void hs_planner_scores2(const uint8_t *cells, int to_move, int move_count, int ai,
                        int64_t *scores225) {
    gz_board b = from_cells(cells, to_move, move_count);
    BB bl = bb_load(b.stones[0]), wh = bb_load(b.stones[1]);
    BB e = bb_empty_cells(bl, wh);
    const BB &a_st = ai == 1 ? bl : wh;
    const BB &o_st = ai == 1 ? wh : bl;
    int64_t s_a = pattern_score(a_st, e);
    BB wb, tb;
    win_three_masks(o_st, e, wb, tb);
    for (int i = 0; i < 225; ++i) {
        int p = cell_to_bit(i);
        scores225[i] = bb_test(e, p) ? planner_score2(b, bl, wh, e, ai, s_a, wb, p) : INT64_MIN;
    }
}

// the same scores through the line-decomposed form the kernel uses (gz_lines.cuh)
void hs_planner_scores2_lines(const uint8_t *cells, int to_move, int move_count, int ai,
                              int64_t *scores225) {
    gz_board b = from_cells(cells, to_move, move_count);
    BB bl = bb_load(b.stones[0]), wh = bb_load(b.stones[1]);
    BB e = bb_empty_cells(bl, wh);
    const BB &a_st = ai == 1 ? bl : wh;
    const BB &o_st = ai == 1 ? wh : bl;
    uint32_t lab[kNumLines];
    int32_t ls[kNumLines];
    int64_t s_a = 0;
    for (int i = 0; i < kNumLines; ++i) {
        int r, c, len;
        line_cell(i, 0, r, c, len);
        const uint32_t xa = line_word(a_st.w, i), xb = line_word(o_st.w, i);
        lab[i] = xa | (xb << 16);
        ls[i] = line_score(xa, ((1u << len) - 1u) & ~(xa | xb));
        s_a += ls[i];
    }
    BB wb, tb;
    win_three_masks(o_st, e, wb, tb);
    int nwb = bb_count(wb);
    for (int i = 0; i < 225; ++i) {
        int p = cell_to_bit(i);
        scores225[i] = bb_test(e, p) ? planner_score2_lines(lab, ls, s_a, nwb, bb_test(wb, p), b.to_move,
                                                            ai, b.move_count, p)
                                     : INT64_MIN;
    }
}

void hs_philox(const uint32_t *ctr, const uint32_t *key, uint32_t *out) {
    Philox4 r = philox4x32_10(ctr[0], ctr[1], ctr[2], ctr[3], key[0], key[1]);
    for (int i = 0; i < 4; ++i) out[i] = r.x[i];
}
}
