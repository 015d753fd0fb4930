// gz_policy.cuh -- rollout policy, planner candidate scoring and the counter-based RNG.
// __host__ __device__ like gz_bitboard.cuh (same desk-check arrangement).
#pragma once
#include "gz_bitboard.cuh"

namespace gz {

// ------------------------------------------------------------------------------------------------
// Philox4x32-10 (Salmon et al., "Parallel random numbers: as easy as 1, 2, 3").  Keyed by the
// caller's seed; the counter is (rollout step, simulation, global game id, stream tag), so a
// playout's random numbers do not depend on which thread, wave or GPU runs it.
struct Philox4 {
    uint32_t x[4];
};

GZ_HD uint32_t mulhi32(uint32_t a, uint32_t b) {
#if defined(__CUDA_ARCH__)
    return __umulhi(a, b);
#else
    return (uint32_t)(((uint64_t)a * (uint64_t)b) >> 32);
#endif
}

GZ_HD Philox4 philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0,
                            uint32_t k1) {
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        uint32_t h0 = mulhi32(0xD2511F53u, c0), l0 = 0xD2511F53u * c0;
        uint32_t h1 = mulhi32(0xCD9E8D57u, c2), l1 = 0xCD9E8D57u * c2;
        uint32_t n0 = h1 ^ c1 ^ k0, n2 = h0 ^ c3 ^ k1;
        c0 = n0; c1 = l1; c2 = n2; c3 = l0;
        k0 += 0x9E3779B9u;
        k1 += 0xBB67AE85u;
    }
    Philox4 o;
    o.x[0] = c0; o.x[1] = c1; o.x[2] = c2; o.x[3] = c3;
    return o;
}

constexpr uint32_t kTagRollout = 0x524F4C4Cu;  // "ROLL"

struct RolloutRng {
    uint32_t seed_lo, seed_hi, game, sim;
    bool deterministic;
};

// random.choice(seq) -> index in [0, n): x[1] of the step's block; deterministic mode -> 0
GZ_HD int rng_choice(const RolloutRng &g, int step, int n) {
    if (g.deterministic) return 0;
    Philox4 r = philox4x32_10((uint32_t)step, g.sim, g.game, kTagRollout, g.seed_lo, g.seed_hi);
    return (int)mulhi32(r.x[1], (uint32_t)n);
}
// random.random(): x[0] of the step's block scaled to [0,1); deterministic mode -> 1.0
GZ_HD double rng_random(const RolloutRng &g, int step) {
    if (g.deterministic) return 1.0;
    Philox4 r = philox4x32_10((uint32_t)step, g.sim, g.game, kTagRollout, g.seed_lo, g.seed_hi);
    return (double)r.x[0] * (1.0 / 4294967296.0);
}

// ------------------------------------------------------------------------------------------------
// Centre masks for tier 5 of the rollout policy (ai_agent.py:338-361): Manhattan distance to (7,7)
// <= 2, and <= 4.
GZ_HD BB centre_mask(int maxdist) {
    BB m = bb_zero();
    for (int r = 7 - maxdist; r <= 7 + maxdist; ++r) {
        int span = maxdist - (r < 7 ? 7 - r : r - 7);
        uint32_t row = ((1u << (2 * span + 1)) - 1u) << (7 - span);
        m.w[r >> 1] |= row << ((r & 1) * 16);
    }
    return m;
}

struct PolicyPick {
    int bit;          // chosen cell as a bit position (16*r+c)
    bool from_choice; // false: tier 1 returned the first winning cell
    BB cand;          // the list handed to random.choice (undefined when !from_choice)
};

// _select_offensive_move (ai_agent.py:306-361) for a LIVE board, mover = b.to_move.
//  1. first empty cell (raster order) that makes five for the mover;
//  2. ("block") cannot fire: the mover's stone never makes the opponent the winner;
//  3. opponent already has a run >= 3 anywhere -> every empty cell qualifies (:363-381 tests the
//     opponent's runs after the MOVER's trial stone, which cannot change them);
//  4. mover already has a run >= 3 -> every empty cell; else the cells that make a run >= 3;
//  5. centre cells (distance <= 2), else extended centre (<= 4), else the rest.
GZ_HD PolicyPick select_offensive_move(const gz_board &b, const RolloutRng &rng, int step) {
    BB black = bb_load(b.stones[0]), white = bb_load(b.stones[1]);
    BB empty = bb_empty_cells(black, white);
    const BB &mine = (b.to_move == 1) ? black : white;
    const BB &theirs = (b.to_move == 1) ? white : black;
    PolicyPick out;
    BB win, three;
    win_three_masks(mine, empty, win, three);
    if (bb_any(win)) {
        out.bit = bb_first(win);
        out.from_choice = false;
        out.cand = bb_zero();
        return out;
    }
    out.from_choice = true;
    if (has_run3(theirs) || has_run3(mine)) out.cand = empty;
    else if (bb_any(three)) out.cand = three;
    else {
        BB c2 = bb_and(empty, centre_mask(2));
        if (bb_any(c2)) out.cand = c2;
        else {
            BB c4 = bb_and(empty, centre_mask(4));
            out.cand = bb_any(c4) ? c4 : empty;
        }
    }
    int n = bb_count(out.cand);
    out.bit = bb_nth(out.cand, rng_choice(rng, step, n));
    return out;
}

// terminal value from the AI's point of view (ai_agent.py:287-304)
GZ_HD double terminal_value(const gz_board &b, int ai_player) {
    if (b.status == 0) return 0.0;
    if (b.status == 3) return 0.1;
    return b.status == ai_player ? 1.0 : -1.0;
}

// offensive-random tail of _simulate (ai_agent.py:274-285): until game over or 100 total steps
GZ_HD double rollout_tail(gz_board &b, int ai_player, const RolloutRng &rng, int &steps,
                          int &tail_steps) {
    while (b.status == 0 && steps < 100) {
        PolicyPick pk = select_offensive_move(b, rng, steps);
        board_place(b, pk.bit);
        ++steps;
        ++tail_steps;
    }
    return terminal_value(b, ai_player);
}

// ------------------------------------------------------------------------------------------------
// Planner candidate score (bg_planner.py:90-131), doubled so that the 0.5-step centre bias stays an
// integer.  `a` is the planner's fixed colour (the AI), the stone placed is the board's mover's.
//   win for A                     ->  2 * 1e6
//   other colour can win next     -> -2 * 1e5   (_opponent_can_win_next, :116-125)
//   else                          ->  2 * pattern_score(temp, A) + max(0, 6 - |r-7| - |c-7|)
// s_a = pattern_score(board, A), win_b = cells that make five for the other colour on `board`.
constexpr int64_t kScoreWin2 = 2000000;
constexpr int64_t kScoreLose2 = -200000;

GZ_HD int64_t planner_score2(const gz_board &b, const BB &black, const BB &white, const BB &empty,
                             int a, int64_t s_a, const BB &win_b, int p) {
    int mover = b.to_move;
    const BB &a_st = (a == 1) ? black : white;
    const BB &o_st = (a == 1) ? white : black;
    BB mv = (mover == 1) ? black : white;
    bb_set(mv, p);
    bool won = five_through(mv, p);
    if (won && mover == a) return kScoreWin2;
    bool opp_wins;
    if (won) {
        // temp is over with winner == other: every make_move on a copy fails and the stale winner
        // satisfies the test at :123 for the first empty cell (there always is one below 200 stones)
        opp_wins = true;
    } else if (b.move_count + 1 >= 200) {
        opp_wins = false;  // temp is a draw: make_move fails, winner is None
    } else if (mover == a) {
        // the other colour's winning cells are unchanged except that p is now taken
        BB rest = win_b;
        rest.w[p >> 5] &= ~(1u << (p & 31));
        opp_wins = bb_any(rest);
    } else {
        // the mover IS the other colour: its winning cells after the new stone
        BB e2 = empty;
        e2.w[p >> 5] &= ~(1u << (p & 31));
        BB w2, t2;
        win_three_masks(mv, e2, w2, t2);
        opp_wins = bb_any(w2);
    }
    (void)o_st;
    if (opp_wins) return kScoreLose2;
    int r = p >> 4, c = p & 15;
    int32_t d = pattern_delta(a_st, empty, r, c, mover == a);
    int dist = (r < 7 ? 7 - r : r - 7) + (c < 7 ? 7 - c : c - 7);
    int bias2 = 6 - dist;
    if (bias2 < 0) bias2 = 0;
    return 2 * (s_a + d) + bias2;
}

}  // namespace gz
