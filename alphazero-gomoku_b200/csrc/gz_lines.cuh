// gz_lines.cuh -- line-decomposed form of the knowledge search (bg_planner.py:90-196).
//
// KnowledgeSearch._pattern_score sums, over every stone of a colour and the four directions, the
// score of the 9-cell window centred on the stone (bg_planner.py:133-196).  A window never leaves
// the board line it lies on (off-board cells are '#', which no pattern contains), so the score is
// a sum over the 88 lines of the board (15 rows, 15 columns, 29 + 29 diagonals) of a function of
// that line's stones alone, and a stone put on cell p changes only the four lines through p.
// A line is at most 15 cells = one 16-bit word per colour; the six pattern tests of every window
// of a line are evaluated together with shifts and ANDs on that word ("which windows contain a
// match" = the match-start mask spread over the window positions), instead of one window at a time.
//
// __host__ __device__ like gz_bitboard.cuh: the same functions run in k_planner_score and in the
// CPU desk check (tests/host_sim).
#pragma once
#include "gz_policy.cuh"

namespace gz {

constexpr int kNumLines = 88;   // 15 rows + 15 columns + 29 diagonals + 29 anti-diagonals

// line through cell (r,c) in direction dir (0 row, 1 column, 2 diagonal (1,1), 3 anti-diagonal
// (1,-1)): index into the 88-entry line arrays, position of the cell on the line, line length
GZ_HD void line_of(int dir, int r, int c, int &idx, int &pos, int &len) {
    if (dir == 0) { idx = r; pos = c; len = 15; }
    else if (dir == 1) { idx = 15 + c; pos = r; len = 15; }
    else if (dir == 2) {
        int k = r - c;
        idx = 30 + k + 14;
        pos = r < c ? r : c;
        len = 15 - (k < 0 ? -k : k);
    } else {
        int a = r + c;
        idx = 59 + a;
        pos = a <= 14 ? r : 14 - c;
        len = a <= 14 ? a + 1 : 29 - a;
    }
}

// cell (r,c) of position `pos` on line `idx` (inverse of line_of), used to build the line words
GZ_HD void line_cell(int idx, int pos, int &r, int &c, int &len) {
    if (idx < 15) { r = idx; c = pos; len = 15; }
    else if (idx < 30) { r = pos; c = idx - 15; len = 15; }
    else if (idx < 59) {
        int k = idx - 30 - 14;
        if (k >= 0) { r = k + pos; c = pos; len = 15 - k; }
        else { r = pos; c = pos - k; len = 15 + k; }
    } else {
        int a = idx - 59;
        if (a <= 14) { r = pos; c = a - pos; len = a + 1; }
        else { r = a - 14 + pos; c = 14 - pos; len = 29 - a; }
    }
}

// stones of bitboard `m` on line idx as a word (bit = position on the line)
GZ_HD uint32_t line_word(const uint32_t *m8, int idx) {
    int r, c, len;
    line_cell(idx, 0, r, c, len);
    if (idx < 15) return (m8[r >> 1] >> ((r & 1) * 16)) & 0x7FFFu;   // a row is half a word
    // consecutive cells of a line are a constant bit distance apart: 16 (column), 17 (diagonal), 15 (anti-diagonal)
    const int step = idx < 30 ? 16 : idx < 59 ? 17 : 15;
    int p = (r << 4) | c;
    uint32_t out = 0;
    for (int pos = 0; pos < len; ++pos, p += step) out |= ((m8[p >> 5] >> (p & 31)) & 1u) << pos;
    return out;
}

// Sum of the window scores (eval_window, gz_bitboard.cuh) of every stone of X on one line.
// X = own stones, E = empty cells of the line (bit = position); cells of the other colour and the
// ends of the line are in neither.  m_* = positions where a pattern STARTS (its lowest cell); a
// pattern of length L lies inside the window of the stone at s iff s-4 <= start <= s+5-L.
GZ_HD int32_t line_score(uint32_t X, uint32_t E) {
    uint32_t x2 = X & (X >> 1), x3 = x2 & (X >> 2), x4 = x3 & (X >> 3), x5 = x4 & (X >> 4);
    uint32_t m_five = x5;                                                          // XXXXX   L5
    uint32_t m_open4 = E & (x4 >> 1) & (E >> 5);                                   // .XXXX.  L6
    uint32_t m_four = (x4 & (E >> 4)) | (E & (x4 >> 1));                           // XXXX. | .XXXX  L5
    uint32_t m_open3 = E & (x3 >> 1) & (E >> 4);                                   // .XXX.   L5
    uint32_t m_split = (x2 & (E >> 2) & (X >> 3)) | (X & (E >> 1) & (x2 >> 2));    // XX.X | X.XX  L4
    uint32_t m_open2 = E & (x2 >> 1) & (E >> 3);                                   // .XX.    L4
    // spread a start mask over the windows that contain the pattern: shifts 4 .. L-5
    #define GZ_SPREAD5(m, out) { uint32_t t_ = (m) | ((m) << 1); t_ |= t_ << 2; out = t_ | ((m) << 4); }
    uint32_t w;
    int32_t v = 0;
    GZ_SPREAD5(m_five, w);  v += 100000 * popc32(w & X);
    { uint32_t t = (m_open4 << 1) | (m_open4 << 2); w = t | (t << 2); }
    v += 10000 * popc32(w & X);
    GZ_SPREAD5(m_four, w);  v += 5000 * popc32(w & X);
    GZ_SPREAD5(m_open3, w); v += 1000 * popc32(w & X);
    GZ_SPREAD5(m_split, w); w |= m_split >> 1; v += 300 * popc32(w & X);
    GZ_SPREAD5(m_open2, w); w |= m_open2 >> 1; v += 50 * popc32(w & X);
    #undef GZ_SPREAD5
    return v;
}

GZ_HD bool line_has_five(uint32_t X) {
    uint32_t x2 = X & (X >> 1), x4 = x2 & (x2 >> 2);
    return (x4 & (X >> 4)) != 0u;
}

// empty cells of the line where a stone of X completes five or more
GZ_HD uint32_t line_win_cells(uint32_t X, uint32_t E) {
    uint32_t r1 = X >> 1, r2 = r1 & (X >> 2), r3 = r2 & (X >> 3), r4 = r3 & (X >> 4);
    uint32_t l1 = X << 1, l2 = l1 & (X << 2), l3 = l2 & (X << 3), l4 = l3 & (X << 4);
    return (r4 | (r3 & l1) | (r2 & l2) | (r1 & l3) | l4) & E;
}

// Planner candidate score (planner_score2 in gz_policy.cuh, same contract) from the line words.
//   lab         : line words of the planner's colour A (low half) and of the other colour B (high half)   [88]
//   ls_old      : line_score(la[i], empties of line i)                            [88]
//   s_a         : sum of ls_old = pattern_score(board, A)
//   n_win_b     : number of empty cells that complete five for B on the board
//   p_wins_b    : p is one of them
GZ_HD int64_t planner_score2_lines(const uint32_t *lab, const int32_t *ls_old,
                                   int64_t s_a, int n_win_b, bool p_wins_b, int mover, int a,
                                   int move_count, int p) {
    const int r = p >> 4, c = p & 15;
    const bool mover_is_a = mover == a;
    bool won = false, new_b_win = false;
    int32_t delta = 0;
    // Two directions per 32-bit word (a line is at most 15 cells: bit 15 of each half is neither a stone nor empty).
    // Every pattern test of line_score / line_has_five / line_win_cells is an AND over a contiguous run of cells, so no
    // match can straddle the separator bit, and what a shift carries across it is cleared by the other factors; the
    // popcounts then add the two lines' scores, which is all the caller needs (the sum over the four directions).
#pragma unroll
    for (int dir = 0; dir < 4; dir += 2) {
        int i0, p0, l0, i1, p1, l1;
        line_of(dir, r, c, i0, p0, l0);
        line_of(dir + 1, r, c, i1, p1, l1);
        const uint32_t w0 = lab[i0], w1 = lab[i1];
#ifdef __CUDA_ARCH__
        uint32_t xa = __byte_perm(w0, w1, 0x5410), xb = __byte_perm(w0, w1, 0x7632);
#else
        uint32_t xa = (w0 & 0xFFFFu) | (w1 << 16), xb = (w0 >> 16) | (w1 & 0xFFFF0000u);
#endif
        const uint32_t bit = (1u << p0) | (0x10000u << p1);
        const uint32_t e2 = (((1u << l0) - 1u) | (((1u << l1) - 1u) << 16)) & ~(xa | xb | bit);
        if (mover_is_a) {
            xa |= bit;
            won = won || line_has_five(xa);
        } else {
            xb |= bit;
            won = won || line_has_five(xb);
            new_b_win = new_b_win || line_win_cells(xb, e2) != 0u;
        }
        delta += line_score(xa, e2) - ls_old[i0] - ls_old[i1];
    }
    if (won && mover_is_a) return kScoreWin2;
    bool opp_wins;
    if (won) opp_wins = true;                        // temp is over, stale winner == B (:123)
    else if (move_count + 1 >= 200) opp_wins = false; // temp is a draw
    else opp_wins = (n_win_b - (p_wins_b ? 1 : 0)) > 0 || new_b_win;
    if (opp_wins) return kScoreLose2;
    int dist = (r < 7 ? 7 - r : r - 7) + (c < 7 ? 7 - c : c - 7);
    int bias2 = 6 - dist;
    if (bias2 < 0) bias2 = 0;
    return 2 * (s_a + delta) + bias2;
}

}  // namespace gz
