// gz_bitboard.cuh -- 15x15 Gomoku rules on 256-bit bitboards (16-column pitch).
//
// Everything here is __host__ __device__ so that the same functions run inside the CUDA kernels
// and inside the CPU desk-check harness under tests/host_sim/ (test infrastructure; the product
// library exposes no CPU path).  Reference semantics: gomoku_board.py (rules), ai_agent.py:306-430
// (rollout policy), bg_planner.py:90-196 (knowledge search).  Citations are file:line into the
// reference tree.
#pragma once
#include <stdint.h>

#include "../../include/gomoku_b200.h"

#if defined(__CUDACC__)
#define GZ_HD __host__ __device__ __forceinline__
#else
#define GZ_HD inline
#endif

namespace gz {

// bit p = 16*row + col, row,col in [0,15); column 15 and row 15 are padding and stay zero.
struct BB {
    uint32_t w[8];
};

GZ_HD uint32_t board_word_mask(int i) { return i == 7 ? 0x00007FFFu : 0x7FFF7FFFu; }

GZ_HD int popc32(uint32_t x) {
#if defined(__CUDA_ARCH__)
    return __popc(x);
#else
    return __builtin_popcount(x);
#endif
}

GZ_HD int ctz32(uint32_t x) {
#if defined(__CUDA_ARCH__)
    return __ffs((int)x) - 1;
#else
    return __builtin_ctz(x);
#endif
}

// index of the n-th (0-based) set bit of x; x must have more than n bits set
GZ_HD int nth_bit32(uint32_t x, int n) {
#if defined(__CUDA_ARCH__)
    return (int)__fns(x, 0, n + 1);
#else
    for (int i = 0; i < n; ++i) x &= x - 1;
    return __builtin_ctz(x);
#endif
}

GZ_HD BB bb_zero() {
    BB r;
#pragma unroll
    for (int i = 0; i < 8; ++i) r.w[i] = 0u;
    return r;
}
GZ_HD BB bb_and(const BB &a, const BB &b) {
    BB r;
#pragma unroll
    for (int i = 0; i < 8; ++i) r.w[i] = a.w[i] & b.w[i];
    return r;
}
GZ_HD BB bb_or(const BB &a, const BB &b) {
    BB r;
#pragma unroll
    for (int i = 0; i < 8; ++i) r.w[i] = a.w[i] | b.w[i];
    return r;
}
GZ_HD BB bb_andnot(const BB &a, const BB &b) {  // a & ~b
    BB r;
#pragma unroll
    for (int i = 0; i < 8; ++i) r.w[i] = a.w[i] & ~b.w[i];
    return r;
}
GZ_HD bool bb_any(const BB &a) {
    uint32_t o = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i) o |= a.w[i];
    return o != 0u;
}
GZ_HD int bb_count(const BB &a) {
    int n = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i) n += popc32(a.w[i]);
    return n;
}
GZ_HD bool bb_test(const BB &a, int p) { return (a.w[p >> 5] >> (p & 31)) & 1u; }
GZ_HD void bb_set(BB &a, int p) {
    // written as a select chain so that the array stays in registers on the device
#pragma unroll
    for (int i = 0; i < 8; ++i)
        if (i == (p >> 5)) a.w[i] |= 1u << (p & 31);
}
// n-th (0-based) set bit in increasing bit order == raster order of cells
GZ_HD int bb_nth(const BB &a, int n) {
    int res = -1;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        int c = popc32(a.w[i]);
        if (res < 0) {
            if (n < c) res = 32 * i + nth_bit32(a.w[i], n);
            else n -= c;
        }
    }
    return res;
}
GZ_HD int bb_first(const BB &a) {
    int res = -1;
#pragma unroll
    for (int i = 7; i >= 0; --i)
        if (a.w[i]) res = 32 * i + ctz32(a.w[i]);
    return res;
}

// out[p] = in[p + D]  (towards lower indices), 0 < D < 32
template <int D>
GZ_HD BB bb_shr(const BB &a) {
    BB r;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        uint32_t lo = a.w[i], hi = (i < 7) ? a.w[i + 1] : 0u;
#if defined(__CUDA_ARCH__)
        r.w[i] = __funnelshift_r(lo, hi, D);
#else
        r.w[i] = (lo >> D) | (hi << (32 - D));
#endif
    }
    return r;
}
// out[p] = in[p - D]
template <int D>
GZ_HD BB bb_shl(const BB &a) {
    BB r;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        uint32_t hi = a.w[i], lo = (i > 0) ? a.w[i - 1] : 0u;
#if defined(__CUDA_ARCH__)
        r.w[i] = __funnelshift_l(lo, hi, D);
#else
        r.w[i] = (hi << D) | (lo >> (32 - D));
#endif
    }
    return r;
}

GZ_HD int cell_to_bit(int cell) { return ((cell / 15) << 4) | (cell % 15); }
GZ_HD int bit_to_cell(int p) { return (p >> 4) * 15 + (p & 15); }

GZ_HD BB bb_load(const uint32_t *src) {
    BB r;
#pragma unroll
    for (int i = 0; i < 8; ++i) r.w[i] = src[i];
    return r;
}
GZ_HD void bb_store(uint32_t *dst, const BB &a) {
#pragma unroll
    for (int i = 0; i < 8; ++i) dst[i] = a.w[i];
}

GZ_HD BB bb_empty_cells(const BB &black, const BB &white) {
    BB r;
#pragma unroll
    for (int i = 0; i < 8; ++i) r.w[i] = ~(black.w[i] | white.w[i]) & board_word_mask(i);
    return r;
}

// pitch-16 bitboard -> 225-bit raster bitset (8 words)
GZ_HD void bb_to_raster(const BB &a, uint32_t *out8) {
#pragma unroll
    for (int i = 0; i < 8; ++i) out8[i] = 0u;
    for (int r = 0; r < 15; ++r) {
        uint32_t row = (a.w[r >> 1] >> ((r & 1) * 16)) & 0x7FFFu;
        int base = r * 15;
        int wi = base >> 5, sh = base & 31;
        out8[wi] |= row << sh;
        if (sh > 17) out8[wi + 1] |= row >> (32 - sh);
    }
}

// ------------------------------------------------------------------------------------------------
// Direction masks.  For a colour bitboard b and direction step D (1 = row, 16 = column,
// 17 = main diagonal, 15 = anti-diagonal):
//   win   |= cells c such that b + {c} has >= 5 contiguous stones through c   (check_win,
//            gomoku_board.py:139-165; overlines count)
//   three |= cells c such that b + {c} has >= 3 contiguous stones through c   (threat level >= 3
//            through the new stone, ai_agent.py:383-430)
// Every term is an AND chain of pure shifts of b starting at distance 1, so a chain that would wrap
// around a board edge first meets the zero padding column/row and dies there.
template <int D>
GZ_HD void dir_masks(const BB &b, BB &win, BB &three) {
    BB s = bb_shr<D>(b);
    BB r1 = s;
    s = bb_shr<D>(s);
    BB r2 = bb_and(r1, s);
    s = bb_shr<D>(s);
    BB r3 = bb_and(r2, s);
    s = bb_shr<D>(s);
    BB r4 = bb_and(r3, s);
    s = bb_shl<D>(b);
    BB l1 = s;
    s = bb_shl<D>(s);
    BB l2 = bb_and(l1, s);
    s = bb_shl<D>(s);
    BB l3 = bb_and(l2, s);
    s = bb_shl<D>(s);
    BB l4 = bb_and(l3, s);
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        win.w[i] |= r4.w[i] | (r3.w[i] & l1.w[i]) | (r2.w[i] & l2.w[i]) | (r1.w[i] & l3.w[i]) | l4.w[i];
        three.w[i] |= r2.w[i] | (r1.w[i] & l1.w[i]) | l2.w[i];
    }
}

// win / three masks of colour bitboard b over the given empty cells
GZ_HD void win_three_masks(const BB &b, const BB &empty, BB &win, BB &three) {
    win = bb_zero();
    three = bb_zero();
    dir_masks<1>(b, win, three);
    dir_masks<16>(b, win, three);
    dir_masks<17>(b, win, three);
    dir_masks<15>(b, win, three);
    win = bb_and(win, empty);
    three = bb_and(three, empty);
}

template <int D>
GZ_HD int dir_maxrun(const BB &b) {  // longest run along D, capped at 5
    BB s = bb_shr<D>(b);
    BB a = bb_and(b, s);
    if (!bb_any(a)) return bb_any(b) ? 1 : 0;
    s = bb_shr<D>(s);
    a = bb_and(a, s);
    if (!bb_any(a)) return 2;
    s = bb_shr<D>(s);
    a = bb_and(a, s);
    if (!bb_any(a)) return 3;
    s = bb_shr<D>(s);
    a = bb_and(a, s);
    return bb_any(a) ? 5 : 4;
}

GZ_HD int maxrun(const BB &b) {
    int m = dir_maxrun<1>(b);
    int t = dir_maxrun<16>(b);
    m = t > m ? t : m;
    t = dir_maxrun<17>(b);
    m = t > m ? t : m;
    t = dir_maxrun<15>(b);
    return t > m ? t : m;
}

// maxrun(b) >= 3, the only question the rollout policy asks (ai_agent.py:378,398)
GZ_HD bool has_run3(const BB &b) {
    BB s1 = bb_shr<1>(b), s16 = bb_shr<16>(b), s17 = bb_shr<17>(b), s15 = bb_shr<15>(b);
    BB t1 = bb_shr<1>(s1), t16 = bb_shr<16>(s16), t17 = bb_shr<17>(s17), t15 = bb_shr<15>(s15);
    uint32_t o = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i)
        o |= b.w[i] & ((s1.w[i] & t1.w[i]) | (s16.w[i] & t16.w[i]) | (s17.w[i] & t17.w[i]) |
                       (s15.w[i] & t15.w[i]));
    return o != 0u;
}

// _evaluate_threat_level (ai_agent.py:403-430): 0 for a longest run <= 1, else min(run, 5)
GZ_HD int threat_level(const BB &b) {
    int m = maxrun(b);
    return m <= 1 ? 0 : m;
}

// does the stone at bit p (already set in b) sit in a run of >= 5?  (check_win)
GZ_HD bool five_through(const BB &b, int p) {
    const int ds[4] = {1, 16, 17, 15};
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        int d = ds[k], cnt = 1;
        for (int q = p + d; q < 256 && bb_test(b, q); q += d) ++cnt;
        for (int q = p - d; q >= 0 && bb_test(b, q); q -= d) ++cnt;
        if (cnt >= 5) return true;
    }
    return false;
}

// any five anywhere (board import: derives game_over / winner from the stones alone)
GZ_HD bool has_five(const BB &b) { return maxrun(b) >= 5; }

// ------------------------------------------------------------------------------------------------
// Board helpers on gz_board.
GZ_HD BB board_bb(const gz_board &b, int colour /*1|2*/) { return bb_load(b.stones[colour - 1]); }

// make_move (gomoku_board.py:84-113) for a bit position that is known to be empty and a live board
GZ_HD void board_place(gz_board &b, int p) {
    int c = b.to_move - 1;
    BB mine = bb_load(b.stones[c]);
    bb_set(mine, p);
    bb_store(b.stones[c], mine);
    b.move_count = (uint16_t)(b.move_count + 1);
    if (five_through(mine, p)) b.status = b.to_move;
    else if (b.move_count >= 200) b.status = 3;  // is_board_full() cannot fire before 200 stones
    b.to_move = (uint8_t)(3 - b.to_move);
}

// full make_move with validity checks; returns make_move's boolean
GZ_HD bool board_make_move(gz_board &b, int cell) {
    if (cell < 0 || cell >= 225 || b.status != 0) return false;
    int p = cell_to_bit(cell);
    if (bb_test(bb_load(b.stones[0]), p) || bb_test(bb_load(b.stones[1]), p)) return false;
    board_place(b, p);
    return true;
}

// ------------------------------------------------------------------------------------------------
// Pattern score (bg_planner.py:133-196).  X = own stones, E = empty cells of a 9-cell window, bit k
// = k-th cell; off-board cells and the other colour are in neither mask (the reference maps them to
// '#'/'O', which no pattern contains).  Each pattern adds its weight at most once per window.
// The "double live three" bonus (:194) needs two non-overlapping ".XXX." in nine cells and cannot
// fire; the oracle restates it literally and the goldens confirm it never does.
GZ_HD int32_t eval_window(uint32_t X, uint32_t E) {
    uint32_t x2 = X & (X >> 1), x3 = x2 & (X >> 2), x4 = x3 & (X >> 3), x5 = x4 & (X >> 4);
    int32_t v = 0;
    if (x5) v += 100000;                                                   // XXXXX
    if (E & (x4 >> 1) & (E >> 5)) v += 10000;                              // .XXXX.
    if ((x4 & (E >> 4)) | (E & (x4 >> 1))) v += 5000;                      // XXXX.  |  .XXXX
    if (E & (x3 >> 1) & (E >> 4)) v += 1000;                               // .XXX.
    if ((x2 & (E >> 2) & (X >> 3)) | (X & (E >> 1) & (x2 >> 2))) v += 300; // XX.X  |  X.XX
    if (E & (x2 >> 1) & (E >> 3)) v += 50;                                 // .XX.
    return v;
}

// bits of mask m along the line through (r,c) in direction (dr,dc), offsets -H..+H -> bits 0..2H
template <int H>
GZ_HD uint32_t line_bits(const BB &m, int r, int c, int dr, int dc) {
    uint32_t out = 0;
#pragma unroll
    for (int k = -H; k <= H; ++k) {
        int rr = r + k * dr, cc = c + k * dc;
        if (rr >= 0 && rr < 15 && cc >= 0 && cc < 15) out |= (uint32_t)bb_test(m, (rr << 4) | cc) << (k + H);
    }
    return out;
}

// sum of the four window scores of the stone at (r,c)
GZ_HD int32_t pattern_score_stone(const BB &own, const BB &empty, int r, int c) {
    int32_t v = 0;
    v += eval_window(line_bits<4>(own, r, c, 1, 0), line_bits<4>(empty, r, c, 1, 0));
    v += eval_window(line_bits<4>(own, r, c, 0, 1), line_bits<4>(empty, r, c, 0, 1));
    v += eval_window(line_bits<4>(own, r, c, 1, 1), line_bits<4>(empty, r, c, 1, 1));
    v += eval_window(line_bits<4>(own, r, c, 1, -1), line_bits<4>(empty, r, c, 1, -1));
    return v;
}

// whole-board pattern score, serial form (the kernels spread the cells over a warp instead)
GZ_HD int64_t pattern_score(const BB &own, const BB &empty) {
    int64_t tot = 0;
    for (int r = 0; r < 15; ++r)
        for (int c = 0; c < 15; ++c)
            if (bb_test(own, (r << 4) | c)) tot += pattern_score_stone(own, empty, r, c);
    return tot;
}

// Change of pattern_score(A) when a stone of colour `mover_is_a ? A : other` is put on the empty
// cell (r,c): windows centred on A-stones within four cells of (r,c) along each direction are
// re-evaluated, and the new stone's own four windows are added when it is A's.
GZ_HD int32_t pattern_delta(const BB &a_stones, const BB &empty, int r, int c, bool mover_is_a) {
    const int drs[4] = {1, 0, 1, 1}, dcs[4] = {0, 1, 1, -1};
    int32_t delta = 0;
#pragma unroll
    for (int d = 0; d < 4; ++d) {
        uint32_t X = line_bits<8>(a_stones, r, c, drs[d], dcs[d]);   // 17 cells, centre = bit 8
        uint32_t E = line_bits<8>(empty, r, c, drs[d], dcs[d]);
        uint32_t X2 = mover_is_a ? (X | (1u << 8)) : X;
        uint32_t E2 = E & ~(1u << 8);
        for (int j = 4; j <= 12; ++j) {
            if (j == 8) continue;
            if (!((X >> j) & 1u)) continue;
            int sh = j - 4;
            delta += eval_window((X2 >> sh) & 0x1FFu, (E2 >> sh) & 0x1FFu) -
                     eval_window((X >> sh) & 0x1FFu, (E >> sh) & 0x1FFu);
        }
        if (mover_is_a) delta += eval_window((X2 >> 4) & 0x1FFu, (E2 >> 4) & 0x1FFu);
    }
    return delta;
}

}  // namespace gz
