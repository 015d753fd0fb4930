// gz_kernels_board.cu -- batched board import/export, make_move and per-board analysis.
// Reference: gomoku_board.py:84-113,139-213; ai_agent.py:383-430; bg_planner.py:133-196.
// HBM-bound byte/bit work: 225 B in -> 80 B out per board on import, 80 B in -> <= 264 B out for
// the analysis kernel; one warp per board, lanes = bit positions of one bitboard word.
#include "gz_engine.h"
#include "gz_warp.cuh"

namespace gz {

// one warp per board: lane l packs cells of bitboard word w for w = 0..7 via ballots
__global__ void __launch_bounds__(256) k_boards_from_cells(const uint8_t *__restrict__ cells,
                                                           const uint8_t *__restrict__ to_move,
                                                           const uint16_t *__restrict__ mc, int G,
                                                           gz_board *__restrict__ out) {
    int g = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    int lane = threadIdx.x & 31;
    if (g >= G) return;
    const uint8_t *c = cells + (size_t)g * kCells;
    uint32_t myb = 0, myw = 0;  // lane w (< 8) keeps word w
#pragma unroll
    for (int w = 0; w < 8; ++w) {
        int p = 32 * w + lane;           // bit position 16*r + col
        int r = p >> 4, col = p & 15;
        uint8_t v = (r < 15 && col < 15) ? c[r * 15 + col] : (uint8_t)0;
        uint32_t bb = __ballot_sync(0xffffffffu, v == 1);
        uint32_t ww = __ballot_sync(0xffffffffu, v == 2);
        if (lane == w) { myb = bb; myw = ww; }
    }
    // every lane needs the full boards for the status test: broadcast the 8 words
    BB black, white;
#pragma unroll
    for (int w = 0; w < 8; ++w) {
        black.w[w] = __shfl_sync(0xffffffffu, myb, w);
        white.w[w] = __shfl_sync(0xffffffffu, myw, w);
    }
    if (lane == 0) {
        gz_board b;
        bb_store(b.stones[0], black);
        bb_store(b.stones[1], white);
        b.move_count = mc[g];
        b.to_move = to_move[g];
        // gomoku_board.py:103-108: over iff a five exists, else draw from 200 stones on
        b.status = has_five(black) ? 1 : has_five(white) ? 2 : (mc[g] >= 200 ? 3 : 0);
        b.reserved[0] = b.reserved[1] = b.reserved[2] = 0;
        out[g] = b;
    }
}

__global__ void __launch_bounds__(256) k_boards_to_cells(const gz_board *__restrict__ b, int G,
                                                         uint8_t *__restrict__ cells) {
    int g = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    int lane = threadIdx.x & 31;
    if (g >= G) return;
    for (int i = lane; i < kCells; i += 32) {
        int p = cell_to_bit(i);
        uint32_t bw = b[g].stones[0][p >> 5], ww = b[g].stones[1][p >> 5];
        cells[(size_t)g * kCells + i] = ((bw >> (p & 31)) & 1u) ? 1 : (((ww >> (p & 31)) & 1u) ? 2 : 0);
    }
}

__global__ void __launch_bounds__(128) k_make_moves(gz_board *__restrict__ b,
                                                    const int32_t *__restrict__ moves, int G,
                                                    uint8_t *__restrict__ ok) {
    int g = blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= G) return;
    gz_board bd = b[g];
    bool r = board_make_move(bd, moves[g]);
    if (r) b[g] = bd;
    if (ok) ok[g] = r ? 1 : 0;
}

// one warp per board
__global__ void __launch_bounds__(128) k_board_ops(const gz_board *__restrict__ boards, int G,
                                                   uint32_t *__restrict__ legal,
                                                   uint32_t *__restrict__ win,
                                                   uint32_t *__restrict__ three,
                                                   int32_t *__restrict__ threat,
                                                   int64_t *__restrict__ pattern) {
    int g = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    int lane = threadIdx.x & 31;
    if (g >= G) return;
    BB st[2];
    st[0] = bb_load(boards[g].stones[0]);
    st[1] = bb_load(boards[g].stones[1]);
    BB empty = bb_empty_cells(st[0], st[1]);
    if (pattern) {
        long long p0 = warp_pattern_score(st[0], empty, lane);
        long long p1 = warp_pattern_score(st[1], empty, lane);
        if (lane == 0) { pattern[2 * g] = p0; pattern[2 * g + 1] = p1; }
    }
    if (lane >= 2) return;
    // lanes 0 and 1 take one colour each
    const BB mine = lane ? st[1] : st[0];
    if (legal && lane == 0) bb_to_raster(empty, legal + 8 * (size_t)g);
    if (win || three) {
        BB w, t;
        win_three_masks(mine, empty, w, t);
        if (has_run3(mine)) t = empty;  // threat level >= 3 already there (ai_agent.py:395-399)
        if (win) bb_to_raster(w, win + 8 * (2 * (size_t)g + lane));
        if (three) bb_to_raster(t, three + 8 * (2 * (size_t)g + lane));
    }
    if (threat) threat[2 * g + lane] = threat_level(mine);
}

static inline int warps_grid(int G, int threads) { return (G * 32 + threads - 1) / threads; }

void launch_boards_from_cells(const uint8_t *cells, const uint8_t *to_move, const uint16_t *mc, int G,
                              gz_board *out, cudaStream_t s) {
    k_boards_from_cells<<<warps_grid(G, 256), 256, 0, s>>>(cells, to_move, mc, G, out);
}
void launch_boards_to_cells(const gz_board *b, int G, uint8_t *cells, cudaStream_t s) {
    k_boards_to_cells<<<warps_grid(G, 256), 256, 0, s>>>(b, G, cells);
}
void launch_make_moves(gz_board *b, const int32_t *moves, int G, uint8_t *ok, cudaStream_t s) {
    k_make_moves<<<(G + 127) / 128, 128, 0, s>>>(b, moves, G, ok);
}
void launch_board_ops(const gz_board *b, int G, uint32_t *legal, uint32_t *win, uint32_t *three,
                      int32_t *threat, int64_t *pattern, cudaStream_t s) {
    k_board_ops<<<warps_grid(G, 128), 128, 0, s>>>(b, G, legal, win, three, threat, pattern);
}

}  // namespace gz
