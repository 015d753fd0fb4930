// gz_kernels_planner.cu -- BGPlannerAI.get_move without the networks (bg_planner.py:232-269):
//   k_planner_score: KnowledgeSearch.top_k_moves -- score every empty cell, stable descending top-k
//                    (one warp per board, lane = bit position inside each bitboard word)
//   k_planner_pick : blend GraphNet's softmax with the DQN output over the k candidates, explore
//                    roll, optional make_move (one thread per board)
// Integer-ALU bound; per board-step 80 B in, <= 72 B out.
#include "gz_engine.h"
#include "gz_warp.cuh"
#include "gz_lines.cuh"

namespace gz {

constexpr long long kNoScore = INT64_MIN;

__global__ void __launch_bounds__(128) k_planner_score(const gz_board *__restrict__ boards,
                                                       const uint8_t *__restrict__ ai_player, int G,
                                                       int k, int32_t *__restrict__ topk,
                                                       int32_t *__restrict__ ntop,
                                                       int64_t *__restrict__ scores2) {
    pdl_trigger();   // the GraphNet kernel behind this one may set itself up while this grid drains (gz_engine.h)
    int g = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    int lane = threadIdx.x & 31;
    if (g >= G) return;
    gz_board b = boards[g];
    if (b.status != 0) {  // _simulate's planner loop stops at game_over (ai_agent.py:266)
        if (lane == 0) ntop[g] = 0;
        if (lane < kMaxTopK) topk[g * kMaxTopK + lane] = -1;
        if (scores2)
            for (int i = lane; i < kCells; i += 32) scores2[(size_t)g * kCells + i] = kNoScore;
        return;
    }
    BB black = bb_load(b.stones[0]), white = bb_load(b.stones[1]);
    BB empty = bb_empty_cells(black, white);
    const int a = ai_player[g];
    const BB o_st = (a == 1) ? white : black;
    // Line words by ballot: lane r holds row r of A, lane 16 + r row r of B (lanes 15 and 31 hold nothing), so ONE vote
    // per line returns A's stones in the low half and B's in the high half, bit = board row; a constant shift turns the
    // row into the position on the line.  73 votes per board (columns and both diagonal families; a row is the lane's own
    // word) instead of a 15-step bit gather per line, colour and lane: that loop was 16 % of the kernel's instructions.
    __shared__ uint32_t s_lab[4][kNumLines];
    __shared__ int32_t s_ls[4][kNumLines];
    const int wi = threadIdx.x >> 5;
    const int rr = lane & 15;
    const uint32_t myrow = (boards[g].stones[(lane >> 4) ^ (a == 1 ? 0 : 1)][rr >> 1] >> ((rr & 1) * 16)) & 0x7FFFu;   // row 15 is empty
    {
        const uint32_t other = __shfl_xor_sync(0xffffffffu, myrow, 16);
        if (lane < 15) s_lab[wi][lane] = myrow | (other << 16);                         // rows: lines 0-14
#pragma unroll
        for (int c = 0; c < 15; ++c) {                                                  // columns: lines 15-29
            const uint32_t w = __ballot_sync(0xffffffffu, (myrow >> c) & 1u);
            if (lane == 0) s_lab[wi][15 + c] = w;
        }
#pragma unroll
        for (int k = -14; k <= 14; ++k) {                                               // diagonals r - c = k: lines 30-58
            const int sh = rr - k;
            const uint32_t w = __ballot_sync(0xffffffffu, sh >= 0 && ((myrow >> sh) & 1u));
            if (lane == 0) s_lab[wi][44 + k] = w >> (k > 0 ? k : 0);
        }
#pragma unroll
        for (int d = 0; d <= 28; ++d) {                                                 // anti-diagonals r + c = d: lines 59-87
            const int sh = d - rr;
            const uint32_t w = __ballot_sync(0xffffffffu, sh >= 0 && ((myrow >> sh) & 1u));
            if (lane == 0) s_lab[wi][59 + d] = w >> (d > 14 ? d - 14 : 0);
        }
    }
    __syncwarp();
    int ls_sum = 0;
    for (int i = lane; i < kNumLines; i += 32) {
        int r0, c0, len;
        line_cell(i, 0, r0, c0, len);
        const uint32_t w = s_lab[wi][i], xa = w & 0xFFFFu, xb = w >> 16;
        int32_t v = line_score(xa, ((1u << len) - 1u) & ~(xa | xb));
        s_ls[wi][i] = v;
        ls_sum += v;
    }
    const long long s_a = warp_sum_ll((long long)ls_sum);   // _pattern_score(board, A), also syncs the warp
    __syncwarp();
    BB win_b, three_b;
    win_three_masks(o_st, empty, win_b, three_b);
    const int n_win_b = bb_count(win_b);
    // Only the empty cells are scored, so they are compacted first: entry r of the list is the r-th empty cell in raster
    // order (a ballot-free prefix count per bitboard word: lane = bit position), and lane l scores entries l, l + 32, ...
    // Scoring the 256 bit positions in place wasted the lanes that sit on stones or on the pitch padding (SIMT efficiency
    // = share of empty cells: 60-70 % in the middle game, 30 % late); the list keeps every lane busy until the last round.
    __shared__ uint8_t s_cell[4][256];
    __shared__ uint32_t s_win[4][8];
    if (lane < 8) s_win[wi][lane] = win_b.w[lane];
    int n_e = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const uint32_t m = empty.w[i];
        if ((m >> lane) & 1u) s_cell[wi][n_e + __popc(m & ((1u << lane) - 1u))] = (uint8_t)(32 * i + lane);
        n_e += __popc(m);
    }
    __syncwarp();
    // keys: biased 31-bit score (|score| < 2^30: at most 2 * pattern score + 6, or the +-2e6 / -2e5 sentinels), 0 = none
    uint32_t key[8];
    int pos[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const int r = 32 * i + lane;
        key[i] = 0u;
        pos[i] = 255;
        if (32 * i < n_e) {                      // warp-uniform: whole rounds past the list are skipped
            if (r < n_e) {
                const int pp = s_cell[wi][r];
                const long long sc = planner_score2_lines(s_lab[wi], s_ls[wi], s_a, n_win_b,
                                                          (s_win[wi][pp >> 5] >> (pp & 31)) & 1u, b.to_move, a,
                                                          b.move_count, pp);
                key[i] = (uint32_t)(sc + (1ll << 30));
                pos[i] = pp;
                if (scores2) scores2[(size_t)g * kCells + bit_to_cell(pp)] = sc;
            }
        }
    }
    if (scores2) {   // cells that are not empty carry no score
        for (int i = lane; i < 256; i += 32)
            if ((i >> 4) < 15 && (i & 15) < 15 && !((empty.w[i >> 5] >> (i & 31)) & 1u))
                scores2[(size_t)g * kCells + bit_to_cell(i)] = kNoScore;
    }
    // stable descending top-k: k rounds of (max score, smallest position).  A round is a per-lane maximum over its
    // entries (they are in raster order, so the strict comparison keeps the lane's smallest position on ties) and two
    // warp-wide redux.max: the biased score, then 255 - position among the lanes that hold it.
    int n = 0;
    for (int j = 0; j < k; ++j) {
        uint32_t lhi = 0u;
        int lp = 255;
#pragma unroll
        for (int i = 0; i < 8; ++i)
            if (key[i] > lhi) { lhi = key[i]; lp = pos[i]; }
        const uint32_t mhi = __reduce_max_sync(0xffffffffu, lhi);
        if (mhi == 0u) break;
        const uint32_t mlo = __reduce_max_sync(0xffffffffu, lhi == mhi ? 255u - (uint32_t)lp : 0u);
        const int bi = 255 - (int)mlo;
#pragma unroll
        for (int i = 0; i < 8; ++i)
            if (pos[i] == bi) key[i] = 0u;
        if (lane == 0) topk[g * kMaxTopK + j] = bit_to_cell(bi);
        ++n;
    }
    if (lane == 0) {
        ntop[g] = n;
        for (int j = n; j < kMaxTopK; ++j) topk[g * kMaxTopK + j] = -1;
    }
}

// Sixteen lanes per board, one candidate each: the gathers of the <= 16 candidates' policy and Q values are one
// round trip instead of a dependent chain per candidate (a thread per board spent 18 us on 4096 boards waiting
// for them).  Lane 0 of the group then rolls the explore dice and plays the move.
__global__ void __launch_bounds__(128) k_planner_pick(gz_board *__restrict__ boards,
                                                      int32_t *__restrict__ steps, int G, SearchCfg c,
                                                      int sim_arg, int step_override,
                                                      const int32_t *__restrict__ topk,
                                                      const int32_t *__restrict__ ntop,
                                                      const float *__restrict__ policy,
                                                      const float *__restrict__ q,
                                                      double *__restrict__ composed,
                                                      int32_t *__restrict__ move_out, int apply_move,
                                                      unsigned long long *__restrict__ counters) {
    static_assert(kMaxTopK == 16, "one lane per candidate");
    const int g = (blockIdx.x * blockDim.x + threadIdx.x) >> 4;
    const int sub = threadIdx.x & 15;
    const int sim = c.d_sim ? *c.d_sim : sim_arg;
    const bool in_range = g < G;
    const int gi = in_range ? g : G - 1;
    int did = 0;
    const int n = ntop[gi];
    const int idx = sub < n ? topk[gi * kMaxTopK + sub] : -1;
    double comp = -1e18;
    if (idx >= 0) {
        const double a = c.alpha, na = 1.0 - c.alpha;
        comp = __dsub_rn(__dmul_rn(a, (double)policy[(size_t)gi * kPolicyPitch + idx]),
                         __dmul_rn(na, (double)q[(size_t)gi * kCells + idx]));
    }
    if (composed && in_range) composed[g * kMaxTopK + sub] = idx >= 0 ? comp : 0.0;
    // first maximum among the group's candidates: larger value, then smaller candidate index
    double best = comp;
    int best_j = idx >= 0 ? sub : 16, best_idx = idx;
#pragma unroll
    for (int o = 8; o > 0; o >>= 1) {
        const double ob = __shfl_xor_sync(0xffffffffu, best, o, 16);
        const int oj = __shfl_xor_sync(0xffffffffu, best_j, o, 16);
        const int oi = __shfl_xor_sync(0xffffffffu, best_idx, o, 16);
        if (oj < 16 && (best_j == 16 || ob > best || (ob == best && oj < best_j))) { best = ob; best_j = oj; best_idx = oi; }
    }
    if (in_range && sub == 0) {
        int mv = -1;
        if (n > 0) {
            const int step = step_override >= 0 ? step_override : steps[g];
            RolloutRng rng = {c.seed_lo, c.seed_hi, c.game_base + (uint32_t)g, (uint32_t)sim, c.deterministic != 0};
            // random.random() is always drawn (bg_planner.py:267)
            if (rng_random(rng, step) < c.explore) mv = topk[g * kMaxTopK + rng_choice(rng, step, n)];
            else mv = best_idx >= 0 ? best_idx : topk[g * kMaxTopK];
            if (apply_move) {
                gz_board b = boards[g];
                board_place(b, cell_to_bit(mv));
                boards[g] = b;
                steps[g] = steps[g] + 1;
                did = 1;
            }
        }
        if (move_out) move_out[g] = mv;
    }
    for (int o = 16; o > 0; o >>= 1) did += __shfl_xor_sync(0xffffffffu, did, o);
    if (counters && (threadIdx.x & 31) == 0 && did) atomicAdd(&counters[3], (unsigned long long)did);
}

void launch_planner_score(const gz_board *b, const uint8_t *ai_player, const int32_t *steps, int G,
                          const SearchCfg &c, int planner_step, int32_t *topk, int32_t *ntop,
                          int64_t *scores2, cudaStream_t s) {
    (void)steps; (void)planner_step;
    int k = c.planner_k < kMaxTopK ? c.planner_k : kMaxTopK;
    k_planner_score<<<(G * 32 + 127) / 128, 128, 0, s>>>(b, ai_player, G, k, topk, ntop, scores2);
}
void launch_planner_pick(gz_board *b, const uint8_t *ai_player, int32_t *steps, int G,
                         const SearchCfg &c, int sim, int planner_step, const int32_t *topk,
                         const int32_t *ntop, const float *gn_policy, const float *dqn_q,
                         double *composed, int32_t *move_out, bool apply_move,
                         unsigned long long *counters, cudaStream_t s) {
    (void)ai_player;
    k_planner_pick<<<(G * 16 + 127) / 128, 128, 0, s>>>(b, steps, G, c, sim, planner_step, topk, ntop,
                                                   gn_policy, dqn_q, composed, move_out,
                                                   apply_move ? 1 : 0, counters);
}

}  // namespace gz
