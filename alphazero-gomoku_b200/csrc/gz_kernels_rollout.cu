// gz_kernels_rollout.cu -- rollout policy and the offensive-random tail of _simulate
// (ai_agent.py:251-430).  One thread per playout: the board lives in 16 registers, every policy
// step is ~1k integer instructions (shift/AND chains), so the kernel is integer-ALU bound; its HBM
// traffic is one 80-byte board in and 8+4 bytes out per playout.
#include "gz_engine.h"
#include "gz_policy.cuh"

namespace gz {

__global__ void __launch_bounds__(128) k_rollout_policy(const gz_board *__restrict__ boards, int G,
                                                        SearchCfg c, int sim, int step,
                                                        int32_t *__restrict__ move,
                                                        uint32_t *__restrict__ cand) {
    int g = blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= G) return;
    gz_board b = boards[g];
    if (b.status != 0) {  // _simulate never asks the policy on a finished board
        move[g] = -1;
        if (cand) for (int i = 0; i < 8; ++i) cand[8 * (size_t)g + i] = 0u;
        return;
    }
    RolloutRng rng = {c.seed_lo, c.seed_hi, c.game_base + (uint32_t)g, (uint32_t)sim, c.deterministic != 0};
    PolicyPick pk = select_offensive_move(b, rng, step);
    move[g] = bit_to_cell(pk.bit);
    if (cand) bb_to_raster(pk.cand, cand + 8 * (size_t)g);
}

// ------------------------------------------------------------------------------------------------
// Tail of _simulate (ai_agent.py:274-285), eight lanes per playout: lane j of a group holds word j of both
// bitboards, a shift by D takes the neighbouring word with one shuffle, and any / count / n-th-bit become
// ballots and an 8-lane prefix sum.  One thread per playout (rollout_tail in gz_policy.cuh, still what
// k_rollout_policy and the CPU desk check run) left 4096 playouts on 128 warps of 32 SMs, each warp as slow
// as its longest game; this form spreads them over 1024 warps and only four games share a warp.
// Same decisions bit for bit: the masks are the ones of dir_masks / has_run3 evaluated word-wise, the
// random numbers are rng_choice's, and "the stone makes five" is membership in the win mask computed before
// the move (five_through counts the same runs through the new stone).
namespace {
struct GroupCtx {
    uint32_t gmask;   // the 8 lanes of this group
    int j;            // word index = lane within the group
};
template <int D> __device__ __forceinline__ uint32_t g_shr(uint32_t w, const GroupCtx &g) {   // out[p] = in[p + D]
    uint32_t hi = __shfl_down_sync(0xffffffffu, w, 1, 8);
    if (g.j == 7) hi = 0u;
    return __funnelshift_r(w, hi, D);
}
template <int D> __device__ __forceinline__ uint32_t g_shl(uint32_t w, const GroupCtx &g) {   // out[p] = in[p - D]
    uint32_t lo = __shfl_up_sync(0xffffffffu, w, 1, 8);
    if (g.j == 0) lo = 0u;
    return __funnelshift_l(lo, w, D);
}
template <int D> __device__ __forceinline__ void g_dir_masks(uint32_t b, const GroupCtx &g, uint32_t &win, uint32_t &three) {
    uint32_t s = g_shr<D>(b, g);
    const uint32_t r1 = s;
    s = g_shr<D>(s, g); const uint32_t r2 = r1 & s;
    s = g_shr<D>(s, g); const uint32_t r3 = r2 & s;
    s = g_shr<D>(s, g); const uint32_t r4 = r3 & s;
    s = g_shl<D>(b, g);
    const uint32_t l1 = s;
    s = g_shl<D>(s, g); const uint32_t l2 = l1 & s;
    s = g_shl<D>(s, g); const uint32_t l3 = l2 & s;
    s = g_shl<D>(s, g); const uint32_t l4 = l3 & s;
    win |= r4 | (r3 & l1) | (r2 & l2) | (r1 & l3) | l4;
    three |= r2 | (r1 & l1) | l2;
}
template <int D> __device__ __forceinline__ uint32_t g_run3(uint32_t b, const GroupCtx &g) {
    const uint32_t s = g_shr<D>(b, g);
    return b & s & g_shr<D>(s, g);
}
__device__ __forceinline__ bool g_any(uint32_t w, const GroupCtx &g) {
    return (__ballot_sync(0xffffffffu, w != 0u) & g.gmask) != 0u;
}
// bit position (32*word + bit) of the n-th set bit of the group's 256-bit value; n < its population count
__device__ __forceinline__ int g_nth(uint32_t w, int n, const GroupCtx &g) {
    const int c = __popc(w);
    int incl = c;
#pragma unroll
    for (int o = 1; o < 8; o <<= 1) {
        const int t = __shfl_up_sync(0xffffffffu, incl, o, 8);
        if (g.j >= o) incl += t;
    }
    const int excl = incl - c;
    const bool own = n >= excl && n < incl;
    const int mine = own ? 32 * g.j + (int)__fns(w, 0, n - excl + 1) : 0;
    const uint32_t who = __ballot_sync(0xffffffffu, own) & g.gmask;
    return __shfl_sync(0xffffffffu, mine, who ? __ffs((int)who) - 1 : 0);
}
__device__ __forceinline__ int g_count(uint32_t w) {
    int c = __popc(w);
#pragma unroll
    for (int o = 1; o < 8; o <<= 1) c += __shfl_xor_sync(0xffffffffu, c, o, 8);
    return c;
}
}  // namespace

__global__ void __launch_bounds__(128) k_rollout_tail(gz_board *__restrict__ boards,
                                                      const uint8_t *__restrict__ ai_player,
                                                      int32_t *__restrict__ steps, int G, SearchCfg c,
                                                      int sim_arg, double *__restrict__ value,
                                                      unsigned long long *__restrict__ counters) {
    const int lane = threadIdx.x & 31;
    const int sim = c.d_sim ? *c.d_sim : sim_arg;
    GroupCtx g;
    g.j = lane & 7;
    g.gmask = 0xFFu << (lane & 24);
    const int game = (blockIdx.x * blockDim.x + threadIdx.x) >> 3;
    const bool in_range = game < G;
    const int gi = in_range ? game : G - 1;
    uint32_t black = boards[gi].stones[0][g.j], white = boards[gi].stones[1][g.j];
    int move_count = boards[gi].move_count, to_move = boards[gi].to_move, status = boards[gi].status;
    int st = steps[gi], tail = 0;
    const uint32_t wmask = board_word_mask(g.j);
    uint32_t centre2 = 0u, centre4 = 0u;
    {
        const BB c2 = centre_mask(2), c4 = centre_mask(4);
#pragma unroll
        for (int i = 0; i < 8; ++i) if (i == g.j) { centre2 = c2.w[i]; centre4 = c4.w[i]; }
    }
    const RolloutRng rng = {c.seed_lo, c.seed_hi, c.game_base + (uint32_t)gi, (uint32_t)sim, c.deterministic != 0};
    bool live = in_range && status == 0 && st < 100;
    while (__any_sync(0xffffffffu, live)) {
        const uint32_t mine = to_move == 1 ? black : white, theirs = to_move == 1 ? white : black;
        const uint32_t empty = ~(black | white) & wmask;
        uint32_t win = 0u, three = 0u;
        g_dir_masks<1>(mine, g, win, three);
        g_dir_masks<16>(mine, g, win, three);
        g_dir_masks<17>(mine, g, win, three);
        g_dir_masks<15>(mine, g, win, three);
        win &= empty;
        three &= empty;
        const uint32_t run3 = g_run3<1>(mine, g) | g_run3<16>(mine, g) | g_run3<17>(mine, g) | g_run3<15>(mine, g) |
                              g_run3<1>(theirs, g) | g_run3<16>(theirs, g) | g_run3<17>(theirs, g) | g_run3<15>(theirs, g);
        // every ballot is taken by the whole warp (groups diverge below, a full-mask ballot must not)
        const bool wins = g_any(win, g), any_run3 = g_any(run3, g), any_three = g_any(three, g);
        const bool any_c2 = g_any(empty & centre2, g), any_c4 = g_any(empty & centre4, g);
        const uint32_t cand = wins ? win                    // tier 1: the first winning cell
                            : any_run3 ? empty              // tiers 3 / 4a
                            : any_three ? three             // tier 4b
                            : any_c2 ? (empty & centre2) : any_c4 ? (empty & centre4) : empty;
        const int n = g_count(cand);
        const int pick = wins || n == 0 ? 0 : rng_choice(rng, st, n);
        const int bit = g_nth(cand, n == 0 ? 0 : pick, g);
        if (live) {
            const uint32_t add = (bit >> 5) == g.j ? 1u << (bit & 31) : 0u;
            if (to_move == 1) black |= add; else white |= add;
            ++move_count;
            if (wins) status = to_move;
            else if (move_count >= 200) status = 3;
            to_move = 3 - to_move;
            ++st;
            if (g.j == 0) ++tail;
            live = status == 0 && st < 100;
        }
    }
    if (in_range) {
        boards[game].stones[0][g.j] = black;
        boards[game].stones[1][g.j] = white;
        if (g.j == 0) {
            boards[game].move_count = (uint16_t)move_count;
            boards[game].to_move = (uint8_t)to_move;
            boards[game].status = (uint8_t)status;
            steps[game] = st;
            const int ai = ai_player[game];
            value[game] = status == 0 ? 0.0 : status == 3 ? 0.1 : status == ai ? 1.0 : -1.0;   // terminal_value
        }
    }
    // one atomic per warp for the step counter
    for (int o = 16; o > 0; o >>= 1) tail += __shfl_xor_sync(0xffffffffu, tail, o);
    if (counters && lane == 0 && tail) atomicAdd(&counters[4], (unsigned long long)tail);
}

void launch_rollout_policy(const gz_board *b, int G, const SearchCfg &c, int sim, int step,
                           int32_t *move, uint32_t *cand, cudaStream_t s) {
    k_rollout_policy<<<(G + 127) / 128, 128, 0, s>>>(b, G, c, sim, step, move, cand);
}
void launch_rollout_tail(gz_board *b, const uint8_t *ai_player, int32_t *steps, int G,
                         const SearchCfg &c, int sim, double *value, unsigned long long *counters,
                         cudaStream_t s) {
    k_rollout_tail<<<(G * 8 + 127) / 128, 128, 0, s>>>(b, ai_player, steps, G, c, sim, value, counters);
}

}  // namespace gz
