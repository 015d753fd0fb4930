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

// Continues every live playout from steps[g] until the game is over or 100 steps were taken, then
// writes the terminal value from the AI's point of view.  Boards are updated in place.
__global__ void __launch_bounds__(128) k_rollout_tail(gz_board *__restrict__ boards,
                                                      const uint8_t *__restrict__ ai_player,
                                                      int32_t *__restrict__ steps, int G, SearchCfg c,
                                                      int sim, double *__restrict__ value,
                                                      unsigned long long *__restrict__ counters) {
    int g = blockIdx.x * blockDim.x + threadIdx.x;
    int tail = 0;
    if (g < G) {
        gz_board b = boards[g];
        int st = steps[g];
        RolloutRng rng = {c.seed_lo, c.seed_hi, c.game_base + (uint32_t)g, (uint32_t)sim, c.deterministic != 0};
        double v = rollout_tail(b, ai_player[g], rng, st, tail);
        boards[g] = b;
        steps[g] = st;
        value[g] = v;
    }
    // one atomic per warp for the step counter
    for (int o = 16; o > 0; o >>= 1) tail += __shfl_xor_sync(0xffffffffu, tail, o);
    if (counters && (threadIdx.x & 31) == 0 && tail) atomicAdd(&counters[4], (unsigned long long)tail);
}

void launch_rollout_policy(const gz_board *b, int G, const SearchCfg &c, int sim, int step,
                           int32_t *move, uint32_t *cand, cudaStream_t s) {
    k_rollout_policy<<<(G + 127) / 128, 128, 0, s>>>(b, G, c, sim, step, move, cand);
}
void launch_rollout_tail(gz_board *b, const uint8_t *ai_player, int32_t *steps, int G,
                         const SearchCfg &c, int sim, double *value, unsigned long long *counters,
                         cudaStream_t s) {
    k_rollout_tail<<<(G + 127) / 128, 128, 0, s>>>(b, ai_player, steps, G, c, sim, value, counters);
}

}  // namespace gz
