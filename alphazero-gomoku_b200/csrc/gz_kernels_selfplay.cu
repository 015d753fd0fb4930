// gz_kernels_selfplay.cu -- the self-play game loop around gz_search (training.py:141-218
// play_one_game, :77-97 SimpleReplay; ai_agent.py:109-166 get_move / _opening_move), device resident:
// boards, per-game ply staging and the replay buffer never leave HBM between plies.
//
// One warp per game.  Lane 0 plays (book plies, epsilon override, make_move, bookkeeping); when a game
// ends the whole warp copies its plies, labelled with the outcome, into the replay buffer.
#include "gz_engine.h"
#include "gz_policy.cuh"

namespace gz {

constexpr uint32_t kTagSelfplay = 0x53454C46u;   // "SELF"

// uniform pick among the set bits of `cand` (random.choice over a raster-ordered list)
__device__ __forceinline__ int sp_choice(const BB &cand, const SelfPlayCfg &c, uint32_t game_id, int ply,
                                         uint32_t which) {
    int n = bb_count(cand);
    if (c.deterministic) return bb_first(cand);
    Philox4 r = philox4x32_10((uint32_t)ply, which, game_id, kTagSelfplay, c.seed_lo, c.seed_hi);
    return bb_nth(cand, (int)mulhi32(r.x[1], (uint32_t)n));
}
__device__ __forceinline__ double sp_random(const SelfPlayCfg &c, uint32_t game_id, int ply) {
    if (c.deterministic) return 1.0;
    Philox4 r = philox4x32_10((uint32_t)ply, 7u, game_id, kTagSelfplay, c.seed_lo, c.seed_hi);
    return (double)r.x[0] * (1.0 / 4294967296.0);
}

__device__ __forceinline__ BB square_mask(int half) {   // cells with |r-7| <= half and |c-7| <= half
    BB m = bb_zero();
    for (int r = 7 - half; r <= 7 + half; ++r) {
        uint32_t row = ((1u << (2 * half + 1)) - 1u) << (7 - half);
        m.w[r >> 1] |= row << ((r & 1) * 16);
    }
    return m;
}

__device__ __forceinline__ void sp_record(gz_replay_record *stage, const gz_board &b, int bit, uint32_t game_id) {
    gz_replay_record &r = stage[b.move_count];
#pragma unroll
    for (int i = 0; i < 8; ++i) { r.stones[0][i] = b.stones[0][i]; r.stones[1][i] = b.stones[1][i]; }
    r.move = (uint8_t)bit_to_cell(bit);
    r.mover = b.to_move;
    r.outcome = 0;
    r.ply = (uint8_t)b.move_count;
    r.game_id = game_id;
}

// one opening-book ply (ai_agent.py:138-166)
__device__ __forceinline__ int sp_book_bit(const BB &empty, int ply, const SelfPlayCfg &c, uint32_t game_id) {
    if (ply == 0) return (7 << 4) | 7;
    BB c3 = bb_and(empty, square_mask(1));
    if (bb_any(c3)) return sp_choice(c3, c, game_id, ply, 0u);
    BB c5 = bb_and(empty, square_mask(2));
    return sp_choice(bb_any(c5) ? c5 : empty, c, game_id, ply, 0u);
}

// side A's colour in the arena: black in even game ordinals (training.py:232,237,266)
__device__ __forceinline__ int sp_side_a(const SelfPlayCfg &c, uint32_t game_id) { return (game_id & 1u) ? 2 : 1; }
// Which games are counted and recorded is fixed by their START ordinal, never by when they finish.  Arena: the
// ordinal is the game id.  Self-play: id = id_base + slot + generation * 2^20, ordinal = generation * G + slot.
__device__ __forceinline__ bool sp_counted(const SelfPlayCfg &c, uint32_t game_id) {
    if (c.games_total <= 0) return true;
    if (c.arena != 0) return (long long)game_id < c.games_total;
    const uint32_t rel = game_id - c.id_base;
    return (long long)(rel >> 20) * c.n_slots + (long long)(rel & 0xFFFFFu) < c.games_total;
}

// lane 0: empty board, then every ply that needs no search: the six opening-book plies
// (ai_agent.py:121-122,138-166); against the RandomAgent (training.py:225-228) its plies are uniform random from
// the first stone on and the loop runs until side A is to move with six stones on the board.
__device__ int sp_start_game(gz_board &b, gz_replay_record *stage, const SelfPlayCfg &c, uint32_t game_id) {
#pragma unroll
    for (int i = 0; i < 8; ++i) b.stones[0][i] = b.stones[1][i] = 0u;
    b.move_count = 0; b.to_move = 1; b.status = 0;
    b.reserved[0] = b.reserved[1] = b.reserved[2] = 0u;
    const int side_a = sp_side_a(c, game_id);
    int plies = 0;
    while (b.status == 0) {
        const int ply = b.move_count;
        const bool random_ply = c.arena == GZ_ARENA_RANDOM && b.to_move != side_a;
        if (!random_ply && ply >= 6) break;
        BB empty = bb_empty_cells(bb_load(b.stones[0]), bb_load(b.stones[1]));
        int bit = random_ply ? sp_choice(empty, c, game_id, ply, 3u) : sp_book_bit(empty, ply, c, game_id);
        sp_record(stage, b, bit, game_id);
        board_place(b, bit);
        ++plies;
    }
    return plies;
}

// arena: the side to move decides how many simulations this position gets (gz_board.reserved[0])
__device__ __forceinline__ void sp_set_budget(gz_board &b, const SelfPlayCfg &c, uint32_t game_id) {
    int n = c.arena == 0 ? 0 : (b.to_move == sp_side_a(c, game_id) ? c.sims_a : c.sims_b);
    b.reserved[0] = (uint32_t)n;
}

__global__ void __launch_bounds__(128) k_sp_begin(SelfPlayArrays sp, int G, SelfPlayCfg c) {
    int g = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (g >= G || (threadIdx.x & 31) != 0) return;
    gz_board b;
    sp.generation[g] = 0;
    sp.game_id[g] = c.id_base + (uint32_t)g;
    int n = sp_start_game(b, sp.staging + (size_t)g * kMaxPlies, c, sp.game_id[g]);
    sp_set_budget(b, c, sp.game_id[g]);
    sp.boards[g] = b;
    sp.ai[g] = b.to_move;
    if (sp_counted(c, sp.game_id[g])) atomicAdd(&sp.ctr[SP_PLIES], (unsigned long long)n);
}

__global__ void __launch_bounds__(128) k_sp_apply(SelfPlayArrays sp, int G, SelfPlayCfg c) {
    const int g = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (g >= G) return;
    gz_replay_record *stage = sp.staging + (size_t)g * kMaxPlies;
    int finished = 0, n_plies = 0, winner = 0, counted = 0;
    gz_board b;
    if (lane == 0) {
        b = sp.boards[g];
        const uint32_t gid = sp.game_id[g];
        const int ply = b.move_count;
        BB empty = bb_empty_cells(bb_load(b.stones[0]), bb_load(b.stones[1]));
        int mv = sp.best[g];
        int bit = mv >= 0 ? cell_to_bit(mv) : sp_choice(empty, c, gid, ply, 1u);   // no children (:203-204)
        if (sp_random(c, gid, ply) < c.exploration) bit = sp_choice(empty, c, gid, ply, 2u);   // (:128-129)
        sp_record(stage, b, bit, gid);
        board_place(b, bit);
        int played = 1;
        if (c.arena == GZ_ARENA_RANDOM && b.status == 0) {   // the RandomAgent's reply (training.py:225-228)
            BB e2 = bb_empty_cells(bb_load(b.stones[0]), bb_load(b.stones[1]));
            int rb = sp_choice(e2, c, gid, b.move_count, 3u);
            sp_record(stage, b, rb, gid);
            board_place(b, rb);
            ++played;
        }
        counted = sp_counted(c, gid) ? 1 : 0;
        if (counted) atomicAdd(&sp.ctr[SP_PLIES], (unsigned long long)played);
        if (b.status != 0) {
            finished = 1;
            n_plies = b.move_count;
            winner = b.status;
            if (counted) {
                atomicAdd(&sp.ctr[SP_GAMES], 1ull);
                atomicAdd(&sp.ctr[winner == 1 ? SP_BLACK : winner == 2 ? SP_WHITE : SP_DRAWS], 1ull);
                if (c.arena != 0 && winner != 3)
                    atomicAdd(&sp.ctr[winner == sp_side_a(c, gid) ? SP_A_WINS : SP_A_LOSSES], 1ull);
            }
        }
    }
    finished = __shfl_sync(0xffffffffu, finished, 0);
    if (finished) {
        n_plies = __shfl_sync(0xffffffffu, n_plies, 0);
        winner = __shfl_sync(0xffffffffu, winner, 0);
        long long base = -1;
        if (lane == 0 && counted) {
            unsigned long long old = atomicAdd(&sp.ctr[SP_CURSOR], (unsigned long long)n_plies);
            if ((long long)(old + n_plies) > sp.out_cap) {
                atomicAdd(&sp.ctr[SP_CURSOR], (unsigned long long)(-(long long)n_plies));
                atomicAdd(&sp.ctr[SP_DROPPED], (unsigned long long)n_plies);
            } else base = (long long)old;
        }
        base = __shfl_sync(0xffffffffu, base, 0);
        __syncwarp();
        if (base >= 0) {   // 72-byte records = 18 words; word 16 holds move | mover | outcome | ply
            const uint32_t *src = reinterpret_cast<const uint32_t *>(stage);
            uint32_t *dst = reinterpret_cast<uint32_t *>(sp.out + base);
            for (int i = lane; i < 18 * n_plies; i += 32) {
                uint32_t w = src[i];
                if (i % 18 == 16) {
                    int mover = (w >> 8) & 0xFF;
                    int outcome = winner == 3 ? 0 : (mover == winner ? 1 : -1);   // training.py:89-94
                    w = (w & 0xFF00FFFFu) | ((uint32_t)(uint8_t)(int8_t)outcome << 16);
                }
                dst[i] = w;
            }
        }
        __syncwarp();
        if (lane == 0) {
            uint32_t gen = sp.generation[g] + 1;
            sp.generation[g] = gen;
            uint32_t gid = c.id_base + (uint32_t)g + gen * (uint32_t)c.id_stride;
            sp.game_id[g] = gid;
            int n = sp_start_game(b, stage, c, gid);
            if (sp_counted(c, gid)) atomicAdd(&sp.ctr[SP_PLIES], (unsigned long long)n);
        }
    }
    if (lane == 0) {
        sp_set_budget(b, c, sp.game_id[g]);
        sp.boards[g] = b;
        sp.ai[g] = b.to_move;
    }
}

void launch_sp_begin(const SelfPlayArrays &sp, int G, const SelfPlayCfg &c, cudaStream_t s) {
    k_sp_begin<<<(G * 32 + 127) / 128, 128, 0, s>>>(sp, G, c);
}
void launch_sp_apply(const SelfPlayArrays &sp, int G, const SelfPlayCfg &c, cudaStream_t s) {
    k_sp_apply<<<(G * 32 + 127) / 128, 128, 0, s>>>(sp, G, c);
}

}  // namespace gz
