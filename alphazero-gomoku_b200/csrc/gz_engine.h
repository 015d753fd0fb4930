// gz_engine.h -- internal state of one engine handle and the kernel launchers (one per .cu file).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <vector>

#include "../../include/gomoku_b200.h"

namespace gz {

constexpr int kCells = 225;
constexpr int kMaxTopK = 16;
constexpr int kLogTableMax = 2048;
constexpr int kTanhTableMax = 4096;
// bf16 activation planes (gz_nn_bf16.cu): [board][channel group][288 positions][8 channels]
constexpr int kPlanePos = 288;
constexpr int kPlaneBytes = kPlanePos * 16;
constexpr int kGuardBytes = 256;
// row pitch (floats) of the engine's internal softmax / prior rows: 225 cells padded for 16-byte stores
constexpr int kPolicyPitch = 256;

// A operand of the head GEMMs (gz_nn_fc.cu): boards in tiles of 128 rows, K in chunks of 64, every chunk a
// bf16 "hi" block followed by a "lo" block (x = hi + lo to ~16 mantissa bits), each block 8 K-groups x 128
// rows x 8 values -- the K-major SWIZZLE_NONE core-matrix order tcgen05.mma reads, so a chunk is one 32 KB
// bulk copy.  The convolution epilogues that produce the values write them straight into this form.
constexpr int kHeadBlockBytes = 8 * 128 * 16;                 // one hi or lo block
constexpr int kHeadChunksP = 8, kHeadChunksV = 4;             // ceil(450 / 64), ceil(225 / 64)
constexpr size_t kHeadTileBytesP = (size_t)kHeadChunksP * 2 * kHeadBlockBytes;
constexpr size_t kHeadTileBytesV = (size_t)kHeadChunksV * 2 * kHeadBlockBytes;
#if defined(__CUDACC__)
// Programmatic dependent launch.  A kernel launched through launch_pdl may begin (barrier set-up, tensor-memory
// allocation, weight loads) while the kernel before it in the stream is still running, and MUST call pdl_wait() before
// it reads or writes anything that kernel touches: pdl_wait() returns once the whole preceding grid has completed and
// its writes are visible.  pdl_trigger() in the preceding kernel lets the next grid's CTAs be scheduled as soon as all
// of this grid's CTAs have started (they take an SM whenever its resources allow); without it they are scheduled when
// this grid exits.  Both are no-ops for kernels launched the ordinary way.
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
bool pdl_enabled();   // GZ_PDL=1 switches the attribute on (gz_api.cu: off by default, it measured no gain)
template <typename... KArgs, typename... Args>
inline void launch_pdl(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t s, Args &&...args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = s;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at;
    cfg.numAttrs = pdl_enabled() ? 1 : 0;
    cudaLaunchKernelEx(&cfg, kern, KArgs(args)...);
}
__host__ __device__ __forceinline__ size_t head_operand_offset(int board, int k, int chunks) {   // of the hi value
    return ((size_t)(board >> 7) * chunks + (k >> 6)) * 2 * kHeadBlockBytes + ((k >> 3) & 7) * 2048 +
           (board & 127) * 16 + (k & 7) * 2;
}
#endif

// fp32 network parameters on the device, BN folded, laid out for the FFMA kernels:
// conv weights as [cin][tap][cout] (cout contiguous), linear weights transposed to [in][out].
struct ConvW {
    float *w;  // [cin*taps][cout]
    float *b;  // [cout]
};
struct PvWeights {
    ConvW conv0;         // 3 -> 128, 3x3 (+bn)
    ConvW tower[4];      // block0.conv1, block0.conv2, block1.conv1, block1.conv2 (+bn each)
    float *pconv_w, *pconv_b;   // [128][2], [2]
    float *pfc_wt, *pfc_b;      // [450][225], [225]
    float *vconv_w, *vconv_b;   // [128], [1]
    float *vfc1_wt, *vfc1_b;    // [225][64], [64]
    float *vfc2_w, *vfc2_b;     // [64], [1]
};
struct GraphNetWeights {
    ConvW embed;         // 3 -> 64, 3x3
    ConvW l3[4];         // 64 -> 64, 3x3
    ConvW l1[4];         // 64 -> 64, 1x1
    float *hconv_w, *hconv_b;   // [64][2], [2]
    float *hfc_wt, *hfc_b;      // [450][225], [225]
};
struct DqnWeights {
    float *w1t, *b1;     // [675][256]
    float *w2t, *b2;     // [256][256]
    float *w3t, *b3;     // [256][225]
};

// bf16 tensor-core path: conv weights packed per K-step as [8 channel groups][cout][8 channels]
// (K-major SWIZZLE_NONE core matrices), BN folded; biases and the first layer reuse the fp32 packs.
struct Bf16Weights {
    float gn_host_bias[4][128];   // host copies for k_gn_block's parameter block: per block b3[64], b1[64]
    float gn_host_head[130];      // GraphNet head 1x1: w[64][2] (device layout of hconv_w), b[2]
    float gn_host_embed_bias[64]; // GraphNet embed bias, for k_gn_block0_embed's parameter block
    uint8_t *gn_embed_i2c_w;      // embed weights as ONE K = 32 GEMM operand: [4 k-groups][64 cout][8 k], k = 3 * (3 * plane + ky) + kx
    uint8_t *pv_tower_w[4];
    uint8_t *gn_l3_w[4];
    uint8_t *gn_l3p_w[4];   // the same 3x3 weights in k_gn_block_pairs' order: 3 x [dx=0 | dx=-1] stacked along N, 3 x dx=+1
    uint8_t *gn_l1_w[4];
    uint8_t *pv_conv0_w, *gn_embed_w;   // input convolutions, [tap][2 cg][cout][8]
    uint8_t *pv_blob, *gn_blob;
    uint8_t *pv_pfc_w, *pv_vfc1_w, *gn_hfc_w;   // head FCs, K-chunks of [8 k-groups][n][8 k] (gz_nn_fc.cu)
    float *pv_pfc_b, *gn_hfc_b;                  // [256] padded biases
    uint8_t *pv_head_blob, *gn_head_blob;
    uint8_t *dqn_w;       // OpponentDQN: 19 K-chunks of [8 k-groups][256 n][8 k] bf16 (gz_nn_fc.cu)
    float *dqn_bias;      // [3][256]
};

// Flat SoA tree storage, tree g owns slots [g*cap, (g+1)*cap); node 0 is the root.
struct TreeArrays {
    int32_t cap;          // max_simulations + 1
    int16_t *parent;      // -1 for the root
    uint8_t *move;        // bit position 16*r+c of the move leading here
    uint8_t *status;      // 0 live, 1/2 winner, 3 draw
    int16_t *nchild;      // children created so far
    int16_t *nempty;      // empty cells of the node's board = len(unexplored) at creation
    int32_t *visits;
    double *value;
    double *bgb;          // beta * tanh(pattern_score/1e4)
    float *vhead;         // value head output of the node (never read by the search, like the reference)
    float *prior;         // optional [cap][kPolicyPitch] renormalised priors (MCTSNode.prior_prob)
    int32_t *n_nodes;     // per tree
    gz_board *root;       // per tree
    uint8_t *ai_player;   // per tree
};

// Per-simulation state of every game (one simulation in flight per tree).
struct SimArrays {
    gz_board *board;      // leaf board, then the rollout board
    int32_t *leaf;        // node index of the leaf
    uint8_t *need_eval;   // leaf was created in this wave and is not terminal
    int32_t *steps;       // rollout steps taken so far (planner steps included)
    double *value;        // rollout result
    // planner scratch
    int32_t *topk;        // [G][16]
    int32_t *ntop;        // [G]
    int64_t *scores2;     // [G][225] (only when requested)
};

constexpr int kMaxLanes = 4;
// gz_search replays its waves from a CUDA graph when the batch is at most this many games (launch-bound regime)
constexpr int kGraphMaxGames = 256;

// ---- self-play state (gz_kernels_selfplay.cu) ---------------------------------------------------
constexpr int kMaxPlies = 200;     // a game ends at the 200th stone at the latest (gomoku_board.py:106)
enum { SP_CURSOR = 0, SP_GAMES, SP_PLIES, SP_BLACK, SP_WHITE, SP_DRAWS, SP_DROPPED, SP_A_WINS, SP_A_LOSSES,
       SP_NCTR = 12 };
struct SelfPlayArrays {
    gz_board *boards;            // [G] current positions
    uint8_t *ai;                 // [G] side to move = the searching AI's colour
    int32_t *best;               // [G] answer of the last search
    uint32_t *game_id;           // [G]
    uint32_t *generation;        // [G] games this slot has finished
    gz_replay_record *staging;   // [G][kMaxPlies] plies of the game in progress
    gz_replay_record *out;       // [out_cap] finished games, labelled
    long long out_cap;
    unsigned long long *ctr;     // [SP_NCTR]
};
struct SelfPlayCfg {
    uint32_t seed_lo, seed_hi, id_base, id_stride;
    int32_t deterministic;
    // evaluation arena (training.py:221-270): 0 = plain self-play; GZ_ARENA_ENGINE / GZ_ARENA_RANDOM: side A's colour
    // alternates with the game ordinal (= game id), games with ordinal >= games_total are played but not counted
    int32_t arena;
    int32_t n_slots;             // G of the session
    long long games_total;       // self-play: 0 = open-ended, else games with ordinal generation * G + slot < this count
    int32_t sims_a, sims_b;      // arena: simulations per search of side A / side B (evaluate_model overrides only B's)
    double exploration;
};

struct Engine {
    // A lane view (gz_api.cu: lane_view) is an Engine whose per-game pointers are offset to a group
    // of games; it owns nothing and forwards profiling / launch counting to `root`.
    Engine *root;
    Engine *lane[kMaxLanes];
    cudaStream_t lane_stream[kMaxLanes];
    cudaEvent_t lane_done[kMaxLanes], lane_go;
    SelfPlayArrays sp;
    SelfPlayCfg sp_cfg;
    gz_search_params sp_params;
    int sp_G;                 // 0 = no self-play session
    long long sp_steps, sp_alloc_games, sp_alloc_out;
    gz_config cfg;
    int sm_count;
    cudaStream_t own_stream;
    bool have_pv, have_planner;
    PvWeights pv;
    GraphNetWeights gn;
    DqnWeights dqn;
    Bf16Weights bw;
    uint8_t *planes[3];   // bf16 activation planes, [G][16][288][8] + guards (precision 1)
    void *weight_blob;    // single allocation backing all weights
    double *log_table;    // device
    double *tanh_table;   // device
    int32_t n_log, n_tanh;
    TreeArrays tree;
    SimArrays sim;
    // network activations (fp32 path): three ping-pong buffers [G][128][225]
    float *act[3];
    float *gn_logits, *gn_policy, *dqn_q;   // [G][225], [G][kPolicyPitch], [G][225]
    // fused head-conv outputs, already in the form the head GEMMs read them (head_operand_offset below):
    // policy 450 inputs -> 8 K-chunks, value 225 -> 4, per tile of 128 boards
    uint8_t *pv_pc, *pv_vc, *gn_pc;
    float *pv_logits, *pv_probs, *pv_value; // [G][225], [G][225], [G]
    gz_board *tmp_boards; // [G]
    uint8_t *tmp_u8;      // [G*232] host-call staging: cells, to_move, ai_player, move_count
    int32_t *io_cv;       // [G][225] host-call result staging
    double *io_cval;      // [G][225]
    int32_t *io_best;     // [G]
    unsigned long long *counters; // device gz_counters-shaped
    int32_t *d_sim;       // simulation index of the wave in flight when waves are replayed from a CUDA graph
    int64_t launches;     // host-side launch counter
    bool searched;
    bool no_graph;        // gz_set_graph(0): never replay waves from a CUDA graph
    int32_t last_simulations;   // simulations per tree of the last search (< num_simulations when the time limit cut it)
    // optional stage timing with CUDA events on the launching stream (gz_set_profiling)
    bool profiling;
    std::vector<cudaEvent_t> ev_pool;
    std::vector<int> ev_stage;      // stage id of the interval starting at event 2*i
    size_t ev_used;
    double stage_ms[GZ_N_STAGES];
    int64_t stage_launches[GZ_N_STAGES];
};

// brackets `n_launches` kernel launches of stage `stage` with events when profiling is on
void prof_begin(Engine &e, int stage, int n_launches, cudaStream_t s);
void prof_end(Engine &e, cudaStream_t s);

struct SearchCfg {
    int32_t S;
    int32_t planner_steps, planner_k, deterministic, eval_nodes;
    double c_puct, beta, alpha, explore;
    uint32_t seed_lo, seed_hi, game_base;
    // simulation index read from device memory instead of the launch argument (waves replayed from a CUDA graph)
    const int32_t *d_sim;
};

// ---- launchers -------------------------------------------------------------------------------
// board kernels (gz_kernels_board.cu)
void launch_boards_from_cells(const uint8_t *cells, const uint8_t *to_move, const uint16_t *mc, int G,
                              gz_board *out, cudaStream_t s);
void launch_boards_to_cells(const gz_board *b, int G, uint8_t *cells, cudaStream_t s);
void launch_make_moves(gz_board *b, const int32_t *moves, int G, uint8_t *ok, cudaStream_t s);
void launch_board_ops(const gz_board *b, int G, uint32_t *legal, uint32_t *win, uint32_t *three,
                      int32_t *threat, int64_t *pattern, cudaStream_t s);

// rollout kernels (gz_kernels_rollout.cu)
void launch_rollout_policy(const gz_board *b, int G, const SearchCfg &c, int sim, int step,
                           int32_t *move, uint32_t *cand, cudaStream_t s);
void launch_rollout_tail(gz_board *b, const uint8_t *ai_player, int32_t *steps, int G,
                         const SearchCfg &c, int sim, double *value, unsigned long long *counters,
                         cudaStream_t s);

// planner kernels (gz_kernels_planner.cu)
void launch_planner_score(const gz_board *b, const uint8_t *ai_player, const int32_t *steps, int G,
                          const SearchCfg &c, int planner_step, int32_t *topk, int32_t *ntop,
                          int64_t *scores2, cudaStream_t s);
void launch_planner_pick(gz_board *b, const uint8_t *ai_player, int32_t *steps, int G,
                         const SearchCfg &c, int sim, int planner_step, const int32_t *topk,
                         const int32_t *ntop, const float *gn_policy, const float *dqn_q,
                         double *composed, int32_t *move_out, bool apply_move,
                         unsigned long long *counters, cudaStream_t s);

// tree kernels (gz_kernels_tree.cu)
void launch_tree_init(const Engine &e, const gz_board *boards, const uint8_t *ai_player, int G,
                      cudaStream_t s);
void launch_tree_step(const Engine &e, int G, const SearchCfg &c, int sim, cudaStream_t s);
void launch_tree_finalize(const Engine &e, int G, const SearchCfg &c, int32_t *child_visits,
                          double *child_value, int32_t *best_move, cudaStream_t s);
void launch_set_sim(int32_t *d_sim, int value, int add, cudaStream_t s);   // *d_sim = add ? *d_sim + value : value

// self-play kernels (gz_kernels_selfplay.cu)
void launch_sp_begin(const SelfPlayArrays &sp, int G, const SelfPlayCfg &c, cudaStream_t s);
void launch_dataset_build(const gz_replay_record *records, long long n, const int32_t *chosen, long long m,
                          const long long *sample_ids, long long first, long long count, bool reference_exact,
                          float *planes, long long *move, float *value, cudaStream_t s);
void launch_sp_apply(const SelfPlayArrays &sp, int G, const SelfPlayCfg &c, cudaStream_t s);

// networks, fp32 path (gz_nn_fp32.cu)
void nn_pv_forward_fp32(Engine &e, const gz_board *boards, const uint8_t *active, int G,
                        float *logits, float *probs, float *value, float *prior_out,
                        const int32_t *prior_slot, cudaStream_t s);
void nn_planner_forward_fp32(Engine &e, const gz_board *boards, int G, float *gn_logits,
                             float *gn_policy, int policy_pitch, float *dqn_q, cudaStream_t s);

// allocation-free fp32 policy-value forward on a file-scope device array (gz_nn_fp32.cu; gz_pv_forward_static)
cudaError_t nn_pv_forward_static(Engine &e, const gz_board *boards, int G, float *logits, float *probs, float *value,
                                 cudaStream_t s);

// networks, bf16 tcgen05 path (gz_nn_bf16.cu)
void nn_pv_forward_bf16(Engine &e, const gz_board *boards, const uint8_t *active, int G,
                        float *logits, float *probs, float *value, float *prior_out,
                        const int32_t *prior_slot, cudaStream_t s);
void nn_graphnet_convs_bf16(Engine &e, const gz_board *boards, const uint8_t *active, int G, cudaStream_t s);
void nn_debug_planes_to_nchw(Engine &e, int which, int C, int G, float *out, cudaStream_t s);
void nn_debug_pv_layer_bf16(Engine &e, const gz_board *boards, int G, int nlayers, float *out, cudaStream_t s);
void nn_debug_pv_layer_fp32(Engine &e, const gz_board *boards, int G, int nlayers, float *out, cudaStream_t s);
// heads and OpponentDQN as tcgen05 GEMMs (gz_nn_fc.cu)
void nn_heads_bf16(Engine &e, bool pv, const gz_board *boards, const uint8_t *active, int G, float *policy,
                   float *logits, float *probs, float *value, float *prior, const int32_t *prior_slot,
                   cudaStream_t s);
void nn_dqn_bf16(Engine &e, const gz_board *boards, int G, float *dqn_q, cudaStream_t s);
void nn_gnhead_dqn_bf16(Engine &e, const gz_board *boards, int G, float *policy, float *dqn_q, cudaStream_t s);
// precision dispatch
void nn_pv_forward(Engine &e, const gz_board *boards, const uint8_t *active, int G, float *logits,
                   float *probs, float *value, float *prior_out, const int32_t *prior_slot, cudaStream_t s);
void nn_planner_forward(Engine &e, const gz_board *boards, int G, float *gn_logits, float *gn_policy,
                        int policy_pitch, float *dqn_q, cudaStream_t s);

}  // namespace gz
