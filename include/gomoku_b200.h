/*
 * gomoku_b200.h -- C ABI of the B200-native batched self-play engine (libgomoku_b200.so).
 *
 * The reference (gitMasterLiujiahui/AlphaZero-Gomoku) is pure Python and has no FFI layer; its
 * boundary for the MCTS hot path is the Python class surface AlphaZeroGomokuAI.get_move /
 * GomokuModel.predict.  The drop-in Python modules in alphazero-gomoku_b200/ keep that surface and
 * call the entry points below through ctypes.  Each entry point cites the reference code it
 * replaces (file:line, relative to the reference root).
 *
 * Conventions
 *   - every function returns 0 on success, a negative gz_status otherwise; gz_last_error() gives
 *     the message of the last failure on the calling thread.  No exception crosses the ABI.
 *   - pointers named d_* are DEVICE pointers owned by the caller (e.g. torch tensors); pointers
 *     named h_* are HOST pointers.  `stream` is a cudaStream_t passed as void* (NULL = default).
 *   - one handle per GPU; a handle is not thread-safe.
 *   - cells are raster order, index = row*15 + col; 0 empty, 1 black, 2 white
 *     (gomoku_board.py:26-29).  Bitsets over cells are 8 x uint32, bit i of the 225-bit raster set
 *     at word i/32, bit i%32.
 *   - there is no CPU fallback: every compute call fails with GZ_ERR_CUDA if no device is present.
 */
#ifndef GOMOKU_B200_H
#define GOMOKU_B200_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct gz_engine *gz_handle;

typedef enum {
    GZ_OK = 0,
    GZ_ERR_ARG = -1,     /* bad argument (null pointer, size out of range, weights not set) */
    GZ_ERR_CUDA = -2,    /* CUDA runtime / driver error, including "no device" */
    GZ_ERR_ALLOC = -3,   /* out of device memory */
    GZ_ERR_STATE = -4    /* call out of order (e.g. gz_get_tree before gz_search) */
} gz_status;

/* Board state in HBM.  Replaces GomokuBoard's int64[15,15] array + history list
 * (gomoku_board.py:40-53): two 256-bit bitboards with a 16-column pitch, bit 16*row+col; column 15
 * and row 15 are always zero so that shifts by 1/15/16/17 never wrap into a real cell. 80 bytes. */
typedef struct {
    uint32_t stones[2][8]; /* [0] black, [1] white */
    uint16_t move_count;   /* len(move_history) */
    uint8_t to_move;       /* current_player: 1 black, 2 white */
    uint8_t status;        /* 0 live, 1 black won, 2 white won, 3 draw (game_over, winner) */
    uint32_t reserved[3];  /* [0]: simulation budget of this position inside gz_search, 0 = p->num_simulations (the
                              arena sets it: side A and side B may search with different budgets); [1..2]: zero */
} gz_board;

typedef struct {
    int32_t device;          /* CUDA ordinal */
    int32_t max_games;       /* capacity G of every batched call */
    int32_t max_simulations; /* capacity S; a tree holds at most S+1 nodes (one node per simulation,
                                ai_agent.py:206-222) */
    int32_t precision;       /* 0 = fp32 FFMA networks (parity mode, 1e-4),
                                1 = bf16 tcgen05 networks (throughput mode, 2e-2) */
    int32_t store_priors;    /* keep MCTSNode.prior_prob per node (ai_agent.py:523,564-582) */
    int32_t lanes;           /* gz_search splits its G games into this many independent groups, each
                                advanced on its own CUDA stream so that one group's integer kernels
                                may overlap another's tensor kernels; 0 = default (1: every kernel fills the
                                GPU by itself, 2-4 lanes measure within 2 % of one).  Lane boundaries fall on
                                multiples of 128 games.  Results do not depend on it (games are independent). */
} gz_config;

/* Mirrors AlphaZeroGomokuAI.params + ctor arguments (ai_agent.py:29-31,65-92) and
 * BGPlannerAI.params (bg_planner.py:215-219); re-read at every call because callers mutate
 * ai.params after construction (training.py:240,246). */
typedef struct {
    int32_t num_simulations;
    int32_t planner_steps;
    int32_t planner_k;       /* 8 / 12 / 16 */
    int32_t deterministic;   /* 1: random.choice -> seq[0], random.random -> 1.0 */
    double c_puct;
    double beta;
    double planner_alpha;    /* 0.5 / 0.65 / 0.75 */
    double planner_explore;  /* 0.2 / 0.1 / 0.05 */
    uint64_t seed;           /* Philox key */
    uint32_t game_base;      /* global id of game 0 of this call (multi-GPU sharding keeps streams
                                independent of the GPU count) */
    int32_t eval_nodes;      /* 1: run the policy-value net for every created non-terminal node, as
                                the reference does (ai_agent.py:522-523); 0 skips it (diagnostic) */
    double time_limit_s;     /* AlphaZeroGomokuAI.time_limit (ai_agent.py:183-189): the search stops starting new
                                simulations once this many wall-clock seconds have passed since the call began.
                                The reference looks at the clock before every simulation; here the host looks every
                                4 waves (all G trees advance in lock-step, so they all stop at the same simulation
                                count).  <= 0 or infinite = no limit (no host synchronisation inside the search). */
} gz_search_params;

typedef struct {
    int64_t simulations;     /* _mcts_simulation calls executed */
    int64_t nodes_created;
    int64_t pv_evals;        /* policy-value forwards (non-terminal nodes created, roots included) */
    int64_t planner_steps;   /* BGPlannerAI.get_move calls executed inside rollouts */
    int64_t rollout_steps;   /* _select_offensive_move calls executed */
    int64_t terminal_leaves; /* simulations whose leaf was already game_over */
    int64_t kernel_launches; /* engine kernels launched since gz_create / last reset */
    int64_t last_search_simulations; /* simulations per tree of the last gz_search (< num_simulations when
                                        time_limit_s cut it short) */
} gz_counters;

/* ---- life cycle --------------------------------------------------------------------------- */
int gz_create(const gz_config *cfg, gz_handle *out);
int gz_destroy(gz_handle h);
const char *gz_last_error(void);
int gz_version(void);

/* ---- weights and tables ------------------------------------------------------------------- */
/* GomokuModel.load_model / `ai.model = ...` (neural_network.py:310-333, training.py:247-248):
 * h_flat is AlphaZeroGomokuNet.state_dict() flattened in state-dict order, fp32, without the
 * num_batches_tracked counters (gz_pv_param_count() floats).  BatchNorm (eval mode, eps 1e-5,
 * neural_network.py:142,87-88) is folded into the convolutions here. */
int gz_set_pv_weights(gz_handle h, const float *h_flat, int64_t n);
/* BGPlannerAI's GraphNet and OpponentDQN (bg_planner.py:210-211), same flattening. */
int gz_set_planner_weights(gz_handle h, const float *h_graphnet, int64_t n_graphnet,
                           const float *h_dqn, int64_t n_dqn);
/* Host-computed np.log(n) and np.tanh(i*50/10000.0) tables, so that the float64 selection score
 * (ai_agent.py:532-562) is bit-identical to numpy's on the calling host.  Optional: without them the
 * device's log()/tanh() are used. */
int gz_set_tables(gz_handle h, const double *h_log, int32_t n_log, const double *h_tanh,
                  int32_t n_tanh);
int64_t gz_pv_param_count(void);
int64_t gz_graphnet_param_count(void);
int64_t gz_dqn_param_count(void);

/* ---- boards (gomoku_board.py) -------------------------------------------------------------- */
/* GomokuBoard -> bitboards.  d_cells u8[G,225]; d_to_move u8[G]; d_move_count u16[G].  status is
 * derived: a five (or more) of either colour, else draw when move_count >= 200
 * (gomoku_board.py:103-108). */
int gz_boards_from_cells(gz_handle h, const uint8_t *d_cells, const uint8_t *d_to_move,
                         const uint16_t *d_move_count, int32_t G, gz_board *d_boards, void *stream);
/* Inverse, for inspection: u8[G,225]. */
int gz_boards_to_cells(gz_handle h, const gz_board *d_boards, int32_t G, uint8_t *d_cells,
                       void *stream);
/* make_move (gomoku_board.py:84-113) for one move per board; d_moves i32[G] raster index;
 * d_ok u8[G] receives make_move's return value. */
int gz_make_moves(gz_handle h, gz_board *d_boards, const int32_t *d_moves, int32_t G, uint8_t *d_ok,
                  void *stream);
/* Per-board analysis, the parity surface of the board kernel.  Any output pointer may be NULL.
 *   d_legal      u32[G,8]    get_valid_moves as a raster bitset       gomoku_board.py:201-213
 *   d_win_cells  u32[G,2,8]  empty cells where a stone of colour p makes five    :139-165
 *   d_three_cells u32[G,2,8] empty cells where a stone of p gives threat level >= 3   ai_agent.py:383-401
 *   d_threat     i32[G,2]    _evaluate_threat_level for black, white  ai_agent.py:403-430
 *   d_pattern    i64[G,2]    KnowledgeSearch._pattern_score           bg_planner.py:133-196 */
int gz_board_ops(gz_handle h, const gz_board *d_boards, int32_t G, uint32_t *d_legal,
                 uint32_t *d_win_cells, uint32_t *d_three_cells, int32_t *d_threat,
                 int64_t *d_pattern, void *stream);

/* ---- networks ------------------------------------------------------------------------------ */
/* GomokuModel.predict batched (neural_network.py:214-252, forward :132-159): planes are generated
 * from the bitboards in-kernel.  d_logits f32[G,225], d_probs f32[G,225] (softmax over all 225
 * cells), d_value f32[G] (tanh).  Any output may be NULL. */
int gz_pv_forward(gz_handle h, const gz_board *d_boards, int32_t G, float *d_logits, float *d_probs,
                  float *d_value, void *stream);
/* GraphNet.forward + softmax and OpponentDQN.forward (bg_planner.py:50-56,74-78,243-250):
 * d_gn_logits f32[G,225], d_gn_policy f32[G,225], d_dqn_q f32[G,225]. */
int gz_planner_nets_forward(gz_handle h, const gz_board *d_boards, int32_t G, float *d_gn_logits,
                            float *d_gn_policy, float *d_dqn_q, void *stream);

/* gz_pv_forward without any allocation.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA-graph capture, tight serving loops, harnesses
 * that forbid it).  Same outputs as gz_pv_forward on a precision-0 handle (the fp32 FFMA kernels, weights of
 * gz_set_pv_weights; a precision-1 handle also computes in fp32 here), but the activation scratch is a file-scope
 * `static __device__` array inside the library instead of the handle's buffers: 3 x GZ_STATIC_MAX_BOARDS x 128 x 225
 * floats = 5 529 600 bytes, held by every GPU that loads the library until exit.  G > GZ_STATIC_MAX_BOARDS is
 * rejected with GZ_ERR_ARG.  The call performs no cudaMalloc / cudaFree and uses no allocating library routine; the
 * caller owns inputs and outputs (any output may be NULL).  The array is shared by all handles and calls on a device
 * and keeps the last call's contents (nothing needs clearing: each layer overwrites what the next one reads); calls
 * that overlap on the GPU (different streams) race on it -- serialise them.  The array's device address is looked up
 * on the first call per device: make that call outside a graph capture. */
#define GZ_STATIC_MAX_BOARDS 16
int gz_pv_forward_static(gz_handle h, const gz_board *d_boards, int32_t G, float *d_logits, float *d_probs,
                         float *d_value, void *stream);

/* Inspection: output of the input convolution (nlayers = 0) or of the first nlayers (1..4) tower
 * convolutions of AlphaZeroGomokuNet in the handle's precision, as fp32 NCHW d_out f32[G,128,225]
 * (neural_network.py:142-146, :85-91).  Used by the per-layer parity tests. */
int gz_debug_pv_layer(gz_handle h, const gz_board *d_boards, int32_t G, int32_t nlayers, float *d_out,
                      void *stream);

/* ---- planner move (bg_planner.py:232-269) --------------------------------------------------- */
/* One BGPlannerAI.get_move per board.  d_ai_player u8[G] is the planner's fixed colour
 * (ai_agent.py:60).  Outputs (any may be NULL):
 *   d_scores2  i64[G,225]  2*score_move for every empty cell (exact integers; occupied = INT64_MIN)
 *   d_topk     i32[G,16]   stable descending top-k, -1 padded              bg_planner.py:108-114
 *   d_composed f64[G,16]   alpha*policy - (1-alpha)*q per candidate        :256-264
 *   d_move     i32[G]      the move returned (-1 if the board has no empty cell)
 * `step` is the rollout step index used to key the explore roll. */
int gz_planner_move(gz_handle h, const gz_board *d_boards, const uint8_t *d_ai_player, int32_t G,
                    const gz_search_params *p, int32_t sim, int32_t step, int64_t *d_scores2,
                    int32_t *d_topk, double *d_composed, int32_t *d_move, void *stream);

/* ---- rollout (ai_agent.py:251-430) ----------------------------------------------------------- */
/* One _select_offensive_move per board: d_move i32[G]; d_cand u32[G,8] = the candidate list handed
 * to random.choice (all zero when tier 1 returned a winning move). */
int gz_rollout_policy(gz_handle h, const gz_board *d_boards, int32_t G, const gz_search_params *p,
                      int32_t sim, int32_t step, int32_t *d_move, uint32_t *d_cand, void *stream);
/* One _simulate per board: up to planner_steps planner moves, then the offensive-random tail, at
 * most 100 steps in total.  d_value f64[G] in {1,-1,0.1,0}; d_steps i32[G]; d_final the final
 * rollout boards (may be NULL). */
int gz_rollout(gz_handle h, const gz_board *d_boards, const uint8_t *d_ai_player, int32_t G,
               const gz_search_params *p, int32_t sim, double *d_value, int32_t *d_steps,
               gz_board *d_final, void *stream);

/* ---- search (ai_agent.py:168-249,441-454,532-562) --------------------------------------------- */
/* One _mcts_search per board, all G trees advanced in lock-step, one simulation in flight per tree.
 *   d_child_visits i32[G,225], d_child_value f64[G,225]: root children indexed by move (0 where the
 *   root has no such child); d_best_move i32[G]: first root child with maximal visits
 *   (ai_agent.py:199-201), -1 if the root has no child. */
int gz_search(gz_handle h, const gz_board *d_boards, const uint8_t *d_ai_player, int32_t G,
              const gz_search_params *p, int32_t *d_child_visits, double *d_child_value,
              int32_t *d_best_move, void *stream);
/* Same through HOST buffers, the call behind AlphaZeroGomokuAI.get_move: copies the positions in,
 * searches, copies the answers out, synchronises.  h_cells u8[G,225], h_to_move u8[G],
 * h_move_count u16[G], h_ai_player u8[G]; outputs as gz_search (any may be NULL). */
int gz_search_host(gz_handle h, const uint8_t *h_cells, const uint8_t *h_to_move,
                   const uint16_t *h_move_count, const uint8_t *h_ai_player, int32_t G,
                   const gz_search_params *p, int32_t *h_child_visits, double *h_child_value,
                   int32_t *h_best_move);
/* Whole tree of game g after gz_search, creation order (node 0 = root): host arrays of `cap`
 * entries; *n_nodes receives the node count. */
int gz_get_tree(gz_handle h, int32_t g, int32_t cap, int32_t *h_parent, int32_t *h_move,
                int32_t *h_visits, double *h_value, int32_t *n_nodes);

/* MCTSNode.prior_prob of node `node` (creation order, 0 = root) of game g after gz_search on a handle created with
 * store_priors = 1 (ai_agent.py:522-523, _get_prior_probability :564-582: GomokuModel.predict's softmax gathered
 * over the node's valid moves and renormalised).  h_prior f32[225] indexed by raster cell, 0 where the cell is
 * occupied (the reference keeps the E values of the empty cells in raster order: compress to compare);
 * *h_value receives the value-head output of the same forward (computed and discarded by the reference,
 * ai_agent.py:570).  Terminal nodes have no prior (ai_agent.py:522): GZ_ERR_STATE.  Either output may be NULL. */
int gz_get_priors(gz_handle h, int32_t g, int32_t node, float *h_prior, float *h_value);

/* ---- self-play (training.py:141-218 play_one_game, :77-97 SimpleReplay) ---------------------------------- */
/* G games are played concurrently on the device, one ply of every game per gz_selfplay_step: the
 * position is searched (gz_search with ai_player = the side to move, as play_one_game alternates
 * ai_black / ai_white), with probability `exploration` the searched move is replaced by a uniform random
 * legal move (ai_agent.py:128-129), the move is played, and the ply is recorded BEFORE the move like
 * SimpleReplay.add.  The first six plies of a game come from the opening book (ai_agent.py:138-166:
 * (7,7), then a random free cell of the centre 3x3, else of the centre 5x5) and are played at once when
 * a game starts, so every step searches G live positions.  When a game ends its plies are labelled
 * (+1 mover won, -1 mover lost, 0 draw; training.py:89-94), appended to the replay buffer, and the slot
 * starts a new game.  The wall-clock time-outs of play_one_game do not exist here. */
typedef struct {
    uint32_t stones[2][8]; /* position before the move, gz_board bit layout ([0] black, [1] white):
                              the reference stores 2700 bytes of float planes instead (training.py:84) */
    uint8_t move;          /* row*15 + col */
    uint8_t mover;         /* 1 black, 2 white */
    int8_t outcome;        /* +1 / -1 / 0 from the mover's point of view */
    uint8_t ply;           /* index of the ply inside its game */
    uint32_t game_id;      /* global id: p->game_base + slot + 2^20 * (games this slot finished before) */
} gz_replay_record;        /* 72 bytes */

typedef struct {
    int64_t games_finished, plies_played, black_wins, white_wins, draws;
    int64_t records_ready;   /* records waiting in the replay buffer */
    int64_t records_dropped; /* records lost because the buffer was full (drain more often) */
    int64_t steps;           /* gz_selfplay_step calls since gz_selfplay_begin */
} gz_selfplay_stats;

/* Starts G (<= max_games) fresh games.  `p` as for gz_search (num_simulations <= max_simulations);
 * p->seed keys every random choice, p->game_base is the id of this handle's first game (rank * G on a
 * multi-GPU run, so ids and random streams are global).  Replay capacity: replay_capacity records
 * (0 = 256 per game).
 * games_total = 0: open-ended session, every finished game is counted and recorded (throughput runs).
 * games_total = N > 0: exactly N complete games, like the reference's `for game in range(N): play_one_game`
 * (training.py:415-420): the counted games are those with start ordinal generation * G + slot < N (a set fixed
 * before any game is played, so game length cannot select what is recorded); later games only keep their slot
 * busy and are neither counted nor recorded.  The session is complete when games_finished == N. */
int gz_selfplay_begin(gz_handle h, int32_t G, const gz_search_params *p, double exploration,
                      int64_t replay_capacity, int64_t games_total, void *stream);
/* One ply of every game. */
int gz_selfplay_step(gz_handle h, void *stream);
/* `ai.params["num_simulations"] = n` between plies of a running self-play session (the reference re-reads params at
 * every get_move, ai_agent.py:181; training.py:240 mutates them): the following steps search with n simulations.
 * bench.py uses it to spread the games over all ages with cheap plies before it times full-budget ones. */
int gz_selfplay_set_simulations(gz_handle h, int32_t num_simulations);
/* Moves up to `cap` finished-game records to d_out (device) and frees their space; *h_n = count. */
int gz_selfplay_drain(gz_handle h, gz_replay_record *d_out, int64_t cap, int64_t *h_n, void *stream);
/* Current boards of the G games (device, for inspection) and the counters. */
int gz_selfplay_boards(gz_handle h, gz_board *d_out, void *stream);
int gz_selfplay_get_stats(gz_handle h, gz_selfplay_stats *out);

/* ---- evaluation arena (training.py:221-270 evaluate_model) ------------------------------- */
/* `games_total` evaluation games of side A against side B, G of them at a time, on the self-play machinery above
 * (gz_selfplay_step plays one A ply of every game; gz_selfplay_drain / gz_selfplay_boards work unchanged).
 * Game ordinals are global: slot s of this handle plays ordinals p->game_base + s + k * ordinal_stride, k = 0, 1, ...
 * (ordinal_stride = G * number of ranks), and the ordinal is the record's game_id.  As in evaluate_model side A
 * ("current") is BLACK in even ordinals and WHITE in odd ones (black_is_current toggles per game, :232,266).
 *   GZ_ARENA_ENGINE  B is the same agent as A playing the other colour (the reference's `baseline` AlphaZeroGomokuAI;
 *                    the two differ only in `.model`, which the search never reads -- SURVEY fact 3): both sides use
 *                    the opening book for the first six stones, then the search with ai_player = the side to move,
 *                    then the epsilon override.
 *   GZ_ARENA_RANDOM  B is evaluate_model's RandomAgent (:225-228): random.choice over the legal moves, from the very
 *                    first stone (no book, no epsilon).  B's reply is played in the same step as A's move, and B's
 *                    plies before A's first searched move are played when the game starts, so every step still
 *                    searches G positions with A to move.
 * Games with ordinal >= games_total keep the slot busy but are neither counted nor recorded. */
enum { GZ_ARENA_ENGINE = 1, GZ_ARENA_RANDOM = 2 };
typedef struct {
    int64_t games_finished;      /* counted games (ordinal < games_total) */
    int64_t wins, losses, draws; /* from side A's point of view (training.py:258-264) */
    int64_t plies_played, steps;
} gz_arena_stats;
/* p->num_simulations is side A's budget; sims_b (GZ_ARENA_ENGINE only, 0 = the same) is side B's: evaluate_model
 * overrides num_simulations of the baseline AI alone when a baseline is given (:241-246), so the current model keeps
 * its preset (100 for "easy") while the baseline searches eval_num_sim.  Both must be <= max_simulations. */
int gz_arena_begin(gz_handle h, int32_t G, const gz_search_params *p, double exploration, int32_t opponent,
                   int32_t sims_b, int64_t games_total, int64_t ordinal_stride, int64_t replay_capacity,
                   void *stream);
int gz_arena_get_stats(gz_handle h, gz_arena_stats *out);

/* ---- dataset build ------------------------------------------------------------ */
/* GomokuSelfPlayDataset + augment_sample (training.py:104-134, 38-70): training samples from packed replay
 * records, decoded on the device.  The virtual dataset has n + 8*m samples:
 *   sample i < n                : record i as recorded;
 *   sample n + 8*j + 2*k + f    : record d_chosen[j] under symmetry (k quarter turns, then a horizontal flip when
 *                                 f = 1), in the order augment_sample emits them (k = 0..3, f = 0, 1).
 * Which records are augmented (random.sample over range(n), training.py:118-119) is the caller's choice.
 * Emits `count` samples: ids d_sample_ids[0..count) when given, else first .. first + count - 1.
 *   d_planes  f32 [count][3][15][15]  black, white, empty (GomokuBoard.get_board_tensor), transformed as
 *                                     _transform_planes does: np.rot90 counter-clockwise k times, then np.flip(axis=2)
 *   d_move    i64 [count]             the played move's index under _transform_index
 *   d_value   f32 [count]             the record's outcome (+1 / -1 / 0 from the mover's point of view)
 * reference_exact != 0 reproduces _transform_index as written: it turns the index CLOCKWISE while the planes turn
 * counter-clockwise, so for k = 1, 3 the label does not point at the played stone (probed);
 * reference_exact = 0 turns the index with the planes. */
int gz_dataset_build(gz_handle h, const gz_replay_record *d_records, int64_t n, const int32_t *d_chosen,
                     int64_t m, const int64_t *d_sample_ids, int64_t first, int64_t count,
                     int32_t reference_exact, float *d_planes, int64_t *d_move, float *d_value, void *stream);

/* ---- stage timing ---------------------------------------------------------------------------------- */
/* CUDA-event timing of the engine's kernel groups on the launching stream.  When enabled, every
 * group of launches inside gz_search / gz_rollout / gz_pv_forward / ... is bracketed by two events;
 * gz_get_stage_times synchronises, accumulates elapsed milliseconds and launch counts per stage and
 * clears the pending events.  Off by default (the events cost a few microseconds per wave). */
enum {
    GZ_STAGE_TREE = 0,         /* select / expand / backup */
    GZ_STAGE_PV_CONV0 = 1,     /* 3 -> 128 input convolution (planes generated in-kernel) */
    GZ_STAGE_PV_TOWER = 2,     /* the four 128 -> 128 3x3 convolutions (dominant FLOPs) */
    GZ_STAGE_PV_HEADS = 3,     /* policy / value heads, softmax, prior */
    GZ_STAGE_PLAN_SCORE = 4,   /* knowledge-search scoring + top-k */
    GZ_STAGE_GN_CONVS = 5,     /* GraphNet convolutions */
    GZ_STAGE_GN_HEAD = 6,      /* GraphNet head + softmax; inside gz_search (bf16) also OpponentDQN: one launch */
    GZ_STAGE_DQN = 7,          /* OpponentDQN when launched alone (gz_planner_nets_forward, fp32 mode) */
    GZ_STAGE_PLAN_PICK = 8,    /* blend, explore roll, make_move */
    GZ_STAGE_ROLLOUT = 9,      /* offensive-random tail */
    GZ_STAGE_MISC = 10,        /* import / export / finalize */
    GZ_N_STAGES = 12
};
int gz_set_profiling(gz_handle h, int32_t on);
/* Number of game groups gz_search advances on separate streams from now on (gz_config.lanes; 0 = automatic,
 * 1 = every kernel on the caller's stream: the setting stage times are meaningful under). */
int gz_set_lanes(gz_handle h, int32_t lanes);
int gz_get_stage_times(gz_handle h, double *h_ms /*[GZ_N_STAGES]*/, int64_t *h_launches /*[GZ_N_STAGES]*/);
/* Batches of at most 256 games are launch-bound (about 50 kernels per wave with microseconds of work each; the
 * get_move call of one interactive game is G = 1): gz_search runs wave 0 eagerly, captures one wave into a CUDA graph
 * whose kernels read the simulation index from device memory, and replays it for the other simulations.  Results are
 * identical either way; on = 0 disables the graph path (stage profiling disables it too). */
int gz_set_graph(gz_handle h, int32_t on);

/* ---- counters ---------------------------------------------------------------------------------- */
int gz_get_counters(gz_handle h, gz_counters *out);
int gz_reset_counters(gz_handle h);

#ifdef __cplusplus
}
#endif
#endif
