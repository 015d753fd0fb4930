/*
 * gomoku_oracle.h -- CPU restatement of the reference's MCTS hot path.  TEST INFRASTRUCTURE ONLY.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
 * load this library, and only as the checker or the timed CPU baseline; the product
 * (alphazero-gomoku_b200/) never links, imports or calls it.
 *
 * Parity status: PINNED.  Every function is checked against fixtures produced by running the
 * unmodified Python reference (tests/golden/make_goldens.py -> tests/golden/ fixtures); see
 * tests/test_oracle_vs_golden.py.  The reference itself ships no golden vectors.
 *
 * Each function cites the reference file:line it follows (paths relative to /root/reference).
 * Arrays, not bitboards, on purpose: the CUDA path uses bitboards, so the two share no code and
 * no tricks.
 */
#ifndef GOMOKU_ORACLE_H
#define GOMOKU_ORACLE_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define GZO_N 15
#define GZO_CELLS 225

/* gomoku_board.py:40-53 */
typedef struct {
    int8_t cells[GZO_CELLS]; /* 0 empty, 1 black, 2 white; raster index r*15+c */
    int32_t to_move;         /* current_player */
    int32_t move_count;      /* len(move_history) */
    int32_t game_over;
    int32_t winner;          /* 0 = None */
} gzo_board;

/* Counter-based RNG shared with the CUDA engine (Philox4x32-10).  mode 0 = deterministic
 * (random.choice -> seq[0], random.random -> 1.0), mode 1 = Philox. */
typedef struct {
    int32_t mode;
    uint32_t seed_lo, seed_hi;
    uint32_t game; /* global game / tree id */
    uint32_t sim;  /* simulation index inside the search */
} gzo_rng;

/* fp32 weights of the three networks, flattened in state-dict order without
 * num_batches_tracked (neural_network.py:115-130; bg_planner.py:34-48,68-72). */
typedef struct {
    const float *pv;       /* 711591 floats + BN stats, see gzo_pv_param_count() */
    const float *graphnet;
    const float *dqn;
} gzo_weights;

typedef struct {
    int32_t num_simulations;
    double c_puct;
    double beta;
    int32_t planner_steps;
    int32_t planner_k;      /* 8 / 12 / 16   bg_planner.py:215-219 */
    double planner_alpha;   /* 0.5 / 0.65 / 0.75 */
    double planner_explore; /* 0.2 / 0.1 / 0.05 */
    int32_t ai_player;
    int32_t eval_nodes;     /* run the policy-value net at node creation (ai_agent.py:522-523) */
    const double *log_table;  /* optional: log_table[n] = np.log(n), n in [0, n_log) */
    int32_t n_log;
    const double *tanh_table; /* optional: tanh_table[i] = np.tanh(i*50/10000.0) */
    int32_t n_tanh;
} gzo_params;

typedef struct {
    int32_t n_nodes;
    int32_t best_move;       /* raster index, -1 if the root has no child */
    int64_t n_pv_evals;
    int64_t n_planner_steps;
    int64_t n_rollout_steps;
    int64_t n_terminal_leaves;
    /* Diagnostics of the planner's blended pick (bg_planner.py:256-264), for the near-tie accounting of
     * the parity tests: over every pick whose argmax was returned (no explore override) the smallest
     * difference between the best and the second-best composed score, and where it happened. */
    double min_gap;          /* +inf when no such pick happened */
    int32_t min_gap_sim, min_gap_step;
    int64_t n_argmax_picks;
} gzo_search_stats;

/* ---- rules ---- */
void gzo_board_init(gzo_board *b);
int gzo_board_from_cells(gzo_board *b, const uint8_t *cells, int to_move, int move_count);
int gzo_make_move(gzo_board *b, int r, int c);
int gzo_check_win(const gzo_board *b, int r, int c);
int gzo_legal_mask(const gzo_board *b, uint8_t *mask225);
int gzo_threat_level(const gzo_board *b, int player);
int64_t gzo_pattern_score(const gzo_board *b, int player);

/* ---- rollout ---- */
int gzo_select_offensive_move(const gzo_board *b, gzo_rng *rng, int step, uint8_t *cand225,
                              int *from_choice);
int gzo_select_offensive_move_literal(const gzo_board *b, gzo_rng *rng, int step,
                                      uint8_t *cand225, int *from_choice);
double gzo_simulate(const gzo_board *start, const gzo_params *p, const gzo_weights *w, gzo_rng *rng,
                    gzo_board *final_board, gzo_search_stats *stats);

/* ---- planner ---- */
void gzo_planner_scores(const gzo_board *b, int ai_player, double *scores225);
int gzo_planner_topk(const gzo_board *b, int ai_player, int k, int32_t *topk);
int gzo_planner_move(const gzo_board *b, const gzo_params *p, const gzo_weights *w, gzo_rng *rng,
                     int step, int32_t *topk_out, double *composed_out);
/* same, also recording the near-tie diagnostics into `stats` (may be NULL) */
int gzo_planner_move_stats(const gzo_board *b, const gzo_params *p, const gzo_weights *w, gzo_rng *rng,
                           int step, int32_t *topk_out, double *composed_out, gzo_search_stats *stats);

/* Optional memo of the planner networks' outputs keyed by the board (GraphNet softmax + DQN q are pure
 * functions of the position): n_entries > 0 allocates a direct-mapped table of that many boards, 0 frees it.
 * Only the fixture generator uses it (1e5 playouts per position share their first planner steps); results
 * are bit-identical with and without it.  Process-global, not thread-safe. */
void gzo_set_planner_cache(int n_entries);

/* ---- networks (plain fp32 loops) ---- */
int64_t gzo_pv_param_count(void);
int64_t gzo_graphnet_param_count(void);
int64_t gzo_dqn_param_count(void);
void gzo_pv_forward(const float *w, const uint8_t *cells, float *logits225, float *probs225,
                    float *value);
void gzo_graphnet_forward(const float *w, const uint8_t *cells, float *logits225);
void gzo_dqn_forward(const float *w, const uint8_t *cells, float *q225);

/* ---- search ---- */
int gzo_search(const gzo_board *root, const gzo_params *p, const gzo_weights *w, uint32_t seed_lo,
               uint32_t seed_hi, uint32_t game, int32_t *node_parent, int32_t *node_move,
               int32_t *node_visits, double *node_value, int32_t max_nodes, gzo_search_stats *stats);

/* Philox4x32-10, exposed so tests can pin it against known-answer vectors. */
void gzo_philox4x32_10(const uint32_t ctr[4], const uint32_t key[2], uint32_t out[4]);

#ifdef __cplusplus
}
#endif
#endif
