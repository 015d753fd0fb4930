// gz_kernels_tree.cu -- MCTS select / expand / backup over flat SoA node arrays, one warp per tree.
// Reference: ai_agent.py:168-249 (search loop), :441-454 (backup, best child), :488-562 (MCTSNode,
// ucb1).  One simulation is in flight per tree, so visit counts and the float64 value sums are
// accumulated in exactly the reference's order and need no atomics.
//
// Children of a node are found by scanning the tree's `parent` array (<= S+1 int16 entries, read
// coalesced by the warp): nodes are stored in creation order, which is also the insertion order of
// every children list, so "first maximal child" == smallest node index among the maxima.
#include <math.h>

#include "gz_engine.h"
#include "gz_warp.cuh"

namespace gz {

struct NodeView {
    int16_t *parent;
    uint8_t *move;
    uint8_t *status;
    int16_t *nchild;
    int16_t *nempty;
    int32_t *visits;
    double *value;
    double *bgb;
};

__device__ __forceinline__ NodeView node_view(const TreeArrays &t, int g) {
    size_t o = (size_t)g * t.cap;
    NodeView v = {t.parent + o, t.move + o, t.status + o, t.nchild + o, t.nempty + o,
                  t.visits + o, t.value + o, t.bgb + o};
    return v;
}

// Board of node `cur`: the root position plus the moves on the path, colours alternating from the
// root's mover (the reference copies the board into every node, ai_agent.py:242-243; replaying
// <= depth stones is cheaper than 80 B per node of HBM).
__device__ __forceinline__ gz_board node_board(const NodeView &v, const gz_board &root, int cur) {
    int depth = 0;
    for (int n = cur; n > 0; n = v.parent[n]) ++depth;
    BB st[2];
    st[0] = bb_load(root.stones[0]);
    st[1] = bb_load(root.stones[1]);
    int d = depth;
    for (int n = cur; n > 0; n = v.parent[n], --d) {
        int colour = (d & 1) ? root.to_move : 3 - root.to_move;   // d-th move from the root
        int p = v.move[n];
        if (colour == 1) bb_set(st[0], p); else bb_set(st[1], p);
    }
    gz_board b;
    bb_store(b.stones[0], st[0]);
    bb_store(b.stones[1], st[1]);
    b.move_count = (uint16_t)(root.move_count + depth);
    b.to_move = (uint8_t)((depth & 1) ? 3 - root.to_move : root.to_move);
    b.status = v.status[cur];
    b.reserved[0] = b.reserved[1] = b.reserved[2] = 0;
    return b;
}

__global__ void __launch_bounds__(128) k_tree_init(TreeArrays t, SimArrays sm,
                                                   const gz_board *__restrict__ boards,
                                                   const uint8_t *__restrict__ ai_player, int G,
                                                   unsigned long long *__restrict__ counters) {
    int g = blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= G) return;
    NodeView v = node_view(t, g);
    gz_board b = boards[g];
    BB e = bb_empty_cells(bb_load(b.stones[0]), bb_load(b.stones[1]));
    v.parent[0] = -1;
    v.move[0] = 0;
    v.status[0] = b.status;
    v.nchild[0] = 0;
    v.nempty[0] = (int16_t)bb_count(e);       // unexplored_moves = get_valid_moves() (:519)
    v.visits[0] = 0;
    v.value[0] = 0.0;
    v.bgb[0] = 0.0;
    t.n_nodes[g] = 1;
    t.root[g] = b;
    t.ai_player[g] = ai_player[g];
    sm.board[g] = b;
    sm.leaf[g] = 0;
    sm.need_eval[g] = b.status == 0;          // the root is evaluated at construction (:522-523)
    sm.steps[g] = 0;
    sm.value[g] = 0.0;
    if (b.status == 0) atomicAdd(&counters[2], 1ull);
}

// ai_agent.py:532-562; evaluation order (exploitation + exploration) + bg_bonus + 0.0, no FMA.
__device__ __forceinline__ double ucb1(double value, int visits, double num, double c_puct, double bgb) {
    if (visits == 0) return INFINITY;
    double exploitation = __ddiv_rn(value, (double)visits);
    double exploration = __dmul_rn(c_puct, __dsqrt_rn(__ddiv_rn(num, (double)visits)));
    return __dadd_rn(__dadd_rn(exploitation, exploration), bgb);
}

__global__ void k_set_sim(int32_t *d_sim, int value, int add) { *d_sim = add ? *d_sim + value : value; }

__global__ void __launch_bounds__(128) k_tree_step(TreeArrays t, SimArrays sm, int G, SearchCfg c,
                                                   int sim_arg, const double *__restrict__ log_table,
                                                   int n_log, const double *__restrict__ tanh_table,
                                                   int n_tanh,
                                                   unsigned long long *__restrict__ counters) {
    pdl_trigger();   // the input convolution behind this kernel may set itself up while this grid drains (gz_engine.h)
    int g = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    int lane = threadIdx.x & 31;
    if (g >= G) return;
    const int sim = c.d_sim ? *c.d_sim : sim_arg;
    NodeView v = node_view(t, g);
    // ---- backup of the previous simulation (:441-448), same sign on the whole path
    if (sim > 0) {
        if (lane == 0) {
            double val = sm.value[g];
            for (int n = sm.leaf[g]; n >= 0; n = v.parent[n]) {
                v.visits[n] += 1;
                v.value[n] = __dadd_rn(v.value[n], val);
            }
        }
        __syncwarp();
    }
    if (sim >= c.S) return;   // the extra launch after the last simulation only backs up
    // per-position budget (gz_board.sim_budget, 0 = c.S): a tree that has run its simulations idles for the
    // rest of the search -- finished dummy leaf, nothing to evaluate, nothing to back up
    const int bud = (int)t.root[g].reserved[0];
    if (bud != 0 && sim >= bud) {
        if (lane == 0) {
            sm.board[g].status = 3;
            sm.leaf[g] = -1;
            sm.steps[g] = 0;
            sm.need_eval[g] = 0;
        }
        return;
    }
    int nn = t.n_nodes[g];
    // ---- select (:224-232)
    int cur = 0;
    while (true) {
        if (v.status[cur] != 0) break;
        int nch = v.nchild[cur];
        if (nch == 0 || nch < v.nempty[cur]) break;
        int pvis = v.visits[cur];
        int la = pvis > 1 ? pvis : 1;
        double lg = (la < n_log) ? log_table[la] : log((double)la);
        double num = lg > 1.0 ? lg : 1.0;
        double best = 0.0;
        int bi = -1;
        for (int j = 1 + lane; j < nn; j += 32) {
            if (v.parent[j] != cur) continue;
            double s = ucb1(v.value[j], v.visits[j], num, c.c_puct, v.bgb[j]);
            if (bi < 0 || s > best) { best = s; bi = j; }
        }
        warp_argmax_first(best, bi);
        cur = bi;
    }
    // ---- expand (:215-216, :234-249)
    const gz_board root = t.root[g];
    gz_board b = node_board(v, root, cur);
    bool created = false;
    if (b.status == 0 && v.visits[cur] > 0 && v.nchild[cur] < v.nempty[cur]) {
        BB black = bb_load(b.stones[0]), white = bb_load(b.stones[1]);
        BB empty = bb_empty_cells(black, white);
        int k = v.nempty[cur] - v.nchild[cur] - 1;   // list.pop(): the LAST unexplored move
        int p = bb_nth(empty, k);
        board_place(b, p);
        int a = t.ai_player[g];
        BB a_st = bb_load(b.stones[a - 1]);
        BB e2 = bb_empty_cells(bb_load(b.stones[0]), bb_load(b.stones[1]));
        long long S = warp_pattern_score(a_st, e2, lane);      // _bg_score (:432-439)
        if (lane == 0) {
            long long q = S / 50;
            double th = (S % 50 == 0 && q < n_tanh) ? tanh_table[q] : tanh((double)S / 10000.0);
            int id = nn;
            v.parent[id] = (int16_t)cur;
            v.move[id] = (uint8_t)p;
            v.status[id] = b.status;
            v.nchild[id] = 0;
            v.nempty[id] = (int16_t)(v.nempty[cur] - 1);
            v.visits[id] = 0;
            v.value[id] = 0.0;
            v.bgb[id] = __dmul_rn(c.beta, th);
            v.nchild[cur] = (int16_t)(v.nchild[cur] + 1);
            t.n_nodes[g] = nn + 1;
        }
        cur = nn;
        created = true;
    }
    if (lane == 0) {
        sm.board[g] = b;
        sm.leaf[g] = cur;
        sm.steps[g] = 0;
        bool ev = created && b.status == 0;
        sm.need_eval[g] = ev ? 1 : 0;
        atomicAdd(&counters[0], 1ull);
        if (created) atomicAdd(&counters[1], 1ull);
        if (ev) atomicAdd(&counters[2], 1ull);
        if (b.status != 0) atomicAdd(&counters[5], 1ull);
    }
}

__global__ void __launch_bounds__(128) k_tree_finalize(TreeArrays t, int G,
                                                       int32_t *__restrict__ child_visits,
                                                       double *__restrict__ child_value,
                                                       int32_t *__restrict__ best_move) {
    int g = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    int lane = threadIdx.x & 31;
    if (g >= G) return;
    NodeView v = node_view(t, g);
    int nn = t.n_nodes[g];
    if (child_visits) for (int i = lane; i < kCells; i += 32) child_visits[(size_t)g * kCells + i] = 0;
    if (child_value) for (int i = lane; i < kCells; i += 32) child_value[(size_t)g * kCells + i] = 0.0;
    __syncwarp();
    double best = 0.0;
    int bi = -1;
    for (int j = 1 + lane; j < nn; j += 32) {
        if (v.parent[j] != 0) continue;
        int cell = bit_to_cell(v.move[j]);
        if (child_visits) child_visits[(size_t)g * kCells + cell] = v.visits[j];
        if (child_value) child_value[(size_t)g * kCells + cell] = v.value[j];
        double s = (double)v.visits[j];
        if (bi < 0 || s > best) { best = s; bi = j; }
    }
    warp_argmax_first(best, bi);      // max(children, key=visits): first maximum (:199-201)
    if (lane == 0 && best_move) best_move[g] = bi >= 0 ? bit_to_cell(v.move[bi]) : -1;
}

void launch_tree_init(const Engine &e, const gz_board *boards, const uint8_t *ai_player, int G,
                      cudaStream_t s) {
    k_tree_init<<<(G + 127) / 128, 128, 0, s>>>(e.tree, e.sim, boards, ai_player, G, e.counters);
}
void launch_tree_step(const Engine &e, int G, const SearchCfg &c, int sim, cudaStream_t s) {
    k_tree_step<<<(G * 32 + 127) / 128, 128, 0, s>>>(e.tree, e.sim, G, c, sim, e.log_table, e.n_log,
                                                     e.tanh_table, e.n_tanh, e.counters);
}
void launch_set_sim(int32_t *d_sim, int value, int add, cudaStream_t s) { k_set_sim<<<1, 1, 0, s>>>(d_sim, value, add); }
void launch_tree_finalize(const Engine &e, int G, const SearchCfg &c, int32_t *child_visits,
                          double *child_value, int32_t *best_move, cudaStream_t s) {
    (void)c;
    k_tree_finalize<<<(G * 32 + 127) / 128, 128, 0, s>>>(e.tree, G, child_visits, child_value, best_move);
}

}  // namespace gz
