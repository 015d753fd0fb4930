// gz_warp.cuh -- warp-cooperative helpers (device only).
#pragma once
#include "gz_policy.cuh"

namespace gz {

__device__ __forceinline__ long long warp_sum_ll(long long v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ int warp_sum_i(int v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// KnowledgeSearch._pattern_score (bg_planner.py:133-155) of the colour `own`, spread over a warp:
// lane l looks at bit positions 32*i + l, i = 0..7 (one bit of every bitboard word).  Integer sum,
// so the reduction order does not matter.  All 32 lanes must call.
__device__ __forceinline__ long long warp_pattern_score(const BB &own, const BB &empty, int lane) {
    int acc = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        if ((own.w[i] >> lane) & 1u) {
            int p = 32 * i + lane;
            acc += pattern_score_stone(own, empty, p >> 4, p & 15);
        }
    }
    return warp_sum_ll((long long)acc);
}

// (score, index) arg-max over a warp: larger score wins, ties go to the SMALLER index -- Python's
// max() keeps the first maximal element (ai_agent.py:200,454).  idx < 0 marks "no candidate".
__device__ __forceinline__ void warp_argmax_first(double &score, int &idx) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        double os = __shfl_xor_sync(0xffffffffu, score, o);
        int oi = __shfl_xor_sync(0xffffffffu, idx, o);
        bool take = (oi >= 0) && (idx < 0 || os > score || (os == score && oi < idx));
        if (take) { score = os; idx = oi; }
    }
}
__device__ __forceinline__ void warp_argmax_first_ll(long long &score, int &idx) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        long long os = __shfl_xor_sync(0xffffffffu, score, o);
        int oi = __shfl_xor_sync(0xffffffffu, idx, o);
        bool take = (oi >= 0) && (idx < 0 || os > score || (os == score && oi < idx));
        if (take) { score = os; idx = oi; }
    }
}

}  // namespace gz
