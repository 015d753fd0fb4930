// gz_nn_fc.cu -- OpponentDQN (bg_planner.py:59-78: Linear 675->256, ReLU, 256->256, ReLU, 256->225)
// as three chained tcgen05 GEMMs, the throughput (bf16) mode of the planner's value net.
//
// Tile = 128 boards (UMMA M = 128), one CTA per tile; N = 256 (the last layer padded from 225).
//   layer 1  A = the 675 one-hot plane bits of each board (black | white | empty, raster order,
//            bg_planner.py:225-230) expanded to bf16 0/1 in shared memory chunk by chunk (K = 64 per
//            chunk, 11 chunks, K padded to 704) -- the 2700-byte float plane tensor never exists;
//   layer 2/3 A = the previous layer's activations: TMEM -> registers -> bias + ReLU -> bf16 ->
//            shared memory in the K-major core-matrix layout the next UMMA reads.
// Weights stream from L2 through a 3-stage ring of 32 KB K-chunks ([8 k-groups][256 n][8 k] bf16,
// packed on the host).  fp32 accumulation in tensor memory (2 x 256 columns).
//   warp 0      weight producer (bulk async copies)
//   warp 1      single-thread tcgen05.mma issuer, owns the TMEM allocation
//   warps 2-5   layer-1 operand generator, then the three epilogues (one board row per thread)
#include <cstdio>
#include "gz_engine.h"
#include "gz_bitboard.cuh"
#include "gz_tcgen05.cuh"

namespace gz {

namespace {
constexpr int kFcN = 256;                       // padded layer width
constexpr int kChunkBytesA = 8 * 128 * 16;      // one K-chunk of A: [8 k-groups][128 rows][8] bf16
constexpr int kChunkBytesW = 8 * kFcN * 16;     // one K-chunk of W: [8 k-groups][256 n][8] bf16
constexpr int kL1Chunks = 11, kL2Chunks = 4, kL3Chunks = 4;
constexpr int kVecBytes = 88;                   // 704 plane bits per board
constexpr int kWStages = 3;
constexpr int kA2Bytes = 32 * 128 * 16;         // 128 rows x 256 k bf16
constexpr int kSmemDqn = 128 * kVecBytes + 2 * kChunkBytesA + kA2Bytes + kWStages * kChunkBytesW +
                         3 * kFcN * 4 + 256;
}  // namespace

// 2^x by one MUFU.EX2 (ex2.approx.ftz: ~2 ulp, results below 2^-126 flush to zero).  exp2f() wraps the same
// instruction in a per-element branch for denormal results -- most of the softmax passes' time.
__device__ __forceinline__ float fast_ex2(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}

namespace {
constexpr int kStageRow = 36;                       // floats per staged row: 32 values + 4 pad (16-byte aligned rows)
constexpr int kStageBytes = 32 * kStageRow * 4;     // per epilogue warp: 32 rows x 32 columns
}  // namespace

// Epilogue stores: a thread owns one board row, so storing its values directly makes every warp store touch
// 32 different rows (16 bytes each).  Instead the warp parks a 32 x 32 block in its staging tile and writes it
// back row by row: 128 contiguous bytes per row, 4 rows per instruction (vector form, 16-byte aligned rows of
// pitch kPolicyPitch) or one row per instruction (scalar form, the [G][225] API outputs).
__device__ __forceinline__ void stage_block(float *stage, int lane, const float *vals) {
    __syncwarp();
#pragma unroll
    for (int j = 0; j < 32; j += 4)
        *reinterpret_cast<float4 *>(stage + lane * kStageRow + j) = make_float4(vals[j], vals[j + 1], vals[j + 2], vals[j + 3]);
    __syncwarp();
}
__device__ __forceinline__ void store_rows_vec(float *stage, int lane, const float *vals, float *row_ptr, int c0) {
    stage_block(stage, lane, vals);
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const int r = 4 * i + (lane >> 3), c4 = (lane & 7) * 4;
        float *dst = reinterpret_cast<float *>(__shfl_sync(0xffffffffu, reinterpret_cast<unsigned long long>(row_ptr), r));
        const float4 x = *reinterpret_cast<const float4 *>(stage + r * kStageRow + c4);
        if (dst) *reinterpret_cast<float4 *>(dst + c0 + c4) = x;
    }
}
__device__ __forceinline__ void store_rows_scalar(float *stage, int lane, const float *vals, float *row_ptr, int c0, int ncols) {
    stage_block(stage, lane, vals);
#pragma unroll 4
    for (int r = 0; r < 32; ++r) {
        float *dst = reinterpret_cast<float *>(__shfl_sync(0xffffffffu, reinterpret_cast<unsigned long long>(row_ptr), r));
        if (dst && c0 + lane < ncols) dst[c0 + lane] = stage[r * kStageRow + lane];
    }
}
// same, for rows that are `pitch` floats apart from `warp_base` (row r of the warp is live iff bit r of row_mask):
// no pointer shuffles, the loop is a shared-memory load and a store per row
__device__ __forceinline__ void store_rows_scalar_affine(float *stage, int lane, const float *vals, float *warp_base,
                                                         int pitch, uint32_t row_mask, int c0, int ncols) {
    stage_block(stage, lane, vals);
    if (c0 + lane < ncols) {
#pragma unroll 8
        for (int r = 0; r < 32; ++r)
            if ((row_mask >> r) & 1u) warp_base[(size_t)r * pitch + c0 + lane] = stage[r * kStageRow + lane];
    }
}

// body of k_dqn_umma for one tile of 128 boards (also one half of k_gnhead_dqn)
__device__ __forceinline__ void dqn_body(const int tile, const gz_board *__restrict__ boards, int G,
                                         const uint8_t *__restrict__ w_packed,
                                         const float *__restrict__ bias /*[3][256]*/, float *__restrict__ q) {
    extern __shared__ __align__(128) uint8_t smem[];
    uint8_t *abuf = smem;                                        // [2][kChunkBytesA]
    uint8_t *a2 = abuf + 2 * kChunkBytesA;                       // [32 k-groups][128][8]
    uint8_t *wbuf = a2 + kA2Bytes;                               // [kWStages][kChunkBytesW]
    uint32_t *vec = reinterpret_cast<uint32_t *>(wbuf + kWStages * kChunkBytesW);   // [128][22]
    float *sbias = reinterpret_cast<float *>(reinterpret_cast<uint8_t *>(vec) + 128 * kVecBytes);
    uint64_t *bars = reinterpret_cast<uint64_t *>(sbias + 3 * kFcN);
    uint64_t *a_full = bars, *a_empty = bars + 2, *b_full = bars + 4, *b_empty = bars + 7;
    uint64_t *d_full = bars + 10, *a2_full = bars + 13;
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(bars + 14);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 0) {
        for (int i = 0; i < 2; ++i) { ptx::mbar_init(&a_full[i], 128); ptx::mbar_init(&a_empty[i], 1); }
        for (int i = 0; i < kWStages; ++i) { ptx::mbar_init(&b_full[i], 1); ptx::mbar_init(&b_empty[i], 1); }
        for (int i = 0; i < 3; ++i) ptx::mbar_init(&d_full[i], 1);
        ptx::mbar_init(a2_full, 128);
        ptx::fence_mbar_init();
    }
    for (int i = threadIdx.x; i < 3 * kFcN; i += blockDim.x) sbias[i] = bias[i];
    if (warp == 1) ptx::tmem_alloc<512>(tmem_slot);
    ptx::tc_fence_before();
    __syncthreads();
    ptx::tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    constexpr int kChunks = kL1Chunks + kL2Chunks + kL3Chunks;

    if (warp == 0) {
        if (lane == 0) {
            for (int i = 0; i < kChunks; ++i) {
                const int st = i % kWStages;
                ptx::mbar_wait(&b_empty[st], ((i / kWStages) & 1) ^ 1);
                ptx::mbar_expect_tx(&b_full[st], kChunkBytesW);
                const uint8_t *src = w_packed + (size_t)i * kChunkBytesW;
                ptx::bulk_g2s(wbuf + st * kChunkBytesW, src, kChunkBytesW / 2, &b_full[st]);
                ptx::bulk_g2s(wbuf + st * kChunkBytesW + kChunkBytesW / 2, src + kChunkBytesW / 2,
                              kChunkBytesW / 2, &b_full[st]);
            }
        }
    } else if (warp == 1) {
        {   // whole warp; one elected lane issues (gz_tcgen05.cuh)
            constexpr uint32_t idesc = ptx::make_idesc_bf16(128, kFcN);
            constexpr uint32_t LBO_A = 128 * 16, LBO_B = kFcN * 16;
            int it = 0;
            auto chunk_mmas = [&](uint32_t a_addr, uint32_t d_col, bool first) {
                const int st = it % kWStages;
                ptx::mbar_wait(&b_full[st], (it / kWStages) & 1);
                ptx::tc_fence_after();
                const uint32_t w_addr = ptx::smem_u32(wbuf + st * kChunkBytesW);
#pragma unroll
                for (int k = 0; k < 4; ++k)
                    ptx::umma_bf16_elect(tmem_base + d_col, ptx::make_desc(a_addr + 2 * k * LBO_A, LBO_A, 128),
                                   ptx::make_desc(w_addr + 2 * k * LBO_B, LBO_B, 128), idesc,
                                   !(first && k == 0));
                ptx::umma_commit_elect(&b_empty[st]);
                ++it;
            };
            for (int c = 0; c < kL1Chunks; ++c) {
                const int ab = c & 1;
                ptx::mbar_wait(&a_full[ab], (c >> 1) & 1);
                ptx::tc_fence_after();
                chunk_mmas(ptx::smem_u32(abuf + ab * kChunkBytesA), 0, c == 0);
                ptx::umma_commit_elect(&a_empty[ab]);
            }
            ptx::umma_commit_elect(&d_full[0]);
            ptx::mbar_wait(a2_full, 0);
            ptx::tc_fence_after();
            for (int c = 0; c < kL2Chunks; ++c) chunk_mmas(ptx::smem_u32(a2) + c * 8 * LBO_A, kFcN, c == 0);
            ptx::umma_commit_elect(&d_full[1]);
            ptx::mbar_wait(a2_full, 1);
            ptx::tc_fence_after();
            for (int c = 0; c < kL3Chunks; ++c) chunk_mmas(ptx::smem_u32(a2) + c * 8 * LBO_A, 0, c == 0);
            ptx::umma_commit_elect(&d_full[2]);
        }
    } else {
        const int qd = warp & 3;                 // TMEM lane quadrant of this warp
        const int m = 32 * qd + lane;            // board row of this thread
        const int b = tile * 128 + m;
#ifdef GZ_HEAD_TIMING
        long long tq[8]; tq[0] = clock64();
#endif
        // ---- 704-bit plane vector of the board: black | white | empty in raster order
        {
            uint32_t v[22];
#pragma unroll
            for (int i = 0; i < 22; ++i) v[i] = 0u;
            if (b < G) {
                BB bl = bb_load(boards[b].stones[0]), wh = bb_load(boards[b].stones[1]);
                BB pl[3] = {bl, wh, bb_empty_cells(bl, wh)};
#pragma unroll
                for (int p = 0; p < 3; ++p) {
#pragma unroll
                    for (int r = 0; r < 15; ++r) {
                        const uint32_t row = (pl[p].w[r >> 1] >> ((r & 1) * 16)) & 0x7FFFu;
                        const int pos = 225 * p + 15 * r, wi = pos >> 5, sh = pos & 31;
                        v[wi] |= row << sh;
                        if (sh > 17) v[wi + 1] |= row >> (32 - sh);
                    }
                }
            }
#pragma unroll
            for (int i = 0; i < 22; ++i) vec[m * 22 + i] = v[i];
        }
#ifdef GZ_HEAD_TIMING
        tq[1] = clock64();
#endif
        // ---- layer-1 operand chunks: byte kg of the vector -> 8 bf16 zeros/ones
        for (int c = 0; c < kL1Chunks; ++c) {
            const int ab = c & 1;
            ptx::mbar_wait(&a_empty[ab], ((c >> 1) & 1) ^ 1);
            const uint32_t w0 = vec[m * 22 + 2 * c], w1 = vec[m * 22 + 2 * c + 1];
            uint8_t *dst = abuf + ab * kChunkBytesA + m * 16;
#pragma unroll
            for (int kg = 0; kg < 8; ++kg) {
                const uint32_t by = ((kg < 4 ? w0 : w1) >> (8 * (kg & 3))) & 0xFFu;
                uint32_t u[4];
#pragma unroll
                for (int j = 0; j < 4; ++j)
                    u[j] = ((by >> (2 * j)) & 1u) * 0x3F80u + ((by >> (2 * j + 1)) & 1u) * 0x3F800000u;
                *reinterpret_cast<uint4 *>(dst + kg * 128 * 16) = make_uint4(u[0], u[1], u[2], u[3]);
            }
            ptx::fence_proxy_async();
            ptx::mbar_arrive(&a_full[ab]);
        }
        // ---- epilogues
#ifdef GZ_HEAD_TIMING
        tq[2] = clock64();
#endif
        const uint32_t trow = tmem_base + ((uint32_t)(32 * qd) << 16);
        for (int layer = 0; layer < 3; ++layer) {
            ptx::mbar_wait(&d_full[layer], 0);
#ifdef GZ_HEAD_TIMING
            tq[3 + layer] = clock64();
#endif
            ptx::tc_fence_after();
            const uint32_t tcol = trow + (layer == 1 ? kFcN : 0);
            const float *bs = sbias + layer * kFcN;
            if (layer == 2) {   // Q values: rows written coalesced through a staging tile (the A buffers are free now)
                float *stage = reinterpret_cast<float *>(abuf) + (warp - 2) * 32 * kStageRow;
                float *q_warp = q + (size_t)(tile * 128 + 32 * qd) * kCells;
                const uint32_t rows = __ballot_sync(0xffffffffu, b < G);
#pragma unroll 1
                for (int c0 = 0; c0 < kFcN; c0 += 32) {
                    float v[32];
                    ptx::tmem_ld16(tcol + c0, v);
                    ptx::tmem_ld16(tcol + c0 + 16, v + 16);
#pragma unroll
                    for (int i = 0; i < 32; ++i) v[i] += bs[c0 + i];
                    store_rows_scalar_affine(stage, lane, v, q_warp, kCells, rows, c0, kCells);
                }
                break;
            }
#pragma unroll 1
            for (int c64 = 0; c64 < kFcN; c64 += 64) {      // hidden layers: bias + ReLU -> bf16 A operand of the next GEMM
                float v[64];
                ptx::tmem_ld64(tcol + c64, v);
#pragma unroll
                for (int c0 = 0; c0 < 64; c0 += 16) {
                    uint32_t pk[8];
#pragma unroll
                    for (int i = 0; i < 8; ++i)
                        pk[i] = ptx::pack_bf16(fmaxf(v[c0 + 2 * i] + bs[c64 + c0 + 2 * i], 0.f),
                                               fmaxf(v[c0 + 2 * i + 1] + bs[c64 + c0 + 2 * i + 1], 0.f));
                    uint8_t *dst = a2 + ((c64 + c0) / 8) * 128 * 16 + m * 16;
                    *reinterpret_cast<uint4 *>(dst) = make_uint4(pk[0], pk[1], pk[2], pk[3]);
                    *reinterpret_cast<uint4 *>(dst + 128 * 16) = make_uint4(pk[4], pk[5], pk[6], pk[7]);
                }
            }
            if (layer < 2) {
                ptx::tc_fence_before();
                ptx::fence_proxy_async();
                ptx::mbar_arrive(a2_full);
            }
        }
#ifdef GZ_HEAD_TIMING
        tq[6] = clock64();
        if (tile == 0 && threadIdx.x == 64)
            printf("dqn: vector %lld, L1 operands %lld, wait L1 %lld, epi1+wait L2 %lld, epi2+wait L3 %lld, epi3 %lld cycles\n",
                   tq[1] - tq[0], tq[2] - tq[1], tq[3] - tq[2], tq[4] - tq[3], tq[5] - tq[4], tq[6] - tq[5]);
#endif
    }
    ptx::tc_fence_before();
    __syncthreads();
    if (warp == 1) ptx::tmem_dealloc<512>(tmem_base);
}

__global__ void __launch_bounds__(192, 1)
k_dqn_umma(const gz_board *__restrict__ boards, int G, const uint8_t *__restrict__ w_packed,
           const float *__restrict__ bias, float *__restrict__ q) {
    dqn_body(blockIdx.x, boards, G, w_packed, bias, q);
}

void nn_dqn_bf16(Engine &e, const gz_board *boards, int G, float *dqn_q, cudaStream_t s) {
    static bool configured = false;
    if (!configured) {
        cudaFuncSetAttribute(k_dqn_umma, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemDqn);
        configured = true;
    }
    prof_begin(e, GZ_STAGE_DQN, 1, s);
    k_dqn_umma<<<(G + 127) / 128, 192, kSmemDqn, s>>>(boards, G, e.bw.dqn_w, e.bw.dqn_bias, dqn_q);
    prof_end(e, s);
}


// ================================================================================================
// Heads after the fused 1x1 head convolutions (gz_nn_bf16.cu epilogues), tile = 128 boards per CTA:
//   policy   Linear 450 -> 225 (+ softmax over all 225 cells; neural_network.py:149-151,248-252 and
//            bg_planner.py:43-56,247) as one tcgen05 GEMM, K = 450 padded to 512, N = 256;
//   value    (PV) Linear 225 -> 64, ReLU, Linear 64 -> 1, tanh (neural_network.py:154-157): the
//            first layer as a second GEMM (K 256, N 64), the 64 -> 1 dot product in the epilogue;
//   prior    (PV) probabilities gathered over the legal cells and renormalised in float64
//            (ai_agent.py:564-582).
// The A operand (fp32 head-conv outputs) is split into bf16 hi + lo parts, both multiplied by the
// same bf16 weights, so the activations keep fp32-like accuracy.  Thread = board row in the epilogue.
namespace {
constexpr int kHdChunksP = kHeadChunksP, kHdChunksV = kHeadChunksV;
constexpr int kHdWBytesP = 8 * 256 * 16, kHdWBytesV = 8 * 64 * 16;
constexpr int kSmemHead = 2 * 2 * kChunkBytesA + kWStages * kHdWBytesP + 4 * kStageBytes + (256 + 64 + 64 + 4) * 4 + 256;
static_assert(kChunkBytesA == kHeadBlockBytes, "head operand tiles are copied chunk by chunk into the A buffers");
}  // namespace

struct HeadArgs {
    const uint8_t *pc;        // policy head conv outputs, bf16 hi/lo operand tiles (gz_engine.h)
    const uint8_t *vc;        // value head conv outputs (PV), likewise
    const gz_board *boards;   // legal mask for the prior (PV)
    const uint8_t *active;    // optional per-board mask
    const uint8_t *w_pol;     // 8 chunks [8 kg][256][8] bf16
    const uint8_t *w_val;     // 4 chunks [8 kg][64][8] bf16 (PV)
    const float *b_pol;       // [256] (225 real)
    const float *vfc1_b, *vfc2_w, *vfc2_b;
    float *policy;            // [G][kPolicyPitch] softmax (internal hot buffer) or nullptr
    float *logits, *probs;    // [G][225] API outputs or nullptr
    float *value;             // [G] or nullptr
    float *prior;             // [node][kPolicyPitch] or nullptr
    const int32_t *prior_slot;
    int prior_cap;
    float *vhead;
    int G;
};

template <bool PV>
__device__ __forceinline__ void head_body(const int tile, const HeadArgs &a) {
    extern __shared__ __align__(128) uint8_t smem[];
    uint8_t *abuf = smem;                                   // [2 buffers][hi, lo][kChunkBytesA]
    uint8_t *wbuf = abuf + 4 * kChunkBytesA;                // [kWStages][kHdWBytesP]
    float *stage_all = reinterpret_cast<float *>(wbuf + kWStages * kHdWBytesP);   // [4 warps][32][kStageRow]
    float *s_bp = stage_all + 4 * 32 * kStageRow;
    float *s_b1 = s_bp + 256, *s_w2 = s_b1 + 64, *s_b2 = s_w2 + 64;
    uint64_t *bars = reinterpret_cast<uint64_t *>(s_b2 + 4);
    uint64_t *a_full = bars, *a_empty = bars + 2, *b_full = bars + 4, *b_empty = bars + 7, *d_full = bars + 10;
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(bars + 12);
    constexpr int kChunks = kHdChunksP + (PV ? kHdChunksV : 0);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 0) {
        for (int i = 0; i < 2; ++i) { ptx::mbar_init(&a_full[i], 1); ptx::mbar_init(&a_empty[i], 1); }
        for (int i = 0; i < kWStages; ++i) { ptx::mbar_init(&b_full[i], 1); ptx::mbar_init(&b_empty[i], 1); }
        ptx::mbar_init(&d_full[0], 1); ptx::mbar_init(&d_full[1], 1);
        ptx::fence_mbar_init();
    }
    for (int i = threadIdx.x; i < 256; i += blockDim.x) s_bp[i] = a.b_pol[i];
    if (PV) {
        if (threadIdx.x < 64) { s_b1[threadIdx.x] = a.vfc1_b[threadIdx.x]; s_w2[threadIdx.x] = a.vfc2_w[threadIdx.x]; }
        if (threadIdx.x == 64) s_b2[0] = a.vfc2_b[0];
    }
    if (warp == 1) ptx::tmem_alloc<512>(tmem_slot);
    ptx::tc_fence_before();
    __syncthreads();
    ptx::tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        if (lane == 0) {
            for (int i = 0; i < kChunks; ++i) {
                const int st = i % kWStages;
                const bool pol = i < kHdChunksP;
                const uint32_t bytes = pol ? kHdWBytesP : kHdWBytesV;
                const uint8_t *src = pol ? a.w_pol + (size_t)i * kHdWBytesP : a.w_val + (size_t)(i - kHdChunksP) * kHdWBytesV;
                ptx::mbar_wait(&b_empty[st], ((i / kWStages) & 1) ^ 1);
                ptx::mbar_expect_tx(&b_full[st], bytes);
                ptx::bulk_g2s(wbuf + st * kHdWBytesP, src, bytes / 2, &b_full[st]);
                ptx::bulk_g2s(wbuf + st * kHdWBytesP + bytes / 2, src + bytes / 2, bytes / 2, &b_full[st]);
                // activations of this chunk: hi block then lo block, written in place by the conv epilogues
                const int ab = i & 1;
                const uint8_t *asrc = pol ? a.pc + (size_t)tile * kHeadTileBytesP + (size_t)i * 2 * kChunkBytesA
                                          : a.vc + (size_t)tile * kHeadTileBytesV + (size_t)(i - kHdChunksP) * 2 * kChunkBytesA;
                ptx::mbar_wait(&a_empty[ab], ((i >> 1) & 1) ^ 1);
                ptx::mbar_expect_tx(&a_full[ab], 2 * kChunkBytesA);
                ptx::bulk_g2s(abuf + 2 * ab * kChunkBytesA, asrc, kChunkBytesA, &a_full[ab]);
                ptx::bulk_g2s(abuf + (2 * ab + 1) * kChunkBytesA, asrc + kChunkBytesA, kChunkBytesA, &a_full[ab]);
            }
        }
    } else if (warp == 1) {
        {   // whole warp; one elected lane issues (gz_tcgen05.cuh)
            constexpr uint32_t idesc_p = ptx::make_idesc_bf16(128, 256), idesc_v = ptx::make_idesc_bf16(128, 64);
            constexpr uint32_t LBO_A = 128 * 16;
            for (int i = 0; i < kChunks; ++i) {
                const int st = i % kWStages, ab = i & 1;
                const bool pol = i < kHdChunksP;
                const uint32_t LBO_B = pol ? 256 * 16 : 64 * 16;
                ptx::mbar_wait(&a_full[ab], (i >> 1) & 1);
                ptx::mbar_wait(&b_full[st], (i / kWStages) & 1);
                ptx::tc_fence_after();
                const uint32_t w_addr = ptx::smem_u32(wbuf + st * kHdWBytesP);
                const uint32_t d = tmem_base + (pol ? 0 : 256);
                const bool first = pol ? i == 0 : i == kHdChunksP;
#pragma unroll
                for (int part = 0; part < 2; ++part) {
                    const uint32_t a_addr = ptx::smem_u32(abuf + (2 * ab + part) * kChunkBytesA);
#pragma unroll
                    for (int k = 0; k < 4; ++k)
                        ptx::umma_bf16_elect(d, ptx::make_desc(a_addr + 2 * k * LBO_A, LBO_A, 128),
                                       ptx::make_desc(w_addr + 2 * k * LBO_B, LBO_B, 128),
                                       pol ? idesc_p : idesc_v, !(first && part == 0 && k == 0));
                }
                ptx::umma_commit_elect(&a_empty[ab]);
                ptx::umma_commit_elect(&b_empty[st]);
                if (i == kHdChunksP - 1) ptx::umma_commit_elect(&d_full[0]);
                if (PV && i == kChunks - 1) ptx::umma_commit_elect(&d_full[1]);
            }
        }
    } else {
        const int qd = warp & 3;
        const int m = 32 * qd + lane;
        const int b = tile * 128 + m;
        const bool live = b < a.G && (!a.active || a.active[b]);
        const int bb = b < a.G ? b : a.G - 1;              // rows past the end read a valid board, write nothing
        const int wq = warp - 2;
#ifdef GZ_HEAD_TIMING
        long long tt[8]; tt[0] = clock64();
#endif
#ifdef GZ_HEAD_TIMING
        tt[1] = clock64();
#endif
        const uint32_t trow = tmem_base + ((uint32_t)(32 * qd) << 16);
        // ---- policy: softmax over the 225 logits of this row (two passes over tensor memory: an online
        // max / sum, then the normalised values; exp through ex2.approx like __expf)
        ptx::mbar_wait(&d_full[0], 0);
        ptx::tc_fence_after();
#ifdef GZ_HEAD_TIMING
        tt[2] = clock64();
#endif
        constexpr float kLog2e = 1.4426950408889634f;
        float mx = -INFINITY, sum = 0.f;
#pragma unroll 1
        for (int c64 = 0; c64 < 256; c64 += 64) {
            float v[64];
            ptx::tmem_ld64(trow + c64, v);
            float cm = -INFINITY;
#pragma unroll
            for (int j = 0; j < 64; ++j) {
                v[j] = c64 + j < kCells ? v[j] + s_bp[c64 + j] : -INFINITY;
                cm = fmaxf(cm, v[j]);
            }
            const float nm = fmaxf(mx, cm);
            float cs = 0.f;
#pragma unroll
            for (int j = 0; j < 64; ++j) cs += fast_ex2((v[j] - nm) * kLog2e);
            sum = sum * fast_ex2((mx - nm) * kLog2e) + cs;
            mx = nm;
        }
#ifdef GZ_HEAD_TIMING
        tt[3] = clock64();
#endif
        const float inv = 1.f / sum;
        uint32_t legal[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        if (PV && a.prior) {
            BB e = bb_empty_cells(bb_load(a.boards[bb].stones[0]), bb_load(a.boards[bb].stones[1]));
            bb_to_raster(e, legal);
        }
        float legal_sum = 0.f;
        float *stage = stage_all + wq * 32 * kStageRow;
        float *pol_row = a.policy && live ? a.policy + (size_t)b * kPolicyPitch : nullptr;
        const uint32_t live_rows = __ballot_sync(0xffffffffu, live);
#pragma unroll 1
        for (int c0 = 0; c0 < 256; c0 += 32) {
            float v[32], pr[32];
            ptx::tmem_ld16(trow + c0, v);
            ptx::tmem_ld16(trow + c0 + 16, v + 16);
            const uint32_t lbits = legal[c0 >> 5];
#pragma unroll
            for (int j = 0; j < 32; ++j) {
                v[j] += s_bp[c0 + j];
                pr[j] = c0 + j < kCells ? fast_ex2((v[j] - mx) * kLog2e) * inv : 0.f;
                if (PV && ((lbits >> j) & 1u)) legal_sum += pr[j];
            }
            if (a.logits) store_rows_scalar_affine(stage, lane, v, a.logits + (size_t)(tile * 128 + 32 * qd) * kCells, kCells, live_rows, c0, kCells);
            if (a.probs) store_rows_scalar_affine(stage, lane, pr, a.probs + (size_t)(tile * 128 + 32 * qd) * kCells, kCells, live_rows, c0, kCells);
            if (a.policy) store_rows_vec(stage, lane, pr, pol_row, c0);
        }
#ifdef GZ_HEAD_TIMING
        tt[4] = clock64();
#endif
        if (PV && a.prior) {   // warp-uniform: tcgen05.ld is .sync.aligned, only the stores are per-row
            float *dst = live ? a.prior + ((size_t)b * a.prior_cap + a.prior_slot[b]) * kPolicyPitch : nullptr;
            const float scale = legal_sum > 0.f ? inv / legal_sum : inv;   // prior = p / sum of legal p (ai_agent.py:573-582)
#pragma unroll 1
            for (int c0 = 0; c0 < 256; c0 += 32) {
                float v[32];
                ptx::tmem_ld16(trow + c0, v);
                ptx::tmem_ld16(trow + c0 + 16, v + 16);
                const uint32_t lbits = legal[c0 >> 5];
#pragma unroll
                for (int j = 0; j < 32; ++j)
                    v[j] = ((lbits >> j) & 1u) ? fast_ex2((v[j] + s_bp[c0 + j] - mx) * kLog2e) * scale : 0.f;
                store_rows_vec(stage, lane, v, dst, c0);
            }
        }
#ifdef GZ_HEAD_TIMING
        tt[5] = clock64();
#endif
        if (PV) {
            ptx::mbar_wait(&d_full[1], 0);
            ptx::tc_fence_after();
            float sacc = 0.f;
#pragma unroll 1
            for (int c0 = 0; c0 < 64; c0 += 16) {
                float v[16];
                ptx::tmem_ld16(trow + 256 + c0, v);
#pragma unroll
                for (int j = 0; j < 16; ++j) sacc = fmaf(s_w2[c0 + j], fmaxf(v[j] + s_b1[c0 + j], 0.f), sacc);
            }
            const float vv = tanhf(sacc + s_b2[0]);
            if (live) {
                if (a.value) a.value[b] = vv;
                if (a.vhead && a.prior_slot) a.vhead[(size_t)b * a.prior_cap + a.prior_slot[b]] = vv;
            }
        }
#ifdef GZ_HEAD_TIMING
        tt[6] = clock64();
        if (tile == 0 && threadIdx.x == 64)
            printf("head PV=%d: operands %lld, wait mma %lld, softmax pass1 %lld, pass2+stores %lld, prior %lld, value %lld cycles\n",
                   (int)PV, tt[1] - tt[0], tt[2] - tt[1], tt[3] - tt[2], tt[4] - tt[3], tt[5] - tt[4], tt[6] - tt[5]);
#endif
    }
    ptx::tc_fence_before();
    __syncthreads();
    if (warp == 1) ptx::tmem_dealloc<512>(tmem_base);
}

template <bool PV>
__global__ void __launch_bounds__(192, 1) k_head_umma(const __grid_constant__ HeadArgs a) {
    head_body<PV>(blockIdx.x, a);
}

// GraphNet head and OpponentDQN of one planner step in ONE launch: both are latency-bound chains on 32 CTAs
// (4096 boards / 128), independent of each other, and k_planner_pick needs both.  CTAs [0, head_tiles) run the head,
// the rest the DQN, so the two chains overlap on different SMs: ~23 us instead of 2 x 23 us per planner step.
__global__ void __launch_bounds__(192, 1)
k_gnhead_dqn(const __grid_constant__ HeadArgs a, int head_tiles, const uint8_t *__restrict__ dqn_w,
             const float *__restrict__ dqn_bias, float *__restrict__ dqn_q) {
    if ((int)blockIdx.x < head_tiles) head_body<false>(blockIdx.x, a);
    else dqn_body((int)blockIdx.x - head_tiles, a.boards, a.G, dqn_w, dqn_bias, dqn_q);
}

void nn_heads_bf16(Engine &e, bool pv, const gz_board *boards, const uint8_t *active, int G, float *policy,
                   float *logits, float *probs, float *value, float *prior, const int32_t *prior_slot,
                   cudaStream_t s) {
    static bool configured = false;
    if (!configured) {
        cudaFuncSetAttribute(k_head_umma<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemHead);
        cudaFuncSetAttribute(k_head_umma<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemHead);
        configured = true;
    }
    HeadArgs a;
    a.pc = pv ? e.pv_pc : e.gn_pc; a.vc = pv ? e.pv_vc : nullptr; a.boards = boards; a.active = active;
    a.w_pol = pv ? e.bw.pv_pfc_w : e.bw.gn_hfc_w; a.w_val = pv ? e.bw.pv_vfc1_w : nullptr;
    a.b_pol = pv ? e.bw.pv_pfc_b : e.bw.gn_hfc_b;
    a.vfc1_b = e.pv.vfc1_b; a.vfc2_w = e.pv.vfc2_w; a.vfc2_b = e.pv.vfc2_b;
    a.policy = policy; a.logits = logits; a.probs = probs; a.value = value; a.prior = prior;
    a.prior_slot = prior_slot; a.prior_cap = e.tree.cap; a.vhead = prior_slot ? e.tree.vhead : nullptr; a.G = G;
    prof_begin(e, pv ? GZ_STAGE_PV_HEADS : GZ_STAGE_GN_HEAD, 1, s);
    if (pv) k_head_umma<true><<<(G + 127) / 128, 192, kSmemHead, s>>>(a);
    else k_head_umma<false><<<(G + 127) / 128, 192, kSmemHead, s>>>(a);
    prof_end(e, s);
}

// GraphNet head (softmax into `policy`, pitch kPolicyPitch) + OpponentDQN, one launch (k_gnhead_dqn)
void nn_gnhead_dqn_bf16(Engine &e, const gz_board *boards, int G, float *policy, float *dqn_q, cudaStream_t s) {
    static bool configured = false;
    constexpr int kSmemBoth = kSmemHead > kSmemDqn ? kSmemHead : kSmemDqn;
    if (!configured) {
        cudaFuncSetAttribute(k_gnhead_dqn, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBoth);
        configured = true;
    }
    HeadArgs a = {};
    a.pc = e.gn_pc; a.boards = boards; a.w_pol = e.bw.gn_hfc_w; a.b_pol = e.bw.gn_hfc_b;
    a.vfc1_b = e.pv.vfc1_b; a.vfc2_w = e.pv.vfc2_w; a.vfc2_b = e.pv.vfc2_b;
    a.policy = policy; a.prior_cap = e.tree.cap; a.G = G;
    const int tiles = (G + 127) / 128;
    prof_begin(e, GZ_STAGE_GN_HEAD, 1, s);
    k_gnhead_dqn<<<2 * tiles, 192, kSmemBoth, s>>>(a, tiles, e.bw.dqn_w, e.bw.dqn_bias, dqn_q);
    prof_end(e, s);
}

}  // namespace gz
