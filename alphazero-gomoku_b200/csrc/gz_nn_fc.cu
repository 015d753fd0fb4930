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

__global__ void __launch_bounds__(192, 1)
k_dqn_umma(const gz_board *__restrict__ boards, int G, const uint8_t *__restrict__ w_packed,
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
    const int tile = blockIdx.x;
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
        const uint32_t trow = tmem_base + ((uint32_t)(32 * qd) << 16);
        for (int layer = 0; layer < 3; ++layer) {
            ptx::mbar_wait(&d_full[layer], 0);
            ptx::tc_fence_after();
            const uint32_t tcol = trow + (layer == 1 ? kFcN : 0);
            const float *bs = sbias + layer * kFcN;
#pragma unroll 1
            for (int c0 = 0; c0 < kFcN; c0 += 16) {
                float v[16];
                ptx::tmem_ld16(tcol + c0, v);
                if (layer < 2) {
                    uint32_t pk[8];
#pragma unroll
                    for (int i = 0; i < 8; ++i)
                        pk[i] = ptx::pack_bf16(fmaxf(v[2 * i] + bs[c0 + 2 * i], 0.f),
                                               fmaxf(v[2 * i + 1] + bs[c0 + 2 * i + 1], 0.f));
                    uint8_t *dst = a2 + (c0 / 8) * 128 * 16 + m * 16;
                    *reinterpret_cast<uint4 *>(dst) = make_uint4(pk[0], pk[1], pk[2], pk[3]);
                    *reinterpret_cast<uint4 *>(dst + 128 * 16) = make_uint4(pk[4], pk[5], pk[6], pk[7]);
                } else if (b < G) {
#pragma unroll
                    for (int i = 0; i < 16; ++i)
                        if (c0 + i < kCells) q[(size_t)b * kCells + c0 + i] = v[i] + bs[c0 + i];
                }
            }
            if (layer < 2) {
                ptx::tc_fence_before();
                ptx::fence_proxy_async();
                ptx::mbar_arrive(a2_full);
            }
        }
    }
    ptx::tc_fence_before();
    __syncthreads();
    if (warp == 1) ptx::tmem_dealloc<512>(tmem_base);
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
constexpr int kHdChunksP = 8, kHdChunksV = 4;
constexpr int kHdWBytesP = 8 * 256 * 16, kHdWBytesV = 8 * 64 * 16;
constexpr int kSmemHead = 2 * 2 * kChunkBytesA + kWStages * kHdWBytesP + (256 + 64 + 64 + 4) * 4 + 256;
}  // namespace

struct HeadArgs {
    const float *pc;          // [G][450] policy head conv outputs
    const float *vc;          // [G][225] value head conv outputs (PV)
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

__device__ __forceinline__ void split_store_hi_lo(uint8_t *hi, uint8_t *lo, const float *x) {
    uint32_t h[4], l[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        h[j] = ptx::pack_bf16(x[2 * j], x[2 * j + 1]);                       // one cvt.rn.bf16x2.f32
        const float r0 = x[2 * j] - __uint_as_float(h[j] << 16);
        const float r1 = x[2 * j + 1] - __uint_as_float(h[j] & 0xFFFF0000u);
        l[j] = ptx::pack_bf16(r0, r1);
    }
    *reinterpret_cast<uint4 *>(hi) = make_uint4(h[0], h[1], h[2], h[3]);
    *reinterpret_cast<uint4 *>(lo) = make_uint4(l[0], l[1], l[2], l[3]);
}

template <bool PV>
__global__ void __launch_bounds__(192, 1) k_head_umma(HeadArgs a) {
    extern __shared__ __align__(128) uint8_t smem[];
    uint8_t *abuf = smem;                                   // [2 buffers][hi, lo][kChunkBytesA]
    uint8_t *wbuf = abuf + 4 * kChunkBytesA;                // [kWStages][kHdWBytesP]
    float *s_bp = reinterpret_cast<float *>(wbuf + kWStages * kHdWBytesP);
    float *s_b1 = s_bp + 256, *s_w2 = s_b1 + 64, *s_b2 = s_w2 + 64;
    uint64_t *bars = reinterpret_cast<uint64_t *>(s_b2 + 4);
    uint64_t *a_full = bars, *a_empty = bars + 2, *b_full = bars + 4, *b_empty = bars + 7, *d_full = bars + 10;
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(bars + 12);
    constexpr int kChunks = kHdChunksP + (PV ? kHdChunksV : 0);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int tile = blockIdx.x;
    if (threadIdx.x == 0) {
        for (int i = 0; i < 2; ++i) { ptx::mbar_init(&a_full[i], 128); ptx::mbar_init(&a_empty[i], 1); }
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
        // ---- operand chunks: 64 inputs of every board row, fp32 -> bf16 hi + lo.  Eight lanes share a row
        // (one k-group of 8 inputs each) so that a warp instruction touches 4 rows x 256 contiguous bytes.
        const int wq = warp - 2;                           // rows 32*wq .. 32*wq + 31 are generated by this warp
        for (int i = 0; i < kChunks; ++i) {
            const int ab = i & 1;
            const bool pol = i < kHdChunksP;
            const int kreal = pol ? 450 : kCells;
            const int k0 = 64 * (pol ? i : i - kHdChunksP) + 8 * (lane & 7);
            ptx::mbar_wait(&a_empty[ab], ((i >> 1) & 1) ^ 1);
#pragma unroll
            for (int t = 0; t < 8; ++t) {
                const int row = 32 * wq + 4 * t + (lane >> 3);
                int gb = tile * 128 + row;
                if (gb >= a.G) gb = a.G - 1;
                const float *src = (pol ? a.pc + (size_t)gb * 450 : a.vc + (size_t)gb * kCells) + k0;
                float x[8];
#pragma unroll
                for (int j = 0; j < 8; j += 2) {
                    if (pol) {                               // 450-float rows are 8-byte aligned
                        float2 v2 = k0 + j < kreal ? __ldg(reinterpret_cast<const float2 *>(src + j)) : make_float2(0.f, 0.f);
                        x[j] = v2.x; x[j + 1] = v2.y;
                    } else {                                 // 225-float rows are only 4-byte aligned
                        x[j] = k0 + j < kreal ? __ldg(src + j) : 0.f;
                        x[j + 1] = k0 + j + 1 < kreal ? __ldg(src + j + 1) : 0.f;
                    }
                }
                uint8_t *hi = abuf + (2 * ab) * kChunkBytesA + (lane & 7) * 128 * 16 + row * 16;
                split_store_hi_lo(hi, hi + kChunkBytesA, x);
            }
            ptx::fence_proxy_async();
            ptx::mbar_arrive(&a_full[ab]);
        }
        const uint32_t trow = tmem_base + ((uint32_t)(32 * qd) << 16);
        // ---- policy: softmax over the 225 logits of this row (two passes over tensor memory: an online
        // max / sum, then the normalised values; exp through ex2.approx like __expf)
        ptx::mbar_wait(&d_full[0], 0);
        ptx::tc_fence_after();
        constexpr float kLog2e = 1.4426950408889634f;
        float mx = -INFINITY, sum = 0.f;
#pragma unroll 1
        for (int c0 = 0; c0 < 240; c0 += 16) {
            float v[16];
            ptx::tmem_ld16(trow + c0, v);
            float cm = -INFINITY;
#pragma unroll
            for (int j = 0; j < 16; ++j) {
                v[j] = c0 + j < kCells ? v[j] + s_bp[c0 + j] : -INFINITY;
                cm = fmaxf(cm, v[j]);
            }
            const float nm = fmaxf(mx, cm);
            float cs = 0.f;
#pragma unroll
            for (int j = 0; j < 16; ++j) cs += exp2f((v[j] - nm) * kLog2e);
            sum = sum * exp2f((mx - nm) * kLog2e) + cs;
            mx = nm;
        }
        const float inv = 1.f / sum;
        uint32_t legal[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        if (PV && a.prior) {
            BB e = bb_empty_cells(bb_load(a.boards[bb].stones[0]), bb_load(a.boards[bb].stones[1]));
            bb_to_raster(e, legal);
        }
        float legal_sum = 0.f;
        float *pol_row = a.policy ? a.policy + (size_t)bb * kPolicyPitch : nullptr;
#pragma unroll 1
        for (int c0 = 0; c0 < 240; c0 += 16) {
            float v[16], pr[16];
            ptx::tmem_ld16(trow + c0, v);
            const uint32_t lbits = (legal[c0 >> 5] >> (c0 & 31)) & 0xFFFFu;
#pragma unroll
            for (int j = 0; j < 16; ++j) {
                const float lg = v[j] + s_bp[c0 + j];
                pr[j] = c0 + j < kCells ? exp2f((lg - mx) * kLog2e) * inv : 0.f;
                if (PV && ((lbits >> j) & 1u)) legal_sum += pr[j];
                if (live && a.logits && c0 + j < kCells) a.logits[(size_t)b * kCells + c0 + j] = lg;
                if (live && a.probs && c0 + j < kCells) a.probs[(size_t)b * kCells + c0 + j] = pr[j];
            }
            if (live && pol_row) {
#pragma unroll
                for (int j = 0; j < 16; j += 4)
                    *reinterpret_cast<float4 *>(pol_row + c0 + j) = make_float4(pr[j], pr[j + 1], pr[j + 2], pr[j + 3]);
            }
        }
        if (PV && a.prior) {   // warp-uniform: tcgen05.ld is .sync.aligned, only the stores are per-row
            float *dst = a.prior + ((size_t)bb * a.prior_cap + (live ? a.prior_slot[bb] : 0)) * kPolicyPitch;
            const float scale = legal_sum > 0.f ? inv / legal_sum : inv;   // prior = p / sum of legal p (ai_agent.py:573-582)
#pragma unroll 1
            for (int c0 = 0; c0 < 240; c0 += 16) {
                float v[16], o[16];
                ptx::tmem_ld16(trow + c0, v);
                const uint32_t lbits = (legal[c0 >> 5] >> (c0 & 31)) & 0xFFFFu;
#pragma unroll
                for (int j = 0; j < 16; ++j)
                    o[j] = ((lbits >> j) & 1u) ? exp2f((v[j] + s_bp[c0 + j] - mx) * kLog2e) * scale : 0.f;
                if (live) {
#pragma unroll
                    for (int j = 0; j < 16; j += 4)
                        *reinterpret_cast<float4 *>(dst + c0 + j) = make_float4(o[j], o[j + 1], o[j + 2], o[j + 3]);
                }
            }
        }
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
    }
    ptx::tc_fence_before();
    __syncthreads();
    if (warp == 1) ptx::tmem_dealloc<512>(tmem_base);
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

}  // namespace gz
