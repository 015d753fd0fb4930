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
        if (lane == 0) {
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
                    ptx::umma_bf16(tmem_base + d_col, ptx::make_desc(a_addr + 2 * k * LBO_A, LBO_A, 128),
                                   ptx::make_desc(w_addr + 2 * k * LBO_B, LBO_B, 128), idesc,
                                   !(first && k == 0));
                ptx::umma_commit(&b_empty[st]);
                ++it;
            };
            for (int c = 0; c < kL1Chunks; ++c) {
                const int ab = c & 1;
                ptx::mbar_wait(&a_full[ab], (c >> 1) & 1);
                ptx::tc_fence_after();
                chunk_mmas(ptx::smem_u32(abuf + ab * kChunkBytesA), 0, c == 0);
                ptx::umma_commit(&a_empty[ab]);
            }
            ptx::umma_commit(&d_full[0]);
            ptx::mbar_wait(a2_full, 0);
            ptx::tc_fence_after();
            for (int c = 0; c < kL2Chunks; ++c) chunk_mmas(ptx::smem_u32(a2) + c * 8 * LBO_A, kFcN, c == 0);
            ptx::umma_commit(&d_full[1]);
            ptx::mbar_wait(a2_full, 1);
            ptx::tc_fence_after();
            for (int c = 0; c < kL3Chunks; ++c) chunk_mmas(ptx::smem_u32(a2) + c * 8 * LBO_A, 0, c == 0);
            ptx::umma_commit(&d_full[2]);
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

}  // namespace gz
