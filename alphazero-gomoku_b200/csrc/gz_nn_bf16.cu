// gz_nn_bf16.cu -- bf16 tensor-core (tcgen05 / TMEM) forward of the convolution stacks: the
// throughput mode of AlphaZeroGomokuNet (neural_network.py:132-159) and GraphNet
// (bg_planner.py:50-56).  fp32 accumulation in tensor memory, bf16 activations between layers.
//
// Activation layout in HBM ("planes"): [board][channel group of 8][288 positions][8 channels] bf16.
// A plane is the 15x15 board on an 18-row x 16-column pitch: plane row 0 and rows 16,17 and column
// 15 are zero padding, board cell (r,c) sits at position 16*(r+1)+c.  With that pitch a 3x3 tap
// (dy,dx) is the constant position offset 16*dy+dx, so the implicit-GEMM A operand of every tap is
// the SAME shared-memory slab read at a different row offset.  One channel group of one position
// is 16 bytes = one row of a UMMA core matrix, and 8 consecutive positions are 128 contiguous bytes
// = one K-major SWIZZLE_NONE core matrix at ANY row offset; no im2col, no tensor maps, no swizzle:
// the slab of a board is fetched with one 1-D bulk copy (UBLKCP) per channel group.
//
// Tile = one board = M 256 (two UMMA M=128 halves: plane positions [16,144) and [144,272)),
// N = all output channels, K = taps x input channels in steps of 64.  Persistent CTAs (one per SM):
//   warp 0    producer: activation slab (double buffered) + weight K-steps (ring, or resident when
//             the whole layer fits in shared memory)
//   warp 1    single-thread tcgen05.mma issuer; accumulators double buffered in TMEM
//   warps 2-5 epilogue: tcgen05.ld, bias (+ residual) + ReLU, zero the padding positions, pack to
//             bf16, 16-byte coalesced stores
#include <cstdio>
#include <cstdlib>
#include <type_traits>
#include "gz_engine.h"
#include "gz_tcgen05.cuh"

namespace gz {

template <int CIN, int COUT, int TAPS, bool RESIDENT>
struct ConvCfg {
    static constexpr int CG = CIN / 8;                       // channel groups of the input
    static constexpr int KSTEP = CIN < 64 ? CIN : 64;        // input channels per K-step
    static constexpr int KK = KSTEP / 16;                    // UMMAs (K = 16) per K-step and M half
    static constexpr int CHUNKS = CIN / KSTEP;               // K chunks per tap
    static constexpr int KSTEPS = TAPS * CHUNKS;
    static constexpr int SLAB_ROWS = TAPS == 9 ? 290 : 256;  // positions [-1,289) or [16,272)
    static constexpr int SLAB_BYTES = CG * SLAB_ROWS * 16;
    static constexpr int WSTEP_BYTES = (KSTEP / 8) * COUT * 16;   // one K-step of weights: [cg][COUT][8]
    // HALVES (the 128-channel streaming layers): the K steps run chunk-major -- nine taps on channels 0-63, then nine on
    // 64-127 -- so the input lives in THREE half slabs of 64 channels (a board's first half is free for the next board
    // after nine K steps) instead of two whole ones, and the 37 KB that frees pays for a weight ring of six 16 KB stages
    // instead of four.  The ring is what paces these layers: the issuer waited 3.5-4 k cycles per board for weights
    // (~220 per K step: a refill from L2 takes ~1800 cycles under load, three full stages cover 1536).
#ifdef GZ_CONV_NO_HALVES      // A/B build (scratch/ab_lib.py): two whole slabs, four weight stages
    static constexpr bool HALVES = false;
#else
    static constexpr bool HALVES = !RESIDENT && CHUNKS == 2 && TAPS == 9;
#endif
    static constexpr int HALF_BYTES = SLAB_BYTES / 2;
    static constexpr int A_BYTES = HALVES ? 3 * HALF_BYTES : 2 * SLAB_BYTES;
    static constexpr int NSTAGE = RESIDENT ? KSTEPS : HALVES ? 6 : 4;
    static constexpr int W_BYTES = NSTAGE * WSTEP_BYTES;
    static constexpr int TCOLS = 4 * COUT;                   // 2 buffers x 2 halves x COUT fp32 columns
    static constexpr int SMEM = A_BYTES + W_BYTES + COUT * 4 + 256 + 128 + (3 * COUT + 4) * 4;
};

// FROM_BOARD: the input is the board itself (CIN = 16: channel 0 black, 1 white, 2 empty, the rest
// zero); the producer warp writes the one-hot slab from the bitboards instead of copying planes.
// HEADS (the last tower convolution of AlphaZeroGomokuNet): the 1x1 head convolutions 128 -> 2 (policy)
// and 128 -> 1 (value), neither with BN nor activation (neural_network.py:149,154), are applied to the
// fp32 epilogue values; only their outputs are written (heads.pc: 450 values per board, K index channel-major
// like the flatten at :150, heads.vc: 225 values; both as bf16 hi/lo operand tiles of the head GEMMs) and the
// 128-channel planes of the last layer are not stored.
struct HeadConv {
    const float *pw, *pb;   // [COUT][2], [2]
    const float *vw, *vb;   // [COUT], [1]
    uint8_t *pc, *vc;       // bf16 hi/lo operand tiles of the head GEMMs (head_operand_offset, gz_engine.h)
};

// x as bf16 hi + lo parts, stored at the value's place in a head operand tile
__device__ __forceinline__ void store_head_operand(uint8_t *tile_base, int board, int k, int chunks, float x) {
    const uint32_t hl = ptx::pack_bf16(x, 0.f) & 0xFFFFu;
    const float r = x - __uint_as_float(hl << 16);
    uint8_t *dst = tile_base + head_operand_offset(board, k, chunks);
    *reinterpret_cast<uint16_t *>(dst) = (uint16_t)hl;
    *reinterpret_cast<uint16_t *>(dst + kHeadBlockBytes) = (uint16_t)(ptx::pack_bf16(r, 0.f) & 0xFFFFu);
}

// Epilogue warps: 4 (each does both halves of the board) when the 144 UMMAs of a 128 -> 128 layer hide the
// epilogue anyway; 8 (one half each) for the 3 -> C input convolutions, whose 18 UMMAs per board leave the
// kernel waiting for TMEM loads and stores (3.4 TB/s of output with 4 warps), and for the layers with a skip
// connection: with 4 warps their epilogue took 12.1 k cycles per board against 9.2 k of UMMAs, and the issuer sat
// 2.9 k cycles per board waiting for a free accumulator (scratch/conv_timing.py).
template <bool FROM_BOARD, bool RESIDUAL> __host__ __device__ constexpr int conv_epi_warps() { return FROM_BOARD || RESIDUAL ? 8 : 4; }

// -DGZ_CONV_TIMING: block 0 of the 128 -> 128 layers prints where its issuer and epilogue warps wait (scratch/conv_timing.py)
#ifdef GZ_CONV_TIMING
#define CV_T(...) __VA_ARGS__
#else
#define CV_T(...)
#endif
template <int CIN, int COUT, int TAPS, bool RESIDENT, bool RESIDUAL, bool FROM_BOARD = false, bool HEADS = false>
__global__ void __launch_bounds__(64 + 32 * conv_epi_warps<FROM_BOARD, RESIDUAL>(), 1)
k_conv_umma(const uint8_t *__restrict__ in_planes, const uint8_t *__restrict__ w_packed,
            const float *__restrict__ bias, const uint8_t *__restrict__ res_planes,
            uint8_t *__restrict__ out_planes, const uint8_t *__restrict__ active, int G, HeadConv heads, int rev) {
    using C = ConvCfg<CIN, COUT, TAPS, RESIDENT>;
    extern __shared__ __align__(128) uint8_t smem[];
    uint8_t *slab = smem;                                   // [2][SLAB_BYTES], or [3][HALF_BYTES] (C::HALVES)
    uint8_t *wbuf = smem + C::A_BYTES;                      // [NSTAGE][WSTEP_BYTES]
    float *sbias = reinterpret_cast<float *>(wbuf + C::W_BYTES);
    uint64_t *bars = reinterpret_cast<uint64_t *>(reinterpret_cast<uint8_t *>(sbias) + COUT * 4);
    uint64_t *a_full = bars, *a_empty = bars + 3, *acc_full = bars + 6, *acc_empty = bars + 8;
    uint64_t *b_full = bars + 10, *b_empty = bars + 16;     // ring (streaming) or b_full[0] = "resident loaded"
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(bars + 22);
    static_assert(C::NSTAGE <= 6 || RESIDENT, "barrier slots");
    float *s_hw = reinterpret_cast<float *>(bars + 32);     // HEADS: [COUT][2] policy, [COUT] value, 3 biases

    const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0), lane = threadIdx.x & 31;   // provably warp-uniform
    // rev: the loop index b walks the boards from the last to the first (memory index ph(b)); consecutive layers
    // alternate, so a layer starts on the boards its predecessor wrote last, the ones still in L2
    auto ph = [&](int b) { return rev ? G - 1 - b : b; };
    if (HEADS) {
        for (int i = threadIdx.x; i < 2 * COUT; i += blockDim.x) s_hw[i] = heads.pw[i];
        for (int i = threadIdx.x; i < COUT; i += blockDim.x) s_hw[2 * COUT + i] = heads.vw[i];
        if (threadIdx.x == 0) { s_hw[3 * COUT] = heads.pb[0]; s_hw[3 * COUT + 1] = heads.pb[1]; s_hw[3 * COUT + 2] = heads.vb[0]; }
    }
    if (threadIdx.x == 0) {
        for (int i = 0; i < 3; ++i) {
            ptx::mbar_init(&a_full[i], 1);
            ptx::mbar_init(&a_empty[i], 1);
        }
        for (int i = 0; i < 2; ++i) {
            ptx::mbar_init(&acc_full[i], 1);
            ptx::mbar_init(&acc_empty[i], conv_epi_warps<FROM_BOARD, RESIDUAL>());
        }
        for (int i = 0; i < 6; ++i) {
            ptx::mbar_init(&b_full[i], 1);
            ptx::mbar_init(&b_empty[i], 1);
        }
        ptx::fence_mbar_init();
    }
    for (int i = threadIdx.x; i < COUT; i += blockDim.x) sbias[i] = bias[i];
    if (FROM_BOARD) {   // channel group 1 and the rows outside the board stay zero for the whole kernel
        for (int i = threadIdx.x; i < 2 * C::SLAB_BYTES / 16; i += blockDim.x)
            reinterpret_cast<uint4 *>(slab)[i] = make_uint4(0u, 0u, 0u, 0u);
        ptx::fence_proxy_async();
    }
    if (warp == 1) ptx::tmem_alloc<C::TCOLS>(tmem_slot);
    ptx::tc_fence_before();
    __syncthreads();
    ptx::tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    // programmatic dependent launch (gz_engine.h): everything above ran beside the tail of the kernel before this one;
    // warp 0 still issues its weight loads first
    pdl_trigger();
    if (warp != 0) pdl_wait();

    if (warp == 0 && FROM_BOARD) {
        // ================= producer: one-hot slab from the bitboards =================
        const gz_board *boards = reinterpret_cast<const gz_board *>(in_planes);
        if (lane == 0) {
            ptx::mbar_expect_tx(&b_full[0], C::KSTEPS * C::WSTEP_BYTES);
            ptx::bulk_g2s(wbuf, w_packed, C::KSTEPS * C::WSTEP_BYTES, &b_full[0]);
        }
        pdl_wait();
        // the next board's bitboards are fetched while this one is expanded (the loop is otherwise one global
        // load latency per board long)
        auto next_board = [&](int b) {
            for (b += gridDim.x; b < G && active && !active[ph(b)]; b += gridDim.x) {}
            return b;
        };
        auto load_stones = [&](int b) { return lane < 16 && b < G ? boards[ph(b)].stones[lane >> 3][lane & 7] : 0u; };
        uint32_t it = 0;
        int b = blockIdx.x < G && active && !active[ph(blockIdx.x)] ? next_board(blockIdx.x) : blockIdx.x;
        uint32_t st = load_stones(b);
        while (b < G) {
            const int ab = it & 1;
            const int bn = next_board(b);
            const uint32_t st_next = load_stones(bn);
            if (lane == 0) ptx::mbar_wait(&a_empty[ab], ((it >> 1) & 1) ^ 1);
            __syncwarp();
            uint8_t *dst = slab + ab * C::SLAB_BYTES;
#pragma unroll
            for (int k = 0; k < 8; ++k) {           // 256 positions [16,272), 8 per lane
                const int pos = 16 + 32 * k + lane, p = pos - 16;
                // bit p of black / white lives in lane (p >> 5) / 8 + (p >> 5)
                const uint32_t wb = __shfl_sync(0xffffffffu, st, p >> 5);
                const uint32_t ww = __shfl_sync(0xffffffffu, st, 8 + (p >> 5));
                const bool valid = (pos & 15) != 15 && (pos >> 4) <= 15;
                const uint32_t bl = (wb >> (p & 31)) & 1u, wh = (ww >> (p & 31)) & 1u;
                const uint32_t em = valid ? 1u - (bl | wh) : 0u;
                *reinterpret_cast<uint4 *>(dst + (pos + 1) * 16) =
                    make_uint4(bl * 0x3F80u + wh * 0x3F800000u, em * 0x3F80u, 0u, 0u);
            }
            ptx::fence_proxy_async();
            __syncwarp();
            if (lane == 0) ptx::mbar_arrive(&a_full[ab]);
            ++it;
            b = bn;
            st = st_next;
        }
    } else if (warp == 0) {
        // ================= producer =================
        if (lane == 0) {
            if (RESIDENT) {
                ptx::mbar_expect_tx(&b_full[0], C::KSTEPS * C::WSTEP_BYTES);
                for (int ks = 0; ks < C::KSTEPS; ++ks)
                    ptx::bulk_g2s(wbuf + ks * C::WSTEP_BYTES, w_packed + (size_t)ks * C::WSTEP_BYTES,
                                  C::WSTEP_BYTES, &b_full[0]);
            }
            pdl_wait();
            uint32_t it = 0, wstage = 0, wphase = 0;
            auto next_board = [&](int b) {
                for (b += gridDim.x; b < G && active && !active[ph(b)]; b += gridDim.x) {}
                return b;
            };
            // slab of the n-th board of this CTA: channel groups [cg0, cg1)
            auto load_slab = [&](int b, uint32_t n, int cg0, int cg1) {
                const int ab = n & 1;
                if (cg0 == 0) {
                    ptx::mbar_wait(&a_empty[ab], ((n >> 1) & 1) ^ 1);
                    ptx::mbar_expect_tx(&a_full[ab], C::SLAB_BYTES);
                }
                const uint8_t *src = in_planes + kGuardBytes + (size_t)ph(b) * C::CG * kPlaneBytes +
                                     (TAPS == 9 ? -16 : 256);
                for (int cg = cg0; cg < cg1; ++cg)
                    ptx::bulk_g2s(slab + ab * C::SLAB_BYTES + cg * C::SLAB_ROWS * 16,
                                  src + (size_t)cg * kPlaneBytes, C::SLAB_ROWS * 16, &a_full[ab]);
            };
            // C::HALVES: channel group cg of half slab number j = 2 * (board number) + chunk, ring of three
            auto load_half = [&](int b, uint32_t j, int cg) {
                const uint32_t hb = j % 3;
                if (cg == 0) {
                    ptx::mbar_wait(&a_empty[hb], ((j / 3) & 1) ^ 1);
                    ptx::mbar_expect_tx(&a_full[hb], C::HALF_BYTES);
                }
                const uint8_t *src = in_planes + kGuardBytes + ((size_t)ph(b) * C::CG + (j & 1) * 8 + cg) * kPlaneBytes - 16;
                ptx::bulk_g2s(slab + hb * C::HALF_BYTES + cg * C::SLAB_ROWS * 16, src, C::SLAB_ROWS * 16, &a_full[hb]);
            };
            int b = blockIdx.x < G && active && !active[ph(blockIdx.x)] ? next_board(blockIdx.x) : blockIdx.x;
            if constexpr (C::HALVES) {
                if (b < G)
                    for (int cg = 0; cg < 8; ++cg) load_half(b, 0, cg);
                while (b < G) {
                    const int bn = next_board(b);
                    // weight steps in chunk-major order; along with them, one 4.6 KB channel group per step, this board's
                    // second half slab (steps 0-7: needed from step 9) and the next board's first (steps 8-15).  Both ring
                    // slots are free by then (the producer runs at most six steps ahead of the UMMAs), so the weight
                    // stream never waits for them; a 37 KB or 74 KB request in one piece delays the weight ring instead.
                    for (int kk = 0; kk < C::KSTEPS; ++kk) {
                        if (kk < 8) load_half(b, 2 * it + 1, kk);
                        else if (kk < 16 && bn < G) load_half(bn, 2 * it + 2, kk - 8);
                        const int ks = (kk % 9) * 2 + kk / 9;           // packed order is tap-major
                        ptx::mbar_wait(&b_empty[wstage], wphase ^ 1);
                        ptx::mbar_expect_tx(&b_full[wstage], C::WSTEP_BYTES);
                        ptx::bulk_g2s(wbuf + wstage * C::WSTEP_BYTES, w_packed + (size_t)ks * C::WSTEP_BYTES,
                                      C::WSTEP_BYTES, &b_full[wstage]);
                        if (++wstage == C::NSTAGE) { wstage = 0; wphase ^= 1; }
                    }
                    ++it;
                    b = bn;
                }
            } else {
            if (b < G) load_slab(b, 0, 0, C::CG);
            while (b < G) {
                const int bn = next_board(b);
                if (RESIDENT) {
                    if (bn < G) load_slab(bn, it + 1, 0, C::CG);
                } else {
                    static_assert(RESIDENT || C::CG <= C::KSTEPS, "one channel group of the next slab per weight step");
                    for (int ks = 0; ks < C::KSTEPS; ++ks) {
                        // The next board's slab rides along with this board's weight stream, one channel group (4.6 KB)
                        // per 16 KB weight step.  Requested behind the 18 weight steps it arrived ~1000 cycles late on
                        // every board; requested in one piece early in the board its 74 KB delayed the weight ring,
                        // which is the latency-critical stream (b_full waits 4.0 k -> 6.4 k cycles per board;
                        // scratch/conv_timing.py).
                        if (bn < G && ks < C::CG) load_slab(bn, it + 1, ks, ks + 1);
                        ptx::mbar_wait(&b_empty[wstage], wphase ^ 1);
                        ptx::mbar_expect_tx(&b_full[wstage], C::WSTEP_BYTES);
                        ptx::bulk_g2s(wbuf + wstage * C::WSTEP_BYTES, w_packed + (size_t)ks * C::WSTEP_BYTES,
                                      C::WSTEP_BYTES, &b_full[wstage]);
                        if (++wstage == C::NSTAGE) { wstage = 0; wphase ^= 1; }
                    }
                }
                ++it;
                b = bn;
            }
            }
        }
    } else if (warp == 1) {
        // ================= MMA issuer (whole warp, one elected lane issues: gz_tcgen05.cuh) =================
        {
            constexpr uint32_t idesc = ptx::make_idesc_bf16(128, COUT);
            constexpr uint32_t LBO_A = C::SLAB_ROWS * 16, LBO_B = COUT * 16;
            if (RESIDENT) ptx::mbar_wait(&b_full[0], 0);
            uint32_t it = 0, wstage = 0, wphase = 0;
            CV_T(const long long tk0 = clock64(); long long w_acc = 0, w_a = 0, w_b = 0;)
            for (int b = blockIdx.x; b < G; b += gridDim.x) {
                if (active && !active[ph(b)]) continue;
                const int ab = it & 1;
                const uint32_t par = (it >> 1) & 1;
                CV_T(const long long t0 = clock64();)
                ptx::mbar_wait(&acc_empty[ab], par ^ 1);
                CV_T(const long long t1 = clock64(); w_acc += t1 - t0;)
                if constexpr (C::HALVES) ptx::mbar_wait(&a_full[(2 * it) % 3], ((2 * it) / 3) & 1);
                else ptx::mbar_wait(&a_full[ab], par);
                CV_T(w_a += clock64() - t1;)
                ptx::tc_fence_after();
                // one base descriptor per operand; everything else is a compile-time constant added to the
                // 16-byte address field (K-step loop fully unrolled)
                const uint32_t acc = tmem_base + ab * 2 * COUT;
                if constexpr (C::HALVES) {
#pragma unroll
                    for (int kk = 0; kk < C::KSTEPS; ++kk) {
                        const int chunk = kk / 9, tap = kk - 9 * chunk;
                        const int rowoff = 17 + 16 * (tap / 3 - 1) + (tap % 3 - 1);
                        const uint32_t j = 2 * it + chunk, hb = j % 3;
                        if (tap == 0 && chunk == 1) {
                            CV_T(const long long t3 = clock64();)
                            ptx::mbar_wait(&a_full[hb], (j / 3) & 1);
                            CV_T(w_a += clock64() - t3;)
                            ptx::tc_fence_after();
                        }
                        const uint64_t da0 = ptx::make_desc(ptx::smem_u32(slab + hb * C::HALF_BYTES), LBO_A, 128);
                        CV_T(const long long t2 = clock64();)
                        ptx::mbar_wait(&b_full[wstage], wphase);
                        CV_T(w_b += clock64() - t2;)
                        ptx::tc_fence_after();
                        const uint64_t db0 = ptx::make_desc(ptx::smem_u32(wbuf + wstage * C::WSTEP_BYTES), LBO_B, 128);
#pragma unroll
                        for (int h = 0; h < 2; ++h) {
#pragma unroll
                            for (int k = 0; k < C::KK; ++k)
                                ptx::umma_bf16_elect(acc + h * COUT,
                                                     da0 + (uint64_t)(2 * k * C::SLAB_ROWS + rowoff + 128 * h),
                                                     db0 + (uint64_t)((2 * k * (int)LBO_B) >> 4), idesc, (kk | k) != 0);
                        }
                        ptx::umma_commit_elect(&b_empty[wstage]);
                        if (++wstage == C::NSTAGE) { wstage = 0; wphase ^= 1; }
                        if (tap == 8) ptx::umma_commit_elect(&a_empty[hb]);
                    }
                    ptx::umma_commit_elect(&acc_full[ab]);
                    ++it;
                    continue;
                }
                const uint64_t da0 = ptx::make_desc(ptx::smem_u32(slab + ab * C::SLAB_BYTES), LBO_A, 128);
#pragma unroll
                for (int ks = 0; ks < C::KSTEPS; ++ks) {
                    const int tap = ks / C::CHUNKS, chunk = ks - tap * C::CHUNKS;
                    constexpr int CGS = C::KSTEP / 8;       // channel groups per K-step
                    const int rowoff = TAPS == 9 ? 17 + 16 * (tap / 3 - 1) + (tap % 3 - 1) : 0;
                    uint64_t db0;
                    if (RESIDENT) {
                        db0 = ptx::make_desc(ptx::smem_u32(wbuf) + ks * C::WSTEP_BYTES, LBO_B, 128);
                    } else {
                        CV_T(const long long t2 = clock64();)
                        ptx::mbar_wait(&b_full[wstage], wphase);
                        CV_T(w_b += clock64() - t2;)
                        ptx::tc_fence_after();
                        db0 = ptx::make_desc(ptx::smem_u32(wbuf + wstage * C::WSTEP_BYTES), LBO_B, 128);
                    }
#pragma unroll
                    for (int h = 0; h < 2; ++h) {
#pragma unroll
                        for (int k = 0; k < C::KK; ++k)
                            ptx::umma_bf16_elect(acc + h * COUT,
                                                 da0 + (uint64_t)((chunk * CGS + 2 * k) * C::SLAB_ROWS + rowoff + 128 * h),
                                                 db0 + (uint64_t)((2 * k * (int)LBO_B) >> 4), idesc, (ks | k) != 0);
                    }
                    if (!RESIDENT) {
                        ptx::umma_commit_elect(&b_empty[wstage]);
                        if (++wstage == C::NSTAGE) { wstage = 0; wphase ^= 1; }
                    }
                }
                ptx::umma_commit_elect(&a_empty[ab]);
                ptx::umma_commit_elect(&acc_full[ab]);
                ++it;
            }
            CV_T(if (blockIdx.x == 0 && lane == 0 && it > 0 && CIN == 128)
                     printf("conv issuer (res %d heads %d): %u boards, total %lld | per board: wait acc_empty %lld, a_full %lld, b_full (18 waits) %lld\n",
                            (int)RESIDUAL, (int)HEADS, it, clock64() - tk0, w_acc / it, w_a / it, w_b / it);)
        }
    } else {
        // ================= epilogue (warps 2..5) =================
        // RESIDUAL: the skip tensor of a position (COUT x 2 B = 16 x 16 B, one per channel group) is fetched
        // into registers a half board ahead, before the wait on the accumulator, so that its latency hides
        // behind the MMAs instead of stalling every 16-channel step.
        const int q = warp & 3;                 // TMEM lane quadrant this warp may read
        constexpr int NCG = COUT / 8;
        uint4 res[conv_epi_warps<FROM_BOARD, RESIDUAL>() == 8 ? 1 : 2][RESIDUAL ? NCG : 1];
        auto fetch_res = [&](int bl, int h, uint4 *r) {
            const int b = ph(bl);
            const uint8_t *src = res_planes + kGuardBytes + (size_t)b * NCG * kPlaneBytes +
                                 (size_t)(16 + 128 * h + 32 * q + lane) * 16;
#pragma unroll
            for (int cg = 0; cg < NCG; ++cg) r[cg] = *reinterpret_cast<const uint4 *>(src + (size_t)cg * kPlaneBytes);
        };
        auto half = [&](int bl, int ab, int h, const uint4 *r) {
            const int b = ph(bl);
            const size_t obase = kGuardBytes + (size_t)b * NCG * kPlaneBytes;
            const int pos = 16 + 128 * h + 32 * q + lane;
            const bool valid = (pos & 15) != 15 && (pos >> 4) <= 15;
            const uint32_t taddr = tmem_base + ((uint32_t)(32 * q) << 16) + ab * 2 * COUT + h * COUT;
            if (FROM_BOARD && COUT == 64) {   // epilogue-bound layer: 64 columns per TMEM round trip (41 -> 35 us; no gain at 128)
#pragma unroll 1
                for (int c64 = 0; c64 < COUT; c64 += 64) {
                    float v[64];
                    ptx::tmem_ld64(taddr + c64, v);
#pragma unroll
                    for (int c0 = 0; c0 < 64; c0 += 16) {
                        uint32_t pk[8];
#pragma unroll
                        for (int i = 0; i < 8; ++i) {
                            const float x0 = fmaxf(v[c0 + 2 * i] + sbias[c64 + c0 + 2 * i], 0.f);
                            const float x1 = fmaxf(v[c0 + 2 * i + 1] + sbias[c64 + c0 + 2 * i + 1], 0.f);
                            pk[i] = valid ? ptx::pack_bf16(x0, x1) : 0u;
                        }
                        const size_t o0 = obase + (size_t)((c64 + c0) / 8) * kPlaneBytes + (size_t)pos * 16;
                        *reinterpret_cast<uint4 *>(out_planes + o0) = make_uint4(pk[0], pk[1], pk[2], pk[3]);
                        *reinterpret_cast<uint4 *>(out_planes + o0 + kPlaneBytes) = make_uint4(pk[4], pk[5], pk[6], pk[7]);
                    }
                }
                return;
            }
            float hp0 = 0.f, hp1 = 0.f, hv0 = 0.f;
            if (HEADS) { hp0 = s_hw[3 * COUT]; hp1 = s_hw[3 * COUT + 1]; hv0 = s_hw[3 * COUT + 2]; }
#pragma unroll(RESIDUAL ? COUT / 16 : 1)
            for (int c0 = 0; c0 < COUT; c0 += 16) {
                float v[16];
                ptx::tmem_ld16(taddr + c0, v);
                const size_t o0 = obase + (size_t)(c0 / 8) * kPlaneBytes + (size_t)pos * 16;
                if (RESIDUAL) {
                    const uint4 r0 = r[c0 / 8], r1 = r[c0 / 8 + 1];
                    const uint32_t rr[8] = {r0.x, r0.y, r0.z, r0.w, r1.x, r1.y, r1.z, r1.w};
#pragma unroll
                    for (int i = 0; i < 8; ++i) {
                        v[2 * i] += __uint_as_float(rr[i] << 16);
                        v[2 * i + 1] += __uint_as_float(rr[i] & 0xFFFF0000u);
                    }
                }
                if (HEADS) {
#pragma unroll
                    for (int i = 0; i < 16; ++i) {
                        const float x = fmaxf(v[i] + sbias[c0 + i], 0.f);
                        hp0 = fmaf(s_hw[2 * (c0 + i)], x, hp0);
                        hp1 = fmaf(s_hw[2 * (c0 + i) + 1], x, hp1);
                        hv0 = fmaf(s_hw[2 * COUT + c0 + i], x, hv0);
                    }
                    continue;
                }
                uint32_t pk[8];
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    float x0 = fmaxf(v[2 * i] + sbias[c0 + 2 * i], 0.f);
                    float x1 = fmaxf(v[2 * i + 1] + sbias[c0 + 2 * i + 1], 0.f);
                    pk[i] = valid ? ptx::pack_bf16(x0, x1) : 0u;
                }
                *reinterpret_cast<uint4 *>(out_planes + o0) = make_uint4(pk[0], pk[1], pk[2], pk[3]);
                *reinterpret_cast<uint4 *>(out_planes + o0 + kPlaneBytes) = make_uint4(pk[4], pk[5], pk[6], pk[7]);
            }
            if (HEADS && valid) {
                const int cell = ((pos >> 4) - 1) * 15 + (pos & 15);
                store_head_operand(heads.pc, b, cell, kHeadChunksP, hp0);
                store_head_operand(heads.pc, b, kCells + cell, kHeadChunksP, hp1);
                store_head_operand(heads.vc, b, cell, kHeadChunksV, hv0);
            }
        };
        auto next_board = [&](int b) {
            for (b += gridDim.x; b < G && active && !active[ph(b)]; b += gridDim.x) {}
            return b;
        };
        uint32_t it = 0;
        CV_T(long long e_wait = 0, e_work = 0;)
        int b = blockIdx.x < G && active && !active[ph(blockIdx.x)] ? next_board(blockIdx.x) : blockIdx.x;
        constexpr bool kEight = conv_epi_warps<FROM_BOARD, RESIDUAL>() == 8;
        const int my_h = (warp - 2) >> 2;       // kEight: the M half of this warp
        if (RESIDUAL && b < G) fetch_res(b, kEight ? my_h : 0, res[0]);
        while (b < G) {
            const int ab = it & 1;
            const int bn = next_board(b);
            CV_T(const long long t0 = clock64();)
            ptx::mbar_wait(&acc_full[ab], (it >> 1) & 1);
            CV_T(const long long t1 = clock64(); e_wait += t1 - t0;)
            ptx::tc_fence_after();
            if constexpr (kEight) {
                half(b, ab, my_h, res[0]);
                // the next board's skip tensor for this half: its latency hides behind that board's UMMAs
                if (RESIDUAL && bn < G) fetch_res(bn, my_h, res[0]);
            } else {
                if (RESIDUAL) fetch_res(b, 1, res[1]);
                half(b, ab, 0, res[0]);
                if (RESIDUAL && bn < G) fetch_res(bn, 0, res[0]);
                half(b, ab, 1, res[1]);
            }
            ptx::tc_fence_before();
            __syncwarp();
            if (lane == 0) ptx::mbar_arrive(&acc_empty[ab]);
            CV_T(e_work += clock64() - t1;)
            ++it;
            b = bn;
        }
        CV_T(if (blockIdx.x == 0 && lane == 0 && warp == 2 && it > 0 && CIN == 128)
                 printf("conv epilogue (res %d heads %d): per board: wait acc_full %lld, work %lld\n", (int)RESIDUAL, (int)HEADS, e_wait / it, e_work / it);)
    }
    ptx::tc_fence_before();
    __syncthreads();
    if (warp == 1) ptx::tmem_dealloc<C::TCOLS>(tmem_base);
}

// bf16 planes -> fp32 NCHW [G][C][225] (debug / parity inspection, and the input of the fp32 heads)
__global__ void __launch_bounds__(256)
k_planes_to_nchw(const uint8_t *__restrict__ planes, int C, const uint8_t *__restrict__ active,
                 float *__restrict__ out) {
    const int b = blockIdx.x;
    if (active && !active[b]) return;
    for (int i = threadIdx.x; i < C * kCells; i += 256) {
        int ch = i / kCells, cell = i - ch * kCells;
        int pos = 16 * (cell / 15 + 1) + cell % 15;
        const uint16_t *p = reinterpret_cast<const uint16_t *>(planes + kGuardBytes +
            ((size_t)b * (C / 8) + ch / 8) * kPlaneBytes + (size_t)pos * 16) + (ch & 7);
        out[(size_t)b * C * kCells + i] = __uint_as_float((uint32_t)(*p) << 16);
    }
}

// ------------------------------------------------------------------------------------------------
// One GraphNet block (bg_planner.py:34-41: conv3x3 64->64 + ReLU, conv1x1 64->64 + ReLU) fused: the
// 3x3 output never leaves the SM.  Per board: 72 UMMAs (3x3) -> TMEM -> bias + ReLU -> bf16 "mid"
// tile in shared memory (the A operand of the 1x1) -> 8 UMMAs -> TMEM -> bias + ReLU -> planes.
// Both weight sets are resident in shared memory; slabs, mid tiles and both accumulators are double
// buffered, and the issuer runs one board ahead with the 3x3 so the tensor pipe never waits for an
// epilogue:  MMA3(0) MMA3(1) MMA1(0) MMA3(2) MMA1(1) ...
//   warp 0     slab producer (bulk copies)          warps 1, 10  tcgen05.mma issuers (M half 0 / 1)
//   warps 2-9  epilogues: quadrant = warp & 3, M half = (warp - 2) >> 2
// -DGZ_GN_TIMING: block 0 prints where its producer / issuer / epilogue warps spend their cycles (scratch/gn_timing.py)
#ifdef GZ_GN_TIMING
#define GN_T(...) __VA_ARGS__
#else
#define GN_T(...)
#endif
struct GnBlockCfg {
    static constexpr int SLAB_ROWS = 290, SLAB_BYTES = 8 * SLAB_ROWS * 16;
    static constexpr int MID_BYTES = 8 * 256 * 16;
    static constexpr int W3_BYTES = 9 * 8 * 64 * 16, W1_BYTES = 8 * 64 * 16;
    static constexpr int SMEM = 2 * SLAB_BYTES + W3_BYTES + W1_BYTES + 2 * 64 * 4 + 256 + 132 * 4;
};

// Biases (and the head's 1x1 weights) travel as a kernel parameter: after unrolling every index is a
// compile-time constant, so the epilogue's FADD/FFMA read them straight from the constant bank / uniform
// registers.  As shared-memory loads they were 17 % of the SM's shared-memory wavefronts, the resource this
// kernel is bound by (tensor operand fetches are the rest; profiles/r01_gn_block_ncu_full.md).
struct GnBlockConsts {
    float b3[64], b1[64];
    float head_w[128], head_b[2];   // FINAL only: [64][2], [2]
    float be[64];                   // k_gn_block0_embed only: the embed convolution's bias
};

// FINAL (the last block): the head's 1x1 convolution 64 -> 2 (bg_planner.py:43,54; no activation) is
// applied to the fp32 epilogue values and only its 450 outputs per board are written (`head_out`:
// bf16 hi/lo operand tiles of the head GEMM, K index channel-major like the flatten at bg_planner.py:55); the 64-channel planes of the
// last layer are never stored.
template <bool FINAL>
__global__ void __launch_bounds__(352, 1)
k_gn_block(const uint8_t *__restrict__ in_planes, const uint8_t *__restrict__ w3, const uint8_t *__restrict__ w1,
           const __grid_constant__ GnBlockConsts kc, uint8_t *__restrict__ out_planes,
           uint8_t *__restrict__ head_out, int G) {
    using C = GnBlockCfg;
    extern __shared__ __align__(128) uint8_t smem[];
    uint8_t *slab = smem;
    uint8_t *sw3 = slab + 2 * C::SLAB_BYTES;
    uint8_t *sw1 = sw3 + C::W3_BYTES;
    float *sb3 = reinterpret_cast<float *>(sw1 + C::W1_BYTES);
    float *sb1 = sb3 + 64;
    uint64_t *bars = reinterpret_cast<uint64_t *>(sb1 + 64);
    uint64_t *a_full = bars, *a_empty = bars + 2, *d1_full = bars + 4, *d1_empty = bars + 6;
    uint64_t *mid_full = bars + 8, *mid_empty = bars + 10, *d2_full = bars + 12, *d2_empty = bars + 14;
    uint64_t *w_full = bars + 16;
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(bars + 17);

    const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0), lane = threadIdx.x & 31;   // provably warp-uniform
    if (FINAL) {
    }
    if (threadIdx.x == 0) {
        for (int i = 0; i < 2; ++i) {   // every tcgen05.commit comes twice: one per issuer warp (M half)
            ptx::mbar_init(&a_full[i], 1);   ptx::mbar_init(&a_empty[i], 2);
            ptx::mbar_init(&d1_full[i], 2);  ptx::mbar_init(&d1_empty[i], 8);
            ptx::mbar_init(&mid_full[i], 8); ptx::mbar_init(&mid_empty[i], 2);
            ptx::mbar_init(&d2_full[i], 2);  ptx::mbar_init(&d2_empty[i], 8);
        }
        ptx::mbar_init(w_full, 1);
        ptx::fence_mbar_init();
    }
    if (warp == 1) ptx::tmem_alloc<512>(tmem_slot);
    ptx::tc_fence_before();
    __syncthreads();
    ptx::tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    const int n_it = blockIdx.x < G ? (G - 1 - blockIdx.x) / gridDim.x + 1 : 0;

    if (warp == 0) {
        if (lane == 0) {
            ptx::mbar_expect_tx(w_full, C::W3_BYTES + C::W1_BYTES);
            for (int t = 0; t < 9; ++t) ptx::bulk_g2s(sw3 + t * 8192, w3 + t * 8192, 8192, w_full);
            ptx::bulk_g2s(sw1, w1, C::W1_BYTES, w_full);
            GN_T(long long w_ae = 0;)
            for (int it = 0; it < n_it; ++it) {
                const int b = blockIdx.x + it * gridDim.x, ab = it & 1;
                GN_T(long long t0 = clock64();)
                ptx::mbar_wait(&a_empty[ab], ((it >> 1) & 1) ^ 1);
                GN_T(w_ae += clock64() - t0;)
                ptx::mbar_expect_tx(&a_full[ab], C::SLAB_BYTES);
                const uint8_t *src = in_planes + kGuardBytes + (size_t)b * 8 * kPlaneBytes - 16;
                for (int cg = 0; cg < 8; ++cg)
                    ptx::bulk_g2s(slab + ab * C::SLAB_BYTES + cg * C::SLAB_ROWS * 16,
                                  src + (size_t)cg * kPlaneBytes, C::SLAB_ROWS * 16, &a_full[ab]);
            }
            GN_T(if (blockIdx.x == 0) printf("gn producer: wait a_empty %lld per board\n", w_ae / (n_it > 0 ? n_it : 1));)
        }
    } else if (warp == 1 || warp == 10) {
        // Two issuer warps, one per M half of the board (rows 0-127 / 128-255, separate accumulators): while one
        // sits in its per-board barrier work (commits, waits, fences) the other keeps the tensor pipe fed.
        const int h = warp == 1 ? 0 : 1;
        {   // whole warp; one elected lane issues (gz_tcgen05.cuh)
            constexpr uint32_t idesc = ptx::make_idesc_bf16(128, 64);
            constexpr uint32_t LBO_A = C::SLAB_ROWS * 16, LBO_B = 64 * 16;
            GN_T(long long tk0 = clock64(); long long w_d1e = 0, w_af = 0, t_i3 = 0, w_mf = 0, w_d2e = 0, t_i1 = 0;)
            ptx::mbar_wait(w_full, 0);
            GN_T(const long long w_w = clock64() - tk0;)
            const uint32_t w3_addr = ptx::smem_u32(sw3), w1_addr = ptx::smem_u32(sw1);
            const uint64_t db0 = ptx::make_desc(w3_addr, LBO_B, 128);
            const uint64_t dw0 = ptx::make_desc(w1_addr, LBO_B, 128);
            // The tensor pipe takes UMMAs from a shallow queue: issuing the 36 instructions of a board's 3x3 blocks this
            // warp for most of their execution time (measured: 2.9 k of 3.9 k cycles per board), and everything it did
            // BETWEEN boards -- four mbarrier waits at ~60-150 cycles each even when already satisfied, the 1x1's four
            // UMMAs, four commits -- was a bubble both issuers hit at the same moment.  So the per-board work is software
            // pipelined into the 3x3's instruction stream, where it costs nothing while the queue is full:
            //   taps 0-3 of board it | 1x1 of board it-1 (its A operand was packed by epilogue 1 long ago) |
            //   taps 4-7 | the waits for board it+1's accumulator and slab | tap 8 | commits
            // and the next board's first UMMA follows the commits directly.
            auto wait_board = [&](int it) {
                const int ab = it & 1;
                const uint32_t par = (it >> 1) & 1;
                GN_T(long long t0 = clock64();)
                ptx::mbar_wait(&d1_empty[ab], par ^ 1);
                GN_T(long long t1 = clock64(); w_d1e += t1 - t0;)
                ptx::mbar_wait(&a_full[ab], par);
                GN_T(w_af += clock64() - t1;)
                ptx::tc_fence_after();
            };
            auto issue_1x1 = [&](int j) {
                const int jb = j & 1;
                const uint32_t par = (j >> 1) & 1;
                GN_T(long long t0 = clock64();)
                ptx::mbar_wait(&mid_full[jb], par);
                GN_T(long long t1 = clock64(); w_mf += t1 - t0;)
                ptx::mbar_wait(&d2_empty[jb], par ^ 1);
                GN_T(long long t2 = clock64(); w_d2e += t2 - t1;)
                ptx::tc_fence_after();
                // A = the 3x3 result packed to bf16 in place in its accumulator columns (epilogue 1): TS form
                const uint32_t acc = tmem_base + 256 + jb * 128;
                const uint32_t amid = tmem_base + jb * 128;
#pragma unroll
                for (int k = 0; k < 4; ++k)
                    ptx::umma_bf16_ts_elect(acc + h * 64, amid + h * 64 + 8 * k,
                                            dw0 + (uint64_t)((2 * k * (int)LBO_B) >> 4), idesc, k != 0);
                ptx::umma_commit_elect(&mid_empty[jb]);
                ptx::umma_commit_elect(&d2_full[jb]);
                GN_T(t_i1 += clock64() - t2;)
            };
            if (n_it > 0) wait_board(0);
            for (int it = 0; it < n_it; ++it) {
                const int ab = it & 1;
                GN_T(long long t2 = clock64();)
                // Descriptors = one base per operand + compile-time constants in the 16-byte address field: anything more
                // per instruction shows at ~40 cycles per UMMA (fully unrolled: 36 instructions, two 64-bit adds each).
                const uint64_t da_h = ptx::make_desc(ptx::smem_u32(slab + ab * C::SLAB_BYTES), LBO_A, 128) + (uint64_t)(128 * h);
                const uint32_t acc_h = tmem_base + ab * 128 + h * 64;
#define GZ_GN_TAPS(T0, T1)                                                                                         \
                _Pragma("unroll") for (int tap = T0; tap < T1; ++tap) {                                            \
                    const int rowoff = 17 + 16 * (tap / 3 - 1) + (tap % 3 - 1);                                    \
                    _Pragma("unroll") for (int k = 0; k < 4; ++k)                                                  \
                        ptx::umma_bf16_elect(acc_h, da_h + (uint64_t)(2 * k * C::SLAB_ROWS + rowoff),              \
                                             db0 + (uint64_t)((tap * 8192 + 2 * k * (int)LBO_B) >> 4), idesc,       \
                                             (tap | k) != 0);                                                      \
                }
                GZ_GN_TAPS(0, 4)
                if (it > 0) issue_1x1(it - 1);
                GZ_GN_TAPS(4, 8)
                if (it + 1 < n_it) wait_board(it + 1);
                GZ_GN_TAPS(8, 9)
#undef GZ_GN_TAPS
                ptx::umma_commit_elect(&a_empty[ab]);
                ptx::umma_commit_elect(&d1_full[ab]);
                GN_T(t_i3 += clock64() - t2;)
            }
            if (n_it > 0) issue_1x1(n_it - 1);
            GN_T(if (blockIdx.x == 0 && lane == 0 && n_it > 0)
                     printf("gn issuer h=%d: %d boards, total %lld | per board: wait d1_empty %lld, a_full %lld, issue3x3 %lld, wait mid_full %lld, d2_empty %lld, issue1x1 %lld | weights %lld\n",
                            h, n_it, clock64() - tk0, w_d1e / n_it, w_af / n_it, t_i3 / n_it, w_mf / n_it, w_d2e / n_it, t_i1 / n_it, w_w);)
        }
    } else {
        const int q = warp & 3, h = (warp - 2) >> 2;
        const int row = 128 * h + 32 * q + lane;            // tile row = plane position - 16
        const int pos = 16 + row;
        const bool valid = (pos & 15) != 15 && (pos >> 4) <= 15;
        const uint32_t tlane = tmem_base + ((uint32_t)(32 * q) << 16) + h * 64;
        GN_T(long long w_d1f = 0, w_me = 0, t_e1 = 0, w_d2f = 0, t_e2 = 0;)
        for (int it = 0; it <= n_it; ++it) {
            // the 1x1 of board it-1 ran in the middle of board it's 3x3: its epilogue first, so that this warp is
            // free the moment the 3x3 completes (epilogue 1 is what the issuers wait for next)
            if (it > 0) {
                const int j = it - 1, jb = j & 1;
                const int b = blockIdx.x + j * gridDim.x;
                GN_T(long long t0 = clock64();)
                ptx::mbar_wait(&d2_full[jb], (j >> 1) & 1);
                GN_T(long long t1 = clock64(); w_d2f += t1 - t0;)
                ptx::tc_fence_after();
                if (FINAL) {
                    float p0 = kc.head_b[0], p1 = kc.head_b[1];
#pragma unroll
                    for (int c0 = 0; c0 < 64; c0 += 16) {
                        float v[16];
                        ptx::tmem_ld16(tlane + 256 + jb * 128 + c0, v);
#pragma unroll
                        for (int i = 0; i < 16; ++i) {
                            const float x = fmaxf(v[i] + kc.b1[c0 + i], 0.f);
                            p0 = fmaf(kc.head_w[2 * (c0 + i)], x, p0);
                            p1 = fmaf(kc.head_w[2 * (c0 + i) + 1], x, p1);
                        }
                    }
                    if (valid) {
                        const int cell = ((pos >> 4) - 1) * 15 + (pos & 15);
                        store_head_operand(head_out, b, cell, kHeadChunksP, p0);
                        store_head_operand(head_out, b, kCells + cell, kHeadChunksP, p1);
                    }
                } else {
                uint8_t *dst = out_planes + kGuardBytes + (size_t)b * 8 * kPlaneBytes + (size_t)pos * 16;
    #pragma unroll
                    for (int c0 = 0; c0 < 64; c0 += 16) {
                        float v[16];
                        ptx::tmem_ld16(tlane + 256 + jb * 128 + c0, v);
                        uint32_t pk[8];
    #pragma unroll
                        for (int i = 0; i < 8; ++i)
                            pk[i] = valid ? ptx::pack_bf16(fmaxf(v[2 * i] + kc.b1[c0 + 2 * i], 0.f),
                                                           fmaxf(v[2 * i + 1] + kc.b1[c0 + 2 * i + 1], 0.f)) : 0u;
                        *reinterpret_cast<uint4 *>(dst + (size_t)(c0 / 8) * kPlaneBytes) = make_uint4(pk[0], pk[1], pk[2], pk[3]);
                        *reinterpret_cast<uint4 *>(dst + (size_t)(c0 / 8 + 1) * kPlaneBytes) = make_uint4(pk[4], pk[5], pk[6], pk[7]);
                    }
                }
                ptx::tc_fence_before();
                __syncwarp();
                if (lane == 0) ptx::mbar_arrive(&d2_empty[jb]);
                GN_T(t_e2 += clock64() - t1;)
            }
            if (it < n_it) {
                const int ab = it & 1;
                const uint32_t par = (it >> 1) & 1;
                GN_T(long long t0 = clock64();)
                ptx::mbar_wait(&d1_full[ab], par);
                GN_T(long long t1 = clock64(); w_d1f += t1 - t0;)
                ptx::mbar_wait(&mid_empty[ab], par ^ 1);
                GN_T(long long t2 = clock64(); w_me += t2 - t1;)
                ptx::tc_fence_after();
                // bias + ReLU + bf16 pack, written back IN PLACE over the accumulator's first 32 columns (the write
                // cursor c0/2 never passes the read cursor c0): the 1x1 reads it as its A operand from TMEM
#pragma unroll
                for (int c0 = 0; c0 < 64; c0 += 16) {
                    float v[16];
                    ptx::tmem_ld16(tlane + ab * 128 + c0, v);
                    uint32_t pk[8];
#pragma unroll
                    for (int i = 0; i < 8; ++i)
                        pk[i] = ptx::pack_bf16(fmaxf(v[2 * i] + kc.b3[c0 + 2 * i], 0.f),
                                               fmaxf(v[2 * i + 1] + kc.b3[c0 + 2 * i + 1], 0.f));
                    ptx::tmem_st8(tlane + ab * 128 + c0 / 2, pk);
                }
                ptx::tmem_wait_st();
                ptx::tc_fence_before();
                __syncwarp();
                if (lane == 0) { ptx::mbar_arrive(&d1_empty[ab]); ptx::mbar_arrive(&mid_full[ab]); }
                GN_T(t_e1 += clock64() - t2;)
            }
        }
        GN_T(if (blockIdx.x == 0 && lane == 0 && (warp == 2 || warp == 9) && n_it > 0)
                 printf("gn epilogue warp %d: per board: wait d1_full %lld, mid_empty %lld, epi1 %lld, wait d2_full %lld, epi2 %lld\n",
                        warp, w_d1f / n_it, w_me / n_it, t_e1 / n_it, w_d2f / n_it, t_e2 / n_it);)
    }
    ptx::tc_fence_before();
    __syncthreads();
    if (warp == 1) ptx::tmem_dealloc<512>(tmem_base);
}

// ------------------------------------------------------------------------------------------------
// k_gn_block_pairs: the same block with TWO taps per UMMA where the geometry allows it.
//
// k_gn_block is bound by the shared-memory operand fetches of its N = 64 UMMAs (A 4 KB + B 2 KB per instruction,
// ~40 cycles measured against 32 of math).  The A tiles of the nine taps are row-shifted views of one slab, so two taps
// can share one A fetch if their weights are stacked along N (N = 128: 8 KB for two taps instead of 12 KB, 64 cycles =
// the math rate): with A read at the offset of tap (dy, dx = 0), columns 0-63 accumulate that tap for output row m,
// and columns 64-127 accumulate W(dy, dx = -1) * slab[m + off(dy, 0)], which is tap (dy, -1)'s contribution to output
// row m + 1.  Epilogue 1 therefore adds D_b[m - 1] to D_a[m]: one lane shuffle, and NO exchange between warps, because
// lane 0 of every epilogue warp sits on board column 0 (tile row = multiple of 32 -> plane position = multiple of
// 16), whose left neighbour is the zero padding column: the missing term is exactly zero.  The dx = +1 taps stay
// single (N = 64).  Per board and M half: 12 N = 128 + 12 N = 64 UMMAs instead of 36 N = 64 (operand wavefronts -22 %).
//
// Tensor memory (512 columns): per (M half, buffer) 128 columns = [D_a | D_b].  The 1x1's accumulator D2 reuses the
// D_b columns of its own buffer once epilogue 1 has consumed them; epilogue 2 zeroes them again after reading D2, so
// that every pair UMMA accumulates (the first single of a board initialises D_a alone: a pair cannot initialise one
// half of its 128 columns and accumulate into the other).  Singles are issued first, which gives epilogue 2 of the
// board two back the time it needs before the pairs touch D_b.
template <int NSLAB>
struct GnPairsCfg {
    static constexpr int SLAB_ROWS = 290, SLAB_BYTES = 8 * SLAB_ROWS * 16;
    static constexpr int PAIR_BYTES = 8 * 128 * 16, SINGLE_BYTES = 8 * 64 * 16;
    static constexpr int W3_BYTES = 3 * PAIR_BYTES + 3 * SINGLE_BYTES, W1_BYTES = 8 * 64 * 16;
    static constexpr int SMEM = NSLAB * SLAB_BYTES + W3_BYTES + W1_BYTES + 2 * 64 * 4 + 256 + 132 * 4;
};

// two x16 loads in flight, one wait
__device__ __forceinline__ void tmem_ld16x2(uint32_t ta, uint32_t tb, float *va, float *vb) {
    uint32_t a[16], b[16];
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
        : "=r"(a[0]), "=r"(a[1]), "=r"(a[2]), "=r"(a[3]), "=r"(a[4]), "=r"(a[5]), "=r"(a[6]), "=r"(a[7]), "=r"(a[8]),
          "=r"(a[9]), "=r"(a[10]), "=r"(a[11]), "=r"(a[12]), "=r"(a[13]), "=r"(a[14]), "=r"(a[15])
        : "r"(ta) : "memory");
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
        : "=r"(b[0]), "=r"(b[1]), "=r"(b[2]), "=r"(b[3]), "=r"(b[4]), "=r"(b[5]), "=r"(b[6]), "=r"(b[7]), "=r"(b[8]),
          "=r"(b[9]), "=r"(b[10]), "=r"(b[11]), "=r"(b[12]), "=r"(b[13]), "=r"(b[14]), "=r"(b[15])
        : "r"(tb) : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;"
                 : "+r"(a[0]), "+r"(a[1]), "+r"(a[2]), "+r"(a[3]), "+r"(a[4]), "+r"(a[5]), "+r"(a[6]), "+r"(a[7]),
                   "+r"(a[8]), "+r"(a[9]), "+r"(a[10]), "+r"(a[11]), "+r"(a[12]), "+r"(a[13]), "+r"(a[14]), "+r"(a[15])
                 : : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;"
                 : "+r"(b[0]), "+r"(b[1]), "+r"(b[2]), "+r"(b[3]), "+r"(b[4]), "+r"(b[5]), "+r"(b[6]), "+r"(b[7]),
                   "+r"(b[8]), "+r"(b[9]), "+r"(b[10]), "+r"(b[11]), "+r"(b[12]), "+r"(b[13]), "+r"(b[14]), "+r"(b[15])
                 : : "memory");
#pragma unroll
    for (int i = 0; i < 16; ++i) { va[i] = __uint_as_float(a[i]); vb[i] = __uint_as_float(b[i]); }
}

// NSLAB input slab buffers (the loads come from HBM: with two, the issuers waited 200-460 cycles per board for a_full).
// rev: boards are walked from the last to the first.  Consecutive layers alternate the direction, so a layer starts on
// the boards its predecessor wrote last -- the part of the 151 MB plane buffer that is still in the 126 MB L2.
// four x16 loads in flight (32 columns of each of two accumulators), one round trip
__device__ __forceinline__ void tmem_ld32x2(uint32_t ta, uint32_t tb, float *va, float *vb) {
    uint32_t r[4][16];
    const uint32_t addr[4] = {ta, ta + 16, tb, tb + 16};
#pragma unroll
    for (int g = 0; g < 4; ++g)
        asm volatile(
            "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
            : "=r"(r[g][0]), "=r"(r[g][1]), "=r"(r[g][2]), "=r"(r[g][3]), "=r"(r[g][4]), "=r"(r[g][5]), "=r"(r[g][6]),
              "=r"(r[g][7]), "=r"(r[g][8]), "=r"(r[g][9]), "=r"(r[g][10]), "=r"(r[g][11]), "=r"(r[g][12]), "=r"(r[g][13]),
              "=r"(r[g][14]), "=r"(r[g][15])
            : "r"(addr[g]) : "memory");
#pragma unroll
    for (int g = 0; g < 4; ++g)
        asm volatile("tcgen05.wait::ld.sync.aligned;"
                     : "+r"(r[g][0]), "+r"(r[g][1]), "+r"(r[g][2]), "+r"(r[g][3]), "+r"(r[g][4]), "+r"(r[g][5]), "+r"(r[g][6]),
                       "+r"(r[g][7]), "+r"(r[g][8]), "+r"(r[g][9]), "+r"(r[g][10]), "+r"(r[g][11]), "+r"(r[g][12]),
                       "+r"(r[g][13]), "+r"(r[g][14]), "+r"(r[g][15])
                     : : "memory");
#pragma unroll
    for (int i = 0; i < 16; ++i) {
        va[i] = __uint_as_float(r[0][i]); va[16 + i] = __uint_as_float(r[1][i]);
        vb[i] = __uint_as_float(r[2][i]); vb[16 + i] = __uint_as_float(r[3][i]);
    }
}

template <bool FINAL, int NSLAB>
__global__ void __launch_bounds__(352, 1)
k_gn_block_pairs(const uint8_t *__restrict__ in_planes, const uint8_t *__restrict__ w3, const uint8_t *__restrict__ w1,
                 const __grid_constant__ GnBlockConsts kc, uint8_t *__restrict__ out_planes,
                 uint8_t *__restrict__ head_out, int G, int rev) {
    using C = GnPairsCfg<NSLAB>;
    extern __shared__ __align__(128) uint8_t smem[];
    uint8_t *slab = smem;
    uint8_t *sw3 = slab + NSLAB * C::SLAB_BYTES;      // [3 pair blocks: 8 cg x 128 rows x 16 B][3 single blocks: 8 cg x 64 x 16 B]
    uint8_t *sw1 = sw3 + C::W3_BYTES;
    uint64_t *bars = reinterpret_cast<uint64_t *>(sw1 + C::W1_BYTES + 2 * 64 * 4);
    static_assert(NSLAB >= 2 && NSLAB <= 4, "slab ring");
    uint64_t *a_full = bars, *a_empty = bars + 4, *d1_full = bars + 8, *d1_empty = bars + 10;
    uint64_t *mid_full = bars + 12, *mid_empty = bars + 14, *d2_full = bars + 16, *d2_empty = bars + 18;
    uint64_t *w_full = bars + 20;
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(bars + 21);

    const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0), lane = threadIdx.x & 31;   // provably warp-uniform
    if (threadIdx.x == 0) {
        for (int i = 0; i < NSLAB; ++i) { ptx::mbar_init(&a_full[i], 1); ptx::mbar_init(&a_empty[i], 2); }
        for (int i = 0; i < 2; ++i) {   // every tcgen05.commit comes twice: one per issuer warp (M half)
            ptx::mbar_init(&d1_full[i], 2);  ptx::mbar_init(&d1_empty[i], 8);
            ptx::mbar_init(&mid_full[i], 8); ptx::mbar_init(&mid_empty[i], 2);
            ptx::mbar_init(&d2_full[i], 2);  ptx::mbar_init(&d2_empty[i], 8);
        }
        ptx::mbar_init(w_full, 1);
        ptx::fence_mbar_init();
    }
    if (warp == 1) ptx::tmem_alloc<512>(tmem_slot);
    ptx::tc_fence_before();
    __syncthreads();
    ptx::tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    const int n_it = blockIdx.x < G ? (G - 1 - blockIdx.x) / gridDim.x + 1 : 0;
    // accumulator block of (buffer ab, M half h): 128 columns [D_a | D_b]
    auto acc_of = [&](int ab, int h) { return tmem_base + (uint32_t)(ab * 256 + h * 128); };
    auto board_of = [&](int it) { const int b = blockIdx.x + it * (int)gridDim.x; return rev ? G - 1 - b : b; };
    pdl_trigger();                 // programmatic dependent launch (gz_engine.h); warp 0 issues its weight loads first
    if (warp != 0) pdl_wait();

    if (warp == 0) {
        if (lane == 0) {
            ptx::mbar_expect_tx(w_full, C::W3_BYTES + C::W1_BYTES);
            for (int t = 0; t < 9; ++t) ptx::bulk_g2s(sw3 + t * 8192, w3 + t * 8192, 8192, w_full);
            ptx::bulk_g2s(sw1, w1, C::W1_BYTES, w_full);
            pdl_wait();
            for (int it = 0; it < n_it; ++it) {
                const int b = board_of(it), ab = it % NSLAB;
                ptx::mbar_wait(&a_empty[ab], ((it / NSLAB) & 1) ^ 1);
                ptx::mbar_expect_tx(&a_full[ab], C::SLAB_BYTES);
                const uint8_t *src = in_planes + kGuardBytes + (size_t)b * 8 * kPlaneBytes - 16;
                for (int cg = 0; cg < 8; ++cg)
                    ptx::bulk_g2s(slab + ab * C::SLAB_BYTES + cg * C::SLAB_ROWS * 16,
                                  src + (size_t)cg * kPlaneBytes, C::SLAB_ROWS * 16, &a_full[ab]);
            }
        }
    } else if (warp == 1 || warp == 10) {
        const int h = warp == 1 ? 0 : 1;
        constexpr uint32_t idesc64 = ptx::make_idesc_bf16(128, 64), idesc128 = ptx::make_idesc_bf16(128, 128);
        constexpr uint32_t LBO_A = C::SLAB_ROWS * 16, LBO_S = 64 * 16, LBO_P = 128 * 16;
        GN_T(long long tk0 = clock64(); long long w_d1e = 0, w_af = 0, t_i3 = 0, w_mf = 0, w_d2e = 0, t_i1 = 0;)
        ptx::mbar_wait(w_full, 0);
        GN_T(const long long w_w = clock64() - tk0;)
        const uint64_t dpair0 = ptx::make_desc(ptx::smem_u32(sw3), LBO_P, 128);
        const uint64_t dsing0 = ptx::make_desc(ptx::smem_u32(sw3) + 3 * C::PAIR_BYTES, LBO_S, 128);
        const uint64_t dw0 = ptx::make_desc(ptx::smem_u32(sw1), LBO_S, 128);
        auto wait_board = [&](int it) {        // accumulator D_a free (epilogue 1 of board it - 2), slab loaded
            const int ab = it & 1;
            const uint32_t par = (it >> 1) & 1;
            GN_T(long long t0 = clock64();)
            ptx::mbar_wait(&d1_empty[ab], par ^ 1);
            GN_T(long long t1 = clock64(); w_d1e += t1 - t0;)
            ptx::mbar_wait(&a_full[it % NSLAB], (it / NSLAB) & 1);
            GN_T(w_af += clock64() - t1;)
            ptx::tc_fence_after();
        };
        auto issue_1x1 = [&](int j) {          // D2 = the D_b columns of board j's own buffer
            const int jb = j & 1;
            GN_T(long long t0 = clock64();)
            ptx::mbar_wait(&mid_full[jb], (j >> 1) & 1);
            GN_T(long long t1 = clock64(); w_mf += t1 - t0;)
            ptx::tc_fence_after();
            const uint32_t acc = acc_of(jb, h);
#pragma unroll
            for (int k = 0; k < 4; ++k)
                ptx::umma_bf16_ts_elect(acc + 64, acc + 8 * k, dw0 + (uint64_t)((2 * k * (int)LBO_S) >> 4), idesc64, k != 0);
            ptx::umma_commit_elect(&mid_empty[jb]);
            ptx::umma_commit_elect(&d2_full[jb]);
            GN_T(t_i1 += clock64() - t1;)
        };
        if (n_it > 0) wait_board(0);
        for (int it = 0; it < n_it; ++it) {
            const int ab = it & 1;
            GN_T(long long t2 = clock64();)
            const uint64_t da_h = ptx::make_desc(ptx::smem_u32(slab + (it % NSLAB) * C::SLAB_BYTES), LBO_A, 128) + (uint64_t)(128 * h);
            const uint32_t acc = acc_of(ab, h);
            // singles first: the dx = +1 taps, N = 64 into D_a; the very first one initialises D_a
#pragma unroll
            for (int ty = 0; ty < 3; ++ty) {
                const int rowoff = 17 + 16 * (ty - 1) + 1;
#pragma unroll
                for (int k = 0; k < 4; ++k)
                    ptx::umma_bf16_elect(acc, da_h + (uint64_t)(2 * k * C::SLAB_ROWS + rowoff),
                                         dsing0 + (uint64_t)((ty * C::SINGLE_BYTES + 2 * k * (int)LBO_S) >> 4), idesc64,
                                         (ty | k) != 0);
            }
            // pairs: A at the offset of tap (dy, 0); columns 0-63 += tap (dy, 0), columns 64-127 += tap (dy, -1) for the
            // NEXT output row.  D_b was zeroed by epilogue 2 of board it - 2 (or at kernel start).
            GN_T(long long t3 = clock64();)
            ptx::mbar_wait(&d2_empty[ab], (it >> 1) & 1);
            GN_T(w_d2e += clock64() - t3;)
            ptx::tc_fence_after();
#define GZ_GN_PAIR(TY, K0, K1)                                                                                     \
            _Pragma("unroll") for (int k = K0; k < K1; ++k)                                                        \
                ptx::umma_bf16_elect(acc, da_h + (uint64_t)(2 * k * C::SLAB_ROWS + 17 + 16 * (TY - 1)),            \
                                     dpair0 + (uint64_t)((TY * C::PAIR_BYTES + 2 * k * (int)LBO_P) >> 4), idesc128, 1u);
            GZ_GN_PAIR(0, 0, 4)
            GZ_GN_PAIR(1, 0, 4)
            GZ_GN_PAIR(2, 0, 2)
            // late enough for epilogue 1 of board it - 1 (two TMEM reads + a shuffle per value: ~1.2 k cycles) to be
            // through: issued after pair row 1 both issuers sat ~600 cycles in this wait
            if (it > 0) issue_1x1(it - 1);
            if (it + 1 < n_it) wait_board(it + 1);
            GZ_GN_PAIR(2, 2, 4)
#undef GZ_GN_PAIR
            ptx::umma_commit_elect(&a_empty[it % NSLAB]);
            ptx::umma_commit_elect(&d1_full[ab]);
            GN_T(t_i3 += clock64() - t2;)
        }
        if (n_it > 0) issue_1x1(n_it - 1);
        GN_T(if (blockIdx.x == 0 && lane == 0 && n_it > 0)
                 printf("gn(pairs) issuer h=%d: %d boards, total %lld | per board: wait d1_empty %lld, a_full %lld, issue3x3 %lld, wait mid_full %lld, d2_empty %lld, issue1x1 %lld | weights %lld\n",
                        h, n_it, clock64() - tk0, w_d1e / n_it, w_af / n_it, t_i3 / n_it, w_mf / n_it, w_d2e / n_it, t_i1 / n_it, w_w);)
    } else {
        const int q = warp & 3, h = (warp - 2) >> 2;
        const int row = 128 * h + 32 * q + lane;            // tile row = plane position - 16
        const int pos = 16 + row;
        const bool valid = (pos & 15) != 15 && (pos >> 4) <= 15;
        const uint32_t tlane = tmem_base + ((uint32_t)(32 * q) << 16) + h * 128;
        GN_T(long long w_d1f = 0, w_me = 0, t_e1 = 0, w_d2f = 0, t_e2 = 0;)
        const uint32_t zeros[8] = {0u, 0u, 0u, 0u, 0u, 0u, 0u, 0u};
        // D_b of both buffers starts at zero (pair UMMAs only accumulate)
#pragma unroll
        for (int ab = 0; ab < 2; ++ab)
#pragma unroll
            for (int c = 0; c < 64; c += 8) ptx::tmem_st8(tlane + ab * 256 + 64 + c, zeros);
        ptx::tmem_wait_st();
        ptx::tc_fence_before();
        __syncwarp();
        if (lane == 0) { ptx::mbar_arrive(&d2_empty[0]); ptx::mbar_arrive(&d2_empty[1]); }
        for (int it = 0; it <= n_it; ++it) {
            if (it > 0) {   // epilogue 2 of board it - 1 (its 1x1 ran in the middle of board it's 3x3)
                const int j = it - 1, jb = j & 1;
                const int b = board_of(j);
                GN_T(long long t0 = clock64();)
                ptx::mbar_wait(&d2_full[jb], (j >> 1) & 1);
                GN_T(long long t1 = clock64(); w_d2f += t1 - t0;)
                ptx::tc_fence_after();
                const uint32_t td2 = tlane + jb * 256 + 64;
                if (FINAL) {
                    float p0 = kc.head_b[0], p1 = kc.head_b[1];
#pragma unroll
                    for (int c0 = 0; c0 < 64; c0 += 16) {
                        float v[16];
                        ptx::tmem_ld16(td2 + c0, v);
#pragma unroll
                        for (int i = 0; i < 16; ++i) {
                            const float x = fmaxf(v[i] + kc.b1[c0 + i], 0.f);
                            p0 = fmaf(kc.head_w[2 * (c0 + i)], x, p0);
                            p1 = fmaf(kc.head_w[2 * (c0 + i) + 1], x, p1);
                        }
                    }
                    if (valid) {
                        const int cell = ((pos >> 4) - 1) * 15 + (pos & 15);
                        store_head_operand(head_out, b, cell, kHeadChunksP, p0);
                        store_head_operand(head_out, b, kCells + cell, kHeadChunksP, p1);
                    }
                } else {
                    uint8_t *dst = out_planes + kGuardBytes + (size_t)b * 8 * kPlaneBytes + (size_t)pos * 16;
#pragma unroll
                    for (int c0 = 0; c0 < 64; c0 += 16) {
                        float v[16];
                        ptx::tmem_ld16(td2 + c0, v);
                        uint32_t pk[8];
#pragma unroll
                        for (int i = 0; i < 8; ++i)
                            pk[i] = valid ? ptx::pack_bf16(fmaxf(v[2 * i] + kc.b1[c0 + 2 * i], 0.f),
                                                           fmaxf(v[2 * i + 1] + kc.b1[c0 + 2 * i + 1], 0.f)) : 0u;
                        *reinterpret_cast<uint4 *>(dst + (size_t)(c0 / 8) * kPlaneBytes) = make_uint4(pk[0], pk[1], pk[2], pk[3]);
                        *reinterpret_cast<uint4 *>(dst + (size_t)(c0 / 8 + 1) * kPlaneBytes) = make_uint4(pk[4], pk[5], pk[6], pk[7]);
                    }
                }
                // hand the columns back as a zeroed D_b
#pragma unroll
                for (int c = 0; c < 64; c += 8) ptx::tmem_st8(td2 + c, zeros);
                ptx::tmem_wait_st();
                ptx::tc_fence_before();
                __syncwarp();
                if (lane == 0) ptx::mbar_arrive(&d2_empty[jb]);
                GN_T(t_e2 += clock64() - t1;)
            }
            if (it < n_it) {   // epilogue 1 of board it: out[m] = D_a[m] + D_b[m - 1], bias, ReLU, bf16, in place
                const int ab = it & 1;
                const uint32_t par = (it >> 1) & 1;
                GN_T(long long t0 = clock64();)
                ptx::mbar_wait(&d1_full[ab], par);
                GN_T(long long t1 = clock64(); w_d1f += t1 - t0;)
                ptx::mbar_wait(&mid_empty[ab], par ^ 1);
                GN_T(long long t2 = clock64(); w_me += t2 - t1;)
                ptx::tc_fence_after();
                const uint32_t td1 = tlane + ab * 256;
#pragma unroll
                for (int c0 = 0; c0 < 64; c0 += 16) {
                    float va[16], vb[16];
#if defined(GZ_EXP_EPI1) && GZ_EXP_EPI1 == 2      // perf experiment (wrong results): no D_b read at all
                    ptx::tmem_ld16(td1 + c0, va);
                    for (int i = 0; i < 16; ++i) vb[i] = 0.f;
#else
                    tmem_ld16x2(td1 + c0, td1 + 64 + c0, va, vb);
#endif
                    uint32_t pk[8];
#pragma unroll
                    for (int i = 0; i < 16; ++i) {
#if defined(GZ_EXP_EPI1)                          // perf experiment (wrong results): no shuffles
                        float up = vb[i];
#else
                        float up = __shfl_up_sync(0xffffffffu, vb[i], 1);
#endif
                        if (lane == 0) up = 0.f;     // board column 0: the row above in the tile is the zero padding column
                        va[i] = fmaxf(va[i] + up + kc.b3[c0 + i], 0.f);
                    }
#pragma unroll
                    for (int i = 0; i < 8; ++i) pk[i] = ptx::pack_bf16(va[2 * i], va[2 * i + 1]);
                    ptx::tmem_st8(td1 + c0 / 2, pk);
                }
                ptx::tmem_wait_st();
                ptx::tc_fence_before();
                __syncwarp();
                if (lane == 0) { ptx::mbar_arrive(&d1_empty[ab]); ptx::mbar_arrive(&mid_full[ab]); }
                GN_T(t_e1 += clock64() - t2;)
            }
        }
        GN_T(if (blockIdx.x == 0 && lane == 0 && (warp == 2 || warp == 9) && n_it > 0)
                 printf("gn(pairs) epilogue warp %d: per board: wait d1_full %lld, mid_empty %lld, epi1 %lld, wait d2_full %lld, epi2 %lld\n",
                        warp, w_d1f / n_it, w_me / n_it, t_e1 / n_it, w_d2f / n_it, t_e2 / n_it);)
    }
    ptx::tc_fence_before();
    __syncthreads();
    if (warp == 1) ptx::tmem_dealloc<512>(tmem_base);
}

// ------------------------------------------------------------------------------------------------
// im2col one-hot tile of an input convolution (3 planes x 3 x 3 = 27 inputs per position, K padded to 32) for
// k_gn_block0_embed.  Tile = [4 k-groups][256 rows][8 k] bf16, row = board bit index 16 * row +
// col (= plane position - 16), K index k = 3 * (3 * plane + ky) + kx (pack_conv0_im2col_bf16 packs the weights in the
// same order).  A warp builds one M half (rows [128 HH, 128 HH + 128)); every lane holds the whole position.
struct I2cPlanes { uint32_t X[3][11]; };   // X[c][1 + j] = word j of plane c (black, white, empty); X[c][0, 9, 10] = 0
__device__ __forceinline__ void im2col_planes(const uint4 (&q)[4], I2cPlanes &P) {
    const uint32_t bw[8] = {q[0].x, q[0].y, q[0].z, q[0].w, q[1].x, q[1].y, q[1].z, q[1].w};
    const uint32_t ww[8] = {q[2].x, q[2].y, q[2].z, q[2].w, q[3].x, q[3].y, q[3].z, q[3].w};
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        P.X[0][1 + j] = bw[j];
        P.X[1][1 + j] = ww[j];
        P.X[2][1 + j] = ~(bw[j] | ww[j]) & (j == 7 ? 0x00007FFFu : 0x7FFF7FFFu);
    }
#pragma unroll
    for (int c = 0; c < 3; ++c) P.X[c][0] = P.X[c][9] = P.X[c][10] = 0u;
}
template <int HH>
__device__ __forceinline__ void im2col_write_half(const I2cPlanes &P, uint8_t *dst, int lane) {
#pragma unroll
    for (int kq = 0; kq < 4; ++kq) {
        constexpr int kbase = 4 * HH;
        const int kk = kbase + kq;              // tile rows 32 kk + lane
        // t[3 c + ky]: the three cells (dx = -1, 0, +1) of plane c in board row + ky - 1 around this lane's cell = bits
        // lane .. lane + 2 of the plane's 64-bit window that starts at bit 32 kk + 16 (ky - 1) - 1
        uint32_t t[9];
#pragma unroll
        for (int c = 0; c < 3; ++c)
#pragma unroll
            for (int ky = 0; ky < 3; ++ky) {
                const int s = 32 * kk + 16 * (ky - 1) - 1 + 32;      // + 32: index into X[c][] (one zero word in front)
                const int j0 = s >> 5, sh = s & 31;
                const uint32_t lo = __funnelshift_r(P.X[c][j0], P.X[c][j0 + 1], sh);
                const uint32_t hi = __funnelshift_r(P.X[c][j0 + 1], P.X[c][j0 + 2], sh);
                t[3 * c + ky] = __funnelshift_r(lo, hi, lane);
            }
        // bf16 1.0 = 0x3F80: a set bit j of t becomes (t & (1 << j)) * (0x3F80 >> j)
        const int m = 32 * kk + lane;
#pragma unroll
        for (int cg = 0; cg < 4; ++cg) {
            uint32_t pk[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const int k0 = 8 * cg + 2 * i, k1 = k0 + 1;
                uint32_t w = 0u;
                if (k0 < 27) w = (t[k0 / 3] & (1u << (k0 % 3))) * (0x3F80u >> (k0 % 3));
                if (k1 < 27) w += (t[k1 / 3] & (1u << (k1 % 3))) * (0x3F800000u >> (k1 % 3));
                pk[i] = w;
            }
            *reinterpret_cast<uint4 *>(dst + ((size_t)cg * 256 + m) * 16) = make_uint4(pk[0], pk[1], pk[2], pk[3]);
        }
    }
}

// ------------------------------------------------------------------------------------------------
// k_gn_block0_embed: GraphNet's embed convolution (3 -> 64, 3x3, ReLU; bg_planner.py:32,51) fused into the first block.
//
// As a launch of its own the embed was 35 us of HBM writes per planner step (151 MB out, read back by block 0).  Here
// it is ONE K = 32 GEMM per board in front of block 0's 3x3: the producer warps expand the bitboards into an im2col
// one-hot tile (row = position, k = 3 * tap + plane, 27 of 32 used), two N = 64 UMMAs per M half put the result into
// tensor memory, eight extra epilogue warps apply bias + ReLU and write bf16 straight into the 3x3's input slab in
// shared memory -- the slab that k_gn_block_pairs fetches from HBM.  Everything after that is k_gn_block_pairs.
//
// Tensor memory has no spare columns (2 buffers x 2 halves x [D_a | D_b]), so the embed result of board m lives in
// D_b of its OWN buffer (m & 1) during the one window in which those columns are idle: after epilogue 1 of board m - 2
// has consumed the pair partials and before the 1x1 of board m - 2 overwrites them with its accumulator.  In the
// issuer's instruction stream of board j:
//     wait epilogue 0 of j | 1x1(j - 2) | singles(j) | embed(j + 1) | pair rows 0, 1, 2
// The embeds of boards 0 and 1 run in the prologue (their D_b columns are zeroed again by epilogue 0, which the pair
// UMMAs rely on).
//   warps 0, 19   im2col producers (M half 0 / 1); warp 0 also loads the weights
//   warps 1, 10   tcgen05.mma issuers
//   warps 2-9     epilogue 1 (3x3 accumulators -> bf16 A operand of the 1x1) and epilogue 0 (embed -> slab)
//   warps 11-18   epilogue 2 (1x1 -> planes): quadrant = warp & 3, M half = (warp - 11) >> 2
// Measured (profiles/r02_gn_block0_embed.md): 84 us against 35 + 66 us for the two launches it replaces.  Per board the
// epilogues now read 256 KB from tensor memory (64 embed + 128 pair partials + 64 1x1) against 192 KB in
// k_gn_block_pairs, and both kernels run at ~56 B of tcgen05.ld per clock: tensor-memory READ bandwidth, not the
// tensor pipe, is what paces them.
struct GnEmbedCfg {
    static constexpr int SLAB_ROWS = 290, SLAB_BYTES = 8 * SLAB_ROWS * 16;
    static constexpr int PAIR_BYTES = 8 * 128 * 16, SINGLE_BYTES = 8 * 64 * 16;
    static constexpr int W3_BYTES = 3 * PAIR_BYTES + 3 * SINGLE_BYTES, W1_BYTES = 8 * 64 * 16;
    static constexpr int I2C_BYTES = 4 * 256 * 16;     // [4 k-groups][256 rows][8 k] bf16
    static constexpr int WE_BYTES = 4 * 64 * 16;       // [4 k-groups][64 cout][8 k]
    static constexpr int THREADS = 640;
    static constexpr int SMEM = 2 * SLAB_BYTES + W3_BYTES + W1_BYTES + 2 * I2C_BYTES + WE_BYTES + 512;
};

__global__ void __launch_bounds__(GnEmbedCfg::THREADS, 1)
k_gn_block0_embed(const gz_board *__restrict__ boards, const uint8_t *__restrict__ we, const uint8_t *__restrict__ w3,
                  const uint8_t *__restrict__ w1, const __grid_constant__ GnBlockConsts kc,
                  uint8_t *__restrict__ out_planes, int G, int rev) {
    using C = GnEmbedCfg;
    extern __shared__ __align__(128) uint8_t smem[];
    uint8_t *slab = smem;                               // [2][SLAB_BYTES]
    uint8_t *sw3 = slab + 2 * C::SLAB_BYTES;
    uint8_t *sw1 = sw3 + C::W3_BYTES;
    uint8_t *i2c = sw1 + C::W1_BYTES;                   // [2][I2C_BYTES]
    uint8_t *swe = i2c + 2 * C::I2C_BYTES;
    uint64_t *bars = reinterpret_cast<uint64_t *>(swe + C::WE_BYTES);
    uint64_t *a_full = bars, *a_empty = bars + 2, *d1_full = bars + 4, *d1_empty = bars + 6;
    uint64_t *mid_full = bars + 8, *mid_empty = bars + 10, *d2_full = bars + 12, *d2_empty = bars + 14;
    uint64_t *e_full = bars + 16, *e_empty = bars + 18, *de_full = bars + 20;
    uint64_t *w_full = bars + 22;
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(bars + 23);

    const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0), lane = threadIdx.x & 31;   // provably warp-uniform
    if (threadIdx.x == 0) {
        for (int i = 0; i < 2; ++i) {   // every tcgen05.commit comes twice: one per issuer warp (M half)
            ptx::mbar_init(&a_full[i], 8);   ptx::mbar_init(&a_empty[i], 2);     // slab: written by the 8 epilogue-0 warps
            ptx::mbar_init(&d1_full[i], 2);  ptx::mbar_init(&d1_empty[i], 8);
            ptx::mbar_init(&mid_full[i], 8); ptx::mbar_init(&mid_empty[i], 2);
            ptx::mbar_init(&d2_full[i], 2);  ptx::mbar_init(&d2_empty[i], 8);
            ptx::mbar_init(&e_full[i], 2);   ptx::mbar_init(&e_empty[i], 2);     // im2col tile: two producer warps
            ptx::mbar_init(&de_full[i], 2);
        }
        ptx::mbar_init(w_full, 1);
        ptx::fence_mbar_init();
    }
    // the slab rows outside the board (and everything the epilogue never writes) stay zero for the whole kernel
    for (int i = threadIdx.x; i < 2 * C::SLAB_BYTES / 16; i += C::THREADS)
        reinterpret_cast<uint4 *>(slab)[i] = make_uint4(0u, 0u, 0u, 0u);
    ptx::fence_proxy_async();
    if (warp == 1) ptx::tmem_alloc<512>(tmem_slot);
    ptx::tc_fence_before();
    __syncthreads();
    ptx::tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    const int n_it = blockIdx.x < G ? (G - 1 - blockIdx.x) / gridDim.x + 1 : 0;
    auto acc_of = [&](int ab, int h) { return tmem_base + (uint32_t)(ab * 256 + h * 128); };   // [D_a | D_b]
    auto board_of = [&](int it) { const int b = blockIdx.x + it * (int)gridDim.x; return rev ? G - 1 - b : b; };
    pdl_trigger();                 // programmatic dependent launch (gz_engine.h); warp 0 issues its weight loads first
    if (warp != 0) pdl_wait();

    if (warp == 0 || warp == 19) {
        // ================= im2col producers: M half hh = tile rows [128 hh, 128 hh + 128) =================
        const int hh = warp == 0 ? 0 : 1;
        if (warp == 0 && lane == 0) {
            ptx::mbar_expect_tx(w_full, C::W3_BYTES + C::W1_BYTES + C::WE_BYTES);
            for (int t = 0; t < 9; ++t) ptx::bulk_g2s(sw3 + t * 8192, w3 + t * 8192, 8192, w_full);
            ptx::bulk_g2s(sw1, w1, C::W1_BYTES, w_full);
            ptx::bulk_g2s(swe, we, C::WE_BYTES, w_full);
        }
        if (warp == 0) pdl_wait();
        // every lane holds the whole position: 8 words per colour (one broadcast load each), fetched one board ahead
        auto load_words = [&](int it, uint4 *q) {
            if (it < n_it) {
                const uint4 *src = reinterpret_cast<const uint4 *>(boards[board_of(it)].stones);
#pragma unroll
                for (int i = 0; i < 4; ++i) q[i] = src[i];
            }
        };
        uint4 q[4];
        load_words(0, q);
        auto produce = [&](auto HH) {       // the M half as a compile-time constant: X[][] stays in registers
            constexpr int hh_c = decltype(HH)::value;
            GN_T(long long p_wait = 0, p_work = 0;)
            for (int it = 0; it < n_it; ++it) {
                const int eb = it & 1;
                I2cPlanes P;
                im2col_planes(q, P);
                load_words(it + 1, q);
                GN_T(const long long tp0 = clock64();)
                if (lane == 0) ptx::mbar_wait(&e_empty[eb], ((it >> 1) & 1) ^ 1);
                __syncwarp();
                GN_T(const long long tp1 = clock64(); p_wait += tp1 - tp0;)
                im2col_write_half<hh_c>(P, i2c + eb * C::I2C_BYTES, lane);
                ptx::fence_proxy_async();
                __syncwarp();
                if (lane == 0) ptx::mbar_arrive(&e_full[eb]);
                GN_T(p_work += clock64() - tp1;)
            }
            GN_T(if (blockIdx.x == 0 && lane == 0 && n_it > 0)
                     printf("gn0 producer %d: per board: wait e_empty %lld, build %lld\n", hh_c, p_wait / n_it, p_work / n_it);)
        };
        if (hh == 0) produce(std::integral_constant<int, 0>()); else produce(std::integral_constant<int, 1>());
    } else if (warp == 1 || warp == 10) {
        // ================= tcgen05.mma issuers, one per M half =================
        const int h = warp == 1 ? 0 : 1;
        constexpr uint32_t idesc64 = ptx::make_idesc_bf16(128, 64), idesc128 = ptx::make_idesc_bf16(128, 128);
        constexpr uint32_t LBO_A = C::SLAB_ROWS * 16, LBO_S = 64 * 16, LBO_P = 128 * 16, LBO_E = 256 * 16;
        GN_T(const long long tk0 = clock64(); long long w_ef = 0, w_mfe = 0, w_af = 0, w_d2e = 0, w_d1e = 0, w_mf1 = 0, t_loop = 0;)
        ptx::mbar_wait(w_full, 0);
        const uint64_t dpair0 = ptx::make_desc(ptx::smem_u32(sw3), LBO_P, 128);
        const uint64_t dsing0 = ptx::make_desc(ptx::smem_u32(sw3) + 3 * C::PAIR_BYTES, LBO_S, 128);
        const uint64_t dw0 = ptx::make_desc(ptx::smem_u32(sw1), LBO_S, 128);
        const uint64_t dwe0 = ptx::make_desc(ptx::smem_u32(swe), LBO_S, 128);
        // embed of board m: D_b of buffer m & 1 = im2col tile (K = 32: two K steps) x embed weights
        auto issue_embed = [&](int m) {
            const int eb = m & 1;
            GN_T(const long long t0 = clock64();)
            ptx::mbar_wait(&e_full[eb], (m >> 1) & 1);
            GN_T(w_ef += clock64() - t0;)
            ptx::tc_fence_after();
            const uint64_t de = ptx::make_desc(ptx::smem_u32(i2c + eb * C::I2C_BYTES), LBO_E, 128) + (uint64_t)(128 * h);
            const uint32_t acc = acc_of(eb, h) + 64;
#pragma unroll
            for (int ks = 0; ks < 2; ++ks)
                ptx::umma_bf16_elect(acc, de + (uint64_t)(2 * ks * 256), dwe0 + (uint64_t)((2 * ks * (int)LBO_S) >> 4), idesc64,
                                     ks != 0);
            ptx::umma_commit_elect(&e_empty[eb]);
            ptx::umma_commit_elect(&de_full[eb]);
        };
        auto issue_1x1 = [&](int j) {          // D2 = the D_b columns of board j's own buffer
            const int jb = j & 1;
            GN_T(const long long t0 = clock64();)
            ptx::mbar_wait(&mid_full[jb], (j >> 1) & 1);
            GN_T(w_mf1 += clock64() - t0;)
            ptx::tc_fence_after();
            const uint32_t acc = acc_of(jb, h);
#pragma unroll
            for (int k = 0; k < 4; ++k)
                ptx::umma_bf16_ts_elect(acc + 64, acc + 8 * k, dw0 + (uint64_t)((2 * k * (int)LBO_S) >> 4), idesc64, k != 0);
            ptx::umma_commit_elect(&mid_empty[jb]);
            ptx::umma_commit_elect(&d2_full[jb]);
        };
        // prologue: the embeds of boards 0 and 1 (their D_b columns were zeroed by the epilogue-2 warps: d2_empty, first phase)
        for (int m = 0; m < 2 && m < n_it; ++m) {
            ptx::mbar_wait(&d2_empty[m], 0);
            ptx::tc_fence_after();
            issue_embed(m);
        }
        for (int it = 0; it < n_it; ++it) {
            const int ab = it & 1;
            GN_T(const long long tl0 = clock64();)
            // slab of board it written by epilogue 0, which has thereby also left the D_b columns of this buffer: the
            // 1x1 of board it - 2 may put its accumulator there.  It goes first so that everything behind it (epilogue 2,
            // which hands D_b back zeroed for this board's pair UMMAs) starts as early as possible; the singles below
            // overwrite its A operand (D_a of this buffer) only after it, the tensor pipe executes in order.
            GN_T(const long long t5 = clock64();)
            ptx::mbar_wait(&a_full[ab], (it >> 1) & 1);
            GN_T(w_af += clock64() - t5;)
            ptx::tc_fence_after();
            if (it >= 2) issue_1x1(it - 2);
            GN_T(const long long t6 = clock64();)
            ptx::mbar_wait(&d1_empty[ab], ((it >> 1) & 1) ^ 1);      // D_a free: epilogue 1 of board it - 2
            GN_T(w_d1e += clock64() - t6;)
            ptx::tc_fence_after();
            const uint64_t da_h = ptx::make_desc(ptx::smem_u32(slab + ab * C::SLAB_BYTES), LBO_A, 128) + (uint64_t)(128 * h);
            const uint32_t acc = acc_of(ab, h);
            // singles first: the dx = +1 taps, N = 64 into D_a; the very first one initialises D_a
#pragma unroll
            for (int ty = 0; ty < 3; ++ty) {
                const int rowoff = 17 + 16 * (ty - 1) + 1;
#pragma unroll
                for (int k = 0; k < 4; ++k)
                    ptx::umma_bf16_elect(acc, da_h + (uint64_t)(2 * k * C::SLAB_ROWS + rowoff),
                                         dsing0 + (uint64_t)((ty * C::SINGLE_BYTES + 2 * k * (int)LBO_S) >> 4), idesc64,
                                         (ty | k) != 0);
            }
            if (it >= 1 && it + 1 < n_it) {
                // embed(it + 1) into D_b of the other buffer: free once epilogue 1 of board it - 1 has read the pair
                // partials (mid_full) and until the 1x1 of board it - 1 (top of the next iteration) writes there
                GN_T(const long long t4 = clock64();)
                ptx::mbar_wait(&mid_full[ab ^ 1], ((it - 1) >> 1) & 1);
                GN_T(w_mfe += clock64() - t4;)
                ptx::tc_fence_after();
                issue_embed(it + 1);
            }
            // pairs accumulate into [D_a | D_b]; D_b was zeroed by epilogue 2 of board it - 2 (epilogue 0 for it < 2)
            GN_T(const long long t3 = clock64();)
            ptx::mbar_wait(&d2_empty[ab], (it >> 1) & 1);
            GN_T(w_d2e += clock64() - t3;)
            ptx::tc_fence_after();
#define GZ_GN_PAIR(TY, K0, K1)                                                                                     \
            _Pragma("unroll") for (int k = K0; k < K1; ++k)                                                        \
                ptx::umma_bf16_elect(acc, da_h + (uint64_t)(2 * k * C::SLAB_ROWS + 17 + 16 * (TY - 1)),            \
                                     dpair0 + (uint64_t)((TY * C::PAIR_BYTES + 2 * k * (int)LBO_P) >> 4), idesc128, 1u);
            GZ_GN_PAIR(0, 0, 4)
            GZ_GN_PAIR(1, 0, 4)
            GZ_GN_PAIR(2, 0, 4)
#undef GZ_GN_PAIR
            ptx::umma_commit_elect(&a_empty[ab]);
            ptx::umma_commit_elect(&d1_full[ab]);
            GN_T(t_loop += clock64() - tl0;)
        }
        if (n_it >= 2) issue_1x1(n_it - 2);
        if (n_it >= 1) issue_1x1(n_it - 1);
        GN_T(if (blockIdx.x == 0 && lane == 0 && n_it > 0)
                 printf("gn0 issuer h=%d: %d boards, total %lld | per board: loop %lld, wait e_full %lld, mid_full(embed) %lld, a_full %lld, d2_empty %lld, d1_empty %lld, mid_full(1x1) %lld\n",
                        h, n_it, clock64() - tk0, t_loop / n_it, w_ef / n_it, w_mfe / n_it, w_af / n_it, w_d2e / n_it, w_d1e / n_it, w_mf1 / n_it);)
    } else if (warp >= 11) {
        // ================= epilogue 2 (1x1 -> planes); quadrant = warp & 3 =================
        // Warps of its own: behind epilogue 2 on the same warps, epilogue 1 -- which gates both the embed's tensor-memory
        // window and the 1x1 -- finished ~1000 cycles later, a stall of both issuers on every board.
        const int q = warp & 3, h0 = (warp - 11) >> 2;
        const int row = 128 * h0 + 32 * q + lane;           // tile row = plane position - 16
        const int pos = 16 + row;
        const bool valid = (pos & 15) != 15 && (pos >> 4) <= 15;
        const uint32_t tlane = tmem_base + ((uint32_t)(32 * q) << 16) + h0 * 128 + 64;      // D_b of buffer 0
        const uint32_t zeros[8] = {0u, 0u, 0u, 0u, 0u, 0u, 0u, 0u};
        GN_T(long long e2_w = 0, e2_work = 0;)
        // D_b of both buffers starts at zero (pair UMMAs only accumulate)
#pragma unroll
        for (int ab = 0; ab < 2; ++ab)
#pragma unroll
            for (int c = 0; c < 64; c += 8) ptx::tmem_st8(tlane + ab * 256 + c, zeros);
        ptx::tmem_wait_st();
        ptx::tc_fence_before();
        __syncwarp();
        if (lane == 0) { ptx::mbar_arrive(&d2_empty[0]); ptx::mbar_arrive(&d2_empty[1]); }
        auto epilogue2 = [&](int j) {
            const int jb = j & 1;
            const int b = board_of(j);
            GN_T(const long long t0 = clock64();)
            ptx::mbar_wait(&d2_full[jb], (j >> 1) & 1);
            GN_T(const long long t1 = clock64(); e2_w += t1 - t0;)
            ptx::tc_fence_after();
            const uint32_t td2 = tlane + jb * 256;
            uint8_t *dst = out_planes + kGuardBytes + (size_t)b * 8 * kPlaneBytes + (size_t)pos * 16;
            float v[64];
            ptx::tmem_ld64(td2, v);
            // hand the columns back as a zeroed D_b before anything else: the pair UMMAs of board j + 2 wait for them
#pragma unroll
            for (int c = 0; c < 64; c += 8) ptx::tmem_st8(td2 + c, zeros);
            ptx::tmem_wait_st();
            ptx::tc_fence_before();
            __syncwarp();
            if (lane == 0) ptx::mbar_arrive(&d2_empty[jb]);
#pragma unroll
            for (int c0 = 0; c0 < 64; c0 += 8) {
                uint32_t pk[4];
#pragma unroll
                for (int i = 0; i < 4; ++i)
                    pk[i] = valid ? ptx::pack_bf16(fmaxf(v[c0 + 2 * i] + kc.b1[c0 + 2 * i], 0.f),
                                                   fmaxf(v[c0 + 2 * i + 1] + kc.b1[c0 + 2 * i + 1], 0.f)) : 0u;
                *reinterpret_cast<uint4 *>(dst + (size_t)(c0 / 8) * kPlaneBytes) = make_uint4(pk[0], pk[1], pk[2], pk[3]);
            }
            GN_T(e2_work += clock64() - t1;)
        };
        for (int j = 0; j < n_it; ++j) epilogue2(j);
        GN_T(if (blockIdx.x == 0 && lane == 0 && (warp == 11 || warp == 18) && n_it > 0)
                 printf("gn0 epilogue-2 warp %d: per board: wait d2_full %lld, epi2 %lld\n", warp, e2_w / n_it, e2_work / n_it);)
    } else {
        // ================= warps 2-9: epilogue 1 (out[m] = D_a[m] + D_b[m - 1], bias, ReLU, bf16, in place) and epilogue 0
        // (embed -> slab).  The two alternate in time -- epilogue 1 of board it, then epilogue 0 of board it + 2, whose
        // embed the issuers launch the moment epilogue 1 signals -- so they never compete for tensor-memory reads.
        const int q = warp & 3, h = (warp - 2) >> 2;
        const int pos = 16 + 128 * h + 32 * q + lane;
        const bool valid = (pos & 15) != 15 && (pos >> 4) <= 15;
        const uint32_t zeros[8] = {0u, 0u, 0u, 0u, 0u, 0u, 0u, 0u};
        const uint32_t tlane = tmem_base + ((uint32_t)(32 * q) << 16) + h * 128;
        GN_T(long long w_d1f = 0, w_me = 0, t_e1 = 0, e_w1 = 0, e_w2 = 0, e_work = 0;)
        auto epilogue0 = [&](int m) {
            const int eb = m & 1;
            GN_T(const long long t0 = clock64();)
            ptx::mbar_wait(&de_full[eb], (m >> 1) & 1);
            GN_T(const long long t1 = clock64(); e_w1 += t1 - t0;)
            ptx::mbar_wait(&a_empty[eb], ((m >> 1) & 1) ^ 1);          // the 3x3 of board m - 2 has read this slab
            GN_T(const long long t2 = clock64(); e_w2 += t2 - t1;)
            ptx::tc_fence_after();
            const uint32_t td = tlane + eb * 256 + 64;             // D_b
            uint8_t *dst = slab + eb * C::SLAB_BYTES + (size_t)(pos + 1) * 16;
            float v[64];
            ptx::tmem_ld64(td, v);
#pragma unroll
            for (int c0 = 0; c0 < 64; c0 += 8) {
                uint32_t pk[4];
#pragma unroll
                for (int i = 0; i < 4; ++i)
                    pk[i] = valid ? ptx::pack_bf16(fmaxf(v[c0 + 2 * i] + kc.be[c0 + 2 * i], 0.f),
                                                   fmaxf(v[c0 + 2 * i + 1] + kc.be[c0 + 2 * i + 1], 0.f)) : 0u;
                *reinterpret_cast<uint4 *>(dst + (size_t)(c0 / 8) * C::SLAB_ROWS * 16) = make_uint4(pk[0], pk[1], pk[2], pk[3]);
            }
            if (m < 2) {   // no 1x1 / epilogue 2 has passed over these columns yet: hand them to the pair UMMAs as zeros
#pragma unroll
                for (int c = 0; c < 64; c += 8) ptx::tmem_st8(td + c, zeros);
                ptx::tmem_wait_st();
            }
            ptx::fence_proxy_async();
            ptx::tc_fence_before();
            __syncwarp();
            if (lane == 0) ptx::mbar_arrive(&a_full[eb]);
            GN_T(e_work += clock64() - t2;)
        };
        if (n_it > 0) epilogue0(0);
        if (n_it > 1) epilogue0(1);
        for (int it = 0; it < n_it; ++it) {
            const int ab = it & 1;
            const uint32_t par = (it >> 1) & 1;
            GN_T(const long long t0 = clock64();)
            ptx::mbar_wait(&d1_full[ab], par);
            GN_T(const long long t1 = clock64(); w_d1f += t1 - t0;)
            ptx::mbar_wait(&mid_empty[ab], par ^ 1);
            GN_T(const long long t2 = clock64(); w_me += t2 - t1;)
            ptx::tc_fence_after();
            const uint32_t td1 = tlane + ab * 256;
            // two round trips of 32 + 32 columns (four were 1950 cycles per board, and this epilogue is on the critical chain
            // epilogue 1 -> embed -> epilogue 0 -> next board's slab)
#pragma unroll
            for (int c0 = 0; c0 < 64; c0 += 32) {
                float va[32], vb[32];
                tmem_ld32x2(td1 + c0, td1 + 64 + c0, va, vb);
#pragma unroll
                for (int i = 0; i < 32; ++i) {
                    float up = __shfl_up_sync(0xffffffffu, vb[i], 1);
                    if (lane == 0) up = 0.f;     // board column 0: the row above in the tile is the zero padding column
                    va[i] = fmaxf(va[i] + up + kc.b3[c0 + i], 0.f);
                }
#pragma unroll
                for (int g = 0; g < 2; ++g) {
                    uint32_t pk[8];
#pragma unroll
                    for (int i = 0; i < 8; ++i) pk[i] = ptx::pack_bf16(va[16 * g + 2 * i], va[16 * g + 2 * i + 1]);
                    ptx::tmem_st8(td1 + c0 / 2 + 8 * g, pk);
                }
            }
            ptx::tmem_wait_st();
            ptx::tc_fence_before();
            __syncwarp();
            if (lane == 0) { ptx::mbar_arrive(&d1_empty[ab]); ptx::mbar_arrive(&mid_full[ab]); }
            GN_T(t_e1 += clock64() - t2;)
            if (it + 2 < n_it) epilogue0(it + 2);
        }
        GN_T(if (blockIdx.x == 0 && lane == 0 && (warp == 2 || warp == 9) && n_it > 0)
                 printf("gn0 epilogue-1/0 warp %d: per board: wait d1_full %lld, mid_empty %lld, epi1 %lld | wait de_full %lld, a_empty %lld, epi0 %lld\n",
                        warp, w_d1f / n_it, w_me / n_it, t_e1 / n_it, e_w1 / n_it, e_w2 / n_it, e_work / n_it);)
    }
    ptx::tc_fence_before();
    __syncthreads();
    if (warp == 1) ptx::tmem_dealloc<512>(tmem_base);
}

static void launch_gn_block(Engine &e, const uint8_t *in, int block, uint8_t *out, uint8_t *head_out, int G,
                            cudaStream_t s, int direction) {
    static bool configured = false;
    if (!configured) {
        cudaFuncSetAttribute(k_gn_block<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, GnBlockCfg::SMEM);
        cudaFuncSetAttribute(k_gn_block<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, GnBlockCfg::SMEM);
        cudaFuncSetAttribute(k_gn_block_pairs<false, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, GnPairsCfg<2>::SMEM);
        cudaFuncSetAttribute(k_gn_block_pairs<true, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, GnPairsCfg<2>::SMEM);
        cudaFuncSetAttribute(k_gn_block_pairs<false, 3>, cudaFuncAttributeMaxDynamicSharedMemorySize, GnPairsCfg<3>::SMEM);
        cudaFuncSetAttribute(k_gn_block_pairs<true, 3>, cudaFuncAttributeMaxDynamicSharedMemorySize, GnPairsCfg<3>::SMEM);
        configured = true;
    }
    GnBlockConsts kc = {};
    memcpy(kc.b3, e.bw.gn_host_bias[block], sizeof(float) * 128);   // b3 then b1
    memcpy(kc.head_w, e.bw.gn_host_head, sizeof(float) * 130);       // head_w then head_b
    const uint8_t *w1 = e.bw.gn_l1_w[block];
    int grid = G < e.sm_count ? G : e.sm_count;
    static const bool pairs = !(getenv("GZ_GN_PAIRS") && atoi(getenv("GZ_GN_PAIRS")) == 0);   // A/B switch for profiling
    static const int nslab = getenv("GZ_GN_SLABS") ? atoi(getenv("GZ_GN_SLABS")) : 3;                // A/B switches
    static const bool alternate = !(getenv("GZ_GN_ALTERNATE") && atoi(getenv("GZ_GN_ALTERNATE")) == 0);
    const int rev = alternate ? direction : 0;   // the layer before walked the other way
    if (pairs) {
        const uint8_t *w3 = e.bw.gn_l3p_w[block];
        if (nslab == 3) {
            if (head_out) launch_pdl(k_gn_block_pairs<true, 3>, dim3(grid), dim3(352), GnPairsCfg<3>::SMEM, s, in, w3, w1, kc, out, head_out, G, rev);
            else launch_pdl(k_gn_block_pairs<false, 3>, dim3(grid), dim3(352), GnPairsCfg<3>::SMEM, s, in, w3, w1, kc, out, (uint8_t *)nullptr, G, rev);
        } else {
            if (head_out) launch_pdl(k_gn_block_pairs<true, 2>, dim3(grid), dim3(352), GnPairsCfg<2>::SMEM, s, in, w3, w1, kc, out, head_out, G, rev);
            else launch_pdl(k_gn_block_pairs<false, 2>, dim3(grid), dim3(352), GnPairsCfg<2>::SMEM, s, in, w3, w1, kc, out, (uint8_t *)nullptr, G, rev);
        }
    } else {
        const uint8_t *w3 = e.bw.gn_l3_w[block];
        if (head_out) k_gn_block<true><<<grid, 352, GnBlockCfg::SMEM, s>>>(in, w3, w1, kc, out, head_out, G);
        else k_gn_block<false><<<grid, 352, GnBlockCfg::SMEM, s>>>(in, w3, w1, kc, out, nullptr, G);
    }
}

// ---- host side -------------------------------------------------------------------------------------
template <int CIN, int COUT, int TAPS, bool RESIDENT, bool RESIDUAL, bool FROM_BOARD = false, bool HEADS = false>
static void launch_conv(Engine &e, const uint8_t *in, const uint8_t *w, const float *bias,
                        const uint8_t *res, uint8_t *out, const uint8_t *active, int G, cudaStream_t s,
                        HeadConv heads = HeadConv(), int rev = 0) {
    using C = ConvCfg<CIN, COUT, TAPS, RESIDENT>;
    auto kern = k_conv_umma<CIN, COUT, TAPS, RESIDENT, RESIDUAL, FROM_BOARD, HEADS>;
    static bool configured = false;
    if (!configured) {
        cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, C::SMEM);
        configured = true;
    }
    int grid = G < e.sm_count ? G : e.sm_count;
    static const bool alternate = !(getenv("GZ_GN_ALTERNATE") && atoi(getenv("GZ_GN_ALTERNATE")) == 0);   // A/B switch
    launch_pdl(kern, dim3(grid), dim3(64 + 32 * conv_epi_warps<FROM_BOARD, RESIDUAL>()), C::SMEM, s, in, w, bias, res, out, active, G,
               heads, alternate ? rev : 0);
}

void nn_pv_forward_bf16(Engine &e, const gz_board *boards, const uint8_t *active, int G,
                        float *logits, float *probs, float *value, float *prior_out,
                        const int32_t *prior_slot, cudaStream_t s) {
    const Bf16Weights &w = e.bw;
    uint8_t *p0 = e.planes[0], *p1 = e.planes[1], *p2 = e.planes[2];
    prof_begin(e, GZ_STAGE_PV_CONV0, 1, s);
    launch_conv<16, 128, 9, true, false, true>(e, reinterpret_cast<const uint8_t *>(boards), w.pv_conv0_w,
                                               e.pv.conv0.b, nullptr, p0, active, G, s);
    prof_end(e, s);
    // board direction alternates layer by layer (the input convolution walks forward): see k_conv_umma's `rev`.
    // The input convolution stays a launch of its own: fused into the first tower layer the way k_gn_block0_embed
    // fuses GraphNet's embed it measured no gain -- the tower's N = 128 UMMAs already take all of the shared-memory
    // bandwidth (8 KB of operands per 64 cycles), and the fused epilogue's slab writes crawled behind them (8.1 k cycles
    // per board): DESIGN.md section 5, profiles/r02_pv_fused_conv0.md.
    prof_begin(e, GZ_STAGE_PV_TOWER, 4, s);
    launch_conv<128, 128, 9, false, false>(e, p0, w.pv_tower_w[0], e.pv.tower[0].b, nullptr, p1, active, G, s, HeadConv(), 1);
    launch_conv<128, 128, 9, false, true>(e, p1, w.pv_tower_w[1], e.pv.tower[1].b, p0, p2, active, G, s, HeadConv(), 0);
    launch_conv<128, 128, 9, false, false>(e, p2, w.pv_tower_w[2], e.pv.tower[2].b, nullptr, p1, active, G, s, HeadConv(), 1);
    HeadConv hc = {e.pv.pconv_w, e.pv.pconv_b, e.pv.vconv_w, e.pv.vconv_b, e.pv_pc, e.pv_vc};
    launch_conv<128, 128, 9, false, true, false, true>(e, p1, w.pv_tower_w[3], e.pv.tower[3].b, p2, p0, active, G, s, hc, 0);
    prof_end(e, s);
    nn_heads_bf16(e, true, boards, active, G, nullptr, logits, probs, value, prior_out, prior_slot, s);
}

void nn_graphnet_convs_bf16(Engine &e, const gz_board *boards, const uint8_t *active, int G, cudaStream_t s) {
    const Bf16Weights &w = e.bw;
    uint8_t *p0 = e.planes[0], *p1 = e.planes[1];
    static const bool fused = !(getenv("GZ_GN_FUSED_EMBED") && atoi(getenv("GZ_GN_FUSED_EMBED")) == 0);   // A/B switch
    if (fused) {
        // embed + block 0 in one launch (k_gn_block0_embed), then blocks 1-3: p1 -> p0 -> p1 -> head operand
        static bool configured = false;
        if (!configured) {
            cudaFuncSetAttribute(k_gn_block0_embed, cudaFuncAttributeMaxDynamicSharedMemorySize, GnEmbedCfg::SMEM);
            configured = true;
        }
        prof_begin(e, GZ_STAGE_GN_CONVS, 4, s);
        GnBlockConsts kc;
        memcpy(kc.b3, w.gn_host_bias[0], sizeof(float) * 128);
        memcpy(kc.head_w, w.gn_host_head, sizeof(float) * 130);
        memcpy(kc.be, w.gn_host_embed_bias, sizeof(float) * 64);
        const int grid = G < e.sm_count ? G : e.sm_count;
        launch_pdl(k_gn_block0_embed, dim3(grid), dim3(GnEmbedCfg::THREADS), GnEmbedCfg::SMEM, s, boards,
                   (const uint8_t *)w.gn_embed_i2c_w, (const uint8_t *)w.gn_l3p_w[0], (const uint8_t *)w.gn_l1_w[0], kc, p1, G, 0);
        for (int i = 1; i < 4; ++i)
            launch_gn_block(e, (i & 1) ? p1 : p0, i, (i & 1) ? p0 : p1, i == 3 ? e.gn_pc : nullptr, G, s, i & 1);
        prof_end(e, s);
        return;
    }
    prof_begin(e, GZ_STAGE_GN_CONVS, 5, s);
    launch_conv<16, 64, 9, true, false, true>(e, reinterpret_cast<const uint8_t *>(boards), w.gn_embed_w,
                                              e.gn.embed.b, nullptr, p0, active, G, s);
    for (int i = 0; i < 4; ++i) {   // p0 -> p1 -> p0 -> p1 -> p0
        launch_gn_block(e, (i & 1) ? p1 : p0, i, (i & 1) ? p0 : p1, i == 3 ? e.gn_pc : nullptr, G, s, (i & 1) ^ 1);
    }
    prof_end(e, s);
}

void nn_debug_planes_to_nchw(Engine &e, int which, int C, int G, float *out, cudaStream_t s) {
    k_planes_to_nchw<<<G, 256, 0, s>>>(e.planes[which], C, nullptr, out);
}

// Output of the first `nlayers` tower convolutions (0 = the input convolution only) as fp32 NCHW.
void nn_debug_pv_layer_bf16(Engine &e, const gz_board *boards, int G, int nlayers, float *out, cudaStream_t s) {
    const Bf16Weights &w = e.bw;
    uint8_t *p0 = e.planes[0], *p1 = e.planes[1], *p2 = e.planes[2];
    launch_conv<16, 128, 9, true, false, true>(e, reinterpret_cast<const uint8_t *>(boards), w.pv_conv0_w,
                                               e.pv.conv0.b, nullptr, p0, nullptr, G, s);
    int last = 0;
    if (nlayers >= 1) { launch_conv<128, 128, 9, false, false>(e, p0, w.pv_tower_w[0], e.pv.tower[0].b, nullptr, p1, nullptr, G, s); last = 1; }
    if (nlayers >= 2) { launch_conv<128, 128, 9, false, true>(e, p1, w.pv_tower_w[1], e.pv.tower[1].b, p0, p2, nullptr, G, s); last = 2; }
    if (nlayers >= 3) { launch_conv<128, 128, 9, false, false>(e, p2, w.pv_tower_w[2], e.pv.tower[2].b, nullptr, p1, nullptr, G, s); last = 1; }
    if (nlayers >= 4) { launch_conv<128, 128, 9, false, true>(e, p1, w.pv_tower_w[3], e.pv.tower[3].b, p2, p0, nullptr, G, s); last = 0; }
    k_planes_to_nchw<<<G, 256, 0, s>>>(e.planes[last], 128, nullptr, out);
}

}  // namespace gz
