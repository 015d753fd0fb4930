// gz_nn_fp32.cu -- fp32 (FFMA) forward of the three networks: the parity-mode path (1e-4 against
// torch fp32).  AlphaZeroGomokuNet (neural_network.py:132-159), GraphNet and OpponentDQN
// (bg_planner.py:50-56,74-78).  BatchNorm is folded into the conv weights on the host
// (gz_api.cu).  Input planes are generated from the bitboards inside the first conv (a5 in
// SURVEY.md 8a), so the 2700-byte float plane tensor of the reference never exists in HBM.
//
// Conv kernel: one CTA per board, thread = (8 output channels) x (one board row of 15 cells),
// 120 accumulators in registers; the input-channel chunk (8 channels, zero halo, 17x17) and its
// weights ([ci][tap][cout], cout contiguous) are staged in shared memory.
#include "gz_engine.h"
#include "gz_bitboard.cuh"

namespace gz {

template <int CIN, int COUT, int KS, bool FROM_BOARD, bool RESIDUAL, bool RELU>
__global__ void __launch_bounds__(COUT * 2)
k_conv(const float *__restrict__ in, const gz_board *__restrict__ boards,
       const uint8_t *__restrict__ active, const float *__restrict__ w,
       const float *__restrict__ bias, const float *__restrict__ res, float *__restrict__ out) {
    constexpr int CC = CIN < 8 ? CIN : 8;
    constexpr int T = KS * KS;
    constexpr int NT = COUT * 2;
    constexpr int PADW = 17;
    __shared__ float sin_[CC][PADW * PADW];
    __shared__ __align__(16) float sw[CC * T * COUT];
    const int b = blockIdx.x;
    if (active && !active[b]) return;
    const int tid = threadIdx.x;
    const int cg = tid >> 4, y = tid & 15;
    for (int i = tid; i < CC * PADW * PADW; i += NT) (&sin_[0][0])[i] = 0.f;
    float acc[8][15];
#pragma unroll
    for (int co = 0; co < 8; ++co) {
        float bv = bias[cg * 8 + co];
#pragma unroll
        for (int x = 0; x < 15; ++x) acc[co][x] = bv;
    }
    for (int ci0 = 0; ci0 < CIN; ci0 += CC) {
        __syncthreads();
        for (int i = tid; i < CC * kCells; i += NT) {
            int ci = i / kCells, pos = i - ci * kCells;
            int r = pos / 15, x = pos - r * 15;
            float v;
            if (FROM_BOARD) {
                int p = (r << 4) | x;
                uint32_t bw = boards[b].stones[0][p >> 5] >> (p & 31) & 1u;
                uint32_t ww = boards[b].stones[1][p >> 5] >> (p & 31) & 1u;
                int plane = ci0 + ci;  // black, white, empty (neural_network.py:226-229)
                v = plane == 0 ? (float)bw : plane == 1 ? (float)ww : (float)(1u - (bw | ww));
            } else {
                v = in[((size_t)b * CIN + ci0 + ci) * kCells + pos];
            }
            sin_[ci][(r + 1) * PADW + x + 1] = v;
        }
        {
            const float4 *src = reinterpret_cast<const float4 *>(w + (size_t)ci0 * T * COUT);
            float4 *dst = reinterpret_cast<float4 *>(sw);
            for (int i = tid; i < CC * T * COUT / 4; i += NT) dst[i] = src[i];
        }
        __syncthreads();
        if (y < 15) {
#pragma unroll 1
            for (int ci = 0; ci < CC; ++ci) {
#pragma unroll
                for (int ky = 0; ky < KS; ++ky) {
                    float iv[17];
                    const float *row = &sin_[ci][(y + ky + (KS == 1 ? 1 : 0)) * PADW];
#pragma unroll
                    for (int x = 0; x < 17; ++x) iv[x] = row[x];
#pragma unroll
                    for (int kx = 0; kx < KS; ++kx) {
                        const float4 *wp = reinterpret_cast<const float4 *>(
                            &sw[((ci * KS + ky) * KS + kx) * COUT + cg * 8]);
                        float4 w0 = wp[0], w1 = wp[1];
                        float wv[8] = {w0.x, w0.y, w0.z, w0.w, w1.x, w1.y, w1.z, w1.w};
#pragma unroll
                        for (int co = 0; co < 8; ++co)
#pragma unroll
                            for (int x = 0; x < 15; ++x)
                                acc[co][x] = fmaf(wv[co], iv[x + kx + (KS == 1 ? 1 : 0)], acc[co][x]);
                    }
                }
            }
        }
    }
    if (y < 15) {
#pragma unroll
        for (int co = 0; co < 8; ++co) {
            size_t o = ((size_t)b * COUT + cg * 8 + co) * kCells + y * 15;
#pragma unroll
            for (int x = 0; x < 15; ++x) {
                float v = acc[co][x];
                if (RESIDUAL) v += res[o + x];
                if (RELU) v = v > 0.f ? v : 0.f;
                out[o + x] = v;
            }
        }
    }
}

// Heads.  NB boards per CTA so that the 450x225 policy matrix is read once per NB boards.
//   policy: conv1x1 C->2 (no BN, no ReLU), flatten channel-major, Linear 450->225   (:149-151)
//   value : conv1x1 C->1, Linear 225->64, ReLU, Linear 64->1, tanh                 (:154-157)
// then softmax over all 225 logits (host numpy in the reference, :248-252) and, when asked, the
// node prior: probabilities gathered over the legal cells and renormalised (ai_agent.py:564-582).
constexpr int kHeadNB = 4;

template <int C, bool VALUE, bool PLANES>
__global__ void __launch_bounds__(256)
k_heads(const float *__restrict__ x, const gz_board *__restrict__ boards,
        const uint8_t *__restrict__ active, int G, const float *__restrict__ pconv_w,
        const float *__restrict__ pconv_b, const float *__restrict__ pfc_wt,
        const float *__restrict__ pfc_b, const float *__restrict__ vconv_w,
        const float *__restrict__ vconv_b, const float *__restrict__ vfc1_wt,
        const float *__restrict__ vfc1_b, const float *__restrict__ vfc2_w,
        const float *__restrict__ vfc2_b, float *__restrict__ logits, float *__restrict__ probs,
        float *__restrict__ value, float *__restrict__ prior, const int32_t *__restrict__ prior_slot,
        int prior_cap, float *__restrict__ vhead, int probs_pitch) {
    __shared__ float spc[kHeadNB][450];
    __shared__ float svc[kHeadNB][kCells];
    __shared__ float slog[kHeadNB][kCells];
    __shared__ float sh[kHeadNB][64];
    const int t = threadIdx.x;
    const int b0 = blockIdx.x * kHeadNB;
    bool any = false;
    for (int nb = 0; nb < kHeadNB; ++nb) {
        int b = b0 + nb;
        if (b < G && (!active || active[b])) any = true;
    }
    if (!any) return;
    if (t < kCells) {
        for (int nb = 0; nb < kHeadNB; ++nb) {
            int b = b0 + nb;
            float p0 = pconv_b[0], p1 = pconv_b[1], v0 = VALUE ? vconv_b[0] : 0.f;
            if (b < G && PLANES) {
                // bf16 planes [board][C/8][288][8]: one 16-byte load gives 8 channels of this cell
                const int pos = 16 * (t / 15 + 1) + t % 15;
                const uint8_t *pb = reinterpret_cast<const uint8_t *>(x) + kGuardBytes +
                                    (size_t)b * (C / 8) * kPlaneBytes + (size_t)pos * 16;
                for (int cg = 0; cg < C / 8; ++cg) {
                    uint4 u = *reinterpret_cast<const uint4 *>(pb + (size_t)cg * kPlaneBytes);
                    const uint32_t uw[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
                    for (int j = 0; j < 8; ++j) {
                        float xv = __uint_as_float((j & 1) ? (uw[j >> 1] & 0xFFFF0000u) : (uw[j >> 1] << 16));
                        int c = cg * 8 + j;
                        p0 = fmaf(pconv_w[2 * c], xv, p0);
                        p1 = fmaf(pconv_w[2 * c + 1], xv, p1);
                        if (VALUE) v0 = fmaf(vconv_w[c], xv, v0);
                    }
                }
            } else if (b < G) {
                const float *xb = x + (size_t)b * C * kCells + t;
                for (int c = 0; c < C; ++c) {
                    float xv = xb[(size_t)c * kCells];
                    p0 = fmaf(pconv_w[2 * c], xv, p0);
                    p1 = fmaf(pconv_w[2 * c + 1], xv, p1);
                    if (VALUE) v0 = fmaf(vconv_w[c], xv, v0);
                }
            }
            spc[nb][t] = p0;
            spc[nb][kCells + t] = p1;
            svc[nb][t] = v0;
        }
    }
    __syncthreads();
    if (t < kCells) {
        float acc[kHeadNB];
        for (int nb = 0; nb < kHeadNB; ++nb) acc[nb] = pfc_b[t];
        for (int i = 0; i < 450; ++i) {
            float wv = pfc_wt[i * kCells + t];
#pragma unroll
            for (int nb = 0; nb < kHeadNB; ++nb) acc[nb] = fmaf(wv, spc[nb][i], acc[nb]);
        }
        for (int nb = 0; nb < kHeadNB; ++nb) slog[nb][t] = acc[nb];
    }
    if (VALUE && t < 64) {
        float acc[kHeadNB];
        for (int nb = 0; nb < kHeadNB; ++nb) acc[nb] = vfc1_b[t];
        for (int i = 0; i < kCells; ++i) {
            float wv = vfc1_wt[i * 64 + t];
#pragma unroll
            for (int nb = 0; nb < kHeadNB; ++nb) acc[nb] = fmaf(wv, svc[nb][i], acc[nb]);
        }
        for (int nb = 0; nb < kHeadNB; ++nb) sh[nb][t] = acc[nb] > 0.f ? acc[nb] : 0.f;
    }
    __syncthreads();
    // one warp per board: value, softmax, prior
    const int warp = t >> 5, lane = t & 31;
    if (warp >= kHeadNB) return;
    const int b = b0 + warp;
    if (b >= G || (active && !active[b])) return;
    if (VALUE) {
        float s = vfc2_w[lane] * sh[warp][lane] + vfc2_w[lane + 32] * sh[warp][lane + 32];
        for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
        float vv = tanhf(s + vfc2_b[0]);
        if (lane == 0) {
            if (value) value[b] = vv;
            if (vhead && prior_slot) vhead[(size_t)b * prior_cap + prior_slot[b]] = vv;
        }
    }
    float mx = -INFINITY;
    for (int i = lane; i < kCells; i += 32) mx = fmaxf(mx, slog[warp][i]);
    for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    float ev[8], sum = 0.f;
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        int i = lane + 32 * k;
        ev[k] = i < kCells ? expf(slog[warp][i] - mx) : 0.f;
        sum += ev[k];
    }
    for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
    double legal_sum = 0.0;
    uint32_t legal_bits = 0;
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        int i = lane + 32 * k;
        if (i >= kCells) continue;
        float pr = ev[k] / sum;
        ev[k] = pr;
        if (logits) logits[(size_t)b * kCells + i] = slog[warp][i];
        if (probs) probs[(size_t)b * probs_pitch + i] = pr;
        if (prior) {
            int p = cell_to_bit(i);
            uint32_t occ = (boards[b].stones[0][p >> 5] | boards[b].stones[1][p >> 5]) >> (p & 31) & 1u;
            if (!occ) { legal_sum += (double)pr; legal_bits |= 1u << k; }
        }
    }
    if (prior) {
        for (int o = 16; o > 0; o >>= 1) legal_sum += __shfl_xor_sync(0xffffffffu, legal_sum, o);
        float *dst = prior + ((size_t)b * prior_cap + prior_slot[b]) * kPolicyPitch;
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            int i = lane + 32 * k;
            if (i >= kCells) continue;
            float pv = 0.f;
            if (legal_bits >> k & 1u) pv = legal_sum > 0.0 ? (float)((double)ev[k] / legal_sum) : ev[k];
            dst[i] = pv;
        }
    }
}

// OpponentDQN: Linear 675->256, ReLU, Linear 256->256, ReLU, Linear 256->225 (bg_planner.py:68-78)
constexpr int kDqnNB = 4;
__global__ void __launch_bounds__(256)
k_dqn(const gz_board *__restrict__ boards, const uint8_t *__restrict__ active, int G, DqnWeights w,
      float *__restrict__ q) {
    __shared__ float sx[kDqnNB][675];
    __shared__ float h1[kDqnNB][256];
    __shared__ float h2[kDqnNB][256];
    const int t = threadIdx.x, b0 = blockIdx.x * kDqnNB;
    bool any = false;
    for (int nb = 0; nb < kDqnNB; ++nb) {
        int b = b0 + nb;
        if (b < G && (!active || active[b])) any = true;
    }
    if (!any) return;
    for (int i = t; i < kDqnNB * kCells; i += 256) {
        int nb = i / kCells, pos = i - nb * kCells, b = b0 + nb;
        float bl = 0.f, wh = 0.f, em = 0.f;
        if (b < G) {
            int p = cell_to_bit(pos);
            uint32_t bw = boards[b].stones[0][p >> 5] >> (p & 31) & 1u;
            uint32_t ww = boards[b].stones[1][p >> 5] >> (p & 31) & 1u;
            bl = (float)bw; wh = (float)ww; em = (float)(1u - (bw | ww));
        }
        sx[nb][pos] = bl; sx[nb][kCells + pos] = wh; sx[nb][2 * kCells + pos] = em;
    }
    __syncthreads();
    float acc[kDqnNB];
    for (int nb = 0; nb < kDqnNB; ++nb) acc[nb] = w.b1[t];
    for (int i = 0; i < 675; ++i) {
        float wv = w.w1t[i * 256 + t];
#pragma unroll
        for (int nb = 0; nb < kDqnNB; ++nb) acc[nb] = fmaf(wv, sx[nb][i], acc[nb]);
    }
    for (int nb = 0; nb < kDqnNB; ++nb) h1[nb][t] = acc[nb] > 0.f ? acc[nb] : 0.f;
    __syncthreads();
    for (int nb = 0; nb < kDqnNB; ++nb) acc[nb] = w.b2[t];
    for (int i = 0; i < 256; ++i) {
        float wv = w.w2t[i * 256 + t];
#pragma unroll
        for (int nb = 0; nb < kDqnNB; ++nb) acc[nb] = fmaf(wv, h1[nb][i], acc[nb]);
    }
    for (int nb = 0; nb < kDqnNB; ++nb) h2[nb][t] = acc[nb] > 0.f ? acc[nb] : 0.f;
    __syncthreads();
    if (t < kCells) {
        for (int nb = 0; nb < kDqnNB; ++nb) acc[nb] = w.b3[t];
        for (int i = 0; i < 256; ++i) {
            float wv = w.w3t[i * kCells + t];
#pragma unroll
            for (int nb = 0; nb < kDqnNB; ++nb) acc[nb] = fmaf(wv, h2[nb][i], acc[nb]);
        }
        for (int nb = 0; nb < kDqnNB; ++nb) {
            int b = b0 + nb;
            if (b < G && (!active || active[b])) q[(size_t)b * kCells + t] = acc[nb];
        }
    }
}

void nn_pv_forward_fp32(Engine &e, const gz_board *boards, const uint8_t *active, int G,
                        float *logits, float *probs, float *value, float *prior_out,
                        const int32_t *prior_slot, cudaStream_t s) {
    const PvWeights &w = e.pv;
    float *a0 = e.act[0], *a1 = e.act[1], *a2 = e.act[2];
    prof_begin(e, GZ_STAGE_PV_CONV0, 1, s);
    k_conv<3, 128, 3, true, false, true><<<G, 256, 0, s>>>(nullptr, boards, active, w.conv0.w, w.conv0.b, nullptr, a0);
    prof_end(e, s);
    prof_begin(e, GZ_STAGE_PV_TOWER, 4, s);
    k_conv<128, 128, 3, false, false, true><<<G, 256, 0, s>>>(a0, nullptr, active, w.tower[0].w, w.tower[0].b, nullptr, a1);
    k_conv<128, 128, 3, false, true, true><<<G, 256, 0, s>>>(a1, nullptr, active, w.tower[1].w, w.tower[1].b, a0, a2);
    k_conv<128, 128, 3, false, false, true><<<G, 256, 0, s>>>(a2, nullptr, active, w.tower[2].w, w.tower[2].b, nullptr, a1);
    k_conv<128, 128, 3, false, true, true><<<G, 256, 0, s>>>(a1, nullptr, active, w.tower[3].w, w.tower[3].b, a2, a0);
    prof_end(e, s);
    prof_begin(e, GZ_STAGE_PV_HEADS, 1, s);
    k_heads<128, true, false><<<(G + kHeadNB - 1) / kHeadNB, 256, 0, s>>>(
        a0, boards, active, G, w.pconv_w, w.pconv_b, w.pfc_wt, w.pfc_b, w.vconv_w, w.vconv_b,
        w.vfc1_wt, w.vfc1_b, w.vfc2_w, w.vfc2_b, logits, probs, value, prior_out, prior_slot,
        e.tree.cap, prior_slot ? e.tree.vhead : nullptr, kCells);
    prof_end(e, s);
}

void nn_planner_forward_fp32(Engine &e, const gz_board *boards, int G, float *gn_logits,
                             float *gn_policy, int policy_pitch, float *dqn_q, cudaStream_t s) {
    const GraphNetWeights &w = e.gn;
    float *a0 = e.act[0], *a1 = e.act[1];
    // the planner nets only see live rollout boards; finished ones are skipped via status
    prof_begin(e, GZ_STAGE_GN_CONVS, 9, s);
    k_conv<3, 64, 3, true, false, true><<<G, 128, 0, s>>>(nullptr, boards, nullptr, w.embed.w, w.embed.b, nullptr, a0);
    for (int i = 0; i < 4; ++i) {
        k_conv<64, 64, 3, false, false, true><<<G, 128, 0, s>>>(a0, nullptr, nullptr, w.l3[i].w, w.l3[i].b, nullptr, a1);
        k_conv<64, 64, 1, false, false, true><<<G, 128, 0, s>>>(a1, nullptr, nullptr, w.l1[i].w, w.l1[i].b, nullptr, a0);
    }
    prof_end(e, s);
    prof_begin(e, GZ_STAGE_GN_HEAD, 1, s);
    k_heads<64, false, false><<<(G + kHeadNB - 1) / kHeadNB, 256, 0, s>>>(
        a0, boards, nullptr, G, w.hconv_w, w.hconv_b, w.hfc_wt, w.hfc_b, nullptr, nullptr, nullptr,
        nullptr, nullptr, nullptr, gn_logits, gn_policy, nullptr, nullptr, nullptr, 0, nullptr, policy_pitch);
    prof_end(e, s);
    prof_begin(e, GZ_STAGE_DQN, 1, s);
    k_dqn<<<(G + kDqnNB - 1) / kDqnNB, 256, 0, s>>>(boards, nullptr, G, e.dqn, dqn_q);
    prof_end(e, s);
}

void nn_debug_pv_layer_fp32(Engine &e, const gz_board *boards, int G, int nlayers, float *out, cudaStream_t s) {
    const PvWeights &w = e.pv;
    float *a0 = e.act[0], *a1 = e.act[1], *a2 = e.act[2], *last = a0;
    k_conv<3, 128, 3, true, false, true><<<G, 256, 0, s>>>(nullptr, boards, nullptr, w.conv0.w, w.conv0.b, nullptr, a0);
    if (nlayers >= 1) { k_conv<128, 128, 3, false, false, true><<<G, 256, 0, s>>>(a0, nullptr, nullptr, w.tower[0].w, w.tower[0].b, nullptr, a1); last = a1; }
    if (nlayers >= 2) { k_conv<128, 128, 3, false, true, true><<<G, 256, 0, s>>>(a1, nullptr, nullptr, w.tower[1].w, w.tower[1].b, a0, a2); last = a2; }
    if (nlayers >= 3) { k_conv<128, 128, 3, false, false, true><<<G, 256, 0, s>>>(a2, nullptr, nullptr, w.tower[2].w, w.tower[2].b, nullptr, a1); last = a1; }
    if (nlayers >= 4) { k_conv<128, 128, 3, false, true, true><<<G, 256, 0, s>>>(a1, nullptr, nullptr, w.tower[3].w, w.tower[3].b, a2, a0); last = a0; }
    cudaMemcpyAsync(out, last, sizeof(float) * (size_t)G * 128 * kCells, cudaMemcpyDeviceToDevice, s);
}

// ---- gz_pv_forward_static: the policy-value forward without any allocation -----------------------------
// Activation scratch of the fp32 forward (three ping-pong buffers [G][128][225] f32) in a file-scope device array
// instead of the handle's cudaMalloc'ed buffers, for callers that must not allocate (CUDA-graph capture, serving
// loops).  kStaticBytes = 5 529 600 bytes for GZ_STATIC_MAX_BOARDS = 16 boards; every GPU that loads the
// library holds it until exit.  ONE array per device shared by every handle and every call: it keeps the last call's
// contents (nothing has to start at zero: every layer writes each element of its output for boards [0, G) before the
// next layer reads it), and two calls that overlap on the GPU (different streams) race on it.
constexpr size_t kStaticActFloats = (size_t)GZ_STATIC_MAX_BOARDS * 128 * kCells;
constexpr size_t kStaticBytes = 3 * kStaticActFloats * sizeof(float);
static __device__ __align__(256) float g_static_act[3 * kStaticActFloats];

// In host code `g_static_act` is a host-side stand-in for the symbol, not a device address: the address is looked up
// once per device, with that device current, and cached.
static float *static_scratch(int device, cudaError_t *err) {
    constexpr int kMaxDevices = 64;
    static float *ptr[kMaxDevices];
    *err = cudaSuccess;
    if (device < 0 || device >= kMaxDevices) { *err = cudaErrorInvalidDevice; return nullptr; }
    if (!ptr[device]) {
        float *p = nullptr;
        *err = cudaGetSymbolAddress((void **)&p, g_static_act);
        if (*err != cudaSuccess) return nullptr;
        ptr[device] = p;
    }
    return ptr[device];
}

// The same six kernels as nn_pv_forward_fp32 (fp32 FFMA, weights of gz_set_pv_weights) on the static scratch.
// e's device must be current.  No stage events (their pool grows by allocation).
cudaError_t nn_pv_forward_static(Engine &e, const gz_board *boards, int G, float *logits, float *probs, float *value,
                                 cudaStream_t s) {
    cudaError_t err;
    float *a0 = static_scratch(e.cfg.device, &err);
    if (!a0) return err;
    float *a1 = a0 + kStaticActFloats, *a2 = a1 + kStaticActFloats;
    const PvWeights &w = e.pv;
    k_conv<3, 128, 3, true, false, true><<<G, 256, 0, s>>>(nullptr, boards, nullptr, w.conv0.w, w.conv0.b, nullptr, a0);
    k_conv<128, 128, 3, false, false, true><<<G, 256, 0, s>>>(a0, nullptr, nullptr, w.tower[0].w, w.tower[0].b, nullptr, a1);
    k_conv<128, 128, 3, false, true, true><<<G, 256, 0, s>>>(a1, nullptr, nullptr, w.tower[1].w, w.tower[1].b, a0, a2);
    k_conv<128, 128, 3, false, false, true><<<G, 256, 0, s>>>(a2, nullptr, nullptr, w.tower[2].w, w.tower[2].b, nullptr, a1);
    k_conv<128, 128, 3, false, true, true><<<G, 256, 0, s>>>(a1, nullptr, nullptr, w.tower[3].w, w.tower[3].b, a2, a0);
    k_heads<128, true, false><<<(G + kHeadNB - 1) / kHeadNB, 256, 0, s>>>(
        a0, boards, nullptr, G, w.pconv_w, w.pconv_b, w.pfc_wt, w.pfc_b, w.vconv_w, w.vconv_b,
        w.vfc1_wt, w.vfc1_b, w.vfc2_w, w.vfc2_b, logits, probs, value, nullptr, nullptr, 0, nullptr, kCells);
    e.launches += 6;
    return cudaGetLastError();
}

void nn_pv_forward(Engine &e, const gz_board *boards, const uint8_t *active, int G, float *logits,
                   float *probs, float *value, float *prior_out, const int32_t *prior_slot, cudaStream_t s) {
    if (e.cfg.precision == 1) nn_pv_forward_bf16(e, boards, active, G, logits, probs, value, prior_out, prior_slot, s);
    else nn_pv_forward_fp32(e, boards, active, G, logits, probs, value, prior_out, prior_slot, s);
}

void nn_planner_forward(Engine &e, const gz_board *boards, int G, float *gn_logits, float *gn_policy,
                        int policy_pitch, float *dqn_q, cudaStream_t s) {
    if (e.cfg.precision == 1) {
        nn_graphnet_convs_bf16(e, boards, nullptr, G, s);      // ends with the fused head conv -> e.gn_pc
        const bool internal = policy_pitch == kPolicyPitch;
        // the search and the planner only read the softmax (pitch kPolicyPitch); logits are an API output
        if (internal) {
            nn_gnhead_dqn_bf16(e, boards, G, gn_policy, dqn_q, s);   // head + DQN in one launch
        } else {
            nn_heads_bf16(e, false, boards, nullptr, G, nullptr, gn_logits, gn_policy, nullptr, nullptr, nullptr, s);
            nn_dqn_bf16(e, boards, G, dqn_q, s);
        }
    } else {
        nn_planner_forward_fp32(e, boards, G, gn_logits, gn_policy, policy_pitch, dqn_q, s);
    }
}

}  // namespace gz
