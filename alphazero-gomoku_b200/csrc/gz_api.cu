// gz_api.cu -- C ABI of libgomoku_b200.so (include/gomoku_b200.h): handle life cycle, weight
// packing (BatchNorm folding, layout change), and the host-side wave scheduler that advances all
// trees one simulation at a time.  No CPU compute path exists here: every entry point that does
// work launches kernels and fails with GZ_ERR_CUDA when there is no device.
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <chrono>
#include <vector>

#include "gz_engine.h"

using namespace gz;

static thread_local char g_err[512] = "";

static int fail(int code, const char *fmt, const char *a = "", const char *b = "") {
    snprintf(g_err, sizeof g_err, fmt, a, b);
    return code;
}
#define CK(call)                                                                             \
    do {                                                                                     \
        cudaError_t _e = (call);                                                             \
        if (_e != cudaSuccess) return fail(GZ_ERR_CUDA, "%s: %s", #call, cudaGetErrorString(_e)); \
    } while (0)
#define CKL()                                                                                \
    do {                                                                                     \
        cudaError_t _e = cudaGetLastError();                                                 \
        if (_e != cudaSuccess) return fail(GZ_ERR_CUDA, "kernel launch: %s%s", cudaGetErrorString(_e)); \
    } while (0)

extern "C" const char *gz_last_error(void) { return g_err; }
extern "C" int gz_version(void) { return 100; }

extern "C" int64_t gz_pv_param_count(void) {
    return 3456 + 128 + 512 + 2 * (2 * (147456 + 128) + 2 * 512) + 256 + 2 + 101250 + 225 + 128 + 1 +
           14400 + 64 + 64 + 1;
}
extern "C" int64_t gz_graphnet_param_count(void) {
    return 64 * 27 + 64 + 4 * (64 * 64 * 9 + 64 + 64 * 64 + 64) + 128 + 2 + 101250 + 225;
}
extern "C" int64_t gz_dqn_param_count(void) { return 256 * 675 + 256 + 256 * 256 + 256 + 225 * 256 + 225; }

template <typename T>
static cudaError_t dalloc(T **p, size_t n) {
    cudaError_t e = cudaMalloc((void **)p, n * sizeof(T));
    if (e == cudaSuccess) e = cudaMemset(*p, 0, n * sizeof(T));
    return e;
}

extern "C" int gz_create(const gz_config *cfg, gz_handle *out) {
    if (!cfg || !out) return fail(GZ_ERR_ARG, "gz_create: null argument");
    if (cfg->max_games < 1 || cfg->max_simulations < 1 || cfg->max_simulations > 32000)
        return fail(GZ_ERR_ARG, "gz_create: max_games/max_simulations out of range");
    int ndev = 0;
    cudaError_t e0 = cudaGetDeviceCount(&ndev);
    if (e0 != cudaSuccess || ndev == 0)
        return fail(GZ_ERR_CUDA, "gz_create: no CUDA device (%s); this library has no CPU path",
                    cudaGetErrorString(e0));
    CK(cudaSetDevice(cfg->device));
    Engine *e = new Engine();   // value-initialised: every POD member is zero
    e->cfg = *cfg;
    cudaDeviceProp prop;
    CK(cudaGetDeviceProperties(&prop, cfg->device));
    e->sm_count = prop.multiProcessorCount;
    CK(cudaStreamCreateWithFlags(&e->own_stream, cudaStreamNonBlocking));
    if (e->cfg.lanes < 0 || e->cfg.lanes > kMaxLanes) { delete e; return fail(GZ_ERR_ARG, "gz_create: lanes out of range"); }
    for (int i = 0; i < kMaxLanes; ++i) {
        e->lane[i] = new Engine();
        CK(cudaStreamCreateWithFlags(&e->lane_stream[i], cudaStreamNonBlocking));
        CK(cudaEventCreateWithFlags(&e->lane_done[i], cudaEventDisableTiming));
    }
    CK(cudaEventCreateWithFlags(&e->lane_go, cudaEventDisableTiming));
    const size_t G = cfg->max_games, cap = (size_t)cfg->max_simulations + 1, N = G * cap;
    TreeArrays &t = e->tree;
    t.cap = (int32_t)cap;
    CK(dalloc(&t.parent, N)); CK(dalloc(&t.move, N)); CK(dalloc(&t.status, N));
    CK(dalloc(&t.nchild, N)); CK(dalloc(&t.nempty, N)); CK(dalloc(&t.visits, N));
    CK(dalloc(&t.value, N)); CK(dalloc(&t.bgb, N)); CK(dalloc(&t.vhead, N));
    if (cfg->store_priors) CK(dalloc(&t.prior, N * kPolicyPitch));
    CK(dalloc(&t.n_nodes, G)); CK(dalloc(&t.root, G)); CK(dalloc(&t.ai_player, G));
    SimArrays &s = e->sim;
    CK(dalloc(&s.board, G)); CK(dalloc(&s.leaf, G)); CK(dalloc(&s.need_eval, G));
    CK(dalloc(&s.steps, G)); CK(dalloc(&s.value, G)); CK(dalloc(&s.topk, G * kMaxTopK));
    CK(dalloc(&s.ntop, G));
    if (cfg->precision == 1) {
        for (int i = 0; i < 3; ++i) CK(dalloc(&e->planes[i], G * 16 * kPlaneBytes + 2 * kGuardBytes));
    } else {
        for (int i = 0; i < 3; ++i) CK(dalloc(&e->act[i], G * 128 * kCells));
    }
    CK(dalloc(&e->gn_logits, G * kCells)); CK(dalloc(&e->gn_policy, G * kPolicyPitch));
    {
        const size_t tiles = (G + 127) / 128;   // zero filled once: K padding and rows past G are never written
        CK(dalloc(&e->pv_pc, tiles * kHeadTileBytesP)); CK(dalloc(&e->pv_vc, tiles * kHeadTileBytesV));
        CK(dalloc(&e->gn_pc, tiles * kHeadTileBytesP));
    }
    CK(dalloc(&e->dqn_q, G * kCells));
    CK(dalloc(&e->pv_logits, G * kCells)); CK(dalloc(&e->pv_probs, G * kCells));
    CK(dalloc(&e->pv_value, G));
    CK(dalloc(&e->tmp_boards, G));
    CK(dalloc(&e->tmp_u8, G * 232));
    CK(dalloc(&e->io_cv, G * kCells)); CK(dalloc(&e->io_cval, G * kCells)); CK(dalloc(&e->io_best, G));
    CK(dalloc(&e->counters, 8));
    CK(dalloc(&e->d_sim, 4));
    CK(dalloc(&e->log_table, kLogTableMax));
    CK(dalloc(&e->tanh_table, kTanhTableMax));
    *out = reinterpret_cast<gz_handle>(e);
    return GZ_OK;
}

static void sp_free(Engine *e);

extern "C" int gz_destroy(gz_handle h) {
    if (!h) return GZ_OK;
    Engine *e = reinterpret_cast<Engine *>(h);
    cudaSetDevice(e->cfg.device);
    cudaDeviceSynchronize();
    TreeArrays &t = e->tree;
    void *ptrs[] = {t.parent, t.move, t.status, t.nchild, t.nempty, t.visits, t.value, t.bgb, t.vhead,
                    t.prior, t.n_nodes, t.root, t.ai_player, e->sim.board, e->sim.leaf,
                    e->sim.need_eval, e->sim.steps, e->sim.value, e->sim.topk, e->sim.ntop, e->act[0],
                    e->act[1], e->act[2], e->gn_logits, e->gn_policy, e->dqn_q, e->pv_logits,
                    e->pv_probs, e->pv_value, e->tmp_boards, e->tmp_u8, e->io_cv, e->io_cval, e->io_best, e->counters, e->d_sim, e->log_table,
                    e->tanh_table, e->weight_blob, e->planes[0], e->planes[1], e->planes[2], e->bw.pv_blob, e->bw.gn_blob, e->bw.dqn_w, e->bw.dqn_bias, e->bw.pv_head_blob, e->bw.gn_head_blob, e->pv_pc, e->pv_vc, e->gn_pc, e->pv.conv0.w, e->gn.embed.w, e->dqn.w1t};
    for (void *p : ptrs) if (p) cudaFree(p);
    sp_free(e);
    for (cudaEvent_t ev : e->ev_pool) cudaEventDestroy(ev);
    for (int i = 0; i < kMaxLanes; ++i) {
        cudaStreamDestroy(e->lane_stream[i]);
        cudaEventDestroy(e->lane_done[i]);
        delete e->lane[i];
    }
    cudaEventDestroy(e->lane_go);
    cudaStreamDestroy(e->own_stream);
    delete e;
    return GZ_OK;
}

// ---- weight packing ---------------------------------------------------------------------------
namespace {
struct Packer {
    std::vector<float> host;
    size_t take(size_t n) {  // 64-float (256 B) aligned slots
        size_t o = host.size();
        host.resize(o + ((n + 63) / 64) * 64, 0.f);
        return o;
    }
};

// conv weight [cout][cin][k][k] (+ optional BN [w,b,mean,var]) -> [cin][tap][cout], folded bias
void pack_conv(Packer &pk, const float *w, const float *b, const float *bn, int cout, int cin, int k,
               size_t &ow, size_t &ob) {
    int T = k * k;
    ow = pk.take((size_t)cin * T * cout);
    ob = pk.take(cout);
    for (int co = 0; co < cout; ++co) {
        double scale = 1.0, shift = 0.0;
        if (bn) {
            scale = (double)bn[co] / sqrt((double)bn[3 * cout + co] + 1e-5);
            shift = (double)bn[cout + co] - (double)bn[2 * cout + co] * scale;
        }
        for (int ci = 0; ci < cin; ++ci)
            for (int t = 0; t < T; ++t)
                pk.host[ow + ((size_t)ci * T + t) * cout + co] =
                    (float)((double)w[((size_t)co * cin + ci) * T + t] * scale);
        pk.host[ob + co] = (float)((double)b[co] * scale + shift);
    }
}
// linear weight [out][in] -> [in][out]
void pack_linear_t(Packer &pk, const float *w, const float *b, int nout, int nin, size_t &ow, size_t &ob) {
    ow = pk.take((size_t)nin * nout);
    ob = pk.take(nout);
    for (int o = 0; o < nout; ++o) {
        for (int i = 0; i < nin; ++i) pk.host[ow + (size_t)i * nout + o] = w[(size_t)o * nin + i];
        pk.host[ob + o] = b[o];
    }
}
// 1x1 head conv [cout][cin] -> [cin][cout]
void pack_head_conv(Packer &pk, const float *w, const float *b, int cout, int cin, size_t &ow, size_t &ob) {
    pack_linear_t(pk, w, b, cout, cin, ow, ob);
}
}  // namespace

static uint16_t f2bf(float f) {   // round to nearest even
    uint32_t u;
    memcpy(&u, &f, 4);
    if ((u & 0x7F800000u) == 0x7F800000u) return (uint16_t)(u >> 16);
    u += 0x7FFFu + ((u >> 16) & 1u);
    return (uint16_t)(u >> 16);
}
// conv weight [cout][cin][k][k] (+ BN) -> per K-step (tap, 64-channel chunk) a block
// [8 channel groups][cout][8 channels] of bf16: K-major SWIZZLE_NONE core matrices for tcgen05.mma
static void pack_conv_bf16(std::vector<uint16_t> &out, const float *w, const float *bn, int cout, int cin, int k) {
    int T = k * k, chunks = cin / 64;
    size_t base = out.size();
    out.resize(base + (size_t)T * cin * cout);
    for (int tap = 0; tap < T; ++tap)
        for (int ch = 0; ch < chunks; ++ch)
            for (int cgl = 0; cgl < 8; ++cgl)
                for (int co = 0; co < cout; ++co) {
                    double scale = bn ? (double)bn[co] / sqrt((double)bn[3 * cout + co] + 1e-5) : 1.0;
                    for (int j = 0; j < 8; ++j) {
                        int ci = ch * 64 + cgl * 8 + j;
                        size_t o = base + ((((size_t)(tap * chunks + ch) * 8 + cgl) * cout + co) * 8 + j);
                        out[o] = f2bf((float)((double)w[((size_t)co * cin + ci) * T + tap] * scale));
                    }
                }
}
// GraphNet 3x3 64 -> 64 weights for k_gn_block_pairs: three pair blocks [8 cg][128 rows][8 ch] -- rows 0-63 = tap
// (ty, dx = 0), rows 64-127 = tap (ty, dx = -1) -- then three single blocks [8 cg][64][8 ch] = tap (ty, dx = +1)
static void pack_gn3x3_pairs(std::vector<uint16_t> &out, const float *w) {
    size_t base = out.size();
    out.resize(base + (size_t)9 * 64 * 64);
    size_t o = base;
    for (int ty = 0; ty < 3; ++ty)
        for (int cg = 0; cg < 8; ++cg)
            for (int r = 0; r < 128; ++r)
                for (int j = 0; j < 8; ++j) {
                    const int co = r & 63, tap = 3 * ty + (r < 64 ? 1 : 0), ci = cg * 8 + j;
                    out[o++] = f2bf(w[((size_t)co * 64 + ci) * 9 + tap]);
                }
    for (int ty = 0; ty < 3; ++ty)
        for (int cg = 0; cg < 8; ++cg)
            for (int co = 0; co < 64; ++co)
                for (int j = 0; j < 8; ++j)
                    out[o++] = f2bf(w[((size_t)co * 64 + cg * 8 + j) * 9 + 3 * ty + 2]);
}
// input convolution [cout][3][3][3] (+ BN) -> per tap [2 channel groups][cout][8 channels] bf16 with
// channels 0..2 = black / white / empty and the other 13 zero (FROM_BOARD mode of k_conv_umma)
static void pack_conv0_bf16(std::vector<uint16_t> &out, const float *w, const float *bn, int cout) {
    size_t base = out.size();
    out.resize(base + (size_t)9 * 2 * cout * 8, 0);
    for (int tap = 0; tap < 9; ++tap)
        for (int co = 0; co < cout; ++co) {
            double scale = bn ? (double)bn[co] / sqrt((double)bn[3 * cout + co] + 1e-5) : 1.0;
            for (int ci = 0; ci < 3; ++ci)
                out[base + (((size_t)tap * 2 + 0) * cout + co) * 8 + ci] =
                    f2bf((float)((double)w[((size_t)co * 3 + ci) * 9 + tap] * scale));
        }
}
// An input convolution [cout][3][3][3] (+ BN) as ONE im2col GEMM operand (k_gn_block0_embed): K = 27
// padded to 32, in the order the kernels' producers build the one-hot tile in: k = 3 * (3 * plane + ky) + kx (plane =
// black / white / empty) -> [4 k-groups][cout][8 k] bf16
static void pack_conv0_im2col_bf16(std::vector<uint16_t> &out, const float *w, const float *bn, int cout) {
    size_t base = out.size();
    out.resize(base + (size_t)4 * cout * 8, 0);
    for (int k = 0; k < 27; ++k) {
        const int plane = k / 9, ky = (k / 3) % 3, kx = k % 3;
        for (int co = 0; co < cout; ++co) {
            double scale = bn ? (double)bn[co] / sqrt((double)bn[3 * cout + co] + 1e-5) : 1.0;
            out[base + ((size_t)(k >> 3) * cout + co) * 8 + (k & 7)] =
                f2bf((float)((double)w[((size_t)co * 3 + plane) * 9 + ky * 3 + kx] * scale));
        }
    }
}
// linear weight [nout][nin] -> K-chunks of 64 inputs: [chunk][8 k-groups][256 n][8 k] bf16, zero padded
static void pack_fc_bf16(std::vector<uint16_t> &out, const float *w, int nout, int nin, int chunks, int npad = 256) {
    size_t base = out.size();
    out.resize(base + (size_t)chunks * 8 * npad * 8, 0);
    for (int c = 0; c < chunks; ++c)
        for (int kg = 0; kg < 8; ++kg)
            for (int n = 0; n < nout; ++n)
                for (int j = 0; j < 8; ++j) {
                    int k = 64 * c + 8 * kg + j;
                    if (k < nin) out[base + ((((size_t)c * 8 + kg) * npad + n) * 8 + j)] = f2bf(w[(size_t)n * nin + k]);
                }
}
static int upload_bytes(const void *src, size_t bytes, uint8_t **dev) {
    if (*dev) { cudaFree(*dev); *dev = nullptr; }
    CK(cudaMalloc((void **)dev, bytes));
    CK(cudaMemcpy(*dev, src, bytes, cudaMemcpyHostToDevice));
    return GZ_OK;
}

static int upload(Packer &pk, float **dev) {
    if (*dev) { cudaFree(*dev); *dev = nullptr; }
    CK(cudaMalloc((void **)dev, pk.host.size() * sizeof(float)));
    CK(cudaMemcpy(*dev, pk.host.data(), pk.host.size() * sizeof(float), cudaMemcpyHostToDevice));
    return GZ_OK;
}

extern "C" int gz_set_pv_weights(gz_handle h, const float *f, int64_t n) {
    if (!h || !f) return fail(GZ_ERR_ARG, "gz_set_pv_weights: null argument");
    if (n != gz_pv_param_count()) return fail(GZ_ERR_ARG, "gz_set_pv_weights: wrong parameter count");
    Engine *e = reinterpret_cast<Engine *>(h);
    CK(cudaSetDevice(e->cfg.device));
    Packer pk;
    size_t o[32];
    const float *p = f;
    pack_conv(pk, p, p + 3456, p + 3456 + 128, 128, 3, 3, o[0], o[1]);
    p += 3456 + 128 + 512;
    for (int blk = 0; blk < 2; ++blk) {
        const float *c1w = p, *c1b = p + 147456, *c2w = c1b + 128, *c2b = c2w + 147456;
        const float *bn1 = c2b + 128, *bn2 = bn1 + 512;
        pack_conv(pk, c1w, c1b, bn1, 128, 128, 3, o[2 + 4 * blk], o[3 + 4 * blk]);
        pack_conv(pk, c2w, c2b, bn2, 128, 128, 3, o[4 + 4 * blk], o[5 + 4 * blk]);
        p = bn2 + 512;
    }
    pack_head_conv(pk, p, p + 256, 2, 128, o[10], o[11]); p += 256 + 2;
    pack_linear_t(pk, p, p + 101250, 225, 450, o[12], o[13]); p += 101250 + 225;
    pack_head_conv(pk, p, p + 128, 1, 128, o[14], o[15]); p += 128 + 1;
    pack_linear_t(pk, p, p + 14400, 64, 225, o[16], o[17]); p += 14400 + 64;
    size_t ow2 = pk.take(64), ob2 = pk.take(1);
    memcpy(&pk.host[ow2], p, 64 * sizeof(float));
    pk.host[ob2] = p[64];
    float *d = e->pv.conv0.w;   // base of the previous PV blob (if any)
    int rc = upload(pk, &d);
    if (rc) return rc;
    PvWeights &w = e->pv;
    w.conv0 = {d + o[0], d + o[1]};
    for (int i = 0; i < 4; ++i) w.tower[i] = {d + o[2 + 2 * i], d + o[3 + 2 * i]};
    w.pconv_w = d + o[10]; w.pconv_b = d + o[11];
    w.pfc_wt = d + o[12]; w.pfc_b = d + o[13];
    w.vconv_w = d + o[14]; w.vconv_b = d + o[15];
    w.vfc1_wt = d + o[16]; w.vfc1_b = d + o[17];
    w.vfc2_w = d + ow2; w.vfc2_b = d + ob2;
    {   // bf16 packs of the four tower convolutions
        std::vector<uint16_t> bf;
        const float *q = f + 3456 + 128 + 512;
        size_t off[4];
        for (int blk = 0; blk < 2; ++blk) {
            const float *c1w = q, *c1b = q + 147456, *c2w = c1b + 128, *c2b = c2w + 147456;
            const float *bn1 = c2b + 128, *bn2 = bn1 + 512;
            off[2 * blk] = bf.size() * 2;
            pack_conv_bf16(bf, c1w, bn1, 128, 128, 3);
            off[2 * blk + 1] = bf.size() * 2;
            pack_conv_bf16(bf, c2w, bn2, 128, 128, 3);
            q = bn2 + 512;
        }
        size_t off0 = bf.size() * 2;
        pack_conv0_bf16(bf, f, f + 3456 + 128, 128);
        rc = upload_bytes(bf.data(), bf.size() * 2, &e->bw.pv_blob);
        if (rc) return rc;
        for (int i = 0; i < 4; ++i) e->bw.pv_tower_w[i] = e->bw.pv_blob + off[i];
        e->bw.pv_conv0_w = e->bw.pv_blob + off0;
    }
    {   // head FCs for gz_nn_fc.cu: policy 450 -> 225 (8 chunks, N 256), value 225 -> 64 (4 chunks, N 64), bias [256]
        const float *q = f + 3456 + 128 + 512 + 2 * (2 * (147456 + 128) + 2 * 512);
        const float *pfc_w = q + 256 + 2, *pfc_b = pfc_w + 101250;
        const float *vfc1_w = pfc_b + 225 + 128 + 1;
        std::vector<uint16_t> bf;
        pack_fc_bf16(bf, pfc_w, 225, 450, 8, 256);
        size_t off_v = bf.size() * 2;
        pack_fc_bf16(bf, vfc1_w, 64, 225, 4, 64);
        size_t off_b = bf.size() * 2;
        bf.resize(bf.size() + 512, 0);
        memcpy(reinterpret_cast<uint8_t *>(bf.data()) + off_b, pfc_b, 225 * sizeof(float));
        rc = upload_bytes(bf.data(), bf.size() * 2, &e->bw.pv_head_blob);
        if (rc) return rc;
        e->bw.pv_pfc_w = e->bw.pv_head_blob;
        e->bw.pv_vfc1_w = e->bw.pv_head_blob + off_v;
        e->bw.pv_pfc_b = reinterpret_cast<float *>(e->bw.pv_head_blob + off_b);
    }
    e->have_pv = true;
    return GZ_OK;
}

extern "C" int gz_set_planner_weights(gz_handle h, const float *gnf, int64_t ngn, const float *dqf,
                                      int64_t ndq) {
    if (!h || !gnf || !dqf) return fail(GZ_ERR_ARG, "gz_set_planner_weights: null argument");
    if (ngn != gz_graphnet_param_count() || ndq != gz_dqn_param_count())
        return fail(GZ_ERR_ARG, "gz_set_planner_weights: wrong parameter count");
    Engine *e = reinterpret_cast<Engine *>(h);
    CK(cudaSetDevice(e->cfg.device));
    {
        Packer pk;
        size_t o[24];
        const float *p = gnf;
        pack_conv(pk, p, p + 1728, nullptr, 64, 3, 3, o[0], o[1]); p += 1728 + 64;
        for (int i = 0; i < 4; ++i) {
            pack_conv(pk, p, p + 36864, nullptr, 64, 64, 3, o[2 + 4 * i], o[3 + 4 * i]); p += 36864 + 64;
            pack_conv(pk, p, p + 4096, nullptr, 64, 64, 1, o[4 + 4 * i], o[5 + 4 * i]); p += 4096 + 64;
        }
        pack_head_conv(pk, p, p + 128, 2, 64, o[18], o[19]); p += 128 + 2;
        pack_linear_t(pk, p, p + 101250, 225, 450, o[20], o[21]);
        float *d = e->gn.embed.w;
        int rc = upload(pk, &d);
        if (rc) return rc;
        GraphNetWeights &w = e->gn;
        w.embed = {d + o[0], d + o[1]};
        for (int i = 0; i < 4; ++i) {
            w.l3[i] = {d + o[2 + 4 * i], d + o[3 + 4 * i]};
            w.l1[i] = {d + o[4 + 4 * i], d + o[5 + 4 * i]};
        }
        w.hconv_w = d + o[18]; w.hconv_b = d + o[19];
        w.hfc_wt = d + o[20]; w.hfc_b = d + o[21];
        for (int i = 0; i < 4; ++i) {   // k_gn_block takes its biases as a kernel parameter
            memcpy(e->bw.gn_host_bias[i], &pk.host[o[3 + 4 * i]], 64 * sizeof(float));
            memcpy(e->bw.gn_host_bias[i] + 64, &pk.host[o[5 + 4 * i]], 64 * sizeof(float));
        }
        memcpy(e->bw.gn_host_head, &pk.host[o[18]], 128 * sizeof(float));
        memcpy(e->bw.gn_host_head + 128, &pk.host[o[19]], 2 * sizeof(float));
        std::vector<uint16_t> bf;
        size_t o3[4], o1[4], o3p[4];
        const float *q = gnf + 1728 + 64;
        for (int i = 0; i < 4; ++i) {
            o3p[i] = bf.size() * 2;
            pack_gn3x3_pairs(bf, q);
            o3[i] = bf.size() * 2;
            pack_conv_bf16(bf, q, nullptr, 64, 64, 3); q += 36864 + 64;
            o1[i] = bf.size() * 2;
            pack_conv_bf16(bf, q, nullptr, 64, 64, 1); q += 4096 + 64;
        }
        size_t off0 = bf.size() * 2;
        pack_conv0_bf16(bf, gnf, nullptr, 64);
        size_t off_i2c = bf.size() * 2;
        pack_conv0_im2col_bf16(bf, gnf, nullptr, 64);
        memcpy(e->bw.gn_host_embed_bias, &pk.host[o[1]], 64 * sizeof(float));
        rc = upload_bytes(bf.data(), bf.size() * 2, &e->bw.gn_blob);
        if (rc) return rc;
        e->bw.gn_embed_i2c_w = e->bw.gn_blob + off_i2c;
        for (int i = 0; i < 4; ++i) {
            e->bw.gn_l3_w[i] = e->bw.gn_blob + o3[i]; e->bw.gn_l1_w[i] = e->bw.gn_blob + o1[i];
            e->bw.gn_l3p_w[i] = e->bw.gn_blob + o3p[i];
        }
        e->bw.gn_embed_w = e->bw.gn_blob + off0;
        {   // head FC 450 -> 225 for gz_nn_fc.cu
            const float *hfc_w = gnf + 1728 + 64 + 4 * (36864 + 64 + 4096 + 64) + 128 + 2, *hfc_b = hfc_w + 101250;
            std::vector<uint16_t> hb;
            pack_fc_bf16(hb, hfc_w, 225, 450, 8, 256);
            size_t off_b = hb.size() * 2;
            hb.resize(hb.size() + 512, 0);
            memcpy(reinterpret_cast<uint8_t *>(hb.data()) + off_b, hfc_b, 225 * sizeof(float));
            rc = upload_bytes(hb.data(), hb.size() * 2, &e->bw.gn_head_blob);
            if (rc) return rc;
            e->bw.gn_hfc_w = e->bw.gn_head_blob;
            e->bw.gn_hfc_b = reinterpret_cast<float *>(e->bw.gn_head_blob + off_b);
        }
    }
    {
        Packer pk;
        size_t o[6];
        const float *p = dqf;
        pack_linear_t(pk, p, p + 256 * 675, 256, 675, o[0], o[1]); p += 256 * 675 + 256;
        pack_linear_t(pk, p, p + 65536, 256, 256, o[2], o[3]); p += 65536 + 256;
        pack_linear_t(pk, p, p + 225 * 256, 225, 256, o[4], o[5]);
        float *d = e->dqn.w1t;
        int rc = upload(pk, &d);
        if (rc) return rc;
        e->dqn = {d + o[0], d + o[1], d + o[2], d + o[3], d + o[4], d + o[5]};
        std::vector<uint16_t> bf;
        std::vector<float> bias(3 * 256, 0.f);
        const float *q = dqf;
        pack_fc_bf16(bf, q, 256, 675, 11); memcpy(&bias[0], q + 256 * 675, 256 * sizeof(float)); q += 256 * 675 + 256;
        pack_fc_bf16(bf, q, 256, 256, 4); memcpy(&bias[256], q + 65536, 256 * sizeof(float)); q += 65536 + 256;
        pack_fc_bf16(bf, q, 225, 256, 4); memcpy(&bias[512], q + 225 * 256, 225 * sizeof(float));
        rc = upload_bytes(bf.data(), bf.size() * 2, &e->bw.dqn_w);
        if (rc) return rc;
        rc = upload_bytes(bias.data(), bias.size() * sizeof(float), reinterpret_cast<uint8_t **>(&e->bw.dqn_bias));
        if (rc) return rc;
    }
    e->have_planner = true;
    return GZ_OK;
}

extern "C" int gz_set_tables(gz_handle h, const double *hl, int32_t nl, const double *ht, int32_t nt) {
    if (!h) return fail(GZ_ERR_ARG, "gz_set_tables: null handle");
    Engine *e = reinterpret_cast<Engine *>(h);
    CK(cudaSetDevice(e->cfg.device));
    if (nl > kLogTableMax) nl = kLogTableMax;
    if (nt > kTanhTableMax) nt = kTanhTableMax;
    if (hl && nl > 0) CK(cudaMemcpy(e->log_table, hl, nl * sizeof(double), cudaMemcpyHostToDevice));
    if (ht && nt > 0) CK(cudaMemcpy(e->tanh_table, ht, nt * sizeof(double), cudaMemcpyHostToDevice));
    e->n_log = hl ? nl : 0;
    e->n_tanh = ht ? nt : 0;
    return GZ_OK;
}

// ---- helpers -------------------------------------------------------------------------------------
static inline Engine *eng(gz_handle h) { return reinterpret_cast<Engine *>(h); }
static inline cudaStream_t strm(void *s) { return reinterpret_cast<cudaStream_t>(s); }

static int check_batch(Engine *e, int G, const char *who) {
    if (!e) return fail(GZ_ERR_ARG, "%s: null handle", who);
    if (G < 1 || G > e->cfg.max_games) return fail(GZ_ERR_ARG, "%s: G out of range (max_games)", who);
    cudaError_t ce = cudaSetDevice(e->cfg.device);
    if (ce != cudaSuccess) return fail(GZ_ERR_CUDA, "%s: %s", who, cudaGetErrorString(ce));
    return GZ_OK;
}

static SearchCfg make_cfg(const gz_search_params *p) {
    SearchCfg c;
    c.S = p->num_simulations;
    c.planner_steps = p->planner_steps < 0 ? 0 : p->planner_steps;
    c.planner_k = p->planner_k;
    c.deterministic = p->deterministic;
    c.eval_nodes = p->eval_nodes;
    c.c_puct = p->c_puct;
    c.beta = p->beta;
    c.alpha = p->planner_alpha;
    c.explore = p->planner_explore;
    c.seed_lo = (uint32_t)(p->seed & 0xFFFFFFFFull);
    c.seed_hi = (uint32_t)(p->seed >> 32);
    c.game_base = p->game_base;
    c.d_sim = nullptr;
    return c;
}

// ---- boards --------------------------------------------------------------------------------------
extern "C" int gz_boards_from_cells(gz_handle h, const uint8_t *d_cells, const uint8_t *d_to_move,
                                    const uint16_t *d_mc, int32_t G, gz_board *d_boards, void *stream) {
    Engine *e = eng(h);
    int rc = check_batch(e, G, "gz_boards_from_cells");
    if (rc) return rc;
    if (!d_cells || !d_to_move || !d_mc || !d_boards) return fail(GZ_ERR_ARG, "gz_boards_from_cells: null pointer");
    launch_boards_from_cells(d_cells, d_to_move, d_mc, G, d_boards, strm(stream));
    e->launches += 1;
    CKL();
    return GZ_OK;
}

extern "C" int gz_boards_to_cells(gz_handle h, const gz_board *d_boards, int32_t G, uint8_t *d_cells,
                                  void *stream) {
    Engine *e = eng(h);
    int rc = check_batch(e, G, "gz_boards_to_cells");
    if (rc) return rc;
    if (!d_boards || !d_cells) return fail(GZ_ERR_ARG, "gz_boards_to_cells: null pointer");
    launch_boards_to_cells(d_boards, G, d_cells, strm(stream));
    e->launches += 1;
    CKL();
    return GZ_OK;
}

extern "C" int gz_make_moves(gz_handle h, gz_board *d_boards, const int32_t *d_moves, int32_t G,
                             uint8_t *d_ok, void *stream) {
    Engine *e = eng(h);
    int rc = check_batch(e, G, "gz_make_moves");
    if (rc) return rc;
    if (!d_boards || !d_moves) return fail(GZ_ERR_ARG, "gz_make_moves: null pointer");
    launch_make_moves(d_boards, d_moves, G, d_ok, strm(stream));
    e->launches += 1;
    CKL();
    return GZ_OK;
}

extern "C" int gz_board_ops(gz_handle h, const gz_board *d_boards, int32_t G, uint32_t *d_legal,
                            uint32_t *d_win, uint32_t *d_three, int32_t *d_threat, int64_t *d_pattern,
                            void *stream) {
    Engine *e = eng(h);
    int rc = check_batch(e, G, "gz_board_ops");
    if (rc) return rc;
    if (!d_boards) return fail(GZ_ERR_ARG, "gz_board_ops: null pointer");
    launch_board_ops(d_boards, G, d_legal, d_win, d_three, d_threat, d_pattern, strm(stream));
    e->launches += 1;
    CKL();
    return GZ_OK;
}

// ---- networks ------------------------------------------------------------------------------------
extern "C" int gz_pv_forward(gz_handle h, const gz_board *d_boards, int32_t G, float *d_logits,
                             float *d_probs, float *d_value, void *stream) {
    Engine *e = eng(h);
    int rc = check_batch(e, G, "gz_pv_forward");
    if (rc) return rc;
    if (!d_boards) return fail(GZ_ERR_ARG, "gz_pv_forward: null pointer");
    if (!e->have_pv) return fail(GZ_ERR_STATE, "gz_pv_forward: gz_set_pv_weights was not called");
    nn_pv_forward(*e, d_boards, nullptr, G, d_logits, d_probs, d_value, nullptr, nullptr, strm(stream));
    CKL();
    return GZ_OK;
}

extern "C" int gz_pv_forward_static(gz_handle h, const gz_board *d_boards, int32_t G, float *d_logits,
                                    float *d_probs, float *d_value, void *stream) {
    Engine *e = eng(h);
    if (!e) return fail(GZ_ERR_ARG, "gz_pv_forward_static: null handle");
    if (!d_boards) return fail(GZ_ERR_ARG, "gz_pv_forward_static: null pointer");
    if (G < 1) return fail(GZ_ERR_ARG, "gz_pv_forward_static: G must be at least 1");
    if (G > GZ_STATIC_MAX_BOARDS)
        return fail(GZ_ERR_ARG, "gz_pv_forward_static: G exceeds GZ_STATIC_MAX_BOARDS (16): the static scratch holds "
                                "no more; use gz_pv_forward");
    if (!e->have_pv) return fail(GZ_ERR_STATE, "gz_pv_forward_static: gz_set_pv_weights was not called");
    CK(cudaSetDevice(e->cfg.device));
    CK(nn_pv_forward_static(*e, d_boards, G, d_logits, d_probs, d_value, strm(stream)));
    return GZ_OK;
}

extern "C" int gz_planner_nets_forward(gz_handle h, const gz_board *d_boards, int32_t G,
                                       float *d_gn_logits, float *d_gn_policy, float *d_dqn_q,
                                       void *stream) {
    Engine *e = eng(h);
    int rc = check_batch(e, G, "gz_planner_nets_forward");
    if (rc) return rc;
    if (!d_boards) return fail(GZ_ERR_ARG, "gz_planner_nets_forward: null pointer");
    if (!e->have_planner) return fail(GZ_ERR_STATE, "gz_planner_nets_forward: planner weights not set");
    nn_planner_forward(*e, d_boards, G, d_gn_logits ? d_gn_logits : e->gn_logits,
                       d_gn_policy ? d_gn_policy : e->gn_policy, d_gn_policy ? kCells : kPolicyPitch,
                       d_dqn_q ? d_dqn_q : e->dqn_q, strm(stream));
    CKL();
    return GZ_OK;
}

extern "C" int gz_debug_pv_layer(gz_handle h, const gz_board *d_boards, int32_t G, int32_t nlayers,
                                 float *d_out, void *stream) {
    Engine *e = eng(h);
    int rc = check_batch(e, G, "gz_debug_pv_layer");
    if (rc) return rc;
    if (!d_boards || !d_out || nlayers < 0 || nlayers > 4) return fail(GZ_ERR_ARG, "gz_debug_pv_layer: bad argument");
    if (!e->have_pv) return fail(GZ_ERR_STATE, "gz_debug_pv_layer: gz_set_pv_weights was not called");
    if (e->cfg.precision == 1) nn_debug_pv_layer_bf16(*e, d_boards, G, nlayers, d_out, strm(stream));
    else nn_debug_pv_layer_fp32(*e, d_boards, G, nlayers, d_out, strm(stream));
    CKL();
    return GZ_OK;
}

// ---- planner move --------------------------------------------------------------------------------
extern "C" int gz_planner_move(gz_handle h, const gz_board *d_boards, const uint8_t *d_ai_player,
                               int32_t G, const gz_search_params *p, int32_t sim, int32_t step,
                               int64_t *d_scores2, int32_t *d_topk, double *d_composed, int32_t *d_move,
                               void *stream) {
    Engine *e = eng(h);
    int rc = check_batch(e, G, "gz_planner_move");
    if (rc) return rc;
    if (!d_boards || !d_ai_player || !p) return fail(GZ_ERR_ARG, "gz_planner_move: null pointer");
    if (!e->have_planner) return fail(GZ_ERR_STATE, "gz_planner_move: planner weights not set");
    SearchCfg c = make_cfg(p);
    cudaStream_t s = strm(stream);
    int32_t *topk = d_topk ? d_topk : e->sim.topk;
    launch_planner_score(d_boards, d_ai_player, nullptr, G, c, step, topk, e->sim.ntop, d_scores2, s);
    nn_planner_forward(*e, d_boards, G, e->gn_logits, e->gn_policy, kPolicyPitch, e->dqn_q, s);
    launch_planner_pick(const_cast<gz_board *>(d_boards), d_ai_player, nullptr, G, c, sim, step, topk,
                        e->sim.ntop, e->gn_policy, e->dqn_q, d_composed, d_move, false, nullptr, s);
    e->launches += 2;
    CKL();
    return GZ_OK;
}

// ---- rollout -------------------------------------------------------------------------------------
extern "C" int gz_rollout_policy(gz_handle h, const gz_board *d_boards, int32_t G,
                                 const gz_search_params *p, int32_t sim, int32_t step, int32_t *d_move,
                                 uint32_t *d_cand, void *stream) {
    Engine *e = eng(h);
    int rc = check_batch(e, G, "gz_rollout_policy");
    if (rc) return rc;
    if (!d_boards || !p || !d_move) return fail(GZ_ERR_ARG, "gz_rollout_policy: null pointer");
    launch_rollout_policy(d_boards, G, make_cfg(p), sim, step, d_move, d_cand, strm(stream));
    e->launches += 1;
    CKL();
    return GZ_OK;
}

// planner prefix + tail on e->sim.board / e->sim.steps (already initialised by the caller)
static void run_playouts(Engine *e, int G, const SearchCfg &c, int sim, cudaStream_t s) {
    for (int k = 0; k < c.planner_steps; ++k) {
        prof_begin(*e, GZ_STAGE_PLAN_SCORE, 1, s);
        launch_planner_score(e->sim.board, e->tree.ai_player, e->sim.steps, G, c, k, e->sim.topk,
                             e->sim.ntop, nullptr, s);
        prof_end(*e, s);
        nn_planner_forward(*e, e->sim.board, G, e->gn_logits, e->gn_policy, kPolicyPitch, e->dqn_q, s);
        prof_begin(*e, GZ_STAGE_PLAN_PICK, 1, s);
        launch_planner_pick(e->sim.board, e->tree.ai_player, e->sim.steps, G, c, sim, -1, e->sim.topk,
                            e->sim.ntop, e->gn_policy, e->dqn_q, nullptr, nullptr, true, e->counters, s);
        prof_end(*e, s);
    }
    prof_begin(*e, GZ_STAGE_ROLLOUT, 1, s);
    launch_rollout_tail(e->sim.board, e->tree.ai_player, e->sim.steps, G, c, sim, e->sim.value,
                        e->counters, s);
    prof_end(*e, s);
}

extern "C" int gz_rollout(gz_handle h, const gz_board *d_boards, const uint8_t *d_ai_player, int32_t G,
                          const gz_search_params *p, int32_t sim, double *d_value, int32_t *d_steps,
                          gz_board *d_final, void *stream) {
    Engine *e = eng(h);
    int rc = check_batch(e, G, "gz_rollout");
    if (rc) return rc;
    if (!d_boards || !d_ai_player || !p) return fail(GZ_ERR_ARG, "gz_rollout: null pointer");
    SearchCfg c = make_cfg(p);
    if (c.planner_steps > 0 && !e->have_planner) return fail(GZ_ERR_STATE, "gz_rollout: planner weights not set");
    cudaStream_t s = strm(stream);
    CK(cudaMemcpyAsync(e->sim.board, d_boards, sizeof(gz_board) * G, cudaMemcpyDeviceToDevice, s));
    CK(cudaMemcpyAsync(e->tree.ai_player, d_ai_player, G, cudaMemcpyDeviceToDevice, s));
    CK(cudaMemsetAsync(e->sim.steps, 0, sizeof(int32_t) * G, s));
    run_playouts(e, G, c, sim, s);
    if (d_value) CK(cudaMemcpyAsync(d_value, e->sim.value, sizeof(double) * G, cudaMemcpyDeviceToDevice, s));
    if (d_steps) CK(cudaMemcpyAsync(d_steps, e->sim.steps, sizeof(int32_t) * G, cudaMemcpyDeviceToDevice, s));
    if (d_final) CK(cudaMemcpyAsync(d_final, e->sim.board, sizeof(gz_board) * G, cudaMemcpyDeviceToDevice, s));
    CKL();
    return GZ_OK;
}

// ---- search --------------------------------------------------------------------------------------
// View of the engine restricted to games [g0, g0 + n): every per-game pointer is offset, weights and
// tables are shared.  Games are independent, so lanes never touch each other's data.
static Engine *lane_view(Engine *e, int l, size_t g0) {
    Engine *v = e->lane[l];
    v->root = e;
    v->cfg = e->cfg;
    v->sm_count = e->sm_count;
    v->have_pv = e->have_pv; v->have_planner = e->have_planner;
    v->pv = e->pv; v->gn = e->gn; v->dqn = e->dqn; v->bw = e->bw;
    v->log_table = e->log_table; v->tanh_table = e->tanh_table; v->n_log = e->n_log; v->n_tanh = e->n_tanh;
    v->counters = e->counters;
    v->d_sim = e->d_sim;
    const size_t cap = e->tree.cap, n0 = g0 * cap;
    const TreeArrays &t = e->tree;
    v->tree = {t.cap, t.parent + n0, t.move + n0, t.status + n0, t.nchild + n0, t.nempty + n0,
               t.visits + n0, t.value + n0, t.bgb + n0, t.vhead + n0,
               t.prior ? t.prior + n0 * kPolicyPitch : nullptr, t.n_nodes + g0, t.root + g0, t.ai_player + g0};
    const SimArrays &m = e->sim;
    v->sim = {m.board + g0, m.leaf + g0, m.need_eval + g0, m.steps + g0, m.value + g0,
              m.topk + g0 * kMaxTopK, m.ntop + g0, nullptr};
    for (int i = 0; i < 3; ++i) {
        v->planes[i] = e->planes[i] ? e->planes[i] + g0 * 16 * kPlaneBytes : nullptr;
        v->act[i] = e->act[i] ? e->act[i] + g0 * 128 * kCells : nullptr;
    }
    v->gn_logits = e->gn_logits + g0 * kCells; v->gn_policy = e->gn_policy + g0 * kPolicyPitch;
    v->pv_pc = e->pv_pc + (g0 >> 7) * kHeadTileBytesP; v->pv_vc = e->pv_vc + (g0 >> 7) * kHeadTileBytesV;   // g0 % 128 == 0
    v->gn_pc = e->gn_pc + (g0 >> 7) * kHeadTileBytesP;
    v->dqn_q = e->dqn_q + g0 * kCells;
    v->pv_logits = e->pv_logits + g0 * kCells; v->pv_probs = e->pv_probs + g0 * kCells;
    v->pv_value = e->pv_value + g0;
    return v;
}

static void search_lane(Engine *v, const gz_board *d_boards, const uint8_t *d_ai, int G, SearchCfg c, cudaStream_t s) {
    prof_begin(*v, GZ_STAGE_TREE, 1, s);
    launch_tree_init(*v, d_boards, d_ai, G, s);
    prof_end(*v, s);
    if (c.eval_nodes)
        nn_pv_forward(*v, v->sim.board, v->sim.need_eval, G, nullptr, nullptr, nullptr, v->tree.prior,
                      v->sim.leaf, s);
}

static int search_device(Engine *e, const gz_board *d_boards, const uint8_t *d_ai, int G,
                         const gz_search_params *p, int32_t *d_cv, double *d_cval, int32_t *d_best,
                         cudaStream_t s) {
    SearchCfg c = make_cfg(p);
    if (c.S < 1 || c.S > e->cfg.max_simulations) return fail(GZ_ERR_ARG, "gz_search: num_simulations out of range");
    if (c.eval_nodes && !e->have_pv) return fail(GZ_ERR_STATE, "gz_search: gz_set_pv_weights was not called");
    if (c.planner_steps > 0 && !e->have_planner) return fail(GZ_ERR_STATE, "gz_search: planner weights not set");
    // automatic: one lane.  Since the tensor kernels issue at full rate each of them fills the machine and
    // two to four lanes measure within 2 % of one (profiles/r01_lanes.md); more lanes stay available on request
    int nl = e->cfg.lanes > 0 ? e->cfg.lanes : 1;
    if (nl > G / 128) nl = G / 128 > 1 ? G / 128 : 1;   // every lane owns at least one whole tile of 128 boards
    // Small batches are launch-bound (~50 kernels per wave, microseconds of work each: AlphaZeroGomokuAI.get_move on
    // one board): wave 0 runs eagerly, one wave is captured into a CUDA graph whose kernels read the simulation index
    // from device memory, and the graph is replayed for the remaining simulations.  The legacy default stream cannot
    // be captured: the search then runs on the engine's own stream, ordered after / before the caller's by events.
    const bool graphed = nl == 1 && !e->profiling && G <= kGraphMaxGames && c.S >= 4 && !e->no_graph;
    cudaStream_t s_user = s;
    const bool redirect = graphed && (s == nullptr || s == cudaStreamLegacy || s == cudaStreamPerThread);
    if (redirect) {
        CK(cudaEventRecord(e->lane_go, s_user));
        s = e->own_stream;
        CK(cudaStreamWaitEvent(s, e->lane_go, 0));
    }
    Engine *v[kMaxLanes];
    cudaStream_t ls[kMaxLanes];
    int g0[kMaxLanes + 1];
    SearchCfg lc[kMaxLanes];
    // lane boundaries on multiples of 128 boards: the head GEMMs work on tiles of 128 rows
    for (int l = 0; l <= nl; ++l) g0[l] = l == nl ? G : (int)((long long)G * l / nl) & ~127;
    if (nl > 1) CK(cudaEventRecord(e->lane_go, s));
    for (int l = 0; l < nl; ++l) {
        v[l] = lane_view(e, l, (size_t)g0[l]);
        ls[l] = nl > 1 ? e->lane_stream[l] : s;
        lc[l] = c;
        lc[l].game_base = c.game_base + (uint32_t)g0[l];   // Philox streams keyed by the global game id
        if (nl > 1) CK(cudaStreamWaitEvent(ls[l], e->lane_go, 0));
        search_lane(v[l], d_boards + g0[l], d_ai + g0[l], g0[l + 1] - g0[l], lc[l], ls[l]);
    }
    // One wave = one simulation of every tree: select/expand (+ backup of the previous wave), policy-value forward of
    // the new nodes, planner-guided playouts.  Waves are issued lane by lane so that the lanes' kernels interleave in
    // the hardware queues.
    auto wave = [&](int sim, const SearchCfg *cfgs) {
        for (int l = 0; l < nl; ++l) {
            const int n = g0[l + 1] - g0[l];
            prof_begin(*v[l], GZ_STAGE_TREE, 1, ls[l]);
            launch_tree_step(*v[l], n, cfgs[l], sim, ls[l]);
            prof_end(*v[l], ls[l]);
            if (c.eval_nodes)
                nn_pv_forward(*v[l], v[l]->sim.board, v[l]->sim.need_eval, n, nullptr, nullptr, nullptr,
                              v[l]->tree.prior, v[l]->sim.leaf, ls[l]);
            run_playouts(v[l], n, cfgs[l], sim, ls[l]);
        }
    };
    // AlphaZeroGomokuAI.time_limit (ai_agent.py:183-189): checked on the host every kTimeCheckWaves waves
    const bool timed = p->time_limit_s > 0.0 && std::isfinite(p->time_limit_s);
    const auto t_start = std::chrono::steady_clock::now();
    constexpr int kTimeCheckWaves = 4;
    auto out_of_time = [&]() -> bool {
        for (int l = 0; l < nl; ++l) cudaStreamSynchronize(ls[l]);
        return std::chrono::duration<double>(std::chrono::steady_clock::now() - t_start).count() > p->time_limit_s;
    };
    int done = 0;   // simulations started
    if (graphed) {
        wave(0, lc);
        done = 1;
        SearchCfg gc = lc[0];
        gc.d_sim = e->d_sim;
        launch_set_sim(e->d_sim, 1, 0, s);
        e->launches += 1;
        const int64_t before = e->launches;
        cudaGraph_t graph = nullptr;
        cudaGraphExec_t exec = nullptr;
        CK(cudaStreamBeginCapture(s, cudaStreamCaptureModeThreadLocal));
        wave(-1, &gc);
        launch_set_sim(e->d_sim, 1, 1, s);
        cudaError_t ce = cudaStreamEndCapture(s, &graph);
        if (ce != cudaSuccess) { cudaGetLastError(); return fail(GZ_ERR_CUDA, "gz_search: graph capture: %s", cudaGetErrorString(ce)); }
        const int64_t per_wave = e->launches - before + 1;   // + the set_sim node
        e->launches = before;                                // nothing ran during the capture
        ce = cudaGraphInstantiate(&exec, graph, 0);
        if (ce != cudaSuccess) { cudaGraphDestroy(graph); return fail(GZ_ERR_CUDA, "gz_search: graph instantiate: %s", cudaGetErrorString(ce)); }
        for (int sim = 1; sim < c.S; ++sim) {
            if (timed && sim % kTimeCheckWaves == 0 && out_of_time()) break;
            ce = cudaGraphLaunch(exec, s);
            if (ce != cudaSuccess) break;
            e->launches += per_wave;
            done = sim + 1;
        }
        cudaGraphExecDestroy(exec);
        cudaGraphDestroy(graph);
        if (ce != cudaSuccess) return fail(GZ_ERR_CUDA, "gz_search: graph launch: %s", cudaGetErrorString(ce));
    } else {
        for (int sim = 0; sim < c.S; ++sim) {
            if (timed && sim > 0 && sim % kTimeCheckWaves == 0 && out_of_time()) break;
            wave(sim, lc);
            done = sim + 1;
        }
    }
    if (done < c.S) {   // cut short by the time limit: the closing launch below backs up simulation done - 1 and stops
        c.S = done;
        for (int l = 0; l < nl; ++l) lc[l].S = done;
    }
    e->last_simulations = done;
    for (int l = 0; l < nl; ++l) {
        const int n = g0[l + 1] - g0[l];
        prof_begin(*v[l], GZ_STAGE_TREE, 2, ls[l]);
        launch_tree_step(*v[l], n, lc[l], c.S, ls[l]);   // backs up the last simulation
        launch_tree_finalize(*v[l], n, lc[l], d_cv ? d_cv + (size_t)g0[l] * kCells : nullptr,
                             d_cval ? d_cval + (size_t)g0[l] * kCells : nullptr,
                             d_best ? d_best + g0[l] : nullptr, ls[l]);
        prof_end(*v[l], ls[l]);
        if (nl > 1) {
            CK(cudaEventRecord(e->lane_done[l], ls[l]));
            CK(cudaStreamWaitEvent(s, e->lane_done[l], 0));
        }
    }
    if (redirect) {
        CK(cudaEventRecord(e->lane_done[0], s));
        CK(cudaStreamWaitEvent(s_user, e->lane_done[0], 0));
    }
    CKL();
    e->searched = true;
    return GZ_OK;
}

extern "C" int gz_search(gz_handle h, const gz_board *d_boards, const uint8_t *d_ai_player, int32_t G,
                         const gz_search_params *p, int32_t *d_child_visits, double *d_child_value,
                         int32_t *d_best_move, void *stream) {
    Engine *e = eng(h);
    int rc = check_batch(e, G, "gz_search");
    if (rc) return rc;
    if (!d_boards || !d_ai_player || !p) return fail(GZ_ERR_ARG, "gz_search: null pointer");
    return search_device(e, d_boards, d_ai_player, G, p, d_child_visits, d_child_value, d_best_move, strm(stream));
}

extern "C" int gz_search_host(gz_handle h, const uint8_t *h_cells, const uint8_t *h_to_move,
                              const uint16_t *h_mc, const uint8_t *h_ai, int32_t G,
                              const gz_search_params *p, int32_t *h_cv, double *h_cval, int32_t *h_best) {
    Engine *e = eng(h);
    int rc = check_batch(e, G, "gz_search_host");
    if (rc) return rc;
    if (!h_cells || !h_to_move || !h_mc || !h_ai || !p) return fail(GZ_ERR_ARG, "gz_search_host: null pointer");
    cudaStream_t s = e->own_stream;
    uint8_t *d_cells = e->tmp_u8, *d_tm = e->tmp_u8 + (size_t)225 * e->cfg.max_games;
    uint8_t *d_ai = d_tm + e->cfg.max_games;
    uint16_t *d_mc = reinterpret_cast<uint16_t *>(e->tmp_u8 + (size_t)228 * e->cfg.max_games);
    CK(cudaMemcpyAsync(d_cells, h_cells, (size_t)G * kCells, cudaMemcpyHostToDevice, s));
    CK(cudaMemcpyAsync(d_tm, h_to_move, G, cudaMemcpyHostToDevice, s));
    CK(cudaMemcpyAsync(d_ai, h_ai, G, cudaMemcpyHostToDevice, s));
    CK(cudaMemcpyAsync(d_mc, h_mc, (size_t)G * 2, cudaMemcpyHostToDevice, s));
    launch_boards_from_cells(d_cells, d_tm, d_mc, G, e->tmp_boards, s);
    e->launches += 1;
    rc = search_device(e, e->tmp_boards, d_ai, G, p, e->io_cv, e->io_cval, e->io_best, s);
    if (rc) return rc;
    if (h_cv) CK(cudaMemcpyAsync(h_cv, e->io_cv, sizeof(int32_t) * G * kCells, cudaMemcpyDeviceToHost, s));
    if (h_cval) CK(cudaMemcpyAsync(h_cval, e->io_cval, sizeof(double) * G * kCells, cudaMemcpyDeviceToHost, s));
    if (h_best) CK(cudaMemcpyAsync(h_best, e->io_best, sizeof(int32_t) * G, cudaMemcpyDeviceToHost, s));
    CK(cudaStreamSynchronize(s));
    return GZ_OK;
}

// ---- dataset build -----------------------------------------------------------------------------------
extern "C" int gz_dataset_build(gz_handle h, const gz_replay_record *d_records, int64_t n, const int32_t *d_chosen,
                                int64_t m, const int64_t *d_sample_ids, int64_t first, int64_t count,
                                int32_t reference_exact, float *d_planes, int64_t *d_move, float *d_value,
                                void *stream) {
    Engine *e = eng(h);
    if (!e) return fail(GZ_ERR_ARG, "gz_dataset_build: null handle");
    if (n < 0 || m < 0 || count < 0 || first < 0) return fail(GZ_ERR_ARG, "gz_dataset_build: negative size");
    if (count == 0) return GZ_OK;
    if (!d_records || !d_planes || !d_move || !d_value) return fail(GZ_ERR_ARG, "gz_dataset_build: null pointer");
    if (m > 0 && !d_chosen) return fail(GZ_ERR_ARG, "gz_dataset_build: m > 0 without d_chosen");
    if (!d_sample_ids && first + count > n + 8 * m) return fail(GZ_ERR_ARG, "gz_dataset_build: range past n + 8 m");
    if (count > (1ll << 26)) return fail(GZ_ERR_ARG, "gz_dataset_build: more than 2^26 samples in one call");
    CK(cudaSetDevice(e->cfg.device));
    launch_dataset_build(d_records, n, d_chosen, m, reinterpret_cast<const long long *>(d_sample_ids), first, count,
                         reference_exact != 0, d_planes, reinterpret_cast<long long *>(d_move), d_value, strm(stream));
    e->launches += 1;
    CKL();
    return GZ_OK;
}

// ---- self-play -----------------------------------------------------------------------------------
static void sp_free(Engine *e) {
    void *ptrs[] = {e->sp.boards, e->sp.ai, e->sp.best, e->sp.game_id, e->sp.generation, e->sp.staging,
                    e->sp.out, e->sp.ctr};
    for (void *p : ptrs) if (p) cudaFree(p);
    e->sp = SelfPlayArrays();
    e->sp_alloc_games = e->sp_alloc_out = 0;
}

static int sp_begin(gz_handle h, int32_t G, const gz_search_params *p, double exploration, int64_t replay_capacity,
                    int32_t arena, int32_t sims_b, int64_t games_total, int64_t ordinal_stride, void *stream) {
    Engine *e = eng(h);
    int rc = check_batch(e, G, "gz_selfplay_begin");
    if (rc) return rc;
    if (!p) return fail(GZ_ERR_ARG, "gz_selfplay_begin: null params");
    if (p->num_simulations < 1 || p->num_simulations > e->cfg.max_simulations)
        return fail(GZ_ERR_ARG, "gz_selfplay_begin: num_simulations out of range");
    if (p->eval_nodes && !e->have_pv) return fail(GZ_ERR_STATE, "gz_selfplay_begin: gz_set_pv_weights was not called");
    if (p->planner_steps > 0 && !e->have_planner) return fail(GZ_ERR_STATE, "gz_selfplay_begin: planner weights not set");
    long long cap = replay_capacity > 0 ? replay_capacity : 256ll * G;
    if (cap < kMaxPlies) cap = kMaxPlies;
    if (e->sp_alloc_games < G || e->sp_alloc_out < cap) {
        CK(cudaDeviceSynchronize());
        sp_free(e);
        CK(dalloc(&e->sp.boards, (size_t)G)); CK(dalloc(&e->sp.ai, (size_t)G)); CK(dalloc(&e->sp.best, (size_t)G));
        CK(dalloc(&e->sp.game_id, (size_t)G)); CK(dalloc(&e->sp.generation, (size_t)G));
        CK(dalloc(&e->sp.staging, (size_t)G * kMaxPlies)); CK(dalloc(&e->sp.out, (size_t)cap));
        CK(dalloc(&e->sp.ctr, (size_t)SP_NCTR));
        e->sp_alloc_games = G; e->sp_alloc_out = cap;
    }
    e->sp.out_cap = cap;
    cudaStream_t s = strm(stream);
    CK(cudaMemsetAsync(e->sp.ctr, 0, SP_NCTR * sizeof(unsigned long long), s));
    e->sp_params = *p;
    e->sp_cfg.seed_lo = (uint32_t)(p->seed & 0xFFFFFFFFull);
    e->sp_cfg.seed_hi = (uint32_t)(p->seed >> 32);
    e->sp_cfg.id_base = p->game_base;
    // self-play: id = game_base + slot + generation * 2^20; arena: id = game ordinal (see gomoku_b200.h)
    e->sp_cfg.id_stride = arena ? (uint32_t)ordinal_stride : 1u << 20;
    e->sp_cfg.arena = arena;
    e->sp_cfg.games_total = games_total;
    e->sp_cfg.n_slots = G;
    e->sp_cfg.sims_a = p->num_simulations;
    e->sp_cfg.sims_b = sims_b > 0 ? sims_b : p->num_simulations;
    if (arena && e->sp_cfg.sims_b > e->cfg.max_simulations)
        return fail(GZ_ERR_ARG, "gz_arena_begin: sims_b exceeds max_simulations");
    // the search runs max(sims_a, sims_b) waves; each position stops at its own budget (gz_board.reserved[0])
    if (arena && e->sp_cfg.sims_b > e->sp_params.num_simulations) e->sp_params.num_simulations = e->sp_cfg.sims_b;
    e->sp_cfg.deterministic = p->deterministic;
    e->sp_cfg.exploration = exploration;
    e->sp_G = G;
    e->sp_steps = 0;
    launch_sp_begin(e->sp, G, e->sp_cfg, s);
    e->launches += 1;
    CKL();
    return GZ_OK;
}

extern "C" int gz_selfplay_begin(gz_handle h, int32_t G, const gz_search_params *p, double exploration,
                                 int64_t replay_capacity, int64_t games_total, void *stream) {
    if (games_total < 0) return fail(GZ_ERR_ARG, "gz_selfplay_begin: games_total must be >= 0");
    return sp_begin(h, G, p, exploration, replay_capacity, 0, 0, games_total, 0, stream);
}

extern "C" int gz_arena_begin(gz_handle h, int32_t G, const gz_search_params *p, double exploration, int32_t opponent,
                              int32_t sims_b, int64_t games_total, int64_t ordinal_stride, int64_t replay_capacity,
                              void *stream) {
    if (opponent != GZ_ARENA_ENGINE && opponent != GZ_ARENA_RANDOM)
        return fail(GZ_ERR_ARG, "gz_arena_begin: opponent must be GZ_ARENA_ENGINE or GZ_ARENA_RANDOM");
    if (games_total < 1 || ordinal_stride < G || ordinal_stride > 0x7FFFFFFFll)
        return fail(GZ_ERR_ARG, "gz_arena_begin: games_total >= 1 and G <= ordinal_stride < 2^31 required");
    return sp_begin(h, G, p, exploration, replay_capacity, opponent, opponent == GZ_ARENA_ENGINE ? sims_b : 0,
                    games_total, ordinal_stride, stream);
}

extern "C" int gz_arena_get_stats(gz_handle h, gz_arena_stats *out) {
    Engine *e = eng(h);
    if (!e || !out) return fail(GZ_ERR_ARG, "gz_arena_get_stats: null argument");
    if (e->sp_G < 1 || e->sp_cfg.arena == 0) return fail(GZ_ERR_STATE, "gz_arena_get_stats: gz_arena_begin was not called");
    CK(cudaSetDevice(e->cfg.device));
    CK(cudaDeviceSynchronize());
    unsigned long long c[SP_NCTR];
    CK(cudaMemcpy(c, e->sp.ctr, sizeof c, cudaMemcpyDeviceToHost));
    out->games_finished = (int64_t)c[SP_GAMES]; out->wins = (int64_t)c[SP_A_WINS];
    out->losses = (int64_t)c[SP_A_LOSSES]; out->draws = (int64_t)c[SP_DRAWS];
    out->plies_played = (int64_t)c[SP_PLIES]; out->steps = e->sp_steps;
    return GZ_OK;
}

extern "C" int gz_selfplay_step(gz_handle h, void *stream) {
    Engine *e = eng(h);
    if (!e) return fail(GZ_ERR_ARG, "gz_selfplay_step: null handle");
    if (e->sp_G < 1) return fail(GZ_ERR_STATE, "gz_selfplay_step: gz_selfplay_begin was not called");
    CK(cudaSetDevice(e->cfg.device));
    cudaStream_t s = strm(stream);
    gz_search_params p = e->sp_params;
    // a fresh Philox key per ply: rollouts of different plies of a game are independent
    p.seed = e->sp_params.seed + 0x9E3779B97F4A7C15ull * (uint64_t)(e->sp_steps + 1);
    int rc = search_device(e, e->sp.boards, e->sp.ai, e->sp_G, &p, nullptr, nullptr, e->sp.best, s);
    if (rc) return rc;
    launch_sp_apply(e->sp, e->sp_G, e->sp_cfg, s);
    e->launches += 1;
    e->sp_steps += 1;
    CKL();
    return GZ_OK;
}

extern "C" int gz_selfplay_set_simulations(gz_handle h, int32_t n) {
    Engine *e = eng(h);
    if (!e) return fail(GZ_ERR_ARG, "gz_selfplay_set_simulations: null handle");
    if (e->sp_G < 1 || e->sp_cfg.arena != 0) return fail(GZ_ERR_STATE, "gz_selfplay_set_simulations: no self-play session");
    if (n < 1 || n > e->cfg.max_simulations) return fail(GZ_ERR_ARG, "gz_selfplay_set_simulations: out of range");
    e->sp_params.num_simulations = n;
    e->sp_cfg.sims_a = e->sp_cfg.sims_b = n;
    return GZ_OK;
}

extern "C" int gz_selfplay_drain(gz_handle h, gz_replay_record *d_out, int64_t cap, int64_t *h_n, void *stream) {
    Engine *e = eng(h);
    if (!e || !h_n) return fail(GZ_ERR_ARG, "gz_selfplay_drain: null argument");
    if (e->sp_G < 1) return fail(GZ_ERR_STATE, "gz_selfplay_drain: gz_selfplay_begin was not called");
    CK(cudaSetDevice(e->cfg.device));
    cudaStream_t s = strm(stream);
    unsigned long long n = 0;
    CK(cudaMemcpyAsync(&n, e->sp.ctr + SP_CURSOR, sizeof n, cudaMemcpyDeviceToHost, s));
    CK(cudaStreamSynchronize(s));
    *h_n = (int64_t)n;
    if (n == 0) return GZ_OK;
    if (!d_out || cap < (int64_t)n) return fail(GZ_ERR_ARG, "gz_selfplay_drain: output buffer too small (see *h_n)");
    CK(cudaMemcpyAsync(d_out, e->sp.out, (size_t)n * sizeof(gz_replay_record), cudaMemcpyDeviceToDevice, s));
    CK(cudaMemsetAsync(e->sp.ctr + SP_CURSOR, 0, sizeof(unsigned long long), s));
    return GZ_OK;
}

extern "C" int gz_selfplay_boards(gz_handle h, gz_board *d_out, void *stream) {
    Engine *e = eng(h);
    if (!e || !d_out) return fail(GZ_ERR_ARG, "gz_selfplay_boards: null argument");
    if (e->sp_G < 1) return fail(GZ_ERR_STATE, "gz_selfplay_boards: gz_selfplay_begin was not called");
    CK(cudaSetDevice(e->cfg.device));
    CK(cudaMemcpyAsync(d_out, e->sp.boards, sizeof(gz_board) * e->sp_G, cudaMemcpyDeviceToDevice, strm(stream)));
    return GZ_OK;
}

extern "C" int gz_selfplay_get_stats(gz_handle h, gz_selfplay_stats *out) {
    Engine *e = eng(h);
    if (!e || !out) return fail(GZ_ERR_ARG, "gz_selfplay_get_stats: null argument");
    if (e->sp_G < 1) return fail(GZ_ERR_STATE, "gz_selfplay_get_stats: gz_selfplay_begin was not called");
    CK(cudaSetDevice(e->cfg.device));
    CK(cudaDeviceSynchronize());
    unsigned long long c[SP_NCTR];
    CK(cudaMemcpy(c, e->sp.ctr, sizeof c, cudaMemcpyDeviceToHost));
    out->games_finished = (int64_t)c[SP_GAMES]; out->plies_played = (int64_t)c[SP_PLIES];
    out->black_wins = (int64_t)c[SP_BLACK]; out->white_wins = (int64_t)c[SP_WHITE];
    out->draws = (int64_t)c[SP_DRAWS]; out->records_ready = (int64_t)c[SP_CURSOR];
    out->records_dropped = (int64_t)c[SP_DROPPED]; out->steps = e->sp_steps;
    return GZ_OK;
}

extern "C" int gz_get_tree(gz_handle h, int32_t g, int32_t cap, int32_t *h_parent, int32_t *h_move,
                           int32_t *h_visits, double *h_value, int32_t *n_nodes) {
    Engine *e = eng(h);
    if (!e || !n_nodes) return fail(GZ_ERR_ARG, "gz_get_tree: null argument");
    if (!e->searched) return fail(GZ_ERR_STATE, "gz_get_tree: no search has run on this handle");
    if (g < 0 || g >= e->cfg.max_games) return fail(GZ_ERR_ARG, "gz_get_tree: game index out of range");
    CK(cudaSetDevice(e->cfg.device));
    CK(cudaDeviceSynchronize());
    int nn = 0;
    CK(cudaMemcpy(&nn, e->tree.n_nodes + g, sizeof(int), cudaMemcpyDeviceToHost));
    *n_nodes = nn;
    int n = nn < cap ? nn : cap;
    size_t o = (size_t)g * e->tree.cap;
    std::vector<int16_t> par(n);
    std::vector<uint8_t> mv(n);
    CK(cudaMemcpy(par.data(), e->tree.parent + o, n * sizeof(int16_t), cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(mv.data(), e->tree.move + o, n, cudaMemcpyDeviceToHost));
    if (h_visits) CK(cudaMemcpy(h_visits, e->tree.visits + o, n * sizeof(int32_t), cudaMemcpyDeviceToHost));
    if (h_value) CK(cudaMemcpy(h_value, e->tree.value + o, n * sizeof(double), cudaMemcpyDeviceToHost));
    for (int i = 0; i < n; ++i) {
        if (h_parent) h_parent[i] = par[i];
        if (h_move) h_move[i] = i == 0 ? -1 : (mv[i] >> 4) * 15 + (mv[i] & 15);
    }
    return GZ_OK;
}

extern "C" int gz_get_priors(gz_handle h, int32_t g, int32_t node, float *h_prior, float *h_value) {
    Engine *e = eng(h);
    if (!e) return fail(GZ_ERR_ARG, "gz_get_priors: null handle");
    if (!e->searched) return fail(GZ_ERR_STATE, "gz_get_priors: no search has run on this handle");
    if (!e->tree.prior) return fail(GZ_ERR_STATE, "gz_get_priors: the handle was created without store_priors");
    if (g < 0 || g >= e->cfg.max_games) return fail(GZ_ERR_ARG, "gz_get_priors: game index out of range");
    CK(cudaSetDevice(e->cfg.device));
    CK(cudaDeviceSynchronize());
    int nn = 0;
    CK(cudaMemcpy(&nn, e->tree.n_nodes + g, sizeof(int), cudaMemcpyDeviceToHost));
    if (node < 0 || node >= nn) return fail(GZ_ERR_ARG, "gz_get_priors: node index out of range");
    const size_t slot = (size_t)g * e->tree.cap + node;
    uint8_t st = 0;
    CK(cudaMemcpy(&st, e->tree.status + slot, 1, cudaMemcpyDeviceToHost));
    if (st != 0) return fail(GZ_ERR_STATE, "gz_get_priors: terminal nodes have no prior (ai_agent.py:522)");
    if (h_prior) CK(cudaMemcpy(h_prior, e->tree.prior + slot * kPolicyPitch, kCells * sizeof(float), cudaMemcpyDeviceToHost));
    if (h_value) CK(cudaMemcpy(h_value, e->tree.vhead + slot, sizeof(float), cudaMemcpyDeviceToHost));
    return GZ_OK;
}

extern "C" int gz_get_counters(gz_handle h, gz_counters *out) {
    Engine *e = eng(h);
    if (!e || !out) return fail(GZ_ERR_ARG, "gz_get_counters: null argument");
    CK(cudaSetDevice(e->cfg.device));
    CK(cudaDeviceSynchronize());
    unsigned long long c[8];
    CK(cudaMemcpy(c, e->counters, sizeof c, cudaMemcpyDeviceToHost));
    out->nodes_created = (int64_t)c[1];
    out->pv_evals = (int64_t)c[2];
    out->planner_steps = (int64_t)c[3];
    out->rollout_steps = (int64_t)c[4];
    out->terminal_leaves = (int64_t)c[5];
    out->simulations = (int64_t)c[0];
    out->kernel_launches = e->launches;
    out->last_search_simulations = e->last_simulations;
    return GZ_OK;
}

extern "C" int gz_reset_counters(gz_handle h) {
    Engine *e = eng(h);
    if (!e) return fail(GZ_ERR_ARG, "gz_reset_counters: null handle");
    CK(cudaSetDevice(e->cfg.device));
    CK(cudaDeviceSynchronize());
    CK(cudaMemset(e->counters, 0, 8 * sizeof(unsigned long long)));
    e->launches = 0;
    return GZ_OK;
}

// ---- stage timing --------------------------------------------------------------------------------
namespace gz {
// Off by default: measured on bench config 3 (profiles/r02_pdl.md) the GraphNet stage of the event-profiled step drops
// by 9 ms, but the timed step does not move (597.2 / 597.7 ms with it, 595.5 / 595.6 without): kernels already queue
// back to back, and under the board's power cap an idle gap that is filled is paid back in clock.  GZ_PDL=1 enables it.
bool pdl_enabled() {
    static const bool on = getenv("GZ_PDL") && atoi(getenv("GZ_PDL")) != 0;
    return on;
}
static cudaEvent_t next_event(Engine &e) {
    if (e.ev_used == e.ev_pool.size()) {
        cudaEvent_t ev;
        cudaEventCreate(&ev);
        e.ev_pool.push_back(ev);
    }
    return e.ev_pool[e.ev_used++];
}
void prof_begin(Engine &e_, int stage, int n_launches, cudaStream_t s) {
    Engine &e = e_.root ? *e_.root : e_;
    e.launches += n_launches;
    if (!e.profiling) return;
    e.ev_stage.push_back(stage);
    e.stage_launches[stage] += n_launches;
    cudaEventRecord(next_event(e), s);
}
void prof_end(Engine &e_, cudaStream_t s) {
    Engine &e = e_.root ? *e_.root : e_;
    if (!e.profiling) return;
    cudaEventRecord(next_event(e), s);
}
}  // namespace gz

extern "C" int gz_set_profiling(gz_handle h, int32_t on) {
    Engine *e = eng(h);
    if (!e) return fail(GZ_ERR_ARG, "gz_set_profiling: null handle");
    e->profiling = on != 0;
    return GZ_OK;
}

extern "C" int gz_set_graph(gz_handle h, int32_t on) {
    Engine *e = eng(h);
    if (!e) return fail(GZ_ERR_ARG, "gz_set_graph: null handle");
    e->no_graph = on == 0;
    return GZ_OK;
}

extern "C" int gz_set_lanes(gz_handle h, int32_t lanes) {
    Engine *e = eng(h);
    if (!e) return fail(GZ_ERR_ARG, "gz_set_lanes: null handle");
    if (lanes < 0 || lanes > kMaxLanes) return fail(GZ_ERR_ARG, "gz_set_lanes: lanes out of range");
    e->cfg.lanes = lanes;
    return GZ_OK;
}

extern "C" int gz_get_stage_times(gz_handle h, double *h_ms, int64_t *h_launches) {
    Engine *e = eng(h);
    if (!e) return fail(GZ_ERR_ARG, "gz_get_stage_times: null handle");
    CK(cudaSetDevice(e->cfg.device));
    CK(cudaDeviceSynchronize());
    for (size_t i = 0; i + 1 < e->ev_used && i / 2 < e->ev_stage.size(); i += 2) {
        float ms = 0.f;
        CK(cudaEventElapsedTime(&ms, e->ev_pool[i], e->ev_pool[i + 1]));
        e->stage_ms[e->ev_stage[i / 2]] += ms;
    }
    e->ev_used = 0;
    e->ev_stage.clear();
    for (int i = 0; i < GZ_N_STAGES; ++i) {
        if (h_ms) h_ms[i] = e->stage_ms[i];
        if (h_launches) h_launches[i] = e->stage_launches[i];
        e->stage_ms[i] = 0.0;
        e->stage_launches[i] = 0;
    }
    return GZ_OK;
}
