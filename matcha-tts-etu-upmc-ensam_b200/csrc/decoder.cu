// decoder.cu -- host orchestration and C ABI of the CFM decoder path (fp32 mode).
//
// Replaces, behind the reference's own interfaces (paths under
// /root/reference/matcha/models/components/):
//   matcha_b200_decoder_forward   <- Decoder.forward(x, mask, mu, t)          decoder.py:271-371
//   matcha_b200_cfm_solve_euler   <- BaseConditionalFlowMatching.solve_ode_euler
//                                                                        flow_matching.py:67-104
// Activations live channels-last [B][T][C] in a caller-provided workspace; the twelve
// "b c t <-> b t c" rearranges of decoder.py:298-352 disappear, the channel concats of
// decoder.py:288,341 and the nearest-neighbour resample of :338-339 become addressing inside the
// implicit-GEMM loader.
#include <stdlib.h>

#include <mutex>
#include <vector>

#include "attention_mma.cuh"
#include "attention_tc.cuh"
#include "ffn_fused.cuh"
#include "gemm_abuild.cuh"
#include "gemm_bf16.cuh"

namespace mb200 {

// ---------------------------------------------------------------- packed weights
// Order and shapes must match matcha_b200/decoder.py::pack_weights.
struct StageW {
    const float *wt, *bt;                   // Resnet1D.mlp[1]            [256][1024], [256]
    const float *wc1, *bc1, *g1, *be1;      // block1 conv [256][3][Cin], bias, GN gamma/beta
    const float *wc2, *bc2, *g2, *be2;      // block2 conv [256][3][256]
    const float *wr, *br;                   // res_conv [256][Cin]
    const float *ln1g, *ln1b, *wqkv;        // norm1, [768][256] = to_q | to_k | to_v
    const float *wo, *bo;                   // to_out.0
    const float *ln3g, *ln3b, *w1, *b1, *w2, *b2;   // norm3, ff.net.0.proj [1024][256], ff.net.2 [256][1024]
    const float *wtail, *wtail2, *btail;    // down/up conv: [256][3][256]; ConvTranspose: even/odd [256][2][256]
    int cin;
};
struct DecW {
    const float *freq, *tw1, *tb1, *tw2, *tb2;
    StageW st[6];
    const float *wf, *bf, *gf, *bef, *wp, *bp;
};

static const int kStageCin[6] = {kInCh, kC, kC, kC, 2 * kC, 2 * kC};

static long long build_weights(const float *blob, DecW *w) {
    long long off = 0;
    auto take = [&](long long n) { const float *p = blob ? blob + off : nullptr; off += n; return p; };
    DecW d;
    d.freq = take(80);
    d.tw1 = take(1024LL * kInCh); d.tb1 = take(1024);
    d.tw2 = take(1024LL * 1024); d.tb2 = take(1024);
    for (int i = 0; i < 6; ++i) {
        StageW &s = d.st[i];
        s.cin = kStageCin[i];
        s.wt = take(256LL * 1024); s.bt = take(256);
        s.wc1 = take(256LL * 3 * s.cin); s.bc1 = take(256); s.g1 = take(256); s.be1 = take(256);
        s.wc2 = take(256LL * 3 * 256); s.bc2 = take(256); s.g2 = take(256); s.be2 = take(256);
        s.wr = take(256LL * s.cin); s.br = take(256);
        s.ln1g = take(256); s.ln1b = take(256); s.wqkv = take(768LL * 256);
        s.wo = take(256LL * 256); s.bo = take(256);
        s.ln3g = take(256); s.ln3b = take(256);
        s.w1 = take(1024LL * 256); s.b1 = take(1024); s.w2 = take(256LL * 1024); s.b2 = take(256);
        s.wtail = s.wtail2 = s.btail = nullptr;
        if (i == 0 || i == 1 || i == 5) { s.wtail = take(256LL * 3 * 256); s.btail = take(256); }
        if (i == 4) { s.wtail = take(256LL * 2 * 256); s.wtail2 = take(256LL * 2 * 256); s.btail = take(256); }
    }
    d.wf = take(256LL * 3 * 256); d.bf = take(256); d.gf = take(256); d.bef = take(256);
    d.wp = take((long long)kMel * 256); d.bp = take(kMel);
    if (w) *w = d;
    return off;
}

// ---------------------------------------------------------------- workspace
struct DecWs {
    float *tb, *th2, *x0, *h, *a, *g, *x1, *ln, *qkv, *att, *x2, *f1, *x3, *s0, *s1, *d, *stats;
    __nv_bfloat16 *x0b, *a1b, *lnb, *attb, *f1b, *inb, *s0b, *s1b, *upb;   // bf16 mode only
    double *gstats;                                                        // bf16 mode: [13 convs][B][8][2] GroupNorm sums
};

constexpr int kGnConvs = 13;      // GroupNorm'd convolutions per eval: 2 per Resnet1D + final_conv

static long long carve_workspace(float *base, int B, int T, int n_rows, int precision, DecWs *ws) {
    const long long M = (long long)B * T, Mu = (long long)B * (T + 1);
    long long off = 0;
    auto take = [&](long long n) { float *p = base ? base + off : nullptr; off += (n + 3) & ~3LL; return p; };
    DecWs w;
    w.tb = take(6LL * n_rows * kC);
    w.th2 = take((long long)n_rows * kTimeDim);
    w.x0 = take(M * kInCh);
    w.h = take(M * kC); w.a = take(M * kC); w.g = take(M * kC); w.x1 = take(M * kC); w.ln = take(M * kC);
    w.qkv = take(M * 3 * kC); w.att = take(M * kC); w.x2 = take(M * kC); w.f1 = take(M * 4 * kC); w.x3 = take(M * kC);
    w.s0 = take(M * kC); w.s1 = take(M * kC); w.d = take(Mu * kC);
    w.stats = take((long long)B * kGroups * 2);
    w.x0b = w.a1b = w.lnb = w.attb = w.f1b = w.inb = w.s0b = w.s1b = w.upb = nullptr;
    w.gstats = nullptr;
    if (precision != MB200_PREC_F32) {
        auto take_h = [&](long long n) { return reinterpret_cast<__nv_bfloat16 *>(take((n + 1) / 2)); };
        w.x0b = take_h(M * kInChPad); w.a1b = take_h(M * kC); w.lnb = take_h(M * kC); w.attb = take_h(M * kC);
        w.f1b = take_h(M * 4 * kC); w.inb = take_h(M * kC); w.s0b = take_h(M * kC); w.s1b = take_h(M * kC);
        w.upb = take_h(M * kC);
        w.gstats = reinterpret_cast<double *>(take(2LL * kGnConvs * B * kGroups * 2));
    }
    if (ws) *ws = w;
    return off * (long long)sizeof(float);
}

// ---------------------------------------------------------------- launch helpers
struct TmapEntry { const void *base; uint64_t d0, d1; uint32_t b1; CUtensorMap map; };
struct Ctx {
    cudaStream_t st;
    const float *mask;   // [B][T] float 0/1, pitch T at every U-Net level (prefix truncation)
    int B, T;
    DecWs ws;
    int launches;
    int err;
    int f16;             // 16-bit operand format of the tensor-core mode: 0 bf16, 1 fp16
    const float *fin_xin = nullptr; float fin_dt = 0.f; float *fin_out = nullptr;   // set around the final-projection GEMM
    // loss mode (matcha_b200_cfm_loss_forward): x is x1, pack_input forms phi_t, the final projection accumulates the MSE sums
    const float *loss_z = nullptr, *loss_t = nullptr; float loss_oms = 0.f; float *loss_phi = nullptr; double *loss_sums = nullptr;
    const float *loss_x1 = nullptr; int fin_loss = 0;
    int debug = 0;
    std::vector<TmapEntry> tmaps;   // tensor maps of this call's bf16 tensors
};

// Every kernel of the path is launched here, with programmatic dependent launch (common.cuh: the kernel's
// prologue overlaps the tail of its predecessor; all of them call pdl_wait() before touching memory).
// MB200_PDL=0 falls back to plain stream order.
static bool pdl_enabled() {
    static int v = -1;
    if (v < 0) { const char *e = getenv("MB200_PDL"); v = (e && e[0] == '0') ? 0 : 1; }
    return v != 0;
}
template <typename... KArgs, typename... Args>
static void launch(Ctx &c, void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, Args &&...args) {
    if (c.err) return;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = c.st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at; cfg.numAttrs = pdl_enabled() ? 1 : 0;
    const cudaError_t e = cudaLaunchKernelEx(&cfg, kern, static_cast<KArgs>(args)...);
    if (e != cudaSuccess) { c.err = cuda_fail(e); return; }
    c.launches++;
}

static void gemm(Ctx &c, const float *a0, int c0, int Ta0, const float *a1, int c1, bool masked, int ntaps,
                 const int *taps, int stride, int Tin, int Tout, const float *w, const float *bias, int N, int act,
                 const float *res, float *out, int out_T, int out_stride, int out_off) {
    GemmArgs g;
    g.a0 = a0; g.a1 = a1; g.c0 = c0; g.c1 = c1; g.Ta0 = Ta0;
    g.mask = masked ? c.mask : nullptr; g.mask_T = c.T;
    g.ntaps = ntaps;
    for (int i = 0; i < 4; ++i) g.tap_off[i] = i < ntaps ? taps[i] : 0;
    g.stride = stride; g.B = c.B; g.Tin = Tin; g.Tout = Tout;
    g.w = w; g.bias = bias; g.N = N; g.K = ntaps * (c0 + c1); g.act = act; g.res = res; g.out = out;
    g.out_T = out_T; g.out_stride = out_stride; g.out_off = out_off; g.out_bf16 = nullptr; g.omask = nullptr; g.o16_scale = 1.0f; g.f16 = 0;
    const int M = c.B * Tout;
    dim3 grid((M + kBM - 1) / kBM, N / kBN);
    launch(c, gemm_f32_kernel, grid, 256, 0, g);
}
static void linear(Ctx &c, const float *a, int K, int T, const float *w, const float *bias, int N, int act,
                   const float *res, float *out) {
    static const int tap0[1] = {0};
    gemm(c, a, K, T, nullptr, 0, false, 1, tap0, 1, T, T, w, bias, N, act, res, out, T, 1, 0);
}
static void gn_stats(Ctx &c, const float *h, int T) {
    launch(c, gn_stats_kernel, c.B * kGroups, 256, 0, h, T, c.ws.stats);
}
static void gn_mish(Ctx &c, const float *h, const float *gamma, const float *beta, const float *tb, int per_sample,
                    float *out, int T) {
    const long long n4 = (long long)c.B * T * (kC / 4);
    int blocks = (int)((n4 + 255) / 256);
    if (blocks > 148 * 16) blocks = 148 * 16;
    launch(c, gn_mish_kernel, blocks, 256, 0, h, c.ws.stats, gamma, beta, c.mask, c.T, tb, per_sample, out, nullptr, 0, 0, c.B, T, nullptr);
}
static void layernorm(Ctx &c, const float *x, const float *g, const float *b, float *out, int T) {
    const size_t rows = (size_t)c.B * T;
    launch(c, layernorm_kernel, (unsigned)((rows + 7) / 8), 256, 0, x, g, b, out, nullptr, 0, rows);
}

// Resnet1D + BasicTransformerBlock of one U-Net stage (decoder.py:80-86, transformer.py:316-368).
// `out` may alias a0: the input is last read by the res_conv GEMM.
static void run_stage(Ctx &c, const StageW &s, const float *a0, int c0, int Ta0, const float *a1, int c1, int T,
                      const float *tb, int tb_per_sample, float *out) {
    static const int taps3[3] = {-1, 0, 1};
    static const int tap0[1] = {0};
    DecWs &w = c.ws;
    gemm(c, a0, c0, Ta0, a1, c1, true, 3, taps3, 1, T, T, s.wc1, s.bc1, kC, 0, nullptr, w.h, T, 1, 0);
    gn_stats(c, w.h, T);
    gn_mish(c, w.h, s.g1, s.be1, tb, tb_per_sample, w.a, T);                 // Mish(GN(.))*mask + time bias
    gemm(c, w.a, kC, T, nullptr, 0, true, 3, taps3, 1, T, T, s.wc2, s.bc2, kC, 0, nullptr, w.h, T, 1, 0);
    gn_stats(c, w.h, T);
    gn_mish(c, w.h, s.g2, s.be2, nullptr, 0, w.g, T);
    gemm(c, a0, c0, Ta0, a1, c1, true, 1, tap0, 1, T, T, s.wr, s.br, kC, 0, w.g, w.x1, T, 1, 0);   // h + res_conv(x*mask)
    layernorm(c, w.x1, s.ln1g, s.ln1b, w.ln, T);
    linear(c, w.ln, kC, T, s.wqkv, nullptr, 3 * kC, 0, nullptr, w.qkv);
    dim3 agrid((T + kAttBQ - 1) / kAttBQ, kHeads, c.B);
    launch(c, attention_f32_kernel, agrid, 128, kAttSmemBytes, w.qkv, c.mask, c.T, w.att, nullptr, 0, T);
    linear(c, w.att, kC, T, s.wo, s.bo, kC, 0, w.x1, w.x2);
    layernorm(c, w.x2, s.ln3g, s.ln3b, w.ln, T);
    linear(c, w.ln, kC, T, s.w1, s.b1, 4 * kC, 1, nullptr, w.f1);
    linear(c, w.f1, 4 * kC, T, s.w2, s.b2, kC, 0, w.x2, out);
}

// One estimator call: velocity (x_in == nullptr) or Euler update out = x_in + dt * v.
static void decoder_eval(Ctx &c, const DecW &W, const float *x, const float *mu, const float *tb_rows, int n_rows,
                         int tb_per_sample, const float *x_in, float dt, float *out) {
    static const int taps3[3] = {-1, 0, 1};
    static const int taps_even[2] = {0, -1};    // ConvTranspose1d(4,2,1): out[2s]   = W[..,1] in[s]   + W[..,3] in[s-1]
    static const int taps_odd[2] = {1, 0};      //                         out[2s+1] = W[..,0] in[s+1] + W[..,2] in[s]
    DecWs &w = c.ws;
    const int T = c.T, T2 = (T + 1) / 2, Tu = 2 * T2;
    auto tb = [&](int st) { return tb_rows + (size_t)st * n_rows * kC; };
    dim3 pgrid((T + 31) / 32, kInCh / 32, c.B);
    launch(c, pack_input_kernel, pgrid, 256, 0, x, mu, w.x0, c.B, T);
    run_stage(c, W.st[0], w.x0, kInCh, T, nullptr, 0, T, tb(0), tb_per_sample, w.s0);
    gemm(c, w.s0, kC, T, nullptr, 0, true, 3, taps3, 2, T, T2, W.st[0].wtail, W.st[0].btail, kC, 0, nullptr, w.d, T2, 1, 0);
    run_stage(c, W.st[1], w.d, kC, T2, nullptr, 0, T2, tb(1), tb_per_sample, w.s1);
    gemm(c, w.s1, kC, T2, nullptr, 0, true, 3, taps3, 1, T2, T2, W.st[1].wtail, W.st[1].btail, kC, 0, nullptr, w.d, T2, 1, 0);
    run_stage(c, W.st[2], w.d, kC, T2, nullptr, 0, T2, tb(2), tb_per_sample, w.d);
    run_stage(c, W.st[3], w.d, kC, T2, nullptr, 0, T2, tb(3), tb_per_sample, w.d);
    run_stage(c, W.st[4], w.d, kC, T2, w.s1, kC, T2, tb(4), tb_per_sample, w.x3);
    gemm(c, w.x3, kC, T2, nullptr, 0, true, 2, taps_even, 1, T2, T2, W.st[4].wtail, W.st[4].btail, kC, 0, nullptr, w.d, Tu, 2, 0);
    gemm(c, w.x3, kC, T2, nullptr, 0, true, 2, taps_odd, 1, T2, T2, W.st[4].wtail2, W.st[4].btail, kC, 0, nullptr, w.d, Tu, 2, 1);
    run_stage(c, W.st[5], w.d, kC, Tu, w.s0, kC, T, tb(5), tb_per_sample, w.x3);
    gemm(c, w.x3, kC, T, nullptr, 0, true, 3, taps3, 1, T, T, W.st[5].wtail, W.st[5].btail, kC, 0, nullptr, w.d, T, 1, 0);
    gemm(c, w.d, kC, T, nullptr, 0, true, 3, taps3, 1, T, T, W.wf, W.bf, kC, 0, nullptr, w.h, T, 1, 0);
    gn_stats(c, w.h, T);
    dim3 fgrid((T + 31) / 32, c.B);
    launch(c, final_proj_kernel, fgrid, 256, 0, w.h, w.stats, W.gf, W.bef, c.mask, W.wp, W.bp, x_in, dt, out, T);
}


// ================================================================== bf16 mode (tcgen05 GEMMs)
typedef __nv_bfloat16 bf16;
// wc1: [512][3][cin_pad] = block1's conv (rows 0..255) stacked on res_conv (rows 256..511, centre tap only, zeros
// elsewhere): both read the stage input, one GEMM with two output blocks computes them (gemm_bf16.cuh, n_split)
struct StageWb { const bf16 *wc1, *wc2, *wqkv, *wo, *w1, *w2, *wtail, *wtail2; int cin_pad; };
struct DecWb { StageWb st[6]; const bf16 *wf, *wp; };
static const int kStageCinPad[6] = {kInChPad, kC, kC, kC, 2 * kC, 2 * kC};

// Order must match pack_bf16 below.
static long long build_weights_bf16(const bf16 *blob, DecWb *w) {
    long long off = 0;
    auto take = [&](long long n) { const bf16 *p = blob ? blob + off : nullptr; off += (n + 7) & ~7LL; return p; };
    DecWb d;
    for (int i = 0; i < 6; ++i) {
        StageWb &s = d.st[i];
        s.cin_pad = kStageCinPad[i];
        s.wc1 = take(512LL * 3 * s.cin_pad); s.wc2 = take(256LL * 3 * 256);
        s.wqkv = take(768LL * 256); s.wo = take(256LL * 256); s.w1 = take(1024LL * 256); s.w2 = take(256LL * 1024);
        s.wtail = s.wtail2 = nullptr;
        if (i == 0) s.wtail = take(256LL * 2 * 512);              // stride-2 conv on the paired view
        if (i == 1 || i == 5) s.wtail = take(256LL * 3 * 256);
        if (i == 4) { s.wtail = take(256LL * 2 * 256); s.wtail2 = take(256LL * 2 * 256); }   // ConvTranspose phases
    }
    d.wf = take(256LL * 3 * 256);
    d.wp = take(256LL * 256);        // final_proj [80][256] zero-padded to the 256-wide tile
    if (w) *w = d;
    return off;
}

static bool tmap_for(Ctx &c, const bf16 *base, int ndim, uint64_t d0, uint64_t d1, uint64_t d2, uint32_t b0,
                     uint32_t b1, CUtensorMap *out) {
    for (auto &e : c.tmaps)
        if (e.base == base && e.d0 == d0 && e.d1 == d1 && e.b1 == b1) { *out = e.map; return true; }
    TmapEntry e;
    e.base = base; e.d0 = d0; e.d1 = d1; e.b1 = b1;
    const uint64_t dims[3] = {d0, d1, d2};
    const uint64_t strides[2] = {d0 * 2, d0 * d1 * 2};
    const uint32_t box[3] = {b0, b1, 1};
    const int rc = encode_tmap_bf16(&e.map, base, ndim, dims, strides, box);
    if (rc != MB200_OK) { c.err = rc; return false; }
    c.tmaps.push_back(e);
    *out = e.map;
    return true;
}

// Output map for the TMA-store epilogue: rows [B][T][pitch] of fp32 or 16-bit elements, row t of the tile grid at
// row t * row_stride of the tensor (ConvTranspose phases write every other row).
static bool tmap_out_for(Ctx &c, const void *base, bool f32, int ncols, int pitch, int T, int out_T, int row_stride, CUtensorMap *out,
                         int box_cols = 32) {
    const uint64_t key1 = (uint64_t)pitch * 8 + (f32 ? 1 : 0) + 2 + (uint64_t)row_stride * 4 + (uint64_t)box_cols * 65536;
    for (auto &e : c.tmaps)
        if (e.base == base && e.d0 == (uint64_t)ncols && e.d1 == (uint64_t)T && e.b1 == 0xffffffffu - (uint32_t)key1) { *out = e.map; return true; }
    TmapEntry e;
    e.base = base; e.d0 = (uint64_t)ncols; e.d1 = (uint64_t)T; e.b1 = 0xffffffffu - (uint32_t)key1;
    const uint64_t es = f32 ? 4 : 2;
    const uint64_t dims[3] = {(uint64_t)ncols, (uint64_t)T, (uint64_t)c.B};
    const uint64_t strides[2] = {(uint64_t)pitch * row_stride * es, (uint64_t)pitch * out_T * es};
    const int rc = encode_tmap_out(&e.map, base, f32 ? 1 : 0, dims, strides, box_cols);
    if (rc != MB200_OK) { c.err = rc; return false; }
    c.tmaps.push_back(e);
    *out = e.map;
    return true;
}

// The epilogue variants that exist (gemm_bf16.cuh: everything about the outputs is a template parameter).
typedef void (*GemmKernel)(CUtensorMap, CUtensorMap, CUtensorMap, CUtensorMap, CUtensorMap, GemmBf16Args);
template <int EPI, int FLAGS>
static GemmKernel gemm_variant(bool f16) { return f16 ? (GemmKernel)gemm_bf16_kernel<EPI, FLAGS, true> : (GemmKernel)gemm_bf16_kernel<EPI, FLAGS, false>; }
static GemmKernel pick_gemm(int epi, int flags, bool f16) {
    switch (epi * 16 + flags) {
        case 0 * 16 + kGF32: return gemm_variant<0, kGF32>(f16);                          // ConvTranspose phases
        case 0 * 16 + (kGF32 | kGStats): return gemm_variant<0, kGF32 | kGStats>(f16);    // conv -> GroupNorm input + its sums
        case 0 * 16 + (kGF32 | kGRes): return gemm_variant<0, kGF32 | kGRes>(f16);        // res_conv, attention out-proj
        case 0 * 16 + kGB16: return gemm_variant<0, kGB16>(f16);                          // q|k|v, down/tail convs
        case 0 * 16 + (kGB16 | kGRes): return gemm_variant<0, kGB16 | kGRes>(f16);        // FFN2 feeding a conv only
        case 0 * 16 + (kGF32 | kGB16): return gemm_variant<0, kGF32 | kGB16>(f16);
        case 0 * 16 + (kGF32 | kGB16 | kGRes): return gemm_variant<0, kGF32 | kGB16 | kGRes>(f16);
        case 1 * 16 + kGB16: return gemm_variant<1, kGB16>(f16);                          // FFN1 + erf-GELU
        case 1 * 16 + (kGF32 | kGB16 | kGRes): return gemm_variant<1, kGF32 | kGB16 | kGRes>(f16);   // tests only
        case 2 * 16 + 0: return gemm_variant<2, 0>(f16);                                  // final_proj + Euler update
        case 3 * 16 + 0: return gemm_variant<3, 0>(f16);                                  // final_proj + loss sums (compute_loss)
    }
    return nullptr;
}
static int gemm_set_attributes() {
    for (int epi = 0; epi < 4; ++epi)
        for (int flags = 0; flags < 16; ++flags)
            for (int f16 = 0; f16 < 2; ++f16) {
                const GemmKernel k = pick_gemm(epi, flags, f16 != 0);
                if (k) MB200_CUDA_TRY(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, kGSmemBytes));
            }
    MB200_CUDA_TRY(cudaFuncSetAttribute(gemm_abuild_kernel<1, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kAbSmemBytes));
    MB200_CUDA_TRY(cudaFuncSetAttribute(gemm_abuild_kernel<1, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, kAbSmemBytes));
    MB200_CUDA_TRY(cudaFuncSetAttribute(gemm_abuild_kernel<2, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kAbSmemBytes));
    MB200_CUDA_TRY(cudaFuncSetAttribute(gemm_abuild_kernel<2, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, kAbSmemBytes));
    MB200_CUDA_TRY(cudaFuncSetAttribute(ffn_fused_kernel<true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, kFfSmemBytes));
    MB200_CUDA_TRY(cudaFuncSetAttribute(ffn_fused_kernel<false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, kFfSmemBytes));
    MB200_CUDA_TRY(cudaFuncSetAttribute(ffn_fused_kernel<true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kFfSmemBytes));
    MB200_CUDA_TRY(cudaFuncSetAttribute(ffn_fused_kernel<false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kFfSmemBytes));
    return MB200_OK;
}

// out = act(A W^T + bias) + res on the tensor cores.  a0/a1: bf16 [B][T][c0|c1] (already x*mask where the
// reference multiplies), w: bf16 [N][ntaps*(c0+c1)].
//
// fp16 range guard.  The residual-stream tensors (stage outputs, the inputs of the convs) are the only 16-bit operands whose size
// the network does not bound itself (everything else sits behind a GroupNorm / LayerNorm / softmax); IEEE half tops out at
// 65504.  In fp16 mode they are stored multiplied by kStreamDown = 2^-4 (exact, but for values below 1e-3 that turn subnormal)
// and the GEMM that consumes them multiplies its fp32 accumulator by 2^4 before the bias (one FMA instead of one add), so the mode
// holds |x| < 1.04e6 on the stream; beyond that the conversion saturates (pack16x2) instead of producing inf.  bf16 mode: scale 1.
// sflags: kSA = the A rows are a stream tensor, kSO = the 16-bit output is one.
constexpr int kSA = 1, kSO = 2;
constexpr float kStreamDown = 1.0f / 16.0f, kStreamUp = 16.0f;
static inline float stream_down(const Ctx &c) { return c.f16 ? kStreamDown : 1.0f; }
static inline float stream_up(const Ctx &c) { return c.f16 ? kStreamUp : 1.0f; }
// 32-row slabs per sample (gemm_bf16.cuh::slab_coords).  MB200_FLAT=0 rounds every sample up to whole 128-row tiles (the round-1
// tiling: tiles never straddle samples).  pad_row: the A-building kernels keep the conv halo in one shared-memory tile and need
// at least one out-of-range (zero) row between two samples.
static bool flat_tiling() {
    static int v = -1;
    if (v < 0) { const char *e = getenv("MB200_FLAT"); v = (e && e[0] == '0') ? 0 : 1; }
    return v != 0;
}
static int slabs_per_sample(int T, bool pad_row) {
    const int rows = T + (pad_row ? 1 : 0);
    return flat_tiling() ? (rows + 31) / 32 : 4 * ((rows + 127) / 128);
}
static void gemm_b(Ctx &c, int sflags, const bf16 *a0, int c0, const bf16 *a1, int c1, int ntaps, const int *taps, int T,
                   const bf16 *w, int N, const float *bias, int act, const float *res, float *out_f32,
                   bf16 *out_b16, bool omask, int out_T = 0, int out_stride = 1, int out_off = 0, double *gstats = nullptr,
                   const float *bias2 = nullptr, float *out2_f32 = nullptr, int kb2_begin = 0, int kb2_count = 0) {
    if (c.err) return;
    CUtensorMap ma0, ma1, mb;
    if (!tmap_for(c, a0, 3, (uint64_t)c0, (uint64_t)T, (uint64_t)c.B, kGK, 32, &ma0)) return;
    ma1 = ma0;
    if (a1 && !tmap_for(c, a1, 3, (uint64_t)c1, (uint64_t)T, (uint64_t)c.B, kGK, 32, &ma1)) return;
    if (!tmap_for(c, w, 2, (uint64_t)ntaps * (c0 + c1), (uint64_t)N, 1, kGK, kGN, &mb)) return;
    GemmBf16Args g;
    g.B = c.B; g.Tout = T; g.spp = slabs_per_sample(T, false); g.n_mtiles = (c.B * g.spp + 3) / 4;
    g.ntaps = ntaps;
    for (int i = 0; i < 4; ++i) g.tap_off[i] = i < ntaps ? taps[i] : 0;
    g.c0_chunks = c0 / kGK; g.c1_chunks = c1 / kGK;
    g.bias = bias; g.act = act; g.res = res; g.out_f32 = out_f32; g.out_bf16 = out_b16;
    g.omask = omask ? c.mask : nullptr; g.omask_T = c.T; g.N = N; g.f16 = c.f16;
    g.acc_scale = (sflags & kSA) ? stream_up(c) : 1.0f; g.o16_scale = (sflags & kSO) ? stream_down(c) : 1.0f;
    g.out_T = out_T > 0 ? out_T : T; g.out_stride = out_T > 0 ? out_stride : 1; g.out_off = out_T > 0 ? out_off : 0;
    g.fin_n = 0; g.fin_xin = nullptr; g.fin_dt = 0.f; g.fin_out = nullptr;
    g.loss_x1 = nullptr; g.loss_z = nullptr; g.loss_oms = 0.f; g.loss_sum = nullptr;
    if (c.fin_out || c.fin_loss) { g.fin_n = kMel; g.fin_xin = c.fin_xin; g.fin_dt = c.fin_dt; g.fin_out = c.fin_out; g.omask = c.mask; }
    if (c.fin_loss) { g.loss_x1 = c.loss_x1; g.loss_z = c.loss_z; g.loss_oms = c.loss_oms; g.loss_sum = c.loss_sums; }
    const long long n_tiles = (long long)g.n_mtiles * (N / kGN);
    const int grid = (int)(n_tiles < kNumSMsB200 ? n_tiles : kNumSMsB200);      // persistent: one CTA per SM
    g.debug = c.debug;
    if ((long long)c.B * g.out_T * N >= 0x7fffffffLL) { c.err = MB200_ERR_UNSUPPORTED; return; }
    g.gstats = gstats;
    g.n_split = out2_f32 ? N / 2 : 0; g.kb2_begin = kb2_begin; g.kb2_count = kb2_count; g.bias2 = bias2; g.out2_f32 = out2_f32;
    const int flags = (res ? kGRes : 0) | (out_f32 ? kGF32 : 0) | (out_b16 ? kGB16 : 0) | (gstats ? kGStats : 0);
    const GemmKernel k = pick_gemm(g.fin_n > 0 ? (c.fin_loss ? 3 : 2) : act, g.fin_n > 0 ? 0 : flags, c.f16 != 0);
    if (!k) { c.err = MB200_ERR_UNSUPPORTED; return; }
    // TMA-store epilogue (one output, no residual): output tensor maps
    CUtensorMap mo = ma0, mo2 = ma0;
    if (g.fin_n == 0 && !res && ((out_f32 != nullptr) != (out_b16 != nullptr))) {
        if (N > kGBiasBytes / 4) { c.err = MB200_ERR_UNSUPPORTED; return; }
        const int pitch = g.n_split ? g.n_split : N;
        const void *ob = out_f32 ? (const void *)(out_f32 + (size_t)g.out_off * pitch) : (const void *)(out_b16 + (size_t)g.out_off * pitch);
        if (!tmap_out_for(c, ob, out_f32 != nullptr, pitch, pitch, T, g.out_T, g.out_stride, &mo)) return;
        if (g.n_split && !tmap_out_for(c, out2_f32, true, pitch, pitch, T, g.out_T, g.out_stride, &mo2)) return;
    }
    launch(c, k, grid, kGThreads, kGSmemBytes, ma0, ma1, mb, mo, mo2, g);
}

// softmax(q k^T / 8 + mask) v on 16-bit q|k|v rows: tcgen05 kernel (attention_tc.cuh); MB200_ATTN=mma selects the
// warp-level mma.sync kernel (attention_mma.cuh).
static bool attn_use_tc() {
    static int v = -1;
    if (v < 0) { const char *e = getenv("MB200_ATTN"); v = (e && e[0] == 'm') ? 0 : 1; }
    return v != 0;
}
static void attention_b(Ctx &c, const bf16 *qkvb, bf16 *out, int T) {
    if (c.err) return;
    if (attn_use_tc()) {
        CUtensorMap mq, mkv;
        if (!tmap_for(c, qkvb, 3, (uint64_t)(3 * kC), (uint64_t)T, (uint64_t)c.B, kDHead, kTaBQ, &mq)) return;
        if (!tmap_for(c, qkvb, 3, (uint64_t)(3 * kC), (uint64_t)T, (uint64_t)c.B, kDHead, kTaBK, &mkv)) return;
        AttnTcArgs a;
        a.mask = c.mask; a.mask_T = c.T; a.out = out; a.B = c.B; a.T = T;
        const int grid = ((T + kTaBQ - 1) / kTaBQ) * kHeads * c.B;
        if (c.f16) launch(c, attention_tc_kernel<true>, grid, kTaThreads, kTaSmemBytes, mq, mkv, a);
        else launch(c, attention_tc_kernel<false>, grid, kTaThreads, kTaSmemBytes, mq, mkv, a);
        return;
    }
    dim3 agrid((T + kFaBQ - 1) / kFaBQ, kHeads, c.B);
    if (c.f16) launch(c, attention_mma_kernel<true>, agrid, 128, kFaSmemBytes, qkvb, c.mask, c.T, out, T);
    else launch(c, attention_mma_kernel<false>, agrid, 128, kFaSmemBytes, qkvb, c.mask, c.T, out, T);
}

// GEMMs that build their A operand from fp32 rows (gemm_abuild.cuh).  MB200_ABUILD=0 keeps the separate GroupNorm / Mish /
// LayerNorm launches; MB200_ABUILD_DESC=1 sets the descriptor base-offset field of the shifted conv views (experiment: wrong results on
// sm_100a -- the swizzle is applied to absolute shared-memory address bits, a row-shifted start address needs no hint).
static int abuild_env(const char *name, int dflt) {
    const char *e = getenv(name);
    return e ? atoi(e) : dflt;
}
static bool abuild_enabled() { static int v = -1; if (v < 0) v = abuild_env("MB200_ABUILD", 1); return v != 0; }
// mode 1: x1 = Mish(GN(h)) * mask + r, out16 = LN(x1) W^T (N = 768).  mode 2: out_f32 = conv3(Mish(GN(h)) * mask + tb) + bias + sums
static void abuild_b(Ctx &c, int mode, const float *h, const double *gsum, const float *gamma, const float *beta, int T,
                     const float *tb, int tb_per_sample, const float *r, const float *ln_g, const float *ln_b, float *x1,
                     const bf16 *w, int N, const float *bias, bf16 *out16, float *out_f32, double *gstats) {
    if (c.err) return;
    static int desc_mode = -1;
    if (desc_mode < 0) desc_mode = abuild_env("MB200_ABUILD_DESC", 0);
    CUtensorMap mb, mo;
    const int K = mode == 2 ? 3 * kC : kC;
    if (!tmap_for(c, w, 2, (uint64_t)K, (uint64_t)N, 1, kGK, kGN, &mb)) return;
    if (mode == 1) { if (!tmap_out_for(c, out16, false, N, N, T, T, 1, &mo, 32)) return; }
    else if (!tmap_out_for(c, out_f32, true, N, N, T, T, 1, &mo, 16)) return;
    AbuildArgs g;
    g.B = c.B; g.T = T; g.tiles_per_sample = (T + kGM - 1) / kGM; g.N = N;    // per-sample tiles: the slab version of this kernel measured slower
    g.h = h; g.gsum = gsum; g.gamma = gamma; g.beta = beta; g.mask = c.mask; g.mask_T = c.T;
    g.tb = tb; g.tb_per_sample = tb_per_sample; g.r = r; g.ln_g = ln_g; g.ln_b = ln_b; g.x1 = x1;
    g.bias = bias; g.gstats = gstats; g.desc_mode = desc_mode;
    if (N > 1024 || (long long)c.B * T * N >= 0x7fffffffLL) { c.err = MB200_ERR_UNSUPPORTED; return; }
    const int n_mtiles = c.B * g.tiles_per_sample;
    const int grid = n_mtiles < kNumSMsB200 ? n_mtiles : kNumSMsB200;
    if (mode == 1) {
        if (c.f16) launch(c, gemm_abuild_kernel<1, true>, grid, kAbThreads, kAbSmemBytes, mb, mo, g);
        else launch(c, gemm_abuild_kernel<1, false>, grid, kAbThreads, kAbSmemBytes, mb, mo, g);
    } else {
        if (c.f16) launch(c, gemm_abuild_kernel<2, true>, grid, kAbThreads, kAbSmemBytes, mb, mo, g);
        else launch(c, gemm_abuild_kernel<2, false>, grid, kAbThreads, kAbSmemBytes, mb, mo, g);
    }
}

// out = res + W2 gelu(W1 a + b1) + b2 in one kernel (ffn_fused.cuh); MB200_FFN=split keeps the two-GEMM path.
static bool ffn_use_fused() {
    static int v = -1;
    if (v < 0) { const char *e = getenv("MB200_FFN"); v = (e && e[0] == 's') ? 0 : 1; }
    return v != 0;
}
// With wo / bo (the OUTPROJ variant): a = the 16-bit attention rows, res = the block's input stream x1; the kernel computes
// x2 = a Wo^T + bo + res, norm3 and the feed-forward, out = x2 + FFN(norm3(x2)).  MB200_OUTPROJ=0 keeps the separate GEMM.
static bool outproj_fused() {
    static int v = -1;
    if (v < 0) { const char *e = getenv("MB200_OUTPROJ"); v = (e && e[0] == '0') ? 0 : 1; }
    return v != 0;
}
static void ffn_b(Ctx &c, const bf16 *a, int T, const bf16 *w1, const float *b1, const bf16 *w2, const float *b2,
                  const float *res, float *out_f32, bf16 *out_b16, bool omask, const float *ln_in = nullptr,
                  const float *ln_g = nullptr, const float *ln_b = nullptr, const bf16 *wo = nullptr, const float *bo = nullptr,
                  bool o_stream = true) {
    if (c.err) return;
    CUtensorMap ma, m1, m2, mo;
    if (!tmap_for(c, a, 3, (uint64_t)kC, (uint64_t)T, (uint64_t)c.B, 64, 32, &ma)) return;
    if (!tmap_for(c, w1, 2, (uint64_t)kC, (uint64_t)kFfHid, 1, 64, kFfChunk, &m1)) return;
    if (!tmap_for(c, w2, 2, (uint64_t)kFfHid, (uint64_t)kC, 1, 64, kC, &m2)) return;
    mo = m2;
    if (wo && !tmap_for(c, wo, 2, (uint64_t)kC, (uint64_t)kC, 1, 64, kC, &mo)) return;
    FfnArgs g;
    g.B = c.B; g.T = T; g.spp = slabs_per_sample(T, false);
    g.b1 = b1; g.b2 = b2; g.res = res; g.out_f32 = out_f32; g.out_b16 = out_b16;
    g.omask = omask ? c.mask : nullptr; g.omask_T = c.T; g.debug = c.debug; g.o16_scale = o_stream ? stream_down(c) : 1.0f;
    g.ln_in = ln_in; g.ln_g = ln_g; g.ln_b = ln_b; g.bo = bo;
    if ((long long)c.B * T * kC >= 0x7fffffffLL) { c.err = MB200_ERR_UNSUPPORTED; return; }
    const int grid = (c.B * g.spp + 3) / 4;
    if (wo) {
        if (!bo || !ln_g || !ln_b || !res) { c.err = MB200_ERR_INVALID_ARG; return; }
        if (c.f16) launch(c, ffn_fused_kernel<true, true>, grid, kFfThreads, kFfSmemBytes, ma, m1, m2, mo, g);
        else launch(c, ffn_fused_kernel<false, true>, grid, kFfThreads, kFfSmemBytes, ma, m1, m2, mo, g);
        return;
    }
    if (c.f16) launch(c, ffn_fused_kernel<true, false>, grid, kFfThreads, kFfSmemBytes, ma, m1, m2, mo, g);
    else launch(c, ffn_fused_kernel<false, false>, grid, kFfThreads, kFfSmemBytes, ma, m1, m2, mo, g);
}

static void run_stage_bf16(Ctx &c, int stage, const StageW &s, const StageWb &sb, const bf16 *in0, int c0, const bf16 *in1,
                           int c1, int T, const float *tb, int tb_per_sample, float *out_f32, bf16 *out_b16) {
    static const int taps3[3] = {-1, 0, 1};
    static const int tap0[1] = {0};
    DecWs &w = c.ws;
    const long long n4 = (long long)c.B * T * (kC / 4);
    int eb = (int)((n4 + 255) / 256); if (eb > 148 * 16) eb = 148 * 16;
    const size_t rows = (size_t)c.B * T;
    // GroupNorm sums come out of the conv GEMMs' epilogues (slot 2*stage, 2*stage+1 of ws.gstats)
    double *gs1 = w.gstats + (size_t)(2 * stage) * c.B * kGroups * 2, *gs2 = gs1 + (size_t)c.B * kGroups * 2;
    // block1's conv and res_conv in one GEMM: h = conv(x*mask) (+ GroupNorm sums), g = res_conv(x*mask) (centre tap only)
    const int chunks = (c0 + c1) / kGK;
    gemm_b(c, kSA, in0, c0, in1, c1, 3, taps3, T, sb.wc1, 2 * kC, s.bc1, 0, nullptr, w.h, nullptr, false, 0, 1, 0, gs1, s.br, w.g, chunks, chunks);
    const unsigned rb = (unsigned)((rows + kGmRowsPerBlock - 1) / kGmRowsPerBlock);
    bf16 *qkvb = reinterpret_cast<bf16 *>(w.qkv);      // 16-bit q|k|v rows in the fp32 qkv scratch
    if (abuild_enabled()) {
        // block2's conv builds its operand Mish(GN(h)) * mask + time bias itself (its output h2 lives in the fp32 FFN scratch, unused on this path),
        // the q|k|v projection builds x1 = Mish(GN(h2)) * mask + res_conv and norm1(x1)
        float *h2 = w.f1;
        abuild_b(c, 2, w.h, gs1, s.g1, s.be1, T, tb, tb_per_sample, nullptr, nullptr, nullptr, nullptr, sb.wc2, kC, s.bc2, nullptr, h2, gs2);
        abuild_b(c, 1, h2, gs2, s.g2, s.be2, T, nullptr, 0, w.g, s.ln1g, s.ln1b, w.x1, sb.wqkv, 3 * kC, nullptr, qkvb, nullptr, nullptr);
    } else {
        launch(c, gn_mish_rows_kernel<0>, rb, 256, 0, w.h, gs1, s.g1, s.be1, c.mask, c.T, tb, tb_per_sample, nullptr, nullptr, nullptr,
               nullptr, w.a1b, c.f16, c.B, T);
        gemm_b(c, 0, w.a1b, kC, nullptr, 0, 3, taps3, T, sb.wc2, kC, s.bc2, 0, nullptr, w.h, nullptr, false, 0, 1, 0, gs2);
        // x1 = Mish(GN(h)) * mask + res_conv, and norm1(x1) as the q|k|v operand, in one pass
        launch(c, gn_mish_rows_kernel<1>, rb, 256, 0, w.h, gs2, s.g2, s.be2, c.mask, c.T, nullptr, 0, w.g, s.ln1g, s.ln1b, w.x1, w.lnb,
               c.f16, c.B, T);
        gemm_b(c, 0, w.lnb, kC, nullptr, 0, 1, tap0, T, sb.wqkv, 3 * kC, nullptr, 0, nullptr, nullptr, qkvb, false);
    }
    attention_b(c, qkvb, w.attb, T);
    if (ffn_use_fused() && outproj_fused()) {      // to_out[0] + residual + norm3 + feed-forward + residual in one kernel
        ffn_b(c, w.attb, T, sb.w1, s.b1, sb.w2, s.b2, w.x1, out_f32, out_b16, true, nullptr, s.ln3g, s.ln3b, sb.wo, s.bo);
        return;
    }
    gemm_b(c, 0, w.attb, kC, nullptr, 0, 1, tap0, T, sb.wo, kC, s.bo, 0, w.x1, w.x2, nullptr, false);
    if (ffn_use_fused()) {      // norm3 is applied while the kernel builds its A operand
        ffn_b(c, w.lnb, T, sb.w1, s.b1, sb.w2, s.b2, w.x2, out_f32, out_b16, true, w.x2, s.ln3g, s.ln3b);
    } else {
        launch(c, layernorm_kernel, (unsigned)((rows + 7) / 8), 256, 0, w.x2, s.ln3g, s.ln3b, nullptr, w.lnb, c.f16, rows);
        gemm_b(c, 0, w.lnb, kC, nullptr, 0, 1, tap0, T, sb.w1, 4 * kC, s.b1, 1, nullptr, nullptr, w.f1b, false);
        gemm_b(c, kSO, w.f1b, 4 * kC, nullptr, 0, 1, tap0, T, sb.w2, kC, s.b2, 0, w.x2, out_f32, out_b16, true);
    }
}

static void decoder_eval_bf16(Ctx &c, const DecW &W, const DecWb &Wb, const float *x, const float *mu,
                              const float *tb_rows, int n_rows, int tb_per_sample, const float *x_in, float dt,
                              float *out) {
    static const int taps3[3] = {-1, 0, 1};
    static const int taps_even[2] = {0, -1};
    static const int taps_odd[2] = {1, 0};
    DecWs &w = c.ws;
    const int T = c.T, T2 = (T + 1) / 2, Tu = 2 * T2;
    auto tb = [&](int st) { return tb_rows + (size_t)st * n_rows * kC; };
    dim3 pgrid((T + 31) / 32, kInChPad / 32, c.B);
    PackLoss pl;
    pl.z = c.loss_z; pl.tvec = c.loss_t; pl.one_minus_sigma = c.loss_oms; pl.phi_out = c.loss_phi; pl.sums = c.loss_sums;
    launch(c, pack_input_bf16_kernel, pgrid, 256, 0, x, mu, c.mask, w.x0b, c.f16, stream_down(c), c.B, T, w.gstats, kGnConvs * c.B * kGroups * 2, pl);
    // down 0 @T; the stride-2 Downsample1D (decoder.py:95) stays on the fp32 implicit-GEMM kernel
    run_stage_bf16(c, 0, W.st[0], Wb.st[0], w.x0b, kInChPad, nullptr, 0, T, tb(0), tb_per_sample, (T & 1) ? w.s0 : nullptr, w.s0b);
    if ((T & 1) == 0) {
        static const int taps_pair[2] = {-1, 0};
        gemm_b(c, kSA | kSO, w.s0b, 2 * kC, nullptr, 0, 2, taps_pair, T2, Wb.st[0].wtail, kC, W.st[0].btail, 0, nullptr, nullptr, w.inb, true);
    } else {
        GemmArgs g;
        g.a0 = w.s0; g.a1 = nullptr; g.c0 = kC; g.c1 = 0; g.Ta0 = T; g.mask = c.mask; g.mask_T = c.T;
        g.ntaps = 3; g.tap_off[0] = -1; g.tap_off[1] = 0; g.tap_off[2] = 1; g.tap_off[3] = 0; g.stride = 2;
        g.B = c.B; g.Tin = T; g.Tout = T2; g.w = W.st[0].wtail; g.bias = W.st[0].btail; g.N = kC; g.K = 3 * kC;
        g.act = 0; g.res = nullptr; g.out = nullptr; g.out_T = T2; g.out_stride = 1; g.out_off = 0;
        g.out_bf16 = w.inb; g.omask = c.mask; g.o16_scale = stream_down(c); g.f16 = c.f16;
        launch(c, gemm_f32_kernel, dim3((c.B * T2 + kBM - 1) / kBM, kC / kBN), 256, 0, g);
    }
    run_stage_bf16(c, 1, W.st[1], Wb.st[1], w.inb, kC, nullptr, 0, T2, tb(1), tb_per_sample, nullptr, w.s1b);
    gemm_b(c, kSA | kSO, w.s1b, kC, nullptr, 0, 3, taps3, T2, Wb.st[1].wtail, kC, W.st[1].btail, 0, nullptr, nullptr, w.inb, true);
    run_stage_bf16(c, 2, W.st[2], Wb.st[2], w.inb, kC, nullptr, 0, T2, tb(2), tb_per_sample, nullptr, w.upb);
    run_stage_bf16(c, 3, W.st[3], Wb.st[3], w.upb, kC, nullptr, 0, T2, tb(3), tb_per_sample, nullptr, w.inb);
    run_stage_bf16(c, 4, W.st[4], Wb.st[4], w.inb, kC, w.s1b, kC, T2, tb(4), tb_per_sample, nullptr, w.upb);
    // ConvTranspose1d(4,2,1) (decoder.py:112) as two phase GEMMs on the fp32 kernel, then the nearest
    // resample (odd T) + x*mask into the bf16 operand of the last Resnet1D
    gemm_b(c, kSA, w.upb, kC, nullptr, 0, 2, taps_even, T2, Wb.st[4].wtail, kC, W.st[4].btail, 0, nullptr, w.d, nullptr, false, Tu, 2, 0);
    gemm_b(c, kSA, w.upb, kC, nullptr, 0, 2, taps_odd, T2, Wb.st[4].wtail2, kC, W.st[4].btail, 0, nullptr, w.d, nullptr, false, Tu, 2, 1);
    {
        const long long n4 = (long long)c.B * T * (kC / 4);
        int eb = (int)((n4 + 255) / 256); if (eb > 148 * 16) eb = 148 * 16;
        launch(c, resample_mask_bf16_kernel, eb, 256, 0, w.d, Tu, c.mask, w.upb, c.f16, stream_down(c), c.B, T);
    }
    run_stage_bf16(c, 5, W.st[5], Wb.st[5], w.upb, kC, w.s0b, kC, T, tb(5), tb_per_sample, nullptr, w.inb);
    gemm_b(c, kSA | kSO, w.inb, kC, nullptr, 0, 3, taps3, T, Wb.st[5].wtail, kC, W.st[5].btail, 0, nullptr, nullptr, w.a1b, true);
    double *gsf = w.gstats + (size_t)12 * c.B * kGroups * 2;
    gemm_b(c, kSA, w.a1b, kC, nullptr, 0, 3, taps3, T, Wb.wf, kC, W.bf, 0, nullptr, w.h, nullptr, false, 0, 1, 0, gsf);
    {   // final_norm + Mish + x*mask into the 16-bit operand, then final_proj + mask + Euler update in the GEMM epilogue
        static const int tap0[1] = {0};
        const long long n4 = (long long)c.B * T * (kC / 4);
        int eb = (int)((n4 + 255) / 256); if (eb > 148 * 16) eb = 148 * 16;
        launch(c, gn_mish_rows_kernel<0>, (unsigned)(((size_t)c.B * T + kGmRowsPerBlock - 1) / kGmRowsPerBlock), 256, 0, w.h, gsf, W.gf,
               W.bef, c.mask, c.T, nullptr, 0, nullptr, nullptr, nullptr, nullptr, w.a1b, c.f16, c.B, T);
        c.fin_xin = x_in; c.fin_dt = dt; c.fin_out = out; c.fin_loss = c.loss_z != nullptr; c.loss_x1 = x;
        gemm_b(c, 0, w.a1b, kC, nullptr, 0, 1, tap0, T, Wb.wp, kC, W.bp, 0, nullptr, nullptr, nullptr, false);
        c.fin_out = nullptr; c.fin_loss = 0;
    }
}

static TimeWeights time_weights(const DecW &W) {
    TimeWeights t;
    t.freq = W.freq; t.w1 = W.tw1; t.b1 = W.tb1; t.w2 = W.tw2; t.b2 = W.tb2;
    for (int i = 0; i < 6; ++i) { t.wt[i] = W.st[i].wt; t.bt[i] = W.st[i].bt; }
    return t;
}

}  // namespace mb200

using namespace mb200;

namespace {
struct TimeVals { float v[64]; };
__global__ void stage_times_kernel(TimeVals tv, int n, float *out) {
    pdl_entry();
    if (threadIdx.x < n) out[threadIdx.x] = tv.v[threadIdx.x];
}

int check_common(const void *a, const void *b, const void *c, const void *d, const void *e, const void *f, int B, int T) {
    if (B <= 0 || T <= 0) return MB200_ERR_INVALID_ARG;
    if (!a || !b || !c || !d || !e || !f) return MB200_ERR_INVALID_ARG;
    if (T < 2) return MB200_ERR_UNSUPPORTED;
    return MB200_OK;
}
int prepare(Ctx &c, DecW &W, DecWb &Wb, const float *blob, const void *blob_bf16, void *ws, long long ws_bytes,
            const float *mask, int B, int T, int n_rows, int precision, void *stream) {
    if (precision != MB200_PREC_F32 && precision != MB200_PREC_BF16 && precision != MB200_PREC_F16) return MB200_ERR_INVALID_ARG;
    if (precision != MB200_PREC_F32 && !blob_bf16) return MB200_ERR_INVALID_ARG;
    build_weights(blob, &W);
    build_weights_bf16(reinterpret_cast<const bf16 *>(blob_bf16), &Wb);
    if (ws_bytes < carve_workspace(nullptr, B, T, n_rows, precision, nullptr)) return MB200_ERR_INVALID_ARG;
    if ((uintptr_t)ws % 16 != 0 || (uintptr_t)blob % 16 != 0 || (uintptr_t)blob_bf16 % 16 != 0) return MB200_ERR_INVALID_ARG;
    carve_workspace((float *)ws, B, T, n_rows, precision, &c.ws);
    c.st = (cudaStream_t)stream; c.mask = mask; c.B = B; c.T = T; c.launches = 0; c.err = 0; c.f16 = precision == MB200_PREC_F16 ? 1 : 0;
    return MB200_OK;
}
void eval_any(Ctx &c, const DecW &W, const DecWb &Wb, int precision, const float *x, const float *mu,
              const float *tb_rows, int n_rows, int per_sample, const float *x_in, float dt, float *out) {
    if (precision != MB200_PREC_F32) decoder_eval_bf16(c, W, Wb, x, mu, tb_rows, n_rows, per_sample, x_in, dt, out);
    else decoder_eval(c, W, x, mu, tb_rows, n_rows, per_sample, x_in, dt, out);
}
}  // namespace

extern "C" long long matcha_b200_decoder_blob_floats(void) { return build_weights(nullptr, nullptr); }
extern "C" long long matcha_b200_decoder_blob_bf16_bytes(void) { return build_weights_bf16(nullptr, nullptr) * 2; }

// The Euler solve runs the batch as up to kMaxSplit (default 4: measured 10.42 / 9.48 / 9.27 / 9.31 ms per solve with 1 / 2 / 4 /
// 8 at B = 32, T = 800) independent sub-batches on as many streams (forked from and joined back into
// the caller's stream, so a CUDA-graph capture of the caller's stream records both branches).  Batch items never
// interact, and at B = 32 the kernels are latency-bound chains of 100-200 CTAs: two interleaved chains fill the SMs one
// chain leaves idle (wave quantisation, launch gaps, single-tile pipelines).  MB200_SPLIT=1 turns it off.
constexpr int kMaxSplit = 8;
static int solve_split(int B) {
    static int v = -1, forced = 0;
    if (v < 0) { const char *e = getenv("MB200_SPLIT"); forced = e != nullptr; v = e ? atoi(e) : 4; if (v < 1) v = 1; if (v > kMaxSplit) v = kMaxSplit; }
    int s = v;
    // from B = 64 on two chains already fill the machine and more only add contention (B = 64, T = 800, n = 10: 15.9 / 16.0 / 16.3 /
    // 16.4 ms with 2 / 3 / 4 / 8; B = 32: 9.19 / 9.00 / 9.00 / 8.95; B = 16: no difference -- profiles/r2_split_sweep.log)
    if (!forced && B >= 48 && s > 2) s = 2;
    while (s > 1 && B < 4 * s) --s;             // small batches stay whole
    return s;
}
static long long ws_align(long long n) { return (n + 1023) & ~1023LL; }

// The only state the decoder entry points keep: per DEVICE, created on first use and cached -- the opt-in shared-memory
// attribute of the kernels (a per-device function attribute) and the side streams + fork / join events of the sub-batch
// split.  DevGuard holds the device's mutex for the whole enqueue of one call, so two host threads driving the same
// device queue their calls one after the other (the event record / stream wait pairs of one call are never interleaved
// with another call's); different devices do not share anything.
constexpr int kMaxDevices = 64;
struct DevRes {
    std::mutex mu;
    bool attr_set = false, split_ready = false;
    cudaStream_t side[kMaxSplit] = {};
    cudaEvent_t ev_fork = nullptr, ev_join[kMaxSplit] = {};
};
static DevRes g_dev[kMaxDevices];
struct DevGuard {
    DevRes *r = nullptr;
    int rc = MB200_OK;
    DevGuard() {
        int dev = -1;
        if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= kMaxDevices) { rc = MB200_ERR_NO_DEVICE; return; }
        r = &g_dev[dev];
        r->mu.lock();
    }
    ~DevGuard() { if (r) r->mu.unlock(); }
};
static int set_attributes(DevRes &r) {
    if (r.attr_set) return MB200_OK;
    MB200_CUDA_TRY(cudaFuncSetAttribute(attention_f32_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kAttSmemBytes));
    { const int rc = gemm_set_attributes(); if (rc != MB200_OK) return rc; }
    MB200_CUDA_TRY(cudaFuncSetAttribute(attention_tc_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kTaSmemBytes));
    MB200_CUDA_TRY(cudaFuncSetAttribute(attention_tc_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, kTaSmemBytes));
    r.attr_set = true;
    return MB200_OK;
}
static int split_resources(DevRes &r) {
    if (r.split_ready) return MB200_OK;
    MB200_CUDA_TRY(cudaEventCreateWithFlags(&r.ev_fork, cudaEventDisableTiming));
    for (int i = 1; i < kMaxSplit; ++i) {
        MB200_CUDA_TRY(cudaStreamCreateWithFlags(&r.side[i], cudaStreamNonBlocking));
        MB200_CUDA_TRY(cudaEventCreateWithFlags(&r.ev_join[i], cudaEventDisableTiming));
    }
    r.split_ready = true;
    return MB200_OK;
}

extern "C" long long matcha_b200_decoder_workspace_bytes(int B, int T, int n_rows, int precision) {
    if (B <= 0 || T <= 0 || n_rows <= 0) return MB200_ERR_INVALID_ARG;
    const int S = solve_split(B);
    long long total = 0;
    for (int i = 0; i < S; ++i) total += ws_align(carve_workspace(nullptr, B / S + (i < B % S ? 1 : 0), T, n_rows, precision, nullptr));
    const long long whole = carve_workspace(nullptr, B, T, n_rows, precision, nullptr);       // decoder_forward uses it unsplit
    return total > whole ? total : whole;
}

extern "C" int matcha_b200_decoder_pack_bf16(const float *blob_dev, void *blob_bf16_dev, int precision, void *stream) {
    if (!blob_dev || !blob_bf16_dev) return MB200_ERR_INVALID_ARG;
    if (precision != MB200_PREC_BF16 && precision != MB200_PREC_F16) return MB200_ERR_INVALID_ARG;
    const int f16 = precision == MB200_PREC_F16 ? 1 : 0;
    DecW W; DecWb Wb;
    build_weights(blob_dev, &W);
    build_weights_bf16(reinterpret_cast<bf16 *>(blob_bf16_dev), &Wb);
    cudaStream_t st = (cudaStream_t)stream;
    auto pack = [&](const float *src, const bf16 *dst, int rows, int cin, int cin_pad, int dst_stride = 0) {
        pack_weight_bf16_kernel<<<256, 256, 0, st>>>(src, const_cast<bf16 *>(dst), rows, cin, cin_pad, f16, dst_stride ? dst_stride : cin_pad);
    };
    for (int i = 0; i < 6; ++i) {
        const StageW &s = W.st[i]; const StageWb &b = Wb.st[i];
        pack(s.wc1, b.wc1, 256 * 3, s.cin, b.cin_pad);
        {   // res_conv rows: zeros, then [256][cin] into the centre-tap columns
            bf16 *wr = const_cast<bf16 *>(b.wc1) + 256LL * 3 * b.cin_pad;
            MB200_CUDA_TRY(cudaMemsetAsync(wr, 0, 256LL * 3 * b.cin_pad * sizeof(bf16), st));
            pack(s.wr, wr + b.cin_pad, 256, s.cin, b.cin_pad, 3 * b.cin_pad);
        }
        pack(s.wc2, b.wc2, 256 * 3, 256, 256);
        pack(s.wqkv, b.wqkv, 768, 256, 256);
        pack(s.wo, b.wo, 256, 256, 256);
        pack(s.w1, b.w1, 1024, 256, 256);
        pack(s.w2, b.w2, 256, 1024, 1024);
        if (i == 0) pack_down_weight_kernel<<<256, 256, 0, st>>>(s.wtail, const_cast<bf16 *>(b.wtail), f16);
        if (i == 1 || i == 5) pack(s.wtail, b.wtail, 256 * 3, 256, 256);
        if (i == 4) { pack(s.wtail, b.wtail, 256 * 2, 256, 256); pack(s.wtail2, b.wtail2, 256 * 2, 256, 256); }
    }
    pack(W.wf, Wb.wf, 256 * 3, 256, 256);
    MB200_CUDA_TRY(cudaMemsetAsync(const_cast<bf16 *>(Wb.wp), 0, 256 * 256 * sizeof(bf16), st));
    pack(W.wp, Wb.wp, kMel, 256, 256);
    MB200_CUDA_TRY(cudaGetLastError());
    return MB200_OK;
}

// Loss mode of the forward (16-bit precisions): x_dev is x1, phi_t and the two loss sums are produced by the first / last kernel.
struct LossFwd { const float *z; float one_minus_sigma; float *phi; double *sums; };
static int forward_impl(const float *x_dev, const float *mu_dev, const float *mask_dev,
                        const float *t_dev, int n_t, float *out_dev, const float *blob_dev,
                        const void *blob_bf16_dev, void *ws_dev, long long ws_bytes, int B, int T,
                        int precision, void *stream, const LossFwd *loss) {
    int rc = check_common(x_dev, mu_dev, mask_dev, t_dev, loss ? (const void *)loss->phi : (const void *)out_dev, blob_dev, B, T);
    if (rc != MB200_OK) return rc;
    if (!ws_dev || (n_t != 1 && n_t != B)) return MB200_ERR_INVALID_ARG;
    if (loss && (precision == MB200_PREC_F32 || n_t != B || !loss->z || !loss->sums)) return MB200_ERR_INVALID_ARG;
    // same sub-batch streams as the Euler solve (one eval is a latency-bound chain of 46 launches at B = 32)
    const int S = solve_split(B);
    DevGuard guard;
    if (guard.rc != MB200_OK) return guard.rc;
    DevRes &dr = *guard.r;
    rc = set_attributes(dr); if (rc != MB200_OK) return rc;
    if (S > 1) { rc = split_resources(dr); if (rc != MB200_OK) return rc; }
    cudaStream_t *g_side = dr.side;
    cudaEvent_t g_ev_fork = dr.ev_fork, *g_ev_join = dr.ev_join;
    Ctx c[kMaxSplit]; DecW W; DecWb Wb;
    int b0[kMaxSplit + 1] = {};
    char *wsp = (char *)ws_dev;
    long long left = ws_bytes;
    for (int i = 0; i < S; ++i) {
        const int Bi = B / S + (i < B % S ? 1 : 0);
        b0[i + 1] = b0[i] + Bi;
        const long long need = S == 1 ? 0 : ws_align(carve_workspace(nullptr, Bi, T, n_t, precision, nullptr));
        rc = prepare(c[i], W, Wb, blob_dev, blob_bf16_dev, wsp, left, mask_dev + (size_t)b0[i] * T, Bi, T, n_t, precision,
                     i == 0 ? stream : (void *)g_side[i]);
        if (rc != MB200_OK) return rc;
        wsp += need; left -= need;
    }
    launch(c[0], time_embed1_kernel, dim3(n_t, 16), 256, 0, time_weights(W), t_dev, c[0].ws.th2);
    launch(c[0], time_embed2_kernel, dim3(n_t, 24), 256, 0, time_weights(W), c[0].ws.th2, n_t, c[0].ws.tb);
    if (S > 1) {
        MB200_CUDA_TRY(cudaEventRecord(g_ev_fork, c[0].st));
        for (int i = 1; i < S; ++i) MB200_CUDA_TRY(cudaStreamWaitEvent(c[i].st, g_ev_fork, 0));
    }
    const int per_sample = n_t == B && B > 1 ? 1 : 0;
    for (int i = 0; i < S; ++i) {
        const size_t off = (size_t)b0[i] * kMel * T;
        const float *tb = c[0].ws.tb + (per_sample ? (size_t)b0[i] * kC : 0);     // rows of this sub-batch's samples
        if (loss) {
            c[i].loss_z = loss->z + off; c[i].loss_t = t_dev + b0[i]; c[i].loss_oms = loss->one_minus_sigma;
            c[i].loss_phi = loss->phi + off; c[i].loss_sums = loss->sums;
        }
        eval_any(c[i], W, Wb, precision, x_dev + off, mu_dev + off, tb, n_t, per_sample, nullptr, 0.f, out_dev ? out_dev + off : nullptr);
    }
    int launches = 0;
    for (int i = 0; i < S; ++i) { if (c[i].err) return c[i].err; launches += c[i].launches; }
    for (int i = 1; i < S; ++i) {
        MB200_CUDA_TRY(cudaEventRecord(g_ev_join[i], c[i].st));
        MB200_CUDA_TRY(cudaStreamWaitEvent(c[0].st, g_ev_join[i], 0));
    }
    MB200_CUDA_TRY(cudaGetLastError());
    return launches;
}

extern "C" int matcha_b200_decoder_forward(const float *x_dev, const float *mu_dev, const float *mask_dev,
                                            const float *t_dev, int n_t, float *out_dev, const float *blob_dev,
                                            const void *blob_bf16_dev, void *ws_dev, long long ws_bytes, int B, int T,
                                            int precision, void *stream) {
    return forward_impl(x_dev, mu_dev, mask_dev, t_dev, n_t, out_dev, blob_dev, blob_bf16_dev, ws_dev, ws_bytes, B, T, precision,
                        stream, nullptr);
}

// BaseConditionalFlowMatching.compute_loss, forward value (flow_matching.py:106-151), in one call: phi_t, Decoder.forward with
// per-sample t, and the two sums of the masked MSE.  16-bit precisions: phi_t is formed inside the input-packing kernel and
// the squared error inside the final projection's epilogue (u_pred and u_target never reach HBM; u_pred_dev may be NULL).
// fp32 precision: the separate mix / mse kernels of align.cu around the fp32 forward (u_pred_dev required).
extern "C" int matcha_b200_cfm_loss_mix(float *, const float *, const float *, const float *, double, int, long long, void *);
extern "C" int matcha_b200_cfm_loss_mse(double *, const float *, const float *, const float *, const float *, double, long long,
                                         long long, void *);
extern "C" int matcha_b200_cfm_loss_forward(const float *x1_dev, const float *z_dev, const float *t_dev, const float *mu_dev,
                                             const float *mask_dev, double sigma_min, float *phi_dev, float *u_pred_dev,
                                             double *sums_dev, const float *blob_dev, const void *blob_bf16_dev, void *ws_dev,
                                             long long ws_bytes, int B, int T, int precision, void *stream) {
    if (!x1_dev || !z_dev || !t_dev || !mu_dev || !mask_dev || !phi_dev || !sums_dev || B < 1 || T < 1) return MB200_ERR_INVALID_ARG;
    if (precision == MB200_PREC_F32) {
        if (!u_pred_dev) return MB200_ERR_INVALID_ARG;
        int rc = matcha_b200_cfm_loss_mix(phi_dev, x1_dev, z_dev, t_dev, sigma_min, B, (long long)kMel * T, stream);
        if (rc != MB200_OK) return rc;
        const int n = forward_impl(phi_dev, mu_dev, mask_dev, t_dev, B, u_pred_dev, blob_dev, blob_bf16_dev, ws_dev, ws_bytes, B, T,
                                   precision, stream, nullptr);
        if (n < 0) return n;
        rc = matcha_b200_cfm_loss_mse(sums_dev, u_pred_dev, x1_dev, z_dev, mask_dev, sigma_min, (long long)B * kMel * T,
                                      (long long)B * T, stream);
        return rc != MB200_OK ? rc : n + 2;
    }
    MB200_CUDA_TRY(cudaMemsetAsync(sums_dev, 0, 2 * sizeof(double), (cudaStream_t)stream));
    LossFwd lf;
    lf.z = z_dev; lf.one_minus_sigma = (float)(1.0 - sigma_min); lf.phi = phi_dev; lf.sums = sums_dev;
    return forward_impl(x1_dev, mu_dev, mask_dev, t_dev, B, u_pred_dev, blob_dev, blob_bf16_dev, ws_dev, ws_bytes, B, T, precision,
                        stream, &lf);
}

extern "C" int matcha_b200_cfm_solve_euler(const float *x0_dev, const float *mu_dev, const float *mask_dev,
                                            float *out_dev, const float *t_host, const float *dt_host, int n_steps,
                                            const float *blob_dev, const void *blob_bf16_dev, void *ws_dev,
                                            long long ws_bytes, int B, int T, int precision, void *stream) {
    int rc = check_common(x0_dev, mu_dev, mask_dev, out_dev, blob_dev, t_host, B, T);
    if (rc != MB200_OK) return rc;
    if (!ws_dev || !dt_host || n_steps < 1 || n_steps > 64) return MB200_ERR_INVALID_ARG;
    const int S = solve_split(B);
    DevGuard guard;
    if (guard.rc != MB200_OK) return guard.rc;
    DevRes &dr = *guard.r;
    rc = set_attributes(dr); if (rc != MB200_OK) return rc;
    if (S > 1) { rc = split_resources(dr); if (rc != MB200_OK) return rc; }
    cudaStream_t *side = dr.side;
    cudaEvent_t ev_fork = dr.ev_fork, *ev_join = dr.ev_join;
    Ctx c[kMaxSplit]; DecW W; DecWb Wb;
    int b0[kMaxSplit + 1] = {};
    char *wsp = (char *)ws_dev;
    long long left = ws_bytes;
    for (int i = 0; i < S; ++i) {
        const int Bi = B / S + (i < B % S ? 1 : 0);
        b0[i + 1] = b0[i] + Bi;
        const long long need = ws_align(carve_workspace(nullptr, Bi, T, n_steps, precision, nullptr));
        rc = prepare(c[i], W, Wb, blob_dev, blob_bf16_dev, wsp, left, mask_dev + (size_t)b0[i] * T, Bi, T, n_steps, precision,
                     i == 0 ? stream : (void *)side[i]);
        if (rc != MB200_OK) return rc;
        wsp += need; left -= need;
    }
    // all time embeddings and per-Resnet time biases of the grid in one launch (shared by the sub-batches)
    TimeVals tv;
    for (int i = 0; i < n_steps; ++i) tv.v[i] = t_host[i];
    float *t_stage = c[0].ws.h;          // free until the first conv writes it
    launch(c[0], stage_times_kernel, 1, 64, 0, tv, n_steps, t_stage);
    launch(c[0], time_embed1_kernel, dim3(n_steps, 16), 256, 0, time_weights(W), t_stage, c[0].ws.th2);
    launch(c[0], time_embed2_kernel, dim3(n_steps, 24), 256, 0, time_weights(W), c[0].ws.th2, n_steps, c[0].ws.tb);
    if (S > 1) {
        MB200_CUDA_TRY(cudaEventRecord(ev_fork, c[0].st));
        for (int i = 1; i < S; ++i) MB200_CUDA_TRY(cudaStreamWaitEvent(c[i].st, ev_fork, 0));
    }
    for (int k = 0; k < n_steps; ++k)
        for (int i = 0; i < S; ++i) {
            const size_t off = (size_t)b0[i] * kMel * T;
            const float *xin = (k == 0 ? x0_dev : out_dev) + off;
            eval_any(c[i], W, Wb, precision, xin, mu_dev + off, c[0].ws.tb + (size_t)k * kC, n_steps, 0, xin, dt_host[k], out_dev + off);
        }
    int launches = 0;
    for (int i = 0; i < S; ++i) { if (c[i].err) return c[i].err; launches += c[i].launches; }
    for (int i = 1; i < S; ++i) {
        MB200_CUDA_TRY(cudaEventRecord(ev_join[i], c[i].st));
        MB200_CUDA_TRY(cudaStreamWaitEvent(c[0].st, ev_join[i], 0));
    }
    MB200_CUDA_TRY(cudaGetLastError());
    return launches;
}

// test aid, not part of the public header: one tcgen05 implicit GEMM on caller-provided bf16 operands
extern "C" int matcha_b200_debug_gemm_bf16(const void *a0, int c0, const void *a1, int c1, int ntaps, const int *taps,
                                            int B, int T, const void *w, int N, const float *bias, int act,
                                            const float *res, float *out_f32, void *out_bf16, const float *mask,
                                            void *stream) {
    Ctx c;
    c.st = (cudaStream_t)stream; c.mask = mask; c.B = B; c.T = T; c.launches = 0; c.err = 0; c.f16 = 0; c.debug = 1;
    { const int rc = gemm_set_attributes(); if (rc != MB200_OK) return rc; }
    gemm_b(c, 0, (const bf16 *)a0, c0, (const bf16 *)a1, c1, ntaps, taps, T, (const bf16 *)w, N, bias, act, res, out_f32,
           (bf16 *)out_bf16, mask != nullptr);
    if (c.err) return c.err;
    MB200_CUDA_TRY(cudaGetLastError());
    return MB200_OK;
}

// test aid, not part of the public header: the fused feed-forward on caller-provided 16-bit operands
extern "C" int matcha_b200_debug_ffn(const void *a16, const void *w1_16, const float *b1, const void *w2_16,
                                      const float *b2, const float *res, float *out_f32, void *out_16,
                                      const float *mask, int B, int T, int f16, void *stream, const float *ln_in,
                                      const float *ln_g, const float *ln_b) {
    if (!a16 || !w1_16 || !w2_16 || !b1 || !b2 || !res || B < 1 || T < 1) return MB200_ERR_INVALID_ARG;
    Ctx c;
    c.st = (cudaStream_t)stream; c.mask = mask; c.B = B; c.T = T; c.launches = 0; c.err = 0; c.f16 = f16; c.debug = 1;
    { const int rc = gemm_set_attributes(); if (rc != MB200_OK) return rc; }
    ffn_b(c, (const bf16 *)a16, T, (const bf16 *)w1_16, b1, (const bf16 *)w2_16, b2, res, out_f32, (bf16 *)out_16, mask != nullptr,
          ln_in, ln_g, ln_b, nullptr, nullptr, false);
    if (c.err) return c.err;
    MB200_CUDA_TRY(cudaGetLastError());
    return MB200_OK;
}

// test aid, not part of the public header: one A-building GEMM (gemm_abuild.cuh) on caller-provided tensors
extern "C" int matcha_b200_debug_abuild(int mode, const float *h, const double *gsum, const float *gamma, const float *beta,
                                         const float *mask, const float *tb, const float *r, const float *ln_g,
                                         const float *ln_b, float *x1, const void *w16, int N, const float *bias, void *out16,
                                         float *out_f32, double *gstats, int B, int T, int f16, void *stream) {
    if (!h || !gsum || !gamma || !beta || !mask || !w16 || B < 1 || T < 1 || (mode != 1 && mode != 2)) return MB200_ERR_INVALID_ARG;
    Ctx c;
    c.st = (cudaStream_t)stream; c.mask = mask; c.B = B; c.T = T; c.launches = 0; c.err = 0; c.f16 = f16;
    { const int rc = gemm_set_attributes(); if (rc != MB200_OK) return rc; }
    abuild_b(c, mode, h, gsum, gamma, beta, T, tb, 0, r, ln_g, ln_b, x1, (const bf16 *)w16, N, bias, (bf16 *)out16, out_f32, gstats);
    if (c.err) return c.err;
    MB200_CUDA_TRY(cudaGetLastError());
    return MB200_OK;
}

// test aid, not part of the public header: the transformer block's tail in one kernel (ffn_fused.cuh, OUTPROJ variant):
// out = x2 + FFN(LayerNorm(x2)), x2 = att16 Wo^T + bo + x1
extern "C" int matcha_b200_debug_ffn_outproj(const void *att16, const void *wo_16, const float *bo, const float *x1,
                                              const float *ln_g, const float *ln_b, const void *w1_16, const float *b1,
                                              const void *w2_16, const float *b2, float *out_f32, void *out_16,
                                              const float *mask, int B, int T, int f16, void *stream) {
    if (!att16 || !wo_16 || !bo || !x1 || !ln_g || !ln_b || !w1_16 || !w2_16 || !b1 || !b2 || B < 1 || T < 1) return MB200_ERR_INVALID_ARG;
    Ctx c;
    c.st = (cudaStream_t)stream; c.mask = mask; c.B = B; c.T = T; c.launches = 0; c.err = 0; c.f16 = f16; c.debug = 1;
    { const int rc = gemm_set_attributes(); if (rc != MB200_OK) return rc; }
    ffn_b(c, (const bf16 *)att16, T, (const bf16 *)w1_16, b1, (const bf16 *)w2_16, b2, x1, out_f32, (bf16 *)out_16, mask != nullptr,
          nullptr, ln_g, ln_b, (const bf16 *)wo_16, bo, false);
    if (c.err) return c.err;
    MB200_CUDA_TRY(cudaGetLastError());
    return MB200_OK;
}

extern "C" int matcha_b200_debug_ffn_stamps(long long *out64) {
    MB200_CUDA_TRY(cudaMemcpyFromSymbol(out64, mb200::g_ffn_dbg, sizeof(long long) * 64));
    return MB200_OK;
}

extern "C" int matcha_b200_debug_gemm_stamps(long long *out8) {
    MB200_CUDA_TRY(cudaMemcpyFromSymbol(out8, mb200::g_gemm_dbg, sizeof(long long) * 8));
    return MB200_OK;
}

// test aid, not part of the public header: one self-attention call on caller-provided 16-bit q|k|v rows
// ([B][T][768]); impl 0 = tcgen05 kernel, 1 = mma.sync kernel.  mask: [B][T] float.
extern "C" int matcha_b200_debug_attention(const void *qkv16, const float *mask, void *out16, int B, int T, int f16,
                                            int impl, void *stream) {
    if (!qkv16 || !mask || !out16 || B < 1 || T < 1) return MB200_ERR_INVALID_ARG;
    Ctx c;
    c.st = (cudaStream_t)stream; c.mask = mask; c.B = B; c.T = T; c.launches = 0; c.err = 0; c.f16 = f16;
    MB200_CUDA_TRY(cudaFuncSetAttribute(attention_tc_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kTaSmemBytes));
    MB200_CUDA_TRY(cudaFuncSetAttribute(attention_tc_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, kTaSmemBytes));
    const bf16 *qkvb = (const bf16 *)qkv16;
    if (impl == 0) {
        CUtensorMap mq, mkv;
        if (!tmap_for(c, qkvb, 3, (uint64_t)(3 * kC), (uint64_t)T, (uint64_t)B, kDHead, kTaBQ, &mq)) return c.err;
        if (!tmap_for(c, qkvb, 3, (uint64_t)(3 * kC), (uint64_t)T, (uint64_t)B, kDHead, kTaBK, &mkv)) return c.err;
        AttnTcArgs a;
        a.mask = mask; a.mask_T = T; a.out = (bf16 *)out16; a.B = B; a.T = T;
        const int grid = ((T + kTaBQ - 1) / kTaBQ) * kHeads * B;
        if (f16) launch(c, attention_tc_kernel<true>, grid, kTaThreads, kTaSmemBytes, mq, mkv, a);
        else launch(c, attention_tc_kernel<false>, grid, kTaThreads, kTaSmemBytes, mq, mkv, a);
    } else {
        dim3 agrid((T + kFaBQ - 1) / kFaBQ, kHeads, B);
        if (f16) launch(c, attention_mma_kernel<true>, agrid, 128, kFaSmemBytes, qkvb, mask, T, (bf16 *)out16, T);
        else launch(c, attention_mma_kernel<false>, agrid, 128, kFaSmemBytes, qkvb, mask, T, (bf16 *)out16, T);
    }
    if (c.err) return c.err;
    MB200_CUDA_TRY(cudaGetLastError());
    return MB200_OK;
}
