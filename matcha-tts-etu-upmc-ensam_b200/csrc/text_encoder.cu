// text_encoder.cu -- the reference's TextEncoder (BASELINE configs[3] "text-encoder mu_x ->
// MAS -> CFM loss") as fp32 sm_100a kernels, channels-last [B][T][C] inside, one C call per forward.
//
// Reference arithmetic restated here (/root/reference/matcha/models/components/text_encoder.py):
//   te_embed_kernel      <- nn.Embedding * sqrt(192), sequence_mask                         :390-392
//   te_gemm_kernel       <- every Conv1d of the module as an implicit GEMM over 32-row x 192/256-column tiles with the
//                           surrounding elementwise work in its prologue / epilogue:
//                             ConvReluNorm: conv(x*mask) -> LayerNorm -> ReLU, (residual + projection) * mask   :48-57
//                             MultiHeadAttention: q|k|v projections (one GEMM), output_conv + residual + norm_layers_1
//                                                                                            :191-200, 313-319
//                             FFN: conv(x*mask) -> ReLU; conv -> *mask + residual + norm_layers_2 (-> *mask)    :257-260, 321-328
//                             DurationPredictor: conv(x*mask) -> ReLU -> LayerNorm, twice, then the 256 -> 1 projection
//                             and both masks folded into the second conv's epilogue                :80-97
//   te_attention_kernel  <- MultiHeadAttention._compute_attention: RoPE on the first 48 of 96 head features of q and k
//                           (cos / sin caches built on the host with the reference's torch ops), scores / sqrt(96),
//                           masked_fill(mask == 0, -1e4), softmax, P.V                        :110-148, 202-229
//   te_mean_proj_kernel  <- mean_projection(encoded) * mask, written channels-first [B][80][T]  :398
// The configuration is the one MatchaTTS builds (matcha_tts.py:122-139); anything else is rejected by the host side.
// Dropout is the identity (inference / eval).  Every GEMM row of the padded [B][T] grid is computed, padded tokens
// included: FFN masks only its first conv's input, so relu(conv1) on the first padded token feeds the second k=3 conv at
// the last valid one (tests/test_oracle_text_encoder.py::test_padding_leaks_through_the_ffn_as_in_the_reference).
#include <cuda_runtime.h>
#include <math.h>
#include <stdlib.h>

#include "common.cuh"

namespace mb200 {
namespace te {

constexpr int kC = 192, kF = 768, kH = 2, kD = 96, kRope = 48, kL = 6, kDp = 256, kFeats = 80;
constexpr int kPreK = 5, kPreL = 3, kK = 3;

// Programmatic dependent launch, the safe way round: wait for the predecessor FIRST, then let the successor be scheduled.
// At most two kernels of the chain are then in flight (this one and its successor, whose launch latency hides under this
// kernel's work).  Triggering before the wait lets three or four of these small kernels become resident at once, and that
// was measured to return wrong results from the second encoder layer on (profiles/r2_te_pdl_race.log): a buffer read by
// kernel N (old contents, cached in that SM's L1), rewritten by kernel N+2 and read by kernel N+3 on the same SM came back
// stale -- the waiting kernel's L1 had been invalidated at ITS launch, before kernel N's read.  For the same reason every
// tensor a previous kernel of the chain wrote is read with ld.global.cg (L2 only) below; weights stay on the default path.
#ifndef TE_PDL_EARLY
#define TE_PDL_EARLY 0
#endif
#ifndef TE_LDCG
#define TE_LDCG 1
#endif
__device__ __forceinline__ void pdl_prologue() {
#if TE_PDL_EARLY
    pdl_prologue();
#else
    pdl_wait();
    pdl_launch_dependents();
#endif
}
__device__ __forceinline__ float ld_act(const float *p) {
#if TE_LDCG
    return __ldcg(p);
#else
    return *p;
#endif
}
__device__ __forceinline__ float2 ld_act2(const float *p) {
#if TE_LDCG
    return __ldcg(reinterpret_cast<const float2 *>(p));
#else
    return *reinterpret_cast<const float2 *>(p);
#endif
}
__device__ __forceinline__ float4 ld_act4(const float *p) {
#if TE_LDCG
    return __ldcg(reinterpret_cast<const float4 *>(p));
#else
    return *reinterpret_cast<const float4 *>(p);
#endif
}

// ------------------------------------------------------------------ embedding + mask
// x0[b][t][:] = E[ids[b][t]][:] * sqrt(192);  mask[b][t] = t < lengths[b]   (one warp per token)
__global__ void __launch_bounds__(256)
te_embed_kernel(const long long *__restrict__ ids, const int *__restrict__ lengths, const float *__restrict__ emb,
                int n_vocab, float *__restrict__ x0, float *__restrict__ mask, int B, int T) {
    pdl_prologue();
    const int lane = threadIdx.x & 31;
    const int row = blockIdx.x * 8 + (threadIdx.x >> 5);
    if (row >= B * T) return;
    const int b = row / T, t = row - b * T;
    long long id = ids[row];
    if (id < 0) id = 0;
    if (id >= n_vocab) id = n_vocab - 1;            // out-of-range ids raise in torch; here they are clamped
    const float4 *src = reinterpret_cast<const float4 *>(emb + (size_t)id * kC);
    float4 *dst = reinterpret_cast<float4 *>(x0 + (size_t)row * kC);
    const float s = 13.856406460551018f;            // float(math.sqrt(192)), text_encoder.py:390
    for (int i = lane; i < kC / 4; i += 32) {
        float4 v = src[i];
        v.x *= s; v.y *= s; v.z *= s; v.w *= s;
        dst[i] = v;
    }
    if (lane == 0) mask[row] = t < lengths[b] ? 1.0f : 0.0f;
}

// ------------------------------------------------------------------ implicit GEMM, 32 rows x NT columns per CTA
// acc[(b,t)][n] = sum_{tap,c} A[b][t + tap - pad][c] * (mask_in ? mask[b][t + tap - pad] : 1) * W[n][tap*Cin + c]
// v = acc + bias[n];  relu_pre: v = max(v, 0);  mask_pre: v *= mask[row];  res: v += res[row][n];
// ln_g: v = LayerNorm_N(v) (needs N == NT);  relu_post;  mask_post: v *= mask[row];  out[row][n] = v
// dp_w: logw[row] = (sum_n v[n] * mask[row] * dp_w[n] + dp_b) * mask[row]   (DurationPredictor.output_projection, :96-97)
struct GemmArgs {
    const float *a; int Cin;
    const float *mask;            // [B][T] 0/1
    int mask_in, ntaps, pad;
    int B, T;
    const float *w, *bias; int N, K;
    int relu_pre, mask_pre;
    const float *res;
    const float *ln_g, *ln_b;
    int relu_post, mask_post;
    float *out;
    const float *dp_w, *dp_b; float *dp_out;
};

template <int NT>
__global__ void __launch_bounds__(256)
te_gemm_kernel(const GemmArgs g) {
    constexpr int V = NT == 192 ? 2 : 4;          // a thread's columns come in J groups of V consecutive ones:
    constexpr int J = NT / (32 * V);              //   column(j, e) = j*32*V + lane*V + e  (conflict-free LDS.64 / LDS.128)
    constexpr int CPT = V * J;                    // 6 or 8 columns x 4 rows per thread
    constexpr int PB = NT == 192 ? 194 : 260;     // pitch of a k-row of the weight tile (transposing stores spread over banks)
    constexpr int PA = 36;
    constexpr int NB4 = NT * 16 / 4 / 256;        // float4 weight loads per thread per 16-wide K chunk
    __shared__ __align__(16) float As[2][16][PA];
    __shared__ __align__(16) float Bs[2][16][PB];
    pdl_prologue();
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int M = g.B * g.T;
    const int m0 = blockIdx.x * 32, n0 = blockIdx.y * NT;

    // loaders: threads 0..127 fetch the A tile (row = tid/4, 4 consecutive k), everyone fetches NB4 float4 of W
    const int lrow = (tid >> 2) & 31, lkq = tid & 3;
    const int lm = m0 + lrow;
    const bool lact = tid < 128 && lm < M;
    const int lb = lact ? lm / g.T : 0, lt = lact ? lm - lb * g.T : 0;
    float4 ra = make_float4(0.f, 0.f, 0.f, 0.f), rb[NB4];
    auto load_regs = [&](int k0) {
        ra = make_float4(0.f, 0.f, 0.f, 0.f);
        if (lact) {
            const int tap = k0 / g.Cin, cc = k0 - tap * g.Cin;
            const int tin = lt + tap - g.pad;
            if (tin >= 0 && tin < g.T) {                               // zero padding of the conv outside [0, T)
                ra = ld_act4(g.a + ((size_t)lb * g.T + tin) * g.Cin + cc + lkq * 4);
                if (g.mask_in) {
                    const float mk = ld_act(g.mask + (size_t)lb * g.T + tin);   // x * x_mask before the conv
                    ra.x *= mk; ra.y *= mk; ra.z *= mk; ra.w *= mk;
                }
            }
        }
#pragma unroll
        for (int i = 0; i < NB4; ++i) {
            const int idx = tid + 256 * i, nl = idx >> 2, kq = idx & 3;
            rb[i] = *reinterpret_cast<const float4 *>(g.w + (size_t)(n0 + nl) * g.K + k0 + kq * 4);
        }
    };
    auto store_smem = [&](int buf) {
        if (tid < 128) {
            As[buf][lkq * 4 + 0][lrow] = ra.x; As[buf][lkq * 4 + 1][lrow] = ra.y;
            As[buf][lkq * 4 + 2][lrow] = ra.z; As[buf][lkq * 4 + 3][lrow] = ra.w;
        }
#pragma unroll
        for (int i = 0; i < NB4; ++i) {
            const int idx = tid + 256 * i, nl = idx >> 2, kq = idx & 3;
            Bs[buf][kq * 4 + 0][nl] = rb[i].x; Bs[buf][kq * 4 + 1][nl] = rb[i].y;
            Bs[buf][kq * 4 + 2][nl] = rb[i].z; Bs[buf][kq * 4 + 3][nl] = rb[i].w;
        }
    };

    float acc[4][CPT];
#pragma unroll
    for (int r = 0; r < 4; ++r)
#pragma unroll
        for (int c = 0; c < CPT; ++c) acc[r][c] = 0.f;

    const int nchunks = g.K / 16;
    load_regs(0);
    store_smem(0);
    __syncthreads();
    for (int ch = 0; ch < nchunks; ++ch) {
        const int buf = ch & 1;
        if (ch + 1 < nchunks) load_regs((ch + 1) * 16);               // global loads in flight under the FMAs
#pragma unroll
        for (int k = 0; k < 16; ++k) {
            const float4 av = *reinterpret_cast<const float4 *>(&As[buf][k][warp * 4]);   // same address for the whole warp
            float bv[CPT];
#pragma unroll
            for (int j = 0; j < J; ++j) {
                if (V == 2) {
                    const float2 t2 = *reinterpret_cast<const float2 *>(&Bs[buf][k][j * 64 + lane * 2]);
                    bv[j * 2] = t2.x; bv[j * 2 + 1] = t2.y;
                } else {
                    const float4 t4 = *reinterpret_cast<const float4 *>(&Bs[buf][k][j * 128 + lane * 4]);
                    bv[j * 4] = t4.x; bv[j * 4 + 1] = t4.y; bv[j * 4 + 2] = t4.z; bv[j * 4 + 3] = t4.w;
                }
            }
            const float a[4] = {av.x, av.y, av.z, av.w};
#pragma unroll
            for (int r = 0; r < 4; ++r)
#pragma unroll
                for (int c = 0; c < CPT; ++c) acc[r][c] = fmaf(a[r], bv[c], acc[r][c]);
        }
        if (ch + 1 < nchunks) store_smem(buf ^ 1);
        __syncthreads();
    }

    // ---- epilogue: warp w owns rows 4w..4w+3 of the tile, its 32 lanes cover all NT columns of those rows
    float bias[CPT], lg[CPT], lb_[CPT], dw[CPT];
#pragma unroll
    for (int j = 0; j < J; ++j)
#pragma unroll
        for (int e = 0; e < V; ++e) {
            const int n = n0 + j * 32 * V + lane * V + e;
            bias[j * V + e] = g.bias ? g.bias[n] : 0.f;
            lg[j * V + e] = g.ln_g ? g.ln_g[n] : 1.f;
            lb_[j * V + e] = g.ln_g ? g.ln_b[n] : 0.f;
            dw[j * V + e] = g.dp_w ? g.dp_w[n] : 0.f;
        }
#pragma unroll
    for (int r = 0; r < 4; ++r) {
        const int m = m0 + warp * 4 + r;                              // warp-uniform
        if (m >= M) continue;
        const float mk = ld_act(g.mask + m);
        float v[CPT];
#pragma unroll
        for (int c = 0; c < CPT; ++c) {
            v[c] = acc[r][c] + bias[c];
            if (g.relu_pre) v[c] = fmaxf(v[c], 0.f);
            if (g.mask_pre) v[c] *= mk;
        }
        if (g.res) {
#pragma unroll
            for (int j = 0; j < J; ++j) {
                const float *p = g.res + (size_t)m * g.N + n0 + j * 32 * V + lane * V;
                if (V == 2) { const float2 t2 = ld_act2(p); v[j * 2] += t2.x; v[j * 2 + 1] += t2.y; }
                else { const float4 t4 = ld_act4(p); v[j * 4] += t4.x; v[j * 4 + 1] += t4.y; v[j * 4 + 2] += t4.z; v[j * 4 + 3] += t4.w; }
            }
        }
        if (g.ln_g) {                                                 // nn.LayerNorm(NT), eps 1e-5, biased variance, two passes
            float s = 0.f;
#pragma unroll
            for (int c = 0; c < CPT; ++c) s += v[c];
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
            const float mean = s * (1.0f / NT);
            float q = 0.f;
#pragma unroll
            for (int c = 0; c < CPT; ++c) { v[c] -= mean; q = fmaf(v[c], v[c], q); }
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) q += __shfl_xor_sync(0xffffffffu, q, o);
            const float rstd = 1.0f / sqrtf(q * (1.0f / NT) + 1e-5f);
#pragma unroll
            for (int c = 0; c < CPT; ++c) v[c] = v[c] * rstd * lg[c] + lb_[c];
        }
#pragma unroll
        for (int c = 0; c < CPT; ++c) {
            if (g.relu_post) v[c] = fmaxf(v[c], 0.f);
            if (g.mask_post) v[c] *= mk;
        }
        if (g.out) {
#pragma unroll
            for (int j = 0; j < J; ++j) {
                float *p = g.out + (size_t)m * g.N + n0 + j * 32 * V + lane * V;
                if (V == 2) *reinterpret_cast<float2 *>(p) = make_float2(v[j * 2], v[j * 2 + 1]);
                else *reinterpret_cast<float4 *>(p) = make_float4(v[j * 4], v[j * 4 + 1], v[j * 4 + 2], v[j * 4 + 3]);
            }
        }
        if (g.dp_w) {
            float d = 0.f;
#pragma unroll
            for (int c = 0; c < CPT; ++c) d = fmaf(v[c] * mk, dw[c], d);
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) d += __shfl_xor_sync(0xffffffffu, d, o);
            if (lane == 0) g.dp_out[m] = (d + g.dp_b[0]) * mk;
        }
    }
}

// ------------------------------------------------------------------ attention with rotary embeddings (fp32 SIMT)
// qkv rows are [q(192) | k(192) | v(192)], head h at columns h*96.  One CTA per (sample, head, 32 queries); four threads per
// query row, each holding the whole rotated q, 8 of the 32 keys of a block and 24 of the 96 output features.
constexpr int kAQ = 32, kAK = 32, kAPK = 100, kAPP = 36;

__device__ __forceinline__ float rope_mix(float x, float c, float partner, float s) {
    return __fadd_rn(__fmul_rn(x, c), __fmul_rn(partner, s));       // (x_rope * cos) + (neg_half * sin), text_encoder.py:146
}

__global__ void __launch_bounds__(128)
te_attention_kernel(const float *__restrict__ qkv, const float *__restrict__ mask, const float *__restrict__ rope_cos,
                    const float *__restrict__ rope_sin, float *__restrict__ out, int T) {
    __shared__ __align__(16) float Ks[kAK][kAPK];
    __shared__ __align__(16) float Vs[kAK][kAPK];
    __shared__ __align__(16) float Ps[kAQ][kAPP];
    pdl_prologue();
    const int b = blockIdx.z, h = blockIdx.y, q0 = blockIdx.x * kAQ;
    const int tid = threadIdx.x, r = tid >> 2, quarter = tid & 3;
    const int tq = q0 + r;
    const float *base = qkv + (size_t)b * T * (3 * kC) + h * kD;
    const float *mrow = mask + (size_t)b * T;

    float q[kD];
#pragma unroll
    for (int d = 0; d < kD; d += 4) {
        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
        if (tq < T) v = ld_act4(base + (size_t)tq * (3 * kC) + d);
        q[d] = v.x; q[d + 1] = v.y; q[d + 2] = v.z; q[d + 3] = v.w;
    }
    if (tq < T) {
        const float *cs = rope_cos + (size_t)tq * kRope, *sn = rope_sin + (size_t)tq * kRope;
#pragma unroll
        for (int d = 0; d < kRope / 2; ++d) {
            const float x0 = q[d], x1 = q[d + kRope / 2];
            q[d] = rope_mix(x0, cs[d], -x1, sn[d]);                                   // [-x[24:48], x[0:24]], :130-133
            q[d + kRope / 2] = rope_mix(x1, cs[d + kRope / 2], x0, sn[d + kRope / 2]);
        }
    }
    const float mq = tq < T ? ld_act(mrow + tq) : 0.f;

    float o[kD / 4];
#pragma unroll
    for (int i = 0; i < kD / 4; ++i) o[i] = 0.f;
    float m_run = -INFINITY, l_run = 0.f;

    for (int k0 = 0; k0 < T; k0 += kAK) {
        __syncthreads();
        for (int i = tid; i < kAK * (kD / 4); i += 128) {
            const int j = i / (kD / 4), d4 = (i - j * (kD / 4)) * 4;
            const int tk = k0 + j;
            float4 kv = make_float4(0.f, 0.f, 0.f, 0.f), vv = kv;
            if (tk < T) {
                const float *p = base + (size_t)tk * (3 * kC);
                kv = ld_act4(p + kC + d4);
                vv = ld_act4(p + 2 * kC + d4);
                if (d4 < kRope) {
                    const bool lo = d4 < kRope / 2;
                    float4 pv = ld_act4(p + kC + (lo ? d4 + kRope / 2 : d4 - kRope / 2));
                    if (lo) { pv.x = -pv.x; pv.y = -pv.y; pv.z = -pv.z; pv.w = -pv.w; }
                    const float4 c4 = *reinterpret_cast<const float4 *>(rope_cos + (size_t)tk * kRope + d4);
                    const float4 s4 = *reinterpret_cast<const float4 *>(rope_sin + (size_t)tk * kRope + d4);
                    kv.x = rope_mix(kv.x, c4.x, pv.x, s4.x); kv.y = rope_mix(kv.y, c4.y, pv.y, s4.y);
                    kv.z = rope_mix(kv.z, c4.z, pv.z, s4.z); kv.w = rope_mix(kv.w, c4.w, pv.w, s4.w);
                }
            }
            *reinterpret_cast<float4 *>(&Ks[j][d4]) = kv;
            *reinterpret_cast<float4 *>(&Vs[j][d4]) = vv;
        }
        __syncthreads();
        float s[kAK / 4];
        float mx = -INFINITY;
#pragma unroll
        for (int jj = 0; jj < kAK / 4; ++jj) {
            const int j = jj * 4 + quarter;
            float acc = 0.f;
#pragma unroll
            for (int d = 0; d < kD; d += 4) {
                const float4 kv = *reinterpret_cast<const float4 *>(&Ks[j][d]);
                acc = fmaf(q[d], kv.x, acc); acc = fmaf(q[d + 1], kv.y, acc);
                acc = fmaf(q[d + 2], kv.z, acc); acc = fmaf(q[d + 3], kv.w, acc);
            }
            const int tk = k0 + j;
            float sc = -INFINITY;                                        // keys beyond T do not exist
            if (tk < T) {
                sc = __fdiv_rn(acc, 9.797958971132712f);                 // / math.sqrt(96), :213
                if (mq * ld_act(mrow + tk) == 0.f) sc = -1e4f;                    // masked_fill(mask == 0, -1e4), :220
            }
            s[jj] = sc;
            mx = fmaxf(mx, sc);
        }
        mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 1));
        mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 2));
        const float m_new = fmaxf(m_run, mx);
        const float corr = expf(m_run - m_new);                          // first block: exp(-inf) = 0
        float ps = 0.f;
#pragma unroll
        for (int jj = 0; jj < kAK / 4; ++jj) {
            const float pv = expf(s[jj] - m_new);
            ps += pv;
            Ps[r][jj * 4 + quarter] = pv;
        }
        ps += __shfl_xor_sync(0xffffffffu, ps, 1);
        ps += __shfl_xor_sync(0xffffffffu, ps, 2);
        l_run = l_run * corr + ps;
        m_run = m_new;
#pragma unroll
        for (int i = 0; i < kD / 4; ++i) o[i] *= corr;
        __syncwarp();
#pragma unroll 4
        for (int j = 0; j < kAK; ++j) {
            const float pv = Ps[r][j];
#pragma unroll
            for (int d = 0; d < kD / 4; d += 4) {
                const float4 vv = *reinterpret_cast<const float4 *>(&Vs[j][quarter * (kD / 4) + d]);
                o[d] = fmaf(pv, vv.x, o[d]); o[d + 1] = fmaf(pv, vv.y, o[d + 1]);
                o[d + 2] = fmaf(pv, vv.z, o[d + 2]); o[d + 3] = fmaf(pv, vv.w, o[d + 3]);
            }
        }
    }
    if (tq < T) {
        const float inv = 1.0f / l_run;
        float *po = out + ((size_t)b * T + tq) * kC + h * kD + quarter * (kD / 4);      // "b h t c -> b (h c) t", :225
#pragma unroll
        for (int d = 0; d < kD / 4; d += 4)
            *reinterpret_cast<float4 *>(po + d) = make_float4(o[d] * inv, o[d + 1] * inv, o[d + 2] * inv, o[d + 3] * inv);
    }
}

// ------------------------------------------------------------------ mean projection, channels-first output
// mu[b][n][t] = (sum_c x[b][t][c] * W[n][c] + bias[n]) * mask[b][t]      text_encoder.py:398
__global__ void __launch_bounds__(256)
te_mean_proj_kernel(const float *__restrict__ x, const float *__restrict__ w, const float *__restrict__ bias,
                    const float *__restrict__ mask, float *__restrict__ mu, int T) {
    __shared__ float A[32][kC + 1];
    pdl_prologue();
    const int b = blockIdx.y, t0 = blockIdx.x * 32, tid = threadIdx.x;
    for (int i = tid; i < 32 * kC; i += 256) {
        const int tt = i / kC, c = i - tt * kC;
        A[tt][c] = (t0 + tt < T) ? ld_act(x + ((size_t)b * T + t0 + tt) * kC + c) : 0.f;
    }
    __syncthreads();
    const int lane = tid & 31, warp = tid >> 5;
    const int t = t0 + lane;
    const float mk = t < T ? ld_act(mask + (size_t)b * T + t) : 0.f;
    for (int n = warp * (kFeats / 8); n < (warp + 1) * (kFeats / 8); ++n) {
        const float *wr = w + (size_t)n * kC;
        float acc = 0.f;
#pragma unroll 8
        for (int c = 0; c < kC; ++c) acc = fmaf(A[lane][c], __ldg(wr + c), acc);
        if (t < T) mu[((size_t)b * kFeats + n) * T + t] = (acc + bias[n]) * mk;
    }
}

// ------------------------------------------------------------------ host side
struct LayerW { const float *wqkv, *bqkv, *wo, *bo, *ln1g, *ln1b, *w1, *b1, *w2, *b2, *ln2g, *ln2b; };
struct EncW {
    const float *emb;
    const float *pw[kPreL], *pb[kPreL], *pg[kPreL], *pbe[kPreL], *ppw, *ppb;
    LayerW layer[kL];
    const float *mw, *mb;
    const float *d1w, *d1b, *d1g, *d1be, *d2w, *d2b, *d2g, *d2be, *dpw, *dpb;
};
// Blob order = matcha_b200/text_encoder.py::pack_weights.  Returns the number of floats.
static long long build_weights(const float *blob, int n_vocab, EncW *w) {
    long long off = 0;
    auto take = [&](long long n) { const float *p = blob ? blob + off : nullptr; off += n; return p; };
    EncW e;
    e.emb = take((long long)n_vocab * kC);
    for (int i = 0; i < kPreL; ++i) { e.pw[i] = take(kC * kPreK * kC); e.pb[i] = take(kC); e.pg[i] = take(kC); e.pbe[i] = take(kC); }
    e.ppw = take(kC * kC); e.ppb = take(kC);
    for (int i = 0; i < kL; ++i) {
        LayerW &l = e.layer[i];
        l.wqkv = take(3 * kC * kC); l.bqkv = take(3 * kC); l.wo = take(kC * kC); l.bo = take(kC);
        l.ln1g = take(kC); l.ln1b = take(kC);
        l.w1 = take(kF * kK * kC); l.b1 = take(kF); l.w2 = take(kC * kK * kF); l.b2 = take(kC);
        l.ln2g = take(kC); l.ln2b = take(kC);
    }
    e.mw = take(kFeats * kC); e.mb = take(kFeats);
    e.d1w = take(kDp * kK * kC); e.d1b = take(kDp); e.d1g = take(kDp); e.d1be = take(kDp);
    e.d2w = take(kDp * kK * kDp); e.d2b = take(kDp); e.d2g = take(kDp); e.d2be = take(kDp);
    e.dpw = take(kDp); e.dpb = take(4);        // the bias is one float; padded so every tensor stays 16-byte aligned
    if (w) *w = e;
    return off;
}

struct EncWs { float *x0, *x, *y, *ha, *hb, *qkv, *att, *f1, *dp1; };
static long long carve_workspace(float *base, int B, int T, EncWs *ws) {
    const long long M = (long long)B * T;
    long long off = 0;
    auto take = [&](long long n) { float *p = base ? base + off : nullptr; off += (n + 63) & ~63LL; return p; };
    EncWs w;
    w.x0 = take(M * kC); w.x = take(M * kC); w.y = take(M * kC); w.ha = take(M * kC); w.hb = take(M * kC);
    w.qkv = take(M * 3 * kC); w.att = take(M * kC); w.f1 = take(M * kF); w.dp1 = take(M * kDp);
    if (ws) *ws = w;
    return off * (long long)sizeof(float);
}

struct Ctx { cudaStream_t st; int B, T; const float *mask; int launches = 0, err = 0; };

static bool pdl_enabled() {
    static int v = -1;
    if (v < 0) { const char *e = getenv("MB200_PDL"); v = (e && e[0] == '0') ? 0 : 1; }
    return v != 0;
}
template <typename... KArgs, typename... Args>
static void launch(Ctx &c, void (*kern)(KArgs...), dim3 grid, dim3 block, Args &&...args) {
    if (c.err) return;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = 0; cfg.stream = c.st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at; cfg.numAttrs = pdl_enabled() ? 1 : 0;
    const cudaError_t e = cudaLaunchKernelEx(&cfg, kern, static_cast<KArgs>(args)...);
    if (e != cudaSuccess) { c.err = cuda_fail(e); return; }
    c.launches++;
}

struct Epi {
    int relu_pre = 0, mask_pre = 0; const float *res = nullptr; const float *ln_g = nullptr, *ln_b = nullptr;
    int relu_post = 0, mask_post = 0; const float *dp_w = nullptr, *dp_b = nullptr; float *dp_out = nullptr;
};
static void gemm(Ctx &c, const float *a, int Cin, bool mask_in, int ntaps, const float *w, const float *bias, int N,
                 const Epi &e, float *out) {
    GemmArgs g;
    g.a = a; g.Cin = Cin; g.mask = c.mask; g.mask_in = mask_in ? 1 : 0; g.ntaps = ntaps; g.pad = ntaps / 2;
    g.B = c.B; g.T = c.T; g.w = w; g.bias = bias; g.N = N; g.K = ntaps * Cin;
    g.relu_pre = e.relu_pre; g.mask_pre = e.mask_pre; g.res = e.res; g.ln_g = e.ln_g; g.ln_b = e.ln_b;
    g.relu_post = e.relu_post; g.mask_post = e.mask_post; g.out = out; g.dp_w = e.dp_w; g.dp_b = e.dp_b; g.dp_out = e.dp_out;
    const int M = c.B * c.T;
    if (N % 192 == 0) launch(c, te_gemm_kernel<192>, dim3((M + 31) / 32, N / 192), dim3(256), g);
    else launch(c, te_gemm_kernel<256>, dim3((M + 31) / 32, N / 256), dim3(256), g);
}

static void forward(Ctx &c, const EncW &W, const EncWs &ws, const long long *ids, const int *lengths, int n_vocab,
                    const float *rope_cos, const float *rope_sin, float *mu, float *logw, float *mask) {
    const int M = c.B * c.T;
    launch(c, te_embed_kernel, dim3((M + 7) / 8), dim3(256), ids, lengths, W.emb, n_vocab, ws.x0, mask, c.B, c.T);
    // prenet (ConvReluNorm, :48-57)
    const float *src = ws.x0;
    float *pp[2] = {ws.ha, ws.hb};
    for (int i = 0; i < kPreL; ++i) {
        Epi e; e.ln_g = W.pg[i]; e.ln_b = W.pbe[i]; e.relu_post = 1;
        gemm(c, src, kC, true, kPreK, W.pw[i], W.pb[i], kC, e, pp[i & 1]);
        src = pp[i & 1];
    }
    { Epi e; e.res = ws.x0; e.mask_post = 1; gemm(c, src, kC, false, 1, W.ppw, W.ppb, kC, e, ws.x); }
    // encoder layers (:307-331); ws.x always holds x * x_mask
    int n_layers = kL;
    if (const char *e = getenv("MB200_TE_LAYERS")) n_layers = atoi(e) < kL ? atoi(e) : kL;     // debugging aid: stop after n layers
    for (int i = 0; i < n_layers; ++i) {
        const LayerW &l = W.layer[i];
        { Epi e; gemm(c, ws.x, kC, false, 1, l.wqkv, l.bqkv, 3 * kC, e, ws.qkv); }
        launch(c, te_attention_kernel, dim3((c.T + kAQ - 1) / kAQ, kH, c.B), dim3(128), ws.qkv, c.mask, rope_cos, rope_sin, ws.att, c.T);
        { Epi e; e.res = ws.x; e.ln_g = l.ln1g; e.ln_b = l.ln1b; gemm(c, ws.att, kC, false, 1, l.wo, l.bo, kC, e, ws.y); }
        { Epi e; e.relu_pre = 1; gemm(c, ws.y, kC, true, kK, l.w1, l.b1, kF, e, ws.f1); }
        { Epi e; e.mask_pre = 1; e.res = ws.y; e.ln_g = l.ln2g; e.ln_b = l.ln2b; e.mask_post = 1;
          gemm(c, ws.f1, kF, false, kK, l.w2, l.b2, kC, e, ws.x); }
    }
    launch(c, te_mean_proj_kernel, dim3((c.T + 31) / 32, c.B), dim3(256), ws.x, W.mw, W.mb, c.mask, mu, c.T);
    // duration predictor (:80-97) on the (detached) encoder output
    { Epi e; e.relu_pre = 1; e.ln_g = W.d1g; e.ln_b = W.d1be; gemm(c, ws.x, kC, true, kK, W.d1w, W.d1b, kDp, e, ws.dp1); }
    { Epi e; e.relu_pre = 1; e.ln_g = W.d2g; e.ln_b = W.d2be; e.dp_w = W.dpw; e.dp_b = W.dpb; e.dp_out = logw;
      gemm(c, ws.dp1, kDp, true, kK, W.d2w, W.d2b, kDp, e, nullptr); }
}

}  // namespace te
}  // namespace mb200

using namespace mb200;

extern "C" long long matcha_b200_text_encoder_blob_floats(int n_vocab) {
    if (n_vocab <= 0) return MB200_ERR_INVALID_ARG;
    return te::build_weights(nullptr, n_vocab, nullptr);
}
extern "C" long long matcha_b200_text_encoder_workspace_bytes(int B, int T) {
    if (B <= 0 || T <= 0) return MB200_ERR_INVALID_ARG;
    return te::carve_workspace(nullptr, B, T, nullptr);
}
extern "C" int matcha_b200_text_encoder_forward(const long long *ids_dev, const int *lengths_dev, const float *blob_dev,
                                                int n_vocab, const float *rope_cos_dev, const float *rope_sin_dev,
                                                float *mu_dev, float *logw_dev, float *mask_dev, int B, int T,
                                                void *workspace_dev, long long workspace_bytes, void *stream) {
    if (B <= 0 || T <= 0 || n_vocab <= 0) return MB200_ERR_INVALID_ARG;
    if (!ids_dev || !lengths_dev || !blob_dev || !rope_cos_dev || !rope_sin_dev || !mu_dev || !logw_dev || !mask_dev || !workspace_dev)
        return MB200_ERR_INVALID_ARG;
    if ((long long)B * T > (1LL << 30) / te::kF) return MB200_ERR_UNSUPPORTED;          // int row arithmetic in the kernels
    if (workspace_bytes < te::carve_workspace(nullptr, B, T, nullptr)) return MB200_ERR_INVALID_ARG;
    if ((uintptr_t)workspace_dev % 16 != 0 || (uintptr_t)blob_dev % 16 != 0 || (uintptr_t)rope_cos_dev % 16 != 0 ||
        (uintptr_t)rope_sin_dev % 16 != 0)
        return MB200_ERR_INVALID_ARG;
    te::EncW W;
    te::build_weights(blob_dev, n_vocab, &W);
    te::EncWs ws;
    te::carve_workspace((float *)workspace_dev, B, T, &ws);
    te::Ctx c;
    c.st = (cudaStream_t)stream; c.B = B; c.T = T; c.mask = mask_dev;
    te::forward(c, W, ws, ids_dev, lengths_dev, n_vocab, rope_cos_dev, rope_sin_dev, mu_dev, logw_dev, mask_dev);
    if (c.err) return c.err;
    return c.launches;
}

// ---- kernel-level entry points for the tests (tests/test_text_encoder_gpu.py); not part of the public header
extern "C" int matcha_b200_debug_te_gemm(const float *a, int Cin, const float *mask, int mask_in, int ntaps, int B, int T,
                                         const float *w, const float *bias, int N, int relu_pre, int mask_pre, const float *res,
                                         const float *ln_g, const float *ln_b, int relu_post, int mask_post, float *out,
                                         const float *dp_w, const float *dp_b, float *dp_out, void *stream) {
    if (!a || !mask || !w || B <= 0 || T <= 0 || (N % 192 != 0 && N % 256 != 0) || Cin % 16 != 0) return MB200_ERR_INVALID_ARG;
    if (ln_g && N != 192 && N != 256) return MB200_ERR_INVALID_ARG;
    te::Ctx c;
    c.st = (cudaStream_t)stream; c.B = B; c.T = T; c.mask = mask;
    te::Epi e;
    e.relu_pre = relu_pre; e.mask_pre = mask_pre; e.res = res; e.ln_g = ln_g; e.ln_b = ln_b; e.relu_post = relu_post;
    e.mask_post = mask_post; e.dp_w = dp_w; e.dp_b = dp_b; e.dp_out = dp_out;
    te::gemm(c, a, Cin, mask_in != 0, ntaps, w, bias, N, e, out);
    return c.err ? c.err : c.launches;
}
extern "C" int matcha_b200_debug_te_attention(const float *qkv, const float *mask, const float *rope_cos, const float *rope_sin,
                                              float *out, int B, int T, void *stream) {
    if (!qkv || !mask || !rope_cos || !rope_sin || !out || B <= 0 || T <= 0) return MB200_ERR_INVALID_ARG;
    te::Ctx c;
    c.st = (cudaStream_t)stream; c.B = B; c.T = T; c.mask = mask;
    te::launch(c, te::te_attention_kernel, dim3((T + te::kAQ - 1) / te::kAQ, te::kH, B), dim3(128), qkv, mask, rope_cos, rope_sin, out, T);
    return c.err ? c.err : c.launches;
}
