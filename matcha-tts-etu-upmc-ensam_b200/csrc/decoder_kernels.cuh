// decoder_kernels.cuh -- fp32 building blocks of the CFM decoder (channels-last [B][T][C]).
//
// Reference arithmetic restated here (paths under /root/reference/matcha/models/components/):
//   time_embed_kernel   <- SinusoidalPosEmb + TimeStepEmbeddingNet + the six Resnet1D.mlp
//                          projections                     decoder.py:22-31, 44-49, 75, 82
//   pack_input_kernel   <- pack([x, mu], "b * t")           decoder.py:288 (to channels-last)
//   gemm_f32_kernel     <- every nn.Conv1d / nn.ConvTranspose1d / nn.Linear of the path as an
//                          implicit GEMM (taps gathered on load, x*mask applied on load)
//                                                           decoder.py:59,78,95,112,192,239,248,251
//   gn_stats_kernel, gn_mish_kernel <- GroupNorm(8) + Mish + mask (+ time bias)
//                                                           decoder.py:60-61,66,83
//   layernorm_kernel    <- nn.LayerNorm(256)                transformer.py:250,285
//   attention_f32_kernel<- diffusers Attention / SDPA with the additive 0/1 key mask
//   final_proj_kernel   <- final_norm + Mish + final_proj + masks + the Euler update
//                                                           decoder.py:367-371, flow_matching.py:95
#pragma once
#include <cuda_bf16.h>
#include <cuda_fp16.h>

#include "common.cuh"

namespace mb200 {

constexpr int kC = 256;        // decoder width
constexpr int kHeads = 4, kDHead = 64;
constexpr int kTimeDim = 1024;
constexpr int kInCh = 160, kMel = 80;
constexpr int kGroups = 8, kGroupCh = 32;

// torch: softplus with threshold 20, then x * tanh(.)   (nn.Mish)
__device__ __forceinline__ float mish_f(float x) {
    const float sp = (x > 20.0f) ? x : log1pf(expf(x));
    return x * tanhf(sp);
}
// Same function in closed form: tanh(log(1+e^x)) = (n^2 + 2n) / (n^2 + 2n + 2), n = e^x  (one exp, one divide;
// no cancellation for either sign of x).  Used on the 16-bit tensor-core path.
__device__ __forceinline__ float mish_fast_f(float x) {
    const float n = __expf(fminf(x, 20.0f));      // beyond 20 the ratio below is exactly 1 in fp32 (torch's threshold)
    const float p = fmaf(n, n, 2.0f * n);
    return x * __fdividef(p, p + 2.0f);
}
// erf-GELU for the 16-bit path's GEMM epilogue, which is issue-bound: 12 FMA-pipe + 2 MUFU instructions.
// With q = erfc(|x|/sqrt2)/2 from Abramowitz & Stegun 7.1.26 (|abs error| < 0.75e-7):
//   gelu(x) = x*Phi(x) = max(x, 0) - |x|*q      (x >= 0: x - x*q;  x < 0: x*q)
__device__ __forceinline__ float gelu_erf_fast_f(float x) {
    const float ax = fabsf(x);
    float t, ex;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(t) : "f"(fmaf(ax, 0.3275911f * 0.70710678118654752440f, 1.0f)));
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(ex) : "f"(x * x * (-0.5f * 1.4426950408889634f)));   // exp(-x^2/2)
    float poly = fmaf(0.5f * 1.061405429f, t, 0.5f * -1.453152027f);
    poly = fmaf(poly, t, 0.5f * 1.421413741f);
    poly = fmaf(poly, t, 0.5f * -0.284496736f);
    poly = fmaf(poly, t, 0.5f * 0.254829592f);
    return fmaxf(x, 0.0f) - ax * (poly * t * ex);
}
// 16-bit GEMM operands: bf16 (f16 == 0) or IEEE fp16 (f16 == 1); both travel as __nv_bfloat16-typed storage
// The fp16 conversions saturate (cvt.rn.satfinite: |x| > 65504 -> +-65504, never inf), so an out-of-range operand costs accuracy on
// that element instead of poisoning every row it meets with NaN.
__device__ __forceinline__ uint32_t pack16x2(float a, float b, int f16) {
    if (f16) { uint32_t r; asm("cvt.rn.satfinite.f16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(b), "f"(a)); return r; }   // first source -> high half
    __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
    return *reinterpret_cast<uint32_t *>(&h);
}
__device__ __forceinline__ __nv_bfloat16 pack16(float a, int f16) {
    if (f16) { unsigned short r; asm("cvt.rn.satfinite.f16.f32 %0, %1;" : "=h"(r) : "f"(a)); return *reinterpret_cast<__nv_bfloat16 *>(&r); }
    return __float2bfloat16_rn(a);
}
__device__ __forceinline__ float silu_f(float x) { return x / (1.0f + expf(-x)); }
__device__ __forceinline__ float gelu_erf_f(float x) { return 0.5f * x * (1.0f + erff(x * 0.70710678118654752440f)); }

// ------------------------------------------------------------------ time embedding
// n_rows time rows (Euler steps, or batch items for per-sample t), two launches spread over 16-24 CTAs per row.
// out_tb[stage][row][256] = Linear_stage(Mish(time_mlp(sinusoidal(t[row]))))
struct TimeWeights {
    const float *freq;            // [80], built on the host exactly as decoder.py:26-27 does
    const float *w1, *b1;         // [1024][160], [1024]
    const float *w2, *b2;         // [1024][1024], [1024]
    const float *wt[6], *bt[6];   // [256][1024], [256]
};

// step 1, grid (n_rows, 16): h2[row][o] = Mish(W2 SiLU(W1 emb(t[row]) + b1) + b2)[o] for 64 outputs per block
__global__ void __launch_bounds__(256)
time_embed1_kernel(TimeWeights w, const float *__restrict__ t, float *__restrict__ h2) {
    pdl_entry();
    __shared__ float emb[kInCh];
    __shared__ float h1[kTimeDim];
    const int row = blockIdx.x, tid = threadIdx.x;
    const float tv = t[row];
    if (tid < 80) {
        const float a = (1000.0f * tv) * w.freq[tid];           // scale * x * emb, decoder.py:28
        emb[tid] = sinf(a);
        emb[80 + tid] = cosf(a);
    }
    __syncthreads();
    const int warp = tid >> 5, lane = tid & 31;
    // One warp per output with coalesced weight rows, EIGHT outputs in flight per warp: a single dot product per iteration is
    // an L2 round trip plus a shuffle tree (~0.7 us) -- 128 of them in a row made this kernel 113 us of every solve / forward.
    for (int o0 = warp * 8; o0 < kTimeDim; o0 += 64) {
        float acc[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const float *wr = w.w1 + (size_t)(o0 + j) * kInCh;
            float a = 0.f;
#pragma unroll
            for (int i = 0; i < kInCh / 32; ++i) a = fmaf(wr[lane + 32 * i], emb[lane + 32 * i], a);
            acc[j] = a;
        }
#pragma unroll
        for (int s = 16; s > 0; s >>= 1)
#pragma unroll
            for (int j = 0; j < 8; ++j) acc[j] += __shfl_xor_sync(0xffffffffu, acc[j], s);
        if (lane == 0) {
#pragma unroll
            for (int j = 0; j < 8; ++j) h1[o0 + j] = silu_f(acc[j] + w.b1[o0 + j]);
        }
    }
    __syncthreads();
    {                                                           // this block's 64 outputs: 8 per warp, all in flight
        const int o0 = blockIdx.y * 64 + warp * 8;
        float acc[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[j] = 0.f;
#pragma unroll 4
        for (int i = lane; i < kTimeDim; i += 32) {
            const float hv = h1[i];
#pragma unroll
            for (int j = 0; j < 8; ++j) acc[j] = fmaf(w.w2[(size_t)(o0 + j) * kTimeDim + i], hv, acc[j]);
        }
#pragma unroll
        for (int s = 16; s > 0; s >>= 1)
#pragma unroll
            for (int j = 0; j < 8; ++j) acc[j] += __shfl_xor_sync(0xffffffffu, acc[j], s);
        if (lane == 0) {
#pragma unroll
            for (int j = 0; j < 8; ++j) h2[(size_t)row * kTimeDim + o0 + j] = mish_f(acc[j] + w.b2[o0 + j]);   // Mish of Resnet1D.mlp[0]
        }
    }
}
// step 2, grid (n_rows, 6 stages x 4): out_tb[stage][row][o] = Linear_stage(h2[row])[o], 64 outputs per block
__global__ void __launch_bounds__(256)
time_embed2_kernel(TimeWeights w, const float *__restrict__ h2, int n_rows, float *__restrict__ out_tb) {
    pdl_entry();
    __shared__ float hs[kTimeDim];
    const int row = blockIdx.x, st = blockIdx.y >> 2, part = blockIdx.y & 3, tid = threadIdx.x;
    for (int i = tid; i < kTimeDim; i += 256) hs[i] = h2[(size_t)row * kTimeDim + i];
    __syncthreads();
    const int warp = tid >> 5, lane = tid & 31;
    for (int oo = warp; oo < 64; oo += 8) {
        const int o = part * 64 + oo;
        const float *wr = w.wt[st] + (size_t)o * kTimeDim;
        float acc = 0.f;
        for (int i = lane; i < kTimeDim; i += 32) acc = fmaf(wr[i], hs[i], acc);
        for (int s = 16; s > 0; s >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, s);
        if (lane == 0) out_tb[((size_t)st * n_rows + row) * kC + o] = acc + w.bt[st][o];
    }
}

// ------------------------------------------------------------------ input packing
// X0[b][t][0:80] = x[b][:, t], X0[b][t][80:160] = mu[b][:, t]   (channels-first -> channels-last)
__global__ void __launch_bounds__(256)
pack_input_kernel(const float *__restrict__ x, const float *__restrict__ mu, float *__restrict__ out,
                  int B, int T) {
    pdl_entry();
    __shared__ float tile[32][33];
    const int b = blockIdx.z, t0 = blockIdx.x * 32, c0 = blockIdx.y * 32;   // c0 over 160 channels
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;                  // 32 x 8
    for (int i = ty; i < 32; i += 8) {
        const int c = c0 + i, t = t0 + tx;
        float v = 0.f;
        if (t < T) v = (c < kMel) ? x[((size_t)b * kMel + c) * T + t] : mu[((size_t)b * kMel + (c - kMel)) * T + t];
        tile[i][tx] = v;
    }
    __syncthreads();
    for (int i = ty; i < 32; i += 8) {
        const int t = t0 + i, c = c0 + tx;
        if (t < T) out[((size_t)b * T + t) * kInCh + c] = tile[tx][i];
    }
}

// ------------------------------------------------------------------ implicit-GEMM (fp32 SIMT)
// out[(b, t*out_stride + out_off)][n] = act(sum_k A[(b,t)][k] * W[n][k] + bias[n]) + res[...][n]
// A[(b,t)][tap*Ct + c] = src(c)[b][t*stride + tap_off[tap]][c] * mask[b][t_in]   (0 outside [0,Tin))
// src 0 holds channels [0,c0) (time length Ta0, nearest-resampled to Tin when they differ,
// decoder.py:338-339), src 1 channels [c0, c0+c1) -- the skip concat of decoder.py:341.
struct GemmArgs {
    const float *a0, *a1;
    int c0, c1, Ta0;
    const float *mask; int mask_T;
    int ntaps, tap_off[4], stride;
    int B, Tin, Tout;
    const float *w, *bias;
    int N, K;
    int act;                      // 0 none, 1 erf-GELU
    const float *res;             // same indexing as out, or nullptr
    float *out;                   // fp32 output or nullptr
    int out_T, out_stride, out_off;
    __nv_bfloat16 *out_bf16;      // optional bf16 copy, multiplied by omask[b][out row] (pitch mask_T) and by o16_scale
    const float *omask;
    float o16_scale;
    int f16;
};

constexpr int kBM = 128, kBN = 64, kBK = 16;

__global__ void __launch_bounds__(256)
gemm_f32_kernel(const GemmArgs g) {
    pdl_entry();
    __shared__ __align__(16) float As[kBK][kBM + 4];
    __shared__ __align__(16) float Bs[kBK][kBN + 4];
    const int tid = threadIdx.x;
    const int m0 = blockIdx.x * kBM, n0 = blockIdx.y * kBN;
    const int M = g.B * g.Tout, Ct = g.c0 + g.c1;

    // loader coordinates
    const int lrow = tid >> 1, lkq = (tid & 1) * 2;
    const int lm = m0 + lrow;
    const bool lvalid = lm < M;
    const int lb = lvalid ? lm / g.Tout : 0, lt = lvalid ? lm % g.Tout : 0;
    const float nn_scale = (float)g.Ta0 / (float)g.Tin;
    const int wn = tid >> 2, wkq = tid & 3;

    float4 ra[2], rb;
    auto load_tile = [&](int k0) {
        const int tap = k0 / Ct, cc = k0 - tap * Ct;
        const int tin = lt * g.stride + g.tap_off[tap];
        const bool ok = lvalid && tin >= 0 && tin < g.Tin;
        float mk = 1.0f;
        if (ok && g.mask) mk = g.mask[(size_t)lb * g.mask_T + tin];
        const float *p = nullptr;
        if (ok) {
            if (cc < g.c0) {
                int ts = tin;
                if (g.Ta0 != g.Tin) { ts = (int)floorf((float)tin * nn_scale); if (ts > g.Ta0 - 1) ts = g.Ta0 - 1; }
                p = g.a0 + ((size_t)lb * g.Ta0 + ts) * g.c0 + cc;
            } else {
                p = g.a1 + ((size_t)lb * g.Tin + tin) * g.c1 + (cc - g.c0);
            }
        }
#pragma unroll
        for (int i = 0; i < 2; ++i) {
            float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
            if (ok) v = *reinterpret_cast<const float4 *>(p + (lkq + i) * 4);
            v.x *= mk; v.y *= mk; v.z *= mk; v.w *= mk;
            ra[i] = v;
        }
        rb = *reinterpret_cast<const float4 *>(g.w + (size_t)(n0 + wn) * g.K + k0 + wkq * 4);
    };
    auto store_tile = [&]() {
#pragma unroll
        for (int i = 0; i < 2; ++i) {
            const int k = (lkq + i) * 4;
            As[k + 0][lrow] = ra[i].x; As[k + 1][lrow] = ra[i].y; As[k + 2][lrow] = ra[i].z; As[k + 3][lrow] = ra[i].w;
        }
        const int k = wkq * 4;
        Bs[k + 0][wn] = rb.x; Bs[k + 1][wn] = rb.y; Bs[k + 2][wn] = rb.z; Bs[k + 3][wn] = rb.w;
    };

    const int ty = tid >> 4, tx = tid & 15;
    float acc[8][4];
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;

    load_tile(0);
    for (int k0 = 0; k0 < g.K; k0 += kBK) {
        store_tile();
        __syncthreads();
        if (k0 + kBK < g.K) load_tile(k0 + kBK);
#pragma unroll
        for (int k = 0; k < kBK; ++k) {
            const float4 a_lo = *reinterpret_cast<const float4 *>(&As[k][ty * 8]);
            const float4 a_hi = *reinterpret_cast<const float4 *>(&As[k][ty * 8 + 4]);
            const float4 bv = *reinterpret_cast<const float4 *>(&Bs[k][tx * 4]);
            const float a[8] = {a_lo.x, a_lo.y, a_lo.z, a_lo.w, a_hi.x, a_hi.y, a_hi.z, a_hi.w};
            const float b[4] = {bv.x, bv.y, bv.z, bv.w};
#pragma unroll
            for (int i = 0; i < 8; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
        }
        __syncthreads();
    }

    const int n = n0 + tx * 4;
    float4 bias = make_float4(0.f, 0.f, 0.f, 0.f);
    if (g.bias) bias = *reinterpret_cast<const float4 *>(g.bias + n);
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const int m = m0 + ty * 8 + i;
        if (m >= M) continue;
        const int b = m / g.Tout, t = m - b * g.Tout;
        const size_t o = ((size_t)b * g.out_T + (size_t)t * g.out_stride + g.out_off) * g.N + n;
        float4 v = make_float4(acc[i][0] + bias.x, acc[i][1] + bias.y, acc[i][2] + bias.z, acc[i][3] + bias.w);
        if (g.act == 1) { v.x = gelu_erf_f(v.x); v.y = gelu_erf_f(v.y); v.z = gelu_erf_f(v.z); v.w = gelu_erf_f(v.w); }
        if (g.res) {
            const float4 r = *reinterpret_cast<const float4 *>(g.res + o);
            v.x += r.x; v.y += r.y; v.z += r.z; v.w += r.w;
        }
        if (g.out) *reinterpret_cast<float4 *>(g.out + o) = v;
        if (g.out_bf16) {
            const float mk = g.o16_scale * (g.omask ? g.omask[(size_t)b * g.mask_T + (size_t)t * g.out_stride + g.out_off] : 1.0f);
            *reinterpret_cast<uint2 *>(g.out_bf16 + o) = make_uint2(pack16x2(v.x * mk, v.y * mk, g.f16), pack16x2(v.z * mk, v.w * mk, g.f16));
        }
    }
}

// ------------------------------------------------------------------ GroupNorm(8, 256)
// stats[(b*8+g)*2] = mean, [..+1] = rstd over 32 channels x T frames (biased variance, eps 1e-5;
// the statistics run over the full padded T).
__global__ void __launch_bounds__(256)
gn_stats_kernel(const float *__restrict__ h, int T, float *__restrict__ stats) {
    pdl_entry();
    const int b = blockIdx.x / kGroups, grp = blockIdx.x % kGroups;
    const float *p = h + (size_t)b * T * kC + grp * kGroupCh;
    double s = 0.0, ss = 0.0;
    const int n4 = T * (kGroupCh / 4);
    for (int i = threadIdx.x; i < n4; i += 256) {
        const int t = i >> 3, c4 = i & 7;
        const float4 v = *reinterpret_cast<const float4 *>(p + (size_t)t * kC + c4 * 4);
        s += (double)v.x + (double)v.y + (double)v.z + (double)v.w;
        ss += (double)v.x * v.x + (double)v.y * v.y + (double)v.z * v.z + (double)v.w * v.w;
    }
    __shared__ double rs[8], rss[8];
    for (int o = 16; o > 0; o >>= 1) {
        s += __shfl_xor_sync(0xffffffffu, s, o);
        ss += __shfl_xor_sync(0xffffffffu, ss, o);
    }
    if ((threadIdx.x & 31) == 0) { rs[threadIdx.x >> 5] = s; rss[threadIdx.x >> 5] = ss; }
    __syncthreads();
    if (threadIdx.x == 0) {
        double a = 0.0, q = 0.0;
        for (int i = 0; i < 8; ++i) { a += rs[i]; q += rss[i]; }
        const double n = (double)T * kGroupCh;
        const double mean = a / n;
        double var = q / n - mean * mean;
        if (var < 0.0) var = 0.0;
        stats[blockIdx.x * 2] = (float)mean;
        stats[blockIdx.x * 2 + 1] = (float)(1.0 / sqrt(var + 1e-5));
    }
}

// out = Mish(GN(h)) * mask (+ tb[row_b][c])          Block1D, decoder.py:60-61,66 (+ :83)
__global__ void __launch_bounds__(256)
gn_mish_kernel(const float *__restrict__ h, const float *__restrict__ stats, const float *__restrict__ gamma,
               const float *__restrict__ beta, const float *__restrict__ mask, int mask_T,
               const float *__restrict__ tb, int tb_per_sample, float *__restrict__ out,
               __nv_bfloat16 *__restrict__ out_bf16, int f16, int fast, int B, int T, const double *__restrict__ gsum) {
    pdl_entry();
    const size_t n4 = (size_t)B * T * (kC / 4);
    const double inv_n = 1.0 / ((double)T * kGroupCh);
    for (size_t i = (size_t)blockIdx.x * 256 + threadIdx.x; i < n4; i += (size_t)gridDim.x * 256) {
        const int c = (int)(i % (kC / 4)) * 4;
        const size_t row = i / (kC / 4);
        const int b = (int)(row / T), t = (int)(row % T);
        float mean, rstd;
        if (gsum) {      // sums from the conv GEMM's epilogue (tensor-core path): biased variance in double
            const double m = gsum[(b * kGroups + c / kGroupCh) * 2] * inv_n;
            double var = gsum[(b * kGroups + c / kGroupCh) * 2 + 1] * inv_n - m * m;
            if (var < 0.0) var = 0.0;
            mean = (float)m;
            rstd = 1.0f / sqrtf((float)var + 1e-5f);
        } else {
            mean = stats[(b * kGroups + c / kGroupCh) * 2];
            rstd = stats[(b * kGroups + c / kGroupCh) * 2 + 1];
        }
        const float mk = mask[(size_t)b * mask_T + t];
        const float4 v = *reinterpret_cast<const float4 *>(h + i * 4);
        const float4 ga = *reinterpret_cast<const float4 *>(gamma + c), be = *reinterpret_cast<const float4 *>(beta + c);
        float4 o;
        if (fast) {
            o.x = mish_fast_f((v.x - mean) * rstd * ga.x + be.x) * mk;
            o.y = mish_fast_f((v.y - mean) * rstd * ga.y + be.y) * mk;
            o.z = mish_fast_f((v.z - mean) * rstd * ga.z + be.z) * mk;
            o.w = mish_fast_f((v.w - mean) * rstd * ga.w + be.w) * mk;
        } else {
            o.x = mish_f((v.x - mean) * rstd * ga.x + be.x) * mk;
            o.y = mish_f((v.y - mean) * rstd * ga.y + be.y) * mk;
            o.z = mish_f((v.z - mean) * rstd * ga.z + be.z) * mk;
            o.w = mish_f((v.w - mean) * rstd * ga.w + be.w) * mk;
        }
        if (tb) {
            const float4 tv = *reinterpret_cast<const float4 *>(tb + (size_t)(tb_per_sample ? b : 0) * kC + c);
            o.x += tv.x; o.y += tv.y; o.z += tv.z; o.w += tv.w;
        }
        if (out) *reinterpret_cast<float4 *>(out + i * 4) = o;
        if (out_bf16) {     // the bf16 copy is the next conv's operand: x*mask applied here (decoder.py:65)
            *reinterpret_cast<uint2 *>(out_bf16 + i * 4) = make_uint2(pack16x2(o.x * mk, o.y * mk, f16), pack16x2(o.z * mk, o.w * mk, f16));
        }
    }
}

// Row-wise variant for the tensor-core path (GroupNorm sums come from the conv GEMM's epilogue): one warp per row,
// four rows per warp in flight, the 8 (mean, rstd) pairs of every sample a block touches computed once into shared
// memory.
//   MODE 0: out16 = (Mish(GN(h)) * mask + tb) * mask                      Block1D + time bias, next conv's operand
//   MODE 1: x = Mish(GN(h)) * mask + r;  out_f32 = x;  out16 = LayerNorm(x)   Block1D + res_conv (decoder.py:85) +
//           norm1 of the transformer block (transformer.py:316) in one pass
constexpr int kGmRowsPerWarp = 4, kGmRowsPerBlock = 8 * kGmRowsPerWarp;
template <int MODE>
__global__ void __launch_bounds__(256)
gn_mish_rows_kernel(const float *__restrict__ h, const double *__restrict__ gsum, const float *__restrict__ gamma,
                    const float *__restrict__ beta, const float *__restrict__ mask, int mask_T,
                    const float *__restrict__ tb, int tb_per_sample, const float *__restrict__ r,
                    const float *__restrict__ ln_g, const float *__restrict__ ln_b, float *__restrict__ out_f32,
                    __nv_bfloat16 *__restrict__ out16, int f16, int B, int T) {
    __shared__ float s_stat[kGmRowsPerBlock][kGroups][2];
    pdl_entry();
    const size_t rows = (size_t)B * T;
    const size_t r0 = (size_t)blockIdx.x * kGmRowsPerBlock;
    const int b0 = (int)(r0 / T);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int c0 = lane * 4, c1 = 128 + lane * 4, g0 = lane >> 3, g1 = 4 + (lane >> 3);
    const float4 ga0 = *reinterpret_cast<const float4 *>(gamma + c0), ga1 = *reinterpret_cast<const float4 *>(gamma + c1);
    const float4 be0 = *reinterpret_cast<const float4 *>(beta + c0), be1 = *reinterpret_cast<const float4 *>(beta + c1);
    float4 v0[kGmRowsPerWarp], v1[kGmRowsPerWarp], x0[kGmRowsPerWarp], x1[kGmRowsPerWarp];
    float mk[kGmRowsPerWarp];
    int bs[kGmRowsPerWarp];
#pragma unroll
    for (int j = 0; j < kGmRowsPerWarp; ++j) {               // every load of the warp's four rows first
        const size_t row = r0 + warp * kGmRowsPerWarp + j;
        const bool ok = row < rows;
        const int b = ok ? (int)(row / T) : b0, t = ok ? (int)(row - (size_t)b * T) : 0;
        bs[j] = b;
        mk[j] = ok ? mask[(size_t)b * mask_T + t] : 0.f;
        v0[j] = ok ? *reinterpret_cast<const float4 *>(h + row * kC + c0) : make_float4(0.f, 0.f, 0.f, 0.f);
        v1[j] = ok ? *reinterpret_cast<const float4 *>(h + row * kC + c1) : make_float4(0.f, 0.f, 0.f, 0.f);
        if (MODE == 1) {
            x0[j] = ok ? *reinterpret_cast<const float4 *>(r + row * kC + c0) : make_float4(0.f, 0.f, 0.f, 0.f);
            x1[j] = ok ? *reinterpret_cast<const float4 *>(r + row * kC + c1) : make_float4(0.f, 0.f, 0.f, 0.f);
        } else if (tb) {
            x0[j] = *reinterpret_cast<const float4 *>(tb + (size_t)(tb_per_sample ? b : 0) * kC + c0);
            x1[j] = *reinterpret_cast<const float4 *>(tb + (size_t)(tb_per_sample ? b : 0) * kC + c1);
        } else {
            x0[j] = x1[j] = make_float4(0.f, 0.f, 0.f, 0.f);
        }
    }
    // the GroupNorm sums are read after the row loads have been issued: both L2 round trips overlap
    {
        const int bb = b0 + (int)(threadIdx.x >> 3), gi = threadIdx.x & 7;
        const size_t r_last = r0 + kGmRowsPerBlock - 1 < rows ? r0 + kGmRowsPerBlock - 1 : rows - 1;
        if (bb <= (int)(r_last / T)) {
            const double inv_n = 1.0 / ((double)T * kGroupCh);
            const double m = gsum[(bb * kGroups + gi) * 2] * inv_n;
            double var = gsum[(bb * kGroups + gi) * 2 + 1] * inv_n - m * m;
            if (var < 0.0) var = 0.0;
            s_stat[bb - b0][gi][0] = (float)m;
            s_stat[bb - b0][gi][1] = 1.0f / sqrtf((float)var + 1e-5f);
        }
    }
    __syncthreads();
    float4 lg0, lg1, lb0, lb1;
    if (MODE == 1) {
        lg0 = *reinterpret_cast<const float4 *>(ln_g + c0); lg1 = *reinterpret_cast<const float4 *>(ln_g + c1);
        lb0 = *reinterpret_cast<const float4 *>(ln_b + c0); lb1 = *reinterpret_cast<const float4 *>(ln_b + c1);
    }
#pragma unroll
    for (int j = 0; j < kGmRowsPerWarp; ++j) {
        const size_t row = r0 + warp * kGmRowsPerWarp + j;
        if (row >= rows) break;
        const float m0 = s_stat[bs[j] - b0][g0][0], rs0 = s_stat[bs[j] - b0][g0][1];
        const float m1 = s_stat[bs[j] - b0][g1][0], rs1 = s_stat[bs[j] - b0][g1][1];
        const float k = mk[j];
        float4 a, b;       // Mish(GN(h)) * mask + (time bias | res_conv)
        a.x = mish_fast_f((v0[j].x - m0) * rs0 * ga0.x + be0.x) * k + x0[j].x;
        a.y = mish_fast_f((v0[j].y - m0) * rs0 * ga0.y + be0.y) * k + x0[j].y;
        a.z = mish_fast_f((v0[j].z - m0) * rs0 * ga0.z + be0.z) * k + x0[j].z;
        a.w = mish_fast_f((v0[j].w - m0) * rs0 * ga0.w + be0.w) * k + x0[j].w;
        b.x = mish_fast_f((v1[j].x - m1) * rs1 * ga1.x + be1.x) * k + x1[j].x;
        b.y = mish_fast_f((v1[j].y - m1) * rs1 * ga1.y + be1.y) * k + x1[j].y;
        b.z = mish_fast_f((v1[j].z - m1) * rs1 * ga1.z + be1.z) * k + x1[j].z;
        b.w = mish_fast_f((v1[j].w - m1) * rs1 * ga1.w + be1.w) * k + x1[j].w;
        if (MODE == 0) {
            *reinterpret_cast<uint2 *>(out16 + row * kC + c0) = make_uint2(pack16x2(a.x * k, a.y * k, f16), pack16x2(a.z * k, a.w * k, f16));
            *reinterpret_cast<uint2 *>(out16 + row * kC + c1) = make_uint2(pack16x2(b.x * k, b.y * k, f16), pack16x2(b.z * k, b.w * k, f16));
        } else {
            *reinterpret_cast<float4 *>(out_f32 + row * kC + c0) = a;
            *reinterpret_cast<float4 *>(out_f32 + row * kC + c1) = b;
            float sum = (a.x + a.y) + (a.z + a.w) + (b.x + b.y) + (b.z + b.w);
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
            const float mean = sum * (1.0f / kC);
            float d[8] = {a.x - mean, a.y - mean, a.z - mean, a.w - mean, b.x - mean, b.y - mean, b.z - mean, b.w - mean};
            float q = 0.f;
#pragma unroll
            for (int i = 0; i < 8; ++i) q = fmaf(d[i], d[i], q);
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) q += __shfl_xor_sync(0xffffffffu, q, o);
            const float rstd = rsqrtf(q * (1.0f / kC) + 1e-5f);
            *reinterpret_cast<uint2 *>(out16 + row * kC + c0) =
                make_uint2(pack16x2(d[0] * rstd * lg0.x + lb0.x, d[1] * rstd * lg0.y + lb0.y, f16),
                           pack16x2(d[2] * rstd * lg0.z + lb0.z, d[3] * rstd * lg0.w + lb0.w, f16));
            *reinterpret_cast<uint2 *>(out16 + row * kC + c1) =
                make_uint2(pack16x2(d[4] * rstd * lg1.x + lb1.x, d[5] * rstd * lg1.y + lb1.y, f16),
                           pack16x2(d[6] * rstd * lg1.z + lb1.z, d[7] * rstd * lg1.w + lb1.w, f16));
        }
    }
}

// ------------------------------------------------------------------ LayerNorm(256), eps 1e-5
// one warp per row
__global__ void __launch_bounds__(256)
layernorm_kernel(const float *__restrict__ x, const float *__restrict__ gamma, const float *__restrict__ beta,
                 float *__restrict__ out, __nv_bfloat16 *__restrict__ out_bf16, int f16, size_t rows) {
    pdl_entry();
    const int lane = threadIdx.x & 31;
    const size_t row = (size_t)blockIdx.x * 8 + (threadIdx.x >> 5);
    if (row >= rows) return;
    const float *p = x + row * kC;
    const float4 a = *reinterpret_cast<const float4 *>(p + lane * 4);
    const float4 b = *reinterpret_cast<const float4 *>(p + 128 + lane * 4);
    float s = a.x + a.y + a.z + a.w + b.x + b.y + b.z + b.w;
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    const float mean = s * (1.0f / kC);
    float d[8] = {a.x - mean, a.y - mean, a.z - mean, a.w - mean, b.x - mean, b.y - mean, b.z - mean, b.w - mean};
    float q = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) q = fmaf(d[i], d[i], q);
    for (int o = 16; o > 0; o >>= 1) q += __shfl_xor_sync(0xffffffffu, q, o);
    const float rstd = rsqrtf(q * (1.0f / kC) + 1e-5f);
    const float4 g0 = *reinterpret_cast<const float4 *>(gamma + lane * 4), g1 = *reinterpret_cast<const float4 *>(gamma + 128 + lane * 4);
    const float4 b0 = *reinterpret_cast<const float4 *>(beta + lane * 4), b1 = *reinterpret_cast<const float4 *>(beta + 128 + lane * 4);
    float4 o0 = make_float4(d[0] * rstd * g0.x + b0.x, d[1] * rstd * g0.y + b0.y, d[2] * rstd * g0.z + b0.z, d[3] * rstd * g0.w + b0.w);
    float4 o1 = make_float4(d[4] * rstd * g1.x + b1.x, d[5] * rstd * g1.y + b1.y, d[6] * rstd * g1.z + b1.z, d[7] * rstd * g1.w + b1.w);
    if (out) {
        *reinterpret_cast<float4 *>(out + row * kC + lane * 4) = o0;
        *reinterpret_cast<float4 *>(out + row * kC + 128 + lane * 4) = o1;
    }
    if (out_bf16) {
        *reinterpret_cast<uint2 *>(out_bf16 + row * kC + lane * 4) = make_uint2(pack16x2(o0.x, o0.y, f16), pack16x2(o0.z, o0.w, f16));
        *reinterpret_cast<uint2 *>(out_bf16 + row * kC + 128 + lane * 4) = make_uint2(pack16x2(o1.x, o1.y, f16), pack16x2(o1.z, o1.w, f16));
    }
}

// ------------------------------------------------------------------ attention (fp32 SIMT)
// softmax_j(q_i . k_j / 8 + mask01[b][j]) v_j per head; qkv rows are [q(256) | k(256) | v(256)],
// head h at columns h*64.  The float 0/1 key mask is ADDED to the logits: padded keys
// are not excluded, only frames beyond T are.
constexpr int kAttBQ = 64, kAttBK = 64, kAttPitch = 68;
constexpr int kAttSmemBytes = (2 * kAttBK * kAttPitch + kAttBQ * kAttPitch) * 4;

__global__ void __launch_bounds__(128)
attention_f32_kernel(const float *__restrict__ qkv, const float *__restrict__ mask, int mask_T,
                     float *__restrict__ out, __nv_bfloat16 *__restrict__ out_bf16, int f16, int T) {
    pdl_entry();
    extern __shared__ __align__(16) float att_smem[];
    float (*Ks)[kAttPitch] = reinterpret_cast<float (*)[kAttPitch]>(att_smem);
    float (*Vs)[kAttPitch] = reinterpret_cast<float (*)[kAttPitch]>(att_smem + kAttBK * kAttPitch);
    float (*Ps)[kAttPitch] = reinterpret_cast<float (*)[kAttPitch]>(att_smem + 2 * kAttBK * kAttPitch);
    const int b = blockIdx.z, h = blockIdx.y, q0 = blockIdx.x * kAttBQ;
    const int tid = threadIdx.x, r = tid >> 1, half = tid & 1;
    const int tq = q0 + r;
    const float *base = qkv + (size_t)b * T * (3 * kC) + h * kDHead;

    float q[kDHead];
#pragma unroll
    for (int d = 0; d < kDHead; d += 4) {
        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
        if (tq < T) v = *reinterpret_cast<const float4 *>(base + (size_t)tq * (3 * kC) + d);
        q[d] = v.x; q[d + 1] = v.y; q[d + 2] = v.z; q[d + 3] = v.w;
    }
    float o[32];
#pragma unroll
    for (int i = 0; i < 32; ++i) o[i] = 0.f;
    float m_run = -INFINITY, l_run = 0.f;

    for (int k0 = 0; k0 < T; k0 += kAttBK) {
        __syncthreads();
        for (int i = tid; i < kAttBK * (kDHead / 4); i += 128) {
            const int j = i >> 4, d4 = (i & 15) * 4;
            float4 kv = make_float4(0.f, 0.f, 0.f, 0.f), vv = kv;
            if (k0 + j < T) {
                const float *p = base + (size_t)(k0 + j) * (3 * kC);
                kv = *reinterpret_cast<const float4 *>(p + kC + d4);
                vv = *reinterpret_cast<const float4 *>(p + 2 * kC + d4);
            }
            *reinterpret_cast<float4 *>(&Ks[j][d4]) = kv;
            *reinterpret_cast<float4 *>(&Vs[j][d4]) = vv;
        }
        __syncthreads();
        float s[32];
        float mx = -INFINITY;
#pragma unroll
        for (int jj = 0; jj < 32; ++jj) {
            const int j = half * 32 + jj;
            float acc = 0.f;
#pragma unroll
            for (int d = 0; d < kDHead; d += 4) {
                const float4 kv = *reinterpret_cast<const float4 *>(&Ks[j][d]);
                acc = fmaf(q[d], kv.x, acc); acc = fmaf(q[d + 1], kv.y, acc);
                acc = fmaf(q[d + 2], kv.z, acc); acc = fmaf(q[d + 3], kv.w, acc);
            }
            const int tk = k0 + j;
            s[jj] = (tk < T) ? fmaf(acc, 0.125f, mask[(size_t)b * mask_T + tk]) : -INFINITY;
            mx = fmaxf(mx, s[jj]);
        }
        mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 1));
        const float m_new = fmaxf(m_run, mx);
        const float corr = expf(m_run - m_new);          // first block: exp(-inf) = 0
        float ps = 0.f;
#pragma unroll
        for (int jj = 0; jj < 32; ++jj) {
            const float pv = expf(s[jj] - m_new);
            ps += pv;
            Ps[r][half * 32 + jj] = pv;
        }
        ps += __shfl_xor_sync(0xffffffffu, ps, 1);
        l_run = l_run * corr + ps;
        m_run = m_new;
#pragma unroll
        for (int i = 0; i < 32; ++i) o[i] *= corr;
        __syncwarp();
        for (int j = 0; j < kAttBK; ++j) {
            const float pv = Ps[r][j];
#pragma unroll
            for (int d = 0; d < 32; d += 4) {
                const float4 vv = *reinterpret_cast<const float4 *>(&Vs[j][half * 32 + d]);
                o[d] = fmaf(pv, vv.x, o[d]); o[d + 1] = fmaf(pv, vv.y, o[d + 1]);
                o[d + 2] = fmaf(pv, vv.z, o[d + 2]); o[d + 3] = fmaf(pv, vv.w, o[d + 3]);
            }
        }
    }
    if (tq < T) {
        const float inv = 1.0f / l_run;
        const size_t oo = ((size_t)b * T + tq) * kC + h * kDHead + half * 32;
#pragma unroll
        for (int d = 0; d < 32; d += 4) {
            if (out) *reinterpret_cast<float4 *>(out + oo + d) = make_float4(o[d] * inv, o[d + 1] * inv, o[d + 2] * inv, o[d + 3] * inv);
            if (out_bf16) {
                *reinterpret_cast<uint2 *>(out_bf16 + oo + d) = make_uint2(pack16x2(o[d] * inv, o[d + 1] * inv, f16), pack16x2(o[d + 2] * inv, o[d + 3] * inv, f16));
            }
        }
    }
}

// ------------------------------------------------------------------ final norm + proj + Euler
// v[b][n][t] = (sum_c Mish(GN(hf))[b][t][c]*mask * Wp[n][c] + bp[n]) * mask     decoder.py:367-371
// out = x_in + dt * v  (flow_matching.py:95) when x_in != nullptr, else out = v.  Channels-first out.
__global__ void __launch_bounds__(256)
final_proj_kernel(const float *__restrict__ hf, const float *__restrict__ stats, const float *__restrict__ gamma,
                  const float *__restrict__ beta, const float *__restrict__ mask, const float *__restrict__ wp,
                  const float *__restrict__ bp, const float *__restrict__ x_in, float dt,
                  float *__restrict__ out, int T) {
    pdl_entry();
    __shared__ float A[32][kC + 1];
    const int b = blockIdx.y, t0 = blockIdx.x * 32, tid = threadIdx.x;
    for (int i = tid; i < 32 * kC; i += 256) {
        const int tt = i >> 8, c = i & 255;
        const int t = t0 + tt;
        float v = 0.f;
        if (t < T) {
            const float mean = stats[(b * kGroups + c / kGroupCh) * 2], rstd = stats[(b * kGroups + c / kGroupCh) * 2 + 1];
            v = mish_f((hf[((size_t)b * T + t) * kC + c] - mean) * rstd * gamma[c] + beta[c]) * mask[(size_t)b * T + t];
        }
        A[tt][c] = v;
    }
    __syncthreads();
    const int lane = tid & 31, warp = tid >> 5;
    const int t = t0 + lane;
    const float mk = (t < T) ? mask[(size_t)b * T + t] : 0.f;
    for (int n = warp * 10; n < warp * 10 + 10; ++n) {
        const float *wr = wp + (size_t)n * kC;
        float acc = 0.f;
#pragma unroll 8
        for (int c = 0; c < kC; ++c) acc = fmaf(A[lane][c], __ldg(wr + c), acc);
        if (t < T) {
            const float v = (acc + bp[n]) * mk;
            const size_t o = ((size_t)b * kMel + n) * T + t;
            out[o] = x_in ? __fadd_rn(x_in[o], __fmul_rn(dt, v)) : v;   // x + dt * v, two roundings as in torch
        }
    }
}

// ------------------------------------------------------------------ bf16-mode helpers
constexpr int kInChPad = 192;      // 160 input channels padded to three 64-wide K chunks

// X0b[b][t][0:80] = x*mask, [80:160] = mu*mask, [160:192] = 0   (decoder.py:288 + the x*mask of :65,85)
// Loss mode (flow_matching.py:130-138, z != nullptr): x holds x1, the kernel forms phi_t = (1 - (1-sigma) t_b) z + t_b x1 in the
// reference's fp32 operation order, stores it channels-first to phi_out (compute_loss returns it) and packs phi_t * mask;
// the blocks of channel slab 0 add sum(mask) into loss_sums[1] (the denominator of flow_matching.py:149).
struct PackLoss { const float *z, *tvec; float one_minus_sigma; float *phi_out; double *sums; };
__global__ void __launch_bounds__(256)
pack_input_bf16_kernel(const float *__restrict__ x, const float *__restrict__ mu, const float *__restrict__ mask,
                       __nv_bfloat16 *__restrict__ out, int f16, float scale, int B, int T, double *__restrict__ gstats, int n_gstats,
                       const PackLoss loss) {
    pdl_entry();
    __shared__ float tile[32][33];
    const int b = blockIdx.z, t0 = blockIdx.x * 32, c0 = blockIdx.y * 32;   // c0 over 192 channels
    if (blockIdx.x == 0 && blockIdx.y == 0)        // first kernel of an eval: clear the GroupNorm accumulators of its 13 convs
        for (int i = blockIdx.z * 256 + threadIdx.x; i < n_gstats; i += gridDim.z * 256) gstats[i] = 0.0;
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    float tb = 0.f, ta = 0.f;
    if (loss.z) { tb = loss.tvec[b]; ta = __fsub_rn(1.f, __fmul_rn(loss.one_minus_sigma, tb)); }
    for (int i = ty; i < 32; i += 8) {
        const int c = c0 + i, t = t0 + tx;
        float v = 0.f;
        if (t < T && c < kInCh) {
            if (c < kMel) {
                const size_t o = ((size_t)b * kMel + c) * T + t;
                v = x[o];
                if (loss.z) { v = __fadd_rn(__fmul_rn(ta, loss.z[o]), __fmul_rn(tb, v)); loss.phi_out[o] = v; }
            } else {
                v = mu[((size_t)b * kMel + (c - kMel)) * T + t];
            }
            v *= mask[(size_t)b * T + t] * scale;
        }
        tile[i][tx] = v;
    }
    if (loss.z && blockIdx.y == 0 && ty == 0) {
        float m = t0 + tx < T ? mask[(size_t)b * T + t0 + tx] : 0.f;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) m += __shfl_xor_sync(0xffffffffu, m, o);
        if (tx == 0) atomicAdd(loss.sums + 1, (double)m);
    }
    __syncthreads();
    for (int i = ty; i < 32; i += 8) {
        const int t = t0 + i, c = c0 + tx;
        if (t < T) out[((size_t)b * T + t) * kInChPad + c] = pack16(tile[tx][i], f16);
    }
}

// out[b][t][:] = bf16(in[b][nearest(t)][:] * mask[b][t]): F.interpolate(mode="nearest") of decoder.py:338-339
// (identity when Tin == T) followed by the x*mask of the next Resnet1D.
__global__ void __launch_bounds__(256)
resample_mask_bf16_kernel(const float *__restrict__ in, int Tin, const float *__restrict__ mask,
                          __nv_bfloat16 *__restrict__ out, int f16, float o16_scale, int B, int T) {
    pdl_entry();
    const size_t n4 = (size_t)B * T * (kC / 4);
    const float scale = (float)Tin / (float)T;
    for (size_t i = (size_t)blockIdx.x * 256 + threadIdx.x; i < n4; i += (size_t)gridDim.x * 256) {
        const int c = (int)(i % (kC / 4)) * 4;
        const size_t row = i / (kC / 4);
        const int b = (int)(row / T), t = (int)(row % T);
        int ts = t;
        if (Tin != T) { ts = (int)floorf((float)t * scale); if (ts > Tin - 1) ts = Tin - 1; }
        const float mk = mask[(size_t)b * T + t] * o16_scale;
        const float4 v = *reinterpret_cast<const float4 *>(in + ((size_t)b * Tin + ts) * kC + c);
        *reinterpret_cast<uint2 *>(out + i * 4) = make_uint2(pack16x2(v.x * mk, v.y * mk, f16), pack16x2(v.z * mk, v.w * mk, f16));
    }
}

// dst[n][tap][0:cin_pad] = bf16(src[n][tap][0:cin]) zero-padded: GEMM weights for the tcgen05 path
__global__ void __launch_bounds__(256)
pack_weight_bf16_kernel(const float *__restrict__ src, __nv_bfloat16 *__restrict__ dst, int rows, int cin, int cin_pad, int f16,
                        int dst_stride) {
    const size_t n = (size_t)rows * cin_pad;
    for (size_t i = (size_t)blockIdx.x * 256 + threadIdx.x; i < n; i += (size_t)gridDim.x * 256) {
        const size_t r = i / cin_pad;
        const int c = (int)(i % cin_pad);
        dst[r * dst_stride + c] = pack16(c < cin ? src[r * cin + c] : 0.f, f16);
    }
}

// Stride-2 Conv1d (Downsample1D, decoder.py:95) on the paired view [B][T/2][512] of an even-T input:
// out[t] = W0 x[2t-1] + W1 x[2t] + W2 x[2t+1] = tap(-1): [0 | W0] . pair[t-1]  +  tap(0): [W1 | W2] . pair[t]
// dst[n][2][512] from src[n][3][256]
__global__ void __launch_bounds__(256)
pack_down_weight_kernel(const float *__restrict__ src, __nv_bfloat16 *__restrict__ dst, int f16) {
    const int n = blockIdx.x;
    for (int i = threadIdx.x; i < 2 * 512; i += 256) {
        const int tap = i >> 9, c = i & 511;
        float v = 0.f;
        if (tap == 0) { if (c >= 256) v = src[((size_t)n * 3 + 0) * 256 + (c - 256)]; }
        else v = src[((size_t)n * 3 + 1 + (c >> 8)) * 256 + (c & 255)];
        dst[(size_t)n * 1024 + i] = pack16(v, f16);
    }
}

}  // namespace mb200
