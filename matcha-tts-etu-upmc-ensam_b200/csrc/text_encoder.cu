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
// Two precisions (MB200_PREC_F32 / MB200_PREC_F16).  fp32: the SIMT kernels above, rel-L2 <= 1e-5 against the reference.
// fp16: every Conv1d runs on the 5th-gen tensor cores (te_gemm_tc_kernel: tcgen05.mma kind::f16, fp32 accumulators in
// TMEM, both operands staged by TMA into 128B-swizzled K-major shared-memory tiles, LayerNorm / ReLU / mask / residual
// / duration projection in the TMEM epilogue); each GEMM's A operand is a 16-bit copy of the activation written by its
// producer with the consumer's x*mask already applied; residual stream, LayerNorm statistics, attention and the mean
// projection stay fp32.  rel-L2 <= 2e-3.
// The configuration is the one MatchaTTS builds (matcha_tts.py:122-139); anything else is rejected by the host side.
// Dropout is the identity (inference / eval).  Every GEMM row of the padded [B][T] grid is computed, padded tokens
// included: FFN masks only its first conv's input, so relu(conv1) on the first padded token feeds the second k=3 conv at
// the last valid one (tests/test_oracle_text_encoder.py::test_padding_leaks_through_the_ffn_as_in_the_reference).
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <math.h>
#include <stdlib.h>

#include <vector>

#include "common.cuh"

namespace mb200 {
namespace te {

constexpr int kC = 192, kF = 768, kH = 2, kD = 96, kRope = 48, kL = 6, kDp = 256, kFeats = 80;
constexpr int kPreK = 5, kPreL = 3, kK = 3;

// Programmatic dependent launch.  Every kernel of the chain runs te_pdl_entry() in every thread at its top: griddepcontrol.wait,
// then griddepcontrol.launch_dependents (TE_PDL_EARLY = 0, the shipped order) -- a kernel lets its successor become resident only
// after its own wait has returned, so at most two kernels of the chain are in flight and the successor's prologue (barriers,
// TMEM allocation, descriptors, parameter loads) overlaps this kernel's body.  Ordering is transitive (a kernel cannot complete
// before its own wait has returned).  VISIBILITY is the delicate part: griddepcontrol.wait only promises that the prerequisite
// grid's writes are visible.  A CTA made resident while kernel N still ran can find, in its SM's L1, a line kernel N read there
// (old contents) after kernel N+2 has rewritten it in L2 -- its own L1 invalidation happened at launch, before kernel N's read.
// A round-2 build that triggered before waiting, with plain loads, returned wrong results from the second encoder layer on, on
// two of six boxes (profiles/r2_te_pdl_race.log: x is read by the out-projection as residual, rewritten by FFN2, read by the next
// q|k|v projection).  Hence the rule of this file: every tensor that a kernel of the chain writes is read through L2 only
// -- ld.global.cg (ld_act*), cp.async.cg, or TMA -- and only weights, biases and the RoPE tables use the default path.
// History: between ece6bc9 and round 3 this helper was called pdl_prologue() and, with TE_PDL_EARLY = 1, called
// ITSELF; nvcc deleted the call, so the shipped kernels carried no griddepcontrol instruction at all although they were launched
// with the PDL attribute.  Round 3 measured the three builds on one box (profiles/r3_te_pdl_ab.log): trigger-then-wait (TE_PDL_EARLY
// = 1, now real) 440 / 539 / 429 us at 32x128 / 8x128 / 1x128 fp16, wait-then-trigger 437 / 413 / 406 us, both bit-stable over 300
// graph replays and equal to the fp64 restatement within tolerance.  The early trigger buys nothing here (three resident waiting
// kernels take SM slots from the one that runs), so the conservative order ships.  tests/test_cabi_cpu.py checks the SASS of
// every te_* kernel for the wait (ACQBULK) so that the instruction cannot silently disappear again.
// TE_PDL_EARLY = 1 and TE_LDCG = 0 are kept as A/B switches.
#ifndef TE_PDL_EARLY
#define TE_PDL_EARLY 0
#endif
#ifndef TE_LDCG
#define TE_LDCG 1
#endif
__device__ __forceinline__ void te_pdl_entry() {    // named apart from every helper in common.cuh: it must never resolve to itself
#if TE_PDL_EARLY
    pdl_launch_dependents();
    pdl_wait();
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
                int n_vocab, float *__restrict__ x0, __half *__restrict__ x0_16, float *__restrict__ mask, int B, int T) {
    te_pdl_entry();
    const int lane = threadIdx.x & 31;
    const int row = blockIdx.x * 8 + (threadIdx.x >> 5);
    if (row >= B * T) return;
    const int b = row / T, t = row - b * T;
    long long id = __ldcg(ids + row);
    if (id < 0) id = 0;
    if (id >= n_vocab) id = n_vocab - 1;            // out-of-range ids raise in torch; here they are clamped
    const float4 *src = reinterpret_cast<const float4 *>(emb + (size_t)id * kC);
    float4 *dst = reinterpret_cast<float4 *>(x0 + (size_t)row * kC);
    const float s = 13.856406460551018f;            // float(math.sqrt(192)), text_encoder.py:390
    const float mk = t < __ldcg(lengths + b) ? 1.0f : 0.0f;
    for (int i = lane; i < kC / 4; i += 32) {
        float4 v = src[i];
        v.x *= s; v.y *= s; v.z *= s; v.w *= s;
        dst[i] = v;
        if (x0_16) {                                // tensor-core mode: the prenet's first conv reads x * x_mask in fp16
            __half2 *d16 = reinterpret_cast<__half2 *>(x0_16 + (size_t)row * kC + 4 * i);
            d16[0] = __floats2half2_rn(v.x * mk, v.y * mk);
            d16[1] = __floats2half2_rn(v.z * mk, v.w * mk);
        }
    }
    if (lane == 0) mask[row] = mk;
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
    te_pdl_entry();
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
__device__ __forceinline__ uint32_t pack_half2_sat(float a, float b) {      // a -> low half; saturating (never inf)
    uint32_t r;
    asm("cvt.rn.satfinite.f16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(b), "f"(a));
    return r;
}
__device__ __forceinline__ long long gtime() { long long t; asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t)); return t; }
constexpr int kAQ = 32, kAK = 32, kAPK = 100, kAPP = 36;
constexpr int kAttTile = kAK * kAPK * 4;                              // one K or V tile: 12.8 KB
constexpr int kAttSmemBytes = 4 * kAttTile + kAQ * kAPP * 4;          // K and V, double-buffered, + P: 55.8 KB (opt-in)

__device__ __forceinline__ float rope_mix(float x, float c, float partner, float s) {
    return __fadd_rn(__fmul_rn(x, c), __fmul_rn(partner, s));       // (x_rope * cos) + (neg_half * sin), text_encoder.py:146
}

// K / V tiles travel with cp.async (16 bytes, L2 only) into two buffers: block kb + 1 is in flight while block kb is
// used.  rope_done = 0 (fp32 mode): q is rotated in registers and every K tile in shared memory once it has landed;
// rope_done = 1 (fp16 mode): the q|k|v projection's epilogue has already rotated both.
__global__ void __launch_bounds__(128)
te_attention_kernel(const float *__restrict__ qkv, const float *__restrict__ mask, const float *__restrict__ rope_cos,
                    const float *__restrict__ rope_sin, int rope_done, float *__restrict__ out, __half *__restrict__ out16, int T,
                    long long *stamps) {
    extern __shared__ __align__(16) uint8_t att_smem[];
    float (*Ps)[kAPP] = reinterpret_cast<float (*)[kAPP]>(att_smem + 4 * kAttTile);
    const uint32_t sm0 = smem_u32(att_smem);
    const bool stamp = stamps && blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0 && threadIdx.x == 0;
    if (stamp) stamps[0] = gtime();
    te_pdl_entry();
    const int b = blockIdx.z, h = blockIdx.y, q0 = blockIdx.x * kAQ;
    const int tid = threadIdx.x, r = tid >> 2, quarter = tid & 3;
    const int tq = q0 + r;
    const float *base = qkv + (size_t)b * T * (3 * kC) + h * kD;
    const float *mrow = mask + (size_t)b * T;

    // tile kb -> buffer kb & 1: K at [buf][0], V at [buf][1]; keys beyond T are zero-filled (src-size 0)
    auto issue_tile = [&](int kb) {
        const uint32_t kdst = sm0 + (uint32_t)(kb & 1) * 2u * kAttTile, vdst = kdst + kAttTile;
        for (int i = tid; i < kAK * (kD / 4); i += 128) {
            const int j = i / (kD / 4), c4 = i - j * (kD / 4);
            const int tk = kb * kAK + j;
            const float *p = base + (size_t)(tk < T ? tk : 0) * (3 * kC) + 4 * c4;
            const uint32_t nbytes = tk < T ? 16u : 0u;
            const uint32_t off = (uint32_t)(j * kAPK + 4 * c4) * 4u;
            asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(kdst + off), "l"(p + kC), "r"(nbytes) : "memory");
            asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(vdst + off), "l"(p + 2 * kC), "r"(nbytes) : "memory");
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
    };
    const int n_kb = (T + kAK - 1) / kAK;
    issue_tile(0);

    float q[kD];
#pragma unroll
    for (int d = 0; d < kD; d += 4) {
        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
        if (tq < T) v = ld_act4(base + (size_t)tq * (3 * kC) + d);
        q[d] = v.x; q[d + 1] = v.y; q[d + 2] = v.z; q[d + 3] = v.w;
    }
    if (tq < T && !rope_done) {
        const float *cs = rope_cos + (size_t)tq * kRope, *sn = rope_sin + (size_t)tq * kRope;
#pragma unroll
        for (int d = 0; d < kRope / 2; ++d) {
            const float x0 = q[d], x1 = q[d + kRope / 2];
            q[d] = rope_mix(x0, cs[d], -x1, sn[d]);                                   // [-x[24:48], x[0:24]], :130-133
            q[d + kRope / 2] = rope_mix(x1, cs[d + kRope / 2], x0, sn[d + kRope / 2]);
        }
    }
    const float mq = tq < T ? ld_act(mrow + tq) : 0.f;
    if (stamp) stamps[1] = gtime();

    float o[kD / 4];
#pragma unroll
    for (int i = 0; i < kD / 4; ++i) o[i] = 0.f;
    float m_run = -INFINITY, l_run = 0.f;

    for (int kb = 0; kb < n_kb; ++kb) {
        const int k0 = kb * kAK;
        if (kb + 1 < n_kb) {
            issue_tile(kb + 1);
            asm volatile("cp.async.wait_group 1;" ::: "memory");
        } else {
            asm volatile("cp.async.wait_group 0;" ::: "memory");
        }
        __syncthreads();
        if (stamp && kb < 3) stamps[2 + 2 * kb] = gtime();
        float (*Ks)[kAPK] = reinterpret_cast<float (*)[kAPK]>(att_smem + (size_t)(kb & 1) * 2 * kAttTile);
        float (*Vs)[kAPK] = reinterpret_cast<float (*)[kAPK]>(att_smem + (size_t)(kb & 1) * 2 * kAttTile + kAttTile);
        if (!rope_done) {
            for (int i = tid; i < kAK * (kRope / 2); i += 128) {
                const int j = i / (kRope / 2), d = i - j * (kRope / 2);
                const int tk = k0 + j;
                if (tk < T) {
                    const float x0 = Ks[j][d], x1 = Ks[j][d + kRope / 2];
                    const float *cs = rope_cos + (size_t)tk * kRope, *sn = rope_sin + (size_t)tk * kRope;
                    Ks[j][d] = rope_mix(x0, __ldg(cs + d), -x1, __ldg(sn + d));
                    Ks[j][d + kRope / 2] = rope_mix(x1, __ldg(cs + d + kRope / 2), x0, __ldg(sn + d + kRope / 2));
                }
            }
            __syncthreads();
        }
        // scores of this thread's 8 keys: feature chunk outermost, so that the 8 shared-memory reads of a chunk are independent
        // and in flight together (key-outermost left one dependent load -> FMA chain per key: 7-8 us per block)
        float s[kAK / 4];
#pragma unroll
        for (int jj = 0; jj < kAK / 4; ++jj) s[jj] = 0.f;
#pragma unroll
        for (int d = 0; d < kD; d += 4) {
            float4 kv[kAK / 4];
#pragma unroll
            for (int jj = 0; jj < kAK / 4; ++jj) kv[jj] = *reinterpret_cast<const float4 *>(&Ks[jj * 4 + quarter][d]);
#pragma unroll
            for (int jj = 0; jj < kAK / 4; ++jj) {
                s[jj] = fmaf(q[d], kv[jj].x, s[jj]); s[jj] = fmaf(q[d + 1], kv[jj].y, s[jj]);
                s[jj] = fmaf(q[d + 2], kv[jj].z, s[jj]); s[jj] = fmaf(q[d + 3], kv[jj].w, s[jj]);
            }
        }
        float mx = -INFINITY;
#pragma unroll
        for (int jj = 0; jj < kAK / 4; ++jj) {
            const int tk = k0 + jj * 4 + quarter;
            float sc = -INFINITY;                                        // keys beyond T do not exist
            if (tk < T) {
                sc = __fdiv_rn(s[jj], 9.797958971132712f);               // / math.sqrt(96), :213
                if (mq * ld_act(mrow + tk) == 0.f) sc = -1e4f;           // masked_fill(mask == 0, -1e4), :220
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
        for (int j0 = 0; j0 < kAK; j0 += 4) {
            const float4 p4 = *reinterpret_cast<const float4 *>(&Ps[r][j0]);
            const float pj[4] = {p4.x, p4.y, p4.z, p4.w};
#pragma unroll
            for (int d = 0; d < kD / 4; d += 4) {
                float4 vv[4];
#pragma unroll
                for (int u = 0; u < 4; ++u) vv[u] = *reinterpret_cast<const float4 *>(&Vs[j0 + u][quarter * (kD / 4) + d]);
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    o[d] = fmaf(pj[u], vv[u].x, o[d]); o[d + 1] = fmaf(pj[u], vv[u].y, o[d + 1]);
                    o[d + 2] = fmaf(pj[u], vv[u].z, o[d + 2]); o[d + 3] = fmaf(pj[u], vv[u].w, o[d + 3]);
                }
            }
        }
        if (stamp && kb < 3) stamps[3 + 2 * kb] = gtime();
        __syncthreads();                                                 // this buffer is refilled by the tile after next
    }
    if (stamp) stamps[8] = gtime();
    if (tq < T) {
        const float inv = 1.0f / l_run;
        const size_t oo = ((size_t)b * T + tq) * kC + h * kD + quarter * (kD / 4);      // "b h t c -> b (h c) t", :225
#pragma unroll
        for (int d = 0; d < kD / 4; d += 4) {
            if (out) *reinterpret_cast<float4 *>(out + oo + d) = make_float4(o[d] * inv, o[d + 1] * inv, o[d + 2] * inv, o[d + 3] * inv);
            if (out16) {
                __half2 *p16 = reinterpret_cast<__half2 *>(out16 + oo + d);
                p16[0] = __floats2half2_rn(o[d] * inv, o[d + 1] * inv);
                p16[1] = __floats2half2_rn(o[d + 2] * inv, o[d + 3] * inv);
            }
        }
    }
}


// ------------------------------------------------------------------ attention on the tensor cores (fp16 mode)
// Same arithmetic as te_attention_kernel with QK^T and PV as mma.sync m16n8k16 (fp16 operands, fp32 accumulate, fp32 online
// softmax in the exp2 domain).  Input: the fp16 q|k|v rows the fused projection's epilogue wrote (RoPE already applied),
// [B*T][576], head h at columns h*96.  One CTA = 4 warps x 16 queries; K / V stream through double-buffered shared-memory tiles
// filled by cp.async (rows of 96 halves at a 256-byte pitch, 16-byte pieces XOR-swizzled by row & 7); fragments by ldmatrix.
// (The SIMT kernel spends its time on shared-memory reads: one 4-FMA group per 16 bytes, 40 us per call; the tensor-core
// form reads each K / V element once per 16 query rows.)
constexpr int kMaBQ = 64, kMaBK = 64, kMaPitch = 256, kMaChunks = kD * 2 / 16;            // 12 sixteen-byte pieces per row
constexpr int kMaTile = kMaBK * kMaPitch;                                                   // 16 KB
constexpr int kMaSmemBytes = kMaBQ * kMaPitch + 4 * kMaTile + 2 * kMaBK * 4;                // Q + 2 x (K, V) + 2 mask tiles = 80.5 KB

__device__ __forceinline__ void ldsm_x4(uint32_t addr, uint32_t &r0, uint32_t &r1, uint32_t &r2, uint32_t &r3) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0, %1, %2, %3}, [%4];"
                 : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(addr));
}
__device__ __forceinline__ void ldsm_x4_t(uint32_t addr, uint32_t &r0, uint32_t &r1, uint32_t &r2, uint32_t &r3) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0, %1, %2, %3}, [%4];"
                 : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(addr));
}
__device__ __forceinline__ void mma16816_f16(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0, %1, %2, %3}, {%4, %5, %6, %7}, {%8, %9}, {%0, %1, %2, %3};"
                 : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3]) : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

__global__ void __launch_bounds__(128)
te_attention_mma_kernel(const __half *__restrict__ qkv, const float *__restrict__ mask, __half *__restrict__ out, int T) {
    extern __shared__ __align__(128) uint8_t ma_smem[];
    const uint32_t sQ = smem_u32(ma_smem);
    const uint32_t sKV = sQ + kMaBQ * kMaPitch;                 // [buf][K | V][64 rows][256 B]
    const float *sMf = reinterpret_cast<const float *>(ma_smem + kMaBQ * kMaPitch + 4 * kMaTile);
    te_pdl_entry();
    const int b = blockIdx.z, h = blockIdx.y, q0 = blockIdx.x * kMaBQ;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const __half *base = qkv + (size_t)b * T * (3 * kC) + h * kD;
    const float *mrow = mask + (size_t)b * T;

    auto load_tile = [&](uint32_t dst, int col_off, int row0) {          // 64 rows x 12 pieces of 16 B; rows beyond T zero-filled
#pragma unroll
        for (int i = 0; i < kMaBK * kMaChunks / 128; ++i) {
            const int idx = tid + i * 128, r = idx / kMaChunks, ch = idx - r * kMaChunks;
            const int t = row0 + r;
            const uint32_t nbytes = t < T ? 16u : 0u;
            const __half *src = base + (size_t)(t < T ? t : 0) * (3 * kC) + col_off + ch * 8;
            asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst + r * kMaPitch + ((ch ^ (r & 7)) << 4)), "l"(src), "r"(nbytes) : "memory");
        }
    };
    // key mask of a block: read through L2 (the rule of this file), parked in shared memory one block ahead
    float *sMw = reinterpret_cast<float *>(ma_smem + kMaBQ * kMaPitch + 4 * kMaTile);
    auto fetch_mask = [&](int row0) { return (tid < kMaBK && row0 + tid < T) ? ld_act(mrow + row0 + tid) : 0.f; };
    load_tile(sQ, 0, q0);
    load_tile(sKV, kC, 0);
    load_tile(sKV + kMaTile, 2 * kC, 0);
    asm volatile("cp.async.commit_group;" ::: "memory");
    if (tid < kMaBK) sMw[tid] = fetch_mask(0);

    const int g = lane >> 2, qd = lane & 3;
    float mq[2];
#pragma unroll
    for (int i = 0; i < 2; ++i) {
        const int t = q0 + warp * 16 + g + 8 * i;
        mq[i] = t < T ? ld_act(mrow + t) : 0.f;
    }
    float o[kD / 8][4];
#pragma unroll
    for (int i = 0; i < kD / 8; ++i) { o[i][0] = o[i][1] = o[i][2] = o[i][3] = 0.f; }
    float m_run[2] = {-INFINITY, -INFINITY}, l_run[2] = {0.f, 0.f};
    uint32_t qa[kD / 16][4];
    const float kLog2e = 1.4426950408889634f;
    const float kScale = 0.10206207261596577f;                  // 1 / sqrt(96)
    const int nkb = (T + kMaBK - 1) / kMaBK;

    for (int kb = 0; kb < nkb; ++kb) {
        const uint32_t sK = sKV + (uint32_t)(kb & 1) * 2u * kMaTile, sV = sK + kMaTile;
        if (kb + 1 < nkb) {                                    // prefetch the next K / V block
            const uint32_t nK = sKV + (uint32_t)((kb + 1) & 1) * 2u * kMaTile;
            load_tile(nK, kC, (kb + 1) * kMaBK);
            load_tile(nK + kMaTile, 2 * kC, (kb + 1) * kMaBK);
            asm volatile("cp.async.commit_group;" ::: "memory");
            asm volatile("cp.async.wait_group 1;" ::: "memory");
        } else {
            asm volatile("cp.async.wait_group 0;" ::: "memory");
        }
        const float m_next = kb + 1 < nkb ? fetch_mask((kb + 1) * kMaBK) : 0.f;      // in flight under this block's MMAs
        __syncthreads();
        if (kb == 0) {
#pragma unroll
            for (int ks = 0; ks < kD / 16; ++ks) {
                const int r = warp * 16 + (lane & 7) + ((lane >> 3) & 1) * 8, ch = 2 * ks + (lane >> 4);
                ldsm_x4(sQ + r * kMaPitch + ((ch ^ (r & 7)) << 4), qa[ks][0], qa[ks][1], qa[ks][2], qa[ks][3]);
            }
        }
        // S = Q K^T  (16 queries x 64 keys per warp)
        float s[8][4];
#pragma unroll
        for (int i = 0; i < 8; ++i) { s[i][0] = s[i][1] = s[i][2] = s[i][3] = 0.f; }
#pragma unroll
        for (int ks = 0; ks < kD / 16; ++ks) {
#pragma unroll
            for (int np = 0; np < 4; ++np) {
                const int r = np * 16 + (lane & 7) + (lane >> 4) * 8, ch = 2 * ks + ((lane >> 3) & 1);
                uint32_t b0, b1, b2, b3;
                ldsm_x4(sK + r * kMaPitch + ((ch ^ (r & 7)) << 4), b0, b1, b2, b3);
                mma16816_f16(s[2 * np], qa[ks], b0, b1);
                mma16816_f16(s[2 * np + 1], qa[ks], b2, b3);
            }
        }
        // logits = masked_fill(s / sqrt(96), mask_q * mask_k == 0, -1e4) (:213-220); keys beyond T do not exist
        float mx[2] = {-INFINITY, -INFINITY};
#pragma unroll
        for (int nt = 0; nt < 8; ++nt) {
            const int kl = nt * 8 + 2 * qd, key = kb * kMaBK + kl;
            const float mk0 = sMf[(kb & 1) * kMaBK + kl], mk1 = sMf[(kb & 1) * kMaBK + kl + 1];
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                const int ky = key + (e & 1);
                const float mkk = (e & 1) ? mk1 : mk0, mqq = mq[e >> 1];
                float v = s[nt][e] * kScale;
                if (mqq * mkk == 0.f) v = -1e4f;
                s[nt][e] = ky < T ? v : -INFINITY;
            }
            mx[0] = fmaxf(mx[0], fmaxf(s[nt][0], s[nt][1]));
            mx[1] = fmaxf(mx[1], fmaxf(s[nt][2], s[nt][3]));
        }
        float corr[2], msc[2];
#pragma unroll
        for (int i = 0; i < 2; ++i) {
            mx[i] = fmaxf(mx[i], __shfl_xor_sync(0xffffffffu, mx[i], 1));
            mx[i] = fmaxf(mx[i], __shfl_xor_sync(0xffffffffu, mx[i], 2));
            const float m_new = fmaxf(m_run[i], mx[i]);
            corr[i] = exp2f((m_run[i] - m_new) * kLog2e);
            m_run[i] = m_new;
            msc[i] = m_new * kLog2e;
        }
        float ps[2] = {0.f, 0.f};
        uint32_t pa[4][4];                                       // P as A fragments, 4 k-steps of 16 keys
#pragma unroll
        for (int nt = 0; nt < 8; ++nt) {
            const float p0 = exp2f(fmaf(s[nt][0], kLog2e, -msc[0])), p1 = exp2f(fmaf(s[nt][1], kLog2e, -msc[0]));
            const float p2 = exp2f(fmaf(s[nt][2], kLog2e, -msc[1])), p3 = exp2f(fmaf(s[nt][3], kLog2e, -msc[1]));
            ps[0] += p0 + p1; ps[1] += p2 + p3;
            pa[nt >> 1][(nt & 1) * 2 + 0] = pack_half2_sat(p0, p1);
            pa[nt >> 1][(nt & 1) * 2 + 1] = pack_half2_sat(p2, p3);
        }
#pragma unroll
        for (int i = 0; i < 2; ++i) {
            ps[i] += __shfl_xor_sync(0xffffffffu, ps[i], 1);
            ps[i] += __shfl_xor_sync(0xffffffffu, ps[i], 2);
            l_run[i] = l_run[i] * corr[i] + ps[i];
        }
#pragma unroll
        for (int dt = 0; dt < kD / 8; ++dt) { o[dt][0] *= corr[0]; o[dt][1] *= corr[0]; o[dt][2] *= corr[1]; o[dt][3] *= corr[1]; }
        // O += P V
#pragma unroll
        for (int j = 0; j < 4; ++j) {
#pragma unroll
            for (int dp = 0; dp < kD / 16; ++dp) {
                const int r = j * 16 + (lane & 7) + ((lane >> 3) & 1) * 8, ch = 2 * dp + (lane >> 4);
                uint32_t b0, b1, b2, b3;
                ldsm_x4_t(sV + r * kMaPitch + ((ch ^ (r & 7)) << 4), b0, b1, b2, b3);
                mma16816_f16(o[2 * dp], pa[j], b0, b1);
                mma16816_f16(o[2 * dp + 1], pa[j], b2, b3);
            }
        }
        if (tid < kMaBK) sMw[((kb + 1) & 1) * kMaBK + tid] = m_next;
        __syncthreads();                                         // the buffer is refilled two iterations later
    }
#pragma unroll
    for (int i = 0; i < 2; ++i) {
        const int t = q0 + warp * 16 + g + 8 * i;
        if (t < T) {
            const float inv = 1.0f / l_run[i];
            uint32_t *po = reinterpret_cast<uint32_t *>(out + ((size_t)b * T + t) * kC + h * kD + 2 * qd);
#pragma unroll
            for (int dt = 0; dt < kD / 8; ++dt) po[dt * 4] = pack_half2_sat(o[dt][2 * i] * inv, o[dt][2 * i + 1] * inv);
        }
    }
}

// ------------------------------------------------------------------ mean projection, channels-first output
// mu[b][n][t] = (sum_c x[b][t][c] * W[n][c] + bias[n]) * mask[b][t]      text_encoder.py:398
__global__ void __launch_bounds__(256)
te_mean_proj_kernel(const float *__restrict__ x, const float *__restrict__ w, const float *__restrict__ bias,
                    const float *__restrict__ mask, float *__restrict__ mu, int T) {
    __shared__ float A[32][kC + 1];
    te_pdl_entry();
    const int b = blockIdx.y, t0 = blockIdx.x * 32, tid = threadIdx.x;
    for (int i = tid; i < 32 * kC; i += 256) {
        const int tt = i / kC, c = i - tt * kC;
        A[tt][c] = (t0 + tt < T) ? ld_act(x + ((size_t)b * T + t0 + tt) * kC + c) : 0.f;
    }
    __syncthreads();
    const int lane = tid & 31, warp = tid >> 5;
    const int t = t0 + lane;
    const float mk = t < T ? ld_act(mask + (size_t)b * T + t) : 0.f;
    // ten outputs per warp, all in flight: one shared-memory read of the row feeds ten FMAs (a single dependent chain of 192 FMAs
    // per output made this kernel 57 us)
    constexpr int kPer = kFeats / 8;
    const int nb = warp * kPer;
    float acc[kPer];
#pragma unroll
    for (int j = 0; j < kPer; ++j) acc[j] = 0.f;
#pragma unroll 4
    for (int c = 0; c < kC; ++c) {
        const float a = A[lane][c];
#pragma unroll
        for (int j = 0; j < kPer; ++j) acc[j] = fmaf(a, __ldg(w + (size_t)(nb + j) * kC + c), acc[j]);
    }
    if (t < T) {
#pragma unroll
        for (int j = 0; j < kPer; ++j) mu[((size_t)b * kFeats + nb + j) * T + t] = (acc[j] + bias[nb + j]) * mk;
    }
}

// ------------------------------------------------------------------ tcgen05 implicit GEMM (fp16 mode)
// D[128 rows x NT columns] (fp32, TMEM) = sum over (tap, 64-channel block) of A[128 x 64] * W[NT x 64]^T, fp16 operands.
//   warp 0      TMA producer: A boxes (64 channels x 128 frames of one sample) from a 3-D map over the fp16 activation
//               [B][T][Cin] -- the conv taps are row offsets, frames outside [0, T) come back as zeros = the conv's zero
//               padding -- and W boxes (64 x NT) from a 2-D map over the fp16 weight [N][K], K = tap * Cin + channel; both
//               land in the 128B-swizzled K-major layout the UMMA shared-memory descriptor describes.  4-stage ring (3 when a
//               residual tile shares the shared memory).
//   warp 1      allocates 256 TMEM columns and issues tcgen05.mma (one lane; UMMA 128 x NT x 16, four per stage);
//               tcgen05.commit frees the stage / publishes the accumulator.
//   warps 2..9  epilogue, thread = output row = TMEM lane, two warps per lane quarter, each owning half of the tile's columns:
//               bias, ReLU, mask, residual; LayerNorm over the row's NT columns in two passes over TMEM (pass 1 forms the
//               pre-norm values, writes them back with tcgen05.st and sums -- the two halves' sums cross through shared
//               memory; pass 2 normalises); RoPE for the q / k tiles of the fused projection (one head per warp); fp32 and / or
//               fp16 stores through a swizzled shared-memory tile (the fp16 copy optionally x mask: the next conv's operand);
//               the duration predictor's 256 -> 1 projection as a per-row dot product.
// One 128-row tile per CTA (grid = B * ceil(T / 128) x N / NT): the encoder's GEMMs have 32-100 tiles, fewer than SMs.
constexpr int kTcMaxStages = 4, kTcM = 128, kTcK = 64, kTcStageA = kTcM * kTcK * 2;
constexpr int kTcThreads = 320;                    // TMA warp, MMA warp, 8 epilogue warps (two per TMEM lane quarter)
constexpr int kTcParamBytes = 4 * 256 * 4;          // bias | ln_g | ln_b | dp_w of the tile's columns, staged once per CTA
// store staging: per epilogue warp a [32 rows][128 B] tile that turns thread = row into row-contiguous stores; it lives in the
// first 16 KB of the TMA ring, which is free once the accumulator is complete (every MMA has read its stage by then)
template <int NT> struct TcSizes {
    static constexpr int kStage = kTcStageA + NT * kTcK * 2;                    // 40 KB (NT = 192) / 48 KB (NT = 256)
    static constexpr int kResPitch = NT * 4 + 16;                               // residual row + 16 B: thread = row reads are conflict-free
    static constexpr int kResBytes = kTcM * kResPitch;                          // 98 KB (NT = 192)
};
// ring | parameters | barriers | [residual tile]; 1 KB slack for the 1024-byte alignment of the ring
template <int NT> constexpr int tc_smem_bytes(int stages, bool res) {
    return stages * TcSizes<NT>::kStage + kTcParamBytes + 256 + (res ? TcSizes<NT>::kResBytes : 0) + 1024;
}

struct TcArgs {
    int B, T, tiles_per_sample;
    int stages;                       // TMA ring depth (3 with a residual tile in shared memory, else 4)
    const float *rope_cos, *rope_sin; // q|k|v projection only: rotate the first 48 features of each q / k head (:135-148)
    int ntaps, pad, chunks;           // chunks = Cin / 64
    int N;
    const float *bias, *mask;
    int relu_pre, mask_pre;
    const float *res;                 // fp32 [B*T][N] or nullptr
    const float *ln_g, *ln_b;         // LayerNorm over the NT columns (N == NT), or nullptr
    int relu_post, mask_post;
    float *out_f32;                   // fp32 [B*T][N] or nullptr
    __half *out16; int mask16;        // fp16 [B*T][N] or nullptr, multiplied by mask[row] when mask16
    const float *dp_w, *dp_b; float *dp_out;
    long long *stamps;                // profiling aid: globaltimer stamps of CTA 0 (nullptr in the product path)
};

__device__ __forceinline__ void tma_load_3d(uint32_t dst, const CUtensorMap *map, uint32_t bar, int c0, int c1, int c2) {
    asm volatile(
        "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
        ::"r"(dst), "l"((uint64_t)map), "r"(bar), "r"(c0), "r"(c1), "r"(c2) : "memory");
}
// K-major, 128B swizzle, 8-row atoms 1024 B apart (SBO), descriptor version 1 (sm_100)
__device__ __forceinline__ uint64_t umma_desc_sw128(uint32_t smem_addr) {
    uint64_t d = 0;
    d |= (uint64_t)((smem_addr >> 4) & 0x3fff);
    d |= (uint64_t)(1024 >> 4) << 32;
    d |= (uint64_t)1 << 46;
    d |= (uint64_t)2 << 61;
    return d;
}
__device__ __forceinline__ void umma_f16(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void umma_commit(uint32_t bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&r)[32]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
          "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
          "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
          "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(taddr));
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_st32(uint32_t taddr, const uint32_t (&r)[32]) {
    asm volatile(
        "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "
        "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, "
        "%17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, %32};"
        ::"r"(taddr), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]),
          "r"(r[8]), "r"(r[9]), "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]),
          "r"(r[16]), "r"(r[17]), "r"(r[18]), "r"(r[19]), "r"(r[20]), "r"(r[21]), "r"(r[22]), "r"(r[23]),
          "r"(r[24]), "r"(r[25]), "r"(r[26]), "r"(r[27]), "r"(r[28]), "r"(r[29]), "r"(r[30]), "r"(r[31]) : "memory");
}

template <int NT>
__global__ void __launch_bounds__(kTcThreads, 1)
te_gemm_tc_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, const TcArgs g) {
    constexpr int kStage = TcSizes<NT>::kStage;
    // kind::f16, fp16 A / B (format 0) K-major, fp32 accumulate, M = 128, N = NT
    constexpr uint32_t kIdesc = (1u << 4) | ((uint32_t)(NT >> 3) << 17) | ((uint32_t)(kTcM >> 4) << 24);
    extern __shared__ uint8_t tc_smem_raw[];
    const uint32_t base = (smem_u32(tc_smem_raw) + 1023u) & ~1023u;
    uint8_t *gbase = tc_smem_raw + (base - smem_u32(tc_smem_raw));
    const uint32_t n_st = (uint32_t)g.stages;
    const uint32_t par_off = n_st * kStage;
    float *const spar = reinterpret_cast<float *>(gbase + par_off);          // [4][256]: bias, ln_g, ln_b, dp_w
    const uint32_t stg_off = 0;                                              // [4 warps][4 KB] store staging = head of the ring
    const uint32_t bar_full = base + par_off + kTcParamBytes;     // [stages] TMA -> MMA
    const uint32_t bar_empty = bar_full + 8u * kTcMaxStages;      // [stages] MMA -> TMA
    const uint32_t bar_tfull = bar_empty + 8u * kTcMaxStages;     // accumulator complete
    const uint32_t tmem_slot = bar_tfull + 8u;
    volatile uint32_t *tmem_slot_p = reinterpret_cast<volatile uint32_t *>(gbase + (tmem_slot - base));
    const uint32_t res_off = par_off + kTcParamBytes + 256;                   // residual tile, only when g.res

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int b = blockIdx.x / g.tiles_per_sample, t0 = (blockIdx.x - b * g.tiles_per_sample) * kTcM;
    const int n0 = blockIdx.y * NT;
    const int n_kb = g.ntaps * g.chunks;

    if (threadIdx.x == 0) {
        for (uint32_t s = 0; s < n_st; ++s) { mbar_init(bar_full + 8u * s, 1u); mbar_init(bar_empty + 8u * s, 1u); }
        mbar_init(bar_tfull, 1u);
        mbar_fence_init();
        tma_prefetch_desc(&tmA); tma_prefetch_desc(&tmB);
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "n"(256) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem_d = *tmem_slot_p;
    const bool stamp = g.stamps && blockIdx.x == 0 && blockIdx.y == 0;
    if (stamp && threadIdx.x == 0) g.stamps[0] = gtime();
    if (warp >= 2) {             // the tile's per-column parameters (weights: no dependency on the predecessor kernel)
        for (int i = threadIdx.x - 64; i < NT; i += 256) {
            spar[i] = g.bias[n0 + i];
            spar[256 + i] = g.ln_g ? g.ln_g[n0 + i] : 1.f;
            spar[512 + i] = g.ln_g ? g.ln_b[n0 + i] : 0.f;
            spar[768 + i] = g.dp_w ? g.dp_w[n0 + i] : 0.f;
        }
    }
    te_pdl_entry();              // barriers, TMEM, descriptors and parameters were set up while the previous kernel drained
    if (stamp && threadIdx.x == 0) g.stamps[1] = gtime();

    if (warp == 0) {
        // ============================ TMA producer ============================
        if (lane == 0) {
            for (int kb = 0; kb < n_kb; ++kb) {
                const uint32_t s = (uint32_t)kb % n_st;
                mbar_wait_parked(bar_empty + 8u * s, (((uint32_t)kb / n_st) & 1u) ^ 1u, 500);
                const int tap = kb / g.chunks, ch = kb - tap * g.chunks;
                const uint32_t sa = base + s * kStage, sb = sa + kTcStageA;
                mbar_arrive_expect_tx(bar_full + 8u * s, kStage);
                tma_load_3d(sa, &tmA, bar_full + 8u * s, ch * kTcK, t0 + tap - g.pad, b);
                tma_load_2d(sb, &tmB, bar_full + 8u * s, kb * kTcK, n0);
            }
        }
    } else if (warp == 1) {
        // ============================ MMA issuer ============================
        if (lane == 0) {
            for (int kb = 0; kb < n_kb; ++kb) {
                const uint32_t s = (uint32_t)kb % n_st;
                mbar_wait_parked(bar_full + 8u * s, ((uint32_t)kb / n_st) & 1u, 500);
                if (stamp && kb == 0) g.stamps[2] = gtime();
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                const uint32_t sa = base + s * kStage, sb = sa + kTcStageA;
                const uint64_t da = umma_desc_sw128(sa), db = umma_desc_sw128(sb);
#pragma unroll
                for (int k = 0; k < kTcK / 16; ++k)      // 32 bytes along K per UMMA_K = 16 step: +2 in 16-byte units
                    umma_f16(tmem_d, da + 2u * k, db + 2u * k, kIdesc, (kb | k) != 0 ? 1u : 0u);
                umma_commit(bar_empty + 8u * s);
            }
            umma_commit(bar_tfull);
        }
    } else {
        // ============================ epilogue: thread = row ============================
        const int q = warp & 3;                          // TMEM lane quarter this warp may touch
        const int half = (warp - 2) >> 2;                // ... and the half of the tile's columns it owns (chunks [jlo, jhi))
        constexpr int NJ = NT / 32;
        const int jlo = half * (NJ / 2), jhi = jlo + NJ / 2;
        const int t = t0 + 32 * q + lane;
        const bool valid = t < g.T;
        const size_t m = (size_t)b * g.T + (valid ? t : 0);
        const float mk = valid ? ld_act(g.mask + m) : 0.f;
        const uint32_t tacc = tmem_d + ((uint32_t)(32 * q) << 16);
        const int trow = 32 * q + lane;                  // row of the tile
        // the residual rows of this warp's 32 rows travel to shared memory (cp.async, L2 only) under the K loop; consecutive
        // lanes copy consecutive 16-byte pieces of a row, so every instruction touches whole 128-byte lines
        const uint32_t res_row = base + res_off + (uint32_t)trow * TcSizes<NT>::kResPitch;
        if (g.res) {
            constexpr int kPieces = NT / 8;                      // 16-byte pieces of this warp's half of a row
            for (int idx = lane; idx < 32 * kPieces; idx += 32) {
                const int rr = idx / kPieces, pc = idx - rr * kPieces + half * kPieces;
                const int tt = t0 + 32 * q + rr;
                if (tt < g.T) {
                    const float *rp = g.res + ((size_t)b * g.T + tt) * g.N + n0 + 4 * pc;
                    const uint32_t dst = base + res_off + (uint32_t)(32 * q + rr) * TcSizes<NT>::kResPitch + 16u * pc;
                    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(rp) : "memory");
                }
            }
            asm volatile("cp.async.commit_group;" ::: "memory");
        }
        asm volatile("bar.sync 1, 256;" ::: "memory");     // parameters staged by all eight epilogue warps
        mbar_wait_parked(bar_tfull, 0u, 1000);
        if (stamp && threadIdx.x == 64) g.stamps[3] = gtime();
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        if (g.res) {                                          // rows were copied by other lanes of this warp
            asm volatile("cp.async.wait_group 0;" ::: "memory");
            __syncwarp();
        }
        const bool ln = g.ln_g != nullptr;
        // pre-norm value of 32 columns: bias, ReLU, mask, residual
        auto pre = [&](int j, uint32_t (&v)[32]) {
            const float4 *bp = reinterpret_cast<const float4 *>(spar + 32 * j);      // broadcast reads
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const float4 b4 = bp[i];
                const float bb[4] = {b4.x, b4.y, b4.z, b4.w};
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                    float x = __uint_as_float(v[4 * i + e]) + bb[e];
                    if (g.relu_pre) x = fmaxf(x, 0.f);
                    if (g.mask_pre) x *= mk;
                    v[4 * i + e] = __float_as_uint(x);
                }
            }
            if (g.res && valid) {
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    const float4 r4 = lds128(res_row + 128u * j + 16u * i);
                    v[4 * i] = __float_as_uint(__uint_as_float(v[4 * i]) + r4.x);
                    v[4 * i + 1] = __float_as_uint(__uint_as_float(v[4 * i + 1]) + r4.y);
                    v[4 * i + 2] = __float_as_uint(__uint_as_float(v[4 * i + 2]) + r4.z);
                    v[4 * i + 3] = __float_as_uint(__uint_as_float(v[4 * i + 3]) + r4.w);
                }
            }
        };
        // fp32 / fp16 stores of 32 finished columns of the warp's 32 rows.  A thread holds one ROW; written straight out, each store
        // instruction would touch 32 different rows (32 sectors per instruction: measured 1.3 us per 32-column chunk).  The warp's
        // 32 x 32 block goes through a 4 KB shared tile (16-byte pieces swizzled by row & 7: conflict-free both ways) and leaves
        // with lane = (row % 4, 16-byte piece): every instruction writes four whole 128-byte row segments.
        const uint32_t stg = base + stg_off + (uint32_t)(warp - 2) * 4096u;
        const int rsub = lane >> 3, pc = lane & 7;
        const float m16 = g.mask16 ? mk : 1.0f;
        auto store = [&](int j, const float (&x)[32]) {
            const uint32_t srow = stg + (uint32_t)lane * 128u;
#pragma unroll
            for (int i = 0; i < 8; ++i)
                sts128(srow + (((uint32_t)i ^ (uint32_t)(lane & 7)) << 4), make_float4(x[4 * i], x[4 * i + 1], x[4 * i + 2], x[4 * i + 3]));
            __syncwarp();
#pragma unroll
            for (int it = 0; it < 8; ++it) {
                const int rr = 4 * it + rsub;                     // row of the warp's block
                const float4 v4 = lds128(stg + (uint32_t)rr * 128u + (((uint32_t)pc ^ (uint32_t)(rr & 7)) << 4));
                const float ms = __shfl_sync(0xffffffffu, m16, rr);
                const int tt = t0 + 32 * q + rr;
                if (tt < g.T) {
                    const size_t o = ((size_t)b * g.T + tt) * g.N + n0 + 32 * j + 4 * pc;
                    if (g.out_f32) *reinterpret_cast<float4 *>(g.out_f32 + o) = v4;
                    if (g.out16)
                        *reinterpret_cast<uint2 *>(g.out16 + o) = make_uint2(pack_half2_sat(v4.x * ms, v4.y * ms), pack_half2_sat(v4.z * ms, v4.w * ms));
                }
            }
            __syncwarp();
        };
        // exchange area of the two warps that share a row (LayerNorm partial sums, duration-projection partial dot products): in the
        // ring too, behind the eight staging tiles
        float2 *const xch = reinterpret_cast<float2 *>(gbase + stg_off + 8 * 4096);
        if (g.rope_cos && n0 < 2 * kC) {
            // q (tile 0) or k (tile 1) of the fused projection: head hh = columns [96 hh, 96 hh + 96) of the tile, of which the
            // first 48 are rotated: x'[d] = x[d] cos[d] - x[d+24] sin[d], x'[d+24] = x[d+24] cos[d+24] + x[d] sin[d+24], d < 24
            // (RotaryPositionalEmbeddings.forward, :135-148, with the reference's two roundings per product pair).
            const float *cs = g.rope_cos + (size_t)(valid ? t : 0) * kRope, *sn = g.rope_sin + (size_t)(valid ? t : 0) * kRope;
            {
                const int hh = half;                            // one head per warp of the pair
                uint32_t v0[32], v1[32];
                tmem_ld32(tacc + 32 * (3 * hh), v0);
                tmem_ld32(tacc + 32 * (3 * hh + 1), v1);
                pre(3 * hh, v0);
                pre(3 * hh + 1, v1);
                float x0[32], x1[32];
#pragma unroll
                for (int i = 0; i < 32; ++i) { x0[i] = __uint_as_float(v0[i]); x1[i] = __uint_as_float(v1[i]); }
#pragma unroll
                for (int d = 0; d < kRope / 2; ++d) {
                    const float c0 = __ldg(cs + d), s0 = __ldg(sn + d), c1 = __ldg(cs + d + kRope / 2), s1 = __ldg(sn + d + kRope / 2);
                    const float a = x0[d];
                    const float bq = (d < 8) ? x0[d + 24] : x1[d - 8];
                    const float ra = rope_mix(a, c0, -bq, s0), rb = rope_mix(bq, c1, a, s1);
                    x0[d] = ra;
                    if (d < 8) x0[d + 24] = rb; else x1[d - 8] = rb;
                }
                store(3 * hh, x0);
                store(3 * hh + 1, x1);
                tmem_ld32(tacc + 32 * (3 * hh + 2), v0);
                pre(3 * hh + 2, v0);
#pragma unroll
                for (int i = 0; i < 32; ++i) x0[i] = __uint_as_float(v0[i]);
                store(3 * hh + 2, x0);
            }
        } else {
            float mean = 0.f, rstd = 1.f;
            if (ln) {
                float s1 = 0.f, s2 = 0.f;
                for (int j = jlo; j < jhi; ++j) {
                    uint32_t v[32];
                    tmem_ld32(tacc + 32 * j, v);
                    pre(j, v);
#pragma unroll
                    for (int i = 0; i < 32; ++i) { const float x = __uint_as_float(v[i]); s1 += x; s2 = fmaf(x, x, s2); }
                    tmem_st32(tacc + 32 * j, v);
                }
                asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
                xch[half * kTcM + trow] = make_float2(s1, s2);              // the other half's sums come from the partner warp
                asm volatile("bar.sync 2, 256;" ::: "memory");
                const float2 oth = xch[(half ^ 1) * kTcM + trow];
                s1 += oth.x; s2 += oth.y;
                mean = s1 * (1.0f / NT);
                const float var = fmaxf(s2 * (1.0f / NT) - mean * mean, 0.f);
                rstd = 1.0f / sqrtf(var + 1e-5f);
            }
            float dp = 0.f;
            for (int j = jlo; j < jhi; ++j) {
                uint32_t v[32];
                tmem_ld32(tacc + 32 * j, v);
                if (!ln) pre(j, v);
                float x[32];
#pragma unroll
                for (int i = 0; i < 32; ++i) {
                    float y = __uint_as_float(v[i]);
                    if (ln) y = (y - mean) * rstd * spar[256 + 32 * j + i] + spar[512 + 32 * j + i];
                    if (g.relu_post) y = fmaxf(y, 0.f);
                    if (g.mask_post) y *= mk;
                    x[i] = y;
                }
                store(j, x);
                if (g.dp_w) {
#pragma unroll
                    for (int i = 0; i < 32; ++i) dp = fmaf(x[i] * mk, spar[768 + 32 * j + i], dp);
                }
            }
            if (g.dp_w) {                                        // the row's dot product = the two halves' partial ones
                float *const dx = reinterpret_cast<float *>(xch + 2 * kTcM);
                if (half == 1) dx[trow] = dp;
                asm volatile("bar.sync 2, 256;" ::: "memory");
                if (half == 0 && valid) g.dp_out[m] = (dp + dx[trow] + __ldg(g.dp_b)) * mk;
            }
        }
    }
    if (stamp && threadIdx.x == 64) g.stamps[4] = gtime();
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (stamp && threadIdx.x == 0) g.stamps[5] = gtime();
    if (warp == 1) {
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_d), "n"(256) : "memory");
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

struct EncWs {
    float *x0, *x, *y, *ha, *hb, *qkv, *att, *f1, *dp1;                       // fp32 (fp16 mode: x0, x, y, qkv only)
    __half *x0h, *xh, *yh, *hah, *hbh, *atth, *f1h, *dp1h, *qkvh;             // fp16 GEMM / attention operands (fp16 mode)
};
static long long carve_workspace(float *base, int B, int T, int precision, EncWs *ws) {
    const long long M = (long long)B * T;
    long long off = 0;
    auto take = [&](long long n) { float *p = base ? base + off : nullptr; off += (n + 63) & ~63LL; return p; };
    auto take_h = [&](long long n) { return reinterpret_cast<__half *>(take((n + 1) / 2)); };
    EncWs w = {};
    w.x0 = take(M * kC); w.x = take(M * kC); w.y = take(M * kC);
    if (precision == MB200_PREC_F32) {
        w.qkv = take(M * 3 * kC);
        w.ha = take(M * kC); w.hb = take(M * kC); w.att = take(M * kC); w.f1 = take(M * kF); w.dp1 = take(M * kDp);
    } else {
        w.x0h = take_h(M * kC); w.xh = take_h(M * kC); w.yh = take_h(M * kC); w.hah = take_h(M * kC); w.hbh = take_h(M * kC);
        w.atth = take_h(M * kC); w.f1h = take_h(M * kF); w.dp1h = take_h(M * kDp); w.qkvh = take_h(M * 3 * kC);
    }
    if (ws) *ws = w;
    return off * (long long)sizeof(float);
}

struct TmapEntry { const void *base; uint64_t d0, d1; uint32_t b1; CUtensorMap map; };
struct Ctx {
    cudaStream_t st; int B, T; const float *mask; int launches = 0, err = 0;
    std::vector<TmapEntry> tmaps;        // tensor maps of this call's fp16 tensors
};

static bool pdl_enabled() {
    static int v = -1;
    if (v < 0) { const char *e = getenv("MB200_PDL"); v = (e && e[0] == '0') ? 0 : 1; }
    return v != 0;
}
template <typename... KArgs, typename... Args>
static void launch_smem(Ctx &c, void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, Args &&...args) {
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
template <typename... KArgs, typename... Args>
static void launch(Ctx &c, void (*kern)(KArgs...), dim3 grid, dim3 block, Args &&...args) {
    launch_smem(c, kern, grid, block, 0, static_cast<Args &&>(args)...);
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

static int set_attributes();
static void forward(Ctx &c, const EncW &W, const EncWs &ws, const long long *ids, const int *lengths, int n_vocab,
                    const float *rope_cos, const float *rope_sin, float *mu, float *logw, float *mask) {
    const int M = c.B * c.T;
    c.err = set_attributes();
    launch(c, te_embed_kernel, dim3((M + 7) / 8), dim3(256), ids, lengths, W.emb, n_vocab, ws.x0, (__half *)nullptr, mask, c.B, c.T);
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
        launch_smem(c, te_attention_kernel, dim3((c.T + kAQ - 1) / kAQ, kH, c.B), dim3(128), (size_t)kAttSmemBytes, ws.qkv, c.mask, rope_cos, rope_sin, 0,
                    ws.att, (__half *)nullptr, c.T, (long long *)nullptr);
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

// ---- fp16 mode: the same network on te_gemm_tc_kernel (weights: the fp32 blob converted to fp16 element for element, so
// every tensor sits at the same element offset; activations: fp16 operand copies written by their producers)
static bool tmap_for(Ctx &c, const void *base, int ndim, uint64_t d0, uint64_t d1, uint64_t d2, uint32_t b0, uint32_t b1,
                     CUtensorMap *out) {
    for (auto &e : c.tmaps)
        if (e.base == base && e.d0 == d0 && e.d1 == d1 && e.b1 == b1) { *out = e.map; return true; }
    TmapEntry e;
    e.base = base; e.d0 = d0; e.d1 = d1; e.b1 = b1;
    const uint64_t dims[3] = {d0, d1, d2};
    const uint64_t strides[2] = {d0 * 2, d0 * d1 * 2};
    const uint32_t box[3] = {b0, b1, 1};
    const int rc = encode_tmap_bf16(&e.map, base, ndim, dims, strides, box);     // 2-byte elements; fp16 travels as such
    if (rc != MB200_OK) { c.err = rc; return false; }
    c.tmaps.push_back(e);
    *out = e.map;
    return true;
}
static int set_attributes() {
    static bool done[64] = {};
    int dev = -1;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return MB200_ERR_NO_DEVICE;
    if (done[dev]) return MB200_OK;
    MB200_CUDA_TRY(cudaFuncSetAttribute(te_attention_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kAttSmemBytes));
    MB200_CUDA_TRY(cudaFuncSetAttribute(te_attention_mma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kMaSmemBytes));
    MB200_CUDA_TRY(cudaFuncSetAttribute(te_gemm_tc_kernel<192>, cudaFuncAttributeMaxDynamicSharedMemorySize, tc_smem_bytes<192>(3, true)));
    MB200_CUDA_TRY(cudaFuncSetAttribute(te_gemm_tc_kernel<256>, cudaFuncAttributeMaxDynamicSharedMemorySize, tc_smem_bytes<256>(4, false)));
    done[dev] = true;
    return MB200_OK;
}
static_assert(tc_smem_bytes<192>(3, true) <= kMaxSmemPerBlock && tc_smem_bytes<192>(4, false) <= kMaxSmemPerBlock &&
              tc_smem_bytes<256>(4, false) <= kMaxSmemPerBlock, "tensor-core GEMM does not fit shared memory");
struct Epi16 {
    Epi e; float *out_f32 = nullptr; __half *out16 = nullptr; int mask16 = 0; const float *rope_cos = nullptr, *rope_sin = nullptr;
    long long *stamps = nullptr;
};
static void gemm_tc(Ctx &c, const __half *a16, int Cin, int ntaps, const __half *w16, const float *bias, int N, const Epi16 &o) {
    if (c.err) return;
    const int K = ntaps * Cin;
    const int NT = (N % 192 == 0) ? 192 : 256;
    if (o.e.res && NT != 192) { c.err = MB200_ERR_UNSUPPORTED; return; }        // the residual tile only fits beside 192-column stages
    CUtensorMap tmA, tmB;
    if (!tmap_for(c, a16, 3, (uint64_t)Cin, (uint64_t)c.T, (uint64_t)c.B, kTcK, kTcM, &tmA)) return;
    if (!tmap_for(c, w16, 2, (uint64_t)K, (uint64_t)N, 1, kTcK, (uint32_t)NT, &tmB)) return;
    TcArgs g;
    g.B = c.B; g.T = c.T; g.tiles_per_sample = (c.T + kTcM - 1) / kTcM;
    g.stages = o.e.res ? 3 : 4; g.rope_cos = o.rope_cos; g.rope_sin = o.rope_sin; g.stamps = o.stamps;
    g.ntaps = ntaps; g.pad = ntaps / 2; g.chunks = Cin / kTcK; g.N = N;
    g.bias = bias; g.mask = c.mask; g.relu_pre = o.e.relu_pre; g.mask_pre = o.e.mask_pre; g.res = o.e.res;
    g.ln_g = o.e.ln_g; g.ln_b = o.e.ln_b; g.relu_post = o.e.relu_post; g.mask_post = o.e.mask_post;
    g.out_f32 = o.out_f32; g.out16 = o.out16; g.mask16 = o.mask16; g.dp_w = o.e.dp_w; g.dp_b = o.e.dp_b; g.dp_out = o.e.dp_out;
    const dim3 grid(c.B * g.tiles_per_sample, N / NT);
    if (NT == 192) launch_smem(c, te_gemm_tc_kernel<192>, grid, dim3(kTcThreads), (size_t)tc_smem_bytes<192>(g.stages, o.e.res != nullptr), tmA, tmB, g);
    else launch_smem(c, te_gemm_tc_kernel<256>, grid, dim3(kTcThreads), (size_t)tc_smem_bytes<256>(g.stages, false), tmA, tmB, g);
}

static void forward_f16(Ctx &c, const EncW &W, const float *blob, const __half *blob16, const EncWs &ws, const long long *ids,
                        const int *lengths, int n_vocab, const float *rope_cos, const float *rope_sin, float *mu, float *logw,
                        float *mask) {
    const int M = c.B * c.T;
    auto h = [&](const float *p) { return blob16 + (p - blob); };            // the fp16 twin of a weight tensor
    c.err = set_attributes();
    launch(c, te_embed_kernel, dim3((M + 7) / 8), dim3(256), ids, lengths, W.emb, n_vocab, ws.x0, ws.x0h, mask, c.B, c.T);
    // prenet (:48-57): conv(x*mask) -> LayerNorm -> ReLU three times, then (residual + projection) * mask
    const __half *src = ws.x0h;
    __half *pp[2] = {ws.hah, ws.hbh};
    for (int i = 0; i < kPreL; ++i) {
        Epi16 o; o.e.ln_g = W.pg[i]; o.e.ln_b = W.pbe[i]; o.e.relu_post = 1; o.out16 = pp[i & 1]; o.mask16 = i + 1 < kPreL;
        gemm_tc(c, src, kC, kPreK, h(W.pw[i]), W.pb[i], kC, o);
        src = pp[i & 1];
    }
    { Epi16 o; o.e.res = ws.x0; o.e.mask_post = 1; o.out_f32 = ws.x; o.out16 = ws.xh; gemm_tc(c, src, kC, 1, h(W.ppw), W.ppb, kC, o); }
    int n_layers = kL;
    if (const char *e = getenv("MB200_TE_LAYERS")) n_layers = atoi(e) < kL ? atoi(e) : kL;     // debugging aid
    for (int i = 0; i < n_layers; ++i) {
        const LayerW &l = W.layer[i];
        { Epi16 o; o.out16 = ws.qkvh; o.rope_cos = rope_cos; o.rope_sin = rope_sin; gemm_tc(c, ws.xh, kC, 1, h(l.wqkv), l.bqkv, 3 * kC, o); }
        launch_smem(c, te_attention_mma_kernel, dim3((c.T + kMaBQ - 1) / kMaBQ, kH, c.B), dim3(128), (size_t)kMaSmemBytes, ws.qkvh, c.mask, ws.atth, c.T);
        { Epi16 o; o.e.res = ws.x; o.e.ln_g = l.ln1g; o.e.ln_b = l.ln1b; o.out_f32 = ws.y; o.out16 = ws.yh; o.mask16 = 1;
          gemm_tc(c, ws.atth, kC, 1, h(l.wo), l.bo, kC, o); }
        { Epi16 o; o.e.relu_pre = 1; o.out16 = ws.f1h; gemm_tc(c, ws.yh, kC, kK, h(l.w1), l.b1, kF, o); }
        { Epi16 o; o.e.mask_pre = 1; o.e.res = ws.y; o.e.ln_g = l.ln2g; o.e.ln_b = l.ln2b; o.e.mask_post = 1; o.out_f32 = ws.x; o.out16 = ws.xh;
          gemm_tc(c, ws.f1h, kF, kK, h(l.w2), l.b2, kC, o); }
    }
    launch(c, te_mean_proj_kernel, dim3((c.T + 31) / 32, c.B), dim3(256), ws.x, W.mw, W.mb, c.mask, mu, c.T);
    { Epi16 o; o.e.relu_pre = 1; o.e.ln_g = W.d1g; o.e.ln_b = W.d1be; o.out16 = ws.dp1h; o.mask16 = 1;
      gemm_tc(c, ws.xh, kC, kK, h(W.d1w), W.d1b, kDp, o); }
    { Epi16 o; o.e.relu_pre = 1; o.e.ln_g = W.d2g; o.e.ln_b = W.d2be; o.e.dp_w = W.dpw; o.e.dp_b = W.dpb; o.e.dp_out = logw;
      gemm_tc(c, ws.dp1h, kDp, kK, h(W.d2w), W.d2b, kDp, o); }
}

}  // namespace te
}  // namespace mb200

using namespace mb200;

extern "C" long long matcha_b200_text_encoder_blob_floats(int n_vocab) {
    if (n_vocab <= 0) return MB200_ERR_INVALID_ARG;
    return te::build_weights(nullptr, n_vocab, nullptr);
}
extern "C" long long matcha_b200_text_encoder_workspace_bytes(int B, int T, int precision) {
    if (B <= 0 || T <= 0 || (precision != MB200_PREC_F32 && precision != MB200_PREC_F16)) return MB200_ERR_INVALID_ARG;
    return te::carve_workspace(nullptr, B, T, precision, nullptr);
}
extern "C" int matcha_b200_text_encoder_forward(const long long *ids_dev, const int *lengths_dev, const float *blob_dev,
                                                const void *blob_f16_dev, int n_vocab, const float *rope_cos_dev,
                                                const float *rope_sin_dev, float *mu_dev, float *logw_dev, float *mask_dev,
                                                int B, int T, int precision, void *workspace_dev, long long workspace_bytes,
                                                void *stream) {
    if (B <= 0 || T <= 0 || n_vocab <= 0) return MB200_ERR_INVALID_ARG;
    if (precision != MB200_PREC_F32 && precision != MB200_PREC_F16) return MB200_ERR_INVALID_ARG;
    if (precision == MB200_PREC_F16 && (!blob_f16_dev || (uintptr_t)blob_f16_dev % 16 != 0)) return MB200_ERR_INVALID_ARG;
    if (!ids_dev || !lengths_dev || !blob_dev || !rope_cos_dev || !rope_sin_dev || !mu_dev || !logw_dev || !mask_dev || !workspace_dev)
        return MB200_ERR_INVALID_ARG;
    if ((long long)B * T > (1LL << 30) / te::kF) return MB200_ERR_UNSUPPORTED;          // int row arithmetic in the kernels
    if (workspace_bytes < te::carve_workspace(nullptr, B, T, precision, nullptr)) return MB200_ERR_INVALID_ARG;
    if ((uintptr_t)workspace_dev % 16 != 0 || (uintptr_t)blob_dev % 16 != 0 || (uintptr_t)rope_cos_dev % 16 != 0 ||
        (uintptr_t)rope_sin_dev % 16 != 0)
        return MB200_ERR_INVALID_ARG;
    te::EncW W;
    te::build_weights(blob_dev, n_vocab, &W);
    te::EncWs ws;
    te::carve_workspace((float *)workspace_dev, B, T, precision, &ws);
    te::Ctx c;
    c.st = (cudaStream_t)stream; c.B = B; c.T = T; c.mask = mask_dev;
    if (precision == MB200_PREC_F32)
        te::forward(c, W, ws, ids_dev, lengths_dev, n_vocab, rope_cos_dev, rope_sin_dev, mu_dev, logw_dev, mask_dev);
    else
        te::forward_f16(c, W, blob_dev, reinterpret_cast<const __half *>(blob_f16_dev), ws, ids_dev, lengths_dev, n_vocab,
                        rope_cos_dev, rope_sin_dev, mu_dev, logw_dev, mask_dev);
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
extern "C" int matcha_b200_debug_te_attention_stamps(const float *qkv, const float *mask, const float *rope_cos, const float *rope_sin,
                                                     float *out, int B, int T, long long *stamps_dev, void *stream);
extern "C" int matcha_b200_debug_te_attention(const float *qkv, const float *mask, const float *rope_cos, const float *rope_sin,
                                              float *out, int B, int T, void *stream) {
    return matcha_b200_debug_te_attention_stamps(qkv, mask, rope_cos, rope_sin, out, B, T, nullptr, stream);
}
extern "C" int matcha_b200_debug_te_attention_stamps(const float *qkv, const float *mask, const float *rope_cos, const float *rope_sin,
                                                     float *out, int B, int T, long long *stamps_dev, void *stream) {
    if (!qkv || !mask || !rope_cos || !rope_sin || !out || B <= 0 || T <= 0) return MB200_ERR_INVALID_ARG;
    te::Ctx c;
    c.st = (cudaStream_t)stream; c.B = B; c.T = T; c.mask = mask;
    c.err = te::set_attributes();
    te::launch_smem(c, te::te_attention_kernel, dim3((T + te::kAQ - 1) / te::kAQ, te::kH, B), dim3(128), (size_t)te::kAttSmemBytes, qkv, mask, rope_cos,
                    rope_sin, 0, out, (__half *)nullptr, T, stamps_dev);
    return c.err ? c.err : c.launches;
}

// fp16 tensor-core GEMM on its own: a16 fp16 [B][T][Cin], w16 fp16 [N][ntaps*Cin]; outputs as in te_gemm_tc_kernel
extern "C" int matcha_b200_debug_te_gemm_tc(const void *a16, int Cin, const float *mask, int ntaps, int B, int T, const void *w16,
                                            const float *bias, int N, int relu_pre, int mask_pre, const float *res, const float *ln_g,
                                            const float *ln_b, int relu_post, int mask_post, float *out_f32, void *out16, int mask16,
                                            const float *dp_w, const float *dp_b, float *dp_out, const float *rope_cos,
                                            const float *rope_sin, long long *stamps_dev, void *stream) {
    if (!a16 || !mask || !w16 || !bias || B <= 0 || T <= 0 || (N % 192 != 0 && N % 256 != 0) || Cin % 64 != 0) return MB200_ERR_INVALID_ARG;
    if (ln_g && N != 192 && N != 256) return MB200_ERR_INVALID_ARG;
    te::Ctx c;
    c.st = (cudaStream_t)stream; c.B = B; c.T = T; c.mask = mask;
    c.err = te::set_attributes();
    te::Epi16 o;
    o.e.relu_pre = relu_pre; o.e.mask_pre = mask_pre; o.e.res = res; o.e.ln_g = ln_g; o.e.ln_b = ln_b; o.e.relu_post = relu_post;
    o.e.mask_post = mask_post; o.e.dp_w = dp_w; o.e.dp_b = dp_b; o.e.dp_out = dp_out;
    o.out_f32 = out_f32; o.out16 = reinterpret_cast<__half *>(out16); o.mask16 = mask16; o.rope_cos = rope_cos; o.rope_sin = rope_sin; o.stamps = stamps_dev;
    te::gemm_tc(c, reinterpret_cast<const __half *>(a16), Cin, ntaps, reinterpret_cast<const __half *>(w16), bias, N, o);
    return c.err ? c.err : c.launches;
}

// fp16 attention on its own: qkv16 fp16 [B][T][576] (RoPE already applied), out16 fp16 [B][T][192]
extern "C" int matcha_b200_debug_te_attention_mma(const void *qkv16, const float *mask, void *out16, int B, int T, void *stream) {
    if (!qkv16 || !mask || !out16 || B <= 0 || T <= 0) return MB200_ERR_INVALID_ARG;
    te::Ctx c;
    c.st = (cudaStream_t)stream; c.B = B; c.T = T; c.mask = mask;
    c.err = te::set_attributes();
    te::launch_smem(c, te::te_attention_mma_kernel, dim3((T + te::kMaBQ - 1) / te::kMaBQ, te::kH, B), dim3(128), (size_t)te::kMaSmemBytes,
                    reinterpret_cast<const __half *>(qkv16), mask, reinterpret_cast<__half *>(out16), T);
    return c.err ? c.err : c.launches;
}
