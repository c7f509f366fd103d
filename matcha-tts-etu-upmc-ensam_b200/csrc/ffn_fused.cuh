// ffn_fused.cuh -- the transformer block's feed-forward as ONE tcgen05 kernel:
//     out = res + W2 . gelu(W1 . a + b1) + b2             FeedForward of transformer.py:105-188 (GELU proj 256->1024,
//                                                         LoRACompatibleLinear 1024->256) + the residual of :366
// The 1024-wide hidden activation never leaves the SM.  One CTA owns 128 rows of one sample:
//   * A = the LayerNorm'd rows, 16-bit [128][256], stays in shared memory (4 K-blocks of the 128B-swizzled K-major
//     layout) for the whole tile.
//   * The hidden dimension is walked in 8 chunks of 128 columns.  FFN1(c): acc1[c&1] (TMEM, 128 columns, double
//     buffered) = A . W1[chunk c]^T.  The epilogue warps read it (thread = row), add b1, apply GELU and write the
//     16-bit result straight into the K-major A-operand layout of H[c&1] in shared memory.  FFN2(c): acc2 (TMEM,
//     256 columns) += H[c&1] . W2[:, chunk c]^T.  The MMA warp issues FFN1(c+1) before FFN2(c), so the tensor
//     pipe works on the next chunk while the epilogue warps run the GELU of this one.
//   * W1 / W2 K-blocks stream through a 3-stage ring of 32 KB TMA stages (one stage = 128 hidden rows x 128 k of W1
//     as two boxes, or 256 output rows x 64 hidden of W2).
//   * Final epilogue: acc2 + b2 + residual -> fp32 rows and/or 16-bit rows * mask (the next conv's operand).
// TMEM: acc2 256 + acc1 2 x 128 = 512 columns.  Shared: A 64 KB + H 2 x 32 KB + ring 96 KB = 224 KB.
//
// OUTPROJ variant (round 2): the attention out-projection and norm3 ride in the same kernel, so the whole tail of the
// transformer block (transformer.py:320-368: to_out[0], the residual add, norm3, FeedForward, the second residual add) is
// ONE launch and the intermediate stream x2 never reaches memory:
//   acc2 = att . Wo^T                       (the 16-bit attention rows arrive by TMA in the A region; Wo streams through
//                                            the weight ring ahead of W1)
//   x2   = acc2 + bo + x1                   (epilogue warps, thread = row; written BACK into acc2 with tcgen05.st)
//   A    = LayerNorm(x2) * g + b            (row statistics from per-thread 32-column partials merged with Chan's
//                                            formula across the four warps of a lane quarter; 16-bit into the A region,
//                                            which GEMM0 has finished reading)
//   acc2 += gelu(A W1^T + b1) W2^T          (FFN2 accumulates ON TOP of x2: the second residual add is free)
//   out  = acc2 + b2
// It removes one launch, one 128x256 fp32 round trip and one CTA life cycle per block from the solve's dependent chain.
#pragma once
#include "gemm_bf16.cuh"

namespace mb200 {

constexpr int kFfM = 128, kFfHid = 1024, kFfChunk = 128, kFfNChunks = kFfHid / kFfChunk;
constexpr int kFfStages = 3, kFfStageBytes = 32768;
constexpr int kFfKb = 16384;                                  // one [128][64] 16-bit K-block
constexpr int kFfOffA = 0, kFfOffH = 4 * kFfKb, kFfOffRing = kFfOffH + 2 * 2 * kFfKb;
constexpr int kFfOffBar = kFfOffRing + kFfStages * kFfStageBytes;
constexpr int kFfSmemBytes = kFfOffBar + 256 + 1024;
constexpr int kFfEpiWarps = 16;
constexpr int kFfThreads = 32 * (2 + kFfEpiWarps);
constexpr int kFfEpiPitch = 36;                               // floats, final-epilogue transpose tiles (in the A/H region)

struct FfnArgs {
    int B, T, spp;                     // spp: 32-row slabs per sample (gemm_bf16.cuh::slab_coords); CTA i owns slabs 4i .. 4i+3
    const float *bo;                   // OUTPROJ: to_out[0] bias [256]; res is then the block's input stream x1
    const float *b1, *b2;              // [1024], [256]
    const float *res;                  // fp32 [B][T][256]
    float *out_f32;                    // fp32 [B][T][256] or nullptr
    __nv_bfloat16 *out_b16;            // 16-bit [B][T][256] or nullptr, rows multiplied by omask
    const float *omask; int omask_T;
    float o16_scale;                   // multiplies the 16-bit output as well (fp16 range guard: stream tensors are stored scaled down)
    // norm3 of the transformer block (transformer.py:352) folded in: when ln_in != nullptr the A operand is
    // LayerNorm(ln_in rows) * ln_g + ln_b, built in shared memory by the epilogue warps (idle until the first FFN1
    // chunk completes) instead of arriving by TMA
    const float *ln_in, *ln_g, *ln_b;
    int debug;
};

// erf-GELU for 16-bit outputs with ONE MUFU op: x * Phi(x) = h + h * tanh(x * p(x^2)), h = x / 2, with p the
// weighted minimax fit of atanh(erf(x / sqrt 2)) / x (|error of the GELU value| < 2.6e-5 before tanh.approx's own
// 2^-11 relative error, i.e. at the level of the 16-bit rounding that follows).  x^2 is clamped where tanh has
// long saturated, which keeps p positive.
__device__ __forceinline__ float gelu_tanh_fit_f(float x) {
    const float x2 = fminf(x * x, 64.0f);
    float p = fmaf(-3.51519810e-04f, x2, 3.70056687e-02f);
    p = fmaf(p, x2, 7.97507852e-01f);
    float t;
    asm("tanh.approx.f32 %0, %1;" : "=f"(t) : "f"(p * x));
    const float h = 0.5f * x;
    return fmaf(h, t, h);
}

constexpr uint32_t kFfIdesc1F16 = (1u << 4) | ((uint32_t)(kFfChunk >> 3) << 17) | ((uint32_t)(kFfM >> 4) << 24);   // M128 N128
constexpr uint32_t kFfIdesc1Bf16 = kFfIdesc1F16 | (1u << 7) | (1u << 10);

// profiling aid (FfnArgs.debug): clock64 stamps of CTA 0 -- [0] start, [1] A landed, [2] acc2 complete, [3] end of the final
// epilogue; [8 + 5c ..] chunk c in epilogue warp 0: acc1 ready, TMEM read, GELU done, H free, H written
__device__ long long g_ffn_dbg[64];

template <bool F16, bool OUTPROJ>
__global__ void __launch_bounds__(kFfThreads, 1)
ffn_fused_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmW1,
                 const __grid_constant__ CUtensorMap tmW2, const __grid_constant__ CUtensorMap tmWo, const FfnArgs g) {
    extern __shared__ uint8_t ffsm_raw[];
    const uint32_t base = (smem_u32(ffsm_raw) + 1023u) & ~1023u;
    uint8_t *gbase = ffsm_raw + (base - smem_u32(ffsm_raw));
    const uint32_t bar0 = base + kFfOffBar;
    const uint32_t bar_a = bar0;                              // A tile landed
    const uint32_t bar_full = bar0 + 8;                       // [3] ring stage landed
    const uint32_t bar_empty = bar_full + 8 * kFfStages;      // [3] ring stage consumed
    const uint32_t bar_a1f = bar_empty + 8 * kFfStages;       // [2] acc1 ready
    const uint32_t bar_a1e = bar_a1f + 16;                    // [2] acc1 read by the epilogue warps
    const uint32_t bar_hf = bar_a1e + 16;                     // [2] H chunk written
    const uint32_t bar_he = bar_hf + 16;                      // [2] H chunk consumed by FFN2
    const uint32_t bar_a2f = bar_he + 16;                     // acc2 complete
    const uint32_t bar_att = bar_a2f + 8;                     // OUTPROJ: attention rows landed
    const uint32_t bar_g0 = bar_att + 8;                      // OUTPROJ: att . Wo^T complete
    const uint32_t tmem_slot = bar_g0 + 8;
    volatile uint32_t *tmem_slot_p = reinterpret_cast<volatile uint32_t *>(gbase + (tmem_slot - base));

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

    if (threadIdx.x == 0) {
        mbar_init(bar_a, (OUTPROJ || g.ln_in) ? (uint32_t)kFfEpiWarps : 1u); mbar_init(bar_a2f, 1u);
        mbar_init(bar_att, 1u); mbar_init(bar_g0, 1u);
        for (int s = 0; s < kFfStages; ++s) { mbar_init(bar_full + 8u * s, 1u); mbar_init(bar_empty + 8u * s, 1u); }
        for (int a = 0; a < 2; ++a) {
            mbar_init(bar_a1f + 8u * a, 1u); mbar_init(bar_a1e + 8u * a, kFfEpiWarps);
            mbar_init(bar_hf + 8u * a, kFfEpiWarps); mbar_init(bar_he + 8u * a, 1u);
        }
        mbar_fence_init();
        tma_prefetch_desc(&tmA); tma_prefetch_desc(&tmW1); tma_prefetch_desc(&tmW2);
        if (OUTPROJ) tma_prefetch_desc(&tmWo);
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "n"(512) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem = *tmem_slot_p;
    const uint32_t tmem_acc2 = tmem, tmem_acc1 = tmem + 256;
    pdl_trigger_at_top();
    const bool dbg = g.debug && blockIdx.x == 0;
    if (dbg && threadIdx.x == 0) g_ffn_dbg[0] = clock64();

    if (warp == 0) {
        // ============================ TMA producer ============================
        if (lane == 0) {
            int sb_[4], st_[4];                               // the tile's four slabs: sample, first frame
#pragma unroll
            for (int i = 0; i < 4; ++i) slab_coords(4 * (int)blockIdx.x + i, g.spp, g.B, g.T, sb_[i], st_[i]);
            uint32_t it = 0;
            auto stage_w1 = [&](int c, int half) {            // W1 rows [128c, 128c+128), k in [128 half, 128 half + 128)
                const uint32_t s = it % kFfStages;
                mbar_wait_parked(bar_empty + 8u * s, ((it / kFfStages) & 1) ^ 1, 500);
                const uint32_t dst = base + kFfOffRing + s * kFfStageBytes;
                mbar_arrive_expect_tx(bar_full + 8u * s, kFfStageBytes);
                tma_load_2d(dst, &tmW1, bar_full + 8u * s, (2 * half) * 64, c * kFfChunk);
                tma_load_2d(dst + kFfKb, &tmW1, bar_full + 8u * s, (2 * half + 1) * 64, c * kFfChunk);
                ++it;
            };
            auto stage_w2 = [&](int c, int kb) {              // W2 all 256 rows, hidden k in [128c + 64kb, +64)
                const uint32_t s = it % kFfStages;
                mbar_wait_parked(bar_empty + 8u * s, ((it / kFfStages) & 1) ^ 1, 500);
                mbar_arrive_expect_tx(bar_full + 8u * s, kFfStageBytes);
                tma_load_2d(base + kFfOffRing + s * kFfStageBytes, &tmW2, bar_full + 8u * s, c * kFfChunk + kb * 64, 0);
                ++it;
            };
            auto stage_wo = [&](int kb) {                      // Wo all 256 rows, k in [64 kb, +64): same stage shape as W2
                const uint32_t s = it % kFfStages;
                mbar_wait_parked(bar_empty + 8u * s, ((it / kFfStages) & 1) ^ 1, 500);
                mbar_arrive_expect_tx(bar_full + 8u * s, kFfStageBytes);
                tma_load_2d(base + kFfOffRing + s * kFfStageBytes, &tmWo, bar_full + 8u * s, kb * 64, 0);
                ++it;
            };
            // the weights do not depend on the predecessor kernel: the first stages are in flight before the wait
            if (OUTPROJ) {
                stage_wo(0); stage_wo(1); stage_wo(2);
                pdl_wait_then_trigger();
                mbar_arrive_expect_tx(bar_att, 4 * kFfKb);
                for (int kb = 0; kb < 4; ++kb)
#pragma unroll
                    for (int i = 0; i < 4; ++i) tma_load_3d(base + kFfOffA + kb * kFfKb + 4096u * i, &tmA, bar_att, kb * 64, st_[i], sb_[i]);
                stage_wo(3);
                stage_w1(0, 0); stage_w1(0, 1); stage_w1(1, 0);
            } else {
                stage_w1(0, 0); stage_w1(0, 1); stage_w1(1, 0);
                pdl_wait_then_trigger();
                if (!g.ln_in) {
                    mbar_arrive_expect_tx(bar_a, 4 * kFfKb);
                    for (int kb = 0; kb < 4; ++kb)
#pragma unroll
                        for (int i = 0; i < 4; ++i) tma_load_3d(base + kFfOffA + kb * kFfKb + 4096u * i, &tmA, bar_a, kb * 64, st_[i], sb_[i]);
                }
            }
            stage_w1(1, 1);
            for (int s = 2; s < kFfNChunks; ++s) {            // same order as the MMA warp: FFN1 runs two chunks ahead of FFN2
                stage_w1(s, 0); stage_w1(s, 1);
                stage_w2(s - 2, 0); stage_w2(s - 2, 1);
            }
            for (int c = kFfNChunks - 2; c < kFfNChunks; ++c) { stage_w2(c, 0); stage_w2(c, 1); }
        }
    } else if (warp == 1) {
        // ============================ MMA issuer ============================
        if (lane == 0) {
            const uint32_t idesc1 = F16 ? kFfIdesc1F16 : kFfIdesc1Bf16;
            const uint32_t idesc2 = F16 ? kIdescF16 : kIdescBf16;
            uint32_t it = 0;
            auto ffn1 = [&](int c) {
                const uint32_t buf = c & 1;
                if (c >= 2) mbar_wait_parked(bar_a1e + 8u * buf, ((c >> 1) - 1) & 1, 200);
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                for (int half = 0; half < 2; ++half, ++it) {
                    const uint32_t s = it % kFfStages;
                    mbar_wait_parked(bar_full + 8u * s, (it / kFfStages) & 1, 200);
                    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
#pragma unroll
                    for (int kb2 = 0; kb2 < 2; ++kb2) {
                        const uint64_t da = umma_desc_sw128(base + kFfOffA + (2 * half + kb2) * kFfKb);
                        const uint64_t db = umma_desc_sw128(base + kFfOffRing + s * kFfStageBytes + kb2 * kFfKb);
#pragma unroll
                        for (int k = 0; k < 4; ++k)
                            umma_bf16(tmem_acc1 + buf * kFfChunk, da + 2u * k, db + 2u * k, idesc1, (half | kb2 | k) != 0 ? 1u : 0u);
                    }
                    umma_commit(bar_empty + 8u * s);
                }
                umma_commit(bar_a1f + 8u * buf);
            };
            auto ffn2 = [&](int c) {
                const uint32_t buf = c & 1;
                mbar_wait_parked(bar_hf + 8u * buf, (c >> 1) & 1, 200);
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                for (int kb = 0; kb < 2; ++kb, ++it) {
                    const uint32_t s = it % kFfStages;
                    mbar_wait_parked(bar_full + 8u * s, (it / kFfStages) & 1, 200);
                    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                    const uint64_t da = umma_desc_sw128(base + kFfOffH + (2 * buf + kb) * kFfKb);
                    const uint64_t db = umma_desc_sw128(base + kFfOffRing + s * kFfStageBytes);
#pragma unroll
                    for (int k = 0; k < 4; ++k)
                        umma_bf16(tmem_acc2, da + 2u * k, db + 2u * k, idesc2, (OUTPROJ || (c | kb | k) != 0) ? 1u : 0u);
                    umma_commit(bar_empty + 8u * s);
                }
                umma_commit(bar_he + 8u * buf);
            };
            if (OUTPROJ) {
                // GEMM0: acc2 = att . Wo^T (K = 256: four K-blocks of the A region against four ring stages)
                mbar_wait_parked(bar_att, 0, 500);
                for (int kb = 0; kb < 4; ++kb, ++it) {
                    const uint32_t s = it % kFfStages;
                    mbar_wait_parked(bar_full + 8u * s, (it / kFfStages) & 1, 200);
                    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                    const uint64_t da = umma_desc_sw128(base + kFfOffA + kb * kFfKb);
                    const uint64_t db = umma_desc_sw128(base + kFfOffRing + s * kFfStageBytes);
#pragma unroll
                    for (int k = 0; k < 4; ++k)
                        umma_bf16(tmem_acc2, da + 2u * k, db + 2u * k, idesc2, (kb | k) != 0 ? 1u : 0u);
                    umma_commit(bar_empty + 8u * s);
                }
                umma_commit(bar_g0);
            }
            mbar_wait_parked(bar_a, 0, 500);
            // FFN1 runs two chunks ahead of FFN2: the GELU of chunk c has FFN2(c-2), FFN1(c+1), FFN2(c-1) and FFN1(c+2)
            // (4 x 1024 tensor-pipe cycles) to finish before FFN2(c) needs H(c), so the epilogue warps are bound by their
            // throughput, not by their latency
            ffn1(0); ffn1(1);
            for (int s = 2; s < kFfNChunks; ++s) { ffn1(s); ffn2(s - 2); }
            ffn2(kFfNChunks - 2); ffn2(kFfNChunks - 1);
            umma_commit(bar_a2f);
        }
    } else {
        // ============================ epilogue warps ============================
        const int ew = warp - 2;
        const int q = warp & 3;                              // TMEM lane quarter this warp may touch
        const int cg = ew >> 2;                              // column group
        const int row = 32 * q + lane;                       // tile row = TMEM lane
        const uint32_t lane_off = (uint32_t)(32 * q) << 16;
        int b, tq;                                           // this warp's lane quarter = slab q of the tile
        slab_coords(4 * (int)blockIdx.x + q, g.spp, g.B, g.T, b, tq);
        if (OUTPROJ) {
            // ---- x2 = att Wo^T + bo + x1 (back into acc2), then A = LayerNorm(x2): thread = row, columns [64 cg, +64)
            pdl_wait();
            const bool ok = tq + lane < g.T;
            const float *x1p = g.res + ((size_t)b * g.T + (ok ? tq + lane : 0)) * kC + cg * 64;
            float4 rv[8];
#pragma unroll
            for (int i = 0; i < 8; ++i) rv[i] = ok ? ld_chain4(x1p + 4 * i) : make_float4(0.f, 0.f, 0.f, 0.f);
            if (dbg && threadIdx.x == 64) g_ffn_dbg[4] = clock64();
            mbar_wait_parked(bar_g0, 0, 200);
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            if (dbg && threadIdx.x == 64) g_ffn_dbg[5] = clock64();
            float pm[2], pq[2];                              // per 32-column half: mean and centred sum of squares
#pragma unroll
            for (int hh = 0; hh < 2; ++hh) {
                uint32_t v[32];
                tmem_ld32(tmem_acc2 + cg * 64 + hh * 32 + lane_off, v);
                const float4 *bo4 = reinterpret_cast<const float4 *>(g.bo + cg * 64 + hh * 32);
                float sum = 0.f;
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    const float4 bb = __ldg(bo4 + i);
                    const float x0 = __uint_as_float(v[4 * i]) + bb.x + rv[i].x, x1 = __uint_as_float(v[4 * i + 1]) + bb.y + rv[i].y;
                    const float x2 = __uint_as_float(v[4 * i + 2]) + bb.z + rv[i].z, x3 = __uint_as_float(v[4 * i + 3]) + bb.w + rv[i].w;
                    v[4 * i] = __float_as_uint(x0); v[4 * i + 1] = __float_as_uint(x1); v[4 * i + 2] = __float_as_uint(x2); v[4 * i + 3] = __float_as_uint(x3);
                    sum += (x0 + x1) + (x2 + x3);
                }
                if (hh == 0) {
#pragma unroll
                    for (int i = 0; i < 8; ++i) rv[i] = ok ? ld_chain4(x1p + 32 + 4 * i) : make_float4(0.f, 0.f, 0.f, 0.f);
                }
                tmem_st32(tmem_acc2 + cg * 64 + hh * 32 + lane_off, v);
                const float m = sum * (1.0f / 32.0f);
                float qq = 0.f;
#pragma unroll
                for (int i = 0; i < 32; ++i) { const float d = __uint_as_float(v[i]) - m; qq = fmaf(d, d, qq); }
                pm[hh] = m; pq[hh] = qq;
            }
            // merge the eight 32-column partials of a row (two here, six in the other three warps of the lane quarter):
            // equal counts, so mean = average of the means and M2 = sum M2_i + 32 sum (mean_i - mean)^2  (Chan et al.)
            float4 *red = reinterpret_cast<float4 *>(gbase + kFfOffH);          // [128 rows][4 column groups], H is idle
            if (dbg && threadIdx.x == 64) g_ffn_dbg[6] = clock64();
            red[row * 4 + cg] = make_float4(pm[0], pq[0], pm[1], pq[1]);
            if (q == 0) asm volatile("bar.sync 1, 128;" ::: "memory");
            else if (q == 1) asm volatile("bar.sync 2, 128;" ::: "memory");
            else if (q == 2) asm volatile("bar.sync 3, 128;" ::: "memory");
            else asm volatile("bar.sync 4, 128;" ::: "memory");
            float mean = 0.f, m2 = 0.f;
            float4 pr[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) { pr[i] = red[row * 4 + i]; mean += pr[i].x + pr[i].z; m2 += pr[i].y + pr[i].w; }
            mean *= 0.125f;
#pragma unroll
            for (int i = 0; i < 4; ++i) { const float d0 = pr[i].x - mean, d1 = pr[i].z - mean; m2 = fmaf(32.0f * d0, d0, m2); m2 = fmaf(32.0f * d1, d1, m2); }
            const float rstd = rsqrtf(m2 * (1.0f / kC) + 1e-5f);
            // the x2 rows written above must be visible to this thread's own re-read
            asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
            if (dbg && threadIdx.x == 64) g_ffn_dbg[7] = clock64();
            const uint32_t arow = base + kFfOffA + (uint32_t)cg * kFfKb + (uint32_t)((row >> 3) * 1024 + (row & 7) * 128);
#pragma unroll
            for (int hh = 0; hh < 2; ++hh) {
                uint32_t v[32];
                tmem_ld32(tmem_acc2 + cg * 64 + hh * 32 + lane_off, v);
                const float4 *g4 = reinterpret_cast<const float4 *>(g.ln_g + cg * 64 + hh * 32);
                const float4 *b4 = reinterpret_cast<const float4 *>(g.ln_b + cg * 64 + hh * 32);
                uint32_t pk[16];
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    const float4 gg = __ldg(g4 + i), bb = __ldg(b4 + i);
                    const float y0 = fmaf((__uint_as_float(v[4 * i]) - mean) * rstd, gg.x, bb.x), y1 = fmaf((__uint_as_float(v[4 * i + 1]) - mean) * rstd, gg.y, bb.y);
                    const float y2 = fmaf((__uint_as_float(v[4 * i + 2]) - mean) * rstd, gg.z, bb.z), y3 = fmaf((__uint_as_float(v[4 * i + 3]) - mean) * rstd, gg.w, bb.w);
                    pk[2 * i] = ok ? pack16x2(y0, y1, F16) : 0u;
                    pk[2 * i + 1] = ok ? pack16x2(y2, y3, F16) : 0u;
                }
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const uint32_t piece = (uint32_t)(hh * 4 + i);
                    asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(arow + ((piece ^ (uint32_t)(row & 7)) << 4)),
                                 "r"(pk[4 * i]), "r"(pk[4 * i + 1]), "r"(pk[4 * i + 2]), "r"(pk[4 * i + 3]) : "memory");
                }
            }
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
            __syncwarp();
            if (lane == 0) mbar_arrive(bar_a);
            if (dbg && threadIdx.x == 64) g_ffn_dbg[1] = clock64();
        } else if (g.ln_in) {
            // ---- A = LayerNorm(256, eps 1e-5) of the tile's rows, one warp per row (8 rows per warp, 4 in flight),
            //      written 16-bit into the four 128B-swizzled K-blocks the UMMA descriptors expect
            pdl_wait();
            int rb, rt0;                                      // warp ew builds tile rows [8 ew, 8 ew + 8): slab ew / 4
            slab_coords(4 * (int)blockIdx.x + (ew >> 2), g.spp, g.B, g.T, rb, rt0);
            const float4 lg0 = *reinterpret_cast<const float4 *>(g.ln_g + lane * 4), lg1 = *reinterpret_cast<const float4 *>(g.ln_g + 128 + lane * 4);
            const float4 lb0 = *reinterpret_cast<const float4 *>(g.ln_b + lane * 4), lb1 = *reinterpret_cast<const float4 *>(g.ln_b + 128 + lane * 4);
#pragma unroll
            for (int half = 0; half < 2; ++half) {
                float4 xa[4], xb[4];
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const int r = ew * 8 + half * 4 + j;
                    const int rt = rt0 + (r & 31);
                    const bool ok = rt < g.T;
                    const float *px = g.ln_in + ((size_t)rb * g.T + (ok ? rt : 0)) * kC;
                    xa[j] = ok ? ld_chain4(px + lane * 4) : make_float4(0.f, 0.f, 0.f, 0.f);
                    xb[j] = ok ? ld_chain4(px + 128 + lane * 4) : make_float4(0.f, 0.f, 0.f, 0.f);
                }
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const int r = ew * 8 + half * 4 + j;
                    float sum = (xa[j].x + xa[j].y) + (xa[j].z + xa[j].w) + (xb[j].x + xb[j].y) + (xb[j].z + xb[j].w);
#pragma unroll
                    for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
                    const float mean = sum * (1.0f / kC);
                    float d[8] = {xa[j].x - mean, xa[j].y - mean, xa[j].z - mean, xa[j].w - mean,
                                  xb[j].x - mean, xb[j].y - mean, xb[j].z - mean, xb[j].w - mean};
                    float qq = 0.f;
#pragma unroll
                    for (int i = 0; i < 8; ++i) qq = fmaf(d[i], d[i], qq);
#pragma unroll
                    for (int o = 16; o > 0; o >>= 1) qq += __shfl_xor_sync(0xffffffffu, qq, o);
                    const float rstd = rsqrtf(qq * (1.0f / kC) + 1e-5f);
                    const bool ok = rt0 + (r & 31) < g.T;   // rows beyond T are zeros, as the TMA path delivers them
                    const uint32_t p0 = ok ? pack16x2(d[0] * rstd * lg0.x + lb0.x, d[1] * rstd * lg0.y + lb0.y, F16) : 0u;
                    const uint32_t p1 = ok ? pack16x2(d[2] * rstd * lg0.z + lb0.z, d[3] * rstd * lg0.w + lb0.w, F16) : 0u;
                    const uint32_t p2 = ok ? pack16x2(d[4] * rstd * lg1.x + lb1.x, d[5] * rstd * lg1.y + lb1.y, F16) : 0u;
                    const uint32_t p3 = ok ? pack16x2(d[6] * rstd * lg1.z + lb1.z, d[7] * rstd * lg1.w + lb1.w, F16) : 0u;
                    // channels 4 lane .. +3: K-block lane / 16, 16-byte piece (lane % 16) / 2 (swizzled), half (lane & 1)
                    const uint32_t dst = base + kFfOffA + (uint32_t)(lane >> 4) * kFfKb + (uint32_t)((r >> 3) * 1024 + (r & 7) * 128) +
                                         (((uint32_t)((lane & 15) >> 1) ^ (uint32_t)(r & 7)) << 4) + (uint32_t)(lane & 1) * 8u;
                    asm volatile("st.shared.v2.b32 [%0], {%1, %2};" ::"r"(dst), "r"(p0), "r"(p1) : "memory");
                    asm volatile("st.shared.v2.b32 [%0], {%1, %2};" ::"r"(dst + 2u * kFfKb), "r"(p2), "r"(p3) : "memory");
                }
            }
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            __syncwarp();
            if (lane == 0) mbar_arrive(bar_a);
        }
        // ---- GELU of every hidden chunk: 32 columns [32 cg, 32 cg + 32) of the chunk per warp
        const uint32_t hrow = (uint32_t)((row >> 3) * 1024 + (row & 7) * 128);
        // b1 is read through L1 (the few KB the 224 KB carve-out leaves): each warp's 128-byte line is prefetched two
        // chunks ahead, so that the loads below do not pay the L2 round trip inside the chunk
        if (lane < 2) asm volatile("prefetch.global.L1 [%0];" ::"l"(g.b1 + lane * kFfChunk + cg * 32));
        for (int c = 0; c < kFfNChunks; ++c) {
            const uint32_t buf = c & 1;
            const float4 *b1v = reinterpret_cast<const float4 *>(g.b1 + c * kFfChunk + cg * 32);
            if (lane == 0 && c + 2 < kFfNChunks) asm volatile("prefetch.global.L1 [%0];" ::"l"(g.b1 + (c + 2) * kFfChunk + cg * 32));
            mbar_wait_parked(bar_a1f + 8u * buf, (c >> 1) & 1, 100);
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            const bool dw = dbg && threadIdx.x == 64;
            if (dw) g_ffn_dbg[8 + 5 * c] = clock64();
            uint32_t v[32];
            tmem_ld32(tmem_acc1 + buf * kFfChunk + cg * 32 + lane_off, v);
            if (dw) g_ffn_dbg[9 + 5 * c] = clock64();
            asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
            __syncwarp();
            if (lane == 0) mbar_arrive(bar_a1e + 8u * buf);
            uint32_t pk[16];
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const float4 bb = __ldg(b1v + i);
                const float y0 = gelu_tanh_fit_f(__uint_as_float(v[4 * i]) + bb.x);
                const float y1 = gelu_tanh_fit_f(__uint_as_float(v[4 * i + 1]) + bb.y);
                const float y2 = gelu_tanh_fit_f(__uint_as_float(v[4 * i + 2]) + bb.z);
                const float y3 = gelu_tanh_fit_f(__uint_as_float(v[4 * i + 3]) + bb.w);
                pk[2 * i] = pack16x2(y0, y1, F16);
                pk[2 * i + 1] = pack16x2(y2, y3, F16);
            }
            if (dw) g_ffn_dbg[10 + 5 * c] = clock64();
            if (c >= 2) mbar_wait_parked(bar_he + 8u * buf, ((c >> 1) - 1) & 1, 100);    // FFN2(c-2) has consumed H[buf]
            if (dw) g_ffn_dbg[11 + 5 * c] = clock64();
            // columns [32 cg, +32) of the chunk = K-block cg / 2, 16-byte pieces 4 (cg & 1) .. +3 of the row
            const uint32_t hdst = base + kFfOffH + (2 * buf + (cg >> 1)) * kFfKb + hrow;
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const uint32_t piece = (uint32_t)((cg & 1) * 4 + i);
                asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(hdst + ((piece ^ (uint32_t)(row & 7)) << 4)),
                             "r"(pk[4 * i]), "r"(pk[4 * i + 1]), "r"(pk[4 * i + 2]), "r"(pk[4 * i + 3]) : "memory");
            }
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            __syncwarp();
            if (lane == 0) mbar_arrive(bar_hf + 8u * buf);
            if (dw) g_ffn_dbg[12 + 5 * c] = clock64();
        }
        // ---- final epilogue: columns [64 cg, 64 cg + 64) of acc2, transposed through shared memory (the A / H
        //      region is free: every MMA has completed) so that global accesses are row-contiguous
        pdl_wait();                                           // g.res is the predecessor chain's output
        float *tb = reinterpret_cast<float *>(gbase) + ew * (32 * kFfEpiPitch);
        const int rsub = lane >> 3, c4 = (lane & 7) * 4;
        int orow[8];
        float mk[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const int t = tq + 4 * i + rsub;
            const bool ok = t < g.T;
            orow[i] = ok ? (int)(((long long)b * g.T + t) * kC) : -1;
            mk[i] = g.o16_scale * ((ok && g.omask) ? g.omask[(size_t)b * g.omask_T + t] : 1.0f);
        }
        float4 rv[8];
        auto load_res = [&](int jj) {
            const int n = cg * 64 + jj * 32 + c4;
#pragma unroll
            for (int i = 0; i < 8; ++i)
                rv[i] = orow[i] >= 0 ? ld_chain4(g.res + orow[i] + n) : make_float4(0.f, 0.f, 0.f, 0.f);
        };
        if (!OUTPROJ) load_res(0);                            // in flight while the last FFN2 chunks complete
        else {
#pragma unroll
            for (int i = 0; i < 8; ++i) rv[i] = make_float4(0.f, 0.f, 0.f, 0.f);   // the residual already sits in acc2
        }
        mbar_wait_parked(bar_a2f, 0, 200);
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        if (dbg && threadIdx.x == 64) g_ffn_dbg[2] = clock64();
#pragma unroll
        for (int jj = 0; jj < 2; ++jj) {
            const int n = cg * 64 + jj * 32 + c4;
            const float4 bias = *reinterpret_cast<const float4 *>(g.b2 + n);
            uint32_t v[32];
            tmem_ld32(tmem_acc2 + cg * 64 + jj * 32 + lane_off, v);
            __syncwarp();
#pragma unroll
            for (int i = 0; i < 8; ++i)
                *reinterpret_cast<float4 *>(tb + lane * kFfEpiPitch + 4 * i) =
                    make_float4(__uint_as_float(v[4 * i]), __uint_as_float(v[4 * i + 1]), __uint_as_float(v[4 * i + 2]), __uint_as_float(v[4 * i + 3]));
            __syncwarp();
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                float4 x = *reinterpret_cast<const float4 *>(tb + (4 * i + rsub) * kFfEpiPitch + c4);
                x.x += bias.x + rv[i].x; x.y += bias.y + rv[i].y; x.z += bias.z + rv[i].z; x.w += bias.w + rv[i].w;
                if (!OUTPROJ && jj == 0 && i == 7) load_res(1);
                if (orow[i] >= 0) {
                    if (g.out_f32) *reinterpret_cast<float4 *>(g.out_f32 + orow[i] + n) = x;
                    if (g.out_b16)
                        *reinterpret_cast<uint2 *>(g.out_b16 + orow[i] + n) =
                            make_uint2(pack16x2(x.x * mk[i], x.y * mk[i], F16), pack16x2(x.z * mk[i], x.w * mk[i], F16));
                }
            }
        }
    }
    if (dbg && threadIdx.x == 64) g_ffn_dbg[3] = clock64();
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 1) {
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "n"(512) : "memory");
    }
}

}  // namespace mb200
