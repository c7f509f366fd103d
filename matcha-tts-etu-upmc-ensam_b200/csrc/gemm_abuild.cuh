// gemm_abuild.cuh -- tcgen05 GEMMs whose A operand is BUILT in shared memory by the CTA itself from fp32 rows,
// instead of arriving by TMA: the GroupNorm / Mish / LayerNorm passes between the decoder's GEMMs disappear as
// separate launches (and with them one 16-bit round trip through L2 each).
//   MODE 1 (q|k|v projection, transformer.py:251-259 after decoder.py:85 and transformer.py:316):
//       x1 = Mish(GN(h)) * mask + r            (Block1D output + res_conv; written back as the fp32 residual stream)
//       A  = LayerNorm(x1) * ln_g + ln_b       (norm1)
//       out16[.., N] = A W^T (+ bias)          N = 768
//   MODE 2 (block2's conv of Resnet1D, decoder.py:59-66,83-84):
//       A  = (Mish(GN(h)) * mask + tb) * mask  rows t0-1 .. t0+128 of the sample (zeros outside [0, T))
//       out[.., 256] = sum over taps -1, 0, +1 of A[t + tap] W_tap^T + bias, fp32, plus its GroupNorm sums
//       The three taps are three row-shifted views of the ONE shared-memory A tile: the UMMA descriptor's start
//       address moves by 128 bytes (one row of the 128B-swizzled K-major layout) per tap.  Measured on sm_100a: the
//       128B swizzle is a function of the absolute shared-memory address, so the shifted start address alone is right
//       and the descriptor's base-offset field must stay 0 (setting it to the row phase gives wrong results).
// One persistent CTA per SM walks M tiles of 128 rows; per tile the 8 epilogue warps first build A (4 K-blocks of
// [rows][64] 16-bit, 128B swizzle), then drain the accumulators (thread = row, swizzled 2 KB staging tiles, TMA stores)
// while the MMA warp already works on the next N tile.  W K-blocks stream through a 3-stage TMA ring and do not depend
// on the predecessor kernel, so they are in flight before griddepcontrol.wait.
#pragma once
#include "gemm_bf16.cuh"

namespace mb200 {

constexpr int kAbStages = 3, kAbStageBytes = 32768;          // one W K-block: [256 rows][64] 16-bit
constexpr int kAbAKb = 17 * 1024;                            // one A K-block: up to 136 rows x 128 B
constexpr int kAbOffA = 0, kAbOffRing = 4 * kAbAKb;
constexpr int kAbOffStg = kAbOffRing + kAbStages * kAbStageBytes;
constexpr int kAbEpiWarps = 8;
constexpr int kAbStgWarp = 4096;                             // two 2 KB staging tiles (32 rows x 64 B, 64B swizzle)
constexpr int kAbOffBias = kAbOffStg + kAbEpiWarps * kAbStgWarp;
constexpr int kAbOffBar = kAbOffBias + 4096;
constexpr int kAbSmemBytes = kAbOffBar + 256 + 1024;
constexpr int kAbThreads = 32 * (2 + kAbEpiWarps);

struct AbuildArgs {
    int B, T, tiles_per_sample, N;
    const float *h;                    // fp32 [B][T][256]: conv output to be normalised
    const double *gsum;                // [B][8][2] GroupNorm sums of h
    const float *gamma, *beta;         // GroupNorm affine
    const float *mask; int mask_T;
    const float *tb; int tb_per_sample;            // MODE 2: time bias rows [.][256] (or nullptr)
    const float *r, *ln_g, *ln_b;      // MODE 1: res_conv output, norm1 affine
    float *x1;                         // MODE 1: fp32 residual stream out
    const float *bias;                 // [N] or nullptr
    double *gstats;                    // MODE 2: GroupNorm sums of the output
    int desc_mode;                     // MODE 2 experiment: 1 = also set the descriptor base-offset field (wrong on sm_100a)
};

// K-major 128B-swizzle descriptor whose start address may sit `rows` (0..7) rows into an 8-row swizzle atom
__device__ __forceinline__ uint64_t umma_desc_sw128_rows(uint32_t smem_addr, uint32_t base_offset) {
    return umma_desc_sw128(smem_addr) | ((uint64_t)(base_offset & 7u) << 49);
}

template <int MODE, bool F16>
__global__ void __launch_bounds__(kAbThreads, 1)
gemm_abuild_kernel(const __grid_constant__ CUtensorMap tmB, const __grid_constant__ CUtensorMap tmO, const AbuildArgs g) {
    constexpr int NTAPS = MODE == 2 ? 3 : 1;
    constexpr int AROWS = MODE == 2 ? 130 : 128;
    extern __shared__ uint8_t absm_raw[];
    const uint32_t base = (smem_u32(absm_raw) + 1023u) & ~1023u;
    uint8_t *gbase = absm_raw + (base - smem_u32(absm_raw));
    float *const sbias = reinterpret_cast<float *>(gbase + kAbOffBias);
    const uint32_t bar0 = base + kAbOffBar;
    const uint32_t bar_full = bar0;                           // [3] W stage landed
    const uint32_t bar_empty = bar_full + 8 * kAbStages;      // [3] W stage consumed
    const uint32_t bar_tfull = bar_empty + 8 * kAbStages;     // [2] accumulator ready
    const uint32_t bar_tempty = bar_tfull + 16;               // [2] accumulator drained
    const uint32_t bar_afull = bar_tempty + 16;               // A tile built
    const uint32_t bar_aempty = bar_afull + 8;                // A tile consumed by every MMA of the M tile
    const uint32_t tmem_slot = bar_aempty + 8;
    volatile uint32_t *tmem_slot_p = reinterpret_cast<volatile uint32_t *>(gbase + (tmem_slot - base));

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int n_mtiles = g.B * g.tiles_per_sample;
    const int n_nt = g.N / kGN;
    const int n_kb = NTAPS * 4;

    if (threadIdx.x == 0) {
        for (int s = 0; s < kAbStages; ++s) { mbar_init(bar_full + 8u * s, 1u); mbar_init(bar_empty + 8u * s, 1u); }
        for (int a = 0; a < 2; ++a) { mbar_init(bar_tfull + 8u * a, 1u); mbar_init(bar_tempty + 8u * a, kAbEpiWarps); }
        mbar_init(bar_afull, kAbEpiWarps); mbar_init(bar_aempty, 1u);
        mbar_fence_init();
        tma_prefetch_desc(&tmB); tma_prefetch_desc(&tmO);
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "n"(512) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem_d = *tmem_slot_p;
    pdl_trigger_at_top();

    if (warp == 0) {
        // ============================ TMA producer: W K-blocks only ============================
        if (lane == 0) {
            uint32_t it = 0;
            for (int tile = blockIdx.x; tile < n_mtiles; tile += gridDim.x)
                for (int nt = 0; nt < n_nt; ++nt)
                    for (int kb = 0; kb < n_kb; ++kb, ++it) {
                        const uint32_t s = it % kAbStages;
                        mbar_wait_parked(bar_empty + 8u * s, ((it / kAbStages) & 1) ^ 1, 500);
                        mbar_arrive_expect_tx(bar_full + 8u * s, kAbStageBytes);
                        tma_load_2d(base + kAbOffRing + s * kAbStageBytes, &tmB, bar_full + 8u * s, kb * kGK, nt * kGN);
                    }
        }
    } else if (warp == 1) {
        // ============================ MMA issuer ============================
        if (lane == 0) {
            const uint32_t idesc = F16 ? kIdescF16 : kIdescBf16;
            uint32_t it = 0, ac = 0, lt = 0;
            for (int tile = blockIdx.x; tile < n_mtiles; tile += gridDim.x, ++lt) {
                mbar_wait_parked(bar_afull, lt & 1, 500);
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                for (int nt = 0; nt < n_nt; ++nt, ++ac) {
                    const uint32_t acc = ac & 1;
                    mbar_wait_parked(bar_tempty + 8u * acc, ((ac >> 1) & 1) ^ 1, 500);
                    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                    for (int kb = 0; kb < n_kb; ++kb, ++it) {
                        const uint32_t s = it % kAbStages;
                        const uint32_t tap = (uint32_t)kb >> 2, kc = (uint32_t)kb & 3u;
                        mbar_wait_parked(bar_full + 8u * s, (it / kAbStages) & 1, 500);
                        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                        // tap views: rows tap .. tap + 127 of the A tile (row 0 = frame t0 - 1 in MODE 2)
                        const uint64_t da = umma_desc_sw128_rows(base + kAbOffA + kc * kAbAKb + tap * 128u, g.desc_mode ? tap : 0u);
                        const uint64_t db = umma_desc_sw128(base + kAbOffRing + s * kAbStageBytes);
#pragma unroll
                        for (int k = 0; k < kGK / 16; ++k)
                            umma_bf16(tmem_d + acc * kGN, da + 2u * k, db + 2u * k, idesc, (kb | k) != 0 ? 1u : 0u);
                        umma_commit(bar_empty + 8u * s);
                    }
                    umma_commit(bar_tfull + 8u * acc);
                }
                umma_commit(bar_aempty);
            }
        }
    } else {
        // ============================ A builders / epilogue warps ============================
        const int ew = warp - 2;
        const int q = warp & 3;                              // TMEM lane quarter
        const int half = ew >> 2;                            // 128-column half of the N tile
        const uint32_t stg = base + kAbOffStg + ew * kAbStgWarp;
        const int c0 = lane * 4, c1 = 128 + lane * 4, g0 = lane >> 3, g1 = 4 + (lane >> 3);
        for (int i = threadIdx.x - 64; i < g.N; i += 32 * kAbEpiWarps) sbias[i] = g.bias ? g.bias[i] : 0.f;
        asm volatile("bar.sync 1, %0;" ::"n"(32 * kAbEpiWarps) : "memory");
        const float4 ga0 = *reinterpret_cast<const float4 *>(g.gamma + c0), ga1 = *reinterpret_cast<const float4 *>(g.gamma + c1);
        const float4 be0 = *reinterpret_cast<const float4 *>(g.beta + c0), be1 = *reinterpret_cast<const float4 *>(g.beta + c1);
        float4 lg0, lg1, lb0, lb1;
        if (MODE == 1) {
            lg0 = *reinterpret_cast<const float4 *>(g.ln_g + c0); lg1 = *reinterpret_cast<const float4 *>(g.ln_g + c1);
            lb0 = *reinterpret_cast<const float4 *>(g.ln_b + c0); lb1 = *reinterpret_cast<const float4 *>(g.ln_b + c1);
        }
        // this lane's two 16-bit destinations inside an A row: channels c0.. in K-block lane / 16, c1.. two K-blocks on
        const uint32_t a_dst0 = base + kAbOffA + (uint32_t)(lane >> 4) * kAbAKb + (uint32_t)(lane & 1) * 8u;
        const uint32_t a_piece = (uint32_t)((lane & 15) >> 1);
        pdl_wait_then_trigger();                             // h, gsum, r come from the predecessor kernels
        uint32_t lt = 0, ac = 0, n_st = 0;
        for (int tile = blockIdx.x; tile < n_mtiles; tile += gridDim.x, ++lt) {
            const int b = tile / g.tiles_per_sample, t0 = (tile - b * g.tiles_per_sample) * kGM;
            // GroupNorm mean / rstd of this lane's two channel groups (sums from the producing conv's epilogue)
            float m0, rs0, m1, rs1;
            {
                const double inv_n = 1.0 / ((double)g.T * kGroupCh);
                const double *gp = g.gsum + (size_t)b * kGroups * 2;
                const double ma = gp[g0 * 2] * inv_n, mb = gp[g1 * 2] * inv_n;
                double va = gp[g0 * 2 + 1] * inv_n - ma * ma, vb = gp[g1 * 2 + 1] * inv_n - mb * mb;
                if (va < 0.0) va = 0.0;
                if (vb < 0.0) vb = 0.0;
                m0 = (float)ma; rs0 = 1.0f / sqrtf((float)va + 1e-5f);
                m1 = (float)mb; rs1 = 1.0f / sqrtf((float)vb + 1e-5f);
            }
            const float *tbp = nullptr;
            if (MODE == 2 && g.tb) tbp = g.tb + (size_t)(g.tb_per_sample ? b : 0) * kC;
            float4 tb0 = make_float4(0.f, 0.f, 0.f, 0.f), tb1 = tb0;
            if (tbp) { tb0 = *reinterpret_cast<const float4 *>(tbp + c0); tb1 = *reinterpret_cast<const float4 *>(tbp + c1); }
            if (lt > 0) mbar_wait_parked(bar_aempty, (lt - 1) & 1, 200);     // the previous tile's MMAs have read A
            // ---- build A: one warp per row (rows ew, ew + 8, ... of the tile), RB rows per batch; the loads of batch
            //      bt + 1 are in flight while batch bt is transformed (the L2 round trip is paid once, not per batch)
            constexpr int RB = MODE == 1 ? 2 : 4;
            constexpr int NBT = ((AROWS + 7) / 8 + RB - 1) / RB;
            float4 v0[2][RB], v1[2][RB], r0[2][RB], r1[2][RB];
            float mk[2][RB];
            auto load_batch = [&](int bt, int set) {
#pragma unroll
                for (int j = 0; j < RB; ++j) {
                    const int i = ew + 8 * (RB * bt + j);
                    const int t = t0 + i - (MODE == 2 ? 1 : 0);
                    const bool ok = i < AROWS && t >= 0 && t < g.T;
                    const size_t row = (size_t)b * g.T + (ok ? t : 0);
                    mk[set][j] = ok ? g.mask[(size_t)b * g.mask_T + t] : 0.f;
                    v0[set][j] = ok ? ld_chain4(g.h + row * kC + c0) : make_float4(0.f, 0.f, 0.f, 0.f);
                    v1[set][j] = ok ? ld_chain4(g.h + row * kC + c1) : make_float4(0.f, 0.f, 0.f, 0.f);
                    if (MODE == 1) {
                        r0[set][j] = ok ? ld_chain4(g.r + row * kC + c0) : make_float4(0.f, 0.f, 0.f, 0.f);
                        r1[set][j] = ok ? ld_chain4(g.r + row * kC + c1) : make_float4(0.f, 0.f, 0.f, 0.f);
                    }
                }
            };
            // (a rolled loop on purpose: fully unrolled, the build alone was ~100 KB of SASS and the warps stalled on
            //  instruction fetch -- ncu: stall_no_inst on top)
            load_batch(0, 0);
#pragma unroll 1
            for (int bt = 0; bt < NBT; ++bt) {
                constexpr int set = 0;
                if (bt + 1 < NBT) load_batch(bt + 1, 1);
#pragma unroll
                for (int j = 0; j < RB; ++j) {
                    const int i = ew + 8 * (RB * bt + j);
                    if (i < AROWS) {
                    const int t = t0 + i - (MODE == 2 ? 1 : 0);
                    const bool ok = t >= 0 && t < g.T;
                    const float k = mk[set][j];
                    const float4 h0 = v0[set][j], h1 = v1[set][j];
                    float4 a, c;
                    a.x = mish_fast_f((h0.x - m0) * rs0 * ga0.x + be0.x) * k; a.y = mish_fast_f((h0.y - m0) * rs0 * ga0.y + be0.y) * k;
                    a.z = mish_fast_f((h0.z - m0) * rs0 * ga0.z + be0.z) * k; a.w = mish_fast_f((h0.w - m0) * rs0 * ga0.w + be0.w) * k;
                    c.x = mish_fast_f((h1.x - m1) * rs1 * ga1.x + be1.x) * k; c.y = mish_fast_f((h1.y - m1) * rs1 * ga1.y + be1.y) * k;
                    c.z = mish_fast_f((h1.z - m1) * rs1 * ga1.z + be1.z) * k; c.w = mish_fast_f((h1.w - m1) * rs1 * ga1.w + be1.w) * k;
                    uint32_t p0, p1, p2, p3;
                    if (MODE == 1) {
                        a.x += r0[set][j].x; a.y += r0[set][j].y; a.z += r0[set][j].z; a.w += r0[set][j].w;
                        c.x += r1[set][j].x; c.y += r1[set][j].y; c.z += r1[set][j].z; c.w += r1[set][j].w;
                        if (ok) {
                            *reinterpret_cast<float4 *>(g.x1 + ((size_t)b * g.T + t) * kC + c0) = a;
                            *reinterpret_cast<float4 *>(g.x1 + ((size_t)b * g.T + t) * kC + c1) = c;
                        }
                        float sum = (a.x + a.y) + (a.z + a.w) + (c.x + c.y) + (c.z + c.w);
#pragma unroll
                        for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
                        const float mean = sum * (1.0f / kC);
                        float d[8] = {a.x - mean, a.y - mean, a.z - mean, a.w - mean, c.x - mean, c.y - mean, c.z - mean, c.w - mean};
                        float qq = 0.f;
#pragma unroll
                        for (int e = 0; e < 8; ++e) qq = fmaf(d[e], d[e], qq);
#pragma unroll
                        for (int o = 16; o > 0; o >>= 1) qq += __shfl_xor_sync(0xffffffffu, qq, o);
                        const float rstd = rsqrtf(qq * (1.0f / kC) + 1e-5f);
                        p0 = pack16x2(d[0] * rstd * lg0.x + lb0.x, d[1] * rstd * lg0.y + lb0.y, F16);
                        p1 = pack16x2(d[2] * rstd * lg0.z + lb0.z, d[3] * rstd * lg0.w + lb0.w, F16);
                        p2 = pack16x2(d[4] * rstd * lg1.x + lb1.x, d[5] * rstd * lg1.y + lb1.y, F16);
                        p3 = pack16x2(d[6] * rstd * lg1.z + lb1.z, d[7] * rstd * lg1.w + lb1.w, F16);
                    } else {
                        p0 = pack16x2((a.x + tb0.x) * k, (a.y + tb0.y) * k, F16); p1 = pack16x2((a.z + tb0.z) * k, (a.w + tb0.w) * k, F16);
                        p2 = pack16x2((c.x + tb1.x) * k, (c.y + tb1.y) * k, F16); p3 = pack16x2((c.z + tb1.z) * k, (c.w + tb1.w) * k, F16);
                    }
                    if (!ok) p0 = p1 = p2 = p3 = 0u;          // rows outside the sample are the conv's zero padding
                    const uint32_t dst = a_dst0 + (uint32_t)i * 128u + ((a_piece ^ (uint32_t)(i & 7)) << 4);
                    asm volatile("st.shared.v2.b32 [%0], {%1, %2};" ::"r"(dst), "r"(p0), "r"(p1) : "memory");
                    asm volatile("st.shared.v2.b32 [%0], {%1, %2};" ::"r"(dst + 2u * kAbAKb), "r"(p2), "r"(p3) : "memory");
                    }
                }
#pragma unroll
                for (int j = 0; j < RB; ++j) {               // rotate the prefetched batch in
                    v0[0][j] = v0[1][j]; v1[0][j] = v1[1][j]; mk[0][j] = mk[1][j];
                    if (MODE == 1) { r0[0][j] = r0[1][j]; r1[0][j] = r1[1][j]; }
                }
            }
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            __syncwarp();
            if (lane == 0) mbar_arrive(bar_afull);

            // ---- epilogues of the tile's N tiles: thread = row (TMEM lane), 2 KB staging tiles with 64-byte rows
            //      (32 16-bit or 16 fp32 columns, 64B swizzle), TMA stores; rows beyond T are clipped by the map
            const int trow0 = t0 + 32 * q;
            const bool valid = trow0 + lane < g.T;
            for (int nt = 0; nt < n_nt; ++nt, ++ac) {
                const uint32_t acc = ac & 1;
                const uint32_t tacc = tmem_d + acc * kGN + ((uint32_t)(32 * q) << 16) + (uint32_t)(half * 128);
                mbar_wait_parked(bar_tfull + 8u * acc, (ac >> 1) & 1, 200);
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                const float *bp = sbias + nt * kGN + half * 128;
                const uint32_t srow = (uint32_t)lane * 64u, sx = (uint32_t)((lane >> 1) & 3);
                if (MODE == 1) {
                    uint32_t v[2][32];
                    tmem_ld32_issue(tacc, v[0]);
#pragma unroll
                    for (int jj = 0; jj < 4; ++jj) {
                        tmem_ld_wait();
                        if (jj + 1 < 4) tmem_ld32_issue(tacc + 32 * (jj + 1), v[(jj + 1) & 1]);
                        else {
                            asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
                            __syncwarp();
                            if (lane == 0) mbar_arrive(bar_tempty + 8u * acc);
                        }
                        const float4 *b4 = reinterpret_cast<const float4 *>(bp + 32 * jj);
                        float x[32];
#pragma unroll
                        for (int i = 0; i < 8; ++i) {
                            const float4 bb = b4[i];
                            x[4 * i] = __uint_as_float(v[jj & 1][4 * i]) + bb.x; x[4 * i + 1] = __uint_as_float(v[jj & 1][4 * i + 1]) + bb.y;
                            x[4 * i + 2] = __uint_as_float(v[jj & 1][4 * i + 2]) + bb.z; x[4 * i + 3] = __uint_as_float(v[jj & 1][4 * i + 3]) + bb.w;
                        }
                        if (lane == 0) tma_store_wait_read<1>();
                        __syncwarp();
                        const uint32_t sbuf = stg + (n_st & 1u) * 2048u;
#pragma unroll
                        for (int i = 0; i < 4; ++i)
                            asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(sbuf + srow + (((uint32_t)i ^ sx) << 4)),
                                         "r"(pack16x2(x[8 * i], x[8 * i + 1], F16)), "r"(pack16x2(x[8 * i + 2], x[8 * i + 3], F16)),
                                         "r"(pack16x2(x[8 * i + 4], x[8 * i + 5], F16)), "r"(pack16x2(x[8 * i + 6], x[8 * i + 7], F16)) : "memory");
                        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                        __syncwarp();
                        if (lane == 0) { tma_store_3d(&tmO, sbuf, nt * kGN + half * 128 + 32 * jj, trow0, b); tma_store_commit(); }
                        ++n_st;
                    }
                } else {
                    uint32_t v[2][16];
                    tmem_ld16_issue(tacc, v[0]);
                    float s1 = 0.f, s2 = 0.f;
#pragma unroll
                    for (int jj = 0; jj < 8; ++jj) {
                        tmem_ld_wait();
                        if (jj + 1 < 8) tmem_ld16_issue(tacc + 16 * (jj + 1), v[(jj + 1) & 1]);
                        else {
                            asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
                            __syncwarp();
                            if (lane == 0) mbar_arrive(bar_tempty + 8u * acc);
                        }
                        const float4 *b4 = reinterpret_cast<const float4 *>(bp + 16 * jj);
                        float x[16];
#pragma unroll
                        for (int i = 0; i < 4; ++i) {
                            const float4 bb = b4[i];
                            x[4 * i] = __uint_as_float(v[jj & 1][4 * i]) + bb.x; x[4 * i + 1] = __uint_as_float(v[jj & 1][4 * i + 1]) + bb.y;
                            x[4 * i + 2] = __uint_as_float(v[jj & 1][4 * i + 2]) + bb.z; x[4 * i + 3] = __uint_as_float(v[jj & 1][4 * i + 3]) + bb.w;
                        }
                        // GroupNorm sums of the output: two 16-column chunks make one 32-channel group
                        {
                            float a1 = 0.f, a2 = 0.f, q1 = 0.f, q2 = 0.f;
#pragma unroll
                            for (int i = 0; i < 16; i += 2) { a1 += x[i]; a2 += x[i + 1]; q1 = fmaf(x[i], x[i], q1); q2 = fmaf(x[i + 1], x[i + 1], q2); }
                            if (valid) { s1 += a1 + a2; s2 += q1 + q2; }
                        }
                        if (jj & 1) {
#pragma unroll
                            for (int o = 16; o > 0; o >>= 1) {
                                s1 += __shfl_xor_sync(0xffffffffu, s1, o);
                                s2 += __shfl_xor_sync(0xffffffffu, s2, o);
                            }
                            if (lane == 0) {
                                double *gp = g.gstats + ((size_t)b * kGroups + (nt * kGN + half * 128 + 16 * jj) / kGroupCh) * 2;
                                atomicAdd(gp, (double)s1);
                                atomicAdd(gp + 1, (double)s2);
                            }
                            s1 = 0.f; s2 = 0.f;
                        }
                        if (lane == 0) tma_store_wait_read<1>();
                        __syncwarp();
                        const uint32_t sbuf = stg + (n_st & 1u) * 2048u;
#pragma unroll
                        for (int i = 0; i < 4; ++i)
                            sts128(sbuf + srow + (((uint32_t)i ^ sx) << 4), make_float4(x[4 * i], x[4 * i + 1], x[4 * i + 2], x[4 * i + 3]));
                        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                        __syncwarp();
                        if (lane == 0) { tma_store_3d(&tmO, sbuf, nt * kGN + half * 128 + 16 * jj, trow0, b); tma_store_commit(); }
                        ++n_st;
                    }
                }
            }
        }
        if (lane == 0) tma_store_wait_read<0>();              // shared memory must outlive the stores' reads
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 1) {
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_d), "n"(512) : "memory");
    }
}

}  // namespace mb200
