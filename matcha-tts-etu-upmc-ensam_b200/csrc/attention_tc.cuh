// attention_tc.cuh -- self-attention of the decoder's transformer blocks on tcgen05 / TMEM.
//
// softmax_j(q_i . k_j / 8 + mask01[b][j]) v_j per head (diffusers Attention + SDPA with an additive float
// key mask; transformer.py:251-259,320-325).  qkv: 16-bit [B][T][768] = [q | k | v], head h at
// columns h*64; out: 16-bit [B][T][256].
//
// One CTA = one (sample, head, 128-query block); two CTAs share an SM (256 TMEM columns and ~100 KB of shared
// memory each), so that one CTA's exp2 phase (MUFU-bound) overlaps the other's logits phase and its
// prologue.  Full query blocks come first in the grid, the ragged last block of every (sample, head) last.
// Warp roles:
//   warp 0      TMA producer: the Q tile once (lane 0), K tiles of 64 keys (lane 0) and V tiles (lane 1)
//               through two 3-stage rings (a K stage is free as soon as its S is computed, a V stage after its
//               PV).  Two 3-D tensor maps over [B][T][768] (128-row and 64-row boxes); rows beyond T are zeros.
//   warp 1      MMA issuer: S_j = Q K_j^T (M128 x N64 x K64, fp32 in TMEM, double-buffered) and
//               O += P_j V_j (M128 x N64 x K64; V is consumed MN-major straight from its [key][d] tile),
//               accumulated in TMEM over all key blocks.
//   warps 2..9  softmax: two threads per query row (TMEM lane), one for keys 0..31 and one for keys 32..63 of every
//               64-key block.  The two halves are two INDEPENDENT online-softmax streams (own reference point m,
//               own row sum, own O accumulator in TMEM: O_lo / O_hi), merged once at the end as in split-KV
//               attention -- no per-block exchange between the warps of a row, and four softmax warps per scheduler
//               (two CTAs per SM) to overlap one warp's exp2 phase (MUFU) with the others' logits / pack phases.
//               Logits, max and exp2 in fp32 registers, P written 16-bit into the 128B-swizzled K-major A-operand
//               layout.  A stream's reference point m of p = exp(z - m) only moves when a block's max exceeds it by
//               more than 8 octaves (p <= 256 is harmless in 16 bits); then its row sum and its O in TMEM are
//               rescaled (rare after the first block).  Softmax is invariant to m, so the result is the reference's.
// The kernel is MUFU-bound by design: one exp2 per (query, key) at 16 per clock per SM.
#pragma once
#include "gemm_bf16.cuh"

namespace mb200 {

constexpr int kTaBQ = 128, kTaBK = 64, kTaStages = 3;
constexpr int kTaSmWarps = 8;                               // softmax warps: lane quarter (warp & 3) x key half
constexpr int kTaThreads = 32 * (2 + kTaSmWarps);
constexpr int kTaTileQ = kTaBQ * 128;                       // bytes of a 128-row x 64-halfword tile (Q, P)
constexpr int kTaTileK = kTaBK * 128;                       // bytes of a 64-key tile (K, V)
constexpr int kTaOffK = kTaTileQ;                           // Q | K x stages | V x stages | P x 2 | mask slices | barriers
constexpr int kTaOffV = kTaOffK + kTaStages * kTaTileK;
constexpr int kTaOffP = kTaOffV + kTaStages * kTaTileK;
constexpr int kTaOffMk = kTaOffP + 2 * kTaTileQ;            // [8 warps][2][32] floats: the block's key-mask slice of the warp's half
constexpr int kTaOffBar = kTaOffMk + 4 * 2 * 64 * 4;
constexpr int kTaSmemBytes = kTaOffBar + 256 + 1024;
constexpr int kTaTmemCols = 256;                            // S: 2 x 64 columns, O_lo / O_hi: 2 x 64 columns
constexpr float kTaRescale = 5.545177444f;                  // 8 ln 2

struct AttnTcArgs {
    const float *mask; int mask_T;      // [B][mask_T] 0/1, added to the logits
    __nv_bfloat16 *out;                 // 16-bit [B][T][256]
    int B, T;
};

// kind::f16 instruction descriptor: fp32 accumulate, M=128, N=n; b_mn: B operand is MN-major
__device__ __forceinline__ uint32_t ta_idesc(bool f16, int n, bool b_mn) {
    uint32_t d = (1u << 4) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
    if (!f16) d |= (1u << 7) | (1u << 10);
    if (b_mn) d |= (1u << 16);
    return d;
}
__device__ __forceinline__ float ex2_approx(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
template <bool F16>
__global__ void __launch_bounds__(kTaThreads, 2)
attention_tc_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmKV, const AttnTcArgs g) {
    extern __shared__ uint8_t tasm_raw[];
    const uint32_t base = (smem_u32(tasm_raw) + 1023u) & ~1023u;
    uint8_t *gbase = tasm_raw + (base - smem_u32(tasm_raw));
    const uint32_t bar0 = base + kTaOffBar;
    const uint32_t bar_q = bar0;                             // Q landed
    const uint32_t bar_kf = bar0 + 8;                        // [3] K landed
    const uint32_t bar_ke = bar_kf + 8 * kTaStages;          // [3] K stage consumed (S committed)
    const uint32_t bar_vf = bar_ke + 8 * kTaStages;          // [3] V landed
    const uint32_t bar_ve = bar_vf + 8 * kTaStages;          // [3] V stage consumed (PV committed)
    const uint32_t bar_sf = bar_ve + 8 * kTaStages;          // [2] S ready
    const uint32_t bar_se = bar_sf + 16;                     // [2] S read by the softmax warps
    const uint32_t bar_pf = bar_se + 16;                     // [2] P written
    // [2] PV of key block j committed, on barrier j & 1.  Two barriers, because a parity wait is only unambiguous
    // when the barrier is at most one phase away from the awaited one: PV_{j-2} is known to be complete once S_j has
    // been read, whereas a single barrier may still be two PVs behind when the MMA warp is slow to issue.
    const uint32_t bar_pv = bar_pf + 16;
    const uint32_t tmem_slot = bar_pv + 16;
    volatile uint32_t *tmem_slot_p = reinterpret_cast<volatile uint32_t *>(gbase + kTaOffBar + 8 + 32 * kTaStages + 64);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int T = g.T;
    // heavy tiles (full query blocks) first, the ragged last block of every (b, h) at the end of the grid
    const int n_full = T / kTaBQ, heavy = n_full * kHeads * g.B;
    int qb, bh;
    if ((int)blockIdx.x < heavy) { bh = blockIdx.x / n_full; qb = blockIdx.x - bh * n_full; }
    else { qb = n_full; bh = blockIdx.x - heavy; }
    const int b = bh / kHeads, h = bh - b * kHeads, q0 = qb * kTaBQ;
    const int n_blk = (T + kTaBK - 1) / kTaBK;
    const int rows_valid = T - q0 < kTaBQ ? T - q0 : kTaBQ;
    const uint32_t n_sm_warps = 2u * (uint32_t)((rows_valid + 31) / 32);     // softmax warps that have rows

    if (threadIdx.x == 0) {
        mbar_init(bar_q, 1u); mbar_init(bar_pv, 1u); mbar_init(bar_pv + 8u, 1u);
        for (int s = 0; s < kTaStages; ++s) {
            mbar_init(bar_kf + 8u * s, 1u); mbar_init(bar_ke + 8u * s, 1u);
            mbar_init(bar_vf + 8u * s, 1u); mbar_init(bar_ve + 8u * s, 1u);
        }
        for (int a = 0; a < 2; ++a) {
            mbar_init(bar_sf + 8u * a, 1u); mbar_init(bar_se + 8u * a, n_sm_warps); mbar_init(bar_pf + 8u * a, n_sm_warps);
        }
        mbar_fence_init();
        tma_prefetch_desc(&tmQ); tma_prefetch_desc(&tmKV);
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "n"(kTaTmemCols) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem = *tmem_slot_p;
    const uint32_t tmem_s = tmem, tmem_o = tmem + 128;
    pdl_entry();

    if (warp == 0) {
        // ============================ TMA producer ============================
        if (lane == 0) {
            mbar_arrive_expect_tx(bar_q, kTaTileQ);
            tma_load_3d(base, &tmQ, bar_q, h * kDHead, q0, b);
            for (int j = 0; j < n_blk; ++j) {
                const uint32_t s = j % kTaStages;
                mbar_wait_parked(bar_ke + 8u * s, ((j / kTaStages) & 1) ^ 1, 500);
                mbar_arrive_expect_tx(bar_kf + 8u * s, kTaTileK);
                tma_load_3d(base + kTaOffK + s * kTaTileK, &tmKV, bar_kf + 8u * s, kC + h * kDHead, j * kTaBK, b);
            }
        } else if (lane == 1) {
            for (int j = 0; j < n_blk; ++j) {
                const uint32_t s = j % kTaStages;
                mbar_wait_parked(bar_ve + 8u * s, ((j / kTaStages) & 1) ^ 1, 500);
                mbar_arrive_expect_tx(bar_vf + 8u * s, kTaTileK);
                tma_load_3d(base + kTaOffV + s * kTaTileK, &tmKV, bar_vf + 8u * s, 2 * kC + h * kDHead, j * kTaBK, b);
            }
        }
    } else if (warp == 1) {
        // ============================ MMA issuer ============================
        if (lane == 0) {
            const uint32_t idesc_pv = ta_idesc(F16, kDHead, true);
            auto issue_pv = [&](int i) {                    // O += P_i V_i, issued one block late (after S_{i+1})
                const int kcnt = T - i * kTaBK < kTaBK ? T - i * kTaBK : kTaBK, n16 = (kcnt + 15) & ~15;
                const uint32_t a = i & 1, s = i % kTaStages;
                mbar_wait_parked(bar_vf + 8u * s, (i / kTaStages) & 1, 500);
                mbar_wait_parked(bar_pf + 8u * a, (i >> 1) & 1, 200);
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                const uint64_t dp = umma_desc_sw128(base + kTaOffP + a * kTaTileQ);
                const uint64_t dv = umma_desc_sw128(base + kTaOffV + s * kTaTileK);
                for (int ks = 0; ks < n16 / 16; ++ks)      // 16 keys per step: P +32 B along its row, V +2048 B; keys 32.. -> O_hi
                    umma_bf16(tmem_o + 64u * (uint32_t)(ks >> 1), dp + (uint64_t)(2 * ks), dv + (uint64_t)(128 * ks), idesc_pv,
                              (i | (ks & 1)) != 0 ? 1u : 0u);
                umma_commit(bar_pv + 8u * (uint32_t)(i & 1));
                umma_commit(bar_ve + 8u * s);
            };
            mbar_wait_parked(bar_q, 0, 500);
            const uint64_t dq = umma_desc_sw128(base);
            for (int j = 0; j < n_blk; ++j) {
                const int kcnt = T - j * kTaBK < kTaBK ? T - j * kTaBK : kTaBK, n16 = (kcnt + 15) & ~15;
                const uint32_t a = j & 1, s = j % kTaStages;
                mbar_wait_parked(bar_kf + 8u * s, (j / kTaStages) & 1, 500);
                mbar_wait_parked(bar_se + 8u * a, ((j >> 1) & 1) ^ 1, 200);
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                const uint64_t dk = umma_desc_sw128(base + kTaOffK + s * kTaTileK);
                const uint32_t idesc = ta_idesc(F16, n16, false);
#pragma unroll
                for (int k = 0; k < kDHead / 16; ++k)
                    umma_bf16(tmem_s + a * 64, dq + 2u * k, dk + 2u * k, idesc, k != 0 ? 1u : 0u);
                umma_commit(bar_sf + 8u * a);
                umma_commit(bar_ke + 8u * s);
                if (j > 0) issue_pv(j - 1);
            }
            issue_pv(n_blk - 1);
        }
    } else {
        // ============================ softmax ============================
        const int q = warp & 3;                              // TMEM lane quarter of this warp
        const int hf = (warp - 2) >> 2;                      // key half of every 64-key block: 0 = keys 0..31, 1 = keys 32..63
        const int row = 32 * q + lane;
        if (32 * q < rows_valid) {
            const uint32_t lane_off = (uint32_t)(32 * q) << 16;
            const uint32_t tmem_oh = tmem_o + 64u * (uint32_t)hf;               // this stream's O accumulator
            const float kLog2e = 1.4426950408889634f;
            float *mk = reinterpret_cast<float *>(gbase + kTaOffMk) + (warp - 2) * 64;
            const float *mrow = g.mask + (size_t)b * g.mask_T + 32 * hf;
            float m_ref = -INFINITY, l_run = 0.f;             // p = exp(z - m_ref); l_run = sum of p so far
            bool seen = false;                                // this stream has had at least one key
            // this warp's copy of the next block's 32 mask values travels global -> registers (one block ahead) -> shared
            float mk0 = 32 * hf + lane < T ? __ldg(mrow + lane) : 0.f;
            for (int j = 0; j < n_blk; ++j) {
                const uint32_t a = j & 1;
                const int k0 = j * kTaBK;
                int kvalid = T - k0 - 32 * hf;                 // keys of this half-block that exist
                kvalid = kvalid < 0 ? 0 : (kvalid > 32 ? 32 : kvalid);
                float *mkb = mk + a * 32;
                mkb[lane] = mk0;
                __syncwarp();
                if (j + 1 < n_blk) mk0 = k0 + kTaBK + 32 * hf + lane < T ? __ldg(mrow + k0 + kTaBK + lane) : 0.f;
                mbar_wait_parked(bar_sf + 8u * a, (j >> 1) & 1, 100);
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                if (kvalid == 0) {
                    // ragged tail: nothing for this stream in this block (the MMA warp skips its key steps too); the
                    // barrier protocol still needs this warp's arrivals
                    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
                    __syncwarp();
                    if (lane == 0) { mbar_arrive(bar_se + 8u * a); mbar_arrive(bar_pf + 8u * a); }
                    continue;
                }
                float z[32];                                   // logits q.k/8 + mask (natural-log domain), in place
                {
                    uint32_t v0[32];
                    tmem_ld32(tmem_s + a * 64 + 32 * hf + lane_off, v0);
                    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
                    __syncwarp();
                    if (lane == 0) mbar_arrive(bar_se + 8u * a);
                    const float4 *mk4 = reinterpret_cast<const float4 *>(mkb);     // broadcast reads
#pragma unroll
                    for (int i = 0; i < 8; ++i) {
                        const float4 m = mk4[i];
                        z[4 * i] = fmaf(__uint_as_float(v0[4 * i]), 0.125f, m.x); z[4 * i + 1] = fmaf(__uint_as_float(v0[4 * i + 1]), 0.125f, m.y);
                        z[4 * i + 2] = fmaf(__uint_as_float(v0[4 * i + 2]), 0.125f, m.z); z[4 * i + 3] = fmaf(__uint_as_float(v0[4 * i + 3]), 0.125f, m.w);
                    }
                    if (kvalid < 32) {                        // ragged last block: stale TMEM columns must not leak in
#pragma unroll
                        for (int i = 0; i < 32; ++i) z[i] = i < kvalid ? z[i] : -INFINITY;
                    }
                }
                float mx[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    mx[0] = fmaxf(mx[0], z[4 * i]); mx[1] = fmaxf(mx[1], z[4 * i + 1]);
                    mx[2] = fmaxf(mx[2], z[4 * i + 2]); mx[3] = fmaxf(mx[3], z[4 * i + 3]);
                }
                const float m_blk = fmaxf(fmaxf(mx[0], mx[1]), fmaxf(mx[2], mx[3]));
                // move the reference point only when this block would overflow the 8-octave headroom
                const bool move = m_blk > m_ref + kTaRescale;          // always true on the stream's first block (m_ref = -inf)
                if (__any_sync(0xffffffffu, move)) {
                    const float alpha = move ? ex2_approx((m_ref - m_blk) * kLog2e) : 1.0f;   // 0 on the first block
                    if (move) m_ref = m_blk;
                    l_run *= alpha;
                    if (seen) {
                        // O_row *= alpha in TMEM: every PV issued so far has finished, and PV_j cannot start before this
                        // warp has arrived on bar_pf below
                        mbar_wait_parked(bar_pv + 8u * (uint32_t)((j - 1) & 1), ((j - 1) >> 1) & 1, 100);
                        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
#pragma unroll 1
                        for (int hcol = 0; hcol < 4; ++hcol) {
                            uint32_t ov[16];
                            tmem_ld16_issue(tmem_oh + 16 * hcol + lane_off, ov);
                            tmem_ld_wait();
#pragma unroll
                            for (int i = 0; i < 16; ++i) ov[i] = __float_as_uint(__uint_as_float(ov[i]) * alpha);
                            tmem_st16(tmem_oh + 16 * hcol + lane_off, ov);
                        }
                        asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
                    }
                }
                seen = true;
                // P = exp(z - m_ref) as 16-bit, this half of row `row` of buffer a: K-major, 128B swizzle
                const float nm = -m_ref * kLog2e;
                float ls[4] = {0.f, 0.f, 0.f, 0.f};
                const uint32_t prow = base + kTaOffP + a * kTaTileQ + row * 128;
#pragma unroll
                for (int ch = 0; ch < 4; ++ch) {
                    float p[8];
#pragma unroll
                    for (int i = 0; i < 8; ++i) { p[i] = ex2_approx(fmaf(z[8 * ch + i], kLog2e, nm)); ls[i & 3] += p[i]; }
                    const uint32_t dst = prow + (((uint32_t)(4 * hf + ch) ^ (uint32_t)(row & 7)) << 4);
                    asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(dst), "r"(pack16x2(p[0], p[1], F16)),
                                 "r"(pack16x2(p[2], p[3], F16)), "r"(pack16x2(p[4], p[5], F16)), "r"(pack16x2(p[6], p[7], F16)) : "memory");
                }
                asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
                __syncwarp();
                if (lane == 0) mbar_arrive(bar_pf + 8u * a);
                l_run += (ls[0] + ls[1]) + (ls[2] + ls[3]);
            }
            // ---- merge the two streams of a row: (m, l) travel through this warp's (now idle) mask slice, the pair of
            //      warps of a lane quarter meets on a named barrier; each warp then finishes 32 of the 64 output columns
            float2 *xch = reinterpret_cast<float2 *>(mk);
            xch[lane] = make_float2(m_ref, l_run);
            if (q == 0) asm volatile("bar.sync 1, 64;" ::: "memory");
            else if (q == 1) asm volatile("bar.sync 2, 64;" ::: "memory");
            else if (q == 2) asm volatile("bar.sync 3, 64;" ::: "memory");
            else asm volatile("bar.sync 4, 64;" ::: "memory");
            const float2 other = reinterpret_cast<const float2 *>(reinterpret_cast<float *>(gbase + kTaOffMk) + ((warp - 2) ^ 4) * 64)[lane];
            const float m_lo = hf == 0 ? m_ref : other.x, l_lo = hf == 0 ? l_run : other.y;
            const float m_hi = hf == 0 ? other.x : m_ref, l_hi = hf == 0 ? other.y : l_run;
            const float m_all = fmaxf(m_lo, m_hi);                             // the lo stream always has keys: finite
            const float a_lo = ex2_approx((m_lo - m_all) * kLog2e);
            const float a_hi = l_hi > 0.f ? ex2_approx((m_hi - m_all) * kLog2e) : 0.f;
            const float inv = 1.0f / fmaf(l_lo, a_lo, l_hi * a_hi);
            const float w_lo = a_lo * inv, w_hi = a_hi * inv;
            // O is complete when the last PV has been committed
            mbar_wait_parked(bar_pv + 8u * (uint32_t)((n_blk - 1) & 1), ((n_blk - 1) >> 1) & 1, 100);
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
#pragma unroll
            for (int part = 0; part < 2; ++part) {             // 16 output columns at a time
                uint32_t olo[16], ohi[16];
                tmem_ld16_issue(tmem_o + 32 * hf + 16 * part + lane_off, olo);
                tmem_ld16_issue(tmem_o + 64 + 32 * hf + 16 * part + lane_off, ohi);
                tmem_ld_wait();
                if (q0 + row < T) {
                    float o[16];
#pragma unroll
                    for (int i = 0; i < 16; ++i) {
                        const float hi = l_hi > 0.f ? __uint_as_float(ohi[i]) : 0.f;      // never written when T <= 32
                        o[i] = fmaf(__uint_as_float(olo[i]), w_lo, hi * w_hi);
                    }
                    uint4 *po = reinterpret_cast<uint4 *>(g.out + ((size_t)b * T + q0 + row) * kC + h * kDHead + 32 * hf + 16 * part);
#pragma unroll
                    for (int i = 0; i < 2; ++i)
                        po[i] = make_uint4(pack16x2(o[8 * i], o[8 * i + 1], F16), pack16x2(o[8 * i + 2], o[8 * i + 3], F16),
                                           pack16x2(o[8 * i + 4], o[8 * i + 5], F16), pack16x2(o[8 * i + 6], o[8 * i + 7], F16));
                }
            }
        }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 1) {
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "n"(kTaTmemCols) : "memory");
    }
}

}  // namespace mb200
