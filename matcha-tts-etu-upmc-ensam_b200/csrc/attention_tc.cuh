// attention_tc.cuh -- self-attention of the decoder's transformer blocks on tcgen05 / TMEM.
//
// softmax_j(q_i . k_j / 8 + mask01[b][j]) v_j per head (diffusers Attention + SDPA with an additive float
// key mask; transformer.py:251-259,320-325).  qkv: 16-bit [B][T][768] = [q | k | v], head h at
// columns h*64; out: 16-bit [B][T][256].
//
// One CTA = one (sample, head, 128-query block).  Warp roles:
//   warp 0      TMA producer: the Q tile once, then (K, V) tiles of 128 keys through a 3-stage ring.  One
//               3-D tensor map over [B][T][768] serves all three; rows beyond T come back as zeros.
//   warp 1      MMA issuer: S_j = Q K_j^T (M128 x N128 x K64, fp32 in TMEM, double-buffered) and
//               O_j = P_j V_j (M128 x N64 x K128; V is consumed MN-major straight from its [key][d] tile).
//   warps 2..9  softmax: thread = one query row x 64 keys of the block (two warps share a TMEM lane quarter
//               and split the keys; the row max is exchanged through shared memory).  exp2 in fp32, P written
//               16-bit into the 128B-swizzled K-major A-operand layout, O_j read back from TMEM and
//               accumulated in registers with the usual online-softmax rescale, so TMEM is never rewritten.
// The block is MUFU-bound by design: 128 x 128 exp2 per key block = 1024 cycles, the two MMAs take 512.
#pragma once
#include "gemm_bf16.cuh"

namespace mb200 {

constexpr int kTaBQ = 128, kTaBK = 128, kTaStages = 3;
constexpr int kTaThreads = 320;
constexpr int kTaMaxT = 2048;                               // mask-bias table in shared memory
constexpr int kTaTile = 128 * 128;                          // bytes of one 128-row x 64-halfword tile
constexpr int kTaOffKV = kTaTile;                           // Q | (K,V) x stages | P x 2 (two 64-key panels each)
constexpr int kTaOffP = kTaOffKV + kTaStages * 2 * kTaTile;
constexpr int kTaOffMask = kTaOffP + 2 * 2 * kTaTile;
constexpr int kTaOffXch = kTaOffMask + kTaMaxT * 4;         // [2][2][128] floats
constexpr int kTaOffBar = kTaOffXch + 2 * 2 * 128 * 4;
constexpr int kTaSmemBytes = kTaOffBar + 256 + 1024;

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
__device__ __forceinline__ void named_bar_sync(int id, int n) { asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(n) : "memory"); }

template <bool F16>
__global__ void __launch_bounds__(kTaThreads, 1)
attention_tc_kernel(const __grid_constant__ CUtensorMap tmQKV, const AttnTcArgs g) {
    extern __shared__ uint8_t tasm_raw[];
    const uint32_t base = (smem_u32(tasm_raw) + 1023u) & ~1023u;
    uint8_t *gbase = tasm_raw + (base - smem_u32(tasm_raw));
    float *maskb = reinterpret_cast<float *>(gbase + kTaOffMask);
    float *xch = reinterpret_cast<float *>(gbase + kTaOffXch);
    const uint32_t bar0 = base + kTaOffBar;
    const uint32_t bar_q = bar0;                             // Q landed
    const uint32_t bar_kvf = bar0 + 8;                       // [3] K,V landed
    const uint32_t bar_kve = bar_kvf + 8 * kTaStages;        // [3] stage consumed (tcgen05.commit)
    const uint32_t bar_sf = bar_kve + 8 * kTaStages;         // [2] S ready
    const uint32_t bar_se = bar_sf + 16;                     // [2] S read by the softmax warps
    const uint32_t bar_pf = bar_se + 16;                     // [2] P written
    const uint32_t bar_of = bar_pf + 16;                     // [2] O_j ready
    const uint32_t bar_oe = bar_of + 16;                     // [2] O_j read
    const uint32_t tmem_slot = bar_oe + 16;
    volatile uint32_t *tmem_slot_p = reinterpret_cast<volatile uint32_t *>(gbase + kTaOffBar + 8 + 16 * kTaStages + 80);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int T = g.T;
    // heavy tiles (full query blocks) first, the ragged last block of every (b, h) at the end of the grid
    const int n_full = T / kTaBQ, heavy = n_full * kHeads * g.B;
    int qb, bh;
    if ((int)blockIdx.x < heavy) { qb = blockIdx.x % n_full; bh = blockIdx.x / n_full; }
    else { qb = n_full; bh = blockIdx.x - heavy; }
    const int b = bh / kHeads, h = bh - b * kHeads, q0 = qb * kTaBQ;
    const int n_blk = (T + kTaBK - 1) / kTaBK;
    const int rows_valid = T - q0 < kTaBQ ? T - q0 : kTaBQ;
    const uint32_t n_sm_warps = 2u * (uint32_t)((rows_valid + 31) / 32);     // softmax warps that have rows

    if (threadIdx.x == 0) {
        mbar_init(bar_q, 1u);
        for (int s = 0; s < kTaStages; ++s) { mbar_init(bar_kvf + 8u * s, 1u); mbar_init(bar_kve + 8u * s, 1u); }
        for (int a = 0; a < 2; ++a) {
            mbar_init(bar_sf + 8u * a, 1u); mbar_init(bar_se + 8u * a, n_sm_warps);
            mbar_init(bar_pf + 8u * a, n_sm_warps);
            mbar_init(bar_of + 8u * a, 1u); mbar_init(bar_oe + 8u * a, n_sm_warps);
        }
        mbar_fence_init();
        tma_prefetch_desc(&tmQKV);
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "n"(512) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    pdl_launch_dependents();
    pdl_wait();
    // additive key mask in the exp2 domain; keys beyond T never win the max and contribute exp2(-inf) = 0
    for (int k = threadIdx.x; k < n_blk * kTaBK; k += kTaThreads)
        maskb[k] = k < T ? g.mask[(size_t)b * g.mask_T + k] * 1.4426950408889634f : -INFINITY;
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem = *tmem_slot_p;
    const uint32_t tmem_s = tmem, tmem_o = tmem + 256;       // S: 2 x 128 columns, O: 2 x 64 columns

    if (warp == 0) {
        // ============================ TMA producer ============================
        if (lane == 0) {
            mbar_arrive_expect_tx(bar_q, kTaTile);
            tma_load_3d(base, &tmQKV, bar_q, h * kDHead, q0, b);
            for (int j = 0; j < n_blk; ++j) {
                const uint32_t s = j % kTaStages;
                mbar_wait_parked(bar_kve + 8u * s, ((j / kTaStages) & 1) ^ 1, 1000);
                const uint32_t sk = base + kTaOffKV + s * 2 * kTaTile, sv = sk + kTaTile;
                mbar_arrive_expect_tx(bar_kvf + 8u * s, 2 * kTaTile);
                tma_load_3d(sk, &tmQKV, bar_kvf + 8u * s, kC + h * kDHead, j * kTaBK, b);
                tma_load_3d(sv, &tmQKV, bar_kvf + 8u * s, 2 * kC + h * kDHead, j * kTaBK, b);
            }
        }
    } else if (warp == 1) {
        // ============================ MMA issuer ============================
        if (lane == 0) {
            auto issue_pv = [&](int i) {
                const int kcnt = T - i * kTaBK < kTaBK ? T - i * kTaBK : kTaBK, n16 = (kcnt + 15) & ~15;
                const uint32_t a = i & 1, s = i % kTaStages;
                mbar_wait_parked(bar_pf + 8u * a, (i >> 1) & 1, 1000);
                mbar_wait_parked(bar_oe + 8u * a, ((i >> 1) & 1) ^ 1, 1000);
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                const uint64_t dp = umma_desc_sw128(base + kTaOffP + a * 2 * kTaTile);
                const uint64_t dv = umma_desc_sw128(base + kTaOffKV + s * 2 * kTaTile + kTaTile);
                const uint32_t idesc = ta_idesc(F16, kDHead, true);
                for (int ks = 0; ks < n16 / 16; ++ks)      // 16 keys per step: P panel ks/4 (+32 B per step), V +2048 B
                    umma_bf16(tmem_o + a * 64, dp + (uint64_t)((ks >> 2) * (kTaTile >> 4) + (ks & 3) * 2),
                              dv + (uint64_t)(ks * 128), idesc, ks != 0 ? 1u : 0u);
                umma_commit(bar_of + 8u * a);
                umma_commit(bar_kve + 8u * s);              // K_i and V_i are no longer needed
            };
            mbar_wait_parked(bar_q, 0, 1000);
            const uint64_t dq = umma_desc_sw128(base);
            for (int j = 0; j < n_blk; ++j) {
                const int kcnt = T - j * kTaBK < kTaBK ? T - j * kTaBK : kTaBK, n16 = (kcnt + 15) & ~15;
                const uint32_t a = j & 1, s = j % kTaStages;
                mbar_wait_parked(bar_kvf + 8u * s, (j / kTaStages) & 1, 1000);
                mbar_wait_parked(bar_se + 8u * a, ((j >> 1) & 1) ^ 1, 1000);
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                const uint64_t dk = umma_desc_sw128(base + kTaOffKV + s * 2 * kTaTile);
                const uint32_t idesc = ta_idesc(F16, n16, false);
#pragma unroll
                for (int k = 0; k < kDHead / 16; ++k)
                    umma_bf16(tmem_s + a * 128, dq + 2u * k, dk + 2u * k, idesc, k != 0 ? 1u : 0u);
                umma_commit(bar_sf + 8u * a);
                if (j > 0) issue_pv(j - 1);
            }
            issue_pv(n_blk - 1);
        }
    } else {
        // ============================ softmax ============================
        const int q = warp & 3, hf = (warp - 2) >> 2;        // TMEM lane quarter, key half of the block
        const int row = 32 * q + lane;
        if (32 * q < rows_valid) {
            const uint32_t lane_off = (uint32_t)(32 * q) << 16;
            const float c = 0.125f * 1.4426950408889634f;     // scale = dim_head^-0.5 (diffusers Attention), log2 e
            float m_run = -INFINITY, l_run = 0.f;
            float o[32];
#pragma unroll
            for (int i = 0; i < 32; ++i) o[i] = 0.f;
            for (int j = 0; j < n_blk; ++j) {
                const uint32_t a = j & 1;
                const int kvalid = T - j * kTaBK - 64 * hf;    // keys of this thread's half that exist
                float t[64];
                mbar_wait_parked(bar_sf + 8u * a, (j >> 1) & 1, 200);
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                {
                    uint32_t v0[32], v1[32];
                    tmem_ld32_issue(tmem_s + a * 128 + 64 * hf + lane_off, v0);
                    tmem_ld32_issue(tmem_s + a * 128 + 64 * hf + 32 + lane_off, v1);
                    tmem_ld_wait();
                    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
                    __syncwarp();
                    if (lane == 0) mbar_arrive(bar_se + 8u * a);
                    const float4 *mb4 = reinterpret_cast<const float4 *>(maskb + j * kTaBK + 64 * hf);
                    if (kvalid >= 64) {
#pragma unroll
                        for (int i = 0; i < 8; ++i) {
                            const float4 m0 = mb4[i], m1 = mb4[8 + i];
                            t[4 * i] = fmaf(__uint_as_float(v0[4 * i]), c, m0.x); t[4 * i + 1] = fmaf(__uint_as_float(v0[4 * i + 1]), c, m0.y);
                            t[4 * i + 2] = fmaf(__uint_as_float(v0[4 * i + 2]), c, m0.z); t[4 * i + 3] = fmaf(__uint_as_float(v0[4 * i + 3]), c, m0.w);
                            t[32 + 4 * i] = fmaf(__uint_as_float(v1[4 * i]), c, m1.x); t[32 + 4 * i + 1] = fmaf(__uint_as_float(v1[4 * i + 1]), c, m1.y);
                            t[32 + 4 * i + 2] = fmaf(__uint_as_float(v1[4 * i + 2]), c, m1.z); t[32 + 4 * i + 3] = fmaf(__uint_as_float(v1[4 * i + 3]), c, m1.w);
                        }
                    } else {                                  // ragged last block: stale TMEM columns must not leak in
#pragma unroll
                        for (int i = 0; i < 32; ++i) {
                            t[i] = i < kvalid ? fmaf(__uint_as_float(v0[i]), c, maskb[j * kTaBK + 64 * hf + i]) : -INFINITY;
                            t[32 + i] = 32 + i < kvalid ? fmaf(__uint_as_float(v1[i]), c, maskb[j * kTaBK + 64 * hf + 32 + i]) : -INFINITY;
                        }
                    }
                }
                float mloc = -INFINITY;
#pragma unroll
                for (int i = 0; i < 64; ++i) mloc = fmaxf(mloc, t[i]);
                xch[(a * 2 + hf) * 128 + row] = mloc;
                named_bar_sync(1 + q, 64);
                const float m_new = fmaxf(m_run, fmaxf(mloc, xch[(a * 2 + (hf ^ 1)) * 128 + row]));
                const float alpha = ex2_approx(m_run - m_new);            // 0 on the first block
                m_run = m_new;
                // P = exp2(t - m) as 16-bit, row `row`, keys 64*hf.. of buffer a: K-major, 128B swizzle
                float lsum = 0.f;
                const uint32_t prow = base + kTaOffP + a * 2 * kTaTile + hf * kTaTile + row * 128;
#pragma unroll
                for (int ch = 0; ch < 8; ++ch) {
                    float p[8];
#pragma unroll
                    for (int i = 0; i < 8; ++i) { p[i] = ex2_approx(t[8 * ch + i] - m_new); lsum += p[i]; }
                    const uint32_t dst = prow + ((ch ^ (row & 7)) << 4);
                    asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(dst), "r"(pack16x2(p[0], p[1], F16)),
                                 "r"(pack16x2(p[2], p[3], F16)), "r"(pack16x2(p[4], p[5], F16)), "r"(pack16x2(p[6], p[7], F16)) : "memory");
                }
                asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                __syncwarp();
                if (lane == 0) mbar_arrive(bar_pf + 8u * a);
                l_run = l_run * alpha + lsum;
                // O_{j-1} (scaled to the previous max) joins the running sum, then both move to the new max
                if (j > 0) {
                    const uint32_t pa = (j - 1) & 1;
                    uint32_t ov[32];
                    mbar_wait_parked(bar_of + 8u * pa, ((j - 1) >> 1) & 1, 200);
                    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                    tmem_ld32(tmem_o + pa * 64 + 32 * hf + lane_off, ov);
                    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
                    __syncwarp();
                    if (lane == 0) mbar_arrive(bar_oe + 8u * pa);
#pragma unroll
                    for (int i = 0; i < 32; ++i) o[i] = (o[i] + __uint_as_float(ov[i])) * alpha;
                }
            }
            {
                const uint32_t pa = (n_blk - 1) & 1;
                uint32_t ov[32];
                mbar_wait_parked(bar_of + 8u * pa, ((n_blk - 1) >> 1) & 1, 200);
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                tmem_ld32(tmem_o + pa * 64 + 32 * hf + lane_off, ov);
                asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
#pragma unroll
                for (int i = 0; i < 32; ++i) o[i] += __uint_as_float(ov[i]);
            }
            // row sum: both halves used the same running max, so the partial sums simply add
            const uint32_t xa = n_blk & 1;
            xch[(xa * 2 + hf) * 128 + row] = l_run;
            named_bar_sync(1 + q, 64);
            const float inv = 1.0f / (l_run + xch[(xa * 2 + (hf ^ 1)) * 128 + row]);
            if (q0 + row < T) {
                uint4 *po = reinterpret_cast<uint4 *>(g.out + ((size_t)b * T + q0 + row) * kC + h * kDHead + 32 * hf);
#pragma unroll
                for (int i = 0; i < 4; ++i)
                    po[i] = make_uint4(pack16x2(o[8 * i] * inv, o[8 * i + 1] * inv, F16), pack16x2(o[8 * i + 2] * inv, o[8 * i + 3] * inv, F16),
                                       pack16x2(o[8 * i + 4] * inv, o[8 * i + 5] * inv, F16), pack16x2(o[8 * i + 6] * inv, o[8 * i + 7] * inv, F16));
            }
        }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 1) {
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "n"(512) : "memory");
    }
}

}  // namespace mb200
