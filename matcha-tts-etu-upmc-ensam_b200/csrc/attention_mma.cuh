// attention_mma.cuh -- self-attention of the decoder's transformer blocks on the tensor cores
// (mma.sync m16n8k16, 16-bit operands, fp32 accumulate and fp32 online softmax).
//
// softmax_j(q_i . k_j / 8 + mask01[b][j]) v_j per head (diffusers Attention + SDPA, additive float
// key mask).  qkv: 16-bit [B*T][768] = [q | k | v], head h at columns h*64; out: 16-bit
// [B*T][256].  One CTA = 4 warps x 16 queries; K/V stream through double-buffered, XOR-swizzled
// shared-memory tiles filled by cp.async; fragments come from ldmatrix.
// Round-1 note: this is the legacy warp-level tensor path (HMMA); moving QK^T / PV to tcgen05 is
// planned.
#pragma once
#include "decoder_kernels.cuh"

namespace mb200 {

constexpr int kFaBQ = 64, kFaBK = 64;
constexpr int kFaSmemBytes = (kFaBQ + 4 * kFaBK) * 128;     // Q + 2x(K,V) tiles of 64 rows x 128 B

__device__ __forceinline__ void cp_async_16(uint32_t dst, const void *src, bool valid) {
    const int sz = valid ? 16 : 0;      // zero-fill rows beyond T
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(sz) : "memory");
}
__device__ __forceinline__ void ldsm_x4(uint32_t addr, uint32_t &r0, uint32_t &r1, uint32_t &r2, uint32_t &r3) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0, %1, %2, %3}, [%4];"
                 : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(addr));
}
__device__ __forceinline__ void ldsm_x4_t(uint32_t addr, uint32_t &r0, uint32_t &r1, uint32_t &r2, uint32_t &r3) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0, %1, %2, %3}, [%4];"
                 : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(addr));
}
template <bool F16>
__device__ __forceinline__ void mma16816(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
    if (F16)
        asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0, %1, %2, %3}, {%4, %5, %6, %7}, {%8, %9}, {%0, %1, %2, %3};"
                     : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3]) : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
    else
        asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0, %1, %2, %3}, {%4, %5, %6, %7}, {%8, %9}, {%0, %1, %2, %3};"
                     : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3]) : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

template <bool F16>
__global__ void __launch_bounds__(128)
attention_mma_kernel(const __nv_bfloat16 *__restrict__ qkv, const float *__restrict__ mask, int mask_T,
                     __nv_bfloat16 *__restrict__ out, int T) {
    pdl_entry();
    extern __shared__ __align__(128) uint8_t fa_smem[];
    const uint32_t sQ = smem_u32(fa_smem);
    const uint32_t sKV = sQ + kFaBQ * 128;                  // [buf][K|V][64 rows][128 B]
    const int b = blockIdx.z, h = blockIdx.y, q0 = blockIdx.x * kFaBQ;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const __nv_bfloat16 *base = qkv + (size_t)b * T * (3 * kC) + h * kDHead;

    auto load_tile = [&](uint32_t dst, int col_off, int row0) {       // 64 rows x 8 chunks of 16 B
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int idx = tid + i * 128, r = idx >> 3, ch = idx & 7;
            const int t = row0 + r;
            const bool ok = t < T;
            const __nv_bfloat16 *src = base + (size_t)(ok ? t : 0) * (3 * kC) + col_off + ch * 8;
            cp_async_16(dst + r * 128 + ((ch ^ (r & 7)) << 4), src, ok);
        }
    };
    load_tile(sQ, 0, q0);
    load_tile(sKV, kC, 0);
    load_tile(sKV + kFaBK * 128, 2 * kC, 0);
    asm volatile("cp.async.commit_group;" ::: "memory");

    float o[8][4];
#pragma unroll
    for (int i = 0; i < 8; ++i) { o[i][0] = o[i][1] = o[i][2] = o[i][3] = 0.f; }
    float m_run[2] = {-INFINITY, -INFINITY}, l_run[2] = {0.f, 0.f};
    uint32_t qa[4][4];
    const float kLog2e = 1.4426950408889634f;
    const int nkb = (T + kFaBK - 1) / kFaBK;

    for (int kb = 0; kb < nkb; ++kb) {
        const uint32_t sK = sKV + (kb & 1) * (2 * kFaBK * 128), sV = sK + kFaBK * 128;
        if (kb + 1 < nkb) {                                    // prefetch the next K/V block
            const uint32_t nK = sKV + ((kb + 1) & 1) * (2 * kFaBK * 128);
            load_tile(nK, kC, (kb + 1) * kFaBK);
            load_tile(nK + kFaBK * 128, 2 * kC, (kb + 1) * kFaBK);
            asm volatile("cp.async.commit_group;" ::: "memory");
            asm volatile("cp.async.wait_group 1;" ::: "memory");
        } else {
            asm volatile("cp.async.wait_group 0;" ::: "memory");
        }
        __syncthreads();
        if (kb == 0) {
#pragma unroll
            for (int ks = 0; ks < 4; ++ks) {
                const int r = warp * 16 + (lane & 7) + ((lane >> 3) & 1) * 8, ch = 2 * ks + (lane >> 4);
                ldsm_x4(sQ + r * 128 + ((ch ^ (r & 7)) << 4), qa[ks][0], qa[ks][1], qa[ks][2], qa[ks][3]);
            }
        }
        // S = Q K^T  (16 queries x 64 keys per warp)
        float s[8][4];
#pragma unroll
        for (int i = 0; i < 8; ++i) { s[i][0] = s[i][1] = s[i][2] = s[i][3] = 0.f; }
#pragma unroll
        for (int ks = 0; ks < 4; ++ks) {
#pragma unroll
            for (int np = 0; np < 4; ++np) {
                const int r = np * 16 + (lane & 7) + (lane >> 4) * 8, ch = 2 * ks + ((lane >> 3) & 1);
                uint32_t b0, b1, b2, b3;
                ldsm_x4(sK + r * 128 + ((ch ^ (r & 7)) << 4), b0, b1, b2, b3);
                mma16816<F16>(s[2 * np], qa[ks], b0, b1);
                mma16816<F16>(s[2 * np + 1], qa[ks], b2, b3);
            }
        }
        // logits = s/8 + mask01[key]; keys beyond T are excluded
        float mx[2] = {-INFINITY, -INFINITY};
#pragma unroll
        for (int nt = 0; nt < 8; ++nt) {
            const int key = kb * kFaBK + nt * 8 + 2 * (lane & 3);
            const float m0 = key < T ? mask[(size_t)b * mask_T + key] : 0.f;
            const float m1 = key + 1 < T ? mask[(size_t)b * mask_T + key + 1] : 0.f;
            s[nt][0] = key < T ? fmaf(s[nt][0], 0.125f, m0) : -INFINITY;
            s[nt][1] = key + 1 < T ? fmaf(s[nt][1], 0.125f, m1) : -INFINITY;
            s[nt][2] = key < T ? fmaf(s[nt][2], 0.125f, m0) : -INFINITY;
            s[nt][3] = key + 1 < T ? fmaf(s[nt][3], 0.125f, m1) : -INFINITY;
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
            pa[nt >> 1][(nt & 1) * 2 + 0] = pack16x2(p0, p1, F16 ? 1 : 0);
            pa[nt >> 1][(nt & 1) * 2 + 1] = pack16x2(p2, p3, F16 ? 1 : 0);
        }
#pragma unroll
        for (int i = 0; i < 2; ++i) {
            ps[i] += __shfl_xor_sync(0xffffffffu, ps[i], 1);
            ps[i] += __shfl_xor_sync(0xffffffffu, ps[i], 2);
            l_run[i] = l_run[i] * corr[i] + ps[i];
        }
#pragma unroll
        for (int dt = 0; dt < 8; ++dt) { o[dt][0] *= corr[0]; o[dt][1] *= corr[0]; o[dt][2] *= corr[1]; o[dt][3] *= corr[1]; }
        // O += P V
#pragma unroll
        for (int j = 0; j < 4; ++j) {
#pragma unroll
            for (int dp = 0; dp < 4; ++dp) {
                const int r = j * 16 + (lane & 7) + ((lane >> 3) & 1) * 8, ch = 2 * dp + (lane >> 4);
                uint32_t b0, b1, b2, b3;
                ldsm_x4_t(sV + r * 128 + ((ch ^ (r & 7)) << 4), b0, b1, b2, b3);
                mma16816<F16>(o[2 * dp], pa[j], b0, b1);
                mma16816<F16>(o[2 * dp + 1], pa[j], b2, b3);
            }
        }
        __syncthreads();                                         // the buffer is refilled two iterations later
    }
    const int g = lane >> 2, qd = lane & 3;
#pragma unroll
    for (int i = 0; i < 2; ++i) {
        const int t = q0 + warp * 16 + g + 8 * i;
        if (t < T) {
            const float inv = 1.0f / l_run[i];
            uint32_t *po = reinterpret_cast<uint32_t *>(out + ((size_t)b * T + t) * kC + h * kDHead + 2 * qd);
#pragma unroll
            for (int dt = 0; dt < 8; ++dt) po[dt * 4] = pack16x2(o[dt][2 * i] * inv, o[dt][2 * i + 1] * inv, F16 ? 1 : 0);
        }
    }
}

}  // namespace mb200
