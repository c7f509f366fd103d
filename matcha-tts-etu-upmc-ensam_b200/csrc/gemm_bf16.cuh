// gemm_bf16.cuh -- tcgen05/TMEM implicit-GEMM for the decoder's convolutions and projections.
//
// D[128 x 256] (fp32, TMEM) = sum over (tap, 64-channel chunk) of A[128 x 64] * W[256 x 64]^T, bf16
// operands staged in shared memory by TMA in the 128B-swizzled K-major layout both TMA and the
// UMMA shared-memory descriptor understand.  Warp roles:
//   warp 0      TMA producer: A boxes from a 3-D map over the bf16 activation [B][T][C] (the conv
//               taps are row offsets; rows outside [0,T) come back as zeros, which is the conv's
//               zero padding, decoder.py:59), W boxes from a 2-D map over [N][K].
//   warp 1      allocates all 512 TMEM columns (two 128x256 fp32 accumulators) and issues tcgen05.mma
//               (one elected lane), releasing stages with tcgen05.commit.
//   warps 2..9  epilogue: tcgen05.ld (lane = output row), transpose through shared memory so that
//               global accesses are row-contiguous, then bias / erf-GELU / residual / row mask and
//               fp32 and/or bf16 stores.
// The CTA is persistent (one per SM, tiles dealt round-robin): while the epilogue warps drain one
// accumulator, the MMA warp already fills the other one with the CTA's next tile.
// Reference arithmetic: nn.Conv1d / nn.Linear of decoder.py:59,78,192,239,248 and the diffusers
// Attention / FeedForward projections (transformer.py:251-259,156,180).
#pragma once
#include <cuda_bf16.h>

#include "decoder_kernels.cuh"

namespace mb200 {

constexpr int kGM = 128, kGN = 256, kGK = 64;          // CTA tile; UMMA 128 x 256 x 16
constexpr int kGStages = 3;                            // TMA ring depth (x 48 KB)
constexpr int kGStageA = kGM * kGK * 2, kGStageB = kGN * kGK * 2;
constexpr int kGStageBytes = kGStageA + kGStageB;      // 48 KB
constexpr int kGEpiWarps = 8;                          // two warps per TMEM lane quarter, four 32-column chunks each
constexpr int kGEpiPitch = 36;                         // floats; 16-byte aligned rows, conflict-free both ways
constexpr int kGEpiWarpBytes = 8192;                   // per epilogue warp: two 4 KB TMA-store staging tiles, or the transpose buffer
constexpr int kGEpiBytes = kGEpiWarps * kGEpiWarpBytes;
constexpr int kGBiasBytes = 4096;                      // bias of every output column (N <= 1024)
constexpr int kGSmemBytes = kGStages * kGStageBytes + kGEpiBytes + kGBiasBytes + 256 + 1024;
constexpr int kGThreads = 32 * (2 + kGEpiWarps);

struct GemmBf16Args {
    int B, Tout;
    int spp, n_mtiles;                 // 32-row slabs per sample and M tiles of the launch (slab_coords below)
    int ntaps, tap_off[4];
    int c0_chunks, c1_chunks;          // 64-channel chunks of source 0 / source 1 (concat)
    const float *bias;                 // [N] or nullptr
    int act;                           // 0 none, 1 erf-GELU
    const float *res;                  // fp32 [B][Tout][N] or nullptr
    float *out_f32;                    // fp32 [B][Tout][N] or nullptr
    __nv_bfloat16 *out_bf16;           // bf16 [B][Tout][N] or nullptr
    const float *omask; int omask_T;   // multiplies the bf16 output rows (x*mask for the next conv)
    int N;
    int out_T, out_stride, out_off;    // output row of tile row t: t*out_stride + out_off (ConvTranspose phases)
    int f16;                           // operands are IEEE fp16 instead of bf16
    int debug;
    // fp16 range guard (decoder.cu kStreamScale): the accumulator is multiplied by acc_scale before the bias (the A rows
    // were stored scaled down), the 16-bit output by o16_scale (it is a residual-stream tensor, stored scaled down)
    float acc_scale, o16_scale;
    // final projection mode (decoder.py:369-371 + flow_matching.py:95): the first fin_n columns are written
    // channels-first, v = (acc + bias) * mask, out = x_in + dt * v (or v when x_in == nullptr)
    int fin_n; const float *fin_xin; float fin_dt; float *fin_out;
    // loss mode of the final projection (flow_matching.py:142-148): *loss_sum += sum over ALL frames t < Tout of
    // (v - (x1 - (1-sigma) z))^2, v the masked prediction (the reference's target is not masked); fin_out may be nullptr
    const float *loss_x1, *loss_z; float loss_oms; double *loss_sum;
    // GroupNorm statistics of the fp32 output (kGStats): [B][8 groups][sum, sum of squares], accumulated with atomics
    double *gstats;
    // Second output block (Resnet1D: block1's conv and res_conv read the same input, decoder.py:81,85): the N tiles at
    // columns >= n_split only walk K-blocks [kb2_begin, kb2_begin + kb2_count) (the centre tap), add bias2, write fp32
    // to out2_f32 (pitch n_split) and take no GroupNorm sums.  n_split == 0: off.  Tiles are then dealt N-major,
    // i.e. all the heavy tiles first.
    int n_split, kb2_begin, kb2_count;
    const float *bias2;
    float *out2_f32;
};

// M tiling.  The rows of a launch are cut into 32-row SLABS: sample b owns slabs [b spp, (b + 1) spp), slab s of a sample covers
// frames [32 s, 32 s + 32), and an M tile of 128 rows is any four consecutive slabs -- it may straddle samples, so a launch has
// ceil(B spp / 4) tiles instead of B ceil(T / 128) (T = 400: 13 slabs = 3.25 tiles per sample instead of 4).  A slab is one
// 32-row TMA box (rows beyond T come back as zeros: the conv padding), one TMEM lane quarter, one epilogue warp's rows; nothing
// in memory changes.  Slabs past the last sample map to rows beyond T of sample 0: loads give zeros, stores are clipped.
__device__ __forceinline__ void slab_coords(int slab, int spp, int B, int T, int &b, int &t0) {
    b = slab / spp;
    t0 = (slab - b * spp) * 32;
    if (b >= B) { b = 0; t0 = T + 64; }
}

__device__ __forceinline__ void tma_load_3d(uint32_t dst, const CUtensorMap *map, uint32_t bar, int c0, int c1, int c2) {
    asm volatile(
        "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
        ::"r"(dst), "l"((uint64_t)map), "r"(bar), "r"(c0), "r"(c1), "r"(c2) : "memory");
}

// K-major, 128B swizzle, 8-row atoms 1024 B apart (SBO), version 1 (sm_100)
__device__ __forceinline__ uint64_t umma_desc_sw128(uint32_t smem_addr) {
    uint64_t d = 0;
    d |= (uint64_t)((smem_addr >> 4) & 0x3fff);
    d |= (uint64_t)(1024 >> 4) << 32;
    d |= (uint64_t)1 << 46;
    d |= (uint64_t)2 << 61;
    return d;
}
// kind::f16, A/B bf16 (format 1) or fp16 (format 0) K-major, fp32 accumulate, M=128, N=256
constexpr uint32_t kIdescF16 = (1u << 4) | ((uint32_t)(kGN >> 3) << 17) | ((uint32_t)(kGM >> 4) << 24);
constexpr uint32_t kIdescBf16 = kIdescF16 | (1u << 7) | (1u << 10);

__device__ __forceinline__ void umma_bf16(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void umma_commit(uint32_t bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void tmem_ld32_issue(uint32_t taddr, uint32_t (&r)[32]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
          "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
          "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
          "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(taddr));
}
__device__ __forceinline__ void tmem_ld16_issue(uint32_t taddr, uint32_t (&r)[16]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
          "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
        : "r"(taddr));
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

__device__ __forceinline__ void tmem_st16(uint32_t taddr, const uint32_t (&r)[16]) {
    asm volatile(
        "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], "
        "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};"
        ::"r"(taddr), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]),
          "r"(r[8]), "r"(r[9]), "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]) : "memory");
}

__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&r)[32]) { tmem_ld32_issue(taddr, r); tmem_ld_wait(); }

// profiling aid: clock64 stamps of CTA 0 (start, first stage landed, first accumulator ready, first tile stored)
__device__ long long g_gemm_dbg[8];

// EPI: 0 plain, 1 erf-GELU, 2 final projection, 3 final projection + squared-error sums (compute_loss); FLAGS: which of residual / fp32 output / 16-bit output exist;
// F16: operand format.  All compile-time: the epilogue is issue-bound, a runtime "if (g.res)" costs slots.
constexpr int kGRes = 1, kGF32 = 2, kGB16 = 4, kGStats = 8;
template <int EPI, int FLAGS, bool F16>
__global__ void __launch_bounds__(kGThreads, 1)
gemm_bf16_kernel(const __grid_constant__ CUtensorMap tmA0, const __grid_constant__ CUtensorMap tmA1,
                 const __grid_constant__ CUtensorMap tmB, const __grid_constant__ CUtensorMap tmO,
                 const __grid_constant__ CUtensorMap tmO2, const GemmBf16Args g) {
    constexpr bool RES = (FLAGS & kGRes) != 0, F32 = (FLAGS & kGF32) != 0, B16 = (FLAGS & kGB16) != 0, STATS = (FLAGS & kGStats) != 0;
    // Variants with exactly one output and no residual stay in the TMEM layout (thread = row) and leave through
    // 128B/64B-swizzled staging tiles and TMA stores; the others transpose through shared memory (below).
    constexpr bool FIN = EPI == 2 || EPI == 3;        // final projection; 3 = the loss variant (flow_matching.py:142-148)
    constexpr bool LOSS = EPI == 3;
    constexpr bool TMA_EPI = !FIN && !RES && (F32 != B16);
    extern __shared__ uint8_t gsm_raw[];
    const uint32_t base = (smem_u32(gsm_raw) + 1023u) & ~1023u;
    uint8_t *gbase = gsm_raw + (base - smem_u32(gsm_raw));
    const uint32_t epi_off = kGStages * kGStageBytes;
    float *const sbias = reinterpret_cast<float *>(gbase + epi_off + kGEpiBytes);
    const uint32_t bar_full = base + epi_off + kGEpiBytes + kGBiasBytes;     // [kGStages]  TMA -> MMA
    const uint32_t bar_empty = bar_full + 8u * kGStages;            // [kGStages]  MMA -> TMA
    const uint32_t bar_tfull = bar_empty + 8u * kGStages;           // [2] accumulator ready   MMA -> epilogue
    const uint32_t bar_tempty = bar_tfull + 16u;                    // [2] accumulator drained epilogue -> MMA
    const uint32_t tmem_slot = bar_tempty + 16u;                    // u32 written by tcgen05.alloc
    volatile uint32_t *tmem_slot_p = reinterpret_cast<volatile uint32_t *>(gbase + (tmem_slot - base));

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int chunks = g.c0_chunks + g.c1_chunks;
    const int n_kb = g.ntaps * chunks;
    const int n_tiles_n = g.N / kGN;
    const int n_mtiles = g.n_mtiles;
    const int n_tiles = n_mtiles * n_tiles_n;
    auto tile_coords = [&](int tile, int &mt, int &n0) {
        if (g.n_split) { const int nt = tile / n_mtiles; mt = tile - nt * n_mtiles; n0 = nt * kGN; }     // N-major
        else { mt = tile / n_tiles_n; n0 = (tile - mt * n_tiles_n) * kGN; }
    };

    if (threadIdx.x == 0) {
        for (int s = 0; s < kGStages; ++s) { mbar_init(bar_full + 8u * s, 1u); mbar_init(bar_empty + 8u * s, 1u); }
        for (int a = 0; a < 2; ++a) { mbar_init(bar_tfull + 8u * a, 1u); mbar_init(bar_tempty + 8u * a, kGEpiWarps); }
        mbar_fence_init();
        tma_prefetch_desc(&tmA0); tma_prefetch_desc(&tmB);
        if (g.c1_chunks) tma_prefetch_desc(&tmA1);
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "n"(512) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem_d = *tmem_slot_p;
    // barriers, TMEM and descriptors are set up while the previous kernel of the stream drains
    pdl_entry();
    const bool dbg = g.debug && blockIdx.x == 0;
    if (dbg && threadIdx.x == 0) g_gemm_dbg[0] = clock64();

    if (warp == 0) {
        // ============================ TMA producer ============================
        if (lane == 0) {
            uint32_t it = 0;
            for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
                int mt, n0;
                tile_coords(tile, mt, n0);
                int sb_[4], st_[4];
#pragma unroll
                for (int i = 0; i < 4; ++i) slab_coords(4 * mt + i, g.spp, g.B, g.Tout, sb_[i], st_[i]);
                const bool light = g.n_split && n0 >= g.n_split;
                const int kb_lo = light ? g.kb2_begin : 0, kb_hi = light ? g.kb2_begin + g.kb2_count : n_kb;
                for (int kb = kb_lo; kb < kb_hi; ++kb, ++it) {
                    const uint32_t s = it % kGStages;
                    mbar_wait_parked(bar_empty + 8u * s, ((it / kGStages) & 1) ^ 1, 1000);
                    const int tap = kb / chunks, ch = kb - tap * chunks;
                    const uint32_t sa = base + s * kGStageBytes, sb = sa + kGStageA;
                    mbar_arrive_expect_tx(bar_full + 8u * s, kGStageBytes);
                    const CUtensorMap *const am = ch < g.c0_chunks ? &tmA0 : &tmA1;
                    const int ck = (ch < g.c0_chunks ? ch : ch - g.c0_chunks) * kGK;
#pragma unroll
                    for (int i = 0; i < 4; ++i)           // four 32-row boxes, one per slab (4 KB apart: whole swizzle atoms)
                        tma_load_3d(sa + 4096u * i, am, bar_full + 8u * s, ck, st_[i] + g.tap_off[tap], sb_[i]);
                    tma_load_2d(sb, &tmB, bar_full + 8u * s, kb * kGK, n0);
                }
            }
        }
    } else if (warp == 1) {
        // ============================ MMA issuer ============================
        if (lane == 0) {
            const uint32_t idesc = F16 ? kIdescF16 : kIdescBf16;
            uint32_t it = 0, lt = 0;
            for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, ++lt) {
                const uint32_t acc = lt & 1;
                int mt, n0;
                tile_coords(tile, mt, n0);
                const bool light = g.n_split && n0 >= g.n_split;
                const int kb_lo = light ? g.kb2_begin : 0, kb_hi = light ? g.kb2_begin + g.kb2_count : n_kb;
                mbar_wait_parked(bar_tempty + 8u * acc, ((lt >> 1) & 1) ^ 1, 1000);     // epilogue has drained it
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                for (int kb = kb_lo; kb < kb_hi; ++kb, ++it) {
                    const uint32_t s = it % kGStages;
                    mbar_wait_parked(bar_full + 8u * s, (it / kGStages) & 1, 1000);
                    if (dbg && it == 0) g_gemm_dbg[1] = clock64();
                    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                    const uint32_t sa = base + s * kGStageBytes, sb = sa + kGStageA;
                    const uint64_t da = umma_desc_sw128(sa), db = umma_desc_sw128(sb);
#pragma unroll
                    for (int k = 0; k < kGK / 16; ++k)     // 32 bytes along K per UMMA_K=16 step: +2 in 16-byte units
                        umma_bf16(tmem_d + acc * kGN, da + 2u * k, db + 2u * k, idesc, (kb != kb_lo || k != 0) ? 1u : 0u);
                    umma_commit(bar_empty + 8u * s);        // frees the stage when these MMAs have read it
                }
                umma_commit(bar_tfull + 8u * acc);          // accumulator complete
            }
        }
    } else {
        // ============================ epilogue ============================
        // TMEM -> registers (lane = row) -> padded shared tile -> registers (lane = 4 columns of one
        // of 4 rows), so that every global access is a 16-byte piece of a 128-byte row segment.
        const int ew = warp - 2;
        const int q = warp & 3;                         // TMEM lane quarter this warp may touch
        const int jhalf = (ew >> 2) * 4;                // this warp's four 32-column chunks
        float *tb = reinterpret_cast<float *>(gbase + epi_off + ew * kGEpiWarpBytes);
        const uint32_t stg = base + epi_off + ew * kGEpiWarpBytes;
        uint32_t n_st = 0;                              // TMA stores issued by this warp
        if (TMA_EPI) {
            // every output column's bias once per CTA (weights: no dependency on the predecessor kernel)
            for (int i = threadIdx.x - 64; i < g.N; i += 32 * kGEpiWarps) {
                float bv = 0.f;
                if (g.n_split && i >= g.n_split) { if (g.bias2) bv = g.bias2[i - g.n_split]; }
                else if (g.bias) bv = g.bias[i];
                sbias[i] = bv;
            }
            asm volatile("bar.sync 1, %0;" ::"n"(32 * kGEpiWarps) : "memory");
            if (lane == 0) { tma_prefetch_desc(&tmO); if (g.n_split) tma_prefetch_desc(&tmO2); }
        }
        const int rsub = lane >> 3, c4 = (lane & 7) * 4;
        uint32_t lt = 0;
        for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, ++lt) {
            int mt, n0;
            tile_coords(tile, mt, n0);
            int b, trow0;                                             // this warp's slab: sample and first frame
            slab_coords(4 * mt + q, g.spp, g.B, g.Tout, b, trow0);
            const bool light = g.n_split && n0 >= g.n_split;
            const int ldo = g.n_split ? g.n_split : g.N;              // output row pitch
            const int ncol0 = light ? n0 - g.n_split : n0;             // column of the tile in its output block
            float *const outp = light ? g.out2_f32 : g.out_f32;
            const float *const biasp = light ? g.bias2 : g.bias;
            const uint32_t acc = lt & 1;
            const uint32_t tacc = tmem_d + acc * kGN + ((uint32_t)(32 * q) << 16);
            mbar_wait_parked(bar_tfull + 8u * acc, (lt >> 1) & 1, 2000);
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            if (dbg && lt == 0 && threadIdx.x == 64) g_gemm_dbg[2] = clock64();
            if (TMA_EPI) {
                const int t = trow0 + lane;                     // this thread's output row (TMEM lane)
                const bool valid = t < g.Tout;
                float mk = g.o16_scale;
                if (B16 && g.omask && valid) mk *= g.omask[(size_t)b * g.omask_T + (size_t)t * g.out_stride + g.out_off];
                const float asc = g.acc_scale;
                const CUtensorMap *const om = light ? &tmO2 : &tmO;
                uint32_t v[2][32];
                tmem_ld32_issue(tacc + 32 * jhalf, v[0]);
#pragma unroll
                for (int jj = 0; jj < 4; ++jj) {
                    const int j = jhalf + jj;
                    tmem_ld_wait();
                    if (jj + 1 < 4) tmem_ld32_issue(tacc + 32 * (j + 1), v[(jj + 1) & 1]);
                    else {
                        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
                        __syncwarp();
                        if (lane == 0) mbar_arrive(bar_tempty + 8u * acc);
                    }
                    const float4 *bp = reinterpret_cast<const float4 *>(sbias + n0 + 32 * j);     // broadcast reads
                    float x[32];
#pragma unroll
                    for (int i = 0; i < 8; ++i) {
                        const float4 bb = bp[i];
                        x[4 * i] = fmaf(__uint_as_float(v[jj & 1][4 * i]), asc, bb.x); x[4 * i + 1] = fmaf(__uint_as_float(v[jj & 1][4 * i + 1]), asc, bb.y);
                        x[4 * i + 2] = fmaf(__uint_as_float(v[jj & 1][4 * i + 2]), asc, bb.z); x[4 * i + 3] = fmaf(__uint_as_float(v[jj & 1][4 * i + 3]), asc, bb.w);
                    }
                    if (EPI == 1) {
#pragma unroll
                        for (int i = 0; i < 32; ++i) x[i] = gelu_erf_fast_f(x[i]);
                    }
                    if (STATS && !light) {
                        // this chunk's 32 columns are GroupNorm group (n0 + 32 j) / 32 of sample b: row sums, then the warp
                        float gs[4] = {0.f, 0.f, 0.f, 0.f}, gq[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
                        for (int i = 0; i < 32; ++i) { gs[i & 3] += x[i]; gq[i & 3] = fmaf(x[i], x[i], gq[i & 3]); }
                        float s1 = valid ? (gs[0] + gs[1]) + (gs[2] + gs[3]) : 0.f, s2 = valid ? (gq[0] + gq[1]) + (gq[2] + gq[3]) : 0.f;
#pragma unroll
                        for (int o = 16; o > 0; o >>= 1) {
                            s1 += __shfl_xor_sync(0xffffffffu, s1, o);
                            s2 += __shfl_xor_sync(0xffffffffu, s2, o);
                        }
                        if (lane == 0) {
                            double *gp = g.gstats + ((size_t)b * kGroups + (n0 + 32 * j) / kGroupCh) * 2;
                            atomicAdd(gp, (double)s1);
                            atomicAdd(gp + 1, (double)s2);
                        }
                    }
                    // staging tile (n_st & 1): free once the store issued two chunks ago has read it
                    if (lane == 0) tma_store_wait_read<1>();
                    __syncwarp();
                    const uint32_t sbuf = stg + (n_st & 1u) * 4096u;
                    if (F32) {                                  // [32 rows][128 B], 16-byte pieces swizzled by row & 7
                        const uint32_t srow = sbuf + (uint32_t)lane * 128u;
#pragma unroll
                        for (int i = 0; i < 8; ++i)
                            sts128(srow + (((uint32_t)i ^ (uint32_t)(lane & 7)) << 4), make_float4(x[4 * i], x[4 * i + 1], x[4 * i + 2], x[4 * i + 3]));
                    } else {                                    // [32 rows][64 B], pieces swizzled by (row >> 1) & 3
                        const uint32_t srow = sbuf + (uint32_t)lane * 64u;
#pragma unroll
                        for (int i = 0; i < 4; ++i)
                            asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(srow + (((uint32_t)i ^ (uint32_t)((lane >> 1) & 3)) << 4)),
                                         "r"(pack16x2(x[8 * i] * mk, x[8 * i + 1] * mk, F16)), "r"(pack16x2(x[8 * i + 2] * mk, x[8 * i + 3] * mk, F16)),
                                         "r"(pack16x2(x[8 * i + 4] * mk, x[8 * i + 5] * mk, F16)), "r"(pack16x2(x[8 * i + 6] * mk, x[8 * i + 7] * mk, F16)) : "memory");
                    }
                    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                    __syncwarp();
                    if (lane == 0) {
                        tma_store_3d(om, sbuf, ncol0 + 32 * j, trow0, b);     // rows >= Tout are clipped by the map
                        tma_store_commit();
                    }
                    ++n_st;
                }
            } else if (FIN) {
                // lane = frame t: for a fixed channel the warp writes 32 consecutive frames (coalesced, no transpose)
                const int t = trow0 + lane;
                const float mk = t < g.Tout ? g.omask[(size_t)b * g.omask_T + t] : 0.f;
                float lsum = 0.f;
                for (int j = jhalf; j < jhalf + 4 && j * 32 < g.fin_n; ++j) {
                    uint32_t v[32];
                    float xin[32];
#pragma unroll
                    for (int i = 0; i < 32; ++i) {              // all loads of the chunk first
                        const int n = 32 * j + i;
                        if (LOSS) {                             // the regression target u = x1 - (1 - sigma) z, not masked
                            const bool ok = n < g.fin_n && t < g.Tout;
                            const size_t o = ((size_t)b * g.fin_n + n) * g.Tout + t;
                            xin[i] = ok ? __fsub_rn(g.loss_x1[o], __fmul_rn(g.loss_oms, g.loss_z[o])) : 0.f;
                        } else {
                            xin[i] = (g.fin_xin && n < g.fin_n && t < g.Tout) ? g.fin_xin[((size_t)b * g.fin_n + n) * g.Tout + t] : 0.f;
                        }
                    }
                    tmem_ld32(tacc + 32 * j, v);
#pragma unroll
                    for (int i = 0; i < 32; ++i) {
                        const int n = 32 * j + i;
                        if (n < g.fin_n && t < g.Tout) {
                            const float vel = (__uint_as_float(v[i]) + g.bias[n]) * mk;
                            if (LOSS) {
                                const float d = __fsub_rn(vel, xin[i]);
                                lsum = fmaf(d, d, lsum);
                                if (g.fin_out) g.fin_out[((size_t)b * g.fin_n + n) * g.Tout + t] = vel;
                            } else {
                                g.fin_out[((size_t)b * g.fin_n + n) * g.Tout + t] = g.fin_xin ? __fadd_rn(xin[i], __fmul_rn(g.fin_dt, vel)) : vel;
                            }
                        }
                    }
                }
                if (LOSS && jhalf * 32 < g.fin_n) {
                    double ds = (double)lsum;
#pragma unroll
                    for (int o = 16; o > 0; o >>= 1) ds += __shfl_xor_sync(0xffffffffu, ds, o);
                    if (lane == 0) atomicAdd(g.loss_sum, ds);
                }
            } else {
                // output element offsets of this lane's 8 rows (column part added per chunk); -1 = row beyond Tout
                int orow[8];                                   // fit 32 bits (B*T*N < 2^31, checked on the host)
                float mk[8];
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    const int t = trow0 + 4 * i + rsub;
                    const bool ok = t < g.Tout;
                    const long long r = (long long)b * g.out_T + (long long)t * g.out_stride + g.out_off;
                    orow[i] = ok ? (int)(r * ldo) : -1;
                    mk[i] = g.o16_scale;
                    if (B16 && g.omask && ok) mk[i] *= g.omask[(size_t)b * g.omask_T + (size_t)t * g.out_stride + g.out_off];
                }
                float4 rv[8];
                auto load_res = [&](int jj) {                    // residual rows of chunk jj (issued one chunk ahead)
                    const int n = ncol0 + 32 * jj + c4;
#pragma unroll
                    for (int i = 0; i < 8; ++i)
                        rv[i] = orow[i] >= 0 ? ld_chain4(g.res + orow[i] + n) : make_float4(0.f, 0.f, 0.f, 0.f);
                };
                if (RES) load_res(jhalf);
                uint32_t v[2][32];                               // TMEM chunk j+1 is in flight while chunk j is processed
                tmem_ld32_issue(tacc + 32 * jhalf, v[0]);
#pragma unroll
                for (int jj = 0; jj < 4; ++jj) {
                    const int j = jhalf + jj;
                    tmem_ld_wait();
                    if (jj + 1 < 4) tmem_ld32_issue(tacc + 32 * (j + 1), v[(jj + 1) & 1]);
                    else {
                        // every TMEM read of this warp has landed: hand the accumulator back to the MMA warp now
                        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
                        __syncwarp();
                        if (lane == 0) mbar_arrive(bar_tempty + 8u * acc);
                    }
                    __syncwarp();
                    if (dbg && lt == 0 && threadIdx.x == 64 && jj == 0) g_gemm_dbg[4] = clock64();
#pragma unroll
                    for (int i = 0; i < 8; ++i)
                        *reinterpret_cast<float4 *>(tb + lane * kGEpiPitch + 4 * i) =
                            make_float4(__uint_as_float(v[jj & 1][4 * i]), __uint_as_float(v[jj & 1][4 * i + 1]),
                                        __uint_as_float(v[jj & 1][4 * i + 2]), __uint_as_float(v[jj & 1][4 * i + 3]));
                    __syncwarp();
                    if (dbg && lt == 0 && threadIdx.x == 64 && jj == 0) g_gemm_dbg[5] = clock64();
                    const int n = ncol0 + 32 * j + c4;
                    float4 bias = make_float4(0.f, 0.f, 0.f, 0.f);
                    if (biasp) bias = *reinterpret_cast<const float4 *>(biasp + n);
                    float4 x[8];
#pragma unroll
                    for (int i = 0; i < 8; ++i) {
                        x[i] = *reinterpret_cast<const float4 *>(tb + (4 * i + rsub) * kGEpiPitch + c4);
                        x[i].x = fmaf(x[i].x, g.acc_scale, bias.x); x[i].y = fmaf(x[i].y, g.acc_scale, bias.y);
                        x[i].z = fmaf(x[i].z, g.acc_scale, bias.z); x[i].w = fmaf(x[i].w, g.acc_scale, bias.w);
                        if (EPI == 1) { x[i].x = gelu_erf_fast_f(x[i].x); x[i].y = gelu_erf_fast_f(x[i].y); x[i].z = gelu_erf_fast_f(x[i].z); x[i].w = gelu_erf_fast_f(x[i].w); }
                        if (RES) { x[i].x += rv[i].x; x[i].y += rv[i].y; x[i].z += rv[i].z; x[i].w += rv[i].w; }
                    }
                    if (RES && jj + 1 < 4) load_res(j + 1);
                    if (STATS && !light) {
                        // the warp holds 32 rows x the 32 channels of GroupNorm group (n0 + 32 j) / 32 of sample b
                        float gs = 0.f, gq = 0.f;
#pragma unroll
                        for (int i = 0; i < 8; ++i) {
                            if (orow[i] >= 0) {
                                gs += (x[i].x + x[i].y) + (x[i].z + x[i].w);
                                gq += (x[i].x * x[i].x + x[i].y * x[i].y) + (x[i].z * x[i].z + x[i].w * x[i].w);
                            }
                        }
#pragma unroll
                        for (int o = 16; o > 0; o >>= 1) {
                            gs += __shfl_xor_sync(0xffffffffu, gs, o);
                            gq += __shfl_xor_sync(0xffffffffu, gq, o);
                        }
                        if (lane == 0) {
                            double *gp = g.gstats + ((size_t)b * kGroups + (n0 + 32 * j) / kGroupCh) * 2;
                            atomicAdd(gp, (double)gs);
                            atomicAdd(gp + 1, (double)gq);
                        }
                    }
                    if (dbg && lt == 0 && threadIdx.x == 64 && jj == 0) g_gemm_dbg[6] = clock64();
#pragma unroll
                    for (int i = 0; i < 8; ++i) {
                        if (orow[i] >= 0) {
                            if (F32) *reinterpret_cast<float4 *>(outp + orow[i] + n) = x[i];
                            if (B16)
                                *reinterpret_cast<uint2 *>(g.out_bf16 + orow[i] + n) =
                                    make_uint2(pack16x2(x[i].x * mk[i], x[i].y * mk[i], F16), pack16x2(x[i].z * mk[i], x[i].w * mk[i], F16));
                        }
                    }
                    if (dbg && lt == 0 && threadIdx.x == 64 && jj == 0) g_gemm_dbg[7] = clock64();
                }
            }
            if (FIN) {
                // this warp's TMEM reads are complete (tcgen05.wait::ld): hand the accumulator back to the MMA warp
                asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
                __syncwarp();
                if (lane == 0) mbar_arrive(bar_tempty + 8u * acc);
            }
            if (dbg && lt == 0 && threadIdx.x == 64) g_gemm_dbg[3] = clock64();
        }
        if (TMA_EPI && lane == 0) tma_store_wait_read<0>();    // shared memory must outlive the stores' reads
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 1) {
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_d), "n"(512) : "memory");
    }
}

}  // namespace mb200
