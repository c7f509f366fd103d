// mas.cu -- monotonic alignment search on sm_100a.
//
// Replaces /root/reference/matcha/utils/monotonic_align/core.pyx:16-128 (forward DP over
// mel frames with the "ties go to token x-1" rule of :73, then the backtrack of :87-92).
// Results are bit-exact with the compiled reference; see DESIGN.md §MAS for the argument.
//
// One persistent CTA per SM slot, utterances dealt round-robin.  Warp roles:
//   warps 0..NDP-1  DP warps, an anti-diagonal wavefront over (row block, 32-frame chunk):
//           warp w owns tokens [32*R*w, 32*R*(w+1)); lane l owns R consecutive tokens.  One
//           step per mel frame: a single shfl_up brings Q[y-1][x0-1] from lane l-1, the other
//           R-1 predecessors are registers.  Warp w runs one chunk behind warp w-1 and gets the
//           boundary row Q[.][32*R*w-1] through a 2-deep shared-memory ring guarded by
//           mbarriers.  The "came from x-1" bit of every cell is packed 32 frames per word and
//           parked in the utterance's own (not yet written) output region, i.e. in L2.
//           Each DP warp also feeds itself: the fp32 score rows it needs for a chunk (only rows
//           inside the DP band) arrive in its own ring of 128B-swizzled stages by TMA
//           (cp.async.bulk.tensor, completion on an mbarrier); the stage it has just finished is
//           re-armed at once with the tile needed nstage chunks later.  When the tensor breaks
//           TMA's 16-byte rules (T_y % 4 != 0 or a misaligned base) the warp fills the same
//           layout with 4-byte cp.async instead.
//   warp NDP        backtrack.  Walks the bit words of the finished utterance chunk by chunk
//           (one coalesced 128-byte read per chunk, the next one prefetched) while the DP
//           warps already run the CTA's next utterance, and emits per-token start frames.
//   last 4 warps    writers.  Turn the per-token start frames found by the backtrack into
//           the dense 0/1 alignment with 16-byte streaming stores, overlapping the DP of
//           the CTA's next utterance.
#include "common.cuh"

#include <stdlib.h>
#include <string.h>

#include <mutex>

namespace mb200 {

constexpr int kChunk = 32;        // mel frames per tile
constexpr int kBoxRows = 32;      // tokens per TMA box
constexpr int kBoxBytes = kBoxRows * kChunk * 4;
constexpr int kWriterWarps = 4;   // 2 writers: 0.869 -> 0.819 of the HBM peak at 400x2000, 1: 0.564 (profiles/r2_exp11_mas_writer_warps.log)
constexpr int kMaxStages = 8;
constexpr int kBndDepth = 4;      // boundary-row ring depth between neighbouring DP warps
constexpr int kBndFloats = 36;    // [0]=Q[32c-1], [1..32]=Q[32c..32c+31], padded to 16 bytes

// most DP warps a CTA may carry for a given rows-per-lane (register budget of the SM)
__host__ __device__ constexpr int max_dp_warps(int R) {
    return R <= 5 ? 8 : (R <= 7 ? 6 : (R <= 9 ? 4 : 3));
}

struct MasParams {
    const float *value;
    const float *mask;
    uint32_t *out;
    const int *t_xs;
    const int *t_ys;
    int *status;
    int *durations;      // optional [B][T_x]: frames owned by each token (0 for x >= t_x)
    int B, T_x, T_y;
    float penalty;
    uint32_t one_bits;   // bit pattern of "1" in the output element type
    int ndp;             // DP warps in the CTA
    int nstage;          // ring depth per DP warp
    int txp;             // T_x rounded up to 32: rows of a stage
    int n_chunks;        // ceil(T_y / 32)
    int use_tma;
    int vec_store;       // output rows are 16-byte aligned
    int early_zero;      // single-wave launches: zero the slab behind the bit words during the DP
    int debug;
};

struct ItemInfo { int tx, ty, valid, pad; };

// profiling aid (MB200_MAS_DEBUG=1): clock64 stamps of CTA 0's first utterance
__device__ long long g_mas_dbg[32];
#define MAS_STAMP(i) do { if (p.debug && blockIdx.x == 0 && k == 0 && lane == 0) g_mas_dbg[i] = clock64(); } while (0)

__device__ __forceinline__ bool lengths_valid(int tx, int ty, int T_x, int T_y) {
    return tx >= 1 && ty >= 1 && tx <= ty && tx <= T_x && ty <= T_y;
}

// chunk range of a row block [r0, r1): first chunk holds its diagonal start, last chunk is
// where its last row leaves the band (or the utterance ends).
__device__ __forceinline__ void rb_chunk_range(int r0, int r1, int tx, int ty, int &c0, int &c1) {
    c0 = r0 / kChunk;
    int y_last = r1 - 1 + ty - tx;
    if (y_last > ty - 1) y_last = ty - 1;
    c1 = y_last / kChunk;
}

// ------------------------------------------------------------------ DP over one chunk
// `stage` points at this warp's first row inside the stage.  has_prev: lane 0 takes
// Q[y-1][x0-1] from the boundary slot `bin` written by the previous warp; put_next: lane 31
// publishes its last row into `bout` for the next warp.
template <int R, bool HAS_MASK, bool DIAG>
__device__ __forceinline__ void dp_chunk(float (&q)[R], uint32_t (&bits)[R], uint32_t stage,
                                         uint32_t mask_off, int rows_avail, int lane, int y_base,
                                         int x0, float penalty, bool has_prev, uint32_t bin,
                                         bool put_next, uint32_t bout) {
    uint32_t row_addr[R], row_sw[R];
#pragma unroll
    for (int r = 0; r < R; ++r) {
        int rho = min(R * lane + r, rows_avail - 1);   // rows past the stage re-read its last row (never used)
        row_addr[r] = stage + rho * 128;
        row_sw[r] = (rho & 7) << 4;
    }
    float *bo = reinterpret_cast<float *>(__cvta_shared_to_generic(bout));
    if (put_next && lane == 31) bo[0] = q[R - 1];
#pragma unroll 1
    for (int g = 0; g < 8; ++g) {
        float4 v[R];
#pragma unroll
        for (int r = 0; r < R; ++r) {
            uint32_t a = row_addr[r] + ((uint32_t)(g << 4) ^ row_sw[r]);
            v[r] = lds128(a);
            if (HAS_MASK) {
                float4 m = lds128(a + mask_off);
                v[r].x *= m.x; v[r].y *= m.y; v[r].z *= m.z; v[r].w *= m.w;
            }
        }
        const int y0 = y_base + 4 * g;
        float4 upv = make_float4(0.f, 0.f, 0.f, 0.f), outv = upv;
        if (has_prev) upv = lds128(bin + 16 * g);   // all lanes (broadcast): keeps the warp converged
        const uint32_t mg = 1u << (4 * g);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int y = y0 + j;
            const uint32_t mj = mg << j;
            float up = __shfl_up_sync(0xffffffffu, q[R - 1], 1);
            if (lane == 0) {
                if (has_prev) up = (j == 0) ? upv.x : (j == 1) ? upv.y : (j == 2) ? upv.z : upv.w;
                else up = (y == 0) ? 0.0f : penalty;             // core.pyx:63-64
            }
            const int dy = y - x0;
#pragma unroll
            for (int r = R - 1; r >= 0; --r) {
                const float prev = (r == 0) ? up : q[r - 1];     // Q[y-1][x-1], core.pyx:66
                const float same = q[r];                         // Q[y-1][x],   core.pyx:71
                const float s = (j == 0) ? v[r].x : (j == 1) ? v[r].y : (j == 2) ? v[r].z : v[r].w;
                bool d = prev >= same;                           // ties -> x-1, core.pyx:73
                if (DIAG) d = d || (r == dy);                    // x == y forces x-1, core.pyx:68-73
                const float best = d ? prev : same;
                q[r] = __fadd_rn(best, s);                       // core.pyx:80
                bits[r] |= d ? mj : 0u;                          // core.pyx:75,78
            }
            if (j == 0) outv.x = q[R - 1]; else if (j == 1) outv.y = q[R - 1];
            else if (j == 2) outv.z = q[R - 1]; else outv.w = q[R - 1];
        }
        if (put_next && lane == 31) {
            float *o = bo + 1 + 4 * g;
            o[0] = outv.x; o[1] = outv.y; o[2] = outv.z; o[3] = outv.w;
        }
    }
}

// ------------------------------------------------------------------ shared-memory carve-up
struct MasSmem {
    uint32_t stages;   // nstage stage sets of txp rows x 128 B (x2 with a mask), 1024-aligned
    uint32_t bnd;      // (ndp-1) * kBndDepth * kBndFloats floats
    uint32_t cs;       // 2 * (txp + 4) ints: start frame of every token (+ end sentinel)
    uint32_t info;     // 2 * ItemInfo
    uint32_t bars;     // full [ndp*nstage], bdfull/bdempty [(ndp-1)*depth],
                       // bfull/bempty [2], dpdone/btdone [2]
    uint32_t stage_bytes;
    uint32_t total;
};

__host__ __device__ inline MasSmem mas_smem_layout(bool has_mask, int ndp, int nstage, int txp) {
    MasSmem s;
    uint32_t off = 0;
    s.stage_bytes = (uint32_t)txp * 128u * (has_mask ? 2u : 1u);
    s.stages = off; off += (uint32_t)nstage * s.stage_bytes;
    s.bnd = off;    off += (uint32_t)(ndp - 1) * kBndDepth * kBndFloats * 4u;
    s.cs = off;     off += 2u * (uint32_t)(txp + 4) * 4u;
    s.info = off;   off += 2u * (uint32_t)sizeof(ItemInfo);
    off = (off + 7u) & ~7u;
    s.bars = off;
    off += ((uint32_t)ndp * (uint32_t)nstage + 2u * (uint32_t)(ndp - 1) * kBndDepth + 8u) * 8u;
    s.total = off + 1024u;   // slack for the manual 1024-byte alignment of the base
    return s;
}

// highest set bit of w at or below bit position `pos` (0..31), or -1
__device__ __forceinline__ int top_bit_upto(uint32_t w, int pos) {
    w &= 0xffffffffu >> (31 - pos);
    return 31 - __clz(w);          // __clz(0) == 32 -> -1
}

// ------------------------------------------------------------------ the kernel
template <int R, bool HAS_MASK>
__global__ void __launch_bounds__(32 * (max_dp_warps(R) + 1 + kWriterWarps), R <= 3 ? 2 : 1)
mas_kernel(const __grid_constant__ CUtensorMap tmap_v, const __grid_constant__ CUtensorMap tmap_m,
           const MasParams p) {
    extern __shared__ uint8_t smem_raw[];
    const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    uint8_t *gbase = smem_raw + (base - smem_u32(smem_raw));
    const MasSmem L = mas_smem_layout(HAS_MASK, p.ndp, p.nstage, p.txp);
    constexpr int RB = 32 * R;
    const int ndp = p.ndp, nstage = p.nstage;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    // barrier addresses
    const uint32_t bar_full = base + L.bars;                               // [w][s]
    const uint32_t bar_bdfull = bar_full + 8u * ndp * nstage;              // [w][d], w < ndp-1
    const uint32_t bar_bdempty = bar_bdfull + 8u * (ndp - 1) * kBndDepth;  // [w][d]
    const uint32_t bar_bfull = bar_bdempty + 8u * (ndp - 1) * kBndDepth;   // [2] backtrack -> writers
    const uint32_t bar_bempty = bar_bfull + 16u;                           // [2] writers -> backtrack
    const uint32_t bar_dpdone = bar_bempty + 16u;                          // [2] DP warps -> backtrack
    const uint32_t bar_btdone = bar_dpdone + 16u;                          // [2] backtrack -> DP warps
    int *cs_g = reinterpret_cast<int *>(gbase + L.cs);
    ItemInfo *info_g = reinterpret_cast<ItemInfo *>(gbase + L.info);
    const uint32_t mask_off = (uint32_t)p.txp * 128u;

    if (threadIdx.x == 0) {
        for (int i = 0; i < ndp * nstage; ++i) {
            mbar_init(bar_full + 8u * i, p.use_tma ? 1u : 32u);
        }
        for (int i = 0; i < (ndp - 1) * kBndDepth; ++i) {
            mbar_init(bar_bdfull + 8u * i, 1u);
            mbar_init(bar_bdempty + 8u * i, 1u);
        }
        for (int i = 0; i < 2; ++i) {
            mbar_init(bar_bfull + 8u * i, 1u);
            mbar_init(bar_bempty + 8u * i, kWriterWarps);
            mbar_init(bar_dpdone + 8u * i, (uint32_t)ndp);
            mbar_init(bar_btdone + 8u * i, 1u);
        }
        mbar_fence_init();
    }
    __syncthreads();
    if (p.debug && blockIdx.x == 0 && threadIdx.x == 0) g_mas_dbg[0] = clock64();

    if (warp < ndp) {
        // ============================ DP wavefront ============================
        const int w = warp;
        const int r0 = w * RB;
        const int x0 = r0 + R * lane;
        if (p.use_tma && w == 0 && lane == 0) {
            tma_prefetch_desc(&tmap_v);
            if (HAS_MASK) tma_prefetch_desc(&tmap_m);
        }
        // Refill cursor: walks the same (utterance, chunk) sequence as the DP loop below, nstage
        // tiles ahead.  The warp refills its own ring: the stage it has just finished reading is
        // immediately re-armed with the tile it will need nstage chunks later (possibly the first
        // tile of the CTA's next utterance), so no producer warp and no "empty" barriers exist.
        int rb = blockIdx.x - (int)gridDim.x, rtx = 0, rty = 0, rc = 0, rc1 = -1;
        uint32_t rit = 0;
        auto refill_one = [&]() {
            while (rc > rc1) {                                  // next utterance this warp works on
                rb += gridDim.x;
                if (rb >= p.B) return;
                rtx = p.t_xs[rb]; rty = p.t_ys[rb];
                if (!lengths_valid(rtx, rty, p.T_x, p.T_y) || r0 >= rtx) continue;
                rb_chunk_range(r0, min(r0 + RB, rtx), rtx, rty, rc, rc1);
            }
            const int c = rc++;
            const int r1 = min(r0 + RB, rtx);
            const uint32_t s = rit % (uint32_t)nstage;
            ++rit;
            const uint32_t stage = base + L.stages + s * L.stage_bytes + (uint32_t)r0 * 128u;
            const uint32_t full = bar_full + 8u * (w * nstage + s);
            // rows of this block inside the band over frames [32c, 32c+32)
            const int lo = max(r0, rtx + kChunk * c - rty);
            const int hi = min(r1, kChunk * c + kChunk);
            const int j0 = (lo - r0) / kBoxRows, j1 = (hi - 1 - r0) / kBoxRows;
            if (p.use_tma) {
                if (lane == 0) {
                    // order this warp's earlier (generic-proxy) reads of the stage before the async-proxy write
                    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                    mbar_arrive_expect_tx(full, (uint32_t)(j1 - j0 + 1) * kBoxBytes * (HAS_MASK ? 2u : 1u));
                    for (int j = j0; j <= j1; ++j) {
                        const int row = rb * p.T_x + r0 + kBoxRows * j;
                        tma_load_2d(stage + j * kBoxBytes, &tmap_v, full, kChunk * c, row);
                        if (HAS_MASK)
                            tma_load_2d(stage + mask_off + j * kBoxBytes, &tmap_m, full, kChunk * c, row);
                    }
                }
            } else {
                const int col = kChunk * c + lane;
                if (col < p.T_y) {
                    const int rho1 = min((j1 + 1) * kBoxRows, r1 - r0);
                    for (int rho = j0 * kBoxRows; rho < rho1; ++rho) {
                        const size_t g = ((size_t)rb * p.T_x + r0 + rho) * (size_t)p.T_y + col;
                        const uint32_t d = stage + rho * 128 + ((((uint32_t)lane >> 2) ^ (rho & 7)) << 4) + (lane & 3) * 4;
                        cp_async_4(d, p.value + g);
                        if (HAS_MASK) cp_async_4(d + mask_off, p.mask + g);
                    }
                }
                cp_async_mbar_arrive_noinc(full);
            }
        };
        for (int i = 0; i < nstage; ++i) refill_one();
        __syncwarp();

        uint32_t it = 0, hin = 0, hout = 0, k = 0;
        for (int b = blockIdx.x; b < p.B; b += gridDim.x, ++k) {
            const int tx = p.t_xs[b], ty = p.t_ys[b];
            const uint32_t buf = k & 1u;
            // the backtrack warp must have consumed utterance k-2 before dpdone[buf] is re-armed
            mbar_wait_parked(bar_btdone + 8u * buf, ((k >> 1) & 1u) ^ 1u, 5000);
            if (lengths_valid(tx, ty, p.T_x, p.T_y) && r0 < tx) {
                // bit words of this utterance: [chunk][token], parked in its own output region
                uint32_t *bits_g = p.out + (size_t)b * p.T_x * p.T_y;
                const int n_act = (tx + RB - 1) / RB;
                const int r1 = min(r0 + RB, tx);
                int c0, c1, c0p = 0, c1p = -1;
                rb_chunk_range(r0, r1, tx, ty, c0, c1);
                if (w > 0) rb_chunk_range(r0 - RB, r0, tx, ty, c0p, c1p);
                const bool has_prev = w > 0, has_next = (w + 1 < n_act);
                // Row r0-1 leaves the band at frame y_e = r0-1+ty-tx; cell (r0, y_e+1) still reads Q[y_e][r0-1]
                // (core.pyx:66).  When y_e is the last frame of chunk c1p that cell opens chunk c1p+1, in which
                // the previous warp no longer runs: it sends one extra "tail" message carrying that value.
                int c0x = 0, c1x = -1;
                if (has_next) rb_chunk_range(r0 + RB, min(r0 + 2 * RB, tx), tx, ty, c0x, c1x);
                const bool tail_out = has_next && (c1 + 1 <= c1x);
                const int c0n = (r0 + RB) / kChunk;     // first chunk of the next warp
                const int cd = (r1 - 1) / kChunk;       // last chunk that holds a diagonal cell
                float q[R];
                uint32_t bits[R];
#pragma unroll
                for (int r = 0; r < R; ++r) q[r] = p.penalty;      // core.pyx:52-53
                MAS_STAMP(1 + 2 * w);
                for (int c = c0; c <= c1; ++c, ++it) {
                    const uint32_t s = it % (uint32_t)nstage, ph = (it / (uint32_t)nstage) & 1u;
                    const uint32_t stage = base + L.stages + s * L.stage_bytes + (uint32_t)r0 * 128u;
                    const bool in_h = has_prev && c <= c1p + 1;   // c1p + 1: the tail message
                    const bool out_h = has_next && c >= c0n;
                    const uint32_t din = hin % kBndDepth, dout = hout % kBndDepth;
                    const uint32_t bin = base + L.bnd + (((w - 1) * kBndDepth + din) * kBndFloats) * 4u;
                    const uint32_t bout = base + L.bnd + ((w * kBndDepth + dout) * kBndFloats) * 4u;
#pragma unroll
                    for (int r = 0; r < R; ++r) bits[r] = 0u;
                    // every lane waits (warp-uniform control flow keeps the shuffles on the fast path)
                    mbar_wait_parked(bar_full + 8u * (w * nstage + s), ph, 2000);
                    if (in_h)
                        mbar_wait_parked(bar_bdfull + 8u * ((w - 1) * kBndDepth + din), (hin / kBndDepth) & 1u, 2000);
                    if (out_h)
                        mbar_wait_parked(bar_bdempty + 8u * (w * kBndDepth + dout), ((hout / kBndDepth) & 1u) ^ 1u, 2000);
                    __syncwarp();
                    if (c <= cd)
                        dp_chunk<R, HAS_MASK, true>(q, bits, stage, mask_off, p.txp - r0, lane, kChunk * c, x0, p.penalty,
                                                    in_h, bin, out_h, bout);
                    else
                        dp_chunk<R, HAS_MASK, false>(q, bits, stage, mask_off, p.txp - r0, lane, kChunk * c, x0, p.penalty,
                                                     in_h, bin, out_h, bout);
                    __syncwarp();                               // all lanes are done reading the stage
                    if (lane == 0 && in_h) mbar_arrive(bar_bdempty + 8u * ((w - 1) * kBndDepth + din));
                    if (lane == 31 && out_h) mbar_arrive(bar_bdfull + 8u * (w * kBndDepth + dout));
                    refill_one();
                    uint32_t *wd = bits_g + (size_t)c * p.T_x + x0;
#pragma unroll
                    for (int r = 0; r < R; ++r)
                        if (x0 + r < p.T_x) wd[r] = bits[r];
                    hin += in_h ? 1u : 0u;
                    hout += out_h ? 1u : 0u;
                }
                if (tail_out) {
                    const uint32_t dout = hout % kBndDepth;
                    mbar_wait_parked(bar_bdempty + 8u * (w * kBndDepth + dout), ((hout / kBndDepth) & 1u) ^ 1u, 2000);
                    __syncwarp();
                    if (lane == 31) {
                        float *bo = reinterpret_cast<float *>(
                            __cvta_shared_to_generic(base + L.bnd + ((w * kBndDepth + dout) * kBndFloats) * 4u));
                        bo[0] = q[R - 1];
                        for (int i = 1; i < 32; ++i) bo[i] = p.penalty;   // row r0-1 is outside the band from here on
                        mbar_arrive(bar_bdfull + 8u * (w * kBndDepth + dout));
                    }
                    ++hout;
                }
            }
            MAS_STAMP(2 + 2 * w);
            __syncwarp();
            if (lane == 0) mbar_arrive(bar_dpdone + 8u * buf);   // release: this warp's bit words are out
        }
    } else if (warp == ndp) {
        // ============================ backtrack ============================
        uint32_t k = 0;
        for (int b = blockIdx.x; b < p.B; b += gridDim.x, ++k) {
            const int tx = p.t_xs[b], ty = p.t_ys[b];
            const bool valid = lengths_valid(tx, ty, p.T_x, p.T_y);
            const uint32_t buf = k & 1u;
            int *cs = cs_g + buf * (p.txp + 4);
            mbar_wait_parked(bar_dpdone + 8u * buf, (k >> 1) & 1u, 5000);
            mbar_wait_parked(bar_bempty + 8u * buf, ((k >> 1) & 1u) ^ 1u, 5000);
            MAS_STAMP(20);
            if (valid) {
                // core.pyx:87-92: token idx owns frames [cs[idx], cs[idx+1]).  Within one 32-frame
                // chunk the path descends at most 32 tokens, so lane l holds the word of token
                // idx_entry - l; the 64 candidate words of the next chunk are prefetched meanwhile.
                const uint32_t *bits_g = p.out + (size_t)b * p.T_x * p.T_y;
                int idx = tx - 1, y = ty - 1, c = y >> 5;
                if (lane == 0) cs[tx] = ty;
                uint32_t cur = 0u, nxa = 0u, nxb = 0u;
                if (idx - lane >= 0) cur = __ldcg(bits_g + (size_t)c * p.T_x + (idx - lane));
                while (idx > 0) {
                    const int ie = idx;                       // token at chunk entry
                    if (c > 0) {
                        const uint32_t *nb = bits_g + (size_t)(c - 1) * p.T_x;
                        if (ie - lane >= 0) nxa = __ldcg(nb + (ie - lane));
                        if (ie - 32 - lane >= 0) nxb = __ldcg(nb + (ie - 32 - lane));
                    }
                    // walk this chunk: frames (c<<5) .. y
                    while (idx > 0) {
                        const uint32_t wv = __shfl_sync(0xffffffffu, cur, ie - idx);
                        const int tb = top_bit_upto(wv, y & 31);
                        if (tb < 0) break;                    // token idx continues into chunk c-1
                        const int yp = (c << 5) + tb;
                        if (lane == 0) cs[idx] = yp;
                        --idx;
                        y = yp - 1;
                        if (tb == 0) break;                   // next token starts in chunk c-1
                    }
                    if (idx > 0) {
                        --c;
                        y = (c << 5) + 31;
                        // re-centre the prefetched words on the new entry token
                        const int sh = ie - idx;              // 0..32
                        const uint32_t a = __shfl_sync(0xffffffffu, nxa, (lane + sh) & 31);
                        const uint32_t bq = __shfl_sync(0xffffffffu, nxb, (lane + sh) & 31);
                        cur = (lane + sh < 32) ? a : bq;
                    }
                }
                if (lane == 0) cs[0] = 0;
            } else if (lane == 0 && p.status != nullptr) {
                atomicAdd(p.status, 1);
            }
            MAS_STAMP(21);
            __syncwarp();
            if (lane == 0) {
                info_g[buf].tx = tx; info_g[buf].ty = ty; info_g[buf].valid = valid ? 1 : 0;
                mbar_arrive(bar_bfull + 8u * buf);
                mbar_arrive(bar_btdone + 8u * buf);
            }
        }
    } else {
        // ============================ writers ============================
        const int ww = warp - ndp - 1;
        uint32_t k = 0;
        const int wt = ww * 32 + lane;                           // 0..127
        constexpr int WT = kWriterWarps * 32;
        for (int b = blockIdx.x; b < p.B; b += gridDim.x, ++k) {
            const uint32_t buf = k & 1u;
            const int *cs = cs_g + buf * (p.txp + 4);
            uint32_t *out = p.out + (size_t)b * p.T_x * p.T_y;
            const uint32_t one = p.one_bits;
            const size_t n = (size_t)p.T_x * p.T_y;
            // The head of the slab holds the utterance's bit words until the backtrack is done;
            // everything behind it is zeroed right away, overlapping the utterance's own DP.
            size_t head = (size_t)p.n_chunks * p.T_x;
            head = (head + 3) & ~(size_t)3;
            if (head > n) head = n;
            auto zero_fill = [&](size_t lo, size_t hi) {         // lo, hi multiples of 4 when vec_store
                if (p.vec_store) {
                    uint4 *o4 = reinterpret_cast<uint4 *>(out);
                    const uint4 z = make_uint4(0u, 0u, 0u, 0u);
                    const size_t hi4 = hi >> 2;
                    size_t i = (lo >> 2) + wt;
                    for (; i + 3 * WT < hi4; i += 4 * WT) {
                        __stcs(o4 + i, z);
                        __stcs(o4 + i + 1 * WT, z);
                        __stcs(o4 + i + 2 * WT, z);
                        __stcs(o4 + i + 3 * WT, z);
                    }
                    for (; i < hi4; i += WT) __stcs(o4 + i, z);
                } else {
                    for (size_t i = lo + wt; i < hi; i += WT) __stcs(out + i, 0u);
                }
            };
            if (ww == 0) MAS_STAMP(22);
            if (p.early_zero) zero_fill(head, n);
            if (ww == 0) MAS_STAMP(23);
            mbar_wait_parked(bar_bfull + 8u * buf, (k >> 1) & 1u, 20000);
            if (ww == 0) MAS_STAMP(24);
            const int tx = info_g[buf].valid ? info_g[buf].tx : 0;
            if (p.durations != nullptr) {                         // compact form of the path (1 int per token)
                int *d = p.durations + (size_t)b * p.T_x;
                for (int x = wt; x < p.T_x; x += WT) d[x] = (x < tx) ? cs[x + 1] - cs[x] : 0;
            }
            if (p.early_zero) {
                zero_fill(0, head);
                // every zero is ordered before the ones dropped on top of it
                asm volatile("bar.sync 2, %0;" ::"n"(kWriterWarps * 32) : "memory");
                for (int x = wt; x < tx; x += WT) {               // one token per lane: its run of frames
                    const int e = cs[x + 1];
                    uint32_t *row = out + (size_t)x * p.T_y;
                    for (int c = cs[x]; c < e; ++c) __stcs(row + c, one);
                }
            } else if (p.vec_store) {
                // steady state (many utterances per CTA): one pass, row by row, paced by the
                // memory system while the DP warps run the next utterance
                for (int x = ww; x < p.T_x; x += kWriterWarps) {
                    int s = 0, e = 0;
                    if (x < tx) { s = cs[x]; e = cs[x + 1]; }
                    uint4 *row = reinterpret_cast<uint4 *>(out + (size_t)x * p.T_y);
                    for (int c = lane * 4; c < p.T_y; c += 128) {
                        uint4 v = make_uint4(0u, 0u, 0u, 0u);
                        if (c + 3 >= s && c < e) {
                            v.x = (c + 0 >= s && c + 0 < e) ? one : 0u;
                            v.y = (c + 1 >= s && c + 1 < e) ? one : 0u;
                            v.z = (c + 2 >= s && c + 2 < e) ? one : 0u;
                            v.w = (c + 3 >= s && c + 3 < e) ? one : 0u;
                        }
                        __stcs(row + (c >> 2), v);
                    }
                }
            } else {
                for (int x = ww; x < p.T_x; x += kWriterWarps) {
                    int s = 0, e = 0;
                    if (x < tx) { s = cs[x]; e = cs[x + 1]; }
                    uint32_t *row = out + (size_t)x * p.T_y;
                    for (int c = lane; c < p.T_y; c += 32)
                        __stcs(row + c, (c >= s && c < e) ? one : 0u);
                }
            }
            if (ww == 0) MAS_STAMP(25);
            __syncwarp();
            if (lane == 0) mbar_arrive(bar_bempty + 8u * buf);
        }
    }
}


// ------------------------------------------------------------------ fallback for tall inputs
// Inputs the wavefront kernel does not take (T_x > 1344 tokens, or a stage ring that does not fit shared memory): one CTA per
// utterance, the DP column in shared memory, one __syncthreads per mel frame, a single-thread backtrack.  The same arithmetic
// (core.pyx:52-92: one __fadd_rn per cell, `prev >= same` ties to the diagonal, forced diagonal at x == y, penalty outside the
// band) and the same bit words parked at the head of the utterance's own output slab; ~1 us per frame, i.e. a correctness path for
// shapes the reference accepts (it takes any T_x), not a fast one.
__global__ void __launch_bounds__(256)
mas_tall_kernel(const MasParams p) {
    extern __shared__ __align__(16) uint8_t tall_smem[];
    float *qa = reinterpret_cast<float *>(tall_smem), *qb = qa + p.T_x;
    uint32_t *bw = reinterpret_cast<uint32_t *>(qb + p.T_x);          // this chunk's "came from x-1" bits, one word per token
    int *cs = reinterpret_cast<int *>(bw + p.T_x);                    // [T_x + 1] first frame of every token
    const int tid = threadIdx.x;
    const size_t n = (size_t)p.T_x * p.T_y;
    const float P = p.penalty;
    for (int b = blockIdx.x; b < p.B; b += gridDim.x) {
        const int tx = p.t_xs[b], ty = p.t_ys[b];
        const bool valid = tx >= 1 && tx <= ty && tx <= p.T_x && ty <= p.T_y;
        uint32_t *out = p.out + (size_t)b * n;
        if (valid) {
            const float *V = p.value + (size_t)b * n;
            const float *Mk = p.mask ? p.mask + (size_t)b * n : nullptr;
            for (int x = tid; x < tx; x += 256) { qa[x] = P; bw[x] = 0u; }
            __syncthreads();
            float *prev = qa, *cur = qb;
            for (int y = 0; y < ty; ++y) {
                const int lo = max(0, tx + y - ty), hi = min(tx, y + 1);
                const bool flush = (y & 31) == 31 || y == ty - 1;
                for (int x = tid; x < tx; x += 256) {                 // a token stays with its thread: bw[x] needs no barrier
                    float q = P;
                    if (x >= lo && x < hi) {
                        const float pv = (x == 0) ? (y == 0 ? 0.0f : P) : prev[x - 1];
                        const float sm = (x == y) ? P : (y > 0 ? prev[x] : P);
                        const bool diag = (pv >= sm) || (x == y);
                        float sc = V[(size_t)x * p.T_y + y];
                        if (Mk) sc = __fmul_rn(sc, Mk[(size_t)x * p.T_y + y]);
                        q = __fadd_rn(diag ? pv : sm, sc);
                        if (diag) bw[x] |= 1u << (y & 31);
                    }
                    cur[x] = q;
                    if (flush) { out[(size_t)(y >> 5) * p.T_x + x] = bw[x]; bw[x] = 0u; }
                }
                __syncthreads();
                float *t = prev; prev = cur; cur = t;
            }
            if (tid == 0) {                                           // core.pyx:87-92
                int idx = tx - 1;
                cs[tx] = ty;
                for (int y = ty - 1; y > 0; --y) {
                    const uint32_t wv = __ldcg(out + (size_t)(y >> 5) * p.T_x + idx);
                    if (idx > 0 && (idx == y || ((wv >> (y & 31)) & 1u))) { cs[idx] = y; --idx; }
                }
                for (int x = idx; x >= 0; --x) cs[x] = 0;             // idx is 0 here for valid lengths (the forced diagonal)
            }
        }
        __syncthreads();
        for (size_t i = tid; i < n; i += 256) __stcs(out + i, 0u);
        __syncthreads();
        if (valid) {
            for (int x = tid; x < tx; x += 256) {
                uint32_t *row = out + (size_t)x * p.T_y;
                for (int c = cs[x]; c < cs[x + 1]; ++c) __stcs(row + c, p.one_bits);
            }
        } else if (tid == 0 && p.status != nullptr) {
            atomicAdd(p.status, 1);
        }
        if (p.durations != nullptr) {
            int *d = p.durations + (size_t)b * p.T_x;
            for (int x = tid; x < p.T_x; x += 256) d[x] = (valid && x < tx) ? cs[x + 1] - cs[x] : 0;
        }
        __syncthreads();
    }
}

// t_xs[b] = (int) sum_x mask[b][x][0], t_ys[b] = (int) sum_y mask[b][0][y]  (__init__.py:52-53)
__global__ void __launch_bounds__(128)
mas_lengths_kernel(const float *__restrict__ mask, int T_x, int T_y, int *__restrict__ t_xs,
                   int *__restrict__ t_ys) {
    const float *m = mask + (size_t)blockIdx.x * T_x * T_y;
    float sx = 0.f, sy = 0.f;
    for (int x = threadIdx.x; x < T_x; x += blockDim.x) sx += m[(size_t)x * T_y];
    for (int y = threadIdx.x; y < T_y; y += blockDim.x) sy += m[y];
    __shared__ float red[2][4];
    for (int o = 16; o > 0; o >>= 1) {
        sx += __shfl_xor_sync(0xffffffffu, sx, o);
        sy += __shfl_xor_sync(0xffffffffu, sy, o);
    }
    if ((threadIdx.x & 31) == 0) { red[0][threadIdx.x >> 5] = sx; red[1][threadIdx.x >> 5] = sy; }
    __syncthreads();
    if (threadIdx.x == 0) {
        t_xs[blockIdx.x] = (int)(red[0][0] + red[0][1] + red[0][2] + red[0][3]);
        t_ys[blockIdx.x] = (int)(red[1][0] + red[1][1] + red[1][2] + red[1][3]);
    }
}

// scores = value * mask (__init__.py:45) for shapes whose fused-mask plan does not fit on chip
__global__ void __launch_bounds__(256)
mas_apply_mask_kernel(float *__restrict__ out, const float *__restrict__ value,
                      const float *__restrict__ mask, size_t n) {
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride)
        out[i] = __fmul_rn(value[i], mask[i]);
}

}  // namespace mb200

// ====================================================================== host side
namespace mb200 {

struct MasPlan {
    int R, ndp, nstage, txp, n_chunks, use_tma, vec_store, threads, early_zero;
    uint32_t smem;
};

static int env_int(const char *name, int dflt) {
    const char *v = getenv(name);
    return (v && *v) ? atoi(v) : dflt;
}

// Picks rows-per-lane R (odd, so the 128B-swizzled LDS.128 pattern is conflict-free), the
// number of DP warps and the ring depth.  Measured on B200: R=3 is the sweet spot (a warp's
// per-frame instruction count stays near its latency floor), R=1 only for <= 2 row groups.
// Tuning overrides (experiments only): MB200_MAS_R, MB200_MAS_NSTAGE.
static int mas_make_plan(int B, int T_x, int T_y, bool has_mask, bool aligned16, MasPlan *pl) {
    const int n_chunks = (T_y + kChunk - 1) / kChunk;
    const int g = (T_x + 31) / 32, txp = 32 * g;
    int R = env_int("MB200_MAS_R", 0);
    if (R <= 0) {
        if (g <= 2) R = 1;
        for (int r = 3; r <= 13 && R <= 0; r += 2)
            if ((g + r - 1) / r <= max_dp_warps(r)) R = r;
        if (R <= 0) return MB200_ERR_UNSUPPORTED;
    }
    if (R < 1 || R > 13 || (R & 1) == 0) return MB200_ERR_INVALID_ARG;
    const int ndp = (g + R - 1) / R;
    if (ndp > max_dp_warps(R)) return MB200_ERR_UNSUPPORTED;
    const int stage_bytes = txp * 128 * (has_mask ? 2 : 1);
    // ring depth: ~64 KB of stages per CTA so that several CTAs share an SM; 3 stages when
    // two CTAs still fit; whatever fits for very tall inputs
    int ns = (64 * 1024) / stage_bytes;
    if (ns > kMaxStages) ns = kMaxStages;
    if (ns < 2) ns = 2;
    if (ns < 3 && mas_smem_layout(has_mask, ndp, 3, txp).total + kSmemBlockReserve <= (uint32_t)kSmemPerSM / 2) ns = 3;
    // at most one utterance per SM: nothing to co-schedule, spend the shared memory on ring depth
    const bool single_wave = env_int("MB200_MAS_SINGLE_WAVE", B <= kNumSMsB200 ? 1 : 0) != 0;
    if (single_wave) { ns = (200 * 1024) / stage_bytes; if (ns > 6) ns = 6; if (ns < 2) ns = 2; }
    const int forced = env_int("MB200_MAS_NSTAGE", 0);
    if (forced >= 1 && forced <= kMaxStages) ns = forced;
    for (; ns >= 1; --ns) {
        MasSmem L = mas_smem_layout(has_mask, ndp, ns, txp);
        if (L.total <= (uint32_t)kMaxSmemPerBlock) {
            pl->R = R; pl->ndp = ndp; pl->nstage = ns; pl->txp = txp; pl->n_chunks = n_chunks;
            pl->use_tma = aligned16 ? 1 : 0; pl->vec_store = 0; pl->smem = L.total;
            pl->threads = 32 * (ndp + 1 + kWriterWarps);
            pl->early_zero = env_int("MB200_MAS_EARLY_ZERO", single_wave ? 1 : 0);
            return MB200_OK;
        }
    }
    return MB200_ERR_UNSUPPORTED;
}

typedef void (*MasKernelFn)(const CUtensorMap, const CUtensorMap, const MasParams);

template <bool HAS_MASK>
static MasKernelFn mas_pick_kernel(int R) {
    switch (R) {
        case 1: return mas_kernel<1, HAS_MASK>;
        case 3: return mas_kernel<3, HAS_MASK>;
        case 5: return mas_kernel<5, HAS_MASK>;
        case 7: return mas_kernel<7, HAS_MASK>;
        case 9: return mas_kernel<9, HAS_MASK>;
        case 11: return mas_kernel<11, HAS_MASK>;
        case 13: return mas_kernel<13, HAS_MASK>;
    }
    return nullptr;
}

// Multi-wave launches (more utterances than resident CTAs) are latency-bound per CTA -- one utterance's DP is a dependent chain --
// so HBM throughput comes from utterances in flight per SM, not from ring depth: shrink the ring (down to 2 stages) whenever that
// lets one more CTA share the SM.  Measured at B = 1024 (profiles/r2_exp8_mas_ctas_per_sm.log): 128x800 3 -> 4 CTAs/SM 0.63 -> 0.69
// of the HBM peak, 200x1000 2 -> 3 CTAs/SM 0.73 -> 0.78; deeper rings at fewer CTAs/SM only lose (6 stages, 2 CTAs/SM: 0.54).
// A fifth CTA per SM needs <= 58 registers: the R = 3 kernel then spills inside the DP loop and drops to 0.585 (r2_exp12, reverted).
static void mas_pick_ring(MasKernelFn fn, bool has_mask, int B, MasPlan *pl) {
    if (env_int("MB200_MAS_NSTAGE", 0) > 0) return;
    int dev = 0, sms = kNumSMsB200, occ = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) { (void)cudaGetLastError(); return; }
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, fn, pl->threads, pl->smem) != cudaSuccess) { (void)cudaGetLastError(); return; }
    if ((long long)sms * occ >= B) return;                  // every utterance already has a CTA of its own
    int best_ns = pl->nstage, best_occ = occ;
    uint32_t best_smem = pl->smem;
    for (int ns = pl->nstage - 1; ns >= 2; --ns) {
        const MasSmem L = mas_smem_layout(has_mask, pl->ndp, ns, pl->txp);
        int o = 0;
        if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&o, fn, pl->threads, L.total) != cudaSuccess) { (void)cudaGetLastError(); break; }
        if (o > best_occ) { best_occ = o; best_ns = ns; best_smem = L.total; }
    }
    pl->nstage = best_ns; pl->smem = best_smem;
}

static int mas_grid(MasKernelFn fn, const MasPlan &pl, int B, int *ctas_per_sm) {
    int dev = 0, sms = kNumSMsB200, occ = 1;
    if (cudaGetDevice(&dev) == cudaSuccess)
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, fn, pl.threads, pl.smem) != cudaSuccess || occ < 1)
        occ = 1;
    if (ctas_per_sm) *ctas_per_sm = occ;
    long long g = (long long)sms * occ;
    return (int)(g < B ? g : B);
}

}  // namespace mb200

using namespace mb200;

static int mas_launch(
    void *paths_dev, int path_dtype, const float *values_dev, const float *mask_dev,
    const int *t_xs_dev, const int *t_ys_dev, int B, int T_x, int T_y, float penalty,
    int *status_dev, int *durations_dev, void *stream) {
    if (B < 0 || T_x <= 0 || T_y <= 0) return MB200_ERR_INVALID_ARG;
    if (B == 0) return MB200_OK;
    if (!paths_dev || !values_dev || !t_xs_dev || !t_ys_dev) return MB200_ERR_INVALID_ARG;
    if (path_dtype != MB200_PATH_I32 && path_dtype != MB200_PATH_F32) return MB200_ERR_INVALID_ARG;
    if ((long long)B * T_x > 0x7fffffffLL) return MB200_ERR_UNSUPPORTED;   // TMA row coordinate is int32

    const bool has_mask = mask_dev != nullptr;
    const bool aligned16 = (T_y % 4 == 0) && ((uintptr_t)values_dev % 16 == 0) &&
                           (!has_mask || (uintptr_t)mask_dev % 16 == 0);
    MasPlan pl;
    int rc = mas_make_plan(B, T_x, T_y, has_mask, aligned16, &pl);
    if (rc == MB200_ERR_UNSUPPORTED) {
        // taller than the wavefront kernel takes: the one-CTA-per-utterance fallback (any T_x whose DP columns fit shared memory)
        const size_t smem = (size_t)(4 * T_x + 4) * 4;
        if (smem > (size_t)kMaxSmemPerBlock) return MB200_ERR_UNSUPPORTED;
        MB200_CUDA_TRY(cudaFuncSetAttribute(mas_tall_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        MasParams p = {};
        p.value = values_dev; p.mask = mask_dev; p.out = (uint32_t *)paths_dev;
        p.t_xs = t_xs_dev; p.t_ys = t_ys_dev; p.status = status_dev; p.durations = durations_dev;
        p.B = B; p.T_x = T_x; p.T_y = T_y; p.penalty = penalty;
        p.one_bits = (path_dtype == MB200_PATH_F32) ? 0x3f800000u : 1u;
        mas_tall_kernel<<<B < 2 * kNumSMsB200 ? B : 2 * kNumSMsB200, 256, smem, (cudaStream_t)stream>>>(p);
        MB200_CUDA_TRY(cudaGetLastError());
        return MB200_OK;
    }
    if (rc != MB200_OK) return rc;
    pl.vec_store = (T_y % 4 == 0) && ((uintptr_t)paths_dev % 16 == 0);

    MasKernelFn fn = has_mask ? mas_pick_kernel<true>(pl.R) : mas_pick_kernel<false>(pl.R);
    MB200_CUDA_TRY(cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)pl.smem));
    mas_pick_ring(fn, has_mask, B, &pl);

    CUtensorMap tm_v, tm_m;
    memset(&tm_v, 0, sizeof(tm_v));
    if (pl.use_tma) {
        rc = encode_tmap_2d_f32(&tm_v, values_dev, (uint64_t)B * T_x, (uint64_t)T_y, kBoxRows, kChunk);
        if (rc != MB200_OK) return rc;
    }
    tm_m = tm_v;
    if (pl.use_tma && has_mask) {
        rc = encode_tmap_2d_f32(&tm_m, mask_dev, (uint64_t)B * T_x, (uint64_t)T_y, kBoxRows, kChunk);
        if (rc != MB200_OK) return rc;
    }

    MasParams p;
    p.value = values_dev; p.mask = mask_dev; p.out = (uint32_t *)paths_dev;
    p.t_xs = t_xs_dev; p.t_ys = t_ys_dev; p.status = status_dev; p.durations = durations_dev;
    p.B = B; p.T_x = T_x; p.T_y = T_y; p.penalty = penalty;
    p.one_bits = (path_dtype == MB200_PATH_F32) ? 0x3f800000u : 1u;
    p.ndp = pl.ndp; p.nstage = pl.nstage; p.txp = pl.txp; p.n_chunks = pl.n_chunks;
    p.use_tma = pl.use_tma; p.vec_store = pl.vec_store; p.early_zero = pl.early_zero;
    p.debug = env_int("MB200_MAS_DEBUG", 0);

    const int grid = mas_grid(fn, pl, B, nullptr);
    fn<<<grid, pl.threads, pl.smem, (cudaStream_t)stream>>>(tm_v, tm_m, p);
    MB200_CUDA_TRY(cudaGetLastError());
    return MB200_OK;
}

extern "C" int matcha_b200_mas_compute_batch_alignments(
    void *paths_dev, int path_dtype, const float *values_dev, const float *mask_dev,
    const int *t_xs_dev, const int *t_ys_dev, int B, int T_x, int T_y, float penalty,
    int *status_dev, void *stream) {
    return mas_launch(paths_dev, path_dtype, values_dev, mask_dev, t_xs_dev, t_ys_dev, B, T_x, T_y, penalty,
                      status_dev, nullptr, stream);
}

extern "C" int matcha_b200_mas_align(
    void *paths_dev, int path_dtype, int *durations_dev, const float *values_dev, const float *mask_dev,
    const int *t_xs_dev, const int *t_ys_dev, int B, int T_x, int T_y, float penalty,
    int *status_dev, void *stream) {
    return mas_launch(paths_dev, path_dtype, values_dev, mask_dev, t_xs_dev, t_ys_dev, B, T_x, T_y, penalty,
                      status_dev, durations_dev, stream);
}

extern "C" int matcha_b200_mas_query_launch(int B, int T_x, int T_y, int has_mask, int aligned16, int out[8]) {
    if (B <= 0 || T_x <= 0 || T_y <= 0 || !out) return MB200_ERR_INVALID_ARG;
    MasPlan pl;
    int rc = mas_make_plan(B, T_x, T_y, has_mask != 0, aligned16 != 0, &pl);
    if (rc != MB200_OK) return rc;
    MasKernelFn fn = has_mask ? mas_pick_kernel<true>(pl.R) : mas_pick_kernel<false>(pl.R);
    int occ = 0;
    if (cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)pl.smem) != cudaSuccess) {
        (void)cudaGetLastError();   // no device: report the plan, leave grid/occupancy at 0
        out[0] = 0;
    } else {
        mas_pick_ring(fn, has_mask != 0, B, &pl);
        out[0] = mas_grid(fn, pl, B, &occ);
    }
    out[1] = pl.threads; out[2] = (int)pl.smem; out[3] = pl.R; out[4] = pl.nstage;
    out[5] = pl.use_tma; out[6] = occ; out[7] = pl.ndp;
    return MB200_OK;
}

// Which kernel mas_launch takes for a shape: the same plan logic, no device needed.
extern "C" int matcha_b200_mas_query_path(int B, int T_x, int T_y, int has_mask, int aligned16) {
    if (B <= 0 || T_x <= 0 || T_y <= 0) return MB200_ERR_INVALID_ARG;
    if ((long long)B * T_x > 0x7fffffffLL) return MB200_ERR_UNSUPPORTED;
    MasPlan pl;
    int rc = mas_make_plan(B, T_x, T_y, has_mask != 0, aligned16 != 0, &pl);
    if (rc == MB200_OK) return MB200_MAS_PATH_WAVEFRONT;
    if (rc != MB200_ERR_UNSUPPORTED) return rc;
    if (has_mask && mas_make_plan(B, T_x, T_y, false, aligned16 != 0, &pl) == MB200_OK) return MB200_MAS_PATH_TALL_MASK_ONLY;
    if ((size_t)(4 * (size_t)T_x + 4) * 4 > (size_t)kMaxSmemPerBlock) return MB200_ERR_UNSUPPORTED;
    return MB200_MAS_PATH_TALL;
}

extern "C" int matcha_b200_mas_lengths_from_mask(const float *mask_dev, int B, int T_x, int T_y,
                                                  int *t_xs_dev, int *t_ys_dev, void *stream) {
    if (B < 0 || T_x <= 0 || T_y <= 0) return MB200_ERR_INVALID_ARG;
    if (B == 0) return MB200_OK;
    if (!mask_dev || !t_xs_dev || !t_ys_dev) return MB200_ERR_INVALID_ARG;
    mas_lengths_kernel<<<B, 128, 0, (cudaStream_t)stream>>>(mask_dev, T_x, T_y, t_xs_dev, t_ys_dev);
    MB200_CUDA_TRY(cudaGetLastError());
    return MB200_OK;
}

extern "C" int matcha_b200_mas_apply_mask(float *scores_dev, const float *values_dev, const float *mask_dev,
                                           long long n, void *stream) {
    if (n < 0) return MB200_ERR_INVALID_ARG;
    if (n == 0) return MB200_OK;
    if (!scores_dev || !values_dev || !mask_dev) return MB200_ERR_INVALID_ARG;
    long long blocks = (n + 255) / 256;
    if (blocks > 148LL * 16) blocks = 148LL * 16;
    mas_apply_mask_kernel<<<(int)blocks, 256, 0, (cudaStream_t)stream>>>(scores_dev, values_dev, mask_dev, (size_t)n);
    MB200_CUDA_TRY(cudaGetLastError());
    return MB200_OK;
}

// ------------------------------------------------------------------ durations without a caller-side slab
// matcha_b200_mas_durations_static: the same kernels as matcha_b200_mas_align, for a caller that wants only the per-token
// durations and has no [B][T_x][T_y] buffer to lend.  The wavefront kernel needs that slab whatever the caller asks for: the
// "came from x-1" bit words of the DP are parked in it until the backtrack has read them, and the dense 0/1 path is laid over
// them afterwards.  Here the slab is a fixed array in the library's own device image, so the call performs no allocation.
// Who needs this: a host thread that runs alignments beside a thread that is capturing one of this library's decoder solves
// into a CUDA graph (cudaMalloc from ANY thread is illegal during a global-mode capture and poisons it); a plain C / ctypes
// binding of the reference's data-prep scripts that shares the GPU with a training process whose caching allocator already
// holds nearly all of the memory, so a 6.5 MB cudaMalloc may fail or stall on the implicit device synchronisation; and
// latency-sensitive serving loops that forbid cudaMalloc / cudaFree on the request path for the same synchronisation.
//
// In HOST code `g_mas_static_slab` is only a handle for the runtime's symbol table, not a device address, so it is never
// handed to a kernel; each device's real address comes from cudaGetSymbolAddress, asked once per device with that device
// current (every device that loads the library has its own copy of the array), and is cached below.
static __device__ __align__(256) uint32_t g_mas_static_slab[MB200_MAS_STATIC_MAX_CELLS];

namespace {
constexpr int kSlabMaxDevices = 64;
std::mutex g_slab_mu;
uint32_t *g_slab_addr[kSlabMaxDevices] = {};   // device address of g_mas_static_slab on device i, nullptr = not asked yet

// Address of the current device's copy of the slab (first call per device: one cudaGetSymbolAddress, no allocation).
int mas_static_slab(uint32_t **out) {
    int dev = -1;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= kSlabMaxDevices) { (void)cudaGetLastError(); return MB200_ERR_NO_DEVICE; }
    std::lock_guard<std::mutex> lock(g_slab_mu);
    if (g_slab_addr[dev] == nullptr) {
        void *ptr = nullptr;
        MB200_CUDA_TRY(cudaGetSymbolAddress(&ptr, g_mas_static_slab));
        if (ptr == nullptr) return MB200_ERR_CUDA;
        g_slab_addr[dev] = static_cast<uint32_t *>(ptr);
    }
    *out = g_slab_addr[dev];
    return MB200_OK;
}
}  // namespace

extern "C" long long matcha_b200_mas_durations_static_max_cells(void) { return (long long)MB200_MAS_STATIC_MAX_CELLS; }

extern "C" int matcha_b200_mas_durations_static(
    int *durations_dev, const float *values_dev, const float *mask_dev, const int *t_xs_dev, const int *t_ys_dev,
    int B, int T_x, int T_y, float penalty, int *status_dev, void *stream) {
    if (B < 0 || T_x <= 0 || T_y <= 0) return MB200_ERR_INVALID_ARG;
    if (B == 0) return MB200_OK;
    if (!durations_dev || !values_dev || !t_xs_dev || !t_ys_dev) return MB200_ERR_INVALID_ARG;
    // the whole batch's slab must fit the fixed array: refuse, never truncate or spill into an allocation
    if ((long long)B * T_x * T_y > (long long)MB200_MAS_STATIC_MAX_CELLS) return MB200_ERR_UNSUPPORTED;
    uint32_t *slab = nullptr;
    const int rc = mas_static_slab(&slab);
    if (rc != MB200_OK) return rc;
    // Nothing in the slab has to start at zero: the kernels write every bit word before the backtrack reads it and overwrite
    // all B*T_x*T_y cells with the dense path, so whatever the previous call left behind is never read.
    return mas_launch(slab, MB200_PATH_I32, values_dev, mask_dev, t_xs_dev, t_ys_dev, B, T_x, T_y, penalty,
                      status_dev, durations_dev, stream);
}

// profiling aid, not part of the public header: copies the clock64 stamps out
extern "C" int matcha_b200_mas_debug_read(long long *out32) {
    MB200_CUDA_TRY(cudaMemcpyFromSymbol(out32, mb200::g_mas_dbg, sizeof(long long) * 32));
    return MB200_OK;
}
