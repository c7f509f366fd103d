// mas.cu -- monotonic alignment search on sm_100a.
//
// Replaces /root/reference/matcha/utils/monotonic_align/core.pyx:16-128 (forward DP over
// mel frames with the "ties go to token x-1" rule of :73, then the backtrack of :87-92).
// Results are bit-exact with the compiled reference; see DESIGN.md §MAS for the argument.
//
// Shape of the kernel (one persistent CTA = 6 warps, utterances dealt round-robin):
//   warp 0  DP warp.  Lane l owns R consecutive tokens (rows) x0..x0+R-1 of the current
//           row block (32*R tokens); one step per mel frame (column): a single shfl_up
//           brings Q[y-1][x0-1] from lane l-1, the other R-1 predecessors are registers.
//           The "came from x-1" bit of every cell is packed 32 frames per word and kept in
//           shared memory, so the backtrack never leaves the SM.
//   warp 1  producer.  Streams the fp32 score tile of each 32-frame chunk (only the rows
//           inside the DP band) into a ring of 128B-swizzled stages with TMA
//           (cp.async.bulk.tensor, completion on an mbarrier); when the tensor breaks
//           TMA's 16-byte rules (T_y % 4 != 0 or a misaligned base) the same warp fills
//           the same layout with 4-byte cp.async.
//   warps 2-5  writers.  Turn the per-token start frames found by the backtrack into the
//           dense 0/1 alignment with 16-byte streaming stores, overlapping the DP of the
//           CTA's next utterance.
#include "common.cuh"

namespace mb200 {

constexpr int kChunk = 32;        // mel frames per tile
constexpr int kBoxRows = 32;      // tokens per TMA box
constexpr int kBoxBytes = kBoxRows * kChunk * 4;
constexpr int kWriterWarps = 4;
constexpr int kThreads = 32 * (2 + kWriterWarps);
constexpr int kMaxStages = 8;

struct MasParams {
    const float *value;
    const float *mask;
    uint32_t *out;
    const int *t_xs;
    const int *t_ys;
    int *status;
    int B, T_x, T_y;
    float penalty;
    uint32_t one_bits;   // bit pattern of "1" in the output element type
    int nstage;
    int txp;             // row pitch of the bit matrix (multiple of 32*R)
    int n_chunks;        // ceil(T_y / 32)
    int use_tma;
    int vec_store;       // output rows are 16-byte aligned
    int multi;           // more than one row block per utterance may occur
};

struct ItemInfo { int tx, ty, valid, pad; };

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
template <int R, bool HAS_MASK, bool DIAG>
__device__ __forceinline__ void dp_chunk(float (&q)[R], uint32_t (&bits)[R], uint32_t stage,
                                         int lane, int y_base, int x0, float penalty,
                                         bool multi, bool first_rb, uint32_t bin, uint32_t bout) {
    uint32_t row_addr[R], row_sw[R];
#pragma unroll
    for (int r = 0; r < R; ++r) {
        int rho = R * lane + r;
        row_addr[r] = stage + rho * 128;
        row_sw[r] = (rho & 7) << 4;
    }
#pragma unroll 1
    for (int g = 0; g < 8; ++g) {
        float4 v[R];
#pragma unroll
        for (int r = 0; r < R; ++r) {
            uint32_t a = row_addr[r] + ((uint32_t)(g << 4) ^ row_sw[r]);
            v[r] = lds128(a);
            if (HAS_MASK) {
                float4 m = lds128(a + R * kBoxBytes);
                v[r].x *= m.x; v[r].y *= m.y; v[r].z *= m.z; v[r].w *= m.w;
            }
        }
        const int y0 = y_base + 4 * g;
        float4 upv = make_float4(0.f, 0.f, 0.f, 0.f), outv = upv;
        if (multi && !first_rb && lane == 0) upv = lds128(bin + y0 * 4);
        const uint32_t mg = 1u << (4 * g);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int y = y0 + j;
            const uint32_t mj = mg << j;
            float up = __shfl_up_sync(0xffffffffu, q[R - 1], 1);
            if (lane == 0) {
                if (multi && !first_rb) up = (j == 0) ? upv.x : (j == 1) ? upv.y : (j == 2) ? upv.z : upv.w;
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
            if (multi) {
                if (j == 0) outv.x = q[R - 1]; else if (j == 1) outv.y = q[R - 1];
                else if (j == 2) outv.z = q[R - 1]; else outv.w = q[R - 1];
            }
        }
        if (multi && lane == 31) {
            // Q[y][last row of the block] is what the next block reads as Q[y'-1][x0-1] at y' = y+1
            float *o = reinterpret_cast<float *>(__cvta_shared_to_generic(bout)) + y0 + 1;
            o[0] = outv.x; o[1] = outv.y; o[2] = outv.z; o[3] = outv.w;
        }
    }
}

// ------------------------------------------------------------------ shared-memory carve-up
struct MasSmem {
    uint32_t stages;   // nstage * stage_bytes, 1024-aligned
    uint32_t bits;     // n_chunks * txp words
    uint32_t qb;       // 2 * n_chunks*32 floats (row-block boundary rows), only if multi
    uint32_t cs;       // 2 * (txp + 4) ints: start frame of every token (+ end sentinel)
    uint32_t info;     // 2 * ItemInfo
    uint32_t bars;     // full[nstage], empty[nstage], bfull[2], bempty[2]
    uint32_t total;
};

__host__ __device__ inline MasSmem mas_smem_layout(int R, bool has_mask, int nstage, int txp,
                                                   int n_chunks, bool multi) {
    MasSmem s;
    uint32_t off = 0;
    const uint32_t stage_bytes = (uint32_t)R * kBoxBytes * (has_mask ? 2 : 1);
    s.stages = off; off += (uint32_t)nstage * stage_bytes;
    s.bits = off;   off += (uint32_t)n_chunks * (uint32_t)txp * 4u;
    s.qb = off;     off += multi ? 2u * (uint32_t)(n_chunks * kChunk + 4) * 4u : 0u;
    s.cs = off;     off += 2u * (uint32_t)(txp + 4) * 4u;
    s.info = off;   off += 2u * (uint32_t)sizeof(ItemInfo);
    off = (off + 7u) & ~7u;
    s.bars = off;   off += (2u * (uint32_t)nstage + 4u) * 8u;
    s.total = off + 1024u;   // slack for the manual 1024-byte alignment of the base
    return s;
}

// ------------------------------------------------------------------ the kernel
template <int R, bool HAS_MASK>
__global__ void __launch_bounds__(kThreads)
mas_kernel(const __grid_constant__ CUtensorMap tmap_v, const __grid_constant__ CUtensorMap tmap_m,
           const MasParams p) {
    extern __shared__ uint8_t smem_raw[];
    const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    uint8_t *gbase = smem_raw + (base - smem_u32(smem_raw));
    const MasSmem L = mas_smem_layout(R, HAS_MASK, p.nstage, p.txp, p.n_chunks, p.multi != 0);
    constexpr int RB = 32 * R;
    constexpr uint32_t kStageBytes = (uint32_t)R * kBoxBytes * (HAS_MASK ? 2 : 1);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t bar_full = base + L.bars;
    const uint32_t bar_empty = bar_full + 8u * p.nstage;
    const uint32_t bar_bfull = bar_empty + 8u * p.nstage;
    const uint32_t bar_bempty = bar_bfull + 16u;
    uint32_t *bits_g = reinterpret_cast<uint32_t *>(gbase + L.bits);
    int *cs_g = reinterpret_cast<int *>(gbase + L.cs);
    ItemInfo *info_g = reinterpret_cast<ItemInfo *>(gbase + L.info);

    if (threadIdx.x == 0) {
        for (int s = 0; s < p.nstage; ++s) {
            mbar_init(bar_full + 8u * s, p.use_tma ? 1u : 32u);
            mbar_init(bar_empty + 8u * s, 1u);
        }
        for (int i = 0; i < 2; ++i) {
            mbar_init(bar_bfull + 8u * i, 1u);
            mbar_init(bar_bempty + 8u * i, kWriterWarps);
        }
        mbar_fence_init();
    }
    __syncthreads();

    if (warp == 1) {
        // ============================ producer ============================
        if (p.use_tma && lane == 0) {
            tma_prefetch_desc(&tmap_v);
            if (HAS_MASK) tma_prefetch_desc(&tmap_m);
        }
        uint32_t it = 0;
        for (int b = blockIdx.x; b < p.B; b += gridDim.x) {
            const int tx = p.t_xs[b], ty = p.t_ys[b];
            if (!lengths_valid(tx, ty, p.T_x, p.T_y)) continue;
            const int n_rb = (tx + RB - 1) / RB;
            for (int rb = 0; rb < n_rb; ++rb) {
                const int r0 = rb * RB, r1 = min(r0 + RB, tx);
                int c0, c1;
                rb_chunk_range(r0, r1, tx, ty, c0, c1);
                for (int c = c0; c <= c1; ++c, ++it) {
                    const uint32_t s = it % (uint32_t)p.nstage, ph = (it / (uint32_t)p.nstage) & 1u;
                    const uint32_t stage = base + L.stages + s * kStageBytes;
                    const uint32_t full = bar_full + 8u * s;
                    // rows of this block inside the band over frames [32c, 32c+32)
                    const int lo = max(r0, tx + kChunk * c - ty);
                    const int hi = min(r1, kChunk * c + kChunk);
                    const int j0 = (lo - r0) / kBoxRows, j1 = (hi - 1 - r0) / kBoxRows;
                    if (lane == 0) mbar_wait(bar_empty + 8u * s, ph ^ 1u);
                    __syncwarp();
                    if (p.use_tma) {
                        if (lane == 0) {
                            mbar_arrive_expect_tx(full, (uint32_t)(j1 - j0 + 1) * kBoxBytes * (HAS_MASK ? 2u : 1u));
                            for (int j = j0; j <= j1; ++j) {
                                const int row = b * p.T_x + r0 + kBoxRows * j;
                                tma_load_2d(stage + j * kBoxBytes, &tmap_v, full, kChunk * c, row);
                                if (HAS_MASK)
                                    tma_load_2d(stage + (R + j) * kBoxBytes, &tmap_m, full, kChunk * c, row);
                            }
                        }
                    } else {
                        const int col = kChunk * c + lane;
                        if (col < p.T_y) {
                            const int rho1 = min((j1 + 1) * kBoxRows, r1 - r0);
                            for (int rho = j0 * kBoxRows; rho < rho1; ++rho) {
                                const size_t g = ((size_t)b * p.T_x + r0 + rho) * (size_t)p.T_y + col;
                                const uint32_t d = stage + rho * 128 + ((((uint32_t)lane >> 2) ^ (rho & 7)) << 4) + (lane & 3) * 4;
                                cp_async_4(d, p.value + g);
                                if (HAS_MASK) cp_async_4(d + R * kBoxBytes, p.mask + g);
                            }
                        }
                        cp_async_mbar_arrive_noinc(full);
                    }
                }
            }
        }
    } else if (warp == 0) {
        // ============================ DP + backtrack ============================
        uint32_t it = 0, k = 0;
        for (int b = blockIdx.x; b < p.B; b += gridDim.x, ++k) {
            const int tx = p.t_xs[b], ty = p.t_ys[b];
            const bool valid = lengths_valid(tx, ty, p.T_x, p.T_y);
            const uint32_t buf = k & 1u;
            int *cs = cs_g + buf * (p.txp + 4);
            if (lane == 0) mbar_wait(bar_bempty + 8u * buf, ((k >> 1) & 1u) ^ 1u);
            __syncwarp();
            if (valid) {
                const int n_rb = (tx + RB - 1) / RB;
                for (int rb = 0; rb < n_rb; ++rb) {
                    const int r0 = rb * RB, r1 = min(r0 + RB, tx);
                    int c0, c1;
                    rb_chunk_range(r0, r1, tx, ty, c0, c1);
                    const int cd = (r1 - 1) / kChunk;      // last chunk that holds a diagonal cell
                    const uint32_t bin = base + L.qb + ((rb & 1) ^ 1) * (uint32_t)(p.n_chunks * kChunk + 4) * 4u;
                    const uint32_t bout = base + L.qb + (rb & 1) * (uint32_t)(p.n_chunks * kChunk + 4) * 4u;
                    const int x0 = r0 + R * lane;
                    float q[R];
                    uint32_t bits[R];
#pragma unroll
                    for (int r = 0; r < R; ++r) q[r] = p.penalty;      // core.pyx:52-53
                    for (int c = c0; c <= c1; ++c, ++it) {
                        const uint32_t s = it % (uint32_t)p.nstage, ph = (it / (uint32_t)p.nstage) & 1u;
                        const uint32_t stage = base + L.stages + s * kStageBytes;
#pragma unroll
                        for (int r = 0; r < R; ++r) bits[r] = 0u;
                        mbar_wait(bar_full + 8u * s, ph);
                        if (c <= cd)
                            dp_chunk<R, HAS_MASK, true>(q, bits, stage, lane, kChunk * c, x0, p.penalty,
                                                        p.multi != 0, rb == 0, bin, bout);
                        else
                            dp_chunk<R, HAS_MASK, false>(q, bits, stage, lane, kChunk * c, x0, p.penalty,
                                                         p.multi != 0, rb == 0, bin, bout);
                        uint32_t *w = bits_g + (size_t)c * p.txp + x0;
#pragma unroll
                        for (int r = 0; r < R; ++r) w[r] = bits[r];
                        __syncwarp();
                        if (lane == 0) mbar_arrive(bar_empty + 8u * s);
                    }
                }
                __syncwarp();
                if (lane == 0) {
                    // backtrack, core.pyx:87-92: token idx owns frames [cs[idx], cs[idx+1])
                    int idx = tx - 1, y = ty - 1;
                    cs[tx] = ty;
                    while (idx > 0) {
                        int c = y >> 5;
                        uint32_t w = bits_g[(size_t)c * p.txp + idx] & (0xffffffffu >> (31 - (y & 31)));
                        while (w == 0u) { --c; w = bits_g[(size_t)c * p.txp + idx]; }
                        const int yp = (c << 5) + 31 - __clz(w);
                        cs[idx] = yp;
                        y = yp - 1;
                        --idx;
                    }
                    cs[0] = 0;
                }
            } else if (lane == 0 && p.status != nullptr) {
                atomicAdd(p.status, 1);
            }
            if (lane == 0) {
                info_g[buf].tx = tx; info_g[buf].ty = ty; info_g[buf].valid = valid ? 1 : 0;
                mbar_arrive(bar_bfull + 8u * buf);
            }
            __syncwarp();
        }
    } else {
        // ============================ writers ============================
        const int ww = warp - 2;
        uint32_t k = 0;
        for (int b = blockIdx.x; b < p.B; b += gridDim.x, ++k) {
            const uint32_t buf = k & 1u;
            const int *cs = cs_g + buf * (p.txp + 4);
            mbar_wait(bar_bfull + 8u * buf, (k >> 1) & 1u);
            const int tx = info_g[buf].valid ? info_g[buf].tx : 0;
            uint32_t *out = p.out + (size_t)b * p.T_x * p.T_y;
            const uint32_t one = p.one_bits;
            if (p.vec_store) {
                for (int x = ww; x < p.T_x; x += kWriterWarps) {
                    int s = 0, e = 0;
                    if (x < tx) { s = cs[x]; e = cs[x + 1]; }
                    uint4 *row = reinterpret_cast<uint4 *>(out + (size_t)x * p.T_y);
                    for (int c4 = lane; c4 * 4 < p.T_y; c4 += 32) {
                        const int c = c4 * 4;
                        uint4 v;
                        v.x = (c + 0 >= s && c + 0 < e) ? one : 0u;
                        v.y = (c + 1 >= s && c + 1 < e) ? one : 0u;
                        v.z = (c + 2 >= s && c + 2 < e) ? one : 0u;
                        v.w = (c + 3 >= s && c + 3 < e) ? one : 0u;
                        __stcs(row + c4, v);
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
            __syncwarp();
            if (lane == 0) mbar_arrive(bar_bempty + 8u * buf);
        }
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

}  // namespace mb200

// ====================================================================== host side
namespace mb200 {

struct MasPlan {
    int R, nstage, txp, n_chunks, multi, use_tma, vec_store;
    uint32_t smem;
};

// Ring depth: enough bytes in flight per CTA (~64 KB) without hogging the SM, so that small
// shapes co-reside several CTAs per SM; at least 3 stages when they fit.
static int stage_cap(int R, bool has_mask) {
    const int stage_bytes = R * kBoxBytes * (has_mask ? 2 : 1);
    int ns = (64 * 1024) / stage_bytes;
    if (ns < 3) ns = 3;
    if (ns > kMaxStages) ns = kMaxStages;
    return ns;
}

// Picks rows-per-lane R (odd, so the 128B-swizzled LDS.128 pattern is conflict-free) and the
// pipeline depth that fit 227 KB next to the bit matrix.  Prefers a single row block.
static int mas_make_plan(int T_x, int T_y, bool has_mask, bool aligned16, MasPlan *pl) {
    const int n_chunks = (T_y + kChunk - 1) / kChunk;
    int r_fit = (T_x + 31) / 32;
    if ((r_fit & 1) == 0) r_fit += 1;
    if (r_fit > 13) r_fit = 13;
    for (int R = r_fit; R >= 1; R -= 2) {
        const int rb = 32 * R;
        const int txp = (T_x + rb - 1) / rb * rb;
        const bool multi = txp > rb;
        for (int ns = stage_cap(R, has_mask); ns >= 2; --ns) {
            MasSmem L = mas_smem_layout(R, has_mask, ns, txp, n_chunks, multi);
            if (L.total <= (uint32_t)kMaxSmemPerBlock) {
                pl->R = R; pl->nstage = ns; pl->txp = txp; pl->n_chunks = n_chunks;
                pl->multi = multi ? 1 : 0; pl->use_tma = aligned16 ? 1 : 0;
                pl->vec_store = 0; pl->smem = L.total;
                return MB200_OK;
            }
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

static int mas_grid(MasKernelFn fn, const MasPlan &pl, int B, int *ctas_per_sm) {
    int dev = 0, sms = kNumSMsB200, occ = 1;
    if (cudaGetDevice(&dev) == cudaSuccess)
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, fn, kThreads, pl.smem) != cudaSuccess || occ < 1)
        occ = 1;
    if (ctas_per_sm) *ctas_per_sm = occ;
    long long g = (long long)sms * occ;
    return (int)(g < B ? g : B);
}

}  // namespace mb200

using namespace mb200;

extern "C" int matcha_b200_mas_compute_batch_alignments(
    void *paths_dev, int path_dtype, const float *values_dev, const float *mask_dev,
    const int *t_xs_dev, const int *t_ys_dev, int B, int T_x, int T_y, float penalty,
    int *status_dev, void *stream) {
    if (B < 0 || T_x <= 0 || T_y <= 0) return MB200_ERR_INVALID_ARG;
    if (B == 0) return MB200_OK;
    if (!paths_dev || !values_dev || !t_xs_dev || !t_ys_dev) return MB200_ERR_INVALID_ARG;
    if (path_dtype != MB200_PATH_I32 && path_dtype != MB200_PATH_F32) return MB200_ERR_INVALID_ARG;
    if ((long long)B * T_x > 0x7fffffffLL) return MB200_ERR_UNSUPPORTED;   // TMA row coordinate is int32

    const bool has_mask = mask_dev != nullptr;
    const bool aligned16 = (T_y % 4 == 0) && ((uintptr_t)values_dev % 16 == 0) &&
                           (!has_mask || (uintptr_t)mask_dev % 16 == 0);
    MasPlan pl;
    int rc = mas_make_plan(T_x, T_y, has_mask, aligned16, &pl);
    if (rc != MB200_OK) return rc;
    pl.vec_store = (T_y % 4 == 0) && ((uintptr_t)paths_dev % 16 == 0);

    MasKernelFn fn = has_mask ? mas_pick_kernel<true>(pl.R) : mas_pick_kernel<false>(pl.R);
    MB200_CUDA_TRY(cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)pl.smem));

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
    p.t_xs = t_xs_dev; p.t_ys = t_ys_dev; p.status = status_dev;
    p.B = B; p.T_x = T_x; p.T_y = T_y; p.penalty = penalty;
    p.one_bits = (path_dtype == MB200_PATH_F32) ? 0x3f800000u : 1u;
    p.nstage = pl.nstage; p.txp = pl.txp; p.n_chunks = pl.n_chunks;
    p.use_tma = pl.use_tma; p.vec_store = pl.vec_store; p.multi = pl.multi;

    const int grid = mas_grid(fn, pl, B, nullptr);
    fn<<<grid, kThreads, pl.smem, (cudaStream_t)stream>>>(tm_v, tm_m, p);
    MB200_CUDA_TRY(cudaGetLastError());
    return MB200_OK;
}

extern "C" int matcha_b200_mas_query_launch(int B, int T_x, int T_y, int has_mask, int aligned16, int out[8]) {
    if (B <= 0 || T_x <= 0 || T_y <= 0 || !out) return MB200_ERR_INVALID_ARG;
    MasPlan pl;
    int rc = mas_make_plan(T_x, T_y, has_mask != 0, aligned16 != 0, &pl);
    if (rc != MB200_OK) return rc;
    MasKernelFn fn = has_mask ? mas_pick_kernel<true>(pl.R) : mas_pick_kernel<false>(pl.R);
    int occ = 0;
    if (cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)pl.smem) != cudaSuccess) {
        (void)cudaGetLastError();   // no device: report the plan, leave grid/occupancy at 0
        out[0] = 0;
    } else {
        out[0] = mas_grid(fn, pl, B, &occ);
    }
    out[1] = kThreads; out[2] = (int)pl.smem; out[3] = pl.R; out[4] = pl.nstage;
    out[5] = pl.use_tma; out[6] = occ; out[7] = pl.multi;
    return MB200_OK;
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
