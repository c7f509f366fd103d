// align.cu -- the callers either side of the two hot paths, sm_100a.
//
//   log_prior_kernel      matcha_tts.py:277-282   mu_x, y  ->  the [B][T_x][T_y] log-likelihood MAS consumes
//   expand_kernel         matcha_tts.py:287, 314-315, 320-321 (training) and :224-229 + model.py:77-114
//                         (inference generate_path): durations -> attn / mu_y / logw_ / prior-loss sum
//   cfm_mix_kernel        flow_matching.py:138     phi_t = (1-(1-sigma)t) z + t x1
//   cfm_mse_kernel        flow_matching.py:142-149 sum (u_pred - (x1-(1-sigma)z))^2, sum(mask)
//
// All of these are HBM-bound byte movers (no tensor-core shape is forced on them): coalesced
// 16-byte accesses where the row pitch allows, grids sized in multiples of the 148 SMs, fp64 for
// the two scalar reductions.  fp32 expressions are written with __fmul_rn/__fadd_rn in the order
// torch evaluates them so that phi_t, attn, mu_y and the durations are bit-identical to the
// reference on the same inputs; only the reductions (loss sums, the K=80 contraction of the
// log-prior) differ by summation order, and the tests state that tolerance.
#include <math.h>
#include <string.h>

#include "common.cuh"

namespace mb200 {

// ------------------------------------------------------------------------------------------
// log_prior[b][x][y] = ((ysq[y] - ymu[x][y]) + musq[x]) + cst           (matcha_tts.py:278-282)
//   ysq[y]    = sum_c (-0.5) * (y[c][y]*y[c][y])         the torch.matmul(factor^T, y**2) row
//   ymu[x][y] = sum_c (2*(-0.5*mu[c][x])) * y[c][y]
//   musq[x]   = sum_c (-0.5) * (mu[c][x]*mu[c][x])
// Tile: 64 tokens x 128 frames per CTA, 256 threads, thread = 8 tokens x 4 consecutive frames;
// both operand tiles sit in shared memory ([C][64] and [C][128] floats, C <= 80+ by chunks of 16).
constexpr int kLpTx = 64, kLpTy = 128, kLpCk = 16;

__global__ void __launch_bounds__(256)
log_prior_kernel(float *__restrict__ out, const float *__restrict__ mu, const float *__restrict__ y,
                 const int *__restrict__ t_xs, const int *__restrict__ t_ys,
                 int C, int T_x, int T_y, float cst) {
    __shared__ __align__(16) float s_mu[kLpCk][kLpTx];
    __shared__ __align__(16) float s_a[kLpCk][kLpTx];
    __shared__ __align__(16) float s_y[kLpCk][kLpTy];
    __shared__ float s_musq[kLpTx];
    __shared__ float s_ysq[kLpTy];
    const int b = blockIdx.z, x0 = blockIdx.y * kLpTx, y0 = blockIdx.x * kLpTy;
    if (t_xs != nullptr && (x0 >= t_xs[b] || y0 >= t_ys[b])) return;   // MAS never reads outside the valid block
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const float *mub = mu + (size_t)b * C * T_x;
    const float *yb = y + (size_t)b * C * T_y;

    float acc[8][4];
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
    float musq = 0.f, ysq = 0.f;      // thread tid < 64 owns token tid; thread tid < 128 owns frame tid

    for (int c0 = 0; c0 < C; c0 += kLpCk) {
        __syncthreads();
        for (int i = tid; i < kLpCk * kLpTx; i += 256) {
            const int c = i / kLpTx, x = i % kLpTx;
            const float m = (c0 + c < C && x0 + x < T_x) ? __ldg(mub + (size_t)(c0 + c) * T_x + x0 + x) : 0.f;
            s_mu[c][x] = m;
            s_a[c][x] = __fmul_rn(2.0f, __fmul_rn(-0.5f, m));          // 2.0 * (factor * mu_x), once per element
        }
        for (int i = tid; i < kLpCk * kLpTy; i += 256) {
            const int c = i / kLpTy, yy = i % kLpTy;
            s_y[c][yy] = (c0 + c < C && y0 + yy < T_y) ? __ldg(yb + (size_t)(c0 + c) * T_y + y0 + yy) : 0.f;
        }
        __syncthreads();
        if (tid < kLpTx) {
#pragma unroll
            for (int c = 0; c < kLpCk; ++c) {
                const float m = s_mu[c][tid];
                musq = __fadd_rn(musq, __fmul_rn(-0.5f, __fmul_rn(m, m)));
            }
        }
        if (tid < kLpTy) {
#pragma unroll
            for (int c = 0; c < kLpCk; ++c) {
                const float v = s_y[c][tid];
                ysq = __fadd_rn(ysq, __fmul_rn(-0.5f, __fmul_rn(v, v)));
            }
        }
#pragma unroll
        for (int c = 0; c < kLpCk; ++c) {
            const float4 yv = *reinterpret_cast<const float4 *>(&s_y[c][lane * 4]);
            const float4 m0 = *reinterpret_cast<const float4 *>(&s_a[c][warp * 8]);
            const float4 m1 = *reinterpret_cast<const float4 *>(&s_a[c][warp * 8 + 4]);
            const float m[8] = {m0.x, m0.y, m0.z, m0.w, m1.x, m1.y, m1.z, m1.w};
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const float a = m[i];
                acc[i][0] = fmaf(a, yv.x, acc[i][0]);
                acc[i][1] = fmaf(a, yv.y, acc[i][1]);
                acc[i][2] = fmaf(a, yv.z, acc[i][2]);
                acc[i][3] = fmaf(a, yv.w, acc[i][3]);
            }
        }
    }
    if (tid < kLpTx) s_musq[tid] = musq;
    if (tid < kLpTy) s_ysq[tid] = ysq;
    __syncthreads();

    float *ob = out + (size_t)b * T_x * T_y;
    const int yc = y0 + lane * 4;
    const bool vec = (T_y % 4 == 0) && (((uintptr_t)out & 15) == 0);
    float q[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) q[j] = s_ysq[lane * 4 + j];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const int x = x0 + warp * 8 + i;
        if (x >= T_x) break;
        const float ms = s_musq[warp * 8 + i];
        float r[4];
#pragma unroll
        for (int j = 0; j < 4; ++j)
            r[j] = __fadd_rn(__fadd_rn(__fsub_rn(q[j], acc[i][j]), ms), cst);
        float *row = ob + (size_t)x * T_y;
        if (vec && yc + 3 < T_y) {
            *reinterpret_cast<float4 *>(row + yc) = make_float4(r[0], r[1], r[2], r[3]);
        } else {
#pragma unroll
            for (int j = 0; j < 4; ++j)
                if (yc + j < T_y) row[yc + j] = r[j];
        }
    }
}

// ------------------------------------------------------------------------------------------
// Block-wide fp64 sum -> one atomicAdd per CTA.
__device__ __forceinline__ void block_atomic_add_f64(double v, double *dst, double *s_red) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
    __syncthreads();
    if (lane == 0) s_red[warp] = v;
    __syncthreads();
    if (warp == 0) {
        double t = lane < nw ? s_red[lane] : 0.0;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
        if (lane == 0) atomicAdd(dst, t);
    }
}

// ------------------------------------------------------------------------------------------
// Duration expansion.  One CTA per (utterance, strip of 256 frames):
//   cum[x]   = (float) sum_{x' <= x} dur[x']  accumulated in fp64, rounded per element -- what
//              torch.cumsum does on fp32 CPU tensors (model.py:96)
//   idx[y]   = first token x < t_x with (float)y < cum[x], for y < t_y; none otherwise.  This is
//              build_alignment_path's  (y < cum[x]) - (y < cum[x-1])  times the mask (model.py:103-113)
//              for durations >= 0 (negative durations are outside the domain: the reference would
//              emit -1 entries there).
//   attn[x][y] = (idx[y] == x)                                       model.py:113 / core.pyx path
//   mu_y[c][y] = mu_x[c][idx[y]] or 0    = (attn^T mu_x^T)^T exactly: every column of attn is one-hot
//                                                                    matcha_tts.py:227-228, 314-315
//   logw[x]  = log(1e-8 + dur[x]) * (x < t_x)                        matcha_tts.py:287
//   prior[0] += sum 0.5*((y - mu_y)^2 + log 2pi) * (y < t_y),  prior[1] += #frames      :320-321
struct ExpandParams {
    const void *dur;       // [B][T_x] int32 or float
    const float *mu_x;     // [B][C][T_x] or null
    const float *y;        // [B][C][T_y] or null (prior loss)
    const int *t_xs, *t_ys;   // [B] or null (= T_x / T_y)
    float *attn;           // [B][T_x][T_y] or null
    float *mu_y;           // [B][C][T_y] or null
    float *logw;           // [B][T_x] or null
    double *prior;         // [2] or null
    int dur_is_int, B, C, T_x, T_y;
};

constexpr int kExpStrip = 256;

__global__ void __launch_bounds__(kExpStrip)
expand_kernel(const ExpandParams p) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    float *cum = reinterpret_cast<float *>(smem_raw);            // [T_x]
    __shared__ double s_red[8];
    __shared__ int s_idx[kExpStrip];
    const int b = blockIdx.y, y0 = blockIdx.x * kExpStrip, tid = threadIdx.x, lane = tid & 31;
    const int tx = p.t_xs ? min(max(p.t_xs[b], 0), p.T_x) : p.T_x;
    const int ty = p.t_ys ? min(max(p.t_ys[b], 0), p.T_y) : p.T_y;
    auto dur_at = [&](int x) -> float {
        return p.dur_is_int ? (float)__ldg(reinterpret_cast<const int *>(p.dur) + (size_t)b * p.T_x + x)
                            : __ldg(reinterpret_cast<const float *>(p.dur) + (size_t)b * p.T_x + x);
    };
    if (tid < 32) {                      // warp 0: chunked fp64 scan, lane = contiguous chunk of tokens
        const int per = (p.T_x + 31) / 32, lo = lane * per, hi = min(lo + per, p.T_x);
        double s = 0.0;
        for (int x = lo; x < hi; ++x) s += (double)dur_at(x);
        double incl = s;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const double t = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += t;
        }
        double run = incl - s;
        for (int x = lo; x < hi; ++x) { run += (double)dur_at(x); cum[x] = (float)run; }
    }
    __syncthreads();

    const int yy = y0 + tid;
    int idx = -1;
    if (yy < ty && tx > 0) {
        const float fy = (float)yy;
        int lo = 0, hi = tx;             // first x in [0, tx) with fy < cum[x]
        while (lo < hi) {
            const int mid = (lo + hi) >> 1;
            if (fy < cum[mid]) hi = mid; else lo = mid + 1;
        }
        if (lo < tx) idx = lo;
    }
    s_idx[tid] = idx;

    if (p.logw != nullptr && blockIdx.x == 0) {
        for (int x = tid; x < p.T_x; x += kExpStrip)
            p.logw[(size_t)b * p.T_x + x] = __fmul_rn(logf(__fadd_rn(1e-8f, dur_at(x))), x < tx ? 1.f : 0.f);
    }

    double acc = 0.0;
    if (p.mu_y != nullptr || p.prior != nullptr) {
        const float *mub = p.mu_x + (size_t)b * p.C * p.T_x;
        const float kLog2Pi = 1.8378770664093453f;             // (float) math.log(2*math.pi)
        if (yy < p.T_y) {
            const bool want_prior = p.prior != nullptr && yy < ty;
            for (int c0 = 0; c0 < p.C; c0 += 8) {                // 8 channels per pass: 16 loads in flight per thread
                float m[8], yv[8];
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    const int c = c0 + j;
                    m[j] = (idx >= 0 && c < p.C) ? __ldg(mub + (size_t)c * p.T_x + idx) : 0.f;
                    yv[j] = (want_prior && c < p.C) ? __ldcs(p.y + ((size_t)b * p.C + c) * p.T_y + yy) : 0.f;
                }
                double dpart = 0.0;
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    const int c = c0 + j;
                    if (c < p.C) {
                        if (p.mu_y != nullptr) __stcs(p.mu_y + ((size_t)b * p.C + c) * p.T_y + yy, m[j]);
                        const float d = __fsub_rn(yv[j], m[j]);
                        dpart += (double)__fmul_rn(0.5f, __fadd_rn(__fmul_rn(d, d), kLog2Pi));
                    }
                }
                if (want_prior) acc += dpart;
            }
        }
    }
    if (p.prior != nullptr) {
        block_atomic_add_f64(acc, p.prior, s_red);
        block_atomic_add_f64((yy < ty) ? 1.0 : 0.0, p.prior + 1, s_red);
    }

    if (p.attn != nullptr) {
        __syncthreads();
        float *ab = p.attn + (size_t)b * p.T_x * p.T_y;
        const int w = min(kExpStrip, p.T_y - y0);              // columns of this strip
        if (w > 0) {
            // warp per row: 32 lanes x 8 passes cover the strip with 128-byte stores
            const int warp = tid >> 5;
            for (int x = warp; x < p.T_x; x += kExpStrip / 32) {
                float *row = ab + (size_t)x * p.T_y + y0;
                for (int c = lane; c < w; c += 32) __stcs(row + c, s_idx[c] == x ? 1.f : 0.f);
            }
        }
    }
}

// ------------------------------------------------------------------------------------------
// flow_matching.py:138  phi_t = (1 - (1 - sigma_min) * t) * z + t * x1     (t per sample, [B,1,1])
__global__ void __launch_bounds__(256)
cfm_mix_kernel(float *__restrict__ phi, const float *__restrict__ x1, const float *__restrict__ z,
               const float *__restrict__ t, float one_minus_sigma, long long per_sample, int B) {
    const long long n = per_sample * B;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        const float tb = __ldg(t + i / per_sample);
        const float a = __fsub_rn(1.f, __fmul_rn(one_minus_sigma, tb));
        phi[i] = __fadd_rn(__fmul_rn(a, __ldg(z + i)), __fmul_rn(tb, __ldg(x1 + i)));
    }
}

// flow_matching.py:142-149  sums[0] += sum (u_pred - (x1 - (1-sigma) z))^2 ; sums[1] += sum mask
__global__ void __launch_bounds__(256)
cfm_mse_kernel(double *__restrict__ sums, const float *__restrict__ u_pred, const float *__restrict__ x1,
               const float *__restrict__ z, const float *__restrict__ mask, float one_minus_sigma,
               long long n, long long n_mask) {
    __shared__ double s_red[8];
    double a = 0.0, m = 0.0;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        const float u = __fsub_rn(__ldg(x1 + i), __fmul_rn(one_minus_sigma, __ldg(z + i)));
        const float d = __fsub_rn(__ldg(u_pred + i), u);
        a += (double)__fmul_rn(d, d);
    }
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n_mask; i += (long long)gridDim.x * blockDim.x)
        m += (double)__ldg(mask + i);
    block_atomic_add_f64(a, sums, s_red);
    block_atomic_add_f64(m, sums + 1, s_red);
}

}  // namespace mb200

using namespace mb200;

extern "C" int matcha_b200_align_log_prior(float *log_prior_dev, const float *mu_x_dev, const float *y_dev,
                                            const int *t_xs_dev, const int *t_ys_dev,
                                            int B, int C, int T_x, int T_y, void *stream) {
    if (B < 0 || C <= 0 || T_x <= 0 || T_y <= 0) return MB200_ERR_INVALID_ARG;
    if (B == 0) return MB200_OK;
    if (!log_prior_dev || !mu_x_dev || !y_dev || ((t_xs_dev == nullptr) != (t_ys_dev == nullptr)))
        return MB200_ERR_INVALID_ARG;
    if (B > 65535 || (T_x + kLpTx - 1) / kLpTx > 65535) return MB200_ERR_UNSUPPORTED;
    const float cst = (float)(-0.5 * log(2.0 * 3.141592653589793) * (double)C);      // matcha_tts.py:277
    dim3 grid((T_y + kLpTy - 1) / kLpTy, (T_x + kLpTx - 1) / kLpTx, B);
    log_prior_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(log_prior_dev, mu_x_dev, y_dev, t_xs_dev, t_ys_dev,
                                                            C, T_x, T_y, cst);
    MB200_CUDA_TRY(cudaGetLastError());
    return MB200_OK;
}

extern "C" int matcha_b200_align_expand(const void *durations_dev, int dur_is_int, const float *mu_x_dev,
                                         const float *y_dev, const int *t_xs_dev, const int *t_ys_dev,
                                         float *attn_dev, float *mu_y_dev, float *logw_dev, double *prior_sums_dev,
                                         int B, int C, int T_x, int T_y, void *stream) {
    if (B < 0 || C <= 0 || T_x <= 0 || T_y <= 0) return MB200_ERR_INVALID_ARG;
    if (B == 0) return MB200_OK;
    if (!durations_dev) return MB200_ERR_INVALID_ARG;
    if ((mu_y_dev || prior_sums_dev) && !mu_x_dev) return MB200_ERR_INVALID_ARG;
    if (prior_sums_dev && !y_dev) return MB200_ERR_INVALID_ARG;
    if (B > 65535 || (size_t)T_x * sizeof(float) > 200 * 1024) return MB200_ERR_UNSUPPORTED;
    ExpandParams p;
    p.dur = durations_dev; p.mu_x = mu_x_dev; p.y = y_dev; p.t_xs = t_xs_dev; p.t_ys = t_ys_dev;
    p.attn = attn_dev; p.mu_y = mu_y_dev; p.logw = logw_dev; p.prior = prior_sums_dev;
    p.dur_is_int = dur_is_int != 0; p.B = B; p.C = C; p.T_x = T_x; p.T_y = T_y;
    const size_t smem = (size_t)T_x * sizeof(float);
    if (smem > 48 * 1024)
        MB200_CUDA_TRY(cudaFuncSetAttribute(expand_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    if (prior_sums_dev) MB200_CUDA_TRY(cudaMemsetAsync(prior_sums_dev, 0, 2 * sizeof(double), (cudaStream_t)stream));
    dim3 grid((T_y + kExpStrip - 1) / kExpStrip, B);
    expand_kernel<<<grid, kExpStrip, smem, (cudaStream_t)stream>>>(p);
    MB200_CUDA_TRY(cudaGetLastError());
    return MB200_OK;
}

static int elementwise_grid(long long n) {
    long long blocks = (n + 255) / 256;
    const long long cap = (long long)kNumSMsB200 * 8;
    return (int)(blocks < cap ? (blocks < 1 ? 1 : blocks) : cap);
}

extern "C" int matcha_b200_cfm_loss_mix(float *phi_dev, const float *x1_dev, const float *z_dev, const float *t_dev,
                                         double sigma_min, int B, long long per_sample, void *stream) {
    if (B < 0 || per_sample <= 0) return MB200_ERR_INVALID_ARG;
    if (B == 0) return MB200_OK;
    if (!phi_dev || !x1_dev || !z_dev || !t_dev) return MB200_ERR_INVALID_ARG;
    cfm_mix_kernel<<<elementwise_grid(per_sample * B), 256, 0, (cudaStream_t)stream>>>(
        phi_dev, x1_dev, z_dev, t_dev, (float)(1.0 - sigma_min), per_sample, B);
    MB200_CUDA_TRY(cudaGetLastError());
    return MB200_OK;
}

extern "C" int matcha_b200_cfm_loss_mse(double *sums_dev, const float *u_pred_dev, const float *x1_dev,
                                         const float *z_dev, const float *mask_dev, double sigma_min,
                                         long long n, long long n_mask, void *stream) {
    if (n < 0 || n_mask < 0) return MB200_ERR_INVALID_ARG;
    if (!sums_dev || !u_pred_dev || !x1_dev || !z_dev || !mask_dev) return MB200_ERR_INVALID_ARG;
    MB200_CUDA_TRY(cudaMemsetAsync(sums_dev, 0, 2 * sizeof(double), (cudaStream_t)stream));
    if (n == 0) return MB200_OK;
    cfm_mse_kernel<<<elementwise_grid(n), 256, 0, (cudaStream_t)stream>>>(
        sums_dev, u_pred_dev, x1_dev, z_dev, mask_dev, (float)(1.0 - sigma_min), n, n_mask);
    MB200_CUDA_TRY(cudaGetLastError());
    return MB200_OK;
}
