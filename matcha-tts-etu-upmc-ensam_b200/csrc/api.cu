// api.cu -- library-wide pieces of the C ABI: error strings, the tensor-map encoder and the
// host-buffer MAS entry point (the numpy round trip of
// /root/reference/matcha/utils/monotonic_align/__init__.py:45-55).
#include <stdlib.h>
#include <string.h>

#include "common.cuh"

namespace mb200 {

thread_local int g_last_cuda_error = 0;

typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *,
                                  const cuuint64_t *, const cuuint64_t *, const cuuint32_t *,
                                  const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode_fn() {
    static EncodeTiledFn fn = nullptr;
    if (!fn) {
        void *p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            fn = (EncodeTiledFn)p;
    }
    return fn;
}

int encode_tmap_2d_f32(CUtensorMap *map, const void *base, uint64_t rows, uint64_t cols,
                       uint32_t box_rows, uint32_t box_cols) {
    EncodeTiledFn fn = get_encode_fn();
    if (!fn) return MB200_ERR_NO_DEVICE;
    cuuint64_t dims[2] = {cols, rows};
    cuuint64_t strides[1] = {cols * sizeof(float)};
    cuuint32_t box[2] = {box_cols, box_rows};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<void *>(base), dims, strides,
                    box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                    CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        g_last_cuda_error = (int)r;
        return MB200_ERR_CUDA;
    }
    return MB200_OK;
}

// bf16 tensor map (2-D or 3-D), 128B swizzle, zero fill outside the tensor.
int encode_tmap_bf16(CUtensorMap *map, const void *base, int ndim, const uint64_t *dims,
                     const uint64_t *strides_bytes, const uint32_t *box) {
    EncodeTiledFn fn = get_encode_fn();
    if (!fn) return MB200_ERR_NO_DEVICE;
    cuuint64_t d[3], st[2];
    cuuint32_t bx[3], es[3] = {1, 1, 1};
    for (int i = 0; i < ndim; ++i) { d[i] = dims[i]; bx[i] = box[i]; }
    for (int i = 0; i + 1 < ndim; ++i) st[i] = strides_bytes[i];
    CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, (cuuint32_t)ndim, const_cast<void *>(base), d, st, bx, es,
                    CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                    CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        g_last_cuda_error = (int)r;
        return MB200_ERR_CUDA;
    }
    return MB200_OK;
}

// Output tensor map for TMA stores from a [32 rows][box_cols] staging tile with 128-byte rows (128B swizzle) or 64-byte
// rows (64B swizzle), fp32 or 16-bit elements; 3-D [d2][d1][d0] with byte strides.
int encode_tmap_out(CUtensorMap *map, const void *base, int is_f32, const uint64_t *dims, const uint64_t *strides_bytes,
                    int box_cols) {
    EncodeTiledFn fn = get_encode_fn();
    if (!fn) return MB200_ERR_NO_DEVICE;
    cuuint64_t d[3] = {dims[0], dims[1], dims[2]}, st[2] = {strides_bytes[0], strides_bytes[1]};
    cuuint32_t bx[3] = {(cuuint32_t)box_cols, 32, 1}, es[3] = {1, 1, 1};
    const bool sw128 = box_cols * (is_f32 ? 4 : 2) == 128;
    CUresult r = fn(map, is_f32 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<void *>(base),
                    d, st, bx, es, CU_TENSOR_MAP_INTERLEAVE_NONE, sw128 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_64B,
                    CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        g_last_cuda_error = (int)r;
        return MB200_ERR_CUDA;
    }
    return MB200_OK;
}

}  // namespace mb200

using namespace mb200;

extern "C" const char *matcha_b200_error_string(int code) {
    switch (code) {
        case MB200_OK: return "ok";
        case MB200_ERR_INVALID_ARG: return "invalid argument (null pointer or non-positive size)";
        case MB200_ERR_UNSUPPORTED: return "shape not supported by the sm_100a kernel";
        case MB200_ERR_CUDA: return "CUDA call failed (see matcha_b200_last_cuda_error)";
        case MB200_ERR_NO_DEVICE: return "no usable CUDA driver/device";
    }
    return code > 0 ? "ok (positive count)" : "unknown error";
}

extern "C" int matcha_b200_last_cuda_error(void) { return g_last_cuda_error; }

extern "C" const char *matcha_b200_version(void) { return "0.1.0"; }

// Host-buffer MAS: the batch is cut into slices that flow through a three-stream pipeline (H2D of slice
// i+1, kernel of slice i, D2H of slice i-1 run concurrently; PCIe is full duplex).  Only the valid
// [t_x][t_y] block of every utterance crosses the bus in either direction, as 2-D copies.
// path[b][x][y] = 1 for the dur[b][x] frames token x owns, in order (core.pyx:87-92 written forwards); the buffer
// is pre-zeroed by the caller (__init__.py:49).  Items outside the DP's domain have all-zero durations.
static void expand_durations_host(int *paths, const int *dur, const int *t_xs, const int *t_ys, int b_lo, int b_hi,
                                  int T_x, int T_y) {
    for (int b = b_lo; b < b_hi; ++b) {
        const int tx = t_xs[b], ty = t_ys[b];
        if (tx < 1 || ty < 1 || tx > ty || tx > T_x || ty > T_y) continue;
        int *p = paths + (size_t)b * T_x * T_y;
        const int *d = dur + (size_t)b * T_x;
        int y = 0;
        for (int x = 0; x < tx; ++x) {
            int *row = p + (size_t)x * T_y;
            const int e = y + d[x] < ty ? y + d[x] : ty;
            for (; y < e; ++y) row[y] = 1;
        }
    }
}

extern "C" int matcha_b200_mas_maximum_path_host(int *paths_host, const float *values_host,
                                                  const int *t_xs_host, const int *t_ys_host,
                                                  int B, int T_x, int T_y, float penalty) {
    if (B < 0 || T_x <= 0 || T_y <= 0) return MB200_ERR_INVALID_ARG;
    if (B == 0) return 0;
    if (!paths_host || !values_host || !t_xs_host || !t_ys_host) return MB200_ERR_INVALID_ARG;
    const size_t item = (size_t)T_x * T_y;
    const char *envm = getenv("MB200_MAS_HOST_MODE");     // experiments: 1 = contiguous [t_x][T_y] row blocks instead of 2-D copies
    const bool rows_only = envm && envm[0] == '1';
    const char *envs = getenv("MB200_MAS_HOST_SLICES");
    int n_slices = B < 16 ? 1 : (envs ? atoi(envs) : 16);
    if (n_slices < 1) n_slices = 1;
    if (item * sizeof(float) * B < (64u << 20)) n_slices = 1;        // small batches: one shot
    const int Bs = (B + n_slices - 1) / n_slices;
    float *d_val[2] = {nullptr, nullptr};
    int *d_path[2] = {nullptr, nullptr}, *d_len = nullptr, *d_dur = nullptr, *h_dur = nullptr;
    cudaStream_t st[3] = {nullptr, nullptr, nullptr};                // h2d, compute, d2h
    cudaEvent_t ev_in[2] = {}, ev_k[2] = {}, ev_out[2] = {};
    int rc = MB200_OK, bad = 0, last_b0 = 0;
    cudaError_t e;
#define HOST_TRY(x) do { e = (x); if (e != cudaSuccess) { rc = cuda_fail(e); goto done; } } while (0)
    for (int i = 0; i < 3; ++i) HOST_TRY(cudaStreamCreateWithFlags(&st[i], cudaStreamNonBlocking));
    for (int i = 0; i < 2; ++i) {
        HOST_TRY(cudaMalloc(&d_val[i], (size_t)Bs * item * sizeof(float)));
        HOST_TRY(cudaMalloc(&d_path[i], (size_t)Bs * item * sizeof(int)));
        HOST_TRY(cudaEventCreateWithFlags(&ev_in[i], cudaEventDisableTiming));
        HOST_TRY(cudaEventCreateWithFlags(&ev_k[i], cudaEventDisableTiming));
        HOST_TRY(cudaEventCreateWithFlags(&ev_out[i], cudaEventDisableTiming));
    }
    HOST_TRY(cudaMalloc(&d_len, (2 * (size_t)B + 1) * sizeof(int)));
    // The alignment crosses PCIe in its compact form: the kernel's per-token durations ([B][T_x] int32, 1.6 KB per
    // utterance instead of up to 3.2 MB of 0/1 ints); the ones are written into the caller's pre-zeroed buffer here,
    // one slice behind the GPU.
    HOST_TRY(cudaMalloc(&d_dur, (size_t)B * T_x * sizeof(int)));
    HOST_TRY(cudaMallocHost(&h_dur, (size_t)B * T_x * sizeof(int)));
    HOST_TRY(cudaMemcpyAsync(d_len, t_xs_host, B * sizeof(int), cudaMemcpyHostToDevice, st[1]));
    HOST_TRY(cudaMemcpyAsync(d_len + B, t_ys_host, B * sizeof(int), cudaMemcpyHostToDevice, st[1]));
    HOST_TRY(cudaMemsetAsync(d_len + 2 * B, 0, sizeof(int), st[1]));
    for (int s = 0, b0 = 0; b0 < B; ++s, b0 += Bs) {
        const int nb = B - b0 < Bs ? B - b0 : Bs, buf = s & 1;
        // the value buffer is free once the kernel that read it (slice s-2) is done
        if (s >= 2) HOST_TRY(cudaStreamWaitEvent(st[0], ev_k[buf], 0));
        for (int i = 0; i < nb; ++i) {
            const int tx = t_xs_host[b0 + i], ty = t_ys_host[b0 + i];
            if (tx < 1 || ty < 1 || tx > T_x || ty > T_y) continue;
            if (rows_only)
                HOST_TRY(cudaMemcpyAsync(d_val[buf] + (size_t)i * item, values_host + (size_t)(b0 + i) * item,
                                         (size_t)tx * T_y * sizeof(float), cudaMemcpyHostToDevice, st[0]));
            else
                HOST_TRY(cudaMemcpy2DAsync(d_val[buf] + (size_t)i * item, (size_t)T_y * sizeof(float),
                                           values_host + (size_t)(b0 + i) * item, (size_t)T_y * sizeof(float),
                                           (size_t)ty * sizeof(float), (size_t)tx, cudaMemcpyHostToDevice, st[0]));
        }
        HOST_TRY(cudaEventRecord(ev_in[buf], st[0]));
        HOST_TRY(cudaStreamWaitEvent(st[1], ev_in[buf], 0));
        // d_path is the kernel's dense output / bit-word scratch; it stays on the device
        rc = matcha_b200_mas_align(d_path[buf], MB200_PATH_I32, d_dur + (size_t)b0 * T_x, d_val[buf], nullptr, d_len + b0,
                                   d_len + B + b0, nb, T_x, T_y, penalty, d_len + 2 * B, st[1]);
        if (rc != MB200_OK) goto done;
        HOST_TRY(cudaEventRecord(ev_k[buf], st[1]));
        HOST_TRY(cudaStreamWaitEvent(st[2], ev_k[buf], 0));
        HOST_TRY(cudaMemcpyAsync(h_dur + (size_t)b0 * T_x, d_dur + (size_t)b0 * T_x, (size_t)nb * T_x * sizeof(int),
                                 cudaMemcpyDeviceToHost, st[2]));
        HOST_TRY(cudaEventRecord(ev_out[buf], st[2]));
        if (s >= 1) {                                              // expand slice s-1 while slice s is in flight
            HOST_TRY(cudaEventSynchronize(ev_out[buf ^ 1]));
            expand_durations_host(paths_host, h_dur, t_xs_host, t_ys_host, b0 - Bs, b0, T_x, T_y);
        }
        last_b0 = b0;
    }
    HOST_TRY(cudaStreamSynchronize(st[2]));
    expand_durations_host(paths_host, h_dur, t_xs_host, t_ys_host, last_b0, B, T_x, T_y);
    HOST_TRY(cudaMemcpyAsync(&bad, d_len + 2 * B, sizeof(int), cudaMemcpyDeviceToHost, st[1]));
    HOST_TRY(cudaStreamSynchronize(st[1]));
#undef HOST_TRY
done:
    for (int i = 0; i < 3; ++i) if (st[i]) cudaStreamSynchronize(st[i]);
    for (int i = 0; i < 2; ++i) {
        if (d_val[i]) cudaFree(d_val[i]);
        if (d_path[i]) cudaFree(d_path[i]);
        if (ev_in[i]) cudaEventDestroy(ev_in[i]);
        if (ev_k[i]) cudaEventDestroy(ev_k[i]);
        if (ev_out[i]) cudaEventDestroy(ev_out[i]);
    }
    if (d_len) cudaFree(d_len);
    if (d_dur) cudaFree(d_dur);
    if (h_dur) cudaFreeHost(h_dur);
    for (int i = 0; i < 3; ++i) if (st[i]) cudaStreamDestroy(st[i]);
    return rc != MB200_OK ? rc : bad;
}
