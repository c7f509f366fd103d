// api.cu -- library-wide pieces of the C ABI: error strings, the tensor-map encoder and the
// host-buffer MAS entry point (the numpy round trip of
// /root/reference/matcha/utils/monotonic_align/__init__.py:45-55).
#include <stdlib.h>
#include <string.h>

#include <mutex>
#include <new>

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

// Scratch of the host-buffer entry, kept across calls: three streams (h2d, compute, d2h), the events between them, two
// value slabs + two path slabs on the device, the lengths and the durations buffers (device + pinned host).  Grown on
// demand, never shrunk; one call at a time per handle (mutex).
struct MasHost {
    std::mutex mu;
    int dev = -1;
    cudaStream_t st[3] = {nullptr, nullptr, nullptr};
    cudaEvent_t ev_in[2] = {}, ev_k[2] = {}, ev_out[2] = {};
    float *d_val[2] = {nullptr, nullptr};
    int *d_path[2] = {nullptr, nullptr};
    size_t slab_bytes = 0;
    int *d_len = nullptr; size_t len_cap = 0;
    int *d_dur = nullptr, *h_dur = nullptr; size_t dur_cap = 0;
    bool ready = false;
};

static void mas_host_free(MasHost &h) {
    for (int i = 0; i < 3; ++i) if (h.st[i]) { cudaStreamSynchronize(h.st[i]); }
    for (int i = 0; i < 2; ++i) {
        if (h.d_val[i]) cudaFree(h.d_val[i]);
        if (h.d_path[i]) cudaFree(h.d_path[i]);
        if (h.ev_in[i]) cudaEventDestroy(h.ev_in[i]);
        if (h.ev_k[i]) cudaEventDestroy(h.ev_k[i]);
        if (h.ev_out[i]) cudaEventDestroy(h.ev_out[i]);
        h.d_val[i] = nullptr; h.d_path[i] = nullptr; h.ev_in[i] = h.ev_k[i] = h.ev_out[i] = nullptr;
    }
    if (h.d_len) cudaFree(h.d_len);
    if (h.d_dur) cudaFree(h.d_dur);
    if (h.h_dur) cudaFreeHost(h.h_dur);
    for (int i = 0; i < 3; ++i) if (h.st[i]) { cudaStreamDestroy(h.st[i]); h.st[i] = nullptr; }
    h.d_len = h.d_dur = h.h_dur = nullptr;
    h.slab_bytes = h.len_cap = h.dur_cap = 0;
    h.ready = false;
}

static int mas_host_ensure(MasHost &h, size_t slab_bytes, size_t n_len, size_t n_dur) {
    int dev = -1;
    MB200_CUDA_TRY(cudaGetDevice(&dev));
    if (h.ready && h.dev != dev) mas_host_free(h);            // the handle follows the caller's current device
    if (!h.ready) {
        h.dev = dev;
        for (int i = 0; i < 3; ++i) MB200_CUDA_TRY(cudaStreamCreateWithFlags(&h.st[i], cudaStreamNonBlocking));
        for (int i = 0; i < 2; ++i) {
            MB200_CUDA_TRY(cudaEventCreateWithFlags(&h.ev_in[i], cudaEventDisableTiming));
            MB200_CUDA_TRY(cudaEventCreateWithFlags(&h.ev_k[i], cudaEventDisableTiming));
            MB200_CUDA_TRY(cudaEventCreateWithFlags(&h.ev_out[i], cudaEventDisableTiming));
        }
        h.ready = true;
    }
    if (slab_bytes > h.slab_bytes) {
        for (int i = 0; i < 2; ++i) {
            if (h.d_val[i]) cudaFree(h.d_val[i]);
            if (h.d_path[i]) cudaFree(h.d_path[i]);
            h.d_val[i] = nullptr; h.d_path[i] = nullptr;
        }
        h.slab_bytes = 0;
        for (int i = 0; i < 2; ++i) {
            MB200_CUDA_TRY(cudaMalloc(&h.d_val[i], slab_bytes));
            MB200_CUDA_TRY(cudaMalloc(&h.d_path[i], slab_bytes));
        }
        h.slab_bytes = slab_bytes;
    }
    if (n_len > h.len_cap) {
        if (h.d_len) cudaFree(h.d_len);
        h.d_len = nullptr; h.len_cap = 0;
        MB200_CUDA_TRY(cudaMalloc(&h.d_len, n_len * sizeof(int)));
        h.len_cap = n_len;
    }
    if (n_dur > h.dur_cap) {
        if (h.d_dur) cudaFree(h.d_dur);
        if (h.h_dur) cudaFreeHost(h.h_dur);
        h.d_dur = h.h_dur = nullptr; h.dur_cap = 0;
        MB200_CUDA_TRY(cudaMalloc(&h.d_dur, n_dur * sizeof(int)));
        MB200_CUDA_TRY(cudaMallocHost(&h.h_dur, n_dur * sizeof(int)));
        h.dur_cap = n_dur;
    }
    return MB200_OK;
}

static int mas_host_run(MasHost &h, int *paths_host, const float *values_host, const int *t_xs_host, const int *t_ys_host,
                        int B, int T_x, int T_y, float penalty) {
    if (B < 0 || T_x <= 0 || T_y <= 0) return MB200_ERR_INVALID_ARG;
    if (B == 0) return 0;
    if (!paths_host || !values_host || !t_xs_host || !t_ys_host) return MB200_ERR_INVALID_ARG;
    std::lock_guard<std::mutex> lock(h.mu);
    const size_t item = (size_t)T_x * T_y;
    const char *envm = getenv("MB200_MAS_HOST_MODE");     // experiments: 1 = contiguous [t_x][T_y] row blocks instead of 2-D copies
    const bool rows_only = envm && envm[0] == '1';
    const char *envs = getenv("MB200_MAS_HOST_SLICES");
    int n_slices = B < 16 ? 1 : (envs ? atoi(envs) : 16);
    if (n_slices < 1) n_slices = 1;
    if (item * sizeof(float) * B < (64u << 20)) n_slices = 1;        // small batches: one shot
    const int Bs = (B + n_slices - 1) / n_slices;
    int rc = mas_host_ensure(h, (size_t)Bs * item * sizeof(float), 2 * (size_t)B + 1, (size_t)B * T_x);
    if (rc != MB200_OK) return rc;
    float *const *d_val = h.d_val;
    int *const *d_path = h.d_path;
    int *d_len = h.d_len, *d_dur = h.d_dur, *h_dur = h.h_dur;
    cudaStream_t *st = h.st;                                         // h2d, compute, d2h
    cudaEvent_t *ev_in = h.ev_in, *ev_k = h.ev_k, *ev_out = h.ev_out;
    int bad = 0, last_b0 = 0;
    cudaError_t e;
#define HOST_TRY(x) do { e = (x); if (e != cudaSuccess) { rc = cuda_fail(e); goto done; } } while (0)
    // The alignment crosses PCIe in its compact form: the kernel's per-token durations ([B][T_x] int32, 1.6 KB per
    // utterance instead of up to 3.2 MB of 0/1 ints); the ones are written into the caller's pre-zeroed buffer here,
    // one slice behind the GPU.
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
    for (int i = 0; i < 3; ++i) if (st[i]) cudaStreamSynchronize(st[i]);     // the call is synchronous, like the reference
    return rc != MB200_OK ? rc : bad;
}

static MasHost &default_mas_host() {
    static MasHost h;
    return h;
}

extern "C" void *matcha_b200_mas_host_create(void) { return new (std::nothrow) MasHost(); }

extern "C" int matcha_b200_mas_host_destroy(void *handle) {
    if (!handle) return MB200_ERR_INVALID_ARG;
    MasHost *h = static_cast<MasHost *>(handle);
    { std::lock_guard<std::mutex> lock(h->mu); mas_host_free(*h); }
    delete h;
    return MB200_OK;
}

extern "C" int matcha_b200_mas_maximum_path_host_h(void *handle, int *paths_host, const float *values_host,
                                                    const int *t_xs_host, const int *t_ys_host,
                                                    int B, int T_x, int T_y, float penalty) {
    if (!handle) return MB200_ERR_INVALID_ARG;
    return mas_host_run(*static_cast<MasHost *>(handle), paths_host, values_host, t_xs_host, t_ys_host, B, T_x, T_y, penalty);
}

// The handle-less form of the reference's call shape: a process-wide default handle, created on first use.
extern "C" int matcha_b200_mas_maximum_path_host(int *paths_host, const float *values_host,
                                                  const int *t_xs_host, const int *t_ys_host,
                                                  int B, int T_x, int T_y, float penalty) {
    return mas_host_run(default_mas_host(), paths_host, values_host, t_xs_host, t_ys_host, B, T_x, T_y, penalty);
}

// Frees the default handle's device / pinned scratch (it is re-created on the next call).
extern "C" int matcha_b200_mas_host_release(void) {
    MasHost &h = default_mas_host();
    std::lock_guard<std::mutex> lock(h.mu);
    mas_host_free(h);
    return MB200_OK;
}
