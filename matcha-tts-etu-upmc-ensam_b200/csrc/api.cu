// api.cu -- library-wide pieces of the C ABI: error strings, the tensor-map encoder and the
// host-buffer MAS entry point (the numpy round trip of
// /root/reference/matcha/utils/monotonic_align/__init__.py:45-55).
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

extern "C" int matcha_b200_mas_maximum_path_host(int *paths_host, const float *values_host,
                                                  const int *t_xs_host, const int *t_ys_host,
                                                  int B, int T_x, int T_y, float penalty) {
    if (B < 0 || T_x <= 0 || T_y <= 0) return MB200_ERR_INVALID_ARG;
    if (B == 0) return 0;
    if (!paths_host || !values_host || !t_xs_host || !t_ys_host) return MB200_ERR_INVALID_ARG;
    const size_t n = (size_t)B * T_x * T_y;
    float *d_val = nullptr;
    int *d_path = nullptr, *d_len = nullptr;
    cudaStream_t st = nullptr;
    int rc = MB200_OK, bad = 0;
    cudaError_t e;
#define HOST_TRY(x) do { e = (x); if (e != cudaSuccess) { rc = cuda_fail(e); goto done; } } while (0)
    HOST_TRY(cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking));
    HOST_TRY(cudaMalloc(&d_val, n * sizeof(float)));
    HOST_TRY(cudaMalloc(&d_path, n * sizeof(int)));
    HOST_TRY(cudaMalloc(&d_len, (2 * (size_t)B + 1) * sizeof(int)));
    HOST_TRY(cudaMemcpyAsync(d_val, values_host, n * sizeof(float), cudaMemcpyHostToDevice, st));
    HOST_TRY(cudaMemcpyAsync(d_len, t_xs_host, B * sizeof(int), cudaMemcpyHostToDevice, st));
    HOST_TRY(cudaMemcpyAsync(d_len + B, t_ys_host, B * sizeof(int), cudaMemcpyHostToDevice, st));
    HOST_TRY(cudaMemsetAsync(d_len + 2 * B, 0, sizeof(int), st));
    rc = matcha_b200_mas_compute_batch_alignments(d_path, MB200_PATH_I32, d_val, nullptr, d_len, d_len + B,
                                                  B, T_x, T_y, penalty, d_len + 2 * B, st);
    if (rc != MB200_OK) goto done;
    HOST_TRY(cudaMemcpyAsync(paths_host, d_path, n * sizeof(int), cudaMemcpyDeviceToHost, st));
    HOST_TRY(cudaMemcpyAsync(&bad, d_len + 2 * B, sizeof(int), cudaMemcpyDeviceToHost, st));
    HOST_TRY(cudaStreamSynchronize(st));
#undef HOST_TRY
done:
    if (d_val) cudaFree(d_val);
    if (d_path) cudaFree(d_path);
    if (d_len) cudaFree(d_len);
    if (st) cudaStreamDestroy(st);
    return rc != MB200_OK ? rc : bad;
}
