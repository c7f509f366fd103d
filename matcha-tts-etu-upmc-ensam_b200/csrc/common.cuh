// common.cuh -- PTX wrappers (mbarrier, TMA, cp.async) and host-side error plumbing shared by
// the sm_100a kernels of libmatcha_b200.so.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/matcha_b200.h"

namespace mb200 {

// ---------------------------------------------------------------- host side
extern thread_local int g_last_cuda_error;

inline int cuda_fail(cudaError_t e) {
    g_last_cuda_error = (int)e;
    return MB200_ERR_CUDA;
}
#define MB200_CUDA_TRY(expr)                                   \
    do {                                                       \
        cudaError_t _e = (expr);                               \
        if (_e != cudaSuccess) return ::mb200::cuda_fail(_e);  \
    } while (0)

constexpr int kNumSMsB200 = 148;
constexpr int kMaxSmemPerBlock = 232448;   // 227 KB opt-in limit on sm_100
constexpr int kSmemPerSM = 233472;         // 228 KB
constexpr int kSmemBlockReserve = 1024;    // per-CTA system reservation

// Encodes a 2-D fp32 row-major tensor map [rows][cols] with a 32x32 box and 128B swizzle.
// Returns MB200_OK or an error; resolves cuTensorMapEncodeTiled through the runtime so the
// library has no link-time dependency on libcuda.
int encode_tmap_2d_f32(CUtensorMap *map, const void *base, uint64_t rows, uint64_t cols,
                       uint32_t box_rows, uint32_t box_cols);

int encode_tmap_bf16(CUtensorMap *map, const void *base, int ndim, const uint64_t *dims,
                     const uint64_t *strides_bytes, const uint32_t *box);

int encode_tmap_out(CUtensorMap *map, const void *base, int is_f32, const uint64_t *dims, const uint64_t *strides_bytes,
                    int box_cols);

// ---------------------------------------------------------------- device side
#ifdef __CUDACC__

__device__ __forceinline__ uint32_t smem_u32(const void *p) {
    return (uint32_t)__cvta_generic_to_shared(p);
}

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_fence_init() {
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint32_t bar, uint32_t parity) {
    uint32_t done;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(done) : "r"(bar), "r"(parity) : "memory");
    return done != 0;
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    while (!mbar_try_wait(bar, parity)) {}
}
// Same, but the hardware may keep the thread suspended for up to `ns` per attempt (it still
// wakes as soon as the phase completes): for waits that are usually long, so that the polling
// loop does not steal issue slots from the warps doing the work.
__device__ __forceinline__ void mbar_wait_parked(uint32_t bar, uint32_t parity, uint32_t ns) {
    uint32_t done;
    do {
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, %3;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(done) : "r"(bar), "r"(parity), "r"(ns) : "memory");
    } while (!done);
}

// TMA: one 2-D box global -> shared, completion counted in bytes on `bar`.
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap *map, uint32_t bar,
                                            int c_inner, int c_outer) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
        ::"r"(dst), "l"((uint64_t)map), "r"(bar), "r"(c_inner), "r"(c_outer) : "memory");
}
// TMA store of one 3-D box shared -> global (bulk async-group completion); rows / columns outside the tensor are clipped.
__device__ __forceinline__ void tma_store_3d(const CUtensorMap *map, uint32_t src, int c0, int c1, int c2) {
    asm volatile("cp.async.bulk.tensor.3d.global.shared::cta.bulk_group [%0, {%2, %3, %4}], [%1];"
                 ::"l"((uint64_t)map), "r"(src), "r"(c0), "r"(c1), "r"(c2) : "memory");
}
__device__ __forceinline__ void tma_store_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void tma_store_wait_read() { asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory"); }
__device__ __forceinline__ void tma_store_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }
__device__ __forceinline__ void tma_prefetch_desc(const CUtensorMap *map) {
    asm volatile("prefetch.tensormap [%0];" ::"l"((uint64_t)map) : "memory");
}

// Ampere-style async copy of one 4-byte word, used where TMA's 16-byte alignment rules fail.
__device__ __forceinline__ void cp_async_4(uint32_t dst, const void *src) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(dst), "l"(src) : "memory");
}
// Arrive on `bar` once all prior cp.async of this thread have landed (does not add to the
// pending count: the barrier's init count must include this arrival).
__device__ __forceinline__ void cp_async_mbar_arrive_noinc(uint32_t bar) {
    asm volatile("cp.async.mbarrier.arrive.noinc.shared::cta.b64 [%0];" ::"r"(bar) : "memory");
}

__device__ __forceinline__ float4 lds128(uint32_t addr) {
    float4 v;
    asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];"
                 : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(addr));
    return v;
}
__device__ __forceinline__ void sts128(uint32_t addr, float4 v) {
    asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}

// One lane of a converged warp (elect.sync).  Single-thread instructions (TMA, tcgen05.mma, tcgen05.commit) are
// issued under it from warp-uniform control flow: the compiler then keeps their operands in uniform registers,
// whereas an `if (lane == 0)` region makes it wrap each of them in a divergence ("waterfall") loop.
__device__ __forceinline__ bool elect_one() {
    uint32_t pred;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "elect.sync _|p, 0xffffffff;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(pred));
    return pred != 0;
}

// Programmatic dependent launch (sm_90+): a kernel launched with the programmatic-serialization attribute may
// start while its predecessor in the stream drains; nothing the predecessor wrote may be read (and nothing it
// reads may be overwritten) before pdl_wait(), which returns once every prerequisite grid has completed and
// flushed.  Every kernel of the decoder path executes it in every thread, so completion stays transitive.
// Both are no-ops for a normal launch.
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
// Order of the two.  MB200_PDL_SAFE = 0 (decoder default): trigger first, so every kernel of a chain may be resident and
// waiting while an earlier one still runs.  griddepcontrol.wait orders a kernel after its prerequisite grid (and, through that
// grid's own wait, after everything before it), but it only promises VISIBILITY of the prerequisite grid's writes: a CTA that
// was made resident while kernel N still ran can find, in its SM's L1, a line kernel N read there and kernel N+1 or N+2 has
// since rewritten.  The text encoder's small kernels hit exactly that on two of six boxes (profiles/r2_te_pdl_race.log) and
// therefore use the other order plus ld.global.cg loads (csrc/text_encoder.cu).  The decoder's tensor-core chain reads every
// re-written activation through L2 as well: operands by TMA, and the residual / A-build inputs of gemm_bf16, gemm_abuild and
// ffn_fused by ld.global.cg (ld_chain4, round 3).  Until round 3 those were plain loads and the chain leaned on its shape (the
// readers take ~200 KB of shared memory, so a waiting successor and the earlier reader never shared an SM) -- an argument
// that two CTAs per SM would have voided.  MB200_PDL_SAFE = 1 compiles the conservative order into the decoder kernels too: a CTA lets the
// successor be scheduled only after its own wait has returned (at most two grids of a chain in flight).  Measured on one box
// (profiles/r2_pdl_order_ab.log): B=32 solve 8.97 -> 9.25 ms (+3 %), B=8 6.61 -> 6.27 ms (-5 %), B=64 15.84 -> 15.73 ms.
#ifndef MB200_PDL_SAFE
#define MB200_PDL_SAFE 0
#endif
// Loads of a tensor that an earlier kernel of a PDL chain wrote (and that a later one rewrites): through L2 only.  A CTA that
// became resident while an older grid still ran may share its SM's L1 with that grid's reads; ld.global.cg never looks there,
// so correctness does not rest on "the reader owns its SM" (1 CTA/SM by shared-memory footprint) any more.
__device__ __forceinline__ float4 ld_chain4(const float *p) { return __ldcg(reinterpret_cast<const float4 *>(p)); }
__device__ __forceinline__ void pdl_entry() {            // kernels whose every thread waits at the top
#if MB200_PDL_SAFE
    pdl_wait();
    pdl_launch_dependents();
#else
    pdl_launch_dependents();
    pdl_wait();
#endif
}
__device__ __forceinline__ void pdl_trigger_at_top() {   // kernels that wait later, per warp role (weights are fetched first)
#if !MB200_PDL_SAFE
    pdl_launch_dependents();
#endif
}
__device__ __forceinline__ void pdl_wait_then_trigger() {   // the first wait of such a kernel's data-dependent warps
    pdl_wait();
#if MB200_PDL_SAFE
    pdl_launch_dependents();
#endif
}

#endif  // __CUDACC__
}  // namespace mb200
