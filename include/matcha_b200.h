/*
 * matcha_b200.h -- C ABI of libmatcha_b200.so (hand-written sm_100a CUDA).
 *
 * Plain pointers and sizes only; no torch types.  Every entry point returns 0 on
 * success or a negative MB200_ERR_* code; matcha_b200_error_string() names it.
 * All *_dev pointers are device pointers on the current CUDA device; `stream` is a
 * cudaStream_t passed as void* (NULL = legacy default stream).  Device entry points
 * are asynchronous on `stream`, thread-safe and allocate no device memory.  The MAS / alignment entry points keep no
 * state at all (matcha_b200_mas_durations_static alone caches, per device, the address of its built-in slab).  The decoder entry points keep, per DEVICE and behind a mutex, the opt-in shared-memory attribute of
 * their kernels and the side streams + events of the sub-batch split (created on first use; a call holds the device's
 * mutex while it enqueues, so two host threads driving one device queue their calls one after the other).
 *
 * Reference interfaces replaced (paths under /root/reference):
 *   matcha_b200_mas_compute_batch_alignments  <- compute_batch_alignments,
 *        matcha/utils/monotonic_align/core.pyx:101-128 (and :16-96 per utterance)
 *   matcha_b200_mas_lengths_from_mask         <- the two mask reductions of
 *        maximum_path, matcha/utils/monotonic_align/__init__.py:52-53
 *   matcha_b200_mas_maximum_path_host         <- the host-buffer round trip of
 *        maximum_path, __init__.py:45-55 (numpy in, numpy out)
 */
#ifndef MATCHA_B200_H
#define MATCHA_B200_H

#ifdef __cplusplus
extern "C" {
#endif

#define MB200_OK                 0
#define MB200_ERR_INVALID_ARG   -1   /* null pointer, non-positive size */
#define MB200_ERR_UNSUPPORTED   -2   /* shape beyond what the kernel handles */
#define MB200_ERR_CUDA          -3   /* a CUDA runtime/driver call failed; see matcha_b200_last_cuda_error */
#define MB200_ERR_NO_DEVICE     -4   /* no sm_100 device current */

/* element type of the alignment written by the MAS kernel (both are 4 bytes wide) */
#define MB200_PATH_I32  0            /* int32 0/1, what core.pyx writes */
#define MB200_PATH_F32  1            /* float 0.0/1.0, what maximum_path returns for fp32 input */

const char *matcha_b200_error_string(int code);
/* cudaError_t of the last failing CUDA call made by this thread through the library */
int matcha_b200_last_cuda_error(void);
/* "major.minor.patch" of the library */
const char *matcha_b200_version(void);

/*
 * Monotonic alignment search over a batch (core.pyx:101-128).
 *   paths_dev   [B][T_x][T_y] int32 or float (path_dtype); EVERY element is written
 *               (0 outside the path, including padded rows/columns), so the buffer need
 *               not be pre-zeroed (the reference requires pre-zeroing, __init__.py:49).
 *   values_dev  [B][T_x][T_y] float log-likelihoods, C-contiguous.  NOT modified (the
 *               reference overwrites it with cumulative scores, core.pyx:83-85; its only
 *               caller passes a private copy, so this is not observable through maximum_path).
 *   mask_dev    NULL, or [B][T_x][T_y] float: scores are value*mask as in __init__.py:45.
 *               With NULL the caller asserts mask==1 on [0,t_x)x[0,t_y), which holds for the
 *               outer product of two prefix masks built at matcha_tts.py:271.
 *   t_xs_dev, t_ys_dev  [B] int32 valid lengths.
 *   penalty     boundary_penalty / max_neg_val (-1e9 in the reference).
 *   Shapes the wavefront kernel does not take (T_x > 1344, or a stage ring beyond shared memory) run on a slower one-CTA-per-
 *   utterance fallback with the same results; MB200_ERR_UNSUPPORTED only when T_x exceeds ~14 000.
 *   status_dev  NULL, or one int32 that receives (atomically added) the number of items
 *               whose lengths violate 1 <= t_x <= t_y, t_x <= T_x, t_y <= T_y.  The reference
 *               has undefined behaviour for those; here their path is all zeros.
 */
int matcha_b200_mas_compute_batch_alignments(
    void *paths_dev, int path_dtype,
    const float *values_dev, const float *mask_dev,
    const int *t_xs_dev, const int *t_ys_dev,
    int B, int T_x, int T_y, float penalty,
    int *status_dev, void *stream);

/* t_xs[b] = (int) sum_x mask[b][x][0],  t_ys[b] = (int) sum_y mask[b][0][y]   (__init__.py:52-53) */
int matcha_b200_mas_lengths_from_mask(
    const float *mask_dev, int B, int T_x, int T_y,
    int *t_xs_dev, int *t_ys_dev, void *stream);

/*
 * Host buffers with the reference's C-level contract (core.pyx:101-107): `paths_host` must be
 * zero-initialised by the caller, exactly as the reference requires (__init__.py:49 passes np.zeros);
 * only the ones of every utterance's path are written.  Values and lengths go to the device in slices
 * through a three-stream pipeline (H2D, kernel and D2H overlap; pinned host memory makes the copies
 * asynchronous); the alignment comes back in its compact form (per-token durations, [B][T_x] int32) and is
 * expanded into paths_host on the host, one slice behind the GPU.  Returns the number of out-of-domain items (>= 0,
 * their paths stay zero) or a negative error.  Synchronous, like the reference call.
 *
 * Scratch (three streams, six events, four device slabs, a pinned durations buffer) lives in a handle and is reused
 * across calls: grown when a larger batch arrives, never shrunk, freed by _destroy.  One call at a time per handle
 * (internal mutex).  The handle-less form uses a process-wide default handle (matcha_b200_mas_host_release frees its
 * scratch; the next call re-creates it).
 */
void *matcha_b200_mas_host_create(void);
int matcha_b200_mas_host_destroy(void *handle);
int matcha_b200_mas_maximum_path_host_h(
    void *handle, int *paths_host, const float *values_host,
    const int *t_xs_host, const int *t_ys_host,
    int B, int T_x, int T_y, float penalty);
int matcha_b200_mas_maximum_path_host(
    int *paths_host, const float *values_host,
    const int *t_xs_host, const int *t_ys_host,
    int B, int T_x, int T_y, float penalty);
int matcha_b200_mas_host_release(void);

/* scores[i] = values[i] * mask[i] (__init__.py:45), for callers whose shape is too large for the
 * fused-mask variant of the MAS kernel (query_launch with has_mask=1 returns MB200_ERR_UNSUPPORTED):
 * form the scores here, then call matcha_b200_mas_compute_batch_alignments with mask_dev = NULL. */
int matcha_b200_mas_apply_mask(float *scores_dev, const float *values_dev, const float *mask_dev,
                               long long n, void *stream);

/* Launch geometry the MAS kernel would use for a shape (for bench/launch accounting):
 * out[0]=grid, out[1]=block, out[2]=dynamic smem bytes, out[3]=rows per lane (R),
 * out[4]=pipeline stages, out[5]=1 if the TMA path is taken, out[6]=CTAs per SM,
 * out[7]=DP warps per CTA. */
int matcha_b200_mas_query_launch(int B, int T_x, int T_y, int has_mask, int aligned16, int out[8]);

/* Which kernel a launch with this shape takes (matcha_b200_mas_query_launch describes the wavefront kernel only and
 * returns MB200_ERR_UNSUPPORTED where matcha_b200_mas_compute_batch_alignments / _align still succeed on the fallback):
 *   MB200_MAS_PATH_WAVEFRONT       the TMA-fed wavefront kernel (the fast path)
 *   MB200_MAS_PATH_TALL            the one-CTA-per-utterance fallback (T_x > 1344): same results, ~1 us per mel frame
 *   MB200_MAS_PATH_TALL_MASK_ONLY  has_mask = 1 and only the fused-mask stage ring does not fit (T_x roughly 900..1344): the
 *                                  launch would run the slow fallback, although the wavefront kernel takes this shape WITHOUT a
 *                                  mask.  A direct C caller should form value*mask with matcha_b200_mas_apply_mask into a buffer
 *                                  of their own and pass mask_dev = NULL (what matcha_b200.monotonic_align.maximum_path does); the
 *                                  library cannot do that itself because it allocates no device memory.
 *   MB200_ERR_UNSUPPORTED          no kernel takes the shape (T_x beyond ~14 000)
 * Needs no device.  (No counterpart in the reference.) */
#define MB200_MAS_PATH_WAVEFRONT       0
#define MB200_MAS_PATH_TALL            1
#define MB200_MAS_PATH_TALL_MASK_ONLY  2
int matcha_b200_mas_query_path(int B, int T_x, int T_y, int has_mask, int aligned16);

/*
 * maximum_path plus the compact form of the result: durations_dev [B][T_x] int32 (may be NULL) receives the
 * number of frames each token owns (= sum_y path[b][x][y], what matcha_tts.py:287 reduces the path to; 0
 * for x >= t_x and for out-of-domain items).  It comes from the backtrack's start frames, so the dense path
 * is not re-read.  All other arguments as matcha_b200_mas_compute_batch_alignments.
 */
int matcha_b200_mas_align(
    void *paths_dev, int path_dtype, int *durations_dev,
    const float *values_dev, const float *mask_dev,
    const int *t_xs_dev, const int *t_ys_dev,
    int B, int T_x, int T_y, float penalty,
    int *status_dev, void *stream);

/*
 * Durations only, with no device buffer beyond the inputs and the [B][T_x] output.  (No counterpart in the reference: its
 * Cython core mallocs its own scratch, core.pyx:39-41; this entry point exists for GPU callers that may not allocate.)
 *
 * The MAS kernel always needs a [B][T_x][T_y] slab of 4-byte cells: it parks the DP's decision bits there and then lays the
 * dense path over them.  matcha_b200_mas_align borrows the caller's paths buffer for that.  This variant uses a fixed array
 * of MB200_MAS_STATIC_MAX_CELLS cells (8 MiB) that is part of the library's device image, so it makes no cudaMalloc / cudaFree
 * call and needs no allocator on the caller's side.  It is meant for
 *   - a host thread that aligns batches while another thread captures a decoder solve into a CUDA graph: cudaMalloc from any
 *     thread breaks a global-mode capture;
 *   - a C / ctypes caller (e.g. a duration-extraction script) sharing the GPU with a trainer whose caching allocator has
 *     already claimed the memory, where a cudaMalloc can fail, and in any case synchronises the whole device;
 *   - serving loops that keep cudaMalloc off the request path.
 * durations_dev, values_dev, mask_dev, t_xs_dev, t_ys_dev, penalty, status_dev, stream: exactly as matcha_b200_mas_align;
 * the result is identical to its durations_dev.  The dense path is not returned.
 *
 * Limits and sharing:
 *   - B*T_x*T_y <= MB200_MAS_STATIC_MAX_CELLS (matcha_b200_mas_durations_static_max_cells() returns it), e.g. 16 x 128 x 800;
 *     anything larger returns MB200_ERR_UNSUPPORTED and launches nothing.  Split the batch, or use matcha_b200_mas_align.
 *   - Every GPU on which the library's kernels are loaded holds its own 8 MiB copy until the process exits.
 *   - The first call on each device looks the array's device address up (cudaGetSymbolAddress with that device current) and
 *     caches it; make that first call outside any stream capture.  Later calls only enqueue the kernel.
 *   - All calls on one device share the one array and it keeps whatever the last call wrote.  Nothing in it has to be
 *     cleared between calls (every cell is rewritten before it is read), but two calls that may run on the GPU at the
 *     same time -- different streams, or different host threads without a common stream -- race on it and both results
 *     are then garbage.  Put such calls on one stream, or order them with events.
 */
#define MB200_MAS_STATIC_MAX_CELLS (2 * 1024 * 1024)
long long matcha_b200_mas_durations_static_max_cells(void);
int matcha_b200_mas_durations_static(
    int *durations_dev, const float *values_dev, const float *mask_dev,
    const int *t_xs_dev, const int *t_ys_dev,
    int B, int T_x, int T_y, float penalty,
    int *status_dev, void *stream);

/* ------------------------------------------------------------------ callers either side of MAS / the decoder
 * (SURVEY.md §8f rows 1-3).  Reference code replaced:
 *   matcha_b200_align_log_prior <- the no_grad block of MatchaTTS.forward            matcha_tts.py:277-282
 *   matcha_b200_align_expand    <- generate_path (utils/model.py:77-114), mu_y = attn^T mu_x
 *                                  (matcha_tts.py:227-228, 314-315), logw_ (:287), prior loss sums (:320-321)
 *   matcha_b200_cfm_loss_mix    <- phi_t of compute_loss                              flow_matching.py:138
 *   matcha_b200_cfm_loss_mse    <- u_target, mse_loss(sum), sum(mask)                 flow_matching.py:142-149
 *
 * log_prior: mu_x [B][C][T_x], y [B][C][T_y] fp32 channels-first -> log_prior [B][T_x][T_y].  With
 * t_xs/t_ys (both or neither) tiles that lie entirely outside an item's valid block are not written
 * (MAS never reads them).
 *
 * expand: durations [B][T_x] int32 (dur_is_int) or fp32, >= 0.  Every output pointer may be NULL:
 *   attn [B][T_x][T_y] fp32 0/1 (fully written), mu_y [B][C][T_y] (fully written, 0 past t_y),
 *   logw [B][T_x], prior_sums double[2] = {sum 0.5*((y-mu_y)^2+log 2pi)*y_mask, sum y_mask}
 *   (needs y_dev; zeroed inside the call).  t_xs/t_ys NULL = full lengths.
 *
 * cfm_loss_mix: phi = (1-(1-sigma_min)*t[b])*z + t[b]*x1 over B samples of per_sample floats.
 * cfm_loss_mse: sums double[2] = {sum (u_pred-(x1-(1-sigma_min)*z))^2 over n, sum mask over n_mask}.
 */
int matcha_b200_align_log_prior(float *log_prior_dev, const float *mu_x_dev, const float *y_dev,
                                const int *t_xs_dev, const int *t_ys_dev,
                                int B, int C, int T_x, int T_y, void *stream);
int matcha_b200_align_expand(const void *durations_dev, int dur_is_int, const float *mu_x_dev,
                             const float *y_dev, const int *t_xs_dev, const int *t_ys_dev,
                             float *attn_dev, float *mu_y_dev, float *logw_dev, double *prior_sums_dev,
                             int B, int C, int T_x, int T_y, void *stream);
int matcha_b200_cfm_loss_mix(float *phi_dev, const float *x1_dev, const float *z_dev, const float *t_dev,
                             double sigma_min, int B, long long per_sample, void *stream);
int matcha_b200_cfm_loss_mse(double *sums_dev, const float *u_pred_dev, const float *x1_dev,
                             const float *z_dev, const float *mask_dev, double sigma_min,
                             long long n, long long n_mask, void *stream);

/* ------------------------------------------------------------------ CFM decoder
 * Reference interfaces replaced (matcha/models/components/):
 *   matcha_b200_decoder_forward  <- Decoder.forward(x, mask, mu, t)               decoder.py:271-371
 *   matcha_b200_cfm_solve_euler  <- BaseConditionalFlowMatching.solve_ode_euler   flow_matching.py:67-104
 * x, mu, out: [B][80][T] fp32 channels-first, as the Python API delivers them; mask: [B][T] (= [B][1][T])
 * fp32 0/1.  blob: the packed weights (matcha_b200.decoder.pack_weights, matcha_b200_decoder_blob_floats()
 * floats, 16-byte aligned).  ws: caller-provided scratch of matcha_b200_decoder_workspace_bytes(B, T, n_rows, precision)
 * bytes, n_rows = number of distinct time rows (n_t, or n_steps).  Nothing is allocated inside.
 * Both return the number of kernels launched (> 0) or a negative error.                           */
#define MB200_PREC_F32   0   /* fp32 SIMT arithmetic: rel-L2 <= 1e-5 against the fp32 reference */
#define MB200_PREC_BF16  1   /* bf16 tensor-core GEMMs, fp32 residual stream and statistics     */
#define MB200_PREC_F16   2   /* same kernels with IEEE fp16 operands (3 more mantissa bits)     */

long long matcha_b200_decoder_blob_floats(void);
/* bf16 mode: GEMM weights re-packed as bf16 (input channels padded to 64) by matcha_b200_decoder_pack_bf16 into a
 * caller-provided buffer of matcha_b200_decoder_blob_bf16_bytes() bytes; pass NULL in fp32 mode. */
long long matcha_b200_decoder_blob_bf16_bytes(void);
int matcha_b200_decoder_pack_bf16(const float *blob_dev, void *blob_bf16_dev, int precision, void *stream);
long long matcha_b200_decoder_workspace_bytes(int B, int T, int n_rows, int precision);

/* t_dev: n_t = 1 (one time for the whole batch, as in the Euler loop) or n_t = B (compute_loss) */
int matcha_b200_decoder_forward(const float *x_dev, const float *mu_dev, const float *mask_dev,
                                const float *t_dev, int n_t, float *out_dev, const float *blob_dev,
                                const void *blob_bf16_dev, void *ws_dev, long long ws_bytes, int B, int T,
                                int precision, void *stream);

/* BaseConditionalFlowMatching.compute_loss, forward value (flow_matching.py:106-151) in one call:
 *   phi = (1 - (1 - sigma_min) t_b) z + t_b x1                   (:138), written to phi_dev [B][80][T]
 *   u_pred = Decoder.forward(phi, mask, mu, t)                    (:145), written to u_pred_dev when it is not NULL
 *   sums_dev[0] = sum (u_pred - (x1 - (1 - sigma_min) z))^2  over all B*80*T elements, sums_dev[1] = sum(mask)   (:142-148)
 * (loss = sums[0] / (sums[1] * 80), :149).  t_dev: [B] per-sample times.  ws: workspace_bytes(B, T, n_rows = B, precision).
 * 16-bit precisions: phi is formed inside the input-packing kernel and the squared error inside the final projection's
 * epilogue, so u_pred / u_target never reach HBM.  MB200_PREC_F32: separate mix / mse kernels, u_pred_dev is required. */
int matcha_b200_cfm_loss_forward(const float *x1_dev, const float *z_dev, const float *t_dev, const float *mu_dev,
                                 const float *mask_dev, double sigma_min, float *phi_dev, float *u_pred_dev,
                                 double *sums_dev, const float *blob_dev, const void *blob_bf16_dev, void *ws_dev,
                                 long long ws_bytes, int B, int T, int precision, void *stream);

/* t_host/dt_host: the n_steps (t_k, dt_k) pairs the reference loop produces in fp32
 * (flow_matching.py:85-102); out_dev may not alias x0_dev.  n_steps <= 64. */
int matcha_b200_cfm_solve_euler(const float *x0_dev, const float *mu_dev, const float *mask_dev,
                                float *out_dev, const float *t_host, const float *dt_host, int n_steps,
                                const float *blob_dev, const void *blob_bf16_dev, void *ws_dev, long long ws_bytes,
                                int B, int T, int precision, void *stream);

/* ---------------------------------------------------------------------------------------------------------------------
 * Text encoder (BASELINE configs[3] "text-encoder mu_x -> MAS -> CFM loss").  Replaces
 *   TextEncoder.forward(text_input, text_lengths) -> (mean_output, log_duration, text_mask)
 *        matcha/models/components/text_encoder.py:377-402 (with ConvReluNorm :48-57, Encoder :307-331,
 *        MultiHeadAttention + RotaryPositionalEmbeddings :110-229, FFN :257-260, DurationPredictor :80-97)
 * for the configuration MatchaTTS builds (matcha_tts.py:122-139: 192 channels, 768 filter channels, 2 heads, 6 layers,
 * kernel 3, prenet; duration predictor 256 x kernel 3; 80 features), inference / eval mode (dropout = identity).
 * precision: MB200_PREC_F32 (SIMT fp32, rel-L2 <= 1e-5) or MB200_PREC_F16 (every Conv1d on tcgen05 tensor cores with fp16
 * operands and fp32 accumulation; residual stream, LayerNorm, attention and the mean projection in fp32; rel-L2 <= 2e-3).
 *   ids_dev      [B][T] int64 token ids (ids outside [0, n_vocab) are clamped; torch raises)
 *   lengths_dev  [B] int32 valid lengths
 *   blob_dev     packed fp32 weights (matcha_b200.text_encoder.pack_weights; matcha_b200_text_encoder_blob_floats(n_vocab)
 *                floats, 16-byte aligned)
 *   blob_f16_dev the same blob converted to IEEE fp16 element for element (MB200_PREC_F16; NULL otherwise)
 *   rope_cos_dev, rope_sin_dev  [>= T][48] fp32: the cos / sin caches of RotaryPositionalEmbeddings._build_cache
 *                (:110-128), built by the caller with the reference's own torch ops
 *   mu_dev [B][80][T], logw_dev [B][1][T], mask_dev [B][1][T]: the three outputs, fp32 (mask_dev is also read by the kernels)
 *   workspace_dev  matcha_b200_text_encoder_workspace_bytes(B, T, precision) bytes, 16-byte aligned; nothing is allocated inside.
 * Returns the number of kernels launched (> 0) or a negative error.  The only state kept is, per device, the opt-in
 * shared-memory attribute of the tensor-core kernel. */
long long matcha_b200_text_encoder_blob_floats(int n_vocab);
long long matcha_b200_text_encoder_workspace_bytes(int B, int T, int precision);
int matcha_b200_text_encoder_forward(const long long *ids_dev, const int *lengths_dev, const float *blob_dev,
                                     const void *blob_f16_dev, int n_vocab, const float *rope_cos_dev,
                                     const float *rope_sin_dev, float *mu_dev, float *logw_dev, float *mask_dev, int B, int T,
                                     int precision, void *workspace_dev, long long workspace_bytes, void *stream);

#ifdef __cplusplus
}
#endif
#endif /* MATCHA_B200_H */
