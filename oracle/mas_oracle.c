/*
 * oracle/mas_oracle.c -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * CPU restatement (plain C, optional OpenMP over the batch) of the reference's
 * monotonic alignment search, used only as the parity checker by tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs.
 * Nothing under matcha-tts-etu-upmc-ensam_b200/ may link, import or call it.
 *
 * Follows /root/reference/matcha/utils/monotonic_align/core.pyx:
 *   mas_oracle_single  <- compute_single_alignment   (core.pyx:16-96)
 *   mas_oracle_batch   <- compute_batch_alignments   (core.pyx:101-128)
 * Parity pinned: checked bit-for-bit against the reference's own core.pyx
 * compiled into oracle/_ref/ (tests/test_oracle_mas.py) and against the golden
 * vectors in tests/golden/ generated from that compiled reference.
 *
 * Deliberate restatement choices (all behaviour-preserving):
 *   - two rolling rows + one byte per cell of "came from x-1" flags, exactly as
 *     core.pyx:39-41; the in-place overwrite of the score matrix with the
 *     cumulative scores (core.pyx:83-85) is reproduced because the reference's
 *     C-level caller can observe it.
 *   - lengths violating 1 <= t_x <= t_y are undefined behaviour in the
 *     reference (core.pyx:37,88,91 with bounds checks off).  The oracle leaves
 *     the path all-zero for such items and reports them in the return value.
 */
#include <stdlib.h>
#include <string.h>

#ifdef _OPENMP
#include <omp.h>
#endif

/* One utterance.  path/score are [t_x_pitch rows][t_y_pitch] row-major slices of
 * the batch tensors; only x < t_x, y < t_y cells are touched.  Returns 0 on
 * success, 1 when the lengths are outside the reference's defined domain,
 * 2 on allocation failure (the reference silently returns, core.pyx:43-50). */
int mas_oracle_single(int *path, float *score, int t_x, int t_y, int t_y_pitch,
                      float penalty)
{
    if (t_x < 1 || t_y < 1 || t_x > t_y)
        return 1;

    float *prev = (float *)malloc((size_t)t_x * sizeof(float));
    float *cur = (float *)malloc((size_t)t_x * sizeof(float));
    unsigned char *diag = (unsigned char *)malloc((size_t)t_x * (size_t)t_y);
    if (!prev || !cur || !diag) {
        free(prev); free(cur); free(diag);
        return 2;
    }

    for (int x = 0; x < t_x; ++x)                       /* core.pyx:52-53 */
        prev[x] = penalty;

    for (int y = 0; y < t_y; ++y) {                     /* core.pyx:55 */
        for (int x = 0; x < t_x; ++x)                   /* core.pyx:56-57 */
            cur[x] = penalty;

        int x_lo = t_x + y - t_y;                       /* core.pyx:59-60 */
        if (x_lo < 0) x_lo = 0;
        int x_hi = y + 1 < t_x ? y + 1 : t_x;

        for (int x = x_lo; x < x_hi; ++x) {
            float from_prev, from_same;
            if (x == 0)                                 /* core.pyx:63-66 */
                from_prev = (y == 0) ? 0.0f : penalty;
            else
                from_prev = prev[x - 1];
            if (x == y)                                 /* core.pyx:68-71 */
                from_same = penalty;
            else
                from_same = (y > 0) ? prev[x] : penalty;

            float best;
            if (from_prev >= from_same || x == y) {     /* core.pyx:73-78: ties go to x-1 */
                best = from_prev;
                diag[(size_t)x * t_y + y] = 1;
            } else {
                best = from_same;
                diag[(size_t)x * t_y + y] = 0;
            }
            /* one fp32 round-to-nearest add, core.pyx:80 */
            volatile float sum = best + score[(size_t)x * t_y_pitch + y];
            cur[x] = sum;
        }

        for (int x = 0; x < t_x; ++x) {                 /* core.pyx:83-85 */
            prev[x] = cur[x];
            score[(size_t)x * t_y_pitch + y] = prev[x];
        }
    }

    int idx = t_x - 1;                                  /* core.pyx:36,87-92 */
    for (int y = t_y - 1; y >= 0; --y) {
        path[(size_t)idx * t_y_pitch + y] = 1;
        if (y == 0)
            break;
        if (idx > 0 && (idx == y || diag[(size_t)idx * t_y + y] == 1))
            idx -= 1;
    }

    free(prev); free(cur); free(diag);
    return 0;
}

/* Whole batch, C-contiguous [B][T_x][T_y] tensors as in core.pyx:101-107.
 * paths must be pre-zeroed by the caller (the reference's wrapper does so,
 * __init__.py:49).  n_threads <= 0 means "all OpenMP threads".
 * Returns the number of items whose lengths were outside the defined domain. */
int mas_oracle_batch(int *paths, float *scores, const int *t_xs, const int *t_ys,
                     int B, int T_x, int T_y, float penalty, int n_threads)
{
    int bad = 0;
    size_t item = (size_t)T_x * (size_t)T_y;
#ifdef _OPENMP
    if (n_threads <= 0) n_threads = omp_get_max_threads();
    #pragma omp parallel for schedule(static) num_threads(n_threads) reduction(+:bad)
#else
    (void)n_threads;
#endif
    for (int b = 0; b < B; ++b) {
        int tx = t_xs[b], ty = t_ys[b];
        if (tx > T_x || ty > T_y) { bad += 1; continue; }
        int rc = mas_oracle_single(paths + b * item, scores + b * item, tx, ty, T_y, penalty);
        if (rc) bad += 1;
    }
    return bad;
}

int mas_oracle_max_threads(void)
{
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}
