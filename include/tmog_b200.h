/*
 * tmog_b200.h -- C ABI of libtmog_b200.so: the B200 (sm_100a) implementation of TransmogrifAI's data-parallel hot
 * path (transmogrify() vectorizers + SanityChecker statistics).
 *
 * This is the boundary a JNI shim binds.  Plain pointers and sizes only; no C++ or torch types.
 * Reference citations are relative to the TransmogrifAI tree:
 *   C/ = core/src/main/scala/com/salesforce/op/        U/ = utils/src/main/scala/com/salesforce/op/utils/
 *
 * Conventions
 *   - every function returns 0 on success, a negative TMOG_E* code otherwise; tmog_last_error() returns the message of
 *     the calling thread's last failure (the Scala side turns it into the IllegalArgumentException / RuntimeException
 *     the reference would have thrown).  No C++ exception crosses this boundary.
 *   - "*_dev" functions take DEVICE pointers and a stream (cudaStream_t passed as void*, NULL = default stream), are
 *     asynchronous and allocate nothing: the caller owns inputs, outputs and scratch.
 *   - functions without the suffix work on HOST memory only (tiny finishing steps on the driver).
 *   - "*_host" functions take (pinned) HOST buffers, stream them through the device in row chunks inside a tmog_ctx
 *     and return host results; these are the calls the reference-facing stages make per Spark partition.
 *   - numeric columns are Arrow-style columnar: column j is values + j*ld (ld = column stride in elements, a multiple
 *     of 2 so that every column is 16-byte aligned); validity is a bitmap per column, bit (r % 64) of 64-bit word
 *     (r / 64), 1 = present, column j at valid + j*ldv; valid == NULL means "no nulls".
 */
#ifndef TMOG_B200_H
#define TMOG_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define TMOG_OK 0
#define TMOG_EINVAL (-1)   /* bad argument (reference: require(...) -> IllegalArgumentException)            */
#define TMOG_ECUDA (-2)    /* CUDA runtime failure                                                          */
#define TMOG_ENOMEM (-3)   /* caller-provided buffer too small                                              */
#define TMOG_ESTATE (-4)   /* call sequence error                                                           */

/* Message of the calling thread's last failing call ("" if none). */
const char* tmog_last_error(void);
/* ABI version (bumped on any signature change). */
int tmog_abi_version(void);

/* ------------------------------------------------------------------------------------------------------------------
 * Real / RealNN / Binary vectorizers fused with the SanityChecker column scan
 * ------------------------------------------------------------------------------------------------------------------
 * One pass over the raw nullable f64 columns and the label produces, per column, the mergeable partial sums from
 * which BOTH the RealVectorizer fill value and every SanityChecker statistic of the vectorized columns follow
 * (SURVEY.md Appendix C).  A partial is TMOG_REAL_NSTAT doubles per column, columns 0..ncols-1 then the label:
 *   [0] n_valid   [1] sum d   [2] sum d^2   [3] sum d*e   [4] sum_valid e   [5] min x   [6] max x   [7] n_rows
 * with d = x - shift[j], e = y - shift[ncols] over the valid rows.  Partials of row chunks / GPUs / Spark partitions
 * that used the same shifts merge by addition (min / max for [5] / [6]).
 */
#define TMOG_REAL_NSTAT 8

/* Pilot shifts: mean of the valid values among the first min(n, 4096) rows per column, then the label's.
 * Keeps the single-pass centred sums well conditioned (replaces the reference's second pass over the data). */
int tmog_real_pilot_shift_dev(const double* values, int64_t ld, const uint64_t* valid, int64_t ldv,
                              const double* label, int64_t n, int ncols, double* shifts /* [ncols+1] */,
                              void* stream);

/* Scratch the fit kernel needs for (ncols) columns: per-CTA partials reduced by a second tiny kernel. */
int tmog_real_fit_scratch_bytes(int ncols, size_t* bytes);

/* Replaces  C/stages/impl/feature/RealVectorizer.scala:85-97  (means via U/spark/SequenceAggregators.scala:76-88)
 * and, for the columns this vectorizer emits, the three data passes of
 * C/stages/impl/preparators/SanityChecker.scala:407 (Statistics.colStats), :464-470 (label correlations) and
 * :265-272 (per-label column sums of the null indicators).
 * partials: [(ncols+1) * TMOG_REAL_NSTAT] doubles, overwritten. */
int tmog_real_fit_partials_dev(const double* values, int64_t ld, const uint64_t* valid, int64_t ldv,
                               const double* label, int64_t n, int ncols, const double* shifts,
                               void* scratch, size_t scratch_bytes, double* partials, void* stream);

/* acc += other for two partial blocks taken with the same shifts (the Aggregator.merge of
 * U/spark/SequenceAggregators.scala:83 and the treeAggregate combOps of SanityChecker's statistics). Host memory. */
int tmog_real_partials_merge(double* acc, const double* other, int ncols);

/* Finish on the host: fill values and the SanityChecker statistics of the vectorized columns.
 *   fill_with_mean != 0: fill = sum/count, 0.0 for an all-null column (SequenceAggregators.scala:84)
 *   else fill = fill_const (RealVectorizer.scala:76-81).
 * Output arrays have one entry per OUTPUT column, width = ncols * (track_nulls ? 2 : 1), interleaved
 * [value_j, isNull_j] as RealVectorizer.scala:112-123 lays them out; any of them may be NULL.
 *   mean/variance/min/max : Statistics.colStats of the filled column / the indicator (variance unbiased, n-1)
 *   cov_label             : sum (f - mean_f)(y - mean_y) as OpStatistics.scala:86-90 accumulates it
 *   label_stats[5]        : {n, mean, variance, min, max} of the label
 *   null_label_sum        : per input column, sum of y over the null rows (feeds the 2 x nLabels contingency table
 *                           of a lone null indicator, SanityChecker.scala:306-322, when the label is 0/1)
 */
int tmog_real_finish(const double* partials, const double* shifts, int ncols, int fill_with_mean, double fill_const,
                     int track_nulls, double* fills /* [ncols] */, double* mean, double* variance, double* min,
                     double* max, double* cov_label, double* label_stats /* [5] */,
                     double* null_label_sum /* [ncols] */);

/* Replaces RealVectorizerModel.transformFn  C/stages/impl/feature/RealVectorizer.scala:112-123  (track_nulls: the
 * interleaved [x or fill, isNull] dense layout) and RealNNVectorizer.transformFn  RealNNVectorizer.scala:52-55
 * (valid == NULL, track_nulls == 0).  Writes columns [out_col0, out_col0 + width) of a row-major dense block with
 * row stride out_ld doubles, so that several vectorizers fill one combined block in place (the dense half of
 * VectorsCombiner, C/stages/impl/feature/VectorsCombiner.scala:84-88). */
int tmog_real_transform_dev(const double* values, int64_t ld, const uint64_t* valid, int64_t ldv, int64_t n,
                            int ncols, const double* fills, int track_nulls, double* out, int64_t out_ld,
                            int64_t out_col0, void* stream);

/* Replaces BinaryVectorizer.transformFn  C/stages/impl/feature/BinaryVectorizer.scala:70-82.
 * values: one byte per cell (0 / non-zero), column stride ld bytes (multiple of 16). */
int tmog_binary_transform_dev(const uint8_t* values, int64_t ld, const uint64_t* valid, int64_t ldv, int64_t n,
                              int ncols, int fill_value, int track_nulls, double* out, int64_t out_ld,
                              int64_t out_col0, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* TMOG_B200_H */
