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

/* tmog_real_fit_partials_dev for callers that cannot allocate device memory.  No counterpart in the reference: this
 * entry point exists for hosts such as a JNI shim in a Spark executor whose GPU memory belongs to another
 * component's pool, so that a side cudaMalloc fails or is not allowed and the shim owns nothing on the device but
 * the partition's input and output buffers.  The per-CTA scratch is a fixed 4 MiB array built into the library
 * (held by every GPU that loads it, until exit), which caps a call at (4 MiB / (64 B x 2 x SMs)) - 1 columns --
 * 220 on a B200; wider blocks return TMOG_EINVAL with a message naming the limit.  Same inputs, outputs and results
 * as tmog_real_fit_partials_dev.  The first call on each device resolves the array's address and should be made
 * outside stream capture; all calls on a device share the array, so calls that overlap on the GPU race -- serialise
 * them.  Never calls cudaMalloc / cudaFree. */
int tmog_real_fit_partials_noalloc_dev(const double* values, int64_t ld, const uint64_t* valid, int64_t ldv,
                                       const double* label, int64_t n, int ncols, const double* shifts,
                                       double* partials, void* stream);

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
/* The same finish with the fill value of every column given (fill_values [ncols]): the statistics of columns filled
 * with their mode (IntegralVectorizer) or any other per-column constant. */
int tmog_real_finish_with_fills(const double* partials, const double* shifts, int ncols, const double* fill_values,
                                int track_nulls, double* mean, double* variance, double* mn, double* mx,
                                double* cov_label, double* label_stats, double* null_label_sum);

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

/* The two transforms above fused with SanityCheckerModel.transformFn -> removeFeatures
 * (C/stages/impl/preparators/DerivedFeatureFilterUtils.scala:291-305): keep_mask (device bitmap, bit i of word i/64)
 * marks the columns the vectorizer would emit (0 .. ncols * (track_nulls ? 2 : 1) - 1) that SanityChecker kept, nkeep
 * of them; they are written packed, in order, into columns [out_col0, out_col0 + nkeep) of the output rows.  Dropped
 * columns are never written. */
int tmog_real_transform_select_dev(const double* values, int64_t ld, const uint64_t* valid, int64_t ldv, int64_t n,
                                   int ncols, const double* fills, int track_nulls, const uint64_t* keep_mask,
                                   int nkeep, double* out, int64_t out_ld, int64_t out_col0, void* stream);
int tmog_binary_transform_select_dev(const uint8_t* values, int64_t ld, const uint64_t* valid, int64_t ldv,
                                     int64_t n, int ncols, int fill_value, int track_nulls,
                                     const uint64_t* keep_mask, int nkeep, double* out, int64_t out_ld,
                                     int64_t out_col0, void* stream);

/* ------------------------------------------------------------------------------------------------------------------
 * Pivot vectorizers (OpTextPivotVectorizer on PickList / Text columns; the shared OpOneHotVectorizer logic)
 * ------------------------------------------------------------------------------------------------------------------
 * Columns arrive dictionary-encoded: codes[ncols][ld] int32 column-major (ld a multiple of 4; code < 0 = null) plus
 * one utf8 dictionary per column.  String work happens once per DICTIONARY ENTRY on the host (the three functions
 * below); the per-row work is integer only. */

/* Replaces TextUtils.cleanString  U/text/TextUtils.scala:39-47  for a batch of strings (Arrow layout: utf8 bytes +
 * int64 offsets [count+1]).  ASCII only: a byte >= 0x80 returns TMOG_EINVAL (Java's Unicode case mapping is not
 * restated; such columns must run with cleanText = false). */
int tmog_clean_strings(const uint8_t* data, const int64_t* offsets, int64_t count, uint8_t* out_data,
                       int64_t out_capacity, int64_t* out_offsets);

/* ids[i] = dense id (first-appearance order) of string i among the distinct strings: raw dictionary entries that
 * clean to the same value share one count, as the Map[String, Int] keys of
 * C/stages/impl/feature/OpOneHotVectorizer.scala:219-225 do. */
int tmog_dedupe_strings(const uint8_t* data, const int64_t* offsets, int64_t count, int32_t* ids, int64_t* ndistinct);

/* Replaces  countOccurrences.filter(_._2 >= minSup).sortBy(v => -v._2 -> v._1).take(topK)
 * C/stages/impl/feature/OpOneHotVectorizer.scala:100-103: count descending, then java.lang.String.compareTo (UTF-16
 * code-unit order).  Entries with count 0 never existed in the reference's map and are skipped.
 * out_ids [top_k] receives the chosen entry indices in vocabulary order, *out_n their number. */
int tmog_pivot_topk(const uint8_t* data, const int64_t* offsets, const int64_t* counts, int64_t count, int top_k,
                    int64_t min_support, int32_t* out_ids, int* out_n);

/* The driver-side finish of ONE pivoted column in one call (thread-safe; run the columns in parallel): cleans the
 * dictionary (tmog_clean_strings), merges raw codes that clean alike (tmog_dedupe_strings), picks the vocabulary
 * (tmog_pivot_topk) and sums the per-label-class counts of every emitted column.
 *   raw_counts [(ncodes + 1) * nclasses]   this column's table from tmog_pivot_count_dev (row 0 = null), as int64
 *   n_top, top_data / top_offsets [top_k+1] the vocabulary strings in order (Arrow layout)
 *   code_to_out [ncodes]                    emitted-column index of every raw code: top position, or n_top = OTHER
 *   out_class_counts [(top_k + 2) * nclasses] rows 0..n_top-1 = top values, n_top = OTHER, n_top + 1 = null */
int tmog_pivot_column_finish(const uint8_t* dict_data, const int64_t* dict_offsets, int64_t ncodes, int clean_text,
                             const int64_t* raw_counts, int nclasses, int top_k, int64_t min_support, int* n_top,
                             int32_t* code_to_out, int64_t* out_class_counts, uint8_t* top_data, int64_t top_capacity,
                             int64_t* top_offsets);

/* The same finish in two halves, so that the half that needs no counts -- cleaning every dictionary entry
 * (TextUtils.cleanString, U/text/TextUtils.scala:39-47) and grouping the entries that clean alike (the key-wise sum of
 * the Algebird map monoid in OpOneHotVectorizer.fitFn, OpOneHotVectorizer.scala:95-97) -- runs on host threads while
 * the count kernel is still busy on the GPU.  prepare() returns an opaque handle (thread-safe, one per column);
 * finish() takes the counts and produces what tmog_pivot_column_finish does; free() releases the handle. */
/* 64-bit digest of a byte range (host).  Ranks that exported their partitions separately compare the digests of their
 * dictionaries (bytes and offsets, two seeds) to see whether the count tables can be summed as they are or must first
 * be re-indexed against the union dictionary -- the bookkeeping Spark's map-monoid reduce does by key. */
int tmog_hash_bytes(const uint8_t* data, int64_t nbytes, uint64_t seed, uint64_t* out);

typedef struct tmog_pivot_dict tmog_pivot_dict;
int tmog_pivot_dict_prepare(const uint8_t* dict_data, const int64_t* dict_offsets, int64_t ncodes, int clean_text,
                            tmog_pivot_dict** out);
int tmog_pivot_dict_finish(const tmog_pivot_dict* dict, const int64_t* raw_counts, int nclasses, int top_k,
                           int64_t min_support, int* n_top, int32_t* code_to_out, int64_t* out_class_counts,
                           uint8_t* top_data, int64_t top_capacity, int64_t* top_offsets);
int tmog_pivot_dict_free(tmog_pivot_dict* dict);

/* code_to_out for a dictionary other than the fitted one (transform of new data): position of each entry's cleaned
 * value in the fitted vocabulary, or n_top (OTHER) -- OneHotModelFun.pivotFn :420-427. */
int tmog_pivot_column_map(const uint8_t* dict_data, const int64_t* dict_offsets, int64_t ncodes, int clean_text,
                          const uint8_t* top_data, const int64_t* top_offsets, int n_top, int32_t* code_to_out);

/* Replaces the value counting of OpOneHotVectorizer.fitFn (:95-97, the RDD.fold of per-row maps) and, in the same
 * read, the per-label column sums SanityChecker.categoricalTests needs for the pivoted columns
 * (C/stages/impl/preparators/SanityChecker.scala:265-272):  counts[count_offset_c + id * nclasses + class] += 1.
 *   col_desc [ncols * 4] int64 (device): per column {nids, slot_offset, count_offset, ncodes};
 *                        id = 1 + slots[slot_offset + code] for code >= 0, id = 0 for a null (nids = values + 1);
 *   slots                concatenated int32 code -> id tables (device), or NULL: id = 1 + code (count per raw code and
 *                        merge the codes that clean alike afterwards, on the host -- no gather in the kernel);
 *   class_ids            one byte per row (tmog_label_classes_dev); ignored when nclasses == 1;
 *   counts               uint32 bins (device; native 32-bit reductions), NOT cleared here, so a call adds to what
 *                        is there; at most 2^32 - 1 rows may be accumulated into one table -- widen to int64 on
 *                        the host when merging row chunks / ranks;
 *   bad_code             device int set to 1 if a code >= ncodes was met. */
int tmog_pivot_count_dev(const int32_t* codes, int64_t ld, int64_t n, int ncols, const uint8_t* class_ids,
                         int nclasses, const int64_t* col_desc, const int32_t* slots, int64_t max_nids,
                         uint32_t* counts, int* bad_code, void* stream);

/* Binary columns (u8 values + validity) as codes -1 null / 0 false / 1 true in codes[ncols][ldc] (ldc a multiple of 4),
 * so that tmog_pivot_count_dev yields the SanityChecker sums of BinaryVectorizer's [value, isNull] columns
 * (C/stages/impl/feature/BinaryVectorizer.scala:70-82) in the same pass as the pivoted columns. */
int tmog_binary_codes_dev(const uint8_t* values, int64_t ld, const uint64_t* valid, int64_t ldv, int64_t n, int ncols,
                          int32_t* codes, int64_t ldc, void* stream);

/* The label side of the covariance of one-hot / null-indicator columns with a CONTINUOUS label
 * (U/stats/OpStatistics.scala:86-90), next to the plain counts of tmog_pivot_count_dev (nclasses = 1):
 * wsum [wsum_offset[c] + id] += label[r] - shift for the row's id in column c (0 = null, 1 + slot(code); col_desc and
 * slots as in tmog_pivot_count_dev; wsum is f64 with nids_c entries per column, NOT cleared here). */
int tmog_pivot_label_sum_dev(const int32_t* codes, int64_t ld, int64_t n, int ncols, const int64_t* col_desc,
                             const int32_t* slots, const int64_t* wsum_offset, const double* label, double shift,
                             double* wsum, void* stream);

/* Replaces OneHotModelFun.pivotFn  C/stages/impl/feature/OpOneHotVectorizer.scala:416-437  for single-valued inputs
 * (exactly one of {top value, OTHER, null indicator} is 1.0 per row and column) fused with reindex
 * (Transmogrifier.scala:437-449) and SanityCheckerModel's removeFeatures:  out[r][out_col0 + c] = pos_c[code], the
 * final column index of the 1.0 inside the output vector, or -1 if SanityChecker dropped that column.
 *   pos                  concatenated int32 tables (device), table c has ncodes_c + 1 entries (the last is the null
 *                        indicator's index) and starts at slot_offset_c + c;
 *   out                  int32 row-major, row stride out_ld. */
int tmog_pivot_transform_dev(const int32_t* codes, int64_t ld, int64_t n, int ncols, const int64_t* col_desc,
                             const int32_t* pos, int32_t* out, int64_t out_ld, int64_t out_col0, void* stream);

/* ------------------------------------------------------------------------------------------------------------------
 * Text columns (SmartTextVectorizer: TextStats, tokeniser, hashing-TF)
 * ------------------------------------------------------------------------------------------------------------------
 * Raw utf8 columns in Arrow layout.  col_desc [ncols * 4] int64 (device): per column {data pointer, offsets pointer
 * (int64 [n+1]), validity bitmap pointer or 0, 0}.  ASCII is the parity contract: a byte >= 0x80 sets flags[0]. */

/* Replaces TextStats of SmartTextVectorizer.fitFn  C/stages/impl/feature/SmartTextVectorizer.scala:79-124,207-299:
 * per column an open-addressing table (`slots` entries, power of two) over the 64-bit hashes of the CLEANED cells
 * (TextUtils.cleanString applied on the fly when clean_text != 0): distinct count, per-label-class counts, first row of
 * every distinct value (the host reads those rows to recover the strings), and sum / sum of squares of the cleaned
 * lengths.  A column that shows more than max_distinct distinct values stops being recorded (only ndistinct >
 * max_distinct is reported): that is all SmartTextVectorizer needs to route it to hashing.
 * `tables` = tmog_text_stats_layout(): head_bytes + ncols * per_column_bytes of device memory; column c's block is at
 * head + c * per_column: hash u64 [slots] | first_row u64 [slots] | counts u32 [slots * nclasses] (padded to 8 B) |
 * ndistinct i32 (+4 pad) | len_sum u64 [2] | nonnull u64. */
int tmog_text_stats_layout(int ncols, int slots, int nclasses, size_t* head_bytes, size_t* per_column_bytes);
int tmog_text_stats_table_bytes(int slots, int nclasses, size_t* bytes);
int tmog_text_stats_dev(const int64_t* col_desc, int ncols, int64_t n, const uint8_t* class_ids, int nclasses,
                        int clean_text, int slots, int max_distinct, void* tables, int32_t* flags, void* stream);
/* The same scan without the length sums of columns that are past the cardinality cap: lengthCounts only feed
 * `lengthStdDev < minLengthStdDev` (SmartTextVectorizer.scala:113), which never holds at the default minLengthStdDev = 0,
 * so such a column needs nothing further from its cells and they are not read: non-ASCII bytes in them are reported
 * by whichever call vectorizes the column afterwards (tmog_text_hash_*_dev, tmog_text_codes_dev: flags[0]).  len_sum
 * and nonnull then cover the cells seen before the cap only. */
int tmog_text_stats_values_dev(const int64_t* col_desc, int ncols, int64_t n, const uint8_t* class_ids, int nclasses,
                        int clean_text, int slots, int max_distinct, void* tables, int32_t* flags, void* stream);

/* TextStats.lengthCounts (SmartTextVectorizer.scala:85,207-299; decides Hash against Ignore through lengthStdDev <
 * minLengthStdDev, :113-115): hist [ncols * nbins] += cells (tokens == 0: by the length of the cleaned cell,
 * TextLengthType.FullEntry) or surviving tokens (tokens != 0: by token length, TextLengthType.Tokens) of every length,
 * lengths >= nbins - 1 in the last bin; first_row [ncols * nbins] = min(row_base + row) over the rows a length was met
 * in (the caller starts it at all-ones).  Neither array is cleared here: row chunks add up.  The reference's fold
 * stops adding to a map that holds more than maxCardinality keys; the caller finds the row where that happens from
 * first_row and asks again for the rows up to it. */
int tmog_text_length_hist_dev(const int64_t* col_desc, int ncols, int64_t n, int64_t row_base, int clean_text,
                              int tokens, int min_token_length, int nbins, uint64_t* hist, uint64_t* first_row,
                              int32_t* flags, void* stream);

/* Dictionary codes for the text columns SmartTextVectorizer pivots: codes[k][r] = table slot of row r's cleaned value
 * in column col_index[k] (-1 null), which tmog_pivot_count_dev / tmog_pivot_transform_dev then consume
 * (SmartTextVectorizerModel.transformFn's categoricalPivotFn, SmartTextVectorizer.scala:333-345).  flags (or NULL):
 * flags[0] is set when clean_text != 0 and a cell holds a non-ASCII byte (its cleaning is then not the reference's). */
int tmog_text_codes_dev(const int64_t* col_desc, const int32_t* col_index, int nsel, int64_t n, int clean_text,
                        int slots, const void* tables, int32_t* codes, int64_t ld, int32_t* flags, void* stream);

/* ---- OpSetVectorizer (MultiPickList) ------------------------------------------------------------------------------
 * Set columns in Arrow List<dictionary> layout (device): int64 offsets [n + 1], int32 codes [offsets[n]]; row r holds
 * the codes [offsets[r], offsets[r + 1]) without duplicates, an empty range is the empty set.  The dictionary strings
 * stay on the host; tmog_clean_strings / tmog_dedupe_strings give the cleaned value (group) of every code.
 *
 * tmog_set_count_dev replaces OpSetVectorizer.convertToSeqOfMaps + the count fold of OpOneHotVectorizer.fitFn
 * (C/stages/impl/feature/OpOneHotVectorizer.scala:75-124,171-177): counts [(1 + ngroups) * nclasses] (u32, NOT cleared
 * here) += rows of each label class holding the cleaned value gid[code] (raw values of a row that clean alike count
 * once); entry 0 counts the empty sets.  *bad is set if a code is outside [0, ncodes). */
int tmog_set_count_dev(const int64_t* offsets, const int32_t* codes, int64_t n, const int32_t* gid, int ncodes,
                       int ngroups, const uint8_t* class_ids, int nclasses, uint32_t* counts, int32_t* bad,
                       void* stream);

/* OneHotModelFun.pivotFn for sets (OpOneHotVectorizer.scala:416-437): with pos [ncodes] = output slot of every code
 * (0 .. k-1 the top values, k = OTHER) row r has the entries (slot, number of its raw values with that slot) and, for an
 * empty set with track_nulls, (k + 1, 1).
 * _stats: SanityChecker sums over k + 2 slots (NOT cleared here): class_sum [slot * nclasses + class] += value,
 *   present [slot * nclasses + class] += 1 (the column clipped to <= 1, SanityChecker.scala:258-268), and for the rare
 *   value > 1: sq_extra [slot] += v^2 - v, aux [3 * slot + {0, 1, 2}] = max v, max ~v, rows with v > 1.
 * _transform: the entries whose newpos [slot] >= 0, as (newpos, value) in slot order.  Call once with row_start = NULL
 *   to get row_nnz [n], prefix-sum it into row_start [n], call again to write out_idx / out_val. */
int tmog_set_stats_dev(const int64_t* offsets, const int32_t* codes, int64_t n, const int32_t* pos, int ncodes, int k,
                       int track_nulls, const uint8_t* class_ids, int nclasses, uint32_t* class_sum, uint32_t* present,
                       uint64_t* sq_extra, uint32_t* aux, const double* label, double label_shift, double* wsum,
                       int32_t* bad, void* stream);
int tmog_set_transform_dev(const int64_t* offsets, const int32_t* codes, int64_t n, const int32_t* pos, int ncodes,
                           int k, int track_nulls, const int32_t* newpos, const int64_t* row_start, int32_t* row_nnz,
                           int32_t* out_idx, int32_t* out_val, int32_t* bad, void* stream);

/* Replaces TextTokenizer.tokenize (C/stages/impl/feature/TextTokenizer.scala:160-220; Lucene 7.3 StandardAnalyzer with
 * the Snowball english stop set, C/utils/text/LuceneTextAnalyzer.scala:103-122,159-166) + HashingFun.hash / prepare
 * (C/stages/impl/feature/OPCollectionHashingVectorizer.scala:244-305; Spark 2.4.5 HashingTF: legacy murmur3_x86_32
 * seed 42, non-negative mod) + the null indicators of SmartTextVectorizerModel (SmartTextVectorizer.scala:363-371).
 *   prefix [ncols * 16] / prefix_len [ncols] (device): the bytes of "<hasher.indexOf(featureName)>_" per column;
 *   params [12] (HOST): numFeatures, sharedHashSpace, binaryFreq, minTokenLength, trackNulls, prependFeatureName,
 *                      null_base (vector index of the first null indicator; >= width of the hash block), nclasses,
 *                      n_hashed (0 = all: columns [0, n_hashed) are hashed, the ones behind them are the columns
 *                      SmartTextVectorizer ignores -- tokenised for their null indicator / text length only,
 *                      SmartTextVectorizer.scala:342-354), len_base (0 = none: vector index of the first text-length
 *                      column, trackTextLen :373-375; value = summed length of the surviving tokens), 0, 0.
 * Vector index of a hit: bucket (+ j * numFeatures for separate spaces); null indicator of column j: null_base + j.
 * text bytes are read with aligned 16-byte loads, so every column's `data` must lie in an allocation whose bounds are
 * 16-byte aligned (true of anything cudaMalloc returns).
 * scratch: tmog_text_hash_scratch_bytes(n) bytes of device memory, 8-byte aligned, contents irrelevant (row counters and
 *   the list of rows with too many distinct buckets for the fast table, which a second launch redoes).
 * flags [4] (device): [0] non-ASCII byte met (unsupported), [1] a run of token characters longer than ~4000 bytes,
 *   [2] a row with more than ~12000 distinct buckets, [3] sparse capacity exceeded.
 *
 * _stats: SanityChecker sums per vector index (NOT cleared here; sized null_base + ncols):
 *   bucket_class_sum [index * nclasses + class] += term frequency tf of every row holding the index;
 *   for the rare tf > 1 only: bucket_dup [index] += tf - 1 (rows holding the index = sum of class sums - dup),
 *   bucket_sq_extra [index] += tf^2 - tf   (so that sum v^2 = sum v + extra), and
 *   bucket_aux [3 * index + {0, 1, 2}] (zeroed by the caller) = max tf, ~(min tf among tf > 1), rows with tf > 1 --
 *   they give the exact column max / min.
 *   Continuous label (more distinct values than label classes are tracked; nclasses = 1): with label (device f64 [n])
 *   and bucket_wsum (f64 per index) given, bucket_wsum [index] += tf * (label[row] - label_shift), which with the sums
 *   above gives the covariance of every column with the label (U/stats/OpStatistics.scala:86-90); both NULL otherwise.
 *   tmog_set_stats_dev's label / wsum [k + 2] and tmog_pivot_label_sum_dev do the same for their columns.
 * _transform: sparse rows in 4-array CSR form: row r owns out_idx / out_val [row_start[r], row_start[r] + row_nnz[r]),
 *   indices ascending; newpos [null_base + ncols] maps a vector index to its output index or -1 (dropped by
 *   SanityChecker); *cursor (device u64, zeroed by the caller) ends as the total number of entries; entries beyond
 *   `capacity` are counted but not written (flags[3]) -- call with capacity 0 to size the buffers. */
int tmog_text_hash_scratch_bytes(int64_t n, size_t* bytes);
/* Scratch that also lets the three calls below take the per-cell path (separate hash spaces of at most 512 buckets per
 * column, the TransmogrifierDefaults layout): a lane tokenises one cell into a 512-bit bucket bitmap, statistics come
 * from warp-transposed bitmaps, and the sparse rows are assembled from 128-byte cell records by an HBM-bound second
 * kernel (sparse_out != 0 adds n * ncols * 128 bytes for the records).  With less scratch than this (but at least
 * tmog_text_hash_scratch_bytes) the calls run the per-row table kernel; ignored columns and text lengths need this one. */
int tmog_text_hash_scratch_bytes2(int64_t n, int ncols, const int32_t* params, int sparse_out, size_t* bytes);
int tmog_text_hash_stats_dev(const int64_t* col_desc, int ncols, int64_t n, const uint8_t* prefix,
                             const int32_t* prefix_len, const int32_t* params, const uint8_t* class_ids,
                             uint32_t* bucket_class_sum, uint32_t* bucket_dup, uint64_t* bucket_sq_extra,
                             uint32_t* bucket_aux, const double* label, double label_shift, double* bucket_wsum,
                             int32_t* flags, void* scratch, size_t scratch_bytes, void* stream);
int tmog_text_hash_transform_dev(const int64_t* col_desc, int ncols, int64_t n, const uint8_t* prefix,
                                 const int32_t* prefix_len, const int32_t* params, const int32_t* newpos,
                                 int64_t* row_start, int32_t* row_nnz, int32_t* out_idx, int32_t* out_val,
                                 uint64_t* cursor, int64_t capacity, int32_t* flags, void* scratch,
                                 size_t scratch_bytes, void* stream);

/* Fit and transform of the hashed text columns in ONE pass over the text (what an OpWorkflow does on its training
 * data: SmartTextVectorizer's model transforms the rows it was just fitted on, FitStagesUtil.scala:261-269): the
 * statistics of tmog_text_hash_stats_dev over all vector indices and the sparse rows of
 * tmog_text_hash_transform_dev (normally with newpos = identity, since SanityChecker has not decided yet).  The
 * rows are then cut down to SanityChecker's kept columns by tmog_sparse_rows_select_dev instead of a second
 * tokenise + hash pass.  Arguments as in the two functions above. */
int tmog_text_hash_fit_transform_dev(const int64_t* col_desc, int ncols, int64_t n, const uint8_t* prefix,
                                     const int32_t* prefix_len, const int32_t* params, const uint8_t* class_ids,
                                     uint32_t* bucket_class_sum, uint32_t* bucket_dup, uint64_t* bucket_sq_extra,
                                     uint32_t* bucket_aux, const double* label, double label_shift,
                                     double* bucket_wsum, const int32_t* newpos, int64_t* row_start, int32_t* row_nnz,
                                     int32_t* out_idx, int32_t* out_val, uint64_t* cursor, int64_t capacity,
                                     int32_t* flags, void* scratch, size_t scratch_bytes, void* stream);

/* Packed cells: the statistics pass of a row chunk can keep what it found in every cell -- the cell's buckets and their
 * term frequencies, ~1.1 bytes per distinct token -- so that the transform of the same rows, which a chunk-streamed
 * fit-then-transform runs after all chunks have been fitted (FitStagesUtil.scala:212-290; the reference tokenises and
 * hashes the rows again, SmartTextVectorizer.scala:333-356), reads ~0.9 KB per row instead of the text and tokenises
 * nothing.  Separate hash spaces of at most 512 buckets per column only (tmog_text_hash_scratch_bytes2's layout).
 *   arena: device memory shared by all chunks, arena_cursor (device u64, zeroed once) counts the bytes handed
 *          out; pack_meta: per chunk, tmog_text_cells_pack_bytes bytes (track_len = the vector has text-length
 *          columns, params[9] > 0); pack_flags[0] is set when the arena was too small for this chunk -- its packed
 *          cells are then unusable and the chunk is transformed from its text (tmog_text_hash_transform_dev).
 *   _stats_pack_dev = tmog_text_hash_stats_dev + the packed cells of the chunk;
 *   _transform_packed_dev = tmog_text_hash_transform_dev's outputs (same entries, same order) from the packed cells. */
int tmog_text_cells_pack_bytes(int64_t n, int ncols, int track_len, size_t* meta_bytes);
int tmog_text_hash_stats_pack_dev(const int64_t* col_desc, int ncols, int64_t n, const uint8_t* prefix,
                                  const int32_t* prefix_len, const int32_t* params, const uint8_t* class_ids,
                                  uint32_t* bucket_class_sum, uint32_t* bucket_dup, uint64_t* bucket_sq_extra,
                                  uint32_t* bucket_aux, const double* label, double label_shift, double* bucket_wsum,
                                  int32_t* flags, void* scratch, size_t scratch_bytes, void* pack_meta, void* arena,
                                  uint64_t arena_bytes, uint64_t* arena_cursor, int32_t* pack_flags, void* stream);
int tmog_text_hash_transform_packed_dev(int ncols, int64_t n, const int32_t* params, void* pack_meta, void* arena,
                                        uint64_t arena_bytes, const int32_t* newpos, int64_t* row_start, int32_t* row_nnz,
                                        int32_t* out_idx, int32_t* out_val, uint64_t* cursor, int64_t capacity,
                                        int32_t* flags, void* stream);

/* Replaces SanityCheckerModel.transformFn -> removeFeatures (C/stages/impl/preparators/DerivedFeatureFilterUtils.scala:
 * 291-305) on sparse rows in 4-array CSR form: row r keeps the entries whose newpos [index] >= 0, as (newpos, value),
 * in order.  Call once with out_start = NULL to get out_nnz [n], prefix-sum it into out_start [n], call again to write
 * out_idx / out_val. */
int tmog_sparse_rows_select_dev(const int64_t* row_start, const int32_t* row_nnz, const int32_t* idx,
                                const int32_t* val, int64_t n, const int32_t* newpos, const int64_t* out_start,
                                int32_t* out_nnz, int32_t* out_idx, int32_t* out_val, void* stream);

/* ------------------------------------------------------------------------------------------------------------------
 * Label classes
 * ------------------------------------------------------------------------------------------------------------------
 * Replaces the keying of  C/stages/impl/preparators/SanityChecker.scala:265-281  (reduceByKey on the label value,
 * distinct-label count, labels sorted ascending).  Distinctness follows java.lang.Double.equals (bit patterns, one
 * NaN).  classes [max_classes] receives the sorted distinct values, *nclasses (DEVICE int) their number, or -1 when
 * there are more than max_classes (<= 255): the label is then treated as continuous.  class_ids (may be NULL)
 * receives one byte per row: the index of the row's label in classes; class_counts [max_classes] (may be NULL) the
 * number of rows per class (the "contingency(d) = 1.0" slot of SanityChecker.scala:270). */
int tmog_label_classes_scratch_bytes(size_t* bytes);
int tmog_label_classes_dev(const double* label, int64_t n, int max_classes, double* classes, int* nclasses,
                           uint8_t* class_ids, uint64_t* class_counts, void* scratch, size_t scratch_bytes,
                           void* stream);

/* Class ids / counts against a GIVEN sorted class list (device arrays classes [*nclasses], *nclasses a device int):
 * the second half of tmog_label_classes_dev, for row shards that must share one class list (multi-GPU: each rank
 * discovers its classes, the union is broadcast, every rank maps its rows against the union). */
int tmog_label_class_ids_dev(const double* label, int64_t n, const double* classes, const int* nclasses,
                             int max_classes, uint8_t* class_ids, uint64_t* class_counts, void* stream);

/* counts[c * nclasses + k] = number of NULL rows of column c whose label class is k: the 2 x nLabels contingency
 * table of a lone null indicator (SanityChecker.scala:306-322) for labels with any number of classes.
 * accumulate != 0 adds to counts (row chunks), else counts is overwritten. */
int tmog_null_class_counts_dev(const uint64_t* valid, int64_t ldv, const uint8_t* class_ids, int64_t n, int ncols,
                               int nclasses, int accumulate, uint64_t* counts, void* stream);

/* ---- IntegralVectorizer (Integral features: int64 values + validity bitmap, device) ---------------------------------
 * Replaces IntegralVectorizer.fitFn / IntegralVectorizerModel.transformFn
 * (C/stages/impl/feature/IntegralVectorizer.scala:47-117) and SequenceAggregators.ModeSeqNullInt
 * (U/spark/SequenceAggregators.scala:100-131): per column the map value -> count over the non-null rows; the mode is
 * the most frequent value, the smallest one among equally frequent values, 0 for a column without values.
 *
 * tmog_integral_mode_dev counts into one open-addressing table per column (cap slots, a power of two >= 64) laid out in
 * `scratch` (tmog_integral_mode_scratch_bytes; tmog_integral_mode_layout gives the byte offsets of keys [ncols][cap]
 * u64, counts [ncols][cap] u64, min_count [ncols] u64 (rows holding INT64_MIN, whose bit pattern marks a free slot),
 * ndistinct [ncols] i32, overflow [ncols] i32 for callers that merge the maps of several GPUs).  A column with more
 * than 3/4 cap distinct values sets overflow[c]: call again with a larger cap.  modes [ncols], mode_counts [ncols]. */
int tmog_integral_mode_scratch_bytes(int ncols, int64_t cap, size_t* bytes);
int tmog_integral_mode_layout(int ncols, int64_t cap, size_t* offsets5);
int tmog_integral_mode_dev(const int64_t* values, int64_t ld, const uint64_t* valid, int64_t ldv, int64_t n, int ncols,
                           int64_t cap, void* scratch, size_t scratch_bytes, int64_t* modes, uint64_t* mode_counts,
                           void* stream);

/* The fused statistics scan and the transform of tmog_real_* over int64 columns (`i.value.getOrElse(m).toDouble`,
 * IntegralVectorizer.scala:103-113): values are converted to double as they are read.  Same arguments, scratch
 * (tmog_real_fit_scratch_bytes) and partials layout; finish with tmog_real_finish_with_fills (fills = the modes, or
 * the constant of setFillWithConstant). */
int tmog_integral_pilot_shift_dev(const int64_t* values, int64_t ld, const uint64_t* valid, int64_t ldv,
                                  const double* label, int64_t n, int ncols, double* shifts, void* stream);
int tmog_integral_fit_partials_dev(const int64_t* values, int64_t ld, const uint64_t* valid, int64_t ldv,
                                   const double* label, int64_t n, int ncols, const double* shifts, void* scratch,
                                   size_t scratch_bytes, double* partials, void* stream);
int tmog_integral_transform_dev(const int64_t* values, int64_t ld, const uint64_t* valid, int64_t ldv, int64_t n,
                                int ncols, const double* fills, int track_nulls, double* out, int64_t out_ld,
                                int64_t out_col0, void* stream);
int tmog_integral_transform_select_dev(const int64_t* values, int64_t ld, const uint64_t* valid, int64_t ldv, int64_t n,
                                       int ncols, const double* fills, int track_nulls, const uint64_t* keep_mask,
                                       int nkeep, double* out, int64_t out_ld, int64_t out_col0, void* stream);

/* ------------------------------------------------------------------------------------------------------------------
 * SanityChecker over a materialised dense OPVector block (row-major f64, row stride ld, d feature columns)
 * ------------------------------------------------------------------------------------------------------------------
 * For vectors that did not come out of the fused vectorizer scans.  The label is treated as column d. */

/* Pilot shifts [d+1]: column means over the first min(n, 4096) rows. */
int tmog_dense_pilot_shift_dev(const double* x, int64_t ld, int64_t n, int d, const double* label, double* shifts,
                               void* stream);
int tmog_dense_stats_scratch_bytes(int64_t n, int d, size_t* bytes);

/* Replaces Statistics.colStats (C/stages/impl/preparators/SanityChecker.scala:407), the label covariances of
 * U/stats/OpStatistics.scala:86-90 and, for the columns flagged in cat_mask [d], the per-label column sums of
 * SanityChecker.scala:265-272.  partials [(d+1) * TMOG_REAL_NSTAT] has the layout of tmog_real_fit_partials_dev
 * (n_valid = n) and is finished by tmog_real_finish(track_nulls = 0); class_sums [nclasses * d] may be NULL. */
int tmog_dense_stats_dev(const double* x, int64_t ld, int64_t n, int d, const double* label,
                         const uint8_t* class_ids, int nclasses, const uint8_t* cat_mask, const double* shifts,
                         void* scratch, size_t scratch_bytes, double* partials, double* class_sums, void* stream);

/* Replaces SanityCheckerModel.transformFn -> removeFeatures
 * (C/stages/impl/preparators/DerivedFeatureFilterUtils.scala:291-305) on a dense vector: out[r][j] = x[r][idx[j]]. */
int tmog_dense_gather_dev(const double* x, int64_t ld, int64_t n, const int32_t* idx, int k, double* out,
                          int64_t out_ld, void* stream);

/* Replaces the Gramian pass of Statistics.corr (SanityChecker.scala:467-470; Spark RowMatrix.computeGramianMatrix):
 * gram [(d+1)^2] += (X - s)^T (X - s), upper triangle (i <= j) only, label as column d.  The caller zeroes gram
 * before the first row chunk; chunks / GPUs add. Fails for d + 1 > 65535 like the reference. */
int tmog_dense_gram_dev(const double* x, int64_t ld, int64_t n, int d, const double* label, const double* shifts,
                        double* gram, void* stream);

/* Sparse OPVector parts written into columns [out_col0, ...) of a dense row-major block that the caller has zeroed
 * (input of tmog_dense_gram_dev for CorrelationLevel.Computed): the one-hot form of tmog_pivot_transform_dev
 * (indices [n][idx_ld], -1 = no entry, value 1.0) and the 4-array sparse rows of tmog_text_hash_transform_dev. */
int tmog_onehot_densify_dev(const int32_t* indices, int64_t idx_ld, int64_t n, int k, double* out, int64_t out_ld,
                            int64_t out_col0, void* stream);
int tmog_sparse_rows_densify_dev(const int64_t* row_start, const int32_t* row_nnz, const int32_t* indices,
                                 const int32_t* values, int64_t n, double* out, int64_t out_ld, int64_t out_col0,
                                 void* stream);

/* Host: Pearson matrix [(D)^2] from the shifted Gramian and the shifted column sums (partials[.][1]) -- Spark's
 * computeCovariance + computeCorrelationMatrixFromCovariance: NaN where |cov_ii| <= 1e-12, diagonal 1.0. */
int tmog_pearson_from_gram(int D, double n, const double* gram, const double* sums, double* corr);

/* ------------------------------------------------------------------------------------------------------------------
 * SanityChecker driver-side finishing (host memory, tiny inputs)
 * ------------------------------------------------------------------------------------------------------------------ */

/* Replaces OpStatistics.contingencyStats  U/stats/OpStatistics.scala:300-326 (filterEmpties :141-175,
 * chiSquaredTestOnFiltered :202-211 = Spark Statistics.chiSqTest without continuity correction, mutualInfo :234-271,
 * maxConfidences :280-290).  matrix is rows x cols ROW-major: rows = feature choices, cols = label values.
 * Outputs (any may be NULL): cramers_v, chi2, mutual_info scalars; pmi [rows*cols] row-major; max_conf, supports
 * [rows].  Returns 1 (not an error) when the matrix is empty after filtering: the reference then yields NaN and empty
 * maps, and every output is NaN. */
int tmog_contingency_stats(const double* matrix, int rows, int cols, double* cramers_v, double* chi2,
                           double* mutual_info, double* pmi, double* max_conf, double* supports);

/* tmog_contingency_stats for every categorical feature group of a vector in one call: matrix g is rows[g] x cols[g]
 * row-major at data + offsets[g]; pmi uses the layout of data; max_conf / supports are packed by row starting at
 * row_offsets[g]; status[g] = 0, or 1 when group g is empty after filtering. */
int tmog_contingency_stats_batch(int ngroups, const int32_t* rows, const int32_t* cols, const int64_t* offsets,
                                 const int64_t* row_offsets, const double* data, double* cramers_v, double* chi2,
                                 double* mutual_info, double* pmi, double* max_conf, double* supports,
                                 int32_t* status);

/* Statistics of hashed term-frequency columns (Statistics.colStats + the label covariance of OpStatistics.scala:86-90)
 * from the per-bucket sums of tmog_text_hash_*_dev: class_sum [w][nclasses] (sum of tf per label class), nnz [w] rows
 * holding the bucket, sq_extra [w] = sum (tf^2 - tf), aux [w][3] = {max tf, ~min tf among tf > 1, rows with tf > 1};
 * classes [nclasses] the label values, n rows, mean_y the label mean.  Host arrays. */
int tmog_bucket_stats(int64_t w, int nclasses, const double* class_sum, const double* nnz, const double* sq_extra,
                      const int64_t* aux, const double* classes, double n, double mean_y, double* mean,
                      double* variance, double* mn, double* mx, double* cov);

/* corr_i = cov_i / (sqrt(var_i) sqrt(var_y) (n - 1)), 0/0 kept as NaN  (U/stats/OpStatistics.scala:92-94). */
int tmog_pearson_from_cov(int d, double n, const double* cov_label, const double* variance, double label_variance,
                          double* corr);

/* Replaces DerivedFeatureFilterUtils.makeColumnStatistics' parent-level maxima
 * C/stages/impl/preparators/DerivedFeatureFilterUtils.scala:112-160, getFeaturesToDrop :234-281 and
 * ColumnStatistics.reasonsToRemove :350-405 with column metadata reduced to integers by the caller:
 *   d                 feature columns (label excluded)
 *   variance[d]
 *   corr_label        correlations with the label, indexed through corr_pos (NULL: none computed)
 *   corr_matrix       corr_dim x corr_dim row-major Pearson matrix incl. the label as last row/col, or NULL
 *                     (CorrelationLevel.Off)
 *   corr_pos[d]       position of column i inside corr_label / corr_matrix, -1 if excluded (hashed text exclusion)
 *   parent_offsets[d+1], parent_ids   CSR list of parent-feature ids per column (parentNamesWithMapKeys)
 *   group_id[d]       feature-group id (featureGroup()) or -1
 *   text_shared_hash[d]  isTextSharedHash :410-414
 *   cramers_v[d], rule_count[d]  rule_count = -1: column is in no categorical group (cramersV = None); else the
 *                     number (0..2) of (confidence, support) pairs in rule_conf / rule_support [2*d]
 *   params[9]         minVariance, minCorrelation, maxCorrelation, maxFeatureCorr, maxCramersV, maxRuleConfidence,
 *                     minRequiredRuleSupport, removeFeatureGroup, protectTextSharedHash
 * Outputs: drop[d] (1 = remove); reasons[d] bit mask (1 variance, 2 minCorr, 4 maxCorr, 8 featureCorr, 16 cramersV,
 * 32 ruleConfidence, 64 group flagged by rule confidence, 128 parent cramersV, 256 parent corr) or NULL;
 * parent_corr / parent_cramers [d] (NaN = none) or NULL. */
int tmog_sanity_decide(int d, const double* variance, const double* corr_label, const double* corr_matrix,
                       int corr_dim, const int32_t* corr_pos, const int32_t* parent_offsets,
                       const int32_t* parent_ids, const int32_t* group_id, const uint8_t* text_shared_hash,
                       const double* cramers_v, const int32_t* rule_count, const double* rule_conf,
                       const double* rule_support, const double* params, uint8_t* drop, uint32_t* reasons,
                       double* parent_corr, double* parent_cramers);

/* ------------------------------------------------------------------------------------------------------------------
 * Host-buffer entry points (what the JNI shim calls per exported Spark partition)
 * ------------------------------------------------------------------------------------------------------------------
 * A tmog_ctx is bound to one device and owns the staging buffers, streams and events that pipeline row chunks
 * H2D -> kernel -> D2H.  One ctx per calling thread (FitStagesUtil runs one task thread per partition,
 * C/utils/stages/FitStagesUtil.scala:96-119); a ctx itself is not thread-safe.  Inputs and outputs should live in
 * pinned memory (tmog_pinned_alloc) so that the copies are asynchronous DMA. */
typedef struct tmog_ctx tmog_ctx;
int tmog_ctx_create(int device, tmog_ctx** ctx);
int tmog_ctx_destroy(tmog_ctx* ctx);
int tmog_pinned_alloc(size_t bytes, void** ptr);
int tmog_pinned_free(void* ptr);

/* tmog_real_fit_partials_dev over host columns, streamed in chunks of chunk_rows (<= 0: default 1 Mi rows).
 *   keep_resident != 0 keeps the uploaded block on the device so that the following tmog_real_transform_host call on
 *                      the same `values` pointer needs no second upload (SequenceEstimator.fit is always followed by
 *                      the model's transform on the same data, C/utils/stages/FitStagesUtil.scala:251-290);
 *   shifts_in          NULL: pilot shifts are taken from the first chunk; else the shifts another partition / rank
 *                      already chose (partials only merge when they share shifts);
 *   shifts_out         [ncols+1] the shifts used;  partials_out  [(ncols+1) * TMOG_REAL_NSTAT];
 *   max_classes        0: skip; else (<= 255) also run tmog_label_classes_dev on the label: classes_out [max_classes],
 *                      *nclasses_out (HOST int; -1 = continuous label), class_counts_out [max_classes] (may be NULL)
 *                      and, when the label is categorical, null_class_counts_out [ncols * nclasses] (may be NULL) as
 *                      tmog_null_class_counts_dev defines it. */
int tmog_real_fit_host(tmog_ctx* ctx, const double* values, int64_t ld, const uint64_t* valid, int64_t ldv,
                       const double* label, int64_t n, int ncols, int64_t chunk_rows, int keep_resident,
                       const double* shifts_in, double* shifts_out, double* partials_out, int max_classes,
                       double* classes_out, int* nclasses_out, uint64_t* class_counts_out,
                       uint64_t* null_class_counts_out);

/* tmog_real_transform_dev over host columns into a host row-major block (same out_ld / out_col0 meaning).
 * keep != NULL (HOST array of nkeep ascending emitted-column indices) writes only the columns that SanityChecker
 * kept, packed (tmog_real_transform_select_dev). */
int tmog_real_transform_host(tmog_ctx* ctx, const double* values, int64_t ld, const uint64_t* valid, int64_t ldv,
                             int64_t n, int ncols, const double* fills, int track_nulls, int64_t chunk_rows,
                             const int32_t* keep, int nkeep, double* out, int64_t out_ld, int64_t out_col0);

/* ---- OPVector -> VectorUDT ----------------------------------------------------------------------------------------
 * The wire format a mapPartitions transform returns to Spark, columnar for the rows of a partition: the reference ends
 * every vectorizer, combiner and SanityCheckerModel with `.compressed` (C/stages/impl/feature/Transmogrifier.scala:466-475,
 * F/utils/spark/RichVector.scala:120-134, C/stages/impl/preparators/DerivedFeatureFilterUtils.scala:291-305): explicit
 * zeros are dropped; a row is a SparseVector when 1.5 (nnz + 1) < size, else a DenseVector.  VectorUDT's struct is
 * (type: byte 0 sparse / 1 dense, size: int, null for dense; indices: int[], null for dense; values: double[]).
 * parts: device array of nparts descriptors, 9 x int64 each, in vector order: {kind, width, col0, ld, k, a, b, c, d}
 *   kind 0 dense   a = f64 [n][ld]
 *   kind 1 one-hot a = i32 [n][ld], k entries per row = index local to the part of a 1.0, -1 = none
 *   kind 2 sparse  a = row_start i64 [n], b = row_nnz i32 [n], c = indices i32 (local to the part), d = values i32
 * _count: type [n], size_out [n] (-1 = null), idx_len / val_len [n] = lengths of the row's two lists; the caller turns
 * the lengths into offsets (exclusive prefix sums) and calls _write, which fills indices / values (ascending indices;
 * dense rows: `size` values). */
int tmog_vector_export_count_dev(const void* parts, int nparts, int64_t n, int size, uint8_t* type, int32_t* size_out,
                                 int64_t* idx_len, int64_t* val_len, void* stream);
int tmog_vector_export_write_dev(const void* parts, int nparts, int64_t n, int size, const uint8_t* type,
                                 const int64_t* idx_off, const int64_t* val_off, int32_t* indices, double* values,
                                 void* stream);

/* Compact return of a transformed partition over PCIe (what the JNI shim copies back instead of full-width arrays; it
 * widens them while it builds the Spark vectors, which it has to touch element by element anyway).
 *   tmog_dense_split_dev: a dense f64 block [n][ld] -> values f64 [n][ldv] of the columns with dest[c] >= 0 (column
 *     dest[c] of the output) and units u8 [n][ldu] of the columns with dest[c] < 0 (column ~dest[c]): the columns that
 *     are 0.0 / 1.0 by construction -- null indicators (OpVectorColumnMetadata.isNullIndicator) and the values of
 *     Binary features (BinaryVectorizer.scala).  flags[0] is set if such a column holds anything else.
 *   tmog_values_narrow_dev: the int32 term frequencies of sparse rows (tmog_text_hash_transform_dev's out_val) as u8; a
 *     value above 254 is written as 255 and listed in over_pos / over_val (over_count counts them, also beyond
 *     over_cap).
 *   tmog_indices_narrow_dev: the int32 column indices of sparse rows whose block has at most 65 536 columns as u16;
 *     flags[1] is set if an index does not fit.
 *   used (device u64, or NULL): the number of entries actually written (the *cursor of the transform call); entries
 *     behind it are not looked at. */
int tmog_dense_split_dev(const double* data, int64_t ld, int64_t n, int ncols, const int32_t* dest, double* values,
                         int64_t ldv, uint8_t* units, int64_t ldu, int32_t* flags, void* stream);
int tmog_indices_narrow_dev(const int32_t* indices, int64_t n, const uint64_t* used, uint16_t* out, int32_t* flags,
                            void* stream);
int tmog_values_narrow_dev(const int32_t* values, int64_t n, const uint64_t* used, uint8_t* out, int64_t* over_pos,
                           int32_t* over_val, int over_cap, uint32_t* over_count, void* stream);

/* ---- string paths over host buffers ------------------------------------------------------------------------------
 * One call per exported partition does what the reference's estimator does in fit(Dataset) and its model in
 * transform(Dataset) (features/src/main/scala/com/salesforce/op/stages/base/sequence/SequenceEstimator.scala:91-106):
 * rows are staged through the device in chunks (the copy of chunk k + 1 under the kernel of chunk k) and the driver-side
 * finishing runs on host threads inside the library -- no allocation, table building or threading is left to the caller.
 *
 * tmog_pivot_fit_host replaces OpTextPivotVectorizer.fitFn (C/stages/impl/feature/OpOneHotVectorizer.scala:75-124,
 * 219-225) fused with SanityChecker.categoricalTests' per-label sums (C/stages/impl/preparators/SanityChecker.scala:
 * 265-272):
 *   codes [ncols][ld] int32 (HOST, pinned for full speed; < 0 = null), per column its dictionary in Arrow utf8 layout
 *   (dict_data[j], dict_offsets[j] int64 [ncodes[j] + 1]); label f64 [n] or NULL (then one class);
 *   max_classes <= 255: labels with more distinct values are refused here (use the *_dev entry points).
 * Outputs: *nclasses_out, classes_out [max_classes], class_counts_out [max_classes] (may be NULL);
 *   per column j: n_top[j], the vocabulary in top_data + j * (top_capacity / ncols) with offsets
 *   top_offsets + j * (top_k + 1) (count desc, then String.compareTo), code_to_out (concatenated, ncodes[j] entries per
 *   column: top position or n_top[j] = OTHER) and out_class_counts + j * (top_k + 2) * nclasses (rows: top values,
 *   OTHER, null) -- everything tmog_pivot_column_finish yields, for all columns, cleaned and ranked on host threads
 *   while the count kernel runs. */
int tmog_pivot_fit_host(tmog_ctx* ctx, const int32_t* codes, int64_t ld, int64_t n, int ncols,
                        const uint8_t* const* dict_data, const int64_t* const* dict_offsets, const int64_t* ncodes,
                        const double* label, int64_t chunk_rows, int clean_text, int top_k, int64_t min_support,
                        int max_classes, int* nclasses_out, double* classes_out, uint64_t* class_counts_out, int* n_top,
                        uint8_t* top_data, int64_t top_capacity, int64_t* top_offsets, int32_t* code_to_out,
                        int64_t* out_class_counts);

/* Replaces OneHotModelFun.pivotFn + reindex (OpOneHotVectorizer.scala:416-437, Transmogrifier.scala:437-449) over host
 * codes: out [n][out_ld] int32 (HOST), columns [out_col0, out_col0 + ncols) = index of each column's 1.0 inside the
 * pivot block (column blocks of n_top[j] + 1 (+ 1 with track_nulls) in order), mapped through newpos [block width]
 * when given (SanityCheckerModel's removeFeatures: -1 = dropped).  The position tables are built here from code_to_out. */
int tmog_pivot_transform_host(tmog_ctx* ctx, const int32_t* codes, int64_t ld, int64_t n, int ncols, const int64_t* ncodes,
                              const int32_t* code_to_out, const int* n_top, int track_nulls, const int32_t* newpos,
                              int64_t chunk_rows, int32_t* out, int64_t out_ld, int64_t out_col0);

/* tmog_text_hash_stats_dev over host text columns (Arrow utf8 per column: data[j], offsets[j] int64 [n + 1] starting at
 * 0, valid[j] bitmap or NULL): the SanityChecker sums of the hashed block and the null indicators of one partition
 * (C/stages/impl/feature/SmartTextVectorizer.scala:333-371, OPCollectionHashingVectorizer.scala:244-272), returned in
 * plain form per vector index (w = null_base + ncols entries): bucket_class_sum [w * nclasses] = sum of term
 * frequencies per label class, bucket_rows [w] = rows holding the index, bucket_sq_sum [w] = sum of squared values,
 * bucket_max / bucket_min [w] = largest / smallest value among the rows holding it.  params as in the _dev call. */
int tmog_text_hash_stats_host(tmog_ctx* ctx, int ncols, int64_t n, const uint8_t* const* data,
                              const int64_t* const* offsets, const uint64_t* const* valid, const uint8_t* prefix,
                              const int32_t* prefix_len, const int32_t* params, const double* label, int max_classes,
                              int* nclasses_out, double* classes_out, uint64_t* class_counts_out,
                              int64_t* bucket_class_sum, int64_t* bucket_rows, int64_t* bucket_sq_sum,
                              int32_t* bucket_max, int32_t* bucket_min);

/* tmog_text_hash_transform_dev over host text columns: the sparse rows of the partition in ROW ORDER as CSR (indptr
 * [n + 1], out_idx / out_val [capacity]; *nnz_out = entries) -- what OpTransformerN.transform's mapPartitions hands
 * back as one sparse vector per row.  capacity 0 sizes the output (nothing written). */
int tmog_text_hash_transform_host(tmog_ctx* ctx, int ncols, int64_t n, const uint8_t* const* data,
                                  const int64_t* const* offsets, const uint64_t* const* valid, const uint8_t* prefix,
                                  const int32_t* prefix_len, const int32_t* params, const int32_t* newpos,
                                  int64_t* indptr, int32_t* out_idx, int32_t* out_val, int64_t capacity,
                                  int64_t* nnz_out);

/* ---- RawFeatureFilter ---------------------------------------------------------------------------
 * Replaces the row maps and reduces of RawFeatureFilter.computeFeatureStats (C/filters/RawFeatureFilter.scala:138-200)
 * for numeric (Real, RealNN, Integral, Binary), Text / TextArea and PickList features, one HBM-bound scan each.
 * Layouts: values [ncols][ld] (dtype 0 = f64, 1 = int64, 2 = uint8), valid = Arrow bitmap words [ncols][ldv] or NULL.
 * Every output ACCUMULATES (row chunks and GPUs add); the caller initialises it.
 *
 * _numeric_summary_dev: Summary of every column (C/filters/Summary.scala:43-65, PreparedFeatures.scala:52-53):
 *   out [ncols][5] = min, max, sum, count, NaN seen (initialise to +Inf, -Inf, 0, 0, 0; java Math.min / max make a
 *   NaN value the min AND the max: the caller applies that).
 * _numeric_hist_dev: FeatureDistribution.histValues for Right(values) (C/filters/FeatureDistribution.scala:328-350) with
 *   NumericBucketizer.bucketize (C/stages/impl/feature/NumericBucketizer.scala:219-264; Inclusion.Left, trackInvalid):
 *   mode [col] 0 = the (training) summary is empty: value_sum [col] += v; 1 = min < max: hist [col][b] += 1 with b the
 *   bucket of v among splits [col][bins] (tmog_rff_numeric_splits), bins - 1 for NaN, +-Inf and values outside;
 *   2 = otherwise: hist [col][0] += (v == same_value [col]), hist [col][1] += (v != same_value [col]).
 *   nonnull [col] += valid rows.  hist u64 [ncols][bins].
 * _rows_hist_dev: FeatureDistribution.histValues for Left(tokens) (:317-321: HashingTF(bins).transform(tokens)) from the
 *   sparse rows tmog_text_hash_transform_dev wrote with separate hash spaces of `bins` buckets per column (index =
 *   column * bins + bucket; larger indices are ignored): hist [column * bins + bucket] += weight * tf; tok [col][3] =
 *   min / max tokens of a non-empty cell, non-empty cells (initialise to INT64_MAX, -1, 0) -- the text Summary
 *   (Summary.scala:61).  valid_ptrs: device array of the columns' bitmap pointers (0 = no empty cell) or NULL;
 *   weight [n] or NULL: rows a row stands for (the rows are a dictionary's entries, weight = tmog_rff_code_count_dev).
 * _code_count_dev: counts [col_off[col] + code] += 1 over dictionary codes [ncols][ld] (< 0 = empty); *bad is set when
 *   a code is >= ncodes [col]; with y (f64 [n]) given, null_y [col] += sum of (y - shift) over the column's empty rows
 *   (what _null_label_sums_dev yields for columns with a bitmap).
 * _null_label_sums_dev: PreparedFeatures.getNullLabelLeakageVector (:62-70) summed: out [c] += sum of (y - shift) over
 *   the rows where column c is empty (c < ncols), out [ncols] += sum (y - shift), out [ncols + 1] += sum (y - shift)^2;
 *   an empty response (y_valid bit 0) counts as 0.0.
 * Host: _numeric_splits (FeatureDistribution.scala:333-334, no fused multiply-add), _js_divergence (:154-163),
 * _null_label_corr (Statistics.corr of [response, null indicators], RawFeatureFilter.scala:184-191: min(|corr|, 1),
 * NaN where a variance is <= 1e-12) from the sums above (n rows, null_count / null_y_sum [ncols]). */
int tmog_rff_numeric_summary_dev(const void* values, int dtype, int64_t ld, const uint64_t* valid, int64_t ldv, int64_t n,
                                 int ncols, double* out, void* stream);
int tmog_rff_numeric_hist_dev(const void* values, int dtype, int64_t ld, const uint64_t* valid, int64_t ldv, int64_t n,
                              int ncols, int bins, const double* splits, const int32_t* mode, const double* same_value,
                              uint64_t* hist, double* value_sum, uint64_t* nonnull, void* stream);
int tmog_rff_rows_hist_dev(const int64_t* row_start, const int32_t* row_nnz, const int32_t* idx, const int32_t* val,
                           int64_t n, int ncols, int bins, const uint64_t* const* valid_ptrs, const uint64_t* weight,
                           uint64_t* hist, int64_t* tok, void* stream);
int tmog_rff_code_count_dev(const int32_t* codes, int64_t ld, int64_t n, int ncols, const int64_t* col_off,
                            const int64_t* ncodes, uint64_t* counts, int32_t* bad, const double* y,
                            const uint64_t* y_valid, double shift, double* null_y, void* stream);
int tmog_rff_null_label_sums_dev(const uint64_t* const* valid_ptrs, int ncols, int64_t n, const double* y,
                                 const uint64_t* y_valid, double shift, double* out, void* stream);
int tmog_rff_numeric_splits(double min, double max, int bins, double* out);
int tmog_rff_js_divergence(const double* a, const double* b, int n, double* out);
int tmog_rff_null_label_corr(double n, double y_sum, double y_sq_sum, int ncols, const double* null_count,
                             const double* null_y_sum, double* out);

/* ---- multi-GPU exchange ---------------------------------------------------------------------------------------------
 * Rows shard across GPUs; only the small results of a fit cross NVLink.  Replaces the reduce side of the reference's
 * aggregations: Aggregator.merge of SequenceAggregators (U/spark/SequenceAggregators.scala:83), the RDD.fold of
 * OpOneHotVectorizer.fitFn (C/stages/impl/feature/OpOneHotVectorizer.scala:95-97), treeAggregate / reduceByKey of
 * SanityChecker (C/stages/impl/preparators/SanityChecker.scala:265-272,407).  One tmog_comm per rank (process or thread
 * with its own GPU): rank 0 makes the 128-byte id (tmog_comm_unique_id) and hands it to the others out of band (the
 * Spark driver broadcasts it), every rank calls tmog_comm_init.  NCCL (libnccl.so.2) is loaded at run time.
 * All-reduce in place of `count` 8-byte elements: dtype 0 = int64 (count tables), 1 = float64 (partial sums, which must
 * share their shifts); op 0 = sum, 1 = max, 2 = min.  _dev: device buffer, asynchronous on `stream`; _host: host
 * buffer, staged through the comm's own device buffer, returns when the result is in place. */
typedef struct tmog_comm tmog_comm;
int tmog_comm_unique_id(uint8_t* id128);
int tmog_comm_init(const uint8_t* id128, int rank, int size, int device, tmog_comm** comm);
int tmog_comm_destroy(tmog_comm* comm);
int tmog_comm_allreduce_dev(tmog_comm* comm, void* buf, int64_t count, int dtype, int op, void* stream);
int tmog_comm_allreduce_host(tmog_comm* comm, void* buf, int64_t count, int dtype, int op);

#ifdef __cplusplus
}
#endif
#endif /* TMOG_B200_H */
