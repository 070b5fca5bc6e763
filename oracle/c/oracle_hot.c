/*
 * oracle_hot.c -- CPU restatement (plain C + OpenMP) of the reference's hot loops.  TEST INFRASTRUCTURE ONLY:
 * used by tests/ as the checker at sizes the Python oracle cannot reach, and by bench.py as the timed CPU baseline
 * ("port": the reference itself is JVM code and cannot run here).  The product (libtmog_b200.so) never links it.
 *
 * It keeps the reference's pass structure -- one pass per Spark job -- and its per-row arithmetic; rows are split
 * into contiguous "partitions" (one per OpenMP thread, like Spark local[N]) whose partial results are merged in
 * partition order.
 *
 * Follows (paths under /root/reference):
 *   utils/src/main/scala/com/salesforce/op/utils/spark/SequenceAggregators.scala:76-88      orc_real_fit
 *   core/src/main/scala/com/salesforce/op/stages/impl/feature/RealVectorizer.scala:112-123  orc_real_transform
 *   core/.../impl/preparators/SanityChecker.scala:397-407 + Spark 2.4.5
 *       mllib.stat.MultivariateOnlineSummarizer (add / merge / mean / variance / min / max) orc_col_stats
 *   utils/src/main/scala/com/salesforce/op/utils/stats/OpStatistics.scala:71-97             orc_label_cov
 *   core/.../impl/preparators/SanityChecker.scala:265-272                                   orc_label_sums
 */
#include <math.h>
#include <omp.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

static inline int bit(const uint64_t* w, int64_t r) { return w == NULL ? 1 : (int)((w[r >> 6] >> (r & 63)) & 1u); }

static void split(int64_t n, int parts, int p, int64_t* lo, int64_t* hi) {
    *lo = n * p / parts;
    *hi = n * (p + 1) / parts;
}

int orc_max_threads(void) { return omp_get_max_threads(); }

/* MeanSeqNullNum: (sum, Int count) per column over the valid values, row order inside a partition; merged in
 * partition order; finish = s / c if c > 0 else s. */
void orc_real_fit(const double* values, int64_t ld, const uint64_t* valid, int64_t ldv, int64_t n, int ncols,
                  int threads, double* fills) {
    if (threads < 1) threads = 1;
    double* sums = calloc((size_t)threads * ncols, sizeof(double));
    int64_t* cnts = calloc((size_t)threads * ncols, sizeof(int64_t));
#pragma omp parallel for num_threads(threads) schedule(static, 1)
    for (int p = 0; p < threads; ++p) {
        int64_t lo, hi;
        split(n, threads, p, &lo, &hi);
        double* s = sums + (size_t)p * ncols;
        int64_t* c = cnts + (size_t)p * ncols;
        /* the reference walks rows; columns are walked per row block here to stay cache friendly */
        for (int64_t r0 = lo; r0 < hi; r0 += 1024) {
            const int64_t r1 = r0 + 1024 < hi ? r0 + 1024 : hi;
            for (int j = 0; j < ncols; ++j) {
                const double* col = values + (int64_t)j * ld;
                const uint64_t* vw = valid ? valid + (int64_t)j * ldv : NULL;
                double sj = s[j];
                int64_t cj = c[j];
                for (int64_t r = r0; r < r1; ++r)
                    if (bit(vw, r)) {
                        sj += col[r];
                        ++cj;
                    }
                s[j] = sj;
                c[j] = cj;
            }
        }
    }
    for (int j = 0; j < ncols; ++j) {
        double s = 0.0;
        int64_t c = 0;
        for (int p = 0; p < threads; ++p) {
            s += sums[(size_t)p * ncols + j];
            c += cnts[(size_t)p * ncols + j];
        }
        fills[j] = c > 0 ? s / (double)c : s;
    }
    free(sums);
    free(cnts);
}

/* RealVectorizerModel.transformFn: dense row [x or fill, isNull] interleaved (track_nulls) */
void orc_real_transform(const double* values, int64_t ld, const uint64_t* valid, int64_t ldv, int64_t n, int ncols,
                        const double* fills, int track_nulls, int threads, double* out, int64_t out_ld) {
    if (threads < 1) threads = 1;
#pragma omp parallel for num_threads(threads) schedule(static)
    for (int64_t r0 = 0; r0 < n; r0 += 256) {
        const int64_t r1 = r0 + 256 < n ? r0 + 256 : n;
        for (int j = 0; j < ncols; ++j) {
            const double* col = values + (int64_t)j * ld;
            const uint64_t* vw = valid ? valid + (int64_t)j * ldv : NULL;
            for (int64_t r = r0; r < r1; ++r) {
                const int v = bit(vw, r);
                if (track_nulls) {
                    out[r * out_ld + 2 * j] = v ? col[r] : fills[j];
                    out[r * out_ld + 2 * j + 1] = v ? 0.0 : 1.0;
                } else {
                    out[r * out_ld + j] = v ? col[r] : fills[j];
                }
            }
        }
    }
}

/* MultivariateOnlineSummarizer over dense rows x[n][ld] (d columns used): Welford over the NON-ZERO entries, Chan
 * merge across partitions, then the zero-aware mean / unbiased variance / min / max. */
typedef struct {
    double *mean, *m2n, *nnz, *mx, *mn;
    double cnt;
} summ_t;

void orc_col_stats(const double* x, int64_t ld, int64_t n, int d, int threads, double* mean, double* variance,
                   double* mn, double* mx) {
    if (threads < 1) threads = 1;
    double* buf = calloc((size_t)threads * d * 5, sizeof(double));
    double* cnts = calloc((size_t)threads, sizeof(double));
#pragma omp parallel for num_threads(threads) schedule(static, 1)
    for (int p = 0; p < threads; ++p) {
        int64_t lo, hi;
        split(n, threads, p, &lo, &hi);
        double* cm = buf + (size_t)p * d * 5;
        double *m2 = cm + d, *nz = cm + 2 * d, *cx = cm + 3 * d, *cn = cm + 4 * d;
        for (int i = 0; i < d; ++i) {
            cx[i] = -INFINITY;  /* Double.MinValue in Spark; only compared */
            cn[i] = INFINITY;
        }
        for (int64_t r = lo; r < hi; ++r) {
            const double* row = x + r * ld;
            for (int i = 0; i < d; ++i) {
                const double v = row[i];
                if (v != 0.0) {
                    if (cx[i] < v) cx[i] = v;
                    if (cn[i] > v) cn[i] = v;
                    const double prev = cm[i];
                    const double diff = v - prev;
                    cm[i] = prev + diff / (nz[i] + 1.0);
                    m2[i] += (v - cm[i]) * diff;
                    nz[i] += 1.0;
                }
            }
        }
        cnts[p] = (double)(hi - lo);
    }
    /* merge in partition order */
    double* cm = buf;
    double *m2 = cm + d, *nz = cm + 2 * d, *cx = cm + 3 * d, *cn = cm + 4 * d;
    double total = cnts[0];
    for (int p = 1; p < threads; ++p) {
        const double* om = buf + (size_t)p * d * 5;
        const double *o2 = om + d, *oz = om + 2 * d, *ox = om + 3 * d, *on = om + 4 * d;
        for (int i = 0; i < d; ++i) {
            const double tn = nz[i] + oz[i];
            if (tn != 0.0) {
                const double delta = om[i] - cm[i];
                cm[i] += delta * oz[i] / tn;
                m2[i] += o2[i] + delta * delta * nz[i] * oz[i] / tn;
            }
            nz[i] = tn;
            if (ox[i] > cx[i]) cx[i] = ox[i];
            if (on[i] < cn[i]) cn[i] = on[i];
        }
        total += cnts[p];
    }
    for (int i = 0; i < d; ++i) {
        mean[i] = total > 0 ? cm[i] * (nz[i] / total) : 0.0;
        const double denom = total - 1.0; /* unit weights: totalWeightSum - weightSquareSum / totalWeightSum */
        double var = 0.0;
        if (denom > 0.0) {
            var = (m2[i] + cm[i] * cm[i] * nz[i] * (total - nz[i]) / total) / denom;
            if (var < 0.0) var = 0.0;
        }
        variance[i] = var;
        double a = cx[i], b = cn[i];
        if (nz[i] < total && a < 0.0) a = 0.0;
        if (nz[i] < total && b > 0.0) b = 0.0;
        if (nz[i] == 0.0 && total == 0.0) a = b = 0.0;
        mx[i] = a;
        mn[i] = b;
    }
    free(buf);
    free(cnts);
}

/* computeCorrelationsWithLabel: cov_i = sum_rows (x_i - mean_i) (y - mean_y), dense loop over every column of
 * every row; label = column d-1 of x. */
void orc_label_cov(const double* x, int64_t ld, int64_t n, int d, const double* mean, int threads, double* cov) {
    if (threads < 1) threads = 1;
    double* buf = calloc((size_t)threads * d, sizeof(double));
#pragma omp parallel for num_threads(threads) schedule(static, 1)
    for (int p = 0; p < threads; ++p) {
        int64_t lo, hi;
        split(n, threads, p, &lo, &hi);
        double* a = buf + (size_t)p * d;
        for (int64_t r = lo; r < hi; ++r) {
            const double* row = x + r * ld;
            const double ey = row[d - 1] - mean[d - 1];
            for (int i = 0; i < d; ++i) a[i] += (row[i] - mean[i]) * ey;
        }
    }
    for (int i = 0; i < d; ++i) {
        double s = 0.0;
        for (int p = 0; p < threads; ++p) s += buf[(size_t)p * d + i];
        cov[i] = s;
    }
    free(buf);
}

/* categoricalTests' reduceByKey: per distinct label value (given sorted in `labels`, nlab of them) the column sums
 * of the feature vector plus the row count in slot d.  y[n] is the label column; x has d feature columns. */
void orc_label_sums(const double* x, int64_t ld, const double* y, int64_t n, int d, const double* labels, int nlab,
                    int threads, double* out /* [nlab][d+1] */) {
    if (threads < 1) threads = 1;
    const size_t w = (size_t)nlab * (d + 1);
    double* buf = calloc((size_t)threads * w, sizeof(double));
#pragma omp parallel for num_threads(threads) schedule(static, 1)
    for (int p = 0; p < threads; ++p) {
        int64_t lo, hi;
        split(n, threads, p, &lo, &hi);
        double* a = buf + (size_t)p * w;
        for (int64_t r = lo; r < hi; ++r) {
            int k = -1;
            for (int q = 0; q < nlab; ++q)
                if (labels[q] == y[r]) {
                    k = q;
                    break;
                }
            if (k < 0) continue;
            double* acc = a + (size_t)k * (d + 1);
            const double* row = x + r * ld;
            for (int i = 0; i < d; ++i) acc[i] += row[i];
            acc[d] += 1.0;
        }
    }
    memset(out, 0, w * sizeof(double));
    for (int p = 0; p < threads; ++p)
        for (size_t i = 0; i < w; ++i) out[i] += buf[(size_t)p * w + i];
    free(buf);
}
