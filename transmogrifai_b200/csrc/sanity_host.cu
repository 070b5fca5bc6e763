// SanityChecker driver-side finishing steps (host code inside the library; tiny inputs):
//   contingency statistics (Cramer's V, mutual information, rule confidence / support) and the removeBadFeatures
//   decision.  Reference file:line for each function is given in include/tmog_b200.h.
#include <math.h>
#include <string.h>

#include <algorithm>
#include <functional>
#include <map>
#include <thread>
#include <vector>

#include "common.cuh"

using namespace tmog;

namespace {

// OpStatistics.filterEmpties: drop all-zero rows, then all-zero columns.  m is rows x cols, row-major.
void filter_empties(const double* m, int rows, int cols, std::vector<double>& out, int& orows, int& ocols) {
    std::vector<int> keep_r, keep_c;
    for (int i = 0; i < rows; ++i) {
        double s = 0.0;
        for (int j = 0; j < cols; ++j) s += m[i * cols + j];
        if (s != 0.0) keep_r.push_back(i);
    }
    for (int j = 0; j < cols; ++j) {
        double s = 0.0;
        for (int i : keep_r) s += m[i * cols + j];
        if (s != 0.0) keep_c.push_back(j);
    }
    orows = (int)keep_r.size();
    ocols = (int)keep_c.size();
    out.assign((size_t)orows * ocols, 0.0);
    for (int a = 0; a < orows; ++a)
        for (int b = 0; b < ocols; ++b) out[(size_t)a * ocols + b] = m[keep_r[a] * cols + keep_c[b]];
}

}  // namespace

extern "C" int tmog_contingency_stats(const double* matrix, int rows, int cols, double* cramers_v, double* chi2,
                                      double* mutual_info, double* pmi, double* max_conf, double* supports) {
    TMOG_REQUIRE(rows >= 0 && cols >= 0 && (matrix != nullptr || rows * cols == 0), "bad arguments");
    const double nan = NAN;
    std::vector<double> f;
    int fr = 0, fc = 0;
    if (rows * cols > 0) filter_empties(matrix, rows, cols, f, fr, fc);
    if (fr * fc == 0) {  // contingencyStats: "filteredMatrix.numActives <= 0" -> all NaN / empty
        if (cramers_v) *cramers_v = nan;
        if (chi2) *chi2 = nan;
        if (mutual_info) *mutual_info = nan;
        if (pmi)
            for (int i = 0; i < rows * cols; ++i) pmi[i] = nan;
        if (max_conf)
            for (int i = 0; i < rows; ++i) max_conf[i] = nan;
        if (supports)
            for (int i = 0; i < rows; ++i) supports[i] = nan;
        return 1;  // "empty": the reference returns empty maps / arrays here
    }
    // chi-squared on the filtered matrix (Spark ChiSqTest.chiSquaredMatrix, no continuity correction)
    double v = nan, stat = nan;
    if (fr > 1 && fc > 1) {
        std::vector<double> rs(fr, 0.0), cs(fc, 0.0);
        double total = 0.0;
        for (int i = 0; i < fr; ++i)
            for (int j = 0; j < fc; ++j) {
                const double x = f[(size_t)i * fc + j];
                rs[i] += x;
                cs[j] += x;
                total += x;
            }
        stat = 0.0;
        for (int j = 0; j < fc; ++j)
            for (int i = 0; i < fr; ++i) {
                const double e = rs[i] * cs[j] / total;
                const double dlt = f[(size_t)i * fc + j] - e;
                stat += dlt * dlt / e;
            }
        const double phi2 = stat / total;
        v = sqrt(phi2 / (double)std::min(fr - 1, fc - 1));
    }
    if (cramers_v) *cramers_v = v;
    if (chi2) *chi2 = stat;
    // mutual information and rule confidences on the UNFILTERED matrix (OpStatistics.scala:314-316)
    std::vector<double> rs(rows, 0.0), cs(cols, 0.0);
    double total = 0.0;
    for (int i = 0; i < rows; ++i)
        for (int j = 0; j < cols; ++j) {
            const double x = matrix[i * cols + j];
            rs[i] += x;
            cs[j] += x;
            total += x;
        }
    double mi = 0.0;
    for (int j = 0; j < cols; ++j)
        for (int i = 0; i < rows; ++i) {
            const double x = matrix[i * cols + j];
            double p = 0.0;
            if (!(x == 0.0 || rs[i] == 0.0 || cs[j] == 0.0))
                p = log(std::max(x, 1e-99) * total / (rs[i] * cs[j])) / log(2.0);
            if (pmi) pmi[i * cols + j] = p;
            mi += p * x / total;
        }
    if (mutual_info) *mutual_info = mi;
    for (int i = 0; i < rows; ++i) {
        if (supports) supports[i] = rs[i] / total;
        if (max_conf) {
            double mx = matrix[i * cols];
            for (int j = 1; j < cols; ++j) mx = std::max(mx, matrix[i * cols + j]);
            max_conf[i] = rs[i] == 0.0 ? 0.0 : mx / rs[i];
        }
    }
    return TMOG_OK;
}

extern "C" int tmog_sanity_decide(int d, const double* variance, const double* corr_label, const double* corr_matrix,
                                  int corr_dim, const int32_t* corr_pos, const int32_t* parent_offsets,
                                  const int32_t* parent_ids, const int32_t* group_id, const uint8_t* text_shared_hash,
                                  const double* cramers_v, const int32_t* rule_count, const double* rule_conf,
                                  const double* rule_support, const double* params, uint8_t* drop,
                                  uint32_t* reasons, double* parent_corr_out, double* parent_cramers_out) {
    TMOG_REQUIRE(d >= 0 && variance && corr_pos && parent_offsets && parent_ids && group_id && text_shared_hash &&
                     cramers_v && rule_count && rule_conf && rule_support && params && drop,
                 "NULL argument");
    TMOG_REQUIRE(corr_matrix == nullptr || corr_dim > 0, "corr_dim must be positive with a correlation matrix");
    const double min_variance = params[0], min_corr = params[1], max_corr = params[2], max_feature_corr = params[3],
                 max_cramers = params[4], max_rule_conf = params[5], min_rule_support = params[6];
    const bool remove_feature_group = params[7] != 0.0, protect_text_shared_hash = params[8] != 0.0;

    // makeColumnStatistics: max |corr| and max Cramer's V per parent feature, NaNs skipped, folded from 0.0.  Parent and
    // group ids are small dense integers, so the two maps are arrays (NaN = no entry).
    int max_parent = -1, max_group = -1;
    {
        int32_t lo = 0, hi = -1;                       // plain min / max loop (vectorises; a hashed block brings millions)
        const int64_t np_total = parent_offsets[d];
        for (int64_t k = 0; k < np_total; ++k) {
            lo = std::min(lo, parent_ids[k]);
            hi = std::max(hi, parent_ids[k]);
        }
        TMOG_REQUIRE(lo >= 0, "negative parent id");
        max_parent = hi;
    }
    for (int i = 0; i < d; ++i) max_group = std::max(max_group, group_id[i]);
    const double none = nan("");
    std::vector<double> corr_parent((size_t)max_parent + 1, none), cramers_parent((size_t)max_parent + 1, none);
    // Columns of a hashed block all carry the same parent list (a shared hash space: every text feature), so the walk
    // over the parents is done once per run of identical lists, not once per column.
    auto same_parents = [&](int a, int b) {
        const int la = parent_offsets[a + 1] - parent_offsets[a], lb = parent_offsets[b + 1] - parent_offsets[b];
        return la == lb && (la == 0 || memcmp(parent_ids + parent_offsets[a], parent_ids + parent_offsets[b],
                                              sizeof(int32_t) * (size_t)la) == 0);
    };
    std::vector<int32_t> run_of((size_t)d, 0);      // first column of the run a column belongs to
    // the pass over the columns of [lo, hi), with its own maxima per parent (a run that crosses a range boundary is
    // simply treated as two runs)
    auto scan_range = [&](int lo, int hi, std::vector<double>& cp, std::vector<double>& vp) {
        for (int i0 = lo; i0 < hi;) {
            int i1 = i0 + 1;
            while (i1 < hi && same_parents(i0, i1)) ++i1;
            bool any = false;
            double run_max = 0.0;
            for (int i = i0; i < i1; ++i) {
                run_of[i] = i0;
                if (corr_pos[i] >= 0 && corr_label) {
                    const double c = corr_label[corr_pos[i]];
                    const double ac = isnan(c) ? 0.0 : fabs(c);
                    run_max = any ? std::max(run_max, ac) : ac;
                    any = true;
                }
                // a column is in cramersVMap iff it belongs to a categorical group (rule_count >= 0); its V may be
                // NaN.  parentMap (.toMap) keeps the LAST parent of a multi-parent column.
                if (rule_count[i] >= 0 && parent_offsets[i + 1] > parent_offsets[i]) {
                    const int last = parent_ids[parent_offsets[i + 1] - 1];
                    const double av = isnan(cramers_v[i]) ? 0.0 : fabs(cramers_v[i]);
                    double& slot = vp[last];
                    slot = isnan(slot) ? av : std::max(slot, av);
                }
            }
            if (any)
                for (int k = parent_offsets[i0]; k < parent_offsets[i0 + 1]; ++k) {
                    double& slot = cp[parent_ids[k]];
                    slot = isnan(slot) ? run_max : std::max(slot, run_max);
                }
            i0 = i1;
        }
    };
    const int nthreads = d >= (1 << 16) ? (int)std::min<unsigned>(8u, std::max(1u, std::thread::hardware_concurrency())) : 1;
    const int chunk = (d + nthreads - 1) / std::max(nthreads, 1);
    if (nthreads == 1) {
        scan_range(0, d, corr_parent, cramers_parent);
    } else {
        std::vector<std::vector<double>> cps((size_t)nthreads, corr_parent), vps((size_t)nthreads, cramers_parent);
        std::vector<std::thread> pool;
        for (int t = 0; t < nthreads; ++t) {
            const int lo = t * chunk, hi = std::min(d, lo + chunk);
            if (lo < hi) pool.emplace_back(scan_range, lo, hi, std::ref(cps[t]), std::ref(vps[t]));
        }
        for (auto& th : pool) th.join();
        for (int t = 0; t < nthreads; ++t)
            for (size_t q = 0; q < corr_parent.size(); ++q) {
                const double a = cps[t][q], v = vps[t][q];
                if (!isnan(a)) corr_parent[q] = isnan(corr_parent[q]) ? a : std::max(corr_parent[q], a);
                if (!isnan(v)) cramers_parent[q] = isnan(cramers_parent[q]) ? v : std::max(cramers_parent[q], v);
            }
    }
    // getFeaturesToDrop: groups flagged by the association-rule check
    std::vector<char> group_flagged((size_t)max_group + 1, 0);
    for (int i = 0; i < d; ++i) {
        if (group_id[i] < 0) continue;
        for (int k = 0; k < std::max(rule_count[i], 0); ++k)
            if (rule_conf[2 * i + k] > max_rule_conf && rule_support[2 * i + k] > min_rule_support) {
                group_flagged[group_id[i]] = 1;
                break;
            }
    }
    // the per-column rules are independent of each other: wide vectors (a hashed block has hundreds of thousands of
    // columns) are split over a few threads
    auto decide_range = [&](int lo, int hi) {
    double run_pc = none, run_pv = none;
    int cur_head = -1;
    for (int i = lo; i < hi; ++i) {
        uint32_t r = 0;
        if (variance[i] <= min_variance) r |= 1u;
        const bool has_corr = corr_pos[i] >= 0 && corr_label != nullptr;
        const double c = has_corr ? corr_label[corr_pos[i]] : none;      // (nan("") is a library call: hoisted)
        if (has_corr && fabs(c) < min_corr) r |= 2u;
        if (has_corr && fabs(c) > max_corr) r |= 4u;
        if (corr_matrix && corr_pos[i] >= 0) {
            // featureCorrs = column corr_pos[i] of the matrix without the label row; .take(index) -- index i itself
            const int lim = std::min(i, corr_dim - 1);
            for (int k = 0; k < lim; ++k)
                if (fabs(corr_matrix[(size_t)k * corr_dim + corr_pos[i]]) > max_feature_corr) {
                    r |= 8u;
                    break;
                }
        }
        if (rule_count[i] >= 0 && cramers_v[i] > max_cramers) r |= 16u;
        for (int k = 0; k < std::max(rule_count[i], 0); ++k)
            if (rule_conf[2 * i + k] > max_rule_conf && rule_support[2 * i + k] > min_rule_support) {
                r |= 32u;
                break;
            }
        if (group_id[i] >= 0 && group_flagged[group_id[i]]) r |= 64u;
        // parent-level values: max over the column's parents that have an entry (once per run of equal parent lists)
        if (run_of[i] != cur_head) {
            cur_head = run_of[i];
            run_pc = run_pv = none;
            for (int k = parent_offsets[cur_head]; k < parent_offsets[cur_head + 1]; ++k) {
                const double a = corr_parent[parent_ids[k]], b = cramers_parent[parent_ids[k]];
                if (!isnan(a)) run_pc = isnan(run_pc) ? a : std::max(run_pc, a);
                if (!isnan(b)) run_pv = isnan(run_pv) ? b : std::max(run_pv, b);
            }
        }
        const double pc = run_pc, pv = run_pv;
        if (parent_corr_out) parent_corr_out[i] = pc;
        if (parent_cramers_out) parent_cramers_out[i] = pv;
        if (remove_feature_group && (!text_shared_hash[i] || !protect_text_shared_hash)) {
            if (!isnan(pv) && pv > max_cramers) r |= 128u;
            if (!isnan(pc) && pc > max_corr) r |= 256u;
        }
        drop[i] = r != 0;
        if (reasons) reasons[i] = r;
    }
    };
    if (nthreads == 1) {
        decide_range(0, d);
    } else {
        std::vector<std::thread> pool;
        for (int t = 0; t < nthreads; ++t) {
            const int lo = t * chunk, hi = std::min(d, lo + chunk);
            if (lo < hi) pool.emplace_back(decide_range, lo, hi);
        }
        for (auto& th : pool) th.join();
    }
    return TMOG_OK;
}

// Column statistics of term-frequency columns from the per-bucket sums the text kernels take:
// sum v = sum_c S_c, sum v^2 = sum v + extra, cov(v, y) = sum_c (y_c - mean_y) S_c; min / max from the rare tf > 1 words.
extern "C" int tmog_bucket_stats(int64_t w, int nclasses, const double* class_sum, const double* nnz,
                                 const double* sq_extra, const int64_t* aux, const double* classes, double n,
                                 double mean_y, double* mean, double* variance, double* mn, double* mx, double* cov) {
    TMOG_REQUIRE(w >= 0 && nclasses >= 1 && class_sum && nnz && sq_extra && aux && classes && mean && variance && mn &&
                     mx && cov,
                 "NULL argument");
    for (int64_t i = 0; i < w; ++i) {
        const double* cs = class_sum + i * nclasses;
        double sv = 0.0, cv = 0.0;
        for (int c = 0; c < nclasses; ++c) {
            sv += cs[c];
            cv += cs[c] * (classes[c] - mean_y);
        }
        const double sv2 = sv + sq_extra[i];
        mean[i] = sv / n;
        const double var = n > 1 ? (sv2 - sv * sv / n) / (n - 1) : 0.0;
        variance[i] = var > 0.0 ? var : 0.0;
        const int64_t a0 = aux[3 * i], a1 = aux[3 * i + 1], a2 = aux[3 * i + 2];
        mx[i] = nnz[i] > 0 ? (double)(a0 > 1 ? a0 : 1) : 0.0;
        const int64_t min_gt1 = (~a1) & 0xFFFFFFFFll;
        mn[i] = nnz[i] >= n ? ((double)a2 >= nnz[i] ? (double)min_gt1 : 1.0) : 0.0;
        cov[i] = cv;
    }
    return TMOG_OK;
}

extern "C" int tmog_pearson_from_cov(int d, double n, const double* cov_label, const double* variance,
                                     double label_variance, double* corr) {
    TMOG_REQUIRE(d >= 0 && cov_label && variance && corr, "NULL argument");
    const double sy = sqrt(label_variance);
    for (int i = 0; i < d; ++i) corr[i] = cov_label[i] / (sqrt(variance[i]) * sy * (n - 1.0));  // 0/0 -> NaN kept
    return TMOG_OK;
}

// Many contingency matrices in one call (one per categorical feature group): matrix g is rows[g] x cols[g] row-major at
// data + offsets[g]; pmi has the layout of data; max_conf / supports are packed by row (row_offsets[g] = sum of rows
// before g).  status[g] = the return value tmog_contingency_stats would give (0, or 1 for "empty after filtering").
extern "C" int tmog_contingency_stats_batch(int ngroups, const int32_t* rows, const int32_t* cols,
                                            const int64_t* offsets, const int64_t* row_offsets, const double* data,
                                            double* cramers_v, double* chi2, double* mutual_info, double* pmi,
                                            double* max_conf, double* supports, int32_t* status) {
    TMOG_REQUIRE(ngroups >= 0 && (ngroups == 0 || (rows && cols && offsets && row_offsets && data)), "NULL argument");
    for (int g = 0; g < ngroups; ++g) {
        const int rc = tmog_contingency_stats(data + offsets[g], rows[g], cols[g], cramers_v ? cramers_v + g : nullptr,
                                              chi2 ? chi2 + g : nullptr, mutual_info ? mutual_info + g : nullptr,
                                              pmi ? pmi + offsets[g] : nullptr,
                                              max_conf ? max_conf + row_offsets[g] : nullptr,
                                              supports ? supports + row_offsets[g] : nullptr);
        if (rc < 0) return rc;
        if (status) status[g] = rc;
    }
    return TMOG_OK;
}
