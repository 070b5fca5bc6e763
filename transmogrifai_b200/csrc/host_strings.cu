// Host-buffer entry points of the string paths: one call per exported Spark partition does what the reference's
// estimator does in fit(Dataset) / what its model does in transform(Dataset)
// (features/src/main/scala/com/salesforce/op/stages/base/sequence/SequenceEstimator.scala:91-106), with no Python or
// torch in between: rows are staged through the device in chunks (double-buffered H2D under the kernels), the
// driver-side finishing (dictionary cleaning, top-K, position tables) runs on host threads inside the library.
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <atomic>
#include <new>
#include <thread>
#include <vector>

#include "host_ctx.cuh"

using namespace tmog;

namespace {

int64_t round_up(int64_t v, int64_t m) { return (v + m - 1) / m * m; }

int ensure_pinned(tmog_ctx* c, size_t bytes) {
    if (bytes <= c->pinned_cap) return TMOG_OK;
    if (c->pinned) cudaFreeHost(c->pinned);
    c->pinned = nullptr;
    c->pinned_cap = 0;
    TMOG_CUDA(cudaHostAlloc(&c->pinned, bytes, cudaHostAllocDefault));
    c->pinned_cap = bytes;
    return TMOG_OK;
}

// run f(i) for i in order[0 .. count) on up to `threads` host threads; -> first non-zero status
template <typename F>
int parallel_for(const std::vector<int>& order, int threads, F&& f) {
    std::atomic<int> next(0), status(0);
    auto work = [&]() {
        for (;;) {
            const int k = next.fetch_add(1);
            if (k >= (int)order.size() || status.load() != 0) return;
            const int rc = f(order[k]);
            if (rc != 0) status.store(rc);
        }
    };
    const int t = std::max(1, std::min<int>(threads, (int)order.size()));
    std::vector<std::thread> pool;
    for (int i = 1; i < t; ++i) pool.emplace_back(work);
    work();
    for (auto& th : pool) th.join();
    return status.load();
}

int host_threads() {
    const unsigned hc = std::thread::hardware_concurrency();
    return (int)std::max(1u, std::min(hc ? hc : 4u, 32u));
}

// label -> class ids on the device (ctx buffers); -> number of classes on the host (-1: continuous)
int label_classes(tmog_ctx* c, const double* label, int64_t n, int max_classes, int* h_ncls, double* classes_out,
                  uint64_t* class_counts_out) {
    size_t lbl_scratch = 0;
    if (int rc = tmog_label_classes_scratch_bytes(&lbl_scratch)) return rc;
    if (int rc = c->scratch.ensure(lbl_scratch)) return rc;
    const int64_t nld = round_up(std::max<int64_t>(n, 1), 2048);
    if (int rc = c->r_label.ensure(sizeof(double) * (size_t)nld)) return rc;
    if (int rc = c->class_ids.ensure((size_t)nld)) return rc;
    if (int rc = c->classes.ensure(sizeof(double) * 256 + sizeof(uint64_t) * 256 + sizeof(int) * 2)) return rc;
    double* d_label = static_cast<double*>(c->r_label.p);
    TMOG_CUDA(cudaMemcpyAsync(d_label, label, sizeof(double) * n, cudaMemcpyHostToDevice, c->s_cmp));
    double* d_cls = static_cast<double*>(c->classes.p);
    uint64_t* d_cnt = reinterpret_cast<uint64_t*>(d_cls + 256);
    int* d_ncls = reinterpret_cast<int*>(d_cnt + 256);
    if (int rc = tmog_label_classes_dev(d_label, n, max_classes, d_cls, d_ncls, static_cast<uint8_t*>(c->class_ids.p),
                                        d_cnt, c->scratch.p, c->scratch.cap, c->s_cmp))
        return rc;
    TMOG_CUDA(cudaMemcpyAsync(h_ncls, d_ncls, sizeof(int), cudaMemcpyDeviceToHost, c->s_cmp));
    TMOG_CUDA(cudaStreamSynchronize(c->s_cmp));
    if (*h_ncls > 0) {
        if (classes_out)
            TMOG_CUDA(cudaMemcpyAsync(classes_out, d_cls, sizeof(double) * *h_ncls, cudaMemcpyDeviceToHost, c->s_cmp));
        if (class_counts_out)
            TMOG_CUDA(cudaMemcpyAsync(class_counts_out, d_cnt, sizeof(uint64_t) * *h_ncls, cudaMemcpyDeviceToHost,
                                      c->s_cmp));
        TMOG_CUDA(cudaStreamSynchronize(c->s_cmp));
    }
    return TMOG_OK;
}

}  // namespace

// ---------------------------------------------------------------------------------------------------------------------
// OpTextPivotVectorizer
// ---------------------------------------------------------------------------------------------------------------------
extern "C" int tmog_pivot_fit_host(tmog_ctx* c, const int32_t* codes, int64_t ld, int64_t n, int ncols,
                                   const uint8_t* const* dict_data, const int64_t* const* dict_offsets,
                                   const int64_t* ncodes, const double* label, int64_t chunk_rows, int clean_text,
                                   int top_k, int64_t min_support, int max_classes, int* nclasses_out,
                                   double* classes_out, uint64_t* class_counts_out, int* n_top, uint8_t* top_data,
                                   int64_t top_capacity, int64_t* top_offsets, int32_t* code_to_out,
                                   int64_t* out_class_counts) {
    TMOG_REQUIRE(c != nullptr, "ctx is NULL");
    TMOG_REQUIRE(n >= 0 && ncols >= 0 && ld >= n && ld % 4 == 0, "bad shape n=%lld ncols=%d ld=%lld", (long long)n, ncols,
                 (long long)ld);
    TMOG_REQUIRE(n < 0xffffffffLL, "at most 2^32 - 1 rows per call (uint32 count bins)");
    TMOG_REQUIRE(top_k > 0 && min_support >= 0, "requirement failed: topK given invalid value %d", top_k);
    TMOG_REQUIRE(max_classes >= 0 && max_classes <= 255, "max_classes must be in [0, 255]");
    TMOG_REQUIRE(ncols == 0 || (codes && dict_data && dict_offsets && ncodes), "NULL input");
    TMOG_REQUIRE(n_top && top_data && top_offsets && code_to_out && out_class_counts && nclasses_out, "NULL output");
    TMOG_REQUIRE(label == nullptr || max_classes > 0, "a label needs max_classes > 0");
    TMOG_CUDA(cudaSetDevice(c->device));
    int ncls = 1;
    *nclasses_out = 0;
    if (label != nullptr && n > 0) {
        int h = 0;
        if (int rc = label_classes(c, label, n, max_classes, &h, classes_out, class_counts_out)) return rc;
        *nclasses_out = h;
        TMOG_REQUIRE(h > 0, "the label has more than %d distinct values: fit this column with the device entry points "
                            "(tmog_pivot_label_sum_dev carries a continuous label)", max_classes);
        ncls = h;
    }
    // dictionary cleaning + grouping starts now, largest dictionary first, beside the count kernel
    std::vector<tmog_pivot_dict*> handles(ncols, nullptr);
    std::vector<int> order(ncols);
    for (int j = 0; j < ncols; ++j) order[j] = j;
    std::sort(order.begin(), order.end(), [&](int a, int b) { return ncodes[a] > ncodes[b]; });
    int prep_rc = 0;
    std::thread prep([&]() {
        prep_rc = parallel_for(order, host_threads(), [&](int j) {
            return tmog_pivot_dict_prepare(dict_data[j], dict_offsets[j], ncodes[j], clean_text, &handles[j]);
        });
    });
    auto free_handles = [&]() {
        for (auto h : handles)
            if (h) tmog_pivot_dict_free(h);
    };
    // count tables on the device
    std::vector<int64_t> desc((size_t)ncols * 4), coff(ncols + 1, 0);
    int64_t max_nids = 1;
    for (int j = 0; j < ncols; ++j) {
        desc[4 * j] = ncodes[j] + 1;
        desc[4 * j + 1] = 0;
        desc[4 * j + 2] = coff[j];
        desc[4 * j + 3] = ncodes[j];
        coff[j + 1] = coff[j] + (ncodes[j] + 1) * ncls;
        max_nids = std::max(max_nids, ncodes[j] + 1);
    }
    const size_t total_bins = (size_t)coff[ncols];
    int rc = TMOG_OK;
    auto fail = [&](int code) {
        prep.join();
        free_handles();
        return code;
    };
    if ((rc = c->s_desc.ensure(sizeof(int64_t) * std::max<size_t>(desc.size(), 4)))) return fail(rc);
    if ((rc = c->s_counts.ensure(sizeof(uint32_t) * std::max<size_t>(total_bins, 1) + 16))) return fail(rc);
    if ((rc = ensure_pinned(c, sizeof(uint32_t) * std::max<size_t>(total_bins, 1) + 16))) return fail(rc);
    uint32_t* d_counts = static_cast<uint32_t*>(c->s_counts.p);
    int* d_bad = reinterpret_cast<int*>(d_counts + total_bins);
    if (cudaMemcpyAsync(c->s_desc.p, desc.data(), sizeof(int64_t) * desc.size(), cudaMemcpyHostToDevice, c->s_cmp) !=
            cudaSuccess ||
        cudaMemsetAsync(d_counts, 0, sizeof(uint32_t) * total_bins + 16, c->s_cmp) != cudaSuccess) {
        set_error("CUDA error while setting up the count tables");
        return fail(TMOG_ECUDA);
    }
    // rows in chunks: the copy of chunk k + 1 runs under the count kernel of chunk k
    if (chunk_rows <= 0) chunk_rows = 4 << 20;
    chunk_rows = std::min<int64_t>(round_up(chunk_rows, 2048), round_up(std::max<int64_t>(n, 1), 2048));
    const int64_t nchunks = n > 0 ? (n + chunk_rows - 1) / chunk_rows : 0;
    for (int b = 0; b < 2 && ncols > 0 && n > 0; ++b)
        if ((rc = c->s_in[b].ensure(sizeof(int32_t) * (size_t)ncols * chunk_rows))) return fail(rc);
    for (int64_t k = 0; k < nchunks; ++k) {
        const int b = (int)(k & 1);
        const int64_t r0 = k * chunk_rows, m = std::min(chunk_rows, n - r0);
        if (k >= 2 && cudaStreamWaitEvent(c->s_h2d, c->ev_cmp[b], 0) != cudaSuccess) return fail(TMOG_ECUDA);
        if (cudaMemcpy2DAsync(c->s_in[b].p, sizeof(int32_t) * chunk_rows, codes + r0, sizeof(int32_t) * ld,
                              sizeof(int32_t) * m, ncols, cudaMemcpyHostToDevice, c->s_h2d) != cudaSuccess) {
            set_error("CUDA error in the upload of a code chunk: %s", cudaGetErrorString(cudaGetLastError()));
            return fail(TMOG_ECUDA);
        }
        cudaEventRecord(c->ev_in[b], c->s_h2d);
        cudaStreamWaitEvent(c->s_cmp, c->ev_in[b], 0);
        if ((rc = tmog_pivot_count_dev(static_cast<const int32_t*>(c->s_in[b].p), chunk_rows, m, ncols,
                                       label ? static_cast<const uint8_t*>(c->class_ids.p) + r0 : nullptr, ncls,
                                       static_cast<const int64_t*>(c->s_desc.p), nullptr, max_nids, d_counts, d_bad,
                                       c->s_cmp)))
            return fail(rc);
        cudaEventRecord(c->ev_cmp[b], c->s_cmp);
    }
    uint32_t* h_counts = static_cast<uint32_t*>(c->pinned);
    if (cudaMemcpyAsync(h_counts, d_counts, sizeof(uint32_t) * total_bins + 16, cudaMemcpyDeviceToHost, c->s_cmp) !=
            cudaSuccess ||
        cudaStreamSynchronize(c->s_cmp) != cudaSuccess) {
        set_error("CUDA error in the count pass: %s", cudaGetErrorString(cudaGetLastError()));
        return fail(TMOG_ECUDA);
    }
    prep.join();
    if (prep_rc != 0) {
        free_handles();
        return prep_rc;
    }
    if (reinterpret_cast<int*>(h_counts + total_bins)[0] != 0) {
        free_handles();
        set_error("a dictionary code outside its column's dictionary was found in the partition");
        return TMOG_EINVAL;
    }
    // per column: top-K of the cleaned values, code -> emitted column, per-class counts of the emitted columns
    std::vector<int64_t> soff(ncols + 1, 0);
    for (int j = 0; j < ncols; ++j) soff[j + 1] = soff[j] + ncodes[j];
    const int64_t per_col_bytes = ncols > 0 ? top_capacity / ncols : 0;
    rc = parallel_for(order, host_threads(), [&](int j) {
        std::vector<int64_t> raw((size_t)(ncodes[j] + 1) * ncls);
        for (size_t q = 0; q < raw.size(); ++q) raw[q] = (int64_t)h_counts[coff[j] + q];
        return tmog_pivot_dict_finish(handles[j], raw.data(), ncls, top_k, min_support, &n_top[j], code_to_out + soff[j],
                                      out_class_counts + (size_t)j * (top_k + 2) * ncls, top_data + (size_t)j * per_col_bytes,
                                      per_col_bytes, top_offsets + (size_t)j * (top_k + 1));
    });
    free_handles();
    return rc;
}

extern "C" int tmog_pivot_transform_host(tmog_ctx* c, const int32_t* codes, int64_t ld, int64_t n, int ncols,
                                         const int64_t* ncodes, const int32_t* code_to_out, const int* n_top,
                                         int track_nulls, const int32_t* newpos, int64_t chunk_rows, int32_t* out,
                                         int64_t out_ld, int64_t out_col0) {
    TMOG_REQUIRE(c != nullptr, "ctx is NULL");
    TMOG_REQUIRE(n >= 0 && ncols >= 0 && ld >= n && ld % 4 == 0 && out_ld >= out_col0 + ncols, "bad shape");
    if (n == 0 || ncols == 0) return TMOG_OK;
    TMOG_REQUIRE(codes && ncodes && code_to_out && n_top && out, "NULL argument");
    TMOG_CUDA(cudaSetDevice(c->device));
    // position tables: entry = index of the 1.0 in the emitted block (column block offsets in order), or newpos[.]
    std::vector<int64_t> desc((size_t)ncols * 4);
    std::vector<int32_t> pos;
    int64_t slot = 0, width = 0, src = 0;
    for (int j = 0; j < ncols; ++j) {
        desc[4 * j] = ncodes[j] + 1;
        desc[4 * j + 1] = slot;
        desc[4 * j + 2] = 0;
        desc[4 * j + 3] = ncodes[j];
        for (int64_t q = 0; q < ncodes[j]; ++q) {
            const int64_t col = width + code_to_out[src + q];
            pos.push_back(newpos ? newpos[col] : (int32_t)col);
        }
        const int64_t null_col = width + n_top[j] + 1;
        pos.push_back(track_nulls ? (newpos ? newpos[null_col] : (int32_t)null_col) : -1);
        slot += ncodes[j];
        src += ncodes[j];
        width += n_top[j] + (track_nulls ? 2 : 1);
    }
    if (int rc = c->s_desc.ensure(sizeof(int64_t) * desc.size())) return rc;
    if (int rc = c->s_pos.ensure(sizeof(int32_t) * pos.size())) return rc;
    TMOG_CUDA(cudaMemcpyAsync(c->s_desc.p, desc.data(), sizeof(int64_t) * desc.size(), cudaMemcpyHostToDevice, c->s_cmp));
    TMOG_CUDA(cudaMemcpyAsync(c->s_pos.p, pos.data(), sizeof(int32_t) * pos.size(), cudaMemcpyHostToDevice, c->s_cmp));
    TMOG_CUDA(cudaStreamSynchronize(c->s_cmp));     // desc / pos go out of scope
    if (chunk_rows <= 0) chunk_rows = 4 << 20;
    chunk_rows = std::min<int64_t>(round_up(chunk_rows, 2048), round_up(n, 2048));
    const int64_t nchunks = (n + chunk_rows - 1) / chunk_rows;
    for (int b = 0; b < 2; ++b) {
        if (int rc = c->s_in[b].ensure(sizeof(int32_t) * (size_t)ncols * chunk_rows)) return rc;
        if (int rc = c->s_out[b].ensure(sizeof(int32_t) * (size_t)ncols * chunk_rows)) return rc;
    }
    for (int64_t k = 0; k < nchunks; ++k) {
        const int b = (int)(k & 1);
        const int64_t r0 = k * chunk_rows, m = std::min(chunk_rows, n - r0);
        if (k >= 2) {
            TMOG_CUDA(cudaStreamWaitEvent(c->s_h2d, c->ev_cmp[b], 0));
            TMOG_CUDA(cudaStreamWaitEvent(c->s_cmp, c->ev_out[b], 0));
        }
        TMOG_CUDA(cudaMemcpy2DAsync(c->s_in[b].p, sizeof(int32_t) * chunk_rows, codes + r0, sizeof(int32_t) * ld,
                                    sizeof(int32_t) * m, ncols, cudaMemcpyHostToDevice, c->s_h2d));
        TMOG_CUDA(cudaEventRecord(c->ev_in[b], c->s_h2d));
        TMOG_CUDA(cudaStreamWaitEvent(c->s_cmp, c->ev_in[b], 0));
        if (int rc = tmog_pivot_transform_dev(static_cast<const int32_t*>(c->s_in[b].p), chunk_rows, m, ncols,
                                              static_cast<const int64_t*>(c->s_desc.p),
                                              static_cast<const int32_t*>(c->s_pos.p),
                                              static_cast<int32_t*>(c->s_out[b].p), ncols, 0, c->s_cmp))
            return rc;
        TMOG_CUDA(cudaEventRecord(c->ev_cmp[b], c->s_cmp));
        TMOG_CUDA(cudaStreamWaitEvent(c->s_d2h, c->ev_cmp[b], 0));
        TMOG_CUDA(cudaMemcpy2DAsync(out + r0 * out_ld + out_col0, sizeof(int32_t) * out_ld, c->s_out[b].p,
                                    sizeof(int32_t) * ncols, sizeof(int32_t) * ncols, m, cudaMemcpyDeviceToHost,
                                    c->s_d2h));
        TMOG_CUDA(cudaEventRecord(c->ev_out[b], c->s_d2h));
    }
    TMOG_CUDA(cudaStreamSynchronize(c->s_d2h));
    TMOG_CUDA(cudaStreamSynchronize(c->s_cmp));
    return TMOG_OK;
}

// ---------------------------------------------------------------------------------------------------------------------
// Hashed text columns
// ---------------------------------------------------------------------------------------------------------------------
namespace {

// upload the text columns of a partition (Arrow utf8 per column) and build the device column descriptors
int upload_text(tmog_ctx* c, int ncols, int64_t n, const uint8_t* const* data, const int64_t* const* offsets,
                const uint64_t* const* valid, int64_t** d_desc) {
    size_t total = 0;
    std::vector<size_t> at_data(ncols), at_off(ncols), at_valid(ncols);
    const size_t vwords = (size_t)(n + 63) / 64;
    for (int j = 0; j < ncols; ++j) {
        const size_t bytes = (size_t)offsets[j][n] - (size_t)offsets[j][0];
        at_data[j] = total;
        total += (bytes + 16 + 255) & ~size_t(255);
        at_off[j] = total;
        total += (sizeof(int64_t) * (size_t)(n + 1) + 255) & ~size_t(255);
        at_valid[j] = total;
        total += valid && valid[j] ? ((sizeof(uint64_t) * vwords + 255) & ~size_t(255)) : 0;
    }
    if (int rc = c->s_in[0].ensure(total + 256)) return rc;
    if (int rc = c->s_desc.ensure(sizeof(int64_t) * 4 * (size_t)std::max(ncols, 1))) return rc;
    uint8_t* base = static_cast<uint8_t*>(c->s_in[0].p);
    std::vector<int64_t> desc((size_t)ncols * 4, 0);
    for (int j = 0; j < ncols; ++j) {
        const int64_t o0 = offsets[j][0];
        const size_t bytes = (size_t)offsets[j][n] - (size_t)o0;
        TMOG_REQUIRE(o0 == 0, "text column %d: offsets must start at 0 (a partition's own buffer)", j);
        TMOG_CUDA(cudaMemcpyAsync(base + at_data[j], data[j], bytes, cudaMemcpyHostToDevice, c->s_cmp));
        TMOG_CUDA(cudaMemcpyAsync(base + at_off[j], offsets[j], sizeof(int64_t) * (size_t)(n + 1), cudaMemcpyHostToDevice,
                                  c->s_cmp));
        desc[4 * j] = reinterpret_cast<int64_t>(base + at_data[j]);
        desc[4 * j + 1] = reinterpret_cast<int64_t>(base + at_off[j]);
        if (valid && valid[j]) {
            TMOG_CUDA(cudaMemcpyAsync(base + at_valid[j], valid[j], sizeof(uint64_t) * vwords, cudaMemcpyHostToDevice,
                                      c->s_cmp));
            desc[4 * j + 2] = reinterpret_cast<int64_t>(base + at_valid[j]);
        }
    }
    TMOG_CUDA(cudaMemcpyAsync(c->s_desc.p, desc.data(), sizeof(int64_t) * desc.size(), cudaMemcpyHostToDevice, c->s_cmp));
    TMOG_CUDA(cudaStreamSynchronize(c->s_cmp));      // desc goes out of scope
    *d_desc = static_cast<int64_t*>(c->s_desc.p);
    return TMOG_OK;
}

int check_text_flags(const int32_t* f) {
    TMOG_REQUIRE(f[0] == 0, "non-ASCII text met: Lucene's Unicode word-break rules are restated for ASCII only");
    TMOG_REQUIRE(f[1] == 0, "a run of letters / digits longer than ~4000 bytes: beyond the tokenizer's text buffer");
    TMOG_REQUIRE(f[2] == 0, "a row holds more than ~12000 distinct hashed tokens: beyond the per-row table");
    return TMOG_OK;
}

}  // namespace

extern "C" int tmog_text_hash_stats_host(tmog_ctx* c, int ncols, int64_t n, const uint8_t* const* data,
                                         const int64_t* const* offsets, const uint64_t* const* valid,
                                         const uint8_t* prefix, const int32_t* prefix_len, const int32_t* params,
                                         const double* label, int max_classes, int* nclasses_out, double* classes_out,
                                         uint64_t* class_counts_out, int64_t* bucket_class_sum, int64_t* bucket_rows,
                                         int64_t* bucket_sq_sum, int32_t* bucket_max, int32_t* bucket_min) {
    TMOG_REQUIRE(c != nullptr && params != nullptr && nclasses_out != nullptr, "NULL argument");
    TMOG_REQUIRE(ncols >= 1 && n >= 1 && data && offsets && prefix && prefix_len, "bad input");
    TMOG_REQUIRE(bucket_class_sum && bucket_rows && bucket_sq_sum && bucket_max && bucket_min, "NULL output");
    TMOG_CUDA(cudaSetDevice(c->device));
    int ncls = 1;
    *nclasses_out = 0;
    if (label != nullptr) {
        int h = 0;
        if (int rc = label_classes(c, label, n, max_classes, &h, classes_out, class_counts_out)) return rc;
        *nclasses_out = h;
        TMOG_REQUIRE(h > 0, "the label has more than %d distinct values", max_classes);
        ncls = h;
    }
    int64_t* d_desc = nullptr;
    if (int rc = upload_text(c, ncols, n, data, offsets, valid, &d_desc)) return rc;
    int32_t p[12];
    memcpy(p, params, sizeof(p));
    p[7] = ncls;
    const int64_t w = (int64_t)p[6] + ncols;                         // null_base + columns
    size_t need = 0;
    if (int rc = tmog_text_hash_scratch_bytes2(n, ncols, p, 0, &need)) return rc;
    if (int rc = c->s_records.ensure(need)) return rc;
    const size_t words = (size_t)w * ncls + (size_t)w + (size_t)w * 2 + (size_t)w * 3 + 64;   // u32 cs | u32 dup | u64 ex | u32 aux | prefix
    if (int rc = c->s_counts.ensure(words * 4 + 16 * (size_t)ncols + 4 * (size_t)ncols + 64)) return rc;
    if (int rc = ensure_pinned(c, words * 4)) return rc;
    uint32_t* d_cs = static_cast<uint32_t*>(c->s_counts.p);
    uint32_t* d_dup = d_cs + (size_t)w * ncls;
    uint64_t* d_ex = reinterpret_cast<uint64_t*>(d_dup + w + (w & 1));
    uint32_t* d_aux = reinterpret_cast<uint32_t*>(d_ex + w);
    int32_t* d_flags = reinterpret_cast<int32_t*>(d_aux + 3 * (size_t)w);
    uint8_t* d_prefix = reinterpret_cast<uint8_t*>(d_flags + 8);
    int32_t* d_plen = reinterpret_cast<int32_t*>(d_prefix + 16 * (size_t)ncols);
    TMOG_CUDA(cudaMemsetAsync(c->s_counts.p, 0, words * 4, c->s_cmp));
    TMOG_CUDA(cudaMemcpyAsync(d_prefix, prefix, 16 * (size_t)ncols, cudaMemcpyHostToDevice, c->s_cmp));
    TMOG_CUDA(cudaMemcpyAsync(d_plen, prefix_len, 4 * (size_t)ncols, cudaMemcpyHostToDevice, c->s_cmp));
    if (int rc = tmog_text_hash_stats_dev(d_desc, ncols, n, d_prefix, d_plen, p,
                                          label ? static_cast<const uint8_t*>(c->class_ids.p) : nullptr, d_cs, d_dup, d_ex,
                                          d_aux, nullptr, 0.0, nullptr, d_flags, c->s_records.p, c->s_records.cap,
                                          c->s_cmp))
        return rc;
    uint32_t* h = static_cast<uint32_t*>(c->pinned);
    TMOG_CUDA(cudaMemcpyAsync(h, c->s_counts.p, words * 4, cudaMemcpyDeviceToHost, c->s_cmp));
    TMOG_CUDA(cudaStreamSynchronize(c->s_cmp));
    const uint32_t* h_cs = h;
    const uint32_t* h_dup = h_cs + (size_t)w * ncls;
    const uint64_t* h_ex = reinterpret_cast<const uint64_t*>(h_dup + w + (w & 1));
    const uint32_t* h_aux = reinterpret_cast<const uint32_t*>(h_ex + w);
    if (int rc = check_text_flags(reinterpret_cast<const int32_t*>(h_aux + 3 * (size_t)w))) return rc;
    for (int64_t i = 0; i < w; ++i) {
        int64_t tot = 0;
        for (int k = 0; k < ncls; ++k) {
            bucket_class_sum[i * ncls + k] = h_cs[i * ncls + k];
            tot += h_cs[i * ncls + k];
        }
        bucket_rows[i] = tot - (int64_t)h_dup[i];
        bucket_sq_sum[i] = tot + (int64_t)h_ex[i];
        bucket_max[i] = bucket_rows[i] > 0 ? (int32_t)std::max<uint32_t>(h_aux[3 * i], 1u) : 0;
        // smallest value among the rows holding the bucket: 1 unless every such row holds it more than once
        bucket_min[i] = bucket_rows[i] == 0 ? 0 : ((int64_t)h_aux[3 * i + 2] >= bucket_rows[i] ? (int32_t)~h_aux[3 * i + 1] : 1);
    }
    return TMOG_OK;
}

extern "C" int tmog_text_hash_transform_host(tmog_ctx* c, int ncols, int64_t n, const uint8_t* const* data,
                                             const int64_t* const* offsets, const uint64_t* const* valid,
                                             const uint8_t* prefix, const int32_t* prefix_len, const int32_t* params,
                                             const int32_t* newpos, int64_t* indptr, int32_t* out_idx, int32_t* out_val,
                                             int64_t capacity, int64_t* nnz_out) {
    TMOG_REQUIRE(c != nullptr && params != nullptr && newpos != nullptr && nnz_out != nullptr, "NULL argument");
    TMOG_REQUIRE(ncols >= 1 && n >= 1 && data && offsets && prefix && prefix_len, "bad input");
    TMOG_REQUIRE(capacity >= 0 && (capacity == 0 || (indptr && out_idx && out_val)), "NULL output");
    TMOG_CUDA(cudaSetDevice(c->device));
    int64_t* d_desc = nullptr;
    if (int rc = upload_text(c, ncols, n, data, offsets, valid, &d_desc)) return rc;
    int32_t p[12];
    memcpy(p, params, sizeof(p));
    p[7] = 1;
    const int64_t w = (int64_t)p[6] + ncols;
    size_t need = 0;
    if (int rc = tmog_text_hash_scratch_bytes2(n, ncols, p, 1, &need)) return rc;
    if (int rc = c->s_records.ensure(need)) return rc;
    // device: newpos | prefix | plen | flags | cursor | row_start | row_nnz | idx | val
    const size_t head = ((size_t)w * 4 + 16 * (size_t)ncols + 4 * (size_t)ncols + 64 + 255) & ~size_t(255);
    const size_t rows_b = ((size_t)n * 12 + 255) & ~size_t(255);
    if (int rc = c->s_out[0].ensure(head + rows_b + (size_t)std::max<int64_t>(capacity, 1) * 8 + 256)) return rc;
    uint8_t* base = static_cast<uint8_t*>(c->s_out[0].p);
    int32_t* d_newpos = reinterpret_cast<int32_t*>(base);
    uint8_t* d_prefix = reinterpret_cast<uint8_t*>(d_newpos + w);
    int32_t* d_plen = reinterpret_cast<int32_t*>(d_prefix + 16 * (size_t)ncols);
    int32_t* d_flags = d_plen + ncols;
    uint64_t* d_cursor = reinterpret_cast<uint64_t*>(base + head - 16);
    int64_t* d_start = reinterpret_cast<int64_t*>(base + head);
    int32_t* d_nnz = reinterpret_cast<int32_t*>(d_start + n);
    int32_t* d_idx = reinterpret_cast<int32_t*>(base + head + rows_b);
    int32_t* d_val = d_idx + std::max<int64_t>(capacity, 1);
    TMOG_CUDA(cudaMemcpyAsync(d_newpos, newpos, (size_t)w * 4, cudaMemcpyHostToDevice, c->s_cmp));
    TMOG_CUDA(cudaMemcpyAsync(d_prefix, prefix, 16 * (size_t)ncols, cudaMemcpyHostToDevice, c->s_cmp));
    TMOG_CUDA(cudaMemcpyAsync(d_plen, prefix_len, 4 * (size_t)ncols, cudaMemcpyHostToDevice, c->s_cmp));
    TMOG_CUDA(cudaMemsetAsync(d_flags, 0, 32, c->s_cmp));
    TMOG_CUDA(cudaMemsetAsync(d_cursor, 0, 16, c->s_cmp));
    if (int rc = tmog_text_hash_transform_dev(d_desc, ncols, n, d_prefix, d_plen, p, d_newpos, d_start, d_nnz,
                                              capacity ? d_idx : nullptr, capacity ? d_val : nullptr, d_cursor, capacity,
                                              d_flags, c->s_records.p, c->s_records.cap, c->s_cmp))
        return rc;
    int32_t flags[8];
    uint64_t total = 0;
    TMOG_CUDA(cudaMemcpyAsync(flags, d_flags, 32, cudaMemcpyDeviceToHost, c->s_cmp));
    TMOG_CUDA(cudaMemcpyAsync(&total, d_cursor, 8, cudaMemcpyDeviceToHost, c->s_cmp));
    TMOG_CUDA(cudaStreamSynchronize(c->s_cmp));
    *nnz_out = (int64_t)total;
    if (int rc = check_text_flags(flags)) return rc;
    if (capacity == 0) return TMOG_OK;                                   // sizing call
    if ((int64_t)total > capacity) {
        set_error("sparse capacity %lld too small: %lld entries", (long long)capacity, (long long)total);
        return TMOG_ENOMEM;
    }
    // rows come back in row order (Spark's sparse vectors, one per row): order them through the pinned buffer
    if (int rc = ensure_pinned(c, (size_t)n * 12 + (size_t)total * 8 + 64)) return rc;
    int64_t* h_start = static_cast<int64_t*>(c->pinned);
    int32_t* h_nnz = reinterpret_cast<int32_t*>(h_start + n);
    int32_t* h_idx = h_nnz + n + (n & 1);
    int32_t* h_val = h_idx + total;
    TMOG_CUDA(cudaMemcpyAsync(h_start, d_start, (size_t)n * 8, cudaMemcpyDeviceToHost, c->s_cmp));
    TMOG_CUDA(cudaMemcpyAsync(h_nnz, d_nnz, (size_t)n * 4, cudaMemcpyDeviceToHost, c->s_cmp));
    TMOG_CUDA(cudaMemcpyAsync(h_idx, d_idx, (size_t)total * 4, cudaMemcpyDeviceToHost, c->s_cmp));
    TMOG_CUDA(cudaMemcpyAsync(h_val, d_val, (size_t)total * 4, cudaMemcpyDeviceToHost, c->s_cmp));
    TMOG_CUDA(cudaStreamSynchronize(c->s_cmp));
    int64_t at = 0;
    for (int64_t r = 0; r < n; ++r) {
        indptr[r] = at;
        memcpy(out_idx + at, h_idx + h_start[r], sizeof(int32_t) * (size_t)h_nnz[r]);
        memcpy(out_val + at, h_val + h_start[r], sizeof(int32_t) * (size_t)h_nnz[r]);
        at += h_nnz[r];
    }
    indptr[n] = at;
    return TMOG_OK;
}
