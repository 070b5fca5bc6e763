// Host-buffer entry points: what the JNI shim calls once per exported Spark partition.  A tmog_ctx owns the device
// staging buffers (grow-only, reused across calls), three streams (H2D, compute, D2H) and the events that pipeline
// row chunks through them.  Inputs should be pinned (tmog_pinned_alloc) so that the copies are asynchronous DMA.
#include <algorithm>
#include <new>
#include <vector>

#include "common.cuh"

#include "host_ctx.cuh"

using namespace tmog;

static int64_t round_up(int64_t v, int64_t m) { return (v + m - 1) / m * m; }

extern "C" int tmog_ctx_create(int device, tmog_ctx** out) {
    TMOG_REQUIRE(out != nullptr, "out is NULL");
    int ndev = 0;
    TMOG_CUDA(cudaGetDeviceCount(&ndev));
    TMOG_REQUIRE(device >= 0 && device < ndev, "device %d out of range (have %d)", device, ndev);
    TMOG_CUDA(cudaSetDevice(device));
    tmog_ctx* c = new (std::nothrow) tmog_ctx();
    TMOG_REQUIRE(c != nullptr, "out of host memory");
    c->device = device;
    TMOG_CUDA(cudaStreamCreateWithFlags(&c->s_h2d, cudaStreamNonBlocking));
    TMOG_CUDA(cudaStreamCreateWithFlags(&c->s_cmp, cudaStreamNonBlocking));
    TMOG_CUDA(cudaStreamCreateWithFlags(&c->s_d2h, cudaStreamNonBlocking));
    for (int i = 0; i < 2; ++i) {
        TMOG_CUDA(cudaEventCreateWithFlags(&c->ev_in[i], cudaEventDisableTiming));
        TMOG_CUDA(cudaEventCreateWithFlags(&c->ev_cmp[i], cudaEventDisableTiming));
        TMOG_CUDA(cudaEventCreateWithFlags(&c->ev_out[i], cudaEventDisableTiming));
    }
    TMOG_CUDA(cudaEventCreateWithFlags(&c->ev_lbl, cudaEventDisableTiming));
    *out = c;
    return TMOG_OK;
}

extern "C" int tmog_ctx_destroy(tmog_ctx* c) {
    if (c == nullptr) return TMOG_OK;
    cudaSetDevice(c->device);
    cudaDeviceSynchronize();
    c->r_values.release();
    c->r_valid.release();
    c->r_label.release();
    c->classes.release();
    c->class_ids.release();
    c->null_counts.release();
    if (c->ev_lbl) cudaEventDestroy(c->ev_lbl);
    for (int i = 0; i < 2; ++i) {
        c->in_values[i].release();
        c->in_valid[i].release();
        c->out_block[i].release();
        if (c->ev_in[i]) cudaEventDestroy(c->ev_in[i]);
        if (c->ev_cmp[i]) cudaEventDestroy(c->ev_cmp[i]);
        if (c->ev_out[i]) cudaEventDestroy(c->ev_out[i]);
    }
    c->scratch.release();
    c->shifts.release();
    c->partials.release();
    c->fills.release();
    c->pos.release();
    for (int i = 0; i < 2; ++i) {
        c->s_in[i].release();
        c->s_aux[i].release();
        c->s_out[i].release();
    }
    c->s_counts.release();
    c->s_desc.release();
    c->s_pos.release();
    c->s_flags.release();
    c->s_records.release();
    if (c->pinned) cudaFreeHost(c->pinned);
    if (c->s_h2d) cudaStreamDestroy(c->s_h2d);
    if (c->s_cmp) cudaStreamDestroy(c->s_cmp);
    if (c->s_d2h) cudaStreamDestroy(c->s_d2h);
    delete c;
    return TMOG_OK;
}

extern "C" int tmog_pinned_alloc(size_t bytes, void** out) {
    TMOG_REQUIRE(out != nullptr, "out is NULL");
    TMOG_CUDA(cudaHostAlloc(out, bytes ? bytes : 1, cudaHostAllocDefault));
    return TMOG_OK;
}

extern "C" int tmog_pinned_free(void* p) {
    if (p) TMOG_CUDA(cudaFreeHost(p));
    return TMOG_OK;
}

static int normalise_chunk(int64_t n, int64_t chunk_rows, int64_t* out) {
    if (chunk_rows <= 0) chunk_rows = 1 << 20;
    chunk_rows = round_up(chunk_rows, 2048);  // keeps every chunk start 16-byte and bitmap-word aligned
    *out = std::min<int64_t>(chunk_rows, std::max<int64_t>(round_up(n, 2048), 2048));
    return TMOG_OK;
}

extern "C" int tmog_real_fit_host(tmog_ctx* c, const double* values, int64_t ld, const uint64_t* valid, int64_t ldv,
                                  const double* label, int64_t n, int ncols, int64_t chunk_rows, int keep_resident,
                                  const double* shifts_in, double* shifts_out, double* partials_out, int max_classes,
                                  double* classes_out, int* nclasses_out, uint64_t* class_counts_out,
                                  uint64_t* null_class_counts_out) {
    TMOG_REQUIRE(c != nullptr, "ctx is NULL");
    TMOG_REQUIRE(n >= 0 && ncols >= 0 && ld >= n, "bad shape n=%lld ncols=%d ld=%lld", (long long)n, ncols,
                 (long long)ld);
    TMOG_REQUIRE((values != nullptr || n == 0 || ncols == 0) && (label != nullptr || n == 0), "NULL input");
    TMOG_REQUIRE(valid == nullptr || ldv >= (n + 63) / 64, "bitmap stride too small");
    TMOG_REQUIRE(partials_out != nullptr, "partials_out is NULL");
    TMOG_REQUIRE(max_classes >= 0 && max_classes <= 255, "max_classes must be in [0, 255]");
    TMOG_REQUIRE(max_classes == 0 || (classes_out && nclasses_out), "classes_out / nclasses_out is NULL");
    TMOG_CUDA(cudaSetDevice(c->device));
    const int NS = TMOG_REAL_NSTAT;
    if (n == 0) {  // empty partition: neutral partials
        for (int col = 0; col <= ncols; ++col) {
            double* q = partials_out + (size_t)col * NS;
            for (int s = 0; s < NS; ++s) q[s] = 0.0;
            q[5] = INFINITY;
            q[6] = -INFINITY;
        }
        if (shifts_out)
            for (int col = 0; col <= ncols; ++col) shifts_out[col] = shifts_in ? shifts_in[col] : 0.0;
        if (nclasses_out) *nclasses_out = 0;
        c->r_key = nullptr;
        return TMOG_OK;
    }
    int64_t chunk;
    normalise_chunk(n, chunk_rows, &chunk);
    const int64_t nchunks = (n + chunk - 1) / chunk;
    const size_t pstride = (size_t)(ncols + 1) * NS;

    size_t scratch_bytes = 0, lbl_scratch = 0;
    if (int rc = tmog_real_fit_scratch_bytes(ncols, &scratch_bytes)) return rc;
    if (int rc = tmog_label_classes_scratch_bytes(&lbl_scratch)) return rc;
    if (int rc = c->scratch.ensure(std::max(scratch_bytes, lbl_scratch))) return rc;
    if (int rc = c->shifts.ensure(sizeof(double) * (ncols + 1))) return rc;
    if (int rc = c->partials.ensure(sizeof(double) * pstride * nchunks)) return rc;

    // the label is small (8 B/row): always uploaded whole, so that its classes are known before the columns stream
    const int64_t nld = round_up(n, 2048);
    if (int rc = c->r_label.ensure(sizeof(double) * (size_t)nld)) return rc;
    double* d_label = static_cast<double*>(c->r_label.p);
    TMOG_CUDA(cudaMemcpyAsync(d_label, label, sizeof(double) * n, cudaMemcpyHostToDevice, c->s_h2d));
    TMOG_CUDA(cudaEventRecord(c->ev_lbl, c->s_h2d));
    TMOG_CUDA(cudaStreamWaitEvent(c->s_cmp, c->ev_lbl, 0));
    int h_ncls = 0;
    const bool want_null_counts = max_classes > 0 && null_class_counts_out != nullptr && valid != nullptr;
    if (max_classes > 0) {
        const size_t cb = sizeof(double) * 256 + sizeof(uint64_t) * 256 + sizeof(int) * 2;
        if (int rc = c->classes.ensure(cb)) return rc;
        if (int rc = c->class_ids.ensure((size_t)nld)) return rc;
        double* d_cls = static_cast<double*>(c->classes.p);
        uint64_t* d_cnt = reinterpret_cast<uint64_t*>(d_cls + 256);
        int* d_ncls = reinterpret_cast<int*>(d_cnt + 256);
        if (int rc = tmog_label_classes_dev(d_label, n, max_classes, d_cls, d_ncls, static_cast<uint8_t*>(c->class_ids.p),
                                            d_cnt, c->scratch.p, c->scratch.cap, c->s_cmp))
            return rc;
        TMOG_CUDA(cudaMemcpyAsync(&h_ncls, d_ncls, sizeof(int), cudaMemcpyDeviceToHost, c->s_cmp));
        TMOG_CUDA(cudaStreamSynchronize(c->s_cmp));  // the class count decides whether null counts are taken
        *nclasses_out = h_ncls;
        if (h_ncls > 0) {
            TMOG_CUDA(cudaMemcpyAsync(classes_out, d_cls, sizeof(double) * h_ncls, cudaMemcpyDeviceToHost, c->s_cmp));
            if (class_counts_out)
                TMOG_CUDA(cudaMemcpyAsync(class_counts_out, d_cnt, sizeof(uint64_t) * h_ncls, cudaMemcpyDeviceToHost,
                                          c->s_cmp));
        }
        if (want_null_counts && h_ncls > 0)
            if (int rc = c->null_counts.ensure(sizeof(uint64_t) * (size_t)std::max(ncols, 1) * h_ncls)) return rc;
    }

    // device layout of a chunk: columns with stride dld (rows rounded to 2048), bitmaps with stride dldv words
    const bool resident = keep_resident != 0;
    const int64_t dld = resident ? nld : chunk;
    const int64_t dldv = dld / 64;
    if (resident) {
        if (int rc = c->r_values.ensure(sizeof(double) * (size_t)dld * std::max(ncols, 1))) return rc;
        if (valid)
            if (int rc = c->r_valid.ensure(sizeof(uint64_t) * (size_t)dldv * std::max(ncols, 1))) return rc;
        c->r_key = values;
        c->r_n = n;
        c->r_ld = dld;
        c->r_ldv = dldv;
        c->r_ncols = ncols;
        c->r_has_valid = valid != nullptr;
    } else {
        c->r_key = nullptr;
        for (int b = 0; b < 2; ++b) {
            if (int rc = c->in_values[b].ensure(sizeof(double) * (size_t)chunk * std::max(ncols, 1))) return rc;
            if (valid)
                if (int rc = c->in_valid[b].ensure(sizeof(uint64_t) * (size_t)(chunk / 64) * std::max(ncols, 1)))
                    return rc;
        }
    }
    double* d_shifts = static_cast<double*>(c->shifts.p);
    double* d_partials = static_cast<double*>(c->partials.p);
    if (shifts_in)
        TMOG_CUDA(cudaMemcpyAsync(d_shifts, shifts_in, sizeof(double) * (ncols + 1), cudaMemcpyHostToDevice, c->s_cmp));

    for (int64_t k = 0; k < nchunks; ++k) {
        const int b = (int)(k & 1);
        const int64_t r0 = k * chunk;
        const int64_t rows = std::min<int64_t>(chunk, n - r0);
        double* dv;
        uint64_t* dm;
        if (resident) {
            dv = static_cast<double*>(c->r_values.p) + r0;
            dm = valid ? static_cast<uint64_t*>(c->r_valid.p) + r0 / 64 : nullptr;
        } else {
            dv = static_cast<double*>(c->in_values[b].p);
            dm = valid ? static_cast<uint64_t*>(c->in_valid[b].p) : nullptr;
            if (k >= 2) TMOG_CUDA(cudaStreamWaitEvent(c->s_h2d, c->ev_cmp[b], 0));  // buffer b free again
        }
        if (ncols > 0)
            TMOG_CUDA(cudaMemcpy2DAsync(dv, sizeof(double) * dld, values + r0, sizeof(double) * ld,
                                        sizeof(double) * rows, ncols, cudaMemcpyHostToDevice, c->s_h2d));
        if (valid && ncols > 0) {
            const int64_t words = (rows + 63) / 64;
            TMOG_CUDA(cudaMemcpy2DAsync(dm, sizeof(uint64_t) * dldv, valid + r0 / 64, sizeof(uint64_t) * ldv,
                                        sizeof(uint64_t) * words, ncols, cudaMemcpyHostToDevice, c->s_h2d));
        }
        TMOG_CUDA(cudaEventRecord(c->ev_in[b], c->s_h2d));
        TMOG_CUDA(cudaStreamWaitEvent(c->s_cmp, c->ev_in[b], 0));
        const double* dy = d_label + r0;
        if (k == 0 && !shifts_in)
            if (int rc = tmog_real_pilot_shift_dev(dv, dld, dm, dldv, dy, rows, ncols, d_shifts, c->s_cmp)) return rc;
        if (int rc = tmog_real_fit_partials_dev(dv, dld, dm, dldv, dy, rows, ncols, d_shifts, c->scratch.p,
                                                c->scratch.cap, d_partials + pstride * k, c->s_cmp))
            return rc;
        if (want_null_counts && h_ncls > 0)
            if (int rc = tmog_null_class_counts_dev(dm, dldv, static_cast<uint8_t*>(c->class_ids.p) + r0, rows, ncols,
                                                    h_ncls, k > 0, static_cast<uint64_t*>(c->null_counts.p),
                                                    c->s_cmp))
                return rc;
        TMOG_CUDA(cudaEventRecord(c->ev_cmp[b], c->s_cmp));
    }
    std::vector<double> host((size_t)pstride * nchunks);
    std::vector<double> hshift(ncols + 1);
    TMOG_CUDA(cudaMemcpyAsync(host.data(), d_partials, sizeof(double) * host.size(), cudaMemcpyDeviceToHost, c->s_cmp));
    TMOG_CUDA(cudaMemcpyAsync(hshift.data(), d_shifts, sizeof(double) * (ncols + 1), cudaMemcpyDeviceToHost, c->s_cmp));
    if (want_null_counts && h_ncls > 0 && ncols > 0)
        TMOG_CUDA(cudaMemcpyAsync(null_class_counts_out, c->null_counts.p, sizeof(uint64_t) * (size_t)ncols * h_ncls,
                                  cudaMemcpyDeviceToHost, c->s_cmp));
    TMOG_CUDA(cudaStreamSynchronize(c->s_cmp));
    std::copy(host.begin(), host.begin() + pstride, partials_out);
    for (int64_t k = 1; k < nchunks; ++k) tmog_real_partials_merge(partials_out, host.data() + pstride * k, ncols);
    if (shifts_out) std::copy(hshift.begin(), hshift.end(), shifts_out);
    return TMOG_OK;
}

extern "C" int tmog_real_transform_host(tmog_ctx* c, const double* values, int64_t ld, const uint64_t* valid,
                                        int64_t ldv, int64_t n, int ncols, const double* fills, int track_nulls,
                                        int64_t chunk_rows, const int32_t* keep, int nkeep, double* out,
                                        int64_t out_ld, int64_t out_col0) {
    TMOG_REQUIRE(c != nullptr, "ctx is NULL");
    TMOG_REQUIRE(n >= 0 && ncols >= 0 && ld >= n, "bad shape");
    TMOG_REQUIRE((values != nullptr && out != nullptr) || n == 0 || ncols == 0, "NULL input / output");
    TMOG_REQUIRE(fills != nullptr || valid == nullptr, "fills is NULL but the columns are nullable");
    const int64_t full_width = (int64_t)ncols * (track_nulls ? 2 : 1);
    TMOG_REQUIRE(keep == nullptr || (nkeep >= 0 && nkeep <= full_width), "bad nkeep %d", nkeep);
    const int64_t width = keep ? nkeep : full_width;
    TMOG_REQUIRE(out_col0 >= 0 && out_ld >= out_col0 + width, "output block exceeds the row stride");
    if (n == 0 || ncols == 0 || width == 0) return TMOG_OK;
    TMOG_CUDA(cudaSetDevice(c->device));
    uint64_t* d_keep = nullptr;
    if (keep) {
        std::vector<uint64_t> mask((size_t)(full_width + 63) / 64, 0ull);
        for (int i = 0; i < nkeep; ++i) {
            TMOG_REQUIRE(keep[i] >= 0 && keep[i] < full_width && (i == 0 || keep[i] > keep[i - 1]),
                         "keep must be ascending indices in [0, %lld)", (long long)full_width);
            mask[keep[i] >> 6] |= 1ull << (keep[i] & 63);
        }
        if (int rc = c->pos.ensure(sizeof(uint64_t) * mask.size())) return rc;
        d_keep = static_cast<uint64_t*>(c->pos.p);
        TMOG_CUDA(cudaMemcpy(d_keep, mask.data(), sizeof(uint64_t) * mask.size(), cudaMemcpyHostToDevice));
    }
    int64_t chunk;
    normalise_chunk(n, chunk_rows, &chunk);
    const int64_t nchunks = (n + chunk - 1) / chunk;
    const bool resident = c->r_key == values && c->r_n == n && c->r_ncols == ncols &&
                          c->r_has_valid == (valid != nullptr);
    const int64_t dw = width;  // dense device rows: padding that is never written costs partial-sector fills
    if (int rc = c->fills.ensure(sizeof(double) * std::max(ncols, 1))) return rc;
    double* d_fills = nullptr;
    if (fills) {
        d_fills = static_cast<double*>(c->fills.p);
        TMOG_CUDA(cudaMemcpyAsync(d_fills, fills, sizeof(double) * ncols, cudaMemcpyHostToDevice, c->s_cmp));
    }
    for (int b = 0; b < 2; ++b) {
        if (int rc = c->out_block[b].ensure(sizeof(double) * (size_t)chunk * dw)) return rc;
        if (!resident) {
            if (int rc = c->in_values[b].ensure(sizeof(double) * (size_t)chunk * ncols)) return rc;
            if (valid)
                if (int rc = c->in_valid[b].ensure(sizeof(uint64_t) * (size_t)(chunk / 64) * ncols)) return rc;
        }
    }
    for (int64_t k = 0; k < nchunks; ++k) {
        const int b = (int)(k & 1);
        const int64_t r0 = k * chunk;
        const int64_t rows = std::min<int64_t>(chunk, n - r0);
        const double* dv;
        const uint64_t* dm;
        int64_t dld, dldv;
        if (resident) {
            dv = static_cast<const double*>(c->r_values.p) + r0;
            dm = valid ? static_cast<const uint64_t*>(c->r_valid.p) + r0 / 64 : nullptr;
            dld = c->r_ld;
            dldv = c->r_ldv;
        } else {
            dld = chunk;
            dldv = chunk / 64;
            if (k >= 2) TMOG_CUDA(cudaStreamWaitEvent(c->s_h2d, c->ev_cmp[b], 0));
            TMOG_CUDA(cudaMemcpy2DAsync(c->in_values[b].p, sizeof(double) * dld, values + r0, sizeof(double) * ld,
                                        sizeof(double) * rows, ncols, cudaMemcpyHostToDevice, c->s_h2d));
            if (valid)
                TMOG_CUDA(cudaMemcpy2DAsync(c->in_valid[b].p, sizeof(uint64_t) * dldv, valid + r0 / 64,
                                            sizeof(uint64_t) * ldv, sizeof(uint64_t) * ((rows + 63) / 64), ncols,
                                            cudaMemcpyHostToDevice, c->s_h2d));
            TMOG_CUDA(cudaEventRecord(c->ev_in[b], c->s_h2d));
            TMOG_CUDA(cudaStreamWaitEvent(c->s_cmp, c->ev_in[b], 0));
            dv = static_cast<const double*>(c->in_values[b].p);
            dm = valid ? static_cast<const uint64_t*>(c->in_valid[b].p) : nullptr;
        }
        if (k >= 2) TMOG_CUDA(cudaStreamWaitEvent(c->s_cmp, c->ev_out[b], 0));  // out_block[b] drained
        double* dout = static_cast<double*>(c->out_block[b].p);
        if (d_keep) {
            if (int rc = tmog_real_transform_select_dev(dv, dld, dm, dldv, rows, ncols, d_fills, track_nulls, d_keep,
                                                        nkeep, dout, dw, 0, c->s_cmp))
                return rc;
        } else if (int rc = tmog_real_transform_dev(dv, dld, dm, dldv, rows, ncols, d_fills, track_nulls, dout, dw, 0,
                                                    c->s_cmp))
            return rc;
        TMOG_CUDA(cudaEventRecord(c->ev_cmp[b], c->s_cmp));
        TMOG_CUDA(cudaStreamWaitEvent(c->s_d2h, c->ev_cmp[b], 0));
        if (out_ld == width && dw == width)  // one contiguous run: plain DMA
            TMOG_CUDA(cudaMemcpyAsync(out + r0 * out_ld, dout, sizeof(double) * (size_t)rows * width,
                                      cudaMemcpyDeviceToHost, c->s_d2h));
        else
            TMOG_CUDA(cudaMemcpy2DAsync(out + r0 * out_ld + out_col0, sizeof(double) * out_ld, dout,
                                        sizeof(double) * dw, sizeof(double) * width, rows, cudaMemcpyDeviceToHost,
                                        c->s_d2h));
        TMOG_CUDA(cudaEventRecord(c->ev_out[b], c->s_d2h));
    }
    TMOG_CUDA(cudaStreamSynchronize(c->s_d2h));
    TMOG_CUDA(cudaStreamSynchronize(c->s_cmp));
    return TMOG_OK;
}
