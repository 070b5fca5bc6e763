// tmog_ctx: the per-thread context of the host-buffer entry points (host.cu, host_strings.cu).
#pragma once
#include "common.cuh"

namespace tmog {

struct DevBuf {
    void* p = nullptr;
    size_t cap = 0;
    int ensure(size_t bytes) {
        if (bytes <= cap) return TMOG_OK;
        if (p) cudaFree(p);
        p = nullptr;
        cap = 0;
        TMOG_CUDA(cudaMalloc(&p, bytes));
        cap = bytes;
        return TMOG_OK;
    }
    void release() {
        if (p) cudaFree(p);
        p = nullptr;
        cap = 0;
    }
};

}  // namespace tmog

using tmog::DevBuf;

struct tmog_ctx {
    int device = 0;
    cudaStream_t s_h2d = nullptr, s_cmp = nullptr, s_d2h = nullptr;
    // resident copy of the last fitted real block (lets transform skip the second upload)
    DevBuf r_values, r_valid, r_label, classes, class_ids, null_counts;
    const void* r_key = nullptr;
    int64_t r_n = 0, r_ld = 0, r_ldv = 0;
    int r_ncols = 0;
    bool r_has_valid = false;
    // streaming buffers
    DevBuf in_values[2], in_valid[2], out_block[2];
    DevBuf scratch, shifts, partials, fills, pos;
    cudaEvent_t ev_in[2] = {nullptr, nullptr}, ev_cmp[2] = {nullptr, nullptr}, ev_out[2] = {nullptr, nullptr};
    cudaEvent_t ev_lbl = nullptr;
    // pivot / text entry points (host_strings.cu): staged inputs, count tables, small tables
    DevBuf s_in[2], s_aux[2], s_out[2], s_counts, s_desc, s_pos, s_flags, s_records;
    void* pinned = nullptr;          // pinned bounce buffer for results that come back through it
    size_t pinned_cap = 0;
};

