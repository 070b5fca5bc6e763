// OpSetVectorizer (MultiPickList) on set columns in Arrow List<dictionary> layout: per column int64 offsets [n + 1] and
// int32 dictionary codes [offsets[n]]; row r holds codes [offsets[r], offsets[r + 1]) (no duplicates: it is a set; an
// empty range is the empty set / a null).  All string work is done per dictionary entry on the host: the kernels see
//   gid [ncodes]  the cleaned value (group) of every dictionary code          (fit)
//   pos [ncodes]  its output slot: 0 .. k-1 = index among the top values, k = OTHER   (statistics / transform)
//
//   set_count_kernel      rows holding each cleaned value, per label class    (OpOneHotVectorizer.fitFn: every row adds
//                         cleaned -> 1 once, however many of its raw values clean alike; OpOneHotVectorizer.scala:171-177)
//   set_rows_kernel       the row's entries (slot, multiplicity) in slot order (OneHotModelFun.pivotFn :416-437: a top
//                         value gets the number of raw values of the row that clean to it, OTHER the number of the
//                         rest, the null slot 1.0 for an empty set), either summed into SanityChecker statistics or
//                         written as sparse rows
//
// Sets are short (a handful of choices), so a thread owns a row and walks its codes once per output slot.
#include <algorithm>

#include "common.cuh"

namespace tmog {

constexpr int SET_THREADS = 256;

__global__ void __launch_bounds__(SET_THREADS)
set_count_kernel(const int64_t* __restrict__ offsets, const int32_t* __restrict__ codes, int64_t n,
                 const int32_t* __restrict__ gid, int ncodes, int ngroups, const uint8_t* __restrict__ class_ids,
                 int nclasses, unsigned int* __restrict__ counts, int* __restrict__ bad) {
    for (int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; r < n; r += (int64_t)gridDim.x * blockDim.x) {
        const int64_t a = offsets[r], b = offsets[r + 1];
        const int cls = nclasses > 1 ? class_ids[r] : 0;
        if (a >= b) {
            atomicAdd(&counts[cls], 1u);                       // id 0: the empty set
            continue;
        }
        for (int64_t i = a; i < b; ++i) {
            const int c = codes[i];
            if (c < 0 || c >= ncodes) {
                *bad = 1;
                continue;
            }
            const int g = gid[c];
            bool first = true;                                  // raw values that clean alike count once per row
            for (int64_t j = a; j < i && first; ++j) {
                const int cj = codes[j];
                if (cj >= 0 && cj < ncodes && gid[cj] == g) first = false;
            }
            if (first && g >= 0 && g < ngroups) atomicAdd(&counts[(int64_t)(1 + g) * nclasses + cls], 1u);
        }
    }
}

// mode 0: statistics; mode 1: row_nnz only; mode 2: write at row_start
template <int MODE>
__global__ void __launch_bounds__(SET_THREADS)
set_rows_kernel(const int64_t* __restrict__ offsets, const int32_t* __restrict__ codes, int64_t n,
                const int32_t* __restrict__ pos, int ncodes, int k, int track_nulls,
                const uint8_t* __restrict__ class_ids, int nclasses,
                // statistics (width = k + 2 slots: top values, OTHER, null)
                unsigned int* __restrict__ class_sum, unsigned int* __restrict__ present,
                unsigned long long* __restrict__ sq_extra, unsigned int* __restrict__ aux,
                const double* __restrict__ label, double label_shift, double* __restrict__ wsum,
                // sparse rows
                const int32_t* __restrict__ newpos, const int64_t* __restrict__ row_start, int32_t* __restrict__ row_nnz,
                int32_t* __restrict__ out_idx, int32_t* __restrict__ out_val, int* __restrict__ bad) {
    for (int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; r < n; r += (int64_t)gridDim.x * blockDim.x) {
        const int64_t a = offsets[r], b = offsets[r + 1];
        const int cls = (MODE == 0 && nclasses > 1) ? class_ids[r] : 0;
        int emitted = 0;
        int64_t at = MODE == 2 ? row_start[r] : 0;
        auto emit = [&](int slot, unsigned v) {
            if (MODE == 0) {
                atomicAdd(&class_sum[(int64_t)slot * nclasses + cls], v);
                atomicAdd(&present[(int64_t)slot * nclasses + cls], 1u);
                if (label) atomicAdd(&wsum[slot], (double)v * (label[r] - label_shift));
                if (v > 1u) {
                    atomicAdd(&sq_extra[slot], (unsigned long long)v * v - v);
                    atomicMax(&aux[3 * slot], v);
                    atomicMax(&aux[3 * slot + 1], ~v);
                    atomicAdd(&aux[3 * slot + 2], 1u);
                }
            } else {
                const int np = newpos[slot];
                if (np < 0) return;
                if (MODE == 2) {
                    out_idx[at + emitted] = np;
                    out_val[at + emitted] = (int32_t)v;
                }
                ++emitted;
            }
        };
        if (a >= b) {
            if (track_nulls) emit(k + 1, 1u);
        } else {
            for (int slot = 0; slot <= k; ++slot) {
                unsigned v = 0u;
                for (int64_t i = a; i < b; ++i) {
                    const int c = codes[i];
                    if (c < 0 || c >= ncodes) {
                        *bad = 1;
                        continue;
                    }
                    v += pos[c] == slot ? 1u : 0u;
                }
                if (v) emit(slot, v);
            }
        }
        if (MODE == 1) row_nnz[r] = emitted;
    }
}

static int set_grid(int64_t n) {
    return (int)std::max<int64_t>(1, std::min<int64_t>((n + SET_THREADS - 1) / SET_THREADS, (int64_t)sm_count() * 8));
}

}  // namespace tmog

using namespace tmog;

extern "C" int tmog_set_count_dev(const int64_t* offsets, const int32_t* codes, int64_t n, const int32_t* gid,
                                  int ncodes, int ngroups, const uint8_t* class_ids, int nclasses, uint32_t* counts,
                                  int32_t* bad, void* stream) {
    TMOG_REQUIRE(n >= 0 && ncodes >= 0 && ngroups >= 0 && nclasses >= 1 && nclasses <= 255, "bad shape");
    TMOG_REQUIRE(counts != nullptr && bad != nullptr && (n == 0 || offsets != nullptr), "NULL argument");
    TMOG_REQUIRE(nclasses == 1 || class_ids != nullptr || n == 0, "class_ids is NULL");
    if (n == 0) return TMOG_OK;
    TMOG_REQUIRE(ncodes == 0 || gid != nullptr, "gid is NULL");
    set_count_kernel<<<set_grid(n), SET_THREADS, 0, (cudaStream_t)stream>>>(offsets, codes, n, gid, ncodes, ngroups,
                                                                           class_ids, nclasses, counts, bad);
    TMOG_LAUNCH_CHECK();
    return TMOG_OK;
}

extern "C" int tmog_set_stats_dev(const int64_t* offsets, const int32_t* codes, int64_t n, const int32_t* pos,
                                  int ncodes, int k, int track_nulls, const uint8_t* class_ids, int nclasses,
                                  uint32_t* class_sum, uint32_t* present, uint64_t* sq_extra, uint32_t* aux,
                                  const double* label, double label_shift, double* wsum, int32_t* bad, void* stream) {
    TMOG_REQUIRE(n >= 0 && ncodes >= 0 && k >= 0 && nclasses >= 1 && nclasses <= 255, "bad shape");
    TMOG_REQUIRE(class_sum && present && sq_extra && aux && bad && (n == 0 || offsets != nullptr), "NULL argument");
    TMOG_REQUIRE(nclasses == 1 || class_ids != nullptr || n == 0, "class_ids is NULL");
    TMOG_REQUIRE((label == nullptr) == (wsum == nullptr), "label and wsum go together");
    if (n == 0) return TMOG_OK;
    TMOG_REQUIRE(ncodes == 0 || pos != nullptr, "pos is NULL");
    set_rows_kernel<0><<<set_grid(n), SET_THREADS, 0, (cudaStream_t)stream>>>(
        offsets, codes, n, pos, ncodes, k, track_nulls, class_ids, nclasses, class_sum, present,
        reinterpret_cast<unsigned long long*>(sq_extra), aux, label, label_shift, wsum, nullptr, nullptr, nullptr,
        nullptr, nullptr, bad);
    TMOG_LAUNCH_CHECK();
    return TMOG_OK;
}

extern "C" int tmog_set_transform_dev(const int64_t* offsets, const int32_t* codes, int64_t n, const int32_t* pos,
                                      int ncodes, int k, int track_nulls, const int32_t* newpos,
                                      const int64_t* row_start, int32_t* row_nnz, int32_t* out_idx, int32_t* out_val,
                                      int32_t* bad, void* stream) {
    TMOG_REQUIRE(n >= 0 && ncodes >= 0 && k >= 0, "bad shape");
    TMOG_REQUIRE(newpos && bad && (n == 0 || offsets != nullptr), "NULL argument");
    TMOG_REQUIRE(row_start != nullptr ? (out_idx && out_val) : row_nnz != nullptr,
                 "pass row_nnz to size the rows, or row_start with out_idx / out_val to write them");
    if (n == 0) return TMOG_OK;
    TMOG_REQUIRE(ncodes == 0 || pos != nullptr, "pos is NULL");
    cudaStream_t st = (cudaStream_t)stream;
    if (row_start == nullptr)
        set_rows_kernel<1><<<set_grid(n), SET_THREADS, 0, st>>>(offsets, codes, n, pos, ncodes, k, track_nulls, nullptr, 1,
                                                                nullptr, nullptr, nullptr, nullptr, nullptr, 0.0,
                                                                nullptr, newpos, nullptr, row_nnz, nullptr, nullptr,
                                                                bad);
    else
        set_rows_kernel<2><<<set_grid(n), SET_THREADS, 0, st>>>(offsets, codes, n, pos, ncodes, k, track_nulls, nullptr, 1,
                                                                nullptr, nullptr, nullptr, nullptr, nullptr, 0.0,
                                                                nullptr, newpos, row_start, row_nnz, out_idx, out_val,
                                                                bad);
    TMOG_LAUNCH_CHECK();
    return TMOG_OK;
}
