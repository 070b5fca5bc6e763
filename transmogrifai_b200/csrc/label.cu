// Label classes: the distinct label values (SanityChecker.categoricalTests keys its reduceByKey on them,
// SanityChecker.scala:265-281) and a one-byte class id per row that the count kernels read instead of the f64 label.
// Also: per-class null counts of nullable columns (the contingency rows of lone null indicators, :306-322).
#include <algorithm>

#include "common.cuh"

namespace tmog {

constexpr int LBL_SLOTS = 1024;                       // open-addressing table, power of two
constexpr unsigned long long LBL_EMPTY = 0xFFF8DEADBEEF0001ull;  // a NaN payload no canonicalised label can have
constexpr unsigned long long CANON_NAN = 0x7FF8000000000000ull;  // java.lang.Double.doubleToLongBits(NaN)

// scratch layout: [0 .. LBL_SLOTS) table of bit patterns, [LBL_SLOTS] distinct counter (as u64)
__device__ __forceinline__ unsigned long long label_bits(double y) {
    return y != y ? CANON_NAN : (unsigned long long)__double_as_longlong(y);
}

__device__ __forceinline__ unsigned slot_of(unsigned long long b) {
    b ^= b >> 33;
    b *= 0xff51afd7ed558ccdull;
    b ^= b >> 33;
    return (unsigned)b & (LBL_SLOTS - 1);
}

__global__ void label_init_kernel(unsigned long long* table) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < LBL_SLOTS) table[i] = LBL_EMPTY;
    if (i == LBL_SLOTS) table[i] = 0ull;
}

__global__ void __launch_bounds__(256)
label_distinct_kernel(const double* __restrict__ label, int64_t n, int max_classes, unsigned long long* table) {
    // the values this thread entered last: a label of a few classes then costs one load and two compares per row
    unsigned long long last = LBL_EMPTY, last2 = LBL_EMPTY;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    auto enter = [&](unsigned long long b) {
        if (b == last || b == last2) return;
        last2 = last;
        last = b;
        // the lanes that meet a new value in the same step (all of them in the first one) agree on one lane per
        // distinct value to carry it to the table: a few thousand table accesses instead of one per thread
        const unsigned peers = __match_any_sync(__activemask(), b);
        if ((int)(__ffs(peers) - 1) != (int)(threadIdx.x & 31)) return;
        if (table[LBL_SLOTS] > (unsigned long long)max_classes) return;  // already "continuous"
        unsigned s = slot_of(b);
        for (int probe = 0; probe < LBL_SLOTS; ++probe) {
            const unsigned long long cur = table[s];
            if (cur == b) break;
            if (cur == LBL_EMPTY) {
                const unsigned long long old = atomicCAS(&table[s], LBL_EMPTY, b);
                if (old == LBL_EMPTY) {
                    atomicAdd(&table[LBL_SLOTS], 1ull);
                    break;
                }
                if (old == b) break;
            }
            s = (s + 1) & (LBL_SLOTS - 1);
        }
    };
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    for (; i + 7 * stride < n; i += 8 * stride) {      // eight independent loads in flight per thread
        double y[8];
#pragma unroll
        for (int u = 0; u < 8; ++u) y[u] = __ldcs(label + i + u * stride);
#pragma unroll
        for (int u = 0; u < 8; ++u) enter(label_bits(y[u]));
    }
    for (; i < n; i += stride) enter(label_bits(label[i]));
}

// one block: compact the table, rank-sort ascending by numeric value (NaN last, -0.0 before 0.0 as
// java.lang.Double.compare orders them)
__global__ void __launch_bounds__(LBL_SLOTS)
label_sort_kernel(const unsigned long long* __restrict__ table, int max_classes, double* __restrict__ classes,
                  int* __restrict__ nclasses) {
    __shared__ unsigned long long vals[LBL_SLOTS];
    __shared__ int count;
    const int t = threadIdx.x;
    if (t == 0) count = 0;
    __syncthreads();
    const unsigned long long mine = table[t];
    if (mine != LBL_EMPTY) vals[atomicAdd(&count, 1)] = mine;
    __syncthreads();
    const int m = count;
    if (t == 0) *nclasses = (m > max_classes || table[LBL_SLOTS] > (unsigned long long)max_classes) ? -1 : m;
    if (m > max_classes) return;
    if (t < m) {
        // total order key: flip so that unsigned comparison == Double.compare
        auto key = [](unsigned long long b) { return (b >> 63) ? ~b : (b | 0x8000000000000000ull); };
        const unsigned long long k = key(vals[t]);
        int rank = 0;
        for (int j = 0; j < m; ++j) rank += key(vals[j]) < k;
        classes[rank] = __longlong_as_double((long long)vals[t]);
    }
}

__global__ void __launch_bounds__(256)
label_class_id_kernel(const double* __restrict__ label, int64_t n, const double* __restrict__ classes,
                      const int* __restrict__ nclasses, uint8_t* __restrict__ ids,
                      unsigned long long* __restrict__ class_counts) {
    __shared__ unsigned long long cls[256];
    __shared__ unsigned int hist[256];
    const int m = *nclasses;
    if (m <= 0) return;
    for (int i = threadIdx.x; i < 256; i += blockDim.x) {
        hist[i] = 0;
        if (i < m) cls[i] = label_bits(classes[i]);
    }
    __syncthreads();
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i0 = (int64_t)blockIdx.x * blockDim.x; i0 < n; i0 += stride) {
        const int64_t i = i0 + threadIdx.x;
        int id = -1;
        if (i < n) {
            const unsigned long long b = label_bits(label[i]);
            id = 0;
            for (int j = 0; j < m; ++j)
                if (cls[j] == b) id = j;
            if (ids) ids[i] = (uint8_t)id;
        }
        // warp-aggregated histogram: one shared atomic per distinct class per warp
        const unsigned active = __ballot_sync(FULL, id >= 0);
        if (id >= 0) {
            const unsigned peers = __match_any_sync(active, id);
            if ((int)(__ffs(peers) - 1) == (int)(threadIdx.x & 31)) atomicAdd(&hist[id], (unsigned)__popc(peers));
        }
    }
    __syncthreads();
    if (class_counts)
        for (int i = threadIdx.x; i < m; i += blockDim.x)
            if (hist[i]) atomicAdd(&class_counts[i], (unsigned long long)hist[i]);
}

// counts[c][k] += number of NULL rows of column c whose label class is k.  A CTA walks tiles of 2048 rows (32 bitmap
// words per column).  Per tile the class ids become one 64-row bit mask per class and word (ballots; classes 0..7),
// after which a column costs one coalesced load of its 32 validity words and, per class, a popcount of
// nulls & mask summed over the warp -- no per-null work.  A tile that holds a class id >= 8 takes the general path
// instead: every null bit is visited and counted in a per-warp shared-memory histogram.
constexpr int NC_FAST = 8;

__global__ void __launch_bounds__(256)
null_class_count_kernel(const uint64_t* __restrict__ valid, int64_t ldv, const uint8_t* __restrict__ ids, int64_t n,
                        int ncols, int nclasses, unsigned long long* __restrict__ counts) {
    extern __shared__ unsigned int nc_smem[];       // [ncols][NC_FAST] accumulators | [8 warps][nclasses] histograms
    __shared__ unsigned int masks[NC_FAST][64];     // class k, 32-row half-word h of the tile
    __shared__ int big;
    unsigned int* acc = nc_smem;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    unsigned int* h = nc_smem + ncols * NC_FAST + warp * nclasses;
    for (int i = tid; i < ncols * NC_FAST; i += 256) acc[i] = 0u;
    const int64_t ntiles = (n + 2047) / 2048;
    for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        __syncthreads();                            // the previous tile's masks are no longer read
        if (tid == 0) big = 0;
        __syncthreads();
        int idv[8];
        int idmax = -1;
#pragma unroll
        for (int it = 0; it < 8; ++it) {               // the tile's ids: eight independent loads per thread
            const int64_t row = tile * 2048 + it * 256 + tid;
            idv[it] = row < n ? (int)ids[row] : -1;
            idmax = max(idmax, idv[it]);
        }
        idmax = __reduce_max_sync(FULL, idmax);
        if (lane == 0 && idmax > 0) atomicMax(&big, idmax);      // big = the largest class id of the tile
        __syncthreads();
        const int kc = min(big + 1, NC_FAST);                    // classes that can occur in this tile
#pragma unroll
        for (int it = 0; it < 8; ++it) {
            const int id = idv[it];
            for (int k = 0; k < kc; ++k) {
                const unsigned b = __ballot_sync(FULL, id == k);
                if (lane == 0) masks[k][it * 8 + warp] = b;
            }
        }
        __syncthreads();
        const int64_t word = tile * 32 + lane;
        const int64_t r0 = word * 64;
        if (big < NC_FAST) {
            const unsigned long long* m64 = reinterpret_cast<const unsigned long long*>(&masks[0][0]);
            unsigned long long mk[NC_FAST];
#pragma unroll
            for (int k = 0; k < NC_FAST; ++k) mk[k] = k < kc ? m64[k * 32 + lane] : 0ull;
            const unsigned long long tail = (r0 < n) ? ((n - r0 < 64) ? (1ull << (n - r0)) - 1ull : ~0ull) : 0ull;
            for (int c0 = warp; c0 < ncols; c0 += 8 * 8) {      // eight columns' words in flight per lane
                unsigned long long nulls[8];
#pragma unroll
                for (int u = 0; u < 8; ++u) {
                    const int c = c0 + 8 * u;
                    nulls[u] = (c < ncols && tail) ? ~__ldcs(valid + (int64_t)c * ldv + word) & tail : 0ull;
                }
#pragma unroll
                for (int u = 0; u < 8; ++u) {
                    const int c = c0 + 8 * u;
                    if (c >= ncols || !__any_sync(FULL, nulls[u] != 0ull)) continue;
#pragma unroll
                    for (int k = 0; k < NC_FAST; ++k) {
                        if (k >= kc) break;
                        const unsigned cnt = __reduce_add_sync(FULL, (unsigned)__popcll(nulls[u] & mk[k]));
                        if (lane == 0 && cnt) acc[c * NC_FAST + k] += cnt;     // column c belongs to this warp alone
                    }
                }
            }
        } else {
            for (int c = warp; c < ncols; c += 8) {
                for (int k = lane; k < nclasses; k += 32) h[k] = 0;
                __syncwarp();
                if (r0 < n) {
                    unsigned long long nulls = ~valid[(int64_t)c * ldv + word];
                    if (n - r0 < 64) nulls &= (1ull << (n - r0)) - 1ull;
                    while (nulls) {
                        const int bit = __ffsll((long long)nulls) - 1;
                        nulls &= nulls - 1;
                        atomicAdd(&h[ids[r0 + bit]], 1u);
                    }
                }
                __syncwarp();
                for (int k = lane; k < nclasses; k += 32)
                    if (h[k]) atomicAdd(&counts[(int64_t)c * nclasses + k], (unsigned long long)h[k]);
                __syncwarp();
            }
        }
    }
    __syncthreads();
    for (int i = tid; i < ncols * NC_FAST; i += 256) {
        const int c = i / NC_FAST, k = i % NC_FAST;
        if (acc[i] && k < nclasses) atomicAdd(&counts[(int64_t)c * nclasses + k], (unsigned long long)acc[i]);
    }
}

}  // namespace tmog

using namespace tmog;

extern "C" int tmog_label_classes_scratch_bytes(size_t* bytes) {
    TMOG_REQUIRE(bytes != nullptr, "bytes is NULL");
    *bytes = sizeof(unsigned long long) * (LBL_SLOTS + 1);
    return TMOG_OK;
}

extern "C" int tmog_label_classes_dev(const double* label, int64_t n, int max_classes, double* classes,
                                      int* nclasses, uint8_t* class_ids, uint64_t* class_counts, void* scratch,
                                      size_t scratch_bytes, void* stream) {
    TMOG_REQUIRE(n >= 0 && (label != nullptr || n == 0), "label is NULL");
    TMOG_REQUIRE(max_classes >= 1 && max_classes <= 255, "max_classes must be in [1, 255], got %d", max_classes);
    TMOG_REQUIRE(classes != nullptr && nclasses != nullptr, "classes / nclasses is NULL");
    if (scratch == nullptr || scratch_bytes < sizeof(unsigned long long) * (LBL_SLOTS + 1)) {
        set_error("scratch too small (tmog_label_classes_scratch_bytes)");
        return TMOG_ENOMEM;
    }
    cudaStream_t st = (cudaStream_t)stream;
    unsigned long long* table = static_cast<unsigned long long*>(scratch);
    label_init_kernel<<<(LBL_SLOTS + 1 + 255) / 256, 256, 0, st>>>(table);
    TMOG_LAUNCH_CHECK();
    if (n > 0) {
        const int grid = (int)std::min<int64_t>((n + 255) / 256, (int64_t)sm_count() * 8);
        label_distinct_kernel<<<grid, 256, 0, st>>>(label, n, max_classes, table);
        TMOG_LAUNCH_CHECK();
    }
    label_sort_kernel<<<1, LBL_SLOTS, 0, st>>>(table, max_classes, classes, nclasses);
    TMOG_LAUNCH_CHECK();
    if (class_counts != nullptr) TMOG_CUDA(cudaMemsetAsync(class_counts, 0, sizeof(uint64_t) * max_classes, st));
    if ((class_ids != nullptr || class_counts != nullptr) && n > 0) {
        const int grid = (int)std::min<int64_t>((n + 255) / 256, (int64_t)sm_count() * 8);
        label_class_id_kernel<<<grid, 256, 0, st>>>(label, n, classes, nclasses, class_ids,
                                                    reinterpret_cast<unsigned long long*>(class_counts));
        TMOG_LAUNCH_CHECK();
    }
    return TMOG_OK;
}

extern "C" int tmog_label_class_ids_dev(const double* label, int64_t n, const double* classes, const int* nclasses,
                                       int max_classes, uint8_t* class_ids, uint64_t* class_counts, void* stream) {
    TMOG_REQUIRE(n >= 0 && (label != nullptr || n == 0), "label is NULL");
    TMOG_REQUIRE(classes != nullptr && nclasses != nullptr, "classes / nclasses is NULL");
    TMOG_REQUIRE(max_classes >= 1 && max_classes <= 255, "max_classes must be in [1, 255]");
    cudaStream_t st = (cudaStream_t)stream;
    if (class_counts != nullptr) TMOG_CUDA(cudaMemsetAsync(class_counts, 0, sizeof(uint64_t) * max_classes, st));
    if (n == 0) return TMOG_OK;
    const int grid = (int)std::min<int64_t>((n + 255) / 256, (int64_t)sm_count() * 8);
    label_class_id_kernel<<<grid, 256, 0, st>>>(label, n, classes, nclasses, class_ids,
                                                reinterpret_cast<unsigned long long*>(class_counts));
    TMOG_LAUNCH_CHECK();
    return TMOG_OK;
}

extern "C" int tmog_null_class_counts_dev(const uint64_t* valid, int64_t ldv, const uint8_t* class_ids, int64_t n,
                                          int ncols, int nclasses, int accumulate, uint64_t* counts, void* stream) {
    TMOG_REQUIRE(n >= 0 && ncols >= 0 && nclasses >= 1 && nclasses <= 255, "bad shape");
    TMOG_REQUIRE(counts != nullptr, "counts is NULL");
    cudaStream_t st = (cudaStream_t)stream;
    if (!accumulate) TMOG_CUDA(cudaMemsetAsync(counts, 0, sizeof(uint64_t) * (size_t)ncols * nclasses, st));
    if (valid == nullptr || n == 0 || ncols == 0) return TMOG_OK;  // no nulls anywhere
    TMOG_REQUIRE(class_ids != nullptr && ldv >= (n + 63) / 64, "class_ids is NULL or ldv too small");
    const int64_t ntiles = (n + 2047) / 2048;
    const size_t smem = sizeof(unsigned) * ((size_t)ncols * NC_FAST + 8 * (size_t)nclasses);
    TMOG_REQUIRE(smem <= 160 * 1024, "too many columns for one null-count call (%d): split the column block", ncols);
    if (smem > 40 * 1024)
        TMOG_CUDA(cudaFuncSetAttribute(null_class_count_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const int grid = (int)std::min<int64_t>(ntiles, (int64_t)sm_count() * 8);
    null_class_count_kernel<<<grid, 256, smem, st>>>(
        valid, ldv, class_ids, n, ncols, nclasses, reinterpret_cast<unsigned long long*>(counts));
    TMOG_LAUNCH_CHECK();
    return TMOG_OK;
}
