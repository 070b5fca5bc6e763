// Pivot vectorizers (OpTextPivotVectorizer / OpSetVectorizer on single-valued PickList / Text columns) on
// dictionary-encoded columns, fused with SanityChecker's per-label column sums.
//
// HBM layout: codes[ncols][ld] int32 column-major (code < 0 = null), one class-id byte per row, and per column a
// slot table slot[code] (int32) that maps a raw dictionary code to the id of its CLEANED value (raw values that clean
// alike share an id); value ids are 1..K_c and the null "value" has id 0, so that the usually frequent null bin
// always sits in the shared-memory part of the histogram.
//
//   fit        counts[col][id][class] (u32 bins per call: at most 2^32 - 1 rows; callers widen them to int64 when they
//              add up row chunks or ranks) -- vocabulary, top-K, contingency matrix, Cramer's V, column moments and
//              label correlations of every one-hot column all derive from this one table
//   transform  out[row][col] = output index of the row's value (top-K position, OTHER, or the null indicator),
//              row-major int32 (one-hot: the value 1.0 is implicit), -1 where SanityChecker dropped the column
//
// Roofline: HBM.  fit reads 4 bytes per cell + 1 byte per row; transform reads and writes 4 bytes per cell.
#include <stdlib.h>

#include <algorithm>

#include "common.cuh"

namespace tmog {

constexpr int PC_THREADS = 1024;
constexpr int PC_SMEM_WORDS = 48 * 1024;  // 192 KB of u32 bins per CTA (1 CTA / SM)
constexpr int64_t PC_MIN_ROWS = 1 << 16;  // rows per work item (grows with the table so that flushes stay amortised)

struct PivotColumn {       // col_desc: 4 x int64 per column
    int64_t nids;          // ids incl. the null id 0 (value ids are 1 .. nids - 1)
    int64_t slot_offset;   // into the concatenated slot tables (= sum of ncodes of the columns before)
    int64_t count_offset;  // into counts (in bins); this column's table is nids x nclasses
    int64_t ncodes;        // raw dictionary size (a code >= ncodes is an export error -> bad_code flag)
};

// Global-memory increment of a cold bin, aggregated over the lanes of the warp that hit the same bin: a value that is
// frequent but sits late in its dictionary would otherwise serialise the whole chip on one L2 address.
__device__ __forceinline__ void cold_add(unsigned int* gcount, int bin) {
    const unsigned active = __activemask();
    const unsigned peers = __match_any_sync(active, bin);
    if ((int)(__ffs(peers) - 1) == (int)(threadIdx.x & 31)) atomicAdd(&gcount[bin], (unsigned)__popc(peers));
}

// One element of the histogram update.  Branch-free on the common path: id = code + 1 with every negative code
// (null) mapped to 0; an id outside the table (a code the dictionary does not have) is skipped and flagged.
template <bool ALL_HOT, bool HAS_SLOT>
__device__ __forceinline__ void pivot_add(unsigned int* hist, unsigned int* gcount, const int32_t* slot, int code,
                                          int cls, int nids, int hot, int nclasses, int* bad_code) {
    int id = max(code, -1) + 1;
    if (HAS_SLOT && id > 0 && id < nids) id = 1 + slot[id - 1];
    const bool ok = (unsigned)id < (unsigned)nids;
    const int bin = id * nclasses + cls;
    if (ALL_HOT) {
        if (ok) atomicAdd(&hist[bin], 1u);
    } else {
        if (ok && id < hot) atomicAdd(&hist[bin], 1u);
        if (ok && id >= hot) cold_add(gcount, bin);
    }
    if (!ok) *bad_code = 1;
}

template <bool ALL_HOT, bool HAS_SLOT>
__device__ __forceinline__ void pivot_count_rows(unsigned int* hist, unsigned int* gcount, const int32_t* slot,
                                                 const int32_t* __restrict__ col, const uint8_t* __restrict__ class_ids,
                                                 int64_t r0, int64_t r1, int nids, int hot, int nclasses,
                                                 int* bad_code) {
    const int tid = threadIdx.x;
    constexpr int64_t STEP = 4ll * PC_THREADS;
    int64_t r = r0 + 4ll * tid;  // r0 is a multiple of 4: 128-bit code loads, 32-bit class-id loads
    for (; r + 3 * STEP + 3 < r1; r += 4 * STEP) {  // four 128-bit loads in flight per thread
        int4 cd[4];
        unsigned cl[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            cd[u] = __ldcs(reinterpret_cast<const int4*>(col + r + u * STEP));
            cl[u] = nclasses > 1 ? *reinterpret_cast<const unsigned*>(class_ids + r + u * STEP) : 0u;
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            pivot_add<ALL_HOT, HAS_SLOT>(hist, gcount, slot, cd[u].x, cl[u] & 0xff, nids, hot, nclasses, bad_code);
            pivot_add<ALL_HOT, HAS_SLOT>(hist, gcount, slot, cd[u].y, (cl[u] >> 8) & 0xff, nids, hot, nclasses, bad_code);
            pivot_add<ALL_HOT, HAS_SLOT>(hist, gcount, slot, cd[u].z, (cl[u] >> 16) & 0xff, nids, hot, nclasses, bad_code);
            pivot_add<ALL_HOT, HAS_SLOT>(hist, gcount, slot, cd[u].w, cl[u] >> 24, nids, hot, nclasses, bad_code);
        }
    }
    for (; r + 3 < r1; r += STEP) {
        const int4 a = __ldcs(reinterpret_cast<const int4*>(col + r));
        const unsigned ca = nclasses > 1 ? *reinterpret_cast<const unsigned*>(class_ids + r) : 0u;
        pivot_add<ALL_HOT, HAS_SLOT>(hist, gcount, slot, a.x, ca & 0xff, nids, hot, nclasses, bad_code);
        pivot_add<ALL_HOT, HAS_SLOT>(hist, gcount, slot, a.y, (ca >> 8) & 0xff, nids, hot, nclasses, bad_code);
        pivot_add<ALL_HOT, HAS_SLOT>(hist, gcount, slot, a.z, (ca >> 16) & 0xff, nids, hot, nclasses, bad_code);
        pivot_add<ALL_HOT, HAS_SLOT>(hist, gcount, slot, a.w, ca >> 24, nids, hot, nclasses, bad_code);
    }
    // tail rows (fewer than 4 left)
    const int64_t tail0 = r0 + ((r1 - r0) & ~3ll);
    for (int64_t q = tail0 + tid; q < r1; q += PC_THREADS)
        pivot_add<ALL_HOT, HAS_SLOT>(hist, gcount, slot, col[q], nclasses > 1 ? class_ids[q] : 0, nids, hot,
                                     nclasses, bad_code);
}

// One CTA per work item (column, row range); a shared-memory histogram holds the first `hot` ids of the column
// (id 0 = null, then dictionary order: first-seen values, i.e. the frequent ones under a skewed distribution),
// the cold tail goes to global atomics.  Shared bins are flushed once per work item.
__global__ void __launch_bounds__(PC_THREADS, 1)
pivot_count_kernel(const int32_t* __restrict__ codes, int64_t ld, int64_t n, int ncols,
                   const uint8_t* __restrict__ class_ids, int nclasses, const PivotColumn* __restrict__ cols,
                   const int32_t* __restrict__ slots, int64_t rows_per_item, int items_per_col,
                   unsigned int* __restrict__ counts, int* __restrict__ bad_code, int hot_words) {
    extern __shared__ unsigned int hist[];
    const int tid = threadIdx.x;
    const int64_t nitems = (int64_t)ncols * items_per_col;
    for (int64_t item = blockIdx.x; item < nitems; item += gridDim.x) {
        const int c = (int)(item / items_per_col);
        const int64_t r0 = (item % items_per_col) * rows_per_item;  // a multiple of 4
        const int64_t r1 = min(n, r0 + rows_per_item);
        if (r0 >= r1) continue;
        const PivotColumn pc = cols[c];
        const int nids = (int)pc.nids;
        const int hot = min(nids, hot_words / nclasses);  // ids [0, hot): shared memory
        const int nbins = hot * nclasses;
        for (int i = tid; i < nbins; i += PC_THREADS) hist[i] = 0;
        __syncthreads();
        const int32_t* col = codes + (int64_t)c * ld;
        const int32_t* slot = slots ? slots + pc.slot_offset : nullptr;
        unsigned int* gcount = counts + pc.count_offset;
        if (slot == nullptr) {
            if (hot == nids)
                pivot_count_rows<true, false>(hist, gcount, slot, col, class_ids, r0, r1, nids, hot, nclasses, bad_code);
            else
                pivot_count_rows<false, false>(hist, gcount, slot, col, class_ids, r0, r1, nids, hot, nclasses, bad_code);
        } else {
            pivot_count_rows<false, true>(hist, gcount, slot, col, class_ids, r0, r1, nids, hot, nclasses, bad_code);
        }
        __syncthreads();
        for (int i = tid; i < nbins; i += PC_THREADS) {
            const unsigned v = hist[i];
            if (v) atomicAdd(&gcount[i], v);
        }
        __syncthreads();
    }
}

// out[r][c] = pos[c][code]  (pos tables hold the final output index -- column offset included -- or -1 when the
// column was dropped; entry ncodes_c of each table is the index for a null).  A CTA owns 64 rows and walks the column
// tiles; per tile every thread first issues its 8 independent code loads, then the 8 dependent table look-ups, then
// fills the padded shared tile; rows leave as contiguous 128-byte runs.
constexpr int PT_MAX_COLS = 4096;

__global__ void __launch_bounds__(256)
pivot_transform_kernel(const int32_t* __restrict__ codes, int64_t ld, int64_t n, int ncols,
                       const PivotColumn* __restrict__ cols, const int32_t* __restrict__ pos,
                       int32_t* __restrict__ out, int64_t out_ld, int64_t out_col0) {
    extern __shared__ long long dyn[];
    __shared__ int tile[32][65];
    long long* s_tab = dyn;                                   // [ncols] start of the column's pos table
    int* s_nc = reinterpret_cast<int*>(dyn + ncols);          // [ncols] dictionary size
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    for (int c = tid; c < ncols; c += 256) {
        s_tab[c] = cols[c].slot_offset + c;                   // every table has ncodes + 1 entries
        s_nc[c] = (int)cols[c].ncodes;
    }
    __syncthreads();
    const int64_t r0 = (int64_t)blockIdx.x * 64;
    const int rows_here = (int)min((int64_t)64, n - r0);
    const int ntiles = (ncols + 31) / 32;
    for (int t = 0; t < ntiles; ++t) {
        const int c0 = t * 32;
        const int tc = min(32, ncols - c0);
        int code[4][2];
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const int cl = warp + 8 * k;
            const int32_t* col = codes + (int64_t)(c0 + min(cl, tc - 1)) * ld + r0;
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                const int rl = lane + 32 * h;
                code[k][h] = (cl < tc && rl < rows_here) ? __ldcs(col + rl) : -1;
            }
        }
        int v[4][2];
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const int cl = min(warp + 8 * k, tc - 1);
            const int nc = s_nc[c0 + cl];
            const int32_t* table = pos + s_tab[c0 + cl];
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                const unsigned idx = (unsigned)code[k][h] < (unsigned)nc ? (unsigned)code[k][h] : (unsigned)nc;
                v[k][h] = __ldg(table + idx);
            }
        }
        if (t > 0) __syncthreads();  // previous tile fully written out before the shared tile is overwritten
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const int cl = warp + 8 * k;
            if (cl < tc) {
                tile[cl][lane] = v[k][0];
                tile[cl][lane + 32] = v[k][1];
            }
        }
        __syncthreads();
        if (lane < tc) {
            int32_t* optr = out + (r0 + warp) * out_ld + out_col0 + c0 + lane;
            const int64_t ostep = 8 * out_ld;
#pragma unroll
            for (int k = 0; k < 8; ++k) {
                const int rl = warp + 8 * k;
                if (rl < rows_here) *optr = tile[lane][rl];
                optr += ostep;
            }
        }
    }
}


// Continuous label (more distinct values than label classes are tracked): one f64 sum per id,
//   wsum [wsum_offset_c + id] += label[r] - shift   for the row's id (0 = null, 1 + slot(code)),
// which together with the plain counts gives the covariance of every one-hot / null-indicator column with the label
// (U/stats/OpStatistics.scala:86-90: sum (x - mean_x)(y - mean_y) = sum_x (y - s) - count (mean_y - s)).
// A CTA takes a stretch of rows of one column; the first PL_HOT ids live in shared memory, the rest go to global
// memory aggregated over the lanes that hit the same id.
constexpr int PL_THREADS = 256;
constexpr int PL_HOT = 4096;
constexpr int PL_ROWS = 1 << 16;

__global__ void __launch_bounds__(PL_THREADS)
pivot_label_sum_kernel(const int32_t* __restrict__ codes, int64_t ld, int64_t n, const PivotColumn* __restrict__ cols,
                       const int32_t* __restrict__ slots, const int64_t* __restrict__ wsum_offset,
                       const double* __restrict__ label, double shift, double* __restrict__ wsum) {
    __shared__ double hot[PL_HOT];
    const int c = blockIdx.y;
    const PivotColumn pc = cols[c];
    const int32_t* col = codes + (int64_t)c * ld;
    const int32_t* sl = slots ? slots + pc.slot_offset : nullptr;
    double* g = wsum + wsum_offset[c];
    const int nhot = (int)min((int64_t)PL_HOT, pc.nids);
    for (int i = threadIdx.x; i < nhot; i += PL_THREADS) hot[i] = 0.0;
    __syncthreads();
    const int64_t r0 = (int64_t)blockIdx.x * PL_ROWS, r1 = min(n, r0 + PL_ROWS);
    for (int64_t rb = r0; rb < r1; rb += PL_THREADS) {
        const int64_t r = rb + threadIdx.x;
        int id = -1;
        double e = 0.0;
        if (r < r1) {
            const int code = col[r];
            if (code < 0) id = 0;
            else if (code < pc.ncodes) id = 1 + (sl ? sl[code] : code);
            e = label[r] - shift;
        }
        if (id >= 0 && id < nhot) {
            atomicAdd(&hot[id], e);
        } else {
            // cold ids: one global atomic per distinct id of the warp
            const unsigned peers = __match_any_sync(__activemask(), id);
            if (id >= 0) {
                double sum = 0.0;
                for (unsigned m = peers; m; m &= m - 1) sum += __shfl_sync(peers, e, __ffs(m) - 1);
                if ((threadIdx.x & 31) == __ffs(peers) - 1) atomicAdd(&g[id], sum);
            }
        }
    }
    __syncthreads();
    for (int i = threadIdx.x; i < nhot; i += PL_THREADS)
        if (hot[i] != 0.0) atomicAdd(&g[i], hot[i]);
}

}  // namespace tmog

using namespace tmog;

static int env_int(const char* name, int dflt) {
    const char* v = getenv(name);
    return v ? atoi(v) : dflt;
}

static int64_t pivot_rows_per_item(int64_t n, int64_t max_nids, int nclasses, int hot_words) {
    const int factor = env_int("TMOG_PIVOT_ROWS_FACTOR", 8);
    int64_t rows = std::max<int64_t>(PC_MIN_ROWS, (int64_t)factor * std::min<int64_t>(max_nids * nclasses, hot_words));
    rows = (rows + 3) & ~3ll;
    return rows;
}

extern "C" int tmog_pivot_count_dev(const int32_t* codes, int64_t ld, int64_t n, int ncols, const uint8_t* class_ids,
                                    int nclasses, const int64_t* col_desc, const int32_t* slots, int64_t max_nids,
                                    uint32_t* counts, int* bad_code, void* stream) {
    TMOG_REQUIRE(n >= 0 && ncols >= 0 && ld >= n, "bad shape n=%lld ncols=%d ld=%lld", (long long)n, ncols,
                 (long long)ld);
    TMOG_REQUIRE(n < (1ll << 32), "at most 2^32 - 1 rows per call (32-bit bins): chunk the batch");
    TMOG_REQUIRE(nclasses >= 1 && nclasses <= 255, "nclasses must be in [1, 255]");
    TMOG_REQUIRE((ld & 3) == 0 && (reinterpret_cast<uintptr_t>(codes) & 15u) == 0,
                 "code columns must be 16-byte aligned (base pointer and ld a multiple of 4)");
    TMOG_REQUIRE(nclasses == 1 || (reinterpret_cast<uintptr_t>(class_ids) & 3u) == 0, "class_ids must be 4-byte aligned");
    if (n == 0 || ncols == 0) return TMOG_OK;
    TMOG_REQUIRE(codes && col_desc && counts && bad_code, "NULL argument");
    TMOG_REQUIRE(nclasses == 1 || class_ids != nullptr, "class_ids is NULL");
    const size_t smem = sizeof(unsigned) * PC_SMEM_WORDS;
    TMOG_CUDA(cudaFuncSetAttribute(pivot_count_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const int hot_words = std::min(PC_SMEM_WORDS, std::max(nclasses, env_int("TMOG_PIVOT_HOT_WORDS", PC_SMEM_WORDS)));
    const int64_t rows_per_item = pivot_rows_per_item(n, max_nids, nclasses, hot_words);
    const int items_per_col = (int)((n + rows_per_item - 1) / rows_per_item);
    const int64_t nitems = (int64_t)items_per_col * ncols;
    const int grid = (int)std::min<int64_t>(nitems, sm_count());
    pivot_count_kernel<<<grid, PC_THREADS, smem, (cudaStream_t)stream>>>(
        codes, ld, n, ncols, class_ids, nclasses, reinterpret_cast<const PivotColumn*>(col_desc), slots,
        rows_per_item, items_per_col, counts, bad_code, hot_words);
    TMOG_LAUNCH_CHECK();
    return TMOG_OK;
}

extern "C" int tmog_pivot_transform_dev(const int32_t* codes, int64_t ld, int64_t n, int ncols,
                                        const int64_t* col_desc, const int32_t* pos, int32_t* out, int64_t out_ld,
                                        int64_t out_col0, void* stream) {
    TMOG_REQUIRE(n >= 0 && ncols >= 0 && ld >= n, "bad shape");
    TMOG_REQUIRE(out_col0 >= 0 && out_ld >= out_col0 + ncols, "output block exceeds the row stride");
    if (n == 0 || ncols == 0) return TMOG_OK;
    TMOG_REQUIRE(codes && col_desc && pos && out, "NULL argument");
    const int64_t gx = (n + 63) / 64;
    TMOG_REQUIRE(gx <= 0x7fffffffLL, "too many rows for one transform call");
    TMOG_REQUIRE(ncols <= PT_MAX_COLS, "at most %d columns per transform call: split the block", PT_MAX_COLS);
    const size_t smem = (size_t)ncols * (sizeof(long long) + sizeof(int));
    pivot_transform_kernel<<<(unsigned)gx, 256, smem, (cudaStream_t)stream>>>(
        codes, ld, n, ncols, reinterpret_cast<const PivotColumn*>(col_desc), pos, out, out_ld, out_col0);
    TMOG_LAUNCH_CHECK();
    return TMOG_OK;
}

// Binary columns as dictionary codes (-1 null, 0 false, 1 true), so that their SanityChecker statistics come out of
// the same count[code][class] table as the pivoted columns.
namespace tmog {
__global__ void __launch_bounds__(256)
binary_codes_kernel(const uint8_t* __restrict__ values, int64_t ld, const uint64_t* __restrict__ valid, int64_t ldv,
                    int64_t n, int32_t* __restrict__ codes, int64_t ldc) {
    const int c = blockIdx.y;
    const uint8_t* col = values + (int64_t)c * ld;
    const uint64_t* vcol = valid ? valid + (int64_t)c * ldv : nullptr;
    int32_t* out = codes + (int64_t)c * ldc;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; r < n; r += stride) {
        const bool ok = vcol ? ((vcol[r >> 6] >> (r & 63)) & 1ull) != 0 : true;
        out[r] = ok ? (col[r] != 0 ? 1 : 0) : -1;
    }
}
}  // namespace tmog

extern "C" int tmog_binary_codes_dev(const uint8_t* values, int64_t ld, const uint64_t* valid, int64_t ldv, int64_t n,
                                     int ncols, int32_t* codes, int64_t ldc, void* stream) {
    TMOG_REQUIRE(n >= 0 && ncols >= 0 && ld >= n && ldc >= n, "bad shape");
    TMOG_REQUIRE(valid == nullptr || ldv >= (n + 63) / 64, "bitmap stride too small");
    if (n == 0 || ncols == 0) return TMOG_OK;
    TMOG_REQUIRE(values && codes, "NULL argument");
    const int gx = (int)std::min<int64_t>((n + 255) / 256, (int64_t)tmog::sm_count() * 8);
    tmog::binary_codes_kernel<<<dim3(gx, ncols), 256, 0, (cudaStream_t)stream>>>(values, ld, valid, ldv, n, codes, ldc);
    TMOG_LAUNCH_CHECK();
    return TMOG_OK;
}

extern "C" int tmog_pivot_label_sum_dev(const int32_t* codes, int64_t ld, int64_t n, int ncols, const int64_t* col_desc,
                                        const int32_t* slots, const int64_t* wsum_offset, const double* label,
                                        double shift, double* wsum, void* stream) {
    TMOG_REQUIRE(n >= 0 && ncols >= 0 && ld >= n, "bad shape");
    TMOG_REQUIRE(ncols == 0 || n == 0 || (codes && col_desc && wsum_offset && label && wsum), "NULL argument");
    if (n == 0 || ncols == 0) return TMOG_OK;
    const int64_t gx = (n + PL_ROWS - 1) / PL_ROWS;
    TMOG_REQUIRE(gx <= 0x7fffffffLL && ncols <= 65535, "too many rows or columns for one call");
    pivot_label_sum_kernel<<<dim3((unsigned)gx, (unsigned)ncols), PL_THREADS, 0, (cudaStream_t)stream>>>(
        codes, ld, n, reinterpret_cast<const PivotColumn*>(col_desc), slots, wsum_offset, label, shift, wsum);
    TMOG_LAUNCH_CHECK();
    return TMOG_OK;
}
