// Pivot vectorizers (OpTextPivotVectorizer / OpSetVectorizer on single-valued PickList / Text columns) on
// dictionary-encoded columns, fused with SanityChecker's per-label column sums.
//
// HBM layout: codes[ncols][ld] int32 column-major (code < 0 = null), one class-id byte per row, and per column a
// slot table slot[code] (int32) that maps a raw dictionary code to the id of its CLEANED value (raw values that clean
// alike share an id); the null "value" has id K_c (one past the last cleaned id).
//
//   fit        counts[col][id][class] (u64) -- vocabulary, top-K, contingency matrix, Cramer's V, column moments and
//              label correlations of every one-hot column all derive from this one table
//   transform  out[row][col] = output index of the row's value (top-K position, OTHER, or the null indicator),
//              row-major int32 (one-hot: the value 1.0 is implicit), -1 where SanityChecker dropped the column
//
// Roofline: HBM.  fit reads 4 bytes per cell + 1 byte per row; transform reads and writes 4 bytes per cell.
#include <algorithm>

#include "common.cuh"

namespace tmog {

constexpr int PC_THREADS = 1024;
constexpr int PC_SMEM_WORDS = 48 * 1024;  // 192 KB of u32 bins per CTA (1 CTA / SM)
constexpr int64_t PC_MIN_ROWS = 1 << 16;  // rows per work item (grows with the table so that flushes stay amortised)

struct PivotColumn {       // col_desc: 4 x int64 per column
    int64_t nids;          // cleaned ids incl. the null id (= nids - 1)
    int64_t slot_offset;   // into the concatenated slot tables (= sum of ncodes of the columns before)
    int64_t count_offset;  // into counts (in bins); this column's table is nids x nclasses
    int64_t ncodes;        // raw dictionary size (a code >= ncodes is an export error -> bad_code flag)
};

// One CTA per work item (column, row range); a shared-memory histogram holds the first `hot` ids of the column
// (ids are dictionary order: first-seen values, i.e. the frequent ones under any skewed distribution, come first),
// the cold tail goes to global atomics.  Shared bins are flushed once per work item.
__global__ void __launch_bounds__(PC_THREADS, 1)
pivot_count_kernel(const int32_t* __restrict__ codes, int64_t ld, int64_t n, int ncols,
                   const uint8_t* __restrict__ class_ids, int nclasses, const PivotColumn* __restrict__ cols,
                   const int32_t* __restrict__ slots, int64_t rows_per_item, int items_per_col,
                   unsigned long long* __restrict__ counts, int* __restrict__ bad_code) {
    extern __shared__ unsigned int hist[];
    const int tid = threadIdx.x;
    const int64_t nitems = (int64_t)ncols * items_per_col;
    for (int64_t item = blockIdx.x; item < nitems; item += gridDim.x) {
        const int c = (int)(item / items_per_col);
        const int64_t r0 = (item % items_per_col) * rows_per_item;
        const int64_t r1 = min(n, r0 + rows_per_item);
        if (r0 >= r1) continue;
        const PivotColumn pc = cols[c];
        const int hot = (int)min(pc.nids, (int64_t)(PC_SMEM_WORDS / nclasses));  // ids [0, hot): shared memory
        const int nbins = hot * nclasses;
        for (int i = tid; i < nbins; i += PC_THREADS) hist[i] = 0;
        __syncthreads();
        const int32_t* col = codes + (int64_t)c * ld;
        const int32_t* slot = slots + pc.slot_offset;
        unsigned long long* gcount = counts + pc.count_offset;
        const int null_id = (int)pc.nids - 1;
        const int ncodes = (int)pc.ncodes;
        // 4 rows per thread per iteration: one 128-bit code load + one 32-bit class-id load
        const int64_t v0 = (r0 + 3) & ~3ll;  // r0 is a multiple of rows_per_item (itself a multiple of 4)
        auto add4 = [&](const int4 cd, const unsigned cls4) {
            const int cc[4] = {cd.x, cd.y, cd.z, cd.w};
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                int id = null_id;
                if (cc[k] >= 0) {
                    if (cc[k] >= ncodes) { *bad_code = 1; continue; }
                    id = slot[cc[k]];
                }
                const int cls = (cls4 >> (8 * k)) & 0xff;
                if (id < hot) atomicAdd(&hist[id * nclasses + cls], 1u);
                else atomicAdd(&gcount[(int64_t)id * nclasses + cls], 1ull);
            }
        };
        int64_t r = v0 + 4ll * tid;
        for (; r + 4ll * PC_THREADS + 3 < r1; r += 8ll * PC_THREADS) {   // two 128-bit loads in flight per thread
            const int4 a = *reinterpret_cast<const int4*>(col + r);
            const int4 b = *reinterpret_cast<const int4*>(col + r + 4ll * PC_THREADS);
            const unsigned ca = nclasses > 1 ? *reinterpret_cast<const unsigned*>(class_ids + r) : 0u;
            const unsigned cb = nclasses > 1 ? *reinterpret_cast<const unsigned*>(class_ids + r + 4ll * PC_THREADS) : 0u;
            add4(a, ca);
            add4(b, cb);
        }
        for (; r + 3 < r1; r += 4ll * PC_THREADS) {
            const int4 a = *reinterpret_cast<const int4*>(col + r);
            const unsigned ca = nclasses > 1 ? *reinterpret_cast<const unsigned*>(class_ids + r) : 0u;
            add4(a, ca);
        }
        // tail rows (fewer than 4 left)
        const int64_t tail0 = v0 + ((r1 - v0) & ~3ll);
        for (int64_t r = tail0 + tid; r < r1; r += PC_THREADS) {
            const int code = col[r];
            int id = null_id;
            if (code >= 0) {
                if (code >= ncodes) { *bad_code = 1; continue; }
                id = slot[code];
            }
            const int cls = nclasses > 1 ? class_ids[r] : 0;
            if (id < hot) atomicAdd(&hist[id * nclasses + cls], 1u);
            else atomicAdd(&gcount[(int64_t)id * nclasses + cls], 1ull);
        }
        __syncthreads();
        for (int i = tid; i < nbins; i += PC_THREADS) {
            const unsigned v = hist[i];
            if (v) atomicAdd(&gcount[i], (unsigned long long)v);
        }
        __syncthreads();
    }
}

// out[r][c] = pos[c][code]  (pos tables hold the final output index -- column offset included -- or -1 when the
// column was dropped; entry ncodes_c of each table is the index for a null).  Tile = 64 rows x 32 columns through
// padded shared memory: 256-byte column runs in, 128-byte row runs out.
__global__ void __launch_bounds__(256)
pivot_transform_kernel(const int32_t* __restrict__ codes, int64_t ld, int64_t n, int ncols,
                       const PivotColumn* __restrict__ cols, const int32_t* __restrict__ pos,
                       int32_t* __restrict__ out, int64_t out_ld, int64_t out_col0) {
    __shared__ int tile[32][65];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int64_t r0 = (int64_t)blockIdx.x * 64;
    const int rows_here = (int)min((int64_t)64, n - r0);
    const int ntiles = (ncols + 31) / 32;
    for (int t = 0; t < ntiles; ++t) {
        const int c0 = t * 32;
        const int tc = min(32, ncols - c0);
        if (t > 0) __syncthreads();
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const int cl = warp + 8 * k;
            if (cl < tc) {
                const PivotColumn pc = cols[c0 + cl];
                const int32_t* col = codes + (int64_t)(c0 + cl) * ld + r0;
                const int32_t* table = pos + pc.slot_offset + (c0 + cl);  // each table has ncodes + 1 entries
#pragma unroll
                for (int h = 0; h < 2; ++h) {
                    const int rl = lane + 32 * h;
                    int v = -1;
                    if (rl < rows_here) {
                        const int code = col[rl];
                        v = table[code < 0 || code >= pc.ncodes ? pc.ncodes : (int64_t)code];
                    }
                    tile[cl][rl] = v;
                }
            }
        }
        __syncthreads();
        if (lane < tc) {
#pragma unroll
            for (int k = 0; k < 8; ++k) {
                const int rl = warp + 8 * k;
                if (rl < rows_here) out[(r0 + rl) * out_ld + out_col0 + c0 + lane] = tile[lane][rl];
            }
        }
    }
}

}  // namespace tmog

using namespace tmog;

static int64_t pivot_rows_per_item(int64_t n, int64_t max_nids, int nclasses) {
    int64_t rows = std::max<int64_t>(PC_MIN_ROWS, 8ll * std::min<int64_t>(max_nids * nclasses, PC_SMEM_WORDS));
    rows = (rows + 3) & ~3ll;
    return rows;
}

extern "C" int tmog_pivot_count_dev(const int32_t* codes, int64_t ld, int64_t n, int ncols, const uint8_t* class_ids,
                                    int nclasses, const int64_t* col_desc, const int32_t* slots, int64_t max_nids,
                                    uint64_t* counts, int* bad_code, void* stream) {
    TMOG_REQUIRE(n >= 0 && ncols >= 0 && ld >= n, "bad shape n=%lld ncols=%d ld=%lld", (long long)n, ncols,
                 (long long)ld);
    TMOG_REQUIRE(nclasses >= 1 && nclasses <= 255, "nclasses must be in [1, 255]");
    TMOG_REQUIRE((ld & 3) == 0 && (reinterpret_cast<uintptr_t>(codes) & 15u) == 0,
                 "code columns must be 16-byte aligned (base pointer and ld a multiple of 4)");
    TMOG_REQUIRE(nclasses == 1 || (reinterpret_cast<uintptr_t>(class_ids) & 3u) == 0, "class_ids must be 4-byte aligned");
    if (n == 0 || ncols == 0) return TMOG_OK;
    TMOG_REQUIRE(codes && col_desc && slots && counts && bad_code, "NULL argument");
    TMOG_REQUIRE(nclasses == 1 || class_ids != nullptr, "class_ids is NULL");
    const size_t smem = sizeof(unsigned) * PC_SMEM_WORDS;
    TMOG_CUDA(cudaFuncSetAttribute(pivot_count_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const int64_t rows_per_item = pivot_rows_per_item(n, max_nids, nclasses);
    const int items_per_col = (int)((n + rows_per_item - 1) / rows_per_item);
    const int64_t nitems = (int64_t)items_per_col * ncols;
    const int grid = (int)std::min<int64_t>(nitems, sm_count());
    pivot_count_kernel<<<grid, PC_THREADS, smem, (cudaStream_t)stream>>>(
        codes, ld, n, ncols, class_ids, nclasses, reinterpret_cast<const PivotColumn*>(col_desc), slots,
        rows_per_item, items_per_col, reinterpret_cast<unsigned long long*>(counts), bad_code);
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
    pivot_transform_kernel<<<(unsigned)gx, 256, 0, (cudaStream_t)stream>>>(
        codes, ld, n, ncols, reinterpret_cast<const PivotColumn*>(col_desc), pos, out, out_ld, out_col0);
    TMOG_LAUNCH_CHECK();
    return TMOG_OK;
}
