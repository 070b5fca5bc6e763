// IntegralVectorizer: the mode of every Integral column (sm_100a).
//
// Path restated here (reference file:line in include/tmog_b200.h): IntegralVectorizer.fitFn with
// SequenceAggregators.ModeSeqNullInt -- per column a map value -> count over the non-null rows, the mode being the
// most frequent value and, among equally frequent ones, the smallest; an all-null column gives 0.  Transform, column
// statistics and null indicators of Integral columns go through the kernels of real.cu (int64 read as double).
//
// HBM layout: column-major int64 values (stride ld), Arrow validity bitmaps.  Per column an open-addressing table in
// global memory (cap slots, a power of two: key u64 + count u64) that the caller provides as scratch.
// Roofline: HBM, 8 bytes + 1 bit per cell read once; the counting itself runs in a per-CTA shared-memory table
// (2048 slots) that absorbs the frequent values, so the global table sees one weighted insert per distinct value and
// CTA, plus the values that find no room in the shared table.
#include <algorithm>

#include "common.cuh"

namespace tmog {

constexpr unsigned long long IM_EMPTY = 0x8000000000000000ull;   // INT64_MIN: counted apart, so it can mark a free slot
constexpr int IM_THREADS = 256;
constexpr int IM_SLOTS = 2048;          // shared-memory table per CTA
constexpr int IM_SMEM_PROBES = 8;

__device__ __forceinline__ unsigned long long im_hash(unsigned long long v) {
    unsigned long long h = v * 0x9E3779B97F4A7C15ull;
    return h ^ (h >> 32);
}

struct IntegralTables {
    unsigned long long* keys;        // [ncols][cap]
    unsigned long long* counts;      // [ncols][cap]
    int* ndistinct;                  // [ncols]
    unsigned long long* min_count;   // [ncols] rows holding INT64_MIN
    int* overflow;                   // [ncols] set when a column's table ran full (its result is then void)
};

__global__ void integral_init_kernel(IntegralTables t, int ncols, int64_t cap) {
    const int64_t total = (int64_t)ncols * cap;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        t.keys[i] = IM_EMPTY;
        t.counts[i] = 0ull;
    }
    if (blockIdx.x == 0)
        for (int c = threadIdx.x; c < ncols; c += blockDim.x) {
            t.ndistinct[c] = 0;
            t.min_count[c] = 0ull;
            t.overflow[c] = 0;
        }
}

// count `v` (weight w) in column c's global table
__device__ __forceinline__ void im_global_add(const IntegralTables& t, int c, int64_t cap, unsigned long long v,
                                              unsigned long long w) {
    unsigned long long* keys = t.keys + (int64_t)c * cap;
    unsigned long long* counts = t.counts + (int64_t)c * cap;
    const int64_t limit = cap - (cap >> 2);     // the table never gets fuller than 3/4
    uint64_t slot = im_hash(v) & (uint64_t)(cap - 1);
    for (int64_t probe = 0; probe < cap; ++probe) {
        unsigned long long cur = keys[slot];
        if (cur == IM_EMPTY) {
            if (t.ndistinct[c] >= limit) break;
            cur = atomicCAS(&keys[slot], IM_EMPTY, v);
            if (cur == IM_EMPTY) {
                atomicAdd(&t.ndistinct[c], 1);
                cur = v;
            }
        }
        if (cur == v) {
            atomicAdd(&counts[slot], w);
            return;
        }
        slot = (slot + 1) & (uint64_t)(cap - 1);
    }
    t.overflow[c] = 1;
}

__global__ void __launch_bounds__(IM_THREADS)
integral_count_kernel(const long long* __restrict__ values, int64_t ld, const uint64_t* __restrict__ valid,
                      int64_t ldv, int64_t n, int64_t cap, IntegralTables t) {
    __shared__ unsigned long long skeys[IM_SLOTS];
    __shared__ unsigned int scnt[IM_SLOTS];
    const int c = blockIdx.y, tid = threadIdx.x;
    for (int i = tid; i < IM_SLOTS; i += IM_THREADS) {
        skeys[i] = IM_EMPTY;
        scnt[i] = 0u;
    }
    __syncthreads();
    const long long* col = values + (int64_t)c * ld;
    const uint64_t* vcol = valid ? valid + (int64_t)c * ldv : nullptr;
    unsigned long long nmin = 0;
    for (int64_t r = (int64_t)blockIdx.x * IM_THREADS + tid; r < n; r += (int64_t)gridDim.x * IM_THREADS) {
        if (vcol && !((vcol[r >> 6] >> (r & 63)) & 1ull)) continue;
        const unsigned long long v = (unsigned long long)__ldcs(col + r);
        if (v == IM_EMPTY) {
            ++nmin;
            continue;
        }
        unsigned slot = (unsigned)im_hash(v) & (IM_SLOTS - 1);
        bool done = false;
#pragma unroll 1
        for (int probe = 0; probe < IM_SMEM_PROBES && !done; ++probe) {
            unsigned long long cur = skeys[slot];
            if (cur == IM_EMPTY) {
                cur = atomicCAS(&skeys[slot], IM_EMPTY, v);
                if (cur == IM_EMPTY) cur = v;
            }
            if (cur == v) {
                atomicAdd(&scnt[slot], 1u);
                done = true;
            }
            slot = (slot + 1) & (IM_SLOTS - 1);
        }
        if (!done) im_global_add(t, c, cap, v, 1ull);
    }
    if (nmin) atomicAdd(&t.min_count[c], nmin);
    __syncthreads();
    for (int i = tid; i < IM_SLOTS; i += IM_THREADS)
        if (skeys[i] != IM_EMPTY) im_global_add(t, c, cap, skeys[i], (unsigned long long)scnt[i]);
}

// mode = the key with the largest count, the smallest such key on ties (keys compared as signed); 0 for an empty map
__global__ void __launch_bounds__(256)
integral_mode_kernel(IntegralTables t, int64_t cap, long long* __restrict__ modes,
                     unsigned long long* __restrict__ mode_counts) {
    __shared__ unsigned long long s_cnt[256];
    __shared__ long long s_key[256];
    const int c = blockIdx.x, tid = threadIdx.x;
    const unsigned long long* keys = t.keys + (int64_t)c * cap;
    const unsigned long long* counts = t.counts + (int64_t)c * cap;
    unsigned long long bc = 0;
    long long bk = 0;
    if (tid == 0 && t.min_count[c] > 0) {
        bc = t.min_count[c];
        bk = (long long)IM_EMPTY;
    }
    for (int64_t i = tid; i < cap; i += 256) {
        const unsigned long long k = keys[i];
        if (k == IM_EMPTY) continue;
        const unsigned long long w = counts[i];
        if (w > bc || (w == bc && w > 0 && (long long)k < bk)) {
            bc = w;
            bk = (long long)k;
        }
    }
    s_cnt[tid] = bc;
    s_key[tid] = bk;
    __syncthreads();
    for (int o = 128; o > 0; o >>= 1) {
        if (tid < o) {
            const unsigned long long w = s_cnt[tid + o];
            const long long k = s_key[tid + o];
            if (w > s_cnt[tid] || (w == s_cnt[tid] && w > 0 && k < s_key[tid])) {
                s_cnt[tid] = w;
                s_key[tid] = k;
            }
        }
        __syncthreads();
    }
    if (tid == 0) {
        modes[c] = s_cnt[0] > 0 ? s_key[0] : 0ll;
        mode_counts[c] = s_cnt[0];
    }
}

static size_t im_align(size_t x) { return (x + 255) & ~size_t(255); }

static IntegralTables im_layout(void* scratch, int ncols, int64_t cap) {
    IntegralTables t;
    unsigned char* p = static_cast<unsigned char*>(scratch);
    const size_t tab = im_align(sizeof(unsigned long long) * (size_t)ncols * (size_t)cap);
    t.keys = reinterpret_cast<unsigned long long*>(p);
    t.counts = reinterpret_cast<unsigned long long*>(p + tab);
    t.min_count = reinterpret_cast<unsigned long long*>(p + 2 * tab);
    t.ndistinct = reinterpret_cast<int*>(p + 2 * tab + im_align(8 * (size_t)ncols));
    t.overflow = reinterpret_cast<int*>(p + 2 * tab + im_align(8 * (size_t)ncols) + im_align(4 * (size_t)ncols));
    return t;
}

}  // namespace tmog

using namespace tmog;

extern "C" int tmog_integral_mode_scratch_bytes(int ncols, int64_t cap, size_t* bytes) {
    TMOG_REQUIRE(ncols >= 0 && bytes != nullptr && cap >= 64 && (cap & (cap - 1)) == 0,
                 "cap must be a power of two >= 64");
    *bytes = 2 * im_align(sizeof(unsigned long long) * (size_t)ncols * (size_t)cap) + im_align(8 * (size_t)ncols) +
             2 * im_align(4 * (size_t)ncols);
    return TMOG_OK;
}

extern "C" int tmog_integral_mode_dev(const int64_t* values, int64_t ld, const uint64_t* valid, int64_t ldv, int64_t n,
                                      int ncols, int64_t cap, void* scratch, size_t scratch_bytes, int64_t* modes,
                                      uint64_t* mode_counts, void* stream) {
    TMOG_REQUIRE(n >= 0 && ncols >= 0 && ld >= n, "bad shape");
    TMOG_REQUIRE(valid == nullptr || ldv >= (n + 63) / 64, "bitmap stride ldv (%lld) too small for n (%lld)",
                 (long long)ldv, (long long)n);
    size_t need = 0;
    if (int rc = tmog_integral_mode_scratch_bytes(ncols, cap, &need)) return rc;
    if (scratch == nullptr || scratch_bytes < need) {
        set_error("scratch too small: %zu bytes given, %zu needed (tmog_integral_mode_scratch_bytes)", scratch_bytes,
                  need);
        return TMOG_ENOMEM;
    }
    if (ncols == 0) return TMOG_OK;
    TMOG_REQUIRE((values != nullptr || n == 0) && modes != nullptr && mode_counts != nullptr, "NULL argument");
    cudaStream_t st = (cudaStream_t)stream;
    const IntegralTables t = im_layout(scratch, ncols, cap);
    integral_init_kernel<<<sm_count() * 4, 256, 0, st>>>(t, ncols, cap);
    TMOG_LAUNCH_CHECK();
    if (n > 0) {
        // enough CTAs per column to fill the chip, few enough that the per-CTA tables see many rows each
        const int64_t per_col = std::max<int64_t>(1, (int64_t)sm_count() * 8 / ncols);
        const int gx = (int)std::min<int64_t>((n + IM_THREADS - 1) / IM_THREADS, per_col);
        integral_count_kernel<<<dim3(gx, ncols), IM_THREADS, 0, st>>>(reinterpret_cast<const long long*>(values), ld,
                                                                      valid, ldv, n, cap, t);
        TMOG_LAUNCH_CHECK();
    }
    integral_mode_kernel<<<ncols, 256, 0, st>>>(t, cap, reinterpret_cast<long long*>(modes),
                                                reinterpret_cast<unsigned long long*>(mode_counts));
    TMOG_LAUNCH_CHECK();
    return TMOG_OK;
}

// Where the pieces of the scratch lie (for a caller that merges the value -> count maps of several GPUs itself):
// byte offsets of keys [ncols][cap] u64, counts [ncols][cap] u64, min_count [ncols] u64, ndistinct [ncols] i32,
// overflow [ncols] i32.
extern "C" int tmog_integral_mode_layout(int ncols, int64_t cap, size_t* offsets5) {
    TMOG_REQUIRE(offsets5 != nullptr && ncols >= 0 && cap >= 64, "bad arguments");
    const IntegralTables t = im_layout(nullptr, ncols, cap);
    offsets5[0] = reinterpret_cast<size_t>(t.keys);
    offsets5[1] = reinterpret_cast<size_t>(t.counts);
    offsets5[2] = reinterpret_cast<size_t>(t.min_count);
    offsets5[3] = reinterpret_cast<size_t>(t.ndistinct);
    offsets5[4] = reinterpret_cast<size_t>(t.overflow);
    return TMOG_OK;
}
