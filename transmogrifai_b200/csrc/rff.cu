// RawFeatureFilter's scans (C/filters/RawFeatureFilter.scala:138-200): the reference maps every row to a
// PreparedFeatures (PreparedFeatures.scala:156-193), reduces per-feature Summaries (Summary.scala:43-65), then maps
// every row again to one FeatureDistribution per feature (FeatureDistribution.scala:217-248, histValues :307-351) and
// reduces those, and finally correlates every predictor's null indicator with the response (RawFeatureFilter.scala:
// 181-191).  Here each of the three is one HBM-bound column scan:
//   * rff_numeric_summary_kernel   min / max / sum / count of the valid values of every numeric column
//   * rff_numeric_hist_kernel      the `bins` histogram of every numeric column over the splits of the TRAINING summary
//                                  (NumericBucketizer.bucketize, Inclusion.Left, trackInvalid: NumericBucketizer.scala:219-264)
//   * rff_rows_hist_kernel         hashed-token distributions: the (bucket, term frequency) rows that the tokenise +
//                                  hash kernels of text.cu wrote are summed per bucket, with the per-row token counts
//                                  (the text Summary); optional row weights serve dictionary-encoded columns
//   * rff_code_count_kernel        rows per dictionary code (PickList columns)
//   * rff_null_label_kernel        sum of the response over every predictor's NULL rows (+ the response's own moments)
// Layouts are the package's: values [ncols][ld] column-major, validity = Arrow bitmap words [ncols][ldv] (bit = valid).
// Grids are sized in multiples of the SM count; every thread block works on one column (blockIdx.y) so that the
// shared-memory histogram of a block belongs to one feature.
#include <algorithm>
#include <cmath>

#include "common.cuh"

namespace tmog {

namespace {

constexpr int RFF_THREADS = 256;
constexpr int RFF_SMEM_BINS = 2048;        // up to here splits (f64) and counts (u32) of a column live in shared memory

template <int DT> struct RffValue;
template <> struct RffValue<0> { typedef double type; };
template <> struct RffValue<1> { typedef long long type; };
template <> struct RffValue<2> { typedef unsigned char type; };

__device__ __forceinline__ bool rff_valid(const unsigned long long* v, int64_t r) {
    return v == nullptr || ((v[r >> 6] >> (r & 63)) & 1ull);
}

__device__ __forceinline__ void atomic_min_f64(double* addr, double v) {
    unsigned long long* a = reinterpret_cast<unsigned long long*>(addr);
    unsigned long long old = *a;
    while (v < __longlong_as_double((long long)old)) {
        const unsigned long long seen = atomicCAS(a, old, (unsigned long long)__double_as_longlong(v));
        if (seen == old) break;
        old = seen;
    }
}
__device__ __forceinline__ void atomic_max_f64(double* addr, double v) {
    unsigned long long* a = reinterpret_cast<unsigned long long*>(addr);
    unsigned long long old = *a;
    while (v > __longlong_as_double((long long)old)) {
        const unsigned long long seen = atomicCAS(a, old, (unsigned long long)__double_as_longlong(v));
        if (seen == old) break;
        old = seen;
    }
}

// ---- staged reads -----------------------------------------------------------------------------------------------------
// The plain kernels below leave it to the compiler how many loads are in flight, and it keeps one or two (it sinks
// every load next to its use).  Here a thread copies RFF_UNROLL 16-byte pieces of its block's stretch into its own
// shared-memory slots with cp.async (issued back to back, no register waits on them), waits for the group, and
// consumes them: 64 bytes in flight per thread whatever the scheduler does.  Needs 16-byte aligned columns
// (values % 16 == 0 and ld * sizeof(T) % 16 == 0: the package's layout); anything else takes the plain kernels.
constexpr int RFF_UNROLL = 4;

__device__ __forceinline__ void rff_cp_async16(void* smem, const void* gmem) {
    const unsigned s = (unsigned)__cvta_generic_to_shared(smem);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(s), "l"(gmem) : "memory");
}
__device__ __forceinline__ void rff_cp_async_wait_all() {
    asm volatile("cp.async.commit_group;" ::: "memory");
    asm volatile("cp.async.wait_group 0;" ::: "memory");
}

// A block's rows: a contiguous stretch whose start is a multiple of 64 (so that a 16-byte piece never straddles a
// validity word and piece / word indices are 32-bit).  -> rows of this block (0 for blocks beyond the end)
__device__ __forceinline__ int rff_stretch(int64_t n, int64_t* r0) {
    const int64_t per = (((n + gridDim.x - 1) / gridDim.x) + 63) & ~(int64_t)63;
    *r0 = (int64_t)blockIdx.x * per;
    const int64_t left = n - *r0;
    return left <= 0 ? 0 : (int)(left < per ? left : per);
}

// f(value as double) for every valid row of the block's stretch
template <typename T, typename F>
__device__ __forceinline__ void rff_for_each_valid(const T* __restrict__ x, const unsigned long long* __restrict__ v,
                                                   int64_t n, F f) {
    constexpr int EPC = 16 / (int)sizeof(T);                          // elements per 16-byte piece
    __shared__ __align__(16) unsigned char stage[RFF_UNROLL][RFF_THREADS][16];
    int64_t r0;
    const int len = rff_stretch(n, &r0);
    const unsigned char* xb = reinterpret_cast<const unsigned char*>(x + r0);
    const unsigned long long* vb = v ? v + (r0 >> 6) : nullptr;
    const int pieces = (len + EPC - 1) / EPC;      // the last piece may read past `len`: still inside the column (ld)
    for (int p0 = 0; p0 < pieces; p0 += RFF_UNROLL * RFF_THREADS) {
#pragma unroll
        for (int u = 0; u < RFF_UNROLL; ++u) {
            const int p = p0 + u * RFF_THREADS + threadIdx.x;
            if (p < pieces) rff_cp_async16(&stage[u][threadIdx.x][0], xb + (size_t)p * 16);
        }
        rff_cp_async_wait_all();
#pragma unroll
        for (int u = 0; u < RFF_UNROLL; ++u) {
            const int p = p0 + u * RFF_THREADS + threadIdx.x;
            if (p >= pieces) continue;
            const int i0 = p * EPC;
            const unsigned long long w = vb ? vb[i0 >> 6] : ~0ull;
            const T* e = reinterpret_cast<const T*>(&stage[u][threadIdx.x][0]);
#pragma unroll
            for (int k = 0; k < EPC; ++k) {
                const int i = i0 + k;
                if (i < len && ((w >> (i & 63)) & 1ull)) f((double)e[k]);
            }
        }
    }
}

// out [ncols][5]: min, max, sum, count, NaN seen (Summary.monoid uses java Math.min / max: a NaN value makes both NaN)
template <int DT>
__global__ void __launch_bounds__(RFF_THREADS)
rff_numeric_summary_kernel(const void* __restrict__ values, int64_t ld, const unsigned long long* __restrict__ valid,
                           int64_t ldv, int64_t n, double* __restrict__ out) {
    typedef typename RffValue<DT>::type T;
    const int col = blockIdx.y;
    const T* x = static_cast<const T*>(values) + (int64_t)col * ld;
    const unsigned long long* v = valid ? valid + (int64_t)col * ldv : nullptr;
    double mn = INFINITY, mx = -INFINITY, sum = 0.0, cnt = 0.0, nan = 0.0;
    // a block's rows: a contiguous stretch; four loads in flight per thread (HBM latency x bandwidth needs ~40 KB per SM)
    const int64_t per = (n + gridDim.x - 1) / gridDim.x;
    const int64_t r0 = (int64_t)blockIdx.x * per, r1 = r0 + per < n ? r0 + per : n;
    for (int64_t base = r0 + threadIdx.x; base < r1; base += 4 * RFF_THREADS) {
        T raw[4];
        bool ok[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const int64_t r = base + (int64_t)u * RFF_THREADS;
            ok[u] = r < r1 && rff_valid(v, r);
            raw[u] = ok[u] ? x[r] : T(0);
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            if (!ok[u]) continue;
            const double d = (double)raw[u];
            if (d != d) nan = 1.0;
            mn = fmin(mn, d);
            mx = fmax(mx, d);
            sum += d;
            cnt += 1.0;
        }
    }
    mn = warp_min(mn);
    mx = warp_max(mx);
    sum = warp_sum(sum);
    cnt = warp_sum(cnt);
    nan = warp_max(nan);
    __shared__ double s[RFF_THREADS / 32][5];
    const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (lane == 0) { s[w][0] = mn; s[w][1] = mx; s[w][2] = sum; s[w][3] = cnt; s[w][4] = nan; }
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int i = 1; i < RFF_THREADS / 32; ++i) {
            mn = fmin(mn, s[i][0]); mx = fmax(mx, s[i][1]); sum += s[i][2]; cnt += s[i][3]; nan = fmax(nan, s[i][4]);
        }
        if (cnt > 0.0) {
            double* o = out + (int64_t)col * 5;
            atomic_min_f64(o + 0, mn);
            atomic_max_f64(o + 1, mx);
            atomicAdd(o + 2, sum);
            atomicAdd(o + 3, cnt);
            if (nan > 0.0) atomic_max_f64(o + 4, 1.0);
        }
    }
}

template <int DT>
__global__ void __launch_bounds__(RFF_THREADS)
rff_numeric_summary_staged_kernel(const void* __restrict__ values, int64_t ld, const unsigned long long* __restrict__ valid,
                                  int64_t ldv, int64_t n, double* __restrict__ out) {
    typedef typename RffValue<DT>::type T;
    const int col = blockIdx.y;
    const T* x = static_cast<const T*>(values) + (int64_t)col * ld;
    const unsigned long long* v = valid ? valid + (int64_t)col * ldv : nullptr;
    double mn = INFINITY, mx = -INFINITY, sum = 0.0, cnt = 0.0, nan = 0.0;
    rff_for_each_valid<T>(x, v, n, [&](double d) {
        if (d != d) nan = 1.0;
        mn = fmin(mn, d);
        mx = fmax(mx, d);
        sum += d;
        cnt += 1.0;
    });
    mn = warp_min(mn);
    mx = warp_max(mx);
    sum = warp_sum(sum);
    cnt = warp_sum(cnt);
    nan = warp_max(nan);
    __shared__ double s[RFF_THREADS / 32][5];
    const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (lane == 0) { s[w][0] = mn; s[w][1] = mx; s[w][2] = sum; s[w][3] = cnt; s[w][4] = nan; }
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int i = 1; i < RFF_THREADS / 32; ++i) {
            mn = fmin(mn, s[i][0]); mx = fmax(mx, s[i][1]); sum += s[i][2]; cnt += s[i][3]; nan = fmax(nan, s[i][4]);
        }
        if (cnt > 0.0) {
            double* o = out + (int64_t)col * 5;
            atomic_min_f64(o + 0, mn);
            atomic_max_f64(o + 1, mx);
            atomicAdd(o + 2, sum);
            atomicAdd(o + 3, cnt);
            if (nan > 0.0) atomic_max_f64(o + 4, 1.0);
        }
    }
}

// mode [col]: 0 = the training summary is empty (FeatureDistribution.scala:329-330: the "distribution" of a row is its
//   value: value_sum [col] += v), 1 = min < max (histogram over splits [col][bins]), 2 = otherwise (:345-348:
//   hist [col][0] += (v == same_value [col]), hist [col][1] += (v != same_value [col])).
// Bucket of v in mode 1 (NumericBucketizer.scala:231-258): the i with splits[i] <= v < splits[i + 1], i < bins - 1;
//   none, NaN or +-Inf -> bins - 1.  The splits are non-decreasing, so i is the last split <= v (binary search).
template <int DT, bool SMEM, bool STAGED>
__global__ void __launch_bounds__(RFF_THREADS)
rff_numeric_hist_kernel(const void* __restrict__ values, int64_t ld, const unsigned long long* __restrict__ valid,
                        int64_t ldv, int64_t n, int bins, const double* __restrict__ splits,
                        const int* __restrict__ mode, const double* __restrict__ same_value,
                        unsigned long long* __restrict__ hist, double* __restrict__ value_sum,
                        unsigned long long* __restrict__ nonnull) {
    typedef typename RffValue<DT>::type T;
    extern __shared__ __align__(16) unsigned char rff_smem[];
    double* s_splits = reinterpret_cast<double*>(rff_smem);
    unsigned* s_hist = reinterpret_cast<unsigned*>(rff_smem + (size_t)(SMEM ? bins : 0) * sizeof(double));
    __shared__ unsigned long long s_cnt;
    __shared__ double s_vsum[RFF_THREADS / 32];
    const int col = blockIdx.y;
    const int md = mode[col];
    const T* x = static_cast<const T*>(values) + (int64_t)col * ld;
    const unsigned long long* v = valid ? valid + (int64_t)col * ldv : nullptr;
    const double* sp = splits + (int64_t)col * bins;
    unsigned long long* gh = hist + (int64_t)col * bins;
    if (threadIdx.x == 0) s_cnt = 0ull;
    if (SMEM) {
        for (int i = threadIdx.x; i < bins; i += blockDim.x) {
            s_splits[i] = sp[i];
            s_hist[i] = 0u;
        }
    }
    __syncthreads();
    const double* lookup = SMEM ? s_splits : sp;
    const double same = same_value[col];
    unsigned cnt = 0;
    double vsum = 0.0;
    // a block's rows: a contiguous stretch (at most 2^32 - 1 rows, so that the shared u32 counts cannot wrap)
    const int64_t per = (n + gridDim.x - 1) / gridDim.x;
    const int64_t r0 = (int64_t)blockIdx.x * per, r1 = r0 + per < n ? r0 + per : n;
    // the bucket is guessed from (d - min) / step and corrected against the splits themselves (the splits decide, so
    // the result is the reference's); a column whose step is not a positive finite number is searched instead
    const double lo0 = lookup[0];
    const double inv = md == 1 ? (double)(bins - 2) / (lookup[bins - 2 > 0 ? bins - 2 : 0] - lo0) : 0.0;
    const bool guess = md == 1 && bins > 2 && inv > 0.0 && !isinf(inv) && !isinf(lo0) && lo0 == lo0;
    auto count = [&](double d) {
        ++cnt;
        int b;
        if (md == 1) {
            if (d != d || isinf(d)) {
                b = bins - 1;
            } else if (guess) {
                const double q = (d - lo0) * inv;
                int g = q > 0.0 ? (q < (double)(bins - 1) ? (int)q : bins - 1) : 0;
                while (g > 0 && d < lookup[g]) --g;
                while (g < bins - 1 && lookup[g + 1] <= d) ++g;
                b = (g >= bins - 1 || d < lookup[g]) ? bins - 1 : g;
            } else {
                int lo = 0, hi = bins;                     // first index with splits[i] > d (or not <= d: NaN)
                while (lo < hi) {
                    const int mid = (lo + hi) >> 1;
                    if (lookup[mid] <= d) lo = mid + 1; else hi = mid;
                }
                b = lo - 1;
                if (b < 0 || b >= bins - 1) b = bins - 1;
            }
        } else if (md == 2) {
            b = d == same ? 0 : 1;
        } else {
            vsum += d;
            return;
        }
        if (SMEM) atomicAdd(&s_hist[b], 1u); else atomicAdd(&gh[b], 1ull);
    };
    if (STAGED) {
        rff_for_each_valid<T>(x, v, n, count);
    } else {
        for (int64_t base = r0 + threadIdx.x; base < r1; base += 4 * RFF_THREADS) {
            T raw[4];
            bool ok[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const int64_t r = base + (int64_t)u * RFF_THREADS;
                ok[u] = r < r1 && rff_valid(v, r);
                raw[u] = ok[u] ? x[r] : T(0);
            }
#pragma unroll
            for (int u = 0; u < 4; ++u)
                if (ok[u]) count((double)raw[u]);
        }
    }
    cnt = __reduce_add_sync(FULL, cnt);
    vsum = warp_sum(vsum);
    const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (lane == 0) {
        if (cnt) atomicAdd(&s_cnt, (unsigned long long)cnt);
        s_vsum[w] = vsum;
    }
    __syncthreads();
    if (SMEM && md != 0)
        for (int i = threadIdx.x; i < bins; i += blockDim.x)
            if (s_hist[i]) atomicAdd(&gh[i], (unsigned long long)s_hist[i]);
    if (threadIdx.x == 0) {
        if (s_cnt) atomicAdd(&nonnull[col], s_cnt);
        if (md == 0) {
            double t = 0.0;
            for (int i = 0; i < RFF_THREADS / 32; ++i) t += s_vsum[i];
            if (s_cnt) atomicAdd(&value_sum[col], t);
        }
    }
}

// Hashed-token distributions from sparse rows (text.cu's 4-array form: row r owns idx / val [row_start[r], +row_nnz[r]),
// indices ascending; index = column * bins + bucket, anything >= ncols * bins is ignored).  Per column:
//   hist [col * bins + bucket] += weight * tf     (FeatureDistribution.scala:317-321: HashingTF.transform(tokens))
//   tok [col] = {min tokens per non-empty cell, max, cells that are not empty}  (Summary.scala:61: Left(v) -> v.size)
// valid_ptrs [col]: the column's validity bitmap (0 = every row holds a value); weight [r] (or NULL = 1): how many data
// rows this row stands for (the rows are a dictionary's entries then; weight 0 = the entry does not occur).
template <bool SMEM>
__global__ void __launch_bounds__(RFF_THREADS)
rff_rows_hist_kernel(const int64_t* __restrict__ row_start, const int* __restrict__ row_nnz,
                     const int* __restrict__ idx, const int* __restrict__ val, int64_t n, int ncols, int bins,
                     const unsigned long long* const* __restrict__ valid_ptrs,
                     const unsigned long long* __restrict__ weight, unsigned long long* __restrict__ hist,
                     long long* __restrict__ tok) {
    extern __shared__ __align__(16) unsigned char rff_smem[];
    unsigned* s_hist = reinterpret_cast<unsigned*>(rff_smem);                       // [ncols * bins] when SMEM
    int* s_min = reinterpret_cast<int*>(rff_smem + (size_t)(SMEM ? ncols * bins : 0) * sizeof(unsigned));
    int* s_max = s_min + ncols;
    unsigned* s_cnt = reinterpret_cast<unsigned*>(s_max + ncols);
    const int width = ncols * bins;
    if (SMEM)
        for (int i = threadIdx.x; i < width; i += blockDim.x) s_hist[i] = 0u;
    for (int i = threadIdx.x; i < ncols; i += blockDim.x) {
        s_min[i] = 0x7fffffff;
        s_max[i] = -1;
        s_cnt[i] = 0u;
    }
    __syncthreads();
    const int64_t per = (n + gridDim.x - 1) / gridDim.x;
    const int64_t r0 = (int64_t)blockIdx.x * per, r1 = r0 + per < n ? r0 + per : n;
    for (int64_t r = r0 + threadIdx.x; r < r1; r += blockDim.x) {
        const unsigned long long w = weight ? weight[r] : 1ull;
        if (w == 0ull) continue;
        int64_t p = row_start[r];
        const int64_t end = p + row_nnz[r];
        for (int c = 0; c < ncols; ++c) {
            const int hi = (c + 1) * bins;
            int t = 0;
            while (p < end) {
                const int i = idx[p];
                if (i >= hi) break;
                const int tf = val[p];
                t += tf;
                if (SMEM) atomicAdd(&s_hist[i], (unsigned)tf);
                else atomicAdd(&hist[i], w * (unsigned long long)tf);
                ++p;
            }
            if (!rff_valid(valid_ptrs ? valid_ptrs[c] : nullptr, r)) continue;
            atomicMin(&s_min[c], t);
            atomicMax(&s_max[c], t);
            if (weight) atomicAdd(reinterpret_cast<unsigned long long*>(&tok[3 * c + 2]), w);
            else atomicAdd(&s_cnt[c], 1u);
        }
    }
    __syncthreads();
    if (SMEM)
        for (int i = threadIdx.x; i < width; i += blockDim.x)
            if (s_hist[i]) atomicAdd(&hist[i], (unsigned long long)s_hist[i]);
    for (int c = threadIdx.x; c < ncols; c += blockDim.x) {
        if (s_max[c] < 0) continue;
        atomicMin(&tok[3 * c + 0], (long long)s_min[c]);
        atomicMax(&tok[3 * c + 1], (long long)s_max[c]);
        if (s_cnt[c]) atomicAdd(reinterpret_cast<unsigned long long*>(&tok[3 * c + 2]), (unsigned long long)s_cnt[c]);
    }
}

// counts [col_off[col] + code] += 1 for every row of a dictionary-encoded column (codes [ncols][ld], < 0 = empty);
// equal codes inside a warp are added once (__match_any_sync: Zipf-distributed codes would serialise otherwise)
__global__ void __launch_bounds__(RFF_THREADS)
rff_code_count_kernel(const int* __restrict__ codes, int64_t ld, int64_t n, const int64_t* __restrict__ col_off,
                      const int64_t* __restrict__ ncodes, unsigned long long* __restrict__ counts, int* __restrict__ bad,
                      const double* __restrict__ y, const unsigned long long* __restrict__ y_valid, double shift,
                      double* __restrict__ null_y) {
    const int col = blockIdx.y;
    const int* c = codes + (int64_t)col * ld;
    unsigned long long* out = counts + col_off[col];
    const int64_t lim = ncodes[col];
    const int lane = threadIdx.x & 31;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    const int64_t rounds = (n + stride - 1) / stride;
    double ny = 0.0;                                   // response summed over this column's empty rows
    for (int64_t k = 0; k < rounds; ++k) {
        const int64_t r = k * stride + (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
        int code = r < n ? c[r] : -1;
        if (code >= lim) {
            atomicOr(bad, 1);
            code = -1;
        }
        if (y && r < n && code < 0) ny += (rff_valid(y_valid, r) ? y[r] : 0.0) - shift;
        const unsigned peers = __match_any_sync(FULL, code);
        if (code >= 0 && lane == __ffs(peers) - 1) atomicAdd(&out[code], (unsigned long long)__popc(peers));
    }
    if (y) {
        ny = warp_sum(ny);
        if (lane == 0 && ny != 0.0) atomicAdd(&null_y[col], ny);
    }
}

// out [ncols + 2]: out[c] += sum over the rows where column c is EMPTY of (y - shift); out[ncols] += sum (y - shift),
// out[ncols + 1] += sum (y - shift)^2 over all rows (blockIdx.y == ncols).  y_valid: the response's own bitmap (an empty
// response counts as 0.0, PreparedFeatures.scala:63-66) or NULL.
__global__ void __launch_bounds__(RFF_THREADS)
rff_null_label_kernel(const unsigned long long* const* __restrict__ valid_ptrs, int ncols, int64_t n,
                      const double* __restrict__ y, const unsigned long long* __restrict__ y_valid, double shift,
                      double* __restrict__ out) {
    const int col = blockIdx.y;
    double a = 0.0, b = 0.0;
    if (col == ncols) {
        for (int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; r < n; r += (int64_t)gridDim.x * blockDim.x) {
            const double d = (rff_valid(y_valid, r) ? y[r] : 0.0) - shift;
            a += d;
            b += d * d;
        }
    } else {
        const unsigned long long* v = valid_ptrs[col];
        if (v == nullptr) return;                                      // never empty
        const int64_t words = (n + 63) >> 6;
        for (int64_t w = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; w < words; w += (int64_t)gridDim.x * blockDim.x) {
            unsigned long long inv = ~v[w];
            if (w == words - 1 && (n & 63)) inv &= (1ull << (n & 63)) - 1ull;
            while (inv) {
                const int bit = __ffsll((long long)inv) - 1;
                inv &= inv - 1;
                const int64_t r = (w << 6) + bit;
                a += (rff_valid(y_valid, r) ? y[r] : 0.0) - shift;
            }
        }
    }
    a = warp_sum(a);
    b = warp_sum(b);
    __shared__ double s[RFF_THREADS / 32][2];
    const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (lane == 0) { s[w][0] = a; s[w][1] = b; }
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int i = 1; i < RFF_THREADS / 32; ++i) { a += s[i][0]; b += s[i][1]; }
        if (col == ncols) {
            atomicAdd(&out[ncols], a);
            atomicAdd(&out[ncols + 1], b);
        } else if (a != 0.0) {
            atomicAdd(&out[col], a);
        }
    }
}

}  // namespace

static int rff_grid_x(int64_t n, int ncols, int per_thread, int blocks_per_sm = 8) {
    // one wave: as many blocks per column as fit the SMs' resident slots (8 blocks of 256 threads per SM) -- one block
    // more per column would leave a second, nearly empty wave -- and never more than the rows need
    const int64_t want = std::max<int64_t>(1, ((int64_t)sm_count() * blocks_per_sm) / std::max(ncols, 1));
    const int64_t need = std::max<int64_t>(1, (n + (int64_t)RFF_THREADS * per_thread - 1) / ((int64_t)RFF_THREADS * per_thread));
    return (int)std::min<int64_t>(std::min(want, need), 65535);
}

}  // namespace tmog

using namespace tmog;

extern "C" int tmog_rff_numeric_summary_dev(const void* values, int dtype, int64_t ld, const uint64_t* valid, int64_t ldv,
                                            int64_t n, int ncols, double* out, void* stream) {
    TMOG_REQUIRE(n >= 0 && ncols >= 0 && ld >= n && dtype >= 0 && dtype <= 2, "bad shape or dtype");
    TMOG_REQUIRE(ncols <= 65535, "at most 65535 columns per call");
    if (n == 0 || ncols == 0) return TMOG_OK;
    TMOG_REQUIRE(values && out && (!valid || ldv * 64 >= n), "NULL argument or validity words too short");
    const dim3 grid(rff_grid_x(n, ncols, 8), ncols);
    const unsigned long long* v = reinterpret_cast<const unsigned long long*>(valid);
    cudaStream_t s = (cudaStream_t)stream;
    const size_t esz = dtype == 2 ? 1 : 8;
    const bool staged = (reinterpret_cast<uintptr_t>(values) & 15u) == 0 && ((size_t)ld * esz) % 16 == 0 &&
                        (n + grid.x - 1) / grid.x < (int64_t)0x7fffff00ll;
    if (staged) {
        if (dtype == 0) rff_numeric_summary_staged_kernel<0><<<grid, RFF_THREADS, 0, s>>>(values, ld, v, ldv, n, out);
        else if (dtype == 1) rff_numeric_summary_staged_kernel<1><<<grid, RFF_THREADS, 0, s>>>(values, ld, v, ldv, n, out);
        else rff_numeric_summary_staged_kernel<2><<<grid, RFF_THREADS, 0, s>>>(values, ld, v, ldv, n, out);
    } else if (dtype == 0) rff_numeric_summary_kernel<0><<<grid, RFF_THREADS, 0, s>>>(values, ld, v, ldv, n, out);
    else if (dtype == 1) rff_numeric_summary_kernel<1><<<grid, RFF_THREADS, 0, s>>>(values, ld, v, ldv, n, out);
    else rff_numeric_summary_kernel<2><<<grid, RFF_THREADS, 0, s>>>(values, ld, v, ldv, n, out);
    TMOG_LAUNCH_CHECK();
    return TMOG_OK;
}

extern "C" int tmog_rff_numeric_hist_dev(const void* values, int dtype, int64_t ld, const uint64_t* valid, int64_t ldv,
                                         int64_t n, int ncols, int bins, const double* splits, const int32_t* mode,
                                         const double* same_value, uint64_t* hist, double* value_sum, uint64_t* nonnull,
                                         void* stream) {
    TMOG_REQUIRE(n >= 0 && ncols >= 0 && ld >= n && dtype >= 0 && dtype <= 2, "bad shape or dtype");
    TMOG_REQUIRE(bins > 1 && bins <= 100000, "Invalid bin size %d, bins must be between 1 and 100000", bins);
    TMOG_REQUIRE(ncols <= 65535, "at most 65535 columns per call");
    if (n == 0 || ncols == 0) return TMOG_OK;
    TMOG_REQUIRE(values && splits && mode && same_value && hist && value_sum && nonnull && (!valid || ldv * 64 >= n),
                 "NULL argument or validity words too short");
    const dim3 grid(rff_grid_x(n, ncols, 16, 6), ncols);       // 38 registers: six blocks of 256 threads per SM
    TMOG_REQUIRE((n + grid.x - 1) / grid.x < (int64_t)0x7fffff00ll, "too many rows per call");
    const unsigned long long* v = reinterpret_cast<const unsigned long long*>(valid);
    unsigned long long* h = reinterpret_cast<unsigned long long*>(hist);
    unsigned long long* nn = reinterpret_cast<unsigned long long*>(nonnull);
    cudaStream_t s = (cudaStream_t)stream;
    const bool smem = bins <= RFF_SMEM_BINS;
    const size_t sh = smem ? (size_t)bins * (sizeof(double) + sizeof(unsigned)) : 0;
    const size_t esz = dtype == 2 ? 1 : 8;
    const bool staged = (reinterpret_cast<uintptr_t>(values) & 15u) == 0 && ((size_t)ld * esz) % 16 == 0;
#define RFF_HIST2(DT, SM, ST)                                                                                            \
    rff_numeric_hist_kernel<DT, SM, ST><<<grid, RFF_THREADS, sh, s>>>(values, ld, v, ldv, n, bins, splits, mode,          \
                                                                     same_value, h, value_sum, nn)
#define RFF_HIST(DT)                                                                                                     \
    do {                                                                                                                 \
        if (smem && staged) RFF_HIST2(DT, true, true);                                                                   \
        else if (smem) RFF_HIST2(DT, true, false);                                                                       \
        else if (staged) RFF_HIST2(DT, false, true);                                                                     \
        else RFF_HIST2(DT, false, false);                                                                                \
    } while (0)
    if (dtype == 0) RFF_HIST(0);
    else if (dtype == 1) RFF_HIST(1);
    else RFF_HIST(2);
#undef RFF_HIST2
#undef RFF_HIST
    TMOG_LAUNCH_CHECK();
    return TMOG_OK;
}

extern "C" int tmog_rff_rows_hist_dev(const int64_t* row_start, const int32_t* row_nnz, const int32_t* idx,
                                      const int32_t* val, int64_t n, int ncols, int bins, const uint64_t* const* valid_ptrs,
                                      const uint64_t* weight, uint64_t* hist, int64_t* tok, void* stream) {
    TMOG_REQUIRE(n >= 0 && ncols >= 0 && bins > 0, "bad shape");
    TMOG_REQUIRE(ncols <= 1024, "at most 1024 hashed columns per call");
    TMOG_REQUIRE((int64_t)ncols * bins < (int64_t)0x7fffffff, "ncols * bins must fit 31 bits");
    if (n == 0 || ncols == 0) return TMOG_OK;
    TMOG_REQUIRE(row_start && row_nnz && idx && val && hist && tok, "NULL argument");
    const int grid = (int)std::max<int64_t>(1, std::min<int64_t>((n + RFF_THREADS * 4 - 1) / (RFF_THREADS * 4),
                                                                   (int64_t)sm_count() * 4));
    TMOG_REQUIRE((n + grid - 1) / grid < (int64_t)(1 << 20), "too many rows per call");   // u32 shared sums of tf <= 4095
    const bool smem = weight == nullptr && (int64_t)ncols * bins <= 8192;
    const size_t sh = (smem ? (size_t)ncols * bins * sizeof(unsigned) : 0) + (size_t)ncols * 3 * sizeof(int);
    const unsigned long long* const* vp = reinterpret_cast<const unsigned long long* const*>(valid_ptrs);
    const unsigned long long* w = reinterpret_cast<const unsigned long long*>(weight);
    unsigned long long* h = reinterpret_cast<unsigned long long*>(hist);
    long long* t = reinterpret_cast<long long*>(tok);
    cudaStream_t s = (cudaStream_t)stream;
    if (smem) rff_rows_hist_kernel<true><<<grid, RFF_THREADS, sh, s>>>(row_start, row_nnz, idx, val, n, ncols, bins, vp, w, h, t);
    else rff_rows_hist_kernel<false><<<grid, RFF_THREADS, sh, s>>>(row_start, row_nnz, idx, val, n, ncols, bins, vp, w, h, t);
    TMOG_LAUNCH_CHECK();
    return TMOG_OK;
}

extern "C" int tmog_rff_code_count_dev(const int32_t* codes, int64_t ld, int64_t n, int ncols, const int64_t* col_off,
                                       const int64_t* ncodes, uint64_t* counts, int32_t* bad, const double* y,
                                       const uint64_t* y_valid, double shift, double* null_y, void* stream) {
    TMOG_REQUIRE(n >= 0 && ncols >= 0 && ld >= n, "bad shape");
    TMOG_REQUIRE(ncols <= 65535, "at most 65535 columns per call");
    if (n == 0 || ncols == 0) return TMOG_OK;
    TMOG_REQUIRE(codes && col_off && ncodes && counts && bad && (!y || null_y), "NULL argument");
    const dim3 grid(rff_grid_x(n, ncols, 8), ncols);
    rff_code_count_kernel<<<grid, RFF_THREADS, 0, (cudaStream_t)stream>>>(
        codes, ld, n, col_off, ncodes, reinterpret_cast<unsigned long long*>(counts), bad, y,
        reinterpret_cast<const unsigned long long*>(y_valid), shift, null_y);
    TMOG_LAUNCH_CHECK();
    return TMOG_OK;
}

extern "C" int tmog_rff_null_label_sums_dev(const uint64_t* const* valid_ptrs, int ncols, int64_t n, const double* y,
                                            const uint64_t* y_valid, double shift, double* out, void* stream) {
    TMOG_REQUIRE(n >= 0 && ncols >= 0, "bad shape");
    TMOG_REQUIRE(ncols < 65535, "at most 65534 columns per call");
    if (n == 0) return TMOG_OK;
    TMOG_REQUIRE(y && out && (ncols == 0 || valid_ptrs), "NULL argument");
    const dim3 grid(rff_grid_x(n, ncols + 1, 8), ncols + 1);
    rff_null_label_kernel<<<grid, RFF_THREADS, 0, (cudaStream_t)stream>>>(
        reinterpret_cast<const unsigned long long* const*>(valid_ptrs), ncols, n, y,
        reinterpret_cast<const unsigned long long*>(y_valid), shift, out);
    TMOG_LAUNCH_CHECK();
    return TMOG_OK;
}

// ---- host side (driver-side finishing, tiny inputs) -------------------------------------------------------------------

extern "C" int tmog_rff_numeric_splits(double min, double max, int bins, double* out) {
    TMOG_REQUIRE(bins > 1 && out, "bad arguments");
    // FeatureDistribution.scala:333-334: step = (max - min) / (bins - 2.0); splits(b) = min + step * b -- a product and
    // a sum, each rounded (the JVM does not contract them into a fused multiply-add; neither may this)
    const double step = (max - min) / ((double)bins - 2.0);
    for (int b = 0; b < bins; ++b) {
        volatile double prod = step * (double)b;
        out[b] = min + prod;
    }
    return TMOG_OK;
}

extern "C" int tmog_rff_js_divergence(const double* a, const double* b, int n, double* out) {
    TMOG_REQUIRE(n >= 0 && out && (n == 0 || (a && b)), "bad arguments");
    // FeatureDistribution.jsDivergence (:154-163): bins empty on both sides are dropped, the rest normalised by the
    // two totals (0 / 0 = NaN as on the JVM), KL terms in log2 = log10(x) / log10(2)
    double ta = 0.0, tb = 0.0;
    for (int i = 0; i < n; ++i)
        if (!(a[i] == 0.0 && b[i] == 0.0)) {
            ta += a[i];
            tb += b[i];
        }
    const double l2 = std::log10(2.0);
    double total = 0.0;
    for (int i = 0; i < n; ++i) {
        if (a[i] == 0.0 && b[i] == 0.0) continue;
        const double pa = a[i] / ta, pb = b[i] / tb;
        const double m = (pa + pb) / 2;
        const double ka = pa == 0.0 ? 0.0 : pa * (std::log10(pa / m) / l2);
        const double kb = pb == 0.0 ? 0.0 : pb * (std::log10(pb / m) / l2);
        total += 0.5 * ka + 0.5 * kb;
    }
    *out = total;
    return TMOG_OK;
}

extern "C" int tmog_rff_null_label_corr(double n, double y_sum, double y_sq_sum, int ncols, const double* null_count,
                                        const double* null_y_sum, double* out) {
    TMOG_REQUIRE(ncols >= 0 && (ncols == 0 || (null_count && null_y_sum && out)), "bad arguments");
    // Statistics.corr over [response, null indicators] (RawFeatureFilter.scala:181-191): Spark's covariance from the
    // Gramian with n - 1, sigma = 0 where the variance is <= 1e-12, NaN where a sigma is 0; then min(|corr|, 1.0).
    // y_sum / y_sq_sum / null_y_sum are sums of the SHIFTED response (any shift: covariances do not see it).
    const double my = y_sum / n;
    const double cyy = (y_sq_sum - n * my * my) / (n - 1.0);
    const double sy = std::fabs(cyy) <= 1e-12 ? 0.0 : std::sqrt(cyy);
    for (int j = 0; j < ncols; ++j) {
        const double mi = null_count[j] / n;
        const double cii = (null_count[j] - n * mi * mi) / (n - 1.0);
        const double si = std::fabs(cii) <= 1e-12 ? 0.0 : std::sqrt(cii);
        const double ciy = (null_y_sum[j] - n * mi * my) / (n - 1.0);
        if (si == 0.0 || sy == 0.0) {
            out[j] = NAN;
        } else {
            const double c = std::fabs(ciy / (si * sy));
            out[j] = c != c ? NAN : std::min(c, 1.0);
        }
    }
    return TMOG_OK;
}
