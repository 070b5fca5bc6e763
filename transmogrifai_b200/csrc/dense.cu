// SanityChecker over an already materialised dense OPVector block (row-major f64): column moments + label
// covariance + per-label-class column sums, the full Gramian for CorrelationLevel.Computed, and the indicesToKeep
// gather of SanityCheckerModel.  These serve vectors that did not come out of the fused vectorizer scans.
#include <math.h>

#include <algorithm>
#include <vector>

#include "common.cuh"

namespace tmog {

constexpr int NS = TMOG_REAL_NSTAT;
constexpr int DS_ROWS_PER_CTA = 4096;
constexpr int PILOT_ROWS = 4096;

// grid (ceil(D/32), row chunks); block (32 columns, 8 row lanes).  Column D-1 == d is the label.
__global__ void __launch_bounds__(256)
dense_pilot_kernel(const double* __restrict__ x, int64_t ld, int64_t n, int d, const double* __restrict__ label,
                   double* __restrict__ shifts) {
    __shared__ double red[8][33];
    const int c = blockIdx.x * 32 + threadIdx.x;
    const int rows = (int)min((int64_t)PILOT_ROWS, n);
    double s = 0.0;
    if (c <= d)
        for (int r = threadIdx.y; r < rows; r += 8) {
            const double v = c == d ? label[r] : x[(int64_t)r * ld + c];
            s += isfinite(v) ? v : 0.0;
        }
    red[threadIdx.y][threadIdx.x] = s;
    __syncthreads();
    if (threadIdx.y == 0 && c <= d) {
        double t = 0.0;
        for (int k = 0; k < 8; ++k) t += red[k][threadIdx.x];
        shifts[c] = rows > 0 ? t / rows : 0.0;
    }
}

__global__ void __launch_bounds__(256)
dense_stats_kernel(const double* __restrict__ x, int64_t ld, int64_t n, int d, const double* __restrict__ label,
                   const uint8_t* __restrict__ class_ids, int nclasses, const uint8_t* __restrict__ cat_mask,
                   const double* __restrict__ shifts, double* __restrict__ scratch, double* __restrict__ class_sums) {
    extern __shared__ double sm[];
    double* red = sm;                    // [8][32][6]
    double* cs = sm + 8 * 32 * 6;        // [nclasses][32] (only when class sums are requested)
    const int tx = threadIdx.x, ty = threadIdx.y;
    const int c = blockIdx.x * 32 + tx;
    const bool is_label = c == d;
    const bool active = c <= d;
    const bool track = class_sums != nullptr && active && !is_label && cat_mask != nullptr && cat_mask[c] != 0;
    if (class_sums != nullptr)
        for (int i = ty * 32 + tx; i < nclasses * 32; i += 256) cs[i] = 0.0;
    __syncthreads();
    const int64_t r0 = (int64_t)blockIdx.y * DS_ROWS_PER_CTA;
    const int64_t r1 = min(n, r0 + DS_ROWS_PER_CTA);
    const double s = active ? shifts[c] : 0.0;
    const double sy = shifts[d];
    double sd = 0, sdd = 0, sde = 0, se = 0, mn = INFINITY, mx = -INFINITY;
    if (active)
        for (int64_t r = r0 + ty; r < r1; r += 8) {
            const double y = label[r];
            const double v = is_label ? y : x[r * ld + c];
            const double dv = v - s, e = y - sy;
            sd += dv;
            sdd = fma(dv, dv, sdd);
            sde = fma(dv, e, sde);
            se += e;
            mn = fmin(mn, v);
            mx = fmax(mx, v);
            if (track && v != 0.0) atomicAdd(&cs[class_ids[r] * 32 + tx], v);
        }
    double* q = red + (ty * 32 + tx) * 6;
    q[0] = sd; q[1] = sdd; q[2] = sde; q[3] = se; q[4] = mn; q[5] = mx;
    __syncthreads();
    if (ty == 0 && active) {
        for (int k = 1; k < 8; ++k) {
            const double* p = red + (k * 32 + tx) * 6;
            sd += p[0]; sdd += p[1]; sde += p[2]; se += p[3];
            mn = fmin(mn, p[4]);
            mx = fmax(mx, p[5]);
        }
        double* o = scratch + ((int64_t)blockIdx.y * (d + 1) + c) * NS;
        o[0] = (double)(r1 - r0); o[1] = sd; o[2] = sdd; o[3] = sde; o[4] = se; o[5] = mn; o[6] = mx;
        o[7] = (double)(r1 - r0);
    }
    if (class_sums != nullptr && c < d) {
        for (int k = ty; k < nclasses; k += 8) {
            const double v = cs[k * 32 + tx];
            if (v != 0.0) atomicAdd(&class_sums[(int64_t)k * d + c], v);
        }
    }
}

__global__ void __launch_bounds__(32)
dense_reduce_kernel(const double* __restrict__ scratch, int nchunks, int D, double* __restrict__ partials) {
    const int c = blockIdx.x, lane = threadIdx.x;
    double s[5] = {0, 0, 0, 0, 0}, mn = INFINITY, mx = -INFINITY, rows = 0;
    for (int b = lane; b < nchunks; b += 32) {
        const double* q = scratch + ((int64_t)b * D + c) * NS;
        for (int k = 0; k < 5; ++k) s[k] += q[k];
        mn = fmin(mn, q[5]);
        mx = fmax(mx, q[6]);
        rows += q[7];
    }
    for (int k = 0; k < 5; ++k) s[k] = warp_sum(s[k]);
    mn = warp_min(mn);
    mx = warp_max(mx);
    rows = warp_sum(rows);
    if (lane == 0) {
        double* o = partials + c * NS;
        for (int k = 0; k < 5; ++k) o[k] = s[k];
        o[5] = mn; o[6] = mx; o[7] = rows;
    }
}

// out[r][j] = x[r][idx[j]]   (DerivedFeatureFilterUtils.removeFeatures on a dense vector)
__global__ void __launch_bounds__(256)
dense_gather_kernel(const double* __restrict__ x, int64_t ld, int64_t n, const int32_t* __restrict__ idx, int k,
                    double* __restrict__ out, int64_t out_ld) {
    const int64_t total = n * (int64_t)k;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) {
        const int64_t r = i / k;
        const int j = (int)(i - r * k);
        out[r * out_ld + j] = x[r * ld + idx[j]];
    }
}

// G[i][j] += sum_r (x[r][i] - s_i)(x[r][j] - s_j), label as column d; fp64 atomics merge the row stretches.
// On the fp64 tensor cores (mma.sync m8n8k4, DMMA): G = sum over row chunks of X_c^T X_c.  A CTA of
// four warps owns a 64 x 64 output tile (upper-triangular tiles only) over one stretch of rows (grid.z); 32 shifted
// rows at a time are staged in shared memory (row stride 68 doubles: the fragment loads of a half-warp hit 16 distinct
// banks), every warp keeps a 32 x 32 block of the tile in 4 x 4 accumulator fragments and issues 16 DMMAs per 8
// fragment loads.  Fragment layout of m8n8k4 (PTX ISA): A[m][k] and B[k][n] are held by lane (m or n) * 4 + k, which for
// X^T X is the same shared-memory index pattern on both sides; C[m][n] by lane m * 4 + n / 2, two values each.
constexpr int GM_TILE = 64;
constexpr int GM_KC = 32;
constexpr int GM_LD = 68;

__device__ __forceinline__ void dmma_m8n8k4(double& c0, double& c1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0, %1}, {%2}, {%3}, {%0, %1};"
                 : "+d"(c0), "+d"(c1)
                 : "d"(a), "d"(b));
}

__device__ __forceinline__ void cp_async8(double* smem_dst, const double* gmem_src) {
    const unsigned d = (unsigned)__cvta_generic_to_shared(smem_dst);
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(d), "l"(gmem_src));
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N)); }

// Raw values are staged with cp.async into two buffers (the copy of the next 32 rows overlaps the DMMAs of the current
// ones); the shifts are subtracted when the fragments are read.  Positions outside the matrix are staged as the
// shift itself, so that they contribute zero.
constexpr int GM_STAGE = GM_KC * GM_LD;          // doubles per staged tile

__global__ void __launch_bounds__(128)
dense_gram_dmma_kernel(const double* __restrict__ x, int64_t ld, int64_t n, int d, const double* __restrict__ label,
                       const double* __restrict__ shifts, int64_t rows_per_cta, double* __restrict__ G) {
    const int ti = blockIdx.x, tj = blockIdx.y;
    if (tj < ti) return;
    extern __shared__ double gm_smem[];          // [2 stages][A | B][GM_KC][GM_LD]
    const bool diag = ti == tj;
    const int D = d + 1;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int wi = warp >> 1, wj = warp & 1;
    const int64_t r0 = (int64_t)blockIdx.z * rows_per_cta;
    const int64_t r1 = min(n, r0 + rows_per_cta);
    double acc[4][4][2];
#pragma unroll
    for (int a = 0; a < 4; ++a)
#pragma unroll
        for (int b = 0; b < 4; ++b) acc[a][b][0] = acc[a][b][1] = 0.0;
    // loader: this thread's column of the staged tiles
    const int cc = tid & 63, rhalf = tid >> 6;           // 2 rows of 64 columns per pass of the CTA
    const int ca = ti * GM_TILE + cc, cb = tj * GM_TILE + cc;
    const double sa = ca < D ? shifts[ca] : 0.0, sb = cb < D ? shifts[cb] : 0.0;
    auto stage = [&](int64_t rb, int buf) {
        double* As = gm_smem + buf * 2 * GM_STAGE;
        double* Bs = As + GM_STAGE;
#pragma unroll 4
        for (int rr = rhalf; rr < GM_KC; rr += 2) {
            const int64_t r = rb + rr;
            if (r < r1 && ca < D) cp_async8(&As[rr * GM_LD + cc], ca == d ? label + r : x + r * ld + ca);
            else As[rr * GM_LD + cc] = sa;
            if (!diag) {
                if (r < r1 && cb < D) cp_async8(&Bs[rr * GM_LD + cc], cb == d ? label + r : x + r * ld + cb);
                else Bs[rr * GM_LD + cc] = sb;
            }
        }
        cp_async_commit();
    };
    // fragments: [k = lane & 3][column = lane >> 2]; shifts of the 4 + 4 fragment columns of this lane
    const int frag = (lane & 3) * GM_LD + (lane >> 2);
    double sha[4], shb[4];
#pragma unroll
    for (int m = 0; m < 4; ++m) {
        const int ia = ti * GM_TILE + wi * 32 + m * 8 + (lane >> 2), ib = tj * GM_TILE + wj * 32 + m * 8 + (lane >> 2);
        sha[m] = ia < D ? shifts[ia] : 0.0;
        shb[m] = ib < D ? shifts[ib] : 0.0;
    }
    int buf = 0;
    if (r0 < r1) stage(r0, 0);
    for (int64_t rb = r0; rb < r1; rb += GM_KC) {
        if (rb + GM_KC < r1) {
            stage(rb + GM_KC, buf ^ 1);
            cp_async_wait<1>();
        } else {
            cp_async_wait<0>();
        }
        __syncthreads();
        const double* As = gm_smem + buf * 2 * GM_STAGE;
        const double* Bt = diag ? As : As + GM_STAGE;
#pragma unroll
        for (int kk = 0; kk < GM_KC / 4; ++kk) {
            double a[4], b[4];
#pragma unroll
            for (int m = 0; m < 4; ++m) a[m] = As[kk * 4 * GM_LD + frag + wi * 32 + m * 8] - sha[m];
#pragma unroll
            for (int q = 0; q < 4; ++q) b[q] = Bt[kk * 4 * GM_LD + frag + wj * 32 + q * 8] - shb[q];
#pragma unroll
            for (int m = 0; m < 4; ++m)
#pragma unroll
                for (int q = 0; q < 4; ++q) dmma_m8n8k4(acc[m][q][0], acc[m][q][1], a[m], b[q]);
        }
        __syncthreads();
        buf ^= 1;
    }
#pragma unroll
    for (int m = 0; m < 4; ++m)
#pragma unroll
        for (int q = 0; q < 4; ++q)
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                const int i = ti * GM_TILE + wi * 32 + m * 8 + (lane >> 2);
                const int j = tj * GM_TILE + wj * 32 + q * 8 + (lane & 3) * 2 + h;
                if (i < D && j < D && i <= j) atomicAdd(&G[(int64_t)i * D + j], acc[m][q][h]);
            }
}

}  // namespace tmog

using namespace tmog;

static int dense_chunks(int64_t n) { return (int)std::max<int64_t>(1, (n + DS_ROWS_PER_CTA - 1) / DS_ROWS_PER_CTA); }

extern "C" int tmog_dense_stats_scratch_bytes(int64_t n, int d, size_t* bytes) {
    TMOG_REQUIRE(n >= 0 && d >= 0 && bytes != nullptr, "bad arguments");
    *bytes = sizeof(double) * (size_t)dense_chunks(n) * (size_t)(d + 1) * NS;
    return TMOG_OK;
}

extern "C" int tmog_dense_pilot_shift_dev(const double* x, int64_t ld, int64_t n, int d, const double* label,
                                          double* shifts, void* stream) {
    TMOG_REQUIRE(n >= 0 && d >= 0 && ld >= d && shifts != nullptr, "bad arguments");
    TMOG_REQUIRE(n == 0 || ((x != nullptr || d == 0) && label != nullptr), "NULL input");
    dense_pilot_kernel<<<(d + 1 + 31) / 32, dim3(32, 8), 0, (cudaStream_t)stream>>>(x, ld, n, d, label, shifts);
    TMOG_LAUNCH_CHECK();
    return TMOG_OK;
}

extern "C" int tmog_dense_stats_dev(const double* x, int64_t ld, int64_t n, int d, const double* label,
                                    const uint8_t* class_ids, int nclasses, const uint8_t* cat_mask,
                                    const double* shifts, void* scratch, size_t scratch_bytes, double* partials,
                                    double* class_sums, void* stream) {
    TMOG_REQUIRE(n >= 0 && d >= 0 && ld >= d, "bad shape n=%lld d=%d ld=%lld", (long long)n, d, (long long)ld);
    TMOG_REQUIRE(n == 0 || ((x != nullptr || d == 0) && label != nullptr), "NULL input");
    TMOG_REQUIRE(shifts != nullptr && partials != nullptr, "shifts / partials is NULL");
    TMOG_REQUIRE(class_sums == nullptr || (class_ids != nullptr && nclasses >= 1 && nclasses <= 255),
                 "class sums need class ids and 1..255 classes");
    const int chunks = dense_chunks(n);
    const size_t need = sizeof(double) * (size_t)chunks * (size_t)(d + 1) * NS;
    if (scratch == nullptr || scratch_bytes < need) {
        set_error("scratch too small: %zu bytes given, %zu needed (tmog_dense_stats_scratch_bytes)", scratch_bytes,
                  need);
        return TMOG_ENOMEM;
    }
    cudaStream_t st = (cudaStream_t)stream;
    if (class_sums) TMOG_CUDA(cudaMemsetAsync(class_sums, 0, sizeof(double) * (size_t)nclasses * d, st));
    const size_t smem = sizeof(double) * (8 * 32 * 6 + (class_sums ? (size_t)nclasses * 32 : 0));
    if (smem > 48 * 1024)
        TMOG_CUDA(cudaFuncSetAttribute(dense_stats_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    dim3 grid((d + 1 + 31) / 32, chunks);
    TMOG_REQUIRE(chunks <= 65535, "too many rows for one dense stats call (%lld): chunk the batch", (long long)n);
    dense_stats_kernel<<<grid, dim3(32, 8), smem, st>>>(x, ld, n, d, label, class_ids, nclasses, cat_mask, shifts,
                                                        static_cast<double*>(scratch), class_sums);
    TMOG_LAUNCH_CHECK();
    dense_reduce_kernel<<<d + 1, 32, 0, st>>>(static_cast<const double*>(scratch), n > 0 ? chunks : 0, d + 1,
                                              partials);
    TMOG_LAUNCH_CHECK();
    return TMOG_OK;
}

extern "C" int tmog_dense_gather_dev(const double* x, int64_t ld, int64_t n, const int32_t* idx, int k, double* out,
                                     int64_t out_ld, void* stream) {
    TMOG_REQUIRE(n >= 0 && k >= 0 && out_ld >= k, "bad shape");
    if (n == 0 || k == 0) return TMOG_OK;
    TMOG_REQUIRE(x != nullptr && idx != nullptr && out != nullptr, "NULL argument");
    const int64_t total = n * (int64_t)k;
    const int grid = (int)std::min<int64_t>((total + 255) / 256, (int64_t)sm_count() * 16);
    dense_gather_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(x, ld, n, idx, k, out, out_ld);
    TMOG_LAUNCH_CHECK();
    return TMOG_OK;
}

extern "C" int tmog_dense_gram_dev(const double* x, int64_t ld, int64_t n, int d, const double* label,
                                   const double* shifts, double* gram, void* stream) {
    TMOG_REQUIRE(n >= 0 && d >= 0 && ld >= d && shifts != nullptr && gram != nullptr, "bad arguments");
    TMOG_REQUIRE(d + 1 <= 65535, "Argument with more than 65535 cols: %d", d + 1);   // Spark RowMatrix limit
    if (n == 0) return TMOG_OK;
    TMOG_REQUIRE((x != nullptr || d == 0) && label != nullptr, "NULL input");
    const int tiles = (d + 1 + GM_TILE - 1) / GM_TILE;
    const int64_t want_ctas = (int64_t)sm_count() * 6;
    int64_t zc = std::max<int64_t>(1, want_ctas / std::max(1, tiles * (tiles + 1) / 2));
    int64_t rows_per = std::max<int64_t>(GM_KC, (n + zc - 1) / zc);
    rows_per = (rows_per + GM_KC - 1) / GM_KC * GM_KC;
    zc = (n + rows_per - 1) / rows_per;
    TMOG_REQUIRE(zc <= 65535 && tiles <= 65535, "grid too large");
    const size_t smem = sizeof(double) * 4 * GM_STAGE;
    TMOG_CUDA(cudaFuncSetAttribute(dense_gram_dmma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    dense_gram_dmma_kernel<<<dim3(tiles, tiles, (unsigned)zc), 128, smem, (cudaStream_t)stream>>>(x, ld, n, d, label,
                                                                                                 shifts, rows_per, gram);
    TMOG_LAUNCH_CHECK();
    return TMOG_OK;
}

// Spark PearsonCorrelation.computeCorrelationMatrixFromCovariance on top of RowMatrix.computeCovariance, from the
// shifted Gramian G (upper triangle valid) and the shifted column sums: cov_ij = (G_ij - n delta_i delta_j)/(n-1)
// with delta = sum_d / n; sigma := 0 where |cov_ii| <= 1e-12 -> NaN row/col; diagonal 1.0.
extern "C" int tmog_pearson_from_gram(int D, double n, const double* gram, const double* sums, double* corr) {
    TMOG_REQUIRE(D >= 0 && gram != nullptr && sums != nullptr && corr != nullptr, "NULL argument");
    TMOG_REQUIRE(n > 1, "Cannot compute the covariance of a RowMatrix with <= 1 row.");
    std::vector<double> sigma(D);
    auto cov = [&](int i, int j) {
        const int a = std::min(i, j), b = std::max(i, j);
        return (gram[(size_t)a * D + b] - sums[a] * sums[b] / n) / (n - 1.0);
    };
    for (int i = 0; i < D; ++i) {
        const double c = cov(i, i);
        sigma[i] = fabs(c) <= 1e-12 ? 0.0 : sqrt(c);
    }
    for (int i = 0; i < D; ++i)
        for (int j = 0; j < D; ++j) {
            if (i == j)
                corr[(size_t)i * D + j] = 1.0;
            else if (sigma[i] == 0.0 || sigma[j] == 0.0)
                corr[(size_t)i * D + j] = NAN;
            else
                corr[(size_t)i * D + j] = cov(i, j) / (sigma[i] * sigma[j]);
        }
    return TMOG_OK;
}

// ---------------------------------------------------------------------------------------------------------------------
// Sparse OPVector parts -> columns of a dense row-major block (for CorrelationLevel.Computed, whose Gramian kernel
// works on dense rows; Spark's RowMatrix densifies the same way through BLAS spr/dspr).  The caller zeroes the block.
// ---------------------------------------------------------------------------------------------------------------------
namespace tmog {
__global__ void __launch_bounds__(256)
onehot_densify_kernel(const int32_t* __restrict__ idx, int64_t idx_ld, int64_t n, int k, double* __restrict__ out,
                      int64_t out_ld, int64_t out_col0) {
    const int64_t total = n * (int64_t)k;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t r = i / k;
        const int v = idx[r * idx_ld + (i - r * k)];
        if (v >= 0) out[r * out_ld + out_col0 + v] = 1.0;
    }
}

__global__ void __launch_bounds__(256)
sparse_rows_densify_kernel(const int64_t* __restrict__ row_start, const int32_t* __restrict__ row_nnz,
                           const int32_t* __restrict__ idx, const int32_t* __restrict__ val, int64_t n,
                           double* __restrict__ out, int64_t out_ld, int64_t out_col0) {
    const int lane = threadIdx.x & 31;
    const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    for (int64_t r = warp; r < n; r += nwarps) {
        const int64_t s = row_start[r];
        const int m = row_nnz[r];
        for (int e = lane; e < m; e += 32) out[r * out_ld + out_col0 + idx[s + e]] = (double)val[s + e];
    }
}
}  // namespace tmog

extern "C" int tmog_onehot_densify_dev(const int32_t* indices, int64_t idx_ld, int64_t n, int k, double* out,
                                       int64_t out_ld, int64_t out_col0, void* stream) {
    TMOG_REQUIRE(n >= 0 && k >= 0 && idx_ld >= k && out_col0 >= 0, "bad shape");
    if (n == 0 || k == 0) return TMOG_OK;
    TMOG_REQUIRE(indices && out, "NULL argument");
    const int grid = (int)std::min<int64_t>((n * (int64_t)k + 255) / 256, (int64_t)sm_count() * 16);
    onehot_densify_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(indices, idx_ld, n, k, out, out_ld, out_col0);
    TMOG_LAUNCH_CHECK();
    return TMOG_OK;
}

extern "C" int tmog_sparse_rows_densify_dev(const int64_t* row_start, const int32_t* row_nnz, const int32_t* indices,
                                            const int32_t* values, int64_t n, double* out, int64_t out_ld,
                                            int64_t out_col0, void* stream) {
    TMOG_REQUIRE(n >= 0 && out_col0 >= 0, "bad shape");
    if (n == 0) return TMOG_OK;
    TMOG_REQUIRE(row_start && row_nnz && out, "NULL argument");
    const int grid = (int)std::min<int64_t>((n * 32 + 255) / 256, (int64_t)sm_count() * 16);
    sparse_rows_densify_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(row_start, row_nnz, indices, values, n, out, out_ld,
                                                                       out_col0);
    TMOG_LAUNCH_CHECK();
    return TMOG_OK;
}
