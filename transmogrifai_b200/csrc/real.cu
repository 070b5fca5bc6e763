// Real / RealNN / Binary vectorizers fused with the SanityChecker column scan (sm_100a).
//
// Path restated here (reference file:line in include/tmog_b200.h):
//   fit        RealVectorizer.fitFn + MeanSeqNullNum, and for the emitted columns SanityChecker's colStats, label
//              correlation and per-label null-indicator sums, all from ONE read of the raw columns
//   transform  RealVectorizerModel / RealNNVectorizer / BinaryVectorizer transformFn into a row-major dense block
//
// HBM layout: column-major f64 values (stride ld), Arrow validity bitmaps (stride ldv words), f64 label.
// Roofline: HBM.  fit reads ncols*8 + ncols/8 + 8 bytes per row once; transform reads the same minus the label and
// writes 16 bytes per (row, column) when null indicators are tracked.
#include <math.h>

#include <atomic>

#include "common.cuh"

namespace tmog {

constexpr int FIT_TILE = 2048;  // rows per tile = 32 lanes x 32 double2
constexpr int FIT_THREADS = 256;
constexpr int FIT_WARPS = FIT_THREADS / 32;
constexpr int FIT_CTAS_PER_SM = 2;
constexpr int NS = TMOG_REAL_NSTAT;
constexpr int PILOT_ROWS = 4096;

struct Acc {
    double sd, sdd, sde, se, mn, mx;
};

__device__ __forceinline__ void acc_init(Acc& a) {
    a.sd = a.sdd = a.sde = a.se = 0.0;
    a.mn = INFINITY;
    a.mx = -INFINITY;
}

__device__ __forceinline__ void acc_add(Acc& a, bool v, double x, double s, double e) {
    const double d = v ? x - s : 0.0;
    const double ev = v ? e : 0.0;
    a.sd += d;
    a.sdd = fma(d, d, a.sdd);
    a.sde = fma(d, ev, a.sde);
    a.se += ev;
    a.mn = fmin(a.mn, v ? x : INFINITY);
    a.mx = fmax(a.mx, v ? x : -INFINITY);
}

__device__ __forceinline__ void kahan_add(double& sum, double& comp, double v) {
    const double y = v - comp;
    const double t = sum + y;
    comp = (t - sum) - y;
    sum = t;
}

// One CTA walks row tiles (persistent, grid = SMs x FIT_CTAS_PER_SM).  Per tile the label is staged once in shared
// memory as e = y - shift_y; each warp then streams whole columns of the tile with 128-bit loads (lane i holds rows
// 2i, 2i+1 of every 64-row group, whose validity is one bitmap word broadcast by shuffle) and keeps six fp64
// accumulators per lane.  A warp owns columns c = warp, warp + 8, ... so the per-CTA accumulators in shared memory
// need no atomics; they are Kahan-compensated across tiles and written once per CTA to `scratch`.
__global__ void __launch_bounds__(FIT_THREADS, FIT_CTAS_PER_SM)
real_fit_kernel(const double* __restrict__ values, int64_t ld, const uint64_t* __restrict__ valid, int64_t ldv,
                const double* __restrict__ label, int64_t n, int ncols, const double* __restrict__ shifts,
                double* __restrict__ scratch) {
    extern __shared__ __align__(16) double sm[];
    double* e_tile = sm;                  // [FIT_TILE]
    double* acc = sm + FIT_TILE;          // [(ncols+1) * NS]
    double* comp = acc + (ncols + 1) * NS;  // [(ncols+1) * 4]  compensation of stats 1..4

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int nacc = (ncols + 1) * NS;
    for (int i = tid; i < nacc; i += FIT_THREADS) {
        const int st = i % NS;
        acc[i] = st == 5 ? INFINITY : (st == 6 ? -INFINITY : 0.0);
    }
    for (int i = tid; i < (ncols + 1) * 4; i += FIT_THREADS) comp[i] = 0.0;

    const double sy = shifts[ncols];
    const int64_t ntiles = (n + FIT_TILE - 1) / FIT_TILE;
    for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        const int64_t r0 = tile * FIT_TILE;
        const int rows = (int)min((int64_t)FIT_TILE, n - r0);
        __syncthreads();  // readers of the previous tile are done (also orders the init above)
        for (int i = tid; i < FIT_TILE; i += FIT_THREADS) e_tile[i] = i < rows ? label[r0 + i] - sy : 0.0;
        __syncthreads();

        for (int c = warp; c <= ncols; c += FIT_WARPS) {
            const bool is_label = (c == ncols);
            const double* col = is_label ? label : values + (int64_t)c * ld;
            const uint64_t* vcol = (is_label || valid == nullptr) ? nullptr : valid + (int64_t)c * ldv;
            const double s = shifts[c];
            Acc a, b;
            acc_init(a);
            acc_init(b);
            unsigned cnt = 0;
            if (rows == FIT_TILE) {
                const unsigned long long myw = vcol ? vcol[(r0 >> 6) + lane] : ~0ull;  // 32 words = this tile
                cnt = __popcll(myw);
                const double2* p = reinterpret_cast<const double2*>(col + r0) + lane;
                const double2* ep = reinterpret_cast<const double2*>(e_tile) + lane;
#pragma unroll 1
                for (int i0 = 0; i0 < 32; i0 += 8) {
                    double2 x[8];
#pragma unroll
                    for (int k = 0; k < 8; ++k) x[k] = ldg_stream(p + (i0 + k) * 32);
#pragma unroll
                    for (int k = 0; k < 8; ++k) {
                        const unsigned long long w = __shfl_sync(FULL, myw, i0 + k);
                        const unsigned bits = (unsigned)(w >> (2 * lane)) & 3u;
                        const double2 e = ep[(i0 + k) * 32];
                        acc_add(a, bits & 1u, x[k].x, s, e.x);
                        acc_add(b, bits & 2u, x[k].y, s, e.y);
                    }
                }
            } else {  // ragged last tile
                for (int i = lane; i < rows; i += 32) {
                    const int64_t r = r0 + i;
                    const bool v = vcol ? ((vcol[r >> 6] >> (r & 63)) & 1ull) != 0 : true;
                    cnt += v;
                    acc_add(a, v, col[r], s, e_tile[i]);
                }
            }
            const double sd = warp_sum(a.sd + b.sd);
            const double sdd = warp_sum(a.sdd + b.sdd);
            const double sde = warp_sum(a.sde + b.sde);
            const double se = warp_sum(a.se + b.se);
            const double mn = warp_min(fmin(a.mn, b.mn));
            const double mx = warp_max(fmax(a.mx, b.mx));
            const unsigned total = __reduce_add_sync(FULL, cnt);
            if (lane == 0) {
                double* q = acc + c * NS;
                double* k = comp + c * 4;
                q[0] += (double)total;
                kahan_add(q[1], k[0], sd);
                kahan_add(q[2], k[1], sdd);
                kahan_add(q[3], k[2], sde);
                kahan_add(q[4], k[3], se);
                q[5] = fmin(q[5], mn);
                q[6] = fmax(q[6], mx);
            }
        }
    }
    __syncthreads();
    double* out = scratch + (int64_t)blockIdx.x * nacc;
    for (int i = tid; i < nacc; i += FIT_THREADS) {
        const int st = i % NS;
        double v = acc[i];
        if (st >= 1 && st <= 4) v -= comp[(i / NS) * 4 + (st - 1)];
        out[i] = v;
    }
}

// ---------------------------------------------------------------------------------------------------------------------
// real_fit_tma_kernel: the same scan with the columns brought in by the bulk-copy engine (cp.async.bulk + mbarrier).
//
// A CTA (one per SM) owns a contiguous run of full sub-tiles of FT_SUB rows (the runs of different CTAs differ by at
// most one sub-tile) and walks it in super-tiles of FT_SUBS sub-tiles.  Work unit = (super-tile, column), owned by
// consumer warp c % FT_WARPS.  Every consumer warp has its own FT_STAGES buffers of one sub-tile: while it works on one
// sub-tile out of shared memory, the bulk copy of its next one (next sub-tile, next column or next super-tile) is in
// flight, so the loads outstanding per SM do not depend on registers or on how the warps are scheduled.  Inside a
// sub-tile lane L owns FT_SUB / 32 consecutive rows: their validity is the lane's own piece of the bitmap (one
// coalesced load, no shuffles; rotated once so that every bit test is an immediate mask), and the lane reads its rows
// as 128-bit pairs in the rotated order p = (k + L) mod PAIRS, which is bank-conflict free.  The label of a super-tile
// is staged once: an extra warp issues its copy into a two-deep ring and turns it into e = y - shift_y in place, one
// super-tile ahead; consumers release a slot through an mbarrier that counts units.  A column's accumulators stay in
// registers over the unit's sub-tiles and are reduced across lanes once per unit; per-CTA accumulators, the scratch
// layout and the fixed-order reduce are those of real_fit_kernel, so the result is deterministic for a given grid.
// Rows past the last full sub-tile are read directly by the last CTA.
__device__ __forceinline__ unsigned smem_u32(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(unsigned bar, unsigned count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(unsigned bar, unsigned bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(unsigned bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned bar, unsigned parity) {
    unsigned ok = 0;
    do {
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(ok)
            : "r"(bar), "r"(parity)
            : "memory");
    } while (!ok);
}
// global -> shared bulk copy (16-byte aligned, size a multiple of 16) that completes on an mbarrier of this CTA
__device__ __forceinline__ void bulk_load(unsigned dst, const void* src, unsigned bytes, unsigned bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst),
                 "l"(src), "r"(bytes), "r"(bar)
                 : "memory");
}

__device__ __forceinline__ void acc_add_lean(Acc& a, bool v, double x, double s, double e) {
    const double d = v ? x - s : 0.0;
    a.sd += d;
    a.sdd = fma(d, d, a.sdd);
    a.sde = fma(d, e, a.sde);  // d == 0 on a null row
    a.se += v ? e : 0.0;
    a.mn = (v & (x < a.mn)) ? x : a.mn;  // branch-free: one predicated compare + select
    a.mx = (v & (x > a.mx)) ? x : a.mx;
}

struct FitChunkIt {  // the sub-tiles of one consumer warp, in the order it works through them
    int64_t st;      // first sub-tile of the super-tile
    int i;           // ordinal of the super-tile within this CTA
    int c;           // column (ncols = the label)
    int j;           // sub-tile within the super-tile
    int nfull;       // sub-tiles of this super-tile
};

// INT64: the value columns hold int64 (Integral features), converted to double as they are read; the label is f64.
template <int FT_SUB, int FT_SUBS, int FT_WARPS, int FT_STAGES, bool INT64>
__global__ void __launch_bounds__((FT_WARPS + 1) * 32, 1)
real_fit_tma_kernel(const double* __restrict__ values, int64_t ld, const uint64_t* __restrict__ valid, int64_t ldv,
                    const double* __restrict__ label, int64_t n, int ncols, const double* __restrict__ shifts,
                    double* __restrict__ scratch) {
    constexpr int FT_SUPER = FT_SUB * FT_SUBS;
    constexpr int FT_THREADS = (FT_WARPS + 1) * 32;
    constexpr int RPL = FT_SUB / 32;   // rows per lane and sub-tile: the lane's own RPL validity bits
    constexpr int PAIRS = RPL / 2;
    static_assert(RPL == 32 || RPL == 16, "a lane owns 16 or 32 rows of a sub-tile");
    extern __shared__ __align__(128) unsigned char ft_smem[];
    double* colbuf = reinterpret_cast<double*>(ft_smem);        // [FT_WARPS][FT_STAGES][FT_SUB]
    double* ybuf = colbuf + FT_WARPS * FT_STAGES * FT_SUB;      // [2][FT_SUPER]
    double* acc = ybuf + 2 * FT_SUPER;                          // [(ncols+1) * NS]
    double* comp = acc + (ncols + 1) * NS;                      // [(ncols+1) * 4]
    unsigned long long* bars = reinterpret_cast<unsigned long long*>(comp + (ncols + 1) * 4);
    const unsigned bar0 = smem_u32(bars);
    // bars: [w * FT_STAGES + stage] column buffer full | label full [2] | label empty [2] | e ready [2]
    const unsigned bar_yfull = bar0 + 8u * (FT_WARPS * FT_STAGES), bar_yempty = bar_yfull + 16u;
    const unsigned bar_eready = bar_yempty + 16u;

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int nacc = (ncols + 1) * NS;
    for (int i = tid; i < nacc; i += FT_THREADS) {
        const int st = i % NS;
        acc[i] = st == 5 ? INFINITY : (st == 6 ? -INFINITY : 0.0);
    }
    for (int i = tid; i < (ncols + 1) * 4; i += FT_THREADS) comp[i] = 0.0;
    if (tid == 0) {
        for (int b = 0; b < FT_WARPS * FT_STAGES + 2; ++b) mbar_init(bar0 + 8u * b, 1);
        mbar_init(bar_yempty, (unsigned)(ncols + 1));
        mbar_init(bar_yempty + 8u, (unsigned)(ncols + 1));
        mbar_init(bar_eready, 1);
        mbar_init(bar_eready + 8u, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    __syncthreads();

    // the full sub-tiles are dealt out in contiguous, equally long runs (they differ by at most one sub-tile)
    const int64_t rows_full = (n / FT_SUB) * FT_SUB;
    const int64_t tiles = n / FT_SUB, t_share = tiles / gridDim.x, t_rem = tiles % gridDim.x;
    const int64_t t_begin = blockIdx.x * t_share + min((int64_t)blockIdx.x, t_rem);
    const int64_t t_end = t_begin + t_share + (blockIdx.x < t_rem ? 1 : 0);
    const double sy = shifts[ncols];

    if (warp == FT_WARPS) {
        // label producer: copies the label of a super-tile into a two-deep ring and turns it into e = y - shift_y
        // in place, one super-tile ahead of the consumers
        const int nsup = (int)((t_end - t_begin + FT_SUBS - 1) / FT_SUBS);
        auto issue_y = [&](int i) {
            if (lane == 0) {
                const int64_t t0 = t_begin + (int64_t)i * FT_SUBS;
                const unsigned bytes = (unsigned)min((int64_t)FT_SUBS, t_end - t0) * (FT_SUB * 8u);
                mbar_expect_tx(bar_yfull + 8u * (i & 1), bytes);
                bulk_load(smem_u32(ybuf + (i & 1) * FT_SUPER), label + t0 * FT_SUB, bytes, bar_yfull + 8u * (i & 1));
            }
        };
        auto convert_y = [&](int i) {
            mbar_wait(bar_yfull + 8u * (i & 1), (i >> 1) & 1);
            const int64_t t0 = t_begin + (int64_t)i * FT_SUBS;
            const int pairs = (int)min((int64_t)FT_SUBS, t_end - t0) * (FT_SUB / 2);
            double2* yb = reinterpret_cast<double2*>(ybuf + (i & 1) * FT_SUPER);
            for (int k = lane; k < pairs; k += 32) {
                double2 v = yb[k];
                v.x -= sy;
                v.y -= sy;
                yb[k] = v;
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(bar_eready + 8u * (i & 1));
        };
        if (nsup > 0) issue_y(0);
        if (nsup > 1) issue_y(1);
        if (nsup > 0) convert_y(0);
        for (int i = 0; i + 1 < nsup; ++i) {
            convert_y(i + 1);
            if (i + 2 < nsup) {
                mbar_wait(bar_yempty + 8u * (i & 1), (i >> 1) & 1);   // every unit of super-tile i is done
                // the slot was last written through the generic proxy (convert_y): order that before the bulk copy
                asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                issue_y(i + 2);
            }
        }
    } else {
        const unsigned my_bar = bar0 + 8u * (warp * FT_STAGES);
        double* my_buf = colbuf + warp * FT_STAGES * FT_SUB;
        auto nfull_of = [&](int64_t t0) { return (int)min((int64_t)FT_SUBS, t_end - t0); };
        auto it_valid = [&](const FitChunkIt& it) { return it.st < t_end && it.c <= ncols; };
        auto it_next = [&](FitChunkIt& it) {
            if (++it.j < it.nfull) return;
            it.j = 0;
            it.c += FT_WARPS;
            if (it.c <= ncols) return;
            it.c = warp;
            it.st += FT_SUBS;
            ++it.i;
            it.nfull = it.st < t_end ? nfull_of(it.st) : 0;
        };
        // issue the bulk copy of chunk `it` into stage q & 1 and fetch the lane's validity word for it
        auto issue = [&](const FitChunkIt& it, unsigned q) -> unsigned {
            const bool is_label = it.c == ncols;
            const int64_t r0 = (it.st + it.j) * FT_SUB;
            const double* src = (is_label ? label : values + (int64_t)it.c * ld) + r0;
            if (lane == 0) {
                const unsigned stg = q % FT_STAGES;
                const unsigned bar = my_bar + 8u * stg;
                mbar_expect_tx(bar, FT_SUB * 8u);
                bulk_load(smem_u32(my_buf + stg * FT_SUB), src, FT_SUB * 8u, bar);
            }
            if (is_label || valid == nullptr) return RPL == 32 ? 0xffffffffu : 0xffffu;
            if (RPL == 32) {
                const unsigned* v32 = reinterpret_cast<const unsigned*>(valid + (int64_t)it.c * ldv);
                return __ldg(v32 + (r0 >> 5) + lane);
            }
            const unsigned short* v16 = reinterpret_cast<const unsigned short*>(valid + (int64_t)it.c * ldv);
            return __ldg(v16 + (r0 >> 4) + lane);
        };

        FitChunkIt cur{t_begin, 0, warp, 0, 0};
        cur.nfull = cur.st < t_end ? nfull_of(cur.st) : 0;
        FitChunkIt nxt = cur;
        // chunk q lives in stage q % FT_STAGES; FT_STAGES - 1 copies are in flight while one chunk is worked on
        unsigned q = 0, qi = 0;
        unsigned wq[FT_STAGES];
#pragma unroll
        for (int k = 0; k < FT_STAGES - 1; ++k) {
            wq[k] = 0;
            if (it_valid(nxt)) {
                wq[k] = issue(nxt, qi++);
                it_next(nxt);
            }
        }
        wq[FT_STAGES - 1] = 0;
        Acc a, b;
        unsigned cnt = 0;
        double s = 0.0;
        while (it_valid(cur)) {
            if (it_valid(nxt)) {  // its stage was released by the __syncwarp that ended chunk q - 1
                wq[FT_STAGES - 1] = issue(nxt, qi++);
                it_next(nxt);
            }
            const unsigned w_cur = wq[0];
            if (cur.j == 0) {
                acc_init(a);
                acc_init(b);
                cnt = 0;
                s = shifts[cur.c];
                mbar_wait(bar_eready + 8u * (cur.i & 1), (cur.i >> 1) & 1);
            }
            mbar_wait(my_bar + 8u * (q % FT_STAGES), (q / FT_STAGES) & 1u);
            {
                const double2* xb =
                    reinterpret_cast<const double2*>(my_buf + (q % FT_STAGES) * FT_SUB) + lane * PAIRS;
                const double2* yb =
                    reinterpret_cast<const double2*>(ybuf + (cur.i & 1) * FT_SUPER + cur.j * FT_SUB) + lane * PAIRS;
                cnt += __popc(w_cur);
                // rotate the lane's bits so that bits 2k, 2k+1 belong to pair (k + lane) % PAIRS
                const unsigned w2 = RPL == 32 ? w_cur : (w_cur | (w_cur << 16));
                const unsigned wr = __funnelshift_r(w2, w2, (2 * lane) & 31);
#pragma unroll
                for (int k = 0; k < PAIRS; ++k) {
                    const int p = (k + lane) & (PAIRS - 1);
                    double2 x = xb[p];
                    const double2 y = yb[p];
                    if (INT64 && cur.c < ncols) {
                        x.x = (double)__double_as_longlong(x.x);
                        x.y = (double)__double_as_longlong(x.y);
                    }
                    acc_add_lean(a, (wr >> (2 * k)) & 1u, x.x, s, y.x);
                    acc_add_lean(b, (wr >> (2 * k + 1)) & 1u, x.y, s, y.y);
                }
            }
            __syncwarp();
            if (cur.j == cur.nfull - 1) {
                const double sd = warp_sum(a.sd + b.sd);
                const double sdd = warp_sum(a.sdd + b.sdd);
                const double sde = warp_sum(a.sde + b.sde);
                const double se = warp_sum(a.se + b.se);
                const double mn = warp_min(fmin(a.mn, b.mn));
                const double mx = warp_max(fmax(a.mx, b.mx));
                const unsigned total = __reduce_add_sync(FULL, cnt);
                if (lane == 0) {
                    double* qa = acc + cur.c * NS;
                    double* kc = comp + cur.c * 4;
                    qa[0] += (double)total;
                    kahan_add(qa[1], kc[0], sd);
                    kahan_add(qa[2], kc[1], sdd);
                    kahan_add(qa[3], kc[2], sde);
                    kahan_add(qa[4], kc[3], se);
                    qa[5] = fmin(qa[5], mn);
                    qa[6] = fmax(qa[6], mx);
                    mbar_arrive(bar_yempty + 8u * (cur.i & 1));
                }
            }
            it_next(cur);
#pragma unroll
            for (int k = 0; k < FT_STAGES - 1; ++k) wq[k] = wq[k + 1];
            ++q;
        }
        // rows past the last full sub-tile: read directly, by the last CTA
        if (blockIdx.x == gridDim.x - 1 && rows_full < n) {
            for (int c = warp; c <= ncols; c += FT_WARPS) {
                const bool is_label = (c == ncols);
                const double* col = is_label ? label : values + (int64_t)c * ld;
                const uint64_t* vcol = (is_label || valid == nullptr) ? nullptr : valid + (int64_t)c * ldv;
                const double sc = shifts[c];
                acc_init(a);
                unsigned tc = 0;
                for (int64_t r = rows_full + lane; r < n; r += 32) {
                    const bool v = vcol ? ((vcol[r >> 6] >> (r & 63)) & 1ull) != 0 : true;
                    tc += v;
                    const double xv = (INT64 && !is_label) ? (double)__double_as_longlong(col[r]) : col[r];
                    acc_add(a, v, xv, sc, label[r] - sy);
                }
                const double sd = warp_sum(a.sd), sdd = warp_sum(a.sdd), sde = warp_sum(a.sde), se = warp_sum(a.se);
                const double mn = warp_min(a.mn), mx = warp_max(a.mx);
                const unsigned total = __reduce_add_sync(FULL, tc);
                if (lane == 0) {
                    double* qa = acc + c * NS;
                    double* kc = comp + c * 4;
                    qa[0] += (double)total;
                    kahan_add(qa[1], kc[0], sd);
                    kahan_add(qa[2], kc[1], sdd);
                    kahan_add(qa[3], kc[2], sde);
                    kahan_add(qa[4], kc[3], se);
                    qa[5] = fmin(qa[5], mn);
                    qa[6] = fmax(qa[6], mx);
                }
            }
        }
    }
    __syncthreads();
    double* out = scratch + (int64_t)blockIdx.x * nacc;
    for (int i = tid; i < nacc; i += FT_THREADS) {
        const int st = i % NS;
        double v = acc[i];
        if (st >= 1 && st <= 4) v -= comp[(i / NS) * 4 + (st - 1)];
        out[i] = v;
    }
}

template <int SUB, int SUBS, int WARPS, int STAGES, bool INT64>
static int launch_fit_tma(const double* values, int64_t ld, const uint64_t* valid, int64_t ldv, const double* label,
                          int64_t n, int ncols, const double* shifts, double* scratch, int nblocks, cudaStream_t st,
                          bool* fits) {
    const size_t smem =
        sizeof(double) * ((size_t)WARPS * STAGES * SUB + 2 * SUB * SUBS + (size_t)(ncols + 1) * (NS + 4)) +
        8 * (size_t)(WARPS * STAGES + 6);
    *fits = smem <= 227 * 1024;
    if (!*fits) return TMOG_OK;
    auto kern = real_fit_tma_kernel<SUB, SUBS, WARPS, STAGES, INT64>;
    TMOG_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    kern<<<nblocks, (WARPS + 1) * 32, smem, st>>>(values, ld, valid, ldv, label, n, ncols, shifts, scratch);
    return TMOG_OK;
}

// Reduce the per-CTA partials in a fixed order (deterministic for a given grid): one warp per column.
__global__ void __launch_bounds__(32)
real_fit_reduce_kernel(const double* __restrict__ scratch, int nblocks, int ncols, int64_t n,
                       double* __restrict__ partials) {
    const int c = blockIdx.x, lane = threadIdx.x;
    const int nacc = (ncols + 1) * NS;
    double s[5] = {0, 0, 0, 0, 0};
    double mn = INFINITY, mx = -INFINITY;
    for (int b = lane; b < nblocks; b += 32) {
        const double* q = scratch + (int64_t)b * nacc + c * NS;
#pragma unroll
        for (int k = 0; k < 5; ++k) s[k] += q[k];
        mn = fmin(mn, q[5]);
        mx = fmax(mx, q[6]);
    }
#pragma unroll
    for (int k = 0; k < 5; ++k) s[k] = warp_sum(s[k]);
    mn = warp_min(mn);
    mx = warp_max(mx);
    if (lane == 0) {
        double* o = partials + c * NS;
#pragma unroll
        for (int k = 0; k < 5; ++k) o[k] = s[k];
        o[5] = mn;
        o[6] = mx;
        o[7] = (double)n;
    }
}

__global__ void __launch_bounds__(128)
real_pilot_shift_kernel(const double* __restrict__ values, int64_t ld, const uint64_t* __restrict__ valid, int64_t ldv,
                        const double* __restrict__ label, int64_t n, int ncols, double* __restrict__ shifts,
                        int values_are_int64) {
    __shared__ double ssum[4];
    __shared__ unsigned scnt[4];
    const int c = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const bool is_label = (c == ncols);
    const double* col = is_label ? label : values + (int64_t)c * ld;
    const uint64_t* vcol = (is_label || valid == nullptr) ? nullptr : valid + (int64_t)c * ldv;
    const int rows = (int)min((int64_t)PILOT_ROWS, n);
    double s = 0.0;
    unsigned cnt = 0;
    for (int r = tid; r < rows; r += 128) {
        const bool v = vcol ? ((vcol[r >> 6] >> (r & 63)) & 1ull) != 0 : true;
        const double x = (values_are_int64 && !is_label) ? (double)__double_as_longlong(col[r]) : col[r];
        if (v && isfinite(x)) {
            s += x;
            ++cnt;
        }
    }
    s = warp_sum(s);
    cnt = __reduce_add_sync(FULL, cnt);
    if (lane == 0) {
        ssum[warp] = s;
        scnt[warp] = cnt;
    }
    __syncthreads();
    if (tid == 0) {
        const double t = ssum[0] + ssum[1] + ssum[2] + ssum[3];
        const unsigned k = scnt[0] + scnt[1] + scnt[2] + scnt[3];
        shifts[c] = k > 0 ? t / (double)k : 0.0;
    }
}

// Column-major -> row-major transposing writers.  A CTA owns 64 rows and walks the column tiles (32 input columns
// each) of those rows, so one CTA writes whole output rows back to back.
//
// vec_transform_kernel: every emitted column is written (the vectorizer's own transformFn).  Per tile the loads are
// 256-byte row runs of one column per warp instruction into a padded shared tile; the stores are contiguous 256-byte
// runs of one output row per warp instruction (lane = emitted column).
template <typename T, bool TRACK>
__global__ void __launch_bounds__(256)
vec_transform_kernel(const T* __restrict__ values, int64_t ld, const uint64_t* __restrict__ valid, int64_t ldv,
                     int64_t n, int ncols, const double* __restrict__ fills, double fill_const,
                     double* __restrict__ out, int64_t out_ld, int64_t out_col0) {
    __shared__ double tile[32][65];
    __shared__ unsigned long long vbits[32];
    constexpr int PER = TRACK ? 2 : 1;  // emitted columns per input column
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int64_t r0 = (int64_t)blockIdx.x * 64;
    const int rows_here = (int)min((int64_t)64, n - r0);
    const int ntiles = (ncols + 31) / 32;
    for (int t = 0; t < ntiles; ++t) {
        const int c0 = t * 32;
        const int tc = min(32, ncols - c0);
        if (t > 0) __syncthreads();  // previous tile fully written before the shared tile is overwritten
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const int cl = warp + 8 * k;
            if (cl < tc) {
                const int c = c0 + cl;
                const T* col = values + (int64_t)c * ld + r0;
                tile[cl][lane] = lane < rows_here ? (double)col[lane] : 0.0;
                tile[cl][lane + 32] = lane + 32 < rows_here ? (double)col[lane + 32] : 0.0;
                if (lane == 0) vbits[cl] = valid ? valid[(int64_t)c * ldv + (r0 >> 6)] : ~0ull;
            }
        }
        __syncthreads();
#pragma unroll 1
        for (int h = 0; h < PER; ++h) {
            const int src = lane + 32 * h;      // emitted column of the tile served by this lane
            if (src >= tc * PER) continue;
            const int cl = TRACK ? (src >> 1) : src;
            const bool is_ind = TRACK && (src & 1);
            const double fill = fills ? fills[c0 + cl] : fill_const;
            const unsigned long long vb = vbits[cl];
            const double* trow = &tile[cl][warp];
            double* optr = out + (r0 + warp) * out_ld + out_col0 + c0 * PER + src;
            const int64_t ostep = 8 * out_ld;
#pragma unroll 4
            for (int k = 0; k < 8; ++k) {
                const int rl = warp + 8 * k;
                if (rl < rows_here) {
                    const bool v = (vb >> rl) & 1ull;
                    double x = trow[8 * k];
                    if (sizeof(T) == 1) x = x != 0.0 ? 1.0 : 0.0;
                    *optr = is_ind ? (v ? 0.0 : 1.0) : (v ? x : fill);
                }
                optr += ostep;
            }
        }
    }
}

// vec_transform_select_kernel: only the columns that survive SanityChecker are written, packed (keep_mask: bit i set
// <=> emitted column i is kept).  Lane = row on the load side; every value goes straight to its OUTPUT position in a
// shared-memory ring of row buffers (the transpose); after each tile the warps flush whole 32-byte-aligned groups of
// every row with contiguous 256-byte stores and carry the (< 4 values) tail to the next tile, so that DRAM only sees
// full-sector writes although a dropped column shifts everything behind it by an odd number of slots.
constexpr int TR_ROWS = 64;
constexpr int TR_RING = 72;     // >= 3 carried + 64 new positions
constexpr int TR_STRIDE = 73;   // odd row stride: lane = row accesses are bank-conflict free

template <typename T, bool TRACK>
__global__ void __launch_bounds__(256)
vec_transform_select_kernel(const T* __restrict__ values, int64_t ld, const uint64_t* __restrict__ valid,
                            int64_t ldv, int64_t n, int ncols, const double* __restrict__ fills, double fill_const,
                            double* __restrict__ out, int64_t out_ld, int64_t out_col0,
                            const unsigned long long* __restrict__ keep_mask) {
    __shared__ double ring[TR_ROWS * TR_STRIDE];
    constexpr int PER = TRACK ? 2 : 1;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int64_t r0 = (int64_t)blockIdx.x * TR_ROWS;
    const int rows_here = (int)min((int64_t)TR_ROWS, n - r0);
    const int ntiles = (ncols + 31) / 32;
    // the output block may start anywhere and rows need not be 32-byte multiples: flush groups are aligned on the
    // actual address of every row
    const int mis0 = (int)(((reinterpret_cast<uintptr_t>(out) >> 3) + (uintptr_t)out_col0) & 3u);
    int plo = 0;    // packed output position of the tile's first kept column (= end of the previous tile)
    int ring0 = 0;  // multiple of TR_RING: slot(p) = p - ring0, minus TR_RING once past the end
    for (int t = 0; t < ntiles; ++t) {
        const int c0 = t * 32;
        const int tc = min(32, ncols - c0);
        const int nbits = tc * PER;
        unsigned long long m = nbits == 64 ? ~0ull : ((1ull << nbits) - 1ull);
        const unsigned long long w = keep_mask[(c0 * PER) >> 6];
        m &= TRACK ? w : (w >> ((c0 * PER) & 63));
        const int cnt = __popcll(m);
        // ---- load + transpose: warp w takes columns w, w+8, w+16, w+24 of the tile; lane = row (and row + 32)
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const int cl = warp + 8 * k;
            if (cl < tc) {
                const int c = c0 + cl;
                const unsigned long long mx = TRACK ? (m >> (2 * cl)) & 3ull : (m >> cl) & 1ull;
                if (mx == 0) continue;  // the whole input column was dropped: never read it
                const T* col = values + (int64_t)c * ld + r0;
                const unsigned long long vb = valid ? valid[(int64_t)c * ldv + (r0 >> 6)] : ~0ull;
                const double fill = fills ? fills[c] : fill_const;
                int px = plo + __popcll(m & ((1ull << (PER * cl)) - 1ull)) - ring0;  // slot of the value column
                int pq = px + (int)(mx & 1ull);                                       // slot of the null indicator
                px -= px >= TR_RING ? TR_RING : 0;
                pq -= pq >= TR_RING ? TR_RING : 0;
#pragma unroll
                for (int h = 0; h < 2; ++h) {
                    const int rl = lane + 32 * h;
                    double x = rl < rows_here ? (double)col[rl] : 0.0;
                    if (sizeof(T) == 1) x = x != 0.0 ? 1.0 : 0.0;
                    const bool v = (vb >> rl) & 1ull;
                    if (mx & 1ull) ring[rl * TR_STRIDE + px] = v ? x : fill;
                    if (TRACK && (mx & 2ull)) ring[rl * TR_STRIDE + pq] = v ? 0.0 : 1.0;
                }
            }
        }
        __syncthreads();
        // ---- flush per row [lo, hi): from where the previous tile stopped up to the last 32-byte boundary of this
        //      row (to the very end on the last tile); the < 4 values behind it wait in the ring for the next tile
        const int end = plo + cnt;
        const bool last = t + 1 == ntiles;
#pragma unroll 2
        for (int k = 0; k < TR_ROWS / 8; ++k) {
            const int rl = warp + 8 * k;
            if (rl < rows_here) {
                const int mis = (mis0 + (int)(((r0 + rl) * out_ld) & 3)) & 3;
                const int lo = t == 0 ? 0 : max(0, ((plo + mis) & ~3) - mis);
                const int hi = last ? end : ((end + mis) & ~3) - mis;
                double* orow = out + (r0 + rl) * out_ld + out_col0;
                const double* rrow = ring + rl * TR_STRIDE;
                for (int p = lo + lane; p < hi; p += 32) {
                    int sl = p - ring0;
                    sl -= sl >= TR_RING ? TR_RING : 0;
                    orow[p] = rrow[sl];
                }
            }
        }
        plo = end;
        ring0 = (max(end - 3, 0) / TR_RING) * TR_RING;  // every unflushed position is >= end - 3
        if (!last) __syncthreads();  // ring slots of flushed positions are reused by the next tile
    }
}

// vec_transform_pack_kernel: the select transform for the common case of dense output rows (out_ld == kept columns,
// no column offset).  A CTA takes 32 rows and ALL columns: every value goes straight to its packed position in a
// shared-memory image of the CTA's stretch of the output (row stride padded to an odd number of doubles, so that
// lane = row accesses are conflict free), and after one barrier the image is copied out as one contiguous run of
// 32 * nkeep doubles -- the stretch starts on a 256-byte multiple, so DRAM sees full sectors only and no row needs
// its own alignment handling.
constexpr int TP_ROWS = 32;
constexpr int TP_THREADS = 384;
constexpr int TP_BATCH = 4;       // loads a warp keeps in flight

template <typename T, bool TRACK>
__global__ void __launch_bounds__(TP_THREADS)
vec_transform_pack_kernel(const T* __restrict__ values, int64_t ld, const uint64_t* __restrict__ valid, int64_t ldv,
                          int64_t n, int ncols, const double* __restrict__ fills, double fill_const,
                          double* __restrict__ out, int nkeep, int stride,
                          const unsigned long long* __restrict__ keep_mask) {
    extern __shared__ double tp_tile[];        // [TP_ROWS][stride], then the packed position of every input column
    int* pos_of = reinterpret_cast<int*>(tp_tile + TP_ROWS * stride);     // -1: the whole column was dropped
    constexpr int PER = TRACK ? 2 : 1;
    constexpr int NW = TP_THREADS / 32;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int64_t r0 = (int64_t)blockIdx.x * TP_ROWS;
    const int rows_here = (int)min((int64_t)TP_ROWS, n - r0);
    for (int c = tid; c < ncols; c += TP_THREADS) {
        const int bit = c * PER;
        const unsigned long long w = keep_mask[bit >> 6];
        const unsigned mx = (unsigned)(w >> (bit & 63)) & (TRACK ? 3u : 1u);
        int pos = __popcll(w & ((1ull << (bit & 63)) - 1ull));
        for (int q = 0; q < (bit >> 6); ++q) pos += __popcll(keep_mask[q]);
        pos_of[c] = mx ? (pos << 2) | (int)mx : -1;
    }
    __syncthreads();
    const int64_t r = r0 + lane;
    const bool row_ok = lane < rows_here;
    const int vshift = (int)(r0 & 63) + lane;
    double* trow = tp_tile + lane * stride;
    for (int cb = warp; cb < ncols; cb += NW * TP_BATCH) {
        double x[TP_BATCH];
        unsigned long long vb[TP_BATCH];
        int pm[TP_BATCH];
#pragma unroll
        for (int k = 0; k < TP_BATCH; ++k) {
            const int c = cb + k * NW;
            pm[k] = c < ncols ? pos_of[c] : -1;
            x[k] = 0.0;
            vb[k] = ~0ull;
            if (pm[k] >= 0) {
                if (row_ok) x[k] = (double)values[(int64_t)c * ld + r];
                if (valid) vb[k] = valid[(int64_t)c * ldv + (r0 >> 6)];
            }
        }
#pragma unroll
        for (int k = 0; k < TP_BATCH; ++k) {
            if (pm[k] < 0) continue;
            const int c = cb + k * NW;
            const int pos = pm[k] >> 2;
            const unsigned mx = (unsigned)pm[k] & 3u;
            double xv = x[k];
            if (sizeof(T) == 1) xv = xv != 0.0 ? 1.0 : 0.0;
            const bool v = (vb[k] >> vshift) & 1ull;
            const double fill = fills ? fills[c] : fill_const;
            if (mx & 1u) trow[pos] = v ? xv : fill;
            if (TRACK && (mx & 2u)) trow[pos + (int)(mx & 1u)] = v ? 0.0 : 1.0;
        }
    }
    // the image is read by the bulk-copy engine below: order the generic-proxy writes before its async-proxy reads
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    __syncthreads();
    double* dst = out + r0 * nkeep;
    const int total = rows_here * nkeep;
    if (stride == nkeep && ((total * (int)sizeof(double)) & 15) == 0) {
        // unpadded image (odd nkeep is conflict free as it is): one TMA bulk store of the whole stretch, issued by one
        // thread -- no per-element instructions, the SM's issue slots go to the other CTAs' loads
        if (tid == 0) {
            const unsigned src = (unsigned)__cvta_generic_to_shared(tp_tile);
            asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst), "r"(src),
                         "r"(total * (int)sizeof(double))
                         : "memory");
            asm volatile("cp.async.bulk.commit_group;" ::: "memory");
            // the CTA may go once the engine has READ the image (its shared memory is free again then); the global
            // writes complete by the end of the grid like any other store
            asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
        }
        return;
    }
    int rl = tid / nkeep, cl = tid - rl * nkeep;           // position of this thread's element; advanced by 512
    const int dr = TP_THREADS / nkeep, dc = TP_THREADS - dr * nkeep;
    for (int q = tid; q < total; q += TP_THREADS) {
        dst[q] = tp_tile[rl * stride + cl];
        rl += dr;
        cl += dc;
        if (cl >= nkeep) {
            cl -= nkeep;
            ++rl;
        }
    }
}

static bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

static int fit_grid() { return sm_count() * FIT_CTAS_PER_SM; }

static size_t fit_smem_bytes(int ncols) { return sizeof(double) * (size_t)(FIT_TILE + (ncols + 1) * (NS + 4)); }

static int check_columns(const void* values, int64_t ld, const uint64_t* valid, int64_t ldv, int64_t n, int ncols,
                         int elem_bytes) {
    TMOG_REQUIRE(n >= 0 && ncols >= 0, "n (%lld) and ncols (%d) must be non-negative", (long long)n, ncols);
    TMOG_REQUIRE(ncols == 0 || n == 0 || values != nullptr, "values is NULL");
    TMOG_REQUIRE(ld >= n, "column stride ld (%lld) is smaller than n (%lld)", (long long)ld, (long long)n);
    TMOG_REQUIRE(((ld * elem_bytes) & 15) == 0 && aligned16(values),
                 "columns must be 16-byte aligned (base pointer and ld * %d bytes)", elem_bytes);
    TMOG_REQUIRE(valid == nullptr || ldv >= (n + 63) / 64, "bitmap stride ldv (%lld) too small for n (%lld)",
                 (long long)ldv, (long long)n);
    return TMOG_OK;
}

}  // namespace tmog

using namespace tmog;

extern "C" int tmog_real_pilot_shift_dev(const double* values, int64_t ld, const uint64_t* valid, int64_t ldv,
                                         const double* label, int64_t n, int ncols, double* shifts, void* stream) {
    if (int rc = check_columns(values, ld, valid, ldv, n, ncols, 8)) return rc;
    TMOG_REQUIRE(shifts != nullptr && (label != nullptr || n == 0), "shifts / label is NULL");
    real_pilot_shift_kernel<<<ncols + 1, 128, 0, (cudaStream_t)stream>>>(values, ld, valid, ldv, label, n, ncols,
                                                                         shifts, 0);
    TMOG_LAUNCH_CHECK();
    return TMOG_OK;
}

extern "C" int tmog_real_fit_scratch_bytes(int ncols, size_t* bytes) {
    TMOG_REQUIRE(ncols >= 0 && bytes != nullptr, "bad arguments");
    *bytes = sizeof(double) * (size_t)fit_grid() * (size_t)(ncols + 1) * NS;
    return TMOG_OK;
}

extern "C" int tmog_real_fit_partials_dev(const double* values, int64_t ld, const uint64_t* valid, int64_t ldv,
                                          const double* label, int64_t n, int ncols, const double* shifts,
                                          void* scratch, size_t scratch_bytes, double* partials, void* stream) {
    if (int rc = check_columns(values, ld, valid, ldv, n, ncols, 8)) return rc;
    TMOG_REQUIRE(label != nullptr || n == 0, "label is NULL");
    TMOG_REQUIRE(aligned16(label), "label must be 16-byte aligned");
    TMOG_REQUIRE(shifts != nullptr && partials != nullptr, "shifts / partials is NULL");
    const int grid = fit_grid();
    const size_t need = sizeof(double) * (size_t)grid * (size_t)(ncols + 1) * NS;
    if (scratch == nullptr || scratch_bytes < need) {
        set_error("scratch too small: %zu bytes given, %zu needed (tmog_real_fit_scratch_bytes)", scratch_bytes, need);
        return TMOG_ENOMEM;
    }
    cudaStream_t st = (cudaStream_t)stream;
    // one CTA per SM, columns streamed through shared memory by the bulk-copy engine: 16 consumer warps, two 4 KB
    // buffers each (sub-tiles of 512 rows), super-tiles of 4096 rows.  Measured at 10 M x 100 (B200): 1.30 ms; 8 warps x
    // 8 KB buffers 1.45 ms, 12 warps x 3 buffers 1.39 ms, super-tiles of 2048 rows 1.37 ms.
    bool fits = false;
    int nblocks = sm_count();
    if (int rc = launch_fit_tma<512, 8, 16, 2, false>(values, ld, valid, ldv, label, n, ncols, shifts,
                                               static_cast<double*>(scratch), nblocks, st, &fits))
        return rc;
    if (!fits) {
        // very wide blocks: the register-streaming kernel, whose shared memory holds only the accumulators
        nblocks = grid;
        const size_t smem = fit_smem_bytes(ncols);
        TMOG_REQUIRE(smem <= 200 * 1024, "too many columns for one fit call (%d): split the column block", ncols);
        if (smem > 48 * 1024) {
            TMOG_CUDA(cudaFuncSetAttribute(real_fit_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        }
        real_fit_kernel<<<grid, FIT_THREADS, smem, st>>>(values, ld, valid, ldv, label, n, ncols, shifts,
                                                         static_cast<double*>(scratch));
    }
    TMOG_LAUNCH_CHECK();
    real_fit_reduce_kernel<<<ncols + 1, 32, 0, st>>>(static_cast<const double*>(scratch), nblocks, ncols, n,
                                                     partials);
    TMOG_LAUNCH_CHECK();
    return TMOG_OK;
}

extern "C" int tmog_real_partials_merge(double* acc, const double* other, int ncols) {
    TMOG_REQUIRE(acc != nullptr && other != nullptr && ncols >= 0, "bad arguments");
    for (int c = 0; c <= ncols; ++c) {
        double* a = acc + c * NS;
        const double* b = other + c * NS;
        for (int k = 0; k < 5; ++k) a[k] += b[k];
        a[5] = fmin(a[5], b[5]);
        a[6] = fmax(a[6], b[6]);
        a[7] += b[7];
    }
    return TMOG_OK;
}

static int real_finish(const double* partials, const double* shifts, int ncols, int fill_with_mean,
                       double fill_const, const double* fill_in, int track_nulls, double* fills, double* mean,
                       double* variance, double* mn, double* mx, double* cov_label, double* label_stats,
                       double* null_label_sum) {
    TMOG_REQUIRE(partials != nullptr && shifts != nullptr && ncols >= 0, "bad arguments");
    const double* L = partials + ncols * NS;
    const double n = L[7];
    const double sy = shifts[ncols];
    // label: e = y - sy over all rows
    const double eps = n > 0 ? L[1] / n : 0.0;  // mean_y - sy
    const double mean_y = sy + eps;
    const double m2y = n > 0 ? L[2] - L[1] * L[1] / n : 0.0;
    if (label_stats) {
        label_stats[0] = n;
        label_stats[1] = n > 0 ? mean_y : 0.0;
        label_stats[2] = n > 1 ? fmax(m2y, 0.0) / (n - 1) : 0.0;
        label_stats[3] = n > 0 ? L[5] : 0.0;
        label_stats[4] = n > 0 ? L[6] : 0.0;
    }
    const int w = track_nulls ? 2 : 1;
    for (int c = 0; c < ncols; ++c) {
        const double* q = partials + c * NS;
        const double v = q[0], sd = q[1], sdd = q[2], sde = q[3], sev = q[4];
        const double z = n - v;
        const double s = shifts[c];
        // SequenceAggregators.scala:84 -- s / c if c > 0 else s (= 0.0)
        const double fill = fill_in ? fill_in[c] : (fill_with_mean ? (v > 0 ? s + sd / v : 0.0) : fill_const);
        if (fills) fills[c] = fill;
        // filled column f = x (valid) | fill (null)
        const double mu = n > 0 ? (fill_with_mean ? fill : (v * s + sd + z * fill) / n) : 0.0;
        const double delta = mu - s;  // x - mu = d - delta
        double m2 = sdd - 2.0 * delta * sd + v * delta * delta + z * (fill - mu) * (fill - mu);
        if (m2 < 0) m2 = 0;
        const double se_null = L[1] - sev;  // sum of e over null rows
        const double cov = (sde - eps * sd - delta * sev + v * delta * eps) + (fill - mu) * (se_null - z * eps);
        const int o = c * w;
        if (mean) mean[o] = mu;
        if (variance) variance[o] = n > 1 ? m2 / (n - 1) : 0.0;
        if (mn) mn[o] = n > 0 ? (z > 0 ? fmin(v > 0 ? q[5] : fill, fill) : q[5]) : 0.0;
        if (mx) mx[o] = n > 0 ? (z > 0 ? fmax(v > 0 ? q[6] : fill, fill) : q[6]) : 0.0;
        if (cov_label) cov_label[o] = cov;
        if (null_label_sum) null_label_sum[c] = se_null + z * sy;
        if (track_nulls) {
            if (mean) mean[o + 1] = n > 0 ? z / n : 0.0;
            if (variance) variance[o + 1] = n > 1 ? (z * (n - z) / n) / (n - 1) : 0.0;
            if (mn) mn[o + 1] = (n > 0 && v == 0) ? 1.0 : 0.0;
            if (mx) mx[o + 1] = z > 0 ? 1.0 : 0.0;
            if (cov_label) cov_label[o + 1] = se_null - z * eps;
        }
    }
    return TMOG_OK;
}

extern "C" int tmog_real_finish(const double* partials, const double* shifts, int ncols, int fill_with_mean,
                                double fill_const, int track_nulls, double* fills, double* mean, double* variance,
                                double* mn, double* mx, double* cov_label, double* label_stats,
                                double* null_label_sum) {
    return real_finish(partials, shifts, ncols, fill_with_mean, fill_const, nullptr, track_nulls, fills, mean, variance,
                       mn, mx, cov_label, label_stats, null_label_sum);
}

extern "C" int tmog_real_finish_with_fills(const double* partials, const double* shifts, int ncols,
                                           const double* fill_values, int track_nulls, double* mean, double* variance,
                                           double* mn, double* mx, double* cov_label, double* label_stats,
                                           double* null_label_sum) {
    TMOG_REQUIRE(fill_values != nullptr || ncols == 0, "fill_values is NULL");
    return real_finish(partials, shifts, ncols, 0, 0.0, fill_values, track_nulls, nullptr, mean, variance, mn, mx,
                       cov_label, label_stats, null_label_sum);
}

template <typename T>
static int launch_transform(const T* values, int64_t ld, const uint64_t* valid, int64_t ldv, int64_t n, int ncols,
                            const double* fills, double fill_const, int track_nulls, double* out, int64_t out_ld,
                            int64_t out_col0, void* stream, const uint64_t* keep = nullptr, int nkeep = 0) {
    if (int rc = check_columns(values, ld, valid, ldv, n, ncols, (int)sizeof(T))) return rc;
    const int64_t width = keep ? nkeep : (int64_t)ncols * (track_nulls ? 2 : 1);
    TMOG_REQUIRE(out != nullptr || n == 0 || width == 0, "out is NULL");
    TMOG_REQUIRE(out_col0 >= 0 && out_ld >= out_col0 + width, "output block [%lld, %lld) exceeds row stride %lld",
                 (long long)out_col0, (long long)(out_col0 + width), (long long)out_ld);
    if (n == 0 || ncols == 0 || width == 0) return TMOG_OK;
    const int64_t gx = (n + 63) / 64;
    TMOG_REQUIRE(gx <= 0x7fffffffLL, "too many rows for one transform call (%lld): chunk the batch", (long long)n);
    cudaStream_t st = (cudaStream_t)stream;
    const unsigned long long* km = reinterpret_cast<const unsigned long long*>(keep);
    if (km == nullptr) {
        if (track_nulls)
            vec_transform_kernel<T, true><<<(unsigned)gx, 256, 0, st>>>(values, ld, valid, ldv, n, ncols, fills,
                                                                         fill_const, out, out_ld, out_col0);
        else
            vec_transform_kernel<T, false><<<(unsigned)gx, 256, 0, st>>>(values, ld, valid, ldv, n, ncols, fills,
                                                                          fill_const, out, out_ld, out_col0);
    } else if (out_col0 == 0 && out_ld == width && (reinterpret_cast<uintptr_t>(out) & 31u) == 0 &&
               (size_t)TP_ROWS * (size_t)(width | 1) * sizeof(double) + 4 * (size_t)ncols <= 100 * 1024) {
        // dense output rows: the packing kernel
        const int stride = (int)width | 1;
        const size_t smem = (size_t)TP_ROWS * stride * sizeof(double) + 4 * (size_t)ncols;
        const int64_t gp = (n + TP_ROWS - 1) / TP_ROWS;
        TMOG_REQUIRE(gp <= 0x7fffffffLL, "too many rows for one transform call (%lld): chunk the batch", (long long)n);
        if (track_nulls) {
            TMOG_CUDA(cudaFuncSetAttribute(vec_transform_pack_kernel<T, true>,
                                           cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            vec_transform_pack_kernel<T, true><<<(unsigned)gp, TP_THREADS, smem, st>>>(
                values, ld, valid, ldv, n, ncols, fills, fill_const, out, (int)width, stride, km);
        } else {
            TMOG_CUDA(cudaFuncSetAttribute(vec_transform_pack_kernel<T, false>,
                                           cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            vec_transform_pack_kernel<T, false><<<(unsigned)gp, TP_THREADS, smem, st>>>(
                values, ld, valid, ldv, n, ncols, fills, fill_const, out, (int)width, stride, km);
        }
    } else {
        if (track_nulls)
            vec_transform_select_kernel<T, true><<<(unsigned)gx, 256, 0, st>>>(
                values, ld, valid, ldv, n, ncols, fills, fill_const, out, out_ld, out_col0, km);
        else
            vec_transform_select_kernel<T, false><<<(unsigned)gx, 256, 0, st>>>(
                values, ld, valid, ldv, n, ncols, fills, fill_const, out, out_ld, out_col0, km);
    }
    TMOG_LAUNCH_CHECK();
    return TMOG_OK;
}

extern "C" int tmog_real_transform_dev(const double* values, int64_t ld, const uint64_t* valid, int64_t ldv,
                                       int64_t n, int ncols, const double* fills, int track_nulls, double* out,
                                       int64_t out_ld, int64_t out_col0, void* stream) {
    TMOG_REQUIRE(fills != nullptr || valid == nullptr, "fills is NULL but the columns are nullable");
    return launch_transform<double>(values, ld, valid, ldv, n, ncols, fills, 0.0, track_nulls, out, out_ld, out_col0,
                                    stream);
}

extern "C" int tmog_binary_transform_dev(const uint8_t* values, int64_t ld, const uint64_t* valid, int64_t ldv,
                                         int64_t n, int ncols, int fill_value, int track_nulls, double* out,
                                         int64_t out_ld, int64_t out_col0, void* stream) {
    return launch_transform<uint8_t>(values, ld, valid, ldv, n, ncols, nullptr, fill_value ? 1.0 : 0.0, track_nulls,
                                     out, out_ld, out_col0, stream);
}

extern "C" int tmog_real_transform_select_dev(const double* values, int64_t ld, const uint64_t* valid, int64_t ldv,
                                              int64_t n, int ncols, const double* fills, int track_nulls,
                                              const uint64_t* keep_mask, int nkeep, double* out, int64_t out_ld,
                                              int64_t out_col0, void* stream) {
    TMOG_REQUIRE(fills != nullptr || valid == nullptr, "fills is NULL but the columns are nullable");
    TMOG_REQUIRE(keep_mask != nullptr && nkeep >= 0, "keep_mask is NULL");
    return launch_transform<double>(values, ld, valid, ldv, n, ncols, fills, 0.0, track_nulls, out, out_ld, out_col0,
                                    stream, keep_mask, nkeep);
}

extern "C" int tmog_binary_transform_select_dev(const uint8_t* values, int64_t ld, const uint64_t* valid, int64_t ldv,
                                                int64_t n, int ncols, int fill_value, int track_nulls,
                                                const uint64_t* keep_mask, int nkeep, double* out,
                                                int64_t out_ld, int64_t out_col0, void* stream) {
    TMOG_REQUIRE(keep_mask != nullptr && nkeep >= 0, "keep_mask is NULL");
    return launch_transform<uint8_t>(values, ld, valid, ldv, n, ncols, nullptr, fill_value ? 1.0 : 0.0, track_nulls,
                                     out, out_ld, out_col0, stream, keep_mask, nkeep);
}

// ---- Integral columns (int64): the same kernels, values converted to double as they are read ---------------------
extern "C" int tmog_integral_pilot_shift_dev(const int64_t* values, int64_t ld, const uint64_t* valid, int64_t ldv,
                                             const double* label, int64_t n, int ncols, double* shifts,
                                             void* stream) {
    if (int rc = check_columns(values, ld, valid, ldv, n, ncols, 8)) return rc;
    TMOG_REQUIRE(shifts != nullptr && (label != nullptr || n == 0), "shifts / label is NULL");
    real_pilot_shift_kernel<<<ncols + 1, 128, 0, (cudaStream_t)stream>>>(reinterpret_cast<const double*>(values), ld,
                                                                         valid, ldv, label, n, ncols, shifts, 1);
    TMOG_LAUNCH_CHECK();
    return TMOG_OK;
}

extern "C" int tmog_integral_fit_partials_dev(const int64_t* values, int64_t ld, const uint64_t* valid, int64_t ldv,
                                              const double* label, int64_t n, int ncols, const double* shifts,
                                              void* scratch, size_t scratch_bytes, double* partials, void* stream) {
    if (int rc = check_columns(values, ld, valid, ldv, n, ncols, 8)) return rc;
    TMOG_REQUIRE(label != nullptr || n == 0, "label is NULL");
    TMOG_REQUIRE(aligned16(label), "label must be 16-byte aligned");
    TMOG_REQUIRE(shifts != nullptr && partials != nullptr, "shifts / partials is NULL");
    const size_t need = sizeof(double) * (size_t)fit_grid() * (size_t)(ncols + 1) * NS;
    if (scratch == nullptr || scratch_bytes < need) {
        set_error("scratch too small: %zu bytes given, %zu needed (tmog_real_fit_scratch_bytes)", scratch_bytes, need);
        return TMOG_ENOMEM;
    }
    cudaStream_t st = (cudaStream_t)stream;
    bool fits = false;
    const int nblocks = sm_count();
    if (int rc = launch_fit_tma<512, 8, 16, 2, true>(reinterpret_cast<const double*>(values), ld, valid, ldv, label, n,
                                                     ncols, shifts, static_cast<double*>(scratch), nblocks, st, &fits))
        return rc;
    TMOG_REQUIRE(fits, "too many columns for one fit call (%d): split the column block", ncols);
    TMOG_LAUNCH_CHECK();
    real_fit_reduce_kernel<<<ncols + 1, 32, 0, st>>>(static_cast<const double*>(scratch), nblocks, ncols, n,
                                                     partials);
    TMOG_LAUNCH_CHECK();
    return TMOG_OK;
}

extern "C" int tmog_integral_transform_dev(const int64_t* values, int64_t ld, const uint64_t* valid, int64_t ldv,
                                           int64_t n, int ncols, const double* fills, int track_nulls, double* out,
                                           int64_t out_ld, int64_t out_col0, void* stream) {
    TMOG_REQUIRE(fills != nullptr || valid == nullptr, "fills is NULL but the columns are nullable");
    return launch_transform<long long>(reinterpret_cast<const long long*>(values), ld, valid, ldv, n, ncols, fills, 0.0,
                                       track_nulls, out, out_ld, out_col0, stream);
}

extern "C" int tmog_integral_transform_select_dev(const int64_t* values, int64_t ld, const uint64_t* valid,
                                                  int64_t ldv, int64_t n, int ncols, const double* fills,
                                                  int track_nulls, const uint64_t* keep_mask, int nkeep, double* out,
                                                  int64_t out_ld, int64_t out_col0, void* stream) {
    TMOG_REQUIRE(fills != nullptr || valid == nullptr, "fills is NULL but the columns are nullable");
    TMOG_REQUIRE(keep_mask != nullptr && nkeep >= 0, "keep_mask is NULL");
    return launch_transform<long long>(reinterpret_cast<const long long*>(values), ld, valid, ldv, n, ncols, fills, 0.0,
                                       track_nulls, out, out_ld, out_col0, stream, keep_mask, nkeep);
}

// ---------------------------------------------------------------------------------------------------------------------
// Allocation-free variant of tmog_real_fit_partials_dev.
//
// Some hosts of this library cannot give us device memory of their own.  The typical one is a JNI shim inside a Spark
// executor whose GPU is already managed by a memory pool (a columnar accelerator plugin, say) that has claimed nearly
// all of HBM up front: a cudaMalloc made on the side fails or is forbidden by policy there, and the shim's only
// device buffers are the partition's inputs and outputs, which the pool handed it.  It has nowhere to get the few
// hundred KiB that the fit kernel wants for its per-CTA partials.  For such callers the scratch lives in this file,
// in a file-scope __device__ array of fixed size that every GPU loading the library holds until the process exits.
//
// Notes for callers
//   * limit: TMOG_NOALLOC_SCRATCH_BYTES (4 MiB).  The kernel needs  CTAs x (ncols + 1) x 64 bytes, CTAs = 2 x SMs
//     (296 on a B200), i.e. at most 220 columns per call on a B200; wider blocks are rejected with TMOG_EINVAL and a
//     message that names the limit -- split the block, partials of column sub-blocks are independent.
//   * the first call on each device looks up the array's device address (cudaGetSymbolAddress, with that device
//     current) and keeps it (an atomic per device: host threads may race on the first call, they store the same
//     value).  Make that first call outside any stream capture.  cudaDeviceReset invalidates the kept address: a
//     process that resets a device must not call this entry point on it afterwards.
//   * every call on a device uses the same array.  Nothing in it has to start at zero (each CTA rewrites its whole
//     slot before the reduce kernel reads it) and a call leaves its partials behind, but two calls that overlap on
//     the GPU -- different streams, or different host threads on the same device -- race on it.  Serialise them.
//   * nothing here calls cudaMalloc / cudaFree or an allocating library routine.
// ---------------------------------------------------------------------------------------------------------------------
#define TMOG_NOALLOC_SCRATCH_BYTES (4u << 20)

static __device__ __align__(256) double g_noalloc_fit_scratch[TMOG_NOALLOC_SCRATCH_BYTES / sizeof(double)];

namespace {
constexpr int NOALLOC_MAX_DEVICES = 64;
std::atomic<double*> g_noalloc_ptr[NOALLOC_MAX_DEVICES];

// Device address of the scratch array on the CURRENT device.  In host code the symbol's name is only a stand-in;
// the kernel must be given the address cudaGetSymbolAddress returns.
int noalloc_scratch(double** out) {
    int dev = 0;
    TMOG_CUDA(cudaGetDevice(&dev));
    TMOG_REQUIRE(dev >= 0 && dev < NOALLOC_MAX_DEVICES, "device index %d not supported", dev);
    double* known = g_noalloc_ptr[dev].load(std::memory_order_relaxed);
    if (known == nullptr) {
        void* p = nullptr;
        TMOG_CUDA(cudaGetSymbolAddress(&p, g_noalloc_fit_scratch));
        TMOG_REQUIRE(p != nullptr, "cudaGetSymbolAddress returned a null address");
        known = static_cast<double*>(p);
        g_noalloc_ptr[dev].store(known, std::memory_order_relaxed);   // racing first calls store the same value
    }
    *out = known;
    return TMOG_OK;
}
}  // namespace

extern "C" int tmog_real_fit_partials_noalloc_dev(const double* values, int64_t ld, const uint64_t* valid,
                                                  int64_t ldv, const double* label, int64_t n, int ncols,
                                                  const double* shifts, double* partials, void* stream) {
    if (int rc = check_columns(values, ld, valid, ldv, n, ncols, 8)) return rc;
    TMOG_REQUIRE(label != nullptr || n == 0, "label is NULL");
    TMOG_REQUIRE(aligned16(label), "label must be 16-byte aligned");
    TMOG_REQUIRE(shifts != nullptr && partials != nullptr, "shifts / partials is NULL");
    const int grid = fit_grid();
    const size_t need = sizeof(double) * (size_t)grid * (size_t)(ncols + 1) * NS;
    const int max_cols = (int)(TMOG_NOALLOC_SCRATCH_BYTES / (sizeof(double) * (size_t)grid * NS)) - 1;
    TMOG_REQUIRE(need <= TMOG_NOALLOC_SCRATCH_BYTES,
                 "tmog_real_fit_partials_noalloc_dev: %d columns need %zu bytes of scratch but the built-in scratch "
                 "is %u bytes (at most %d columns per call on this device): split the column block or use "
                 "tmog_real_fit_partials_dev with caller-provided scratch",
                 ncols, need, (unsigned)TMOG_NOALLOC_SCRATCH_BYTES, max_cols);
    const size_t smem = fit_smem_bytes(ncols);
    TMOG_REQUIRE(smem <= 48 * 1024, "too many columns for one allocation-free fit call (%d)", ncols);
    double* scratch = nullptr;
    if (int rc = noalloc_scratch(&scratch)) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    real_fit_kernel<<<grid, FIT_THREADS, smem, st>>>(values, ld, valid, ldv, label, n, ncols, shifts, scratch);
    TMOG_LAUNCH_CHECK();
    real_fit_reduce_kernel<<<ncols + 1, 32, 0, st>>>(scratch, grid, ncols, n, partials);
    TMOG_LAUNCH_CHECK();
    return TMOG_OK;
}
