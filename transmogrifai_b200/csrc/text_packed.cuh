// The transform of rows whose cells the fit pass left in packed form (TsPack, text_cells.cuh): sorted sparse rows of the
// kept columns without the text and without tokenising or hashing anything -- the second pass over a chunk-streamed
// dataset (SmartTextVectorizer.scala:333-356 applied to the rows it was fitted on) reads ~0.9 KB per row instead of ~7 KB
// of text.  Same output as text_cells_kernel / text_span_kernel + text_rows_kernel, entry for entry.
//
//   text_rows_packed_kernel   a warp per row, a lane per cell: the lane walks its cell's entries (ascending buckets),
//                             maps every bucket to its column among the kept ones and stages (column, term frequency)
//                             in shared memory at the cell's place in the row; the warp then takes the row's place in
//                             the output from the global cursor and copies the staged entries out with consecutive lanes
//                             writing consecutive entries, squeezing out the dropped columns on the way.  Text lengths
//                             and null indicators follow as in text_rows_kernel.  Rows with more entries than the stage
//                             holds take a two-pass path without it.
#pragma once

namespace tmog {

constexpr int TP_WARPS = 8;
constexpr int TP_STAGE = 1152;            // entries of a row the stage holds (4 CTAs of 8 warps fit an SM)
constexpr int TP_WARP_BYTES = TP_STAGE * 4 + TP_STAGE;

struct TpCell {
    unsigned long long off;               // arena offset of the cell's bytes (of its slot when slot is set)
    int cnt;                              // stage places the cell takes: its bytes (an upper bound of its entries), or
                                          // for a slot the buckets that hold something
    bool slot, nohit;
    unsigned len_sum;
};

// the cell (row, j): where its bytes are and how many
__device__ __forceinline__ TpCell tp_cell(const TsPack& pk, int ncols, int64_t n, int64_t row, int j, bool want_len) {
    TpCell c;
    const unsigned r = pk.rel[row * ncols + j];
    const TsGroup g = pk.grp[(row >> 5) * ncols + j];
    const unsigned lo = r & TS_REL_MASK;
    const unsigned hi = ((row & 31) < 31 && row + 1 < n) ? (pk.rel[(row + 1) * ncols + j] & TS_REL_MASK) : g.total;
    c.off = g.off + lo;
    c.cnt = (int)(hi - lo);
    c.slot = (r & TS_REL_SLOT) != 0u;
    c.nohit = (r & TS_REL_NOHIT) != 0u;
    c.len_sum = (want_len && pk.lens) ? pk.lens[row * ncols + j] : 0u;
    if (c.slot) {
        const unsigned char* at = pk.arena + c.off;
        unsigned long long s = 0ull;
        for (int q = 0; q < 8; ++q) s |= (unsigned long long)at[q] << (8 * q);
        c.off = s;
        const unsigned short* sl = reinterpret_cast<const unsigned short*>(pk.arena + s);
        int m = 0;
        for (int q = 0; q < TC_MAXF; ++q) m += sl[q] != 0;
        c.cnt = m;
        c.len_sum = want_len ? ((unsigned)sl[TC_MAXF] | ((unsigned)sl[TC_MAXF + 1] << 16)) : 0u;
        c.nohit = sl[TC_MAXF + 2] != 0;
    }
    return c;
}

// the entries of a cell in ascending bucket order: f(k, bucket, term frequency) -> number of entries.  The bytes are
// read sixteen at a time (loads in flight), then decoded in order (TsPack's format); for a slot the buckets that hold
// something are found by walking (slots are rare)
template <typename F>
__device__ __forceinline__ int tp_for_entries(const TsPack& pk, const TpCell& c, F&& f) {
    int k = 0;
    if (!c.slot) {
        const unsigned char* __restrict__ e = pk.arena + c.off;
        const int nbytes = c.cnt;
        int prev = -1, tused = 0;
        for (int k0 = 0; k0 < nbytes; k0 += 16) {
            unsigned v[16];
#pragma unroll
            for (int i = 0; i < 16; ++i) v[i] = k0 + i < nbytes ? (unsigned)e[k0 + i] : 0xFCu;
#pragma unroll
            for (int i = 0; i < 16; ++i) {
                const unsigned b = v[i], g = b >> 2;
                // a byte past the stream (the trailer shrinks it from behind) is no entry; gap 63 skips 63 buckets
                const bool entry = g != 63u && k0 + i < nbytes - tused;
                prev += g == 63u ? 63 : (int)g + 1;
                if (entry) {
                    int tf = (int)(b & 3u) + 1;
                    if (tf == 4) tf = (int)e[nbytes - 1 - tused++] + 4;
                    f(k++, prev, tf);
                }
            }
        }
    } else {
        const unsigned short* sl = reinterpret_cast<const unsigned short*>(pk.arena + c.off);
        for (int q = 0; q < TC_MAXF; ++q)
            if (sl[q]) f(k++, q, (int)sl[q]);
    }
    return k;
}

__global__ void __launch_bounds__(TP_WARPS * 32)
text_rows_packed_kernel(int ncols, int64_t n, TcParams p, TsPack pk, const int32_t* __restrict__ newpos,
                        int64_t* __restrict__ row_start, int32_t* __restrict__ row_nnz, int32_t* __restrict__ out_idx,
                        int32_t* __restrict__ out_val, unsigned long long* __restrict__ cursor, int64_t capacity,
                        int* __restrict__ flags) {
    extern __shared__ __align__(16) unsigned char tp_smem[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    int* sidx = reinterpret_cast<int*>(tp_smem + (size_t)warp * TP_WARP_BYTES);
    unsigned char* sval = reinterpret_cast<unsigned char*>(sidx + TP_STAGE);
    const int F = p.num_features;
    const unsigned lt = (1u << lane) - 1u;
    const bool want_len = p.len_base >= 0;

    for (int64_t row = (int64_t)blockIdx.x * TP_WARPS + warp; row < n; row += (int64_t)gridDim.x * TP_WARPS) {
        // the cells of the first 32 columns stay in registers; beyond that they are looked up again when needed
        TpCell first;
        first.cnt = 0;
        first.slot = first.nohit = false;
        first.len_sum = 0u;
        first.off = 0ull;
        if (lane < ncols) first = tp_cell(pk, ncols, n, row, lane, want_len);
        auto cell_at = [&](int j0) -> TpCell {
            if (j0 == 0) return first;
            TpCell c;
            c.cnt = 0;
            c.slot = c.nohit = false;
            c.len_sum = 0u;
            c.off = 0ull;
            if (j0 + lane < ncols) c = tp_cell(pk, ncols, n, row, j0 + lane, want_len);
            return c;
        };
        // ---- a lane per cell: the cell's entries mapped to the kept columns and staged at the cell's place in the row
        int kept = 0, n_len = 0, n_null = 0, raw_total = 0;
        bool any_slot = false;
        for (int j0 = 0; j0 < ncols; j0 += 32) {
            const int j = j0 + lane;
            const TpCell c = cell_at(j0);
            int incl = c.cnt;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int up = __shfl_up_sync(FULL, incl, o);
                if (lane >= o) incl += up;
            }
            const int so = raw_total + incl - c.cnt;
            raw_total += __shfl_sync(FULL, incl, 31);
            if (j < ncols && j < p.n_hashed) {
                const int32_t* np = newpos + (int64_t)j * F;
                // a column none of whose buckets was dropped maps bucket b to (its first output column) + b
                const int lo = np[0], hi = np[F - 1];
                const bool whole = lo >= 0 && hi - lo == F - 1;
                // (a cell that does not fit the stage means a row that does not: the stage is not used then)
                const bool cell_fits = so + c.cnt <= TP_STAGE;
                int* si = sidx + so;
                unsigned char* sv = sval + so;
                const int used = tp_for_entries(pk, c, [&](int k, int bucket, int tf) {
                    const int v = whole ? lo + bucket : np[bucket];
                    kept += v >= 0;
                    if (cell_fits) {
                        si[k] = v;
                        sv[k] = (unsigned char)(p.binary_freq ? 1 : min(tf, 255));
                    }
                });
                if (cell_fits)
                    for (int k = used; k < c.cnt; ++k) si[k] = -1;   // fewer entries than bytes: the rest is empty
            }
            const bool has_len = j < ncols && want_len && c.len_sum > 0u && newpos[p.len_base + j] >= 0;
            const bool has_null = j < ncols && p.track_nulls && c.nohit && newpos[p.null_base + j] >= 0;
            n_len += __popc(__ballot_sync(FULL, has_len));
            n_null += __popc(__ballot_sync(FULL, has_null));
            any_slot |= __any_sync(FULL, c.slot);
        }
        const bool fits = raw_total <= TP_STAGE;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) kept += __shfl_xor_sync(FULL, kept, o);
        const int total = kept + n_len + n_null;
        unsigned long long base = 0;
        if (lane == 0) {
            base = atomicAdd(cursor, (unsigned long long)total);
            row_start[row] = (int64_t)base;
            row_nnz[row] = total;
            if ((int64_t)base + total > capacity) flags[3] = 1;
        }
        base = __shfl_sync(FULL, base, 0);
        if ((int64_t)base + total > capacity) continue;          // sizing call (capacity 0) or too small: nothing written
        __syncwarp();
        int run = 0;
        if (fits) {
            // ---- the staged entries go out, consecutive lanes to consecutive places, dropped columns squeezed out
            for (int i = 0; i < raw_total; i += 32) {
                const int k = i + lane;
                const int v = k < raw_total ? sidx[k] : -1;
                const unsigned bal = __ballot_sync(FULL, v >= 0);
                if (v >= 0) {
                    const int64_t at = (int64_t)base + run + __popc(bal & lt);
                    out_idx[at] = v;
                    out_val[at] = (int32_t)sval[k];
                }
                run += __popc(bal);
            }
        } else {
            // ---- a row with more entries than the stage holds: every lane writes its cells' kept entries directly
            for (int j0 = 0; j0 < ncols; j0 += 32) {
                const int j = j0 + lane;
                int mine = 0;
                const bool on = j < ncols && j < p.n_hashed;
                const TpCell c = cell_at(j0);
                const int32_t* np = newpos + (int64_t)(on ? j : 0) * F;
                if (on) tp_for_entries(pk, c, [&](int, int bucket, int) { mine += np[bucket] >= 0; });
                int incl = mine;
#pragma unroll
                for (int o = 1; o < 32; o <<= 1) {
                    const int up = __shfl_up_sync(FULL, incl, o);
                    if (lane >= o) incl += up;
                }
                int64_t at = (int64_t)base + run + incl - mine;
                run += __shfl_sync(FULL, incl, 31);
                if (on)
                    tp_for_entries(pk, c, [&](int, int bucket, int tf) {
                        const int v = np[bucket];
                        if (v >= 0) {
                            out_idx[at] = v;
                            out_val[at] = p.binary_freq ? 1 : tf;
                            ++at;
                        }
                    });
            }
        }
        // ---- term frequencies above 255 do not fit the stage's bytes; only a slot can hold one: patched from the slot
        if (fits && any_slot && !p.binary_freq) {
            int before = 0;                                       // kept entries of the cells before the one at hand
            for (int j = 0; j < p.n_hashed; ++j) {
                const TpCell c = tp_cell(pk, ncols, n, row, j, false);
                const int32_t* np = newpos + (int64_t)j * F;
                int mine = 0;
                tp_for_entries(pk, c, [&](int, int bucket, int tf) {
                    if (np[bucket] >= 0) {
                        if (c.slot && tf > 255 && lane == 0) out_val[(int64_t)base + before + mine] = tf;
                        ++mine;
                    }
                });
                before += mine;
            }
        }
        __syncwarp();
        // ---- text lengths, then null indicators
        int at = kept;
        for (int j0 = 0; j0 < ncols; j0 += 32) {
            const int j = j0 + lane;
            const TpCell c = cell_at(j0);
            const bool has = j < ncols && want_len && c.len_sum > 0u && newpos[p.len_base + j] >= 0;
            const unsigned bl = __ballot_sync(FULL, has);
            if (has) {
                out_idx[(int64_t)base + at + __popc(bl & lt)] = newpos[p.len_base + j];
                out_val[(int64_t)base + at + __popc(bl & lt)] = (int32_t)c.len_sum;
            }
            at += __popc(bl);
        }
        for (int j0 = 0; j0 < ncols; j0 += 32) {
            const int j = j0 + lane;
            const TpCell c = cell_at(j0);
            const bool has = j < ncols && p.track_nulls && c.nohit && newpos[p.null_base + j] >= 0;
            const unsigned bl = __ballot_sync(FULL, has);
            if (has) {
                out_idx[(int64_t)base + at + __popc(bl & lt)] = newpos[p.null_base + j];
                out_val[(int64_t)base + at + __popc(bl & lt)] = 1;
            }
            at += __popc(bl);
        }
        __syncwarp();
    }
}

}  // namespace tmog
