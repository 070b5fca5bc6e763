// Tokenise + hash for SEPARATE hash spaces of at most 512 buckets per column (TransmogrifierDefaults: 512 hashes per
// text column, OPCollectionHashingVectorizer.scala:191-199) -- included by text.cu, same semantics as text_hash_kernel.
//
// With separate spaces a cell (one row of one text column) owns its own slice of the vector, so nothing has to be
// merged across the cells of a row.  The work is cut the other way round from text_hash_kernel:
//
//   text_cells_kernel   a LANE owns a cell; a warp takes 32 consecutive rows of ONE column (their bytes are contiguous
//                       in the Arrow buffer).  The lane reads its cell 16 bytes at a time, lower-cases and classifies
//                       the bytes in registers (SWAR; UAX#29 Mid* rules from the neighbouring bytes), keeps the last 64
//                       bytes in a private ring in shared memory, and at every token end reads the token back from the
//                       ring for the stop-word probe and murmur3.  Term frequencies of the cell: a 512-bit bitmap
//                       (bucket present) plus a short list of the buckets met more than once -- both private to the
//                       lane, so there are no atomics and no per-row table, and the buckets come out sorted for free.
//                       Fit statistics: the 32 bitmaps of a warp are transposed with shuffles and every lane keeps the
//                       per-class counts of "its" 16 buckets in registers until the CTA moves to another column.
//                       Transform: every cell leaves a 128-byte record (bitmap, list, length sum, flags).
//   text_cells_slow_kernel   the statistics of the few cells the fast path sets aside (a token longer than the ring
//                       holds, more repeated buckets than the list holds): a thread per cell, byte by byte.
//   text_rows_kernel    HBM-bound: a warp per row turns the row's records into the sorted sparse row (kept columns
//                       only), popcounts + prefix sums instead of a sort.
#pragma once

namespace tmog {

#ifndef TC_WARPS_V
#define TC_WARPS_V 8
#endif
#ifndef TC_ROUND_LANES_V
#define TC_ROUND_LANES_V 26
#endif
#ifndef TC_PEND_V
#define TC_PEND_V 12
#endif
constexpr int TC_WARPS = TC_WARPS_V;
constexpr int TC_FW = 16;                 // bitmap words of a cell: numFeatures <= 512
constexpr int TC_MAXF = TC_FW * 32;
constexpr int TC_DUPS = 12;               // buckets with term frequency > 1 a cell record holds (13 slots)
constexpr int TC_RING = 16;               // words of recent text a lane keeps (64 bytes)
constexpr int TC_RING_ROWS = TC_RING + 5; // + the first five words again, behind the last (no wrap inside a token)
constexpr int TC_BACK = 44;               // a token may start this many bytes before the chunk it ends in
constexpr int TC_ITEM_ROWS = 8192;        // rows of one column a CTA takes at a time
constexpr int TC_ROUND_LANES = TC_ROUND_LANES_V;        // a round runs when this many lanes have a token waiting
constexpr int TC_WARP_ROWS = TC_RING_ROWS + 4 + TC_FW + TC_DUPS;   // shared-memory rows (32 words) of a warp
constexpr int TC_STAT_ROWS = TC_FW;       // + per-class bucket counts of the warp's column (u16 x 2 classes), fit only

// record of one cell, 32 words: bitmap [16] | header | length sum | extra | dup [13] (bucket | count << 16)
constexpr int TC_REC_WORDS = 32;
constexpr unsigned TC_H_NOHIT = 0x100u;   // no token survived (null, empty, all stop words): the null indicator is 1
constexpr unsigned TC_H_DEFER = 0x200u;   // set aside by the fast path: whoever reads the record tokenises the cell

struct TcParams {
    int num_features;
    int binary_freq;
    int min_token_length;
    int track_nulls;
    int prepend;
    int null_base;         // vector index of the first null indicator
    int len_base;          // vector index of the first text-length column, or -1
    int n_hashed;          // columns [0, n_hashed) are hashed, the rest only give null indicators / lengths
    int nclasses;
    int pow2_mask;
    int force_defer;       // testing: every cell takes the slow path
};

struct TcScratch {
    unsigned long long item_counter;
    unsigned long long row_counter;
    unsigned long long pad[6];
};

// ---- packed cells: what the fit pass found in every cell (its buckets and their term frequencies), kept in HBM so that
//      the transform of the same rows needs neither the text nor a second tokenise + hash pass (text_packed.cuh reads
//      them).  A group = 32 consecutive rows of one column = one span of text_span_kernel (text_span.cuh):
//        arena   bytes.  A cell is a stream of bytes in ascending bucket order: (gap << 2 | tf code) with gap = buckets
//                skipped since the entry before (0 .. 62) and tf code 0 .. 2 = term frequency 1 .. 3, code 3 = the term
//                frequency (- 4) is the next byte of the cell's trailer, which is read backwards from the cell's last
//                byte; gap 63 (byte >= 0xFC) is no entry: it skips 63 buckets.  ~1.1 bytes per distinct token.  The
//                cells of a group follow one another; a group starts at a multiple of 16.  A cell the fast path set
//                aside holds eight bytes instead: the arena offset of its slot (TS_SLOT_ENTRIES u16: the term frequency
//                of every bucket, the summed token length, the no-token flag), written by the slow kernel.
//        grp     [ceil(n / 32)][ncols]: arena offset of the group's bytes (~0: the arena was full), their number
//        rel     [n][ncols] u16: where the cell's bytes begin inside the group | no token survived << 14 | slot << 15
//        lens    [n][ncols] u32 summed token lengths (only when the vector has text-length columns)
constexpr int TS_SLOT_ENTRIES = TC_MAXF + 8;          // 512 term frequencies | length sum (2) | no-token flag | pad
constexpr unsigned TS_REL_MASK = 0x3fffu, TS_REL_NOHIT = 0x4000u, TS_REL_SLOT = 0x8000u;
struct TsGroup {
    unsigned long long off;
    unsigned total, pad;
};
struct TsPack {
    unsigned char* arena;
    unsigned long long arena_bytes;
    unsigned long long* cursor;            // bytes handed out so far
    TsGroup* grp;
    unsigned short* rel;
    unsigned* lens;                        // or NULL
    int* flags;                            // [0] = 1: the arena was too small for this chunk
};

// ---- exact scalar tokeniser (the definition the fast path must agree with; oracle/text.py standard_tokenize) --------
template <typename F>
__device__ void tc_scalar_tokens(const uint8_t* s, int64_t len, int min_len, bool& non_ascii, F&& emit) {
    int64_t i = 0;
    auto low = [&](int64_t k) -> unsigned { return to_lower(s[k]); };
    while (i < len) {
        unsigned c = low(i);
        if (c >= 0x80u) non_ascii = true;
        if (!(is_letter(c) || is_digit(c) || c == '_')) {
            ++i;
            continue;
        }
        int64_t j = i;
        bool alnum = false;
        while (j < len) {
            c = low(j);
            if (c >= 0x80u) non_ascii = true;
            if (is_letter(c) || is_digit(c)) {
                alnum = true;
                ++j;
            } else if (c == '_') {
                ++j;
            } else if (j > i && j + 1 < len) {
                const unsigned p = low(j - 1), q = low(j + 1);
                const bool mid_l = c == ':' || c == '.' || c == '\'';
                const bool mid_n = c == ',' || c == ';' || c == '.' || c == '\'';
                if ((mid_l && is_letter(p) && is_letter(q)) || (mid_n && is_digit(p) && is_digit(q))) ++j;
                else break;
            } else {
                break;
            }
        }
        if (alnum)
            for (int64_t k = i; k < j; k += TH_MAX_TOKEN) {
                const int plen = (int)min((int64_t)TH_MAX_TOKEN, j - k);
                if (plen >= min_len) emit(k, plen);
            }
        i = j;
    }
}

// stop-word test of the lower-cased bytes s[k .. k + plen) against the constant-memory table
__device__ inline bool tc_scalar_is_stop(const uint8_t* s, int64_t k, int plen) {
    if (plen > STOP_MAXLEN) return false;
    unsigned w[3] = {0u, 0u, 0u};
    for (int q = 0; q < plen; ++q) w[q >> 2] |= to_lower(s[k + q]) << (8 * (q & 3));
    const unsigned h1 = stop_slot_hash(w[0], w[1], w[2]), h2 = stop_slot_hash2(w[0], w[1], w[2]);
    for (int t = 0; t < 2; ++t) {
        const StopEntry& e = c_stop[t == 0 ? h1 : h2];
        if (e.len != plen) continue;
        bool same = true;
        for (int q = 0; q < plen; ++q) same &= (unsigned char)e.w[q] == to_lower(s[k + q]);
        if (same) return true;
    }
    return false;
}

// term frequencies of one cell into counts [num_features] (u16, zeroed by the caller): -> no token survived;
// len_sum = summed length of the surviving tokens
__device__ inline bool tc_scalar_cell(const uint8_t* s, int64_t len, const TcParams& p, const uint8_t* prefix, int plen0,
                                      bool hashed, unsigned short* counts, unsigned& len_sum, bool& non_ascii,
                                      bool& count_ovf) {
    bool nohit = true;
    unsigned ls = 0u;
    tc_scalar_tokens(s, len, p.min_token_length, non_ascii, [&](int64_t k, int plen) {
        if (tc_scalar_is_stop(s, k, plen)) return;
        nohit = false;
        ls += (unsigned)plen;
        if (!hashed) return;
        Murmur m;
        m.init();
        for (int q = 0; q < plen0; ++q) m.push(prefix[q]);
        for (int q = 0; q < plen; ++q) m.push(to_lower(s[k + q]));
        const unsigned hv = m.finish();
        const int b = p.pow2_mask >= 0 ? (int)(hv & (unsigned)p.pow2_mask) : non_negative_mod(hv, p.num_features);
        if (counts[b] == 0xffffu) count_ovf = true;
        else ++counts[b];
    });
    len_sum = ls;
    return nohit;
}

// ---- 32 x 32 bit transpose across a warp: lane k ends up with bit i = (lane i's word >> k) & 1 -------------------
__device__ __forceinline__ unsigned tc_transpose32(unsigned x, int lane) {
#pragma unroll
    for (int j = 16; j >= 1; j >>= 1) {
        const unsigned m = j == 16 ? 0x0000ffffu : j == 8 ? 0x00ff00ffu : j == 4 ? 0x0f0f0f0fu
                           : j == 2 ? 0x33333333u : 0x55555555u;
        const unsigned y = __shfl_xor_sync(FULL, x, j);
        x = (lane & j) ? ((x & ~m) | ((y >> j) & m)) : ((x & m) | ((y << j) & ~m));
    }
    return x;
}

__device__ __forceinline__ unsigned tc_ring_byte(const unsigned* ring, int lane, int pos) {
    return (ring[((pos >> 2) & (TC_RING - 1)) * 32 + lane] >> ((pos & 3) * 8)) & 0xffu;
}

// stop-word probe of the token [ps, ps + plen) (plen <= STOP_MAXLEN) read from the lane's ring
__device__ __forceinline__ bool tc_is_stop(const unsigned* ring, int lane, int ps, int plen) {
    const int w0 = ps >> 2, sh = (ps & 3) * 8;
    const unsigned a = ring[((w0) & (TC_RING - 1)) * 32 + lane], b = ring[((w0 + 1) & (TC_RING - 1)) * 32 + lane];
    const unsigned c = ring[((w0 + 2) & (TC_RING - 1)) * 32 + lane], d = ring[((w0 + 3) & (TC_RING - 1)) * 32 + lane];
    const unsigned l0 = __funnelshift_r(a, b, sh) & byte_range_mask(0, plen);
    const unsigned l1 = __funnelshift_r(b, c, sh) & byte_range_mask(0, plen - 4);
    const unsigned l2 = __funnelshift_r(c, d, sh) & byte_range_mask(0, plen - 8);
    // few tokens get here (the pre-filter knows first letter, second character and length of every stop word), so the
    // table stays in constant memory
    const StopSlot e1 = c_stop_packed[stop_slot_hash(l0, l1, l2)];
    const StopSlot e2 = c_stop_packed[stop_slot_hash2(l0, l1, l2)];
    const unsigned d1 = (e1.len ^ (unsigned)plen) | (e1.w0 ^ l0) | (e1.w1 ^ l1) | (e1.w2 ^ l2);
    const unsigned d2 = (e2.len ^ (unsigned)plen) | (e2.w0 ^ l0) | (e2.w1 ^ l1) | (e2.w2 ^ l2);
    return d1 == 0u || d2 == 0u;
}

// murmur3 of "<prefix><token>" continued from the state after the prefix; the token is read from the ring (rows
// 16 .. 20 repeat rows 0 .. 4).  Tokens of up to 15 bytes with the prefix remainder take a fixed number of steps, the
// same for every lane; longer ones loop.
__device__ __forceinline__ unsigned tc_hash_token(const unsigned* ring, int lane, int ps, int plen, unsigned ph,
                                                  unsigned pw, unsigned pn) {
    const int r = (int)(pn & 3u);
    const unsigned keep_hi = ~((1u << (8 * r)) - 1u);
    const int tot = r + plen, nfull = tot >> 2, tail = tot & 3;
    const int q0 = ps - r, sh = (q0 & 3) * 8;
    const int w0 = q0 >> 2;                       // arithmetic shift: q0 may be slightly negative
    unsigned h = ph, wd;
    if (tot <= 15) {
        const unsigned* rp = ring + (w0 & (TC_RING - 1)) * 32 + lane;
        const unsigned a0 = rp[0], a1 = rp[32], a2 = rp[64], a3 = rp[96], a4 = rp[128];
        const unsigned s0 = (__funnelshift_r(a0, a1, sh) & keep_hi) | pw, s1 = __funnelshift_r(a1, a2, sh);
        const unsigned s2 = __funnelshift_r(a2, a3, sh), s3 = __funnelshift_r(a3, a4, sh);
        unsigned t = mix_h1(h, mix_k1(s0));
        h = nfull > 0 ? t : h;
        t = mix_h1(h, mix_k1(s1));
        h = nfull > 1 ? t : h;
        t = mix_h1(h, mix_k1(s2));
        h = nfull > 2 ? t : h;
        wd = nfull == 0 ? s0 : nfull == 1 ? s1 : nfull == 2 ? s2 : s3;
    } else {
        unsigned nxt = ring[(w0 & (TC_RING - 1)) * 32 + lane];
        for (int q = 0; q < nfull; ++q) {
            const unsigned cur = nxt;
            nxt = ring[((w0 + q + 1) & (TC_RING - 1)) * 32 + lane];
            unsigned x = __funnelshift_r(cur, nxt, sh);
            if (q == 0) x = (x & keep_hi) | pw;
            h = mix_h1(h, mix_k1(x));
        }
        wd = __funnelshift_r(nxt, ring[((w0 + nfull + 1) & (TC_RING - 1)) * 32 + lane], sh);
    }
    // the (< 4) bytes left are mixed one by one, sign-extended (Spark's legacy hashUnsafeBytes)
    unsigned t = mix_h1(h, mix_k1((unsigned)(int)(signed char)(wd & 0xffu)));
    h = tail > 0 ? t : h;
    t = mix_h1(h, mix_k1((unsigned)(int)(signed char)((wd >> 8) & 0xffu)));
    h = tail > 1 ? t : h;
    t = mix_h1(h, mix_k1((unsigned)(int)(signed char)((wd >> 16) & 0xffu)));
    h = tail > 2 ? t : h;
    return fmix32(h, pn + (unsigned)plen);
}

// bit 7 of every byte that is a token character, for the four words of a chunk (lower-cased, bytes outside the cell
// zero); cl / cd = letter / digit bits of the chunk before (carry in, carry out), nxt0 = first word of the next chunk
__device__ __forceinline__ unsigned tc_classify(const uint4& v, unsigned& cl, unsigned& cd, unsigned nxt0,
                                                unsigned& non_ascii) {
    const unsigned W[4] = {v.x, v.y, v.z, v.w};
    unsigned fl[4], fd[4], tc[4], rng_any = 0u;
#pragma unroll
    for (int q = 0; q < 4; ++q) {
        const unsigned h7 = W[q] & 0x7f7f7f7fu, ascii = ~W[q] & 0x80808080u;
        non_ascii |= W[q] & 0x80808080u;
        // bit 7 of (h7 + (0x80 - x) * 0x01010101) is set iff the byte is >= x
        fl[q] = (h7 + 0x1f1f1f1fu) & ~(h7 + 0x05050505u) & ascii;                 // 'a' .. 'z'
        fd[q] = (h7 + 0x50505050u) & ~(h7 + 0x46464646u) & ascii;                 // '0' .. '9'
        const unsigned ne_us = (h7 ^ 0x5f5f5f5fu) + 0x7f7f7f7fu;                   // bit 7 set iff the byte is not '_'
        tc[q] = (fl[q] | fd[q] | (~ne_us & ascii)) & 0x80808080u;
        // the Mid* characters . ' : , ; all lie in 0x27 .. 0x3b
        rng_any |= (h7 + 0x59595959u) & ~(h7 + 0x44444444u) & ~fd[q] & ascii & 0x80808080u;
    }
    if (rng_any) {
        const unsigned n7 = nxt0 & 0x7fu, nasc = (~nxt0 & 0x80u);
        const unsigned nl = ((n7 + 0x1fu) & ~(n7 + 0x05u)) & nasc;                // next chunk's first byte: letter
        const unsigned nd = ((n7 + 0x50u) & ~(n7 + 0x46u)) & nasc;                // ... digit  (bit 7)
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const unsigned h7 = W[q] & 0x7f7f7f7fu, ascii = ~W[q] & 0x80808080u;
            const unsigned ne_dot = (h7 ^ 0x2e2e2e2eu) + 0x7f7f7f7fu;
            const unsigned ne_apo = (h7 ^ 0x27272727u) + 0x7f7f7f7fu;
            const unsigned ne_col = (h7 ^ 0x3a3a3a3au) + 0x7f7f7f7fu;
            const unsigned ne_com = (h7 ^ 0x2c2c2c2cu) + 0x7f7f7f7fu;
            const unsigned ne_sem = (h7 ^ 0x3b3b3b3bu) + 0x7f7f7f7fu;
            const unsigned mid_l = ~(ne_dot & ne_apo & ne_col);          // MidLetter | MidNumLet | '
            const unsigned mid_n = ~(ne_dot & ne_apo & ne_com & ne_sem); // MidNum | MidNumLet | '
            const unsigned pl = q == 0 ? cl : fl[q - 1], pd = q == 0 ? cd : fd[q - 1];
            const unsigned xl = q == 3 ? nl : fl[q + 1], xd = q == 3 ? nd : fd[q + 1];
            // class of the byte before / after every byte, in that byte's bit 7
            const unsigned left_l = __funnelshift_l(pl, fl[q], 8), left_d = __funnelshift_l(pd, fd[q], 8);
            const unsigned right_l = __funnelshift_r(fl[q], xl, 8), right_d = __funnelshift_r(fd[q], xd, 8);
            tc[q] |= ((mid_l & left_l & right_l) | (mid_n & left_d & right_d)) & ascii & 0x80808080u;
        }
    }
    cl = fl[3];
    cd = fd[3];
    // bits 7, 15, 23, 31 of each word -> a nibble; four nibbles -> the 16 token bits of the chunk
    unsigned m = 0u;
#pragma unroll
    for (int q = 0; q < 4; ++q) m |= ((((tc[q] >> 7) * 0x00204081u) >> 21) & 0xfu) << (4 * q);
    return m;
}

// STATS: per-bucket statistics of the cells (fit);  CELLS: a record per cell (transform)
template <bool STATS, bool CELLS>
__global__ void __launch_bounds__(TC_WARPS * 32, STATS ? 3 : 4)
text_cells_kernel(const TextColumn* __restrict__ cols, int ncols, int64_t n, const uint8_t* __restrict__ prefix,
                  const int* __restrict__ prefix_len, TcParams p, const uint8_t* __restrict__ class_ids,
                  TcScratch* __restrict__ scratch, unsigned* __restrict__ defer_bits,
                  unsigned int* __restrict__ bucket_class_sum, unsigned int* __restrict__ bucket_dup,
                  unsigned long long* __restrict__ bucket_sq_extra, unsigned int* __restrict__ bucket_aux,
                  const double* __restrict__ label, double label_shift, double* __restrict__ bucket_wsum,
                  unsigned* __restrict__ records, int* __restrict__ flags) {
    extern __shared__ __align__(16) unsigned char tc_smem[];
    unsigned short* stop_pre = reinterpret_cast<unsigned short*>(tc_smem);
    unsigned* item_info = reinterpret_cast<unsigned*>(tc_smem + TH_STOP_PRE_BYTES);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    constexpr int WARP_WORDS = (TC_WARP_ROWS + (STATS ? TC_STAT_ROWS : 0)) * 32;
    unsigned* wbase = reinterpret_cast<unsigned*>(tc_smem + TH_STOP_PRE_BYTES + 64) + (size_t)warp * WARP_WORDS;
    unsigned* ring = wbase;                         // [TC_RING_ROWS][32]
    unsigned* masks = ring + TC_RING_ROWS * 32;     // [4][32] token end | start << 16 masks of the last chunks
    unsigned* bm = masks + 4 * 32;                  // [TC_FW][32]
    unsigned* dups = bm + TC_FW * 32;               // [TC_DUPS][32]: bucket | count << 16
    // fit: rows of each label class holding bucket 32 w + lane of the warp's column, [class][TC_FW][32] u16
    unsigned short* hist = reinterpret_cast<unsigned short*>(dups + TC_DUPS * 32);

    for (int i = tid; i < STOP_PRE; i += TC_WARPS * 32) stop_pre[i] = c_stop_pre[i];
    __syncthreads();

    const int64_t items_per_col = (n + TC_ITEM_ROWS - 1) / TC_ITEM_ROWS;
    const int64_t nitems = items_per_col * ncols;
    const int F = p.num_features;
    const bool two_classes = p.nclasses <= 2;
    unsigned non_ascii = 0u;

    for (;;) {
        __syncthreads();
        if (tid == 0) {
            const unsigned long long it = atomicAdd(&scratch->item_counter, 1ull);
            item_info[0] = (unsigned)(it & 0xffffffffu);
            item_info[1] = (unsigned)(it >> 32);
            if ((int64_t)it < nitems) {
                const int j = (int)((int64_t)it / items_per_col);
                Murmur pm;
                pm.init();
                const int plen = (p.prepend && j < p.n_hashed) ? prefix_len[j] : 0;   // ignored columns have no prefix
                for (int q = 0; q < plen; ++q) pm.push(prefix[j * TH_PREFIX + q]);
                item_info[2] = pm.h;
                item_info[3] = pm.word;
                item_info[4] = pm.n;
            }
        }
        __syncthreads();
        const int64_t item = (int64_t)(((unsigned long long)item_info[1] << 32) | item_info[0]);
        if (item >= nitems) break;
        const int j = (int)(item / items_per_col);
        const int64_t r_lo = (item % items_per_col) * TC_ITEM_ROWS, r_hi = min(n, r_lo + TC_ITEM_ROWS);
        const unsigned ph = item_info[2], pw = item_info[3], pn = item_info[4];
        const TextColumn tc = cols[j];
        const bool hashed = j < p.n_hashed;
        // statistics of this column's buckets: lane k holds bucket 32 w + k of word w, classes 0 / 1 (a warp sees at
        // most TC_ITEM_ROWS / TC_WARPS cells of an item, so 16 bits hold a count)
        unsigned null0 = 0u, null1 = 0u;
        if (STATS) {
#pragma unroll
            for (int w = 0; w < 2 * TC_FW; ++w) hist[w * 32 + lane] = 0;
        }

        for (int64_t g0 = r_lo + (int64_t)warp * 32; g0 < r_hi; g0 += TC_WARPS * 32) {
            const int64_t row = g0 + lane;
            const bool in_rows = row < r_hi;
            int64_t a = 0, b = 0;
            if (in_rows && cell_valid(tc, row)) {
                a = tc.offsets[row];
                b = tc.offsets[row + 1];
            }
            const int64_t len = b - a;
            const uint8_t* ptr = tc.data + a;
            const int sh = (int)(reinterpret_cast<uintptr_t>(ptr) & 15u);
            const uint8_t* base = ptr - sh;
            // chunks through the zero byte after the cell; a cell without bytes has none
            const int nch = len > 0 ? (int)((sh + len + 16) >> 4) : 0;
            const int end = sh + (int)min(len, (int64_t)0x7fffff00);       // cells are < 2 GiB (int32 positions)
#pragma unroll
            for (int w = 0; w < TC_FW; ++w) bm[w * 32 + lane] = 0u;
            bool deferred = p.force_defer != 0 && len > 0;
            if (len > 0x7fffff00ll) deferred = true;
            bool nohit = true, in_tok = false;
            int tok_start = 0, ndup = 0;
            unsigned len_sum = 0u, cl = 0u, cd = 0u;

            // the first chunk is masked at both ends; of the later ones only the chunk holding the cell's last byte
            // needs a mask, prepared here once (chunks behind it are the zero terminator)
            const int last_real = (end - 1) >> 4;
            unsigned lm0 = ~0u, lm1 = ~0u, lm2 = ~0u, lm3 = ~0u;
            {
                const int hi = end - 16 * last_real;
                lm0 = byte_range_mask(0, hi);
                lm1 = byte_range_mask(0, hi - 4);
                lm2 = byte_range_mask(0, hi - 8);
                lm3 = byte_range_mask(0, hi - 12);
            }
            auto load0 = [&]() -> uint4 {
                uint4 v = __ldg(reinterpret_cast<const uint4*>(base));
                const int hi = min(16, end);
                v.x &= byte_range_mask(sh, hi);
                v.y &= byte_range_mask(sh - 4, hi - 4);
                v.z &= byte_range_mask(sh - 8, hi - 8);
                v.w &= byte_range_mask(sh - 12, hi - 12);
                return make_uint4(lower4(v.x), lower4(v.y), lower4(v.z), lower4(v.w));
            };
            auto load = [&](int ci) -> uint4 {                     // ci >= 1
                if (ci > last_real) return make_uint4(0u, 0u, 0u, 0u);
                uint4 v = __ldg(reinterpret_cast<const uint4*>(base + 16 * (int64_t)ci));
                const unsigned all = ci == last_real ? 0u : ~0u;
                return make_uint4(lower4(v.x & (lm0 | all)), lower4(v.y & (lm1 | all)), lower4(v.z & (lm2 | all)),
                                  lower4(v.w & (lm3 | all)));
            };

            const int nch_eff = deferred ? 0 : nch;
            uint4 cur = nch_eff > 0 ? load0() : make_uint4(0u, 0u, 0u, 0u);
            int max_ch = nch_eff;
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) max_ch = max(max_ch, __shfl_xor_sync(FULL, max_ch, o));
            // Finding the tokens is cheap and stays with the chunk: two 16-bit masks per chunk (where a token ends,
            // where one starts) wait in a four-deep FIFO beside the ring.  The expensive part -- stop-word probe,
            // murmur3, bucket -- runs in ROUNDS in which every lane with a token waiting takes one, so that it
            // executes with most lanes busy instead of once per lane and token end.
            bool c_in = false;                   // classification side: a token is open at the end of the last chunk
            int c_start = 0;
            int hc = 0, p_start = 0, pend = 0;   // processing side: chunk of the next token end, open token, ends waiting
            unsigned h_end = 0u, h_start = 0u;
            bool p_carry = false;
            unsigned* ringl = ring + lane;

            for (int ci = 0;; ++ci) {
                const bool last = ci >= max_ch;
                const bool active = ci < nch_eff && !deferred;
                // ---- rounds: tokens whose bytes chunk ci is about to overwrite must go first; otherwise a round
                //      runs while most lanes have a token waiting
                for (;;) {
                    const bool must = pend > 0 && (last || (active && (p_carry ? (p_start >> 4) : hc) <= ci - 4));
                    const unsigned have = __ballot_sync(FULL, pend > 0);
                    if (!(__any_sync(FULL, must || pend >= TC_PEND_V) || __popc(have) >= TC_ROUND_LANES)) break;
                    if (pend > 0) {
                        while (h_end == 0u) {                      // on to the next chunk that holds a token end
                            if (h_start) {
                                p_start = (hc << 4) + __ffs(h_start) - 1;
                                p_carry = true;
                            }
                            ++hc;
                            const unsigned w = masks[(hc & 3) * 32 + lane];
                            h_end = w & 0xffffu;
                            h_start = w >> 16;
                        }
                        const int pe = __ffs(h_end) - 1;
                        h_end &= h_end - 1u;
                        int ps;
                        if (p_carry) {
                            ps = p_start;
                            p_carry = false;
                        } else {
                            ps = (hc << 4) + __ffs(h_start) - 1;
                            h_start &= h_start - 1u;
                        }
                        const int plen = (hc << 4) + pe - ps;
                        --pend;
                        if (plen > TC_BACK) {
                            deferred = true;                       // longer than the ring holds
                        } else if (plen >= p.min_token_length) {
                            const unsigned* rs = ringl + ((ps >> 2) & (TC_RING - 1)) * 32;
                            const unsigned first4 = __funnelshift_r(rs[0], rs[32], (ps & 3) * 8);
                            const unsigned c0 = first4 & 0xffu;
                            bool keep = true;
                            if (c0 == '_') {                       // a run of '_' only is not a token
                                keep = false;
                                for (int q = ps; q < ps + plen; ++q) keep |= tc_ring_byte(ring, lane, q) != '_';
                            }
                            if (keep && plen <= STOP_MAXLEN && (c0 - 'a') < 26u) {
                                const unsigned c1 = plen > 1 ? (first4 >> 8) & 0xffu : 0u;
                                if ((stop_pre[stop_pre_index(c0, c1)] >> plen) & 1u)
                                    keep = !tc_is_stop(ring, lane, ps, plen);
                            }
                            if (keep) {
                                nohit = false;
                                len_sum += (unsigned)plen;
                                if (hashed) {
                                    const unsigned hv = tc_hash_token(ring, lane, ps, plen, ph, pw, pn);
                                    const int bucket = p.pow2_mask >= 0 ? (int)(hv & (unsigned)p.pow2_mask)
                                                                        : non_negative_mod(hv, F);
                                    const unsigned bit = 1u << (bucket & 31);
                                    const int wi = (bucket >> 5) * 32 + lane;
                                    const unsigned old = bm[wi];
                                    if (!(old & bit)) {
                                        bm[wi] = old | bit;
                                    } else {
                                        // met before in this cell: count it in the list of repeated buckets
                                        int at = -1;
                                        for (int q = 0; q < ndup; ++q)
                                            if ((dups[q * 32 + lane] & 0xffffu) == (unsigned)bucket) at = q;
                                        if (at >= 0) {
                                            const unsigned d = dups[at * 32 + lane];
                                            if ((d >> 16) == 0xffffu) deferred = true;
                                            else dups[at * 32 + lane] = d + 0x10000u;
                                        } else if (ndup < TC_DUPS) {
                                            dups[ndup * 32 + lane] = (unsigned)bucket | (2u << 16);
                                            ++ndup;
                                        } else {
                                            deferred = true;
                                        }
                                    }
                                }
                            }
                        }
                        if (deferred) pend = 0;                    // nothing of a cell that is set aside counts
                    }
                }
                if (last) break;
                uint4 nxt = make_uint4(0u, 0u, 0u, 0u);
                if (active && ci + 1 < nch_eff) nxt = load(ci + 1);
                if (active) {
                    const unsigned m = tc_classify(cur, cl, cd, nxt.x, non_ascii);
                    // the ring: chunk c lives in rows 4 (c & 3) ..; rows 16 .. 20 repeat rows 0 .. 4, so that a token
                    // is read from consecutive rows whatever its position
                    const int slot = (ci & 3) * 4;
                    unsigned* rw = ringl + slot * 32;
                    rw[0] = cur.x;
                    rw[32] = cur.y;
                    rw[64] = cur.z;
                    rw[96] = cur.w;
                    if (slot == 0) {
                        ringl[16 * 32] = cur.x;
                        ringl[17 * 32] = cur.y;
                        ringl[18 * 32] = cur.z;
                        ringl[19 * 32] = cur.w;
                    } else if (slot == 4) {
                        ringl[20 * 32] = cur.x;
                    }
                    // position p ends a token when byte p - 1 is a token byte (or a token was open at the chunk's
                    // start, p = 0) and byte p is not; a token starts where a token byte follows a byte that is none
                    const unsigned ext = (m << 1) | (c_in ? 1u : 0u);
                    const unsigned endm = ext & ~m & 0xffffu, startm = m & ~ext & 0xffffu;
                    if (pend == 0) {
                        hc = ci;
                        h_end = endm;
                        h_start = startm;
                        p_carry = c_in;
                        p_start = c_start;
                    } else {
                        masks[(ci & 3) * 32 + lane] = endm | (startm << 16);
                    }
                    pend += __popc(endm);
                    c_in = (m >> 15) & 1u;
                    if (c_in && startm) c_start = (ci << 4) + 31 - __clz(startm);
                }
                cur = nxt;
            }
            __syncwarp();

            // ---- end of the 32 cells
            const int cls = (p.nclasses > 1 && class_ids && in_rows) ? class_ids[row] : 0;
            if (deferred && in_rows) {
                const int64_t cell = row * ncols + j;
                atomicOr(&defer_bits[cell >> 5], 1u << (cell & 31));
            }
            if (CELLS && in_rows) {
                // the record of the cell: 8 x 16 bytes
                unsigned* rec = records + ((size_t)row * ncols + j) * TC_REC_WORDS;
                const unsigned head = deferred ? TC_H_DEFER : ((unsigned)ndup | (nohit ? TC_H_NOHIT : 0u));
                uint4* r4 = reinterpret_cast<uint4*>(rec);
#pragma unroll
                for (int q = 0; q < 4; ++q)
                    r4[q] = make_uint4(bm[(4 * q) * 32 + lane], bm[(4 * q + 1) * 32 + lane], bm[(4 * q + 2) * 32 + lane],
                                       bm[(4 * q + 3) * 32 + lane]);
                unsigned t[16];
                t[0] = head;
                t[1] = len_sum;
                t[2] = 0u;
#pragma unroll
                for (int q = 0; q < TC_DUPS; ++q) t[3 + q] = q < ndup ? dups[q * 32 + lane] : 0u;
#pragma unroll
                for (int q = 0; q < 4; ++q) r4[4 + q] = make_uint4(t[4 * q], t[4 * q + 1], t[4 * q + 2], t[4 * q + 3]);
            }
            if (STATS) {
                const bool counted = in_rows && !deferred;
                if (deferred) {
#pragma unroll
                    for (int w = 0; w < TC_FW; ++w) bm[w * 32 + lane] = 0u;
                }
                const double e_row = (label && counted) ? label[row] - label_shift : 0.0;
                if (two_classes) {
                    const unsigned c1 = __ballot_sync(FULL, counted && cls == 1);
                    const unsigned nh = __ballot_sync(FULL, counted && nohit);
                    null0 += __popc(nh & ~c1);
                    null1 += __popc(nh & c1);
#pragma unroll
                    for (int w = 0; w < TC_FW; ++w) {
                        const unsigned x = tc_transpose32(bm[w * 32 + lane], lane);
                        hist[(TC_FW + w) * 32 + lane] += (unsigned short)__popc(x & c1);
                        hist[w * 32 + lane] += (unsigned short)__popc(x & ~c1);
                    }
                    if (label && counted) {
                        for (int w = 0; w < TC_FW; ++w) {
                            unsigned x = bm[w * 32 + lane];
                            while (x) {
                                const int bkt = w * 32 + __ffs(x) - 1;
                                x &= x - 1u;
                                if (bkt < F) atomicAdd(&bucket_wsum[(int64_t)j * F + bkt], e_row);
                            }
                        }
                    }
                } else if (counted) {
                    // many label classes: straight to the global tables
                    for (int w = 0; w < TC_FW; ++w) {
                        unsigned x = bm[w * 32 + lane];
                        while (x) {
                            const int bkt = w * 32 + __ffs(x) - 1;
                            x &= x - 1u;
                            atomicAdd(&bucket_class_sum[((int64_t)j * F + bkt) * p.nclasses + cls], 1u);
                            if (label) atomicAdd(&bucket_wsum[(int64_t)j * F + bkt], e_row);
                        }
                    }
                    if (nohit) atomicAdd(&bucket_class_sum[(int64_t)(p.null_base + j) * p.nclasses + cls], 1u);
                }
                if (counted) {
                    if (label && nohit) atomicAdd(&bucket_wsum[p.null_base + j], e_row);
                    // buckets met more than once: everything beyond "present once"
                    if (!p.binary_freq)
                        for (int q = 0; q < ndup; ++q) {
                            const unsigned e = dups[q * 32 + lane];
                            const unsigned v = (unsigned)((int64_t)j * F) + (e & 0xffffu), cnt = e >> 16;
                            atomicAdd(&bucket_class_sum[(int64_t)v * p.nclasses + cls], cnt - 1u);
                            if (label) atomicAdd(&bucket_wsum[v], (double)(cnt - 1u) * e_row);
                            th_rare_update(make_uint2(v, cnt), bucket_dup, bucket_sq_extra, bucket_aux);
                        }
                    if (p.len_base >= 0 && len_sum > 0u) {
                        const unsigned v = (unsigned)(p.len_base + j);
                        atomicAdd(&bucket_class_sum[(int64_t)v * p.nclasses + cls], len_sum);
                        if (label) atomicAdd(&bucket_wsum[v], (double)len_sum * e_row);
                        if (len_sum > 1u) th_rare_update(make_uint2(v, len_sum), bucket_dup, bucket_sq_extra, bucket_aux);
                    }
                }
            }
            __syncwarp();
        }

        if (STATS && two_classes) {
            // this warp's share of the column's statistics
#pragma unroll
            for (int w = 0; w < TC_FW; ++w) {
                const int bkt = w * 32 + lane;
                if (bkt < F && hashed) {
                    const int64_t v = (int64_t)j * F + bkt;
                    const unsigned a0 = hist[w * 32 + lane], a1 = hist[(TC_FW + w) * 32 + lane];
                    if (a0) atomicAdd(&bucket_class_sum[v * p.nclasses], a0);
                    if (a1) atomicAdd(&bucket_class_sum[v * p.nclasses + 1], a1);
                }
            }
            if (lane == 0) {
                const int64_t v = p.null_base + j;
                if (null0) atomicAdd(&bucket_class_sum[v * p.nclasses], null0);
                if (null1) atomicAdd(&bucket_class_sum[v * p.nclasses + 1], null1);
            }
        }
    }
    if (non_ascii) flags[0] = 1;
}

// the statistics of the cells the fast path set aside: a thread per cell
__global__ void __launch_bounds__(128)
text_cells_slow_kernel(const TextColumn* __restrict__ cols, int ncols, int64_t n, const uint8_t* __restrict__ prefix,
                       const int* __restrict__ prefix_len, TcParams p, const uint8_t* __restrict__ class_ids,
                       const unsigned* __restrict__ defer_bits, unsigned int* __restrict__ bucket_class_sum,
                       unsigned int* __restrict__ bucket_dup, unsigned long long* __restrict__ bucket_sq_extra,
                       unsigned int* __restrict__ bucket_aux, const double* __restrict__ label, double label_shift,
                       double* __restrict__ bucket_wsum, int* __restrict__ flags, TsPack pk) {
    const int64_t ncells = n * ncols;
    for (int64_t w = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; w * 32 < ncells;
         w += (int64_t)gridDim.x * blockDim.x) {
        unsigned bits = defer_bits[w];
        while (bits) {
            const int64_t cell = w * 32 + __ffs(bits) - 1;
            bits &= bits - 1u;
            const int64_t row = cell / ncols;
            const int j = (int)(cell % ncols);
            const TextColumn tc = cols[j];
            const int64_t a = tc.offsets[row], b = tc.offsets[row + 1];
            unsigned short counts[TC_MAXF];
            for (int q = 0; q < p.num_features; ++q) counts[q] = 0;
            unsigned len_sum = 0u;
            bool na = false, ovf = false;
            const bool hashed = j < p.n_hashed;
            const bool nohit = tc_scalar_cell(tc.data + a, b - a, p, prefix + j * TH_PREFIX,
                                              (p.prepend && hashed) ? prefix_len[j] : 0, hashed, counts, len_sum, na, ovf);
            if (na) flags[0] = 1;
            if (ovf) flags[2] = 1;
            if (pk.arena != nullptr) {
                // packed cells: this cell gets a slot of its own; its four entries in the group point at it
                const TsGroup g = pk.grp[(row >> 5) * ncols + j];
                if (g.off != ~0ull) {
                    const unsigned long long slot = atomicAdd(pk.cursor, (unsigned long long)TS_SLOT_ENTRIES * 2);
                    if (slot + TS_SLOT_ENTRIES * 2 > pk.arena_bytes) {
                        pk.flags[0] = 1;
                    } else {
                        unsigned short* sl = reinterpret_cast<unsigned short*>(pk.arena + slot);
                        for (int q = 0; q < TC_MAXF; ++q) sl[q] = hashed && q < p.num_features ? counts[q] : (unsigned short)0;
                        sl[TC_MAXF] = (unsigned short)(len_sum & 0xffffu);
                        sl[TC_MAXF + 1] = (unsigned short)(len_sum >> 16);
                        sl[TC_MAXF + 2] = nohit ? 1 : 0;
                        unsigned char* at = pk.arena + g.off + (pk.rel[row * ncols + j] & TS_REL_MASK);
                        for (int q = 0; q < 8; ++q) at[q] = (unsigned char)((slot >> (8 * q)) & 0xffull);
                    }
                }
            }
            const int cls = (p.nclasses > 1 && class_ids) ? class_ids[row] : 0;
            const double e_row = label ? label[row] - label_shift : 0.0;
            auto add = [&](unsigned v, unsigned cnt) {
                atomicAdd(&bucket_class_sum[(int64_t)v * p.nclasses + cls], cnt);
                if (label) atomicAdd(&bucket_wsum[v], (double)cnt * e_row);
                if (cnt > 1u) th_rare_update(make_uint2(v, cnt), bucket_dup, bucket_sq_extra, bucket_aux);
            };
            if (hashed)
                for (int q = 0; q < p.num_features; ++q)
                    if (counts[q]) add((unsigned)((int64_t)j * p.num_features + q), p.binary_freq ? 1u : counts[q]);
            if (nohit) add((unsigned)(p.null_base + j), 1u);
            if (p.len_base >= 0 && len_sum > 0u) add((unsigned)(p.len_base + j), len_sum);
        }
    }
}

// TextStats.lengthCounts (SmartTextVectorizer.scala:207-299): per column a histogram of the cell lengths -- the length of
// the cleaned cell (TextLengthType.FullEntry) or the lengths of the cell's tokens (Tokens) -- and the first row every
// length was met in (the reference's fold stops adding once the map holds more than maxCardinality lengths: the
// caller finds that row from the first rows and asks again for the rows up to it).  Lengths >= nbins - 1 share the
// last bin.  A thread per cell.
__global__ void __launch_bounds__(256)
text_length_hist_kernel(const TextColumn* __restrict__ cols, int ncols, int64_t n, int64_t row_base, int clean, int tokens,
                        int min_token_length, int nbins, unsigned long long* __restrict__ hist,
                        unsigned long long* __restrict__ first_row, int* __restrict__ flags) {
    const int c = blockIdx.y;
    const TextColumn tc = cols[c];
    unsigned long long* h = hist + (size_t)c * nbins;
    unsigned long long* f = first_row + (size_t)c * nbins;
    bool non_ascii = false;
    for (int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; r < n; r += (int64_t)gridDim.x * blockDim.x) {
        if (!cell_valid(tc, r)) continue;
        const int64_t a = tc.offsets[r], b = tc.offsets[r + 1];
        const uint8_t* s = tc.data + a;
        if (!tokens) {
            const int m = clean_length_cell(s, b - a, clean != 0, non_ascii);
            const int bin = min(m, nbins - 1);
            atomicAdd(&h[bin], 1ull);
            atomicMin(&f[bin], (unsigned long long)(row_base + r));
        } else {
            tc_scalar_tokens(s, b - a, min_token_length, non_ascii, [&](int64_t k, int plen) {
                if (tc_scalar_is_stop(s, k, plen)) return;
                const int bin = min(plen, nbins - 1);
                atomicAdd(&h[bin], 1ull);
                atomicMin(&f[bin], (unsigned long long)(row_base + r));
            });
        }
    }
    if (non_ascii) flags[0] = 1;
}

// keepw [j * TC_FW + k]: the buckets of bitmap word k of column j whose vector column SanityChecker kept
__global__ void tc_keep_words_kernel(const int32_t* __restrict__ newpos, int n_hashed, int F, unsigned* __restrict__ keepw) {
    const int w = blockIdx.x * blockDim.x + threadIdx.x;
    if (w >= n_hashed * TC_FW) return;
    const int j = w / TC_FW, k = w % TC_FW;
    unsigned x = 0u;
    for (int q = 0; q < 32; ++q) {
        const int bkt = k * 32 + q;
        if (bkt < F && newpos[(int64_t)j * F + bkt] >= 0) x |= 1u << q;
    }
    keepw[w] = x;
}

// records -> sorted sparse rows of the kept columns: a warp per row
__global__ void __launch_bounds__(256)
text_rows_kernel(const TextColumn* __restrict__ cols, int ncols, int64_t n, const uint8_t* __restrict__ prefix,
                 const int* __restrict__ prefix_len, TcParams p, unsigned* records,
                 const unsigned* __restrict__ keepw, const int32_t* __restrict__ newpos,
                 int64_t* __restrict__ row_start, int32_t* __restrict__ row_nnz, int32_t* __restrict__ out_idx,
                 int32_t* __restrict__ out_val, unsigned long long* __restrict__ cursor, int64_t capacity,
                 int* __restrict__ flags) {
    extern __shared__ __align__(16) unsigned char tr_smem[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int nwarps = blockDim.x >> 5;
    // per warp: kept bitmap words [ncols * TC_FW] u32 | counts of a set-aside cell [TC_MAXF] u16 | word offsets u16
    const size_t per_warp = ((size_t)ncols * TC_FW * 4 + TC_MAXF * 2 + (size_t)ncols * TC_FW * 2 + 15) & ~size_t(15);
    unsigned* kept_words = reinterpret_cast<unsigned*>(tr_smem + per_warp * warp);
    unsigned short* counts = reinterpret_cast<unsigned short*>(kept_words + (size_t)ncols * TC_FW);
    unsigned short* woff = counts + TC_MAXF;

    const int F = p.num_features;
    const unsigned lt = (1u << lane) - 1u;
    for (int64_t row = (int64_t)blockIdx.x * nwarps + warp; row < n; row += (int64_t)gridDim.x * nwarps) {
        unsigned* rrow = records + (size_t)row * ncols * TC_REC_WORDS;
        // ---- cells the fast path set aside are tokenised here, by one lane, and the bitmap of their record rebuilt
        bool any_deferred = false;
        for (int j0 = 0; j0 < ncols; j0 += 32) {
            const int j = j0 + lane;
            const bool d = j < ncols && (rrow[j * TC_REC_WORDS + TC_FW] & TC_H_DEFER);
            any_deferred |= __any_sync(FULL, d);
        }
        if (any_deferred) {
            for (int j = 0; j < ncols; ++j) {
                const unsigned head = rrow[j * TC_REC_WORDS + TC_FW];
                if (!(head & TC_H_DEFER)) continue;
                for (int q = lane; q < TC_MAXF; q += 32) counts[q] = 0;
                __syncwarp();
                unsigned len_sum = 0u, nohit = 0u;
                if (lane == 0) {
                    const TextColumn tc = cols[j];
                    const int64_t a = tc.offsets[row], b = tc.offsets[row + 1];
                    bool na = false, ovf = false;
                    nohit = tc_scalar_cell(tc.data + a, b - a, p, prefix + j * TH_PREFIX,
                                           (p.prepend && j < p.n_hashed) ? prefix_len[j] : 0,
                                           j < p.n_hashed, counts, len_sum, na, ovf) ? 1u : 0u;
                    if (na) flags[0] = 1;
                    if (ovf) flags[2] = 1;
                }
                __syncwarp();
                len_sum = __shfl_sync(FULL, len_sum, 0);
                nohit = __shfl_sync(FULL, nohit, 0);
                unsigned* rec = rrow + j * TC_REC_WORDS;
                if (lane < TC_FW) {
                    unsigned x = 0u;
                    for (int q = 0; q < 32; ++q) x |= (counts[lane * 32 + q] ? 1u : 0u) << q;
                    rec[lane] = x;
                }
                if (lane == 0) {
                    rec[TC_FW] = TC_H_DEFER | (nohit ? TC_H_NOHIT : 0u);
                    rec[TC_FW + 1] = len_sum;
                }
                __syncwarp();
            }
        }
        // ---- a lane per cell loads the record (vector loads) and leaves the kept bits of its bitmap in shared memory
        for (int j0 = 0; j0 < p.n_hashed; j0 += 32) {
            const int j = j0 + lane;
            if (j < p.n_hashed) {
                const uint4* r4 = reinterpret_cast<const uint4*>(rrow + j * TC_REC_WORDS);
                const uint4* k4 = reinterpret_cast<const uint4*>(keepw + j * TC_FW);
                uint4* o4 = reinterpret_cast<uint4*>(kept_words + j * TC_FW);
#pragma unroll
                for (int q = 0; q < TC_FW / 4; ++q) {
                    const uint4 v = r4[q], kq = k4[q];
                    o4[q] = make_uint4(v.x & kq.x, v.y & kq.y, v.z & kq.z, v.w & kq.w);
                }
            }
        }
        __syncwarp();
        // ---- a lane per bitmap word: entries before every word (consecutive lanes then write consecutive entries)
        const int nwords = p.n_hashed * TC_FW;
        int hash_total = 0;
        for (int w0 = 0; w0 < nwords; w0 += 32) {
            const int w = w0 + lane;
            const int c = w < nwords ? __popc(kept_words[w]) : 0;
            int incl = c;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int up = __shfl_up_sync(FULL, incl, o);
                if (lane >= o) incl += up;
            }
            if (w < nwords) woff[w] = (unsigned short)(hash_total + incl - c);
            hash_total += __shfl_sync(FULL, incl, 31);
        }
        int total = hash_total;
        for (int j0 = 0; j0 < ncols; j0 += 32) {
            const int j = j0 + lane;
            const bool has = j < ncols && p.len_base >= 0 && rrow[j * TC_REC_WORDS + TC_FW + 1] > 0u &&
                             newpos[p.len_base + j] >= 0;
            total += __popc(__ballot_sync(FULL, has));
        }
        for (int j0 = 0; j0 < ncols; j0 += 32) {
            const int j = j0 + lane;
            const bool has = j < ncols && p.track_nulls && (rrow[j * TC_REC_WORDS + TC_FW] & TC_H_NOHIT) &&
                             newpos[p.null_base + j] >= 0;
            total += __popc(__ballot_sync(FULL, has));
        }
        unsigned long long base = 0;
        if (lane == 0) {
            base = atomicAdd(cursor, (unsigned long long)total);
            row_start[row] = (int64_t)base;
            row_nnz[row] = total;
            if ((int64_t)base + total > capacity) flags[3] = 1;
        }
        base = __shfl_sync(FULL, base, 0);
        if ((int64_t)base + total > capacity) continue;      // sizing call (capacity 0) or too small: nothing written
        __syncwarp();
        for (int w0 = 0; w0 < nwords; w0 += 32) {
            const int w = w0 + lane;
            if (w < nwords) {
                unsigned x = kept_words[w];
                int64_t at = (int64_t)base + woff[w];
                const int32_t* np = newpos + (int64_t)(w / TC_FW) * F + (w % TC_FW) * 32;
                while (x) {
                    const int bq = __ffs(x) - 1;
                    x &= x - 1u;
                    out_idx[at] = np[bq];
                    out_val[at] = 1;
                    ++at;
                }
            }
        }
        __syncwarp();
        // ---- term frequencies > 1, a lane per cell from the record's list (a set-aside cell is patched below)
        if (!p.binary_freq) {
            for (int j0 = 0; j0 < p.n_hashed; j0 += 32) {
                const int j = j0 + lane;
                if (j < p.n_hashed) {
                    const unsigned head = rrow[j * TC_REC_WORDS + TC_FW];
                    const int nd = (head & TC_H_DEFER) ? 0 : (int)(head & 0xffu);
                    for (int d = 0; d < nd; ++d) {
                        const unsigned e = rrow[j * TC_REC_WORDS + TC_FW + 3 + d];
                        const int bkt = (int)(e & 0xffffu), w = j * TC_FW + (bkt >> 5);
                        const unsigned wword = kept_words[w];
                        if ((wword >> (bkt & 31)) & 1u)
                            out_val[(int64_t)base + woff[w] + __popc(wword & ((1u << (bkt & 31)) - 1u))] = (int32_t)(e >> 16);
                    }
                }
            }
        }
        __syncwarp();
        // ---- set-aside cells: their term frequencies > 1 need the counts (tokenised again; rare)
        if (!p.binary_freq && any_deferred) {
            for (int j = 0; j < p.n_hashed; ++j) {
                const unsigned head = rrow[j * TC_REC_WORDS + TC_FW];
                if (!(head & TC_H_DEFER)) continue;
                for (int q = lane; q < TC_MAXF; q += 32) counts[q] = 0;
                __syncwarp();
                if (lane == 0) {
                    const TextColumn tc = cols[j];
                    const int64_t a = tc.offsets[row], b = tc.offsets[row + 1];
                    bool na = false, ovf = false;
                    unsigned ls = 0u;
                    tc_scalar_cell(tc.data + a, b - a, p, prefix + j * TH_PREFIX, p.prepend ? prefix_len[j] : 0, true,
                                   counts, ls, na, ovf);
                }
                __syncwarp();
                for (int bkt = lane; bkt < F; bkt += 32) {
                    const unsigned cnt = counts[bkt];
                    const int w = j * TC_FW + (bkt >> 5);
                    const unsigned wword = kept_words[w];
                    if (cnt > 1u && ((wword >> (bkt & 31)) & 1u))
                        out_val[(int64_t)base + woff[w] + __popc(wword & ((1u << (bkt & 31)) - 1u))] = (int32_t)cnt;
                }
                __syncwarp();
            }
        }
        // ---- text lengths, then null indicators
        int at = hash_total;
        for (int j0 = 0; j0 < ncols; j0 += 32) {
            const int j = j0 + lane;
            unsigned lv = 0u;
            bool has = false;
            if (j < ncols && p.len_base >= 0) {
                lv = rrow[j * TC_REC_WORDS + TC_FW + 1];
                has = lv > 0u && newpos[p.len_base + j] >= 0;
            }
            const unsigned bl = __ballot_sync(FULL, has);
            if (has) {
                out_idx[(int64_t)base + at + __popc(bl & lt)] = newpos[p.len_base + j];
                out_val[(int64_t)base + at + __popc(bl & lt)] = (int32_t)lv;
            }
            at += __popc(bl);
        }
        for (int j0 = 0; j0 < ncols; j0 += 32) {
            const int j = j0 + lane;
            const bool has = j < ncols && p.track_nulls && (rrow[j * TC_REC_WORDS + TC_FW] & TC_H_NOHIT) &&
                             newpos[p.null_base + j] >= 0;
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
