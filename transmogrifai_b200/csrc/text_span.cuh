// Tokenise + hash for separate hash spaces of at most 512 buckets per column, third cut of the work (included by text.cu
// after text_cells.cuh; same records, same statistics arrays, same slow path for the cells set aside).
//
//   text_span_kernel    a warp takes 32 consecutive rows of ONE column and treats their bytes -- one contiguous SPAN of
//                       the Arrow buffer -- as a single stream: every step the 32 lanes load one 512-byte slab (16 bytes
//                       per lane, fully coalesced), lower-case and classify it in registers (SWAR; the UAX#29 Mid* rules
//                       look at the neighbouring lanes' bytes through shuffles), and turn it into token starts and token
//                       ends; cell borders are just positions in the stream at which a token is cut.  Starts and ends go
//                       to a queue in shared memory (prefix sums over the lanes give every lane its place), the slab
//                       itself to a 2 KB ring.  Whenever 32 tokens wait, every lane takes ONE: stop-word filter, murmur3
//                       from the cached prefix state, bucket -- all 32 lanes busy whatever the cells' lengths are.  The
//                       cell a token belongs to travels with its start (rank of the cell among the span's non-empty
//                       cells); term frequencies go to the cell's 512-bit bitmap with a shared-memory atomicOr, a bucket
//                       met again is appended to the cell's short list.  At the end of the span lane k is cell k again:
//                       it collapses its list and the warp proceeds exactly like text_cells_kernel (transposed bitmaps
//                       for the fit statistics, a 128-byte record per cell for the transform).
//
// Against text_cells_kernel (a lane per cell): no lane idles because its cell is shorter than its neighbours', rounds do
// not follow the longest cell of the 32, and the per-lane FIFO / ring bookkeeping is gone.
#pragma once

namespace tmog {

#ifndef TS_WARPS_V
#define TS_WARPS_V 8
#endif
#ifndef TS_ITEM_ROWS_V
#define TS_ITEM_ROWS_V 2048
#endif
constexpr int TS_WARPS = TS_WARPS_V;
constexpr int TS_MIN_CTAS = 24 / TS_WARPS;   // 24 warps per SM at 80 registers
constexpr int TS_ITEM_ROWS = TS_ITEM_ROWS_V;   // rows of one column a CTA takes at a time
constexpr int TS_SLAB = 512;               // bytes of a step
constexpr int TS_RING_WORDS = 512;         // the last four slabs
constexpr int TS_RING_PAD = 8;             // + the first words again behind the last (a token is read without a wrap)
constexpr int TS_POS_MASK = TS_RING_WORDS * 4 - 1;
constexpr int TS_Q = 512;                  // token queue: a slab can add 256 tokens to the < 32 left waiting
constexpr int TS_APP = 16;                 // repeated buckets a cell may append before it is set aside
constexpr int TS_BM_STRIDE = TC_FW + 1;    // bitmap words of a cell + 1: conflict free by cell and by word
constexpr int TS_MAX_SPAN = (1 << 26) - 2048;   // bytes of 32 cells the fast path takes (positions are packed with a rank)
constexpr int TS_WARP_BYTES = (TS_RING_WORDS + TS_RING_PAD) * 4 + 32 * TS_BM_STRIDE * 4 + 5 * 32 * 4 + TS_APP * 32 * 2 + TS_Q * 4;
constexpr int TS_BLOOM_WORDS = 128;        // second stop-word filter: a bit per (first four bytes, length) of a stop word
static_assert(TS_WARP_BYTES % 16 == 0, "per-warp shared memory must stay 16-byte aligned");

__device__ __forceinline__ unsigned ts_ring_byte(const unsigned* ring, int pos) {
    return (ring[(pos & TS_POS_MASK) >> 2] >> ((pos & 3) * 8)) & 0xffu;
}

// stop-word probe of the token [ps, ps + plen) (plen <= STOP_MAXLEN), ps a ring position
__device__ __forceinline__ bool ts_is_stop(const unsigned* ring, int ps, int plen) {
    const unsigned* rs = ring + (ps >> 2);
    const int sh = (ps & 3) * 8;
    const unsigned a = rs[0], b = rs[1], c = rs[2], d = rs[3];
    const unsigned l0 = __funnelshift_r(a, b, sh) & byte_range_mask(0, plen);
    const unsigned l1 = __funnelshift_r(b, c, sh) & byte_range_mask(0, plen - 4);
    const unsigned l2 = __funnelshift_r(c, d, sh) & byte_range_mask(0, plen - 8);
    const StopSlot e1 = c_stop_packed[stop_slot_hash(l0, l1, l2)];
    const StopSlot e2 = c_stop_packed[stop_slot_hash2(l0, l1, l2)];
    const unsigned d1 = (e1.len ^ (unsigned)plen) | (e1.w0 ^ l0) | (e1.w1 ^ l1) | (e1.w2 ^ l2);
    const unsigned d2 = (e2.len ^ (unsigned)plen) | (e2.w0 ^ l0) | (e2.w1 ^ l1) | (e2.w2 ^ l2);
    return d1 == 0u || d2 == 0u;
}

// murmur3 of "<prefix><token>" continued from the state after the prefix (tc_hash_token over the linear ring)
__device__ __forceinline__ unsigned ts_hash_token(const unsigned* ring, int ps, int plen, unsigned ph, unsigned pw,
                                                  unsigned pn) {
    const int r = (int)(pn & 3u);
    const unsigned keep_hi = ~((1u << (8 * r)) - 1u);
    const int tot = r + plen, nfull = tot >> 2, tail = tot & 3;
    const int q0 = ps - r, sh = (q0 & 3) * 8;
    const int w0 = (q0 >> 2) & (TS_RING_WORDS - 1);      // q0 may be slightly negative: the bytes there are replaced
    unsigned h = ph, wd;
    if (tot <= 15) {
        const unsigned* rp = ring + w0;
        const unsigned a0 = rp[0], a1 = rp[1], a2 = rp[2], a3 = rp[3], a4 = rp[4];
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
        unsigned nxt = ring[w0];
        for (int q = 0; q < nfull; ++q) {
            const unsigned cur = nxt;
            nxt = ring[(w0 + q + 1) & (TS_RING_WORDS - 1)];
            unsigned x = __funnelshift_r(cur, nxt, sh);
            if (q == 0) x = (x & keep_hi) | pw;
            h = mix_h1(h, mix_k1(x));
        }
        wd = __funnelshift_r(nxt, ring[(w0 + nfull + 1) & (TS_RING_WORDS - 1)], sh);
    }
    unsigned t = mix_h1(h, mix_k1((unsigned)(int)(signed char)(wd & 0xffu)));
    h = tail > 0 ? t : h;
    t = mix_h1(h, mix_k1((unsigned)(int)(signed char)((wd >> 8) & 0xffu)));
    h = tail > 1 ? t : h;
    t = mix_h1(h, mix_k1((unsigned)(int)(signed char)((wd >> 16) & 0xffu)));
    h = tail > 2 ? t : h;
    return fmix32(h, pn + (unsigned)plen);
}

// the same for a token of at most 15 bytes with the prefix remainder, from the four words that follow the position
// (token start - prefix remainder), already loaded and aligned: no loads, a fixed number of steps
__device__ __forceinline__ unsigned ts_hash_words(unsigned s0r, unsigned s1, unsigned s2, unsigned s3, int plen, unsigned ph,
                                                  unsigned pw, unsigned pn) {
    const int r = (int)(pn & 3u);
    const unsigned keep_hi = ~((1u << (8 * r)) - 1u);
    const int tot = r + plen, nfull = tot >> 2, tail = tot & 3;
    const unsigned s0 = (s0r & keep_hi) | pw;
    unsigned h = ph;
    unsigned t = mix_h1(h, mix_k1(s0));
    h = nfull > 0 ? t : h;
    t = mix_h1(h, mix_k1(s1));
    h = nfull > 1 ? t : h;
    t = mix_h1(h, mix_k1(s2));
    h = nfull > 2 ? t : h;
    const unsigned wd = nfull == 0 ? s0 : nfull == 1 ? s1 : nfull == 2 ? s2 : s3;
    t = mix_h1(h, mix_k1((unsigned)(int)(signed char)(wd & 0xffu)));
    h = tail > 0 ? t : h;
    t = mix_h1(h, mix_k1((unsigned)(int)(signed char)((wd >> 8) & 0xffu)));
    h = tail > 1 ? t : h;
    t = mix_h1(h, mix_k1((unsigned)(int)(signed char)((wd >> 16) & 0xffu)));
    h = tail > 2 ? t : h;
    return fmix32(h, pn + (unsigned)plen);
}

// bits 7, 15, 23, 31 of four words -> 16 bits
__device__ __forceinline__ unsigned ts_compress(const unsigned (&t)[4]) {
    unsigned m = 0u;
#pragma unroll
    for (int q = 0; q < 4; ++q) m |= ((((t[q] >> 7) & 0x01010101u) * 0x00204081u) >> 21 & 0xfu) << (4 * q);
    return m;
}

constexpr int TS_FIXED_SMEM = TH_STOP_PRE_BYTES + 64 + TS_BLOOM_WORDS * 4;

template <bool STATS, bool CELLS, bool PACK>
__global__ void __launch_bounds__(TS_WARPS * 32, TS_MIN_CTAS)
text_span_kernel(const TextColumn* __restrict__ cols, int ncols, int64_t n, const uint8_t* __restrict__ prefix,
                 const int* __restrict__ prefix_len, TcParams p, const uint8_t* __restrict__ class_ids,
                 TcScratch* __restrict__ scratch, unsigned* __restrict__ defer_bits,
                 unsigned int* __restrict__ bucket_class_sum, unsigned int* __restrict__ bucket_dup,
                 unsigned long long* __restrict__ bucket_sq_extra, unsigned int* __restrict__ bucket_aux,
                 const double* __restrict__ label, double label_shift, double* __restrict__ bucket_wsum,
                 unsigned* __restrict__ records, int* __restrict__ flags, TsPack pk) {
    extern __shared__ __align__(16) unsigned char ts_smem[];
    unsigned short* stop_pre = reinterpret_cast<unsigned short*>(ts_smem);
    unsigned* item_info = reinterpret_cast<unsigned*>(ts_smem + TH_STOP_PRE_BYTES);
    unsigned* bloom = item_info + 16;
    // fit: rows of each label class holding each bucket of the CTA's column, [class][TC_MAXF]
    unsigned* hist = bloom + TS_BLOOM_WORDS;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    unsigned char* wmem = ts_smem + TS_FIXED_SMEM + (STATS ? 2 * TC_MAXF * 4 : 0) + (size_t)warp * TS_WARP_BYTES;
    unsigned* ring = reinterpret_cast<unsigned*>(wmem);          // [TS_RING_WORDS + TS_RING_PAD]
    unsigned* bm = ring + TS_RING_WORDS + TS_RING_PAD;           // [32][TS_BM_STRIDE]: word w of cell c at c * 17 + w
    unsigned* lens = bm + 32 * TS_BM_STRIDE;                     // summed length of the cell's surviving tokens
    unsigned* napp = lens + 32;                                  // repeated buckets appended
    unsigned* dflag = napp + 32;                                 // the cell is set aside
    unsigned* tbl = dflag + 32;                                  // rank among the non-empty cells -> cell
    unsigned* bndw = tbl + 32;                                   // cell borders of the slab, 16 bits per lane
    unsigned* tq = bndw + 32;                                    // [TS_Q] tokens: ring position | rank << 11 | length << 16
    unsigned short* app = reinterpret_cast<unsigned short*>(tq + TS_Q);   // [TS_APP][32]: bucket | count << 9

    for (int i = tid; i < STOP_PRE; i += TS_WARPS * 32) stop_pre[i] = c_stop_pre[i];
    for (int i = tid; i < TS_BLOOM_WORDS; i += TS_WARPS * 32) bloom[i] = c_stop_bloom[i];

    const int64_t items_per_col = (n + TS_ITEM_ROWS - 1) / TS_ITEM_ROWS;
    const int64_t nitems = items_per_col * ncols;
    const int F = p.num_features;
    const bool two_classes = p.nclasses <= 2;
    const unsigned lt = (1u << lane) - 1u;
    const int min_len = max(1, p.min_token_length);
    unsigned non_ascii = 0u;

    for (;;) {
        __syncthreads();
        if (tid == 0) {
            const unsigned long long it = atomicAdd(&scratch->item_counter, 1ull);
            item_info[0] = (unsigned)(it & 0xffffffffu);
            item_info[1] = (unsigned)(it >> 32);
            if ((int64_t)it < nitems) {
                const int j = (int)((int64_t)it % ncols);
                Murmur pm;
                pm.init();
                const int plen = (p.prepend && j < p.n_hashed) ? prefix_len[j] : 0;   // ignored columns have no prefix
                for (int q = 0; q < plen; ++q) pm.push(prefix[j * TH_PREFIX + q]);
                item_info[2] = pm.h;
                item_info[3] = pm.word;
                item_info[4] = pm.n;
            }
        }
        if (STATS)
            for (int i = tid; i < 2 * TC_MAXF; i += TS_WARPS * 32) hist[i] = 0u;
        __syncthreads();
        const int64_t item = (int64_t)(((unsigned long long)item_info[1] << 32) | item_info[0]);
        if (item >= nitems) break;
        // items run over the columns first: the CTAs at work at any time are spread over all columns, and so are their
        // atomics on the columns' statistics
        const int j = (int)(item % ncols);
        const int64_t r_lo = (item / ncols) * TS_ITEM_ROWS, r_hi = min(n, r_lo + TS_ITEM_ROWS);
        const unsigned ph = item_info[2], pw = item_info[3], pn = item_info[4];
        const TextColumn tc = cols[j];
        const bool hashed = j < p.n_hashed;
        // the summed token length of a cell is needed for the text-length column and, where nothing is hashed, to
        // know whether any token survived
        const bool need_lens = p.len_base >= 0 || !hashed;
        unsigned null0 = 0u, null1 = 0u;

        for (int64_t g0 = r_lo + (int64_t)warp * 32; g0 < r_hi; g0 += TS_WARPS * 32) {
            const int64_t row = g0 + lane;
            const bool in_rows = row < r_hi;
            const int64_t o_lo = tc.offsets[min(row, r_hi)], o_hi = tc.offsets[min(row + 1, r_hi)];
            const bool valid = in_rows && cell_valid(tc, row);
            const int64_t len = o_hi - o_lo;
            const int64_t span_a = __shfl_sync(FULL, o_lo, 0), span_b = __shfl_sync(FULL, o_hi, 31);
            const int64_t span = span_b - span_a;
            const uint8_t* ptr = tc.data + span_a;
            const int a_rel = (int)(reinterpret_cast<uintptr_t>(ptr) & 15u);
            const uint8_t* base = ptr - a_rel;
            const bool skip = span > TS_MAX_SPAN || p.force_defer != 0;
            const int b_rel = a_rel + (int)(skip ? 0 : span);
            const int st = a_rel + (int)min(o_lo - span_a, (int64_t)TS_MAX_SPAN);        // where the cell starts
            const bool ne = len > 0;
            const unsigned ne_mask = __ballot_sync(FULL, ne), valid_mask = __ballot_sync(FULL, valid);
            const int cls = (p.nclasses > 1 && class_ids && in_rows) ? class_ids[row] : 0;
            const unsigned c1_mask = __ballot_sync(FULL, cls == 1);
            if (ne) tbl[__popc(ne_mask & lt)] = (unsigned)lane;
#pragma unroll
            for (int w = 0; w < TS_BM_STRIDE; ++w) bm[w * 32 + lane] = 0u;
            lens[lane] = 0u;
            napp[lane] = 0u;
            dflag[lane] = (skip && valid && ne) ? 1u : 0u;
            __syncwarp();

            const int nslab = (!skip && span > 0) ? (b_rel >> 9) + 1 : 0;                // through the byte behind the span
            auto load_seg = [&](int s) -> uint4 {
                const int pos = s * TS_SLAB + 16 * lane;
                if (pos >= b_rel) return make_uint4(0u, 0u, 0u, 0u);
                uint4 v = __ldg(reinterpret_cast<const uint4*>(base + pos));
                if (pos < a_rel || pos + 16 > b_rel) {
                    const int lo = a_rel - pos, hi = b_rel - pos;
                    v.x &= byte_range_mask(lo, hi);
                    v.y &= byte_range_mask(lo - 4, hi - 4);
                    v.z &= byte_range_mask(lo - 8, hi - 8);
                    v.w &= byte_range_mask(lo - 12, hi - 12);
                }
                return make_uint4(lower4(v.x), lower4(v.y), lower4(v.z), lower4(v.w));
            };

            unsigned q_head = 0u, q_ends = 0u;      // tokens taken / queued (warp-uniform)
            // every lane takes one waiting token
            auto batch = [&](int ntake) {
                const unsigned e = lane < ntake ? tq[(q_head + (unsigned)lane) & (TS_Q - 1)] : 0u;
                const int plen = (int)(e >> 16), ps = (int)(e & (unsigned)TS_POS_MASK);
                const int cell = (int)tbl[(e >> 11) & 31u];
                bool keep = plen >= min_len && ((valid_mask >> cell) & 1u);       // plen == 0: no token
                // the words that hold the token, from (start - prefix remainder) on: read once, for the stop-word filter
                // and for murmur3
                const int pr = (int)(pn & 3u), q0 = ps - pr, sh = (q0 & 3) * 8;
                const unsigned* rp = ring + ((q0 >> 2) & (TS_RING_WORDS - 1));
                const unsigned a0 = rp[0], a1 = rp[1], a2 = rp[2], a3 = rp[3], a4 = rp[4];
                const unsigned s0r = __funnelshift_r(a0, a1, sh), s1 = __funnelshift_r(a1, a2, sh);
                const unsigned s2 = __funnelshift_r(a2, a3, sh), s3 = __funnelshift_r(a3, a4, sh);
                if (keep) {
                    const unsigned first4 = __funnelshift_r(s0r, s1, 8 * pr);
                    const unsigned c0 = first4 & 0xffu;
                    if (c0 == '_') {                                 // a run of '_' only is not a token
                        keep = false;
                        for (int q = ps; q < ps + plen; ++q) keep |= ts_ring_byte(ring, q) != '_';
                    }
                    if (keep && plen <= STOP_MAXLEN && (c0 - 'a') < 26u) {
                        const unsigned c1 = plen > 1 ? (first4 >> 8) & 0xffu : 0u;
                        if ((stop_pre[stop_pre_index(c0, c1)] >> plen) & 1u) {
                            const unsigned b = stop_bloom_index(first4 & byte_range_mask(0, plen), (unsigned)plen);
                            if ((bloom[b >> 5] >> (b & 31u)) & 1u) keep = !ts_is_stop(ring, ps, plen);
                        }
                    }
                }
                if (need_lens) {
                    // the lanes of one cell add their lengths up first: one shared-memory atomic per cell and batch
                    const unsigned grp = __match_any_sync(FULL, keep ? cell : 32);
                    const unsigned sum = __reduce_add_sync(grp, keep ? (unsigned)plen : 0u);
                    if (keep && lane == __ffs(grp) - 1) atomicAdd(&lens[cell], sum);
                }
                if (keep && hashed) {
                    const unsigned hv = pr + plen <= 15 ? ts_hash_words(s0r, s1, s2, s3, plen, ph, pw, pn)
                                                        : ts_hash_token(ring, ps, plen, ph, pw, pn);
                    const int bucket = p.pow2_mask >= 0 ? (int)(hv & (unsigned)p.pow2_mask) : non_negative_mod(hv, F);
                    const unsigned bit = 1u << (bucket & 31);
                    const unsigned old = atomicOr(&bm[cell * TS_BM_STRIDE + (bucket >> 5)], bit);
                    if (STATS && two_classes && !(old & bit))        // fit: the bucket's first token in this cell
                        atomicAdd(&hist[((c1_mask >> cell) & 1u) * TC_MAXF + bucket], 1u);
                    if ((old & bit) && !p.binary_freq) {
                        const unsigned k = atomicAdd(&napp[cell], 1u);
                        if (k < (unsigned)TS_APP) app[k * 32 + cell] = (unsigned short)bucket;
                        else dflag[cell] = 1u;
                    }
                }
                q_head += (unsigned)ntake;
            };

            uint4 cur = nslab > 0 ? load_seg(0) : make_uint4(0u, 0u, 0u, 0u);
            unsigned p_prev = 0u;          // letter / digit flags of this lane's last byte in the slab before (bits 31, 30)
            unsigned carry_tok = 0u;       // the slab before ended inside a token
            unsigned open = 0u;            // the last token start so far: 1 << 31 | position << 5 | rank of its cell

            for (int s = 0;; ++s) {
                const bool more = s < nslab;
                uint4 nxt = make_uint4(0u, 0u, 0u, 0u);
                if (more) {
                    const int slab0 = s * TS_SLAB, pos = slab0 + 16 * lane;
                    if (s + 1 < nslab) nxt = load_seg(s + 1);
                    // the ring: slab s lives in words 128 (s & 3) ..; words 512 .. 519 repeat words 0 .. 7
                    {
                        const int slot = (s & 3) * 128 + lane * 4;
                        *reinterpret_cast<uint4*>(ring + slot) = cur;
                        if ((s & 3) == 0 && lane < 2) *reinterpret_cast<uint4*>(ring + TS_RING_WORDS + lane * 4) = cur;
                    }
                    // ---- cell borders inside this slab: bit i of bnd = a non-empty cell starts at byte i of this lane
                    bndw[lane] = 0u;
                    __syncwarp();
                    if (ne && (st >> 9) == s) atomicOr(&bndw[(st >> 4) & 31], 1u << (st & 15));
                    const int c_slab0 = __popc(__ballot_sync(FULL, ne && (st >> 9) < s));   // non-empty cells before
                    __syncwarp();
                    const unsigned bnd = bndw[lane];
                    // ---- classes of the 16 bytes (bit 7 of every byte)
                    const unsigned W[4] = {cur.x, cur.y, cur.z, cur.w};
                    unsigned fl[4], fd[4], tcw[4], rng[4], rng_any = 0u;
#pragma unroll
                    for (int q = 0; q < 4; ++q) {
                        const unsigned h7 = W[q] & 0x7f7f7f7fu, ascii = ~W[q] & 0x80808080u;
                        non_ascii |= W[q] & 0x80808080u;
                        fl[q] = (h7 + 0x1f1f1f1fu) & ~(h7 + 0x05050505u) & ascii;                 // 'a' .. 'z'
                        fd[q] = (h7 + 0x50505050u) & ~(h7 + 0x46464646u) & ascii;                 // '0' .. '9'
                        const unsigned ne_us = (h7 ^ 0x5f5f5f5fu) + 0x7f7f7f7fu;                   // bit 7 set iff not '_'
                        tcw[q] = (fl[q] | fd[q] | (~ne_us & ascii)) & 0x80808080u;
                        // the Mid* characters . ' : , ; all lie in 0x27 .. 0x3b
                        rng[q] = (h7 + 0x59595959u) & ~(h7 + 0x44444444u) & ~fd[q] & ascii & 0x80808080u;
                        rng_any |= rng[q];
                    }
                    unsigned m = ts_compress(tcw);
                    const unsigned p_own = (fl[3] & 0x80000000u) | ((fd[3] & 0x80000000u) >> 1);
                    if (__any_sync(FULL, rng_any != 0u)) {
                        // a Mid* character counts when the bytes on both sides are letters (MidLetter, MidNumLet, ')
                        // or digits (MidNum, MidNumLet, ') of the same cell: neighbours across the lanes by shuffles
                        unsigned left = __shfl_up_sync(FULL, p_own, 1);
                        const unsigned left0 = __shfl_sync(FULL, p_prev, 31);
                        if (lane == 0) left = left0;
                        const unsigned f_own = (fl[0] & 0x80u) | ((fd[0] & 0x80u) >> 1);
                        const unsigned n7 = nxt.x & 0x7fu, nasc = ~nxt.x & 0x80u;
                        const unsigned f_nxt =
                            (((n7 + 0x1fu) & ~(n7 + 0x05u)) & nasc) | ((((n7 + 0x50u) & ~(n7 + 0x46u)) & nasc) >> 1);
                        unsigned right = __shfl_down_sync(FULL, f_own, 1);
                        const unsigned right31 = __shfl_sync(FULL, f_nxt, 0);
                        if (lane == 31) right = right31;
                        const unsigned cl = left & 0x80000000u, cd = (left << 1) & 0x80000000u;
                        const unsigned nl = right & 0x80u, nd = (right << 1) & 0x80u;
                        unsigned mid[4], bl[4], bd[4], mid_any = 0u;
#pragma unroll
                        for (int q = 0; q < 4; ++q) {
                            const unsigned pl = q == 0 ? cl : fl[q - 1], pd = q == 0 ? cd : fd[q - 1];
                            const unsigned xl = q == 3 ? nl : fl[q + 1], xd = q == 3 ? nd : fd[q + 1];
                            bl[q] = __funnelshift_l(pl, fl[q], 8) & __funnelshift_r(fl[q], xl, 8);
                            bd[q] = __funnelshift_l(pd, fd[q], 8) & __funnelshift_r(fd[q], xd, 8);
                            mid[q] = rng[q] & (bl[q] | bd[q]);
                            mid_any |= mid[q];
                        }
                        if (__any_sync(FULL, mid_any != 0u)) {
                            if (mid_any) {
#pragma unroll
                                for (int q = 0; q < 4; ++q) {
                                    const unsigned h7 = W[q] & 0x7f7f7f7fu;
                                    const unsigned ne_dot = (h7 ^ 0x2e2e2e2eu) + 0x7f7f7f7fu;
                                    const unsigned ne_apo = (h7 ^ 0x27272727u) + 0x7f7f7f7fu;
                                    const unsigned ne_col = (h7 ^ 0x3a3a3a3au) + 0x7f7f7f7fu;
                                    const unsigned ne_com = (h7 ^ 0x2c2c2c2cu) + 0x7f7f7f7fu;
                                    const unsigned ne_sem = (h7 ^ 0x3b3b3b3bu) + 0x7f7f7f7fu;
                                    const unsigned mid_l = ~(ne_dot & ne_apo & ne_col);          // MidLetter | MidNumLet | '
                                    const unsigned mid_n = ~(ne_dot & ne_apo & ne_com & ne_sem); // MidNum | MidNumLet | '
                                    mid[q] &= (mid_l & bl[q]) | (mid_n & bd[q]);
                                }
                            }
                            // not across a cell border: neither the first byte of a cell nor its last
                            unsigned before = __shfl_down_sync(FULL, bnd, 1) & 1u;
                            const bool next_slab_border = __any_sync(FULL, ne && st == slab0 + TS_SLAB);
                            if (lane == 31) before = next_slab_border ? 1u : 0u;
                            m |= ts_compress(mid) & ~bnd & ~((bnd >> 1) | (before << 15));
                        }
                    }
                    p_prev = p_own;
                    // ---- token ends and starts: position i ends a token when byte i - 1 is a token byte and byte i
                    //      is none or begins a cell; a token starts where a token byte follows none or begins a cell
                    unsigned ct = (__shfl_up_sync(FULL, m, 1) >> 15) & 1u;
                    const unsigned last_tok = (__shfl_sync(FULL, m, 31) >> 15) & 1u;
                    if (lane == 0) ct = carry_tok;
                    carry_tok = last_tok;
                    const unsigned ext = (m << 1) | ct;
                    const unsigned endm = ext & (~m | bnd) & 0xffffu, startm = m & (~ext | bnd) & 0xffffu;
                    // one scan for two counts: ends | borders << 16
                    const unsigned cnt = (unsigned)__popc(endm) | ((unsigned)__popc(bnd) << 16);
                    unsigned incl = cnt;
#pragma unroll
                    for (int o = 1; o < 32; o <<= 1) {
                        const unsigned up = __shfl_up_sync(FULL, incl, o);
                        if (lane >= o) incl += up;
                    }
                    const unsigned total = __shfl_sync(FULL, incl, 31), excl = incl - cnt;
                    // rank of the cell of a start at bit b of this lane: r_base + borders up to and including b
                    const int r_base = c_slab0 + (int)(excl >> 16) - 1;
                    // the token an end closes began at the last start before it: in this lane, or the last start of
                    // the lanes below, or the one left open by the slabs before
                    unsigned own = 0u;
                    if (startm) {
                        const int sb = 31 - __clz(startm);
                        own = 0x80000000u | ((unsigned)(pos + sb) << 5) | (unsigned)(r_base + __popc(bnd & ((2u << sb) - 1u)));
                    }
                    const unsigned has_start = __ballot_sync(FULL, startm != 0u);
                    const unsigned below = has_start & lt;
                    const unsigned from_below = __shfl_sync(FULL, own, below ? 31 - __clz(below) : 0);
                    const unsigned carried = below ? from_below : open;
                    if (has_start) open = __shfl_sync(FULL, own, 31 - __clz(has_start));
                    {
                        unsigned em = endm, sm = startm;
                        unsigned ei = q_ends + (excl & 0xffffu);
                        // at most one end of a lane closes a token begun below the lane's bytes: the lowest, when no
                        // start lies below it
                        if (em) {
                            const int b = __ffs(em) - 1;
                            if (!(sm & ((1u << b) - 1u))) {
                                em &= em - 1u;
                                const int sp = (int)((carried & 0x7fffffffu) >> 5), plen = pos + b - sp;
                                const unsigned rk = carried & 31u;
                                unsigned entry = ((unsigned)sp & (unsigned)TS_POS_MASK) | (rk << 11) | ((unsigned)plen << 16);
                                if (plen > TC_BACK) {                // longer than the fast path reads back
                                    dflag[tbl[rk]] = 1u;
                                    entry = 0u;
                                }
                                tq[ei & (TS_Q - 1)] = entry;
                                ++ei;
                            }
                        }
                        // the others begin and end inside the lane's bytes: k-th start left with k-th end left
                        const unsigned posr = (unsigned)pos & (unsigned)TS_POS_MASK;
                        while (__any_sync(FULL, em != 0u)) {
                            if (em) {
                                const int b = __ffs(em) - 1, sb = __ffs(sm) - 1;
                                em &= em - 1u;
                                sm &= sm - 1u;
                                const unsigned rk = (unsigned)(r_base + __popc(bnd & ((2u << sb) - 1u)));
                                tq[ei & (TS_Q - 1)] = (posr | (unsigned)sb) | (rk << 11) | ((unsigned)(b - sb) << 16);
                                ++ei;
                            }
                        }
                    }
                    q_ends += total & 0xffffu;
                }
                __syncwarp();
                // ---- tokens: full batches; what is left waits for the next slab, but no longer
                bool ran = false;
                for (;;) {
                    const unsigned avail = q_ends - q_head;
                    const int ntake = avail >= 32u ? 32 : ((!more || !ran) ? (int)avail : 0);
                    if (ntake == 0) break;
                    batch(ntake);
                    ran = true;
                }
                if (!more) break;
                cur = nxt;
            }
            __syncwarp();

            // ---- end of the span: lane k is cell k again
            bool deferred = valid && ne && dflag[lane] != 0u;
            unsigned* mybm = bm + lane * TS_BM_STRIDE;
            unsigned len_sum = lens[lane];
            if (!need_lens) {
                unsigned any = 0u;
#pragma unroll
                for (int w = 0; w < TC_FW; ++w) any |= mybm[w];
                len_sum = any ? 1u : 0u;              // only "did a token survive" is asked
            }
            const bool nohit = len_sum == 0u;
            int ndup = 0;
            const int na = (int)min(napp[lane], (unsigned)TS_APP);
            if (__any_sync(FULL, na > 0)) {
                // the appended buckets collapse, in place, to (bucket | count << 9) of the distinct ones.  Every appended
                // bucket has its bit in the cell's bitmap: the first time it is met here the bit is cleared (and set
                // again below), so that only a bucket met three times or more has to be looked up in the list
                for (int q = 0; q < na; ++q) {
                    const unsigned b = app[q * 32 + lane];
                    const unsigned bit = 1u << (b & 31u);
                    const unsigned w = mybm[b >> 5];
                    if (w & bit) {
                        if (ndup < TC_DUPS) {
                            mybm[b >> 5] = w & ~bit;
                            app[ndup * 32 + lane] = (unsigned short)(b | (2u << 9));
                            ++ndup;
                        } else {
                            deferred = true;                         // more distinct repeated buckets than a record holds
                        }
                    } else {
                        for (int r = 0; r < ndup; ++r)
                            if ((app[r * 32 + lane] & 0x1ffu) == b) app[r * 32 + lane] += 0x200u;
                    }
                }
                for (int q = 0; q < ndup; ++q) {
                    const unsigned b = app[q * 32 + lane] & 0x1ffu;
                    mybm[b >> 5] |= 1u << (b & 31u);
                }
            }
            if (deferred && in_rows) {
                const int64_t cell = row * ncols + j;
                atomicOr(&defer_bits[cell >> 5], 1u << (cell & 31));
            }
            if (PACK) {
                // ---- the cells of this span in packed form (TsPack).  The ring and the queue are free now: the ring
                //      holds, per bitmap word of a cell, the bytes of the cell's stream before it; the queue stages the
                //      group's bytes
                unsigned* before = ring;              // [TC_FW][32]: bytes before word w's first entry byte
                int nbytes = 0, cnt = 0, n4 = 0;
                unsigned words = 0u;                  // the bitmap words that hold a bucket
                const bool plain = in_rows && hashed && !deferred;
                if (plain) {
                    int zrun = 0;                     // buckets skipped since the last one met
#pragma unroll
                    for (int w = 0; w < TC_FW; ++w) {
                        const unsigned x = mybm[w];
                        if (x) {
                            words |= 1u << w;
                            nbytes += (zrun + __ffs(x) - 1) / 63;     // "skip 63" bytes before the word's first entry
                            before[w * 32 + lane] = (unsigned)nbytes;
                            cnt += __popc(x);
                            nbytes += __popc(x);
                            zrun = __clz(x);
                        } else {
                            zrun += 32;
                        }
                    }
                    if (!p.binary_freq)
                        for (int q = 0; q < ndup; ++q) n4 += (app[q * 32 + lane] >> 9) >= 4u;
                    nbytes += n4;
                }
                if (in_rows && deferred) nbytes = 8;
                int incl = nbytes;
#pragma unroll
                for (int o = 1; o < 32; o <<= 1) {
                    const int up = __shfl_up_sync(FULL, incl, o);
                    if (lane >= o) incl += up;
                }
                const int excl = incl - nbytes, total = __shfl_sync(FULL, incl, 31), alloc = (total + 15) & ~15;
                unsigned long long off = 0ull;
                if (lane == 0 && total > 0) {
                    off = atomicAdd(pk.cursor, (unsigned long long)alloc);
                    // a cell's place inside its group has 14 bits: a group beyond that makes the chunk unusable
                    if (off + (unsigned long long)alloc > pk.arena_bytes || total > (int)TS_REL_MASK) {
                        pk.flags[0] = 1;
                        off = ~0ull;
                    }
                }
                off = __shfl_sync(FULL, off, 0);
                if (in_rows) {
                    pk.rel[row * ncols + j] = (unsigned short)(((unsigned)excl & TS_REL_MASK) | (nohit ? TS_REL_NOHIT : 0u) |
                                                               (deferred ? TS_REL_SLOT : 0u));
                    if (pk.lens) pk.lens[row * ncols + j] = need_lens ? len_sum : 0u;
                }
                if (lane == 0) {
                    TsGroup g;
                    g.off = off;
                    g.total = (unsigned)total;
                    g.pad = 0u;
                    pk.grp[(g0 >> 5) * ncols + j] = g;
                }
                if (total > 0 && off != ~0ull) {
                    const bool staged = alloc <= TS_Q * 4;
                    unsigned char* stage = reinterpret_cast<unsigned char*>(tq);
                    unsigned char* dst = (staged ? stage : pk.arena + off) + excl;
                    if (in_rows && deferred) {
                        for (int q = 0; q < 8; ++q) dst[q] = 0;                 // the slow kernel leaves the slot here
                    } else if (plain && cnt > 0) {
                        // one loop over the cell's buckets (not one per bitmap word: the lanes would wait for the
                        // fullest word of the 32 cells sixteen times)
                        int w = 0, prev = -1, k = 0;
                        unsigned x = 0u;
                        for (int i = 0; i < cnt; ++i) {
                            if (!x) {                                // on to the next word that holds a bucket
                                w = __ffs(words) - 1;
                                words &= words - 1u;
                                x = mybm[w];
                            }
                            const int bucket = w * 32 + __ffs(x) - 1;
                            int gap = bucket - prev - 1;
                            x &= x - 1u;
                            prev = bucket;
                            // one in twenty buckets is far (more than 62 skipped): a "skip 63" byte first -- written
                            // without a branch, the lanes would take it in most rounds otherwise
                            const bool far = gap > 62;
                            gap -= far ? 63 : 0;
                            if (gap > 62) {                          // further than 125: rare
                                dst[k++] = 0xFC;
                                for (; gap > 125; gap -= 63) dst[k++] = 0xFC;
                                gap -= 63;
                            }
                            dst[k] = far ? (unsigned char)0xFC : (unsigned char)(gap << 2);
                            if (far) dst[k + 1] = (unsigned char)(gap << 2);
                            k += far ? 2 : 1;
                        }
                        // term frequencies above 1: the code goes into the entry's byte, 4 and more to the trailer
                        if (!p.binary_freq)
                            for (int q = 0; q < ndup; ++q) {
                                const unsigned e = app[q * 32 + lane], bkt = e & 0x1ffu, tf = e >> 9;
                                const unsigned below = mybm[bkt >> 5] & ((1u << (bkt & 31u)) - 1u);
                                const int at = (int)before[(bkt >> 5) * 32 + lane] + __popc(below);
                                dst[at] = (unsigned char)(dst[at] | (tf <= 3u ? tf - 1u : 3u));
                                if (tf >= 4u) {
                                    int order = 0;                   // its place among the trailer's entries: by bucket
                                    for (int r = 0; r < ndup; ++r) {
                                        const unsigned e2 = app[r * 32 + lane];
                                        order += (e2 >> 9) >= 4u && (e2 & 0x1ffu) < bkt;
                                    }
                                    dst[nbytes - 1 - order] = (unsigned char)(tf - 4u);
                                }
                            }
                    }
                    if (staged) {
                        __syncwarp();
                        for (int i = lane * 16; i < alloc; i += 512)
                            *reinterpret_cast<uint4*>(pk.arena + off + i) = *reinterpret_cast<const uint4*>(stage + i);
                    }
                }
                __syncwarp();
            }
            if (CELLS && in_rows) {
                unsigned* rec = records + ((size_t)row * ncols + j) * TC_REC_WORDS;
                const unsigned head = deferred ? TC_H_DEFER : ((unsigned)ndup | (nohit ? TC_H_NOHIT : 0u));
                uint4* r4 = reinterpret_cast<uint4*>(rec);
#pragma unroll
                for (int q = 0; q < 4; ++q) r4[q] = make_uint4(mybm[4 * q], mybm[4 * q + 1], mybm[4 * q + 2], mybm[4 * q + 3]);
                unsigned t[16];
                t[0] = head;
                t[1] = p.len_base >= 0 ? len_sum : 0u;
                t[2] = 0u;
#pragma unroll
                for (int q = 0; q < TC_DUPS; ++q) {
                    const unsigned e = q < ndup ? app[q * 32 + lane] : 0u;
                    t[3 + q] = (e & 0x1ffu) | ((e >> 9) << 16);
                }
                t[15] = 0u;
#pragma unroll
                for (int q = 0; q < 4; ++q) r4[4 + q] = make_uint4(t[4 * q], t[4 * q + 1], t[4 * q + 2], t[4 * q + 3]);
            }
            if (STATS) {
                const bool counted = in_rows && !deferred;
                if (deferred) {
                    // what the cell's tokens added to the CTA's histogram while it was still on the fast path
                    for (int w = 0; w < TC_FW; ++w) {
                        unsigned x = mybm[w];
                        mybm[w] = 0u;
                        while (two_classes && x) {
                            const int bkt = w * 32 + __ffs(x) - 1;
                            x &= x - 1u;
                            atomicSub(&hist[(cls == 1 ? TC_MAXF : 0) + bkt], 1u);
                        }
                    }
                }
                const double e_row = (label && counted) ? label[row] - label_shift : 0.0;
                if (two_classes) {
                    const unsigned c1 = __ballot_sync(FULL, counted && cls == 1);
                    const unsigned nh = __ballot_sync(FULL, counted && nohit);
                    null0 += __popc(nh & ~c1);
                    null1 += __popc(nh & c1);
                    if (label && counted) {
                        for (int w = 0; w < TC_FW; ++w) {
                            unsigned x = mybm[w];
                            while (x) {
                                const int bkt = w * 32 + __ffs(x) - 1;
                                x &= x - 1u;
                                if (bkt < F) atomicAdd(&bucket_wsum[(int64_t)j * F + bkt], e_row);
                            }
                        }
                    }
                } else if (counted) {
                    for (int w = 0; w < TC_FW; ++w) {
                        unsigned x = mybm[w];
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
                    if (!p.binary_freq)
                        for (int q = 0; q < ndup; ++q) {
                            const unsigned e = app[q * 32 + lane];
                            const unsigned v = (unsigned)((int64_t)j * F) + (e & 0x1ffu), cnt = e >> 9;
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
            __syncthreads();
            if (hashed)
                for (int i = tid; i < 2 * TC_MAXF; i += TS_WARPS * 32) {
                    const int c = i / TC_MAXF, bkt = i % TC_MAXF;
                    const unsigned a = hist[i];
                    if (a && bkt < F) atomicAdd(&bucket_class_sum[((int64_t)j * F + bkt) * p.nclasses + c], a);
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

}  // namespace tmog
