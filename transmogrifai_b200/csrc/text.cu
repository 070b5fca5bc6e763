// Text path of SmartTextVectorizer / OPCollectionHashingVectorizer on raw utf8 columns (Arrow layout: int64 offsets +
// bytes + validity bitmap per column):
//
//   text_stats_kernel   TextStats of SmartTextVectorizer.fitFn: per row a 64-bit hash and the length of the CLEANED
//                       cell (TextUtils.cleanString applied on the fly), counted in a small open-addressing table per
//                       column (distinct values, per-label-class counts, first row of every value) -- enough to decide
//                       pivot / hash and to build the pivot vocabulary of low-cardinality text columns
//   text_code_kernel    dictionary codes of a low-cardinality text column from that table (feeds the pivot kernels)
//   text_hash_kernel    tokenise (ASCII restriction of Lucene's StandardTokenizer: UAX#29 word breaks) -> lower-case ->
//                       Snowball english stop filter -> min-length filter -> "<nameHash>_<token>" -> Spark's legacy
//                       murmur3_x86_32 (seed 42) -> non-negative mod numFeatures -> per-row term frequencies, plus the
//                       null indicators; either accumulated into per-bucket statistics for SanityChecker (fit) or
//                       written as sorted sparse rows (transform)
//
// The tokenise + hash kernel is instruction-bound, not HBM-bound; its design is described above the kernel and in
// DESIGN.md.
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <type_traits>
#include <vector>

#include "common.cuh"

namespace tmog {

// ------------------------------------------------------------------------------------------------------------------
// character classes (ASCII)
// ------------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ bool is_letter(unsigned c) { return ((c | 0x20u) - 'a') < 26u; }
__device__ __forceinline__ bool is_digit(unsigned c) { return (c - '0') < 10u; }
__device__ __forceinline__ bool is_punct(unsigned c) {  // java \p{Punct}: the 32 ASCII punctuation characters
    return (c - 0x21u) < 15u || (c - 0x3Au) < 7u || (c - 0x5Bu) < 6u || (c - 0x7Bu) < 4u;
}
__device__ __forceinline__ unsigned to_lower(unsigned c) { return (c - 'A') < 26u ? c + 32u : c; }

// bytes [lo, hi) of a word kept (lo, hi relative to the word, any range)
__device__ __forceinline__ unsigned byte_range_mask(int lo, int hi) {
    lo = max(lo, 0);
    hi = min(hi, 4);
    if (hi <= lo) return 0u;
    const unsigned upto_hi = hi == 4 ? 0xffffffffu : (1u << (8 * hi)) - 1u;
    return upto_hi & ~((1u << (8 * lo)) - 1u);
}

// ------------------------------------------------------------------------------------------------------------------
// Spark 2.4.5 mllib HashingTF: Murmur3_x86_32.hashUnsafeBytes (legacy tail handling), seed 42
// ------------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ unsigned rotl32(unsigned x, int r) { return __funnelshift_l(x, x, r); }
__device__ __forceinline__ unsigned mix_k1(unsigned k) {
    k *= 0xcc9e2d51u;
    k = rotl32(k, 15);
    return k * 0x1b873593u;
}
__device__ __forceinline__ unsigned mix_h1(unsigned h, unsigned k) {
    h ^= k;
    h = rotl32(h, 13);
    return h * 5u + 0xe6546b64u;
}
__device__ __forceinline__ unsigned fmix32(unsigned h, unsigned len) {
    h ^= len;
    h ^= h >> 16;
    h *= 0x85ebca6bu;
    h ^= h >> 13;
    h *= 0xc2b2ae35u;
    h ^= h >> 16;
    return h;
}
// streaming form: bytes are pushed one at a time; full little-endian words are mixed as they complete, the (< 4)
// bytes left at the end are mixed one by one, sign-extended
struct Murmur {
    unsigned h, word, n;
    __device__ __forceinline__ void init() { h = 42u; word = 0u; n = 0u; }
    __device__ __forceinline__ void push(unsigned b) {
        word |= b << (8 * (n & 3u));
        ++n;
        if ((n & 3u) == 0u) {
            h = mix_h1(h, mix_k1(word));
            word = 0u;
        }
    }
    __device__ __forceinline__ unsigned finish() {
        const unsigned tail = n & 3u;
        for (unsigned i = 0; i < tail; ++i) {
            const int sb = (int)(signed char)((word >> (8 * i)) & 0xffu);
            h = mix_h1(h, mix_k1((unsigned)sb));
        }
        return fmix32(h, n);
    }
};
__device__ __forceinline__ int non_negative_mod(unsigned h, int n) {
    const int r = (int)h % n;
    return r < 0 ? r + n : r;
}

// ------------------------------------------------------------------------------------------------------------------
// Snowball english stop words (Lucene EnglishAnalyzer / StandardAnalyzer(englishStopwords) as the reference loads
// them, LuceneTextAnalyzer.scala:159-161): open-addressing table in constant memory, filled by the host
// ------------------------------------------------------------------------------------------------------------------
constexpr int STOP_SLOTS = 512;
constexpr int STOP_MAXLEN = 12;
struct StopEntry {
    unsigned char len;
    char w[STOP_MAXLEN];
};
__constant__ StopEntry c_stop[STOP_SLOTS];
// stop-word pre-filter: entry (first letter, second character class) holds a bit per word length, so that the two-slot
// probe runs only for the few tokens that share first letter, second character and length with a stop word.
// Second character classes: 0..25 letters, 26 the apostrophe, 27 anything else / past the end of the token.
constexpr int STOP_PRE = 26 * 28;
__constant__ unsigned short c_stop_pre[STOP_PRE];
__host__ __device__ inline int stop_pre_index(unsigned c0, unsigned c1) {   // c0 must be 'a' .. 'z'
    const unsigned k1 = (c1 - 'a') < 26u ? c1 - 'a' : (c1 == '\'' ? 26u : 27u);
    return (int)((c0 - 'a') * 28u + k1);
}
// second filter (text_span_kernel): one bit per (first four bytes, length) of a stop word, 4096 bits
constexpr int STOP_BLOOM_WORDS = 128;
__constant__ unsigned c_stop_bloom[STOP_BLOOM_WORDS];
__host__ __device__ inline unsigned stop_bloom_index(unsigned first4, unsigned len) {   // first4: bytes past len are 0
    return ((first4 * 0x9E3779B1u) ^ (len * 0x85EBCA6Bu)) >> 20;
}

// ------------------------------------------------------------------------------------------------------------------
// column descriptors
// ------------------------------------------------------------------------------------------------------------------
struct TextColumn {             // col_desc: 4 x int64 per column (device)
    const uint8_t* data;
    const int64_t* offsets;     // [n + 1]
    const uint64_t* valid;      // bitmap or NULL
    int64_t reserved;
};

__device__ __forceinline__ bool cell_valid(const TextColumn& tc, int64_t r) {
    return tc.valid == nullptr || ((tc.valid[r >> 6] >> (r & 63)) & 1ull);
}

// ------------------------------------------------------------------------------------------------------------------
// TextStats: hash + length of the cleaned cell
// ------------------------------------------------------------------------------------------------------------------
// cleanString == cut the lower-cased string at runs of {space, ASCII punctuation}, capitalise each piece, concatenate
// (see strings_host.cu).  Returns the 64-bit FNV-1a hash of the cleaned bytes and their number.
__device__ __forceinline__ void clean_hash_cell(const uint8_t* s, int64_t len, bool clean, unsigned long long& hash,
                                                int& out_len, bool& non_ascii) {
    unsigned long long h = 1469598103934665603ull;
    int m = 0;
    bool start = true;
    for (int64_t i = 0; i < len; ++i) {
        unsigned c = s[i];
        if (c >= 0x80u) non_ascii = true;
        if (clean) {
            if (c == ' ' || is_punct(c)) {
                start = true;
                continue;
            }
            c = to_lower(c);
            if (start && (c - 'a') < 26u) c -= 32u;
            start = false;
        }
        h = (h ^ c) * 1099511628211ull;
        ++m;
    }
    // avalanche so that the table index bits depend on every byte
    h ^= h >> 32;
    h *= 0xd6e8feb86659fd93ull;
    h ^= h >> 32;
    hash = h;
    out_len = m;
}

// Length of the cleaned cell only (what TextStats still needs once the column is past the cardinality cap): the bytes
// that are neither ' ' nor punctuation, counted 16 bytes per load, four bytes per step
__device__ __forceinline__ int clean_length_cell(const uint8_t* s, int64_t len, bool clean, bool& non_ascii) {
    if (!clean) {
        // no cleaning: every byte counts; only the non-ASCII check needs the bytes, and it is not needed either
        return (int)len;
    }
    const int sh = (int)(reinterpret_cast<uintptr_t>(s) & 15u);
    const uint8_t* base = s - sh;
    const int64_t nchunk = (sh + len + 15) >> 4;
    int kept = 0;
    unsigned na = 0u;
    for (int64_t c = 0; c < nchunk; ++c) {
        const uint4 v = __ldg(reinterpret_cast<const uint4*>(base + 16 * c));
        const int lo = c == 0 ? sh : 0;
        const int hi = (int)min((int64_t)16, sh + len - 16 * c);
        const unsigned w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            unsigned in = 0x80808080u;                       // bit 7 of the bytes that belong to the cell
            if (lo != 0 || hi != 16) in &= byte_range_mask(lo - 4 * q, hi - 4 * q);
            const unsigned h7 = w[q] & 0x7f7f7f7fu;
            // bit 7 of (h7 + (0x80 - x) * 0x01010101) is set iff the byte is >= x
            const unsigned drop = ((h7 + 0x60606060u) & ~(h7 + 0x50505050u)) |      // ' ' and 0x21 .. 0x2f
                                  ((h7 + 0x46464646u) & ~(h7 + 0x3f3f3f3fu)) |      // ':' .. '@'
                                  ((h7 + 0x25252525u) & ~(h7 + 0x1f1f1f1fu)) |      // '[' .. '`'
                                  ((h7 + 0x05050505u) & ~(h7 + 0x01010101u));       // '{' .. '~'
            na |= w[q] & in;
            kept += __popc(in & ~(drop & ~w[q]));
        }
    }
    if (na) non_ascii = true;
    return kept;
}

constexpr unsigned long long TS_EMPTY = 0xFFFFFFFFFFFFFFFFull;

// per column: table of `slots` entries {hash, first_row}, counts [slots][nclasses] u32, ndistinct, overflow flag
struct TextStatsTable {
    unsigned long long* hash;        // [slots]
    unsigned long long* first_row;   // [slots]
    unsigned int* counts;            // [slots * nclasses]
    int* ndistinct;                  // [1]
    unsigned long long* len_sum;     // [2] sum of lengths, sum of squares (over non-null cells)
    unsigned long long* nonnull;     // [1]
};

__device__ __forceinline__ int ts_insert(const TextStatsTable& t, int slots, int max_distinct, unsigned long long h,
                                         int64_t row, int cls, int nclasses) {
    // once a column is known to exceed the cardinality cap nothing more is recorded for it: only the fact matters
    // (SmartTextVectorizer then hashes the column; TextStats.additionHelper stops counting there as well)
    if (*t.ndistinct > max_distinct) return -1;
    if (h == TS_EMPTY) h = 0x1234567ull;
    unsigned s = (unsigned)(h >> 20) & (unsigned)(slots - 1);
    for (int probe = 0; probe < slots; ++probe) {
        unsigned long long cur = t.hash[s];
        if (cur == TS_EMPTY) {
            if (*t.ndistinct > max_distinct) return -1;
            const unsigned long long old = atomicCAS(&t.hash[s], TS_EMPTY, h);
            if (old == TS_EMPTY) {
                atomicAdd(t.ndistinct, 1);
                cur = h;
            } else {
                cur = old;
            }
        }
        if (cur == h) {
            atomicMin(&t.first_row[s], (unsigned long long)row);
            atomicAdd(&t.counts[(int64_t)s * nclasses + cls], 1u);
            return (int)s;
        }
        s = (s + 1) & (unsigned)(slots - 1);
    }
    return -1;
}

// one thread per cell (cells are short strings; consecutive threads take consecutive rows of one column)
__global__ void __launch_bounds__(256)
text_stats_kernel(const TextColumn* __restrict__ cols, int ncols, int64_t r_begin, int64_t n,
                  const uint8_t* __restrict__ class_ids, int nclasses, int clean, int want_lengths, int slots,
                  int max_distinct, const TextStatsTable* __restrict__ tables, int* __restrict__ flags) {
    const int c = blockIdx.y;
    const TextColumn tc = cols[c];
    const TextStatsTable t = tables[c];
    unsigned long long lsum = 0, lsq = 0, nn = 0;
    bool non_ascii = false;
    for (int64_t r = r_begin + (int64_t)blockIdx.x * blockDim.x + threadIdx.x; r < n;
         r += (int64_t)gridDim.x * blockDim.x) {
        // Past the cardinality cap nothing more is recorded for the column but the lengths, and those only matter to
        // a positive minLengthStdDev (SmartTextVectorizer.scala:113).  Without it the scan of the column ends here:
        // whoever vectorizes the column reads every cell anyway and reports non-ASCII bytes then
        // (tmog_text_hash_*_dev, tmog_text_codes_dev).
        const bool capped = *t.ndistinct > max_distinct;
        if (capped && !want_lengths) break;
        if (!cell_valid(tc, r)) continue;
        const int64_t a = tc.offsets[r], b = tc.offsets[r + 1];
        int m;
        if (capped) {
            m = clean_length_cell(tc.data + a, b - a, clean != 0, non_ascii);
        } else {
            unsigned long long h;
            clean_hash_cell(tc.data + a, b - a, clean != 0, h, m, non_ascii);
            ts_insert(t, slots, max_distinct, h, r, nclasses > 1 ? class_ids[r] : 0, nclasses);
        }
        lsum += (unsigned long long)m;
        lsq += (unsigned long long)m * (unsigned long long)m;
        ++nn;
    }
    // block-level reduction of the length sums (three atomics per warp)
    for (int o = 16; o > 0; o >>= 1) {
        lsum += __shfl_xor_sync(FULL, lsum, o);
        lsq += __shfl_xor_sync(FULL, lsq, o);
        nn += __shfl_xor_sync(FULL, nn, o);
    }
    if ((threadIdx.x & 31) == 0 && nn) {
        atomicAdd(&t.len_sum[0], lsum);
        atomicAdd(&t.len_sum[1], lsq);
        atomicAdd(t.nonnull, nn);
    }
    if (non_ascii && clean) flags[0] = 1;
}

// codes[c][r] = slot of the row's cleaned value in the column's table (-1 for a null; -2 if the value is not in the
// table, which cannot happen for a column whose cardinality stayed under the cap)
__global__ void __launch_bounds__(256)
text_code_kernel(const TextColumn* __restrict__ cols, const int* __restrict__ col_index, int nsel, int64_t n, int clean,
                 int slots, const TextStatsTable* __restrict__ tables, int32_t* __restrict__ codes, int64_t ld,
                 int* __restrict__ flags) {
    const int k = blockIdx.y;
    const int c = col_index[k];
    const TextColumn tc = cols[c];
    const TextStatsTable t = tables[c];
    for (int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; r < n; r += (int64_t)gridDim.x * blockDim.x) {
        int code = -1;
        if (cell_valid(tc, r)) {
            const int64_t a = tc.offsets[r], b = tc.offsets[r + 1];
            unsigned long long h;
            int m;
            bool na = false;
            clean_hash_cell(tc.data + a, b - a, clean != 0, h, m, na);
            if (na && clean && flags) flags[0] = 1;
            if (h == TS_EMPTY) h = 0x1234567ull;
            code = -2;
            unsigned s = (unsigned)(h >> 20) & (unsigned)(slots - 1);
            for (int probe = 0; probe < slots; ++probe) {
                const unsigned long long cur = t.hash[s];
                if (cur == h) { code = (int)s; break; }
                if (cur == TS_EMPTY) break;
                s = (s + 1) & (unsigned)(slots - 1);
            }
        }
        codes[(int64_t)k * ld + r] = code;
    }
}

// ------------------------------------------------------------------------------------------------------------------
// tokenise + hash
// ------------------------------------------------------------------------------------------------------------------
// One warp owns one row at a time (rows are handed out through a global counter).  The work on a row is split into
// phases that each keep all 32 lanes busy:
//
//   stage     the bytes of a cell are copied to the warp's text buffer in shared memory with 16-byte loads (lower-cased
//             four bytes at a time, bytes outside the cell zeroed); cells are packed back to back on 32-byte boundaries
//             with the same alignment (mod 32) as in global memory
//   classify  one lane per byte: a byte belongs to a token if it is a letter, a digit, '_' or a Mid* character between
//             two letters / two digits (UAX#29 WB6/7, WB11/12 for ASCII -- the decision only needs the two
//             neighbours); __ballot_sync gives 32 bits of the token mask, token starts (a set bit after a clear one)
//             are compacted into a list
//   hash      one lane per TOKEN of the list: its end is found in the mask, the stop-word probe and murmur3 (continued
//             from the state after the "<nameHash>_" prefix of its column) read the staged bytes as unaligned words,
//             and the bucket goes into the row's open-addressing table (key | count in one slot) in shared memory
//   flush     at the end of the row the table is emitted: per-bucket statistics with global atomics (fit) or compacted,
//             bitonic-sorted and written as a sparse row (transform)
//
// Rows that do not fit the table of the fast variant (too many distinct buckets, or a count that overflows the packed
// slot) are put on a list and redone by a second launch whose table is 64-bit slots in nearly all of shared memory.
constexpr int TH_PREFIX = 16;        // bytes reserved for "<nameHash>_"
constexpr int TH_COLCACHE = 64;      // columns whose murmur3 prefix state is cached in shared memory
constexpr int TH_GRAB = 4;           // rows taken from the counter at once
constexpr int TH_MAX_TOKEN = 255;    // StandardAnalyzer.DEFAULT_MAX_TOKEN_LENGTH
constexpr int TH_STOP_PRE_BYTES = (26 * 28 * 2 + 15) / 16 * 16;
// shared memory in front of the per-warp regions: stop-word table, cached prefix states, stop-word pre-filter
constexpr int TH_FIXED_SMEM = 512 * 16 + TH_COLCACHE * 12 + TH_STOP_PRE_BYTES;

struct TextHashParams {
    int num_features;      // hash buckets per space
    int shared;            // 1: one space for all columns, 0: column j owns [j * num_features, (j+1) * num_features)
    int binary_freq;
    int min_token_length;
    int track_nulls;
    int prepend;           // prefix bytes are used
    int hash_width;        // total width of the hash block (num_features or ncols * num_features)
    int null_base;         // index of the first null-indicator column (= hash_width + length columns)
    int nclasses;
    int pow2_mask;         // num_features - 1 if it is a power of two, else -1
    int count_bits;        // low bits of a table slot that hold the count (the key sits above)
    int max_token_bound;   // rows with more possible tokens (bytes / 2) than this go to the next level right away
};

struct StopSlot {          // stop-word table in shared memory: length + the word packed little-endian in 3 words
    unsigned len, w0, w1, w2;
};

__constant__ StopSlot c_stop_packed[STOP_SLOTS];      // the same table, words packed (read by the per-cell kernels)

// two slot choices per word (cuckoo hashing): a probe is two loads, no loop
__host__ __device__ inline unsigned stop_slot_hash(unsigned w0, unsigned w1, unsigned w2) {
    return ((w0 * 0x9E3779B1u) ^ (w1 * 0x85EBCA6Bu) ^ (w2 * 0xC2B2AE35u)) >> 23;   // 9 bits: STOP_SLOTS == 512
}
__host__ __device__ inline unsigned stop_slot_hash2(unsigned w0, unsigned w1, unsigned w2) {
    return ((w0 * 0x85EBCA6Bu) ^ (w1 * 0xC2B2AE35u) ^ (w2 * 0x27D4EB2Fu) ^ 0x165667B1u) >> 23;
}

template <typename Slot> struct SlotTraits;
template <> struct SlotTraits<unsigned> { static constexpr unsigned EMPTY = 0xffffffffu; };
template <> struct SlotTraits<unsigned long long> { static constexpr unsigned long long EMPTY = ~0ull; };

// lower-case four ASCII bytes at once
__device__ __forceinline__ unsigned lower4(unsigned w) {
    const unsigned h = w & 0x7f7f7f7fu;
    const unsigned ge_a = h + 0x3f3f3f3fu;      // bit 7 of a byte set iff it is >= 'A'
    const unsigned gt_z = h + 0x25252525u;      // bit 7 set iff > 'Z'
    return w | ((ge_a & ~gt_z & ~w & 0x80808080u) >> 2);
}

constexpr int th_pow2_floor(int x) {
    int p = 1;
    while (2 * p <= x) p *= 2;
    return p;
}

// bytes of one warp's working set in shared memory
template <typename Slot, int T, int TB>
struct ThLayout {
    using Tok = typename std::conditional<(TB <= 2048), unsigned short, unsigned>::type;
    static constexpr int BUF = 0;                               // 16 slack + TB text + 32 slack
    static constexpr int TOKS = BUF + 16 + TB + 32;             // TB / 2 token records
    static constexpr int MASK = TOKS + TB / 2 * (int)sizeof(Tok);            // TB / 32 words: bit b of word w = byte 32 w + b
    static constexpr int WCELL = MASK + (TB / 32 * 4 + 15) / 16 * 16;        // cell of every 16 bytes (u8)
    static constexpr int TABLE = WCELL + (TB / 16 + 15) / 16 * 16;
    // text + token list double as the ordered scatter table of the sparse-row flush
    static constexpr int T2 = th_pow2_floor((MASK - BUF) / (int)sizeof(Slot));
    static constexpr int BYTES = TABLE + T * (int)sizeof(Slot);
};

// stop-word probe of the staged (lower-cased) token [ps, ps + plen), plen <= STOP_MAXLEN
__device__ __forceinline__ bool th_is_stop(const StopSlot* stop, const unsigned* bufw, int ps, int plen) {
    const int q0 = ps + 16, sh = (q0 & 3) * 8;
    const unsigned* bw = bufw + (q0 >> 2);
    const unsigned a = bw[0], b = bw[1], c = bw[2], d = bw[3];
    const unsigned l0 = __funnelshift_r(a, b, sh) & byte_range_mask(0, plen);
    const unsigned l1 = __funnelshift_r(b, c, sh) & byte_range_mask(0, plen - 4);
    const unsigned l2 = __funnelshift_r(c, d, sh) & byte_range_mask(0, plen - 8);
    const uint4 e1 = reinterpret_cast<const uint4*>(stop)[stop_slot_hash(l0, l1, l2)];
    const uint4 e2 = reinterpret_cast<const uint4*>(stop)[stop_slot_hash2(l0, l1, l2)];
    const unsigned d1 = (e1.x ^ (unsigned)plen) | (e1.y ^ l0) | (e1.z ^ l1) | (e1.w ^ l2);
    const unsigned d2 = (e2.x ^ (unsigned)plen) | (e2.y ^ l0) | (e2.z ^ l1) | (e2.w ^ l2);
    return d1 == 0u || d2 == 0u;
}

// murmur3 of "<prefix><token>" continued from the state (ph, pw, pn) after the prefix: the r = pn & 3 bytes the prefix
// left pending are the low bytes of the first word; the token is read from the staged text as unaligned words
__device__ __forceinline__ unsigned th_hash_piece(const unsigned* bufw, int ps, int plen, unsigned ph, unsigned pw,
                                                  unsigned pn) {
    const int r = (int)(pn & 3u);
    const unsigned keep_hi = ~((1u << (8 * r)) - 1u);
    const int tot = r + plen, nfull = tot >> 2, tail = tot & 3;
    const int q0 = ps - r + 16, sh = (q0 & 3) * 8;
    const unsigned* bw = bufw + (q0 >> 2);
    unsigned h = ph, nxt = bw[0];
    for (int q = 0; q < nfull; ++q) {
        const unsigned cur = nxt;
        nxt = bw[q + 1];
        unsigned wd = __funnelshift_r(cur, nxt, sh);
        if (q == 0) wd = (wd & keep_hi) | pw;
        h = mix_h1(h, mix_k1(wd));
    }
    unsigned wd = __funnelshift_r(nxt, bw[nfull + 1], sh);
    if (nfull == 0) wd = (wd & keep_hi) | pw;
    // the (< 4) bytes left are mixed one by one, sign-extended (Spark's legacy hashUnsafeBytes)
    if (tail > 0) h = mix_h1(h, mix_k1((unsigned)(int)(signed char)(wd & 0xffu)));
    if (tail > 1) h = mix_h1(h, mix_k1((unsigned)(int)(signed char)((wd >> 8) & 0xffu)));
    if (tail > 2) h = mix_h1(h, mix_k1((unsigned)(int)(signed char)((wd >> 16) & 0xffu)));
    return fmix32(h, pn + (unsigned)plen);
}

__device__ __forceinline__ void th_col_state(const unsigned* colstate, const uint8_t* prefix, const int* prefix_len,
                                             const TextHashParams& p, int j, unsigned& ph, unsigned& pw, unsigned& pn) {
    if (j < TH_COLCACHE) {
        ph = colstate[3 * j];
        pw = colstate[3 * j + 1];
        pn = colstate[3 * j + 2];
    } else {
        Murmur pm;
        pm.init();
        const int plen = p.prepend ? prefix_len[j] : 0;
        for (int q = 0; q < plen; ++q) pm.push(prefix[j * TH_PREFIX + q]);
        ph = pm.h;
        pw = pm.word;
        pn = pm.n;
    }
}

// a bucket met cnt > 1 times in a row: everything beyond "present once" goes to the auxiliary arrays
__device__ __forceinline__ void th_rare_update(uint2 e, unsigned* bucket_dup, unsigned long long* bucket_sq_extra,
                                               unsigned* bucket_aux) {
    const unsigned v = e.x, cnt = e.y;
    atomicAdd(&bucket_dup[v], cnt - 1u);
    atomicAdd(&bucket_sq_extra[v], (unsigned long long)cnt * cnt - cnt);
    atomicMax(&bucket_aux[3 * (int64_t)v], cnt);
    atomicMax(&bucket_aux[3 * (int64_t)v + 1], ~cnt);      // the largest complement = the smallest count
    atomicAdd(&bucket_aux[3 * (int64_t)v + 2], 1u);
}

// count `key` in the row's table (1 << tbits slots in use, the count in the low cb bits of a slot)
template <typename Slot>
__device__ __forceinline__ void th_insert(Slot* table, int tbits, int cb, unsigned key, bool& newd, bool& ovf) {
    constexpr Slot EMPTY = SlotTraits<Slot>::EMPTY;
    const Slot cmask = (Slot(1) << cb) - Slot(1);
    const unsigned tmask = (1u << tbits) - 1u;
    unsigned slot = (key * 0x9E3779B1u) >> (32 - tbits);
    // double hashing: an odd step visits every slot of the power-of-two table, and keys that collide on their first
    // slot part ways afterwards -- at a load of 7/8 linear probing would build long runs that a few lanes of the warp
    // walk while the others wait
    const unsigned step = ((key * 0x85EBCA6Bu) >> (32 - tbits)) | 1u;
    for (unsigned probe = 0; probe <= tmask; ++probe) {
        Slot cur = table[slot];
        if (cur == EMPTY) {
            cur = atomicCAS(&table[slot], EMPTY, (Slot(key) << cb) | Slot(1));
            if (cur == EMPTY) {
                newd = true;
                return;
            }
        }
        if ((unsigned)(cur >> cb) == key) {
            const Slot old = atomicAdd(&table[slot], Slot(1));
            if ((old & cmask) == cmask) ovf = true;       // the count ran into the key
            return;
        }
        slot = (slot + step) & tmask;
    }
    ovf = true;                                           // table full
}

// flags: [0] non-ASCII byte met, [1] a run of token characters longer than the text buffer, [2] a row has more
//        distinct buckets than the large table holds, [3] sparse output capacity exceeded
template <typename Slot, int T, int TB, int WARPS, bool SPARSE_OUT>
__global__ void __launch_bounds__(WARPS * 32, WARPS >= 8 ? 2 : 1)
text_hash_kernel(const TextColumn* __restrict__ cols, int ncols, int64_t n, const uint8_t* __restrict__ prefix,
                 const int* __restrict__ prefix_len, TextHashParams p, const uint8_t* __restrict__ class_ids,
                 // rows to work on: all of [0, n), or row_list[0 .. *row_list_n) (rows deferred by the fast variant)
                 const int32_t* __restrict__ row_list, const unsigned* __restrict__ row_list_n,
                 unsigned long long* __restrict__ row_counter,
                 // rows this launch cannot hold are appended here (NULL: set flags[2] instead)
                 int32_t* __restrict__ defer_list, unsigned* __restrict__ defer_n,
                 // statistics mode
                 unsigned int* __restrict__ bucket_class_sum, unsigned int* __restrict__ bucket_dup,
                 unsigned long long* __restrict__ bucket_sq_extra, unsigned int* __restrict__ bucket_aux,
                 // continuous label (or NULL): bucket_wsum [index] += tf * (label[row] - label_shift)
                 const double* __restrict__ label, double label_shift, double* __restrict__ bucket_wsum,
                 // sparse-output mode
                 const int32_t* __restrict__ newpos, int64_t* __restrict__ row_start, int32_t* __restrict__ row_nnz,
                 int32_t* __restrict__ out_idx, int32_t* __restrict__ out_val, unsigned long long* __restrict__ cursor,
                 int64_t capacity, int* __restrict__ flags) {
    using L = ThLayout<Slot, T, TB>;
    constexpr Slot EMPTY = SlotTraits<Slot>::EMPTY;
    extern __shared__ __align__(16) unsigned char th_smem[];
    StopSlot* stop = reinterpret_cast<StopSlot*>(th_smem);                                   // [STOP_SLOTS]
    unsigned* colstate = reinterpret_cast<unsigned*>(th_smem + STOP_SLOTS * sizeof(StopSlot));  // [TH_COLCACHE][3]
    unsigned short* stop_pre = reinterpret_cast<unsigned short*>(th_smem + TH_FIXED_SMEM - TH_STOP_PRE_BYTES);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    static_assert(L::BYTES % 16 == 0 && (TH_COLCACHE * 12) % 16 == 0, "per-warp regions must stay 16-byte aligned");
    unsigned char* wbase = th_smem + TH_FIXED_SMEM + (size_t)warp * L::BYTES;
    unsigned char* buf = wbase + L::BUF + 16;                      // text; buf[-16 ..) and (.. TB + 32) are slack
    const unsigned* bufw = reinterpret_cast<const unsigned*>(wbase + L::BUF);
    unsigned* mask = reinterpret_cast<unsigned*>(wbase + L::MASK);
    using Tok = typename L::Tok;
    constexpr int SB = TB <= 2048 ? 11 : 12;            // bits of a buffer position in a token record
    static_assert(TB == 1024 || TB == 2048 || TB == 4096, "token records are packed for these buffer sizes");
    Tok* toks = reinterpret_cast<Tok*>(wbase + L::TOKS);
    unsigned char* wcell = wbase + L::WCELL;
    Slot* table = reinterpret_cast<Slot*>(wbase + L::TABLE);

    for (int i = tid; i < STOP_SLOTS; i += WARPS * 32) {
        const StopEntry e = c_stop[i];
        StopSlot sl = {e.len, 0u, 0u, 0u};
        for (int q = 0; q < e.len; ++q) {
            const unsigned c = (unsigned char)e.w[q];
            if (q < 4) sl.w0 |= c << (8 * q);
            else if (q < 8) sl.w1 |= c << (8 * (q - 4));
            else sl.w2 |= c << (8 * (q - 8));
        }
        stop[i] = sl;
    }
    for (int i = tid; i < STOP_PRE; i += WARPS * 32) stop_pre[i] = c_stop_pre[i];
    for (int j = tid; j < min(ncols, TH_COLCACHE); j += WARPS * 32) {
        Murmur pm;
        pm.init();
        const int plen = p.prepend ? prefix_len[j] : 0;
        for (int q = 0; q < plen; ++q) pm.push(prefix[j * TH_PREFIX + q]);
        colstate[3 * j] = pm.h;
        colstate[3 * j + 1] = pm.word;
        colstate[3 * j + 2] = pm.n;
    }
    for (int i = lane; i < T; i += 32) table[i] = EMPTY;
    for (int i = lane; i < (L::TABLE - L::BUF) / 4; i += 32) reinterpret_cast<unsigned*>(wbase)[i] = 0u;
    __syncthreads();

    const int cb = p.count_bits;
    const Slot cmask = (Slot(1) << cb) - Slot(1);
    const unsigned lt_mask = (1u << lane) - 1u;
    const int64_t nrows = row_list ? (int64_t)*row_list_n : n;

    for (;;) {
        unsigned long long grab = 0;
        if (lane == 0) grab = atomicAdd(row_counter, (unsigned long long)TH_GRAB);
        grab = __shfl_sync(FULL, grab, 0);
        if ((int64_t)grab >= nrows) break;
        const int64_t grab_end = min((int64_t)grab + TH_GRAB, nrows);
        for (int64_t ri = (int64_t)grab; ri < grab_end; ++ri) {
            const int64_t row = row_list ? (int64_t)row_list[ri] : ri;

            // ---- size of the table for this row: a power of two >= 2 x (upper bound of the hits), at most T
            int64_t bytes_bound = 0;
            for (int c0 = 0; c0 < ncols; c0 += 32) {
                const int j = c0 + lane;
                if (j < ncols) {
                    const TextColumn tc = cols[j];
                    if (cell_valid(tc, row)) bytes_bound += (tc.offsets[row + 1] - tc.offsets[row] + 1) >> 1;
                }
            }
            for (int o = 16; o > 0; o >>= 1) bytes_bound += __shfl_xor_sync(FULL, bytes_bound, o);
            bytes_bound += ncols;
            if (defer_list && bytes_bound > p.max_token_bound) {      // too long for this level's table
                if (lane == 0) defer_list[atomicAdd(defer_n, 1u)] = (int32_t)row;
                continue;
            }
            int tbits = 6;
            while (tbits < 31 && (1ll << tbits) < 2 * bytes_bound && (1 << tbits) < T) ++tbits;
            const int tsize = 1 << tbits;
            // the table never gets fuller than 7/8 (+ 32): linear probing stays short while it fills (about four
            // probes per insert on average up to that load), and rows of ~850 distinct buckets still fit the
            // 1024-slot table of the level that runs 24 warps per SM
            const int dist_limit = tsize - (tsize >> 3);
            int ndist = 0;
            bool row_bad = false;

            for (int c0 = 0; c0 < ncols && !row_bad; c0 += 32) {
                // ---- lane k holds cell k of this pass
                const int ncells = min(32, ncols - c0);
                const uint8_t* my_ptr = nullptr;
                int64_t my_len = 0;
                if (lane < ncells) {
                    const TextColumn tc = cols[c0 + lane];
                    if (cell_valid(tc, row)) {
                        const int64_t a = tc.offsets[row];
                        my_len = tc.offsets[row + 1] - a;
                        my_ptr = tc.data + a;
                    }
                }
                unsigned emitted = 0u;       // bit k: a token of cell k survived (per lane, OR-reduced below)
                int k = 0;
                int64_t off = 0;             // bytes of cell k already consumed (cells longer than the buffer)
                for (;;) {
                    // ---- stage as many cells as the buffer holds: 16-byte chunks, global and shared addresses
                    //      congruent mod 16; every byte of [0, fill) that is not a cell byte is zero.  The chunks of all
                    //      the cells of the batch are numbered through and taken one per lane.
                    int fill = 0;
                    int piece_pos0 = -1, piece_cnt = 0;      // set when the buffer holds one piece of a long cell
                    {
                        const uint8_t* cptr = my_ptr + (lane == k ? off : 0);
                        const int64_t crem = my_len - (lane == k ? off : 0);
                        const int csh = (int)(reinterpret_cast<uintptr_t>(cptr) & 15u);
                        const bool mine = lane >= k && lane < ncells && crem > 0;
                        // chunks through the zero byte after the cell (a cell longer than the buffer: more than fit)
                        const int need = !mine ? 0 : (crem > TB ? TB : (int)((csh + crem + 16) >> 4));
                        int incl = need;
                        for (int o = 1; o < 32; o <<= 1) {
                            const int up = __shfl_up_sync(FULL, incl, o);
                            if (lane >= o) incl += up;
                        }
                        const int cstart = incl - need;          // first chunk of this lane's cell
                        // incl is monotone, so the cells that fit are a run starting at k
                        const int kend = min(ncells, k + __popc(__ballot_sync(FULL, lane >= k && (incl << 4) <= TB)));
                        unsigned na = 0u;
                        if (kend > k) {
                            const int nchunks = __shfl_sync(FULL, incl, kend - 1);
                            for (int f0 = 0; f0 < nchunks; f0 += 32) {
                                const int f = min(f0 + lane, nchunks - 1);
                                int lo_c = k, hi_c = kend - 1;       // the last cell whose first chunk is <= f
#pragma unroll
                                for (int it = 0; it < 5; ++it) {
                                    const int mid = (lo_c + hi_c + 1) >> 1;
                                    const int cs = __shfl_sync(FULL, cstart, mid);
                                    if (cs <= f) lo_c = mid;
                                    else hi_c = mid - 1;
                                }
                                const int c = f - __shfl_sync(FULL, cstart, lo_c);
                                const uint8_t* g = reinterpret_cast<const uint8_t*>(
                                    __shfl_sync(FULL, reinterpret_cast<unsigned long long>(cptr), lo_c));
                                const int cnt = (int)__shfl_sync(FULL, (int)min(crem, (int64_t)TB), lo_c);
                                if (f0 + lane < nchunks) {
                                    const int sh16 = (int)(reinterpret_cast<uintptr_t>(g) & 15u);
                                    const int lo = c == 0 ? sh16 : 0;
                                    const int hi = min(16, sh16 + cnt - 16 * c);
                                    uint4 v = make_uint4(0u, 0u, 0u, 0u);
                                    if (hi > 0) {
                                        v = __ldg(reinterpret_cast<const uint4*>(g - sh16 + 16 * c));
                                        if (lo != 0 || hi != 16) {
                                            v.x &= byte_range_mask(lo, hi);
                                            v.y &= byte_range_mask(lo - 4, hi - 4);
                                            v.z &= byte_range_mask(lo - 8, hi - 8);
                                            v.w &= byte_range_mask(lo - 12, hi - 12);
                                        }
                                        na |= (v.x | v.y | v.z | v.w) & 0x80808080u;
                                        v.x = lower4(v.x);
                                        v.y = lower4(v.y);
                                        v.z = lower4(v.z);
                                        v.w = lower4(v.w);
                                    }
                                    *reinterpret_cast<uint4*>(buf + 16 * f) = v;
                                    wcell[f] = (unsigned char)lo_c;
                                }
                            }
                            fill = nchunks << 4;
                            k = kend;
                            off = 0;
                        } else {
                            // cell k alone is longer than the buffer: stage a piece of it
                            const uint8_t* g = reinterpret_cast<const uint8_t*>(
                                __shfl_sync(FULL, reinterpret_cast<unsigned long long>(cptr), k));
                            const int sh16 = (int)(reinterpret_cast<uintptr_t>(g) & 15u);
                            const int cnt = TB - 17;
                            piece_pos0 = sh16;
                            piece_cnt = cnt;
                            const int nchunks = (sh16 + cnt + 16) >> 4;
                            for (int c = lane; c < nchunks; c += 32) {
                                const int lo = c == 0 ? sh16 : 0;
                                const int hi = min(16, sh16 + cnt - 16 * c);
                                uint4 v = make_uint4(0u, 0u, 0u, 0u);
                                if (hi > 0) {
                                    v = __ldg(reinterpret_cast<const uint4*>(g - sh16 + 16 * c));
                                    v.x &= byte_range_mask(lo, hi);
                                    v.y &= byte_range_mask(lo - 4, hi - 4);
                                    v.z &= byte_range_mask(lo - 8, hi - 8);
                                    v.w &= byte_range_mask(lo - 12, hi - 12);
                                    na |= (v.x | v.y | v.z | v.w) & 0x80808080u;
                                    v.x = lower4(v.x);
                                    v.y = lower4(v.y);
                                    v.z = lower4(v.z);
                                    v.w = lower4(v.w);
                                }
                                *reinterpret_cast<uint4*>(buf + 16 * c) = v;
                                wcell[c] = (unsigned char)k;
                            }
                            fill = nchunks << 4;
                        }
                        if (na) flags[0] = 1;
                    }
                    __syncwarp();

                    // ---- classify: four bytes per lane (SWAR), 128 mask bits per step; token starts are listed
                    int ntok = 0;
                    {
                        unsigned carry = 0u;         // token bit of the byte before the window
                        const int nwin = (fill + 127) >> 7;
                        for (int w = 0; w < nwin; ++w) {
                            const unsigned* bp = bufw + 4 + (w << 5) + lane;
                            const unsigned W = bp[0];
                            const unsigned h7 = W & 0x7f7f7f7fu, ascii = ~W & 0x80808080u;
                            // bit 7 of (h7 + (0x80 - x) * 0x01010101) is set iff the byte is >= x
                            const unsigned fl = (h7 + 0x1f1f1f1fu) & ~(h7 + 0x05050505u) & ascii;     // 'a' .. 'z'
                            const unsigned fd = (h7 + 0x50505050u) & ~(h7 + 0x46464646u) & ascii;     // '0' .. '9'
                            // bit 7 of ((h7 ^ c) + 0x7f) is set iff the byte differs from c
                            const unsigned ne_us = (h7 ^ 0x5f5f5f5fu) + 0x7f7f7f7fu;
                            unsigned tcf = (fl | fd | (~ne_us & ascii)) & 0x80808080u;
                            // the Mid* characters . ' : , ; all lie in 0x27 .. 0x3b: windows without a non-digit byte
                            // of that range (most of them) skip the neighbour tests
                            const unsigned rng = (h7 + 0x59595959u) & ~(h7 + 0x44444444u) & ~fd & ascii;
                            if (__any_sync(FULL, rng != 0u)) {
                                const unsigned plast = bp[-1] >> 24, nfirst = bp[1] & 0xffu;
                                const unsigned ne_dot = (h7 ^ 0x2e2e2e2eu) + 0x7f7f7f7fu;
                                const unsigned ne_apo = (h7 ^ 0x27272727u) + 0x7f7f7f7fu;
                                const unsigned ne_col = (h7 ^ 0x3a3a3a3au) + 0x7f7f7f7fu;
                                const unsigned ne_com = (h7 ^ 0x2c2c2c2cu) + 0x7f7f7f7fu;
                                const unsigned ne_sem = (h7 ^ 0x3b3b3b3bu) + 0x7f7f7f7fu;
                                const unsigned mid_l = ~(ne_dot & ne_apo & ne_col);          // MidLetter | MidNumLet | '
                                const unsigned mid_n = ~(ne_dot & ne_apo & ne_com & ne_sem); // MidNum | MidNumLet | '
                                const unsigned pl = (plast - 'a') < 26u ? 0x80u : 0u, pd = (plast - '0') < 10u ? 0x80u : 0u;
                                const unsigned nl = (nfirst - 'a') < 26u ? 0x80000000u : 0u;
                                const unsigned nd = (nfirst - '0') < 10u ? 0x80000000u : 0u;
                                const unsigned joined = (mid_l & ((fl << 8) | pl) & ((fl >> 8) | nl)) |
                                                        (mid_n & ((fd << 8) | pd) & ((fd >> 8) | nd));
                                tcf |= joined & ascii & 0x80808080u;
                            }
                            if ((((w << 5) + lane) << 2) >= fill) tcf = 0u;
                            // bits 7, 15, 23, 31 -> a nibble, placed by lane; eight lanes make a mask word
                            const unsigned nib = (((tcf >> 7) * 0x00204081u) >> 21) & 0xfu;
                            const unsigned mine = nib << ((lane & 7) << 2);
                            const int q = lane >> 3;
                            uint4 m;
                            m.x = __reduce_or_sync(FULL, q == 0 ? mine : 0u);
                            m.y = __reduce_or_sync(FULL, q == 1 ? mine : 0u);
                            m.z = __reduce_or_sync(FULL, q == 2 ? mine : 0u);
                            m.w = __reduce_or_sync(FULL, q == 3 ? mine : 0u);
                            if (lane == 0) reinterpret_cast<uint4*>(mask)[w] = m;
                            // token starts (a token byte after a non-token byte): at most two per lane
                            const unsigned up = __shfl_up_sync(FULL, nib, 1);
                            const unsigned before = ((nib << 1) | (lane == 0 ? carry : (up >> 3))) & 0xfu;
                            const unsigned sn = nib & ~before;
                            const int sc = __popc(sn);
                            const unsigned s1 = __ballot_sync(FULL, sc >= 1), s2 = __ballot_sync(FULL, sc == 2);
                            const int at = ntok + __popc(s1 & lt_mask) + __popc(s2 & lt_mask);
                            const int pos4 = (w << 7) + (lane << 2);
                            if (sc >= 1) toks[at] = (Tok)(pos4 + __ffs(sn) - 1);
                            if (sc == 2) toks[at + 1] = (Tok)(pos4 + 31 - __clz(sn));
                            ntok += __popc(s1) + __popc(s2);
                            carry = m.w >> 31;
                        }
                    }
                    __syncwarp();
                    if (piece_pos0 >= 0) {
                        // a piece of a long cell: everything after its last break is redone with the next piece.  The
                        // last byte does not count as a break (its right neighbour is not staged).
                        const int end = piece_pos0 + piece_cnt;
                        int zpos = -1;
                        for (int wb = ((end - 2) >> 5) & ~31; wb >= 0 && zpos < 0; wb -= 32) {
                            const int wi = wb + lane;
                            unsigned inv = 0u;
                            if ((wi << 5) < end - 1) {
                                inv = ~mask[wi];
                                const int lo = piece_pos0 - (wi << 5), hi = end - 1 - (wi << 5);
                                if (lo > 0) inv &= lo >= 32 ? 0u : ~((1u << lo) - 1u);
                                if (hi < 32) inv &= (1u << hi) - 1u;
                            }
                            const unsigned any = __ballot_sync(FULL, inv != 0u);
                            if (any) {
                                const int src = 31 - __clz(any);
                                const unsigned iv = __shfl_sync(FULL, inv, src);
                                zpos = ((wb + src) << 5) + 31 - __clz(iv);
                            }
                        }
                        if (zpos < 0) {                  // no break in the whole piece: the variant with the larger
                            if (defer_list) row_bad = true;              // buffer redoes the row, or gives up
                            else if (lane == 0) flags[1] = 1;
                            off += piece_cnt;
                        } else {
                            if (ntok > 0 && (int)toks[ntok - 1] > zpos) --ntok;
                            off += zpos - piece_pos0 + 1;
                        }
                    }

                    // ---- pass 1, one lane per token: find its end, drop stop words and short tokens; the survivors of
                    //      at most 32 bytes are compacted to the front of the list as start | (length - 1) << SB, the
                    //      rare longer runs to its back
                    int nsurv = 0, nlong = 0;
                    for (int t0 = 0; t0 < ntok; t0 += 32) {
                        const int t = t0 + lane;
                        int kind = 0;
                        unsigned rec = 0u;
                        if (t < ntok) {
                            const int s = toks[t];
                            int w = s >> 5;
                            unsigned inv = ~mask[w] >> (s & 31);
                            int e;
                            if (inv) {
                                e = s + __ffs(inv) - 1;
                            } else {
                                do inv = ~mask[++w]; while (!inv);
                                e = (w << 5) + __ffs(inv) - 1;
                            }
                            const int len = e - s;
                            if (len > 32) {
                                kind = 2;
                                rec = (unsigned)s;
                            } else if (len >= p.min_token_length) {
                                bool keep = true;
                                const unsigned c0 = buf[s];
                                if (c0 == '_') {              // a run of '_' only is not a token
                                    keep = false;
                                    for (int q = s; q < e; ++q) keep |= buf[q] != '_';
                                }
                                // stop words start with a letter; the pre-filter knows (first letter, second
                                // character, length) of every one of them
                                if (keep && len <= STOP_MAXLEN && (c0 - 'a') < 26u &&
                                    ((stop_pre[stop_pre_index(c0, len > 1 ? buf[s + 1] : 0u)] >> len) & 1u))
                                    keep = !th_is_stop(stop, bufw, s, len);
                                if (keep) {
                                    kind = 1;
                                    rec = (unsigned)s | ((unsigned)(len - 1) << SB);
                                    emitted |= 1u << wcell[s >> 4];
                                }
                            }
                        }
                        const unsigned b1 = __ballot_sync(FULL, kind == 1), b2 = __ballot_sync(FULL, kind == 2);
                        __syncwarp();
                        if (kind == 1) toks[nsurv + __popc(b1 & lt_mask)] = (Tok)rec;
                        if (kind == 2) toks[TB / 2 - 1 - (nlong + __popc(b2 & lt_mask))] = (Tok)rec;
                        nsurv += __popc(b1);
                        nlong += __popc(b2);
                    }
                    __syncwarp();
                    // ---- pass 2, one lane per surviving token: murmur3 and the table
                    for (int t0 = 0; t0 < nsurv && !row_bad; t0 += 32) {
                        const int t = t0 + lane;
                        bool newd = false, ovf = false;
                        if (t < nsurv) {
                            const unsigned rec = toks[t];
                            const int s = (int)(rec & (unsigned)(TB - 1)), len = (int)(rec >> SB) + 1;
                            const int j = c0 + wcell[s >> 4];
                            unsigned ph, pw, pn;
                            th_col_state(colstate, prefix, prefix_len, p, j, ph, pw, pn);
                            const unsigned hv = th_hash_piece(bufw, s, len, ph, pw, pn);
                            const int bucket = p.pow2_mask >= 0 ? (int)(hv & (unsigned)p.pow2_mask)
                                                                : non_negative_mod(hv, p.num_features);
                            const unsigned key = (unsigned)((p.shared ? 0 : j * p.num_features) + bucket);
                            th_insert<Slot>(table, tbits, cb, key, newd, ovf);
                        }
                        ndist += __popc(__ballot_sync(FULL, newd));
                        if (__any_sync(FULL, ovf) || ndist > dist_limit) row_bad = true;
                    }
                    // ---- runs longer than 32 bytes: emitted in pieces of maxTokenLength
                    for (int t0 = 0; t0 < nlong && !row_bad; t0 += 32) {
                        const int t = t0 + lane;
                        bool ovf = false;
                        int nnew = 0;
                        if (t < nlong) {
                            const int s = toks[TB / 2 - 1 - t];
                            int w = s >> 5;
                            unsigned inv = ~mask[w] >> (s & 31);
                            int e;
                            if (inv) {
                                e = s + __ffs(inv) - 1;
                            } else {
                                do inv = ~mask[++w]; while (!inv);
                                e = (w << 5) + __ffs(inv) - 1;
                            }
                            const int kc = wcell[s >> 4];
                            const int j = c0 + kc;
                            unsigned ph, pw, pn;
                            th_col_state(colstate, prefix, prefix_len, p, j, ph, pw, pn);
                            bool has_alnum = false;
                            for (int q = s; q < e; ++q) has_alnum |= buf[q] != '_';
                            for (int ps = s; has_alnum && ps < e; ps += TH_MAX_TOKEN) {
                                const int plen = min(TH_MAX_TOKEN, e - ps);
                                if (plen < p.min_token_length) continue;
                                if (plen <= STOP_MAXLEN && th_is_stop(stop, bufw, ps, plen)) continue;
                                const unsigned hv = th_hash_piece(bufw, ps, plen, ph, pw, pn);
                                const int bucket = p.pow2_mask >= 0 ? (int)(hv & (unsigned)p.pow2_mask)
                                                                    : non_negative_mod(hv, p.num_features);
                                const unsigned key = (unsigned)((p.shared ? 0 : j * p.num_features) + bucket);
                                emitted |= 1u << kc;
                                bool nd1 = false;
                                th_insert<Slot>(table, tbits, cb, key, nd1, ovf);
                                nnew += nd1 ? 1 : 0;
                            }
                        }
                        ndist += __reduce_add_sync(FULL, nnew);
                        if (__any_sync(FULL, ovf) || ndist > dist_limit) row_bad = true;
                    }
                    __syncwarp();
                    if (k >= ncells || row_bad) break;
                }
                // ---- null indicators of this pass: no token of the cell survived (null, empty or all stop words)
                emitted = __reduce_or_sync(FULL, emitted);
                if (p.track_nulls && !row_bad) {
                    bool newd = false;
                    if (lane < ncells && !((emitted >> lane) & 1u)) {
                        const unsigned key = (unsigned)(p.null_base + c0 + lane);
                        unsigned slot = (key * 0x9E3779B1u) >> (32 - tbits);
                        for (;;) {
                            const Slot cur = atomicCAS(&table[slot], EMPTY, (Slot(key) << cb) | Slot(1));
                            if (cur == EMPTY) {
                                newd = true;
                                break;
                            }
                            slot = (slot + 1) & (tsize - 1);
                        }
                    }
                    ndist += __popc(__ballot_sync(FULL, newd));
                    if (ndist > dist_limit) row_bad = true;
                }
                __syncwarp();
            }

            // ---- flush the row's table
            if (row_bad) {
                for (int i = lane; i < tsize; i += 32) table[i] = EMPTY;
                if (lane == 0) {
                    if (defer_list) defer_list[atomicAdd(defer_n, 1u)] = (int32_t)row;
                    else flags[2] = 1;
                    if (SPARSE_OUT && !defer_list) {
                        row_start[row] = 0;
                        row_nnz[row] = 0;
                    }
                }
                __syncwarp();
                continue;
            }
            if (!SPARSE_OUT) {
                const int cls = (p.nclasses > 1 && class_ids) ? class_ids[row] : 0;
                const double e_row = label ? label[row] - label_shift : 0.0;
                // every lane looks at slots lane, lane + 32, ... and remembers the occupied ones in a bitmap, so that the
                // emitting loop below runs with most lanes busy
                // the rare term frequencies > 1 are set aside (in the token list, free by now) and handled afterwards
                // with all lanes busy
                uint2* rare = reinterpret_cast<uint2*>(toks);
                constexpr int RARE_CAP = TB / 2 * (int)sizeof(Tok) / (int)sizeof(uint2);
                int nrare = 0;
                for (int i0 = 0; i0 < tsize; i0 += 1024) {
                    const int lim = min(32, (tsize - i0) >> 5);
                    unsigned occ = 0u;
                    for (int b = 0; b < lim; ++b) occ |= (table[i0 + (b << 5) + lane] != EMPTY ? 1u : 0u) << b;
                    while (__any_sync(FULL, occ != 0u)) {
                        unsigned v = 0u, cnt = 0u;
                        if (occ) {
                            const int i = i0 + ((__ffs(occ) - 1) << 5) + lane;
                            occ &= occ - 1u;
                            const Slot sv = table[i];
                            table[i] = EMPTY;
                            v = (unsigned)(sv >> cb);
                            cnt = (p.binary_freq && v < (unsigned)p.hash_width) ? 1u : (unsigned)(sv & cmask);
                            atomicAdd(&bucket_class_sum[(int64_t)v * p.nclasses + cls], cnt);
                            if (label) atomicAdd(&bucket_wsum[v], (double)cnt * e_row);
                        }
                        const unsigned rb = __ballot_sync(FULL, cnt > 1u);
                        if (rb) {
                            if (nrare + 32 > RARE_CAP) {          // no room (never with the shipped sizes): drain first
                                __syncwarp();
                                for (int i = lane; i < nrare; i += 32) th_rare_update(rare[i], bucket_dup, bucket_sq_extra, bucket_aux);
                                nrare = 0;
                                __syncwarp();
                            }
                            if (cnt > 1u) rare[nrare + __popc(rb & lt_mask)] = make_uint2(v, cnt);
                            nrare += __popc(rb);
                        }
                    }
                }
                __syncwarp();
                for (int i = lane; i < nrare; i += 32) th_rare_update(rare[i], bucket_dup, bucket_sq_extra, bucket_aux);
            } else {
                // The kept buckets go out in ascending order of their new position.  They are first compacted to the
                // front of the table as (new position | value).  Fast way to order them: an ordered scatter -- every
                // entry is pushed into a second table at a slot proportional to its new position, each slot keeping
                // the smaller of what it holds and what arrives and passing the larger on, so that reading that table
                // front to back gives the sorted row.  It lives in the text buffer and the token list (free by now); a
                // crowded row takes several phases, each over a slice of the positions.  When the scatter runs off the
                // end of its table (skewed positions) the compacted list is bitonic-sorted in place instead.
                Slot* t2 = reinterpret_cast<Slot*>(wbase + L::BUF);
                constexpr int T2 = L::T2;
                int kept = 0, np_min = 0x7fffffff, np_max = -1;
                // fit + transform in one pass: with bucket_class_sum given, the statistics of ALL buckets (kept or
                // not) are taken here as well
                const int cls_s = (bucket_class_sum && p.nclasses > 1 && class_ids) ? class_ids[row] : 0;
                const double e_s = (bucket_class_sum && label) ? label[row] - label_shift : 0.0;
                for (int i0 = 0; i0 < tsize; i0 += 32) {
                    const Slot sv = table[i0 + lane];
                    int np = -1;
                    unsigned val = 0u;
                    if (sv != EMPTY) {
                        const unsigned v = (unsigned)(sv >> cb);
                        np = newpos[v];
                        val = (p.binary_freq && v < (unsigned)p.hash_width) ? 1u : (unsigned)(sv & cmask);
                        if (bucket_class_sum) {
                            atomicAdd(&bucket_class_sum[(int64_t)v * p.nclasses + cls_s], val);
                            if (label) atomicAdd(&bucket_wsum[v], (double)val * e_s);
                            if (val > 1u) th_rare_update(make_uint2(v, val), bucket_dup, bucket_sq_extra, bucket_aux);
                        }
                    }
                    const unsigned kb = __ballot_sync(FULL, np >= 0);
                    table[i0 + lane] = EMPTY;
                    __syncwarp();
                    if (np >= 0) {
                        table[kept + __popc(kb & lt_mask)] = (Slot((unsigned)np) << cb) | Slot(val);
                        np_min = min(np_min, np);
                        np_max = max(np_max, np);
                    }
                    kept += __popc(kb);
                    __syncwarp();
                }
                np_min = __reduce_min_sync(FULL, np_min);
                np_max = __reduce_max_sync(FULL, np_max);
                unsigned long long base = 0;
                if (lane == 0) {
                    base = atomicAdd(cursor, (unsigned long long)kept);
                    row_start[row] = (int64_t)base;
                    row_nnz[row] = kept;
                    if ((int64_t)base + kept > capacity) flags[3] = 1;
                }
                base = __shfl_sync(FULL, base, 0);
                // Narrow outputs (separate hash spaces of a few hundred buckets per column): order by a bitmap over the
                // new positions instead.  Per slice of BM_WORDS * 32 positions every entry sets its bit, a scan of the
                // words gives the number of set bits before each word, and an entry's place in the row is that prefix
                // plus the set bits below its own -- no displacement chains, every lane busy.
                constexpr int BM_WORDS = 256;
                static_assert(BM_WORDS * 6 <= L::MASK - L::BUF, "bitmap + word prefixes live in the text buffer");
                bool placed = false;
                if (kept > 0 && (long long)np_max - np_min < 8ll * BM_WORDS * 32) {
                    unsigned* bm = reinterpret_cast<unsigned*>(wbase + L::BUF);
                    unsigned short* pre = reinterpret_cast<unsigned short*>(bm + BM_WORDS);
                    int done = 0;
                    for (int lo_p = np_min; lo_p <= np_max; lo_p += BM_WORDS * 32) {
                        for (int i = lane; i < BM_WORDS; i += 32) bm[i] = 0u;
                        __syncwarp();
                        for (int i = lane; i < kept; i += 32) {
                            const unsigned q = (unsigned)((int)(table[i] >> cb) - lo_p);
                            if (q < (unsigned)(BM_WORDS * 32)) atomicOr(&bm[q >> 5], 1u << (q & 31u));
                        }
                        __syncwarp();
                        int run = 0;
                        for (int w0 = 0; w0 < BM_WORDS; w0 += 32) {
                            const int c = __popc(bm[w0 + lane]);
                            int incl = c;
#pragma unroll
                            for (int o = 1; o < 32; o <<= 1) {
                                const int up = __shfl_up_sync(FULL, incl, o);
                                if (lane >= o) incl += up;
                            }
                            pre[w0 + lane] = (unsigned short)(run + incl - c);
                            run += __shfl_sync(FULL, incl, 31);
                        }
                        __syncwarp();
                        for (int i = lane; i < kept; i += 32) {
                            const Slot x = table[i];
                            const unsigned q = (unsigned)((int)(x >> cb) - lo_p);
                            if (q < (unsigned)(BM_WORDS * 32)) {
                                const unsigned w = q >> 5;
                                const int64_t at = (int64_t)base + done + pre[w] +
                                                   __popc(bm[w] & ((1u << (q & 31u)) - 1u));
                                if (at < capacity) {
                                    out_idx[at] = (int32_t)(x >> cb);
                                    out_val[at] = (int32_t)(x & cmask);
                                }
                            }
                        }
                        done += run;
                        __syncwarp();
                    }
                    for (int i = lane; i < kept; i += 32) table[i] = EMPTY;
                    placed = true;
                }
                const int phases = max(1, (kept + T2 / 2 - 1) / (T2 / 2));
                bool ordered = !placed && kept > 0 && phases <= 8;
                if (ordered) {
                    const unsigned long long span = (unsigned long long)(np_max - np_min) + 1ull;
                    int done = 0;
                    for (int ph = 0; ph < phases; ++ph) {
                        const int lo_p = np_min + (int)(span * ph / phases);
                        const int hi_p = np_min + (int)(span * (ph + 1) / phases);
                        const float scale = (float)(T2 - 64) / (float)(hi_p - lo_p);
                        for (int i = lane; i < T2; i += 32) t2[i] = EMPTY;
                        __syncwarp();
                        bool fail = false;
                        for (int i = lane; i < kept; i += 32) {
                            Slot x = table[i];
                            const int np = (int)(x >> cb);
                            if (np < lo_p || np >= hi_p) continue;
                            int slot = min((int)((float)(np - lo_p) * scale), T2 - 64);
                            for (;;) {
                                const Slot old = atomicMin(&t2[slot], x);
                                if (old == EMPTY) break;
                                x = old > x ? old : x;
                                if (++slot >= T2) {
                                    fail = true;
                                    break;
                                }
                            }
                        }
                        __syncwarp();
                        if (__any_sync(FULL, fail)) {
                            ordered = false;
                            break;
                        }
                        for (int i0 = 0; i0 < T2 && done < kept; i0 += 32) {
                            const Slot sv = t2[i0 + lane];
                            const unsigned ob = __ballot_sync(FULL, sv != EMPTY);
                            const int64_t at = (int64_t)base + done + __popc(ob & lt_mask);
                            if (sv != EMPTY && at < capacity) {
                                out_idx[at] = (int32_t)(sv >> cb);
                                out_val[at] = (int32_t)(sv & cmask);
                            }
                            done += __popc(ob);
                        }
                        __syncwarp();
                    }
                    if (ordered)
                        for (int i = lane; i < kept; i += 32) table[i] = EMPTY;
                }
                __syncwarp();
                if (lane < 4) reinterpret_cast<unsigned*>(wbase + L::BUF)[lane] = 0u;       // the slack before the text
                if (!ordered && !placed) {
                    int m = 32;
                    while (m < kept) m <<= 1;
                    for (int k2 = 2; k2 <= m; k2 <<= 1) {
                        for (int jj = k2 >> 1; jj > 0; jj >>= 1) {
                            for (int t = lane; t < (m >> 1); t += 32) {
                                const int i = ((t & ~(jj - 1)) << 1) | (t & (jj - 1));
                                const int ixj = i | jj;
                                const Slot x = table[i], y = table[ixj];
                                const bool up = (i & k2) == 0;
                                if ((x > y) == up) {
                                    table[i] = y;
                                    table[ixj] = x;
                                }
                            }
                            __syncwarp();
                        }
                    }
                    for (int i = lane; i < kept; i += 32) {
                        const Slot sv = table[i];
                        table[i] = EMPTY;
                        const int64_t at = (int64_t)base + i;
                        if (at < capacity) {
                            out_idx[at] = (int32_t)(sv >> cb);
                            out_val[at] = (int32_t)(sv & cmask);
                        }
                    }
                }
            }
            __syncwarp();
        }
    }
}


// Column selection on sparse rows (SanityCheckerModel's removeFeatures on the rows the fit pass already wrote): a warp
// per row keeps the entries whose new position is >= 0, in order.  WRITE = false only counts them.
template <bool WRITE>
__global__ void __launch_bounds__(256)
sparse_rows_select_kernel(const int64_t* __restrict__ row_start, const int32_t* __restrict__ row_nnz,
                          const int32_t* __restrict__ idx, const int32_t* __restrict__ val, int64_t n,
                          const int32_t* __restrict__ newpos, const int64_t* __restrict__ out_start,
                          int32_t* __restrict__ out_nnz, int32_t* __restrict__ out_idx, int32_t* __restrict__ out_val) {
    const int lane = threadIdx.x & 31;
    const unsigned lt = (1u << lane) - 1u;
    const int64_t warps = (int64_t)gridDim.x * (blockDim.x >> 5);
    for (int64_t r = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); r < n; r += warps) {
        const int64_t a = row_start[r];
        const int m = row_nnz[r];
        const int64_t o = WRITE ? out_start[r] : 0;
        int kept = 0;
        for (int i0 = 0; i0 < m; i0 += 32) {
            const int i = i0 + lane;
            const int np = i < m ? newpos[idx[a + i]] : -1;
            const unsigned k = __ballot_sync(FULL, np >= 0);
            if (WRITE && np >= 0) {
                const int64_t at = o + kept + __popc(k & lt);
                out_idx[at] = np;
                out_val[at] = val[a + i];
            }
            kept += __popc(k);
        }
        if (!WRITE && lane == 0) out_nnz[r] = kept;
    }
}

}  // namespace tmog

#include "text_cells.cuh"
#include "text_span.cuh"
#include "text_packed.cuh"

using namespace tmog;
// Snowball english_stop.txt (174 words) as loaded by Lucene's StandardAnalyzer in the reference
// (core/src/main/scala/com/salesforce/op/utils/text/LuceneTextAnalyzer.scala:159-161)
static const char* const kStopWords[] = {
    "i", "me", "my", "myself", "we", "our", "ours", "ourselves", "you", "your", "yours", "yourself", "yourselves", "he",
    "him", "his", "himself", "she", "her", "hers", "herself", "it", "its", "itself", "they", "them", "their", "theirs",
    "themselves", "what", "which", "who", "whom", "this", "that", "these", "those", "am", "is", "are", "was", "were",
    "be", "been", "being", "have", "has", "had", "having", "do", "does", "did", "doing", "would", "should", "could",
    "ought", "i'm", "you're", "he's", "she's", "it's", "we're", "they're", "i've", "you've", "we've", "they've", "i'd",
    "you'd", "he'd", "she'd", "we'd", "they'd", "i'll", "you'll", "he'll", "she'll", "we'll", "they'll", "isn't",
    "aren't", "wasn't", "weren't", "hasn't", "haven't", "hadn't", "doesn't", "don't", "didn't", "won't", "wouldn't",
    "shan't", "shouldn't", "can't", "cannot", "couldn't", "mustn't", "let's", "that's", "who's", "what's", "here's",
    "there's", "when's", "where's", "why's", "how's", "a", "an", "the", "and", "but", "if", "or", "because", "as",
    "until", "while", "of", "at", "by", "for", "with", "about", "against", "between", "into", "through", "during",
    "before", "after", "above", "below", "to", "from", "up", "down", "in", "out", "on", "off", "over", "under", "again",
    "further", "then", "once", "here", "there", "when", "where", "why", "how", "all", "any", "both", "each", "few",
    "more", "most", "other", "some", "such", "no", "nor", "not", "only", "own", "same", "so", "than", "too", "very"};

static int upload_stop_words() {
    static bool done[64] = {false};
    int dev = 0;
    TMOG_CUDA(cudaGetDevice(&dev));
    if (dev >= 0 && dev < 64 && done[dev]) return TMOG_OK;
    static StopEntry table[STOP_SLOTS];
    memset(table, 0, sizeof(table));
    const int nwords = (int)(sizeof(kStopWords) / sizeof(kStopWords[0]));
    for (int i = 0; i < nwords; ++i) {      // cuckoo insertion: every word ends in one of its two slots
        StopEntry cur;
        memset(&cur, 0, sizeof(cur));
        cur.len = (unsigned char)strlen(kStopWords[i]);
        memcpy(cur.w, kStopWords[i], cur.len);
        unsigned pos = ~0u;
        bool placed = false;
        for (int kick = 0; kick < 256 && !placed; ++kick) {
            unsigned pw[3] = {0u, 0u, 0u};
            for (int q = 0; q < cur.len; ++q) pw[q >> 2] |= (unsigned)(unsigned char)cur.w[q] << (8 * (q & 3));
            const unsigned a = stop_slot_hash(pw[0], pw[1], pw[2]), b = stop_slot_hash2(pw[0], pw[1], pw[2]);
            pos = (pos == a) ? b : a;
            if (table[pos].len == 0) {
                table[pos] = cur;
                placed = true;
            } else {
                std::swap(table[pos], cur);
            }
        }
        TMOG_REQUIRE(placed, "stop-word table: cuckoo insertion failed");
    }
    TMOG_CUDA(cudaMemcpyToSymbol(c_stop, table, sizeof(table)));
    static StopSlot packed[STOP_SLOTS];
    memset(packed, 0, sizeof(packed));
    for (int i = 0; i < STOP_SLOTS; ++i) {
        packed[i].len = table[i].len;
        for (int q = 0; q < table[i].len; ++q) {
            const unsigned c = (unsigned char)table[i].w[q];
            unsigned& w = q < 4 ? packed[i].w0 : q < 8 ? packed[i].w1 : packed[i].w2;
            w |= c << (8 * (q & 3));
        }
    }
    TMOG_CUDA(cudaMemcpyToSymbol(c_stop_packed, packed, sizeof(packed)));
    static unsigned short pre[STOP_PRE];
    memset(pre, 0, sizeof(pre));
    for (int i = 0; i < nwords; ++i) {
        const char* w = kStopWords[i];
        const size_t len = strlen(w);
        TMOG_REQUIRE(len >= 1 && len <= STOP_MAXLEN && w[0] >= 'a' && w[0] <= 'z', "stop word '%s' breaks the pre-filter", w);
        pre[stop_pre_index((unsigned char)w[0], len > 1 ? (unsigned char)w[1] : 0u)] |= (unsigned short)(1u << len);
    }
    TMOG_CUDA(cudaMemcpyToSymbol(c_stop_pre, pre, sizeof(pre)));
    static unsigned bloom[STOP_BLOOM_WORDS];
    memset(bloom, 0, sizeof(bloom));
    for (int i = 0; i < nwords; ++i) {
        const char* w = kStopWords[i];
        const size_t len = strlen(w);
        unsigned first4 = 0u;
        for (size_t q = 0; q < len && q < 4; ++q) first4 |= (unsigned)(unsigned char)w[q] << (8 * q);
        const unsigned b = stop_bloom_index(first4, (unsigned)len);
        bloom[b >> 5] |= 1u << (b & 31);
    }
    TMOG_CUDA(cudaMemcpyToSymbol(c_stop_bloom, bloom, sizeof(bloom)));
    if (dev >= 0 && dev < 64) done[dev] = true;
    return TMOG_OK;
}

extern "C" int tmog_text_stats_table_bytes(int slots, int nclasses, size_t* bytes) {
    TMOG_REQUIRE(bytes != nullptr && slots >= 64 && (slots & (slots - 1)) == 0 && nclasses >= 1, "bad arguments");
    // hash [slots] u64 | first_row [slots] u64 | counts [slots*nclasses] u32 (padded to 8) | ndistinct+pad | len_sum[2] | nonnull
    *bytes = sizeof(uint64_t) * (size_t)slots * 2 + (((size_t)slots * nclasses * 4 + 7) & ~size_t(7)) + 8 + 16 + 8;
    return TMOG_OK;
}

static int text_stats_launch(const int64_t* col_desc, int ncols, int64_t n, const uint8_t* class_ids, int nclasses,
                             int clean_text, int want_lengths, int slots, int max_distinct, void* tables,
                             int32_t* flags, void* stream) {
    TMOG_REQUIRE(ncols >= 0 && n >= 0 && nclasses >= 1 && nclasses <= 255, "bad shape");
    TMOG_REQUIRE(slots >= 64 && (slots & (slots - 1)) == 0, "slots must be a power of two >= 64");
    TMOG_REQUIRE(max_distinct >= 0 && max_distinct + 1 < slots / 2, "max_distinct %d too large for %d slots", max_distinct,
                 slots);
    if (ncols == 0) return TMOG_OK;
    TMOG_REQUIRE(col_desc && tables && flags, "NULL argument");
    TMOG_REQUIRE(nclasses == 1 || class_ids != nullptr, "class_ids is NULL");
    cudaStream_t st = (cudaStream_t)stream;
    size_t per = 0;
    tmog_text_stats_table_bytes(slots, nclasses, &per);
    // lay the per-column tables out inside `tables`: first ncols TextStatsTable structs, then the storage
    const size_t head = (sizeof(TextStatsTable) * (size_t)ncols + 255) & ~size_t(255);
    std::vector<TextStatsTable> host(ncols);
    uint8_t* base = static_cast<uint8_t*>(tables) + head;
    for (int c = 0; c < ncols; ++c) {
        uint8_t* q = base + per * c;
        host[c].hash = reinterpret_cast<unsigned long long*>(q);
        host[c].first_row = host[c].hash + slots;
        host[c].counts = reinterpret_cast<unsigned int*>(host[c].first_row + slots);
        uint8_t* tail = q + sizeof(uint64_t) * (size_t)slots * 2 + (((size_t)slots * nclasses * 4 + 7) & ~size_t(7));
        host[c].ndistinct = reinterpret_cast<int*>(tail);
        host[c].len_sum = reinterpret_cast<unsigned long long*>(tail + 8);
        host[c].nonnull = reinterpret_cast<unsigned long long*>(tail + 24);
    }
    TMOG_CUDA(cudaMemsetAsync(base, 0, per * ncols, st));
    for (int c = 0; c < ncols; ++c)   // hash + first_row start as all-ones (empty / "no row yet")
        TMOG_CUDA(cudaMemsetAsync(host[c].hash, 0xFF, sizeof(uint64_t) * (size_t)slots * 2, st));
    TMOG_CUDA(cudaMemcpyAsync(tables, host.data(), sizeof(TextStatsTable) * ncols, cudaMemcpyHostToDevice, st));
    TMOG_CUDA(cudaStreamSynchronize(st));  // `host` goes out of scope
    if (n == 0) return TMOG_OK;
    // The first rows go through a small launch of their own: a free-text column passes the cardinality cap within its
    // first few hundred cells, and the threads of the main launch then see that at once -- instead of a whole wave of
    // them hammering a table of a few hundred slots with their first cells.
    const int64_t pilot = std::min<int64_t>(n, 4096);
    text_stats_kernel<<<dim3((unsigned)((pilot + 255) / 256), ncols), 256, 0, st>>>(
        reinterpret_cast<const TextColumn*>(col_desc), ncols, 0, pilot, class_ids, nclasses, clean_text, want_lengths,
        slots, max_distinct, static_cast<const TextStatsTable*>(tables), flags);
    TMOG_LAUNCH_CHECK();
    if (n > pilot) {
        const int gx = (int)std::min<int64_t>((n - pilot + 255) / 256, (int64_t)sm_count() * 4);
        text_stats_kernel<<<dim3(gx, ncols), 256, 0, st>>>(
            reinterpret_cast<const TextColumn*>(col_desc), ncols, pilot, n, class_ids, nclasses, clean_text,
            want_lengths, slots, max_distinct, static_cast<const TextStatsTable*>(tables), flags);
        TMOG_LAUNCH_CHECK();
    }
    return TMOG_OK;
}

extern "C" int tmog_text_stats_dev(const int64_t* col_desc, int ncols, int64_t n, const uint8_t* class_ids, int nclasses,
                                   int clean_text, int slots, int max_distinct, void* tables, int32_t* flags,
                                   void* stream) {
    return text_stats_launch(col_desc, ncols, n, class_ids, nclasses, clean_text, 1, slots, max_distinct, tables, flags,
                             stream);
}

extern "C" int tmog_text_stats_values_dev(const int64_t* col_desc, int ncols, int64_t n, const uint8_t* class_ids,
                                          int nclasses, int clean_text, int slots, int max_distinct, void* tables,
                                          int32_t* flags, void* stream) {
    return text_stats_launch(col_desc, ncols, n, class_ids, nclasses, clean_text, 0, slots, max_distinct, tables, flags,
                             stream);
}

extern "C" int tmog_text_stats_layout(int ncols, int slots, int nclasses, size_t* head_bytes, size_t* per_column_bytes) {
    TMOG_REQUIRE(head_bytes && per_column_bytes, "NULL argument");
    *head_bytes = (sizeof(TextStatsTable) * (size_t)ncols + 255) & ~size_t(255);
    return tmog_text_stats_table_bytes(slots, nclasses, per_column_bytes);
}

extern "C" int tmog_text_codes_dev(const int64_t* col_desc, const int32_t* col_index, int nsel, int64_t n,
                                   int clean_text, int slots, const void* tables, int32_t* codes, int64_t ld,
                                   int32_t* flags, void* stream) {
    TMOG_REQUIRE(nsel >= 0 && n >= 0 && ld >= n, "bad shape");
    if (nsel == 0 || n == 0) return TMOG_OK;
    TMOG_REQUIRE(col_desc && col_index && tables && codes, "NULL argument");
    const int gx = (int)std::min<int64_t>((n + 255) / 256, (int64_t)sm_count() * 4);
    text_code_kernel<<<dim3(gx, nsel), 256, 0, (cudaStream_t)stream>>>(
        reinterpret_cast<const TextColumn*>(col_desc), col_index, nsel, n, clean_text, slots,
        static_cast<const TextStatsTable*>(tables), codes, ld, flags);
    TMOG_LAUNCH_CHECK();
    return TMOG_OK;
}

// scratch of the tokenise + hash entry points: a row counter per level, the lengths of the two lists of deferred rows,
// the lists
struct ThScratch {
    unsigned long long row_counter[3];
    unsigned defer_n[2];
};

extern "C" int tmog_text_hash_scratch_bytes(int64_t n, size_t* bytes) {
    TMOG_REQUIRE(bytes != nullptr && n >= 0 && n < 0x7fffffffLL, "bad arguments");
    *bytes = sizeof(ThScratch) + 2 * sizeof(int32_t) * (size_t)n;
    return TMOG_OK;
}

struct ThArgs {
    const TextColumn* cols;
    int ncols;
    int64_t n;
    const uint8_t* prefix;
    const int32_t* prefix_len;
    TextHashParams p;
    const uint8_t* class_ids;
    uint32_t *class_sum, *dup;
    unsigned long long* sq_extra;
    uint32_t* aux;
    const double* label;
    double label_shift;
    double* wsum;
    const int32_t* newpos;
    int64_t* row_start;
    int32_t *row_nnz, *out_idx, *out_val;
    unsigned long long* cursor;
    int64_t capacity;
    int32_t* flags;
    ThScratch* scratch;
    size_t scratch_bytes;
    int n_hashed;          // columns [0, n_hashed) are hashed; the rest only give null indicators / text lengths
    int len_base;          // vector index of the first text-length column, or -1
    cudaStream_t stream;
    TsPack pack;           // arena == NULL: the cells are not kept
};

// level 0 reads all rows, level 1 the rows level 0 deferred, level 2 the rows level 1 deferred
template <typename Slot, int T, int TB, int WARPS, bool SPARSE_OUT>
static int th_launch(const ThArgs& a, int level, int max_token_bound) {
    auto kernel = text_hash_kernel<Slot, T, TB, WARPS, SPARSE_OUT>;
    const size_t smem = TH_FIXED_SMEM + 32 + (size_t)WARPS * ThLayout<Slot, T, TB>::BYTES;
    TMOG_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int per_sm = 0;
    TMOG_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, WARPS * 32, smem));
    TMOG_REQUIRE(per_sm >= 1, "text hash kernel does not fit an SM");
    int64_t grid = (int64_t)sm_count() * per_sm;
    if (level == 0) grid = std::max<int64_t>(1, std::min<int64_t>(grid, (a.n + WARPS - 1) / WARPS));
    int32_t* lists = reinterpret_cast<int32_t*>(a.scratch + 1);
    TextHashParams p = a.p;
    p.max_token_bound = max_token_bound;
    kernel<<<(int)grid, WARPS * 32, smem, a.stream>>>(
        a.cols, a.ncols, a.n, a.prefix, a.prefix_len, p, a.class_ids,
        level == 0 ? nullptr : lists + (size_t)(level - 1) * a.n, level == 0 ? nullptr : &a.scratch->defer_n[level - 1],
        &a.scratch->row_counter[level],
        level == 2 ? nullptr : lists + (size_t)level * a.n, level == 2 ? nullptr : &a.scratch->defer_n[level],
        a.class_sum, a.dup, a.sq_extra, a.aux, a.label, a.label_shift, a.wsum, a.newpos, a.row_start, a.row_nnz, a.out_idx, a.out_val, a.cursor,
        a.capacity, a.flags);
    TMOG_LAUNCH_CHECK();
    return TMOG_OK;
}

// A row is sent to a level by an upper bound of its tokens (half its bytes).  The bound is loose -- ordinary text has
// a token per 6 to 8 bytes and repeats words -- so a level takes rows whose bound is about five times the distinct
// buckets its table holds; the few rows that do overflow are redone by the next level.
// Three launches, by the size of the row's table: rows whose text is short enough for about 900 distinct buckets run
// with 32 warps per SM (a 1 KB text buffer: cells are staged four or five at a time), rows up to about 1800 with 16 warps per SM, the rest (and whatever overflowed a smaller table)
// with 64-bit slots in most of the shared memory of an SM.  Packed 32-bit slots need >= 8 bits left for the count.
template <bool SPARSE_OUT>
static int th_run(ThArgs& a) {
    TMOG_CUDA(cudaMemsetAsync(a.scratch, 0, sizeof(ThScratch), a.stream));
    const int64_t max_key = (int64_t)a.p.null_base + a.ncols;      // keys (and new positions) are < max_key
    int key_bits = 1;
    while ((1ll << key_bits) - 1 <= max_key) ++key_bits;
    if (key_bits <= 24) {
        a.p.count_bits = 32 - key_bits;
        if (int rc = th_launch<unsigned, 1024, 1024, 16, SPARSE_OUT>(a, 0, 4608)) return rc;
        if (int rc = th_launch<unsigned, 2048, 2048, 8, SPARSE_OUT>(a, 1, 9216)) return rc;
    } else {
        a.p.count_bits = 32;
        if (int rc = th_launch<unsigned long long, 512, 2048, 12, SPARSE_OUT>(a, 0, 2304)) return rc;
        if (int rc = th_launch<unsigned long long, 1024, 2048, 8, SPARSE_OUT>(a, 1, 4608)) return rc;
    }
    a.p.count_bits = 32;
    return th_launch<unsigned long long, 16384, 4096, 1, SPARSE_OUT>(a, 2, 0x7fffffff);
}

static int text_hash_common(const int64_t* col_desc, int ncols, int64_t n, const uint8_t* prefix,
                            const int32_t* prefix_len, const int32_t* params, void* scratch, size_t scratch_bytes,
                            ThArgs* a) {
    TMOG_REQUIRE(ncols >= 0 && n >= 0 && n < 0x7fffffffLL, "bad shape");
    TMOG_REQUIRE(col_desc && prefix && prefix_len && params, "NULL argument");
    TextHashParams& p = a->p;
    p.num_features = params[0];
    p.shared = params[1];
    p.binary_freq = params[2];
    p.min_token_length = params[3];
    p.track_nulls = params[4];
    p.prepend = params[5];
    p.null_base = params[6];
    p.nclasses = params[7];
    TMOG_REQUIRE(p.num_features > 0 && p.num_features <= 10000000,
                 "requirement failed: numFeatures must be in (0, 10000000], got %d", p.num_features);
    TMOG_REQUIRE(p.nclasses >= 1 && p.nclasses <= 255, "nclasses must be in [1, 255]");
    const int nh = params[8] > 0 ? params[8] : ncols;
    const int64_t hw = p.shared ? p.num_features : (int64_t)p.num_features * nh;
    TMOG_REQUIRE(hw + 2 * (int64_t)ncols < 0x7fffffffLL, "hash block too wide");
    p.hash_width = (int)hw;
    TMOG_REQUIRE(p.null_base >= p.hash_width && (int64_t)p.null_base + ncols < 0x7fffffffLL,
                 "null_base must not overlap the hash block");
    p.pow2_mask = (p.num_features & (p.num_features - 1)) == 0 ? p.num_features - 1 : -1;
    p.count_bits = 32;
    a->n_hashed = params[8] > 0 ? params[8] : ncols;
    a->len_base = params[9] > 0 ? params[9] : -1;
    TMOG_REQUIRE(a->n_hashed <= ncols, "n_hashed %d exceeds the %d columns", a->n_hashed, ncols);
    if (a->n_hashed < ncols || a->len_base >= 0) {
        const int64_t hw2 = p.shared ? p.num_features : (int64_t)p.num_features * a->n_hashed;
        p.hash_width = (int)hw2;
        TMOG_REQUIRE(a->len_base < 0 || a->len_base >= p.hash_width, "len_base must not overlap the hash block");
        TMOG_REQUIRE(p.null_base >= p.hash_width + (a->len_base >= 0 ? ncols : 0), "null_base overlaps");
    }
    a->scratch_bytes = scratch_bytes;
    size_t need = 0;
    if (int rc = tmog_text_hash_scratch_bytes(n, &need)) return rc;
    TMOG_REQUIRE(scratch != nullptr && scratch_bytes >= need, "scratch too small: %zu bytes needed", need);
    TMOG_REQUIRE((reinterpret_cast<uintptr_t>(scratch) & 7u) == 0, "scratch must be 8-byte aligned");
    a->cols = reinterpret_cast<const TextColumn*>(col_desc);
    a->ncols = ncols;
    a->n = n;
    a->prefix = prefix;
    a->prefix_len = prefix_len;
    a->scratch = reinterpret_cast<ThScratch*>(scratch);
    return upload_stop_words();
}

// ---- the per-cell path (text_cells.cuh): separate hash spaces of at most 512 buckets ---------------------------------
static bool tc_eligible(const TextHashParams& p, int ncols, int n_hashed) {
    const char* force = getenv("TMOG_TEXT_PATH");
    if (force && strcmp(force, "rows") == 0) return false;
    if (ncols < 1 || p.num_features > TC_MAXF) return false;
    return !p.shared || n_hashed <= 1;
}

static size_t tc_scratch_need(int64_t n, int ncols, int n_hashed, bool sparse_out) {
    const size_t keepw = ((size_t)n_hashed * TC_FW * 4 + 255) & ~size_t(255);
    const size_t bits = ((((size_t)n * ncols + 31) / 32) * 4 + 255) & ~size_t(255);
    return 256 + keepw + bits + (sparse_out ? (size_t)n * ncols * TC_REC_WORDS * 4 : 0);
}

template <bool STATS, bool CELLS>
static int tc_run(const ThArgs& a) {
    TcParams p;
    p.num_features = a.p.num_features;
    p.binary_freq = a.p.binary_freq;
    p.min_token_length = a.p.min_token_length;
    p.track_nulls = a.p.track_nulls;
    p.prepend = a.p.prepend;
    p.null_base = a.p.null_base;
    p.len_base = a.len_base;
    p.n_hashed = a.n_hashed;
    p.nclasses = a.p.nclasses;
    p.pow2_mask = a.p.pow2_mask;
    const char* slow = getenv("TMOG_TEXT_FORCE_SLOW");
    p.force_defer = (slow && slow[0] == '1') ? 1 : 0;
    uint8_t* base = reinterpret_cast<uint8_t*>(a.scratch);
    TcScratch* sc = reinterpret_cast<TcScratch*>(base);
    const size_t keepw_bytes = ((size_t)a.n_hashed * TC_FW * 4 + 255) & ~size_t(255);
    const size_t bits_bytes = ((((size_t)a.n * a.ncols + 31) / 32) * 4 + 255) & ~size_t(255);
    unsigned* keepw = reinterpret_cast<unsigned*>(base + 256);
    unsigned* defer_bits = reinterpret_cast<unsigned*>(base + 256 + keepw_bytes);
    unsigned* records = reinterpret_cast<unsigned*>(base + 256 + keepw_bytes + bits_bytes);
    TMOG_CUDA(cudaMemsetAsync(base, 0, 256, a.stream));
    TMOG_CUDA(cudaMemsetAsync(defer_bits, 0, bits_bytes, a.stream));
    // TMOG_TEXT_PATH=cells: the lane-per-cell kernel (kept for comparison); default: the span kernel
    const char* path = getenv("TMOG_TEXT_PATH");
    const bool lane_cells = path && strcmp(path, "cells") == 0;
    const bool pack = a.pack.arena != nullptr;
    TMOG_REQUIRE(!pack || (STATS && !CELLS && !lane_cells), "packed cells come from the statistics pass of the span kernel");
    const int threads = lane_cells ? TC_WARPS * 32 : TS_WARPS * 32;
    const size_t smem = lane_cells
        ? TH_STOP_PRE_BYTES + 64 + (size_t)TC_WARPS * (TC_WARP_ROWS + (STATS ? TC_STAT_ROWS : 0)) * 32 * 4
        : TS_FIXED_SMEM + (STATS ? 2 * TC_MAXF * 4 : 0) + (size_t)TS_WARPS * TS_WARP_BYTES;
    const int item_rows = lane_cells ? TC_ITEM_ROWS : TS_ITEM_ROWS;
    const int64_t items = ((a.n + item_rows - 1) / item_rows) * a.ncols;
    auto launch = [&](auto kernel, auto... extra) -> int {
        TMOG_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        int per_sm = 0;
        TMOG_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, threads, smem));
        TMOG_REQUIRE(per_sm >= 1, "text cell kernel does not fit an SM");
        const int grid = (int)std::max<int64_t>(1, std::min<int64_t>((int64_t)sm_count() * per_sm, items));
        kernel<<<grid, threads, smem, a.stream>>>(a.cols, a.ncols, a.n, a.prefix, a.prefix_len, p, a.class_ids, sc,
                                                  defer_bits, a.class_sum, a.dup, a.sq_extra, a.aux, a.label,
                                                  a.label_shift, a.wsum, records, a.flags, extra...);
        return TMOG_OK;
    };
    int rc;
    if (lane_cells) rc = launch(text_cells_kernel<STATS, CELLS>);
    else if (STATS && !CELLS && pack) rc = launch(text_span_kernel<STATS, CELLS, STATS && !CELLS>, a.pack);
    else rc = launch(text_span_kernel<STATS, CELLS, false>, TsPack{});
    if (rc) return rc;
    TMOG_LAUNCH_CHECK();
    if (STATS) {
        const int64_t words = ((int64_t)a.n * a.ncols + 31) / 32;
        // a thread per 32 cells: the few cells that were set aside are spread over as many threads as possible (a thread
        // tokenises its cells one after the other)
        const int sgrid = (int)std::max<int64_t>(1, std::min<int64_t>((words + 127) / 128, (int64_t)1 << 20));
        text_cells_slow_kernel<<<sgrid, 128, 0, a.stream>>>(a.cols, a.ncols, a.n, a.prefix, a.prefix_len, p, a.class_ids,
                                                            defer_bits, a.class_sum, a.dup, a.sq_extra, a.aux, a.label,
                                                            a.label_shift, a.wsum, a.flags, a.pack);
        TMOG_LAUNCH_CHECK();
    }
    if (CELLS) {
        const int nw = a.n_hashed * TC_FW;
        if (nw > 0) {
            tc_keep_words_kernel<<<(nw + 127) / 128, 128, 0, a.stream>>>(a.newpos, a.n_hashed, p.num_features, keepw);
            TMOG_LAUNCH_CHECK();
        }
        const size_t per_warp = ((size_t)a.ncols * TC_FW * 4 + TC_MAXF * 2 + (size_t)a.ncols * TC_FW * 2 + 15) & ~size_t(15);
        const size_t rsmem = per_warp * 8;
        TMOG_REQUIRE(rsmem <= 200 * 1024, "too many text columns for the row assembly (%d)", a.ncols);
        TMOG_CUDA(cudaFuncSetAttribute(text_rows_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)rsmem));
        const int rgrid = (int)std::max<int64_t>(1, std::min<int64_t>((a.n + 7) / 8, (int64_t)sm_count() * 8));
        text_rows_kernel<<<rgrid, 256, rsmem, a.stream>>>(a.cols, a.ncols, a.n, a.prefix, a.prefix_len, p, records, keepw,
                                                          a.newpos, a.row_start, a.row_nnz, a.out_idx, a.out_val, a.cursor,
                                                          a.capacity, a.flags);
        TMOG_LAUNCH_CHECK();
    }
    return TMOG_OK;
}

// which path serves a call: the per-cell path when the layout allows it and the caller's scratch holds its records
template <bool STATS, bool CELLS>
static int text_hash_dispatch(ThArgs& a) {
    if (tc_eligible(a.p, a.ncols, a.n_hashed) && a.scratch_bytes >= tc_scratch_need(a.n, a.ncols, a.n_hashed, CELLS))
        return tc_run<STATS, CELLS>(a);
    TMOG_REQUIRE(a.n_hashed == a.ncols && a.len_base < 0,
                 "ignored text columns / text lengths need separate hash spaces of at most %d buckets per column and "
                 "the scratch of tmog_text_hash_scratch_bytes2", TC_MAXF);
    return CELLS ? th_run<true>(a) : th_run<false>(a);
}

extern "C" int tmog_text_length_hist_dev(const int64_t* col_desc, int ncols, int64_t n, int64_t row_base, int clean_text,
                                         int tokens, int min_token_length, int nbins, uint64_t* hist,
                                         uint64_t* first_row, int32_t* flags, void* stream) {
    TMOG_REQUIRE(ncols >= 0 && n >= 0 && nbins >= 2 && row_base >= 0, "bad shape");
    if (ncols == 0 || n == 0) return TMOG_OK;
    TMOG_REQUIRE(col_desc && hist && first_row && flags, "NULL argument");
    if (int rc = upload_stop_words()) return rc;
    const int gx = (int)std::max<int64_t>(1, std::min<int64_t>((n + 255) / 256, (int64_t)sm_count() * 8));
    text_length_hist_kernel<<<dim3(gx, ncols), 256, 0, (cudaStream_t)stream>>>(
        reinterpret_cast<const TextColumn*>(col_desc), ncols, n, row_base, clean_text, tokens, min_token_length, nbins,
        reinterpret_cast<unsigned long long*>(hist), reinterpret_cast<unsigned long long*>(first_row), flags);
    TMOG_LAUNCH_CHECK();
    return TMOG_OK;
}

extern "C" int tmog_text_hash_scratch_bytes2(int64_t n, int ncols, const int32_t* params, int sparse_out, size_t* bytes) {
    TMOG_REQUIRE(bytes != nullptr && params != nullptr && n >= 0 && n < 0x7fffffffLL && ncols >= 0, "bad arguments");
    size_t need = 0;
    if (int rc = tmog_text_hash_scratch_bytes(n, &need)) return rc;
    TextHashParams p = {};
    p.num_features = params[0];
    p.shared = params[1];
    const int nh = params[8] > 0 ? params[8] : ncols;
    if (tc_eligible(p, ncols, nh)) need = std::max(need, tc_scratch_need(n, ncols, nh, sparse_out != 0));
    *bytes = need;
    return TMOG_OK;
}

extern "C" int tmog_text_hash_stats_dev(const int64_t* col_desc, int ncols, int64_t n, const uint8_t* prefix,
                                        const int32_t* prefix_len, const int32_t* params, const uint8_t* class_ids,
                                        uint32_t* bucket_class_sum, uint32_t* bucket_dup, uint64_t* bucket_sq_extra,
                                        uint32_t* bucket_aux, const double* label, double label_shift,
                                        double* bucket_wsum, int32_t* flags, void* scratch, size_t scratch_bytes,
                                        void* stream) {
    ThArgs a = {};
    if (int rc = text_hash_common(col_desc, ncols, n, prefix, prefix_len, params, scratch, scratch_bytes, &a)) return rc;
    TMOG_REQUIRE(bucket_class_sum && bucket_dup && bucket_sq_extra && bucket_aux && flags, "NULL output");
    TMOG_REQUIRE(a.p.nclasses == 1 || class_ids != nullptr, "class_ids is NULL");
    if (n == 0 || ncols == 0) return TMOG_OK;
    a.class_ids = class_ids;
    a.class_sum = bucket_class_sum;
    a.dup = bucket_dup;
    a.sq_extra = reinterpret_cast<unsigned long long*>(bucket_sq_extra);
    a.aux = bucket_aux;
    TMOG_REQUIRE((label == nullptr) == (bucket_wsum == nullptr), "label and bucket_wsum go together");
    a.label = label;
    a.label_shift = label_shift;
    a.wsum = bucket_wsum;
    a.flags = flags;
    a.stream = (cudaStream_t)stream;
    return text_hash_dispatch<true, false>(a);
}

// ---- packed cells (text_cells.cuh TsPack, text_packed.cuh): the fit pass keeps every cell's buckets, the transform of the
//      same rows reads them back instead of the text
static size_t pack_align(size_t b) { return (b + 255) & ~size_t(255); }

extern "C" int tmog_text_cells_pack_bytes(int64_t n, int ncols, int track_len, size_t* meta_bytes) {
    TMOG_REQUIRE(meta_bytes != nullptr && n >= 0 && n < 0x7fffffffLL && ncols >= 0, "bad arguments");
    const size_t groups = (size_t)((n + 31) / 32) * ncols, cells = (size_t)n * ncols;
    *meta_bytes = pack_align(groups * sizeof(TsGroup)) + pack_align((cells + 1) * 2) + (track_len ? pack_align(cells * 4) : 0);
    return TMOG_OK;
}

static int pack_view(void* meta, int64_t n, int ncols, bool track_len, void* arena, uint64_t arena_bytes, uint64_t* cursor,
                     int32_t* pack_flags, TsPack* pk) {
    TMOG_REQUIRE(meta && arena && cursor && pack_flags, "NULL packed-cell argument");
    TMOG_REQUIRE((reinterpret_cast<uintptr_t>(meta) & 15u) == 0 && (reinterpret_cast<uintptr_t>(arena) & 15u) == 0,
                 "packed-cell buffers must be 16-byte aligned");
    const size_t groups = (size_t)((n + 31) / 32) * ncols, cells = (size_t)n * ncols;
    uint8_t* m = static_cast<uint8_t*>(meta);
    pk->grp = reinterpret_cast<TsGroup*>(m);
    pk->rel = reinterpret_cast<unsigned short*>(m + pack_align(groups * sizeof(TsGroup)));
    pk->lens = track_len ? reinterpret_cast<unsigned*>(m + pack_align(groups * sizeof(TsGroup)) + pack_align((cells + 1) * 2))
                         : nullptr;
    pk->arena = static_cast<unsigned char*>(arena);
    pk->arena_bytes = arena_bytes;
    pk->cursor = reinterpret_cast<unsigned long long*>(cursor);
    pk->flags = pack_flags;
    return TMOG_OK;
}

extern "C" int tmog_text_hash_stats_pack_dev(const int64_t* col_desc, int ncols, int64_t n, const uint8_t* prefix,
                                             const int32_t* prefix_len, const int32_t* params, const uint8_t* class_ids,
                                             uint32_t* bucket_class_sum, uint32_t* bucket_dup, uint64_t* bucket_sq_extra,
                                             uint32_t* bucket_aux, const double* label, double label_shift,
                                             double* bucket_wsum, int32_t* flags, void* scratch, size_t scratch_bytes,
                                             void* pack_meta, void* arena, uint64_t arena_bytes, uint64_t* arena_cursor,
                                             int32_t* pack_flags, void* stream) {
    ThArgs a = {};
    if (int rc = text_hash_common(col_desc, ncols, n, prefix, prefix_len, params, scratch, scratch_bytes, &a)) return rc;
    TMOG_REQUIRE(bucket_class_sum && bucket_dup && bucket_sq_extra && bucket_aux && flags, "NULL output");
    TMOG_REQUIRE(a.p.nclasses == 1 || class_ids != nullptr, "class_ids is NULL");
    if (n == 0 || ncols == 0) return TMOG_OK;
    const char* path = getenv("TMOG_TEXT_PATH");
    TMOG_REQUIRE(tc_eligible(a.p, ncols, a.n_hashed) && !(path && strcmp(path, "cells") == 0) &&
                     a.scratch_bytes >= tc_scratch_need(n, ncols, a.n_hashed, false),
                 "packed cells need separate hash spaces of at most %d buckets per column", TC_MAXF);
    if (int rc = pack_view(pack_meta, n, ncols, a.len_base >= 0, arena, arena_bytes, arena_cursor, pack_flags, &a.pack))
        return rc;
    a.class_ids = class_ids;
    a.class_sum = bucket_class_sum;
    a.dup = bucket_dup;
    a.sq_extra = reinterpret_cast<unsigned long long*>(bucket_sq_extra);
    a.aux = bucket_aux;
    TMOG_REQUIRE((label == nullptr) == (bucket_wsum == nullptr), "label and bucket_wsum go together");
    a.label = label;
    a.label_shift = label_shift;
    a.wsum = bucket_wsum;
    a.flags = flags;
    a.stream = (cudaStream_t)stream;
    return tc_run<true, false>(a);
}

extern "C" int tmog_text_hash_transform_packed_dev(int ncols, int64_t n, const int32_t* params, void* pack_meta, void* arena,
                                                   uint64_t arena_bytes, const int32_t* newpos, int64_t* row_start,
                                                   int32_t* row_nnz, int32_t* out_idx, int32_t* out_val, uint64_t* cursor,
                                                   int64_t capacity, int32_t* flags, void* stream) {
    TMOG_REQUIRE(ncols >= 0 && n >= 0 && n < 0x7fffffffLL && params != nullptr, "bad arguments");
    TMOG_REQUIRE(newpos && row_start && row_nnz && cursor && flags, "NULL argument");
    TMOG_REQUIRE(capacity >= 0 && (capacity == 0 || (out_idx && out_val)), "NULL output");
    if (n == 0 || ncols == 0) return TMOG_OK;
    TcParams p = {};
    p.num_features = params[0];
    p.binary_freq = params[2];
    p.track_nulls = params[4];
    p.null_base = params[6];
    p.n_hashed = params[8] > 0 ? params[8] : ncols;
    p.len_base = params[9] > 0 ? params[9] : -1;
    TMOG_REQUIRE(p.num_features > 0 && p.num_features <= TC_MAXF && !params[1] && p.n_hashed <= ncols,
                 "packed cells need separate hash spaces of at most %d buckets per column", TC_MAXF);
    TsPack pk = {};
    uint64_t unused_cursor = 0;
    if (int rc = pack_view(pack_meta, n, ncols, p.len_base >= 0, arena, arena_bytes, &unused_cursor, flags, &pk)) return rc;
    pk.cursor = nullptr;
    const size_t smem = (size_t)TP_WARPS * TP_WARP_BYTES;
    TMOG_CUDA(cudaFuncSetAttribute(text_rows_packed_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int per_sm = 0;
    TMOG_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, text_rows_packed_kernel, TP_WARPS * 32, smem));
    TMOG_REQUIRE(per_sm >= 1, "packed row kernel does not fit an SM");
    const int grid = (int)std::max<int64_t>(1, std::min<int64_t>((n + TP_WARPS - 1) / TP_WARPS, (int64_t)sm_count() * per_sm));
    text_rows_packed_kernel<<<grid, TP_WARPS * 32, smem, (cudaStream_t)stream>>>(
        ncols, n, p, pk, newpos, row_start, row_nnz, out_idx, out_val, reinterpret_cast<unsigned long long*>(cursor),
        capacity, flags);
    TMOG_LAUNCH_CHECK();
    return TMOG_OK;
}

extern "C" int tmog_text_hash_transform_dev(const int64_t* col_desc, int ncols, int64_t n, const uint8_t* prefix,
                                            const int32_t* prefix_len, const int32_t* params, const int32_t* newpos,
                                            int64_t* row_start, int32_t* row_nnz, int32_t* out_idx, int32_t* out_val,
                                            uint64_t* cursor, int64_t capacity, int32_t* flags, void* scratch,
                                            size_t scratch_bytes, void* stream) {
    ThArgs a = {};
    if (int rc = text_hash_common(col_desc, ncols, n, prefix, prefix_len, params, scratch, scratch_bytes, &a)) return rc;
    TMOG_REQUIRE(newpos && row_start && row_nnz && cursor && flags, "NULL argument");
    TMOG_REQUIRE(capacity >= 0 && (capacity == 0 || (out_idx && out_val)), "NULL output");
    if (n == 0 || ncols == 0) return TMOG_OK;
    a.newpos = newpos;
    a.row_start = row_start;
    a.row_nnz = row_nnz;
    a.out_idx = out_idx;
    a.out_val = out_val;
    a.cursor = reinterpret_cast<unsigned long long*>(cursor);
    a.capacity = capacity;
    a.flags = flags;
    a.stream = (cudaStream_t)stream;
    return text_hash_dispatch<false, true>(a);
}

extern "C" int tmog_text_hash_fit_transform_dev(const int64_t* col_desc, int ncols, int64_t n, const uint8_t* prefix,
                                                const int32_t* prefix_len, const int32_t* params,
                                                const uint8_t* class_ids, uint32_t* bucket_class_sum,
                                                uint32_t* bucket_dup, uint64_t* bucket_sq_extra, uint32_t* bucket_aux,
                                                const double* label, double label_shift, double* bucket_wsum,
                                                const int32_t* newpos, int64_t* row_start, int32_t* row_nnz,
                                                int32_t* out_idx, int32_t* out_val, uint64_t* cursor, int64_t capacity,
                                                int32_t* flags, void* scratch, size_t scratch_bytes, void* stream) {
    ThArgs a = {};
    if (int rc = text_hash_common(col_desc, ncols, n, prefix, prefix_len, params, scratch, scratch_bytes, &a)) return rc;
    TMOG_REQUIRE(bucket_class_sum && bucket_dup && bucket_sq_extra && bucket_aux, "NULL statistics output");
    TMOG_REQUIRE(a.p.nclasses == 1 || class_ids != nullptr, "class_ids is NULL");
    TMOG_REQUIRE((label == nullptr) == (bucket_wsum == nullptr), "label and bucket_wsum go together");
    TMOG_REQUIRE(newpos && row_start && row_nnz && cursor && flags, "NULL argument");
    TMOG_REQUIRE(capacity >= 0 && (capacity == 0 || (out_idx && out_val)), "NULL output");
    if (n == 0 || ncols == 0) return TMOG_OK;
    a.class_ids = class_ids;
    a.class_sum = bucket_class_sum;
    a.dup = bucket_dup;
    a.sq_extra = reinterpret_cast<unsigned long long*>(bucket_sq_extra);
    a.aux = bucket_aux;
    a.label = label;
    a.label_shift = label_shift;
    a.wsum = bucket_wsum;
    a.newpos = newpos;
    a.row_start = row_start;
    a.row_nnz = row_nnz;
    a.out_idx = out_idx;
    a.out_val = out_val;
    a.cursor = reinterpret_cast<unsigned long long*>(cursor);
    a.capacity = capacity;
    a.flags = flags;
    a.stream = (cudaStream_t)stream;
    return text_hash_dispatch<true, true>(a);
}

extern "C" int tmog_sparse_rows_select_dev(const int64_t* row_start, const int32_t* row_nnz, const int32_t* idx,
                                           const int32_t* val, int64_t n, const int32_t* newpos,
                                           const int64_t* out_start, int32_t* out_nnz, int32_t* out_idx,
                                           int32_t* out_val, void* stream) {
    TMOG_REQUIRE(n >= 0 && (n == 0 || (row_start && row_nnz && newpos)), "NULL argument");
    TMOG_REQUIRE(out_start != nullptr ? (out_idx && out_val) : out_nnz != nullptr,
                 "pass out_nnz to size the rows, or out_start with out_idx / out_val to write them");
    if (n == 0) return TMOG_OK;
    const int grid = (int)std::max<int64_t>(1, std::min<int64_t>((n + 7) / 8, (int64_t)sm_count() * 16));
    if (out_start == nullptr)
        sparse_rows_select_kernel<false><<<grid, 256, 0, (cudaStream_t)stream>>>(row_start, row_nnz, idx, val, n, newpos,
                                                                                 nullptr, out_nnz, nullptr, nullptr);
    else
        sparse_rows_select_kernel<true><<<grid, 256, 0, (cudaStream_t)stream>>>(row_start, row_nnz, idx, val, n, newpos,
                                                                                out_start, out_nnz, out_idx, out_val);
    TMOG_LAUNCH_CHECK();
    return TMOG_OK;
}
