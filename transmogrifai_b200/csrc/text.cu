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
// One CTA (4 warps) works on one row at a time: warps take columns, lanes take tokens; the row's bucket indices are
// collected in shared memory, bitonic-sorted and run-length encoded.  This path is instruction-bound, not HBM-bound
// (~100 instructions per token); see DESIGN.md.
#include <string.h>

#include <algorithm>
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

__host__ __device__ inline unsigned stop_hash(const unsigned char* s, int len) {
    unsigned h = 2166136261u;
    for (int i = 0; i < len; ++i) h = (h ^ s[i]) * 16777619u;
    return h;
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
text_stats_kernel(const TextColumn* __restrict__ cols, int ncols, int64_t n, const uint8_t* __restrict__ class_ids,
                  int nclasses, int clean, int slots, int max_distinct, const TextStatsTable* __restrict__ tables,
                  int* __restrict__ flags) {
    const int c = blockIdx.y;
    const TextColumn tc = cols[c];
    const TextStatsTable t = tables[c];
    unsigned long long lsum = 0, lsq = 0, nn = 0;
    bool non_ascii = false;
    for (int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; r < n; r += (int64_t)gridDim.x * blockDim.x) {
        if (!cell_valid(tc, r)) continue;
        const int64_t a = tc.offsets[r], b = tc.offsets[r + 1];
        unsigned long long h;
        int m;
        clean_hash_cell(tc.data + a, b - a, clean != 0, h, m, non_ascii);
        ts_insert(t, slots, max_distinct, h, r, nclasses > 1 ? class_ids[r] : 0, nclasses);
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
                 int slots, const TextStatsTable* __restrict__ tables, int32_t* __restrict__ codes, int64_t ld) {
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
constexpr int TH_THREADS = 128;
constexpr int TH_WARPS = TH_THREADS / 32;
constexpr int TH_TOK_CAP = 512;    // tokens per cell handled per warp
constexpr int TH_LIST_CAP = 4096;  // bucket hits per row
constexpr int TH_PREFIX = 16;      // bytes reserved for "<nameHash>_"

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
};

// flags: [0] non-ASCII byte met, [1] a cell had more than TH_TOK_CAP tokens, [2] a row had more than TH_LIST_CAP hits,
//        [3] sparse output capacity exceeded
template <bool SPARSE_OUT>
__global__ void __launch_bounds__(TH_THREADS)
text_hash_kernel(const TextColumn* __restrict__ cols, int ncols, int64_t n, const uint8_t* __restrict__ prefix,
                 const int* __restrict__ prefix_len, TextHashParams p, const uint8_t* __restrict__ class_ids,
                 // statistics mode
                 unsigned int* __restrict__ bucket_class_sum, unsigned int* __restrict__ bucket_nnz,
                 unsigned long long* __restrict__ bucket_sq_extra, unsigned int* __restrict__ bucket_aux,
                 // sparse-output mode
                 const int32_t* __restrict__ newpos, int64_t* __restrict__ row_start, int32_t* __restrict__ row_nnz,
                 int32_t* __restrict__ out_idx, int32_t* __restrict__ out_val, unsigned long long* __restrict__ cursor,
                 int64_t capacity, int* __restrict__ flags) {
    __shared__ int list[TH_LIST_CAP];
    __shared__ int tok_start[TH_WARPS][TH_TOK_CAP];
    __shared__ int tok_end[TH_WARPS][TH_TOK_CAP];
    __shared__ int list_len;
    __shared__ unsigned long long out_base;
    __shared__ int scan_tmp[TH_WARPS];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const unsigned lt_mask = (1u << lane) - 1u;

    for (int64_t row = blockIdx.x; row < n; row += gridDim.x) {
        if (tid == 0) list_len = 0;
        __syncthreads();
        for (int j = warp; j < ncols; j += TH_WARPS) {
            const TextColumn tc = cols[j];
            int emitted = 0;
            if (cell_valid(tc, row)) {
                const int64_t a = tc.offsets[row];
                const int len = (int)(tc.offsets[row + 1] - a);
                const uint8_t* s = tc.data + a;
                // ---- pass 1: token boundaries, 32 bytes per step
                int ntok = 0, nend = 0;
                unsigned carry_t = 0;                 // previous byte was a token character
                bool carry_l = false, carry_d = false;  // previous byte was a letter / digit
                for (int base = 0; base < len; base += 32) {
                    const int pos = base + lane;
                    const unsigned b = pos < len ? s[pos] : ' ';
                    const unsigned nb = pos + 1 < len ? s[pos + 1] : ' ';
                    if (b >= 0x80u) flags[0] = 1;
                    const bool L = is_letter(b), D = is_digit(b);
                    bool pl = __shfl_up_sync(FULL, L, 1), pd = __shfl_up_sync(FULL, D, 1);
                    if (lane == 0) { pl = carry_l; pd = carry_d; }
                    const bool nl = is_letter(nb), nd = is_digit(nb);
                    const bool mid_letter = b == ':' || b == '.' || b == '\'';   // MidLetter | MidNumLet | SingleQuote
                    const bool mid_num = b == ',' || b == ';' || b == '.' || b == '\'';  // MidNum | MidNumLet | SQ
                    const bool tokc = L || D || b == '_' || (mid_letter && pl && nl) || (mid_num && pd && nd);
                    const unsigned T = __ballot_sync(FULL, tokc);
                    const unsigned prevT = (T << 1) | carry_t;
                    const unsigned S = T & ~prevT, E = ~T & prevT;
                    if ((S >> lane) & 1u) {
                        const int ord = ntok + __popc(S & lt_mask);
                        if (ord < TH_TOK_CAP) tok_start[warp][ord] = pos;
                    }
                    if ((E >> lane) & 1u) {
                        const int ord = nend + __popc(E & lt_mask);
                        if (ord < TH_TOK_CAP) tok_end[warp][ord] = pos;
                    }
                    ntok += __popc(S);
                    nend += __popc(E);
                    carry_t = T >> 31;
                    carry_l = __shfl_sync(FULL, L, 31);
                    carry_d = __shfl_sync(FULL, D, 31);
                }
                if (carry_t && lane == 0 && nend < TH_TOK_CAP) tok_end[warp][nend] = len;
                if (ntok > TH_TOK_CAP) {
                    if (lane == 0) flags[1] = 1;
                    ntok = TH_TOK_CAP;
                }
                __syncwarp();
                // ---- pass 2: one token per lane: lower-case, filters, murmur3 of prefix + token
                const int plen = p.prepend ? prefix_len[j] : 0;
                const uint8_t* pre = prefix + j * TH_PREFIX;
                const int space_off = p.shared ? 0 : j * p.num_features;
                for (int k0 = 0; k0 < ntok; k0 += 32) {
                    const int k = k0 + lane;
                    int hits = 0;
                    if (k < ntok) {
                        const int ts = tok_start[warp][k], te = tok_end[warp][k];
                        // tokens longer than 255 characters are chopped into 255-character pieces (Lucene >= 6.5)
                        for (int ps = ts; ps < te; ps += 255) {
                            const int pe = min(te, ps + 255);
                            const int tl = pe - ps;
                            bool alnum = false;
                            unsigned char low[STOP_MAXLEN];
                            Murmur mm;
                            mm.init();
                            for (int q = 0; q < plen; ++q) mm.push(pre[q]);
                            for (int q = 0; q < tl; ++q) {
                                const unsigned c = to_lower(s[ps + q]);
                                alnum |= is_letter(c) || is_digit(c);
                                if (q < STOP_MAXLEN) low[q] = (unsigned char)c;
                                mm.push(c);
                            }
                            if (!alnum || tl < p.min_token_length) continue;
                            bool stop = false;
                            if (tl <= STOP_MAXLEN) {
                                unsigned sl = stop_hash(low, tl) & (STOP_SLOTS - 1);
                                for (int probe = 0; probe < STOP_SLOTS; ++probe) {
                                    const int el = c_stop[sl].len;
                                    if (el == 0) break;
                                    if (el == tl) {
                                        bool eq = true;
                                        for (int q = 0; q < tl; ++q) eq &= c_stop[sl].w[q] == (char)low[q];
                                        if (eq) { stop = true; break; }
                                    }
                                    sl = (sl + 1) & (STOP_SLOTS - 1);
                                }
                            }
                            if (stop) continue;
                            const int idx = space_off + non_negative_mod(mm.finish(), p.num_features);
                            const int at = atomicAdd(&list_len, 1);
                            if (at < TH_LIST_CAP) list[at] = idx;
                            ++hits;
                        }
                    }
                    emitted += __reduce_add_sync(FULL, hits);
                }
            }
            // null indicator of this text column: no token survived (null, empty or all stop words)
            if (p.track_nulls && emitted == 0 && lane == 0) {
                const int at = atomicAdd(&list_len, 1);
                if (at < TH_LIST_CAP) list[at] = p.null_base + j;
            }
        }
        __syncthreads();
        int len = list_len;
        if (len > TH_LIST_CAP) {
            if (tid == 0) flags[2] = 1;
            len = TH_LIST_CAP;
        }
        // ---- sort the row's hits (bitonic, padded to a power of two with INT_MAX)
        int m = 32;
        while (m < len) m <<= 1;
        for (int i = len + tid; i < m; i += TH_THREADS) list[i] = 0x7fffffff;
        __syncthreads();
        for (int k2 = 2; k2 <= m; k2 <<= 1) {
            for (int jj = k2 >> 1; jj > 0; jj >>= 1) {
                for (int i = tid; i < m; i += TH_THREADS) {
                    const int ixj = i ^ jj;
                    if (ixj > i) {
                        const int x = list[i], y = list[ixj];
                        const bool up = (i & k2) == 0;
                        if ((x > y) == up) {
                            list[i] = y;
                            list[ixj] = x;
                        }
                    }
                }
                __syncthreads();
            }
        }
        // ---- run-length encode: a thread owning a run head emits (index, count)
        const int cls = (p.nclasses > 1 && class_ids) ? class_ids[row] : 0;
        if (!SPARSE_OUT) {
            for (int i = tid; i < len; i += TH_THREADS) {
                const int v = list[i];
                if (i == 0 || list[i - 1] != v) {
                    int e = i + 1;
                    while (e < len && list[e] == v) ++e;
                    const unsigned cnt = p.binary_freq && v < p.hash_width ? 1u : (unsigned)(e - i);
                    atomicAdd(&bucket_class_sum[(int64_t)v * p.nclasses + cls], cnt);
                    atomicAdd(&bucket_nnz[v], 1u);
                    if (cnt > 1u) {   // rare: everything beyond "present once" goes to the auxiliary arrays
                        atomicAdd(&bucket_sq_extra[v], (unsigned long long)cnt * cnt - cnt);
                        atomicMax(&bucket_aux[3 * (int64_t)v], cnt);          // largest frequency
                        atomicMin(&bucket_aux[3 * (int64_t)v + 1], ~cnt);     // smallest frequency > 1, stored negated
                        atomicAdd(&bucket_aux[3 * (int64_t)v + 2], 1u);       // rows with a frequency > 1
                    }
                }
            }
        } else {
            // heads per thread -> block exclusive scan -> packed output (kept columns only, renumbered)
            // contiguous chunk of the sorted list per thread, so that the packed output keeps the order
            const int per = (len + TH_THREADS - 1) / TH_THREADS;
            const int lo = min(len, tid * per), hi = min(len, lo + per);
            int heads = 0;
            for (int i = lo; i < hi; ++i) {
                const int v = list[i];
                if ((i == 0 || list[i - 1] != v) && newpos[v] >= 0) ++heads;
            }
            int incl = heads;
            for (int o = 1; o < 32; o <<= 1) {
                const int t2 = __shfl_up_sync(FULL, incl, o);
                if (lane >= o) incl += t2;
            }
            if (lane == 31) scan_tmp[warp] = incl;
            __syncthreads();
            int warp_off = 0, total = 0;
            for (int w = 0; w < TH_WARPS; ++w) {
                if (w < warp) warp_off += scan_tmp[w];
                total += scan_tmp[w];
            }
            if (tid == 0) {
                out_base = atomicAdd(cursor, (unsigned long long)total);
                row_start[row] = (int64_t)out_base;
                row_nnz[row] = total;
                if ((int64_t)out_base + total > capacity) flags[3] = 1;
            }
            __syncthreads();
            int64_t w0 = (int64_t)out_base + warp_off + incl - heads;
            for (int i = lo; i < hi; ++i) {
                const int v = list[i];
                if ((i == 0 || list[i - 1] != v) && newpos[v] >= 0) {
                    int e = i + 1;
                    while (e < len && list[e] == v) ++e;
                    if (w0 < capacity) {
                        out_idx[w0] = newpos[v];
                        out_val[w0] = p.binary_freq && v < p.hash_width ? 1 : e - i;
                    }
                    ++w0;
                }
            }
        }
        __syncthreads();
    }
}

}  // namespace tmog

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
    for (int i = 0; i < nwords; ++i) {
        const int len = (int)strlen(kStopWords[i]);
        unsigned s = stop_hash((const unsigned char*)kStopWords[i], len) & (STOP_SLOTS - 1);
        while (table[s].len != 0) s = (s + 1) & (STOP_SLOTS - 1);
        table[s].len = (unsigned char)len;
        memcpy(table[s].w, kStopWords[i], len);
    }
    TMOG_CUDA(cudaMemcpyToSymbol(c_stop, table, sizeof(table)));
    if (dev >= 0 && dev < 64) done[dev] = true;
    return TMOG_OK;
}

extern "C" int tmog_text_stats_table_bytes(int slots, int nclasses, size_t* bytes) {
    TMOG_REQUIRE(bytes != nullptr && slots >= 64 && (slots & (slots - 1)) == 0 && nclasses >= 1, "bad arguments");
    // hash [slots] u64 | first_row [slots] u64 | counts [slots*nclasses] u32 (padded to 8) | ndistinct+pad | len_sum[2] | nonnull
    *bytes = sizeof(uint64_t) * (size_t)slots * 2 + (((size_t)slots * nclasses * 4 + 7) & ~size_t(7)) + 8 + 16 + 8;
    return TMOG_OK;
}

extern "C" int tmog_text_stats_dev(const int64_t* col_desc, int ncols, int64_t n, const uint8_t* class_ids, int nclasses,
                                   int clean_text, int slots, int max_distinct, void* tables, int32_t* flags,
                                   void* stream) {
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
    const int gx = (int)std::min<int64_t>((n + 255) / 256, (int64_t)sm_count() * 4);
    text_stats_kernel<<<dim3(gx, ncols), 256, 0, st>>>(reinterpret_cast<const TextColumn*>(col_desc), ncols, n, class_ids,
                                                       nclasses, clean_text, slots, max_distinct,
                                                       static_cast<const TextStatsTable*>(tables), flags);
    TMOG_LAUNCH_CHECK();
    return TMOG_OK;
}

extern "C" int tmog_text_stats_layout(int ncols, int slots, int nclasses, size_t* head_bytes, size_t* per_column_bytes) {
    TMOG_REQUIRE(head_bytes && per_column_bytes, "NULL argument");
    *head_bytes = (sizeof(TextStatsTable) * (size_t)ncols + 255) & ~size_t(255);
    return tmog_text_stats_table_bytes(slots, nclasses, per_column_bytes);
}

extern "C" int tmog_text_codes_dev(const int64_t* col_desc, const int32_t* col_index, int nsel, int64_t n,
                                   int clean_text, int slots, const void* tables, int32_t* codes, int64_t ld,
                                   void* stream) {
    TMOG_REQUIRE(nsel >= 0 && n >= 0 && ld >= n, "bad shape");
    if (nsel == 0 || n == 0) return TMOG_OK;
    TMOG_REQUIRE(col_desc && col_index && tables && codes, "NULL argument");
    const int gx = (int)std::min<int64_t>((n + 255) / 256, (int64_t)sm_count() * 4);
    text_code_kernel<<<dim3(gx, nsel), 256, 0, (cudaStream_t)stream>>>(
        reinterpret_cast<const TextColumn*>(col_desc), col_index, nsel, n, clean_text, slots,
        static_cast<const TextStatsTable*>(tables), codes, ld);
    TMOG_LAUNCH_CHECK();
    return TMOG_OK;
}

static int text_hash_common(const int64_t* col_desc, int ncols, int64_t n, const uint8_t* prefix,
                            const int32_t* prefix_len, const int32_t* params, TextHashParams* p) {
    TMOG_REQUIRE(ncols >= 0 && n >= 0, "bad shape");
    TMOG_REQUIRE(col_desc && prefix && prefix_len && params, "NULL argument");
    p->num_features = params[0];
    p->shared = params[1];
    p->binary_freq = params[2];
    p->min_token_length = params[3];
    p->track_nulls = params[4];
    p->prepend = params[5];
    p->null_base = params[6];
    p->nclasses = params[7];
    TMOG_REQUIRE(p->num_features > 0 && p->num_features <= 10000000,
                 "requirement failed: numFeatures must be in (0, 10000000], got %d", p->num_features);
    TMOG_REQUIRE(p->nclasses >= 1 && p->nclasses <= 255, "nclasses must be in [1, 255]");
    const int64_t hw = p->shared ? p->num_features : (int64_t)p->num_features * ncols;
    TMOG_REQUIRE(hw + ncols < 0x7fffffffLL, "hash block too wide");
    p->hash_width = (int)hw;
    TMOG_REQUIRE(p->null_base >= p->hash_width, "null_base must not overlap the hash block");
    return upload_stop_words();
}

extern "C" int tmog_text_hash_stats_dev(const int64_t* col_desc, int ncols, int64_t n, const uint8_t* prefix,
                                        const int32_t* prefix_len, const int32_t* params, const uint8_t* class_ids,
                                        uint32_t* bucket_class_sum, uint32_t* bucket_nnz, uint64_t* bucket_sq_extra,
                                        uint32_t* bucket_aux, int32_t* flags, void* stream) {
    TextHashParams p;
    if (int rc = text_hash_common(col_desc, ncols, n, prefix, prefix_len, params, &p)) return rc;
    TMOG_REQUIRE(bucket_class_sum && bucket_nnz && bucket_sq_extra && bucket_aux && flags, "NULL output");
    TMOG_REQUIRE(p.nclasses == 1 || class_ids != nullptr, "class_ids is NULL");
    if (n == 0 || ncols == 0) return TMOG_OK;
    const int grid = (int)std::min<int64_t>(n, (int64_t)sm_count() * 8);
    text_hash_kernel<false><<<grid, TH_THREADS, 0, (cudaStream_t)stream>>>(
        reinterpret_cast<const TextColumn*>(col_desc), ncols, n, prefix, prefix_len, p, class_ids, bucket_class_sum,
        bucket_nnz, reinterpret_cast<unsigned long long*>(bucket_sq_extra), bucket_aux, nullptr, nullptr, nullptr, nullptr,
        nullptr, nullptr, 0, flags);
    TMOG_LAUNCH_CHECK();
    return TMOG_OK;
}

extern "C" int tmog_text_hash_transform_dev(const int64_t* col_desc, int ncols, int64_t n, const uint8_t* prefix,
                                            const int32_t* prefix_len, const int32_t* params, const int32_t* newpos,
                                            int64_t* row_start, int32_t* row_nnz, int32_t* out_idx, int32_t* out_val,
                                            uint64_t* cursor, int64_t capacity, int32_t* flags, void* stream) {
    TextHashParams p;
    if (int rc = text_hash_common(col_desc, ncols, n, prefix, prefix_len, params, &p)) return rc;
    TMOG_REQUIRE(newpos && row_start && row_nnz && cursor && flags, "NULL argument");
    TMOG_REQUIRE(capacity >= 0 && (capacity == 0 || (out_idx && out_val)), "NULL output");
    if (n == 0 || ncols == 0) return TMOG_OK;
    const int grid = (int)std::min<int64_t>(n, (int64_t)sm_count() * 8);
    text_hash_kernel<true><<<grid, TH_THREADS, 0, (cudaStream_t)stream>>>(
        reinterpret_cast<const TextColumn*>(col_desc), ncols, n, prefix, prefix_len, p, nullptr, nullptr, nullptr,
        nullptr, nullptr, newpos, row_start, row_nnz, out_idx, out_val, reinterpret_cast<unsigned long long*>(cursor), capacity,
        flags);
    TMOG_LAUNCH_CHECK();
    return TMOG_OK;
}
