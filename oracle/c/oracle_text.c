/*
 * oracle_text.c -- CPU restatement (plain C + OpenMP) of the reference's string paths.  TEST INFRASTRUCTURE ONLY (see
 * oracle_hot.c): checked against the Python oracle in tests/test_oracle_c_port.py, timed by bench.py as the CPU
 * baseline of the text / pick-list workloads ("port": the reference is JVM code and cannot run here).
 *
 * Follows (paths under /root/reference):
 *   utils/src/main/scala/com/salesforce/op/utils/text/TextUtils.scala:39-47                      orc_clean_string
 *   core/src/main/scala/com/salesforce/op/stages/impl/feature/TextTokenizer.scala:160-220,
 *   core/.../utils/text/LuceneTextAnalyzer.scala:103-122,159-166 (Lucene 7.3.0 StandardAnalyzer + Snowball
 *       english stop set, restated for ASCII)                                                    tokenizer below
 *   core/.../impl/feature/OPCollectionHashingVectorizer.scala:244-305 + Spark 2.4.5 HashingTF
 *       (legacy Murmur3_x86_32.hashUnsafeBytes, seed 42, non-negative mod)                        orc_text_hash
 *   core/.../impl/feature/SmartTextVectorizer.scala:207-299 (TextStats)                           orc_text_stats
 *   core/.../impl/feature/OpOneHotVectorizer.scala:75-124,416-437 on dictionary codes            orc_pivot_counts
 *   core/.../impl/preparators/SanityChecker.scala:265-272,397-407 (column sums per label, colStats sums)
 *
 * Rows are split into contiguous partitions, one per OpenMP thread (Spark local[N]); partial sums are merged in
 * partition order.
 */
#include <omp.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

static inline int tbit(const uint64_t* w, int64_t r) { return w == NULL ? 1 : (int)((w[r >> 6] >> (r & 63)) & 1u); }
static inline int is_l(unsigned c) { return (c | 0x20u) - 'a' < 26u; }
static inline int is_d(unsigned c) { return c - '0' < 10u; }
static inline int is_punct(unsigned c) {
    return (c - 0x21u) < 15u || (c - 0x3Au) < 7u || (c - 0x5Bu) < 6u || (c - 0x7Bu) < 4u;
}
static inline unsigned lower(unsigned c) { return (c - 'A') < 26u ? c + 32u : c; }

/* ---- murmur3 (Spark's legacy hashUnsafeBytes: each tail byte is mixed as its own sign-extended block) ------------- */
static inline uint32_t rotl(uint32_t x, int r) { return (x << r) | (x >> (32 - r)); }
static inline uint32_t mix_k1(uint32_t k) { k *= 0xcc9e2d51u; k = rotl(k, 15); return k * 0x1b873593u; }
static inline uint32_t mix_h1(uint32_t h, uint32_t k) { h ^= k; h = rotl(h, 13); return h * 5u + 0xe6546b64u; }
static uint32_t murmur3_legacy(const uint8_t* p, int len, uint32_t seed) {
    uint32_t h = seed;
    int i = 0;
    for (; i + 4 <= len; i += 4) {
        uint32_t k = (uint32_t)p[i] | ((uint32_t)p[i + 1] << 8) | ((uint32_t)p[i + 2] << 16) | ((uint32_t)p[i + 3] << 24);
        h = mix_h1(h, mix_k1(k));
    }
    for (; i < len; ++i) h = mix_h1(h, mix_k1((uint32_t)(int32_t)(int8_t)p[i]));
    h ^= (uint32_t)len;
    h ^= h >> 16; h *= 0x85ebca6bu; h ^= h >> 13; h *= 0xc2b2ae35u; h ^= h >> 16;
    return h;
}
static inline int nonneg_mod(uint32_t h, int n) { int r = (int32_t)h % n; return r < 0 ? r + n : r; }

/* ---- Snowball english stop words (Lucene StandardAnalyzer(englishStopwords) as the reference loads them) --------- */
static const char* const STOP[] = {
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
#define NSTOP ((int)(sizeof(STOP) / sizeof(STOP[0])))
#define STOP_SLOTS 1024
static int16_t stop_table[STOP_SLOTS];
static int stop_ready = 0;
static uint32_t fnv(const uint8_t* p, int len) {
    uint32_t h = 2166136261u;
    for (int i = 0; i < len; ++i) h = (h ^ p[i]) * 16777619u;
    return h;
}
static void stop_init(void) {
#pragma omp critical(orc_stop_init)
    if (!stop_ready) {
        for (int i = 0; i < STOP_SLOTS; ++i) stop_table[i] = -1;
        for (int i = 0; i < NSTOP; ++i) {
            uint32_t s = fnv((const uint8_t*)STOP[i], (int)strlen(STOP[i])) & (STOP_SLOTS - 1);
            while (stop_table[s] >= 0) s = (s + 1) & (STOP_SLOTS - 1);
            stop_table[s] = (int16_t)i;
        }
        stop_ready = 1;
    }
}
static int is_stop(const uint8_t* p, int len) {
    if (len > 10) return 0;
    uint32_t s = fnv(p, len) & (STOP_SLOTS - 1);
    while (stop_table[s] >= 0) {
        const char* w = STOP[stop_table[s]];
        if ((int)strlen(w) == len && memcmp(w, p, (size_t)len) == 0) return 1;
        s = (s + 1) & (STOP_SLOTS - 1);
    }
    return 0;
}

/* ---- per-row term frequencies --------------------------------------------------------------------------------- */
typedef struct {
    int32_t* key;   /* -1 = empty */
    int32_t* cnt;
    int32_t* used;  /* occupied slots, for the walk and the reset */
    int cap, nused;
} RowMap;

static void rowmap_reserve(RowMap* m, int want) {
    int cap = m->cap ? m->cap : 256;
    while (cap < 2 * want + 2) cap <<= 1;
    if (cap == m->cap) return;
    free(m->key); free(m->cnt); free(m->used);
    m->key = (int32_t*)malloc(sizeof(int32_t) * (size_t)cap);
    m->cnt = (int32_t*)malloc(sizeof(int32_t) * (size_t)cap);
    m->used = (int32_t*)malloc(sizeof(int32_t) * (size_t)cap);
    for (int i = 0; i < cap; ++i) m->key[i] = -1;
    m->cap = cap;
    m->nused = 0;
}
static void rowmap_add(RowMap* m, int32_t key) {
    uint32_t s = ((uint32_t)key * 0x9E3779B1u) & (uint32_t)(m->cap - 1);
    for (;;) {
        if (m->key[s] < 0) {
            m->key[s] = key;
            m->cnt[s] = 1;
            m->used[m->nused++] = (int32_t)s;
            return;
        }
        if (m->key[s] == key) {
            ++m->cnt[s];
            return;
        }
        s = (s + 1) & (uint32_t)(m->cap - 1);
    }
}
static int cmp_i32(const void* a, const void* b) {
    const int32_t x = *(const int32_t*)a, y = *(const int32_t*)b;
    return (x > y) - (x < y);
}

/* Tokens of one cell -> bucket hits in `m`.  Returns the number of surviving tokens.  `low` is scratch of >= len bytes;
 * `tok` scratch of >= 16 + 255 bytes holding the "<nameHash>_" prefix in its first plen bytes. */
static int hash_cell(const uint8_t* s, int64_t len, uint8_t* low, uint8_t* tok, int plen, int space_off,
                     int num_features, int min_token_length, RowMap* m) {
    for (int64_t i = 0; i < len; ++i) low[i] = (uint8_t)lower(s[i]);
    int emitted = 0;
    int64_t i = 0;
    while (i < len) {
        unsigned c = low[i];
        if (!(is_l(c) || is_d(c) || c == '_')) {
            ++i;
            continue;
        }
        int64_t j = i;
        int has_alnum = 0;
        while (j < len) {
            c = low[j];
            if (is_l(c) || is_d(c)) {
                has_alnum = 1;
                ++j;
            } else if (c == '_') {
                ++j;
            } else if (j > i && j + 1 < len &&
                       (((c == ':' || c == '.' || c == '\'') && is_l(low[j - 1]) && is_l(low[j + 1])) ||
                        ((c == ',' || c == ';' || c == '.' || c == '\'') && is_d(low[j - 1]) && is_d(low[j + 1])))) {
                ++j;
            } else {
                break;
            }
        }
        if (has_alnum) {
            for (int64_t k = i; k < j; k += 255) {          /* maxTokenLength pieces */
                const int plen_tok = (int)((j - k) < 255 ? (j - k) : 255);
                if (plen_tok < min_token_length || is_stop(low + k, plen_tok)) continue;
                memcpy(tok + plen, low + k, (size_t)plen_tok);
                const uint32_t h = murmur3_legacy(tok, plen + plen_tok, 42u);
                rowmap_add(m, space_off + nonneg_mod(h, num_features));
                ++emitted;
            }
        }
        i = j > i ? j : i + 1;
    }
    return emitted;
}

/*
 * SmartTextVectorizer's hashed columns for all rows: tokenise, hash, term frequencies, null indicators (index
 * null_base + j when no token of column j survived).  Column j's data / offsets / validity are given as addresses
 * (data_ptr[j] -> uint8 bytes, offs_ptr[j] -> int64 [n + 1], valid_ptr[j] -> uint64 bitmap words or 0).
 *   statistics (class_sum != NULL), arrays of width = null_base + ncols, ADDED to:
 *       class_sum [width * nclasses] sum of tf per label class; sq_sum [width] sum tf^2; nnz [width] rows holding the
 *       index; vmax [width] max tf; vmin [width] min tf over the rows holding it (caller initialises to INT32_MAX)
 *   rows (want_rows): row_start / row_nnz [n], and *out_idx / *out_val malloc'ed here (orc_free them), indices
 *       ascending per row, *total entries.
 * Returns 0, or -1 if memory ran out.
 */
int orc_text_hash(const uintptr_t* data_ptr, const uintptr_t* offs_ptr, const uintptr_t* valid_ptr, int ncols,
                  int64_t n, const uint8_t* prefix, const int32_t* prefix_len, int num_features, int shared,
                  int binary_freq, int min_token_length, int track_nulls, int null_base, const uint8_t* class_ids,
                  int nclasses, int threads, int64_t* class_sum, int64_t* sq_sum, int64_t* nnz, int32_t* vmax,
                  int32_t* vmin, int want_rows, int64_t* row_start, int32_t* row_nnz, int32_t** out_idx,
                  int32_t** out_val, int64_t* total) {
    stop_init();
    const int width = null_base + ncols;
    const int hash_width = shared ? num_features : num_features * ncols;
    if (threads < 1) threads = 1;
    int failed = 0;
    int32_t** t_idx = (int32_t**)calloc((size_t)threads, sizeof(int32_t*));
    int32_t** t_val = (int32_t**)calloc((size_t)threads, sizeof(int32_t*));
    int64_t* t_cnt = (int64_t*)calloc((size_t)threads, sizeof(int64_t));
    int64_t** t_stats = (int64_t**)calloc((size_t)threads, sizeof(int64_t*));
#pragma omp parallel for schedule(static, 1) num_threads(threads)
    for (int p = 0; p < threads; ++p) {
        const int64_t lo = n * p / threads, hi = n * (p + 1) / threads;
        RowMap m = {0};
        uint8_t tok[16 + 256];
        int64_t low_cap = 1024;
        uint8_t* low = (uint8_t*)malloc((size_t)low_cap);
        int64_t cap = want_rows ? 1 << 16 : 0, cnt = 0;
        int32_t* idx = want_rows ? (int32_t*)malloc(sizeof(int32_t) * (size_t)cap) : NULL;
        int32_t* val = want_rows ? (int32_t*)malloc(sizeof(int32_t) * (size_t)cap) : NULL;
        /* private sums: [class_sum | sq_sum | nnz | vmax | vmin] */
        int64_t* st = NULL;
        if (class_sum) {
            st = (int64_t*)calloc((size_t)width * (size_t)(nclasses + 4), sizeof(int64_t));
            if (st) for (int i = 0; i < width; ++i) st[(size_t)width * (nclasses + 3) + i] = INT32_MAX;
        }
        int32_t* order = NULL;
        int order_cap = 0;
        int ok = low && (!want_rows || (idx && val)) && (!class_sum || st);
        for (int64_t r = lo; r < hi && ok; ++r) {
            int64_t bound = ncols;
            for (int j = 0; j < ncols; ++j) {
                const int64_t* off = (const int64_t*)offs_ptr[j];
                if (tbit((const uint64_t*)valid_ptr[j], r)) bound += (off[r + 1] - off[r] + 1) / 2;
            }
            rowmap_reserve(&m, (int)bound);
            for (int j = 0; j < ncols; ++j) {
                const int64_t* off = (const int64_t*)offs_ptr[j];
                int emitted = 0;
                if (tbit((const uint64_t*)valid_ptr[j], r)) {
                    const int64_t len = off[r + 1] - off[r];
                    if (len > low_cap) {
                        low_cap = len * 2;
                        free(low);
                        low = (uint8_t*)malloc((size_t)low_cap);
                        if (!low) { ok = 0; break; }
                    }
                    const int plen = prefix_len ? prefix_len[j] : 0;
                    if (plen) memcpy(tok, prefix + 16 * j, (size_t)plen);
                    emitted = hash_cell((const uint8_t*)data_ptr[j] + off[r], len, low, tok, plen,
                                        shared ? 0 : j * num_features, num_features, min_token_length, &m);
                }
                if (track_nulls && emitted == 0) rowmap_add(&m, null_base + j);
            }
            if (!ok) break;
            /* the row's entries in index order */
            if (m.nused > order_cap) {
                order_cap = m.nused * 2;
                free(order);
                order = (int32_t*)malloc(sizeof(int32_t) * (size_t)order_cap);
                if (!order) { ok = 0; break; }
            }
            for (int i = 0; i < m.nused; ++i) order[i] = m.key[m.used[i]];
            qsort(order, (size_t)m.nused, sizeof(int32_t), cmp_i32);
            if (want_rows) {
                if (cnt + m.nused > cap) {
                    while (cnt + m.nused > cap) cap *= 2;
                    idx = (int32_t*)realloc(idx, sizeof(int32_t) * (size_t)cap);
                    val = (int32_t*)realloc(val, sizeof(int32_t) * (size_t)cap);
                    if (!idx || !val) { ok = 0; break; }
                }
                row_start[r] = cnt;           /* thread-local for now */
                row_nnz[r] = m.nused;
            }
            const int cls = nclasses > 1 ? class_ids[r] : 0;
            for (int i = 0; i < m.nused; ++i) {
                const int32_t key = order[i];
                uint32_t s = ((uint32_t)key * 0x9E3779B1u) & (uint32_t)(m.cap - 1);
                while (m.key[s] != key) s = (s + 1) & (uint32_t)(m.cap - 1);
                const int32_t tf = (binary_freq && key < hash_width) ? 1 : m.cnt[s];
                if (want_rows) {
                    idx[cnt] = key;
                    val[cnt] = tf;
                    ++cnt;
                }
                if (st) {
                    st[(size_t)key * nclasses + cls] += tf;
                    st[(size_t)width * nclasses + key] += (int64_t)tf * tf;
                    st[(size_t)width * (nclasses + 1) + key] += 1;
                    if (tf > st[(size_t)width * (nclasses + 2) + key]) st[(size_t)width * (nclasses + 2) + key] = tf;
                    if (tf < st[(size_t)width * (nclasses + 3) + key]) st[(size_t)width * (nclasses + 3) + key] = tf;
                }
            }
            for (int i = 0; i < m.nused; ++i) m.key[m.used[i]] = -1;
            m.nused = 0;
        }
        if (!ok) {
#pragma omp atomic write
            failed = 1;
        }
        t_idx[p] = idx;
        t_val[p] = val;
        t_cnt[p] = cnt;
        t_stats[p] = st;
        free(low);
        free(order);
        free(m.key); free(m.cnt); free(m.used);
    }
    /* merge in partition order */
    if (!failed && class_sum) {
        for (int p = 0; p < threads; ++p) {
            const int64_t* st = t_stats[p];
            if (!st) continue;
            for (size_t i = 0; i < (size_t)width * nclasses; ++i) class_sum[i] += st[i];
            for (int i = 0; i < width; ++i) {
                sq_sum[i] += st[(size_t)width * nclasses + i];
                nnz[i] += st[(size_t)width * (nclasses + 1) + i];
                if (st[(size_t)width * (nclasses + 2) + i] > vmax[i]) vmax[i] = (int32_t)st[(size_t)width * (nclasses + 2) + i];
                if (st[(size_t)width * (nclasses + 3) + i] < vmin[i]) vmin[i] = (int32_t)st[(size_t)width * (nclasses + 3) + i];
            }
        }
    }
    if (!failed && want_rows) {
        int64_t sum = 0;
        for (int p = 0; p < threads; ++p) sum += t_cnt[p];
        int32_t* oi = (int32_t*)malloc(sizeof(int32_t) * (size_t)(sum ? sum : 1));
        int32_t* ov = (int32_t*)malloc(sizeof(int32_t) * (size_t)(sum ? sum : 1));
        if (!oi || !ov) {
            failed = 1;
            free(oi); free(ov);
        } else {
            int64_t base = 0;
            for (int p = 0; p < threads; ++p) {
                if (t_cnt[p]) {
                    memcpy(oi + base, t_idx[p], sizeof(int32_t) * (size_t)t_cnt[p]);
                    memcpy(ov + base, t_val[p], sizeof(int32_t) * (size_t)t_cnt[p]);
                }
                const int64_t lo = n * p / threads, hi = n * (p + 1) / threads;
                for (int64_t r = lo; r < hi; ++r) row_start[r] += base;
                base += t_cnt[p];
            }
            *out_idx = oi;
            *out_val = ov;
            *total = sum;
        }
    }
    for (int p = 0; p < threads; ++p) { free(t_idx[p]); free(t_val[p]); free(t_stats[p]); }
    free(t_idx); free(t_val); free(t_cnt); free(t_stats);
    return failed ? -1 : 0;
}

void orc_free(void* p) { free(p); }

/* TextUtils.cleanString for ASCII input: lower-case, punctuation -> break, words capitalised and concatenated
 * (only ' ' splits words: Java's split(" ")).  out needs len bytes; returns the cleaned length. */
int64_t orc_clean_string(const uint8_t* s, int64_t len, uint8_t* out) {
    int64_t m = 0;
    int start = 1;
    for (int64_t i = 0; i < len; ++i) {
        unsigned c = s[i];
        if (c == ' ' || is_punct(c)) {
            start = 1;
            continue;
        }
        c = lower(c);
        if (start && (c - 'a') < 26u) c -= 32u;
        start = 0;
        out[m++] = (uint8_t)c;
    }
    return m;
}

/* TextStats of SmartTextVectorizer.fitFn per column: distinct cleaned values, counted until there are more than
 * max_card of them (the reference's map stops growing there), the non-null count and the sums of the cleaned lengths.
 * distinct [ncols] (max_card + 1 means "more"), nonnull [ncols], len_sum / len_sq [ncols].  Columns in parallel. */
int orc_text_stats(const uintptr_t* data_ptr, const uintptr_t* offs_ptr, const uintptr_t* valid_ptr, int ncols,
                   int64_t n, int clean, int max_card, int threads, int32_t* distinct, int64_t* nonnull,
                   int64_t* len_sum, int64_t* len_sq) {
    int failed = 0;
#pragma omp parallel for schedule(dynamic, 1) num_threads(threads < 1 ? 1 : threads)
    for (int j = 0; j < ncols; ++j) {
        const uint8_t* data = (const uint8_t*)data_ptr[j];
        const int64_t* off = (const int64_t*)offs_ptr[j];
        const uint64_t* valid = (const uint64_t*)valid_ptr[j];
        int cap = 64;
        while (cap < 4 * (max_card + 2)) cap <<= 1;
        uint64_t* seen = (uint64_t*)calloc((size_t)cap, sizeof(uint64_t));   /* 0 = empty */
        int64_t buf_cap = 1024;
        uint8_t* buf = (uint8_t*)malloc((size_t)buf_cap);
        int nd = 0;
        int64_t nn = 0, ls = 0, lq = 0;
        int ok = seen && buf;
        for (int64_t r = 0; r < n && ok; ++r) {
            if (!tbit(valid, r)) continue;
            const int64_t len = off[r + 1] - off[r];
            if (len > buf_cap) {
                buf_cap = 2 * len;
                free(buf);
                buf = (uint8_t*)malloc((size_t)buf_cap);
                if (!buf) { ok = 0; break; }
            }
            int64_t m = len;
            const uint8_t* v = data + off[r];
            if (clean) {
                m = orc_clean_string(v, len, buf);
                v = buf;
            }
            ++nn;
            ls += m;
            lq += m * m;
            if (nd <= max_card) {
                uint64_t h = 1469598103934665603ull;
                for (int64_t i = 0; i < m; ++i) h = (h ^ v[i]) * 1099511628211ull;
                h |= 1ull;
                uint64_t s = (h >> 20) & (uint64_t)(cap - 1);
                while (seen[s] != 0 && seen[s] != h) s = (s + 1) & (uint64_t)(cap - 1);
                if (seen[s] == 0) {
                    seen[s] = h;
                    ++nd;
                }
            }
        }
        if (!ok) {
#pragma omp atomic write
            failed = 1;
        }
        distinct[j] = nd;
        nonnull[j] = nn;
        len_sum[j] = ls;
        len_sq[j] = lq;
        free(seen);
        free(buf);
    }
    return failed ? -1 : 0;
}

/* OpOneHotVectorizer.fitFn's counts on dictionary codes: counts [count_off[c] + (1 + code) * nclasses + class] += 1
 * (code < 0: the null id 0).  codes [ncols][ld] int32.  Rows in partitions, merged in partition order. */
int orc_pivot_counts(const int32_t* codes, int64_t ld, int64_t n, int ncols, const int64_t* count_off,
                     int64_t total_bins, const uint8_t* class_ids, int nclasses, int threads, int64_t* counts) {
    if (threads < 1) threads = 1;
    int64_t** part = (int64_t**)calloc((size_t)threads, sizeof(int64_t*));
    int failed = 0;
#pragma omp parallel for schedule(static, 1) num_threads(threads)
    for (int p = 0; p < threads; ++p) {
        int64_t* mine = (int64_t*)calloc((size_t)total_bins, sizeof(int64_t));
        part[p] = mine;
        if (!mine) {
#pragma omp atomic write
            failed = 1;
            continue;
        }
        const int64_t lo = n * p / threads, hi = n * (p + 1) / threads;
        for (int c = 0; c < ncols; ++c) {
            const int32_t* col = codes + (int64_t)c * ld;
            int64_t* bins = mine + count_off[c];
            for (int64_t r = lo; r < hi; ++r) {
                const int32_t code = col[r];
                bins[(int64_t)(code < 0 ? 0 : 1 + code) * nclasses + (nclasses > 1 ? class_ids[r] : 0)] += 1;
            }
        }
    }
    for (int p = 0; p < threads; ++p) {
        if (part[p] && !failed)
            for (int64_t i = 0; i < total_bins; ++i) counts[i] += part[p][i];
        free(part[p]);
    }
    free(part);
    return failed ? -1 : 0;
}

/* OneHotModelFun.pivotFn on dictionary codes: out [n][ncols] = pos [pos_off[c] + (code < 0 ? ncodes_c : code)], the
 * vector index of the row's single 1.0 in column c's block. */
void orc_pivot_transform(const int32_t* codes, int64_t ld, int64_t n, int ncols, const int32_t* pos,
                         const int64_t* pos_off, const int32_t* ncodes, int threads, int32_t* out) {
#pragma omp parallel for schedule(static) num_threads(threads < 1 ? 1 : threads)
    for (int64_t r = 0; r < n; ++r)
        for (int c = 0; c < ncols; ++c) {
            const int32_t code = codes[(int64_t)c * ld + r];
            out[r * ncols + c] = pos[pos_off[c] + (code < 0 ? ncodes[c] : code)];
        }
}

/* cleanString over a whole dictionary (Arrow utf8 in, Arrow utf8 out; out_data needs offsets[count] bytes). */
void orc_clean_strings(const uint8_t* data, const int64_t* offsets, int64_t count, uint8_t* out_data,
                       int64_t* out_offsets) {
    int64_t w = 0;
    out_offsets[0] = 0;
    for (int64_t i = 0; i < count; ++i) {
        w += orc_clean_string(data + offsets[i], offsets[i + 1] - offsets[i], out_data + w);
        out_offsets[i + 1] = w;
    }
}
