// Dictionary-side string work of the pivot vectorizers (host code, inputs are dictionaries: at most one entry per
// distinct raw value, never per row): TextUtils.cleanString, grouping of raw values that clean alike, and the
// top-K selection with the reference's ordering.  Reference file:line per function in include/tmog_b200.h.
#include <string.h>

#include <algorithm>
#include <functional>
#include <memory>
#include <new>
#include <string>
#include <thread>
#include <unordered_map>
#include <vector>

#include "common.cuh"

using namespace tmog;

namespace {

inline bool is_ascii_punct(unsigned char c) {  // java.util.regex \p{Punct} (no UNICODE_CHARACTER_CLASS): 32 chars
    return (c >= 0x21 && c <= 0x2F) || (c >= 0x3A && c <= 0x40) || (c >= 0x5B && c <= 0x60) || (c >= 0x7B && c <= 0x7E);
}

// raw.toLowerCase.replaceAll("[\\p{Punct}]", " ").replaceAll(" +", " ").split(" ").map(_.capitalize).mkString("")
// == cut the lower-cased string at every run of {space, ASCII punctuation}, capitalise each piece, concatenate.
// ASCII only: returns false when a byte >= 0x80 is met (Java's Unicode case mapping is not restated here).
bool clean_ascii(const uint8_t* s, int64_t len, std::string& out) {
    out.clear();
    bool start = true;
    for (int64_t i = 0; i < len; ++i) {
        unsigned char c = s[i];
        if (c >= 0x80) return false;
        if (c == ' ' || is_ascii_punct(c)) {
            start = true;
            continue;
        }
        if (c >= 'A' && c <= 'Z') c = (unsigned char)(c + 32);
        if (start && c >= 'a' && c <= 'z') c = (unsigned char)(c - 32);
        start = false;
        out.push_back((char)c);
    }
    return true;
}

// java.lang.String.compareTo == UTF-16 code unit order.  For well-formed UTF-8 that is byte order except that
// supplementary characters (4-byte sequences, surrogate pairs 0xD800..0xDFFF in UTF-16) sort BEFORE U+E000..U+FFFF.
inline uint32_t utf16_key_of_lead(const unsigned char* p, const unsigned char* end, int& adv) {
    const unsigned char c = *p;
    if (c < 0x80) { adv = 1; return c; }
    if (c < 0xE0 && p + 1 < end) { adv = 2; return ((c & 0x1F) << 6) | (p[1] & 0x3F); }
    if (c < 0xF0 && p + 2 < end) { adv = 3; return ((c & 0x0F) << 12) | ((p[1] & 0x3F) << 6) | (p[2] & 0x3F); }
    if (p + 3 < end) {
        adv = 4;
        const uint32_t cp = ((c & 0x07) << 18) | ((p[1] & 0x3F) << 12) | ((p[2] & 0x3F) << 6) | (p[3] & 0x3F);
        return 0x110000u + (cp - 0x10000u);  // marker: handled by the caller as a surrogate pair
    }
    adv = 1;
    return c;
}

int java_compare(const std::string& a, const std::string& b) {
    const unsigned char *pa = (const unsigned char*)a.data(), *ea = pa + a.size();
    const unsigned char *pb = (const unsigned char*)b.data(), *eb = pb + b.size();
    // expand to UTF-16 code units lazily
    auto next_units = [](const unsigned char*& p, const unsigned char* e, uint16_t out[2]) -> int {
        int adv = 1;
        const uint32_t k = utf16_key_of_lead(p, e, adv);
        p += adv;
        if (k >= 0x110000u) {
            const uint32_t v = k - 0x110000u;
            out[0] = (uint16_t)(0xD800 + (v >> 10));
            out[1] = (uint16_t)(0xDC00 + (v & 0x3FF));
            return 2;
        }
        out[0] = (uint16_t)k;
        return 1;
    };
    uint16_t ua[2], ub[2];
    int na = 0, nb = 0, ia = 0, ib = 0;
    for (;;) {
        if (ia == na) { if (pa == ea) na = 0; else na = next_units(pa, ea, ua); ia = 0; }
        if (ib == nb) { if (pb == eb) nb = 0; else nb = next_units(pb, eb, ub); ib = 0; }
        if (na == 0 || nb == 0) return (na == 0 ? 0 : 1) - (nb == 0 ? 0 : 1);
        if (ua[ia] != ub[ib]) return ua[ia] < ub[ib] ? -1 : 1;
        ++ia;
        ++ib;
    }
}

}  // namespace

extern "C" int tmog_clean_strings(const uint8_t* data, const int64_t* offsets, int64_t count, uint8_t* out_data,
                                  int64_t out_capacity, int64_t* out_offsets) {
    TMOG_REQUIRE(count >= 0 && offsets != nullptr && out_offsets != nullptr, "NULL argument");
    TMOG_REQUIRE(count == 0 || data != nullptr || offsets[count] == 0, "data is NULL");
    std::string tmp;
    int64_t w = 0;
    out_offsets[0] = 0;
    for (int64_t i = 0; i < count; ++i) {
        const int64_t len = offsets[i + 1] - offsets[i];
        TMOG_REQUIRE(len >= 0, "offsets must be non-decreasing");
        if (!clean_ascii(data + offsets[i], len, tmp)) {
            set_error("cleanString: entry %lld holds non-ASCII text; Java's Unicode case mapping is not restated on "
                      "this path (set cleanText = false or pre-clean in the reference)", (long long)i);
            return TMOG_EINVAL;
        }
        if (w + (int64_t)tmp.size() > out_capacity) {
            set_error("out_data too small (%lld bytes)", (long long)out_capacity);
            return TMOG_ENOMEM;
        }
        if (!tmp.empty()) memcpy(out_data + w, tmp.data(), tmp.size());
        w += (int64_t)tmp.size();
        out_offsets[i + 1] = w;
    }
    return TMOG_OK;
}

extern "C" int tmog_dedupe_strings(const uint8_t* data, const int64_t* offsets, int64_t count, int32_t* ids,
                                   int64_t* ndistinct) {
    TMOG_REQUIRE(count >= 0 && offsets != nullptr && ids != nullptr && ndistinct != nullptr, "NULL argument");
    std::unordered_map<std::string, int32_t> seen;
    seen.reserve((size_t)count * 2);
    for (int64_t i = 0; i < count; ++i) {
        std::string key((const char*)data + offsets[i], (size_t)(offsets[i + 1] - offsets[i]));
        auto it = seen.find(key);
        if (it == seen.end()) it = seen.emplace(std::move(key), (int32_t)seen.size()).first;
        ids[i] = it->second;
    }
    *ndistinct = (int64_t)seen.size();
    return TMOG_OK;
}

extern "C" int tmog_pivot_topk(const uint8_t* data, const int64_t* offsets, const int64_t* counts, int64_t count,
                               int top_k, int64_t min_support, int32_t* out_ids, int* out_n) {
    TMOG_REQUIRE(count >= 0 && offsets != nullptr && counts != nullptr && out_ids != nullptr && out_n != nullptr,
                 "NULL argument");
    TMOG_REQUIRE(top_k > 0, "requirement failed: topK must be positive, got %d", top_k);
    std::vector<int32_t> cand;
    for (int64_t i = 0; i < count; ++i)
        if (counts[i] >= min_support && counts[i] > 0) cand.push_back((int32_t)i);
    std::vector<std::string> strs(cand.size());
    for (size_t k = 0; k < cand.size(); ++k)
        strs[k].assign((const char*)data + offsets[cand[k]], (size_t)(offsets[cand[k] + 1] - offsets[cand[k]]));
    std::vector<int32_t> order(cand.size());
    for (size_t k = 0; k < order.size(); ++k) order[k] = (int32_t)k;
    auto less = [&](int32_t a, int32_t b) {  // sortBy(v => -count -> value)
        const int64_t ca = counts[cand[a]], cb = counts[cand[b]];
        if (ca != cb) return ca > cb;
        return java_compare(strs[a], strs[b]) < 0;
    };
    const size_t k = std::min<size_t>((size_t)top_k, order.size());
    std::partial_sort(order.begin(), order.begin() + k, order.end(), less);
    for (size_t i = 0; i < k; ++i) out_ids[i] = cand[order[i]];
    *out_n = (int)k;
    return TMOG_OK;
}

// ---------------------------------------------------------------------------------------------------------------------
// Whole driver-side finish of one pivoted column in one call (what OpOneHotVectorizer.fitFn does after its fold):
// clean every dictionary entry, merge the raw codes whose cleaned values coincide, pick the top-K vocabulary, and
// produce the per-label-class counts of every emitted column (top values, OTHER, null).  Thread-safe; callers run
// the columns of a vectorizer in parallel.
// ---------------------------------------------------------------------------------------------------------------------
namespace {

struct StrRef {
    const char* p;
    int64_t len;
};

// 64-bit string hash, eight bytes per step (murmur3-style mixing; only used to bucket strings -- equality is always
// checked on the bytes)
inline uint64_t fnv64(const char* p, int64_t len) {
    auto rotl = [](uint64_t x, int r) { return (x << r) | (x >> (64 - r)); };
    uint64_t h = 0x9E3779B97F4A7C15ull ^ ((uint64_t)len * 0x87c37b91114253d5ull);
    while (len > 0) {
        uint64_t w = 0;
        memcpy(&w, p, (size_t)(len >= 8 ? 8 : len));
        w *= 0x87c37b91114253d5ull;
        w = rotl(w, 31);
        w *= 0x4cf5ad432745937full;
        h ^= w;
        h = rotl(h, 27) * 5 + 0x52dce729;
        p += 8;
        len -= 8;
    }
    h ^= h >> 33;
    h *= 0xff51afd7ed558ccdull;
    h ^= h >> 33;
    h *= 0xc4ceb9fe1a85ec53ull;
    h ^= h >> 33;
    return h;
}

// cleaned strings of a dictionary, stored back to back; returns false on non-ASCII input when cleaning.  Cleaning never
// lengthens a string, so the pool is sized once and written through a raw pointer.
bool clean_dictionary(const uint8_t* data, const int64_t* offsets, int64_t count, bool clean, std::string& pool,
                      std::vector<int64_t>& pool_off, int64_t* bad_entry) {
    pool_off.resize((size_t)count + 1);
    pool_off[0] = 0;
    const int64_t total = count > 0 ? offsets[count] - offsets[0] : 0;
    pool.resize((size_t)total);
    if (!clean) {
        if (total > 0) memcpy(&pool[0], data + offsets[0], (size_t)total);
        for (int64_t i = 0; i < count; ++i) pool_off[i + 1] = offsets[i + 1] - offsets[0];
        return true;
    }
    // clean_ascii through byte tables: kind 0 = separator (space / punctuation), 1 = kept, 2 = non-ASCII;
    // lower[c] the lower-cased byte, upper[c] the byte that starts a piece (capitalised lower-case letter)
    static unsigned char kind[256], lower[256], upper[256];
    static const bool tables_ready = [] {
        for (int c = 0; c < 256; ++c) {
            kind[c] = c >= 0x80 ? 2 : ((c == ' ' || is_ascii_punct((unsigned char)c)) ? 0 : 1);
            const int lo = (c >= 'A' && c <= 'Z') ? c + 32 : c;
            lower[c] = (unsigned char)lo;
            upper[c] = (unsigned char)((lo >= 'a' && lo <= 'z') ? lo - 32 : lo);
        }
        return true;
    }();
    (void)tables_ready;
    char* out = total > 0 ? &pool[0] : nullptr;
    int64_t w = 0;
    for (int64_t i = 0; i < count; ++i) {
        const uint8_t* s = data + offsets[i];
        const int64_t len = offsets[i + 1] - offsets[i];
        bool start = true;
        for (int64_t q = 0; q < len; ++q) {
            const unsigned char c = s[q];
            const unsigned char k = kind[c];
            if (k == 1) {
                out[w++] = (char)(start ? upper[c] : lower[c]);
                start = false;
            } else if (k == 0) {
                start = true;
            } else {
                *bad_entry = i;
                return false;
            }
        }
        pool_off[i + 1] = w;
    }
    pool.resize((size_t)w);
    return true;
}

// ids[i] = dense id of cleaned string i (first-appearance order); open addressing over (hash, id): a slot keeps the
// hash of its string, so the strings themselves are only compared when the hashes agree
int64_t group_strings(const std::string& pool, const std::vector<int64_t>& off, std::vector<int32_t>& ids,
                      std::vector<int32_t>& first) {
    const int64_t count = (int64_t)off.size() - 1;
    size_t cap = 16;
    while (cap < (size_t)count * 2 + 2) cap <<= 1;
    struct Slot {
        uint64_t hash;
        int32_t id;
    };
    std::vector<Slot> slot(cap, Slot{0, -1});
    ids.resize((size_t)count);
    first.clear();
    first.reserve((size_t)count);
    for (int64_t i = 0; i < count; ++i) {
        const char* p = pool.data() + off[i];
        const int64_t len = off[i + 1] - off[i];
        const uint64_t h = fnv64(p, len);
        size_t s = (size_t)h & (cap - 1);
        for (;;) {
            Slot& sl = slot[s];
            if (sl.id < 0) {
                sl.hash = h;
                sl.id = (int32_t)first.size();
                ids[i] = sl.id;
                first.push_back((int32_t)i);
                break;
            }
            if (sl.hash == h) {
                const int32_t j = first[sl.id];
                if (off[j + 1] - off[j] == len && memcmp(pool.data() + off[j], p, (size_t)len) == 0) {
                    ids[i] = sl.id;
                    break;
                }
            }
            s = (s + 1) & (cap - 1);
        }
    }
    return (int64_t)first.size();
}

// group_strings for a large dictionary on several host threads; same ids and first-entry list.  Hashes are taken in
// index chunks; then thread t owns the strings whose hash falls in its share of the hash range and finds, in index
// order, the first equal string of each of them (rep[i]); a sequential sweep numbers the groups by first appearance.
int64_t group_strings_mt(const std::string& pool, const std::vector<int64_t>& off, std::vector<int32_t>& ids,
                         std::vector<int32_t>& first, int nthreads) {
    const int64_t count = (int64_t)off.size() - 1;
    std::vector<uint64_t> hash((size_t)count);
    std::vector<int32_t> rep((size_t)count);
    auto run = [&](const std::function<void(int)>& body) {
        std::vector<std::thread> th;
        th.reserve((size_t)nthreads - 1);
        for (int t = 1; t < nthreads; ++t) th.emplace_back(body, t);
        body(0);
        for (auto& x : th) x.join();
    };
    run([&](int t) {
        const int64_t lo = count * t / nthreads, hi = count * (t + 1) / nthreads;
        for (int64_t i = lo; i < hi; ++i) hash[(size_t)i] = fnv64(pool.data() + off[i], off[i + 1] - off[i]);
    });
    auto owner = [&](uint64_t h) { return (int)(((h >> 32) * (uint64_t)nthreads) >> 32); };   // by the high bits
    run([&](int t) {
        int64_t mine = 0;
        for (int64_t i = 0; i < count; ++i) mine += owner(hash[(size_t)i]) == t;
        size_t cap = 16;
        while (cap < (size_t)mine * 2 + 2) cap <<= 1;
        struct Slot {
            uint64_t hash;
            int32_t idx;
        };
        std::vector<Slot> slot(cap, Slot{0, -1});
        for (int64_t i = 0; i < count; ++i) {
            const uint64_t h = hash[(size_t)i];
            if (owner(h) != t) continue;
            const char* p = pool.data() + off[i];
            const int64_t len = off[i + 1] - off[i];
            size_t s = (size_t)h & (cap - 1);                                               // by the low bits
            for (;;) {
                Slot& sl = slot[s];
                if (sl.idx < 0) {
                    sl.hash = h;
                    sl.idx = (int32_t)i;
                    rep[(size_t)i] = (int32_t)i;
                    break;
                }
                if (sl.hash == h) {
                    const int32_t j = sl.idx;
                    if (off[j + 1] - off[j] == len && memcmp(pool.data() + off[j], p, (size_t)len) == 0) {
                        rep[(size_t)i] = j;
                        break;
                    }
                }
                s = (s + 1) & (cap - 1);
            }
        }
    });
    ids.resize((size_t)count);
    first.clear();
    first.reserve((size_t)count);
    for (int64_t i = 0; i < count; ++i) {
        if (rep[(size_t)i] == (int32_t)i) {
            ids[(size_t)i] = (int32_t)first.size();
            first.push_back((int32_t)i);
        } else {
            ids[(size_t)i] = ids[(size_t)rep[(size_t)i]];
        }
    }
    return (int64_t)first.size();
}

// dictionaries from this size on are grouped by several threads (the largest dictionary of a vectorizer is the
// critical path of its fit: the other columns' dictionaries are done long before it)
constexpr int64_t GROUP_MT_MIN = 32768;

}  // namespace

// The counts-independent half of the finish (cleaning + grouping of the dictionary entries), so that a caller can run
// it while the count kernel is still busy on the GPU.
struct tmog_pivot_dict {
    std::string pool;               // cleaned entries back to back
    std::vector<int64_t> off;
    std::vector<int32_t> ids;       // cleaned-value id of every dictionary entry
    std::vector<int32_t> first;     // first entry of every cleaned value
    int64_t ncodes = 0;
};

extern "C" int tmog_pivot_dict_prepare(const uint8_t* dict_data, const int64_t* dict_offsets, int64_t ncodes,
                                       int clean_text, tmog_pivot_dict** out) {
    TMOG_REQUIRE(ncodes >= 0 && dict_offsets && out, "NULL argument");
    std::unique_ptr<tmog_pivot_dict> d(new (std::nothrow) tmog_pivot_dict());
    TMOG_REQUIRE(d != nullptr, "out of host memory");
    int64_t bad = -1;
    if (!clean_dictionary(dict_data, dict_offsets, ncodes, clean_text != 0, d->pool, d->off, &bad)) {
        set_error("cleanString: dictionary entry %lld holds non-ASCII text; Java's Unicode case mapping is not restated "
                  "on this path (set cleanText = false)", (long long)bad);
        return TMOG_EINVAL;
    }
    const int hw = (int)std::thread::hardware_concurrency();
    const int nthreads = (int)std::min<int64_t>(std::min(8, std::max(hw, 1)), ncodes / (GROUP_MT_MIN / 2));
    if (ncodes >= GROUP_MT_MIN && nthreads >= 2)
        group_strings_mt(d->pool, d->off, d->ids, d->first, nthreads);
    else
        group_strings(d->pool, d->off, d->ids, d->first);
    d->ncodes = ncodes;
    *out = d.release();
    return TMOG_OK;
}

extern "C" int tmog_pivot_dict_free(tmog_pivot_dict* d) {
    delete d;
    return TMOG_OK;
}

extern "C" int tmog_pivot_dict_finish(const tmog_pivot_dict* d, const int64_t* raw_counts, int nclasses, int top_k,
                                      int64_t min_support, int* n_top, int32_t* code_to_out,
                                      int64_t* out_class_counts, uint8_t* top_data, int64_t top_capacity,
                                      int64_t* top_offsets) {
    TMOG_REQUIRE(d && raw_counts && n_top && code_to_out && out_class_counts && top_offsets, "NULL argument");
    TMOG_REQUIRE(top_k > 0 && nclasses >= 1, "requirement failed: topK must be positive");
    const std::string& pool = d->pool;
    const std::vector<int64_t>& off = d->off;
    const std::vector<int32_t>&ids = d->ids, &first = d->first;
    const int64_t ncodes = d->ncodes, nd = (int64_t)first.size();
    // rows per cleaned value (raw_counts row 0 is the null bin, row 1 + code the raw code)
    std::vector<int64_t> total((size_t)nd, 0);
    for (int64_t i = 0; i < ncodes; ++i) {
        int64_t v = 0;
        for (int k = 0; k < nclasses; ++k) v += raw_counts[(i + 1) * nclasses + k];
        total[ids[i]] += v;
    }
    // top-K: count >= minSupport (and > 0), sort by (-count, java string order)
    std::vector<int32_t> cand;
    for (int64_t v = 0; v < nd; ++v)
        if (total[v] >= min_support && total[v] > 0) cand.push_back((int32_t)v);
    auto str_of = [&](int32_t v) {
        const int32_t i = first[v];
        return std::string(pool.data() + off[i], (size_t)(off[i + 1] - off[i]));
    };
    auto less = [&](int32_t a, int32_t b) {
        if (total[a] != total[b]) return total[a] > total[b];
        return java_compare(str_of(a), str_of(b)) < 0;
    };
    const size_t k = std::min<size_t>((size_t)top_k, cand.size());
    if (cand.size() > 4 * k && k > 0) {
        // only values whose count reaches the k-th largest count can make the vocabulary: cut the candidates down
        // with a selection on the counts alone before the (string-comparing) sort
        std::vector<int64_t> cnts(cand.size());
        for (size_t i = 0; i < cand.size(); ++i) cnts[i] = total[cand[i]];
        std::nth_element(cnts.begin(), cnts.begin() + (k - 1), cnts.end(), std::greater<int64_t>());
        const int64_t kth = cnts[k - 1];
        size_t w2 = 0;
        for (size_t i = 0; i < cand.size(); ++i)
            if (total[cand[i]] >= kth) cand[w2++] = cand[i];
        cand.resize(w2);
    }
    std::partial_sort(cand.begin(), cand.begin() + k, cand.end(), less);
    std::vector<int32_t> pos((size_t)nd, (int32_t)k);  // OTHER = k
    int64_t w = 0;
    top_offsets[0] = 0;
    for (size_t t = 0; t < k; ++t) {
        pos[cand[t]] = (int32_t)t;
        const std::string sv = str_of(cand[t]);
        if (w + (int64_t)sv.size() > top_capacity) {
            set_error("top_data too small");
            return TMOG_ENOMEM;
        }
        if (!sv.empty()) memcpy(top_data + w, sv.data(), sv.size());
        w += (int64_t)sv.size();
        top_offsets[t + 1] = w;
    }
    *n_top = (int)k;
    for (int64_t i = 0; i < ncodes; ++i) code_to_out[i] = pos[ids[i]];
    // emitted columns: k top values, OTHER, null
    for (size_t i = 0; i < (k + 2) * (size_t)nclasses; ++i) out_class_counts[i] = 0;
    for (int64_t i = 0; i < ncodes; ++i) {
        int64_t* o = out_class_counts + (size_t)pos[ids[i]] * nclasses;
        for (int c = 0; c < nclasses; ++c) o[c] += raw_counts[(i + 1) * nclasses + c];
    }
    for (int c = 0; c < nclasses; ++c) out_class_counts[(k + 1) * nclasses + c] = raw_counts[c];
    return TMOG_OK;
}

// 64-bit digest of a byte range (eight bytes per step): what the ranks compare to find out whether they encoded their
// rows against the same dictionary.
extern "C" int tmog_hash_bytes(const uint8_t* data, int64_t nbytes, uint64_t seed, uint64_t* out) {
    TMOG_REQUIRE(nbytes >= 0 && out != nullptr && (data != nullptr || nbytes == 0), "NULL argument");
    auto rotl = [](uint64_t x, int r) { return (x << r) | (x >> (64 - r)); };
    auto mix = [&](uint64_t h, uint64_t w) {
        w *= 0x87c37b91114253d5ull;
        w = rotl(w, 31);
        w *= 0x4cf5ad432745937full;
        h ^= w;
        return rotl(h, 27) * 5 + 0x52dce729;
    };
    // four independent lanes (32 bytes per step) so that the multiplies overlap; the lanes are folded at the end
    uint64_t h0 = seed ^ ((uint64_t)nbytes * 0x9E3779B97F4A7C15ull), h1 = h0 ^ 0x6a09e667f3bcc909ull,
             h2 = h0 ^ 0xbb67ae8584caa73bull, h3 = h0 ^ 0x3c6ef372fe94f82bull;
    const uint8_t* p = data;
    int64_t left = nbytes;
    while (left >= 32) {
        uint64_t w[4];
        memcpy(w, p, 32);
        h0 = mix(h0, w[0]);
        h1 = mix(h1, w[1]);
        h2 = mix(h2, w[2]);
        h3 = mix(h3, w[3]);
        p += 32;
        left -= 32;
    }
    while (left > 0) {
        uint64_t w = 0;
        memcpy(&w, p, (size_t)(left >= 8 ? 8 : left));
        h0 = mix(h0, w);
        p += 8;
        left -= 8;
    }
    uint64_t h = mix(mix(mix(h0, h1), h2), h3);
    h ^= h >> 33;
    h *= 0xff51afd7ed558ccdull;
    h ^= h >> 33;
    h *= 0xc4ceb9fe1a85ec53ull;
    h ^= h >> 33;
    *out = h;
    return TMOG_OK;
}

extern "C" int tmog_pivot_column_finish(const uint8_t* dict_data, const int64_t* dict_offsets, int64_t ncodes,
                                        int clean_text, const int64_t* raw_counts, int nclasses, int top_k,
                                        int64_t min_support, int* n_top, int32_t* code_to_out,
                                        int64_t* out_class_counts, uint8_t* top_data, int64_t top_capacity,
                                        int64_t* top_offsets) {
    TMOG_REQUIRE(ncodes >= 0 && dict_offsets && raw_counts && n_top && code_to_out && out_class_counts && top_offsets,
                 "NULL argument");
    TMOG_REQUIRE(top_k > 0 && nclasses >= 1, "requirement failed: topK must be positive");
    tmog_pivot_dict* d = nullptr;
    if (int rc = tmog_pivot_dict_prepare(dict_data, dict_offsets, ncodes, clean_text, &d)) return rc;
    const int rc = tmog_pivot_dict_finish(d, raw_counts, nclasses, top_k, min_support, n_top, code_to_out,
                                          out_class_counts, top_data, top_capacity, top_offsets);
    tmog_pivot_dict_free(d);
    return rc;
}

// code_to_out[i] = position of dictionary entry i's cleaned value in the fitted vocabulary `top` (n_top strings), or
// n_top (OTHER): the table OneHotModelFun.pivotFn needs for a batch whose dictionary differs from the fitted one.
extern "C" int tmog_pivot_column_map(const uint8_t* dict_data, const int64_t* dict_offsets, int64_t ncodes,
                                     int clean_text, const uint8_t* top_data, const int64_t* top_offsets, int n_top,
                                     int32_t* code_to_out) {
    TMOG_REQUIRE(ncodes >= 0 && dict_offsets && top_offsets && code_to_out && n_top >= 0, "NULL argument");
    std::unordered_map<std::string, int32_t> top;
    for (int t = 0; t < n_top; ++t)
        top.emplace(std::string((const char*)top_data + top_offsets[t], (size_t)(top_offsets[t + 1] - top_offsets[t])), t);
    std::string tmp;
    for (int64_t i = 0; i < ncodes; ++i) {
        const int64_t len = dict_offsets[i + 1] - dict_offsets[i];
        if (clean_text) {
            if (!clean_ascii(dict_data + dict_offsets[i], len, tmp)) {
                set_error("cleanString: dictionary entry %lld holds non-ASCII text", (long long)i);
                return TMOG_EINVAL;
            }
        } else {
            tmp.assign((const char*)dict_data + dict_offsets[i], (size_t)len);
        }
        auto it = top.find(tmp);
        code_to_out[i] = it == top.end() ? n_top : it->second;
    }
    return TMOG_OK;
}
