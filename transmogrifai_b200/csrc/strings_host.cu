// Dictionary-side string work of the pivot vectorizers (host code, inputs are dictionaries: at most one entry per
// distinct raw value, never per row): TextUtils.cleanString, grouping of raw values that clean alike, and the
// top-K selection with the reference's ordering.  Reference file:line per function in include/tmog_b200.h.
#include <string.h>

#include <algorithm>
#include <string>
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
