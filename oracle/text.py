"""Oracle: string cleaning, tokenisation and Spark-legacy murmur3 hashing (TEST INFRASTRUCTURE ONLY).

Follows
  * utils/src/main/scala/com/salesforce/op/utils/text/TextUtils.scala:39-47 (cleanString)
  * core/src/main/scala/com/salesforce/op/stages/impl/feature/TextTokenizer.scala:160-220
  * core/src/main/scala/com/salesforce/op/utils/text/LuceneTextAnalyzer.scala:103-122,159-166
  * core/src/main/scala/com/salesforce/op/stages/impl/feature/OPCollectionHashingVectorizer.scala:191-305
and restates the third-party pieces that are not under /root/reference:
  * Spark 2.4.5 mllib.feature.HashingTF.murmur3Hash / Murmur3_x86_32.hashUnsafeBytes (legacy tail handling)
  * Lucene 7.3.0 StandardTokenizer (UAX#29) restricted to ASCII + Snowball english_stop.txt
"""
from __future__ import annotations

import string
from typing import Iterable, List, Optional, Sequence

M32 = 0xFFFFFFFF

# java.util.regex \p{Punct} without UNICODE_CHARACTER_CLASS = the 32 ASCII punctuation characters
PUNCT = set(string.punctuation)
assert len(PUNCT) == 32

# Snowball english_stop.txt as loaded at LuceneTextAnalyzer.scala:159-161 (list recalled, SURVEY Appendix B;
# membership of the words exercised by the reference tests is pinned in tests/test_oracle_golden.py)
ENGLISH_STOP = frozenset("""
i me my myself we our ours ourselves you your yours yourself yourselves he him his himself she her hers herself it its
itself they them their theirs themselves what which who whom this that these those am is are was were be been being have
has had having do does did doing would should could ought i'm you're he's she's it's we're they're i've you've we've
they've i'd you'd he'd she'd we'd they'd i'll you'll he'll she'll we'll they'll isn't aren't wasn't weren't hasn't
haven't hadn't doesn't don't didn't won't wouldn't shan't shouldn't can't cannot couldn't mustn't let's that's who's
what's here's there's when's where's why's how's a an the and but if or because as until while of at by for with about
against between into through during before after above below to from up down in out on off over under again further
then once here there when where why how all any both each few more most other some such no nor not only own same so
than too very
""".split())
assert len(ENGLISH_STOP) == 174


# ----------------------------------------------------------------------------------------------------------------
# cleanString  (TextUtils.scala:39-47)
# ----------------------------------------------------------------------------------------------------------------
def _capitalize(w: str) -> str:
    # scala 2.11 StringOps.capitalize: upper-case the first UTF-16 char only
    if not w:
        return w
    c = w[0]
    u = c.upper()
    if len(u) != 1:  # Character.toUpperCase is a 1:1 char mapping (e.g. 'ß' stays)
        u = c
    return u + w[1:]


def clean_string(raw: str) -> str:
    """``raw.toLowerCase.replaceAll("[\\p{Punct}]"," ").replaceAll(" +"," ").split(" ").map(capitalize).mkString("")``.

    Equivalent formulation: the lower-cased string is cut at every run of {space, ASCII punctuation}; each piece is
    capitalised and the pieces are concatenated (leading/trailing empties from Java's split contribute nothing).
    """
    low = raw.lower()
    out: List[str] = []
    word: List[str] = []
    for ch in low:
        if ch == " " or ch in PUNCT:
            if word:
                out.append(_capitalize("".join(word)))
                word = []
        else:
            word.append(ch)
    if word:
        out.append(_capitalize("".join(word)))
    return "".join(out)


def clean_text_fn(s: str, should_clean: bool) -> str:
    """Transmogrifier.scala:527"""
    return clean_string(s) if should_clean else s


# ----------------------------------------------------------------------------------------------------------------
# Tokeniser: toLowerCase -> StandardTokenizer -> LowerCaseFilter -> StopFilter -> length filter
# ----------------------------------------------------------------------------------------------------------------
MAX_TOKEN_LENGTH = 255  # StandardAnalyzer.DEFAULT_MAX_TOKEN_LENGTH

_MID_LETTER = set(":.'")     # MidLetter | MidNumLet | Single_Quote  (joins letter-letter)
_MID_NUM = set(",;.'")       # MidNum | MidNumLet | Single_Quote     (joins digit-digit)


def _is_letter(c: str) -> bool:
    return ("a" <= c <= "z") or ("A" <= c <= "Z") or (ord(c) > 127 and c.isalpha())


def _is_digit(c: str) -> bool:
    return ("0" <= c <= "9") or (ord(c) > 127 and c.isdigit())


def standard_tokenize(s: str) -> List[str]:
    """ASCII restriction of Lucene 7.3 StandardTokenizer (UAX#29 word breaks; only segments holding a letter or
    digit are emitted).  Greedy scan == JFlex longest match because every continuation decision is local."""
    toks: List[str] = []
    n = len(s)
    i = 0
    while i < n:
        c = s[i]
        if not (_is_letter(c) or _is_digit(c) or c == "_"):
            i += 1
            continue
        j = i
        has_alnum = False
        while j < n:
            c = s[j]
            if _is_letter(c) or _is_digit(c):
                has_alnum = True
                j += 1
            elif c == "_":
                j += 1
            elif j > i and j + 1 < n and (
                (c in _MID_LETTER and _is_letter(s[j - 1]) and _is_letter(s[j + 1]))
                or (c in _MID_NUM and _is_digit(s[j - 1]) and _is_digit(s[j + 1]))
            ):
                j += 2 - 1  # consume the mid char; the next char (alnum) is consumed by the next iteration
            else:
                break
        if has_alnum:
            tok = s[i:j]
            # tokens longer than maxTokenLength are chopped into maxTokenLength pieces (Lucene >= 6.5)
            for k in range(0, len(tok), MAX_TOKEN_LENGTH):
                toks.append(tok[k:k + MAX_TOKEN_LENGTH])
        i = j if j > i else i + 1
    return toks


def tokenize(text: Optional[str], to_lowercase: bool = True, min_token_length: int = 1) -> List[str]:
    """TextTokenizer.tokenize with defaults (Language.Unknown, no language detection, no html stripping)."""
    if text is None:
        return []
    low = text.lower() if to_lowercase else text
    toks = standard_tokenize(low)
    toks = [t.lower() for t in toks]                 # LowerCaseFilter
    toks = [t for t in toks if t not in ENGLISH_STOP]  # StopFilter
    return [t for t in toks if len(t) >= min_token_length]


# ----------------------------------------------------------------------------------------------------------------
# Spark 2.4.5 HashingTF (legacy murmur3_x86_32, seed 42)
# ----------------------------------------------------------------------------------------------------------------
def _rotl(x: int, r: int) -> int:
    return ((x << r) | (x >> (32 - r))) & M32


def _mix_k1(k: int) -> int:
    k = (k * 0xCC9E2D51) & M32
    k = _rotl(k, 15)
    return (k * 0x1B873593) & M32


def _mix_h1(h: int, k: int) -> int:
    h ^= k
    h = _rotl(h, 13)
    return (h * 5 + 0xE6546B64) & M32


def _fmix(h: int, n: int) -> int:
    h ^= n & M32
    h ^= h >> 16
    h = (h * 0x85EBCA6B) & M32
    h ^= h >> 13
    h = (h * 0xC2B2AE35) & M32
    h ^= h >> 16
    return h


def murmur3_legacy(data: bytes, seed: int = 42) -> int:
    """Murmur3_x86_32.hashUnsafeBytes: aligned 4-byte LE blocks, then EACH tail byte sign-extended as its own
    block, then fmix(len).  Returns the signed 32-bit value."""
    n = len(data)
    h = seed & M32
    nb = n - (n % 4)
    for i in range(0, nb, 4):
        k = int.from_bytes(data[i:i + 4], "little")
        h = _mix_h1(h, _mix_k1(k))
    for i in range(nb, n):
        b = data[i]
        if b >= 128:
            b -= 256
        h = _mix_h1(h, _mix_k1(b & M32))
    h = _fmix(h, n)
    return h - (1 << 32) if h >= (1 << 31) else h


def hashing_tf_index(term: str, num_features: int) -> int:
    """HashingTF.indexOf for a String term: Utils.nonNegativeMod(murmur3Hash(term), numFeatures)."""
    h = murmur3_legacy(term.encode("utf-8"))
    # Java % truncates toward zero; ((h % n) + n) % n
    r = abs(h) % num_features
    if h < 0:
        r = -r
    return (r + num_features) % num_features


def hashing_tf_transform(terms: Iterable[str], num_features: int, binary: bool = False):
    """HashingTF.transform: term-frequency map -> (sorted indices, values)."""
    tf = {}
    for t in terms:
        i = hashing_tf_index(t, num_features)
        tf[i] = 1.0 if binary else tf.get(i, 0.0) + 1.0
    idx = sorted(tf)
    return idx, [tf[i] for i in idx]
