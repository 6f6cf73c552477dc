// similarity.cpp — CPU restatement of the reference's similarity-edge compare.
// TEST INFRASTRUCTURE ONLY: only tests/ may load this library; the product never links or calls it.
//
// PARITY UNPINNED: the reference ships no tests or vectors for this step and cannot be run here.
// `J/` = /root/reference/src/main/java/org/uu/nl/embedding/.  Restated from
//   J/convert/Rdf2GrphConverter.java:127-186 (one CompareJob per source literal, both sides as arrays)
//   J/compare/CompareJob.java:21-53           (startIndex, own-vertex skip, >= threshold, (float) cast)
//   J/util/similarity/*.java                  (the metrics; cited at each function)
// Third-party code on this path that is NOT under /root/reference: info.debatty
// java-string-similarity (pom.xml:50-51, version "RELEASE" = unpinned; 2.0.0 is the last release).
// Its ShingleBased.getProfile (k-shingles of the string after collapsing every whitespace run to one
// blank, with occurrence counts), JaroWinkler and NormalizedLevenshtein are restated from the
// published 2.0.0 source.
// JDK pieces restated from their Javadoc: String.trim, Integer.parseInt (ASCII digits only here),
// DateTimeFormatter.BASIC_ISO_DATE (strict; optional offset Z / +HH / +HHMM / +HHMMss),
// LocalDate.toEpochDay / monthsUntil.  Strings are UTF-16 code units, as in Java.
//
// Build: -O2 -ffp-contract=off -fno-fast-math.
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <map>
#include <set>
#include <string>
#include <vector>

namespace {

using ustr = std::u16string;
using Profile = std::map<ustr, int32_t>;  // HashMap<String,Integer>: only order-independent uses below

enum Method { NGRAM_COSINE = 0, NGRAM_JACCARD, TOKEN_COSINE, TOKEN_JACCARD, NUMERIC, DATE_DAYS, DATE_MONTHS, DATE_YEARS, JAROWINKLER, LEVENSHTEIN };
enum TimeDir { BIDIRECTIONAL = 0, BACKWARDS = 1, FORWARDS = 2 };

struct JobDied {};  // an exception that CompareJob.call does not catch

// ---- info.debatty ShingleBased.getProfile(k) ------------------------------------------------
bool java_space(char16_t c) { return c == u' ' || c == u'\t' || c == u'\n' || c == 0x0B || c == u'\f' || c == u'\r'; }  // regex \s

Profile ngram_profile(const ustr& s, int k) {
    ustr t;  // SPACE_REG.matcher(string).replaceAll(" ")
    for (size_t i = 0; i < s.size();) {
        if (java_space(s[i])) { t.push_back(u' '); while (i < s.size() && java_space(s[i])) i++; }
        else t.push_back(s[i++]);
    }
    Profile p;
    for (int64_t i = 0; i + k <= (int64_t)t.size(); i++) p[t.substr((size_t)i, (size_t)k)]++;
    return p;
}

// ---- TokenBased.getProfile (J/util/similarity/TokenBased.java:20-29, Tokenator :33-73) ------
ustr java_trim(const ustr& s) {  // String.trim: code units <= U+0020 off both ends
    size_t a = 0, b = s.size();
    while (a < b && s[a] <= u' ') a++;
    while (b > a && s[b - 1] <= u' ') b--;
    return s.substr(a, b - a);
}

bool legal_token(const ustr& t) {  // :43-48
    static const char16_t* const STOP[] = {u"the", u"of", u"and", u"a", u"an", u"to", u"in", u"is", u"you", u"that", u"it",
                                            u"for", u"on", u"from", u"are", u"as", u"with", u"at", u"or", u"by", u"but", u"if"};
    if (t.size() <= 1) return false;
    for (const char16_t* w : STOP) if (t == w) return false;
    return true;
}

Profile token_profile(const ustr& s) {
    Profile p;
    size_t from = 0;  // tokenCharIndex
    for (size_t pos = 0; pos < s.size(); pos++) {  // :53-67: a token ends at a blank or at the last character
        if (s[pos] == u' ' || pos == s.size() - 1) {
            const ustr tok = java_trim(s.substr(from, pos + 1 - from));
            from = pos + 1;
            if (legal_token(tok)) p[tok]++;  // shingles.merge(token, 1, Integer::sum)
        }
    }
    return p;
}

// ---- profile metrics -------------------------------------------------------------------------
double jaccard(const Profile& a, const Profile& b) {  // PreComputedNgramJaccard.java:41-47, PreComputedTokenJaccard.java:45-51
    std::set<ustr> uni;
    for (auto& e : a) uni.insert(e.first);
    for (auto& e : b) uni.insert(e.first);
    const int32_t inter = (int32_t)a.size() + (int32_t)b.size() - (int32_t)uni.size();
    return inter / (double)uni.size();
}

double norm(const Profile& p) {  // TokenCosine.java:30-32
    double s = 0;
    for (auto& e : p) s += std::pow((double)e.second, 2);
    return std::sqrt(s);
}

int32_t dot_product(const Profile& a, const Profile& b) {  // TokenCosine.java:34-40: walk the smaller map, int sum
    const Profile &small = a.size() > b.size() ? b : a, &big = a.size() > b.size() ? a : b;
    uint32_t s = 0;
    for (auto& e : small) {
        auto it = big.find(e.first);
        s += (uint32_t)(it == big.end() ? 0 : it->second) * (uint32_t)e.second;
    }
    return (int32_t)s;
}

double cosine(const Profile& a, const Profile& b) { return dot_product(a, b) / (norm(a) * norm(b)); }

// ---- info.debatty NormalizedLevenshtein / Levenshtein (published 2.0.0 source) ---------------
// Levenshtein.distance: two-row DP, unit costs, returned as double; NormalizedLevenshtein.distance =
// that / max(len1, len2) (0 for equal strings or two empty ones); similarity = 1.0 - distance.
double levenshtein_distance(const ustr& s1, const ustr& s2) {
    if (s1 == s2) return 0;
    if (s1.empty()) return (double)s2.size();
    if (s2.empty()) return (double)s1.size();
    std::vector<int32_t> v0(s2.size() + 1), v1(s2.size() + 1);
    for (size_t i = 0; i < v0.size(); i++) v0[i] = (int32_t)i;
    for (size_t i = 0; i < s1.size(); i++) {
        v1[0] = (int32_t)i + 1;
        for (size_t j = 0; j < s2.size(); j++) {
            const int32_t cost = s1[i] == s2[j] ? 0 : 1;
            v1[j + 1] = std::min(v1[j] + 1, std::min(v0[j + 1] + 1, v0[j] + cost));
        }
        std::swap(v0, v1);
    }
    return (double)v0[s2.size()];
}

double normalized_levenshtein(const ustr& s1, const ustr& s2) {
    if (s1 == s2) return 1.0 - 0;
    const int32_t m_len = (int32_t)std::max(s1.size(), s2.size());
    if (m_len == 0) return 1.0 - 0;
    return 1.0 - levenshtein_distance(s1, s2) / m_len;
}

// ---- info.debatty JaroWinkler (published 2.0.0 source; threshold 0.7, JW_COEF 0.1) -----------
double jaro_winkler(const ustr& s1, const ustr& s2) {
    if (s1 == s2) return 1;
    // matches(): {matches, transpositions / 2, prefix, max.length()}
    const ustr &mx = s1.size() > s2.size() ? s1 : s2, &mn = s1.size() > s2.size() ? s2 : s1;
    const int32_t range = std::max((int32_t)mx.size() / 2 - 1, 0);
    std::vector<int32_t> match_indexes(mn.size(), -1);
    std::vector<char> match_flags(mx.size(), 0);
    int32_t matches = 0;
    for (int32_t mi = 0; mi < (int32_t)mn.size(); mi++) {
        const char16_t c1 = mn[mi];
        for (int32_t xi = std::max(mi - range, 0), xn = std::min(mi + range + 1, (int32_t)mx.size()); xi < xn; xi++) {
            if (!match_flags[xi] && c1 == mx[xi]) { match_indexes[mi] = xi; match_flags[xi] = 1; matches++; break; }
        }
    }
    ustr ms1, ms2;
    for (size_t i = 0; i < mn.size(); i++) if (match_indexes[i] != -1) ms1.push_back(mn[i]);
    for (size_t i = 0; i < mx.size(); i++) if (match_flags[i]) ms2.push_back(mx[i]);
    int32_t transpositions = 0;
    for (size_t k = 0; k < ms1.size(); k++) if (ms1[k] != ms2[k]) transpositions++;
    int32_t prefix = 0;
    for (size_t k = 0; k < mn.size(); k++) { if (s1[k] == s2[k]) prefix++; else break; }
    const int32_t mtp[4] = {matches, transpositions / 2, prefix, (int32_t)mx.size()};
    // similarity(): the Jaro term is float arithmetic (m is a float), widened afterwards
    const float m = (float)mtp[0];
    if (m == 0) return 0.0f;
    const double j = ((m / (float)s1.size() + m / (float)s2.size() + (m - (float)mtp[1]) / m)) / (float)3;
    double jw = j;
    if (j > 0.7) jw = j + std::min(0.1, 1.0 / mtp[3]) * mtp[2] * (1 - j);
    return jw;
}

// ---- Integer.parseInt ------------------------------------------------------------------------
bool parse_int(const ustr& s, int32_t* out) {
    if (s.empty()) return false;
    size_t i = 0;
    bool neg = false;
    if (s[0] == u'-' || s[0] == u'+') { neg = s[0] == u'-'; i = 1; if (s.size() == 1) return false; }
    int64_t v = 0;
    for (; i < s.size(); i++) {
        if (s[i] < u'0' || s[i] > u'9') return false;
        v = v * 10 + (s[i] - u'0');
        if (v > 2147483648LL) return false;
    }
    if (neg) v = -v;
    if (v > 2147483647LL) return false;
    *out = (int32_t)v;
    return true;
}

double smooth_pow(double base, double distance, double smooth) {  // Numeric.java:39, Date.java:60
    return std::pow(std::fabs(base - distance) + 1, smooth - 1);
}

// ---- Numeric.similarity (J/util/similarity/Numeric.java:18-44) -------------------------------
double numeric(ustr s1, ustr s2, double alpha, double distance) {
    if (s1.empty() || s2.empty()) return 0;
    if (s1 == s2) return 1;
    const size_t h1 = s1.find(u'^');
    const size_t h2 = s1.find(u'^');  // :30 — the reference looks for '^' in s1 again, not in s2
    if (h1 != ustr::npos) s1 = s1.substr(0, h1);
    if (h2 != ustr::npos) {
        if (h2 > s2.size()) throw JobDied();  // String.substring: StringIndexOutOfBoundsException, outside the try
        s2 = s2.substr(0, h2);
    }
    int32_t a, b;
    if (!parse_int(s1, &a) || !parse_int(s2, &b)) return 0;  // NumberFormatException caught :40-43
    const uint32_t d = (uint32_t)a - (uint32_t)b;            // int arithmetic wraps
    int32_t ad = (int32_t)d;
    if (ad < 0) ad = (int32_t)(0u - d);                      // Math.abs(Integer.MIN_VALUE) stays negative
    return smooth_pow((double)ad, distance, alpha);
}

// ---- LocalDate pieces ------------------------------------------------------------------------
struct Ymd { int64_t y; int m, d; };

bool leap(int64_t y) { return (y % 4 == 0) && (y % 100 != 0 || y % 400 == 0); }
int month_len(int64_t y, int m) { static const int L[12] = {31, 28, 31, 30, 31, 30, 31, 31, 30, 31, 30, 31}; return m == 2 && leap(y) ? 29 : L[m - 1]; }

int64_t epoch_day(const Ymd& t) {  // LocalDate.toEpochDay (JDK Javadoc algorithm, proleptic ISO)
    const int64_t y = t.y, m = t.m;
    int64_t total = 365 * y;
    if (y >= 0) total += (y + 3) / 4 - (y + 99) / 100 + (y + 399) / 400;
    else total -= y / -4 - y / -100 + y / -400;
    total += (367 * m - 362) / 12;
    total += t.d - 1;
    if (m > 2) { total--; if (!leap(y)) total--; }
    return total - 719528;  // DAYS_0000_TO_1970
}

bool parse_basic_iso(const ustr& s, Ymd* out) {  // DateTimeFormatter.BASIC_ISO_DATE, ResolverStyle.STRICT
    if (s.size() < 8) return false;
    int v[8];
    for (int i = 0; i < 8; i++) { if (s[i] < u'0' || s[i] > u'9') return false; v[i] = s[i] - u'0'; }
    out->y = v[0] * 1000 + v[1] * 100 + v[2] * 10 + v[3];
    out->m = v[4] * 10 + v[5];
    out->d = v[6] * 10 + v[7];
    if (out->m < 1 || out->m > 12 || out->d < 1 || out->d > month_len(out->y, out->m)) return false;
    const ustr rest = s.substr(8);  // optional offset: Z | [+-]HH | [+-]HHMM | [+-]HHMMss
    if (rest.empty()) return true;
    if (rest == u"Z" || rest == u"z") return true;
    if (rest[0] != u'+' && rest[0] != u'-') return false;
    const size_t nd = rest.size() - 1;
    if (nd != 2 && nd != 4 && nd != 6) return false;
    int f[3] = {0, 0, 0};
    for (size_t i = 0; i < nd; i++) {
        if (rest[1 + i] < u'0' || rest[1 + i] > u'9') return false;
        f[i / 2] = f[i / 2] * 10 + (rest[1 + i] - u'0');
    }
    if (f[0] > 18 || f[1] > 59 || f[2] > 59 || (f[0] == 18 && (f[1] || f[2]))) return false;  // ZoneOffset range
    return true;
}

// ---- Date.similarity (J/util/similarity/Date.java:30-66) -------------------------------------
double date(ustr s1, ustr s2, int unit, double smooth, double distance, int dir) {
    if (s1.empty() || s2.empty()) return 0;
    if (s1 == s2) return 1;
    const size_t h1 = s1.find(u'^'), h2 = s2.find(u'^');
    if (h1 != ustr::npos) s1 = s1.substr(0, h1);
    if (h2 != ustr::npos) s2 = s2.substr(0, h2);
    Ymd d1, d2;
    if (!parse_basic_iso(s1, &d1) || !parse_basic_iso(s2, &d2)) return 0;  // DateTimeParseException caught :62-65
    const int64_t e1 = epoch_day(d1), e2 = epoch_day(d2);
    if (dir == BACKWARDS && e1 > e2) return 0;  // d1.isAfter(d2)
    if (dir == FORWARDS && e1 < e2) return 0;   // d1.isBefore(d2)
    int64_t between;
    if (unit == DATE_DAYS) between = e2 - e1;   // ChronoUnit.DAYS.between = d1.until(d2, DAYS)
    else {
        const int64_t p1 = (d1.y * 12 + d1.m - 1) * 32 + d1.d, p2 = (d2.y * 12 + d2.m - 1) * 32 + d2.d;
        between = (p2 - p1) / 32;               // LocalDate.monthsUntil
        if (unit == DATE_YEARS) between /= 12;
    }
    return smooth_pow(std::fabs((double)between), distance, smooth);
}

}  // namespace

extern "C" {

// labels: n strings as UTF-16 code units chars[off[k] .. off[k+1]); vert[k] = vertex of label k.
// source / target: label indices in group order.  Returns the pair count (pairs beyond cap are
// counted, not stored); order = (source index, target index).  iso dates only.
int64_t orc_sim_pairs(int32_t method, int32_t ngram, double threshold, double smooth, double distance, int32_t time_dir,
                      int32_t upper, int32_t n, const uint16_t* chars, const int64_t* off, const int32_t* vert,
                      int32_t n_source, const int32_t* source, int32_t n_target, const int32_t* target, int64_t cap,
                      int32_t* out_v, int32_t* out_o, float* out_s) {
    std::vector<ustr> label((size_t)n);
    for (int32_t k = 0; k < n; k++) label[k] = ustr((const char16_t*)chars + off[k], (size_t)(off[k + 1] - off[k]));
    std::vector<Profile> prof;
    if (method <= TOKEN_JACCARD) {  // preCompute while the triples are read (Rdf2GrphConverter.java:103,109)
        prof.resize((size_t)n);
        for (int32_t k = 0; k < n; k++)
            prof[k] = (method == NGRAM_COSINE || method == NGRAM_JACCARD) ? ngram_profile(label[k], ngram) : token_profile(label[k]);
    }
    int64_t count = 0;
    for (int32_t i = 0; i < n_source; i++) {  // one CompareJob each
        const int32_t a = source[i], va = vert[a];
        std::vector<int32_t> ov;
        std::vector<float> os;
        bool died = false;
        for (int32_t j = upper ? i + 1 : 0; j < n_target; j++) {  // CompareJob.java:28,37
            const int32_t b = target[j], vb = vert[b];
            if (vb == va) continue;  // :40
            const ustr &s1 = label[a], &s2 = label[b];
            double sim;
            try {
                switch (method) {
                    case NGRAM_JACCARD: case TOKEN_JACCARD: sim = s1 == s2 ? 1 : jaccard(prof[a], prof[b]); break;
                    case NGRAM_COSINE:  // PreComputedNgramCosine.java:25-31
                        sim = s1 == s2 ? 1 : ((int64_t)s1.size() >= ngram && (int64_t)s2.size() >= ngram ? cosine(prof[a], prof[b]) : 0);
                        break;
                    case TOKEN_COSINE: sim = s1 == s2 ? 1 : cosine(prof[a], prof[b]); break;
                    case NUMERIC: sim = numeric(s1, s2, smooth, distance); break;
                    case JAROWINKLER: sim = jaro_winkler(s1, s2); break;
                    case LEVENSHTEIN: sim = normalized_levenshtein(s1, s2); break;
                    default: sim = date(s1, s2, method, smooth, distance, time_dir); break;
                }
            } catch (const JobDied&) { died = true; break; }
            if (sim >= threshold) { ov.push_back(vb); os.push_back((float)sim); }  // :46-49
        }
        if (died) continue;  // the future throws, the consumer loop gets no result for this literal (:176-178)
        for (size_t k = 0; k < ov.size(); k++, count++)
            if (count < cap) { out_v[count] = va; out_o[count] = ov[k]; out_s[count] = os[k]; }
    }
    return count;
}

// pieces exposed for direct known-answer tests
int32_t orc_sim_parse_int(const uint16_t* s, int32_t len, int32_t* out) { return parse_int(ustr((const char16_t*)s, (size_t)len), out) ? 1 : 0; }
int32_t orc_sim_parse_iso(const uint16_t* s, int32_t len, int64_t* epoch, int64_t* packed) {
    Ymd t;
    if (!parse_basic_iso(ustr((const char16_t*)s, (size_t)len), &t)) return 0;
    *epoch = epoch_day(t);
    *packed = (t.y * 12 + t.m - 1) * 32 + t.d;
    return 1;
}
// profile of one label as (key string, count) pairs flattened: returns the number of entries;
// keys are written back to back as UTF-16 into key_chars with key_off[0..n]
int32_t orc_sim_profile(int32_t method, int32_t ngram, const uint16_t* s, int32_t len, int32_t cap_entries, int32_t cap_chars,
                        uint16_t* key_chars, int32_t* key_off, int32_t* cnt) {
    const ustr str((const char16_t*)s, (size_t)len);
    const Profile p = (method == NGRAM_COSINE || method == NGRAM_JACCARD) ? ngram_profile(str, ngram) : token_profile(str);
    int32_t k = 0, at = 0;
    for (auto& e : p) {
        if (k >= cap_entries || at + (int32_t)e.first.size() > cap_chars) return -1;
        key_off[k] = at;
        for (char16_t c : e.first) key_chars[at++] = (uint16_t)c;
        cnt[k++] = e.second;
    }
    key_off[k] = at;
    return k;
}

}  // extern "C"
