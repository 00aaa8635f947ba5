// Host-side batched text front-end (no device code): HTS full-context labels -> (phoneme ids, accent ids) padded batches.
//
// Stands behind the Japanese branch of the reference's tokenisation, which runs per sentence in Python with ~7 regex
// searches per label and becomes the bottleneck once synthesis runs at > 10^4 audio-s/s:
//   prepare_tg_accent_jsut.py:18-70  pp_symbols            full-context label -> phoneme + accent mark ('0' '[' ']' '#')
//   convert_label.py:52-59           openjtalk2julius      A/I/U/E/O -> lower case, cl -> q, pau -> sp
//   synthesize.py:101,217-218        id lookup             symbol_to_id[p] (last occurrence of a symbol wins) and
//                                                          {'0':0, '[':1, ']':2, '#':3}
// Utterances are independent: a small pool of std::threads splits the batch.  The scanner below restates the regular
// expressions as single forward passes over the label (each rule cites the expression it replaces).
#include "common.cuh"
#include <cstring>
#include <thread>
#include <atomic>

struct masg_frontend {
    std::unordered_map<std::string, int64_t> sym;
    std::string err;
};

namespace {

constexpr int NOT_FOUND = -50;                   // numeric_feature_by_regex's "no match" value (prepare_tg_accent_jsut.py:13-17)

struct Span { const char* p; size_t n; };

inline bool is_digit(char c) { return c >= '0' && c <= '9'; }

// re.search(r"\-(.*?)\+", s).group(1): text between the first '-' and the first '+' after it
inline bool field_p3(Span s, Span& out) {
    const char* a = (const char*)memchr(s.p, '-', s.n);
    if (!a) return false;
    const char* b = (const char*)memchr(a + 1, '+', s.n - (a + 1 - s.p));
    if (!b) return false;
    out.p = a + 1; out.n = (size_t)(b - a - 1);
    return true;
}
inline int to_int(const char* p, const char* e) {             // int(match.group(1)); digits with optional leading '-'
    bool neg = false; long v = 0;
    if (p < e && *p == '-') { neg = true; ++p; }
    for (; p < e; ++p) v = v * 10 + (*p - '0');
    return (int)(neg ? -v : v);
}
// re.search(r"/A:([0-9\-]+)\+", s)
inline int field_a1(Span s) {
    for (size_t i = 0; i + 3 < s.n; ++i) {
        if (s.p[i] != '/' || s.p[i + 1] != 'A' || s.p[i + 2] != ':') continue;
        size_t j = i + 3;
        while (j < s.n && (is_digit(s.p[j]) || s.p[j] == '-')) ++j;
        // greedy [0-9\-]+ then '+': backtracking cannot help (the class excludes '+'), so the run must end at a '+'
        if (j > i + 3 && j < s.n && s.p[j] == '+') {
            // Python's int() rejects strings like "1-2" or "-": the reference would raise there; labels never hold them
            return to_int(s.p + i + 3, s.p + j);
        }
    }
    return NOT_FOUND;
}
// re.search(r"\+(\d+)" + term, s): first '+', one or more digits, then `term`
inline int field_plus_digits(Span s, char term) {
    for (size_t i = 0; i + 2 < s.n; ++i) {
        if (s.p[i] != '+') continue;
        size_t j = i + 1;
        while (j < s.n && is_digit(s.p[j])) ++j;
        if (j > i + 1 && j < s.n && s.p[j] == term) return to_int(s.p + i + 1, s.p + j);
    }
    return NOT_FOUND;
}
// re.search(r"/F:(\d+)_", s)
inline int field_f1(Span s) {
    for (size_t i = 0; i + 3 < s.n; ++i) {
        if (s.p[i] != '/' || s.p[i + 1] != 'F' || s.p[i + 2] != ':') continue;
        size_t j = i + 3;
        while (j < s.n && is_digit(s.p[j])) ++j;
        if (j > i + 3 && j < s.n && s.p[j] == '_') return to_int(s.p + i + 3, s.p + j);
    }
    return NOT_FOUND;
}

struct UttResult { int status = 0; std::string msg; };   // 0 ok, 1 assertion (sil inside), 2 unknown symbol, 3 malformed, 4 too long

// one utterance: labels[0..n) -> ids/accents rows (zero padded to L_cap); returns the phoneme count
int64_t tokenize_one(const masg_frontend* fe, const std::vector<Span>& lab, size_t beg, size_t end, int L_cap, int64_t* ids,
                     int64_t* acc, int drop_unvoiced, UttResult& res) {
    const size_t N = end - beg;
    int64_t len = 0;
    std::string key;
    for (size_t n = 0; n < N; ++n) {
        const Span cur = lab[beg + n];
        Span p3;
        if (!field_p3(cur, p3)) { res.status = 3; res.msg = "label without a '-phoneme+' field: " + std::string(cur.p, std::min<size_t>(cur.n, 60)); return 0; }
        key.assign(p3.p, p3.n);
        // `p3 in "AEIOU"` is a substring test (the empty string included): lower-case unvoiced vowels
        if (drop_unvoiced && std::string("AEIOU").find(key) != std::string::npos)
            for (auto& ch : key) ch = (char)tolower((unsigned char)ch);
        int mark = 0;
        if (key == "sil") {
            if (!(n == 0 || n == N - 1)) { res.status = 1; res.msg = "'sil' inside an utterance (pp_symbols asserts n == 0 or n == N-1)"; return 0; }
            continue;                                           // the "" it may append is dropped by `if p != ''` (synthesize.py:100)
        } else if (key == "pau") {
            key = "sp"; mark = 0;                               // PP.append("sp"); accent.append('0')
        } else {
            if (n + 1 >= N) { res.status = 3; res.msg = "utterance does not end in 'sil' (labels[n + 1] would raise IndexError)"; return 0; }
            const int a1 = field_a1(cur), a2 = field_plus_digits(cur, '+'), a3 = field_plus_digits(cur, '/'), f1 = field_f1(cur);
            const int a2_next = field_plus_digits(lab[beg + n + 1], '+');
            if (a3 == 1 && a2_next == 1) mark = 3;                                  // '#': accent-phrase boundary
            else if (a1 == 0 && a2_next == a2 + 1 && a2 != f1) mark = 2;            // ']': pitch fall (accent nucleus)
            else if (a2 == 1 && a2_next == 2) mark = 1;                             // '[': pitch rise
            else mark = 0;
            if (key.empty()) continue;                          // '' is filtered from the phonemes ... (its accent mark is not:
                                                                // cannot happen on well-formed labels, checked by the caller's len test)
        }
        // openjtalk2julius (convert_label.py:52-59)
        if (key.size() == 1 && strchr("AIUEO", key[0])) key[0] = (char)tolower((unsigned char)key[0]);
        else if (key == "cl") key = "q";
        else if (key == "pau") key = "sp";
        auto it = fe->sym.find(key);
        if (it == fe->sym.end()) { res.status = 2; res.msg = "unknown symbol '" + key + "' (KeyError in the reference)"; return 0; }
        if (len >= L_cap) { res.status = 4; res.msg = "utterance longer than L_cap"; return 0; }
        ids[len] = it->second; acc[len] = mark; ++len;
    }
    for (int64_t i = len; i < L_cap; ++i) { ids[i] = 0; acc[i] = 0; }
    return len;
}

thread_local std::string g_fe_err;

}  // namespace

extern "C" {

int masg_frontend_create(const char* const* symbols, int n_symbols, masg_frontend** out) {
    try {
        if (!symbols || n_symbols <= 0 || !out) { g_fe_err = "masg_frontend_create: bad argument"; return 1; }
        masg_frontend* fe = new masg_frontend();
        for (int i = 0; i < n_symbols; ++i) fe->sym[std::string(symbols[i])] = i;     // later duplicates win, as the dict comprehension does
        *out = fe;
        return 0;
    } catch (const std::exception& e) { g_fe_err = e.what(); return 2; }
}

void masg_frontend_destroy(masg_frontend* fe) { delete fe; }

const char* masg_frontend_last_error(masg_frontend* fe) { return fe ? fe->err.c_str() : g_fe_err.c_str(); }

int masg_frontend_tokenize(masg_frontend* fe, const char* blob, int64_t blob_bytes, const int64_t* labels_per_utt, int n_utts,
                           int L_cap, int drop_unvoiced_vowels, int n_threads, int64_t* texts, int64_t* accents, int64_t* src_lens) {
    if (!fe) { g_fe_err = "null masg_frontend"; return 1; }
    try {
        if (!blob || !labels_per_utt || !texts || !accents || !src_lens || n_utts < 0 || L_cap <= 0) { fe->err = "masg_frontend_tokenize: bad argument"; return 1; }
        // split the blob into lines ('\n'-separated, one label per line; a trailing '\r' is dropped)
        std::vector<Span> lab;
        lab.reserve(1024);
        const char* p = blob; const char* e = blob + blob_bytes;
        while (p < e) {
            const char* nl = (const char*)memchr(p, '\n', (size_t)(e - p));
            const char* q = nl ? nl : e;
            size_t n = (size_t)(q - p);
            if (n && p[n - 1] == '\r') --n;
            lab.push_back(Span{p, n});
            p = nl ? nl + 1 : e;
        }
        std::vector<size_t> start((size_t)n_utts + 1, 0);
        for (int u = 0; u < n_utts; ++u) {
            if (labels_per_utt[u] < 0) { fe->err = "negative label count"; return 1; }
            start[u + 1] = start[u] + (size_t)labels_per_utt[u];
        }
        if (start[n_utts] != lab.size()) {
            fe->err = "label counts (" + std::to_string(start[n_utts]) + ") do not match the lines of the blob (" + std::to_string(lab.size()) + ")";
            return 1;
        }
        std::vector<UttResult> res((size_t)n_utts);
        std::atomic<int> next(0);
        auto work = [&]() {
            for (;;) {
                const int u = next.fetch_add(1);
                if (u >= n_utts) break;
                src_lens[u] = tokenize_one(fe, lab, start[u], start[u + 1], L_cap, texts + (size_t)u * L_cap, accents + (size_t)u * L_cap,
                                           drop_unvoiced_vowels, res[u]);
            }
        };
        int nt = std::max(1, std::min(n_threads, n_utts));
        if (nt == 1) work();
        else {
            std::vector<std::thread> th;
            for (int i = 0; i < nt; ++i) th.emplace_back(work);
            for (auto& t : th) t.join();
        }
        for (int u = 0; u < n_utts; ++u)
            if (res[u].status) { fe->err = "utterance " + std::to_string(u) + ": " + res[u].msg; return 10 + res[u].status; }
        return 0;
    } catch (const std::exception& ex) { fe->err = ex.what(); return 2; }
}

}  // extern "C"
