// the sparse ENCODER the reference runs before it talks to the vector store --
// fastembed SparseTextEmbedding("Qdrant/bm25").embed([text]) at embedding_handler.py:41,123-138.  Host-only code
// (no kernel here): the hot path takes the (index, value) lists this produces.
//
// The algorithm lives in fastembed (not vendored in /root/reference, version unpinned); restated from its published
// behaviour: text -> non-word characters to spaces -> lower-case -> whitespace split   [done by the Python host with
// the same regular expressions, rag_application_b200/bm25_encoder.py]   -> drop punctuation tokens, English stop
// words and tokens longer than 40 characters -> Snowball English (Porter2) stem -> id = |murmur3_x86_32(stem, 0)| ->
// value = tf*(k1+1) / (tf + k1*(1 - b + b*len/avgdl)), len = number of kept tokens; ids in first-occurrence order.
// The stemmer works on bytes: exact for ASCII words; a non-ASCII letter counts as one consonant per UTF-8 byte.
// Parity: unpinned (no fastembed here); known-answer tests in tests/test_bm25_encoder.py use the published Porter2
// and MurmurHash3 vectors.
#include <stdint.h>
#include <string.h>

#include <string>
#include <unordered_map>
#include <unordered_set>
#include <vector>

#include "../../include/brag.h"

namespace {

// ------------------------------------------------------------------ MurmurHash3 x86_32
uint32_t rotl32(uint32_t x, int r) { return (x << r) | (x >> (32 - r)); }

uint32_t murmur3_32(const uint8_t* data, int len, uint32_t seed) {
    const uint32_t c1 = 0xcc9e2d51u, c2 = 0x1b873593u;
    uint32_t h = seed;
    const int nblocks = len / 4;
    for (int i = 0; i < nblocks; ++i) {
        uint32_t k;
        memcpy(&k, data + 4 * i, 4);
        k *= c1; k = rotl32(k, 15); k *= c2;
        h ^= k; h = rotl32(h, 13); h = h * 5 + 0xe6546b64u;
    }
    const uint8_t* tail = data + 4 * nblocks;
    uint32_t k = 0;
    switch (len & 3) {
        case 3: k ^= (uint32_t)tail[2] << 16;  // fallthrough
        case 2: k ^= (uint32_t)tail[1] << 8;   // fallthrough
        case 1: k ^= tail[0]; k *= c1; k = rotl32(k, 15); k *= c2; h ^= k;
    }
    h ^= (uint32_t)len;
    h ^= h >> 16; h *= 0x85ebca6bu; h ^= h >> 13; h *= 0xc2b2ae35u; h ^= h >> 16;
    return h;
}

// ------------------------------------------------------------------ Snowball English (Porter2)
struct Stemmer {
    std::string w;
    size_t r1 = 0, r2 = 0;

    static bool vowel(char c) { return c == 'a' || c == 'e' || c == 'i' || c == 'o' || c == 'u' || c == 'y'; }
    bool ends(const char* s) const {
        const size_t n = strlen(s);
        return w.size() >= n && w.compare(w.size() - n, n, s) == 0;
    }
    bool ends_in(const char* s, size_t region) const { return ends(s) && w.size() - strlen(s) >= region; }
    void chop(size_t n) { w.erase(w.size() - n); }
    void replace_end(size_t n, const char* s) { w.erase(w.size() - n); w += s; }
    bool has_vowel_before(size_t end) const {
        for (size_t i = 0; i < end; ++i) if (vowel(w[i])) return true;
        return false;
    }
    // short syllable ending at position `end` (exclusive)
    bool short_syllable_at(size_t end) const {
        if (end == 2) return vowel(w[0]) && !vowel(w[1]);
        if (end >= 3) {
            const char c = w[end - 1];
            return !vowel(w[end - 3]) && vowel(w[end - 2]) && !vowel(c) && c != 'w' && c != 'x' && c != 'Y';
        }
        return false;
    }
    bool is_short() const { return r1 >= w.size() && short_syllable_at(w.size()); }

    void mark_regions() {
        const size_t n = w.size();
        r1 = n;
        static const char* pre[] = {"gener", "commun", "arsen"};
        bool special = false;
        for (const char* p : pre) {
            const size_t l = strlen(p);
            if (w.compare(0, l, p) == 0) { r1 = l; special = true; break; }
        }
        if (!special) {
            for (size_t i = 1; i < n; ++i)
                if (!vowel(w[i]) && vowel(w[i - 1])) { r1 = i + 1; break; }
        }
        r2 = n;
        for (size_t i = r1 + 1; i < n; ++i)
            if (!vowel(w[i]) && vowel(w[i - 1])) { r2 = i + 1; break; }
    }

    void step0() {
        if (ends("'s'")) chop(3);
        else if (ends("'s")) chop(2);
        else if (ends("'")) chop(1);
    }
    void step1a() {
        if (ends("sses")) { chop(2); return; }
        if (ends("ied") || ends("ies")) { if (w.size() > 4) chop(2); else chop(1); return; }
        if (ends("us") || ends("ss")) return;
        if (ends("s")) {
            // delete if the preceding word part contains a vowel not immediately before the s
            if (w.size() >= 2 && has_vowel_before(w.size() - 2)) chop(1);
        }
    }
    void step1b() {
        if (ends("eedly")) { if (ends_in("eedly", r1)) replace_end(5, "ee"); return; }
        if (ends("eed")) { if (ends_in("eed", r1)) replace_end(3, "ee"); return; }
        size_t n = 0;
        if (ends("ingly")) n = 5;
        else if (ends("edly")) n = 4;
        else if (ends("ing")) n = 3;
        else if (ends("ed")) n = 2;
        if (n == 0 || !has_vowel_before(w.size() - n)) return;
        chop(n);
        if (ends("at") || ends("bl") || ends("iz")) { w += 'e'; return; }
        static const char* dbl[] = {"bb", "dd", "ff", "gg", "mm", "nn", "pp", "rr", "tt"};
        for (const char* d : dbl) if (ends(d)) { chop(1); return; }
        if (is_short()) w += 'e';
    }
    void step1c() {
        const size_t n = w.size();
        if (n > 2 && (w[n - 1] == 'y' || w[n - 1] == 'Y') && !vowel(w[n - 2])) w[n - 1] = 'i';
    }
    void step2() {
        struct R { const char* s; const char* r; };
        // longest match first among suffixes sharing an ending
        static const R rules[] = {
            {"ization", "ize"}, {"ational", "ate"}, {"fulness", "ful"}, {"ousness", "ous"}, {"iveness", "ive"},
            {"tional", "tion"}, {"biliti", "ble"}, {"lessli", "less"}, {"entli", "ent"}, {"ation", "ate"},
            {"alism", "al"}, {"aliti", "al"}, {"ousli", "ous"}, {"iviti", "ive"}, {"fulli", "ful"},
            {"enci", "ence"}, {"anci", "ance"}, {"abli", "able"}, {"izer", "ize"}, {"ator", "ate"},
            {"alli", "al"}, {"bli", "ble"}};
        for (const R& r : rules) {
            if (ends(r.s)) { if (ends_in(r.s, r1)) replace_end(strlen(r.s), r.r); return; }
        }
        if (ends("ogi")) { if (ends_in("ogi", r1) && w.size() >= 4 && w[w.size() - 4] == 'l') chop(1); return; }
        if (ends("li")) {
            if (ends_in("li", r1) && w.size() >= 3 && strchr("cdeghkmnrt", w[w.size() - 3])) chop(2);
        }
    }
    void step3() {
        struct R { const char* s; const char* r; };
        static const R rules[] = {{"ational", "ate"}, {"tional", "tion"}, {"alize", "al"}, {"icate", "ic"},
                                  {"iciti", "ic"}, {"ical", "ic"}, {"ness", ""}, {"ful", ""}};
        for (const R& r : rules) {
            if (ends(r.s)) { if (ends_in(r.s, r1)) replace_end(strlen(r.s), r.r); return; }
        }
        if (ends("ative")) { if (ends_in("ative", r1) && ends_in("ative", r2)) chop(5); }
    }
    void step4() {
        static const char* sfx[] = {"ement", "ance", "ence", "able", "ible", "ment", "ant", "ent", "ism", "ate",
                                    "iti", "ous", "ive", "ize", "al", "er", "ic"};
        // "ion" competes by length with the 3-letter suffixes: none of them ends in "ion", so order is free
        for (const char* s : sfx) {
            if (ends(s)) { if (ends_in(s, r2)) chop(strlen(s)); return; }
        }
        if (ends("ion")) {
            if (ends_in("ion", r2) && w.size() >= 4 && (w[w.size() - 4] == 's' || w[w.size() - 4] == 't')) chop(3);
        }
    }
    void step5() {
        const size_t n = w.size();
        if (n == 0) return;
        if (w[n - 1] == 'e') {
            if (n - 1 >= r2 || (n - 1 >= r1 && !short_syllable_at(n - 1))) chop(1);
        } else if (w[n - 1] == 'l') {
            if (n - 1 >= r2 && n >= 2 && w[n - 2] == 'l') chop(1);
        }
    }

    std::string stem(const std::string& in) {
        w = in;
        if (w.size() <= 2) return w;
        static const std::unordered_map<std::string, std::string> exc1 = {
            {"skis", "ski"}, {"skies", "sky"}, {"dying", "die"}, {"lying", "lie"}, {"tying", "tie"},
            {"idly", "idl"}, {"gently", "gentl"}, {"ugly", "ugli"}, {"early", "earli"}, {"only", "onli"},
            {"singly", "singl"}, {"sky", "sky"}, {"news", "news"}, {"howe", "howe"}, {"atlas", "atlas"},
            {"cosmos", "cosmos"}, {"bias", "bias"}, {"andes", "andes"}};
        auto it = exc1.find(w);
        if (it != exc1.end()) return it->second;
        if (w[0] == '\'') w.erase(0, 1);
        if (w.empty()) return w;
        if (w[0] == 'y') w[0] = 'Y';
        for (size_t i = 1; i < w.size(); ++i)
            if (w[i] == 'y' && vowel(w[i - 1])) w[i] = 'Y';
        mark_regions();
        step0();
        step1a();
        static const std::unordered_set<std::string> exc2 = {"inning", "outing", "canning", "herring", "earring",
                                                             "proceed", "exceed", "succeed"};
        if (!exc2.count(w)) {
            step1b();
            step1c();
            step2();
            step3();
            step4();
            step5();
        }
        for (char& c : w) if (c == 'Y') c = 'y';
        return w;
    }
};

const std::unordered_set<std::string>& stopwords() {
    // NLTK English stop words, the list fastembed ships with Qdrant/bm25
    static const std::unordered_set<std::string> s = {
        "i", "me", "my", "myself", "we", "our", "ours", "ourselves", "you", "you're", "you've", "you'll", "you'd",
        "your", "yours", "yourself", "yourselves", "he", "him", "his", "himself", "she", "she's", "her", "hers",
        "herself", "it", "it's", "its", "itself", "they", "them", "their", "theirs", "themselves", "what", "which",
        "who", "whom", "this", "that", "that'll", "these", "those", "am", "is", "are", "was", "were", "be", "been",
        "being", "have", "has", "had", "having", "do", "does", "did", "doing", "a", "an", "the", "and", "but", "if",
        "or", "because", "as", "until", "while", "of", "at", "by", "for", "with", "about", "against", "between",
        "into", "through", "during", "before", "after", "above", "below", "to", "from", "up", "down", "in", "out",
        "on", "off", "over", "under", "again", "further", "then", "once", "here", "there", "when", "where", "why",
        "how", "all", "any", "both", "each", "few", "more", "most", "other", "some", "such", "no", "nor", "not",
        "only", "own", "same", "so", "than", "too", "very", "s", "t", "can", "will", "just", "don", "don't",
        "should", "should've", "now", "d", "ll", "m", "o", "re", "ve", "y", "ain", "aren", "aren't", "couldn",
        "couldn't", "didn", "didn't", "doesn", "doesn't", "hadn", "hadn't", "hasn", "hasn't", "haven", "haven't",
        "isn", "isn't", "ma", "mightn", "mightn't", "mustn", "mustn't", "needn", "needn't", "shan", "shan't",
        "shouldn", "shouldn't", "wasn", "wasn't", "weren", "weren't", "won", "won't", "wouldn", "wouldn't"};
    return s;
}

bool is_punct_token(const std::string& t) {
    if (t.size() != 1) return false;
    return strchr("!\"#$%&'()*+,-./:;<=>?@[\\]^_`{|}~", t[0]) != nullptr;
}

int utf8_chars(const std::string& t) {
    int n = 0;
    for (unsigned char c : t) n += (c & 0xC0) != 0x80;
    return n;
}

}  // namespace

extern "C" {

uint32_t brag_murmur3_32(const void* data, int32_t len, uint32_t seed) {
    return murmur3_32(reinterpret_cast<const uint8_t*>(data), len, seed);
}

int brag_bm25_stem(const char* word, char* out, int32_t cap) {
    if (!word || !out || cap < 1) return 1;
    Stemmer st;
    const std::string r = st.stem(word);
    if ((int)r.size() + 1 > cap) return 1;
    memcpy(out, r.c_str(), r.size() + 1);
    return 0;
}

int brag_bm25_encode(const char* tokens, const int32_t* tok_off, int32_t n_tokens, double k1, double b, double avgdl,
                     int32_t max_token_len, int32_t query_mode, int64_t* out_ids, double* out_vals, int32_t* out_n) {
    if (!tok_off || !out_ids || !out_vals || !out_n || n_tokens < 0 || (n_tokens > 0 && !tokens)) return 1;
    Stemmer st;
    std::vector<std::string> order;
    std::unordered_map<std::string, int> tf;
    int doc_len = 0;
    for (int i = 0; i < n_tokens; ++i) {
        const std::string tok(tokens + tok_off[i], tokens + tok_off[i + 1]);
        if (tok.empty() || is_punct_token(tok) || stopwords().count(tok) || utf8_chars(tok) > max_token_len) continue;
        const std::string s = st.stem(tok);
        if (s.empty()) continue;
        ++doc_len;
        auto it = tf.find(s);
        if (it == tf.end()) { tf.emplace(s, 1); order.push_back(s); }
        else ++it->second;
    }
    int n = 0;
    for (const std::string& s : order) {
        const int32_t h = (int32_t)murmur3_32(reinterpret_cast<const uint8_t*>(s.data()), (int)s.size(), 0u);
        const int64_t id = h < 0 ? -(int64_t)h : (int64_t)h;
        const double c = (double)tf[s];
        // several stems can hash to one id: like a Python dict, the last one written wins, position of the first
        int at = n;
        for (int j = 0; j < n; ++j) if (out_ids[j] == id) { at = j; break; }
        out_ids[at] = id;
        out_vals[at] = query_mode ? 1.0 : c * (k1 + 1.0) / (c + k1 * (1.0 - b + b * (double)doc_len / avgdl));
        if (at == n) ++n;
    }
    *out_n = n;
    return 0;
}

}  // extern "C"
