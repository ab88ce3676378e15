// tokenizer.cu -- see tokenizer.cuh.  Host code only.
//
// What has to equal the reference's output, and where it comes from there:
//   * SPM: bigram merging by score, ties to the leftmost pair, bytes <0xXX> for what is left (llama.cpp:5970-6110)
//   * BPE: GPT-2 pre-tokenizer as a hand-rolled state machine with its quirks (contractions keep the collecting state,
//     the last character is always glued to the pending word; llama.cpp:6249-6386), byte -> printable-codepoint
//     alphabet (unicode.h:405-431), merging by rank, ties to the leftmost pair (llama.cpp:6101-6219)
//   * special tokens: a token is "special" when no split of its text into two vocabulary entries exists; the text is
//     cut at their occurrences, one special token after the other in the iteration order of the hash map they are
//     cached in (llama.cpp:2538-2625, 6431-6524).  That order is libstdc++'s: the two maps below are the same container
//     types, filled in the same order, so the order is the same.
#include "tokenizer.cuh"

#include <limits.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <map>
#include <queue>
#include <unordered_map>

#include "engine.cuh"

struct pib_vocab {
    int type = PIB_VOCAB_SPM;
    int bos = -1, linefeed = -1;
    const pib_model * m = nullptr;
    std::unordered_map<std::string, int32_t>            index;      // piece text -> id (the reference's token_to_id)
    std::unordered_map<std::string, int32_t>            special;    // special-token cache
    std::map<std::pair<std::string, std::string>, int>  ranks;      // BPE merge -> rank
    std::string                                         byte_sym[256];    // byte -> its printable stand-in (UTF-8)
    std::unordered_map<uint32_t, uint8_t>               sym_byte;         // stand-in codepoint -> byte
};

// ---- UTF-8 ------------------------------------------------------------------------------------------------------------
// length a lead byte announces; continuation bytes count 1, 0xF8.. count 4 (the reference's utf8_len, llama.cpp:233-237)
static inline size_t lead_len(unsigned char c) { return c < 0xC0 ? 1 : c < 0xE0 ? 2 : c < 0xF0 ? 3 : 4; }

static void put_utf8(uint32_t cp, std::string & s) {
    if (cp < 0x80) { s += (char) cp; return; }
    if (cp < 0x800) { s += (char) (0xC0 | (cp >> 6)); s += (char) (0x80 | (cp & 0x3F)); return; }
    if (cp < 0x10000) { s += (char) (0xE0 | (cp >> 12)); s += (char) (0x80 | ((cp >> 6) & 0x3F)); s += (char) (0x80 | (cp & 0x3F)); return; }
    s += (char) (0xF0 | ((cp >> 18) & 7)); s += (char) (0x80 | ((cp >> 12) & 0x3F)); s += (char) (0x80 | ((cp >> 6) & 0x3F)); s += (char) (0x80 | (cp & 0x3F));
}

// strict about structure (lead byte, continuation bytes present), lenient about overlong forms -- like unicode.h:259-291
static bool get_utf8(const std::string & s, std::vector<uint32_t> & cps) {
    for (size_t i = 0; i < s.size();) {
        const unsigned char c = (unsigned char) s[i];
        int n; uint32_t cp;
        if (c < 0x80) { n = 0; cp = c; }
        else if (c < 0xC0) return false;
        else if (c < 0xE0) { n = 1; cp = c & 0x1F; }
        else if (c < 0xF0) { n = 2; cp = c & 0x0F; }
        else if (c < 0xF8) { n = 3; cp = c & 0x07; }
        else return false;
        if (n && i + n >= s.size()) return false;
        for (int k = 1; k <= n; k++) {
            const unsigned char t = (unsigned char) s[i + k];
            if ((t & 0xC0) != 0x80) return false;
            cp = (cp << 6) | (t & 0x3F);
        }
        cps.push_back(cp);
        i += n + 1;
    }
    return true;
}

// ---- codepoint classes (generated table, see tools/gen_unicode_classes.py) --------------------------------------------
enum { CL_OTHER = 0, CL_DIGIT = 1, CL_LETTER = 2, CL_SPACE = 3 };
struct cp_run { uint32_t first, last; int cls; };
static const cp_run cp_runs[] = {
#include "unicode_classes.inc"
};

static int cp_class(uint32_t cp) {
    const cp_run * lo = cp_runs, * hi = cp_runs + sizeof(cp_runs) / sizeof(cp_runs[0]);
    while (lo < hi) {
        const cp_run * mid = lo + (hi - lo) / 2;
        if (cp > mid->last) lo = mid + 1; else hi = mid;
    }
    return lo < cp_runs + sizeof(cp_runs) / sizeof(cp_runs[0]) && cp >= lo->first ? lo->cls : CL_OTHER;
}

extern "C" int pib_tokenizer_codepoint_class(uint32_t cp) { return cp_class(cp); }      // include/pipeinfer_llama.h

// ---- building the vocabulary ------------------------------------------------------------------------------------------
static void tokenize_raw(const pib_vocab * v, const std::string & text, bool first_fragment, std::vector<int32_t> & out, bool & ok);

// GPT-2's byte alphabet: printable Latin-1 bytes stand for themselves, the other 68 take U+0100.. in byte order
static void build_byte_alphabet(pib_vocab * v) {
    int next = 256;
    for (int b = 0; b < 256; b++) {
        const bool keeps = (b >= '!' && b <= '~') || (b >= 0xA1 && b <= 0xAC) || (b >= 0xAE && b <= 0xFF);
        const uint32_t cp = keeps ? (uint32_t) b : (uint32_t) next++;
        put_utf8(cp, v->byte_sym[b]);
        v->sym_byte[cp] = (uint8_t) b;
    }
}

// no way to write the token's text as two vocabulary entries (cut only at character boundaries) and more than one
// character long => the tokenizer can never produce it from text => it is a special token (llama.cpp:2556-2617)
static bool is_special_text(const pib_vocab * v, const std::string & t) {
    if (t.size() < 2) return false;
    for (size_t cut = 1; cut < t.size();) {
        const size_t step = lead_len((unsigned char) t[cut - 1]);       // the reference looks at the byte before the cut
        if (step == 1) {
            if (v->index.count(t.substr(0, cut)) && v->index.count(t.substr(cut))) return false;
            cut++;
        } else {
            cut += step - 1;
        }
    }
    size_t n_chars = 0;
    for (size_t i = 0; i < t.size(); i += lead_len((unsigned char) t[i])) n_chars++;
    return n_chars > 1;
}

pib_vocab * pib_vocab_build(const pib_model * m) {
    pib_vocab * v = new pib_vocab();
    v->m    = m;
    v->type = m->tok_model == "gpt2" ? PIB_VOCAB_BPE : PIB_VOCAB_SPM;      // anything else: SPM, like llama.cpp:2462-2467
    v->bos  = m->tok_bos;
    for (size_t i = 0; i < m->tok_text.size(); i++) v->index[m->tok_text[i]] = (int32_t) i;
    if (v->type == PIB_VOCAB_BPE) {
        build_byte_alphabet(v);
        for (size_t r = 0; r < m->tok_merges.size(); r++) {
            const std::string & line = m->tok_merges[r];
            const size_t sp = line.find(' ', 1);
            std::pair<std::string, std::string> key;
            if (sp != std::string::npos) key = { line.substr(0, sp), line.substr(sp + 1) };
            v->ranks.emplace(key, (int) r);                              // the first rank of a repeated pair wins
        }
        // the newline token: whatever "Ċ" tokenizes to first (llama.cpp:2491-2493)
        std::vector<int32_t> ids;
        bool ok = true;
        tokenize_raw(v, "\xC4\x8A", true, ids, ok);
        v->linefeed = ok && !ids.empty() ? ids[0] : -1;
    } else {
        auto it = v->index.find("<0x0A>");
        v->linefeed = it == v->index.end() ? -1 : it->second;
    }
    for (const auto & e : v->index) if (is_special_text(v, e.first)) v->special[e.first] = e.second;
    return v;
}

void pib_vocab_free(pib_vocab * v) { delete v; }
int  pib_vocab_type(const pib_vocab * v) { return v->type; }
int  pib_vocab_linefeed(const pib_vocab * v) { return v->linefeed; }
int  pib_vocab_n_special(const pib_vocab * v) { return (int) v->special.size(); }

// ---- merging: one doubly linked list of symbols over a word, a heap of candidate pairs --------------------------------
struct tk_sym { size_t off, n; int prev, next; };

static void split_chars(const std::string & s, std::vector<tk_sym> & syms) {
    syms.clear();
    for (size_t off = 0; off < s.size();) {
        const size_t n = std::min(lead_len((unsigned char) s[off]), s.size() - off);
        syms.push_back({ off, n, (int) syms.size() - 1, off + n >= s.size() ? -1 : (int) syms.size() + 1 });
        off += n;
    }
}

// SentencePiece flavour.  s is already escaped (U+2581 for ' ').
static void spm_word(const pib_vocab * v, const std::string & s, std::vector<int32_t> & out, bool & ok) {
    const pib_model * m = v->m;
    std::vector<tk_sym> syms;
    split_chars(s, syms);
    struct cand { int left, right; float score; size_t size; };
    auto worse = [](const cand & a, const cand & b) { return a.score < b.score || (a.score == b.score && a.left > b.left); };
    std::priority_queue<cand, std::vector<cand>, decltype(worse)> heap(worse);
    auto offer = [&](int l, int r) {
        if (l < 0 || r < 0) return;
        auto it = v->index.find(s.substr(syms[l].off, syms[l].n + syms[r].n));
        if (it == v->index.end() || (size_t) it->second >= m->tok_text.size()) return;
        heap.push({ l, r, (size_t) it->second < m->tok_score.size() ? m->tok_score[it->second] : 0.f, syms[l].n + syms[r].n });
    };
    for (int i = 1; i < (int) syms.size(); i++) offer(i - 1, i);
    while (!heap.empty()) {
        const cand c = heap.top();
        heap.pop();
        tk_sym & L = syms[c.left];
        tk_sym & R = syms[c.right];
        if (L.n == 0 || R.n == 0 || L.n + R.n != c.size) continue;        // one side was merged away since
        L.n += R.n; R.n = 0;
        L.next = R.next;
        if (R.next >= 0) syms[R.next].prev = c.left;
        offer(L.prev, c.left);
        offer(c.left, L.next);
    }
    for (int i = syms.empty() ? -1 : 0; i != -1; i = syms[i].next) {
        const std::string piece = s.substr(syms[i].off, syms[i].n);
        auto it = v->index.find(piece);
        if (it != v->index.end()) { out.push_back(it->second); continue; }
        for (unsigned char ch : piece) {                                  // byte tokens <0xXX>
            char name[8];
            snprintf(name, sizeof(name), "<0x%02X>", ch);
            auto bt = v->index.find(name);
            if (bt == v->index.end()) { pib_set_error("tokenize: the vocabulary has no token %s", name); ok = false; return; }
            out.push_back(bt->second);
        }
    }
}

// GPT-2 pre-tokenizer: cuts the text where the regex  's|'t|'re|'ve|'m|'ll|'d| ?\p{L}+| ?\p{N}+| ?[^\s\p{L}\p{N}]+|\s+(?!\S)|\s+
// would, AS THE REFERENCE'S STATE MACHINE DOES IT (llama.cpp:6249-6375), including where that differs from the regex:
//   - after a contraction the machine stays in the run it was in (the next character usually ends that run at once),
//   - the last character of the text is appended to the pending word before the word is closed, whatever its class.
// Words come back as [first, last) ranges of codepoint indices.
static void gpt2_words(const std::vector<uint32_t> & cp, std::vector<std::pair<int, int>> & words) {
    enum { IDLE, LETTERS, DIGITS, OTHERS, SPACES } run = IDLE;
    const int n = (int) cp.size();
    std::vector<uint8_t> cls(n + 1, CL_OTHER);                            // cls[n]: "no next character" has class 0
    for (int i = 0; i < n; i++) cls[i] = (uint8_t) cp_class(cp[i]);
    auto is = [&](int i, char c) { return i < n && cp[i] == (uint32_t) (unsigned char) c; };
    int start = 0;                                                        // the pending word is [start, i)
    for (int i = 0; i < n; i++) {
        if (is(i, '\'')) {
            int len = 0;
            if (i + 1 < n && (is(i + 1, 's') || is(i + 1, 't') || is(i + 1, 'm') || is(i + 1, 'd'))) len = 2;
            else if (i + 2 < n && ((is(i + 1, 'r') && is(i + 2, 'e')) || (is(i + 1, 'v') && is(i + 2, 'e')) || (is(i + 1, 'l') && is(i + 2, 'l')))) len = 3;
            if (len) {
                if (start < i) words.push_back({ start, i });
                words.push_back({ i, i + len });
                i += len - 1;
                start = i + 1;
                continue;                                                 // `run` is deliberately left as it was
            }
        }
        const int  c = cls[i], nx = cls[i + 1];
        const bool lead_space = start == i && is(i, ' ');                 // a single space may open a word of the next class
        bool cut = false;
        if (run == IDLE) {
            if (c == CL_LETTER || (lead_space && nx == CL_LETTER)) run = LETTERS;
            else if (c == CL_DIGIT || (lead_space && nx == CL_DIGIT)) run = DIGITS;
            else if (c == CL_OTHER || (lead_space && nx == CL_OTHER)) run = OTHERS;
            else if (nx == CL_SPACE) run = SPACES;                        // c is whitespace from here on
            else cut = true;
        } else {
            cut = (run == LETTERS && c != CL_LETTER) || (run == DIGITS && c != CL_DIGIT) || (run == OTHERS && c != CL_OTHER) ||
                  (run == SPACES && (nx == CL_LETTER || nx == CL_DIGIT));
        }
        if (i + 1 == n) {                                                 // the last character joins the pending word
            words.push_back({ start, n });
            break;
        }
        if (cut) {
            if (start < i) words.push_back({ start, i });
            start = i;
            run   = IDLE;
        }
    }
}

// byte-level BPE over one raw-text fragment
static void bpe_fragment(const pib_vocab * v, const std::string & text, std::vector<int32_t> & out, bool & ok) {
    std::vector<uint32_t> cp;
    if (!get_utf8(text, cp)) { pib_set_error("tokenize: the text is not valid UTF-8"); ok = false; return; }
    std::vector<std::pair<int, int>> words;
    gpt2_words(cp, words);
    struct cand { int left, right, rank; std::string text; };
    auto later = [](const cand & a, const cand & b) { return a.rank > b.rank || (a.rank == b.rank && a.left > b.left); };
    std::vector<tk_sym> syms;
    std::string raw, w;
    for (const auto & range : words) {
        raw.clear();
        for (int i = range.first; i < range.second; i++) put_utf8(cp[i], raw);
        w.clear();
        for (unsigned char b : raw) w += v->byte_sym[b];
        split_chars(w, syms);
        std::priority_queue<cand, std::vector<cand>, decltype(later)> heap(later);
        auto text_of = [&](int i) { return w.substr(syms[i].off, syms[i].n); };
        auto offer = [&](int l, int r) {
            if (l < 0 || r < 0) return;
            auto it = v->ranks.find({ text_of(l), text_of(r) });
            if (it != v->ranks.end()) heap.push({ l, r, it->second, text_of(l) + text_of(r) });
        };
        for (int i = 1; i < (int) syms.size(); i++) offer(i - 1, i);
        while (!heap.empty()) {
            const cand c = heap.top();
            heap.pop();
            tk_sym & L = syms[c.left];
            tk_sym & R = syms[c.right];
            if (L.n == 0 || R.n == 0 || text_of(c.left) + text_of(c.right) != c.text) continue;     // stale candidate
            L.n += R.n; R.n = 0;
            L.next = R.next;
            if (R.next >= 0) syms[R.next].prev = c.left;
            offer(L.prev, c.left);
            offer(c.left, L.next);
        }
        for (const tk_sym & s : syms) {
            if (s.n == 0) continue;
            const std::string piece = w.substr(s.off, s.n);
            auto it = v->index.find(piece);
            if (it != v->index.end()) { out.push_back(it->second); continue; }
            for (char ch : piece) {                                       // byte by byte, as the reference does
                auto bt = v->index.find(std::string(1, ch));
                if (bt == v->index.end()) { pib_set_error("tokenize: byte 0x%02X of an unknown piece has no token", (unsigned char) ch); ok = false; return; }
                out.push_back(bt->second);
            }
        }
    }
}

static void tokenize_raw(const pib_vocab * v, const std::string & text, bool first_fragment, std::vector<int32_t> & out, bool & ok) {
    if (v->type == PIB_VOCAB_BPE) { bpe_fragment(v, text, out, ok); return; }
    std::string s;                                                        // ' ' -> U+2581; a leading one for the first fragment only
    if (first_fragment) s += "\xe2\x96\x81";
    for (char ch : text) { if (ch == ' ') s += "\xe2\x96\x81"; else s += ch; }
    spm_word(v, s, out, ok);
}

int pib_vocab_tokenize(const pib_vocab * v, const std::string & text, bool add_bos, bool special, std::vector<int32_t> & out) {
    if (add_bos && v->bos != -1) out.push_back(v->bos);
    if (text.empty()) return 0;
    struct frag { int32_t token; size_t off, len; };                      // token < 0: raw text [off, off+len)
    std::vector<frag> frags { { -1, 0, text.size() } };
    if (special) {
        for (const auto & st : v->special) {
            const std::string & name = st.first;
            std::vector<frag> next;
            for (const frag & f : frags) {
                if (f.token >= 0) { next.push_back(f); continue; }
                size_t pos = f.off;
                const size_t end = f.off + f.len;
                while (pos < end) {
                    const size_t hit = text.find(name, pos);
                    if (hit == std::string::npos || hit + name.size() > end) break;
                    if (hit > pos) next.push_back({ -1, pos, hit - pos });
                    next.push_back({ st.second, 0, 0 });
                    pos = hit + name.size();
                }
                if (pos < end) next.push_back({ -1, pos, end - pos });
            }
            frags.swap(next);
        }
    }
    bool ok = true;
    for (size_t i = 0; i < frags.size() && ok; i++) {
        if (frags[i].token >= 0) out.push_back(frags[i].token);
        else tokenize_raw(v, text.substr(frags[i].off, frags[i].len), i == 0, out, ok);
    }
    return ok ? 0 : PIB_TOKENIZE_INVALID;
}

bool pib_vocab_piece(const pib_vocab * v, int32_t token, std::string & out) {
    const pib_model * m = v->m;
    out.clear();
    if (token < 0 || (size_t) token >= m->tok_text.size()) return false;
    const int kind = (size_t) token < m->tok_type.size() ? m->tok_type[token] : 1;      // llama.h:75-83; 1 = normal
    const std::string & t = m->tok_text[token];
    if (v->type == PIB_VOCAB_BPE) {
        if (kind != 1) return false;
        std::vector<uint32_t> cp;
        if (!get_utf8(t, cp)) return false;
        for (uint32_t c : cp) {
            auto it = v->sym_byte.find(c);
            if (it == v->sym_byte.end()) { pib_set_error("token_to_piece: token %d is not written in the byte alphabet", token); out.clear(); return false; }
            out += (char) it->second;
        }
        return true;
    }
    if (kind == 1) {                                                      // U+2581 -> ' '
        for (size_t i = 0; i < t.size();) {
            if (t.compare(i, 3, "\xe2\x96\x81") == 0) { out += ' '; i += 3; }
            else out += t[i++];
        }
        return true;
    }
    if (kind == 2) { out = "\xe2\x96\x85"; return true; }                 // unknown
    if (kind == 6 && t.size() >= 5) { out = std::string(1, (char) strtol(t.substr(3, 2).c_str(), nullptr, 16)); return true; }       // <0xXX>
    return false;
}
