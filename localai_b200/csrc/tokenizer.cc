// tokenizer.cc -- see tokenizer.h.  Algorithms follow llama.cpp's llm_tokenizer_spm / llm_tokenizer_bpe and the
// "llama-bpe" pre-tokeniser (un-vendored upstream, pin /root/reference/Makefile:11); checked in tests/ against
// the sentencepiece and HF tokenizers libraries on the committed golden vectors (tests/golden/tok_*.json).
#include "tokenizer.h"

#include <algorithm>
#include <cstring>
#include <queue>

#include "common.h"
#include "unicode_tables.h"

namespace b2 {

// ------------------------------------------------------------------------------------------ unicode helpers
static bool in_ranges(const CpRange *r, int n, uint32_t cp) {
    int lo = 0, hi = n - 1;
    while (lo <= hi) {
        const int mid = (lo + hi) / 2;
        if (cp < r[mid].lo) hi = mid - 1;
        else if (cp > r[mid].hi) lo = mid + 1;
        else return true;
    }
    return false;
}
static bool cp_is_letter(uint32_t c) { return in_ranges(kLetterRanges, kLetterRangesCount, c); }
static bool cp_is_number(uint32_t c) { return in_ranges(kNumberRanges, kNumberRangesCount, c); }
static bool cp_is_space(uint32_t c) { return in_ranges(kWhitespaceRanges, kWhitespaceRangesCount, c); }

static size_t utf8_len(uint8_t c) {
    static const size_t lookup[] = {1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 2, 2, 3, 4};
    return lookup[c >> 4];
}

// lenient decoder: invalid sequences yield U+FFFD and consume one byte
static std::vector<uint32_t> utf8_to_cpts(const std::string &s, std::vector<uint32_t> *byte_off = nullptr) {
    std::vector<uint32_t> out;
    size_t i = 0;
    const size_t n = s.size();
    while (i < n) {
        if (byte_off) byte_off->push_back((uint32_t)i);
        const uint8_t c = (uint8_t)s[i];
        auto cont = [&](size_t k) { return i + k < n && (((uint8_t)s[i + k]) & 0xC0) == 0x80; };
        if (c < 0x80) { out.push_back(c); i += 1; }
        else if ((c & 0xE0) == 0xC0 && cont(1)) { out.push_back(((c & 0x1F) << 6) | (s[i + 1] & 0x3F)); i += 2; }
        else if ((c & 0xF0) == 0xE0 && cont(1) && cont(2)) {
            out.push_back(((c & 0x0F) << 12) | ((s[i + 1] & 0x3F) << 6) | (s[i + 2] & 0x3F));
            i += 3;
        } else if ((c & 0xF8) == 0xF0 && cont(1) && cont(2) && cont(3)) {
            out.push_back(((c & 0x07) << 18) | ((s[i + 1] & 0x3F) << 12) | ((s[i + 2] & 0x3F) << 6) | (s[i + 3] & 0x3F));
            i += 4;
        } else { out.push_back(0xFFFD); i += 1; }
    }
    if (byte_off) byte_off->push_back((uint32_t)n);
    return out;
}

static std::string cpt_to_utf8(uint32_t cp) {
    std::string r;
    if (cp < 0x80) r += (char)cp;
    else if (cp < 0x800) { r += (char)(0xC0 | (cp >> 6)); r += (char)(0x80 | (cp & 0x3F)); }
    else if (cp < 0x10000) { r += (char)(0xE0 | (cp >> 12)); r += (char)(0x80 | ((cp >> 6) & 0x3F)); r += (char)(0x80 | (cp & 0x3F)); }
    else { r += (char)(0xF0 | (cp >> 18)); r += (char)(0x80 | ((cp >> 12) & 0x3F)); r += (char)(0x80 | ((cp >> 6) & 0x3F)); r += (char)(0x80 | (cp & 0x3F)); }
    return r;
}

// GPT-2 byte <-> unicode tables
struct ByteMap {
    std::string b2u[256];
    std::unordered_map<std::string, uint8_t> u2b;
    ByteMap() {
        bool direct[256] = {false};
        for (int c = '!'; c <= '~'; c++) direct[c] = true;
        for (int c = 0xA1; c <= 0xAC; c++) direct[c] = true;
        for (int c = 0xAE; c <= 0xFF; c++) direct[c] = true;
        int n = 0;
        for (int b = 0; b < 256; b++) {
            const uint32_t cp = direct[b] ? (uint32_t)b : (uint32_t)(256 + n++);
            b2u[b] = cpt_to_utf8(cp);
            u2b[b2u[b]] = (uint8_t)b;
        }
    }
};
static const ByteMap &byte_map() {
    static const ByteMap m;
    return m;
}

// ------------------------------------------------------------------------------------------ llama-3 pre-tokeniser
// "(?i:'s|'t|'re|'ve|'m|'ll|'d)|[^\r\n\p{L}\p{N}]?\p{L}+|\p{N}{1,3}| ?[^\s\p{L}\p{N}]+[\r\n]*|\s*[\r\n]+|\s+(?!\S)|\s+"
// The qwen2 pre-tokeniser is the same expression with "\p{N}" (one digit per word) in place of "\p{N}{1,3}": `max_digits` 1.
std::vector<std::string> llama3_pretokenize(const std::string &text, int max_digits) {
    std::vector<uint32_t> off;
    const std::vector<uint32_t> cpts = utf8_to_cpts(text, &off);
    const size_t n = cpts.size();
    static const uint32_t OOR = 0xFFFFFFFFu;
    auto cpt = [&](size_t p) -> uint32_t { return p < n ? cpts[p] : OOR; };
    auto is_l = [&](size_t p) { return p < n && cp_is_letter(cpts[p]); };
    auto is_n = [&](size_t p) { return p < n && cp_is_number(cpts[p]); };
    auto is_s = [&](size_t p) { return p < n && cp_is_space(cpts[p]); };
    auto lower = [](uint32_t c) { return (c >= 'A' && c <= 'Z') ? c + 32 : c; };
    std::vector<std::string> words;
    size_t prev_end = 0;
    auto add = [&](size_t end) -> size_t {
        const size_t len = end - prev_end;
        if (len > 0) words.push_back(text.substr(off[prev_end], off[end] - off[prev_end]));
        prev_end = end;
        return len;
    };
    for (size_t pos = 0; pos < n;) {
        const uint32_t c = cpt(pos);
        if (c == '\'' && pos + 1 < n) { // contractions, case-insensitive
            const uint32_t c1 = lower(cpt(pos + 1));
            if (c1 == 's' || c1 == 't' || c1 == 'm' || c1 == 'd') { pos += add(pos + 2); continue; }
            if (pos + 2 < n) {
                const uint32_t c2 = lower(cpt(pos + 2));
                if ((c1 == 'r' && c2 == 'e') || (c1 == 'v' && c2 == 'e') || (c1 == 'l' && c2 == 'l')) { pos += add(pos + 3); continue; }
            }
        }
        if (!(c == '\r' || c == '\n' || is_n(pos))) { // [^\r\n\p{L}\p{N}]?\p{L}+
            if (is_l(pos) || is_l(pos + 1)) {
                pos++;
                while (is_l(pos)) pos++;
                add(pos);
                continue;
            }
        }
        if (is_n(pos)) { // \p{N}{1,max_digits}
            size_t ini = pos;
            while (is_n(pos)) {
                if (++pos - ini >= (size_t)max_digits) { add(pos); ini = pos; }
            }
            add(pos);
            continue;
        }
        { // " ?[^\s\p{L}\p{N}]+[\r\n]*"
            size_t p2 = (c == ' ') ? pos + 1 : pos;
            auto other = [&](size_t p) { return p < n && !is_s(p) && !is_l(p) && !is_n(p); };
            if (other(p2)) {
                pos = p2;
                while (other(pos)) pos++;
                while (cpt(pos) == '\r' || cpt(pos) == '\n') pos++;
                add(pos);
                continue;
            }
        }
        size_t nws = 0, last_rn = 0;
        while (is_s(pos + nws)) {
            const uint32_t c2 = cpt(pos + nws);
            if (c2 == '\r' || c2 == '\n') last_rn = pos + nws + 1;
            nws++;
        }
        if (last_rn > 0) { pos = last_rn; add(pos); continue; }                         // \s*[\r\n]+
        if (nws > 1 && cpt(pos + nws) != OOR) { pos += nws - 1; add(pos); continue; }     // \s+(?!\S)
        if (nws > 0) { pos += nws; add(pos); continue; }                                 // \s+
        add(++pos);
    }
    return words;
}

// ------------------------------------------------------------------------------------------ construction
Tokenizer::Tokenizer(const GGUFFile &f) {
    model_ = f.get_s("tokenizer.ggml.model", "");
    pre_ = f.get_s("tokenizer.ggml.pre", "default");
    const GGUFValue *toks = f.get("tokenizer.ggml.tokens");
    if (!toks || toks->arr_s.empty()) { // a model without a vocabulary can still be evaluated on token ids
        kind_ = NONE;
        return;
    }
    tokens_ = toks->arr_s;
    const size_t n = tokens_.size();
    scores_.assign(n, 0.f);
    types_.assign(n, TOK_NORMAL);
    if (const GGUFValue *s = f.get("tokenizer.ggml.scores"))
        for (size_t i = 0; i < n && i < s->arr_f.size(); i++) scores_[i] = (float)s->arr_f[i];
    if (const GGUFValue *t = f.get("tokenizer.ggml.token_type"))
        for (size_t i = 0; i < n && i < t->arr_i.size(); i++) types_[i] = (int)t->arr_i[i];
    token_to_id_.reserve(n * 2);
    for (size_t i = 0; i < n; i++) token_to_id_.emplace(tokens_[i], (int32_t)i); // first id wins for duplicates
    if (model_ == "llama") {
        kind_ = SPM;
        bos_ = 1; eos_ = 2; unk_ = 0;
        add_bos_ = true;
    } else if (model_ == "gpt2") {
        kind_ = BPE;
        bos_ = 11; eos_ = 11;
        add_bos_ = false;
        if (pre_ == "smaug-bpe") pre_ = "llama-bpe"; // llama.cpp maps it onto the llama-3 pre-tokeniser as well
        B2_CHECK(pre_ == "llama-bpe" || pre_ == "llama3" || pre_ == "llama-v3" || pre_ == "default" || pre_ == "qwen2", B2_EUNSUPPORTED,
                 "unsupported BPE pre-tokeniser '" + pre_ + "' (llama-bpe, qwen2)");
        // llama-3 emits a word that is itself a token without merging and adds a BOS; qwen2 does neither (llama.cpp's
        // per-pre-tokeniser defaults; the add_bos_token key below still overrides)
        ignore_merges_ = pre_ != "default" && pre_ != "qwen2";
        if (ignore_merges_) add_bos_ = true;
        if (const GGUFValue *m = f.get("tokenizer.ggml.merges")) {
            bpe_ranks_.reserve(m->arr_s.size() * 2);
            for (size_t i = 0; i < m->arr_s.size(); i++) bpe_ranks_.emplace(m->arr_s[i], (int)i);
        }
    } else {
        throw Error(B2_EUNSUPPORTED, "unsupported tokenizer model '" + model_ + "'");
    }
    bos_ = (int32_t)f.get_u("tokenizer.ggml.bos_token_id", (uint64_t)(int64_t)bos_);
    eos_ = (int32_t)f.get_u("tokenizer.ggml.eos_token_id", (uint64_t)(int64_t)eos_);
    unk_ = (int32_t)f.get_u("tokenizer.ggml.unknown_token_id", (uint64_t)(int64_t)unk_);
    eot_ = f.has("tokenizer.ggml.eot_token_id") ? (int32_t)f.get_u("tokenizer.ggml.eot_token_id", 0) : -1;
    if (eot_ < 0) {
        auto it = token_to_id_.find("<|eot_id|>");
        if (it != token_to_id_.end()) eot_ = it->second;
    }
    nl_ = byte_to_token((uint8_t)'\n');
    if (nl_ == unk_) nl_ = -1;
    add_bos_ = f.get_b("tokenizer.ggml.add_bos_token", add_bos_);
    add_eos_ = f.get_b("tokenizer.ggml.add_eos_token", false);
    add_space_prefix_ = f.get_b("tokenizer.ggml.add_space_prefix", true);
    for (size_t i = 0; i < n; i++)
        if (types_[i] == TOK_CONTROL || types_[i] == TOK_USER_DEFINED || types_[i] == TOK_UNKNOWN) special_.push_back((int32_t)i);
    std::stable_sort(special_.begin(), special_.end(),
                     [&](int32_t a, int32_t b) { return tokens_[a].size() > tokens_[b].size(); });
}

int32_t Tokenizer::byte_to_token(uint8_t b) const {
    if (kind_ == SPM) {
        static const char *hex = "0123456789ABCDEF";
        const char buf[7] = {'<', '0', 'x', hex[b >> 4], hex[b & 15], '>', 0};
        auto it = token_to_id_.find(buf);
        if (it != token_to_id_.end()) return it->second;
        auto it2 = token_to_id_.find(std::string(1, (char)b));
        return it2 != token_to_id_.end() ? it2->second : unk_;
    }
    auto it = token_to_id_.find(byte_map().b2u[b]);
    return it != token_to_id_.end() ? it->second : unk_;
}

// ------------------------------------------------------------------------------------------ SPM
namespace {
struct Sym {
    int prev, next;
    const char *text;
    size_t n;
};
struct SpmBigram {
    int left, right;
    float score;
    size_t size;
    bool operator<(const SpmBigram &o) const { return score < o.score || (score == o.score && left > o.left); }
};
} // namespace

void Tokenizer::spm_tokenize(const std::string &text, std::vector<int32_t> &out) const {
    std::vector<Sym> syms;
    std::priority_queue<SpmBigram> work;
    std::map<std::string, std::pair<int, int>> rev_merge;
    size_t offs = 0;
    int index = 0;
    while (offs < text.size()) {
        Sym s;
        const size_t len = std::min(text.size() - offs, utf8_len((uint8_t)text[offs]));
        s.text = text.c_str() + offs;
        s.n = len;
        offs += len;
        s.prev = index - 1;
        s.next = offs == text.size() ? -1 : index + 1;
        index++;
        syms.push_back(s);
    }
    auto try_add = [&](int left, int right) {
        if (left == -1 || right == -1) return;
        const std::string t(syms[left].text, syms[left].n + syms[right].n);
        auto it = token_to_id_.find(t);
        if (it == token_to_id_.end()) return;
        SpmBigram b{left, right, scores_[it->second], t.size()};
        work.push(b);
        rev_merge[t] = std::make_pair(left, right);
    };
    for (int i = 1; i < (int)syms.size(); i++) try_add(i - 1, i);
    while (!work.empty()) {
        const SpmBigram b = work.top();
        work.pop();
        Sym &l = syms[b.left];
        Sym &r = syms[b.right];
        if (l.n == 0 || r.n == 0 || l.n + r.n != b.size) continue;
        l.n += r.n;
        r.n = 0;
        l.next = r.next;
        if (r.next >= 0) syms[r.next].prev = b.left;
        try_add(l.prev, b.left);
        try_add(b.left, l.next);
    }
    // resegment: symbols that are not tokens fall back to their merge parts, then to bytes
    std::vector<const Sym *> stack;
    for (int i = 0; i != -1 && !syms.empty(); i = syms[i].next) {
        stack.clear();
        stack.push_back(&syms[i]);
        // explicit recursion on (text, n) spans
        struct Span { const char *p; size_t n; };
        std::vector<Span> st{{syms[i].text, syms[i].n}};
        while (!st.empty()) {
            const Span sp = st.back();
            st.pop_back();
            const std::string t(sp.p, sp.n);
            auto it = token_to_id_.find(t);
            if (it != token_to_id_.end()) { out.push_back(it->second); continue; }
            auto pm = rev_merge.find(t);
            if (pm == rev_merge.end()) {
                for (size_t j = 0; j < sp.n; j++) out.push_back(byte_to_token((uint8_t)sp.p[j]));
                continue;
            }
            // the merge parts are the symbols' ORIGINAL spans: left starts at sp.p, right follows
            const Sym &ls = syms[pm->second.first];
            const size_t ln = (size_t)(syms[pm->second.second].text - ls.text);
            st.push_back({sp.p + ln, sp.n - ln}); // right, processed after left
            st.push_back({sp.p, ln});
        }
    }
}

// ------------------------------------------------------------------------------------------ BPE
int Tokenizer::bpe_rank(const std::string &l, const std::string &r) const {
    auto it = bpe_ranks_.find(l + " " + r);
    return it == bpe_ranks_.end() ? -1 : it->second;
}

namespace {
struct BpeBigram {
    int left, right;
    std::string text;
    int rank;
    size_t size;
    bool operator<(const BpeBigram &o) const { return rank > o.rank || (rank == o.rank && left > o.left); }
};
} // namespace

void Tokenizer::bpe_tokenize(const std::string &text, std::vector<int32_t> &out) const {
    const ByteMap &bm = byte_map();
    for (const std::string &raw : llama3_pretokenize(text, pre_ == "qwen2" ? 1 : 3)) {
        std::string word;
        for (unsigned char c : raw) word += bm.b2u[c];
        std::vector<Sym> syms;
        std::priority_queue<BpeBigram> work;
        size_t offset = 0;
        int index = 0;
        if (ignore_merges_ && token_to_id_.count(word)) {
            syms.push_back(Sym{-1, -1, word.c_str(), word.size()});
            offset = word.size();
        }
        while (offset < word.size()) {
            Sym s;
            const size_t len = std::min(word.size() - offset, utf8_len((uint8_t)word[offset]));
            s.text = word.c_str() + offset;
            s.n = len;
            offset += len;
            s.prev = index - 1;
            s.next = offset == word.size() ? -1 : index + 1;
            index++;
            syms.push_back(s);
        }
        auto add = [&](int left, int right) {
            if (left == -1 || right == -1) return;
            const std::string lt(syms[left].text, syms[left].n), rt(syms[right].text, syms[right].n);
            const int rank = bpe_rank(lt, rt);
            if (rank < 0) return;
            work.push(BpeBigram{left, right, lt + rt, rank, lt.size() + rt.size()});
        };
        for (int i = 1; i < (int)syms.size(); i++) add(i - 1, i);
        while (!work.empty()) {
            const BpeBigram b = work.top();
            work.pop();
            Sym &l = syms[b.left];
            Sym &r = syms[b.right];
            if (l.n == 0 || r.n == 0) continue;
            if (std::string(l.text, l.n) + std::string(r.text, r.n) != b.text) continue; // outdated
            l.n += r.n;
            r.n = 0;
            l.next = r.next;
            if (r.next >= 0) syms[r.next].prev = b.left;
            add(l.prev, b.left);
            add(b.left, l.next);
        }
        for (const Sym &s : syms) {
            if (s.n == 0) continue;
            const std::string t(s.text, s.n);
            auto it = token_to_id_.find(t);
            if (it != token_to_id_.end()) { out.push_back(it->second); continue; }
            for (size_t j = 0; j < t.size();) { // unknown piece: per-character fallback
                const size_t len = std::min(t.size() - j, utf8_len((uint8_t)t[j]));
                auto ib = token_to_id_.find(t.substr(j, len));
                if (ib != token_to_id_.end()) out.push_back(ib->second);
                j += len;
            }
        }
    }
}

// ------------------------------------------------------------------------------------------ front end
std::vector<int32_t> Tokenizer::tokenize(const std::string &text, bool add_bos, bool parse_special) const {
    B2_CHECK(kind_ != NONE, B2_EUNSUPPORTED, "this GGUF file carries no tokenizer vocabulary");
    // split the text at special tokens (tokenizer_st_partition): fragments are raw text or a token id
    struct Frag { bool is_tok; int32_t tok; size_t off, len; };
    std::vector<Frag> frags;
    if (!text.empty()) frags.push_back({false, 0, 0, text.size()}); // empty text -> no fragment -> just BOS
    for (int32_t sid : special_) {
        const int ty = types_[sid];
        if (!parse_special && (ty == TOK_CONTROL || ty == TOK_UNKNOWN)) continue;
        const std::string &st = tokens_[sid];
        if (st.empty()) continue;
        std::vector<Frag> next;
        for (const Frag &fr : frags) {
            if (fr.is_tok) { next.push_back(fr); continue; }
            size_t base = fr.off, end = fr.off + fr.len;
            while (base < end) {
                const size_t m = text.find(st, base);
                if (m == std::string::npos || m + st.size() > end) break;
                if (m > base) next.push_back({false, 0, base, m - base});
                next.push_back({true, sid, m, st.size()});
                base = m + st.size();
            }
            if (base < end) next.push_back({false, 0, base, end - base});
        }
        frags.swap(next);
    }
    std::vector<int32_t> out;
    if (kind_ == SPM) {
        bool prev_special = true;
        if (add_bos && add_bos_ && bos_ >= 0) out.push_back(bos_);
        for (const Frag &fr : frags) {
            if (fr.is_tok) { out.push_back(fr.tok); prev_special = true; continue; }
            std::string raw = text.substr(fr.off, fr.len);
            if (add_space_prefix_ && prev_special) raw = " " + raw;
            std::string esc;
            for (char c : raw) {
                if (c == ' ') esc += "\xe2\x96\x81";
                else esc += c;
            }
            spm_tokenize(esc, out);
            prev_special = false;
        }
        if (add_bos && add_eos_ && eos_ >= 0) out.push_back(eos_);
    } else {
        if (add_bos && add_bos_ && bos_ >= 0) out.push_back(bos_);
        for (const Frag &fr : frags) {
            if (fr.is_tok) out.push_back(fr.tok);
            else bpe_tokenize(text.substr(fr.off, fr.len), out);
        }
        if (add_bos && add_eos_ && eos_ >= 0) out.push_back(eos_);
    }
    return out;
}

std::string Tokenizer::token_to_piece(int32_t id, bool render_special) const {
    if (id < 0 || id >= (int32_t)tokens_.size()) return "";
    const int ty = types_[id];
    const std::string &t = tokens_[id];
    if (ty == TOK_USER_DEFINED) return t;
    if (ty == TOK_CONTROL || ty == TOK_UNKNOWN || ty == TOK_UNUSED) return render_special ? t : "";
    if (kind_ == SPM) {
        if (ty == TOK_BYTE) {
            if (t.size() == 6 && t[0] == '<' && t[1] == '0' && t[2] == 'x' && t[5] == '>')
                return std::string(1, (char)strtol(t.substr(3, 2).c_str(), nullptr, 16));
            return t;
        }
        std::string r;
        for (size_t i = 0; i < t.size();) {
            if (t.compare(i, 3, "\xe2\x96\x81") == 0) { r += ' '; i += 3; }
            else r += t[i++];
        }
        return r;
    }
    // BPE: unicode-mapped characters back to raw bytes
    const ByteMap &bm = byte_map();
    std::string r;
    for (size_t i = 0; i < t.size();) {
        const size_t len = std::min(t.size() - i, utf8_len((uint8_t)t[i]));
        auto it = bm.u2b.find(t.substr(i, len));
        if (it != bm.u2b.end()) r += (char)it->second;
        else r += t.substr(i, len);
        i += len;
    }
    return r;
}

} // namespace b2
