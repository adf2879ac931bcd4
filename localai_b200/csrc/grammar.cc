// grammar.cc -- see grammar.h.  Part 1: UTF-8 and the GBNF parser.
#include "grammar.h"

#include <algorithm>
#include <cstring>
#include <map>

#include "common.h"

namespace b2 {

// ---- UTF-8 -----------------------------------------------------------------------------------------------------------
static inline int utf8_len_from_lead(uint8_t b) {
    // by high nibble: 0xxx -> 1, 10xx -> continuation (0 = invalid as a first byte), 110x -> 2, 1110 -> 3, 1111 -> 4
    static const int8_t len[16] = {1, 1, 1, 1, 1, 1, 1, 1, 0, 0, 0, 0, 2, 2, 3, 4};
    return len[b >> 4];
}

void utf8_decode_piece(const std::string &bytes, Utf8Tail &tail, std::vector<uint32_t> &out) {
    size_t i = 0;
    uint32_t value = tail.value;
    int n_remain = tail.n_remain;
    // finish the sequence the previous piece left open
    while (i < bytes.size() && n_remain > 0) {
        const uint8_t b = (uint8_t)bytes[i];
        if ((b >> 6) != 2) { // not a continuation byte: the sequence is broken
            out.push_back(0);
            tail = {0, -1};
            return;
        }
        value = (value << 6) | (b & 0x3F);
        i++;
        if (--n_remain == 0) out.push_back(value);
    }
    while (i < bytes.size()) {
        const uint8_t b = (uint8_t)bytes[i];
        n_remain = utf8_len_from_lead(b) - 1;
        if (n_remain < 0) { // a continuation byte where a lead byte must be
            out.clear();
            out.push_back(0);
            tail = {0, -1};
            return;
        }
        static const uint8_t lead_mask[4] = {0x7F, 0x1F, 0x0F, 0x07};
        value = b & lead_mask[n_remain];
        i++;
        while (i < bytes.size() && n_remain > 0) {
            value = (value << 6) | ((uint8_t)bytes[i] & 0x3F);
            i++;
            n_remain--;
        }
        if (n_remain == 0) out.push_back(value);
    }
    tail = n_remain > 0 ? Utf8Tail{value, n_remain} : Utf8Tail{0, 0};
}

// ---- parser ------------------------------------------------------------------------------------------------------------
namespace {

struct Item { // an element while parsing: its ranges are still local
    GrammarElem::Kind kind = GrammarElem::CHARS;
    bool negated = false;
    uint32_t ref = 0;
    std::vector<std::pair<uint32_t, uint32_t>> ranges;
};
using Seq = std::vector<Item>;

struct Parser {
    const char *p, *end;
    std::map<std::string, uint32_t> ids;
    std::vector<std::string> names;
    std::vector<std::vector<Seq>> rules; // rule id -> alternatives
    std::vector<uint8_t> defined;

    [[noreturn]] void fail(const std::string &what) const {
        std::string at(p, std::min<size_t>(end - p, 24));
        throw Error(B2_EINVAL, "grammar: " + what + " at \"" + at + "\"");
    }
    bool eof() const { return p >= end; }
    char peek(size_t k = 0) const { return p + k < end ? p[k] : '\0'; }

    uint32_t symbol(const std::string &name) {
        auto it = ids.find(name);
        if (it != ids.end()) return it->second;
        const uint32_t id = (uint32_t)names.size();
        ids.emplace(name, id);
        names.push_back(name);
        rules.emplace_back();
        defined.push_back(0);
        return id;
    }
    uint32_t fresh_symbol(const std::string &base) { // helper rules of groups and repetitions
        const uint32_t id = (uint32_t)names.size();
        return symbol(base + "_" + std::to_string(id));
    }
    void define(uint32_t id, std::vector<Seq> &&alts) {
        rules[id] = std::move(alts);
        defined[id] = 1;
    }

    void skip_space(bool newline_ok) {
        while (!eof()) {
            const char c = *p;
            if (c == '#') {
                while (!eof() && *p != '\r' && *p != '\n') p++;
            } else if (c == ' ' || c == '\t' || (newline_ok && (c == '\r' || c == '\n'))) {
                p++;
            } else {
                break;
            }
        }
    }
    static bool word_char(char c) { return (c >= 'a' && c <= 'z') || (c >= 'A' && c <= 'Z') || c == '-' || (c >= '0' && c <= '9'); }
    std::string name() {
        const char *b = p;
        while (!eof() && word_char(*p)) p++;
        if (p == b) fail("expecting a name");
        return std::string(b, p);
    }
    uint32_t hex(int n) {
        uint32_t v = 0;
        for (int i = 0; i < n; i++, p++) {
            const char c = peek();
            int d;
            if (c >= '0' && c <= '9') d = c - '0';
            else if (c >= 'a' && c <= 'f') d = c - 'a' + 10;
            else if (c >= 'A' && c <= 'F') d = c - 'A' + 10;
            else fail("expecting " + std::to_string(n) + " hex digits");
            v = (v << 4) | (uint32_t)d;
        }
        return v;
    }
    uint32_t one_char() { // a literal code point or an escape
        if (eof()) fail("unexpected end of input");
        if (*p == '\\') {
            const char c = peek(1);
            p += 2;
            switch (c) {
            case 'x': return hex(2);
            case 'u': return hex(4);
            case 'U': return hex(8);
            case 't': return '\t';
            case 'r': return '\r';
            case 'n': return '\n';
            case '\\': case '"': case '[': case ']': return (uint8_t)c;
            default: p -= 2; fail("unknown escape");
            }
        }
        const uint8_t b = (uint8_t)*p;
        int n = utf8_len_from_lead(b);
        if (n == 0) n = 1; // a stray continuation byte stands for itself
        static const uint8_t lead_mask[5] = {0, 0x7F, 0x1F, 0x0F, 0x07};
        uint32_t v = b & lead_mask[n];
        p++;
        for (int i = 1; i < n && !eof(); i++, p++) v = (v << 6) | ((uint8_t)*p & 0x3F);
        return v;
    }
    int integer() {
        const char *b = p;
        long v = 0;
        while (!eof() && *p >= '0' && *p <= '9') {
            v = v * 10 + (*p - '0');
            if (v > 100000) fail("repetition count too large");
            p++;
        }
        if (p == b) fail("expecting an integer");
        return (int)v;
    }

    static Item ref_item(uint32_t id) {
        Item it;
        it.kind = GrammarElem::REF;
        it.ref = id;
        return it;
    }

    // X{min,max} on the last symbol of `seq` (which starts at `from`); max < 0 = unbounded.
    //   X{2,4} -> X X (X (X)?)?      X* -> R  with  R ::= X R |      X+ -> X R
    void repeat(Seq &seq, size_t from, const std::string &rule_name, int min_times, int max_times) {
        if (from >= seq.size()) fail("expecting a preceding item to repeat");
        const Seq sym(seq.begin() + from, seq.end());
        if (min_times == 0) seq.resize(from);
        for (int i = 1; i < min_times; i++) seq.insert(seq.end(), sym.begin(), sym.end());
        const int n_opt = max_times < 0 ? 1 : max_times - min_times;
        uint32_t last = 0;
        for (int i = 0; i < n_opt; i++) {
            const uint32_t id = fresh_symbol(rule_name);
            Seq body = sym;
            if (max_times < 0) body.push_back(ref_item(id));
            else if (i > 0) body.push_back(ref_item(last));
            std::vector<Seq> alts;
            alts.push_back(std::move(body));
            alts.emplace_back(); // | (nothing)
            define(id, std::move(alts));
            last = id;
        }
        if (n_opt > 0) seq.push_back(ref_item(last));
    }

    void sequence(const std::string &rule_name, Seq &seq, bool nested) {
        size_t last_start = seq.size();
        while (!eof()) {
            const char c = *p;
            if (c == '"') {
                p++;
                last_start = seq.size();
                while (peek() != '"') {
                    if (eof()) fail("unexpected end of input in a literal");
                    Item it;
                    const uint32_t cp = one_char();
                    it.ranges.emplace_back(cp, cp);
                    seq.push_back(std::move(it));
                }
                p++;
                skip_space(nested);
            } else if (c == '[') {
                p++;
                Item it;
                if (peek() == '^') {
                    p++;
                    it.negated = true;
                }
                last_start = seq.size();
                while (peek() != ']') {
                    if (eof()) fail("unexpected end of input in a character class");
                    const uint32_t lo = one_char();
                    uint32_t hi = lo;
                    if (peek() == '-' && peek(1) != ']') {
                        p++;
                        if (eof()) fail("unexpected end of input in a character class");
                        hi = one_char();
                    }
                    it.ranges.emplace_back(lo, hi);
                }
                p++;
                seq.push_back(std::move(it));
                skip_space(nested);
            } else if (word_char(c)) {
                const uint32_t id = symbol(name());
                skip_space(nested);
                last_start = seq.size();
                seq.push_back(ref_item(id));
            } else if (c == '(') {
                p++;
                skip_space(true);
                const uint32_t id = fresh_symbol(rule_name);
                alternates(rule_name, id, true);
                last_start = seq.size();
                seq.push_back(ref_item(id));
                if (peek() != ')') fail("expecting ')'");
                p++;
                skip_space(nested);
            } else if (c == '.') {
                last_start = seq.size();
                Item it;
                it.kind = GrammarElem::ANY;
                seq.push_back(std::move(it));
                p++;
                skip_space(nested);
            } else if (c == '*' || c == '+' || c == '?') {
                p++;
                skip_space(nested);
                repeat(seq, last_start, rule_name, c == '+' ? 1 : 0, c == '?' ? 1 : -1);
                last_start = seq.size(); // a second operator needs a new item in front of it
            } else if (c == '{') {
                p++;
                skip_space(nested);
                const int lo = integer();
                int hi = lo;
                skip_space(nested);
                if (peek() == ',') {
                    p++;
                    skip_space(nested);
                    if (peek() >= '0' && peek() <= '9') {
                        hi = integer();
                        skip_space(nested);
                    } else {
                        hi = -1;
                    }
                }
                if (peek() != '}') fail("expecting '}'");
                p++;
                if (hi >= 0 && hi < lo) fail("repetition maximum below its minimum");
                skip_space(nested);
                repeat(seq, last_start, rule_name, lo, hi);
                last_start = seq.size();
            } else {
                break;
            }
        }
    }

    void alternates(const std::string &rule_name, uint32_t id, bool nested) {
        std::vector<Seq> alts(1);
        sequence(rule_name, alts.back(), nested);
        while (peek() == '|') {
            p++;
            skip_space(true);
            alts.emplace_back();
            sequence(rule_name, alts.back(), nested);
        }
        define(id, std::move(alts));
    }

    void rule() {
        const std::string nm = name();
        skip_space(false);
        const uint32_t id = symbol(nm);
        if (!(peek() == ':' && peek(1) == ':' && peek(2) == '=')) fail("expecting ::=");
        p += 3;
        skip_space(true);
        alternates(nm, id, false);
        if (peek() == '\r') p += peek(1) == '\n' ? 2 : 1;
        else if (peek() == '\n') p++;
        else if (!eof()) fail("expecting a newline or the end of the grammar");
        skip_space(true);
    }
};

} // namespace

// ---- flattening + checks ----------------------------------------------------------------------------------------------
std::shared_ptr<const GrammarRules> grammar_parse(const std::string &text) {
    Parser ps;
    ps.p = text.data();
    ps.end = text.data() + text.size();
    ps.skip_space(true);
    while (!ps.eof()) ps.rule();

    auto out = std::make_shared<GrammarRules>();
    const size_t n = ps.names.size();
    B2_CHECK(ps.ids.count("root"), B2_EINVAL, "grammar: there is no root rule");
    out->root = ps.ids["root"];
    for (size_t r = 0; r < n; r++)
        B2_CHECK(ps.defined[r], B2_EINVAL, "grammar: rule '" + ps.names[r] + "' is used but not defined");
    out->names = ps.names;
    out->alts.resize(n);
    for (size_t r = 0; r < n; r++) {
        B2_CHECK(!ps.rules[r].empty(), B2_EINVAL, "grammar: rule '" + ps.names[r] + "' is empty");
        for (const Seq &alt : ps.rules[r]) {
            out->alts[r].push_back((uint32_t)out->elems.size());
            for (const Item &it : alt) {
                GrammarElem e;
                e.kind = it.kind;
                e.negated = it.negated;
                e.ref = it.ref;
                e.r0 = (uint32_t)out->ranges.size();
                out->ranges.insert(out->ranges.end(), it.ranges.begin(), it.ranges.end());
                e.r1 = (uint32_t)out->ranges.size();
                out->elems.push_back(e);
            }
            out->elems.push_back(GrammarElem{}); // END
        }
    }

    // Left recursion would make the stack expansion endless.  nullable[r]: r can produce the empty string;
    // r "starts with" q if q is reachable at the front of an alternative through nullable prefixes.
    std::vector<uint8_t> nullable(n, 0);
    for (bool changed = true; changed;) {
        changed = false;
        for (size_t r = 0; r < n; r++) {
            if (nullable[r]) continue;
            for (uint32_t a : out->alts[r]) {
                bool all = true;
                for (uint32_t i = a; out->elems[i].kind != GrammarElem::END; i++) {
                    const GrammarElem &e = out->elems[i];
                    if (e.kind != GrammarElem::REF || !nullable[e.ref]) {
                        all = false;
                        break;
                    }
                }
                if (all) {
                    nullable[r] = 1;
                    changed = true;
                    break;
                }
            }
        }
    }
    // front[r]: the rules that can stand leftmost in r (behind nullable rules only); a cycle in that graph is a
    // left recursion
    std::vector<std::vector<uint32_t>> front(n);
    for (size_t r = 0; r < n; r++)
        for (uint32_t a : out->alts[r])
            for (uint32_t i = a; out->elems[i].kind == GrammarElem::REF; i++) {
                front[r].push_back(out->elems[i].ref);
                if (!nullable[out->elems[i].ref]) break;
            }
    std::vector<uint8_t> state(n, 0); // 0 = unseen, 1 = on the path, 2 = done
    std::vector<std::pair<uint32_t, size_t>> path; // (rule, next edge)
    for (size_t r0 = 0; r0 < n; r0++) {
        if (state[r0]) continue;
        state[r0] = 1;
        path.emplace_back((uint32_t)r0, 0);
        while (!path.empty()) {
            const uint32_t r = path.back().first;
            if (path.back().second == front[r].size()) {
                state[r] = 2;
                path.pop_back();
                continue;
            }
            const uint32_t q = front[r][path.back().second++];
            B2_CHECK(state[q] != 1, B2_EINVAL, "grammar: left recursion through rule '" + out->names[q] + "'");
            if (state[q] == 0) {
                state[q] = 1;
                path.emplace_back(q, 0);
            }
        }
    }
    return out;
}

// ---- matcher -----------------------------------------------------------------------------------------------------------
Grammar::Grammar(std::shared_ptr<const GrammarRules> rules) : rules_(std::move(rules)) {
    for (uint32_t a : rules_->alts[rules_->root]) {
        Stack st;
        if (rules_->elems[a].kind != GrammarElem::END) st.push_back(a);
        advance(st, stacks_);
    }
}

// Expands rule references on top of `st` until every resulting stack is empty or has a character element on top.
void Grammar::advance(Stack &st, std::vector<Stack> &out) const {
    if (st.empty()) {
        if (std::find(out.begin(), out.end(), st) == out.end()) out.push_back(st);
        return;
    }
    const uint32_t pos = st.back();
    const GrammarElem &e = rules_->elems[pos];
    if (e.kind != GrammarElem::REF) {
        if (std::find(out.begin(), out.end(), st) == out.end()) out.push_back(st);
        return;
    }
    for (uint32_t a : rules_->alts[e.ref]) {
        Stack nx(st.begin(), st.end() - 1);
        if (rules_->elems[pos + 1].kind != GrammarElem::END) nx.push_back(pos + 1); // what follows the reference
        if (rules_->elems[a].kind != GrammarElem::END) nx.push_back(a);             // the alternative itself
        advance(nx, out);
    }
}

bool Grammar::elem_matches(const GrammarElem &e, uint32_t cp) const {
    if (e.kind == GrammarElem::ANY) return true;
    for (uint32_t i = e.r0; i < e.r1; i++)
        if (rules_->ranges[i].first <= cp && cp <= rules_->ranges[i].second) return !e.negated;
    return e.negated;
}

// Could the unfinished sequence `t` still become a code point this element takes?  It stands for the whole interval
// of code points sharing its leading bits.
bool Grammar::elem_matches_partial(const GrammarElem &e, const Utf8Tail &t) const {
    if (t.n_remain < 0 || (t.n_remain == 1 && t.value < 2)) return false; // invalid, or an overlong 7-bit form
    uint32_t lo = t.value << (t.n_remain * 6);
    const uint32_t hi = lo | ((1u << (t.n_remain * 6)) - 1);
    if (lo == 0) { // shortest-form rule: 3- and 4-byte sequences start above the shorter forms
        if (t.n_remain == 2) lo = 1u << 11;
        else if (t.n_remain == 3) lo = 1u << 16;
    }
    if (e.kind == GrammarElem::ANY) return true;
    for (uint32_t i = e.r0; i < e.r1; i++)
        if (rules_->ranges[i].first <= hi && lo <= rules_->ranges[i].second) return !e.negated;
    return e.negated;
}

void Grammar::step(const std::vector<Stack> &from, uint32_t cp, std::vector<Stack> &to) const {
    to.clear();
    for (const Stack &st : from) {
        if (st.empty()) continue;
        const uint32_t pos = st.back();
        if (!elem_matches(rules_->elems[pos], cp)) continue;
        Stack nx(st.begin(), st.end() - 1);
        if (rules_->elems[pos + 1].kind != GrammarElem::END) nx.push_back(pos + 1);
        advance(nx, to);
    }
}

int Grammar::intern(std::vector<Stack> &&stacks) const {
    auto ins = state_ids_.try_emplace(std::move(stacks), (int)states_.size());
    if (ins.second) states_.push_back(&ins.first->first);
    return ins.first->second;
}

int Grammar::next_state(int id, uint32_t cp) const {
    const uint64_t key = ((uint64_t)(uint32_t)id << 32) | cp;
    auto it = trans_.find(key);
    if (it != trans_.end()) return it->second;
    std::vector<Stack> nxt;
    step(*states_[id], cp, nxt);
    const int to = nxt.empty() ? -1 : intern(std::move(nxt));
    if (trans_.size() > (1u << 20)) trans_.clear(); // bounded: a grammar over a huge alphabet cannot grow it forever
    trans_.emplace(key, to);
    return to;
}

bool Grammar::fits(const std::vector<uint32_t> &cps, const Utf8Tail &tail) const {
    if (tail.n_remain < 0) return false;
    if (cps.empty() && tail.n_remain == 0) return false;
    if (cur_id_ < 0) cur_id_ = intern(std::vector<Stack>(stacks_));
    int id = cur_id_;
    for (uint32_t cp : cps) {
        id = next_state(id, cp);
        if (id < 0) return false;
    }
    if (tail.n_remain == 0) return true;
    for (const Stack &st : *states_[id])
        if (!st.empty() && elem_matches_partial(rules_->elems[st.back()], tail)) return true;
    return false;
}

bool Grammar::can_stop() const {
    for (const Stack &st : stacks_)
        if (st.empty()) return true;
    return false;
}

bool Grammar::token_fits(const GrammarVocab &v, int32_t tok) const {
    if (tok < 0 || (size_t)tok >= v.cps.size()) return false;
    if (v.is_eog[tok]) return can_stop();
    if (v.is_empty[tok]) return false;
    if (tail_.n_remain == 0) return fits(v.cps[tok], v.tails[tok]);
    std::vector<uint32_t> cps; // the previous token ended inside a sequence: decode this piece as its continuation
    Utf8Tail t = tail_;
    utf8_decode_piece(v.pieces[tok], t, cps);
    return fits(cps, t);
}

void Grammar::token_mask(const GrammarVocab &v, uint32_t *bits) const {
    const size_t n = v.cps.size();
    std::memset(bits, 0, ((n + 31) / 32) * sizeof(uint32_t));
    const bool stop_ok = can_stop();
    for (size_t t = 0; t < n; t++) {
        bool ok;
        if (v.is_eog[t]) {
            ok = stop_ok;
        } else if (v.is_empty[t]) {
            ok = false;
        } else if (tail_.n_remain == 0) {
            ok = fits(v.cps[t], v.tails[t]);
        } else {
            ok = token_fits(v, (int32_t)t);
        }
        if (ok) bits[t >> 5] |= 1u << (t & 31);
    }
}

void Grammar::accept_token(const GrammarVocab &v, int32_t tok) {
    B2_CHECK(tok >= 0 && (size_t)tok < v.cps.size(), B2_EINVAL, "grammar: token id out of range");
    if (v.is_eog[tok]) {
        B2_CHECK(can_stop(), B2_ESTATE, "grammar: end of generation before the grammar is complete");
        return;
    }
    std::vector<uint32_t> cps;
    Utf8Tail t = tail_;
    utf8_decode_piece(v.pieces[tok], t, cps);
    std::vector<Stack> cur = stacks_, nxt;
    for (uint32_t cp : cps) {
        step(cur, cp, nxt);
        cur.swap(nxt);
    }
    B2_CHECK(!cur.empty() && t.n_remain >= 0, B2_ESTATE, "grammar: accepted token does not fit the grammar");
    stacks_.swap(cur);
    tail_ = t;
    cur_id_ = -1;
}

std::string Grammar::state_key() const {
    std::string k;
    for (const Stack &st : stacks_) {
        k.append((const char *)st.data(), st.size() * sizeof(uint32_t));
        const uint32_t sep = 0xFFFFFFFFu;
        k.append((const char *)&sep, sizeof(sep));
    }
    k.append((const char *)&tail_.value, sizeof(tail_.value));
    k.append((const char *)&tail_.n_remain, sizeof(tail_.n_remain));
    return k;
}

bool Grammar::accept_bytes(const std::string &s) {
    std::vector<uint32_t> cps;
    Utf8Tail t = tail_;
    utf8_decode_piece(s, t, cps);
    if (t.n_remain < 0) return false;
    std::vector<Stack> cur = stacks_, nxt;
    for (uint32_t cp : cps) {
        step(cur, cp, nxt);
        if (nxt.empty()) return false;
        cur.swap(nxt);
    }
    stacks_.swap(cur);
    tail_ = t;
    cur_id_ = -1;
    return true;
}

} // namespace b2
