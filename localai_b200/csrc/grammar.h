// grammar.h -- GBNF grammar-constrained sampling: parser + pushdown matcher, host side.
// Replaces the grammar leg of gpt_sampler_{init,sample,accept} as the reference drives it
// (backend/cpp/llama/grpc-server.cpp:655 `sparams.grammar`, :2114 `data["grammar"]`, :890 init, :1946-1948
// sample/accept); the grammar text is what pkg/functions/grammars/ produces for tools / function calls
// (core/http/endpoints/openai/chat.go:216-269).  The matcher itself lives upstream in llama.cpp's
// llama-grammar.cpp (not vendored); this is a restatement of its published behaviour with its own data layout:
// every alternative is a run of elements in one flat array, a stack is a vector of indices into that array.
//
// The device never sees the grammar: the sampled id is checked here, and only when it does not fit is a
// vocabulary bit mask built and the row re-sampled under it (the order gpt_sampler_sample uses).
#pragma once
#include <cstdint>
#include <map>
#include <memory>
#include <string>
#include <unordered_map>
#include <vector>

namespace b2 {

struct Utf8Tail { // an unfinished UTF-8 sequence at the end of a piece
    uint32_t value = 0;
    int n_remain = 0; // bytes still missing; -1 = invalid sequence
};

// decodes `bytes` (continuing `tail`) into code points; the unfinished rest goes back into `tail`
void utf8_decode_piece(const std::string &bytes, Utf8Tail &tail, std::vector<uint32_t> &out);

struct GrammarElem {
    enum Kind : uint8_t { END, CHARS, ANY, REF };
    Kind kind = END;
    bool negated = false;    // CHARS: [^...]
    uint32_t ref = 0;        // REF: rule id
    uint32_t r0 = 0, r1 = 0; // CHARS: [r0, r1) into GrammarRules::ranges
};

struct GrammarRules {
    std::vector<GrammarElem> elems;                 // all alternatives back to back, each closed by an END element
    std::vector<std::pair<uint32_t, uint32_t>> ranges; // inclusive code point ranges of the CHARS elements
    std::vector<std::vector<uint32_t>> alts;        // rule id -> start index of each alternative in `elems`
    std::vector<std::string> names;                 // rule id -> name
    uint32_t root = 0;
};

// Throws b2::Error(B2_EINVAL, ...) on a syntax error, an undefined rule, a missing root or left recursion.
std::shared_ptr<const GrammarRules> grammar_parse(const std::string &text);

// The vocabulary as the matcher sees it, decoded once per model.
struct GrammarVocab {
    std::vector<std::vector<uint32_t>> cps; // token id -> code points of its piece
    std::vector<Utf8Tail> tails;            // token id -> unfinished trailing sequence of the piece (from a clean start)
    std::vector<uint8_t> is_eog;            // end-of-generation tokens: allowed only when the grammar may stop
    std::vector<uint8_t> is_empty;          // control / empty pieces: never allowed
    std::vector<std::string> pieces;
};

class Grammar {
  public:
    using Stack = std::vector<uint32_t>; // indices into rules->elems; back() is the element to match next

    explicit Grammar(std::shared_ptr<const GrammarRules> rules);
    Grammar(const Grammar &) = delete; // the state table below points into its own map
    Grammar &operator=(const Grammar &) = delete;

    // does the current state accept this token?  (llama_grammar_apply on a single candidate)
    bool token_fits(const GrammarVocab &v, int32_t tok) const;
    // bit i set = token i fits; `bits` holds (n_vocab + 31) / 32 words
    void token_mask(const GrammarVocab &v, uint32_t *bits) const;
    // advance over an accepted token (llama_grammar_accept); throws if it does not fit
    void accept_token(const GrammarVocab &v, int32_t tok);

    // byte-string interface (tests, tools)
    bool accept_bytes(const std::string &s); // false (state unchanged) if some code point has no continuation
    bool can_stop() const;                   // some stack is empty: the text so far is a complete sentence
    bool dead() const { return stacks_.empty(); }
    size_t n_stacks() const { return stacks_.size(); }
    // identifies the matcher state (stacks + open UTF-8 sequence): equal keys have equal token masks
    std::string state_key() const;

  private:
    void advance(Stack &st, std::vector<Stack> &out) const;
    void step(const std::vector<Stack> &from, uint32_t cp, std::vector<Stack> &to) const;
    bool elem_matches(const GrammarElem &e, uint32_t cp) const;
    bool elem_matches_partial(const GrammarElem &e, const Utf8Tail &t) const;
    bool fits(const std::vector<uint32_t> &cps, const Utf8Tail &tail) const;
    // The automaton is determinised lazily: every set of stacks met so far has a number, and (state, code point) ->
    // state is remembered, so a pass over the vocabulary costs one hash lookup per code point once the few states
    // around the current position are known (inside a JSON string almost every character leads back to the same one).
    int intern(std::vector<Stack> &&stacks) const;
    int next_state(int id, uint32_t cp) const; // -1 = no stack survives

    std::shared_ptr<const GrammarRules> rules_;
    std::vector<Stack> stacks_;
    Utf8Tail tail_;
    mutable std::map<std::vector<Stack>, int> state_ids_;
    mutable std::vector<const std::vector<Stack> *> states_; // id -> key of state_ids_ (map nodes do not move)
    mutable std::unordered_map<uint64_t, int> trans_;
    mutable int cur_id_ = -1; // number of stacks_, -1 = not looked up since the last accept
};

} // namespace b2
