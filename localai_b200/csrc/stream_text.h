// stream_text.h -- what a client may see of a generation, and how a prompt is made to fit a slot.
//
// Host-side text bookkeeping of the request path.  Its observable behaviour is that of the reference's slot engine
// (backend/cpp/llama/grpc-server.cpp:1010-1123 for the text, :1694-1749 for the prompt), which SURVEY.md 8(a9) asks to
// match byte for byte; the organisation is this engine's own: a small value type per concern instead of fields of the slot.
#pragma once
#include <algorithm>
#include <cstdint>
#include <string>
#include <vector>

namespace b2 {

// Bytes generated so far, and how many of them have been handed to the client.  A byte is handed over as soon as it can no
// longer turn out to be half a UTF-8 character or the beginning of a stop word.
class StreamText {
  public:
    struct Step {
        std::string release; // bytes to hand to the client now (streaming requests)
        bool stop_word = false;
    };

    void reset(const std::vector<std::string> *stop_words) {
        stops_ = stop_words;
        kept_.clear();
        released_ = 0;
    }
    const std::string &text() const { return kept_; }

    // `piece`: the bytes of the token just sampled
    Step append(const std::string &piece) {
        Step out;
        kept_ += piece;
        if (ends_inside_utf8(kept_)) return out; // wait for the rest of the character
        const size_t from = std::min(released_, kept_.size());
        const std::string pending = kept_.substr(from);
        // a stop word can only have been completed by this piece: look no further back than its length plus the piece's
        size_t cut = std::string::npos;
        for (const std::string &w : words()) {
            const size_t reach = w.size() + piece.size();
            const size_t at = pending.find(w, pending.size() > reach ? pending.size() - reach : 0);
            if (at != std::string::npos) cut = std::min(cut, at);
        }
        if (cut != std::string::npos) { // the stop word and everything after it never reach the client
            kept_.erase(from + cut);
            out.stop_word = true;
            return out;
        }
        // otherwise the tail may be the beginning of a stop word: hold it back from its earliest possible start
        for (const std::string &w : words()) cut = std::min(cut, prefix_overlap(pending, w));
        if (cut == std::string::npos) {
            out.release = pending;
            released_ += pending.size();
        }
        return out;
    }

    // true while the last character of `s` is missing continuation bytes
    static bool ends_inside_utf8(const std::string &s) {
        size_t back = 0; // continuation bytes seen from the end
        while (back < 4 && back < s.size()) {
            const unsigned char c = (unsigned char)s[s.size() - 1 - back];
            if ((c & 0xC0) != 0x80) { // a lead byte (or ASCII): how long does it say the character is?
                const size_t want = (c & 0xE0) == 0xC0 ? 2 : (c & 0xF0) == 0xE0 ? 3 : (c & 0xF8) == 0xF0 ? 4 : 1;
                return want > 1 && back + 1 < want;
            }
            back++;
        }
        return false;
    }

    // position in `text` where its longest suffix that is a prefix of `word` begins (npos: none)
    static size_t prefix_overlap(const std::string &text, const std::string &word) {
        for (size_t n = std::min(word.size(), text.size()); n > 0; n--)
            if (text.compare(text.size() - n, n, word, 0, n) == 0) return text.size() - n;
        return std::string::npos;
    }

  private:
    const std::vector<std::string> &words() const {
        static const std::vector<std::string> none;
        return stops_ ? *stops_ : none;
    }
    const std::vector<std::string> *stops_ = nullptr;
    std::string kept_;
    size_t released_ = 0;
};

// A prompt longer than the slot's context keeps its first `n_keep` tokens and its tail; whole blocks of half the
// remaining window are dropped from the middle.  Returns true if tokens were dropped.
inline bool fit_prompt_to_context(std::vector<int32_t> &tokens, int n_ctx, int n_keep) {
    if ((int)tokens.size() < n_ctx) return false;
    const int window = n_ctx - n_keep;
    const int block = std::max(1, window / 2);
    const int drop = ((int)tokens.size() - n_keep - block) / block * block;
    tokens.erase(tokens.begin() + n_keep, tokens.begin() + n_keep + drop);
    while ((int)tokens.size() >= n_ctx) tokens.erase(tokens.begin() + n_keep); // (cannot happen for n_keep <= n_ctx - 4; the prompt must fit)
    return true;
}

// the request's n_keep as the truncation uses it: negative = the whole prompt, never more than n_ctx - 4 (grpc-server.cpp:1694-1698)
inline int effective_n_keep(int n_keep, int n_prompt, int n_ctx) {
    if (n_keep < 0) n_keep = n_prompt;
    n_keep = std::min(n_ctx - 4, n_keep);
    return n_keep < 0 ? 0 : n_keep;
}

// tokens of `prompt` whose K/V the slot still holds from its previous request (never the whole cached sequence: its last
// token is evaluated again)
inline int reusable_prefix(const std::vector<int32_t> &cached, const std::vector<int32_t> &prompt) {
    size_t n = 0;
    while (n < cached.size() && n < prompt.size() && cached[n] == prompt[n]) n++;
    if (n > 0 && n == cached.size()) n--;
    return (int)n;
}

} // namespace b2
