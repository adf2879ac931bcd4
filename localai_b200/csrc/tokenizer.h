// tokenizer.h -- SentencePiece-style ("llama") and byte-level BPE ("gpt2", pre-tokeniser "llama-bpe")
// tokenisers driven by the vocabulary stored in the GGUF file.
// Replaces ::llama_tokenize / llama_token_to_piece as called from the reference engine
// (backend/cpp/llama/grpc-server.cpp:559-607, :116-124, :1012).
#pragma once
#include <cstdint>
#include <map>
#include <string>
#include <unordered_map>
#include <vector>

#include "gguf.h"

namespace b2 {

enum TokenAttr { TOK_NORMAL = 1, TOK_UNKNOWN = 2, TOK_CONTROL = 3, TOK_USER_DEFINED = 4, TOK_UNUSED = 5, TOK_BYTE = 6 };

class Tokenizer {
  public:
    explicit Tokenizer(const GGUFFile &f);

    // text -> ids.  add_bos: prepend BOS if the vocabulary says so (tokenizer.ggml.add_bos_token).
    // parse_special: control tokens written out in the text ("<|eot_id|>") become their ids.
    std::vector<int32_t> tokenize(const std::string &text, bool add_bos, bool parse_special) const;
    // id -> bytes (control tokens render as "", like llama_token_to_piece with special=false)
    std::string token_to_piece(int32_t id, bool render_special = false) const;

    int n_vocab() const { return (int)tokens_.size(); }
    int32_t bos_id() const { return bos_; }
    int32_t eos_id() const { return eos_; }
    int32_t eot_id() const { return eot_; }
    int32_t nl_id() const { return nl_; } // llama_token_nl: the byte / byte-level token of '\n'
    bool add_bos() const { return add_bos_; }
    bool is_eog(int32_t id) const { return id >= 0 && (id == eos_ || id == eot_); }
    const std::string &model_name() const { return model_; }

  private:
    enum Kind { SPM, BPE, NONE };
    void spm_tokenize(const std::string &text, std::vector<int32_t> &out) const;
    void bpe_tokenize(const std::string &text, std::vector<int32_t> &out) const;
    int32_t byte_to_token(uint8_t b) const;
    int bpe_rank(const std::string &l, const std::string &r) const;

    Kind kind_ = NONE;
    std::string model_, pre_;
    std::vector<std::string> tokens_;
    std::vector<float> scores_;
    std::vector<int> types_;
    std::unordered_map<std::string, int32_t> token_to_id_;
    std::unordered_map<std::string, int> bpe_ranks_; // "left right" -> rank
    std::vector<int32_t> special_;                   // ids matched verbatim in text, longest first
    int32_t bos_ = -1, eos_ = -1, eot_ = -1, unk_ = -1, nl_ = -1;
    bool add_bos_ = true, add_eos_ = false, add_space_prefix_ = true, ignore_merges_ = false;
};

// exposed for tests
std::vector<std::string> llama3_pretokenize(const std::string &text, int max_digits = 3);

} // namespace b2
