// stop_harness.cc -- C face of two helpers of the reference's server, compiled FROM THE REFERENCE'S OWN SOURCE.
//
// Test infrastructure only.  backend/cpp/llama/grpc-server.cpp as a whole needs llama.cpp, gRPC and protobuf and does not
// build here, but `common_part` (how much of a prompt a slot's cache already holds, :1740), `ends_with` and
// `find_partial_stop_string` (the rule that decides which tail of the generated text is held back because it may be the
// beginning of a stop word, called from process_token at :1056-1089) use nothing but std::string / std::vector.
// oracle/Makefile cuts those functions out of the file where it lies under /root/reference into
// oracle/_ref/stop_extract.inc (a build artefact, deleted after the compile; nothing of the reference is committed) and
// this file wraps them, so that tests/test_capi_cpu.py can hold the product's own rules (stream_text.h) to the
// reference's code rather than to a restatement of it.
#include <cstddef>
#include <cstdint>
#include <string>
#include <vector>

typedef int32_t llama_token; // llama.h

#include "_ref/stop_extract.inc"

extern "C" int64_t ref_common_part(const int32_t *a, size_t na, const int32_t *b, size_t nb) {
    return (int64_t)common_part(std::vector<llama_token>(a, a + na), std::vector<llama_token>(b, b + nb));
}

extern "C" int64_t ref_find_partial_stop_string(const char *stop, size_t n_stop, const char *text, size_t n_text) {
    const size_t at = find_partial_stop_string(std::string(stop, n_stop), std::string(text, n_text));
    return at == std::string::npos ? (int64_t)-1 : (int64_t)at;
}

// The prompt truncation of update_slots (:1692-1721) is a block inside a 400-line function: the Makefile cuts the block
// out (from "slot.num_prompt_tokens = prompt_tokens.size();" to the closing brace after its GGML_ASSERT) and it is compiled
// here inside a scope that provides the four names it uses.
#include <algorithm>
#include <cstdlib>
#define LOG_VERBOSE(...)
#define GGML_ASSERT(x) do { if (!(x)) abort(); } while (0)
namespace {
struct ref_params { int n_keep; };
struct ref_slot { int num_prompt_tokens; int n_ctx; ref_params params; bool truncated; };
}
extern "C" int ref_truncate_prompt(int32_t *tokens, int n, int n_ctx, int n_keep, int *truncated) {
    ref_slot slot{0, n_ctx, {n_keep}, false};
    std::vector<llama_token> prompt_tokens(tokens, tokens + n);
#include "_ref/truncate_block.inc"
    std::copy(prompt_tokens.begin(), prompt_tokens.end(), tokens);
    if (truncated) *truncated = slot.truncated ? 1 : 0;
    return slot.num_prompt_tokens;
}

// The check for an incomplete UTF-8 character at the end of the generated text (process_token, :1027-1054), cut out the same way.
namespace {
struct ref_text_slot { std::string generated_text; };
}
extern "C" int ref_utf8_incomplete(const char *text, size_t n) {
    ref_text_slot slot{std::string(text, n)};
#include "_ref/utf8_block.inc"
    return incomplete ? 1 : 0;
}

// The text half of process_token as a whole (:1027-1096: incomplete UTF-8, full and partial stop words, what is sent), with
// find_stopping_strings (:976-1008) and the stop_type enum it takes, driven piece by piece.  The scope below provides the
// names the block uses: the slot's text fields, a result with text_to_send, and two no-op callees.
#include "_ref/stop_type.inc"
namespace {
struct ref_stream_params { std::vector<std::string> antiprompt; bool stream = true; };
struct completion_token_output { std::string text_to_send; };
struct llama_client_slot {
    std::string generated_text;
    size_t sent_count = 0;
    bool has_next_token = true, stopped_word = false;
    std::string stopping_word;
    ref_stream_params params;
    void add_token_string(const completion_token_output &) {}
};
void send_partial_response(llama_client_slot &, const completion_token_output &) {}
}
#include "_ref/find_stopping.inc"

extern "C" int ref_stream_text(const char *pieces, const int *lens, int n, const char *const *stops, int n_stops, char *streamed, int cap_s,
                               char *final_text, int cap_f, int *n_used) {
    llama_client_slot slot;
    for (int i = 0; i < n_stops; i++) slot.params.antiprompt.emplace_back(stops[i]);
    std::string out;
    int used = 0;
    size_t off = 0;
    for (int i = 0; i < n; i++) {
        completion_token_output result;
        const std::string token_str(pieces + off, (size_t)lens[i]);
        off += (size_t)lens[i];
        slot.generated_text += token_str; // (:1016-1017)
        slot.has_next_token = true;
#include "_ref/process_block.inc"
        out += result.text_to_send;
        used++;
        if (!slot.has_next_token) break;
    }
    if ((int)out.size() >= cap_s || (int)slot.generated_text.size() >= cap_f) return -1;
    std::copy(out.begin(), out.end(), streamed);
    streamed[out.size()] = 0;
    std::copy(slot.generated_text.begin(), slot.generated_text.end(), final_text);
    final_text[slot.generated_text.size()] = 0;
    if (n_used) *n_used = used;
    return slot.stopped_word ? 1 : 0;
}
