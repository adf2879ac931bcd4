// options_harness.cc -- parse_options of the reference's server (backend/cpp/llama/grpc-server.cpp:2068-2137), compiled FROM
// THE REFERENCE'S OWN SOURCE, behind a C face.
//
// Test infrastructure only.  parse_options turns a backend.PredictOptions message into the json the slot is launched with;
// it needs nothing of llama.cpp or gRPC except the message's getters and nlohmann's json.hpp, which the reference vendors.
// oracle/Makefile cuts the function out of the file where it lies (oracle/_ref/parse_options.inc, deleted after the
// compile), json.hpp is included from /root/reference, and the class below stands in for the generated protobuf class:
// same getter names, values from a plain struct.  tests/test_capi_cpu.py compares the json with what pyshim's
// request_kwargs (the Python twin of backend/go/llm/b200/llm.go's mapping) makes of the same message.
#include <cstdint>
#include <cstring>
#include <string>
#include <vector>

#include "json.hpp" // -I/root/reference/backend/cpp/llama
using json = nlohmann::json;

extern "C" struct ref_predict_options {
    int32_t tokens, topk, repeat, mirostat, nkeep, seed;
    float topp, tailfreesamplingz, typicalp, temperature, penalty, frequencypenalty, presencepenalty, mirostattau, mirostateta;
    int32_t promptcacheall, penalizenl, ignoreeos, embeddings;
    const char *grammar, *prompt, *correlationid;
    const char *const *stopprompts;
    int32_t n_stopprompts;
};

namespace backend {
class PredictOptions {
  public:
    explicit PredictOptions(const ref_predict_options &o) : o_(o) {
        for (int i = 0; i < o.n_stopprompts; i++) stops_.emplace_back(o.stopprompts[i]);
    }
    bool promptcacheall() const { return o_.promptcacheall != 0; }
    int32_t tokens() const { return o_.tokens; }
    int32_t topk() const { return o_.topk; }
    float topp() const { return o_.topp; }
    float tailfreesamplingz() const { return o_.tailfreesamplingz; }
    float typicalp() const { return o_.typicalp; }
    float temperature() const { return o_.temperature; }
    int32_t repeat() const { return o_.repeat; }
    float penalty() const { return o_.penalty; }
    float frequencypenalty() const { return o_.frequencypenalty; }
    float presencepenalty() const { return o_.presencepenalty; }
    int32_t mirostat() const { return o_.mirostat; }
    float mirostattau() const { return o_.mirostattau; }
    float mirostateta() const { return o_.mirostateta; }
    bool penalizenl() const { return o_.penalizenl != 0; }
    int32_t nkeep() const { return o_.nkeep; }
    int32_t seed() const { return o_.seed; }
    std::string grammar() const { return o_.grammar ? o_.grammar : ""; }
    std::string prompt() const { return o_.prompt ? o_.prompt : ""; }
    bool ignoreeos() const { return o_.ignoreeos != 0; }
    bool embeddings() const { return o_.embeddings != 0; }
    std::string correlationid() const { return o_.correlationid ? o_.correlationid : ""; }
    int images_size() const { return 0; }
    std::string images(int) const { return ""; }
    const std::vector<std::string> &stopprompts() const { return stops_; }

  private:
    ref_predict_options o_;
    std::vector<std::string> stops_;
};
} // namespace backend
struct llama_server_context {};

#include "_ref/parse_options.inc"

extern "C" int ref_parse_options(int streaming, const ref_predict_options *o, char *out, int cap) {
    llama_server_context llama;
    const backend::PredictOptions p(*o);
    const std::string s = parse_options(streaming != 0, &p, llama).dump();
    if ((int)s.size() >= cap) return -1;
    memcpy(out, s.c_str(), s.size() + 1);
    return (int)s.size();
}
