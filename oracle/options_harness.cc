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

// params_parse (:2205-2300): ModelOptions -> gpt_params, compiled the same way.  gpt_params below has the members the
// function assigns, initialised to llama.cpp's defaults where the function only overrides on non-zero input (common.h @
// d5cb868: rope_freq_base / rope_freq_scale 0 = from the model, yarn_ext_factor -1, yarn_attn_factor 1, beta_fast 32,
// beta_slow 1, n_parallel 1); the enum values are llama.h's.
#include <cstdlib>
#include <regex>
#define GGML_ASSERT(x) do { if (!(x)) abort(); } while (0)
enum llama_rope_scaling_type { LLAMA_ROPE_SCALING_TYPE_UNSPECIFIED = -1, LLAMA_ROPE_SCALING_TYPE_NONE = 0, LLAMA_ROPE_SCALING_TYPE_LINEAR = 1, LLAMA_ROPE_SCALING_TYPE_YARN = 2 };
static size_t llama_max_devices() { return 16; }
struct ref_lora { std::string path; float scale; };
struct ref_cpuparams { int n_threads = -1; };
struct gpt_params {
    std::string model, mmproj, model_alias, rpc_servers;
    int n_ctx = 0, n_gpu_layers = -1, n_batch = 2048, n_parallel = 1, main_gpu = 0;
    bool cont_batching = true, use_mlock = false, use_mmap = true, flash_attn = false, no_kv_offload = false, embedding = false;
    ref_cpuparams cpuparams;
    float tensor_split[128] = {0};
    std::vector<ref_lora> lora_adapters;
    llama_rope_scaling_type rope_scaling_type = LLAMA_ROPE_SCALING_TYPE_UNSPECIFIED;
    float yarn_ext_factor = -1.0f, yarn_attn_factor = 1.0f, yarn_beta_fast = 32.0f, yarn_beta_slow = 1.0f;
    float rope_freq_base = 0.0f, rope_freq_scale = 0.0f;
};

extern "C" struct ref_model_options {
    const char *modelfile, *mmproj, *tensorsplit, *maingpu, *loraadapter, *lorabase, *ropescaling;
    int32_t contextsize, threads, ngpulayers, nbatch, mlock, mmap, flashattention, nokvoffload, embeddings;
    float lorascale, yarnextfactor, yarnattnfactor, yarnbetafast, yarnbetaslow, ropefreqbase, ropefreqscale;
};
namespace backend {
class ModelOptions {
  public:
    explicit ModelOptions(const ref_model_options &o) : o_(o) {}
    std::string modelfile() const { return s(o_.modelfile); }
    std::string mmproj() const { return s(o_.mmproj); }
    std::string tensorsplit() const { return s(o_.tensorsplit); }
    std::string maingpu() const { return s(o_.maingpu); }
    std::string loraadapter() const { return s(o_.loraadapter); }
    std::string lorabase() const { return s(o_.lorabase); }
    std::string ropescaling() const { return s(o_.ropescaling); }
    int32_t contextsize() const { return o_.contextsize; }
    int32_t threads() const { return o_.threads; }
    int32_t ngpulayers() const { return o_.ngpulayers; }
    int32_t nbatch() const { return o_.nbatch; }
    bool mlock() const { return o_.mlock != 0; }
    bool mmap() const { return o_.mmap != 0; }
    bool flashattention() const { return o_.flashattention != 0; }
    bool nokvoffload() const { return o_.nokvoffload != 0; }
    bool embeddings() const { return o_.embeddings != 0; }
    float lorascale() const { return o_.lorascale; }
    float yarnextfactor() const { return o_.yarnextfactor; }
    float yarnattnfactor() const { return o_.yarnattnfactor; }
    float yarnbetafast() const { return o_.yarnbetafast; }
    float yarnbetaslow() const { return o_.yarnbetaslow; }
    float ropefreqbase() const { return o_.ropefreqbase; }
    float ropefreqscale() const { return o_.ropefreqscale; }

  private:
    static std::string s(const char *p) { return p ? p : ""; }
    ref_model_options o_;
};
} // namespace backend

#include "_ref/params_parse.inc"

extern "C" int ref_params_parse(const ref_model_options *o, char *out, int cap) {
    gpt_params p;
    const backend::ModelOptions m(*o);
    params_parse(&m, p);
    json j = {{"model", p.model}, {"n_ctx", p.n_ctx}, {"n_batch", p.n_batch}, {"n_parallel", p.n_parallel}, {"main_gpu", p.main_gpu},
              {"embedding", p.embedding}, {"rope_scaling_type", (int)p.rope_scaling_type}, {"yarn_ext_factor", p.yarn_ext_factor},
              {"yarn_attn_factor", p.yarn_attn_factor}, {"yarn_beta_fast", p.yarn_beta_fast}, {"yarn_beta_slow", p.yarn_beta_slow},
              {"rope_freq_base", p.rope_freq_base}, {"rope_freq_scale", p.rope_freq_scale}, {"flash_attn", p.flash_attn},
              {"use_mmap", p.use_mmap}, {"use_mlock", p.use_mlock}, {"n_threads", p.cpuparams.n_threads}, {"n_gpu_layers", p.n_gpu_layers}};
    const std::string s = j.dump();
    if ((int)s.size() >= cap) return -1;
    memcpy(out, s.c_str(), s.size() + 1);
    return (int)s.size();
}
