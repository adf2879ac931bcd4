// gguf.h -- minimal GGUF v2/v3 container reader (host side of LoadModel).
// Replaces what llama_init_from_gpt_params does with the file (grpc-server.cpp:461): header, metadata KVs,
// tensor infos, 32-byte aligned data section, mmap'ed.
#pragma once
#include <cstdint>
#include <map>
#include <string>
#include <vector>

namespace b2 {

enum GGUFType : uint32_t {
    GT_U8 = 0, GT_I8, GT_U16, GT_I16, GT_U32, GT_I32, GT_F32, GT_BOOL, GT_STR, GT_ARR, GT_U64, GT_I64, GT_F64
};

struct GGUFValue {
    GGUFType type = GT_U32;
    GGUFType elem_type = GT_U32; // for arrays
    uint64_t u = 0;              // integers / bool
    double f = 0;                // floats
    std::string s;
    std::vector<std::string> arr_s;
    std::vector<double> arr_f;
    std::vector<int64_t> arr_i;
};

struct GGUFTensorInfo {
    std::string name;
    std::vector<int64_t> ne; // ne[0] = innermost
    int type = 0;
    uint64_t offset = 0; // relative to data section
    const uint8_t *data = nullptr;
    int64_t n_bytes = 0;
};

class GGUFFile {
  public:
    explicit GGUFFile(const std::string &path);
    ~GGUFFile();
    GGUFFile(const GGUFFile &) = delete;

    bool has(const std::string &k) const { return kv.count(k) != 0; }
    uint64_t get_u(const std::string &k, uint64_t def) const;
    double get_f(const std::string &k, double def) const;
    std::string get_s(const std::string &k, const std::string &def) const;
    bool get_b(const std::string &k, bool def) const { return get_u(k, def ? 1 : 0) != 0; }
    const GGUFValue *get(const std::string &k) const;
    const GGUFTensorInfo *tensor(const std::string &name) const;

    std::map<std::string, GGUFValue> kv;
    std::vector<GGUFTensorInfo> tensors;
    uint32_t version = 0;
    size_t file_size = 0;

  private:
    int fd_ = -1;
    const uint8_t *map_ = nullptr;
    std::map<std::string, size_t> index_;
};

} // namespace b2
