// gguf.cc -- GGUF container reader (mmap).  See gguf.h.
#include "gguf.h"

#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

#include <cstring>

#include "common.h"

namespace b2 {

namespace {
struct Cursor {
    const uint8_t *p, *end;
    template <typename T> T rd() {
        if (p + sizeof(T) > end) throw Error(B2_EIO, "gguf: truncated file");
        T v;
        memcpy(&v, p, sizeof(T));
        p += sizeof(T);
        return v;
    }
    std::string str() {
        const uint64_t n = rd<uint64_t>();
        if (n > (uint64_t)(end - p)) throw Error(B2_EIO, "gguf: bad string length");
        std::string s((const char *)p, (size_t)n);
        p += n;
        return s;
    }
};

void read_scalar(Cursor &c, GGUFType t, GGUFValue &v) {
    switch (t) {
        case GT_U8: v.u = c.rd<uint8_t>(); v.f = (double)v.u; break;
        case GT_I8: { int8_t x = c.rd<int8_t>(); v.u = (uint64_t)(int64_t)x; v.f = x; } break;
        case GT_U16: v.u = c.rd<uint16_t>(); v.f = (double)v.u; break;
        case GT_I16: { int16_t x = c.rd<int16_t>(); v.u = (uint64_t)(int64_t)x; v.f = x; } break;
        case GT_U32: v.u = c.rd<uint32_t>(); v.f = (double)v.u; break;
        case GT_I32: { int32_t x = c.rd<int32_t>(); v.u = (uint64_t)(int64_t)x; v.f = x; } break;
        case GT_F32: v.f = c.rd<float>(); v.u = (uint64_t)v.f; break;
        case GT_BOOL: v.u = c.rd<uint8_t>() ? 1 : 0; v.f = (double)v.u; break;
        case GT_STR: v.s = c.str(); break;
        case GT_U64: v.u = c.rd<uint64_t>(); v.f = (double)v.u; break;
        case GT_I64: { int64_t x = c.rd<int64_t>(); v.u = (uint64_t)x; v.f = (double)x; } break;
        case GT_F64: v.f = c.rd<double>(); v.u = (uint64_t)v.f; break;
        default: throw Error(B2_EIO, "gguf: unknown value type");
    }
}
} // namespace

GGUFFile::GGUFFile(const std::string &path) {
    fd_ = open(path.c_str(), O_RDONLY);
    if (fd_ < 0) throw Error(B2_EIO, "gguf: cannot open " + path + ": " + strerror(errno));
    struct stat st;
    if (fstat(fd_, &st) != 0 || st.st_size < 24) {
        close(fd_);
        throw Error(B2_EIO, "gguf: cannot stat / file too small: " + path);
    }
    file_size = (size_t)st.st_size;
    void *m = mmap(nullptr, file_size, PROT_READ, MAP_PRIVATE, fd_, 0);
    if (m == MAP_FAILED) {
        close(fd_);
        throw Error(B2_EIO, "gguf: mmap failed: " + path);
    }
    map_ = (const uint8_t *)m;
    try {
        Cursor c{map_, map_ + file_size};
        if (c.rd<uint32_t>() != 0x46554747u) throw Error(B2_EIO, "gguf: bad magic (not a GGUF file): " + path);
        version = c.rd<uint32_t>();
        if (version < 2 || version > 3) throw Error(B2_EIO, "gguf: unsupported version " + std::to_string(version));
        const uint64_t n_tensors = c.rd<uint64_t>();
        const uint64_t n_kv = c.rd<uint64_t>();
        for (uint64_t i = 0; i < n_kv; i++) {
            std::string key = c.str();
            GGUFValue v;
            v.type = (GGUFType)c.rd<uint32_t>();
            if (v.type == GT_ARR) {
                v.elem_type = (GGUFType)c.rd<uint32_t>();
                const uint64_t n = c.rd<uint64_t>();
                // an array element takes at least one byte of the file (strings: 8): a count beyond what is left of
                // the file is a corrupt header, not something to reserve memory for
                if (n > (uint64_t)(c.end - c.p)) throw Error(B2_EIO, "gguf: array length exceeds the file size (key " + key + ")");
                if (v.elem_type == GT_STR) {
                    v.arr_s.reserve(n);
                    for (uint64_t k = 0; k < n; k++) v.arr_s.push_back(c.str());
                } else if (v.elem_type == GT_F32 || v.elem_type == GT_F64) {
                    v.arr_f.reserve(n);
                    for (uint64_t k = 0; k < n; k++) {
                        GGUFValue e;
                        read_scalar(c, v.elem_type, e);
                        v.arr_f.push_back(e.f);
                    }
                } else {
                    v.arr_i.reserve(n);
                    for (uint64_t k = 0; k < n; k++) {
                        GGUFValue e;
                        read_scalar(c, v.elem_type, e);
                        v.arr_i.push_back((int64_t)e.u);
                    }
                }
            } else {
                read_scalar(c, v.type, v);
            }
            kv.emplace(std::move(key), std::move(v));
        }
        if (n_tensors > (uint64_t)(c.end - c.p) / 24) throw Error(B2_EIO, "gguf: tensor count exceeds the file size"); // >= 24 bytes per info
        tensors.resize(n_tensors);
        for (uint64_t i = 0; i < n_tensors; i++) {
            GGUFTensorInfo &t = tensors[i];
            t.name = c.str();
            const uint32_t nd = c.rd<uint32_t>();
            if (nd > 4) throw Error(B2_EIO, "gguf: tensor with more than 4 dims");
            for (uint32_t d = 0; d < nd; d++) t.ne.push_back((int64_t)c.rd<uint64_t>());
            t.type = (int)c.rd<uint32_t>();
            t.offset = c.rd<uint64_t>();
        }
        const uint64_t align = get_u("general.alignment", 32);
        if (align == 0 || (align & (align - 1)) != 0 || align > (1u << 20)) throw Error(B2_EIO, "gguf: general.alignment must be a power of two");
        const size_t meta = (size_t)(c.p - map_);
        const size_t data_off = (meta + align - 1) / align * align;
        if (data_off > file_size) throw Error(B2_EIO, "gguf: header exceeds file size");
        for (size_t i = 0; i < tensors.size(); i++) {
            GGUFTensorInfo &t = tensors[i];
            int64_t n = 1;
            for (auto e : t.ne) { // the element count must not overflow
                if (e <= 0 || n > (int64_t)(INT64_MAX / 4) / e) throw Error(B2_EIO, "gguf: tensor " + t.name + " has an invalid shape");
                n *= e;
            }
            const int be = type_block_elems(t.type), bb = type_block_bytes(t.type);
            if (be == 0) {
                t.n_bytes = -1; // unsupported type: only an error if the model needs this tensor
            } else {
                if (t.ne.empty() || t.ne[0] % be != 0) throw Error(B2_EIO, "gguf: tensor " + t.name + " row not a block multiple");
                t.n_bytes = n / be * bb;
                // overflow-safe: offset <= file_size - data_off - n_bytes
                const uint64_t room = (uint64_t)(file_size - data_off);
                if ((uint64_t)t.n_bytes > room || t.offset > room - (uint64_t)t.n_bytes)
                    throw Error(B2_EIO, "gguf: tensor " + t.name + " exceeds file size");
            }
            t.data = map_ + data_off + t.offset;
            index_[t.name] = i;
        }
    } catch (...) {
        munmap((void *)map_, file_size);
        close(fd_);
        map_ = nullptr;
        fd_ = -1;
        throw;
    }
}

GGUFFile::~GGUFFile() {
    if (map_) munmap((void *)map_, file_size);
    if (fd_ >= 0) close(fd_);
}

const GGUFValue *GGUFFile::get(const std::string &k) const {
    auto it = kv.find(k);
    return it == kv.end() ? nullptr : &it->second;
}
uint64_t GGUFFile::get_u(const std::string &k, uint64_t def) const {
    auto *v = get(k);
    if (!v) return def;
    if (v->type == GT_F32 || v->type == GT_F64) return (uint64_t)v->f;
    return v->u;
}
double GGUFFile::get_f(const std::string &k, double def) const {
    auto *v = get(k);
    return v ? v->f : def;
}
std::string GGUFFile::get_s(const std::string &k, const std::string &def) const {
    auto *v = get(k);
    return v && v->type == GT_STR ? v->s : def;
}
const GGUFTensorInfo *GGUFFile::tensor(const std::string &name) const {
    auto it = index_.find(name);
    return it == index_.end() ? nullptr : &tensors[it->second];
}

} // namespace b2
