// Minimal GGUF v2/v3 reader + single-tensor writer (host side, no ggml).
//
// Format constants follow /root/reference/llama.cpp/ggml/include/gguf.h:41-46,53-68 and the
// parser in /root/reference/llama.cpp/ggml/src/gguf.cpp:319-480 (magic "GGUF", version,
// n_tensors, n_kv, KV pairs, tensor infos, data section aligned to general.alignment, default 32).
// Only what the MioCodec decode path needs: scalar/string/array KVs, F32/F16/I32 tensors.
#pragma once
#include <cstdint>
#include <map>
#include <string>
#include <vector>

namespace mio {

enum gguf_dtype : uint32_t { GGUF_F32 = 0, GGUF_F16 = 1, GGUF_I32 = 26 };

struct gguf_tensor {
    std::string name;
    std::vector<int64_t> ne;   // ggml order: ne[0] is the contiguous dimension
    uint32_t type = 0;
    uint64_t offset = 0;       // relative to the data section
    const uint8_t * data = nullptr;
    int64_t nelements() const { int64_t n = 1; for (auto d : ne) n *= d; return n; }
    // element count with every dimension > 0 and no 64-bit overflow of count * elem_size; false otherwise
    bool checked_bytes(size_t elem_size, uint64_t & bytes) const {
        uint64_t n = 1;
        for (auto d : ne) {
            if (d <= 0 || (uint64_t) d > UINT64_MAX / n) return false;
            n *= (uint64_t) d;
        }
        if (elem_size == 0 || n > UINT64_MAX / elem_size) return false;
        bytes = n * elem_size;
        return true;
    }
};

struct gguf_kv {
    uint32_t type = 0;         // gguf value type
    uint64_t u = 0;            // integers / bool
    double f = 0.0;            // floats
    std::string s;             // strings
};

class gguf_file {
public:
    // Reads the whole file into memory. Returns false and sets err on malformed input.
    bool load(const std::string & path, std::string & err);

    const gguf_tensor * find(const std::string & name) const;
    bool get_u32(const std::string & key, uint32_t & out) const;
    bool get_f32(const std::string & key, float & out) const;
    const std::vector<gguf_tensor> & tensors() const { return tensors_; }
    uint32_t version() const { return version_; }

private:
    std::vector<uint8_t> blob_;
    std::vector<gguf_tensor> tensors_;
    std::map<std::string, size_t> index_;
    std::map<std::string, gguf_kv> kv_;
    uint32_t version_ = 0;
};

// Writes a GGUF v3 file holding string/u32 KVs and one F32 1-D tensor (the layout
// gguf_write_to_file produces for /root/reference/src/mio-tts-lib.cpp:253-310).
bool gguf_write_f32_vector(const std::string & path,
                           const std::vector<std::pair<std::string, std::string>> & str_kvs,
                           const std::vector<std::pair<std::string, uint32_t>> & u32_kvs,
                           const std::string & tensor_name,
                           const float * data, size_t n, std::string & err);

} // namespace mio
