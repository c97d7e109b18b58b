#include "gguf_io.h"

#include <cstdio>
#include <cstring>

namespace mio {

namespace {

struct cursor {
    const uint8_t * p;
    const uint8_t * end;
    bool ok = true;
    template <typename T> T get() {
        T v{};
        if ((size_t) (end - p) < sizeof(T)) { ok = false; return v; }
        std::memcpy(&v, p, sizeof(T));
        p += sizeof(T);
        return v;
    }
    std::string str(bool v1_len32 = false) {
        uint64_t n = v1_len32 ? get<uint32_t>() : get<uint64_t>();
        if (!ok || (uint64_t) (end - p) < n) { ok = false; return {}; }
        std::string s((const char *) p, (size_t) n);
        p += n;
        return s;
    }
};

// gguf value types, gguf.h:53-68
enum { T_U8, T_I8, T_U16, T_I16, T_U32, T_I32, T_F32, T_BOOL, T_STR, T_ARR, T_U64, T_I64, T_F64 };

size_t scalar_size(uint32_t t) {
    switch (t) {
        case T_U8: case T_I8: case T_BOOL: return 1;
        case T_U16: case T_I16: return 2;
        case T_U32: case T_I32: case T_F32: return 4;
        case T_U64: case T_I64: case T_F64: return 8;
        default: return 0;
    }
}

bool read_scalar(cursor & c, uint32_t t, gguf_kv & kv) {
    switch (t) {
        case T_U8:  kv.u = c.get<uint8_t>(); break;
        case T_I8:  kv.u = (uint64_t) (int64_t) c.get<int8_t>(); break;
        case T_U16: kv.u = c.get<uint16_t>(); break;
        case T_I16: kv.u = (uint64_t) (int64_t) c.get<int16_t>(); break;
        case T_U32: kv.u = c.get<uint32_t>(); break;
        case T_I32: kv.u = (uint64_t) (int64_t) c.get<int32_t>(); break;
        case T_F32: kv.f = c.get<float>(); break;
        case T_BOOL: kv.u = c.get<uint8_t>() != 0; break;
        case T_U64: kv.u = c.get<uint64_t>(); break;
        case T_I64: kv.u = (uint64_t) c.get<int64_t>(); break;
        case T_F64: kv.f = c.get<double>(); break;
        default: return false;
    }
    return c.ok;
}

size_t type_size(uint32_t t) {
    switch (t) {
        case GGUF_F32: case GGUF_I32: return 4;
        case GGUF_F16: return 2;
        default: return 0;
    }
}

} // namespace

bool gguf_file::load(const std::string & path, std::string & err) {
    tensors_.clear(); index_.clear(); kv_.clear(); blob_.clear();
    FILE * f = std::fopen(path.c_str(), "rb");
    if (!f) { err = "failed to load GGUF"; return false; }
    std::fseek(f, 0, SEEK_END);
    long sz = std::ftell(f);
    std::fseek(f, 0, SEEK_SET);
    if (sz < 24) { std::fclose(f); err = "failed to load GGUF"; return false; }
    blob_.resize((size_t) sz);
    size_t got = std::fread(blob_.data(), 1, (size_t) sz, f);
    std::fclose(f);
    if (got != (size_t) sz) { err = "failed to load GGUF"; return false; }

    cursor c{blob_.data(), blob_.data() + blob_.size()};
    if (std::memcmp(c.p, "GGUF", 4) != 0) { err = "failed to load GGUF"; return false; }
    c.p += 4;
    version_ = c.get<uint32_t>();
    if (version_ < 2 || version_ > 3) { err = "failed to load GGUF"; return false; }
    const uint64_t n_tensors = c.get<uint64_t>();
    const uint64_t n_kv = c.get<uint64_t>();
    if (!c.ok || n_tensors > (1u << 20) || n_kv > (1u << 20)) { err = "failed to load GGUF"; return false; }

    size_t alignment = 32;
    for (uint64_t i = 0; i < n_kv; ++i) {
        std::string key = c.str();
        gguf_kv kv;
        kv.type = c.get<uint32_t>();
        if (!c.ok) break;
        if (kv.type == T_STR) {
            kv.s = c.str();
        } else if (kv.type == T_ARR) {
            const uint32_t et = c.get<uint32_t>();
            const uint64_t n = c.get<uint64_t>();
            if (et == T_STR) {
                for (uint64_t j = 0; j < n && c.ok; ++j) c.str();
            } else {
                const size_t es = scalar_size(et);
                if (es == 0 || n > (uint64_t) (c.end - c.p) / es) { c.ok = false; break; }
                c.p += n * es;
            }
        } else if (!read_scalar(c, kv.type, kv)) {
            c.ok = false;
        }
        if (!c.ok) break;
        if (key == "general.alignment" && kv.type == T_U32 && kv.u > 0) alignment = (size_t) kv.u;
        kv_[key] = kv;
    }
    if (!c.ok) { err = "failed to load GGUF"; return false; }

    tensors_.resize((size_t) n_tensors);
    for (uint64_t i = 0; i < n_tensors; ++i) {
        gguf_tensor & t = tensors_[(size_t) i];
        t.name = c.str();
        const uint32_t nd = c.get<uint32_t>();
        if (!c.ok || nd > 4) { err = "failed to load GGUF"; return false; }
        for (uint32_t d = 0; d < nd; ++d) t.ne.push_back((int64_t) c.get<uint64_t>());
        t.type = c.get<uint32_t>();
        t.offset = c.get<uint64_t>();
        if (!c.ok) { err = "failed to load GGUF"; return false; }
    }
    size_t data_start = (size_t) (c.p - blob_.data());
    data_start = (data_start + alignment - 1) / alignment * alignment;
    for (size_t i = 0; i < tensors_.size(); ++i) {
        gguf_tensor & t = tensors_[i];
        const size_t es = type_size(t.type);
        if (es == 0) { err = "unsupported tensor type in GGUF: " + t.name; return false; }
        uint64_t bytes = 0;
        if (!t.checked_bytes(es, bytes) || data_start > blob_.size() || t.offset > blob_.size() - data_start ||
            bytes > blob_.size() - data_start - t.offset) {
            err = "failed to load GGUF";
            return false;
        }
        t.data = blob_.data() + data_start + t.offset;
        index_[t.name] = i;
    }
    return true;
}

const gguf_tensor * gguf_file::find(const std::string & name) const {
    auto it = index_.find(name);
    return it == index_.end() ? nullptr : &tensors_[it->second];
}

bool gguf_file::get_u32(const std::string & key, uint32_t & out) const {
    // get_u32_kv in /root/reference/src/miocodec-decoder.cpp:356-372 accepts exactly UINT32
    auto it = kv_.find(key);
    if (it == kv_.end() || it->second.type != T_U32) return false;
    out = (uint32_t) it->second.u;
    return true;
}

bool gguf_file::get_f32(const std::string & key, float & out) const {
    auto it = kv_.find(key);
    if (it == kv_.end() || it->second.type != T_F32) return false;
    out = (float) it->second.f;
    return true;
}

bool gguf_write_f32_vector(const std::string & path,
                           const std::vector<std::pair<std::string, std::string>> & str_kvs,
                           const std::vector<std::pair<std::string, uint32_t>> & u32_kvs,
                           const std::string & tensor_name,
                           const float * data, size_t n, std::string & err) {
    std::vector<uint8_t> out;
    auto put = [&](const void * p, size_t k) { const uint8_t * b = (const uint8_t *) p; out.insert(out.end(), b, b + k); };
    auto put_u32 = [&](uint32_t v) { put(&v, 4); };
    auto put_u64 = [&](uint64_t v) { put(&v, 8); };
    auto put_str = [&](const std::string & s) { put_u64(s.size()); put(s.data(), s.size()); };
    put("GGUF", 4);
    put_u32(3);
    put_u64(1);
    put_u64(str_kvs.size() + u32_kvs.size());
    for (auto & kv : str_kvs) { put_str(kv.first); put_u32(T_STR); put_str(kv.second); }
    for (auto & kv : u32_kvs) { put_str(kv.first); put_u32(T_U32); put_u32(kv.second); }
    put_str(tensor_name);
    put_u32(1);
    put_u64(n);
    put_u32(GGUF_F32);
    put_u64(0);
    while (out.size() % 32 != 0) out.push_back(0);
    put(data, n * sizeof(float));
    while (out.size() % 32 != 0) out.push_back(0);
    FILE * f = std::fopen(path.c_str(), "wb");
    if (!f) { err = "failed to write embedding GGUF"; return false; }
    const size_t w = std::fwrite(out.data(), 1, out.size(), f);
    const int rc = std::fclose(f);
    if (w != out.size() || rc != 0) { err = "failed to write embedding GGUF"; return false; }
    return true;
}

} // namespace mio
