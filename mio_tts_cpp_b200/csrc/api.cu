// C ABI of libmio_tts_b200.so (include/mio_tts_b200.h).  Host-side behaviour -- argument checks, their order,
// error strings, malloc ownership -- follows /root/reference/src/mio-tts-lib.cpp function by function
// (cited below as lib.cpp:LINE); the decode itself runs on the GPU through mio::engine, with no CPU fallback.
#include "../../include/mio_tts_b200.h"

#include "engine.h"
#include "gguf_io.h"

#include <dlfcn.h>

#include <algorithm>
#include <cctype>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <limits>
#include <optional>
#include <unordered_map>

using namespace mio;

struct mio_tts_context {
    std::shared_ptr<model> mdl;
    std::unique_ptr<engine> eng;
    std::string vocoder_path, wavlm_path;
};

struct mio_tts_vocab_map {
    std::unordered_map<llama_token, int32_t> token_to_code;
};

namespace {

constexpr const char * k_embedding_tensor = "mio.global_embedding";   // lib.cpp:29
constexpr int32_t k_code_min = 0, k_code_max = 12799;                 // lib.cpp:30-31

void set_err(char * err, size_t err_size, const std::string & msg) {   // lib.cpp:65-69
    if (err != nullptr && err_size > 0) std::snprintf(err, err_size, "%s", msg.c_str());
}

// "CUDA3" / "cuda:3" / "3" -> 3; "" -> -1 (keep); anything else -> -2
int parse_device(const std::string & name) {
    if (name.empty()) return -1;
    std::string s;
    for (char c : name) s.push_back((char) std::toupper((unsigned char) c));
    if (s == "CPU") return -3;
    size_t i = 0;
    if (s.rfind("CUDA", 0) == 0) i = 4;
    else if (s.rfind("GPU", 0) == 0) i = 3;
    if (i < s.size() && s[i] == ':') ++i;
    if (i >= s.size()) return i == 0 ? -2 : 0;
    for (size_t j = i; j < s.size(); ++j) if (!std::isdigit((unsigned char) s[j])) return -2;
    return std::atoi(s.c_str() + i);
}

int default_device() {   // MIO_TTS_BACKEND_DEVICE, /root/reference/src/miocodec-decoder.cpp:68-72
    const char * e = std::getenv("MIO_TTS_BACKEND_DEVICE");
    if (e != nullptr && e[0] != '\0') {
        const int d = parse_device(e);
        if (d >= 0) return d;
    }
    return 0;
}

std::shared_ptr<model> load_model(const std::string & path, int device, std::string & err) {
    int n_dev = 0;
    if (cudaGetDeviceCount(&n_dev) != cudaSuccess || n_dev <= 0) {
        cudaGetLastError();
        err = "no CUDA device available (the B200 decode library has no CPU fallback)";
        return nullptr;
    }
    if (device >= n_dev) { err = "failed to initialize backend device: CUDA" + std::to_string(device); return nullptr; }
    auto m = std::make_shared<model>();
    if (!m->load(path, device, err)) return nullptr;
    return m;
}

// ---- text codes (lib.cpp:71-113, 225-286) ----
std::string strip_ascii_punct(std::string s) {
    while (!s.empty() && std::ispunct((unsigned char) s.front()) && s.front() != '<' && s.front() != '-') s.erase(s.begin());
    while (!s.empty() && std::ispunct((unsigned char) s.back()) && s.back() != '>' && !std::isdigit((unsigned char) s.back())) s.pop_back();
    return s;
}

bool parse_code_token(const std::string & raw, int32_t & out) {
    std::string token = strip_ascii_punct(raw);
    if (token.empty()) return false;
    if (token.rfind("<|s_", 0) == 0 && token.size() > 6 && token.substr(token.size() - 2) == "|>") token = token.substr(4, token.size() - 6);
    try {
        size_t used = 0;
        const long long v = std::stoll(token, &used, 10);
        if (used != token.size()) return false;
        if (v < (long long) std::numeric_limits<int32_t>::min() || v > (long long) std::numeric_limits<int32_t>::max()) return false;
        out = (int32_t) v;
        return true;
    } catch (...) {
        return false;
    }
}

bool parse_codec_piece(const std::string & piece, int32_t & code) {   // lib.cpp:115-138
    if (piece.size() < 6 || piece.compare(0, 4, "<|s_") != 0) return false;
    if (piece[piece.size() - 2] != '|' || piece.back() != '>') return false;
    const std::string num = piece.substr(4, piece.size() - 6);
    if (num.empty()) return false;
    char * end = nullptr;
    const long value = std::strtol(num.c_str(), &end, 10);
    if (end == nullptr || *end != '\0' || value < 0 || value > std::numeric_limits<int32_t>::max()) return false;
    code = (int32_t) value;
    return true;
}

// ---- libllama tokenizer entry points, resolved from the host process (lib.cpp:140-177 uses them directly) ----
typedef int32_t (*llama_token_to_piece_fn)(const struct llama_vocab *, llama_token, char *, int32_t, int32_t, bool);
typedef int32_t (*llama_tokenize_fn)(const struct llama_vocab *, const char *, int32_t, llama_token *, int32_t, bool, bool);

bool token_to_piece(llama_token_to_piece_fn fn, const llama_vocab * vocab, llama_token token, std::string & out) {
    out.clear();
    int cap = 32;
    for (int i = 0; i < 8; ++i) {
        std::vector<char> buf((size_t) cap);
        const int n = fn(vocab, token, buf.data(), (int32_t) buf.size(), 0, true);
        if (n >= 0) { out.assign(buf.data(), (size_t) n); return true; }
        cap = std::max(cap * 2, -n + 1);
    }
    return false;
}

std::optional<llama_token> single_token_id(llama_tokenize_fn tok, llama_token_to_piece_fn t2p, const llama_vocab * vocab,
                                           const std::string & s) {
    const int n = tok(vocab, s.c_str(), (int32_t) s.size(), nullptr, 0, false, true);
    if (n >= 0 || -n != 1) return std::nullopt;
    llama_token t = -1;
    if (tok(vocab, s.c_str(), (int32_t) s.size(), &t, 1, false, true) != 1) return std::nullopt;
    std::string piece;
    if (!token_to_piece(t2p, vocab, t, piece) || piece != s) return std::nullopt;
    return t;
}

template <typename F> bool guarded(char * err, size_t err_size, F && f) {
    try {
        f();
        return true;
    } catch (const std::exception & e) {
        set_err(err, err_size, e.what());
    } catch (...) {
        set_err(err, err_size, "unknown error");
    }
    cudaGetLastError();
    return false;
}

} // namespace

extern "C" {

struct mio_tts_params mio_tts_default_params(void) {
    struct mio_tts_params p;
    p.n_threads = 0;
    p.max_reference_seconds = 20.0f;
    p.wavlm_flash_attn_type = LLAMA_FLASH_ATTN_TYPE_AUTO;
    p.miocodec_flash_attn_type = LLAMA_FLASH_ATTN_TYPE_AUTO;
    return p;
}

struct mio_tts_context * mio_tts_init_from_file(const char * model_vocoder, const char * wavlm_model, char * err,
                                                size_t err_size) {
    if (model_vocoder == nullptr || model_vocoder[0] == '\0') {
        set_err(err, err_size, "model_vocoder path is empty");
        return nullptr;
    }
    mio_tts_context * ctx = nullptr;
    const bool ok = guarded(err, err_size, [&] {
        std::string load_err;
        auto m = load_model(model_vocoder, default_device(), load_err);
        if (!m) throw std::runtime_error("failed to load MioCodec GGUF: " + load_err);
        ctx = new mio_tts_context();
        ctx->mdl = m;
        ctx->eng.reset(new engine(m));
        ctx->vocoder_path = model_vocoder;
        if (wavlm_model != nullptr && wavlm_model[0] != '\0') ctx->wavlm_path = wavlm_model;
    });
    if (!ok) { delete ctx; return nullptr; }
    return ctx;
}

struct mio_tts_context * mio_tts_context_clone(const struct mio_tts_context * ctx, char * err, size_t err_size) {
    if (ctx == nullptr || !ctx->mdl) {
        set_err(err, err_size, "source context is null");
        return nullptr;
    }
    mio_tts_context * out = nullptr;
    const bool ok = guarded(err, err_size, [&] {
        out = new mio_tts_context();
        out->mdl = ctx->mdl;
        out->eng.reset(new engine(ctx->mdl));
        out->vocoder_path = ctx->vocoder_path;
        out->wavlm_path = ctx->wavlm_path;
    });
    if (!ok) { delete out; return nullptr; }
    return out;
}

void mio_tts_free(struct mio_tts_context * ctx) { delete ctx; }

bool mio_tts_context_is_dynamic_global(const struct mio_tts_context * ctx) {
    return ctx != nullptr && ctx->mdl && ctx->mdl->hp.dynamic_global;
}

int32_t mio_tts_context_sample_rate(const struct mio_tts_context * ctx) {
    return (ctx == nullptr || !ctx->mdl) ? 0 : ctx->mdl->hp.sample_rate;
}

int32_t mio_tts_context_samples_per_token(const struct mio_tts_context * ctx) {
    return (ctx == nullptr || !ctx->mdl) ? 0 : ctx->mdl->hp.samples_per_token;
}

bool mio_tts_context_set_backend_device(struct mio_tts_context * ctx, const char * backend_name, char * err,
                                        size_t err_size) {
    if (ctx == nullptr || !ctx->mdl) {
        set_err(err, err_size, "context is null");
        return false;
    }
    const std::string name = backend_name != nullptr ? backend_name : "";
    const int dev = parse_device(name);
    if (dev == -3) {
        set_err(err, err_size, "decoder backend setup failed: the B200 decode library has no CPU fallback: " + name);
        return false;
    }
    if (dev == -2) {
        set_err(err, err_size, "decoder backend setup failed: failed to initialize backend device: " + name);
        return false;
    }
    if (dev == -1 || dev == ctx->mdl->device) return true;
    return guarded(err, err_size, [&] {
        std::lock_guard<std::mutex> lk(ctx->eng->mtx);
        std::string load_err;
        auto m = load_model(ctx->vocoder_path, dev, load_err);
        if (!m) throw std::runtime_error("decoder backend setup failed: " + load_err);
        ctx->eng.reset();
        ctx->mdl = m;
        ctx->eng.reset(new engine(m));
    });
}

struct mio_tts_vocab_map * mio_tts_vocab_map_init(const struct llama_vocab * vocab, char * err, size_t err_size) {
    if (vocab == nullptr) {
        set_err(err, err_size, "vocab is null");
        return nullptr;
    }
    auto t2p = (llama_token_to_piece_fn) dlsym(RTLD_DEFAULT, "llama_token_to_piece");
    auto tok = (llama_tokenize_fn) dlsym(RTLD_DEFAULT, "llama_tokenize");
    if (t2p == nullptr || tok == nullptr) {
        set_err(err, err_size, "libllama tokenizer (llama_tokenize, llama_token_to_piece) is not loaded in this process");
        return nullptr;
    }
    // build_mio_audio_token_map, lib.cpp:179-223
    const auto first = single_token_id(tok, t2p, vocab, "<|s_0|>");
    const auto last = single_token_id(tok, t2p, vocab, "<|s_12799|>");
    if (!first.has_value() || !last.has_value()) {
        set_err(err, err_size, "MioTTS audio token range not found (<|s_0|>..<|s_12799|>)");
        return nullptr;
    }
    if (*first > *last) {
        set_err(err, err_size, "invalid MioTTS token range");
        return nullptr;
    }
    auto * map = new mio_tts_vocab_map();
    for (llama_token id = *first; id <= *last; ++id) {
        std::string piece;
        int32_t code = -1;
        if (!token_to_piece(t2p, vocab, id, piece) || !parse_codec_piece(piece, code)) continue;
        if (code < k_code_min || code > k_code_max) continue;
        map->token_to_code[id] = code;
    }
    if (map->token_to_code.empty()) {
        set_err(err, err_size, "failed to build MioTTS token->code map");
        delete map;
        return nullptr;
    }
    return map;
}

void mio_tts_vocab_map_free(struct mio_tts_vocab_map * map) { delete map; }

bool mio_tts_tokens_to_codes(const struct mio_tts_vocab_map * map, const llama_token * tokens, size_t n_tokens,
                             int32_t ** codes_out, size_t * n_codes_out, char * err, size_t err_size) {
    if (map == nullptr) { set_err(err, err_size, "vocab map is null"); return false; }
    if (tokens == nullptr || n_tokens == 0) { set_err(err, err_size, "input tokens are empty"); return false; }
    if (codes_out == nullptr || n_codes_out == nullptr) { set_err(err, err_size, "output pointers are null"); return false; }
    std::vector<int32_t> codes;
    codes.reserve(n_tokens);
    for (size_t i = 0; i < n_tokens; ++i) {
        auto it = map->token_to_code.find(tokens[i]);
        if (it != map->token_to_code.end()) codes.push_back(it->second);
    }
    if (codes.empty()) { set_err(err, err_size, "no Mio audio codes were found in token sequence"); return false; }
    int32_t * out = (int32_t *) std::malloc(codes.size() * sizeof(int32_t));
    if (out == nullptr) { set_err(err, err_size, "failed to allocate output codes"); return false; }
    std::memcpy(out, codes.data(), codes.size() * sizeof(int32_t));
    *codes_out = out;
    *n_codes_out = codes.size();
    return true;
}

bool mio_tts_token_to_code(const struct mio_tts_vocab_map * map, llama_token token, int32_t * code_out) {
    if (map == nullptr || code_out == nullptr) return false;
    auto it = map->token_to_code.find(token);
    if (it == map->token_to_code.end()) return false;
    *code_out = it->second;
    return true;
}

bool mio_tts_codes_load(const char * path, int32_t ** codes_out, size_t * n_codes_out, char * err, size_t err_size) {
    if (codes_out == nullptr || n_codes_out == nullptr) { set_err(err, err_size, "output pointers are null"); return false; }
    if (path == nullptr || path[0] == '\0') { set_err(err, err_size, "codes input path is empty"); return false; }
    std::ifstream file(path);
    if (!file) { set_err(err, err_size, "failed to open codes input file"); return false; }
    std::vector<int32_t> codes;
    std::string tok;
    while (file >> tok) {
        int32_t code = 0;
        if (!parse_code_token(tok, code)) { set_err(err, err_size, "failed to parse code token: " + tok); return false; }
        if (code < k_code_min || code > k_code_max) { set_err(err, err_size, "code id out of range in input file"); return false; }
        codes.push_back(code);
    }
    if (codes.empty()) { set_err(err, err_size, "codes input file is empty"); return false; }
    int32_t * out = (int32_t *) std::malloc(codes.size() * sizeof(int32_t));
    if (out == nullptr) { set_err(err, err_size, "failed to allocate output codes"); return false; }
    std::memcpy(out, codes.data(), codes.size() * sizeof(int32_t));
    *codes_out = out;
    *n_codes_out = codes.size();
    return true;
}

bool mio_tts_codes_save(const char * path, const int32_t * codes, size_t n_codes, char * err, size_t err_size) {
    if (path == nullptr || path[0] == '\0') { set_err(err, err_size, "codes output path is empty"); return false; }
    if (codes == nullptr || n_codes == 0) { set_err(err, err_size, "codes are empty"); return false; }
    std::ofstream file(path, std::ios::trunc);
    if (!file) { set_err(err, err_size, "failed to open codes output file"); return false; }
    for (size_t i = 0; i < n_codes; ++i) file << codes[i] << "\n";
    if (!file.good()) { set_err(err, err_size, "failed while writing codes output file"); return false; }
    return true;
}

void mio_tts_codes_free(int32_t * codes) { std::free(codes); }

bool mio_tts_reference_to_embedding(struct mio_tts_context * ctx, const char * reference_audio, struct mio_tts_params,
                                    float ** embedding_out, size_t * n_embedding_out, char * err, size_t err_size) {
    if (ctx == nullptr || !ctx->mdl) { set_err(err, err_size, "context is null"); return false; }
    if (reference_audio == nullptr || reference_audio[0] == '\0') { set_err(err, err_size, "reference_audio path is empty"); return false; }
    if (embedding_out == nullptr || n_embedding_out == nullptr) { set_err(err, err_size, "output pointers are null"); return false; }
    if (!ctx->mdl->hp.dynamic_global) { set_err(err, err_size, "reference embedding requires dynamic-global MioCodec"); return false; }
    // WavLM + global encoder are not part of the decode path this library accelerates
    set_err(err, err_size, "WavLM model is not loaded");
    return false;
}

bool mio_tts_embedding_load_gguf(const char * path, float ** embedding_out, size_t * n_embedding_out, char * err,
                                 size_t err_size) {
    if (embedding_out == nullptr || n_embedding_out == nullptr) { set_err(err, err_size, "output pointers are null"); return false; }
    if (path == nullptr || path[0] == '\0') { set_err(err, err_size, "embedding input path is empty"); return false; }
    gguf_file g;
    std::string e;
    if (!g.load(path, e)) { set_err(err, err_size, "failed to open embedding GGUF"); return false; }
    const gguf_tensor * t = g.find(k_embedding_tensor);
    if (t == nullptr) {
        if (g.tensors().size() == 1) t = &g.tensors()[0];
        else { set_err(err, err_size, "embedding GGUF is missing tensor 'mio.global_embedding'"); return false; }
    }
    if (t->data == nullptr) { set_err(err, err_size, "failed to load embedding tensor data from GGUF"); return false; }
    if (t->type != GGUF_F32) { set_err(err, err_size, "embedding tensor must be f32"); return false; }
    const int64_t n = t->nelements();
    if (n <= 0) { set_err(err, err_size, "embedding tensor is empty"); return false; }
    float * out = (float *) std::malloc((size_t) n * sizeof(float));
    if (out == nullptr) { set_err(err, err_size, "failed to allocate output embedding"); return false; }
    std::memcpy(out, t->data, (size_t) n * sizeof(float));
    *embedding_out = out;
    *n_embedding_out = (size_t) n;
    return true;
}

bool mio_tts_embedding_save_gguf(const char * path, const float * embedding, size_t n_embedding, char * err,
                                 size_t err_size) {
    if (embedding == nullptr || n_embedding == 0) { set_err(err, err_size, "embedding is empty"); return false; }
    if (path == nullptr || path[0] == '\0') { set_err(err, err_size, "embedding output path is empty"); return false; }
    std::string e;
    // KVs written by save_embedding_gguf_vec, lib.cpp:268-270
    if (!gguf_write_f32_vector(path, {{"general.architecture", "mio-embedding"}, {"general.type", "embedding"}},
                               {{"mio.embedding.dim", (uint32_t) n_embedding}}, k_embedding_tensor, embedding, n_embedding, e)) {
        set_err(err, err_size, e);
        return false;
    }
    return true;
}

void mio_tts_embedding_free(float * embedding) { std::free(embedding); }

bool mio_tts_b200_synthesize_batch(struct mio_tts_context * ctx, const int32_t * const * codes, const size_t * n_codes,
                                   const float * const * embedding, size_t n_embedding, size_t batch,
                                   struct mio_tts_params, float ** audio_out, size_t * n_audio_out,
                                   int32_t * sample_rate_out, char * err, size_t err_size) {
    if (ctx == nullptr || !ctx->mdl) { set_err(err, err_size, "context is null"); return false; }
    if (codes == nullptr || n_codes == nullptr || batch == 0) { set_err(err, err_size, "codes are empty"); return false; }
    for (size_t u = 0; u < batch; ++u)
        if (codes[u] == nullptr || n_codes[u] == 0) { set_err(err, err_size, "codes are empty"); return false; }
    if (audio_out == nullptr || n_audio_out == nullptr || sample_rate_out == nullptr) {
        set_err(err, err_size, "output pointers are null");
        return false;
    }
    std::vector<float *> bufs(batch, nullptr);
    const bool ok = guarded(err, err_size, [&] {
        std::lock_guard<std::mutex> lk(ctx->eng->mtx);
        engine & E = *ctx->eng;
        E.prepare(codes, n_codes, embedding, n_embedding, batch);
        E.run(true);
        for (size_t u = 0; u < batch; ++u) {
            bufs[u] = result_alloc(E.audio_len(u));
            if (bufs[u] == nullptr) throw std::runtime_error("failed to allocate output audio");
            E.fetch_audio(u, bufs[u]);
        }
        E.sync();
        for (size_t u = 0; u < batch; ++u) { audio_out[u] = bufs[u]; n_audio_out[u] = E.audio_len(u); }
        *sample_rate_out = ctx->mdl->hp.sample_rate;
    });
    if (!ok) {
        try { ctx->eng->sync(); } catch (...) {}
        for (float * p : bufs) if (p && !result_free(p)) std::free(p);
    }
    return ok;
}

bool mio_tts_synthesize(struct mio_tts_context * ctx, const int32_t * codes, size_t n_codes, const float * embedding,
                        size_t n_embedding, struct mio_tts_params params, float ** audio_out, size_t * n_audio_out,
                        int32_t * sample_rate_out, char * err, size_t err_size) {
    if (ctx == nullptr || !ctx->mdl) { set_err(err, err_size, "context is null"); return false; }
    if (codes == nullptr || n_codes == 0) { set_err(err, err_size, "codes are empty"); return false; }
    if (audio_out == nullptr || n_audio_out == nullptr || sample_rate_out == nullptr) {
        set_err(err, err_size, "output pointers are null");
        return false;
    }
    const bool has_emb = embedding != nullptr && n_embedding > 0;
    const float * emb_arr[1] = {embedding};
    return mio_tts_b200_synthesize_batch(ctx, &codes, &n_codes, has_emb ? emb_arr : nullptr, has_emb ? n_embedding : 0, 1,
                                         params, audio_out, n_audio_out, sample_rate_out, err, err_size);
}

bool mio_tts_reserve_workspace(struct mio_tts_context * ctx, int32_t n_codes, char * err, size_t err_size) {
    if (ctx == nullptr || !ctx->mdl) { set_err(err, err_size, "context is null"); return false; }
    const int n = std::max<int32_t>(1, n_codes);
    if (!guarded(err, err_size, [&] {
            std::lock_guard<std::mutex> lk(ctx->eng->mtx);
            ctx->eng->reserve(n, 1);
        })) {
        set_err(err, err_size, "failed to reserve decode workspace");
        return false;
    }
    return true;
}

bool mio_tts_estimate_workspace_bytes(struct mio_tts_context * ctx, int32_t n_codes, uint64_t * out_bytes, char * err,
                                      size_t err_size) {
    if (ctx == nullptr || !ctx->mdl) { set_err(err, err_size, "context is null"); return false; }
    if (out_bytes == nullptr) { set_err(err, err_size, "out_bytes is null"); return false; }
    // Closed form of what engine::prepare allocates for one utterance of n codes (band attention: no N^2 term,
    // unlike estimate_decode_workspace_bytes, /root/reference/src/miocodec-decoder.cpp:238-282).
    const hparams & hp = ctx->mdl->hp;
    const uint64_t T = (uint64_t) std::max<int32_t>(1, n_codes);
    const uint64_t stft = std::max<uint64_t>(1, T * hp.samples_per_token / std::max(1, hp.hop_length));
    const uint64_t F = std::max<uint64_t>(1, ctx->mdl->up_total > 1 ? stft / ctx->mdl->up_total : stft);
    auto pr = [](uint64_t r) { return (r + 127) / 128 * 128 + 128; };
    const uint64_t RT = pr(T), RF = pr(F + 2), RS = pr(stft + 2);
    uint64_t f32e = std::max(RF * hp.decoder_dim, RT * 2 * hp.decoder_dim);
    uint64_t a16e = std::max(RT * hp.prenet_dim, RF * hp.decoder_dim);
    uint64_t rows = RF;
    for (const upstage & s : ctx->mdl->ups) {
        const uint64_t ro = pr(rows * s.factor);
        f32e = std::max(f32e, std::max(rows * s.ksize * s.cout, ro * s.cout));
        a16e = std::max(a16e, std::max(rows * s.cin, ro * s.cout));
        rows = ro;
    }
    a16e = std::max(a16e, rows * ctx->mdl->head_in);
    uint64_t b = RT * hp.prenet_dim * 4 + 3 * f32e * 4 + a16e * 2;
    b += std::max(RT * 3 * hp.prenet_dim, RF * 3 * hp.decoder_dim) * 2;
    b += std::max(RT * hp.prenet_dim, RF * hp.decoder_dim) * 2;
    b += std::max(RT * hp.prenet_ff, RF * hp.decoder_ff) * 2;
    b += RS * (uint64_t) (hp.n_fft + 4) * 4;
    b += stft * hp.hop_length * 4 + (T + 2 * F + 2 * stft) * 16;
    *out_bytes = b + b / 8 + (16ull << 20);
    return true;
}

bool mio_tts_reserve_reference_workspace(struct mio_tts_context * ctx, float, char * err, size_t err_size) {
    if (ctx == nullptr || !ctx->mdl) { set_err(err, err_size, "context is null"); return false; }
    set_err(err, err_size, "WavLM model is not loaded");
    return false;
}

bool mio_tts_estimate_reference_workspace_bytes(struct mio_tts_context * ctx, float, uint64_t * out_bytes, char * err,
                                                size_t err_size) {
    if (ctx == nullptr || !ctx->mdl) { set_err(err, err_size, "context is null"); return false; }
    if (out_bytes == nullptr) { set_err(err, err_size, "out_bytes is null"); return false; }
    set_err(err, err_size, "WavLM model is not loaded");
    return false;
}

bool mio_tts_audio_save_wav16(const char * path, const float * audio, size_t n_audio, int32_t sample_rate, char * err,
                              size_t err_size) {
    if (audio == nullptr || n_audio == 0) { set_err(err, err_size, "audio is empty"); return false; }
    if (path == nullptr || path[0] == '\0') { set_err(err, err_size, "output wav path is empty"); return false; }
    std::ofstream file(path, std::ios::binary);
    if (!file) { set_err(err, err_size, "failed to open output wav file"); return false; }
    // 44-byte PCM header, lib.cpp:49-63, 770-777
    const uint32_t data_size = (uint32_t) (n_audio * 2);
    uint8_t h[44];
    auto u32 = [&](int o, uint32_t v) { std::memcpy(h + o, &v, 4); };
    auto u16 = [&](int o, uint16_t v) { std::memcpy(h + o, &v, 2); };
    std::memcpy(h, "RIFF", 4); u32(4, 36 + data_size); std::memcpy(h + 8, "WAVE", 4); std::memcpy(h + 12, "fmt ", 4);
    u32(16, 16); u16(20, 1); u16(22, 1); u32(24, (uint32_t) sample_rate); u32(28, (uint32_t) sample_rate * 2);
    u16(32, 2); u16(34, 16); std::memcpy(h + 36, "data", 4); u32(40, data_size);
    file.write((const char *) h, 44);
    std::vector<int16_t> pcm(n_audio);
    for (size_t i = 0; i < n_audio; ++i) {
        const float x = std::min(1.0f, std::max(-1.0f, audio[i]));
        pcm[i] = (int16_t) std::lrintf(x * 32767.0f);
    }
    file.write((const char *) pcm.data(), (std::streamsize) (pcm.size() * 2));
    if (!file.good()) { set_err(err, err_size, "failed while writing output wav file"); return false; }
    return true;
}

void mio_tts_audio_free(float * audio) {
    if (audio == nullptr) return;
    if (!result_free(audio)) std::free(audio);
}

// ---------------------------------------------------------------------------------------------------
// extensions
// ---------------------------------------------------------------------------------------------------
bool mio_tts_b200_decode_spec(struct mio_tts_context * ctx, const int32_t * codes, size_t n_codes, const float * embedding,
                              size_t n_embedding, float ** spec_out, int32_t * n_frames_out, int32_t * n_bins_out,
                              char * err, size_t err_size) {
    if (ctx == nullptr || !ctx->mdl) { set_err(err, err_size, "context is null"); return false; }
    if (codes == nullptr || n_codes == 0) { set_err(err, err_size, "empty token sequence"); return false; }
    if (spec_out == nullptr || n_frames_out == nullptr || n_bins_out == nullptr) { set_err(err, err_size, "output pointers are null"); return false; }
    float * out = nullptr;
    const bool ok = guarded(err, err_size, [&] {
        std::lock_guard<std::mutex> lk(ctx->eng->mtx);
        engine & E = *ctx->eng;
        const bool has_emb = embedding != nullptr && n_embedding > 0;
        const float * emb_arr[1] = {embedding};
        E.prepare(&codes, &n_codes, has_emb ? emb_arr : nullptr, has_emb ? n_embedding : 0, 1);
        E.run(false);
        const size_t n = (size_t) E.spec_frames(0) * E.spec_bins();
        out = (float *) std::malloc(std::max<size_t>(n, 1) * sizeof(float));
        if (out == nullptr) throw std::runtime_error("failed to allocate output spectrogram");
        E.fetch_spec(0, out);
        E.sync();
        *spec_out = out;
        *n_frames_out = E.spec_frames(0);
        *n_bins_out = E.spec_bins();
    });
    if (!ok) std::free(out);
    return ok;
}

bool mio_tts_b200_debug_stage(struct mio_tts_context * ctx, const int32_t * codes, size_t n_codes, const float * embedding,
                              size_t n_embedding, int32_t stage, float ** out, int32_t * rows_out, int32_t * cols_out,
                              char * err, size_t err_size) {
    if (ctx == nullptr || !ctx->mdl) { set_err(err, err_size, "context is null"); return false; }
    if (out == nullptr || rows_out == nullptr || cols_out == nullptr) { set_err(err, err_size, "output pointers are null"); return false; }
    return guarded(err, err_size, [&] {
        std::lock_guard<std::mutex> lk(ctx->eng->mtx);
        engine & E = *ctx->eng;
        const bool has_emb = embedding != nullptr && n_embedding > 0;
        const float * emb_arr[1] = {embedding};
        E.prepare(&codes, &n_codes, has_emb ? emb_arr : nullptr, has_emb ? n_embedding : 0, 1);
        E.run(false, nullptr, stage);
        std::vector<float> v;
        int r = 0, c = 0;
        E.fetch_debug(v, r, c);
        float * o = (float *) std::malloc(std::max<size_t>(v.size(), 1) * sizeof(float));
        if (o == nullptr) throw std::runtime_error("failed to allocate debug output");
        std::memcpy(o, v.data(), v.size() * sizeof(float));
        *out = o; *rows_out = r; *cols_out = c;
    });
}

bool mio_tts_b200_gather(struct mio_tts_context * ctx, const int32_t * codes, size_t n_codes, float * out, char * err,
                         size_t err_size) {
    if (ctx == nullptr || !ctx->mdl) { set_err(err, err_size, "context is null"); return false; }
    if (out == nullptr) { set_err(err, err_size, "output pointers are null"); return false; }
    return guarded(err, err_size, [&] {
        std::lock_guard<std::mutex> lk(ctx->eng->mtx);
        ctx->eng->gather(codes, n_codes, out);
    });
}

int32_t mio_tts_b200_prenet_dim(const struct mio_tts_context * ctx) {
    return (ctx == nullptr || !ctx->mdl) ? 0 : ctx->mdl->hp.prenet_dim;
}

bool mio_tts_b200_istft(struct mio_tts_context * ctx, const float * spec, int32_t n_frames, int32_t n_fft, int32_t hop,
                        float ** audio_out, size_t * n_audio_out, char * err, size_t err_size) {
    if (ctx == nullptr || !ctx->mdl) { set_err(err, err_size, "context is null"); return false; }
    if (audio_out == nullptr || n_audio_out == nullptr) { set_err(err, err_size, "output pointers are null"); return false; }
    return guarded(err, err_size, [&] {
        std::lock_guard<std::mutex> lk(ctx->eng->mtx);
        std::vector<float> a;
        ctx->eng->istft_host(spec, n_frames, n_fft, hop, a);
        float * out = (float *) std::malloc(std::max<size_t>(a.size(), 1) * sizeof(float));
        if (out == nullptr) throw std::runtime_error("failed to allocate output audio");
        if (!a.empty()) std::memcpy(out, a.data(), a.size() * sizeof(float));
        *audio_out = out;
        *n_audio_out = a.size();
    });
}

void mio_tts_b200_free(void * p) { std::free(p); }

void mio_tts_b200_set_generic_attention(struct mio_tts_context * ctx, int32_t enable) {
    if (ctx != nullptr && ctx->eng) ctx->eng->force_generic_attention = enable != 0;
}

bool mio_tts_b200_prepare(struct mio_tts_context * ctx, const int32_t * const * codes, const size_t * n_codes,
                          const float * const * embedding, size_t n_embedding, size_t batch, char * err, size_t err_size) {
    if (ctx == nullptr || !ctx->mdl) { set_err(err, err_size, "context is null"); return false; }
    return guarded(err, err_size, [&] {
        std::lock_guard<std::mutex> lk(ctx->eng->mtx);
        ctx->eng->prepare(codes, n_codes, embedding, n_embedding, batch);
        ctx->eng->sync();
    });
}

float mio_tts_b200_run_resident(struct mio_tts_context * ctx, int32_t iters, char * err, size_t err_size) {
    if (ctx == nullptr || !ctx->mdl) { set_err(err, err_size, "context is null"); return -1.0f; }
    float ms = -1.0f;
    guarded(err, err_size, [&] {
        std::lock_guard<std::mutex> lk(ctx->eng->mtx);
        engine & E = *ctx->eng;
        if (E.batch() == 0) throw std::runtime_error("no prepared batch");
        cudaEvent_t a, b;
        MIO_CUDA(cudaEventCreate(&a));
        MIO_CUDA(cudaEventCreate(&b));
        MIO_CUDA(cudaEventRecord(a, E.stream()));
        for (int i = 0; i < std::max(1, iters); ++i) E.run(true);
        MIO_CUDA(cudaEventRecord(b, E.stream()));
        E.sync();
        float t = 0;
        MIO_CUDA(cudaEventElapsedTime(&t, a, b));
        cudaEventDestroy(a);
        cudaEventDestroy(b);
        ms = t / (float) std::max(1, iters);
    });
    return ms;
}

bool mio_tts_b200_profile_resident(struct mio_tts_context * ctx, struct mio_tts_b200_profile * out, char * err,
                                   size_t err_size) {
    if (ctx == nullptr || !ctx->mdl) { set_err(err, err_size, "context is null"); return false; }
    if (out == nullptr) { set_err(err, err_size, "output pointers are null"); return false; }
    return guarded(err, err_size, [&] {
        std::lock_guard<std::mutex> lk(ctx->eng->mtx);
        engine & E = *ctx->eng;
        if (E.batch() == 0) throw std::runtime_error("no prepared batch");
        stage_times t;
        E.run(true, &t);
        out->prenet_ms = t.prenet; out->upsample_prior_ms = t.upsample_prior; out->decoder_ms = t.decoder;
        out->post_head_ms = t.post_head; out->istft_ms = t.istft; out->total_ms = t.total;
        out->gemm_ms = t.gemm_ms; out->gemm_flops = t.gemm_flops; out->gemm_launches = t.gemm_launches;
        out->attn_ms = t.attn_ms; out->attn_launches = t.attn_launches;
    });
}

uint64_t mio_tts_b200_resident_samples(struct mio_tts_context * ctx) {
    if (ctx == nullptr || !ctx->eng) return 0;
    uint64_t n = 0;
    for (size_t u = 0; u < ctx->eng->batch(); ++u) n += ctx->eng->audio_len(u);
    return n;
}

float mio_tts_b200_istft_bench(struct mio_tts_context * ctx, int32_t n_utt, int32_t frames_per_utt, int32_t n_fft,
                               int32_t hop, int32_t iters, uint64_t seed, double * checksum, char * err, size_t err_size) {
    if (ctx == nullptr || !ctx->mdl) { set_err(err, err_size, "context is null"); return -1.0f; }
    float ms = -1.0f;
    guarded(err, err_size, [&] {
        std::lock_guard<std::mutex> lk(ctx->eng->mtx);
        ms = ctx->eng->istft_bench(n_utt, frames_per_utt, n_fft, hop, iters, seed, checksum);
    });
    return ms;
}

uint64_t mio_tts_b200_launch_count(void) { return (uint64_t) g_launch_count.load(); }

uint64_t mio_tts_b200_workspace_bytes(struct mio_tts_context * ctx) {
    return (ctx == nullptr || !ctx->eng) ? 0 : (uint64_t) ctx->eng->workspace_bytes();
}

uint64_t mio_tts_b200_weight_bytes(struct mio_tts_context * ctx) {
    return (ctx == nullptr || !ctx->mdl) ? 0 : (uint64_t) ctx->mdl->weight_bytes;
}

uint64_t mio_tts_b200_last_h2d_bytes(struct mio_tts_context * ctx) {
    return (ctx == nullptr || !ctx->eng) ? 0 : (uint64_t) ctx->eng->h2d_bytes();
}

} // extern "C"
