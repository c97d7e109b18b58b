// C ABI of libmio_tts_b200.so (include/mio_tts_b200.h).  Host-side behaviour -- argument checks, their order,
// error strings, malloc ownership -- follows /root/reference/src/mio-tts-lib.cpp function by function
// (cited below as lib.cpp:LINE); the decode itself runs on the GPU through mio::engine, with no CPU fallback.
#include "../../include/mio_tts_b200.h"

#include "batcher.h"
#include "engine.h"

#include <atomic>
#include <chrono>
#include "gguf_io.h"
#include "refenc.h"

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
    std::unique_ptr<engine> eng2;       // second stream + workspace for large single-caller batches (created on first use)
    std::mutex eng2_mtx;
    uint64_t last_h2d = 0;              // bytes uploaded by the last synthesize call on this context (both halves)
    std::shared_ptr<refmodel> ref;      // WavLM + global-encoder weights (reference-audio path), shared by clones
    std::unique_ptr<refenc> renc;       // per-context stream + workspace of that path
    std::shared_ptr<batcher> bat;       // micro-batcher shared with the clones (null: calls decode on this context's engine)
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

// MIO_TTS_B200_BATCHING="max_batch[,max_wait_us[,workers]]": contexts created by mio_tts_init_from_file start with the
// micro-batcher on, so an unmodified server gets coalesced slots (batcher.h)
bool env_batching(int out[3]) {
    const char * e = std::getenv("MIO_TTS_B200_BATCHING");
    if (e == nullptr || e[0] == '\0' || (e[0] == '0' && e[1] == '\0')) return false;
    out[0] = 64; out[1] = 300; out[2] = 4;
    std::sscanf(e, "%d,%d,%d", &out[0], &out[1], &out[2]);
    return out[0] > 0;
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

// WavLM GGUF (optional) + the global-encoder tensors of the MioCodec GGUF (optional), lib.cpp:833-840
std::shared_ptr<refmodel> load_refmodel(const std::string & vocoder_path, const std::string & wavlm_path, int device, int out_dim,
                                        std::string & err) {
    auto r = std::make_shared<refmodel>();
    {
        gguf_file g;
        std::string e;
        if (!g.load(vocoder_path, e)) { err = "failed to load MioCodec GGUF: " + e; return nullptr; }
        if (!r->load_global_encoder(g, device, out_dim, e)) { err = "failed to load MioCodec GGUF: " + e; return nullptr; }
    }
    if (!wavlm_path.empty()) {
        std::string e;
        if (!r->load_wavlm(wavlm_path, device, e)) { err = "failed to load WavLM GGUF: " + e; return nullptr; }
    }
    return r;
}

// PCM WAV file -> mono float32 (stereo averaged) + sample rate.  The reference decodes through miniaudio
// (/root/reference/src/wavlm-extractor.cpp:153-203), which also reads MP3/FLAC and resamples to the codec rate; this
// library reads RIFF/WAVE PCM 16/24/32-bit and float32 and resamples other rates linearly.
bool read_wav_mono(const std::string & path, std::vector<float> & out, int & sample_rate, std::string & err) {
    std::ifstream f(path, std::ios::binary);
    if (!f) { err = "ma_decoder_init_file failed"; return false; }
    std::vector<uint8_t> d((std::istreambuf_iterator<char>(f)), std::istreambuf_iterator<char>());
    if (d.size() < 12 || std::memcmp(d.data(), "RIFF", 4) != 0 || std::memcmp(d.data() + 8, "WAVE", 4) != 0) {
        err = "ma_decoder_init_file failed (only RIFF/WAVE PCM is supported by the B200 library)";
        return false;
    }
    auto u16 = [&](size_t o) { return (uint32_t) d[o] | ((uint32_t) d[o + 1] << 8); };
    auto u32 = [&](size_t o) { return u16(o) | (u16(o + 2) << 16); };
    size_t pos = 12, data_off = 0, data_len = 0;
    uint32_t tag = 0, ch = 0, sr = 0, bits = 0;
    while (pos + 8 <= d.size()) {
        const uint32_t size = u32(pos + 4);
        const size_t body = pos + 8;
        if (std::memcmp(d.data() + pos, "fmt ", 4) == 0 && body + 16 <= d.size()) {
            tag = u16(body); ch = u16(body + 2); sr = u32(body + 4); bits = u16(body + 14);
            if (tag == 0xFFFE && size >= 26 && body + 26 <= d.size()) tag = u16(body + 24);   // WAVE_FORMAT_EXTENSIBLE
        } else if (std::memcmp(d.data() + pos, "data", 4) == 0) {
            data_off = body;
            data_len = std::min<size_t>(size, d.size() - body);
        }
        pos = body + size + (size & 1);
    }
    if (ch == 0 || sr == 0 || data_off == 0) { err = "ma_decoder_init_file failed (no fmt/data chunk)"; return false; }
    const size_t bps = bits / 8;
    if (!((tag == 1 && (bits == 16 || bits == 24 || bits == 32)) || (tag == 3 && bits == 32))) {
        err = "ma_decoder_init_file failed (unsupported WAV sample format)";
        return false;
    }
    const size_t frames = data_len / (bps * ch);
    if (frames == 0) { err = "ma_decoder_get_length_in_pcm_frames failed"; return false; }
    out.resize(frames);
    const uint8_t * p = d.data() + data_off;
    for (size_t i = 0; i < frames; ++i) {
        float acc = 0.0f;
        for (uint32_t c = 0; c < ch; ++c) {
            const uint8_t * q = p + (i * ch + c) * bps;
            float v;
            if (tag == 3) { std::memcpy(&v, q, 4); }
            else if (bits == 16) { v = (float) (int16_t) (q[0] | (q[1] << 8)) / 32768.0f; }
            else if (bits == 24) { v = (float) ((int32_t) ((uint32_t) q[0] << 8 | (uint32_t) q[1] << 16 | (uint32_t) q[2] << 24) >> 8) / 8388608.0f; }
            else { int32_t t; std::memcpy(&t, q, 4); v = (float) ((double) t / 2147483648.0); }
            acc += v;
        }
        out[i] = ch > 1 ? acc / (float) ch : acc;
    }
    sample_rate = (int) sr;
    return true;
}

// ---- text codes ----
// Spec (behaviour, not code): /root/reference/src/mio-tts-lib.cpp:79-113 (a code written as text: optional ASCII
// punctuation around it, optional <|s_N|> wrapper, then what std::stoll accepts in full, within int32) and :115-138
// (a vocabulary piece: exactly <|s_N|>, N what strtol accepts up to the first NUL, 0 <= N <= INT32_MAX).
// One pointer scanner serves both: [space]* [+-]? digit+ to the end of the range, no int64 overflow.
bool scan_decimal(const char * p, const char * end, long long & value) {
    while (p < end && (*p == ' ' || (*p >= '\t' && *p <= '\r'))) ++p;
    bool neg = false;
    if (p < end && (*p == '+' || *p == '-')) neg = *p++ == '-';
    if (p == end) return false;
    const unsigned long long lim = neg ? 9223372036854775808ull : 9223372036854775807ull;
    unsigned long long mag = 0;
    for (; p < end; ++p) {
        const unsigned d = (unsigned) (unsigned char) *p - (unsigned) '0';
        if (d > 9 || mag > (lim - d) / 10) return false;
        mag = mag * 10 + d;
    }
    value = neg ? (long long) (0ull - mag) : (long long) mag;
    return true;
}

bool parse_code_token(const std::string & raw, int32_t & out) {
    auto punct = [](unsigned char c) {      // "C" locale ispunct
        return (c >= 33 && c <= 47) || (c >= 58 && c <= 64) || (c >= 91 && c <= 96) || (c >= 123 && c <= 126);
    };
    const char * b = raw.data(), * e = b + raw.size();
    while (b < e && punct((unsigned char) *b) && *b != '<' && *b != '-') ++b;
    while (e > b && punct((unsigned char) e[-1]) && e[-1] != '>') --e;
    if (b == e) return false;
    if (e - b > 6 && std::memcmp(b, "<|s_", 4) == 0 && e[-2] == '|' && e[-1] == '>') { b += 4; e -= 2; }
    long long v = 0;
    if (!scan_decimal(b, e, v) || v < INT32_MIN || v > INT32_MAX) return false;
    out = (int32_t) v;
    return true;
}

bool parse_codec_piece(const std::string & piece, int32_t & code) {
    const size_t n = piece.size();
    if (n < 7 || std::memcmp(piece.data(), "<|s_", 4) != 0 || piece[n - 2] != '|' || piece[n - 1] != '>') return false;
    const char * b = piece.data() + 4;
    const char * e = (const char *) std::memchr(b, 0, n - 6);
    if (e == nullptr) e = piece.data() + n - 2;
    long long v = 0;
    if (!scan_decimal(b, e, v) || v < 0 || v > INT32_MAX) return false;
    code = (int32_t) v;
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
        const std::string wavlm = (wavlm_model != nullptr && wavlm_model[0] != '\0') ? wavlm_model : "";
        auto r = load_refmodel(model_vocoder, wavlm, m->device, m->hp.decoder_adanorm_dim, load_err);
        if (!r) throw std::runtime_error(load_err);
        ctx = new mio_tts_context();
        ctx->mdl = m;
        ctx->eng.reset(new engine(m));
        ctx->ref = r;
        ctx->renc.reset(new refenc(r));
        ctx->vocoder_path = model_vocoder;
        ctx->wavlm_path = wavlm;
        int bp[3];
        if (env_batching(bp)) ctx->bat = std::make_shared<batcher>(m, bp[0], bp[1], bp[2]);
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
        out->ref = ctx->ref;
        out->bat = ctx->bat;     // the slots of one device feed one batcher
        if (ctx->ref) out->renc.reset(new refenc(ctx->ref));
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
        // The objects being replaced own the mutex that guards them, so they are parked in locals declared BEFORE the
        // guard: they are destroyed after it has unlocked (and only once the new ones are in place).
        std::unique_ptr<engine> old_eng, old_eng2;
        std::unique_ptr<refenc> old_renc;
        std::shared_ptr<batcher> old_bat;
        std::string load_err;
        auto m = load_model(ctx->vocoder_path, dev, load_err);
        if (!m) throw std::runtime_error("decoder backend setup failed: " + load_err);
        auto r = load_refmodel(ctx->vocoder_path, ctx->wavlm_path, dev, m->hp.decoder_adanorm_dim, load_err);
        if (!r) throw std::runtime_error("wavlm backend setup failed: " + load_err);
        std::unique_ptr<engine> new_eng(new engine(m));
        std::unique_ptr<refenc> new_renc(new refenc(r));
        std::shared_ptr<batcher> new_bat;
        if (ctx->bat) new_bat = std::make_shared<batcher>(m, ctx->bat->max_batch, ctx->bat->max_wait_us, ctx->bat->workers);
        {
            std::lock_guard<std::mutex> lk(ctx->eng->mtx);
            std::lock_guard<std::mutex> lk2(ctx->eng2_mtx);
            old_eng = std::move(ctx->eng);
            old_eng2 = std::move(ctx->eng2);
            old_renc = std::move(ctx->renc);
            old_bat = std::move(ctx->bat);
            ctx->mdl = m;
            ctx->ref = r;
            ctx->eng = std::move(new_eng);
            ctx->renc = std::move(new_renc);
            ctx->bat = new_bat;
        }
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

// checks of mio_tts_reference_to_embedding, lib.cpp:1056-1083, in the reference's order
static bool reference_checks(struct mio_tts_context * ctx, char * err, size_t err_size) {
    if (!ctx->mdl->hp.dynamic_global) { set_err(err, err_size, "reference embedding requires dynamic-global MioCodec"); return false; }
    if (!ctx->ref || !ctx->ref->has_global_encoder()) {
        set_err(err, err_size, "reference embedding requires global_encoder tensors in MioCodec GGUF");
        return false;
    }
    if (ctx->wavlm_path.empty() || !ctx->ref->has_wavlm()) { set_err(err, err_size, "WavLM model is not loaded"); return false; }
    return true;
}

bool mio_tts_b200_reference_pcm_to_embedding_batch(struct mio_tts_context * ctx, const float * const * pcm, const size_t * n_samples,
                                                   int32_t sample_rate, size_t batch, struct mio_tts_params params,
                                                   float ** embedding_out, size_t * n_embedding_out, char * err, size_t err_size) {
    if (ctx == nullptr || !ctx->mdl) { set_err(err, err_size, "context is null"); return false; }
    if (pcm == nullptr || n_samples == nullptr || batch == 0) { set_err(err, err_size, "reference audio is empty"); return false; }
    if (embedding_out == nullptr || n_embedding_out == nullptr) { set_err(err, err_size, "output pointers are null"); return false; }
    if (!reference_checks(ctx, err, err_size)) return false;
    std::vector<float *> bufs(batch, nullptr);
    const bool ok = guarded(err, err_size, [&] {
        std::lock_guard<std::mutex> lk(ctx->renc->mtx);
        refenc & R = *ctx->renc;
        try {
            R.prepare(pcm, n_samples, sample_rate, params.max_reference_seconds, batch);
            R.run_wavlm();
        } catch (const std::exception & e) {
            throw std::runtime_error(std::string("failed to extract WavLM SSL features: ") + e.what());
        }
        try {
            R.run_global();
        } catch (const std::exception & e) {
            throw std::runtime_error(std::string("failed to encode global embedding: ") + e.what());
        }
        const size_t n_out = (size_t) ctx->ref->ge_out;
        for (size_t c = 0; c < batch; ++c) {
            bufs[c] = (float *) std::malloc(n_out * sizeof(float));      // lib.cpp:1115, freed by mio_tts_embedding_free
            if (bufs[c] == nullptr) throw std::runtime_error("failed to allocate output embedding");
            R.fetch_embedding(c, bufs[c]);
        }
        R.sync();
        for (size_t c = 0; c < batch; ++c) { embedding_out[c] = bufs[c]; n_embedding_out[c] = n_out; }
    });
    if (!ok) {
        try { ctx->renc->sync(); } catch (...) {}
        for (float * p : bufs) std::free(p);
    }
    return ok;
}

bool mio_tts_reference_to_embedding(struct mio_tts_context * ctx, const char * reference_audio, struct mio_tts_params params,
                                    float ** embedding_out, size_t * n_embedding_out, char * err, size_t err_size) {
    if (ctx == nullptr || !ctx->mdl) { set_err(err, err_size, "context is null"); return false; }
    if (reference_audio == nullptr || reference_audio[0] == '\0') { set_err(err, err_size, "reference_audio path is empty"); return false; }
    if (embedding_out == nullptr || n_embedding_out == nullptr) { set_err(err, err_size, "output pointers are null"); return false; }
    if (!reference_checks(ctx, err, err_size)) return false;
    std::vector<float> pcm;
    int sr = 0;
    std::string e;
    if (!read_wav_mono(reference_audio, pcm, sr, e)) {
        set_err(err, err_size, "failed to extract WavLM SSL features: failed to decode reference audio: " + e);
        return false;
    }
    // the reference asks miniaudio for the codec sample rate, then resamples to WavLM's 16 kHz itself; a file at
    // another rate goes through one linear resampling here (wavlm-extractor.cpp:218-240) instead of miniaudio's filter
    const float * p = pcm.data();
    const size_t n = pcm.size();
    return mio_tts_b200_reference_pcm_to_embedding_batch(ctx, &p, &n, sr, 1, params, embedding_out, n_embedding_out, err, err_size);
}

// parity hooks: WavLM features of PCM clips, and the global encoder alone on caller-supplied features
bool mio_tts_b200_extract_ssl(struct mio_tts_context * ctx, const float * pcm, size_t n_samples, int32_t sample_rate,
                              float max_reference_seconds, float ** ssl_out, int32_t * n_frames_out, int32_t * dim_out, char * err,
                              size_t err_size) {
    if (ctx == nullptr || !ctx->mdl) { set_err(err, err_size, "context is null"); return false; }
    if (ssl_out == nullptr || n_frames_out == nullptr || dim_out == nullptr) { set_err(err, err_size, "output pointers are null"); return false; }
    if (!ctx->ref || !ctx->ref->has_wavlm()) { set_err(err, err_size, "WavLM model is not loaded"); return false; }
    float * buf = nullptr;
    const bool ok = guarded(err, err_size, [&] {
        std::lock_guard<std::mutex> lk(ctx->renc->mtx);
        refenc & R = *ctx->renc;
        R.prepare(&pcm, &n_samples, sample_rate, max_reference_seconds, 1);
        R.run_wavlm();
        const size_t n = (size_t) R.frames(0) * ctx->ref->hp.embed_dim;
        buf = (float *) std::malloc(std::max<size_t>(n, 1) * sizeof(float));
        if (buf == nullptr) throw std::runtime_error("failed to allocate ssl features");
        R.fetch_ssl(0, buf);
        R.sync();
        *ssl_out = buf; *n_frames_out = R.frames(0); *dim_out = ctx->ref->hp.embed_dim;
    });
    if (!ok) std::free(buf);
    return ok;
}

bool mio_tts_b200_ssl_to_embedding(struct mio_tts_context * ctx, const float * ssl, int32_t n_frames, int32_t channels,
                                   float ** embedding_out, size_t * n_embedding_out, char * err, size_t err_size) {
    if (ctx == nullptr || !ctx->mdl) { set_err(err, err_size, "context is null"); return false; }
    if (embedding_out == nullptr || n_embedding_out == nullptr) { set_err(err, err_size, "output pointers are null"); return false; }
    if (!ctx->ref || !ctx->ref->has_global_encoder()) { set_err(err, err_size, "global encoder tensors are not available in this GGUF"); return false; }
    if (ssl == nullptr || n_frames <= 0) { set_err(err, err_size, "invalid ssl frame length"); return false; }
    if (channels != ctx->ref->ge_in) { set_err(err, err_size, "ssl feature shape mismatch"); return false; }
    float * buf = nullptr;
    const bool ok = guarded(err, err_size, [&] {
        std::lock_guard<std::mutex> lk(ctx->renc->mtx);
        refenc & R = *ctx->renc;
        const int fr = n_frames;
        R.prepare_ssl(&ssl, &fr, 1);
        R.run_global();
        buf = (float *) std::malloc((size_t) ctx->ref->ge_out * sizeof(float));
        if (buf == nullptr) throw std::runtime_error("failed to allocate output embedding");
        R.fetch_embedding(0, buf);
        R.sync();
        *embedding_out = buf; *n_embedding_out = (size_t) ctx->ref->ge_out;
    });
    if (!ok) std::free(buf);
    return ok;
}

// resident timing of the reference path: `batch` copies of one clip, `iters` passes, milliseconds per batch
float mio_tts_b200_reference_bench(struct mio_tts_context * ctx, const float * pcm, size_t n_samples, int32_t sample_rate,
                                   int32_t batch, int32_t iters, char * err, size_t err_size) {
    if (ctx == nullptr || !ctx->mdl) { set_err(err, err_size, "context is null"); return -1.0f; }
    if (!reference_checks(ctx, err, err_size)) return -1.0f;
    float ms = -1.0f;
    guarded(err, err_size, [&] {
        std::lock_guard<std::mutex> lk(ctx->renc->mtx);
        std::vector<const float *> pp((size_t) std::max(1, batch), pcm);
        std::vector<size_t> nn((size_t) std::max(1, batch), n_samples);
        ctx->renc->prepare(pp.data(), nn.data(), sample_rate, 0.0f, pp.size());
        ms = ctx->renc->timed(std::max(1, iters), true);
    });
    return ms;
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

// MIO_TTS_B200_PIPELINE_CODES: smallest batch (in codes) that is decoded as two pipelined halves; 0 = never
static size_t pipeline_min_codes() {
    static const size_t v = [] {
        const char * e = std::getenv("MIO_TTS_B200_PIPELINE_CODES");
        const long long x = e ? std::atoll(e) : 20000;
        return x <= 0 ? (size_t) -1 : (size_t) x;
    }();
    return v;
}

// shared body of the batch entry points: fp32 waveforms (wav16 = false) or RIFF/WAVE 16-bit buffers (wav16 = true)
static bool synthesize_batch_any(struct mio_tts_context * ctx, const int32_t * const * codes, const size_t * n_codes,
                                 const float * const * embedding, size_t n_embedding, size_t batch, bool wav16, void ** out,
                                 size_t * n_out, int32_t * sample_rate_out, char * err, size_t err_size) {
    if (ctx == nullptr || !ctx->mdl) { set_err(err, err_size, "context is null"); return false; }
    if (codes == nullptr || n_codes == nullptr || batch == 0) { set_err(err, err_size, "codes are empty"); return false; }
    for (size_t u = 0; u < batch; ++u)
        if (codes[u] == nullptr || n_codes[u] == 0) { set_err(err, err_size, "codes are empty"); return false; }
    if (out == nullptr || n_out == nullptr || (!wav16 && sample_rate_out == nullptr)) {   // the WAV header carries the rate
        set_err(err, err_size, "output pointers are null");
        return false;
    }
    const bool has_emb = embedding != nullptr && n_embedding > 0;
    std::vector<decode_request> reqs(batch);
    std::vector<decode_request *> rp(batch);
    for (size_t u = 0; u < batch; ++u) {
        reqs[u].codes = codes[u]; reqs[u].n_codes = n_codes[u];
        reqs[u].emb = has_emb ? embedding[u] : nullptr; reqs[u].n_emb = has_emb ? n_embedding : 0;
        reqs[u].wav16 = wav16;
        rp[u] = &reqs[u];
    }
    if (batch == 1 && ctx->bat) {
        ctx->bat->submit(reqs[0]);             // coalesced with whatever the other slots have queued
        if (!reqs[0].ok) { set_err(err, err_size, reqs[0].err); return false; }
    } else if (!guarded(err, err_size, [&] {
                   // A batch large enough that each half still fills the GPU (>= ~10 k codes per half) runs as two
                   // pipelined halves: device -> host copies of the first under the kernels of the second.
                   size_t total = 0;
                   for (size_t u = 0; u < batch; ++u) total += n_codes[u];
                   // Only for a lone caller: with other batch calls in flight (server slots, each on its own clone) the
                   // copies of one call already hide under the kernels of another, and four interleaved streams cost 3 %.
                   // "Lone" = no two batch calls overlapped during the last 200 ms (slots that run in lockstep leave short
                   // windows with nothing in flight, so the instantaneous count alone is not enough).
                   static std::atomic<int> in_flight{0};
                   static std::atomic<long long> last_overlap_ns{-(1ll << 62)};
                   struct flight { int n; flight() : n(++in_flight) {} ~flight() { --in_flight; } } fl;
                   const long long now_ns = std::chrono::duration_cast<std::chrono::nanoseconds>(
                                                std::chrono::steady_clock::now().time_since_epoch()).count();
                   if (fl.n >= 2) last_overlap_ns.store(now_ns);
                   const bool lone = fl.n == 1 && now_ns - last_overlap_ns.load() > 200000000ll;
                   if (batch >= 2 && total >= pipeline_min_codes() && lone) {
                       {
                           std::lock_guard<std::mutex> lk(ctx->eng2_mtx);
                           if (!ctx->eng2) ctx->eng2.reset(new engine(ctx->mdl));
                       }
                       decode_batch_pipelined(*ctx->eng, *ctx->eng2, rp.data(), batch);
                       ctx->last_h2d = ctx->eng->h2d_bytes() + ctx->eng2->h2d_bytes();
                   } else {
                       decode_batch(*ctx->eng, rp.data(), batch);
                       ctx->last_h2d = ctx->eng->h2d_bytes();
                   }
               })) {
        return false;
    }
    for (size_t u = 0; u < batch; ++u) { out[u] = reqs[u].out; n_out[u] = wav16 ? 44 + 2 * reqs[u].n_samples : reqs[u].n_samples; }
    if (sample_rate_out) *sample_rate_out = ctx->mdl->hp.sample_rate;
    return true;
}

bool mio_tts_b200_synthesize_batch(struct mio_tts_context * ctx, const int32_t * const * codes, const size_t * n_codes,
                                   const float * const * embedding, size_t n_embedding, size_t batch,
                                   struct mio_tts_params, float ** audio_out, size_t * n_audio_out,
                                   int32_t * sample_rate_out, char * err, size_t err_size) {
    return synthesize_batch_any(ctx, codes, n_codes, embedding, n_embedding, batch, false, (void **) audio_out, n_audio_out,
                                sample_rate_out, err, err_size);
}

bool mio_tts_b200_synthesize_wav16_batch(struct mio_tts_context * ctx, const int32_t * const * codes, const size_t * n_codes,
                                         const float * const * embedding, size_t n_embedding, size_t batch,
                                         struct mio_tts_params, uint8_t ** wav_out, size_t * wav_bytes_out,
                                         int32_t * sample_rate_out, char * err, size_t err_size) {
    return synthesize_batch_any(ctx, codes, n_codes, embedding, n_embedding, batch, true, (void **) wav_out, wav_bytes_out,
                                sample_rate_out, err, err_size);
}

void mio_tts_b200_wav_free(uint8_t * wav) {
    if (wav == nullptr) return;
    if (!result_free(wav)) std::free(wav);
}

bool mio_tts_b200_context_enable_batching(struct mio_tts_context * ctx, int32_t max_batch, int32_t max_wait_us, int32_t workers,
                                          char * err, size_t err_size) {
    if (ctx == nullptr || !ctx->mdl) { set_err(err, err_size, "context is null"); return false; }
    if (max_batch <= 0) { ctx->bat.reset(); return true; }
    return guarded(err, err_size, [&] {
        ctx->bat = std::make_shared<batcher>(ctx->mdl, max_batch, max_wait_us, workers <= 0 ? 4 : workers);
    });
}

bool mio_tts_b200_batching_stats(struct mio_tts_context * ctx, uint64_t * requests, uint64_t * batches, uint64_t * largest,
                                 uint64_t * retried_single) {
    if (ctx == nullptr || !ctx->bat) return false;
    const batcher_stats st = ctx->bat->stats();
    if (requests) *requests = st.requests;
    if (batches) *batches = st.batches;
    if (largest) *largest = st.largest;
    if (retried_single) *retried_single = st.retried_single;
    return true;
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
            if (ctx->bat) ctx->bat->reserve(n);   // batching on: every worker engine, for max_batch such utterances
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

// wavlm_conv_stack_out_len over llround(rate * seconds) source samples (wavlm-extractor.cpp:242-255, 331-340)
static long long reference_samples(const struct mio_tts_context * ctx, float max_reference_seconds) {
    const int sr = std::max(1, ctx->mdl->hp.sample_rate);
    const float sec = max_reference_seconds > 0.0f ? max_reference_seconds : 20.0f;   // lib.cpp:1388
    return std::max<long long>(1, std::llround((double) sr * (double) sec));
}

bool mio_tts_reserve_reference_workspace(struct mio_tts_context * ctx, float max_reference_seconds, char * err, size_t err_size) {
    if (ctx == nullptr || !ctx->mdl) { set_err(err, err_size, "context is null"); return false; }
    if (ctx->wavlm_path.empty() || !ctx->ref || !ctx->ref->has_wavlm()) { set_err(err, err_size, "WavLM model is not loaded"); return false; }
    if (!ctx->ref->has_global_encoder()) { set_err(err, err_size, "global encoder tensors are not available in this GGUF"); return false; }
    const size_t n = (size_t) reference_samples(ctx, max_reference_seconds);
    if (!guarded(err, err_size, [&] {
            std::lock_guard<std::mutex> lk(ctx->renc->mtx);
            std::vector<float> z(n, 0.0f);
            const float * p = z.data();
            ctx->renc->prepare(&p, &n, ctx->mdl->hp.sample_rate, 0.0f, 1);   // sizes every buffer for this length
            ctx->renc->sync();
        })) {
        set_err(err, err_size, "failed to reserve reference workspace");
        return false;
    }
    return true;
}

bool mio_tts_estimate_reference_workspace_bytes(struct mio_tts_context * ctx, float max_reference_seconds, uint64_t * out_bytes,
                                                char * err, size_t err_size) {
    if (ctx == nullptr || !ctx->mdl) { set_err(err, err_size, "context is null"); return false; }
    if (out_bytes == nullptr) { set_err(err, err_size, "out_bytes is null"); return false; }
    if (ctx->wavlm_path.empty() || !ctx->ref || !ctx->ref->has_wavlm()) { set_err(err, err_size, "WavLM model is not loaded"); return false; }
    if (!ctx->ref->has_global_encoder()) { set_err(err, err_size, "global encoder tensors are not available in this GGUF"); return false; }
    // closed form of what refenc::prepare allocates for one clip (no seq^2 attention term, unlike
    // wavlm_extractor::estimate_workspace_bytes, wavlm-extractor.cpp:342-372)
    const refmodel & M = *ctx->ref;
    const uint64_t n_src = (uint64_t) reference_samples(ctx, max_reference_seconds);
    const uint64_t n_wav = std::max<uint64_t>(1, (uint64_t) std::llround((double) n_src * (double) M.hp.sample_rate / (double) std::max(1, ctx->mdl->hp.sample_rate)));
    const uint64_t l0 = n_wav / (uint64_t) std::max(1, M.hp.conv_stride[0]) + 256;
    const uint64_t seq = l0 / 64 + 192;
    uint64_t b = (n_src + n_wav) * 4 + l0 * (uint64_t) M.conv_ch * 2 * 2 + (l0 / 128 + 1) * (uint64_t) M.conv_ch * 16;
    b += seq * ((uint64_t) M.hp.embed_dim * (4 + 4 + 2 + 2 + 6 + 2 + 4) + (uint64_t) M.ffn * 2);
    b += seq * ((uint64_t) M.ge_dim * 8 + (uint64_t) std::max(M.ge_dim, M.ge_ff) * 4);
    *out_bytes = b + b / 8 + (32ull << 20);
    return true;
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
        out->gemm_bytes = t.gemm_bytes;
    });
}

uint64_t mio_tts_b200_resident_samples(struct mio_tts_context * ctx) {
    if (ctx == nullptr || !ctx->eng) return 0;
    uint64_t n = 0;
    for (size_t u = 0; u < ctx->eng->batch(); ++u) n += ctx->eng->audio_len(u);
    return n;
}

bool mio_tts_b200_audio_to_pcm16(struct mio_tts_context * ctx, const float * const * audio, const size_t * n_audio, size_t batch,
                                 int16_t ** pcm_out, char * err, size_t err_size) {
    if (ctx == nullptr || !ctx->mdl) { set_err(err, err_size, "context is null"); return false; }
    if (audio == nullptr || n_audio == nullptr || pcm_out == nullptr || batch == 0) { set_err(err, err_size, "audio is empty"); return false; }
    std::vector<int16_t *> bufs(batch, nullptr);
    const bool ok = guarded(err, err_size, [&] {
        for (size_t u = 0; u < batch; ++u) {
            if (audio[u] == nullptr || n_audio[u] == 0) throw std::runtime_error("audio is empty");
            bufs[u] = (int16_t *) std::malloc(n_audio[u] * 2);
            if (bufs[u] == nullptr) throw std::runtime_error("failed to allocate output audio");
        }
        std::lock_guard<std::mutex> lk(ctx->eng->mtx);
        ctx->eng->pcm16_host(audio, n_audio, batch, bufs.data());
    });
    if (!ok) { for (int16_t * p : bufs) std::free(p); return false; }
    for (size_t u = 0; u < batch; ++u) pcm_out[u] = bufs[u];
    return true;
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

float mio_tts_b200_gather_bench(struct mio_tts_context * ctx, int32_t iters, char * err, size_t err_size) {
    if (ctx == nullptr || !ctx->mdl) { set_err(err, err_size, "context is null"); return -1.0f; }
    float ms = -1.0f;
    guarded(err, err_size, [&] {
        std::lock_guard<std::mutex> lk(ctx->eng->mtx);
        ms = ctx->eng->gather_bench(iters);
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
    return (ctx == nullptr || !ctx->eng) ? 0 : (ctx->last_h2d ? ctx->last_h2d : (uint64_t) ctx->eng->h2d_bytes());
}

} // extern "C"
