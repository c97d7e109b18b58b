// oracle/ref_driver.cpp -- TEST INFRASTRUCTURE ONLY (never linked by the product library).
//
// Thin C driver over the UNMODIFIED reference.  The reference translation unit
// /root/reference/src/mio-tts-lib.cpp is #included (not copied) so that its file-static
// iSTFT `embd_to_audio` (src/mio-tts-lib.cpp:625-756) is reachable next to the public
// mio_tts_* C API it defines.  `miocodec_decoder::decode_tokens`
// (src/miocodec-decoder.cpp:1079) is a public method and is called directly to obtain the
// intermediate spectrogram.
#include "mio-tts-lib.cpp"   // resolved through -I$(REF)/src

#include "ggml-cpu.h"

extern "C" {

// --- decoder-only handle: spectrogram dump --------------------------------------------
void * ref_decoder_open(const char * gguf_path, char * err, size_t err_size) {
    auto * d = new miocodec_decoder();
    std::string e;
    if (!d->load(gguf_path, e)) { set_err(err, err_size, e); delete d; return nullptr; }
    std::string be;
    d->set_backend_device("CPU", be);   // force the CPU (ggml-cpu) path = the parity oracle
    return d;
}
void ref_decoder_close(void * h) { delete (miocodec_decoder *) h; }

// spec_out: malloc'd [n_frames][n_bins] (bins contiguous per frame).  flash: -1 auto, 0 off, 1 on.
bool ref_decoder_decode(void * h, const int32_t * codes, size_t n_codes, const float * emb, size_t n_emb,
                        int n_threads, int flash, float ** spec_out, int * n_frames_out, int * n_bins_out,
                        char * err, size_t err_size) {
    auto * d = (miocodec_decoder *) h;
    std::vector<int32_t> c(codes, codes + n_codes);
    std::vector<float> e;
    if (emb != nullptr && n_emb > 0) e.assign(emb, emb + n_emb);
    std::vector<float> spec; int nf = 0; std::string de;
    if (!d->decode_tokens(c, e, n_threads, (llama_flash_attn_type) flash, spec, nf, de)) {
        set_err(err, err_size, de); return false;
    }
    float * out = (float *) std::malloc(spec.size() * sizeof(float));
    std::memcpy(out, spec.data(), spec.size() * sizeof(float));
    *spec_out = out; *n_frames_out = nf; *n_bins_out = nf > 0 ? (int) (spec.size() / (size_t) nf) : 0;
    return true;
}

// --- iSTFT only: the reference's own embd_to_audio -------------------------------------
// spec: [n_frames][n_fft+2]; returns malloc'd audio, *n_out samples.
float * ref_istft(const float * spec, int n_frames, int n_fft, int hop, int n_threads, size_t * n_out) {
    std::vector<float> a = embd_to_audio(spec, n_frames, n_fft + 2, n_threads, n_fft, hop);
    float * out = (float *) std::malloc(std::max<size_t>(1, a.size()) * sizeof(float));
    std::memcpy(out, a.data(), a.size() * sizeof(float));
    *n_out = a.size();
    return out;
}

// --- reference-audio path --------------------------------------
// wavlm_extractor::extract_ssl_features (src/wavlm-extractor.cpp:543): audio file -> ssl [n_frames][embed_dim]
void * ref_wavlm_open(const char * gguf_path, char * err, size_t err_size) {
    auto * w = new wavlm_extractor();
    std::string e;
    if (!w->load(gguf_path, e)) { set_err(err, err_size, e); delete w; return nullptr; }
    std::string be;
    w->set_backend_device("CPU", be);
    return w;
}
void ref_wavlm_close(void * h) { delete (wavlm_extractor *) h; }

bool ref_wavlm_extract(void * h, const char * audio_path, int source_sample_rate, float max_seconds, int n_threads,
                       int flash, float ** ssl_out, int * n_frames_out, int * dim_out, char * err, size_t err_size) {
    auto * w = (wavlm_extractor *) h;
    std::vector<float> ssl; int nf = 0; std::string e;
    if (!w->extract_ssl_features(audio_path, source_sample_rate, max_seconds, n_threads, (llama_flash_attn_type) flash,
                                 ssl, nf, e)) {
        set_err(err, err_size, e); return false;
    }
    if (!e.empty()) set_err(err, err_size, e);   // fallback notices are reported, the call still succeeds
    float * out = (float *) std::malloc(std::max<size_t>(1, ssl.size()) * sizeof(float));
    std::memcpy(out, ssl.data(), ssl.size() * sizeof(float));
    *ssl_out = out; *n_frames_out = nf; *dim_out = nf > 0 ? (int) (ssl.size() / (size_t) nf) : 0;
    return true;
}

// miocodec_decoder::encode_global_embedding_from_ssl (src/miocodec-decoder.cpp:824): ssl [n_frames][C] -> emb
bool ref_decoder_encode_global(void * h, const float * ssl, int n_frames, int channels, int n_threads, float ** emb_out,
                               int * n_emb_out, char * err, size_t err_size) {
    auto * d = (miocodec_decoder *) h;
    std::vector<float> s(ssl, ssl + (size_t) n_frames * channels), emb; std::string e;
    if (!d->encode_global_embedding_from_ssl(s, n_frames, n_threads, emb, e)) { set_err(err, err_size, e); return false; }
    if (!e.empty()) set_err(err, err_size, e);
    float * out = (float *) std::malloc(std::max<size_t>(1, emb.size()) * sizeof(float));
    std::memcpy(out, emb.data(), emb.size() * sizeof(float));
    *emb_out = out; *n_emb_out = (int) emb.size();
    return true;
}

void ref_free(void * p) { std::free(p); }

int ref_hw_threads(void) { return (int) std::thread::hardware_concurrency(); }

} // extern "C"
