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

void ref_free(void * p) { std::free(p); }

int ref_hw_threads(void) { return (int) std::thread::hardware_concurrency(); }

} // extern "C"
