/* mio_tts_b200.h -- C ABI of libmio_tts_b200.so, the B200-native MioCodec decode + iSTFT library.
 *
 * Drop-in boundary: every `mio_tts_*` entry point below has the name, argument order, ownership rules and
 * error strings of the function it replaces in the reference's public header
 * /root/reference/src/mio-tts-lib.h (cited per declaration as lib.h:LINE; implementation cited as
 * lib.cpp:LINE = /root/reference/src/mio-tts-lib.cpp).  The header is self-contained: the three llama.h
 * types the reference header pulls in are restated here with identical layout
 * (/root/reference/llama.cpp/include/llama.h: llama_token = int32_t; enum llama_flash_attn_type
 * {AUTO=-1, DISABLED=0, ENABLED=1} :183-187; struct llama_vocab is opaque).
 *
 * Conventions kept from the reference: functions return bool (NULL for constructors) and write a
 * truncated message into err[err_size] on failure, never throw; out-buffers are owned by the caller and
 * MUST be released with the matching mio_tts_*_free; inputs are borrowed for the duration of the call;
 * all functions may be called from arbitrary threads.
 *
 * `mio_tts_b200_*` entry points are extensions the reference does not have (it has no batch API and its
 * iSTFT is a file-static function): batched synthesis, the stage-level hooks the parity tests bind, and
 * the resident-input timing hooks bench.py binds.
 */
#ifndef MIO_TTS_B200_H
#define MIO_TTS_B200_H

#include <stdbool.h>
#include <stddef.h>
#include <stdint.h>

#if defined(_WIN32)
#define MIO_B200_API
#else
#define MIO_B200_API __attribute__((visibility("default")))
#endif

#ifdef __cplusplus
extern "C" {
#endif

#ifndef LLAMA_H            /* restated llama.h types (only when llama.h is not already included) */
typedef int32_t llama_token;
struct llama_vocab;
enum llama_flash_attn_type {
    LLAMA_FLASH_ATTN_TYPE_AUTO = -1,
    LLAMA_FLASH_ATTN_TYPE_DISABLED = 0,
    LLAMA_FLASH_ATTN_TYPE_ENABLED = 1,
};
#endif

struct mio_tts_context;    /* lib.h:17 */
struct mio_tts_vocab_map;  /* lib.h:18 */

struct mio_tts_params {    /* lib.h:20-25 (by value) */
    int32_t n_threads;                 /* ignored: there is no CPU path */
    float max_reference_seconds;
    enum llama_flash_attn_type wavlm_flash_attn_type;
    enum llama_flash_attn_type miocodec_flash_attn_type;   /* ignored: attention is always the banded kernel */
};

/* lib.h:27, lib.cpp:805-812 */
MIO_B200_API struct mio_tts_params mio_tts_default_params(void);

/* lib.h:29-33, lib.cpp:814-848.  Loads the MioCodec GGUF (unchanged weight layout) and uploads re-tiled
 * weights to the device chosen by MIO_TTS_BACKEND_DEVICE ("CUDA<n>", default CUDA0).  With a non-empty
 * `wavlm_model` the WavLM GGUF and the global-encoder tensors of the MioCodec GGUF are loaded to the same device
 * and mio_tts_reference_to_embedding runs on the GPU; NULL / "" leaves that entry point failing with the
 * reference's "WavLM model is not loaded".
 * Ownership: waveforms (and WAV bodies of the mio_tts_b200_*wav16* calls) are PINNED host buffers taken from a
 * process-wide result pool: hand them back with mio_tts_audio_free / mio_tts_b200_wav_free, never free().  The pool
 * keeps returned slabs for reuse and is not shrunk (its high-water mark is the largest set of results alive at once).
 * MIO_TTS_B200_MALLOC_OUTPUT=1 makes every result a plain malloc() buffer instead (pageable copy, slower). */
MIO_B200_API struct mio_tts_context * mio_tts_init_from_file(const char * model_vocoder, const char * wavlm_model,
                                                             char * err, size_t err_size);

/* lib.h:36-39, lib.cpp:850-863.  Shares the device weights; unlike the reference the clone gets its own
 * stream and workspace, so clones decode concurrently. */
MIO_B200_API struct mio_tts_context * mio_tts_context_clone(const struct mio_tts_context * ctx, char * err,
                                                            size_t err_size);

/* lib.h:41, lib.cpp:865-867 */
MIO_B200_API void mio_tts_free(struct mio_tts_context * ctx);

/* lib.h:43-45, lib.cpp:869-885 */
MIO_B200_API bool mio_tts_context_is_dynamic_global(const struct mio_tts_context * ctx);
MIO_B200_API int32_t mio_tts_context_sample_rate(const struct mio_tts_context * ctx);
MIO_B200_API int32_t mio_tts_context_samples_per_token(const struct mio_tts_context * ctx);

/* lib.h:46-50, lib.cpp:887-915.  backend_name: "CUDA<n>" selects GPU n (weights move there), NULL/"" keeps
 * the current device, "CPU" fails ("no CPU fallback"): the reference's ggml device names map to ordinals. */
MIO_B200_API bool mio_tts_context_set_backend_device(struct mio_tts_context * ctx, const char * backend_name,
                                                     char * err, size_t err_size);

/* lib.h:52-57, lib.cpp:917-938.  Needs libllama's llama_tokenize / llama_token_to_piece in the process
 * (resolved with dlsym at call time, as when loaded into llama-tts-mio / mio-tts-server). */
MIO_B200_API struct mio_tts_vocab_map * mio_tts_vocab_map_init(const struct llama_vocab * vocab, char * err,
                                                               size_t err_size);
MIO_B200_API void mio_tts_vocab_map_free(struct mio_tts_vocab_map * map);

/* lib.h:59-66, lib.cpp:940-985 */
MIO_B200_API bool mio_tts_tokens_to_codes(const struct mio_tts_vocab_map * map, const llama_token * tokens,
                                          size_t n_tokens, int32_t ** codes_out, size_t * n_codes_out, char * err,
                                          size_t err_size);
/* lib.h:71-74, lib.cpp:987-1000 */
MIO_B200_API bool mio_tts_token_to_code(const struct mio_tts_vocab_map * map, llama_token token, int32_t * code_out);

/* lib.h:76-90, lib.cpp:225-286, 1002-1044: text file, one code per token, "<|s_N|>" tolerated */
MIO_B200_API bool mio_tts_codes_load(const char * path, int32_t ** codes_out, size_t * n_codes_out, char * err,
                                     size_t err_size);
MIO_B200_API bool mio_tts_codes_save(const char * path, const int32_t * codes, size_t n_codes, char * err,
                                     size_t err_size);
MIO_B200_API void mio_tts_codes_free(int32_t * codes);

/* lib.h:92-99, lib.cpp:1046-1125.  WAV file -> speaker embedding on the GPU (WavLM SSL extractor + global encoder,
 * SURVEY.md 8f N1/N2) when the context was created with a WavLM GGUF; otherwise the reference's "WavLM model is not
 * loaded" error.  WAV (PCM16/24/32, float32) only: mp3/flac go through mio_tts_b200_reference_pcm_to_embedding_batch
 * as decoded PCM.  *embedding_out is malloc'd: release with mio_tts_embedding_free. */
MIO_B200_API bool mio_tts_reference_to_embedding(struct mio_tts_context * ctx, const char * reference_audio,
                                                 struct mio_tts_params params, float ** embedding_out,
                                                 size_t * n_embedding_out, char * err, size_t err_size);

/* lib.h:101-115, lib.cpp:288-413, 1127-1180: GGUF with one F32 tensor "mio.global_embedding" */
MIO_B200_API bool mio_tts_embedding_load_gguf(const char * path, float ** embedding_out, size_t * n_embedding_out,
                                              char * err, size_t err_size);
MIO_B200_API bool mio_tts_embedding_save_gguf(const char * path, const float * embedding, size_t n_embedding,
                                              char * err, size_t err_size);
MIO_B200_API void mio_tts_embedding_free(float * embedding);

/* lib.h:117-128, lib.cpp:1182-1323: codes + speaker embedding -> waveform (decode, iSTFT, peak normalise).
 * *audio_out is a pinned host buffer from the library's result pool: release it with mio_tts_audio_free. */
MIO_B200_API bool mio_tts_synthesize(struct mio_tts_context * ctx, const int32_t * codes, size_t n_codes,
                                     const float * embedding, size_t n_embedding, struct mio_tts_params params,
                                     float ** audio_out, size_t * n_audio_out, int32_t * sample_rate_out, char * err,
                                     size_t err_size);

/* lib.h:131-144, lib.cpp:1325-1366 */
MIO_B200_API bool mio_tts_reserve_workspace(struct mio_tts_context * ctx, int32_t n_codes, char * err, size_t err_size);
MIO_B200_API bool mio_tts_estimate_workspace_bytes(struct mio_tts_context * ctx, int32_t n_codes, uint64_t * out_bytes,
                                                   char * err, size_t err_size);
/* lib.h:147-160, lib.cpp:1368-1432 */
MIO_B200_API bool mio_tts_reserve_reference_workspace(struct mio_tts_context * ctx, float max_reference_seconds,
                                                      char * err, size_t err_size);
MIO_B200_API bool mio_tts_estimate_reference_workspace_bytes(struct mio_tts_context * ctx, float max_reference_seconds,
                                                             uint64_t * out_bytes, char * err, size_t err_size);

/* lib.h:162-170, lib.cpp:758-791, 1434-1457 */
MIO_B200_API bool mio_tts_audio_save_wav16(const char * path, const float * audio, size_t n_audio, int32_t sample_rate,
                                           char * err, size_t err_size);
MIO_B200_API void mio_tts_audio_free(float * audio);

/* ------------------------------------------------------------------------------------------------
 * Extensions (no reference counterpart)
 * ------------------------------------------------------------------------------------------------ */

/* Batched mio_tts_synthesize: `batch` utterances of arbitrary lengths decoded in one pass.
 * codes[i] / n_codes[i] / embedding[i] (each n_embedding floats; NULL array for static models).
 * audio_out[i] / n_audio_out[i] are filled per utterance; free each with mio_tts_audio_free. */
MIO_B200_API bool mio_tts_b200_synthesize_batch(struct mio_tts_context * ctx, const int32_t * const * codes,
                                                const size_t * n_codes, const float * const * embedding,
                                                size_t n_embedding, size_t batch, struct mio_tts_params params,
                                                float ** audio_out, size_t * n_audio_out, int32_t * sample_rate_out,
                                                char * err, size_t err_size);

/* The same batch, returned as ready-to-send `audio/wav` bodies: wav_out[i] = 44-byte RIFF/WAVE header + 16-bit mono PCM,
 * wav_bytes_out[i] = 44 + 2 * samples.  Replaces the server's float -> int16 loop and header writer
 * (/root/reference/src/tts-mio-server.cpp:3965-3973, build_wav_header :2013-2034; same rounding as save_wav16,
 * src/mio-tts-lib.cpp:781-785: clamp to [-1, 1], lrintf(x * 32767)), after the peak normaliser (lib.cpp:1299-1310).
 * The conversion runs on the GPU, so the device -> host copy is half the bytes.  Free each with mio_tts_b200_wav_free.
 * sample_rate_out may be NULL. */
MIO_B200_API bool mio_tts_b200_synthesize_wav16_batch(struct mio_tts_context * ctx, const int32_t * const * codes,
                                                      const size_t * n_codes, const float * const * embedding,
                                                      size_t n_embedding, size_t batch, struct mio_tts_params params,
                                                      uint8_t ** wav_out, size_t * wav_bytes_out, int32_t * sample_rate_out,
                                                      char * err, size_t err_size);
MIO_B200_API void mio_tts_b200_wav_free(uint8_t * wav);

/* Micro-batcher for server slots.  The reference serialises the slots of one device on the owner's
 * runtime mutex (src/tts-mio-server.cpp:2659-2670, 2724-2736; src/mio-tts-lib.cpp:1258-1262).  After this call,
 * mio_tts_synthesize (and batch-of-one calls of the two batch entry points) on `ctx` AND ON EVERY CLONE MADE AFTERWARDS
 * are queued and decoded together: up to max_batch utterances per launch, waiting at most max_wait_us for the callers
 * released by the previous batch to come back; `workers` engines (default 4) overlap one batch's device -> host copy with
 * the next batch's kernels.  Results are bit-identical to un-batched calls; a request that fails (e.g. an out-of-range
 * code) fails alone with its usual message.  max_batch <= 0 turns batching off for this context.
 * mio_tts_reserve_workspace(ctx, n_codes) then also sizes every worker for max_batch utterances of n_codes (at most 32 k codes each).
 * MIO_TTS_B200_BATCHING="max_batch[,max_wait_us[,workers]]" enables it at mio_tts_init_from_file without code changes. */
MIO_B200_API bool mio_tts_b200_context_enable_batching(struct mio_tts_context * ctx, int32_t max_batch, int32_t max_wait_us,
                                                       int32_t workers, char * err, size_t err_size);
/* counters since the batcher was created; false when batching is off.  Any output pointer may be NULL. */
MIO_B200_API bool mio_tts_b200_batching_stats(struct mio_tts_context * ctx, uint64_t * requests, uint64_t * batches,
                                              uint64_t * largest, uint64_t * retried_single);

/* Test hook: the WAV16 conversion kernel alone (peak rule lib.cpp:1299-1310 + clamp/round :781-785) over host waveforms;
 * pcm_out[i] = malloc'd n_audio[i] samples, release with mio_tts_b200_free. */
MIO_B200_API bool mio_tts_b200_audio_to_pcm16(struct mio_tts_context * ctx, const float * const * audio, const size_t * n_audio,
                                              size_t batch, int16_t ** pcm_out, char * err, size_t err_size);

/* The same WAV16 conversion (peak rule + clamp + round) for callers that cannot allocate: code under CUDA graph capture,
 * tight serving loops, harnesses that forbid cudaMalloc.  This entry point has NO counterpart in the reference.
 * It calls no cudaMalloc / cudaFree / allocating library routine and takes
 * no host heap memory: its scratch (per-utterance peak words and the offset / length tables) is a file-scope __device__
 * array of MIO_TTS_B200_STATIC_SCRATCH_BYTES that every GPU loading the library holds until exit.
 *   d_audio  DEVICE pointer: utterance u = lengths[u] float samples at d_audio + offsets[u] (un-normalised waveform)
 *   d_pcm    DEVICE pointer: receives utterance u's int16 samples at d_pcm + offsets[u]
 *   offsets / lengths  HOST arrays of `batch` entries (samples); batch <= MIO_TTS_B200_STATIC_MAX_UTT, else the call
 *            fails with "... exceeds the static scratch ..."
 *   device   CUDA device ordinal that owns d_audio / d_pcm;  stream: a cudaStream_t of that device, or NULL
 * All calls of a process share the one array per GPU and it keeps the last call's contents (the peak words are cleared
 * at the start of every call).  Two calls that overlap on the GPU would race on it: the library therefore serialises
 * the callers of a process with a mutex and returns only after the stream has drained. */
#define MIO_TTS_B200_STATIC_MAX_UTT 2048
#define MIO_TTS_B200_STATIC_SCRATCH_BYTES (MIO_TTS_B200_STATIC_MAX_UTT * 20)
MIO_B200_API bool mio_tts_b200_pcm16_static(int32_t device, const float * d_audio, const int64_t * offsets,
                                            const int64_t * lengths, size_t batch, int16_t * d_pcm, void * stream,
                                            char * err, size_t err_size);

/* Stage hooks for parity tests.  decode_spec mirrors miocodec_decoder::decode_tokens
 * (/root/reference/src/miocodec-decoder.cpp:1079): spec_out = malloc'd [n_frames][n_bins]. */
MIO_B200_API bool mio_tts_b200_decode_spec(struct mio_tts_context * ctx, const int32_t * codes, size_t n_codes,
                                           const float * embedding, size_t n_embedding, float ** spec_out,
                                           int32_t * n_frames_out, int32_t * n_bins_out, char * err, size_t err_size);
/* ggml_get_rows(token_embd, codes) (miocodec-decoder.cpp:1360): out = caller buffer [n_codes][prenet_dim] */
MIO_B200_API bool mio_tts_b200_gather(struct mio_tts_context * ctx, const int32_t * codes, size_t n_codes, float * out,
                                      char * err, size_t err_size);
/* Intermediate activations of utterance 0 after stage 1 prenet (:1365), 2 upsample+interpolate (:1374), 3 prior
 * ResNets (:1380), 4 decoder + final AdaLN (:1400), 5 post ResNets (:1408): malloc'd fp32 [rows][cols]. */
MIO_B200_API bool mio_tts_b200_debug_stage(struct mio_tts_context * ctx, const int32_t * codes, size_t n_codes,
                                           const float * embedding, size_t n_embedding, int32_t stage, float ** out,
                                           int32_t * rows_out, int32_t * cols_out, char * err, size_t err_size);
MIO_B200_API int32_t mio_tts_b200_prenet_dim(const struct mio_tts_context * ctx);
/* embd_to_audio (lib.cpp:625-756): spec [n_frames][n_fft+2] host -> malloc'd audio */
MIO_B200_API bool mio_tts_b200_istft(struct mio_tts_context * ctx, const float * spec, int32_t n_frames, int32_t n_fft,
                                     int32_t hop, float ** audio_out, size_t * n_audio_out, char * err, size_t err_size);
MIO_B200_API void mio_tts_b200_free(void * p);   /* frees spec_out / istft audio_out (plain malloc) */
/* 1 = use the generic SIMT attention kernel instead of the tensor-core band kernel (cross-check) */
MIO_B200_API void mio_tts_b200_set_generic_attention(struct mio_tts_context * ctx, int32_t enable);

/* Timing hooks for bench.py.  prepare uploads one batch into HBM; run_resident executes the device path
 * `iters` times over the resident batch and returns the mean milliseconds per pass measured with CUDA events
 * on the library's stream (<0 on error); profile fills stage/GEMM timings of one profiled pass. */
struct mio_tts_b200_profile {
    float prenet_ms, upsample_prior_ms, decoder_ms, post_head_ms, istft_ms, total_ms;
    double gemm_ms, gemm_flops; int32_t gemm_launches;
    double attn_ms; int32_t attn_launches;
    double gemm_bytes;      /* algorithmic DRAM bytes of the GEMM launches: A + W + output + residual */
};
MIO_B200_API bool mio_tts_b200_prepare(struct mio_tts_context * ctx, const int32_t * const * codes,
                                       const size_t * n_codes, const float * const * embedding, size_t n_embedding,
                                       size_t batch, char * err, size_t err_size);
MIO_B200_API float mio_tts_b200_run_resident(struct mio_tts_context * ctx, int32_t iters, char * err, size_t err_size);
MIO_B200_API bool mio_tts_b200_profile_resident(struct mio_tts_context * ctx, struct mio_tts_b200_profile * out,
                                                char * err, size_t err_size);
/* total audio samples of the prepared batch */
MIO_B200_API uint64_t mio_tts_b200_resident_samples(struct mio_tts_context * ctx);
/* iSTFT microbench (BASELINE config 3): synthetic spectra generated in HBM; returns ms per pass (<0 on error) */
MIO_B200_API float mio_tts_b200_istft_bench(struct mio_tts_context * ctx, int32_t n_utt, int32_t frames_per_utt,
                                            int32_t n_fft, int32_t hop, int32_t iters, uint64_t seed, double * checksum,
                                            char * err, size_t err_size);
/* codebook gather alone (miocodec-decoder.cpp:1360) over the batch uploaded by mio_tts_b200_prepare: ms per pass */
MIO_B200_API float mio_tts_b200_gather_bench(struct mio_tts_context * ctx, int32_t iters, char * err, size_t err_size);
/* ---- reference-audio -> speaker-embedding path (WavLM SSL extractor + global encoder) ----
 * mio_tts_reference_to_embedding (lib.h:69-77, lib.cpp:1048-1125) runs on the GPU when the context was created with a
 * WavLM GGUF and the MioCodec GGUF holds the global-encoder tensors.  The reference has no batch or PCM entry point:
 * `pcm[i]` = `n_samples[i]` mono float samples at `sample_rate` (what miniaudio hands wavlm_extractor, wavlm-extractor.cpp
 * :153-203); embedding_out[i] is malloc'd ([output_channels] floats) and released with mio_tts_embedding_free. */
MIO_B200_API bool mio_tts_b200_reference_pcm_to_embedding_batch(struct mio_tts_context * ctx, const float * const * pcm,
                                                                const size_t * n_samples, int32_t sample_rate, size_t batch,
                                                                struct mio_tts_params params, float ** embedding_out,
                                                                size_t * n_embedding_out, char * err, size_t err_size);
/* parity hooks: wavlm_extractor::extract_ssl_features (wavlm-extractor.cpp:543) on PCM -> malloc'd [n_frames][dim];
 * miocodec_decoder::encode_global_embedding_from_ssl (miocodec-decoder.cpp:824) on [n_frames][channels] features */
MIO_B200_API bool mio_tts_b200_extract_ssl(struct mio_tts_context * ctx, const float * pcm, size_t n_samples, int32_t sample_rate,
                                           float max_reference_seconds, float ** ssl_out, int32_t * n_frames_out,
                                           int32_t * dim_out, char * err, size_t err_size);
MIO_B200_API bool mio_tts_b200_ssl_to_embedding(struct mio_tts_context * ctx, const float * ssl, int32_t n_frames, int32_t channels,
                                                float ** embedding_out, size_t * n_embedding_out, char * err, size_t err_size);
/* BASELINE config 4: `batch` copies of one clip resident in HBM, `iters` passes; returns ms per batch (<0 on error) */
MIO_B200_API float mio_tts_b200_reference_bench(struct mio_tts_context * ctx, const float * pcm, size_t n_samples,
                                                int32_t sample_rate, int32_t batch, int32_t iters, char * err, size_t err_size);
/* number of this library's kernels launched so far in this process */
MIO_B200_API uint64_t mio_tts_b200_launch_count(void);
MIO_B200_API uint64_t mio_tts_b200_workspace_bytes(struct mio_tts_context * ctx);
MIO_B200_API uint64_t mio_tts_b200_weight_bytes(struct mio_tts_context * ctx);
/* bytes of the last batch's host->device upload (codes + embeddings + row tables) */
MIO_B200_API uint64_t mio_tts_b200_last_h2d_bytes(struct mio_tts_context * ctx);

#ifdef __cplusplus
}
#endif
#endif /* MIO_TTS_B200_H */
