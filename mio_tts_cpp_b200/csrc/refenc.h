// Reference-audio -> speaker-embedding path on the B200 (SURVEY.md section 8f rows N1 + N2):
//   WavLM SSL extractor    wavlm_extractor::extract_ssl_features      /root/reference/src/wavlm-extractor.cpp:543-1081
//   global encoder         miocodec_decoder::encode_global_embedding_from_ssl      src/miocodec-decoder.cpp:824-1066
// i.e. the body of mio_tts_reference_to_embedding (src/mio-tts-lib.cpp:1048-1125), batched over clips.
//
// Layout: every activation is [rows, channels] row-major.  The conv stack keeps clip c at row offset row0[c] >> level
// (level-0 extents are padded to 128 rows) so a stride-2 convolution is ONE implicit GEMM over the pair view
// [rows/2, 2C] of its input for the whole batch; from the feature projection on, clips sit in a padded layout with 64
// zero gap rows between them, which gives the k=128 grouped positional conv and the k=7 embed conv their zero padding
// through plain row-shifted TMA loads.  All dense contractions run on gemm_tcgen05.cu.
#pragma once
#include "common.h"
#include "model.h"

#include <array>
#include <map>
#include <memory>
#include <mutex>
#include <string>
#include <vector>

namespace mio {

struct gguf_file;

struct wavlm_hparams {   // struct defaults: /root/reference/src/wavlm-extractor.h:13-25
    int sample_rate = 16000, n_layers = 2, n_heads = 12, head_dim = 64, embed_dim = 768, num_buckets = 320, max_distance = 800;
    float norm_eps = 1e-5f;
    int conv_kernel[7] = {10, 3, 3, 3, 3, 2, 2};
    int conv_stride[7] = {5, 2, 2, 2, 2, 2, 2};
};

struct wavlm_layer {
    dev_linear in_proj, out_proj, w1, w2;
    float * gru_w = nullptr, * gru_b = nullptr, * gru_const = nullptr;
    dev_norm norm1, norm2;
};

struct ge_block {
    float * dw_w = nullptr, * dw_b = nullptr;   // [dim][KW] fp32 (rounded to fp16 in the kernel like the reference)
    int dw_k = 7;
    dev_norm norm;
    dev_linear pw1, pw2;
    float * gamma = nullptr;
};

// weights of both networks, resident in HBM (WavLM GGUF arch "wavlm-ssl"; global-encoder tensors of the MioCodec GGUF)
class refmodel {
public:
    ~refmodel();
    bool load_wavlm(const std::string & path, int device, std::string & err);                   // wavlm-extractor.cpp:454-540
    bool load_global_encoder(const gguf_file & g, int device, int out_dim, std::string & err);   // miocodec-decoder.cpp:713-744
    bool has_wavlm() const { return wavlm_loaded; }
    bool has_global_encoder() const { return ge_loaded; }

    int device = 0, num_sms = 148;
    int linear_dtype = OP_BF16, conv_dtype = OP_F16;
    // WavLM
    bool wavlm_loaded = false;
    wavlm_hparams hp;
    int conv_ch = 512, ffn = 3072, pos_k = 128, pos_groups = 16, pos_kp = 64;
    float * conv0_w = nullptr; dev_norm conv0_norm;
    dev_linear conv[7];                       // [1..6]: [C, k*C] tap-major fp16
    dev_norm proj_norm, tr_norm;
    dev_linear proj;
    std::vector<dev_linear> pos;              // one per group: N = out/group, K = pos_kp per tap, taps = pos_k
    float * pos_bias = nullptr;
    std::vector<float> rel_embed;             // host [num_buckets][heads]
    std::vector<wavlm_layer> layers;
    // global encoder
    bool ge_loaded = false;
    int ge_in = 768, ge_dim = 384, ge_ff = 1152, ge_out = 128, ge_attn = 128, ge_embed_k = 7;
    dev_linear ge_embed;
    dev_norm ge_norm, ge_final_norm, ge_pool_norm;
    std::vector<ge_block> ge_blocks;
    dev_linear ge_attn0, ge_attn2;
    float * ge_proj_w = nullptr, * ge_proj_b = nullptr;
    size_t weight_bytes = 0;

private:
    std::vector<void *> allocs_;
    void * dalloc(size_t bytes);
    float * upload_f32(const float * src, size_t n);
    void make_linear(dev_linear & L, const std::vector<float> & rows, int N, int K, int taps, int dtype,
                     const std::vector<float> * bias);
    friend struct refloader;
};

struct ref_times { float wavlm_ms = 0, global_ms = 0, total_ms = 0; double gemm_flops = 0; };

// one batch of clips at a time on one stream; workspace grows to the high-water mark
class refenc {
public:
    explicit refenc(std::shared_ptr<refmodel> m);
    ~refenc();
    std::shared_ptr<refmodel> mdl;
    std::mutex mtx;

    // pcm[c]: n[c] mono float samples at `sample_rate`; max_seconds <= 0: no limit (src/wavlm-extractor.cpp:562-600)
    void prepare(const float * const * pcm, const size_t * n, int sample_rate, float max_seconds, size_t batch);
    void run_wavlm();                                   // waveform -> ssl (device)
    void run_global();                                  // ssl (device) -> embeddings (device)
    // ssl given by the caller instead of run_wavlm (parity hook for the global encoder alone): [frames[c]][channels]
    void prepare_ssl(const float * const * ssl, const int * frames, size_t batch);
    size_t batch() const { return (size_t) n_clips_; }
    int frames(size_t c) const { return seq_[c]; }
    void fetch_ssl(size_t c, float * dst);              // [frames(c)][embed_dim], async on the stream
    void fetch_embedding(size_t c, float * dst);        // [ge_out], async
    void sync();
    float timed(int iters, bool with_global);           // ms per batch over the prepared (resident) batch
    size_t workspace_bytes() const;
    size_t h2d_bytes() const { return h2d_bytes_; }
    cudaStream_t stream() const { return st_; }

private:
    struct dbuf { void * p = nullptr; size_t cap = 0; };
    cudaStream_t st_ = nullptr;
    cudaEvent_t ev_[2];
    int n_clips_ = 0;
    bool have_wave_ = false;
    std::vector<int> n_src_, n_wav_, seq_, row0_l0_, prow0_, gate_off_;
    std::vector<std::array<int, 7>> len_;      // conv-stack output lengths per clip
    int rows0_ = 0, prows_ = 0, max_seq_ = 0, total_seq_ = 0, n_chunks_ = 0, src_rate_ = 0;
    size_t h2d_bytes_ = 0;
    // device metadata (one blob)
    std::vector<uint8_t> blob_;
    void * h_blob_ = nullptr; size_t h_blob_cap_ = 0;
    dbuf d_blob_;
    const long long * d_src_off_ = nullptr, * d_wav_off_ = nullptr;
    const int * d_n_src_ = nullptr, * d_n_wav_ = nullptr, * d_len0_ = nullptr, * d_row0_l0_ = nullptr, * d_seq_ = nullptr,
              * d_prow0_ = nullptr, * d_gate_off_ = nullptr, * d_map_pad_ = nullptr, * d_map_valid_ = nullptr,
              * d_map_compact_ = nullptr, * d_row_clip_ = nullptr, * d_row_pos_ = nullptr;
    const int2 * d_work_ = nullptr; int n_work_ = 0;
    // workspace
    dbuf src_, wav_, peak_, part_, mr_, lvl_[7], c6_, a16_, x32_, xp16_, x16_, qkv16_, att16_, ff16_, sum32_, gate_, rel_,
         g32_, g16_, ga16_, gl32_, stat_, emb_, sslc_;
    int rel_max_seq_ = 0;
    std::map<std::array<uint64_t, 3>, CUtensorMap> amaps_;
    std::map<std::array<uint64_t, 5>, CUtensorMap> omaps_;
    double flops_ = 0;

    void ensure(dbuf & b, size_t bytes);
    size_t blob_add(const void * src, size_t bytes);
    void upload_blob();
    void layout_from_seq();
    const CUtensorMap & amap(const void * p, uint64_t width, uint64_t rows);
    const CUtensorMap & omap(const void * p, int rows, int cols, int ld, int elem_bytes);
    struct gopt { int tap_row0 = 0, tap_shift = 0, a_col0 = 0, act = 0; const float * bias = nullptr; const float * gate = nullptr;
                  bool residual = false; int out_dtype = OP_BF16; };
    void gemm(const void * a, uint64_t a_width, uint64_t a_rows, const dev_linear & w, int M, int epilogue, void * out, int ldo,
              int out_cols, const gopt & o);
};

// ---- kernels (ref_ops.cu, attention_full.cu) ----
int wav_conv0_chunk_frames();
void launch_wav_prepare(const float * src, const long long * src_off, const int * n_src, float * wav, const long long * wav_off,
                        const int * n_wav, unsigned int * peak_bits, int rate_in, int rate_out, int n_clips, int max_n_src,
                        int max_n_wav, cudaStream_t st);
// statistics through the clip's autocorrelation (K + K*K sums per clip) instead of a convolution pass
void launch_wav_conv0_moments(const float * wav, const long long * wav_off, const int * n_frames, const float * w, int C, int K,
                              int S, double * partial, int n_chunks, float eps, float2 * mean_rstd, int n_clips, cudaStream_t st);
void launch_wav_conv0_apply(const float * wav, const long long * wav_off, const int * n_frames, const int * row0,
                            const float * w, int C, int K, int S, int n_chunks, const float2 * mean_rstd, const float * gw,
                            const float * gb, void * out16, int n_clips, cudaStream_t st);
void launch_relayout(const float * src, const int * map, int rows, int dim, float scale, float * dst32, void * dst16,
                     int dtype, cudaStream_t st);
void launch_accumulate(const float * x, float * acc, size_t n, bool first, cudaStream_t st);
void launch_wavlm_gate(const float * x, const int * row0, const int * seq_len, const int * gate_off, int n_clips, int max_seq,
                       int H, int hd, const float * gw, const float * gb, const float * gconst, float * gate, cudaStream_t st);
void launch_dwconv_ln(const float * x, const int * row_clip, const int * row_pos, const int * seq_len, int rows, int dim, int KW,
                      const float * w, const float * b, const float * nw, const float * nb, float eps, void * out16, int dtype,
                      cudaStream_t st);
void launch_attn_stats_pool(const float * logits, const float * x, const int * row0, const int * seq_len, int n_clips, int dim,
                            float * stat, cudaStream_t st);
void launch_pool_proj_ln(const float * stat, int in_dim, const float * w, const float * b, const float * nw, const float * nb,
                         float eps, int out_dim, float * emb, int n_clips, cudaStream_t st);
void launch_attention_relbias(const void * qkv16, void * out16, int dtype, int dim, int heads, const int * row0,
                              const int * seq_len, const int * gate_off, const float * gate, const float * rel_d, int max_seq,
                              const int2 * work, int n_work, cudaStream_t st);

} // namespace mio
