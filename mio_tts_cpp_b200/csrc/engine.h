// Batched MioCodec decode + iSTFT executor: one CUDA stream, one workspace arena, one batch at a time.
//
// Replaces the body of mio_tts_synthesize (/root/reference/src/mio-tts-lib.cpp:1182-1323), i.e.
// miocodec_decoder::decode_tokens (src/miocodec-decoder.cpp:1079-1664) + embd_to_audio (:625-756) + the peak
// normaliser (:1299-1310), for B utterances of arbitrary lengths packed along rows (the reference decodes one
// utterance per call under a mutex).  No ggml graph, no per-call allocation: kernels are launched back to back
// on the engine's stream over buffers sized once per high-water mark.
#pragma once
#include "model.h"

#include <array>
#include <map>
#include <memory>
#include <mutex>
#include <vector>

namespace mio {

struct seq_layout {
    int n_utt = 0, gap = 0, rows = 0, valid = 0, max_len = 0, n_tiles = 0, n_tiles128 = 0;
    std::vector<int> len, row0;
    const int * d_len = nullptr, * d_row0 = nullptr, * d_row_utt = nullptr, * d_row_pos = nullptr;
    const row_tile * d_tiles = nullptr, * d_tiles128 = nullptr;   // runs of <= 64 / <= 128 rows inside one utterance
    size_t off_len = 0, off_row0 = 0, off_row_utt = 0, off_row_pos = 0, off_tiles = 0, off_tiles128 = 0;   // offsets in the metadata blob
};

struct ln_fuse {      // the LayerNorm an EPI_STORE32_LN GEMM emits: out16 = LN(x) * (plus1 + a[utt * ld + n]) + b[utt * ld + n]
    void * out16 = nullptr; int out_dtype = 0;
    const float * a = nullptr, * b = nullptr; int ld = 0; float plus1 = 0.0f;
};

struct dbuf {
    void * p = nullptr;
    size_t cap = 0;
};

struct stage_times {   // milliseconds of the last profiled run (CUDA events on the engine's stream)
    float prenet = 0, upsample_prior = 0, decoder = 0, post_head = 0, istft = 0, total = 0;
    double gemm_ms = 0, gemm_flops = 0; int gemm_launches = 0;
    double gemm_bytes = 0;     // algorithmic DRAM bytes of those launches (A + W + out + residual)
    double attn_ms = 0; int attn_launches = 0;
};

class engine {
public:
    explicit engine(std::shared_ptr<model> m);
    ~engine();

    std::shared_ptr<model> mdl;

    // Validates inputs like mio_tts_synthesize (same error strings), builds the row layouts and uploads
    // codes / embeddings / tables.  Throws std::runtime_error with the user-visible message.
    void prepare(const int32_t * const * codes, const size_t * n_codes, const float * const * emb, size_t n_emb,
                 size_t batch);
    // Device path over the prepared batch: decode -> spec (-> iSTFT -> peak normalise when with_istft).
    // stop_stage (debug): 1 prenet, 2 upsample, 3 prior, 4 decoder, 5 post -> stop there and expose that buffer
    // (small repeated batch shapes are replayed from a captured CUDA graph: MIO_TTS_B200_GRAPHS=0 turns that off)
    void run(bool with_istft, stage_times * prof = nullptr, int stop_stage = 0);
    // dense fp32 copy of utterance 0 of the buffer exposed by a stopped run: [rows][cols]
    void fetch_debug(std::vector<float> & out, int & rows, int & cols);
    void sync();
    // results of the prepared batch
    size_t batch() const { return (size_t) n_utt_; }
    size_t audio_len(size_t u) const { return (size_t) audio_len_[u]; }
    int spec_frames(size_t u) const { return spec_->len[u]; }
    int spec_bins() const { return mdl->hp.n_fft + 2; }
    void fetch_audio(size_t u, float * dst);   // async D2H on the engine's stream (call sync() afterwards)
    void fetch_pcm16(size_t u, int16_t * dst); // after a run with out_pcm16 (async D2H)
    void fetch_spec(size_t u, float * dst);    // dense [frames, n_bins]
    size_t h2d_bytes() const { return h2d_bytes_; }

    // stand-alone stages for the parity tests / microbench (host in, host out)
    void gather(const int32_t * codes, size_t n, float * out);
    void istft_host(const float * spec, int n_frames, int n_fft, int hop, std::vector<float> & audio);
    void pcm16_host(const float * const * audio, const size_t * n, size_t batch, int16_t * const * out);
    // resident iSTFT microbench: `n_utt` utterances x `frames` frames of synthetic spectra already in HBM
    float istft_bench(int n_utt, int frames, int n_fft, int hop, int iters, unsigned long long seed, double * checksum);
    float gather_bench(int iters);

    size_t workspace_bytes() const;
    void reserve(int n_codes, int batch);
    bool force_generic_attention = false;
    bool out_pcm16 = false;     // run(): emit clamp/round int16 samples (pcm16_kernel) instead of normalising the fp32 waveform
    cudaStream_t stream() const { return st_; }
    std::mutex mtx;

private:
    cudaStream_t st_ = nullptr;
    cudaEvent_t ev_[8];
    std::vector<cudaEvent_t> gemm_ev_;
    bool prof_gemm_ = false;
    stage_times * prof_ = nullptr;
    std::vector<std::pair<size_t, double>> gemm_ev_flops_;   // (event pair index, flops)
    double gemm_bytes_ = 0;
    std::vector<size_t> attn_ev_;
    std::vector<std::array<int, 5>> gemm_ev_shape_;   // M, N, K, taps, epilogue

    // batch state
    int n_utt_ = 0;
    seq_layout tok_, frm_;
    std::vector<seq_layout> ups_;
    seq_layout * spec_ = nullptr;
    std::vector<long long> audio_off_, audio_len_;
    long long audio_total_ = 0;
    int n_gn_slots_ = 0, gn_groups_max_ = 1;
    const float * snake_a_ = nullptr, * snake_ib_ = nullptr;   // SnakeBeta vectors of the next gemm() call (act 3)
    size_t h2d_bytes_ = 0;
    // metadata blob (pinned host + device)
    std::vector<uint8_t> blob_;
    void * h_blob_ = nullptr; size_t h_blob_cap_ = 0;
    dbuf d_blob_;
    const int * d_codes_ = nullptr; const float * d_emb_ = nullptr;
    const int * d_tok2_row0_ = nullptr, * d_tok2_len_ = nullptr;
    const long long * d_audio_off_ = nullptr, * d_audio_len_ = nullptr;

    // workspace
    dbuf xt_, fx_, fy_, ft_, act16_, qkv16_, att16_, ff16_, cond_, gn_stats_, spec_buf_, audio_, peak_, pcm_;
    int spec_pitch_ = 0;
    const void * dbg_ptr_ = nullptr; int dbg_cols_ = 0, dbg_is16_ = 0, dbg_dtype_ = 0; const seq_layout * dbg_lay_ = nullptr;
    float2 * rope_p_ = nullptr, * rope_d_ = nullptr; int rope_p_pos_ = 0, rope_d_pos_ = 0;
    std::map<std::pair<const void *, int>, CUtensorMap> tmaps_;
    const CUtensorMap & xmap(const void * p, int rows, int cols, int box_rows);
    std::map<std::array<uint64_t, 5>, CUtensorMap> omaps_;   // (ptr, rows, cols, ld, elem bytes) -> output map
    std::map<std::pair<int, int>, istft_plan> istft_plans_;

    struct gentry { cudaGraphExec_t exec = nullptr; int seen = 0; unsigned long long kernels = 0, last_use = 0; };
    std::map<std::vector<int>, gentry> graphs_;      // (code counts..., output format) -> captured launch sequence
    bool graphs_ok_ = true;
    size_t graph_max_codes_ = 4096;                  // larger batches are not launch-bound
    unsigned long long graph_clock_ = 0;
    void drop_graphs();
    void run_direct(bool with_istft, stage_times * prof, int stop_stage);

    void ensure(dbuf & b, size_t bytes);
    void ensure_rope(int prenet_pos, int decoder_pos);
    const CUtensorMap & amap(const dbuf & b, int K);
    const CUtensorMap & omap(const void * p, int rows, int cols, int ld, int elem_bytes);
    void build_layout(seq_layout & L, const std::vector<int> & len, int gap);
    void append_layout(seq_layout & L);
    void bind_layout(seq_layout & L);
    size_t blob_add(const void * src, size_t bytes);

    void gemm(const dbuf & a, const dev_linear & w, int M, int epilogue, void * out, int ldo, int out_dtype,
              const float * bias, const float * res, int ldr, const float * gate, const seq_layout * lay,
              const float2 * rope = nullptr, int heads = 0, double * gn_stats = nullptr, int gn_groups = 0,
              const ln_fuse * ln = nullptr);
    bool run_tblock(float * x, const seq_layout & L, const tblock & blk, int dim, int heads, int ff, int window,
                    const float2 * rope, bool cond, bool ln1_done, const tblock * next);
    bool run_resblock(float *& x, float *& y, const seq_layout & L, const resblock & rb, int & gn_slot, bool in_ready,
                      bool feeds_next);
    void group_norm_silu(const float * x, const seq_layout & L, int channels, int groups, const dev_norm & n, int slot,
                         bool stats_ready);
    void flip();
    int dir_ = 0;
    double * gn_slot_ptr(int slot) { return (double *) gn_stats_.p + (size_t) slot * n_utt_ * gn_groups_max_ * 2; }
};

// pinned result buffers handed to callers by the C API and returned through mio_tts_audio_free
float * result_alloc(size_t n_floats);
bool result_free(void * p);   // false if p did not come from result_alloc

} // namespace mio
