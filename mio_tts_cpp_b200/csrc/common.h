// Shared host/device declarations for the B200 MioCodec decode + iSTFT path.
#pragma once
#include <cuda_runtime.h>
#include <cuda.h>
#include <atomic>
#include <cstdint>
#include <stdexcept>
#include <string>

namespace mio {

struct cuda_error : std::runtime_error {
    explicit cuda_error(const std::string & m) : std::runtime_error(m) {}
};

inline void cuda_check(cudaError_t e, const char * what, const char * file, int line) {
    if (e != cudaSuccess) {
        throw cuda_error(std::string("CUDA error: ") + cudaGetErrorString(e) + " (" + what + ") at " + file + ":" +
                         std::to_string(line));
    }
}
#define MIO_CUDA(x) ::mio::cuda_check((x), #x, __FILE__, __LINE__)

// every launch of one of OUR kernels bumps this (bench.py reports it as gpu_launches)
extern std::atomic<unsigned long long> g_launch_count;
inline void count_launch(int n = 1) { g_launch_count.fetch_add((unsigned long long) n, std::memory_order_relaxed); }

// ---------------------------------------------------------------------------------------------
// Programmatic dependent launch.  A decode step is ~140 short kernels back to back on one stream; each is launched
// with programmaticStreamSerialization so its CTAs are scheduled (and run their prologue: barrier init, TMEM
// allocation, index math) while the previous kernel drains.  Every kernel starts with MIO_PDL_PROLOGUE():
// launch_dependents lets the NEXT kernel be scheduled early, wait blocks until the PREVIOUS grid has completed and
// its writes are visible (completion is transitive because every kernel waits).  MIO_TTS_B200_PDL=0 disables it.
// ---------------------------------------------------------------------------------------------
bool pdl_enabled();
#ifdef __CUDACC__
#define MIO_PDL_TRIGGER() asm volatile("griddepcontrol.launch_dependents;" ::: "memory")
#define MIO_PDL_WAIT() asm volatile("griddepcontrol.wait;" ::: "memory")
#define MIO_PDL_PROLOGUE() do { MIO_PDL_TRIGGER(); MIO_PDL_WAIT(); } while (0)
template <typename... KArgs, typename... Args>
inline void launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args &&... args) {
    cudaLaunchConfig_t cfg;
    cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = pdl_enabled() ? 1 : 0;
    cfg.attrs = attr; cfg.numAttrs = 1;
    MIO_CUDA(cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...));
}
#endif

// ---------------------------------------------------------------------------------------------
// Ping-pong traversal.  Every kernel of the decode chain streams [rows, channels] activations that its predecessor has
// just written, and the working set of one stage (100-260 MB) is larger than the 126 MB L2: if both walk the rows in
// the same direction the consumer starts with the rows that left the L2 longest ago and misses all the way.  The
// engine therefore flips the direction from launch to launch: a kernel launched "reversed" maps its block / tile index
// b to (count - 1 - b), so it begins with the rows the previous kernel wrote LAST.  Results do not depend on the order
// (no kernel mixes rows).  traversal_reverse() is the flag the launchers read; MIO_TTS_B200_PINGPONG=0 pins it to 0.
// ---------------------------------------------------------------------------------------------
int traversal_reverse();
void set_traversal_reverse(int rev);
#define MIO_BLOCK_X(rev) ((rev) ? gridDim.x - 1 - blockIdx.x : blockIdx.x)

// 16-bit operand formats of the tensor-core GEMMs (tcgen05 kind::f16 a/b_format field)
enum op16 : int { OP_F16 = 0, OP_BF16 = 1 };
#ifdef __CUDACC__
// (lo, hi) -> packed fp16 pair, SATURATING: a value beyond +-65504 becomes +-65504 instead of inf (one F2FP.SATFINITE).
// fp16 operands are the default for the linears (10 mantissa bits, like the TF32 the reference's own CUDA path uses); a
// checkpoint with an activation outlier then clips one value instead of turning the utterance into NaNs.
__device__ __forceinline__ uint32_t pack_f16x2_sat(float lo, float hi) {
    uint32_t r;
    asm("cvt.rn.satfinite.f16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
    return r;
}
#endif

// ---------------------------------------------------------------------------------------------
// Row layouts.  Activations are [rows, channels] row-major (channel-contiguous).  A layout packs
// the utterances of a batch along rows; `gap` zero rows separate utterances (and precede the
// first / follow the last) so a k=3 convolution can read row+-1 without per-utterance branches.
// ---------------------------------------------------------------------------------------------
struct row_tile {        // a run of <= 64 rows inside ONE utterance (attention / statistics work unit)
    int row0;            // first row (layout coordinates)
    int n;               // rows in this tile
    int utt;             // utterance index
    int pos0;            // position of row0 inside the utterance
    int utt_row0;        // first row of the utterance
    int utt_len;         // rows of the utterance
};

// ---------------------------------------------------------------------------------------------
// GEMM (gemm_tcgen05.cu):  D[M,N] = A[M, taps*K] * W[N, taps*K]^T with fused epilogues
// ---------------------------------------------------------------------------------------------
enum gemm_epilogue : int {
    EPI_STORE32 = 0,   // out32[r,n] = (acc + bias[n]) * gate[utt(r),n] + res[r,n]   (each term optional)
    EPI_STORE16 = 1,   // out16[r,n] = acc + bias[n]
    EPI_SWIGLU16 = 2,  // out16[r,n/2] = silu(acc[r,2j]) * acc[r,2j+1]   (weights interleaved gate/up)
    EPI_ROPE16 = 3,    // out16[r,n] = rope(acc) for n < rope_cols (q,k); plain for v
    // EPI_STORE32 with the in-place residual, FOLLOWED BY THE LAYERNORM THAT CONSUMES IT, in one kernel (N == 512 only):
    //   x[r,:] = (acc + bias) * gate + x[r,:]  (fp32, map_out);  out16_2[r,:] = LN(x[r,:]) * A + B  (16-bit, map_out2)
    // A CTA pair owns whole rows (both 256-column accumulators of its TMEM), writes x back into TMEM, takes the row
    // statistics there and emits the next GEMM's A operand: the LayerNorm launch and its re-read of x disappear.
    EPI_STORE32_LN = 4,
};

struct gemm_args {
    CUtensorMap map_a;       // A: [rows, K] 16-bit, box 64 x 128, swizzle 128B
    CUtensorMap map_b;       // W: [N, taps*K] 16-bit, box 64 x gemm_w_box_rows(), swizzle 128B
    CUtensorMap map_b_tail;  // the same tensor with gemm_w_tail_box_rows()-row boxes: the column sub-tiles of a split last wave
    CUtensorMap map_out;     // output (and in-place residual): box 32 x 32, swizzle 128B (fp32) / 64B (16-bit)
    int M, N, K, taps;       // K = reduction length PER TAP; W column of (tap, k) = tap*K + k
    // A coordinates of tap t:  row = tile row + tap_row0 + (t >> tap_shift),  column = a_col0 + (t & ((1<<tap_shift)-1)) * K.
    //   plain GEMM: taps 1.  k=3 "same" conv: tap_row0 -1.  stride-2 conv over the PAIR view [rows/2, 2C] of the input:
    //   tap_shift 1 (taps 0,1 = the two frames of pair row r, tap 2 = first frame of pair row r+1).  grouped conv:
    //   a_col0 = first input channel of the group, tap_row0 = -padding.
    int tap_row0, tap_shift, a_col0;
    int act;                 // activation after bias: 0 none, 1 GELU (ggml: tanh form through fp16), 2 tanh,
                             // 3 SnakeBeta x + snake_ib[n] * sin^2(snake_a[n] * x) (EPI_STORE32 without gate / residual, EPI_STORE16)
    const float * snake_a; const float * snake_ib;   // per output column, padded to a multiple of 32 columns
    int dtype;               // op16 of A and W
    int out_dtype;           // op16 of a 16-bit output
    int epilogue;
    void * out; int ldo;
    const float * bias;
    const float * res; int ldr;          // fp32 residual: nullptr or == out (read through map_out)
    int red_add;                         // residual added by the L2 (TMA reduce-add store) instead of being read back: set by gemm_launch
    int split_from, split_shift;         // tiles >= split_from (the partial last wave) are cut into 1 << split_shift column sub-tiles: set by gemm_launch
    const float * gate; int gate_ld;     // gate[utt * gate_ld + n]  (gate_ld 0: one per-column scale for all rows)
    const int * row_utt;                 // row -> utterance (needed with gate_ld != 0)
    const int * row_pos;                 // row -> position in utterance (needed with rope)
    const float2 * rope;                 // [pos, head_dim/2] (cos, sin)
    int rope_cols, head_dim;             // columns [0, rope_cols) = q,k: rotated in (even, odd) pairs per head
    // EPI_STORE32 only: the GroupNorm that consumes this output gets its per-(utterance, group) sum / sum of squares
    // from the epilogue (double atomics on gn_stats[(utt * gn_groups + g) * 2 + {0,1}], gap rows excluded) instead of
    // a separate pass over the output.  Needs row_utt, N % 32 == 0 and N / gn_groups == 16 or a multiple of 32.
    double * gn_stats; int gn_groups;
    int reverse_m;                       // walk the row tiles from the last to the first (ping-pong traversal): set by gemm_launch
    // EPI_STORE32_LN: y = (x - mean) * rstd, out2 = y * (ln_plus1 + A[n]) + B[n] with A = ln_a + utt * ln_ld, B = ln_b + utt * ln_ld
    // (AdaLN: per-utterance scale / shift, ln_plus1 = 1; plain LayerNorm: ln_ld = 0, ln_plus1 = 0); gap rows (utt < 0) -> 0
    CUtensorMap map_out2;                // [rows, N] 16-bit, box 32 x 32, swizzle 64B
    const float * ln_a; const float * ln_b; int ln_ld; float ln_plus1; float ln_eps; int ln_out_dtype;
    int dbg;                             // timing experiments (MIO_TTS_B200_GEMM_DBG), 0 in production
    int ln_prefetch;                     // L1 prefetch of the item's per-column vectors (MIO_TTS_B200_LN_PREFETCH=0 turns it off): set by gemm_launch
};
inline bool gemm_gn_stats_ok(int N, int groups) {
    if (groups <= 0 || N % groups != 0 || N % 32 != 0) return false;
    const int cpg = N / groups;
    return cpg == 16 || cpg % 32 == 0;
}

void gemm_launch(const gemm_args & a, int num_sms, cudaStream_t st);
// A-operand / weight tensor maps (row-major [rows, cols] 16-bit, cols contiguous)
CUtensorMap make_tmap_16(const void * ptr, uint64_t rows, uint64_t cols, uint32_t box_rows);
// output tensor map: [rows, cols] of 2- or 4-byte elements, row pitch ld_elems (stores clip to rows x cols)
CUtensorMap make_tmap_out(const void * ptr, uint64_t rows, uint64_t cols, uint64_t ld_elems, int elem_bytes);
int gemm_block_n(int N);   // tile width the launcher will use for this N (weight rows are padded to it)
int gemm_w_box_rows();      // rows per TMA box of a weight tensor map (whole tiles)
int gemm_w_tail_box_rows(); // ... of its second map (split-tail sub-tiles)

// ---------------------------------------------------------------------------------------------
// elementwise / normalisation kernels (ops.cu)
// ---------------------------------------------------------------------------------------------
void launch_gather(const float * table, const int * codes, float * out, int n_rows, int dim, cudaStream_t st);
void launch_cond(const float * w, const float * b, const float * emb, float * out, int n_out, int adim, int n_utt,
                 cudaStream_t st);
// LayerNorm over channels; affine (w,b) or AdaLN (shift,scale per utterance at cond + utt*cond_ld + off)
void launch_layernorm(const float * x, int rows, int dim, float eps, const float * w, const float * b,
                      const float * cond, int cond_ld, int shift_off, int scale_off, const int * row_utt,
                      void * out16, int out_dtype, float * out32, cudaStream_t st);
void launch_cast16(const float * x, void * out16, int out_dtype, size_t n, cudaStream_t st);
// the same with rows that carry no data (row_utt < 0) written as zeros
void launch_cast16_rows(const float * x, void * out16, int out_dtype, int rows, int cols, const int * row_utt, cudaStream_t st);
void launch_gn_stats(const float * x, int dim, int groups, const row_tile * tiles, int n_tiles, double * stats,
                     cudaStream_t st);
// (sum, sumsq) -> (mean, rstd) per (utterance, group), then normalise + affine + SiLU -> 16-bit conv operand; gap rows
// of the layout are written as zeros
void launch_gn_apply(const float * x, int rows, int dim, int groups, const double * stats, const int * utt_len, float eps,
                     const float * w, const float * b, const int * row_utt, void * out16, int out_dtype, cudaStream_t st);
// ggml_interpolate(bilinear, width only) from the x2 transposed-conv output to the decoder frame layout
void launch_interp(const float * src, const int * src_row0, const int * src_len, float * dst, int rows, int dim,
                   const int * row_utt, const int * row_pos, const int * utt_len, cudaStream_t st);
void launch_rope_table(float2 * table, int n_pos, int half_dim, float theta_scale, cudaStream_t st);
// wave-upsampler: col2im of a transposed conv GEMM output + bias + crop + SnakeBeta
void launch_convt_fold_snake(const float * y, int ld_y, const int * src_row0, const int * src_len, float * dst,
                             int rows, int cout, int ksize, int stride, int pad, const float * bias,
                             const float * snake_a, const float * snake_inv_b, const int * row_utt,
                             const int * row_pos, cudaStream_t st);
void launch_snake(float * x, int rows, int dim, const float * snake_a, const float * snake_inv_b, cudaStream_t st);

// ---------------------------------------------------------------------------------------------
// banded attention (attention.cu): qkv16 [rows, 3*dim] (RoPE + scale already applied) -> out16 [rows, dim]
// ---------------------------------------------------------------------------------------------
void launch_attention(const void * qkv16, void * out16, int dtype, int dim, int heads, int window,
                      const row_tile * tiles, int n_tiles, bool force_generic, cudaStream_t st);
// persistent TMA-fed form (head_dim 64, window <= 65): map_qkv = make_tmap_16(qkv, rows, 3*dim, 64),
// map_out = make_tmap_16(out, rows, dim, 16); same tiles as launch_attention
bool attention_tma_ok(int dim, int heads, int window);
// tcgen05 form (attention_tc.cu), the default: 128-row tiles, maps = make_tmap_16(qkv, rows, 3*dim, 128 / 192) and
// make_tmap_16(out, rows, dim, 32).  MIO_TTS_B200_ATTN = v1 | v2 select the mma.sync kernels instead.
int attention_mode();
bool attention_tc_ok(int dim, int heads, int window);
void launch_attention_tc(const CUtensorMap & map_q, const CUtensorMap & map_kv, const CUtensorMap & map_out, void * out16,
                         int dtype, int dim, int heads, int window, int rows, const row_tile * tiles128, int n_tiles,
                         int num_sms, cudaStream_t st);
void launch_attention_tma(const CUtensorMap & map_qkv, const CUtensorMap & map_out, void * out16, int dtype, int dim, int heads,
                          int window, int rows, const row_tile * tiles, int n_tiles, int num_sms, cudaStream_t st);

// ---------------------------------------------------------------------------------------------
// iSTFT (istft.cu)
// ---------------------------------------------------------------------------------------------
struct istft_plan {
    int n_fft = 0, hop = 0;
    int n_radix = 0;
    int radix[16] = {0};
    float2 * d_tw_m = nullptr;    // exp(+2 pi i j / M), j < M = n_fft/2
    float2 * d_tw_n = nullptr;    // exp(+2 pi i k / n_fft), k <= M/2
    float * d_hann = nullptr;     // periodic Hann, host cosf (src/mio-tts-lib.cpp:415-420)
    size_t smem_bytes = 0;
};
bool istft_plan_init(istft_plan & p, int n_fft, int hop, std::string & err);
void istft_plan_free(istft_plan & p);
// spec: rows in a layout (utterance u: rows spec_row0[u] .. +n_frames[u]), pitch floats per row.
// audio: utterance u at audio + audio_off[u], n_frames[u]*hop samples (when n_fft-hop is even).
void launch_istft(const istft_plan & p, const float * spec, int pitch, const int * spec_row0, const int * n_frames,
                  const long long * audio_off, float * audio, unsigned int * peak_bits, int n_utt, int max_frames,
                  cudaStream_t st);
void launch_peak_normalize(float * audio, const long long * audio_off, const long long * audio_len,
                           const unsigned int * peak_bits, int n_utt, long long max_len, cudaStream_t st);
void launch_pcm16(const float * audio, const long long * audio_off, const long long * audio_len, const unsigned int * peak_bits,
                  int16_t * pcm, int n_utt, long long max_len, cudaStream_t st);

} // namespace mio
