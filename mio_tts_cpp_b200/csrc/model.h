// MioCodec decoder weights, resident in HBM in the layouts the sm_100a kernels consume.
//
// Mirrors miocodec_decoder::load (/root/reference/src/miocodec-decoder.cpp:392-822): same GGUF KVs, same tensor
// names, same required/optional rules and error strings.  The GGUF layout (scripts/convert_miocodec_to_gguf.py:
// 351-588) is kept unchanged; weights are re-tiled ONCE at load: linears -> 16-bit [N, K] K-major for TMA/tcgen05
// (QKV fused, SwiGLU gate/up row-interleaved), k=3 convs -> [Cout, 3*Cin] tap-major fp16, transposed convs ->
// [k*Cout, Cin], AdaLN condition linears concatenated (fp32, they are GEMVs over a 128-vector).
#pragma once
#include "common.h"

#include <string>
#include <vector>

namespace mio {

struct hparams {   // struct defaults: /root/reference/src/miocodec-decoder.h:12-49
    int model_type = 0;
    int sample_rate = 24000, n_fft = 1920, hop_length = 480, n_mels = 0, samples_per_token = 960;
    int prenet_layers = 6, prenet_dim = 768, prenet_heads = 12, prenet_ff = 2048, prenet_window = 65;
    int decoder_layers = 8, decoder_dim = 512, decoder_heads = 8, decoder_ff = 1536, decoder_window = 65;
    int decoder_adanorm_dim = 128;
    int resnet_blocks = 2, resnet_groups = 32;
    bool dynamic_global = false;
    float rope_theta = 10000.0f, norm_eps = 1e-5f, group_norm_eps = 1e-6f;
};

struct dev_linear {
    void * w16 = nullptr;       // [N, taps*K], K contiguous
    float * bias = nullptr;     // [N] or null
    int N = 0, K = 0, taps = 1;
    int dtype = OP_BF16;
    CUtensorMap map_b, map_b_tail;   // 128-row boxes (whole tiles) / 32-row boxes (split-tail sub-tiles)
};

struct dev_norm { float * w = nullptr; float * b = nullptr; };

struct tblock {
    dev_norm ln1, ln2;                      // affine LayerNorm (prenet / static decoder)
    int cond_attn_off = -1, cond_ffn_off = -1;   // AdaLN: offsets of [shift|scale|gate] in the condition vector
    dev_linear qkv, o, gate_up, down;
};

struct resblock {
    dev_norm n1, n2;
    dev_linear c1, c2;
    int channels = 0, groups = 1;
};

struct upstage {
    dev_linear up;              // transposed conv as GEMM: N = k*Cout, bias applied in the fold kernel
    float * up_bias = nullptr;
    float * snake_a = nullptr, * snake_inv_b = nullptr;
    resblock rb;
    int factor = 1, ksize = 1, pad = 0, cin = 0, cout = 0;
    // Polyphase form (set when ksize - 2 pad == factor and every tap lands on input row t-1, t or t+1): output row
    // t*factor + p of an utterance is a 3-tap "same" convolution of the INPUT rows with the kernel taps j = p + pad - s*factor
    // (s = -1, 0, +1), so the whole stage is ONE implicit GEMM [rows, 3*cin] x [factor*cout, 3*cin]^T whose output, seen as
    // [rows*factor, cout], is the upsampled sequence -- bias and SnakeBeta in the epilogue, no k*cout intermediate, no fold.
    bool poly = false;
    dev_linear up_poly;         // N = factor*cout, K = cin, taps = 3 (zero blocks where a phase has no tap)
    float * poly_bias = nullptr, * poly_snake_a = nullptr, * poly_snake_inv_b = nullptr;   // [factor*cout], padded to 32
};

class model {
public:
    ~model();
    bool load(const std::string & path, int device, std::string & err);

    hparams hp;
    int device = 0;
    int num_sms = 148;
    int linear_dtype = OP_BF16;   // operand type of the fp32-reference linears (MIO_TTS_B200_LINEAR_DTYPE)
    int conv_dtype = OP_F16;      // the reference's convs round operands to fp16

    float * tok_embd = nullptr; int n_codebook = 0;
    std::vector<tblock> prenet, decoder;
    dev_norm prenet_norm, decoder_norm;
    dev_linear prenet_out, upsample;
    std::vector<resblock> prior, post;
    float * cond_w = nullptr, * cond_b = nullptr; int cond_total = 0, cond_final_off = -1;
    std::vector<upstage> ups; int up_total = 1;
    float * out_snake_a_pad = nullptr, * out_snake_inv_b_pad = nullptr;   // out_snake vectors padded to 32 columns (GEMM epilogue)
    dev_linear up_out_proj; float * out_snake_a = nullptr, * out_snake_inv_b = nullptr;
    dev_linear head;
    int head_in = 0;              // channels feeding the iSTFT head
    istft_plan istft;
    size_t weight_bytes = 0;

private:
    std::vector<void *> allocs_;
    void * dalloc(size_t bytes);
    float * upload_f32(const float * src, size_t n);
    friend struct loader;
};

} // namespace mio
