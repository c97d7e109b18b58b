#include "model.h"

#include "gguf_io.h"

#include <cuda_bf16.h>
#include <cuda_fp16.h>

#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <cstring>

namespace mio {

model::~model() {
    istft_plan_free(istft);
    for (void * p : allocs_) cudaFree(p);
}

void * model::dalloc(size_t bytes) {
    void * p = nullptr;
    MIO_CUDA(cudaMalloc(&p, std::max<size_t>(bytes, 16)));
    allocs_.push_back(p);
    weight_bytes += bytes;
    return p;
}

float * model::upload_f32(const float * src, size_t n) {
    float * d = (float *) dalloc(n * sizeof(float));
    MIO_CUDA(cudaMemcpy(d, src, n * sizeof(float), cudaMemcpyHostToDevice));
    return d;
}

struct loader {
    model & m;
    const gguf_file & g;
    std::string & err;

    static uint16_t to16(float v, int dtype) {
        if (dtype == OP_BF16) { __nv_bfloat16 h = __float2bfloat16_rn(v); uint16_t u; std::memcpy(&u, &h, 2); return u; }
        __half h = __float2half_rn(v);
        uint16_t u; std::memcpy(&u, &h, 2); return u;
    }

    // tensor as fp32 host vector (F32, or F16 widened); empty + err on failure
    bool get(const std::string & name, std::vector<float> & out, size_t expect, bool required = true) {
        const gguf_tensor * t = g.find(name);
        if (t == nullptr) {
            if (required && err.empty()) err = "missing tensor: " + name;
            return false;
        }
        const size_t n = (size_t) t->nelements();
        if (expect != 0 && n != expect) { if (err.empty()) err = "unexpected shape for tensor: " + name; return false; }
        out.resize(n);
        if (t->type == GGUF_F32) {
            std::memcpy(out.data(), t->data, n * sizeof(float));
        } else if (t->type == GGUF_F16) {
            const __half * h = reinterpret_cast<const __half *>(t->data);
            for (size_t i = 0; i < n; ++i) out[i] = __half2float(h[i]);
        } else {
            if (err.empty()) err = "unsupported tensor type for: " + name;
            return false;
        }
        return true;
    }

    bool norm(const std::string & prefix, int dim, dev_norm & out, bool required = true) {
        std::vector<float> w, b;
        if (!get(prefix + ".weight", w, (size_t) dim, required) || !get(prefix + ".bias", b, (size_t) dim, required)) return false;
        out.w = m.upload_f32(w.data(), w.size());
        out.b = m.upload_f32(b.data(), b.size());
        return true;
    }

    // rows: host fp32 [N, taps*K] already in GEMM order
    void make_linear(dev_linear & L, const std::vector<float> & rows, int N, int K, int taps, int dtype,
                     const std::vector<float> * bias) {
        L.N = N; L.K = K; L.taps = taps; L.dtype = dtype;
        std::vector<uint16_t> h(rows.size());
        for (size_t i = 0; i < rows.size(); ++i) h[i] = to16(rows[i], dtype);
        // pad the row count to the GEMM tile so the TMA box never exceeds the allocation
        const int bn = gemm_block_n(N);
        const size_t n_pad = (size_t) ((N + bn - 1) / bn) * bn;
        L.w16 = m.dalloc(n_pad * (size_t) taps * K * 2);
        MIO_CUDA(cudaMemset(L.w16, 0, n_pad * (size_t) taps * K * 2));
        MIO_CUDA(cudaMemcpy(L.w16, h.data(), h.size() * 2, cudaMemcpyHostToDevice));
        L.map_b = make_tmap_16(L.w16, (uint64_t) N, (uint64_t) taps * K, (uint32_t) gemm_w_box_rows());
        L.map_b_tail = make_tmap_16(L.w16, (uint64_t) N, (uint64_t) taps * K, (uint32_t) gemm_w_tail_box_rows());
        if (bias) {
            std::vector<float> bp(((size_t) N + 3) / 4 * 4, 0.0f);
            std::copy(bias->begin(), bias->end(), bp.begin());
            L.bias = m.upload_f32(bp.data(), bp.size());
        }
    }

    bool linear(const std::string & wname, const std::string & bname, int N, int K, dev_linear & L) {
        std::vector<float> w, b;
        if (!get(wname, w, (size_t) N * K)) return false;
        if (!bname.empty() && !get(bname, b, (size_t) N)) return false;
        make_linear(L, w, N, K, 1, m.linear_dtype, bname.empty() ? nullptr : &b);
        return true;
    }

    bool tblock_load(const std::string & p, int dim, int ff, int heads, tblock & blk) {
        std::vector<float> q, k, v, o, w1, w2, w3;
        if (!get(p + "attn_q.weight", q, (size_t) dim * dim) || !get(p + "attn_k.weight", k, (size_t) dim * dim) ||
            !get(p + "attn_v.weight", v, (size_t) dim * dim) || !get(p + "attn_output.weight", o, (size_t) dim * dim) ||
            !get(p + "ffn_gate.weight", w1, (size_t) ff * dim) || !get(p + "ffn_down.weight", w2, (size_t) dim * ff) ||
            !get(p + "ffn_up.weight", w3, (size_t) ff * dim))
            return false;
        // softmax(q k^T * 1/sqrt(head_dim)) (src/miocodec-decoder.cpp:1251-1257): the scale is folded into Wq
        // (exact for head_dim = 64: a power of two), so the QKV epilogue only rotates
        const float q_scale = 1.0f / sqrtf((float) (dim / std::max(1, heads)));
        for (float & x : q) x *= q_scale;
        std::vector<float> qkv;
        qkv.reserve(3 * q.size());
        qkv.insert(qkv.end(), q.begin(), q.end());
        qkv.insert(qkv.end(), k.begin(), k.end());
        qkv.insert(qkv.end(), v.begin(), v.end());
        make_linear(blk.qkv, qkv, 3 * dim, dim, 1, m.linear_dtype, nullptr);
        make_linear(blk.o, o, dim, dim, 1, m.linear_dtype, nullptr);
        std::vector<float> gu((size_t) 2 * ff * dim);   // row 2j = gate_j (w1), row 2j+1 = up_j (w3)
        for (int j = 0; j < ff; ++j) {
            std::memcpy(&gu[(size_t) (2 * j) * dim], &w1[(size_t) j * dim], dim * sizeof(float));
            std::memcpy(&gu[(size_t) (2 * j + 1) * dim], &w3[(size_t) j * dim], dim * sizeof(float));
        }
        make_linear(blk.gate_up, gu, 2 * ff, dim, 1, m.linear_dtype, nullptr);
        make_linear(blk.down, w2, dim, ff, 1, m.linear_dtype, nullptr);
        return true;
    }

    // torch Conv1d weight [Cout, Cin, 3] -> [Cout, tap*Cin + ci]
    bool conv3(const std::string & prefix, int c, dev_linear & L) {
        std::vector<float> w, b;
        if (!get(prefix + ".weight", w, 0) || !get(prefix + ".bias", b, (size_t) c)) return false;
        if (w.size() % ((size_t) c * c) != 0) { err = "unexpected shape for tensor: " + prefix + ".weight"; return false; }
        const int kk = (int) (w.size() / ((size_t) c * c));
        if (kk != 3) { err = "unsupported conv kernel size (expected 3): " + prefix + ".weight"; return false; }
        std::vector<float> r((size_t) c * 3 * c);
        for (int co = 0; co < c; ++co)
            for (int ci = 0; ci < c; ++ci)
                for (int t = 0; t < 3; ++t) r[(size_t) co * 3 * c + (size_t) t * c + ci] = w[((size_t) co * c + ci) * 3 + t];
        make_linear(L, r, c, c, 3, m.conv_dtype, &b);
        return true;
    }

    static int choose_groups(int resnet_groups, int channels) {   // src/miocodec-decoder.cpp:1345-1351
        int g = std::max(1, std::min(resnet_groups, channels));
        while (g > 1 && channels % g != 0) --g;
        return std::max(1, g);
    }

    bool resblock_load(const std::string & prefix, int c, resblock & rb) {
        rb.channels = c;
        rb.groups = choose_groups(m.hp.resnet_groups, c);
        return norm(prefix + ".norm1", c, rb.n1) && conv3(prefix + ".conv1", c, rb.c1) &&
               norm(prefix + ".norm2", c, rb.n2) && conv3(prefix + ".conv2", c, rb.c2);
    }

    // torch ConvTranspose1d weight [Cin, Cout, k] -> GEMM rows [kk*Cout + co][ci]
    bool convt(const std::string & prefix, int cin, int k_expected, int dtype, dev_linear & L, std::vector<float> & bias,
               int & cout, int & ksize) {
        const gguf_tensor * t = g.find(prefix + ".weight");
        if (t == nullptr) { err = "missing tensor: " + prefix + ".weight"; return false; }
        if (t->ne.size() != 3 || (int) t->ne[2] != cin) { err = "unexpected shape for tensor: " + prefix + ".weight"; return false; }
        ksize = (int) t->ne[0];
        cout = (int) t->ne[1];
        if (k_expected > 0 && ksize != k_expected) { err = "unexpected kernel size for tensor: " + prefix + ".weight"; return false; }
        std::vector<float> w;
        if (!get(prefix + ".weight", w, (size_t) cin * cout * ksize) || !get(prefix + ".bias", bias, (size_t) cout)) return false;
        std::vector<float> r((size_t) ksize * cout * cin);
        for (int ci = 0; ci < cin; ++ci)
            for (int co = 0; co < cout; ++co)
                for (int kk = 0; kk < ksize; ++kk)
                    r[((size_t) kk * cout + co) * cin + ci] = w[((size_t) ci * cout + co) * ksize + kk];
        make_linear(L, r, ksize * cout, cin, 1, dtype, nullptr);
        w_last_convt = std::move(w);
        return true;
    }
    std::vector<float> w_last_convt;   // [cin][cout][k] of the transposed convolution loaded last (polyphase repack)

    // Polyphase weights of a transposed convolution (model.h, upstage): out[t*f + p] = sum_j x[t + (p + pad - j)/f] W[:, :, j]
    // over the taps j with (p + pad - j) divisible by f.  Returns false when a tap falls outside rows t-1 .. t+1 or the
    // output is not exactly f rows per input row.
    bool convt_poly(upstage & st, const std::vector<float> & bias, const std::vector<float> & al, const std::vector<float> & ib,
                    int dtype) {
        const int f = st.factor, k = st.ksize, pad = st.pad, cin = st.cin, cout = st.cout;
        if (k - 2 * pad != f || (cin % 64) != 0 || ((f * cout) % 8) != 0) return false;
        for (int p = 0; p < f; ++p)
            for (int j = 0; j < k; ++j)
                if ((p + pad - j) % f == 0 && std::abs((p + pad - j) / f) > 1) return false;
        const int N = f * cout;
        std::vector<float> r((size_t) N * 3 * cin, 0.0f);
        for (int p = 0; p < f; ++p)
            for (int j = 0; j < k; ++j) {
                if ((p + pad - j) % f != 0) continue;
                const int tap = (p + pad - j) / f + 1;          // GEMM tap 0 / 1 / 2 reads input row t-1 / t / t+1
                for (int co = 0; co < cout; ++co)
                    for (int ci = 0; ci < cin; ++ci)
                        r[((size_t) (p * cout + co) * 3 + tap) * cin + ci] = w_last_convt[((size_t) ci * cout + co) * k + j];
            }
        make_linear(st.up_poly, r, N, cin, 3, dtype, nullptr);
        const size_t np = ((size_t) N + 31) / 32 * 32;
        std::vector<float> b(np, 0.0f), a(np, 0.0f), i2(np, 0.0f);
        for (int n = 0; n < N; ++n) { b[n] = bias[n % cout]; a[n] = al[n % cout]; i2[n] = ib[n % cout]; }
        st.poly_bias = m.upload_f32(b.data(), b.size());
        st.poly_snake_a = m.upload_f32(a.data(), a.size());
        st.poly_snake_inv_b = m.upload_f32(i2.data(), i2.size());
        st.poly = true;
        return true;
    }

    std::vector<float> snake_al, snake_ib;   // host copies of the vectors loaded last (exp(alpha), 1 / (exp(beta) + 1e-9))
    bool snake(const std::string & prefix, int c, float *& a, float *& inv_b) {
        std::vector<float> al, be;
        if (!get(prefix + ".alpha", al, (size_t) c) || !get(prefix + ".beta", be, (size_t) c)) return false;
        // snake_beta_apply: x + sin^2(exp(alpha) x) / (exp(beta) + 1e-9), src/miocodec-decoder.cpp:1332-1343
        for (int i = 0; i < c; ++i) { al[i] = expf(al[i]); be[i] = 1.0f / (expf(be[i]) + 1e-9f); }
        a = m.upload_f32(al.data(), al.size());
        inv_b = m.upload_f32(be.data(), be.size());
        snake_al = al; snake_ib = be;
        return true;
    }
};

bool model::load(const std::string & path, int dev, std::string & err) {
    device = dev;
    MIO_CUDA(cudaSetDevice(device));
    cudaDeviceProp prop;
    MIO_CUDA(cudaGetDeviceProperties(&prop, device));
    if (prop.major != 10) {
        err = "this build runs on sm_100a (B200) only; device reports sm_" + std::to_string(prop.major * 10 + prop.minor);
        return false;
    }
    num_sms = prop.multiProcessorCount;
    // fp16 operands for the linears by default (saturating conversions, common.h): same tensor rate as bf16, 17 dB closer
    // to the reference's fp32 output at 400 codes.  MIO_TTS_B200_LINEAR_DTYPE=bf16 trades the 3 extra
    // mantissa bits for range.
    linear_dtype = OP_F16;
    if (const char * e = std::getenv("MIO_TTS_B200_LINEAR_DTYPE")) {
        if (std::strcmp(e, "bf16") == 0) linear_dtype = OP_BF16;
        else if (std::strcmp(e, "f16") != 0 && std::strcmp(e, "fp16") != 0) { err = "MIO_TTS_B200_LINEAR_DTYPE must be f16 or bf16"; return false; }
    }

    gguf_file g;
    if (!g.load(path, err)) return false;

    uint32_t u = 0;
    float f = 0.0f;
    // KV reads: /root/reference/src/miocodec-decoder.cpp:458-497
    if (g.get_u32("miocodec.model_type", u)) hp.model_type = (int) u;
    if (g.get_u32("miocodec.sample_rate", u)) hp.sample_rate = (int) u;
    if (g.get_u32("miocodec.n_fft", u)) hp.n_fft = (int) u;
    if (g.get_u32("miocodec.hop_length", u)) hp.hop_length = (int) u;
    if (g.get_u32("miocodec.n_mels", u)) hp.n_mels = (int) u;
    if (g.get_u32("miocodec.samples_per_token", u)) hp.samples_per_token = (int) u;
    if (g.get_u32("miocodec.prenet_layers", u)) hp.prenet_layers = (int) u;
    if (g.get_u32("miocodec.prenet_dim", u)) hp.prenet_dim = (int) u;
    if (g.get_u32("miocodec.prenet_heads", u)) hp.prenet_heads = (int) u;
    if (g.get_u32("miocodec.prenet_ff", u)) hp.prenet_ff = (int) u;
    if (g.get_u32("miocodec.prenet_window", u)) hp.prenet_window = (int) u;
    if (g.get_u32("miocodec.decoder_layers", u)) hp.decoder_layers = (int) u;
    if (g.get_u32("miocodec.decoder_dim", u)) hp.decoder_dim = (int) u;
    if (g.get_u32("miocodec.decoder_heads", u)) hp.decoder_heads = (int) u;
    if (g.get_u32("miocodec.decoder_ff", u)) hp.decoder_ff = (int) u;
    if (g.get_u32("miocodec.decoder_window", u)) hp.decoder_window = (int) u;
    if (g.get_u32("miocodec.decoder_adanorm_dim", u)) hp.decoder_adanorm_dim = (int) u;
    if (g.get_u32("miocodec.resnet_blocks", u)) hp.resnet_blocks = (int) u;
    if (g.get_u32("miocodec.resnet_groups", u)) hp.resnet_groups = (int) u;
    int up_layers = 0;
    if (g.get_u32("miocodec.wave_upsampler_layers", u)) up_layers = (int) u;
    if (g.get_u32("miocodec.dynamic_global", u)) hp.dynamic_global = u != 0;
    if (g.get_f32("miocodec.rope_theta", f)) hp.rope_theta = f;
    if (g.get_f32("miocodec.norm_eps", f)) hp.norm_eps = f;
    if (g.get_f32("miocodec.group_norm_eps", f)) hp.group_norm_eps = f;

    if (hp.model_type != 0) {
        // decode_mel_to_audio (mel-mode MioVocoder) is outside this path: SURVEY.md section 2 row 3
        err = "mel-mode MioCodec (model_type != 0) is not supported by the B200 wave/iSTFT path";
        return false;
    }
    if (hp.prenet_heads <= 0 || hp.decoder_heads <= 0 || hp.prenet_dim % hp.prenet_heads || hp.decoder_dim % hp.decoder_heads ||
        ((hp.prenet_dim / hp.prenet_heads) & 1) || ((hp.decoder_dim / hp.decoder_heads) & 1) || hp.hop_length <= 0 ||
        hp.samples_per_token <= 0 || hp.prenet_layers < 0 || hp.decoder_layers < 0) {
        err = "invalid MioCodec hyper-parameters";
        return false;
    }
    if ((hp.prenet_dim % 8) || (hp.decoder_dim % 8) || (hp.prenet_ff % 8) || (hp.decoder_ff % 8)) {
        err = "MioCodec dims must be multiples of 8 for the tensor-core path";
        return false;
    }

    loader L{*this, g, err};
    const int pd = hp.prenet_dim, dd = hp.decoder_dim, ad = hp.decoder_adanorm_dim;

    {   // token_embd [n_codebook, prenet_dim]
        const gguf_tensor * t = g.find("token_embd");
        if (t == nullptr) { err = "missing tensor: token_embd"; return false; }
        if (t->ne.size() != 2 || (int) t->ne[0] != pd) { err = "unexpected shape for tensor: token_embd"; return false; }
        std::vector<float> e;
        if (!L.get("token_embd", e, 0)) return false;
        n_codebook = (int) t->ne[1];
        tok_embd = upload_f32(e.data(), e.size());
    }

    prenet.resize((size_t) hp.prenet_layers);
    for (int i = 0; i < hp.prenet_layers; ++i) {
        const std::string p = "wave_prenet.blk." + std::to_string(i) + ".";
        if (!L.norm(p + "attn_norm", pd, prenet[i].ln1) || !L.norm(p + "ffn_norm", pd, prenet[i].ln2) ||
            !L.tblock_load(p, pd, hp.prenet_ff, hp.prenet_heads, prenet[i]))
            return false;
    }
    if (!L.norm("wave_prenet.norm", pd, prenet_norm)) return false;
    if (!L.linear("wave_prenet.output.weight", "wave_prenet.output.bias", dd, pd, prenet_out)) return false;

    {   // wave_upsample: ConvTranspose1d k=2, s=2 (stride hard-coded at src/miocodec-decoder.cpp:1370), F32 kernel
        std::vector<float> b;
        int cout = 0, ks = 0;
        if (!L.convt("wave_upsample", dd, 2, linear_dtype, upsample, b, cout, ks)) return false;
        if (cout != dd) { err = "unexpected shape for tensor: wave_upsample.weight"; return false; }
        std::vector<float> b2((size_t) 2 * dd);
        for (int kk = 0; kk < 2; ++kk) std::copy(b.begin(), b.end(), b2.begin() + (size_t) kk * dd);
        upsample.bias = upload_f32(b2.data(), b2.size());
    }

    prior.resize((size_t) hp.resnet_blocks);
    post.resize((size_t) hp.resnet_blocks);
    for (int i = 0; i < hp.resnet_blocks; ++i) {
        if (!L.resblock_load("wave_prior." + std::to_string(i), dd, prior[i])) return false;
        if (!L.resblock_load("wave_post." + std::to_string(i), dd, post[i])) return false;
    }

    // decoder blocks + AdaLN condition linears (dynamic vs static tensors: :662-711)
    std::vector<float> cw, cb;
    auto add_cond = [&](const std::string & name, int n_out, int & off) -> bool {
        std::vector<float> w, b;
        const bool ok = L.get(name + ".weight", w, (size_t) n_out * ad, false) && L.get(name + ".bias", b, (size_t) n_out, false);
        if (!ok) return false;
        off = (int) cb.size();
        cw.insert(cw.end(), w.begin(), w.end());
        cb.insert(cb.end(), b.begin(), b.end());
        return true;
    };
    decoder.resize((size_t) hp.decoder_layers);
    for (int i = 0; i < hp.decoder_layers; ++i) {
        const std::string p = "wave_decoder.blk." + std::to_string(i) + ".";
        if (!L.tblock_load(p, dd, hp.decoder_ff, hp.decoder_heads, decoder[i])) return false;
        if (hp.dynamic_global) {
            if (!add_cond(p + "attn_cond", 3 * dd, decoder[i].cond_attn_off) ||
                !add_cond(p + "ffn_cond", 3 * dd, decoder[i].cond_ffn_off)) {
                if (err.empty()) err = "dynamic_global=1 but decoder conditioning tensors are missing";
                return false;
            }
        } else {
            if (!L.norm(p + "attn_norm", dd, decoder[i].ln1, false) || !L.norm(p + "ffn_norm", dd, decoder[i].ln2, false)) {
                if (err.empty()) err = "static decoder tensors are missing";
                return false;
            }
        }
    }
    if (hp.dynamic_global) {
        if (!add_cond("wave_decoder.norm_cond", 2 * dd, cond_final_off)) {
            if (err.empty()) err = "dynamic_global=1 but decoder norm conditioning tensors are missing";
            return false;
        }
        cond_total = (int) cb.size();
        cond_w = upload_f32(cw.data(), cw.size());
        cond_b = upload_f32(cb.data(), cb.size());
    } else if (!L.norm("wave_decoder.norm", dd, decoder_norm, false)) {
        if (err.empty()) err = "static decoder norm tensors are missing";
        return false;
    }

    // optional wave upsampler (:571-638)
    head_in = dd;
    up_total = 1;
    if (up_layers > 0) {
        const gguf_tensor * ft = g.find("miocodec.wave_upsampler.factors");
        const gguf_tensor * kt = g.find("miocodec.wave_upsampler.kernel_sizes");
        if (ft == nullptr) { err = "missing tensor: miocodec.wave_upsampler.factors"; return false; }
        if (kt == nullptr) { err = "missing tensor: miocodec.wave_upsampler.kernel_sizes"; return false; }
        if (ft->type != GGUF_I32 || kt->type != GGUF_I32) { err = "wave upsampler factors/kernel sizes must be i32 tensors"; return false; }
        if ((int) ft->nelements() != up_layers || (int) kt->nelements() != up_layers) { err = "wave upsampler tensor size mismatch"; return false; }
        const int32_t * fv = reinterpret_cast<const int32_t *>(ft->data);
        const int32_t * kv = reinterpret_cast<const int32_t *>(kt->data);
        ups.resize((size_t) up_layers);
        int cin = dd;
        for (int i = 0; i < up_layers; ++i) {
            upstage & st = ups[(size_t) i];
            st.factor = fv[i];
            st.ksize = kv[i];
            if (st.factor <= 0 || st.ksize <= 0) { err = "wave upsampler factor/kernel must be positive"; return false; }
            up_total *= st.factor;
            st.pad = std::max(0, (st.ksize - st.factor) / 2);
            st.cin = cin;
            std::vector<float> b;
            int ks = 0;
            const std::string si = std::to_string(i);
            // the reference casts this kernel to F16 and the F16 op rounds the input to F16 (:1418-1421)
            if (!L.convt("wave_upsampler.up." + si, cin, st.ksize, conv_dtype, st.up, b, st.cout, ks)) return false;
            if (st.cout % 4) { err = "wave upsampler channel counts must be multiples of 4"; return false; }
            st.up_bias = upload_f32(b.data(), b.size());
            if (!L.snake("wave_upsampler.snake." + si, st.cout, st.snake_a, st.snake_inv_b)) return false;
            static const bool poly_env = [] { const char * e = std::getenv("MIO_TTS_B200_CONVT_POLY"); return !(e && e[0] == '0'); }();
            if (poly_env) L.convt_poly(st, b, L.snake_al, L.snake_ib, conv_dtype);
            if (!L.resblock_load("wave_upsampler.resblk." + si, st.cout, st.rb)) return false;
            cin = st.cout;
        }
        const gguf_tensor * pw = g.find("wave_upsampler.out_proj.weight");
        if (pw == nullptr) { err = "missing tensor: wave_upsampler.out_proj.weight"; return false; }
        if (pw->ne.size() != 2 || (int) pw->ne[0] != cin) { err = "unexpected shape for tensor: wave_upsampler.out_proj.weight"; return false; }
        const int out_dim = (int) pw->ne[1];
        if (out_dim % 8) { err = "wave upsampler out_proj width must be a multiple of 8"; return false; }
        if (!L.linear("wave_upsampler.out_proj.weight", "wave_upsampler.out_proj.bias", out_dim, cin, up_out_proj)) return false;
        if (!L.snake("wave_upsampler.out_snake", out_dim, out_snake_a, out_snake_inv_b)) return false;
        {
            std::vector<float> a2(((size_t) out_dim + 31) / 32 * 32, 0.0f), i2(a2.size(), 0.0f);
            std::copy(L.snake_al.begin(), L.snake_al.end(), a2.begin());
            std::copy(L.snake_ib.begin(), L.snake_ib.end(), i2.begin());
            out_snake_a_pad = upload_f32(a2.data(), a2.size());
            out_snake_inv_b_pad = upload_f32(i2.data(), i2.size());
        }
        head_in = out_dim;
    }

    if (!L.linear("istft_head.out.weight", "istft_head.out.bias", hp.n_fft + 2, head_in, head)) return false;

    if (!istft_plan_init(istft, hp.n_fft, hp.hop_length, err)) return false;
    MIO_CUDA(cudaDeviceSynchronize());
    return true;
}

} // namespace mio
