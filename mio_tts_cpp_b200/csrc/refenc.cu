#include "refenc.h"

#include "gguf_io.h"

#include <cuda_bf16.h>
#include <cuda_fp16.h>

#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <cstring>

namespace mio {

namespace {
size_t round_up(size_t v, size_t a) { return (v + a - 1) / a * a; }
constexpr int GAP = 64;      // zero rows between clips in the padded layout (>= pos-conv padding 64, embed-conv padding)

uint16_t to16(float v, int dtype) {
    if (dtype == OP_BF16) { __nv_bfloat16 h = __float2bfloat16_rn(v); uint16_t u; std::memcpy(&u, &h, 2); return u; }
    __half h = __float2half_rn(v);
    uint16_t u; std::memcpy(&u, &h, 2); return u;
}
} // namespace

// -------------------------------------------------------------------------------------------------
// weights
// -------------------------------------------------------------------------------------------------
refmodel::~refmodel() { for (void * p : allocs_) cudaFree(p); }

void * refmodel::dalloc(size_t bytes) {
    void * p = nullptr;
    MIO_CUDA(cudaMalloc(&p, std::max<size_t>(bytes, 16)));
    allocs_.push_back(p);
    weight_bytes += bytes;
    return p;
}

float * refmodel::upload_f32(const float * src, size_t n) {
    float * d = (float *) dalloc(std::max<size_t>(n, 4) * sizeof(float));
    MIO_CUDA(cudaMemcpy(d, src, n * sizeof(float), cudaMemcpyHostToDevice));
    return d;
}

// rows: host fp32 [N, taps*K] in GEMM order
void refmodel::make_linear(dev_linear & L, const std::vector<float> & rows, int N, int K, int taps, int dtype,
                           const std::vector<float> * bias) {
    if (K % 64) throw cuda_error("reference path: GEMM reduction lengths must be multiples of 64");
    L.N = N; L.K = K; L.taps = taps; L.dtype = dtype;
    std::vector<uint16_t> h(rows.size());
    for (size_t i = 0; i < rows.size(); ++i) h[i] = to16(rows[i], dtype);
    const int bn = gemm_block_n(N);
    const size_t n_pad = (size_t) ((N + bn - 1) / bn) * bn;
    L.w16 = dalloc(n_pad * (size_t) taps * K * 2);
    MIO_CUDA(cudaMemset(L.w16, 0, n_pad * (size_t) taps * K * 2));
    MIO_CUDA(cudaMemcpy(L.w16, h.data(), h.size() * 2, cudaMemcpyHostToDevice));
    L.map_b = make_tmap_16(L.w16, (uint64_t) N, (uint64_t) taps * K, (uint32_t) gemm_w_box_rows());
    L.map_b_tail = make_tmap_16(L.w16, (uint64_t) N, (uint64_t) taps * K, (uint32_t) gemm_w_tail_box_rows());
    if (bias) {
        std::vector<float> bp(((size_t) N + 3) / 4 * 4, 0.0f);
        std::copy(bias->begin(), bias->end(), bp.begin());
        L.bias = upload_f32(bp.data(), bp.size());
    }
}

struct refloader {
    refmodel & m;
    const gguf_file & g;
    std::string & err;

    bool get(const std::string & name, std::vector<float> & out, size_t expect) {
        const gguf_tensor * t = g.find(name);
        if (t == nullptr) { if (err.empty()) err = "missing tensor: " + name; return false; }
        const size_t n = (size_t) t->nelements();
        if (expect != 0 && n != expect) { if (err.empty()) err = "unexpected shape for tensor: " + name; return false; }
        out.resize(n);
        if (t->type == GGUF_F32) std::memcpy(out.data(), t->data, n * sizeof(float));
        else if (t->type == GGUF_F16) {
            const __half * h = reinterpret_cast<const __half *>(t->data);
            for (size_t i = 0; i < n; ++i) out[i] = __half2float(h[i]);
        } else { if (err.empty()) err = "unsupported tensor type for: " + name; return false; }
        return true;
    }
    bool vec(const std::string & name, int n, float *& out) {
        std::vector<float> v;
        if (!get(name, v, (size_t) n)) return false;
        out = m.upload_f32(v.data(), v.size());
        return true;
    }
    bool norm(const std::string & prefix, int dim, dev_norm & out) {
        return vec(prefix + ".weight", dim, out.w) && vec(prefix + ".bias", dim, out.b);
    }
    bool linear(const std::string & prefix, int N, int K, int dtype, dev_linear & L, float row_scale_first = 1.0f, int scaled_rows = 0) {
        std::vector<float> w, b;
        if (!get(prefix + ".weight", w, (size_t) N * K) || !get(prefix + ".bias", b, (size_t) N)) return false;
        for (int r = 0; r < scaled_rows; ++r) {
            for (int k = 0; k < K; ++k) w[(size_t) r * K + k] *= row_scale_first;
            b[(size_t) r] *= row_scale_first;
        }
        m.make_linear(L, w, N, K, 1, dtype, &b);
        return true;
    }
    // torch Conv1d weight [Cout, Cin, k] (ggml ne = [k, Cin, Cout]) -> GEMM rows [Cout][tap*Cin + ci]; both operands of
    // the reference's ggml_conv_1d are rounded to fp16 (ggml.c:4409-4426)
    bool conv(const std::string & wname, const std::string & bname, int & cout, int & cin, int & k, dev_linear & L) {
        const gguf_tensor * t = g.find(wname);
        if (t == nullptr) { if (err.empty()) err = "missing tensor: " + wname; return false; }
        if (t->ne.size() != 3) { if (err.empty()) err = "unexpected shape for tensor: " + wname; return false; }
        k = (int) t->ne[0]; cin = (int) t->ne[1]; cout = (int) t->ne[2];
        std::vector<float> w, b;
        if (!get(wname, w, 0)) return false;
        if (!bname.empty() && !get(bname, b, (size_t) cout)) return false;
        std::vector<float> rows((size_t) cout * k * cin);
        for (int o = 0; o < cout; ++o)
            for (int ci = 0; ci < cin; ++ci)
                for (int j = 0; j < k; ++j)
                    rows[((size_t) o * k + j) * cin + ci] = w[((size_t) o * cin + ci) * k + j];
        m.make_linear(L, rows, cout, cin, k, m.conv_dtype, bname.empty() ? nullptr : &b);
        return true;
    }
};

bool refmodel::load_wavlm(const std::string & path, int dev, std::string & err) {
    device = dev;
    MIO_CUDA(cudaSetDevice(device));
    cudaDeviceProp prop;
    MIO_CUDA(cudaGetDeviceProperties(&prop, device));
    num_sms = prop.multiProcessorCount;
    gguf_file g;
    if (!g.load(path, err)) { if (err.empty()) err = "failed to load GGUF"; return false; }
    uint32_t u = 0; float f = 0.0f;
    if (g.get_u32("wavlm.sample_rate", u)) hp.sample_rate = (int) u;
    if (g.get_u32("wavlm.n_layers", u)) hp.n_layers = (int) u;
    if (g.get_u32("wavlm.n_heads", u)) hp.n_heads = (int) u;
    if (g.get_u32("wavlm.head_dim", u)) hp.head_dim = (int) u;
    if (g.get_u32("wavlm.embed_dim", u)) hp.embed_dim = (int) u;
    if (g.get_u32("wavlm.num_buckets", u)) hp.num_buckets = (int) u;
    if (g.get_u32("wavlm.max_distance", u)) hp.max_distance = (int) u;
    if (g.get_f32("wavlm.layer_norm_eps", f)) hp.norm_eps = f;
    for (int i = 0; i < 7; ++i) {
        if (g.get_u32("wavlm.feat.conv" + std::to_string(i) + ".kernel", u)) hp.conv_kernel[i] = (int) u;
        if (g.get_u32("wavlm.feat.conv" + std::to_string(i) + ".stride", u)) hp.conv_stride[i] = (int) u;
    }
    const int E = hp.embed_dim, H = hp.n_heads, hd = hp.head_dim;
    if (E != H * hd || hd != 64) { err = "WavLM: the B200 path needs head_dim 64 and embed_dim = n_heads * head_dim"; return false; }
    for (int i = 1; i < 7; ++i) {
        const int k = hp.conv_kernel[i], s = hp.conv_stride[i];
        if (s != 2 || (k != 2 && k != 3)) { err = "WavLM: conv layers 1..6 must have stride 2 and kernel 2 or 3"; return false; }
    }
    refloader L{*this, g, err};
    {   // conv0: [C, 1, K]
        const gguf_tensor * t = g.find("wavlm.feat.conv0.weight");
        if (t == nullptr) { err = "missing tensor: wavlm.feat.conv0.weight"; return false; }
        if (t->ne.size() != 3 || (int) t->ne[0] != hp.conv_kernel[0] || t->ne[1] != 1) { err = "unexpected shape for tensor: wavlm.feat.conv0.weight"; return false; }
        conv_ch = (int) t->ne[2];
        if (!L.vec("wavlm.feat.conv0.weight", conv_ch * hp.conv_kernel[0], conv0_w)) return false;
        if (!L.norm("wavlm.feat.conv0.norm", conv_ch, conv0_norm)) return false;
    }
    if (conv_ch % 64) { err = "WavLM: conv channels must be a multiple of 64"; return false; }
    for (int i = 1; i < 7; ++i) {
        int co = 0, ci = 0, k = 0;
        if (!L.conv("wavlm.feat.conv" + std::to_string(i) + ".weight", "", co, ci, k, conv[i])) return false;
        if (co != conv_ch || ci != conv_ch || k != hp.conv_kernel[i]) { err = "unexpected shape for tensor: wavlm.feat.conv" + std::to_string(i) + ".weight"; return false; }
    }
    if (!L.norm("wavlm.proj.norm", conv_ch, proj_norm)) return false;
    if (!L.linear("wavlm.proj", E, conv_ch, linear_dtype, proj)) return false;
    {   // grouped positional conv [E, E/groups, k] -> one GEMM weight per group, input channels padded to 64 per tap
        const gguf_tensor * t = g.find("wavlm.pos_conv.weight");
        if (t == nullptr) { err = "missing tensor: wavlm.pos_conv.weight"; return false; }
        if (t->ne.size() != 3 || (int) t->ne[2] != E) { err = "unexpected shape for tensor: wavlm.pos_conv.weight"; return false; }
        pos_k = (int) t->ne[0];
        const int ipg = (int) t->ne[1];
        pos_groups = 16;    // src/wavlm-extractor.cpp:693
        if (E % pos_groups || ipg != E / pos_groups) { err = "unexpected grouped positional-conv shape"; return false; }
        if ((ipg * 2) % 16 || (ipg * 4) % 16) { err = "WavLM: positional-conv group width must be a multiple of 8"; return false; }
        pos_kp = (int) round_up((size_t) ipg, 64);
        std::vector<float> w, b;
        if (!L.get("wavlm.pos_conv.weight", w, 0) || !L.get("wavlm.pos_conv.bias", b, (size_t) E)) return false;
        b.resize(round_up((size_t) E, 4) + 64, 0.0f);
        pos_bias = upload_f32(b.data(), b.size());
        pos.resize((size_t) pos_groups);
        const int opg = ipg;
        for (int gi = 0; gi < pos_groups; ++gi) {
            std::vector<float> rows((size_t) opg * pos_k * pos_kp, 0.0f);
            for (int o = 0; o < opg; ++o)
                for (int ci = 0; ci < ipg; ++ci)
                    for (int j = 0; j < pos_k; ++j)
                        rows[((size_t) o * pos_k + j) * pos_kp + ci] = w[((size_t) (gi * opg + o) * ipg + ci) * pos_k + j];
            make_linear(pos[(size_t) gi], rows, opg, pos_kp, pos_k, conv_dtype, nullptr);
        }
    }
    if (!L.norm("wavlm.transformer.norm", E, tr_norm)) return false;
    {
        const gguf_tensor * t = g.find("wavlm.layer.0.attn.rel_embed.weight");
        if (t == nullptr) { err = "missing tensor: wavlm.layer.0.attn.rel_embed.weight"; return false; }
        std::vector<float> re;
        if (!L.get("wavlm.layer.0.attn.rel_embed.weight", re, (size_t) hp.num_buckets * H)) return false;
        // both layouts of src/wavlm-extractor.cpp:750-756 -> host [bucket][head]
        rel_embed.assign((size_t) hp.num_buckets * H, 0.0f);
        const bool buckets_fast = t->ne.size() == 2 && (int) t->ne[0] == hp.num_buckets && (int) t->ne[1] == H;
        const bool heads_fast = t->ne.size() == 2 && (int) t->ne[0] == H && (int) t->ne[1] == hp.num_buckets;
        if (!buckets_fast && !heads_fast) { err = "unexpected wavlm rel_embed tensor shape"; return false; }
        for (int bk = 0; bk < hp.num_buckets; ++bk)
            for (int h = 0; h < H; ++h)
                rel_embed[(size_t) bk * H + h] = heads_fast ? re[(size_t) h + (size_t) H * bk] : re[(size_t) bk + (size_t) hp.num_buckets * h];
    }
    layers.resize((size_t) hp.n_layers);
    const float q_scale = 1.0f / std::sqrt((float) hd);
    for (int i = 0; i < hp.n_layers; ++i) {
        wavlm_layer & l = layers[(size_t) i];
        const std::string p = "wavlm.layer." + std::to_string(i) + ".";
        // softmax(q k^T * scale + mask): the scale is folded into the q rows of in_proj (weights AND bias)
        if (!L.linear(p + "attn.in_proj", 3 * E, E, linear_dtype, l.in_proj, q_scale, E)) return false;
        if (!L.linear(p + "attn.out_proj", E, E, linear_dtype, l.out_proj)) return false;
        if (!L.vec(p + "attn.gru.weight", 8 * hd, l.gru_w) || !L.vec(p + "attn.gru.bias", 8, l.gru_b) ||
            !L.vec(p + "attn.gru_const", H, l.gru_const)) return false;
        if (!L.norm(p + "norm1", E, l.norm1) || !L.norm(p + "norm2", E, l.norm2)) return false;
        const gguf_tensor * t = g.find(p + "ffn.w1.weight");
        if (t == nullptr || t->ne.size() != 2 || (int) t->ne[0] != E) { err = "unexpected shape for tensor: " + p + "ffn.w1.weight"; return false; }
        ffn = (int) t->ne[1];
        if (ffn % 64) { err = "WavLM: ffn width must be a multiple of 64"; return false; }
        if (!L.linear(p + "ffn.w1", ffn, E, linear_dtype, l.w1) || !L.linear(p + "ffn.w2", E, ffn, linear_dtype, l.w2)) return false;
    }
    MIO_CUDA(cudaDeviceSynchronize());
    wavlm_loaded = true;
    return true;
}

bool refmodel::load_global_encoder(const gguf_file & g, int dev, int out_dim, std::string & err) {
    device = dev;
    MIO_CUDA(cudaSetDevice(device));
    cudaDeviceProp prop;
    MIO_CUDA(cudaGetDeviceProperties(&prop, device));
    num_sms = prop.multiProcessorCount;
    if (g.find("global_encoder.backbone.embed.weight") == nullptr) return true;   // optional tensors: not present
    uint32_t u = 0;
    int layers_n = 4;
    if (g.get_u32("miocodec.global_encoder.input_channels", u)) ge_in = (int) u;
    if (g.get_u32("miocodec.global_encoder.output_channels", u)) ge_out = (int) u;
    if (g.get_u32("miocodec.global_encoder.dim", u)) ge_dim = (int) u;
    if (g.get_u32("miocodec.global_encoder.intermediate_dim", u)) ge_ff = (int) u;
    if (g.get_u32("miocodec.global_encoder.num_layers", u)) layers_n = (int) u;
    (void) out_dim;
    refloader L{*this, g, err};
    int co = 0, ci = 0, k = 0;
    if (!L.conv("global_encoder.backbone.embed.weight", "global_encoder.backbone.embed.bias", co, ci, k, ge_embed)) return false;
    if (co != ge_dim || ci != ge_in || (k & 1) == 0) { err = "unexpected shape for tensor: global_encoder.backbone.embed.weight"; return false; }
    ge_embed_k = k;
    if (ge_dim % 64 || ge_ff % 64 || ge_in % 64) { err = "global encoder dims must be multiples of 64"; return false; }
    if (!L.norm("global_encoder.backbone.norm", ge_dim, ge_norm)) return false;
    ge_blocks.resize((size_t) layers_n);
    for (int i = 0; i < layers_n; ++i) {
        ge_block & b = ge_blocks[(size_t) i];
        const std::string p = "global_encoder.backbone.blk." + std::to_string(i) + ".";
        const gguf_tensor * t = g.find(p + "dwconv.weight");
        if (t == nullptr) { err = "missing tensor: " + p + "dwconv.weight"; return false; }
        if (t->ne.size() != 3 || t->ne[1] != 1 || (int) t->ne[2] != ge_dim || ((int) t->ne[0] & 1) == 0) { err = "unexpected shape for tensor: " + p + "dwconv.weight"; return false; }
        b.dw_k = (int) t->ne[0];
        if (!L.vec(p + "dwconv.weight", ge_dim * b.dw_k, b.dw_w) || !L.vec(p + "dwconv.bias", ge_dim, b.dw_b)) return false;
        if (!L.norm(p + "norm", ge_dim, b.norm)) return false;
        if (!L.linear(p + "pw1", ge_ff, ge_dim, linear_dtype, b.pw1) || !L.linear(p + "pw2", ge_dim, ge_ff, linear_dtype, b.pw2)) return false;
        if (!L.vec(p + "gamma", ge_dim, b.gamma)) return false;
    }
    if (!L.norm("global_encoder.backbone.final_norm", ge_dim, ge_final_norm)) return false;
    if (!L.conv("global_encoder.pool.attn0.weight", "global_encoder.pool.attn0.bias", co, ci, k, ge_attn0)) return false;
    if (ci != ge_dim || k != 1 || co % 64) { err = "unexpected shape for tensor: global_encoder.pool.attn0.weight"; return false; }
    ge_attn = co;
    if (!L.conv("global_encoder.pool.attn2.weight", "global_encoder.pool.attn2.bias", co, ci, k, ge_attn2)) return false;
    if (ci != ge_attn || co != ge_dim || k != 1) { err = "unexpected shape for tensor: global_encoder.pool.attn2.weight"; return false; }
    if (!L.vec("global_encoder.pool.proj.weight", ge_out * 2 * ge_dim, ge_proj_w) || !L.vec("global_encoder.pool.proj.bias", ge_out, ge_proj_b)) return false;
    if (!L.norm("global_encoder.pool.norm", ge_out, ge_pool_norm)) return false;
    MIO_CUDA(cudaDeviceSynchronize());
    ge_loaded = true;
    return true;
}

// -------------------------------------------------------------------------------------------------
// executor
// -------------------------------------------------------------------------------------------------
refenc::refenc(std::shared_ptr<refmodel> m) : mdl(std::move(m)) {
    MIO_CUDA(cudaSetDevice(mdl->device));
    MIO_CUDA(cudaStreamCreateWithFlags(&st_, cudaStreamNonBlocking));
    for (auto & e : ev_) MIO_CUDA(cudaEventCreate(&e));
}

refenc::~refenc() {
    cudaSetDevice(mdl->device);
    if (st_) cudaStreamSynchronize(st_);
    for (dbuf * b : {&src_, &wav_, &peak_, &part_, &mr_, &lvl_[0], &lvl_[1], &lvl_[2], &lvl_[3], &lvl_[4], &lvl_[5], &lvl_[6], &c6_,
                     &a16_, &x32_, &xp16_, &x16_, &qkv16_, &att16_, &ff16_, &sum32_, &gate_, &rel_, &g32_, &g16_, &ga16_, &gl32_,
                     &stat_, &emb_, &sslc_, &d_blob_})
        if (b->p) cudaFree(b->p);
    if (h_blob_) cudaFreeHost(h_blob_);
    for (auto & e : ev_) cudaEventDestroy(e);
    if (st_) cudaStreamDestroy(st_);
}

void refenc::sync() { MIO_CUDA(cudaStreamSynchronize(st_)); }

void refenc::ensure(dbuf & b, size_t bytes) {
    bytes += 4096;
    if (bytes <= b.cap) return;
    MIO_CUDA(cudaStreamSynchronize(st_));
    if (b.p) { MIO_CUDA(cudaFree(b.p)); b.p = nullptr; b.cap = 0; }
    const size_t cap = round_up(bytes + bytes / 8, 1 << 20);
    MIO_CUDA(cudaMalloc(&b.p, cap));
    MIO_CUDA(cudaMemsetAsync(b.p, 0, cap, st_));
    b.cap = cap;
    amaps_.clear();
    omaps_.clear();
}

size_t refenc::workspace_bytes() const {
    size_t s = 0;
    for (const dbuf * b : {&src_, &wav_, &peak_, &part_, &mr_, &lvl_[0], &lvl_[1], &lvl_[2], &lvl_[3], &lvl_[4], &lvl_[5], &lvl_[6],
                           &c6_, &a16_, &x32_, &xp16_, &x16_, &qkv16_, &att16_, &ff16_, &sum32_, &gate_, &rel_, &g32_, &g16_, &ga16_,
                           &gl32_, &stat_, &emb_, &sslc_, &d_blob_})
        s += b->cap;
    return s;
}

size_t refenc::blob_add(const void * src, size_t bytes) {
    const size_t off = round_up(blob_.size(), 256);
    blob_.resize(off + std::max<size_t>(bytes, 4));
    if (bytes) std::memcpy(blob_.data() + off, src, bytes);
    return off;
}

const CUtensorMap & refenc::amap(const void * p, uint64_t width, uint64_t rows) {
    const std::array<uint64_t, 3> key = {(uint64_t) (uintptr_t) p, width, rows};
    auto it = amaps_.find(key);
    if (it != amaps_.end()) return it->second;
    if (amaps_.size() > 512) amaps_.clear();
    return amaps_.emplace(key, make_tmap_16(p, rows, width, 128)).first->second;
}

const CUtensorMap & refenc::omap(const void * p, int rows, int cols, int ld, int elem_bytes) {
    const std::array<uint64_t, 5> key = {(uint64_t) (uintptr_t) p, (uint64_t) rows, (uint64_t) cols, (uint64_t) ld, (uint64_t) elem_bytes};
    auto it = omaps_.find(key);
    if (it != omaps_.end()) return it->second;
    if (omaps_.size() > 512) omaps_.clear();
    return omaps_.emplace(key, make_tmap_out(p, (uint64_t) rows, (uint64_t) cols, (uint64_t) ld, elem_bytes)).first->second;
}

void refenc::gemm(const void * a, uint64_t a_width, uint64_t a_rows, const dev_linear & w, int M, int epilogue, void * out, int ldo,
                  int out_cols, const gopt & o) {
    gemm_args g;
    std::memset(&g, 0, sizeof(g));
    g.map_a = amap(a, a_width, a_rows);
    g.map_b = w.map_b;
    g.map_b_tail = w.map_b_tail;
    g.map_out = omap(out, M, out_cols, ldo, epilogue == EPI_STORE32 ? 4 : 2);
    g.M = M; g.N = w.N; g.K = w.K; g.taps = w.taps;
    g.tap_row0 = o.tap_row0; g.tap_shift = o.tap_shift; g.a_col0 = o.a_col0; g.act = o.act;
    g.dtype = w.dtype; g.out_dtype = o.out_dtype; g.epilogue = epilogue;
    g.out = out; g.ldo = ldo; g.bias = o.bias;
    g.res = o.residual ? (const float *) out : nullptr; g.ldr = ldo;
    g.gate = o.gate; g.gate_ld = 0;
    gemm_launch(g, mdl->num_sms, st_);
    flops_ += 2.0 * (double) M * w.N * w.K * w.taps;
}

// -------------------------------------------------------------------------------------------------
void refenc::upload_blob() {
    ensure(d_blob_, blob_.size() + 256);
    if (blob_.size() > h_blob_cap_) {
        MIO_CUDA(cudaStreamSynchronize(st_));
        if (h_blob_) MIO_CUDA(cudaFreeHost(h_blob_));
        h_blob_cap_ = round_up(blob_.size() * 2, 1 << 16);
        MIO_CUDA(cudaHostAlloc(&h_blob_, h_blob_cap_, cudaHostAllocDefault));
    } else {
        MIO_CUDA(cudaStreamSynchronize(st_));
    }
    std::memcpy(h_blob_, blob_.data(), blob_.size());
    MIO_CUDA(cudaMemcpyAsync(d_blob_.p, h_blob_, blob_.size(), cudaMemcpyHostToDevice, st_));
    h2d_bytes_ += blob_.size();
}

// padded layout + per-row tables + attention work list from seq_[]; appends to blob_ and binds device pointers
void refenc::layout_from_seq() {
    const refmodel & M = *mdl;
    const int n = n_clips_;
    prow0_.assign((size_t) n, 0);
    gate_off_.assign((size_t) n, 0);
    int r = GAP, go = 0;
    max_seq_ = 0; total_seq_ = 0;
    for (int c = 0; c < n; ++c) {
        prow0_[(size_t) c] = r;
        r += seq_[(size_t) c] + GAP;
        gate_off_[(size_t) c] = go;
        go += seq_[(size_t) c] * M.hp.n_heads;
        max_seq_ = std::max(max_seq_, seq_[(size_t) c]);
        total_seq_ += seq_[(size_t) c];
    }
    prows_ = r;
    std::vector<int> map_pad((size_t) prows_, -1), map_valid((size_t) prows_, -1), row_clip((size_t) prows_, -1), row_pos((size_t) prows_, 0),
        map_compact((size_t) std::max(1, total_seq_), 0);
    std::vector<int2> work;
    int cr = 0;
    for (int c = 0; c < n; ++c) {
        const int src0 = have_wave_ ? (row0_l0_[(size_t) c] >> 6) : cr;
        for (int t = 0; t < seq_[(size_t) c]; ++t) {
            const int pr = prow0_[(size_t) c] + t;
            map_pad[(size_t) pr] = src0 + t;
            map_valid[(size_t) pr] = pr;
            row_clip[(size_t) pr] = c;
            row_pos[(size_t) pr] = t;
            map_compact[(size_t) cr + t] = pr;
        }
        for (int q0 = 0; q0 < seq_[(size_t) c]; q0 += 64) work.push_back(make_int2(c, q0));
        cr += seq_[(size_t) c];
    }
    n_work_ = (int) work.size();
    const size_t o_seq = blob_add(seq_.data(), (size_t) n * 4), o_pr = blob_add(prow0_.data(), (size_t) n * 4),
                 o_go = blob_add(gate_off_.data(), (size_t) n * 4), o_mp = blob_add(map_pad.data(), map_pad.size() * 4),
                 o_mv = blob_add(map_valid.data(), map_valid.size() * 4), o_mc = blob_add(map_compact.data(), map_compact.size() * 4),
                 o_rc = blob_add(row_clip.data(), row_clip.size() * 4), o_rp = blob_add(row_pos.data(), row_pos.size() * 4),
                 o_wk = blob_add(work.data(), work.size() * sizeof(int2));
    ensure(d_blob_, blob_.size() + 256);
    const uint8_t * base = (const uint8_t *) d_blob_.p;
    d_seq_ = (const int *) (base + o_seq); d_prow0_ = (const int *) (base + o_pr); d_gate_off_ = (const int *) (base + o_go);
    d_map_pad_ = (const int *) (base + o_mp); d_map_valid_ = (const int *) (base + o_mv); d_map_compact_ = (const int *) (base + o_mc);
    d_row_clip_ = (const int *) (base + o_rc); d_row_pos_ = (const int *) (base + o_rp); d_work_ = (const int2 *) (base + o_wk);

    const int E = std::max(M.hp.embed_dim, M.ge_loaded ? M.ge_in : 0);
    const size_t PR = round_up((size_t) prows_, 256) + 256;
    ensure(x32_, PR * E * 4);
    ensure(sum32_, PR * E * 4);
    ensure(xp16_, PR * E * 2);
    if (M.wavlm_loaded) {
        ensure(x16_, PR * E * 2);
        ensure(qkv16_, PR * 3 * E * 2);
        ensure(att16_, PR * E * 2);
        ensure(ff16_, PR * M.ffn * 2);
        ensure(gate_, (size_t) std::max(1, go) * 4);
        // relative-position embedding per distance d = k - q, |d| < max_seq (src/wavlm-extractor.cpp:257-279, :893-912)
        if (max_seq_ > rel_max_seq_) {
            const int H = M.hp.n_heads, nb = M.hp.num_buckets, md = M.hp.max_distance;
            const int cap = std::max(max_seq_, std::max(1024, rel_max_seq_ * 2));
            std::vector<float> tab((size_t) H * (2 * cap - 1));
            for (int d = -(cap - 1); d <= cap - 1; ++d) {
                int bucket = 0;
                const int half = nb / 2;
                if (d > 0) bucket += half;
                const int nn = std::abs(d);
                const int max_exact = half / 2;
                if (nn < max_exact) bucket += nn;
                else {
                    const float n_f = (float) nn, me = (float) max_exact;
                    const float denom = std::log((float) md / me);
                    int val = max_exact + (int) std::floor(std::log(n_f / me) / denom * (half - max_exact));
                    bucket += std::min(val, half - 1);
                }
                for (int h = 0; h < H; ++h) tab[(size_t) h * (2 * cap - 1) + (size_t) (d + cap - 1)] = M.rel_embed[(size_t) bucket * H + h];
            }
            ensure(rel_, tab.size() * 4);
            MIO_CUDA(cudaMemcpyAsync(rel_.p, tab.data(), tab.size() * 4, cudaMemcpyHostToDevice, st_));
            MIO_CUDA(cudaStreamSynchronize(st_));
            rel_max_seq_ = cap;
        }
    }
    if (M.ge_loaded) {
        ensure(g32_, PR * M.ge_dim * 4);
        ensure(g16_, PR * std::max(M.ge_dim, M.ge_ff) * 2);
        ensure(ga16_, PR * std::max(M.ge_ff, M.ge_attn) * 2);
        ensure(gl32_, PR * M.ge_dim * 4);
        ensure(stat_, (size_t) n * 2 * M.ge_dim * 4);
        ensure(emb_, (size_t) n * M.ge_out * 4);
    }
    ensure(sslc_, (size_t) std::max(1, total_seq_) * E * 4);
}

void refenc::prepare(const float * const * pcm, const size_t * n, int sample_rate, float max_seconds, size_t batch) {
    const refmodel & M = *mdl;
    MIO_CUDA(cudaSetDevice(M.device));
    if (!M.wavlm_loaded) throw std::runtime_error("WavLM model is not loaded");
    if (batch == 0 || pcm == nullptr || n == nullptr) throw std::runtime_error("reference audio is empty");
    if (sample_rate <= 0) throw std::runtime_error("invalid source sample rate");
    n_clips_ = (int) batch;
    have_wave_ = true;
    src_rate_ = sample_rate;
    h2d_bytes_ = 0;
    // src/wavlm-extractor.cpp:562-567: decode at most llround(rate * max_seconds) source samples
    size_t max_src = 0;
    if (max_seconds > 0.0f) max_src = (size_t) std::max<long long>(1, std::llround((double) sample_rate * (double) max_seconds));
    n_src_.assign(batch, 0); n_wav_.assign(batch, 0); seq_.assign(batch, 0); row0_l0_.assign(batch, 0);
    len_.assign(batch, {});
    std::vector<long long> src_off(batch), wav_off(batch);
    std::vector<int> len0(batch);
    long long so = 0, wo = 0;
    int r0 = 0, max_nsrc = 0, max_nwav = 0, max_len0 = 0;
    const int chunk = wav_conv0_chunk_frames();
    for (size_t c = 0; c < batch; ++c) {
        if (pcm[c] == nullptr || n[c] == 0) throw std::runtime_error("reference audio is empty");
        size_t ns = n[c];
        if (max_src > 0) ns = std::min(ns, max_src);
        if (ns > (size_t) 1 << 28) throw std::runtime_error("reference audio is too large");
        n_src_[c] = (int) ns;
        long long nw = (long long) ns;
        if (sample_rate != M.hp.sample_rate) {
            const double ratio = (double) M.hp.sample_rate / (double) sample_rate;
            nw = std::max<long long>(1, std::llround((double) ns * ratio));
        }
        n_wav_[c] = (int) nw;
        long long L = nw;
        for (int i = 0; i < 7; ++i) {   // wavlm_conv_stack_out_len, src/wavlm-extractor.cpp:242-255, without its clamps:
            const int k = M.hp.conv_kernel[i], s = M.hp.conv_stride[i];
            if (L < k) throw std::runtime_error("reference audio is too short for the WavLM feature extractor");
            L = (L - k) / s + 1;
            len_[c][(size_t) i] = (int) L;
        }
        seq_[c] = len_[c][6];
        len0[c] = len_[c][0];
        src_off[c] = so; so += (long long) round_up(ns, 4);
        wav_off[c] = wo; wo += (long long) round_up((size_t) nw, 4);
        row0_l0_[c] = r0;
        r0 += (int) round_up((size_t) len_[c][0], (size_t) std::max(chunk, 64));
        max_nsrc = std::max(max_nsrc, n_src_[c]); max_nwav = std::max(max_nwav, n_wav_[c]); max_len0 = std::max(max_len0, len0[c]);
    }
    rows0_ = (int) round_up((size_t) r0, 128);
    n_chunks_ = (max_len0 + chunk - 1) / chunk;
    // conv0 apply writes whole chunks: every clip extent is a chunk multiple, so n_chunks_ chunks of the LONGEST clip may
    // overrun a shorter clip's extent -> the apply kernel is launched per clip extent (grid.x covers the longest; shorter
    // clips' extra chunks would clobber the next clip).  Guarded in the kernel launch below through per-clip chunk counts.
    blob_.clear();
    const size_t o_so = blob_add(src_off.data(), batch * 8), o_wo = blob_add(wav_off.data(), batch * 8),
                 o_ns = blob_add(n_src_.data(), batch * 4), o_nw = blob_add(n_wav_.data(), batch * 4),
                 o_l0 = blob_add(len0.data(), batch * 4), o_r0 = blob_add(row0_l0_.data(), batch * 4);
    layout_from_seq();
    const int C = M.conv_ch;
    ensure(src_, (size_t) so * 4);
    ensure(wav_, (size_t) wo * 4 + 256);
    ensure(peak_, batch * 4);
    ensure(part_, batch * (size_t) n_chunks_ * (16 * 16 + 16) * sizeof(double));   // (s, R) partials of the conv0 statistics
    ensure(mr_, batch * (size_t) C * sizeof(float2));
    for (int i = 0; i < 6; ++i) ensure(lvl_[i], ((size_t) (rows0_ >> i) + 256) * C * 2);
    ensure(c6_, ((size_t) (rows0_ >> 6) + 256) * C * 4);
    ensure(a16_, ((size_t) (rows0_ >> 6) + 256) * C * 2);
    const uint8_t * base = (const uint8_t *) d_blob_.p;
    d_src_off_ = (const long long *) (base + o_so); d_wav_off_ = (const long long *) (base + o_wo);
    d_n_src_ = (const int *) (base + o_ns); d_n_wav_ = (const int *) (base + o_nw);
    d_len0_ = (const int *) (base + o_l0); d_row0_l0_ = (const int *) (base + o_r0);
    upload_blob();
    for (size_t c = 0; c < batch; ++c) {
        MIO_CUDA(cudaMemcpyAsync((float *) src_.p + src_off[c], pcm[c], (size_t) n_src_[c] * 4, cudaMemcpyHostToDevice, st_));
        h2d_bytes_ += (size_t) n_src_[c] * 4;
    }
    (void) max_nsrc; (void) max_nwav;
}

void refenc::prepare_ssl(const float * const * ssl, const int * frames, size_t batch) {
    const refmodel & M = *mdl;
    MIO_CUDA(cudaSetDevice(M.device));
    if (batch == 0 || ssl == nullptr || frames == nullptr) throw std::runtime_error("invalid ssl frame length");
    n_clips_ = (int) batch;
    have_wave_ = false;
    h2d_bytes_ = 0;
    seq_.assign(batch, 0);
    for (size_t c = 0; c < batch; ++c) {
        if (ssl[c] == nullptr || frames[c] <= 0) throw std::runtime_error("invalid ssl frame length");
        seq_[c] = frames[c];
    }
    blob_.clear();
    layout_from_seq();
    upload_blob();
    // stage the compact ssl rows, then spread them into the padded layout (sum32_ holds "layer-output sum", scale 1)
    const int E = M.ge_loaded ? M.ge_in : M.hp.embed_dim;
    size_t off = 0;
    for (size_t c = 0; c < batch; ++c) {
        MIO_CUDA(cudaMemcpyAsync((float *) sslc_.p + off * E, ssl[c], (size_t) seq_[c] * E * 4, cudaMemcpyHostToDevice, st_));
        off += (size_t) seq_[c];
        h2d_bytes_ += (size_t) seq_[c] * E * 4;
    }
    launch_relayout((const float *) sslc_.p, d_map_pad_, prows_, E, 1.0f, (float *) sum32_.p, nullptr, 0, st_);
}

void refenc::run_wavlm() {
    const refmodel & M = *mdl;
    MIO_CUDA(cudaSetDevice(M.device));
    if (!have_wave_) throw std::runtime_error("no waveform batch prepared");
    const wavlm_hparams & hp = M.hp;
    const int C = M.conv_ch, E = hp.embed_dim, H = hp.n_heads, ldt = M.linear_dtype, cdt = M.conv_dtype;
    const int n = n_clips_;
    int max_nsrc = 0, max_nwav = 0;
    for (int c = 0; c < n; ++c) { max_nsrc = std::max(max_nsrc, n_src_[(size_t) c]); max_nwav = std::max(max_nwav, n_wav_[(size_t) c]); }
    // front end + conv0 + GroupNorm over time + GELU -> level 0 (fp16)
    launch_wav_prepare((const float *) src_.p, d_src_off_, d_n_src_, (float *) wav_.p, d_wav_off_, d_n_wav_, (unsigned int *) peak_.p,
                       src_rate_, hp.sample_rate, n, max_nsrc, max_nwav, st_);
    launch_wav_conv0_moments((const float *) wav_.p, d_wav_off_, d_len0_, M.conv0_w, C, hp.conv_kernel[0], hp.conv_stride[0],
                             (double *) part_.p, n_chunks_, hp.norm_eps, (float2 *) mr_.p, n, st_);
    launch_wav_conv0_apply((const float *) wav_.p, d_wav_off_, d_len0_, d_row0_l0_, M.conv0_w, C, hp.conv_kernel[0], hp.conv_stride[0],
                           n_chunks_, (const float2 *) mr_.p, M.conv0_norm.w, M.conv0_norm.b, lvl_[0].p, n, st_);
    // conv1..6: stride-2 convolutions as implicit GEMMs over the pair view of the previous level, GELU in the epilogue
    for (int i = 1; i < 7; ++i) {
        const int rows_out = rows0_ >> i, rows_in = rows0_ >> (i - 1);
        gopt o; o.tap_shift = 1; o.act = 1; o.out_dtype = cdt;
        if (i < 6) gemm(lvl_[i - 1].p, (uint64_t) 2 * C, (uint64_t) rows_in / 2, M.conv[i], rows_out, EPI_STORE16, lvl_[i].p, C, C, o);
        else gemm(lvl_[i - 1].p, (uint64_t) 2 * C, (uint64_t) rows_in / 2, M.conv[i], rows_out, EPI_STORE32, c6_.p, C, C, o);
    }
    const int rows6 = rows0_ >> 6;
    // LayerNorm + projection (src/wavlm-extractor.cpp:686-687)
    launch_layernorm((const float *) c6_.p, rows6, C, hp.norm_eps, M.proj_norm.w, M.proj_norm.b, nullptr, 0, 0, 0, nullptr, a16_.p, ldt,
                     nullptr, st_);
    {
        gopt o; o.bias = M.proj.bias;
        // projection output goes to sum32_ as scratch (compact level-6 layout), then into the padded layout
        gemm(a16_.p, (uint64_t) C, (uint64_t) rows6, M.proj, rows6, EPI_STORE32, sum32_.p, E, E, o);
    }
    launch_relayout((const float *) sum32_.p, d_map_pad_, prows_, E, 1.0f, (float *) x32_.p, xp16_.p, cdt, st_);
    // grouped positional conv (k=128, pad 64) + bias + GELU, added onto x in place (:689-737): one GEMM per group
    const int ipg = E / M.pos_groups;
    for (int g = 0; g < M.pos_groups; ++g) {
        gopt o; o.tap_row0 = -64; o.a_col0 = g * ipg; o.act = 1; o.bias = M.pos_bias + g * ipg; o.residual = true;
        float * out = (float *) x32_.p + g * ipg;
        gemm(xp16_.p, (uint64_t) E, (uint64_t) prows_, M.pos[(size_t) g], prows_, EPI_STORE32, out, E, E - g * ipg, o);
    }
    launch_layernorm((const float *) x32_.p, prows_, E, hp.norm_eps, M.tr_norm.w, M.tr_norm.b, nullptr, 0, 0, 0, nullptr, x16_.p, ldt,
                     (float *) x32_.p, st_);
    // transformer layers (wavlm_layer_eval :762-851), post-LN
    for (int li = 0; li < hp.n_layers; ++li) {
        const wavlm_layer & l = M.layers[(size_t) li];
        launch_wavlm_gate((const float *) x32_.p, d_prow0_, d_seq_, d_gate_off_, n, max_seq_, H, hp.head_dim, l.gru_w, l.gru_b, l.gru_const,
                          (float *) gate_.p, st_);
        { gopt o; o.bias = l.in_proj.bias; o.out_dtype = ldt;
          gemm(x16_.p, (uint64_t) E, (uint64_t) prows_, l.in_proj, prows_, EPI_STORE16, qkv16_.p, 3 * E, 3 * E, o); }
        launch_attention_relbias(qkv16_.p, att16_.p, ldt, E, H, d_prow0_, d_seq_, d_gate_off_, (const float *) gate_.p,
                                 (const float *) rel_.p, rel_max_seq_, d_work_, n_work_, st_);
        { gopt o; o.bias = l.out_proj.bias; o.residual = true;
          gemm(att16_.p, (uint64_t) E, (uint64_t) prows_, l.out_proj, prows_, EPI_STORE32, x32_.p, E, E, o); }      // h = x + attn
        launch_layernorm((const float *) x32_.p, prows_, E, hp.norm_eps, l.norm1.w, l.norm1.b, nullptr, 0, 0, 0, nullptr, x16_.p, ldt,
                         nullptr, st_);
        { gopt o; o.bias = l.w1.bias; o.act = 1; o.out_dtype = ldt;
          gemm(x16_.p, (uint64_t) E, (uint64_t) prows_, l.w1, prows_, EPI_STORE16, ff16_.p, M.ffn, M.ffn, o); }
        { gopt o; o.bias = l.w2.bias; o.residual = true;
          gemm(ff16_.p, (uint64_t) M.ffn, (uint64_t) prows_, l.w2, prows_, EPI_STORE32, x32_.p, E, E, o); }         // h + ffn
        launch_layernorm((const float *) x32_.p, prows_, E, hp.norm_eps, l.norm2.w, l.norm2.b, nullptr, 0, 0, 0, nullptr, x16_.p, ldt,
                         (float *) x32_.p, st_);
        launch_accumulate((const float *) x32_.p, (float *) sum32_.p, (size_t) prows_ * E, li == 0, st_);
    }
}

void refenc::run_global() {
    const refmodel & M = *mdl;
    MIO_CUDA(cudaSetDevice(M.device));
    if (!M.ge_loaded) throw std::runtime_error("global encoder tensors are not available in this GGUF");
    const int E = M.ge_in, D = M.ge_dim, ldt = M.linear_dtype, cdt = M.conv_dtype;
    if (have_wave_ && M.hp.embed_dim != E) throw std::runtime_error("ssl feature shape mismatch");
    const float scale = have_wave_ ? 1.0f / (float) std::max(1, M.hp.n_layers) : 1.0f;
    // ssl (mean of the layer outputs) -> fp16 conv operand, gap rows zero
    launch_relayout((const float *) sum32_.p, d_map_valid_, prows_, E, scale, nullptr, xp16_.p, cdt, st_);
    { gopt o; o.tap_row0 = -(M.ge_embed_k / 2); o.bias = M.ge_embed.bias;
      gemm(xp16_.p, (uint64_t) E, (uint64_t) prows_, M.ge_embed, prows_, EPI_STORE32, g32_.p, D, D, o); }
    launch_layernorm((const float *) g32_.p, prows_, D, 1e-6f, M.ge_norm.w, M.ge_norm.b, nullptr, 0, 0, 0, nullptr, nullptr, 0,
                     (float *) g32_.p, st_);
    for (const ge_block & b : M.ge_blocks) {
        launch_dwconv_ln((const float *) g32_.p, d_row_clip_, d_row_pos_, d_seq_, prows_, D, b.dw_k, b.dw_w, b.dw_b, b.norm.w, b.norm.b,
                         1e-6f, g16_.p, ldt, st_);
        { gopt o; o.bias = b.pw1.bias; o.act = 1; o.out_dtype = ldt;
          gemm(g16_.p, (uint64_t) D, (uint64_t) prows_, b.pw1, prows_, EPI_STORE16, ga16_.p, M.ge_ff, M.ge_ff, o); }
        { gopt o; o.bias = b.pw2.bias; o.gate = b.gamma; o.residual = true;
          gemm(ga16_.p, (uint64_t) M.ge_ff, (uint64_t) prows_, b.pw2, prows_, EPI_STORE32, g32_.p, D, D, o); }
    }
    launch_layernorm((const float *) g32_.p, prows_, D, 1e-6f, M.ge_final_norm.w, M.ge_final_norm.b, nullptr, 0, 0, 0, nullptr, g16_.p, cdt,
                     (float *) g32_.p, st_);
    // attentive statistics pooling: k=1 convs (fp16 operands), tanh, softmax over time per channel
    { gopt o; o.bias = M.ge_attn0.bias; o.act = 2; o.out_dtype = cdt;
      gemm(g16_.p, (uint64_t) D, (uint64_t) prows_, M.ge_attn0, prows_, EPI_STORE16, ga16_.p, M.ge_attn, M.ge_attn, o); }
    { gopt o; o.bias = M.ge_attn2.bias;
      gemm(ga16_.p, (uint64_t) M.ge_attn, (uint64_t) prows_, M.ge_attn2, prows_, EPI_STORE32, gl32_.p, D, D, o); }
    launch_attn_stats_pool((const float *) gl32_.p, (const float *) g32_.p, d_prow0_, d_seq_, n_clips_, D, (float *) stat_.p, st_);
    launch_pool_proj_ln((const float *) stat_.p, 2 * D, M.ge_proj_w, M.ge_proj_b, M.ge_pool_norm.w, M.ge_pool_norm.b, 1e-5f, M.ge_out,
                        (float *) emb_.p, n_clips_, st_);
}

void refenc::fetch_ssl(size_t c, float * dst) {
    const refmodel & M = *mdl;
    const int E = M.hp.embed_dim;
    const float scale = 1.0f / (float) std::max(1, M.hp.n_layers);
    size_t off = 0;
    for (size_t i = 0; i < c; ++i) off += (size_t) seq_[i];
    if (c == 0 || true) launch_relayout((const float *) sum32_.p, d_map_compact_, total_seq_, E, scale, (float *) sslc_.p, nullptr, 0, st_);
    MIO_CUDA(cudaMemcpyAsync(dst, (const float *) sslc_.p + off * E, (size_t) seq_[c] * E * 4, cudaMemcpyDeviceToHost, st_));
}

void refenc::fetch_embedding(size_t c, float * dst) {
    MIO_CUDA(cudaMemcpyAsync(dst, (const float *) emb_.p + c * (size_t) mdl->ge_out, (size_t) mdl->ge_out * 4, cudaMemcpyDeviceToHost, st_));
}

float refenc::timed(int iters, bool with_global) {
    MIO_CUDA(cudaSetDevice(mdl->device));
    run_wavlm();
    if (with_global) run_global();
    flops_ = 0;
    MIO_CUDA(cudaEventRecord(ev_[0], st_));
    for (int i = 0; i < iters; ++i) { run_wavlm(); if (with_global) run_global(); }
    MIO_CUDA(cudaEventRecord(ev_[1], st_));
    MIO_CUDA(cudaStreamSynchronize(st_));
    float ms = 0;
    MIO_CUDA(cudaEventElapsedTime(&ms, ev_[0], ev_[1]));
    return ms / (float) std::max(1, iters);
}

} // namespace mio
