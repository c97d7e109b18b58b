#include "engine.h"

#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <cstdio>
#include <cstring>
#include <unordered_map>

namespace mio {

void launch_fill_spec(float * spec, long long rows, int pitch, int n_freq, unsigned long long seed, cudaStream_t st);
double device_sum(const float * x, long long n, cudaStream_t st);

namespace {

constexpr int k_code_min = 0, k_code_max = 12799;   // /root/reference/src/mio-tts-lib.cpp:30-31

size_t round_up(size_t v, size_t a) { return (v + a - 1) / a * a; }
size_t pad_rows(int rows) { return round_up((size_t) std::max(rows, 1), 128) + 128; }

} // namespace

// -------------------------------------------------------------------------------------------------
// pinned result pool
// -------------------------------------------------------------------------------------------------
namespace {
struct result_pool {
    std::mutex mtx;
    std::unordered_map<void *, size_t> live;                       // ptr -> class bytes
    std::unordered_map<size_t, std::vector<void *>> free_list;     // class bytes -> buffers
};
result_pool & pool() { static result_pool * p = new result_pool(); return *p; }
bool use_malloc_results() {
    static int v = -1;
    if (v < 0) { const char * e = std::getenv("MIO_TTS_B200_MALLOC_OUTPUT"); v = (e && e[0] == '1') ? 1 : 0; }
    return v == 1;
}
} // namespace

float * result_alloc(size_t n_floats) {
    const size_t bytes = std::max<size_t>(n_floats * sizeof(float), 4);
    if (use_malloc_results()) return (float *) std::malloc(bytes);
    size_t cls = 64 * 1024;
    while (cls < bytes) cls <<= 1;
    result_pool & P = pool();
    {
        std::lock_guard<std::mutex> lk(P.mtx);
        auto & fl = P.free_list[cls];
        if (!fl.empty()) {
            void * p = fl.back();
            fl.pop_back();
            P.live[p] = cls;
            return (float *) p;
        }
    }
    // Grow by slabs: cudaHostAlloc takes driver-wide locks (it stalls the launches of every other slot), so one call
    // brings in 32 MB worth of buffers of this class instead of one buffer.
    const size_t n_buf = std::max<size_t>(1, (32u << 20) / cls);
    void * p = nullptr;
    if (cudaHostAlloc(&p, cls * n_buf, cudaHostAllocPortable) != cudaSuccess) {
        cudaGetLastError();
        if (n_buf == 1 || cudaHostAlloc(&p, cls, cudaHostAllocPortable) != cudaSuccess) { cudaGetLastError(); return nullptr; }
        std::lock_guard<std::mutex> lk(P.mtx);
        P.live[p] = cls;
        return (float *) p;
    }
    std::lock_guard<std::mutex> lk(P.mtx);
    auto & fl = P.free_list[cls];
    for (size_t i = 1; i < n_buf; ++i) fl.push_back((uint8_t *) p + i * cls);
    P.live[p] = cls;
    return (float *) p;
}

bool result_free(void * p) {
    if (p == nullptr) return true;
    result_pool & P = pool();
    std::lock_guard<std::mutex> lk(P.mtx);
    auto it = P.live.find(p);
    if (it == P.live.end()) return false;
    P.free_list[it->second].push_back(p);
    P.live.erase(it);
    return true;
}

// -------------------------------------------------------------------------------------------------
engine::engine(std::shared_ptr<model> m) : mdl(std::move(m)) {
    MIO_CUDA(cudaSetDevice(mdl->device));
    MIO_CUDA(cudaStreamCreateWithFlags(&st_, cudaStreamNonBlocking));
    for (auto & e : ev_) MIO_CUDA(cudaEventCreate(&e));
    if (const char * e = std::getenv("MIO_TTS_B200_GRAPHS")) graphs_ok_ = !(e[0] == '0' && e[1] == '\0');
    if (const char * e = std::getenv("MIO_TTS_B200_GRAPH_MAX_CODES")) graph_max_codes_ = (size_t) std::max(0, std::atoi(e));
}

engine::~engine() {
    cudaSetDevice(mdl->device);
    if (st_) cudaStreamSynchronize(st_);
    for (dbuf * b : {&xt_, &fx_, &fy_, &ft_, &act16_, &qkv16_, &att16_, &ff16_, &cond_, &gn_stats_, &spec_buf_,
                     &audio_, &peak_, &pcm_, &d_blob_})
        if (b->p) cudaFree(b->p);
    if (h_blob_) cudaFreeHost(h_blob_);
    if (rope_p_) cudaFree(rope_p_);
    if (rope_d_) cudaFree(rope_d_);
    drop_graphs();
    for (auto & kv : istft_plans_) istft_plan_free(kv.second);
    for (auto & e : ev_) cudaEventDestroy(e);
    for (auto & e : gemm_ev_) cudaEventDestroy(e);
    if (st_) cudaStreamDestroy(st_);
}

void engine::sync() { MIO_CUDA(cudaStreamSynchronize(st_)); }

void engine::ensure(dbuf & b, size_t bytes) {
    if (bytes <= b.cap) return;
    MIO_CUDA(cudaStreamSynchronize(st_));
    // first allocation: the request + 12.5 %; a buffer that has to grow again grows by at least half, so a server whose
    // longest request creeps upward pays this (stream sync, free, malloc, clear, dropped tensor maps and graphs) O(log) times
    const size_t regrow = b.p ? b.cap + b.cap / 2 : 0;
    if (b.p) { MIO_CUDA(cudaFree(b.p)); b.p = nullptr; b.cap = 0; }
    const size_t cap = round_up(std::max(bytes + bytes / 8, regrow), 1 << 20);
    MIO_CUDA(cudaMalloc(&b.p, cap));
    MIO_CUDA(cudaMemsetAsync(b.p, 0, cap, st_));
    b.cap = cap;
    tmaps_.clear();   // cached tensor maps may point at the freed buffer
    omaps_.clear();
    drop_graphs();    // ... and so do the captured launch sequences
}

size_t engine::workspace_bytes() const {
    size_t s = 0;
    for (const dbuf * b : {&xt_, &fx_, &fy_, &ft_, &act16_, &qkv16_, &att16_, &ff16_, &cond_, &gn_stats_,
                           &spec_buf_, &audio_, &peak_, &pcm_, &d_blob_})
        s += b->cap;
    return s;
}

void engine::ensure_rope(int prenet_pos, int decoder_pos) {
    const hparams & hp = mdl->hp;
    auto grow = [&](float2 *& tab, int & have, int need, int head_dim) {
        if (need <= have) return;
        const int cap = std::max(need, std::max(2048, have * 2));
        MIO_CUDA(cudaStreamSynchronize(st_));
        if (tab) { MIO_CUDA(cudaFree(tab)); tab = nullptr; }
        drop_graphs();
        MIO_CUDA(cudaMalloc(&tab, (size_t) cap * (head_dim / 2) * sizeof(float2)));
        // theta_scale = powf(freq_base, -2/n_dims), computed on the host like ggml
        // (/root/reference/llama.cpp/ggml/src/ggml-cpu/ops.cpp:5616-5625, 5699-5716)
        const float theta_scale = powf(hp.rope_theta, -2.0f / (float) head_dim);
        launch_rope_table(tab, cap, head_dim / 2, theta_scale, st_);
        have = cap;
    };
    grow(rope_p_, rope_p_pos_, prenet_pos, hp.prenet_dim / hp.prenet_heads);
    grow(rope_d_, rope_d_pos_, decoder_pos, hp.decoder_dim / hp.decoder_heads);
}

const CUtensorMap & engine::amap(const dbuf & b, int K) {
    auto key = std::make_pair((const void *) b.p, K);
    auto it = tmaps_.find(key);
    if (it != tmaps_.end()) return it->second;
    const uint64_t rows = b.cap / ((size_t) K * 2);
    auto ins = tmaps_.emplace(key, make_tmap_16(b.p, rows, (uint64_t) K, 128));
    return ins.first->second;
}

// 16-bit [rows, cols] map with a 64-column x box_rows box (the attention kernel's q/k/v loads and output stores)
const CUtensorMap & engine::xmap(const void * p, int rows, int cols, int box_rows) {
    const std::array<uint64_t, 5> key = {(uint64_t) (uintptr_t) p, (uint64_t) rows, (uint64_t) cols, (uint64_t) box_rows, 16};
    auto it = omaps_.find(key);
    if (it != omaps_.end()) return it->second;
    if (omaps_.size() > 512) omaps_.clear();
    auto ins = omaps_.emplace(key, make_tmap_16(p, (uint64_t) rows, (uint64_t) cols, (uint32_t) box_rows));
    return ins.first->second;
}

const CUtensorMap & engine::omap(const void * p, int rows, int cols, int ld, int elem_bytes) {
    const std::array<uint64_t, 5> key = {(uint64_t) (uintptr_t) p, (uint64_t) rows, (uint64_t) cols, (uint64_t) ld,
                                         (uint64_t) elem_bytes};
    auto it = omaps_.find(key);
    if (it != omaps_.end()) return it->second;
    if (omaps_.size() > 512) omaps_.clear();   // batches of many different shapes: bound the cache
    auto ins = omaps_.emplace(key, make_tmap_out(p, (uint64_t) rows, (uint64_t) cols, (uint64_t) ld, elem_bytes));
    return ins.first->second;
}

// -------------------------------------------------------------------------------------------------
// layouts + metadata blob
// -------------------------------------------------------------------------------------------------
size_t engine::blob_add(const void * src, size_t bytes) {
    const size_t off = round_up(blob_.size(), 256);
    blob_.resize(off + bytes);
    if (bytes) std::memcpy(blob_.data() + off, src, bytes);
    return off;
}

void engine::build_layout(seq_layout & L, const std::vector<int> & len, int gap) {
    L = seq_layout();
    L.n_utt = (int) len.size();
    L.gap = gap;
    L.len = len;
    L.row0.resize(len.size());
    int r = gap;
    for (size_t u = 0; u < len.size(); ++u) {
        L.row0[u] = r;
        r += len[u] + gap;
        L.max_len = std::max(L.max_len, len[u]);
    }
    L.rows = r;
    for (int l : len) L.valid += l;   // rows that carry data (algorithmic work excludes the gap rows)
}

void engine::append_layout(seq_layout & L) {
    std::vector<int> row_utt((size_t) L.rows, -1), row_pos((size_t) L.rows, 0);
    std::vector<row_tile> tiles, tiles128;
    for (int u = 0; u < L.n_utt; ++u) {
        for (int p = 0; p < L.len[u]; ++p) { row_utt[(size_t) L.row0[u] + p] = u; row_pos[(size_t) L.row0[u] + p] = p; }
        for (int span : {64, 128}) {
            for (int s = 0; s < L.len[u]; s += span) {
                row_tile t;
                t.row0 = L.row0[u] + s; t.n = std::min(span, L.len[u] - s); t.utt = u; t.pos0 = s;
                t.utt_row0 = L.row0[u]; t.utt_len = L.len[u];
                (span == 64 ? tiles : tiles128).push_back(t);
            }
        }
    }
    L.n_tiles = (int) tiles.size();
    L.n_tiles128 = (int) tiles128.size();
    L.off_tiles128 = blob_add(tiles128.data(), tiles128.size() * sizeof(row_tile));
    L.off_len = blob_add(L.len.data(), L.len.size() * sizeof(int));
    L.off_row0 = blob_add(L.row0.data(), L.row0.size() * sizeof(int));
    L.off_row_utt = blob_add(row_utt.data(), row_utt.size() * sizeof(int));
    L.off_row_pos = blob_add(row_pos.data(), row_pos.size() * sizeof(int));
    L.off_tiles = blob_add(tiles.data(), tiles.size() * sizeof(row_tile));
}

void engine::bind_layout(seq_layout & L) {
    const uint8_t * base = (const uint8_t *) d_blob_.p;
    L.d_len = (const int *) (base + L.off_len);
    L.d_row0 = (const int *) (base + L.off_row0);
    L.d_row_utt = (const int *) (base + L.off_row_utt);
    L.d_row_pos = (const int *) (base + L.off_row_pos);
    L.d_tiles = (const row_tile *) (base + L.off_tiles);
    L.d_tiles128 = (const row_tile *) (base + L.off_tiles128);
}

void engine::prepare(const int32_t * const * codes, const size_t * n_codes, const float * const * emb, size_t n_emb,
                     size_t batch) {
    const hparams & hp = mdl->hp;
    MIO_CUDA(cudaSetDevice(mdl->device));
    if (batch == 0 || codes == nullptr || n_codes == nullptr) throw std::runtime_error("codes are empty");
    // same checks, same order, same messages as mio_tts_synthesize (src/mio-tts-lib.cpp:1198-1234)
    for (size_t u = 0; u < batch; ++u)
        if (codes[u] == nullptr || n_codes[u] == 0) throw std::runtime_error("codes are empty");
    for (size_t u = 0; u < batch; ++u)
        for (size_t i = 0; i < n_codes[u]; ++i) {
            const int32_t c = codes[u][i];
            if (c < k_code_min || c > k_code_max || c >= mdl->n_codebook) throw std::runtime_error("code id out of range");
        }
    bool has_embedding = emb != nullptr && n_emb > 0;
    if (has_embedding)
        for (size_t u = 0; u < batch; ++u) if (emb[u] == nullptr) has_embedding = false;
    if (hp.dynamic_global && !has_embedding) throw std::runtime_error("dynamic-global MioCodec requires embedding");
    if (!hp.dynamic_global && has_embedding) throw std::runtime_error("static MioCodec does not accept external embedding");
    if (has_embedding && (int) n_emb != hp.decoder_adanorm_dim) throw std::runtime_error("embedding dimension mismatch");

    n_utt_ = (int) batch;
    std::vector<int> tlen(batch), flen(batch);
    size_t total_codes = 0;
    for (size_t u = 0; u < batch; ++u) {
        if (n_codes[u] > (size_t) (1 << 24)) throw std::runtime_error("too many codes");
        tlen[u] = (int) n_codes[u];
        // frame counts: src/miocodec-decoder.cpp:1096-1103
        const long long stft = std::max<long long>(1, (long long) tlen[u] * hp.samples_per_token / std::max(1, hp.hop_length));
        flen[u] = (int) std::max<long long>(1, mdl->up_total > 1 ? stft / mdl->up_total : stft);
        total_codes += n_codes[u];
    }
    blob_.clear();
    build_layout(tok_, tlen, 0);
    build_layout(frm_, flen, 1);
    ups_.assign(mdl->ups.size(), seq_layout());
    {
        std::vector<int> cur = flen;
        for (size_t i = 0; i < mdl->ups.size(); ++i) {
            const upstage & s = mdl->ups[i];
            for (auto & l : cur) {
                const long long o = (long long) (l - 1) * s.factor + s.ksize - 2LL * s.pad;
                if (o <= 0) throw std::runtime_error("MioCodec decode failed: wave upsampler crop became non-positive");
                l = (int) o;
            }
            if (s.poly) {
                // polyphase stage: the output is the GEMM's [rows, factor*cout] result seen as [rows*factor, cout], so the
                // next layout is the input layout scaled by the factor (utterance u at factor*row0[u]; gaps stay >= 1 row)
                const seq_layout & in = i == 0 ? frm_ : ups_[i - 1];
                seq_layout & L = ups_[i];
                L = seq_layout();
                L.n_utt = in.n_utt; L.gap = in.gap * s.factor; L.len = cur; L.row0.resize(cur.size());
                for (size_t u = 0; u < cur.size(); ++u) {
                    if (cur[u] != in.len[u] * s.factor) throw std::runtime_error("MioCodec decode failed: polyphase upsampler length mismatch");
                    L.row0[u] = in.row0[u] * s.factor;
                    L.max_len = std::max(L.max_len, cur[u]);
                    L.valid += cur[u];
                }
                L.rows = in.rows * s.factor;
            } else {
                build_layout(ups_[i], cur, 1);
            }
        }
    }
    spec_ = ups_.empty() ? &frm_ : &ups_.back();

    // audio extents: embd_to_audio, src/mio-tts-lib.cpp:638-641, 742-744
    audio_off_.assign(batch, 0);
    audio_len_.assign(batch, 0);
    audio_total_ = 0;
    for (size_t u = 0; u < batch; ++u) {
        const long long S = spec_->len[u];
        const long long n_full = (S - 1) * hp.hop_length + hp.n_fft;
        const long long n_pad = std::max(0, (hp.n_fft - hp.hop_length) / 2);
        const long long n = std::max<long long>(n_pad, n_full - n_pad) - n_pad;
        audio_off_[u] = audio_total_;
        audio_len_[u] = n;
        audio_total_ += round_up((size_t) n, 64);
        if (n <= 0) throw std::runtime_error("audio is empty");
    }

    // ---- metadata blob ----
    std::vector<int32_t> all_codes(total_codes);
    {
        size_t o = 0;
        for (size_t u = 0; u < batch; ++u) { std::memcpy(&all_codes[o], codes[u], n_codes[u] * sizeof(int32_t)); o += n_codes[u]; }
    }
    const size_t off_codes = blob_add(all_codes.data(), all_codes.size() * sizeof(int32_t));
    size_t off_emb = 0;
    if (has_embedding) {
        std::vector<float> e(batch * n_emb);
        for (size_t u = 0; u < batch; ++u) std::memcpy(&e[u * n_emb], emb[u], n_emb * sizeof(float));
        off_emb = blob_add(e.data(), e.size() * sizeof(float));
    }
    std::vector<int> t2r(batch), t2l(batch);
    for (size_t u = 0; u < batch; ++u) { t2r[u] = 2 * tok_.row0[u]; t2l[u] = 2 * tok_.len[u]; }
    const size_t off_t2r = blob_add(t2r.data(), batch * sizeof(int));
    const size_t off_t2l = blob_add(t2l.data(), batch * sizeof(int));
    const size_t off_aoff = blob_add(audio_off_.data(), batch * sizeof(long long));
    const size_t off_alen = blob_add(audio_len_.data(), batch * sizeof(long long));
    append_layout(tok_);
    append_layout(frm_);
    for (auto & L : ups_) append_layout(L);

    // ---- workspace ----
    const int pd = hp.prenet_dim, dd = hp.decoder_dim;
    const size_t RT = pad_rows(tok_.rows), RF = pad_rows(frm_.rows);
    size_t f32_elems = std::max(RF * dd, RT * 2 * dd);
    size_t a16_elems = std::max(RT * pd, RF * dd);
    int gmax = 1;
    size_t up_head16 = 0;       // the fused out_proj + SnakeBeta writes the head's 16-bit operand into ff16_
    for (auto & rb : mdl->prior) gmax = std::max(gmax, rb.groups);
    for (auto & rb : mdl->post) gmax = std::max(gmax, rb.groups);
    {
        size_t rows_in = RF;
        for (size_t i = 0; i < mdl->ups.size(); ++i) {
            const upstage & s = mdl->ups[i];
            const size_t rows_out = pad_rows(ups_[i].rows);
            f32_elems = std::max(f32_elems, rows_in * (size_t) s.ksize * s.cout);
            f32_elems = std::max(f32_elems, rows_out * (size_t) s.cout);
            a16_elems = std::max(a16_elems, rows_in * (size_t) s.cin);
            a16_elems = std::max(a16_elems, rows_out * (size_t) s.cout);
            gmax = std::max(gmax, s.rb.groups);
            rows_in = rows_out;
        }
        if (!mdl->ups.empty()) { f32_elems = std::max(f32_elems, rows_in * (size_t) mdl->head_in); up_head16 = rows_in * (size_t) mdl->head_in; }
        a16_elems = std::max(a16_elems, rows_in * (size_t) mdl->head_in);
    }
    const size_t RS = pad_rows(spec_->rows);
    spec_pitch_ = (int) round_up((size_t) hp.n_fft + 2, 4);
    n_gn_slots_ = 4 * hp.resnet_blocks + 2 * (int) mdl->ups.size();
    gn_groups_max_ = gmax;

    ensure(xt_, RT * pd * 4);
    ensure(fx_, f32_elems * 4);
    ensure(fy_, f32_elems * 4);
    ensure(ft_, f32_elems * 4);
    ensure(act16_, a16_elems * 2);
    ensure(qkv16_, std::max(RT * 3 * pd, RF * 3 * dd) * 2);
    ensure(att16_, std::max(RT * pd, RF * dd) * 2);
    ensure(ff16_, std::max(std::max(std::max(RT * hp.prenet_ff, RF * hp.decoder_ff), RT * dd), up_head16) * 2);
    ensure(cond_, std::max<size_t>(16, batch * (size_t) mdl->cond_total * 4));
    ensure(gn_stats_, std::max<size_t>(16, (size_t) n_gn_slots_ * batch * gmax * 2 * sizeof(double)));
    ensure(spec_buf_, RS * spec_pitch_ * 4);
    ensure(audio_, (size_t) audio_total_ * 4 + 256);
    ensure(peak_, batch * 4);
    ensure(d_blob_, blob_.size() + 256);
    ensure_rope(tok_.max_len, frm_.max_len);

    if (blob_.size() > h_blob_cap_) {
        MIO_CUDA(cudaStreamSynchronize(st_));
        if (h_blob_) MIO_CUDA(cudaFreeHost(h_blob_));
        h_blob_cap_ = round_up(blob_.size() * 2, 1 << 16);
        MIO_CUDA(cudaHostAlloc(&h_blob_, h_blob_cap_, cudaHostAllocDefault));
    } else {
        MIO_CUDA(cudaStreamSynchronize(st_));   // the previous batch's H2D must have consumed the staging buffer
    }
    std::memcpy(h_blob_, blob_.data(), blob_.size());
    MIO_CUDA(cudaMemcpyAsync(d_blob_.p, h_blob_, blob_.size(), cudaMemcpyHostToDevice, st_));
    h2d_bytes_ = blob_.size();

    const uint8_t * base = (const uint8_t *) d_blob_.p;
    d_codes_ = (const int *) (base + off_codes);
    d_emb_ = has_embedding ? (const float *) (base + off_emb) : nullptr;
    d_tok2_row0_ = (const int *) (base + off_t2r);
    d_tok2_len_ = (const int *) (base + off_t2l);
    d_audio_off_ = (const long long *) (base + off_aoff);
    d_audio_len_ = (const long long *) (base + off_alen);
    bind_layout(tok_);
    bind_layout(frm_);
    for (auto & L : ups_) bind_layout(L);
}

void engine::reserve(int n_codes, int batch) {
    std::vector<int32_t> c((size_t) std::max(1, n_codes), 0);
    std::vector<float> e((size_t) mdl->hp.decoder_adanorm_dim, 0.0f);
    std::vector<const int32_t *> cp((size_t) batch, c.data());
    std::vector<size_t> cn((size_t) batch, c.size());
    std::vector<const float *> ep((size_t) batch, e.data());
    prepare(cp.data(), cn.data(), mdl->hp.dynamic_global ? ep.data() : nullptr, mdl->hp.dynamic_global ? e.size() : 0,
            (size_t) batch);
    sync();
}

// -------------------------------------------------------------------------------------------------
// kernels in sequence
// -------------------------------------------------------------------------------------------------
// every kernel of the chain walks the rows in the opposite direction of its predecessor (ping-pong traversal, common.h)
void engine::flip() { set_traversal_reverse(dir_ ^= 1); }

void engine::gemm(const dbuf & a, const dev_linear & w, int M, int epilogue, void * out, int ldo, int out_dtype,
                  const float * bias, const float * res, int ldr, const float * gate, const seq_layout * lay,
                  const float2 * rope, int heads, double * gn_stats, int gn_groups, const ln_fuse * ln) {
    gemm_args g;
    std::memset(&g, 0, sizeof(g));
    if (ln != nullptr) {        // EPI_STORE32_LN: the LayerNorm that consumes this GEMM's output, emitted as 16-bit rows
        g.map_out2 = omap(ln->out16, M, w.N, w.N, 2);
        g.ln_a = ln->a; g.ln_b = ln->b; g.ln_ld = ln->ld; g.ln_plus1 = ln->plus1; g.ln_eps = mdl->hp.norm_eps;
        g.ln_out_dtype = ln->out_dtype;
    }
    g.map_a = amap(a, w.K);
    g.map_b = w.map_b;
    g.map_b_tail = w.map_b_tail;
    g.map_out = omap(out, M, epilogue == EPI_SWIGLU16 ? w.N / 2 : w.N, ldo, (epilogue == EPI_STORE32 || epilogue == EPI_STORE32_LN) ? 4 : 2);
    g.M = M; g.N = w.N; g.K = w.K; g.taps = w.taps;
    g.tap_row0 = -(w.taps >> 1);   // k=3 "same" convolution over rows (taps == 1: plain GEMM)
    g.dtype = w.dtype; g.out_dtype = out_dtype; g.epilogue = epilogue;
    g.out = out; g.ldo = ldo; g.bias = bias; g.res = res; g.ldr = ldr;
    g.gate = gate; g.gate_ld = mdl->cond_total;
    g.row_utt = lay ? lay->d_row_utt : nullptr;
    g.row_pos = lay ? lay->d_row_pos : nullptr;
    g.gn_stats = gn_stats; g.gn_groups = gn_groups;
    if (snake_a_ != nullptr) { g.act = 3; g.snake_a = snake_a_; g.snake_ib = snake_ib_; }
    if (epilogue == EPI_ROPE16) {
        const int dim = w.N / 3;
        g.rope = rope; g.rope_cols = 2 * dim; g.head_dim = dim / heads;   // (the 1/sqrt(head_dim) q scale lives in Wq)
    }
    size_t evi = 0;
    if (prof_gemm_) {
        evi = gemm_ev_flops_.size() * 2 + attn_ev_.size() * 2;
        while (gemm_ev_.size() < evi + 2) { cudaEvent_t e; MIO_CUDA(cudaEventCreate(&e)); gemm_ev_.push_back(e); }
        MIO_CUDA(cudaEventRecord(gemm_ev_[evi], st_));
    }
    flip();
    gemm_launch(g, mdl->num_sms, st_);
    if (prof_gemm_) {
        MIO_CUDA(cudaEventRecord(gemm_ev_[evi + 1], st_));
        gemm_ev_flops_.push_back({evi, 2.0 * (double) (lay ? lay->valid : M) * w.N * w.K * w.taps});
        // algorithmic DRAM bytes of this launch: A once (the taps re-read the same rows), W once, the output, the residual
        const double out_cols = epilogue == EPI_SWIGLU16 ? w.N / 2 : w.N;
        gemm_bytes_ += (double) M * w.K * 2 + (double) w.N * w.K * w.taps * 2 +
                       (double) M * out_cols * ((epilogue == EPI_STORE32 || epilogue == EPI_STORE32_LN) ? 4 : 2) +
                       (res ? (double) M * w.N * 4 : 0.0) + (ln ? (double) M * w.N * 2 : 0.0);
        gemm_ev_shape_.push_back({M, w.N, w.K, w.taps, epilogue});
    }
}

// One transformer block.  ln1_done: the block's first LayerNorm already sits in act16_ (emitted by the previous block's
// down-projection).  next: the block whose first LayerNorm this block's down-projection should emit (nullptr: none).
// With dim == 512 (the AdaLN decoder) the O-projection emits the block's second LayerNorm the same way, so a fused block
// launches no LayerNorm kernel at all (MIO_TTS_B200_LN_FUSE=0 restores the separate kernels).
bool engine::run_tblock(float * x, const seq_layout & L, const tblock & blk, int dim, int heads, int ff, int window,
                        const float2 * rope, bool cond, bool ln1_done, const tblock * next) {
    static const bool fuse_env = [] { const char * e = std::getenv("MIO_TTS_B200_LN_FUSE"); return !(e && e[0] == '0'); }();
    const hparams & hp = mdl->hp;
    const int rows = L.rows;
    const int ldt = mdl->linear_dtype;
    const float * cv = (const float *) cond_.p;
    const bool fuse = fuse_env && dim == 512 && (blk.o.K % 64) == 0 && (blk.down.K % 64) == 0;
    // LayerNorm parameters as the fused epilogue wants them: AdaLN (scale, shift) rows of the utterance, or (weight, bias)
    auto ln_params = [&](const tblock & b, bool second) {
        ln_fuse f;
        f.out16 = act16_.p; f.out_dtype = ldt;
        if (cond) {
            const int off = second ? b.cond_ffn_off : b.cond_attn_off;
            f.a = cv + off + dim; f.b = cv + off; f.ld = mdl->cond_total; f.plus1 = 1.0f;
        } else {
            const dev_norm & n = second ? b.ln2 : b.ln1;
            f.a = n.w; f.b = n.b; f.ld = 0; f.plus1 = 0.0f;
        }
        return f;
    };
    auto ln_launch = [&](const tblock & b, bool second) {
        flip();
        if (cond) {
            const int off = second ? b.cond_ffn_off : b.cond_attn_off;
            launch_layernorm(x, rows, dim, hp.norm_eps, nullptr, nullptr, cv, mdl->cond_total, off, off + dim, L.d_row_utt, act16_.p,
                             ldt, nullptr, st_);
        } else {
            const dev_norm & n = second ? b.ln2 : b.ln1;
            launch_layernorm(x, rows, dim, hp.norm_eps, n.w, n.b, nullptr, 0, 0, 0, L.d_row_utt, act16_.p, ldt, nullptr, st_);
        }
    };
    if (!ln1_done) ln_launch(blk, false);
    gemm(act16_, blk.qkv, rows, EPI_ROPE16, qkv16_.p, 3 * dim, ldt, nullptr, nullptr, 0, nullptr, &L, rope, heads);
    size_t evi = 0;
    if (prof_gemm_) {
        evi = gemm_ev_flops_.size() * 2 + attn_ev_.size() * 2;
        while (gemm_ev_.size() < evi + 2) { cudaEvent_t e; MIO_CUDA(cudaEventCreate(&e)); gemm_ev_.push_back(e); }
        MIO_CUDA(cudaEventRecord(gemm_ev_[evi], st_));
    }
    flip();
    if (!force_generic_attention && attention_tc_ok(dim, heads, window))
        launch_attention_tc(xmap(qkv16_.p, rows, 3 * dim, 128), xmap(qkv16_.p, rows, 3 * dim, 192), xmap(att16_.p, rows, dim, 32),
                            att16_.p, ldt, dim, heads, window, rows, L.d_tiles128, L.n_tiles128, mdl->num_sms, st_);
    else if (!force_generic_attention && attention_tma_ok(dim, heads, window))
        launch_attention_tma(xmap(qkv16_.p, rows, 3 * dim, 64), xmap(att16_.p, rows, dim, 16), att16_.p, ldt, dim, heads, window,
                             rows, L.d_tiles, L.n_tiles, mdl->num_sms, st_);
    else
        launch_attention(qkv16_.p, att16_.p, ldt, dim, heads, window, L.d_tiles, L.n_tiles, force_generic_attention, st_);
    if (prof_gemm_) { MIO_CUDA(cudaEventRecord(gemm_ev_[evi + 1], st_)); attn_ev_.push_back(evi); }
    const float * gate1 = cond ? cv + blk.cond_attn_off + 2 * dim : nullptr;
    const float * gate2 = cond ? cv + blk.cond_ffn_off + 2 * dim : nullptr;
    if (fuse) {
        const ln_fuse f2 = ln_params(blk, true);
        gemm(att16_, blk.o, rows, EPI_STORE32_LN, x, dim, 0, nullptr, x, dim, gate1, &L, nullptr, 0, nullptr, 0, &f2);
    } else {
        gemm(att16_, blk.o, rows, EPI_STORE32, x, dim, 0, nullptr, x, dim, gate1, &L);
        ln_launch(blk, true);
    }
    gemm(act16_, blk.gate_up, rows, EPI_SWIGLU16, ff16_.p, ff, ldt, nullptr, nullptr, 0, nullptr, &L);
    if (fuse && next != nullptr) {
        const ln_fuse f1 = ln_params(*next, false);
        gemm(ff16_, blk.down, rows, EPI_STORE32_LN, x, dim, 0, nullptr, x, dim, gate2, &L, nullptr, 0, nullptr, 0, &f1);
        return true;
    }
    gemm(ff16_, blk.down, rows, EPI_STORE32, x, dim, 0, nullptr, x, dim, gate2, &L);
    return false;
}

// GroupNorm + SiLU -> 16-bit conv operand.  The per-(utterance, group) sums of slot `slot` either were accumulated by the
// epilogue of the GEMM that produced x (stats_ready) or come from one pass over x here.
void engine::group_norm_silu(const float * x, const seq_layout & L, int channels, int groups, const dev_norm & n,
                             int slot, bool stats_ready) {
    double * stats = gn_slot_ptr(slot);
    if (!stats_ready) { flip(); launch_gn_stats(x, channels, groups, L.d_tiles, L.n_tiles, stats, st_); }
    { flip(); launch_gn_apply(x, L.rows, channels, groups, stats, L.d_len, mdl->hp.group_norm_eps, n.w, n.b, L.d_row_utt, act16_.p,
                    mdl->conv_dtype, st_); }
}

// resnet_block_eval, /root/reference/src/miocodec-decoder.cpp:1293-1330 (x updated in place, y is scratch).
// in_ready: the producer of x already accumulated norm1's statistics into slot gn_slot; feeds_next: the next kernel is
// another resnet block's norm1 on this block's output, so conv2's epilogue accumulates them (returned as in_ready).
bool engine::run_resblock(float *& x, float *& y, const seq_layout & L, const resblock & rb, int & gn_slot, bool in_ready,
                          bool feeds_next) {
    static const bool fuse_env = [] { const char * e = std::getenv("MIO_TTS_B200_GN_FUSE"); return !(e && e[0] == '0'); }();
    const int c = rb.channels;
    const bool fuse = fuse_env && gemm_gn_stats_ok(c, rb.groups);
    const int s1 = gn_slot++, s2 = gn_slot++;
    group_norm_silu(x, L, c, rb.groups, rb.n1, s1, in_ready);
    gemm(act16_, rb.c1, L.rows, EPI_STORE32, y, c, 0, rb.c1.bias, nullptr, 0, nullptr, &L, nullptr, 0,
         fuse ? gn_slot_ptr(s2) : nullptr, rb.groups);
    group_norm_silu(y, L, c, rb.groups, rb.n2, s2, fuse);
    const bool next_ready = fuse && feeds_next;
    gemm(act16_, rb.c2, L.rows, EPI_STORE32, x, c, 0, rb.c2.bias, x, c, nullptr, &L, nullptr, 0,
         next_ready ? gn_slot_ptr(gn_slot) : nullptr, rb.groups);
    return next_ready;
}

// Small batches are launch-bound (141 launches, ~9 us of host work each, against ~5 us of GPU work): the second time the
// same batch shape (list of code counts + output format) comes by, its launch sequence is captured into a CUDA graph
// and replayed from then on.  Kernel arguments depend only on the shape and on workspace addresses, so growing any
// buffer drops the cached graphs; the per-batch tables reach the device through prepare()'s copy, outside the graph.
void engine::drop_graphs() {
    for (auto & kv : graphs_) if (kv.second.exec) cudaGraphExecDestroy(kv.second.exec);
    graphs_.clear();
}

void engine::run(bool with_istft, stage_times * prof, int stop_stage) {
    size_t codes = 0;
    for (int l : tok_.len) codes += (size_t) l;
    if (!graphs_ok_ || !with_istft || prof != nullptr || stop_stage != 0 || codes > graph_max_codes_ || force_generic_attention) {
        run_direct(with_istft, prof, stop_stage);
        return;
    }
    MIO_CUDA(cudaSetDevice(mdl->device));
    std::vector<int> key(tok_.len);
    key.push_back(out_pcm16 ? 1 : 0);
    gentry & g = graphs_[key];
    g.last_use = ++graph_clock_;
    if (g.exec != nullptr) {
        MIO_CUDA(cudaGraphLaunch(g.exec, st_));
        count_launch((int) g.kernels);
        return;
    }
    if (g.seen++ == 0) {             // first sight: plain launches (this also sizes every buffer the shape needs)
        if (graphs_.size() > 64) {   // forget the least recently used half
            std::vector<std::pair<unsigned long long, std::vector<int>>> order;
            for (auto & kv : graphs_) order.push_back({kv.second.last_use, kv.first});
            std::sort(order.begin(), order.end());
            for (size_t i = 0; i < order.size() / 2; ++i) {
                auto it = graphs_.find(order[i].second);
                if (it->second.exec) cudaGraphExecDestroy(it->second.exec);
                graphs_.erase(it);
            }
        }
        run_direct(with_istft, nullptr, 0);
        return;
    }
    const unsigned long long n0 = g_launch_count.load();
    cudaGraph_t graph = nullptr;
    cudaGraphExec_t exec = nullptr;
    bool ok = cudaStreamBeginCapture(st_, cudaStreamCaptureModeThreadLocal) == cudaSuccess;
    if (ok) {
        try {
            run_direct(with_istft, nullptr, 0);
        } catch (...) {
            ok = false;
        }
        if (cudaStreamEndCapture(st_, &graph) != cudaSuccess || graph == nullptr) ok = false;
    }
    if (ok && cudaGraphInstantiate(&exec, graph, 0) != cudaSuccess) ok = false;
    if (graph) cudaGraphDestroy(graph);
    if (!ok) {                       // capture is not available here: stay on plain launches for good
        cudaGetLastError();
        if (exec) cudaGraphExecDestroy(exec);
        graphs_ok_ = false;
        drop_graphs();
        g_launch_count.store(n0);
        run_direct(with_istft, nullptr, 0);
        return;
    }
    gentry & ge = graphs_[key];      // (run_direct cannot have dropped the map: every buffer already had its size)
    ge.exec = exec;
    ge.kernels = g_launch_count.load() - n0;
    MIO_CUDA(cudaGraphLaunch(exec, st_));
}

void engine::run_direct(bool with_istft, stage_times * prof, int stop_stage) {
    const hparams & hp = mdl->hp;
    MIO_CUDA(cudaSetDevice(mdl->device));
    prof_ = prof;
    prof_gemm_ = prof != nullptr;
    gemm_ev_flops_.clear();
    gemm_bytes_ = 0;
    gemm_ev_shape_.clear();
    attn_ev_.clear();
    const int pd = hp.prenet_dim, dd = hp.decoder_dim;
    const int ldt = mdl->linear_dtype;
    auto mark = [&](int i) { if (prof) MIO_CUDA(cudaEventRecord(ev_[i], st_)); };
    auto stop = [&](int stage, const void * p, int cols, int is16, int dtype, const seq_layout * lay) {
        if (stop_stage != stage) return false;
        dbg_ptr_ = p; dbg_cols_ = cols; dbg_is16_ = is16; dbg_dtype_ = dtype; dbg_lay_ = lay;
        prof_gemm_ = false; prof_ = nullptr;
        return true;
    };

    mark(0);
    if (hp.dynamic_global)
        launch_cond(mdl->cond_w, mdl->cond_b, d_emb_, (float *) cond_.p, mdl->cond_total, hp.decoder_adanorm_dim, n_utt_, st_);
    if (n_gn_slots_ > 0)
        MIO_CUDA(cudaMemsetAsync(gn_stats_.p, 0, (size_t) n_gn_slots_ * n_utt_ * gn_groups_max_ * 2 * sizeof(double), st_));
    MIO_CUDA(cudaMemsetAsync(peak_.p, 0, (size_t) n_utt_ * 4, st_));

    // ---- prenet: gather -> transformer blocks -> LN -> linear (src/miocodec-decoder.cpp:1357-1365) ----
    float * xt = (float *) xt_.p;
    launch_gather(mdl->tok_embd, d_codes_, xt, tok_.rows, pd, st_);
    {
        bool ln_done = false;
        for (size_t i = 0; i < mdl->prenet.size(); ++i)
            ln_done = run_tblock(xt, tok_, mdl->prenet[i], pd, hp.prenet_heads, hp.prenet_ff, hp.prenet_window, rope_p_, false, ln_done,
                                 i + 1 < mdl->prenet.size() ? &mdl->prenet[i + 1] : nullptr);
    }
    flip();
    launch_layernorm(xt, tok_.rows, pd, hp.norm_eps, mdl->prenet_norm.w, mdl->prenet_norm.b, nullptr, 0, 0, 0,
                     tok_.d_row_utt, act16_.p, ldt, nullptr, st_);
    gemm(act16_, mdl->prenet_out, tok_.rows, EPI_STORE16, ff16_.p, dd, ldt, mdl->prenet_out.bias, nullptr, 0, nullptr, &tok_);
    mark(1);
    if (stop(1, ff16_.p, dd, 1, ldt, &tok_)) return;

    // ---- x2 transposed conv as one GEMM + interpolate to the decoder frame count (:1368-1374) ----
    float * fx = (float *) fx_.p, * fy = (float *) fy_.p, * ft = (float *) ft_.p;
    gemm(ff16_, mdl->upsample, tok_.rows, EPI_STORE32, ft, 2 * dd, 0, mdl->upsample.bias, nullptr, 0, nullptr, &tok_);
    flip();
    launch_interp(ft, d_tok2_row0_, d_tok2_len_, fx, frm_.rows, dd, frm_.d_row_utt, frm_.d_row_pos, frm_.d_len, st_);
    if (stop(2, fx, dd, 0, 0, &frm_)) return;
    int gn_slot = 0;
    bool gn_ready = false;
    for (size_t j = 0; j < mdl->prior.size(); ++j)
        gn_ready = run_resblock(fx, fy, frm_, mdl->prior[j], gn_slot, gn_ready, j + 1 < mdl->prior.size());
    mark(2);
    if (stop(3, fx, dd, 0, 0, &frm_)) return;

    // ---- AdaLN-Zero decoder (:1382-1400) ----
    {
        bool ln_done = false;
        for (size_t i = 0; i < mdl->decoder.size(); ++i)
            ln_done = run_tblock(fx, frm_, mdl->decoder[i], dd, hp.decoder_heads, hp.decoder_ff, hp.decoder_window, rope_d_,
                                 hp.dynamic_global, ln_done, i + 1 < mdl->decoder.size() ? &mdl->decoder[i + 1] : nullptr);
    }
    if (hp.dynamic_global)
        { flip(); launch_layernorm(fx, frm_.rows, dd, hp.norm_eps, nullptr, nullptr, (const float *) cond_.p, mdl->cond_total,
                         mdl->cond_final_off, mdl->cond_final_off + dd, frm_.d_row_utt, nullptr, 0, fy, st_); }
    else
        { flip(); launch_layernorm(fx, frm_.rows, dd, hp.norm_eps, mdl->decoder_norm.w, mdl->decoder_norm.b, nullptr, 0, 0, 0,
                         frm_.d_row_utt, nullptr, 0, fy, st_); }
    std::swap(fx, fy);
    mark(3);
    if (stop(4, fx, dd, 0, 0, &frm_)) return;

    // ---- post ResNets, optional wave upsampler, iSTFT head (:1403-1461) ----
    gn_ready = false;
    for (size_t j = 0; j < mdl->post.size(); ++j)
        gn_ready = run_resblock(fx, fy, frm_, mdl->post[j], gn_slot, gn_ready, j + 1 < mdl->post.size());
    if (stop(5, fx, dd, 0, 0, &frm_)) return;
    const seq_layout * cur = &frm_;
    int cur_c = dd;
    const dbuf * head_in16 = &act16_;
    for (size_t i = 0; i < mdl->ups.size(); ++i) {
        const upstage & s = mdl->ups[i];
        const seq_layout & nxt = ups_[i];
        flip();
        if (s.poly) {
            // one implicit GEMM: 3 row-shifted taps of the input (gap rows zeroed: the taps read them as padding),
            // bias + SnakeBeta in the epilogue, output rows written straight into the upsampled layout
            launch_cast16_rows(fx, act16_.p, mdl->conv_dtype, cur->rows, cur_c, cur->d_row_utt, st_);
            snake_a_ = s.poly_snake_a; snake_ib_ = s.poly_snake_inv_b;
            gemm(act16_, s.up_poly, cur->rows, EPI_STORE32, fy, s.factor * s.cout, 0, s.poly_bias, nullptr, 0, nullptr, cur);
            snake_a_ = snake_ib_ = nullptr;
        } else {
            launch_cast16(fx, act16_.p, mdl->conv_dtype, (size_t) cur->rows * cur_c, st_);
            gemm(act16_, s.up, cur->rows, EPI_STORE32, ft, s.ksize * s.cout, 0, nullptr, nullptr, 0, nullptr, cur);
            launch_convt_fold_snake(ft, s.ksize * s.cout, cur->d_row0, cur->d_len, fy, nxt.rows, s.cout, s.ksize, s.factor,
                                    s.pad, s.up_bias, s.snake_a, s.snake_inv_b, nxt.d_row_utt, nxt.d_row_pos, st_);
        }
        std::swap(fx, fy);
        run_resblock(fx, fy, nxt, s.rb, gn_slot, false, false);
        cur = &nxt;
        cur_c = s.cout;
    }
    if (!mdl->ups.empty()) {
        flip();
        launch_cast16(fx, act16_.p, ldt, (size_t) cur->rows * cur_c, st_);
        static const bool fuse_snake = [] { const char * e = std::getenv("MIO_TTS_B200_CONVT_POLY"); return !(e && e[0] == '0'); }();
        if (fuse_snake) {
            // out_proj + bias + SnakeBeta -> the head's 16-bit operand in one epilogue (no fp32 round trip, no snake / cast passes)
            snake_a_ = mdl->out_snake_a_pad; snake_ib_ = mdl->out_snake_inv_b_pad;
            gemm(act16_, mdl->up_out_proj, cur->rows, EPI_STORE16, ff16_.p, mdl->head_in, ldt, mdl->up_out_proj.bias, nullptr, 0,
                 nullptr, cur);
            snake_a_ = snake_ib_ = nullptr;
            head_in16 = &ff16_;
        } else {
            gemm(act16_, mdl->up_out_proj, cur->rows, EPI_STORE32, ft, mdl->head_in, 0, mdl->up_out_proj.bias, nullptr, 0,
                 nullptr, cur);
            launch_snake(ft, cur->rows, mdl->head_in, mdl->out_snake_a, mdl->out_snake_inv_b, st_);
            flip();
            launch_cast16(ft, act16_.p, ldt, (size_t) cur->rows * mdl->head_in, st_);
        }
    } else {
        flip();
        launch_cast16(fx, act16_.p, ldt, (size_t) cur->rows * cur_c, st_);
    }
    gemm(*head_in16, mdl->head, cur->rows, EPI_STORE32, spec_buf_.p, spec_pitch_, 0, mdl->head.bias, nullptr, 0, nullptr, cur);
    mark(4);

    if (with_istft) {
        launch_istft(mdl->istft, (const float *) spec_buf_.p, spec_pitch_, spec_->d_row0, spec_->d_len, d_audio_off_,
                     (float *) audio_.p, (unsigned int *) peak_.p, n_utt_, spec_->max_len, st_);
        long long max_len = 0;
        for (auto l : audio_len_) max_len = std::max(max_len, l);
        if (out_pcm16) {   // WAV16 samples for the server's response instead of the normalised fp32 waveform
            ensure(pcm_, (size_t) audio_total_ * 2 + 256);
            launch_pcm16((const float *) audio_.p, d_audio_off_, d_audio_len_, (const unsigned int *) peak_.p,
                         (int16_t *) pcm_.p, n_utt_, max_len, st_);
        } else {
            launch_peak_normalize((float *) audio_.p, d_audio_off_, d_audio_len_, (const unsigned int *) peak_.p, n_utt_,
                                  max_len, st_);
        }
    }
    mark(5);

    if (prof) {
        MIO_CUDA(cudaStreamSynchronize(st_));
        auto ms = [&](int a, int b) { float t = 0; cudaEventElapsedTime(&t, ev_[a], ev_[b]); return t; };
        prof->prenet = ms(0, 1); prof->upsample_prior = ms(1, 2); prof->decoder = ms(2, 3);
        prof->post_head = ms(3, 4); prof->istft = ms(4, 5); prof->total = ms(0, 5);
        prof->gemm_ms = 0; prof->gemm_flops = 0; prof->gemm_bytes = gemm_bytes_; prof->gemm_launches = (int) gemm_ev_flops_.size();
        const bool log = std::getenv("MIO_TTS_B200_GEMM_LOG") != nullptr;
        for (size_t gi = 0; gi < gemm_ev_flops_.size(); ++gi) {
            auto & ef = gemm_ev_flops_[gi];
            float t = 0;
            cudaEventElapsedTime(&t, gemm_ev_[ef.first], gemm_ev_[ef.first + 1]);
            if (log) {
                const auto & sh = gemm_ev_shape_[gi];
                std::fprintf(stderr, "gemm M=%d N=%d K=%d taps=%d epi=%d  %.1f us  %.0f TFLOP/s\n", sh[0], sh[1], sh[2],
                             sh[3], sh[4], t * 1e3, ef.second / (t * 1e-3) / 1e12);
            }
            prof->gemm_ms += t;
            prof->gemm_flops += ef.second;
        }
        prof->attn_ms = 0; prof->attn_launches = (int) attn_ev_.size();
        for (size_t e : attn_ev_) { float t = 0; cudaEventElapsedTime(&t, gemm_ev_[e], gemm_ev_[e + 1]); prof->attn_ms += t; }
    }
    prof_gemm_ = false;
    prof_ = nullptr;
}

void engine::fetch_audio(size_t u, float * dst) {
    MIO_CUDA(cudaMemcpyAsync(dst, (const float *) audio_.p + audio_off_[u], (size_t) audio_len_[u] * sizeof(float),
                             cudaMemcpyDeviceToHost, st_));
}

void engine::fetch_pcm16(size_t u, int16_t * dst) {
    if (pcm_.p == nullptr) throw std::runtime_error("no WAV16 samples: the last run produced fp32 audio");
    MIO_CUDA(cudaMemcpyAsync(dst, (const int16_t *) pcm_.p + audio_off_[u], (size_t) audio_len_[u] * sizeof(int16_t),
                             cudaMemcpyDeviceToHost, st_));
}

// Test hook for pcm16_kernel alone: host waveforms in (un-normalised), int16 out; the peak that the iSTFT kernel
// would have left in peak_ is taken on the host here.
void engine::pcm16_host(const float * const * audio, const size_t * n, size_t batch, int16_t * const * out) {
    MIO_CUDA(cudaSetDevice(mdl->device));
    std::vector<long long> off(batch), len(batch);
    std::vector<unsigned int> peak(batch);
    long long total = 0, max_len = 0;
    for (size_t u = 0; u < batch; ++u) {
        off[u] = total; len[u] = (long long) n[u];
        total += (long long) round_up(n[u], 64);
        max_len = std::max(max_len, len[u]);
        float pk = 0.0f;
        for (size_t i = 0; i < n[u]; ++i) { const float a = std::fabs(audio[u][i]); if (std::isfinite(a)) pk = std::max(pk, a); }
        std::memcpy(&peak[u], &pk, 4);
    }
    ensure(audio_, (size_t) total * 4 + 256);
    ensure(pcm_, (size_t) total * 2 + 256);
    ensure(peak_, batch * 4 + 256);
    blob_.clear();
    const size_t o_off = blob_add(off.data(), batch * 8), o_len = blob_add(len.data(), batch * 8);
    ensure(d_blob_, blob_.size() + 256);
    MIO_CUDA(cudaMemcpyAsync(d_blob_.p, blob_.data(), blob_.size(), cudaMemcpyHostToDevice, st_));
    MIO_CUDA(cudaMemcpyAsync(peak_.p, peak.data(), batch * 4, cudaMemcpyHostToDevice, st_));
    for (size_t u = 0; u < batch; ++u)
        MIO_CUDA(cudaMemcpyAsync((float *) audio_.p + off[u], audio[u], n[u] * 4, cudaMemcpyHostToDevice, st_));
    const uint8_t * b = (const uint8_t *) d_blob_.p;
    launch_pcm16((const float *) audio_.p, (const long long *) (b + o_off), (const long long *) (b + o_len),
                 (const unsigned int *) peak_.p, (int16_t *) pcm_.p, (int) batch, max_len, st_);
    for (size_t u = 0; u < batch; ++u)
        MIO_CUDA(cudaMemcpyAsync(out[u], (const int16_t *) pcm_.p + off[u], n[u] * 2, cudaMemcpyDeviceToHost, st_));
    MIO_CUDA(cudaStreamSynchronize(st_));
}

void engine::fetch_debug(std::vector<float> & out, int & rows, int & cols) {
    if (dbg_ptr_ == nullptr || dbg_lay_ == nullptr) throw std::runtime_error("no debug stage captured");
    rows = dbg_lay_->len[0];
    cols = dbg_cols_;
    out.resize((size_t) rows * cols);
    const size_t es = dbg_is16_ ? 2 : 4;
    std::vector<uint8_t> raw((size_t) rows * cols * es);
    MIO_CUDA(cudaMemcpyAsync(raw.data(), (const uint8_t *) dbg_ptr_ + (size_t) dbg_lay_->row0[0] * cols * es, raw.size(),
                             cudaMemcpyDeviceToHost, st_));
    sync();
    if (!dbg_is16_) {
        std::memcpy(out.data(), raw.data(), raw.size());
    } else {
        const uint16_t * h = (const uint16_t *) raw.data();
        for (size_t i = 0; i < out.size(); ++i) {
            if (dbg_dtype_ == OP_BF16) { uint32_t u = (uint32_t) h[i] << 16; std::memcpy(&out[i], &u, 4); }
            else {   // fp16 -> fp32
                const uint32_t s = (h[i] >> 15) & 1, e = (h[i] >> 10) & 0x1F, m = h[i] & 0x3FF;
                float v;
                if (e == 0) v = std::ldexp((float) m, -24);
                else if (e == 31) v = m ? NAN : INFINITY;
                else v = std::ldexp((float) (m | 0x400), (int) e - 25);
                out[i] = s ? -v : v;
            }
        }
    }
}

void engine::fetch_spec(size_t u, float * dst) {
    const int bins = spec_bins();
    MIO_CUDA(cudaMemcpy2DAsync(dst, (size_t) bins * 4, (const float *) spec_buf_.p + (size_t) spec_->row0[u] * spec_pitch_,
                               (size_t) spec_pitch_ * 4, (size_t) bins * 4, (size_t) spec_->len[u], cudaMemcpyDeviceToHost,
                               st_));
}

// -------------------------------------------------------------------------------------------------
// stand-alone stages
// -------------------------------------------------------------------------------------------------
void engine::gather(const int32_t * codes, size_t n, float * out) {
    MIO_CUDA(cudaSetDevice(mdl->device));
    if (codes == nullptr || n == 0) throw std::runtime_error("codes are empty");
    for (size_t i = 0; i < n; ++i)
        if (codes[i] < k_code_min || codes[i] > k_code_max || codes[i] >= mdl->n_codebook) throw std::runtime_error("code id out of range");
    const int pd = mdl->hp.prenet_dim;
    ensure(xt_, pad_rows((int) n) * pd * 4);
    ensure(d_blob_, n * 4 + 256);
    MIO_CUDA(cudaMemcpyAsync(d_blob_.p, codes, n * 4, cudaMemcpyHostToDevice, st_));
    launch_gather(mdl->tok_embd, (const int *) d_blob_.p, (float *) xt_.p, (int) n, pd, st_);
    MIO_CUDA(cudaMemcpyAsync(out, xt_.p, n * pd * 4, cudaMemcpyDeviceToHost, st_));
    sync();
}

void engine::istft_host(const float * spec, int n_frames, int n_fft, int hop, std::vector<float> & audio) {
    MIO_CUDA(cudaSetDevice(mdl->device));
    audio.clear();
    // embd_to_audio guards, src/mio-tts-lib.cpp:632-637
    if (spec == nullptr || n_frames <= 0 || n_fft <= 0 || hop <= 0) return;
    auto key = std::make_pair(n_fft, hop);
    if (!istft_plans_.count(key)) {
        istft_plan p;
        std::string e;
        if (!istft_plan_init(p, n_fft, hop, e)) throw std::runtime_error(e);
        istft_plans_[key] = p;
    }
    const istft_plan & p = istft_plans_[key];
    const int bins = n_fft + 2;
    const int pitch = (int) round_up((size_t) bins, 4);
    const long long n_full = (long long) (n_frames - 1) * hop + n_fft;
    const long long n_pad = std::max(0, (n_fft - hop) / 2);
    const long long n_out = std::max<long long>(n_pad, n_full - n_pad) - n_pad;
    if (n_out <= 0) return;
    ensure(spec_buf_, (size_t) n_frames * pitch * 4);
    ensure(audio_, (size_t) n_out * 4 + 256);
    ensure(d_blob_, 1024);
    struct { int row0; int frames; long long off; } meta = {0, n_frames, 0};
    MIO_CUDA(cudaMemcpyAsync(d_blob_.p, &meta, sizeof(meta), cudaMemcpyHostToDevice, st_));
    MIO_CUDA(cudaMemcpy2DAsync(spec_buf_.p, (size_t) pitch * 4, spec, (size_t) bins * 4, (size_t) bins * 4, (size_t) n_frames,
                               cudaMemcpyHostToDevice, st_));
    const uint8_t * b = (const uint8_t *) d_blob_.p;
    launch_istft(p, (const float *) spec_buf_.p, pitch, (const int *) b, (const int *) (b + 4), (const long long *) (b + 8),
                 (float *) audio_.p, nullptr, 1, n_frames, st_);
    audio.resize((size_t) n_out);
    MIO_CUDA(cudaMemcpyAsync(audio.data(), audio_.p, (size_t) n_out * 4, cudaMemcpyDeviceToHost, st_));
    sync();
}

// the codebook gather alone over the prepared batch (north_star: achieved HBM GB/s of the gather)
float engine::gather_bench(int iters) {
    MIO_CUDA(cudaSetDevice(mdl->device));
    if (n_utt_ == 0) throw std::runtime_error("no prepared batch");
    auto go = [&]() { launch_gather(mdl->tok_embd, d_codes_, (float *) xt_.p, tok_.rows, mdl->hp.prenet_dim, st_); };
    go();
    MIO_CUDA(cudaEventRecord(ev_[6], st_));
    for (int i = 0; i < iters; ++i) go();
    MIO_CUDA(cudaEventRecord(ev_[7], st_));
    MIO_CUDA(cudaStreamSynchronize(st_));
    float ms = 0;
    MIO_CUDA(cudaEventElapsedTime(&ms, ev_[6], ev_[7]));
    return ms / (float) std::max(1, iters);
}

float engine::istft_bench(int n_utt, int frames, int n_fft, int hop, int iters, unsigned long long seed, double * checksum) {
    MIO_CUDA(cudaSetDevice(mdl->device));
    auto key = std::make_pair(n_fft, hop);
    if (!istft_plans_.count(key)) {
        istft_plan p;
        std::string e;
        if (!istft_plan_init(p, n_fft, hop, e)) throw std::runtime_error(e);
        istft_plans_[key] = p;
    }
    const istft_plan & p = istft_plans_[key];
    const int bins = n_fft + 2;
    const int pitch = (int) round_up((size_t) bins, 4);
    const long long rows = (long long) n_utt * frames;
    const long long n_pad = std::max(0, (n_fft - hop) / 2);
    const long long n_out = (long long) (frames - 1) * hop + n_fft - 2 * n_pad;
    const long long out_stride = (long long) round_up((size_t) n_out, 64);
    ensure(spec_buf_, (size_t) rows * pitch * 4);
    ensure(audio_, (size_t) out_stride * n_utt * 4 + 256);
    std::vector<int> row0((size_t) n_utt), len((size_t) n_utt, frames);
    std::vector<long long> off((size_t) n_utt);
    for (int u = 0; u < n_utt; ++u) { row0[(size_t) u] = u * frames; off[(size_t) u] = (long long) u * out_stride; }
    blob_.clear();
    const size_t o_r = blob_add(row0.data(), row0.size() * 4);
    const size_t o_l = blob_add(len.data(), len.size() * 4);
    const size_t o_o = blob_add(off.data(), off.size() * 8);
    ensure(d_blob_, blob_.size() + 256);
    MIO_CUDA(cudaMemcpyAsync(d_blob_.p, blob_.data(), blob_.size(), cudaMemcpyHostToDevice, st_));
    MIO_CUDA(cudaStreamSynchronize(st_));
    launch_fill_spec((float *) spec_buf_.p, rows, pitch, bins / 2, seed, st_);
    const uint8_t * b = (const uint8_t *) d_blob_.p;
    auto go = [&]() {
        launch_istft(p, (const float *) spec_buf_.p, pitch, (const int *) (b + o_r), (const int *) (b + o_l),
                     (const long long *) (b + o_o), (float *) audio_.p, nullptr, n_utt, frames, st_);
    };
    go();   // warm-up
    MIO_CUDA(cudaEventRecord(ev_[6], st_));
    for (int i = 0; i < iters; ++i) go();
    MIO_CUDA(cudaEventRecord(ev_[7], st_));
    MIO_CUDA(cudaStreamSynchronize(st_));
    float ms = 0;
    MIO_CUDA(cudaEventElapsedTime(&ms, ev_[6], ev_[7]));
    if (checksum) *checksum = device_sum((const float *) audio_.p, out_stride * n_utt, st_);
    return ms / (float) std::max(1, iters);
}

} // namespace mio
