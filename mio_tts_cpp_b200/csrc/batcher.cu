// Micro-batcher (see batcher.h) and the shared "one batch -> pinned result buffers" step of the C API.
#include "batcher.h"

#include <algorithm>
#include <chrono>
#include <functional>
#include <cstring>

namespace mio {

void wav16_header(uint8_t * p, uint32_t sample_rate, uint32_t pcm_bytes) {
    auto u32 = [&](int o, uint32_t v) { std::memcpy(p + o, &v, 4); };
    auto u16 = [&](int o, uint16_t v) { std::memcpy(p + o, &v, 2); };
    std::memcpy(p, "RIFF", 4); u32(4, 36 + pcm_bytes); std::memcpy(p + 8, "WAVE", 4); std::memcpy(p + 12, "fmt ", 4);
    u32(16, 16); u16(20, 1); u16(22, 1); u32(24, sample_rate); u32(28, sample_rate * 2); u16(32, 2); u16(34, 16);
    std::memcpy(p + 36, "data", 4); u32(40, pcm_bytes);
}

void decode_batch(engine & E, decode_request * const * reqs, size_t batch, const std::function<void()> & kernels_done) {
    std::vector<const int32_t *> codes(batch);
    std::vector<size_t> n(batch);
    std::vector<const float *> embs(batch);
    const bool has_emb = reqs[0]->emb != nullptr && reqs[0]->n_emb > 0;
    const bool wav16 = reqs[0]->wav16;
    for (size_t u = 0; u < batch; ++u) { codes[u] = reqs[u]->codes; n[u] = reqs[u]->n_codes; embs[u] = reqs[u]->emb; }
    std::vector<void *> bufs(batch, nullptr);
    std::lock_guard<std::mutex> lk(E.mtx);
    try {
        E.out_pcm16 = wav16;
        E.prepare(codes.data(), n.data(), has_emb ? embs.data() : nullptr, has_emb ? reqs[0]->n_emb : 0, batch);
        E.run(true);
        if (kernels_done) {          // the batcher hands the GPU to its other worker here: our D2H runs under its kernels
            E.sync();
            kernels_done();
        }
        for (size_t u = 0; u < batch; ++u) {
            const size_t ns = E.audio_len(u);
            bufs[u] = wav16 ? (void *) result_alloc((44 + 2 * ns + 3) / 4) : (void *) result_alloc(ns);
            if (bufs[u] == nullptr) throw std::runtime_error("failed to allocate output audio");
            if (wav16) {
                wav16_header((uint8_t *) bufs[u], (uint32_t) E.mdl->hp.sample_rate, (uint32_t) (2 * ns));
                E.fetch_pcm16(u, (int16_t *) ((uint8_t *) bufs[u] + 44));
            } else {
                E.fetch_audio(u, (float *) bufs[u]);
            }
        }
        E.sync();
        E.out_pcm16 = false;
    } catch (...) {
        E.out_pcm16 = false;
        try { E.sync(); } catch (...) {}
        if (kernels_done) kernels_done();
        for (void * p : bufs) if (p && !result_free(p)) std::free(p);
        throw;
    }
    for (size_t u = 0; u < batch; ++u) { reqs[u]->out = bufs[u]; reqs[u]->n_samples = E.audio_len(u); reqs[u]->ok = true; }
}

// One large batch from ONE caller, as two parts on two engines (streams + workspaces): the second half's kernels are
// ordered behind the first half's by an event, and the first half's device -> host copies run on its own stream under
// them.  The cut sits at the utterance where 70 % of the codes is reached; no kernel mixes utterances, so every
// waveform is bit-identical to the single-batch path.
void decode_batch_pipelined(engine & E1, engine & E2, decode_request * const * reqs, size_t batch) {
    if (batch < 2) { decode_batch(E1, reqs, batch); return; }
    size_t total = 0, acc = 0, cut = batch / 2;
    for (size_t u = 0; u < batch; ++u) total += reqs[u]->n_codes;
    // the first part's copies hide under the second part's kernels, the second part's copies are exposed: the cut sits where
    // `split` percent of the codes is reached (MIO_TTS_B200_PIPELINE_SPLIT, default 70: single caller 101.5 k -> 105.8 k
    // audio-s/s against an even cut, alternated in one call; 70 / 80 / 85 are equal within noise)
    static const size_t split = [] { const char * e = std::getenv("MIO_TTS_B200_PIPELINE_SPLIT"); const int v = e ? std::atoi(e) : 70;
                                     return (size_t) (v >= 10 && v <= 90 ? v : 70); }();
    for (size_t u = 0; u < batch; ++u) { acc += reqs[u]->n_codes; if (100 * acc >= split * total) { cut = u + 1; break; } }
    cut = std::min(std::max<size_t>(cut, 1), batch - 1);
    std::vector<const int32_t *> codes(batch);
    std::vector<size_t> n(batch);
    std::vector<const float *> embs(batch);
    const bool has_emb = reqs[0]->emb != nullptr && reqs[0]->n_emb > 0;
    const bool wav16 = reqs[0]->wav16;
    for (size_t u = 0; u < batch; ++u) { codes[u] = reqs[u]->codes; n[u] = reqs[u]->n_codes; embs[u] = reqs[u]->emb; }
    std::vector<void *> bufs(batch, nullptr);
    std::vector<size_t> ns(batch, 0);
    std::lock(E1.mtx, E2.mtx);
    std::lock_guard<std::mutex> lk1(E1.mtx, std::adopt_lock), lk2(E2.mtx, std::adopt_lock);
    engine * eng[2] = {&E1, &E2};
    const size_t lo[2] = {0, cut}, hi[2] = {cut, batch};
    cudaEvent_t ev = nullptr;
    try {
        MIO_CUDA(cudaSetDevice(E1.mdl->device));
        MIO_CUDA(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
        for (int h = 0; h < 2; ++h) {
            engine & E = *eng[h];
            E.out_pcm16 = wav16;
            E.prepare(codes.data() + lo[h], n.data() + lo[h], has_emb ? embs.data() + lo[h] : nullptr, has_emb ? reqs[0]->n_emb : 0,
                      hi[h] - lo[h]);
            if (h == 1) MIO_CUDA(cudaStreamWaitEvent(E.stream(), ev, 0));     // kernels of half 2 behind the kernels of half 1
            E.run(true);
            if (h == 0) MIO_CUDA(cudaEventRecord(ev, E.stream()));            // ... but not behind its copies
            for (size_t u = lo[h]; u < hi[h]; ++u) {
                ns[u] = E.audio_len(u - lo[h]);
                bufs[u] = wav16 ? (void *) result_alloc((44 + 2 * ns[u] + 3) / 4) : (void *) result_alloc(ns[u]);
                if (bufs[u] == nullptr) throw std::runtime_error("failed to allocate output audio");
                if (wav16) {
                    wav16_header((uint8_t *) bufs[u], (uint32_t) E.mdl->hp.sample_rate, (uint32_t) (2 * ns[u]));
                    E.fetch_pcm16(u - lo[h], (int16_t *) ((uint8_t *) bufs[u] + 44));
                } else {
                    E.fetch_audio(u - lo[h], (float *) bufs[u]);
                }
            }
        }
        E1.sync();
        E2.sync();
        E1.out_pcm16 = E2.out_pcm16 = false;
        cudaEventDestroy(ev);
    } catch (...) {
        E1.out_pcm16 = E2.out_pcm16 = false;
        try { E1.sync(); } catch (...) {}
        try { E2.sync(); } catch (...) {}
        if (ev) cudaEventDestroy(ev);
        for (void * p : bufs) if (p && !result_free(p)) std::free(p);
        throw;
    }
    for (size_t u = 0; u < batch; ++u) { reqs[u]->out = bufs[u]; reqs[u]->n_samples = ns[u]; reqs[u]->ok = true; }
}

batcher::batcher(std::shared_ptr<model> m, int max_batch_, int max_wait_us_, int workers_)
    : max_batch(std::max(1, max_batch_)), max_wait_us(std::max(0, max_wait_us_)), workers(std::min(8, std::max(1, workers_))),
      mdl(std::move(m)) {
    for (int w = 0; w < workers; ++w) eng_.emplace_back(new engine(mdl));
    for (int w = 0; w < workers; ++w) th_.emplace_back([this, w] { loop(w); });
}

batcher::~batcher() {
    {
        std::lock_guard<std::mutex> lk(m_);
        stop_ = true;
    }
    cv_work_.notify_all();
    for (auto & t : th_) t.join();
}

void batcher::reserve(int n_codes) {
    // at most 32 k codes per worker up front (~3 GB each): longer batches still grow the workspace on demand
    const int n = std::max(1, n_codes);
    const int batch = std::max(1, std::min(max_batch, 32768 / n));
    for (auto & e : eng_) {
        std::lock_guard<std::mutex> lk(e->mtx);
        e->reserve(n, batch);
    }
}

batcher_stats batcher::stats() {
    std::lock_guard<std::mutex> lk(m_);
    return st_;
}

void batcher::submit(decode_request & r) {
    std::unique_lock<std::mutex> lk(m_);
    if (stop_) { r.err = "batcher is shutting down"; r.done = true; return; }
    q_.push_back(&r);
    cv_work_.notify_all();
    cv_done_.wait(lk, [&] { return r.done; });
}

void batcher::loop(int w) {
    engine & E = *eng_[(size_t) w];
    std::vector<decode_request *> take;
    for (;;) {
        take.clear();
        size_t codes = 0;
        {
            // Admission: a worker starts a batch only while the kernels already running cover fewer than sat_codes_ codes.
            // Small batches are launch-latency bound and overlap well on the GPU, so they run side by side (never slower
            // than plain clones); once the GPU is saturated, requests pile up in the queue and the next admitted worker
            // takes all of them -- the batch size follows the load with no timer involved.
            std::unique_lock<std::mutex> lk(m_);
            cv_work_.wait(lk, [&] { return stop_ || (!q_.empty() && busy_codes_ < sat_codes_); });
            if (q_.empty()) { if (stop_) return; continue; }
            // Callers released together by a finished batch come back within microseconds of each other: give them a
            // bounded moment so the first one back is not peeled off alone.
            const size_t want = std::min<size_t>((size_t) max_batch, std::max<size_t>(expect_, 1));
            if (max_wait_us > 0 && q_.size() < want && !stop_)
                cv_work_.wait_for(lk, std::chrono::microseconds(max_wait_us), [&] { return stop_ || q_.size() >= want; });
            if (q_.empty()) continue;      // another worker took them during the window
            decode_request * head = q_.front();
            const bool he = head->emb != nullptr && head->n_emb > 0;
            for (auto it = q_.begin(); it != q_.end() && take.size() < (size_t) max_batch;) {
                decode_request * r = *it;
                const bool re = r->emb != nullptr && r->n_emb > 0;
                const bool fits = take.empty() || codes + r->n_codes <= max_codes_;
                if (re == he && (!re || r->n_emb == head->n_emb) && r->wav16 == head->wav16 && fits) {
                    take.push_back(r);
                    codes += r->n_codes;
                    it = q_.erase(it);
                } else {
                    ++it;
                }
            }
            st_.requests += take.size();
            st_.batches += 1;
            st_.largest = std::max<uint64_t>(st_.largest, take.size());
            busy_codes_ += codes;
            expect_ = 1;                   // until a batch finishes and tells how many callers are on their way back
        }
        bool released = false;
        auto release = [&] {               // kernels done: admit the next batch; our D2H copies run under its kernels
            if (released) return;
            released = true;
            { std::lock_guard<std::mutex> lk(m_); busy_codes_ -= codes; }
            cv_work_.notify_all();
        };
        bool whole = true;
        try {
            decode_batch(E, take.data(), take.size(), release);
        } catch (const std::exception & e) {
            whole = false;
            if (take.size() == 1) take[0]->err = e.what();
        } catch (...) {
            whole = false;
            if (take.size() == 1) take[0]->err = "unknown error";
        }
        release();
        if (!whole && take.size() > 1) {
            // one bad request (e.g. an out-of-range code) must not fail its neighbours: decode each on its own so
            // every caller gets exactly the error (or the audio) an un-batched call would have given
            for (decode_request * r : take) {
                try {
                    decode_batch(E, &r, 1);
                } catch (const std::exception & e) {
                    r->err = e.what();
                } catch (...) {
                    r->err = "unknown error";
                }
            }
            std::lock_guard<std::mutex> lk(m_);
            st_.retried_single += take.size();
        }
        {
            std::lock_guard<std::mutex> lk(m_);
            for (decode_request * r : take) r->done = true;
            expect_ = take.size();
        }
        cv_done_.notify_all();
    }
}

} // namespace mio
