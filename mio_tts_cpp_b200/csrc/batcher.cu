// Micro-batcher (see batcher.h) and the shared "one batch -> pinned result buffers" step of the C API.
#include "batcher.h"

#include <algorithm>
#include <chrono>
#include <cstring>

namespace mio {

void wav16_header(uint8_t * p, uint32_t sample_rate, uint32_t pcm_bytes) {
    auto u32 = [&](int o, uint32_t v) { std::memcpy(p + o, &v, 4); };
    auto u16 = [&](int o, uint16_t v) { std::memcpy(p + o, &v, 2); };
    std::memcpy(p, "RIFF", 4); u32(4, 36 + pcm_bytes); std::memcpy(p + 8, "WAVE", 4); std::memcpy(p + 12, "fmt ", 4);
    u32(16, 16); u16(20, 1); u16(22, 1); u32(24, sample_rate); u32(28, sample_rate * 2); u16(32, 2); u16(34, 16);
    std::memcpy(p + 36, "data", 4); u32(40, pcm_bytes);
}

void decode_batch(engine & E, decode_request * const * reqs, size_t batch) {
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
        for (void * p : bufs) if (p && !result_free(p)) std::free(p);
        throw;
    }
    for (size_t u = 0; u < batch; ++u) { reqs[u]->out = bufs[u]; reqs[u]->n_samples = E.audio_len(u); reqs[u]->ok = true; }
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

batcher_stats batcher::stats() {
    std::lock_guard<std::mutex> lk(m_);
    return st_;
}

void batcher::submit(decode_request & r) {
    std::unique_lock<std::mutex> lk(m_);
    if (stop_) { r.err = "batcher is shutting down"; r.done = true; return; }
    q_.push_back(&r);
    cv_work_.notify_one();
    cv_done_.wait(lk, [&] { return r.done; });
}

void batcher::loop(int w) {
    engine & E = *eng_[(size_t) w];
    std::vector<decode_request *> take;
    for (;;) {
        take.clear();
        {
            std::unique_lock<std::mutex> lk(m_);
            cv_work_.wait(lk, [&] { return stop_ || !q_.empty(); });
            if (q_.empty()) return;   // stop_ and drained
            // Gathering window: callers released by the previous batch come back within microseconds of each other;
            // wait (bounded) until as many as last time are queued instead of peeling the first one off alone.
            const size_t want = std::min<size_t>((size_t) max_batch, std::max<size_t>(expect_, 1));
            if (max_wait_us > 0 && q_.size() < want)
                cv_work_.wait_for(lk, std::chrono::microseconds(max_wait_us), [&] { return stop_ || q_.size() >= want; });
            if (q_.empty()) continue;   // another worker took them during the window
            decode_request * head = q_.front();
            const bool he = head->emb != nullptr && head->n_emb > 0;
            size_t codes = 0;
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
            expect_ = take.size();
            if (!q_.empty()) cv_work_.notify_one();   // incompatible or overflow requests: next worker
        }
        bool whole = true;
        try {
            decode_batch(E, take.data(), take.size());
        } catch (const std::exception & e) {
            whole = false;
            if (take.size() == 1) take[0]->err = e.what();
        } catch (...) {
            whole = false;
            if (take.size() == 1) take[0]->err = "unknown error";
        }
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
        }
        cv_done_.notify_all();
    }
}

} // namespace mio
