// Micro-batcher in front of the batched decode.
//
// The reference server gives every slot on a device a clone of one context and serialises their mio_tts_synthesize calls
// on the owner's runtime mutex (/root/reference/src/tts-mio-server.cpp:2659-2670, 2724-2736; src/mio-tts-lib.cpp:1258-1262),
// so `--parallel N` on one GPU decodes N utterances one after another.  Here the clones of a context share ONE batcher:
// concurrent mio_tts_synthesize calls are queued, a worker thread takes everything that is waiting (up to max_batch
// utterances), decodes it as one batch on its own engine (stream + workspace) and hands every caller its own buffer.
// A worker is admitted only while the kernels already running cover fewer than ~4 k codes: small batches overlap on the
// GPU like plain clones would; once it is saturated, arrivals pile up and the next admitted worker takes them all, so the
// batch size follows the load without a timer.
// Results are bit-identical to un-batched calls (no kernel mixes utterances: tests/test_parity_gpu.py batch invariance).
// Four workers by default: the D2H copy of one batch hides under the kernels of the next.
#pragma once
#include "engine.h"

#include <condition_variable>
#include <deque>
#include <functional>
#include <string>
#include <thread>

namespace mio {

struct decode_request {
    const int32_t * codes = nullptr; size_t n_codes = 0;
    const float * emb = nullptr; size_t n_emb = 0;
    bool wav16 = false;         // false: fp32 waveform; true: 44-byte RIFF/WAVE header + int16 samples
    void * out = nullptr;       // result-pool buffer (release with result_free / mio_tts_audio_free)
    size_t n_samples = 0;
    bool ok = false, done = false;
    std::string err;
};

// One batch on one engine: prepare -> run -> pinned result buffers.  Throws std::runtime_error (nothing is leaked).
// kernels_done (optional) is called once the batch's kernels have finished, before the device -> host copies start.
void decode_batch(engine & E, decode_request * const * reqs, size_t batch, const std::function<void()> & kernels_done = nullptr);
// The same for one large batch of a single caller, as two halves on two engines so that the first half's device -> host
// copies hide under the second half's kernels (bit-identical results; see batcher.cu)
void decode_batch_pipelined(engine & E1, engine & E2, decode_request * const * reqs, size_t batch);
void wav16_header(uint8_t * p, uint32_t sample_rate, uint32_t pcm_bytes);   // tts-mio-server.cpp:2013-2034, lib.cpp:34-48

struct batcher_stats { uint64_t requests = 0, batches = 0, largest = 0, retried_single = 0; };

class batcher {
public:
    batcher(std::shared_ptr<model> m, int max_batch, int max_wait_us, int workers);
    ~batcher();
    void submit(decode_request & r);          // blocks until r.done; r.ok / r.err / r.out say how it went
    batcher_stats stats();
    void reserve(int n_codes);                // size every worker's workspace for max_batch utterances of n_codes
    const int max_batch, max_wait_us, workers;
    const std::shared_ptr<model> mdl;

private:
    void loop(int w);
    std::mutex m_;
    std::condition_variable cv_work_, cv_done_;
    std::deque<decode_request *> q_;
    std::vector<std::unique_ptr<engine>> eng_;
    std::vector<std::thread> th_;
    batcher_stats st_;
    size_t expect_ = 1;          // callers released by the batch that finished last: a gathering worker waits (bounded) for them
    size_t busy_codes_ = 0;      // codes of the batches whose kernels are running
    size_t sat_codes_ = 4096;    // admission threshold: ~10 utterances of 400 codes keep the GPU busy
    size_t max_codes_ = 1 << 16; // bound on the codes of one batch (workspace ~ 100 KB per code: 6.5 GB per worker at most)
    bool stop_ = false;
};

} // namespace mio
