// Server-slot load generator over the C ABI (SURVEY 8f N3 / BASELINE config 5 on one GPU): `slots` host threads, each
// with its own context clone like the reference server's workers (/root/reference/src/tts-mio-server.cpp:2724-2736),
// call mio_tts_synthesize (or the WAV16 entry point) on 400 codes in a closed loop.
//
//   g++ -O2 -std=c++17 -Iinclude tools/slots_bench.cpp -o gpurun_out/slots_bench -Lmio_tts_cpp_b200 -lmio_tts_b200 \
//       -Wl,-rpath,$PWD/mio_tts_cpp_b200 -lpthread
//   gpurun_out/slots_bench model.gguf [--devices N] [slots ...]
#include "mio_tts_b200.h"

#include <atomic>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <random>
#include <string>
#include <thread>
#include <vector>

// owners: one context per device; slot i clones owners[i % n_owners] (the reference server's round-robin, :2667-2670)
static double run(const std::vector<mio_tts_context *> & owners, int n_slots, int rounds, bool wav16, double * ms_per_request) {
    char err[512] = {0};
    std::vector<mio_tts_context *> slots((size_t) n_slots);
    std::vector<std::vector<int32_t>> work((size_t) n_slots);
    std::vector<float> emb(128);
    std::mt19937 rng(7);
    for (auto & e : emb) e = std::normal_distribution<float>(0.0f, 0.1f)(rng);
    for (int i = 0; i < n_slots; ++i) {
        slots[(size_t) i] = mio_tts_context_clone(owners[(size_t) i % owners.size()], err, sizeof(err));
        if (!slots[(size_t) i]) { std::fprintf(stderr, "clone: %s\n", err); std::exit(1); }
        work[(size_t) i].resize(400);
        for (auto & c : work[(size_t) i]) c = (int32_t) (rng() % 12800);
    }
    std::atomic<int> ready{0}, failures{0};
    std::atomic<bool> go{false};
    const mio_tts_params params = mio_tts_default_params();
    auto call = [&](int i) {
        char e[256];
        const int32_t * codes = work[(size_t) i].data();
        const size_t n = work[(size_t) i].size();
        if (wav16) {
            const float * ep = emb.data();
            uint8_t * wav = nullptr; size_t bytes = 0;
            if (!mio_tts_b200_synthesize_wav16_batch(slots[(size_t) i], &codes, &n, &ep, emb.size(), 1, params, &wav, &bytes, nullptr, e, sizeof(e))) { failures++; return; }
            mio_tts_b200_wav_free(wav);
        } else {
            float * audio = nullptr; size_t na = 0; int32_t sr = 0;
            if (!mio_tts_synthesize(slots[(size_t) i], codes, n, emb.data(), emb.size(), params, &audio, &na, &sr, e, sizeof(e))) { failures++; return; }
            mio_tts_audio_free(audio);
        }
    };
    std::vector<std::thread> th;
    for (int i = 0; i < n_slots; ++i)
        th.emplace_back([&, i] {
            for (int r = 0; r < 5; ++r) call(i);      // warm-up: workspaces of the engines grow to their high-water mark
            ready++;
            while (!go.load()) std::this_thread::yield();
            for (int r = 0; r < rounds; ++r) call(i);
        });
    while (ready.load() < n_slots) std::this_thread::yield();
    const auto t0 = std::chrono::steady_clock::now();
    go = true;
    for (auto & t : th) t.join();
    const double dt = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    for (auto * s : slots) mio_tts_free(s);
    if (failures.load()) { std::fprintf(stderr, "%d calls failed\n", failures.load()); std::exit(1); }
    *ms_per_request = 1e3 * dt / rounds;
    return (double) n_slots * rounds * 16.0 / dt;    // 400 codes = 16 s of audio at 24 kHz
}

int main(int argc, char ** argv) {
    if (argc < 2) { std::fprintf(stderr, "usage: %s model.gguf [--devices N] [slots ...]\n", argv[0]); return 2; }
    char err[512] = {0};
    int n_dev = 1;
    std::vector<int> counts;
    for (int i = 2; i < argc; ++i) {
        if (std::string(argv[i]) == "--devices" && i + 1 < argc) n_dev = std::max(1, std::atoi(argv[++i]));
        else counts.push_back(std::atoi(argv[i]));
    }
    const auto t0 = std::chrono::steady_clock::now();
    std::vector<mio_tts_context *> owners;
    for (int d = 0; d < n_dev; ++d) {
        mio_tts_context * ctx = mio_tts_init_from_file(argv[1], nullptr, err, sizeof(err));
        if (!ctx) { std::fprintf(stderr, "init: %s\n", err); return 1; }
        const std::string dev = "CUDA" + std::to_string(d);
        if (!mio_tts_context_set_backend_device(ctx, dev.c_str(), err, sizeof(err))) { std::fprintf(stderr, "%s: %s\n", dev.c_str(), err); return 1; }
        owners.push_back(ctx);
    }
    const double load_s = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    if (counts.empty()) counts = {1, 2, 4, 8, 16, 32, 64};
    auto batching = [&](int max_batch) {
        for (auto * c : owners) {
            mio_tts_b200_context_enable_batching(c, max_batch, 300, 4, err, sizeof(err));
            if (max_batch > 0) mio_tts_reserve_workspace(c, 400, err, sizeof(err));
        }
    };
    std::printf("{\"load_s\": %.3f, \"devices\": %d, \"rows\": [\n", load_s, n_dev);
    for (size_t ci = 0; ci < counts.size(); ++ci) {
        const int n = counts[ci], rounds = std::max(50, 2000 / n);
        double ms[3], v[3];
        batching(0);
        v[0] = run(owners, n, rounds, false, &ms[0]);
        batching(64);
        v[1] = run(owners, n, rounds, false, &ms[1]);
        uint64_t rq = 0, bt = 0, lg = 0, rs = 0;
        for (auto * c : owners) {
            uint64_t a = 0, b = 0, l = 0, r = 0;
            mio_tts_b200_batching_stats(c, &a, &b, &l, &r);
            rq += a; bt += b; lg = std::max(lg, l); rs += r;
        }
        batching(64);   // fresh counters
        v[2] = run(owners, n, rounds, true, &ms[2]);
        std::printf(" {\"slots\": %d, \"overlap_only\": %.0f, \"overlap_only_ms\": %.3f, \"batched\": %.0f, \"batched_ms\": %.3f, "
                    "\"batched_wav16\": %.0f, \"batched_wav16_ms\": %.3f, \"mean_batch\": %.2f, \"largest\": %llu}%s\n",
                    n, v[0], ms[0], v[1], ms[1], v[2], ms[2], bt ? (double) rq / (double) bt : 0.0, (unsigned long long) lg,
                    ci + 1 < counts.size() ? "," : "");
        std::fflush(stdout);
    }
    std::printf("], \"unit\": \"audio-seconds/s (ms = per request)\"}\n");
    for (auto * c : owners) mio_tts_free(c);
    return 0;
}
