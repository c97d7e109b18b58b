// mio_tts_b200_pcm16_static: the peak rule + WAV16 sample conversion (pcm16_kernel of istft.cu) for callers that must
// not allocate: code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc.  It has no
// counterpart in the reference.
//
// The per-utterance peak words (atomicMax targets, must start at zero) and the offset / length tables that the pooled
// path keeps in engine buffers live in ONE file-scope __device__ array of MIO_TTS_B200_STATIC_SCRATCH_BYTES.  Every GPU
// that loads the library holds that array until the process exits.  All calls share it and it keeps the last call's
// contents: the peak words are cleared at the start of each call, the tables are overwritten.  Calls that overlap on
// one GPU race on it; the host side therefore serialises callers of this process with a mutex and returns only after
// the stream has drained.  Nothing here calls cudaMalloc / cudaFree / an allocating library routine, and no host heap
// memory is taken either (fixed-size locals only).
#include <cmath>
#include <cstdio>
#include <cstring>
#include <mutex>

#include "../../include/mio_tts_b200.h"
#include "common.h"

namespace {

constexpr int kMaxUtt = MIO_TTS_B200_STATIC_MAX_UTT;
// layout of the scratch: [kMaxUtt] long long offsets | [kMaxUtt] long long lengths | [kMaxUtt] unsigned peak words
constexpr size_t kOffBytes = (size_t) kMaxUtt * 8, kLenBytes = (size_t) kMaxUtt * 8, kPeakBytes = (size_t) kMaxUtt * 4;
static_assert(kOffBytes + kLenBytes + kPeakBytes == MIO_TTS_B200_STATIC_SCRATCH_BYTES, "scratch layout and header disagree");

// In HOST code this name is only a stand-in for the symbol, never a device address: the address each device gave its
// own copy comes from cudaGetSymbolAddress (scratch_for below).  Device code may name it directly.
static __device__ __align__(256) unsigned char g_static_scratch[MIO_TTS_B200_STATIC_SCRATCH_BYTES];

constexpr int kMaxDevices = 64;
unsigned char * g_scratch_ptr[kMaxDevices] = {nullptr};   // per device, filled on first use with that device current
std::mutex g_static_mtx;

// max |x| over the finite samples of each utterance (what engine::pcm16_host takes on the host, and what the iSTFT
// kernel leaves in peak_bits on the pooled path) -> atomicMax on the utterance's peak word
__global__ void static_peak_kernel(const float * __restrict__ audio, const long long * __restrict__ off,
                                   const long long * __restrict__ len, unsigned int * __restrict__ peak_bits) {
    const int utt = blockIdx.y;
    const float * p = audio + off[utt];
    const long long n = len[utt];
    float peak = 0.0f;
    for (long long i = (long long) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long) gridDim.x * blockDim.x) {
        const float a = fabsf(p[i]);
        if (a <= 3.402823466e+38f) peak = fmaxf(peak, a);   // false for inf and NaN
    }
    for (int o = 16; o > 0; o >>= 1) peak = fmaxf(peak, __shfl_xor_sync(0xffffffffu, peak, o));
    if ((threadIdx.x & 31) == 0 && peak > 0.0f) atomicMax(peak_bits + utt, __float_as_uint(peak));
}

// the same arithmetic as pcm16_kernel (istft.cu; /root/reference/src/mio-tts-lib.cpp:1299-1310 and :781-785):
// gain = (float)(0.95 / peak) when peak > 0.98, y = clamp(x * gain, -1, 1), pcm = (int16) lrintf(y * 32767)
__global__ void static_pcm16_kernel(const float * __restrict__ audio, const long long * __restrict__ off,
                                    const long long * __restrict__ len, const unsigned int * __restrict__ peak_bits,
                                    short * __restrict__ pcm) {
    const int utt = blockIdx.y;
    const double peak = (double) __uint_as_float(peak_bits[utt]);
    const bool scale = peak > 0.98;
    const float gain = scale ? (float) (0.95 / peak) : 1.0f;
    const float * p = audio + off[utt];
    short * q = pcm + off[utt];
    const long long n = len[utt];
    for (long long i = (long long) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long) gridDim.x * blockDim.x) {
        float x = p[i];
        if (scale) x = __fmul_rn(x, gain);
        x = fminf(fmaxf(x, -1.0f), 1.0f);
        q[i] = (short) __float2int_rn(__fmul_rn(x, 32767.0f));
    }
}

bool fail(char * err, size_t err_size, const char * what, cudaError_t e = cudaSuccess) {
    if (err != nullptr && err_size > 0) {
        if (e != cudaSuccess) std::snprintf(err, err_size, "%s: %s", what, cudaGetErrorString(e));
        else std::snprintf(err, err_size, "%s", what);
    }
    return false;
}

// set-up, once per device, with that device current: the address of THIS device's copy of the array
bool scratch_for(int device, unsigned char ** out, char * err, size_t err_size) {
    if (g_scratch_ptr[device] == nullptr) {
        void * p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&p, g_static_scratch);
        if (e != cudaSuccess || p == nullptr) return fail(err, err_size, "cudaGetSymbolAddress(static scratch) failed", e);
        g_scratch_ptr[device] = (unsigned char *) p;
    }
    *out = g_scratch_ptr[device];
    return true;
}

} // namespace

extern "C" MIO_B200_API bool mio_tts_b200_pcm16_static(int32_t device, const float * d_audio, const int64_t * offsets,
                                                       const int64_t * lengths, size_t batch, int16_t * d_pcm,
                                                       void * stream, char * err, size_t err_size) {
    if (d_audio == nullptr || d_pcm == nullptr || offsets == nullptr || lengths == nullptr || batch == 0)
        return fail(err, err_size, "audio is empty");
    if (batch > (size_t) kMaxUtt) {
        if (err != nullptr && err_size > 0)
            std::snprintf(err, err_size, "batch of %zu utterances exceeds the static scratch (at most %d per call)", batch, kMaxUtt);
        return false;
    }
    if (device < 0 || device >= kMaxDevices) return fail(err, err_size, "device index out of range");
    long long max_len = 0;
    for (size_t u = 0; u < batch; ++u) {
        if (offsets[u] < 0 || lengths[u] <= 0) return fail(err, err_size, "audio is empty");
        if (lengths[u] > max_len) max_len = lengths[u];
    }
    std::lock_guard<std::mutex> lk(g_static_mtx);
    int prev = -1;
    cudaError_t e = cudaGetDevice(&prev);
    if (e != cudaSuccess) return fail(err, err_size, "cudaGetDevice failed", e);
    e = cudaSetDevice(device);
    if (e != cudaSuccess) return fail(err, err_size, "cudaSetDevice failed", e);
    bool ok = false;
    do {
        unsigned char * s = nullptr;
        if (!scratch_for(device, &s, err, err_size)) break;
        cudaStream_t st = (cudaStream_t) stream;
        long long * d_off = (long long *) s;
        long long * d_len = (long long *) (s + kOffBytes);
        unsigned int * d_peak = (unsigned int *) (s + kOffBytes + kLenBytes);
        static_assert(sizeof(long long) == sizeof(int64_t), "offset tables are 64-bit");
        // the array keeps the last call's contents: the peak words must start at zero, the tables are overwritten
        if ((e = cudaMemsetAsync(d_peak, 0, batch * 4, st)) != cudaSuccess) { fail(err, err_size, "clearing the peak words failed", e); break; }
        if ((e = cudaMemcpyAsync(d_off, offsets, batch * 8, cudaMemcpyHostToDevice, st)) != cudaSuccess ||
            (e = cudaMemcpyAsync(d_len, lengths, batch * 8, cudaMemcpyHostToDevice, st)) != cudaSuccess) {
            fail(err, err_size, "uploading the offset tables failed", e); break;
        }
        const int threads = 256;
        const long long want = (max_len + threads * 8 - 1) / (threads * 8);
        const dim3 grid((unsigned) (want < 1 ? 1 : (want > 592 ? 592 : want)), (unsigned) batch);
        static_peak_kernel<<<grid, threads, 0, st>>>(d_audio, d_off, d_len, d_peak);
        static_pcm16_kernel<<<grid, threads, 0, st>>>(d_audio, d_off, d_len, d_peak, (short *) d_pcm);
        if ((e = cudaGetLastError()) != cudaSuccess) { fail(err, err_size, "kernel launch failed", e); break; }
        mio::count_launch(2);
        if ((e = cudaStreamSynchronize(st)) != cudaSuccess) { fail(err, err_size, "static WAV16 conversion failed", e); break; }
        ok = true;
    } while (false);
    if (prev >= 0 && prev != device) cudaSetDevice(prev);
    return ok;
}
