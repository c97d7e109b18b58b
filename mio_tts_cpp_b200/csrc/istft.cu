// Fused batched iSTFT: exp/cos/sin -> inverse real FFT -> synthesis window -> overlap-add -> envelope
// normalisation -> crop, one kernel, shared-memory staged, deterministic (no atomics on samples).
//
// Reference: embd_to_audio (/root/reference/src/mio-tts-lib.cpp:625-756) with irfft_with_plan (:518-623):
//   mag = min(exp(logmag), 100);  X_k = mag (cos phi + i sin phi), k = 0..N/2
//   y[t] = Re( sum_{k=0}^{N/2} X_k e^{+2 pi i k t / N} ) / (N/2 + 1)        (NOT a Hermitian irfft)
//   audio = OLA(y * hann) / OLA(hann^2) where > 1e-12, cropped by (N - hop)/2 per side.
//
// Algorithm.  With M = N/2, w = e^{2 pi i / N}:  y[2m] = Re IDFT_M(X_k)[m] + Re X_M,
// y[2m+1] = Re IDFT_M(X_k w^k)[m] - Re X_M.  Both real parts come from ONE M-point complex inverse FFT of
// Z_k = E_k + i O_k, E_k = (X_k + conj X_{M-k})/2, O_k = (X_k w^k + conj(X_{M-k} w^{M-k}))/2, so that
// z[m] = y[2m] - Re X_M + i (y[2m+1] + Re X_M).  The M-point FFT is a mixed-radix Stockham autosort
// (radices 4/2/3/5 hard-coded, other primes <= 16 via the root table) executed by ONE WARP per frame in its
// own shared-memory ping-pong buffers; eight warps = eight frames in flight per CTA.  After each round of
// eight frames the CTA overlap-adds them in ascending frame order (the reference's 1-thread summation order)
// on top of a carry of the previous rounds, divides by the envelope and streams finished samples out.
#include "common.h"

#include <cmath>
#include <cstdlib>
#include <vector>

namespace mio {
namespace {

constexpr int ISTFT_WARPS = 8;
constexpr int ISTFT_THREADS = ISTFT_WARPS * 32;

struct istft_params {
    int n_fft, hop, M, n_freq, pad, overlap;       // overlap = ceil(n_fft / hop)
    int frames_per_chunk;                          // frames a CTA is responsible for
    int n_radix;
    int radix[16];
    const float2 * tw_m;
    const float2 * tw_n;
    const float * hann;
    const float * spec; int pitch;
    const int * spec_row0; const int * n_frames;
    const long long * audio_off;
    float * audio;
    unsigned int * peak_bits;
    float scale;                                   // 1 / n_freq
};

__device__ __forceinline__ float2 cmul(float2 a, float2 b) { return make_float2(a.x * b.x - a.y * b.y, a.x * b.y + a.y * b.x); }
__device__ __forceinline__ float2 cadd(float2 a, float2 b) { return make_float2(a.x + b.x, a.y + b.y); }
__device__ __forceinline__ float2 csub(float2 a, float2 b) { return make_float2(a.x - b.x, a.y - b.y); }
__device__ __forceinline__ float2 cmul_i(float2 a) { return make_float2(-a.y, a.x); }   // i * a

// in-register inverse DFT of size R (kernel e^{+2 pi i p q / R})
template <int R> __device__ __forceinline__ void dft_r(float2 (&v)[R], const float2 * tw_m, int M) {
    if constexpr (R == 2) {
        const float2 a = v[0], b = v[1];
        v[0] = cadd(a, b); v[1] = csub(a, b);
    } else if constexpr (R == 4) {
        const float2 t0 = cadd(v[0], v[2]), t1 = csub(v[0], v[2]);
        const float2 t2 = cadd(v[1], v[3]), t3 = cmul_i(csub(v[1], v[3]));
        v[0] = cadd(t0, t2); v[1] = cadd(t1, t3); v[2] = csub(t0, t2); v[3] = csub(t1, t3);
    } else if constexpr (R == 3) {
        const float s3 = 0.86602540378443864676f;
        const float2 s = cadd(v[1], v[2]), d = csub(v[1], v[2]);
        const float2 m = make_float2(v[0].x - 0.5f * s.x, v[0].y - 0.5f * s.y);
        const float2 id = make_float2(-s3 * d.y, s3 * d.x);     // i * s3 * d
        v[0] = cadd(v[0], s); v[1] = cadd(m, id); v[2] = csub(m, id);
    } else if constexpr (R == 5) {
        const float c1 = 0.30901699437494742410f, c2 = -0.80901699437494742410f;
        const float s1 = 0.95105651629515357212f, s2 = 0.58778525229247312917f;
        const float2 a1 = cadd(v[1], v[4]), a2 = cadd(v[2], v[3]);
        const float2 b1 = csub(v[1], v[4]), b2 = csub(v[2], v[3]);
        const float2 m1 = make_float2(v[0].x + c1 * a1.x + c2 * a2.x, v[0].y + c1 * a1.y + c2 * a2.y);
        const float2 m2 = make_float2(v[0].x + c2 * a1.x + c1 * a2.x, v[0].y + c2 * a1.y + c1 * a2.y);
        const float2 n1 = make_float2(s1 * b1.x + s2 * b2.x, s1 * b1.y + s2 * b2.y);
        const float2 n2 = make_float2(s2 * b1.x - s1 * b2.x, s2 * b1.y - s1 * b2.y);
        v[0] = cadd(v[0], cadd(a1, a2));
        v[1] = cadd(m1, cmul_i(n1)); v[4] = csub(m1, cmul_i(n1));
        v[2] = cadd(m2, cmul_i(n2)); v[3] = csub(m2, cmul_i(n2));
    } else {
        float2 o[R];
        const int step = M / R;
#pragma unroll
        for (int p = 0; p < R; ++p) {
            float2 acc = v[0];
#pragma unroll
            for (int q = 1; q < R; ++q) acc = cadd(acc, cmul(v[q], tw_m[((p * q) % R) * step]));
            o[p] = acc;
        }
#pragma unroll
        for (int p = 0; p < R; ++p) v[p] = o[p];
    }
}

// one Stockham pass executed by a warp: x -> y, sub-transform length so far Ls
template <int R>
__device__ __forceinline__ void stockham_pass(const float2 * x, float2 * y, int M, int Ls, const float2 * tw_m, int lane) {
    const int m = M / R;
    const int tw_step = M / (Ls * R);
    for (int j = lane; j < m; j += 32) {
        const int k = j % Ls;
        float2 v[R];
#pragma unroll
        for (int q = 0; q < R; ++q) v[q] = x[j + q * m];
        if (k != 0) {
#pragma unroll
            for (int q = 1; q < R; ++q) v[q] = cmul(v[q], tw_m[q * k * tw_step]);
        }
        dft_r<R>(v, tw_m, M);
        const int base = (j - k) * R + k;
#pragma unroll
        for (int p = 0; p < R; ++p) y[base + p * Ls] = v[p];
    }
}

__global__ void __launch_bounds__(ISTFT_THREADS, 1) istft_kernel(const istft_params P) {
    MIO_PDL_PROLOGUE();
    extern __shared__ __align__(16) uint8_t smem[];
    const int N = P.n_fft, H = P.hop, M = P.M;
    const int carry_len = N - H;
    float2 * s_twm = reinterpret_cast<float2 *>(smem);                 // [M]
    float2 * s_twn = s_twm + M;                                         // [M/2 + 1]
    float2 * s_buf = s_twn + (M / 2 + 1);                               // [WARPS][2][M]
    float * s_hann = reinterpret_cast<float *>(s_buf + (size_t) ISTFT_WARPS * 2 * M);   // [N]
    float * s_carry = s_hann + N;                                       // [2][carry_len]
    float * s_xm = s_carry + 2 * carry_len;                             // [WARPS]
    int * s_valid = reinterpret_cast<int *>(s_xm + ISTFT_WARPS);        // [WARPS]

    const int utt = blockIdx.y;
    const int S = P.n_frames[utt];
    const long long n_full = (long long) (S - 1) * H + N;
    const long long chunk = (long long) P.frames_per_chunk * H;
    const long long a = (long long) blockIdx.x * chunk;
    if (S <= 0 || a >= n_full) return;
    const long long b = a + chunk < n_full ? a + chunk : n_full;
    const long long out_len = n_full - 2 * P.pad;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int i = threadIdx.x; i < M; i += blockDim.x) s_twm[i] = P.tw_m[i];
    for (int i = threadIdx.x; i <= M / 2; i += blockDim.x) s_twn[i] = P.tw_n[i];
    for (int i = threadIdx.x; i < N; i += blockDim.x) s_hann[i] = P.hann[i];
    for (int i = threadIdx.x; i < 2 * carry_len; i += blockDim.x) s_carry[i] = 0.0f;

    // first / last frame that touches [a, b)
    long long lf = a - N >= 0 ? (a - N) / H + 1 : 0;
    if (lf < 0) lf = 0;
    long long ll = (b - 1) / H;
    if (ll > S - 1) ll = S - 1;

    float2 * bufA = s_buf + (size_t) warp * 2 * M;
    float2 * bufB = bufA + M;
    const float * spec_u = P.spec + (size_t) P.spec_row0[utt] * P.pitch;
    float * audio_u = P.audio + P.audio_off[utt];
    float peak = 0.0f;
    int cur = 0;
    __syncthreads();

    for (long long lr0 = lf; lr0 * H < b; lr0 += ISTFT_WARPS) {   // trailing rounds only flush the carry
        // ---------------- phase 1: one frame per warp ----------------
        const long long l = lr0 + warp;
        const bool valid = l <= ll;
        const float2 * res = bufA;
        if (valid) {
            const float * fr = spec_u + (size_t) l * P.pitch;
            for (int k = lane; k <= M / 2; k += 32) {
                const int k2 = M - k;        // partner bin (k = 0 pairs with the Nyquist bin M)
                float sk, ck, s2, c2;
                const float mk = fminf(expf(fr[k]), 100.0f);
                sincosf(fr[P.n_freq + k], &sk, &ck);
                const float m2 = fminf(expf(fr[k2]), 100.0f);
                sincosf(fr[P.n_freq + k2], &s2, &c2);
                const float2 Xk = make_float2(mk * ck, mk * sk);
                const float2 X2 = make_float2(m2 * c2, m2 * s2);
                if (k == 0) {
                    bufA[0] = make_float2(Xk.x, Xk.x);            // Z_0 = Re X_0 (1 + i)
                    s_xm[warp] = X2.x;                             // Re X_M
                } else {
                    const float2 w = s_twn[k];
                    const float2 Bk = cmul(Xk, w);
                    const float2 B2 = cmul(X2, make_float2(-w.x, w.y));          // w^(M-k) = -conj(w^k)
                    const float2 E = make_float2(0.5f * (Xk.x + X2.x), 0.5f * (Xk.y - X2.y));
                    const float2 O = make_float2(0.5f * (Bk.x + B2.x), 0.5f * (Bk.y - B2.y));
                    bufA[k] = make_float2(E.x - O.y, E.y + O.x);                 // E + iO
                    if (k2 != k) bufA[k2] = make_float2(E.x + O.y, O.x - E.y);   // conj(E) + i conj(O)
                }
            }
            __syncwarp();
            float2 * x = bufA;
            float2 * y = bufB;
            int Ls = 1;
            for (int ps = 0; ps < P.n_radix; ++ps) {
                const int r = P.radix[ps];
                switch (r) {
                    case 2: stockham_pass<2>(x, y, M, Ls, s_twm, lane); break;
                    case 3: stockham_pass<3>(x, y, M, Ls, s_twm, lane); break;
                    case 4: stockham_pass<4>(x, y, M, Ls, s_twm, lane); break;
                    case 5: stockham_pass<5>(x, y, M, Ls, s_twm, lane); break;
                    case 7: stockham_pass<7>(x, y, M, Ls, s_twm, lane); break;
                    case 11: stockham_pass<11>(x, y, M, Ls, s_twm, lane); break;
                    case 13: stockham_pass<13>(x, y, M, Ls, s_twm, lane); break;
                    default: break;
                }
                Ls *= r;
                float2 * tmp = x; x = y; y = tmp;
                __syncwarp();
            }
            res = x;
        }
        if (lane == 0) s_valid[warp] = valid ? (int) (res == bufA ? 1 : 2) : 0;
        __syncthreads();

        // ---------------- phase 2: overlap-add the round, emit finished samples ----------------
        const long long base = lr0 * H;
        const float * carry_in = s_carry + cur * carry_len;
        float * carry_out = s_carry + (cur ^ 1) * carry_len;
        const int span = ISTFT_WARPS * H;
        for (int x = threadIdx.x; x < span + carry_len; x += blockDim.x) {
            float acc = (x < carry_len) ? carry_in[x] : 0.0f;
            const bool to_carry = x >= span;
            if (to_carry) acc = 0.0f;
            int w_lo = (x - N + H) / H;            // smallest w with x - w*H < N
            if (x - N + 1 <= 0) w_lo = 0;
            int w_hi = x / H;
            if (w_hi > ISTFT_WARPS - 1) w_hi = ISTFT_WARPS - 1;
            for (int w = w_lo; w <= w_hi; ++w) {
                const int v = s_valid[w];
                if (v == 0) continue;
                const int t = x - w * H;
                const float * zf = reinterpret_cast<const float *>(s_buf + (size_t) w * 2 * M + (v == 2 ? M : 0));
                const float xm = s_xm[w];
                const float yv = (zf[t] + ((t & 1) ? -xm : xm)) * P.scale;
                acc += yv * s_hann[t];
            }
            if (to_carry) {
                carry_out[x - span] = acc;
            } else {
                const long long i = base + x;
                const long long o = i - P.pad;
                if (i >= a && i < b && o >= 0 && o < out_len) {
                    // envelope: sum of hann^2 over the frames covering sample i, ascending frame order
                    long long e_lo = i - N + 1 > 0 ? (i - N + H) / H : 0;
                    long long e_hi = i / H;
                    if (e_hi > S - 1) e_hi = S - 1;
                    float env = 0.0f;
                    for (long long f = e_lo; f <= e_hi; ++f) {
                        const float hv = s_hann[(int) (i - f * H)];
                        env += hv * hv;
                    }
                    float yv = acc;
                    if (env > 1e-12f) yv /= env;
                    audio_u[o] = yv;
                    const float av = fabsf(yv);
                    if (av <= 3.402823466e38f) peak = fmaxf(peak, av);    // finite samples only
                }
            }
        }
        cur ^= 1;
        __syncthreads();
    }

    if (P.peak_bits) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) peak = fmaxf(peak, __shfl_xor_sync(0xffffffffu, peak, o));
        if (lane == 0 && peak > 0.0f) atomicMax(P.peak_bits + utt, __float_as_uint(peak));
    }
}

// mio_tts_synthesize tail (src/mio-tts-lib.cpp:1299-1310): if peak > 0.98, scale by (float)(0.95 / peak)
__global__ void peak_normalize_kernel(float * __restrict__ audio, const long long * __restrict__ audio_off,
                                      const long long * __restrict__ audio_len,
                                      const unsigned int * __restrict__ peak_bits) {
    MIO_PDL_PROLOGUE();
    const int utt = blockIdx.y;
    const double peak = (double) __uint_as_float(peak_bits[utt]);
    if (!(peak > 0.98)) return;
    const float gain = (float) (0.95 / peak);
    float * p = audio + audio_off[utt];
    const long long n = audio_len[utt];
    for (long long i = (long long) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long) gridDim.x * blockDim.x)
        p[i] *= gain;
}

// Peak normaliser + WAV16 sample emission in one pass (server response path: /root/reference/src/tts-mio-server.cpp:3965-3973,
// same rule as save_wav16, src/mio-tts-lib.cpp:781-785): y = clamp(x * gain, -1, 1); pcm = (int16) lrintf(y * 32767).
// The fp32 waveform is left untouched (un-normalised); each product is rounded separately like the reference's two loops.
__global__ void pcm16_kernel(const float * __restrict__ audio, const long long * __restrict__ audio_off,
                             const long long * __restrict__ audio_len, const unsigned int * __restrict__ peak_bits,
                             short * __restrict__ pcm) {
    MIO_PDL_PROLOGUE();
    const int utt = blockIdx.y;
    const double peak = (double) __uint_as_float(peak_bits[utt]);
    const bool scale = peak > 0.98;
    const float gain = scale ? (float) (0.95 / peak) : 1.0f;
    const float * p = audio + audio_off[utt];           // offsets are multiples of 64 samples
    short * q = pcm + audio_off[utt];
    const long long n = audio_len[utt], n4 = n >> 2;
    auto cvt = [&](float x) {
        if (scale) x = __fmul_rn(x, gain);
        x = fminf(fmaxf(x, -1.0f), 1.0f);
        return (short) __float2int_rn(__fmul_rn(x, 32767.0f));
    };
    const long long stride = (long long) gridDim.x * blockDim.x, t = (long long) blockIdx.x * blockDim.x + threadIdx.x;
    for (long long i = t; i < n4; i += stride) {
        const float4 v = reinterpret_cast<const float4 *>(p)[i];
        short4 o;
        o.x = cvt(v.x); o.y = cvt(v.y); o.z = cvt(v.z); o.w = cvt(v.w);
        reinterpret_cast<short4 *>(q)[i] = o;
    }
    for (long long i = (n4 << 2) + t; i < n; i += stride) q[i] = cvt(p[i]);
}

// synthetic spectra for the iSTFT microbench (BASELINE config 3): log-magnitude ~ N(-2, 1), phase ~ U(-pi, pi)
__device__ __forceinline__ unsigned long long splitmix64(unsigned long long x) {
    x += 0x9E3779B97F4A7C15ull;
    x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
    x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
    return x ^ (x >> 31);
}

__global__ void fill_spec_kernel(float * __restrict__ spec, long long rows, int pitch, int n_freq, unsigned long long seed) {
    MIO_PDL_PROLOGUE();
    const long long i = (long long) blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= rows * n_freq) return;
    const long long r = i / n_freq;
    const int k = (int) (i % n_freq);
    const unsigned long long h = splitmix64(seed ^ (unsigned long long) i);
    const unsigned long long h2 = splitmix64(h);
    const float u1 = ((float) (h >> 40) + 1.0f) * (1.0f / 16777217.0f);
    const float u2 = (float) (h2 >> 40) * (1.0f / 16777216.0f);
    const float u3 = (float) ((h2 >> 16) & 0xFFFFFF) * (1.0f / 16777216.0f);
    const float nrm = sqrtf(-2.0f * logf(u1)) * cospif(2.0f * u2);
    spec[r * pitch + k] = -2.0f + nrm;
    spec[r * pitch + n_freq + k] = (2.0f * u3 - 1.0f) * 3.14159265358979f;
}

__global__ void sum_kernel(const float * __restrict__ x, long long n, double * __restrict__ out) {
    MIO_PDL_PROLOGUE();
    double acc = 0.0;
    for (long long i = (long long) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long) gridDim.x * blockDim.x)
        acc += (double) x[i];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    if ((threadIdx.x & 31) == 0) atomicAdd(out, acc);
}

} // namespace

void launch_fill_spec(float * spec, long long rows, int pitch, int n_freq, unsigned long long seed, cudaStream_t st) {
    const long long n = rows * n_freq;
    if (n <= 0) return;
    launch_pdl(fill_spec_kernel, dim3((unsigned) ((n + 255) / 256)), dim3(256), 0, st, spec, rows, pitch, n_freq, seed);
    MIO_CUDA(cudaGetLastError());
    count_launch();
}

double device_sum(const float * x, long long n, cudaStream_t st) {
    double * d = nullptr;
    MIO_CUDA(cudaMalloc(&d, sizeof(double)));
    MIO_CUDA(cudaMemsetAsync(d, 0, sizeof(double), st));
    launch_pdl(sum_kernel, dim3(1024), dim3(256), 0, st, x, n, d);
    MIO_CUDA(cudaGetLastError());
    count_launch();
    double h = 0.0;
    MIO_CUDA(cudaMemcpyAsync(&h, d, sizeof(double), cudaMemcpyDeviceToHost, st));
    MIO_CUDA(cudaStreamSynchronize(st));
    cudaFree(d);
    return h;
}

bool istft_plan_init(istft_plan & p, int n_fft, int hop, std::string & err) {
    istft_plan_free(p);
    if (n_fft <= 0 || hop <= 0 || (n_fft & 1) || hop > n_fft) { err = "unsupported n_fft/hop for iSTFT"; return false; }
    if (n_fft > 9 * hop) { err = "unsupported n_fft/hop ratio for iSTFT (n_fft > 9*hop)"; return false; }
    const int M = n_fft / 2;
    int rem = M, nr = 0;
    const int cand[] = {4, 2, 3, 5, 7, 11, 13};
    for (int r : cand) {
        while (rem % r == 0 && rem > 1) {
            if (nr >= 16) { err = "unsupported n_fft for iSTFT"; return false; }
            p.radix[nr++] = r;
            rem /= r;
        }
    }
    if (rem != 1) { err = "unsupported n_fft for iSTFT (prime factor > 13)"; return false; }
    if (nr == 0) { p.radix[nr++] = 1; }
    p.n_radix = nr;
    p.n_fft = n_fft;
    p.hop = hop;
    p.smem_bytes = (size_t) M * 8 + (size_t) (M / 2 + 1) * 8 + (size_t) ISTFT_WARPS * 2 * M * 8 + (size_t) n_fft * 4 +
                   (size_t) 2 * (n_fft - hop) * 4 + ISTFT_WARPS * 8 + 64;
    if (p.smem_bytes > 227 * 1024) { err = "n_fft too large for the fused iSTFT kernel"; return false; }

    std::vector<float2> twm((size_t) M), twn((size_t) M / 2 + 1);
    for (int j = 0; j < M; ++j) {
        const double ang = 2.0 * M_PI * (double) j / (double) M;
        twm[(size_t) j] = make_float2((float) std::cos(ang), (float) std::sin(ang));
    }
    for (int k = 0; k <= M / 2; ++k) {
        const double ang = 2.0 * M_PI * (double) k / (double) n_fft;
        twn[(size_t) k] = make_float2((float) std::cos(ang), (float) std::sin(ang));
    }
    std::vector<float> hann((size_t) n_fft);
    for (int i = 0; i < n_fft; ++i)   // fill_hann_window(periodic), src/mio-tts-lib.cpp:415-420
        hann[(size_t) i] = 0.5f * (1.0f - cosf((2.0f * (float) M_PI * (float) i) / (float) n_fft));
    MIO_CUDA(cudaMalloc(&p.d_tw_m, twm.size() * sizeof(float2)));
    MIO_CUDA(cudaMalloc(&p.d_tw_n, twn.size() * sizeof(float2)));
    MIO_CUDA(cudaMalloc(&p.d_hann, hann.size() * sizeof(float)));
    MIO_CUDA(cudaMemcpy(p.d_tw_m, twm.data(), twm.size() * sizeof(float2), cudaMemcpyHostToDevice));
    MIO_CUDA(cudaMemcpy(p.d_tw_n, twn.data(), twn.size() * sizeof(float2), cudaMemcpyHostToDevice));
    MIO_CUDA(cudaMemcpy(p.d_hann, hann.data(), hann.size() * sizeof(float), cudaMemcpyHostToDevice));
    return true;
}

void istft_plan_free(istft_plan & p) {
    if (p.d_tw_m) cudaFree(p.d_tw_m);
    if (p.d_tw_n) cudaFree(p.d_tw_n);
    if (p.d_hann) cudaFree(p.d_hann);
    p = istft_plan();
}

bool launch_istft_fast(const istft_plan & p, const float * spec, int pitch, const int * spec_row0, const int * n_frames,
                       const long long * audio_off, float * audio, unsigned int * peak_bits, int n_utt, int max_frames,
                       cudaStream_t st);

void launch_istft(const istft_plan & p, const float * spec, int pitch, const int * spec_row0, const int * n_frames,
                  const long long * audio_off, float * audio, unsigned int * peak_bits, int n_utt, int max_frames,
                  cudaStream_t st) {
    if (n_utt <= 0 || max_frames <= 0) return;
    static const bool no_fast = std::getenv("MIO_TTS_B200_ISTFT_GENERIC") != nullptr;
    if (!no_fast && launch_istft_fast(p, spec, pitch, spec_row0, n_frames, audio_off, audio, peak_bits, n_utt, max_frames, st))
        return;
    istft_params P;
    P.n_fft = p.n_fft; P.hop = p.hop; P.M = p.n_fft / 2; P.n_freq = p.n_fft / 2 + 1;
    P.pad = (p.n_fft - p.hop) / 2;
    P.overlap = (p.n_fft + p.hop - 1) / p.hop;
    // frames needed by a chunk = frames_per_chunk + overlap - 1: make that a multiple of the 8 warps
    P.frames_per_chunk = 64 - (P.overlap - 1);
    P.n_radix = p.n_radix;
    for (int i = 0; i < 16; ++i) P.radix[i] = p.radix[i];
    if (P.n_radix == 1 && P.radix[0] == 1) P.n_radix = 0;
    P.tw_m = p.d_tw_m; P.tw_n = p.d_tw_n; P.hann = p.d_hann;
    P.spec = spec; P.pitch = pitch; P.spec_row0 = spec_row0; P.n_frames = n_frames;
    P.audio_off = audio_off; P.audio = audio; P.peak_bits = peak_bits;
    P.scale = 1.0f / (float) P.n_freq;
    const long long n_full_max = (long long) (max_frames - 1) * p.hop + p.n_fft;
    const long long chunk = (long long) P.frames_per_chunk * p.hop;
    dim3 grid((unsigned) ((n_full_max + chunk - 1) / chunk), (unsigned) n_utt);
    {   // opt-in dynamic shared memory is a per-device function attribute: raise it to the high-water mark
        static size_t smem_set[64] = {0};
        int dev = 0;
        MIO_CUDA(cudaGetDevice(&dev));
        if (p.smem_bytes > smem_set[dev & 63]) {
            MIO_CUDA(cudaFuncSetAttribute(istft_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) p.smem_bytes));
            smem_set[dev & 63] = p.smem_bytes;
        }
    }
    launch_pdl(istft_kernel, dim3(grid), dim3(ISTFT_THREADS), p.smem_bytes, st, P);
    MIO_CUDA(cudaGetLastError());
    count_launch();
}

void launch_peak_normalize(float * audio, const long long * audio_off, const long long * audio_len,
                           const unsigned int * peak_bits, int n_utt, long long max_len, cudaStream_t st) {
    if (n_utt <= 0 || max_len <= 0) return;
    long long bx = (max_len + 256 * 8 - 1) / (256 * 8);
    if (bx > 1024) bx = 1024;
    dim3 grid((unsigned) bx, (unsigned) n_utt);
    launch_pdl(peak_normalize_kernel, dim3(grid), dim3(256), 0, st, audio, audio_off, audio_len, peak_bits);
    MIO_CUDA(cudaGetLastError());
    count_launch();
}

void launch_pcm16(const float * audio, const long long * audio_off, const long long * audio_len, const unsigned int * peak_bits,
                  int16_t * pcm, int n_utt, long long max_len, cudaStream_t st) {
    long long bx = (max_len / 4 + 255) / 256;
    if (bx > 1024) bx = 1024;
    if (bx < 1) bx = 1;
    dim3 grid((unsigned) bx, (unsigned) n_utt);
    launch_pdl(pcm16_kernel, dim3(grid), dim3(256), 0, st, audio, audio_off, audio_len, peak_bits, (short *) pcm);
    MIO_CUDA(cudaGetLastError());
    count_launch();
}

} // namespace mio
