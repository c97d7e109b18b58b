// Fast path of the fused iSTFT for the common frame sizes (n_fft = 4*hop, n_fft/2 in {960, 1024, 640}).
//
// Same math and summation order as istft.cu (see its header for the derivation and the reference lines,
// /root/reference/src/mio-tts-lib.cpp:518-756), restructured for throughput:
//   * the radix schedule is a template pack, so every index (j % Ls, twiddle strides, butterfly counts) is a
//     compile-time constant and each pass is fully unrolled;
//   * passes run IN PLACE: a warp pulls all butterfly inputs of the pass into registers (<= 32 complex per lane),
//     __syncwarp, then writes the autosorted outputs -- one M-point buffer per frame instead of a ping-pong pair,
//     which lets 16 warps (16 frames in flight) share an SM;
//   * exp / sin / cos use the SFU (ex2.approx, sin.approx, cos.approx) after a two-constant Cody-Waite reduction of
//     the phase to [-pi, pi]: absolute error ~4e-7 per bin, far inside the 1e-3 / 40 dB gate;
//   * the steady-state envelope (sum of hann^2 over the 4 overlapping frames) is tabulated once per CTA.
#include "common.h"

namespace mio {
namespace {

constexpr int FW = 16;                 // warps (= frames in flight) per CTA
constexpr int FTHREADS = FW * 32;

struct fast_params {
    int n_freq, pad, frames_per_chunk;
    const float2 * tw_m; const float2 * tw_n; const float * hann;
    const float * spec; int pitch;
    const int * spec_row0; const int * n_frames;
    const long long * audio_off;
    float * audio;
    unsigned int * peak_bits;
    float scale;
};

__device__ __forceinline__ float2 cmul(float2 a, float2 b) { return make_float2(a.x * b.x - a.y * b.y, a.x * b.y + a.y * b.x); }
__device__ __forceinline__ float2 cadd(float2 a, float2 b) { return make_float2(a.x + b.x, a.y + b.y); }
__device__ __forceinline__ float2 csub(float2 a, float2 b) { return make_float2(a.x - b.x, a.y - b.y); }
__device__ __forceinline__ float2 cmul_i(float2 a) { return make_float2(-a.y, a.x); }

template <int R> __device__ __forceinline__ void dft_small(float2 (&v)[R]) {   // inverse kernel e^{+2 pi i p q / R}
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
        const float2 id = make_float2(-s3 * d.y, s3 * d.x);
        v[0] = cadd(v[0], s); v[1] = cadd(m, id); v[2] = csub(m, id);
    } else {
        static_assert(R == 5, "radix");
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
    }
}

// one in-place Stockham pass by one warp (reads x[j + q*m], writes x[(j-k)*R + k + p*Ls], k = j % Ls)
template <int M, int R, int LS>
__device__ __forceinline__ void pass_inplace(float2 * buf, const float2 * tw, int lane) {
    constexpr int m = M / R;
    constexpr int B = (m + 31) / 32;
    constexpr int TS = M / (LS * R);
    float2 v[B][R];
#pragma unroll
    for (int i = 0; i < B; ++i) {
        const int j = lane + 32 * i;
        if (j < m) {
#pragma unroll
            for (int q = 0; q < R; ++q) v[i][q] = buf[j + q * m];
        }
    }
    __syncwarp();
#pragma unroll
    for (int i = 0; i < B; ++i) {
        const int j = lane + 32 * i;
        if (j < m) {
            const int k = j % LS;
            if constexpr (LS > 1) {
#pragma unroll
                for (int q = 1; q < R; ++q) v[i][q] = cmul(v[i][q], tw[q * k * TS]);
            }
            dft_small<R>(v[i]);
            const int base = (j - k) * R + k;
#pragma unroll
            for (int p = 0; p < R; ++p) buf[base + p * LS] = v[i][p];
        }
    }
    __syncwarp();
}

template <int M, int LS> __device__ __forceinline__ void run_passes(float2 *, const float2 *, int) {}
template <int M, int LS, int R, int... Rest>
__device__ __forceinline__ void run_passes(float2 * buf, const float2 * tw, int lane) {
    pass_inplace<M, R, LS>(buf, tw, lane);
    run_passes<M, LS * R, Rest...>(buf, tw, lane);
}

// X_k = min(exp(logmag), 100) * (cos phi, sin phi) on the SFU
__device__ __forceinline__ float2 polar_fast(float logmag, float phase) {
    const float mag = fminf(__expf(logmag), 100.0f);
    const float kf = rintf(phase * 0.15915494309189535f);
    float r = fmaf(-kf, 6.2831854820251465f, phase);
    r = fmaf(-kf, -1.7484555e-7f, r);
    float s, c;
    __sincosf(r, &s, &c);
    return make_float2(mag * c, mag * s);
}

template <int N, int... RADIX>
__global__ void __launch_bounds__(FTHREADS, 1) istft_fast_kernel(const fast_params P) {
    MIO_PDL_PROLOGUE();
    constexpr int M = N / 2, H = N / 4, CARRY = N - H, SPAN = FW * H;
    extern __shared__ __align__(16) uint8_t smem[];
    float2 * s_twm = reinterpret_cast<float2 *>(smem);                 // [M]
    float2 * s_twn = s_twm + M;                                         // [M/2 + 1]
    float2 * s_buf = s_twn + (M / 2 + 1);                               // [FW][M]
    float * s_hann = reinterpret_cast<float *>(s_buf + (size_t) FW * M);   // [N]
    float * s_env = s_hann + N;                                         // [H]
    float * s_carry = s_env + H;                                        // [2][CARRY]
    float * s_xm = s_carry + 2 * CARRY;                                 // [FW]
    int * s_valid = reinterpret_cast<int *>(s_xm + FW);                 // [FW]

    const int utt = blockIdx.y;
    const int S = P.n_frames[utt];
    const long long n_full = (long long) (S - 1) * H + N;
    const long long chunk = (long long) P.frames_per_chunk * H;
    const long long a = (long long) blockIdx.x * chunk;
    if (S <= 0 || a >= n_full) return;
    const long long b = a + chunk < n_full ? a + chunk : n_full;
    const long long out_len = n_full - 2 * P.pad;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int i = threadIdx.x; i < M; i += FTHREADS) s_twm[i] = P.tw_m[i];
    for (int i = threadIdx.x; i <= M / 2; i += FTHREADS) s_twn[i] = P.tw_n[i];
    for (int i = threadIdx.x; i < N; i += FTHREADS) s_hann[i] = P.hann[i];
    for (int i = threadIdx.x; i < 2 * CARRY; i += FTHREADS) s_carry[i] = 0.0f;
    __syncthreads();
    for (int r = threadIdx.x; r < H; r += FTHREADS) {   // 4 overlapping frames, ascending frame = descending offset
        float e = 0.0f;
#pragma unroll
        for (int q = 3; q >= 0; --q) { const float hv = s_hann[r + q * H]; e += hv * hv; }
        s_env[r] = e;
    }

    long long lf = a - N >= 0 ? (a - N) / H + 1 : 0;
    long long ll = (b - 1) / H;
    if (ll > S - 1) ll = S - 1;

    float2 * buf = s_buf + (size_t) warp * M;
    const float * spec_u = P.spec + (size_t) P.spec_row0[utt] * P.pitch;
    float * audio_u = P.audio + P.audio_off[utt];
    float peak = 0.0f;
    int cur = 0;
    __syncthreads();

    for (long long lr0 = lf; lr0 * H < b; lr0 += FW) {
        // ---------------- phase 1: one frame per warp ----------------
        const long long l = lr0 + warp;
        const bool valid = l <= ll;
        if (valid) {
            const float * fr = spec_u + (size_t) l * P.pitch;
            constexpr int ITER = (M / 2) / 32 + 1;
#pragma unroll 4
            for (int i = 0; i < ITER; ++i) {
                const int k = lane + 32 * i;
                if (k <= M / 2) {
                    const int k2 = M - k;
                    const float2 Xk = polar_fast(fr[k], fr[P.n_freq + k]);
                    const float2 X2 = polar_fast(fr[k2], fr[P.n_freq + k2]);
                    if (k == 0) {
                        buf[0] = make_float2(Xk.x, Xk.x);
                        s_xm[warp] = X2.x;
                    } else {
                        const float2 w = s_twn[k];
                        const float2 Bk = cmul(Xk, w);
                        const float2 B2 = cmul(X2, make_float2(-w.x, w.y));
                        const float2 E = make_float2(0.5f * (Xk.x + X2.x), 0.5f * (Xk.y - X2.y));
                        const float2 O = make_float2(0.5f * (Bk.x + B2.x), 0.5f * (Bk.y - B2.y));
                        buf[k] = make_float2(E.x - O.y, E.y + O.x);
                        if (k2 != k) buf[k2] = make_float2(E.x + O.y, O.x - E.y);
                    }
                }
            }
            __syncwarp();
            run_passes<M, 1, RADIX...>(buf, s_twm, lane);
        }
        if (lane == 0) s_valid[warp] = valid ? 1 : 0;
        __syncthreads();

        // ---------------- phase 2: overlap-add the round, emit finished samples ----------------
        const long long base = lr0 * H;
        const float * carry_in = s_carry + cur * CARRY;
        float * carry_out = s_carry + (cur ^ 1) * CARRY;
        for (int x = threadIdx.x; x < SPAN + CARRY; x += FTHREADS) {
            const bool to_carry = x >= SPAN;
            float acc = (!to_carry && x < CARRY) ? carry_in[x] : 0.0f;
            const int q = x / H;
            const int w_lo = q - 3 > 0 ? q - 3 : 0;
            const int w_hi = q < FW - 1 ? q : FW - 1;
            for (int w = w_lo; w <= w_hi; ++w) {
                if (s_valid[w] == 0) continue;
                const int t = x - w * H;
                const float * zf = reinterpret_cast<const float *>(s_buf + (size_t) w * M);
                const float xm = s_xm[w];
                const float yv = (zf[t] + ((t & 1) ? -xm : xm)) * P.scale;
                acc += yv * s_hann[t];
            }
            if (to_carry) {
                carry_out[x - SPAN] = acc;
            } else {
                const long long i = base + x;
                const long long o = i - P.pad;
                if (i >= a && i < b && o >= 0 && o < out_len) {
                    const long long fi = i / H;
                    float env;
                    if (fi >= 3 && fi <= S - 1) {
                        env = s_env[(int) (i - fi * H)];
                    } else {
                        const long long e_lo = fi - 3 > 0 ? fi - 3 : 0;
                        const long long e_hi = fi < S - 1 ? fi : S - 1;
                        env = 0.0f;
                        for (long long f = e_lo; f <= e_hi; ++f) { const float hv = s_hann[(int) (i - f * H)]; env += hv * hv; }
                    }
                    float yv = acc;
                    if (env > 1e-12f) yv /= env;
                    audio_u[o] = yv;
                    const float av = fabsf(yv);
                    if (av <= 3.402823466e38f) peak = fmaxf(peak, av);
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

template <int N, int... RADIX>
void launch_fast_t(const fast_params & P, int n_utt, int max_frames, cudaStream_t st) {
    constexpr int M = N / 2, H = N / 4;
    const size_t smem = (size_t) M * 8 + (size_t) (M / 2 + 1) * 8 + (size_t) FW * M * 8 + (size_t) N * 4 + (size_t) H * 4 +
                        (size_t) 2 * (N - H) * 4 + FW * 8 + 64;
    static bool configured[64] = {false};
    int dev = 0;
    MIO_CUDA(cudaGetDevice(&dev));
    if (!configured[dev & 63]) {
        MIO_CUDA(cudaFuncSetAttribute(istft_fast_kernel<N, RADIX...>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem));
        configured[dev & 63] = true;
    }
    const long long n_full_max = (long long) (max_frames - 1) * H + N;
    const long long chunk = (long long) P.frames_per_chunk * H;
    dim3 grid((unsigned) ((n_full_max + chunk - 1) / chunk), (unsigned) n_utt);
    launch_pdl(istft_fast_kernel<N, RADIX...>, dim3(grid), dim3(FTHREADS), smem, st, P);
    MIO_CUDA(cudaGetLastError());
    count_launch();
}

} // namespace

// returns false when the plan is outside the fast path (the generic kernel in istft.cu then runs)
bool launch_istft_fast(const istft_plan & p, const float * spec, int pitch, const int * spec_row0, const int * n_frames,
                       const long long * audio_off, float * audio, unsigned int * peak_bits, int n_utt, int max_frames,
                       cudaStream_t st) {
    if (p.hop * 4 != p.n_fft) return false;
    if (p.n_fft != 1920 && p.n_fft != 2048 && p.n_fft != 1280) return false;
    fast_params P;
    P.n_freq = p.n_fft / 2 + 1;
    P.pad = (p.n_fft - p.hop) / 2;
    P.frames_per_chunk = 64 - 3;      // + 3 halo frames = 4 rounds of 16
    P.tw_m = p.d_tw_m; P.tw_n = p.d_tw_n; P.hann = p.d_hann;
    P.spec = spec; P.pitch = pitch; P.spec_row0 = spec_row0; P.n_frames = n_frames;
    P.audio_off = audio_off; P.audio = audio; P.peak_bits = peak_bits;
    P.scale = 1.0f / (float) P.n_freq;
    switch (p.n_fft) {
        case 1920: launch_fast_t<1920, 4, 4, 4, 3, 5>(P, n_utt, max_frames, st); break;
        case 2048: launch_fast_t<2048, 4, 4, 4, 4, 4>(P, n_utt, max_frames, st); break;
        default: launch_fast_t<1280, 4, 4, 4, 2, 5>(P, n_utt, max_frames, st); break;
    }
    return true;
}

} // namespace mio
