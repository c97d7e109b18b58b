// Fast path of the fused iSTFT for the common frame sizes (n_fft = 4*hop; n_fft in {1920, 2048, 1280, 392, 1024, 512, 800, 640,
// 400, 960, 1200, 1600}: three Stockham passes of radix <= 16 each).
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

#include <cstdlib>
#include <type_traits>
#include <utility>

namespace mio {
namespace {

// FW = warps (= frames in flight) per CTA, a template parameter: 16 with one CTA per SM, 8 with two, 4 with four.  A CTA's
// warps move through the load / FFT phase and the overlap-add phase together, so CTAs that share an SM are what overlaps
// one CTA's global loads with another's shared-memory work.

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

// ---- complex arithmetic on the packed fp32x2 pipe (FADD2 / FMUL2 / FFMA2, sm_100): one instruction per complex add and
// two per complex multiply; ptxas folds the pack / swap / broadcast / negate below into operand modifiers
// (.F32x2.LO_HI.NP, .F32), so the helpers cost nothing beyond the arithmetic itself.
typedef unsigned long long u64;
__device__ __forceinline__ u64 pk2(float lo, float hi) { u64 r; asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi)); return r; }
__device__ __forceinline__ float2 up2(u64 v) { float2 r; asm("mov.b64 {%0, %1}, %2;" : "=f"(r.x), "=f"(r.y) : "l"(v)); return r; }
__device__ __forceinline__ float2 cadd(float2 a, float2 b) {
    u64 d; asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(pk2(a.x, a.y)), "l"(pk2(b.x, b.y))); return up2(d);
}
__device__ __forceinline__ float2 csub(float2 a, float2 b) {
    u64 d; asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(pk2(a.x, a.y)), "l"(pk2(b.x, b.y))); return up2(d);
}
__device__ __forceinline__ float2 mul2(float2 a, float2 b) {
    u64 d; asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(pk2(a.x, a.y)), "l"(pk2(b.x, b.y))); return up2(d);
}
__device__ __forceinline__ float2 fma2(float2 a, float2 b, float2 c) {
    u64 d; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(pk2(a.x, a.y)), "l"(pk2(b.x, b.y)), "l"(pk2(c.x, c.y))); return up2(d);
}
__device__ __forceinline__ float2 cmul_i(float2 a) { return make_float2(-a.y, a.x); }
__device__ __forceinline__ float2 bc(float s) { return make_float2(s, s); }
// a * w = a.x * (w.x, w.y) + a.y * (-w.y, w.x)
__device__ __forceinline__ float2 cmul(float2 a, float2 w) { return fma2(bc(a.y), make_float2(-w.y, w.x), mul2(bc(a.x), w)); }
// the same with the rotated twiddle already in the table: W = (w.x, w.y, -w.y, w.x)
__device__ __forceinline__ float2 cmul_pre(float2 a, float4 W) {
    return fma2(bc(a.y), make_float2(W.z, W.w), mul2(bc(a.x), make_float2(W.x, W.y)));
}

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
        const float2 m = fma2(bc(-0.5f), s, v[0]);
        const float2 id = mul2(bc(s3), cmul_i(d));
        v[0] = cadd(v[0], s); v[1] = cadd(m, id); v[2] = csub(m, id);
    } else if constexpr (R == 7) {
        // X[m] = v0 + sum_k cos(2 pi m k / 7) (v[k] + v[7-k]) + i sum_k sin(2 pi m k / 7) (v[k] - v[7-k]),  k = 1..3
        const float c1 = 0.62348980185873353053f, c2 = -0.22252093395631440429f, c3 = -0.90096886790241912624f;
        const float s1 = 0.78183148246802980871f, s2 = 0.97492791218182360702f, s3 = 0.43388373911755812048f;
        const float2 a1 = cadd(v[1], v[6]), a2 = cadd(v[2], v[5]), a3 = cadd(v[3], v[4]);
        const float2 b1 = csub(v[1], v[6]), b2 = csub(v[2], v[5]), b3 = csub(v[3], v[4]);
        const float2 m1 = fma2(bc(c3), a3, fma2(bc(c2), a2, fma2(bc(c1), a1, v[0])));
        const float2 m2 = fma2(bc(c1), a3, fma2(bc(c3), a2, fma2(bc(c2), a1, v[0])));
        const float2 m3 = fma2(bc(c2), a3, fma2(bc(c1), a2, fma2(bc(c3), a1, v[0])));
        const float2 n1 = cmul_i(fma2(bc(s3), b3, fma2(bc(s2), b2, mul2(bc(s1), b1))));
        const float2 n2 = cmul_i(fma2(bc(-s1), b3, fma2(bc(-s3), b2, mul2(bc(s2), b1))));
        const float2 n3 = cmul_i(fma2(bc(s2), b3, fma2(bc(-s1), b2, mul2(bc(s3), b1))));
        v[0] = cadd(v[0], cadd(a1, cadd(a2, a3)));
        v[1] = cadd(m1, n1); v[6] = csub(m1, n1);
        v[2] = cadd(m2, n2); v[5] = csub(m2, n2);
        v[3] = cadd(m3, n3); v[4] = csub(m3, n3);
    } else {
        static_assert(R == 5, "radix");
        const float c1 = 0.30901699437494742410f, c2 = -0.80901699437494742410f;
        const float s1 = 0.95105651629515357212f, s2 = 0.58778525229247312917f;
        const float2 a1 = cadd(v[1], v[4]), a2 = cadd(v[2], v[3]);
        const float2 b1 = csub(v[1], v[4]), b2 = csub(v[2], v[3]);
        const float2 m1 = fma2(bc(c2), a2, fma2(bc(c1), a1, v[0]));
        const float2 m2 = fma2(bc(c1), a2, fma2(bc(c2), a1, v[0]));
        const float2 n1 = cmul_i(fma2(bc(s2), b2, mul2(bc(s1), b1)));
        const float2 n2 = cmul_i(fma2(bc(-s1), b2, mul2(bc(s2), b1)));
        v[0] = cadd(v[0], cadd(a1, a2));
        v[1] = cadd(m1, n1); v[4] = csub(m1, n1);
        v[2] = cadd(m2, n2); v[3] = csub(m2, n2);
    }
}

// ---- compile-time loops (indices must be constant expressions so the small-DFT twiddles fold into immediates) ----
template <int... I, typename F>
__device__ __forceinline__ void static_for_impl(std::integer_sequence<int, I...>, F && f) { (f(std::integral_constant<int, I>{}), ...); }
template <int N, typename F> __device__ __forceinline__ void static_for(F && f) { static_for_impl(std::make_integer_sequence<int, N>{}, f); }

template <int R> struct root_tab;      // e^{+2 pi i a / R}, a = 0..R-1
template <> struct root_tab<16> {
    static constexpr float c[16] = {1.0f, 0.923879533f, 0.707106781f, 0.382683432f, 0.0f, -0.382683432f, -0.707106781f, -0.923879533f, -1.0f, -0.923879533f, -0.707106781f, -0.382683432f, 0.0f, 0.382683432f, 0.707106781f, 0.923879533f};
    static constexpr float s[16] = {0.0f, 0.382683432f, 0.707106781f, 0.923879533f, 1.0f, 0.923879533f, 0.707106781f, 0.382683432f, 0.0f, -0.382683432f, -0.707106781f, -0.923879533f, -1.0f, -0.923879533f, -0.707106781f, -0.382683432f};
};
template <> struct root_tab<8> {
    static constexpr float c[8] = {1.0f, 0.707106781f, 0.0f, -0.707106781f, -1.0f, -0.707106781f, 0.0f, 0.707106781f};
    static constexpr float s[8] = {0.0f, 0.707106781f, 1.0f, 0.707106781f, 0.0f, -0.707106781f, -1.0f, -0.707106781f};
};
template <> struct root_tab<15> {
    static constexpr float c[15] = {1.0f, 0.913545458f, 0.669130606f, 0.309016994f, -0.104528463f, -0.5f, -0.809016994f, -0.978147601f, -0.978147601f, -0.809016994f, -0.5f, -0.104528463f, 0.309016994f, 0.669130606f, 0.913545458f};
    static constexpr float s[15] = {0.0f, 0.406736643f, 0.743144825f, 0.951056516f, 0.994521895f, 0.866025404f, 0.587785252f, 0.207911691f, -0.207911691f, -0.587785252f, -0.866025404f, -0.994521895f, -0.951056516f, -0.743144825f, -0.406736643f};
};
template <> struct root_tab<10> {
    static constexpr float c[10] = {1.0f, 0.809016994f, 0.309016994f, -0.309016994f, -0.809016994f, -1.0f, -0.809016994f, -0.309016994f, 0.309016994f, 0.809016994f};
    static constexpr float s[10] = {0.0f, 0.587785252f, 0.951056516f, 0.951056516f, 0.587785252f, 0.0f, -0.587785252f, -0.951056516f, -0.951056516f, -0.587785252f};
};

// DFT of composite size R1*R2 in registers (Cooley-Tukey, natural order in and out):
//   X[k1 + R1 k2] = sum_n2 W_R^{n2 k1} W_R2^{n2 k2} sum_n1 x[n2 + R2 n1] W_R1^{n1 k1}
template <int R1, int R2> __device__ __forceinline__ void dft_ct(float2 (&v)[R1 * R2]) {
    constexpr int R = R1 * R2;
    float2 y[R2][R1];
    static_for<R2>([&](auto N2) {
        constexpr int n2 = decltype(N2)::value;
        float2 t[R1];
        static_for<R1>([&](auto N1) { constexpr int n1 = decltype(N1)::value; t[n1] = v[n2 + R2 * n1]; });
        dft_small<R1>(t);
        static_for<R1>([&](auto K1) {
            constexpr int k1 = decltype(K1)::value;
            constexpr int a = (n2 * k1) % R;
            constexpr float wc = root_tab<R>::c[a], ws = root_tab<R>::s[a];   // folded by the front end: immediates
            if constexpr (a == 0) y[n2][k1] = t[k1];
            else y[n2][k1] = cmul(t[k1], make_float2(wc, ws));
        });
    });
    static_for<R1>([&](auto K1) {
        constexpr int k1 = decltype(K1)::value;
        float2 u[R2];
        static_for<R2>([&](auto N2) { constexpr int n2 = decltype(N2)::value; u[n2] = y[n2][k1]; });
        dft_small<R2>(u);
        static_for<R2>([&](auto K2) { constexpr int k2 = decltype(K2)::value; v[k1 + R1 * k2] = u[k2]; });
    });
}

template <int R> __device__ __forceinline__ void dft_any(float2 (&v)[R]) {
    if constexpr (R <= 5 || R == 7) dft_small<R>(v);
    else if constexpr (R == 16) dft_ct<4, 4>(v);
    else if constexpr (R == 8) dft_ct<2, 4>(v);
    else if constexpr (R == 15) dft_ct<3, 5>(v);
    else { static_assert(R == 10, "radix"); dft_ct<2, 5>(v); }
}

// Shared-memory index of complex element i while a frame is between passes: one pad slot per 16 elements, so the
// stride-R stores of the first pass (and the stride-16 ones after it) fall on distinct banks.
__device__ __forceinline__ int padi(int i) { return i + (i >> 4); }
template <int M> struct padded_len { static constexpr int value = (M + (M >> 4) + 1) & ~1; };   // even: 16-byte frame buffers

// One in-place Stockham pass by one warp: reads x[j + q*m] (q < R), writes x[(j-k)*R + k + p*Ls], k = j % Ls.
// All inputs of the pass sit in registers before the first store, so the pass runs in place; the LAST pass stores to
// the un-padded (linear) positions, which is what the overlap-add reads with 16-byte loads.
// tw: this pass's twiddles w = W_{Ls*R}^{q k}, compact and conflict-free at [(q-1)*Ls + k].  When Ls divides 32 every
// butterfly of a lane has the same k, so the R-1 twiddles are loaded once per frame and shared by the lane's butterflies.
template <int LS> struct tw_shared { static constexpr bool value = LS > 1 && (32 % LS) == 0; };
template <int R, int LS> struct tw_units { static constexpr int value = LS > 1 ? (R - 1) * LS : 0; };

template <int M, int R, int LS, bool LAST>
__device__ __forceinline__ void pass_inplace(float2 * buf, const float2 * tw, int lane) {
    constexpr int m = M / R;
    constexpr int B = (m + 31) / 32;
    float2 v[B][R];
#pragma unroll
    for (int i = 0; i < B; ++i) {
        const int j = lane + 32 * i;
        if (j < m) {
#pragma unroll
            for (int q = 0; q < R; ++q) v[i][q] = buf[padi(j + q * m)];
        }
    }
    float2 ws[tw_shared<LS>::value ? R : 1];
    if constexpr (tw_shared<LS>::value) {
#pragma unroll
        for (int q = 1; q < R; ++q) ws[q] = tw[(q - 1) * LS + lane % LS];
    }
    __syncwarp();
#pragma unroll
    for (int i = 0; i < B; ++i) {
        const int j = lane + 32 * i;
        if (j < m) {
            const int k = j % LS;
            if constexpr (tw_shared<LS>::value) {
#pragma unroll
                for (int q = 1; q < R; ++q) v[i][q] = cmul(v[i][q], ws[q]);
            } else if constexpr (LS > 1) {
#pragma unroll
                for (int q = 1; q < R; ++q) v[i][q] = cmul(v[i][q], tw[(q - 1) * LS + k]);
            }
            dft_any<R>(v[i]);
            const int base = (j - k) * R + k;
#pragma unroll
            for (int p = 0; p < R; ++p) {
                const int idx = base + p * LS;
                buf[LAST ? idx : padi(idx)] = v[i][p];
            }
        }
    }
    __syncwarp();
}

template <int M, int LS> __device__ __forceinline__ void run_passes(float2 *, const float2 *, int) {}
template <int M, int LS, int R, int... Rest>
__device__ __forceinline__ void run_passes(float2 * buf, const float2 * tw, int lane) {
    pass_inplace<M, R, LS, sizeof...(Rest) == 0>(buf, tw, lane);
    run_passes<M, LS * R, Rest...>(buf, tw + tw_units<R, LS>::value, lane);
}

// per-pass twiddle tables in float2 units, laid out in pass order (the first pass has none)
template <int M, int LS, int... Rs> struct tw_table_len { static constexpr int value = 0; };
template <int M, int LS, int R, int... Rest> struct tw_table_len<M, LS, R, Rest...> {
    static constexpr int value = tw_units<R, LS>::value + tw_table_len<M, LS * R, Rest...>::value;
};
template <int M, int LS> __device__ __forceinline__ void fill_tw(float2 *, const float2 *, int, int) {}
template <int M, int LS, int R, int... Rest>
__device__ __forceinline__ void fill_tw(float2 * dst, const float2 * __restrict__ tw_m, int tid, int nthreads) {
    if constexpr (LS > 1) {
        constexpr int TS = M / (LS * R);
        for (int e = tid; e < (R - 1) * LS; e += nthreads) {
            const int q = e / LS + 1, k = e - (q - 1) * LS;
            const float2 w = tw_m[q * k * TS];
            dst[e] = w;
        }
    }
    fill_tw<M, LS * R, Rest...>(dst + tw_units<R, LS>::value, tw_m, tid, nthreads);
}

// X_k = min(exp(logmag), 100) * (cos phi, sin phi) on the SFU
__device__ __forceinline__ float2 polar_fast(float logmag, float phase) {
    const float mag = fminf(__expf(logmag), 100.0f);
    const float kf = rintf(phase * 0.15915494309189535f);
    float r = fmaf(-kf, 6.2831854820251465f, phase);
    r = fmaf(-kf, -1.7484555e-7f, r);
    float s, c;
    __sincosf(r, &s, &c);
    return mul2(bc(mag), make_float2(c, s));
}

template <int FW, int N, int... RADIX>
__global__ void __launch_bounds__(FW * 32, 16 / FW) istft_fast_kernel(const fast_params P) {
    MIO_PDL_PROLOGUE();
    constexpr int FTHREADS = FW * 32;
    constexpr int M = N / 2, H = N / 4, CARRY = N - H, SPAN = FW * H, MP = padded_len<M>::value;
    constexpr int TWL = (tw_table_len<M, 1, RADIX...>::value + 1) & ~1;
    static_assert(H % 2 == 0 && MP % 2 == 0 && TWL % 2 == 0, "8-byte rows (16-byte when the hop is a multiple of 4)");
    extern __shared__ __align__(16) uint8_t smem[];
    float2 * s_buf = reinterpret_cast<float2 *>(smem);                  // [FW][MP]
    float2 * s_tw = s_buf + (size_t) FW * MP;                           // per-pass twiddles (TWL float2 units, even)
    float2 * s_twn = s_tw + TWL;                                        // [M/2 + 1]
    float * s_hs = reinterpret_cast<float *>(s_twn + (M / 2 + 2));      // [N]  hann * 1/n_freq
    float * s_env = s_hs + N;                                           // [H]  1 / sum of hann^2 over the 4 overlapping frames
    float * s_carry = s_env + H;                                        // [2][CARRY]
    float * s_xm = s_carry + 2 * CARRY;                                 // [FW]

    const int utt = blockIdx.y;
    const int S = P.n_frames[utt];
    const long long n_full = (long long) (S - 1) * H + N;
    const long long chunk = (long long) P.frames_per_chunk * H;
    const long long a = (long long) blockIdx.x * chunk;
    if (S <= 0 || a >= n_full) return;
    const long long b = a + chunk < n_full ? a + chunk : n_full;
    const long long out_len = n_full - 2 * P.pad;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    fill_tw<M, 1, RADIX...>(s_tw, P.tw_m, threadIdx.x, FTHREADS);
    for (int i = threadIdx.x; i <= M / 2; i += FTHREADS) { const float2 w = P.tw_n[i]; s_twn[i] = make_float2(0.5f * w.x, 0.5f * w.y); }
    for (int i = threadIdx.x; i < N; i += FTHREADS) s_hs[i] = P.hann[i] * P.scale;
    for (int i = threadIdx.x; i < 2 * CARRY; i += FTHREADS) s_carry[i] = 0.0f;
    for (int r = threadIdx.x; r < H; r += FTHREADS) {   // 4 overlapping frames, ascending frame = descending offset
        float e = 0.0f;
#pragma unroll
        for (int q = 3; q >= 0; --q) { const float hv = P.hann[r + q * H]; e += hv * hv; }
        s_env[r] = 1.0f / e;     // steady-state envelope (>= 1.5 for the periodic Hann at 4x overlap): multiply instead of divide
    }

    long long lf = a - N >= 0 ? (a - N) / H + 1 : 0;
    long long ll = (b - 1) / H;
    if (ll > S - 1) ll = S - 1;

    float2 * buf = s_buf + (size_t) warp * MP;
    const float * spec_u = P.spec + (size_t) P.spec_row0[utt] * P.pitch;
    // samples [ri_lo, ri_hi) relative to `a` are this CTA's to emit: inside the chunk and inside the cropped output
    const int ri_lo = (int) (P.pad - a > 0 ? P.pad - a : 0);
    const int ri_hi = (int) ((b < out_len + P.pad ? b : out_len + P.pad) - a);
    float * out_base = P.audio + P.audio_off[utt] + (a - P.pad);
    float peak = 0.0f;
    int cur = 0;
    __syncthreads();

    for (long long lr0 = lf; lr0 * H < b; lr0 += FW) {
        // ---------------- phase 1: one frame per warp ----------------
        const long long l = lr0 + warp;
        const bool valid = l <= ll;
        if (valid) {
            const float * fr = spec_u + (size_t) l * P.pitch;
            // The CTA's warps move through the phases together, so the memory system would idle during the FFT and
            // the overlap-add: pull the NEXT round's spectrum row into L2 now (one 128-byte line per lane, twice).
            if (l + FW <= ll) {
                const char * nx = reinterpret_cast<const char *>(fr + (size_t) FW * P.pitch);
#pragma unroll
                for (int c = 0; c < (N + 2) * 4; c += 32 * 128)
                    if (c + lane * 128 < (N + 2) * 4) asm volatile("prefetch.global.L2 [%0];" ::"l"(nx + c + lane * 128));
            }
            constexpr int ITER = (M / 2) / 32 + 1;
#pragma unroll 8
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
                        // E = (X_k + conj X_{M-k}) / 2, O = w_k (X_k - conj X_{M-k}) / 2 (s_twn holds w / 2):
                        // z_k = E + iO, z_{M-k} = conj(E - iO)
                        const float2 X2c = make_float2(X2.x, -X2.y);
                        const float2 iO = cmul_i(cmul(csub(Xk, X2c), s_twn[k]));
                        const float2 E = mul2(bc(0.5f), cadd(Xk, X2c));
                        buf[padi(k)] = cadd(E, iO);
                        if (k2 != k) { const float2 t = csub(E, iO); buf[padi(k2)] = make_float2(t.x, -t.y); }
                    }
                }
            }
            __syncwarp();
            run_passes<M, 1, RADIX...>(buf, s_tw, lane);
        }
        __syncthreads();

        // ---------------- phase 2: overlap-add the round four samples at a time, emit finished samples ----------------
        const int n_valid = (int) (ll - lr0 + 1 < FW ? ll - lr0 + 1 : FW);     // frames of this round that exist (a prefix)
        const int rel0 = (int) (lr0 * H - a);                                    // sample index of the round, relative to a
        const float * carry_in = s_carry + cur * CARRY;
        float * carry_out = s_carry + (cur ^ 1) * CARRY;
        // A thread keeps one 4-sample column r of the hop and walks the hops q = grp, grp + GROUPS, ...: its window
        // values hs[r + d*H] and its envelope are the same for every hop, so they are loaded once per round.
        if constexpr (H % 8 != 0) {
            // the crop (n_fft - hop) / 2 = 3 hop / 2 is not a multiple of 4 (hop 98 of the 44.1 kHz checkpoints: 147; hop 100:
            // 150), so 16-byte stores of emitted samples would be misaligned: two-sample columns, samples bounds-checked and
            // stored one by one
            constexpr int H2 = H / 2, GROUPS = FTHREADS / H2;
            const int grp = threadIdx.x / H2, r = (threadIdx.x - grp * H2) * 2;
            if (grp < GROUPS) {
                float2 hsd[4];
#pragma unroll
                for (int d = 0; d < 4; ++d) hsd[d] = *reinterpret_cast<const float2 *>(s_hs + r + d * H);
                const float2 ie = *reinterpret_cast<const float2 *>(s_env + r);
                for (int q = grp; q < FW + 3; q += GROUPS) {
                    const int x = q * H + r;
                    float2 a0 = make_float2(0.0f, 0.0f);
                    if (q < 3) a0 = *reinterpret_cast<const float2 *>(carry_in + x);
#pragma unroll
                    for (int d = 3; d >= 0; --d) {
                        const int w = q - d;
                        if (w >= 0 && w < n_valid) {
                            const float2 z = *reinterpret_cast<const float2 *>(reinterpret_cast<const float *>(s_buf + (size_t) w * MP) + r + d * H);
                            const float xm = s_xm[w];
                            a0 = fma2(cadd(z, make_float2(xm, -xm)), hsd[d], a0);   // r + d*H is even
                        }
                    }
                    if (q >= FW) { *reinterpret_cast<float2 *>(carry_out + (x - SPAN)) = a0; continue; }
                    const int ri = rel0 + x;
                    if (ri + 1 < ri_lo || ri >= ri_hi) continue;
                    const int fi = (int) lr0 + q;
                    float y[2];
                    if (fi >= 3 && fi <= S - 1) {
                        const float2 y0 = mul2(a0, ie);
                        y[0] = y0.x; y[1] = y0.y;
                    } else {
                        const int e_lo = fi - 3 > 0 ? fi - 3 : 0;
                        const int e_hi = fi < S - 1 ? fi : S - 1;
                        float e[2] = {0.0f, 0.0f};
                        for (int f = e_lo; f <= e_hi; ++f) {
#pragma unroll
                            for (int c = 0; c < 2; ++c) { const float hv = P.hann[r + c + (fi - f) * H]; e[c] += hv * hv; }
                        }
                        y[0] = e[0] > 1e-12f ? a0.x / e[0] : a0.x;
                        y[1] = e[1] > 1e-12f ? a0.y / e[1] : a0.y;
                    }
#pragma unroll
                    for (int c = 0; c < 2; ++c) {
                        if (ri + c >= ri_lo && ri + c < ri_hi) {
                            out_base[ri + c] = y[c];
                            const float av = fabsf(y[c]);
                            peak = fmaxf(peak, av <= 3.402823466e38f ? av : 0.0f);
                        }
                    }
                }
            }
        } else {
        constexpr int H4 = H / 4, GROUPS = FTHREADS / H4;
        const int grp = threadIdx.x / H4, r = (threadIdx.x - grp * H4) * 4;
        if (grp < GROUPS) {
            float4 hsd[4];
#pragma unroll
            for (int d = 0; d < 4; ++d) hsd[d] = *reinterpret_cast<const float4 *>(s_hs + r + d * H);
            const float4 ie = *reinterpret_cast<const float4 *>(s_env + r);
            for (int q = grp; q < FW + 3; q += GROUPS) {         // hops FW .. FW+2 are the carry into the next round
                const int x = q * H + r;
                float2 a0 = make_float2(0.0f, 0.0f), a1 = a0;
                if (q < 3) { const float4 c = *reinterpret_cast<const float4 *>(carry_in + x); a0 = make_float2(c.x, c.y); a1 = make_float2(c.z, c.w); }
#pragma unroll
                for (int d = 3; d >= 0; --d) {        // frame w = q - d: ascending frames, like the reference's accumulation order
                    const int w = q - d;
                    if (w >= 0 && w < n_valid) {
                        const float4 z = *reinterpret_cast<const float4 *>(reinterpret_cast<const float *>(s_buf + (size_t) w * MP) + r + d * H);
                        const float xm = s_xm[w];
                        const float2 nyq = make_float2(xm, -xm);        // the Nyquist bin alternates sign; r + d*H is a multiple of 4
                        a0 = fma2(cadd(make_float2(z.x, z.y), nyq), make_float2(hsd[d].x, hsd[d].y), a0);
                        a1 = fma2(cadd(make_float2(z.z, z.w), nyq), make_float2(hsd[d].z, hsd[d].w), a1);
                    }
                }
                if (q >= FW) {
                    *reinterpret_cast<float4 *>(carry_out + (x - SPAN)) = make_float4(a0.x, a0.y, a1.x, a1.y);
                    continue;
                }
                const int ri = rel0 + x;
                if (ri < ri_lo || ri >= ri_hi) continue;
                const int fi = (int) lr0 + q;                             // hop index of these four samples
                float4 y;
                if (fi >= 3 && fi <= S - 1) {
                    const float2 y0 = mul2(a0, make_float2(ie.x, ie.y)), y1 = mul2(a1, make_float2(ie.z, ie.w));
                    y = make_float4(y0.x, y0.y, y1.x, y1.y);
                } else {                                                  // first / last three hops: partial envelope
                    const int e_lo = fi - 3 > 0 ? fi - 3 : 0;
                    const int e_hi = fi < S - 1 ? fi : S - 1;
                    float e[4] = {0.0f, 0.0f, 0.0f, 0.0f};
                    for (int f = e_lo; f <= e_hi; ++f) {
#pragma unroll
                        for (int c = 0; c < 4; ++c) { const float hv = P.hann[r + c + (fi - f) * H]; e[c] += hv * hv; }
                    }
                    y = make_float4(e[0] > 1e-12f ? a0.x / e[0] : a0.x, e[1] > 1e-12f ? a0.y / e[1] : a0.y,
                                    e[2] > 1e-12f ? a1.x / e[2] : a1.x, e[3] > 1e-12f ? a1.y / e[3] : a1.y);
                }
                *reinterpret_cast<float4 *>(out_base + ri) = y;
                const float av[4] = {fabsf(y.x), fabsf(y.y), fabsf(y.z), fabsf(y.w)};
#pragma unroll
                for (int c = 0; c < 4; ++c) peak = fmaxf(peak, av[c] <= 3.402823466e38f ? av[c] : 0.0f);
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

template <int FW, int N, int... RADIX>
void launch_fast_t(fast_params P, const int * h_frames, int n_utt, int max_frames, cudaStream_t st) {
    constexpr int FTHREADS = FW * 32;
    constexpr int M = N / 2, H = N / 4, MP = padded_len<M>::value;
    constexpr int TWL = (tw_table_len<M, 1, RADIX...>::value + 1) & ~1;
    const size_t smem = (size_t) FW * MP * 8 + (size_t) TWL * 8 + (size_t) (M / 2 + 2) * 8 + (size_t) N * 4 + (size_t) H * 4 +
                        (size_t) 2 * (N - H) * 4 + FW * 8 + 64;
    static bool configured[64] = {false};
    static int n_sm[64] = {0};
    int dev = 0;
    MIO_CUDA(cudaGetDevice(&dev));
    if (!configured[dev & 63]) {
        MIO_CUDA(cudaFuncSetAttribute(istft_fast_kernel<FW, N, RADIX...>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem));
        MIO_CUDA(cudaDeviceGetAttribute(&n_sm[dev & 63], cudaDevAttrMultiProcessorCount, dev));
        configured[dev & 63] = true;
    }
    // Chunk length = rounds of FW frames minus the 3 halo frames a chunk recomputes.  16 / FW CTAs per SM: pick the round
    // count whose CTA total wastes the least in the last wave (cost = waves * rounds).
    const int sms = std::max(1, n_sm[dev & 63]) * (16 / FW);
    long long best_cost = -1;
    int best_rounds = 4;
    for (int rounds = 2 * (16 / FW); rounds <= 10 * (16 / FW); ++rounds) {
        const int fpc = rounds * FW - 3;
        long long ctas = 0;
        if (h_frames) for (int u = 0; u < n_utt; ++u) ctas += ((long long) h_frames[u] + 3 + fpc - 1) / fpc;
        else ctas = (long long) n_utt * ((max_frames + 3 + fpc - 1) / fpc);
        const long long cost = ((ctas + sms - 1) / sms) * rounds;
        if (best_cost < 0 || cost < best_cost) { best_cost = cost; best_rounds = rounds; }
    }
    P.frames_per_chunk = best_rounds * FW - 3;
    const long long n_full_max = (long long) (max_frames - 1) * H + N;
    const long long chunk = (long long) P.frames_per_chunk * H;
    dim3 grid((unsigned) ((n_full_max + chunk - 1) / chunk), (unsigned) n_utt);
    launch_pdl(istft_fast_kernel<FW, N, RADIX...>, dim3(grid), dim3(FTHREADS), smem, st, P);
    MIO_CUDA(cudaGetLastError());
    count_launch();
}

} // namespace

// returns false when the plan is outside the fast path (the generic kernel in istft.cu then runs)
bool launch_istft_fast(const istft_plan & p, const float * spec, int pitch, const int * spec_row0, const int * n_frames,
                       const long long * audio_off, float * audio, unsigned int * peak_bits, int n_utt, int max_frames,
                       cudaStream_t st) {
    if (p.hop * 4 != p.n_fft) return false;
    switch (p.n_fft) {
        case 1920: case 2048: case 1280: case 392: case 1024: case 512: case 800: case 640: case 400: case 960: case 1200: case 1600: break;
        default: return false;
    }
    fast_params P;
    P.n_freq = p.n_fft / 2 + 1;
    P.pad = (p.n_fft - p.hop) / 2;
    P.frames_per_chunk = 64 - 3;      // (set by launch_fast_t)
    static const int fw = [] { const char * e = std::getenv("MIO_TTS_B200_ISTFT_FW"); const int v = e ? std::atoi(e) : 8; return v == 16 || v == 4 ? v : 8; }();
    P.tw_m = p.d_tw_m; P.tw_n = p.d_tw_n; P.hann = p.d_hann;
    P.spec = spec; P.pitch = pitch; P.spec_row0 = spec_row0; P.n_frames = n_frames;
    P.audio_off = audio_off; P.audio = audio; P.peak_bits = peak_bits;
    P.scale = 1.0f / (float) P.n_freq;
    // the other common STFT sizes (n_fft = 4 hop) run with 8 frames per CTA only: one instantiation each
    switch (p.n_fft) {
        case 392: launch_fast_t<8, 392, 4, 7, 7>(P, nullptr, n_utt, max_frames, st); return true;
        case 1024: launch_fast_t<8, 1024, 16, 8, 4>(P, nullptr, n_utt, max_frames, st); return true;
        case 512: launch_fast_t<8, 512, 16, 4, 4>(P, nullptr, n_utt, max_frames, st); return true;
        case 800: launch_fast_t<8, 800, 16, 5, 5>(P, nullptr, n_utt, max_frames, st); return true;
        case 640: launch_fast_t<8, 640, 16, 5, 4>(P, nullptr, n_utt, max_frames, st); return true;
        case 400: launch_fast_t<8, 400, 10, 5, 4>(P, nullptr, n_utt, max_frames, st); return true;
        case 960: launch_fast_t<8, 960, 16, 10, 3>(P, nullptr, n_utt, max_frames, st); return true;
        case 1200: launch_fast_t<8, 1200, 15, 10, 4>(P, nullptr, n_utt, max_frames, st); return true;
        case 1600: launch_fast_t<8, 1600, 16, 10, 5>(P, nullptr, n_utt, max_frames, st); return true;
        default: break;
    }
#define MIO_ISTFT_CASE(FWV)                                                                               \
    switch (p.n_fft) {                                                                                    \
        case 1920: launch_fast_t<FWV, 1920, 16, 15, 4>(P, nullptr, n_utt, max_frames, st); break;          \
        case 2048: launch_fast_t<FWV, 2048, 16, 16, 4>(P, nullptr, n_utt, max_frames, st); break;          \
        default: launch_fast_t<FWV, 1280, 16, 10, 4>(P, nullptr, n_utt, max_frames, st); break;           \
    }
    if (fw == 16) { MIO_ISTFT_CASE(16) } else if (fw == 4) { MIO_ISTFT_CASE(4) } else { MIO_ISTFT_CASE(8) }
#undef MIO_ISTFT_CASE
    return true;
}

} // namespace mio
