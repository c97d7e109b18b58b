// SIMT kernels of the reference-audio -> speaker-embedding path:
// WavLM conv0 + per-channel GroupNorm over time, relayout/cast, gated relative-position gate, layer-output
// accumulation, ConvNeXt depthwise conv + LayerNorm, attentive statistics pooling and the final projection.
// The dense contractions of this path run on the tcgen05 GEMM (gemm_tcgen05.cu); attention is attention_full.cu.
// Reference semantics are cited per kernel (paths relative to /root/reference).
#include "refenc.h"

#include <cuda_bf16.h>
#include <cuda_fp16.h>

#include <algorithm>

namespace mio {
namespace {

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}
__device__ __forceinline__ float f16r(float x) { return __half2float(__float2half_rn(x)); }
__device__ __forceinline__ uint16_t to16(float v, int dtype) {
    if (dtype == OP_BF16) { __nv_bfloat16 h = __float2bfloat16_rn(v); return *reinterpret_cast<uint16_t *>(&h); }
    __half h = __float2half_rn(v);
    return *reinterpret_cast<uint16_t *>(&h);
}
// ggml_gelu on the CPU backend: fp16 table (see act_apply in gemm_tcgen05.cu)
__device__ __forceinline__ float gelu_ggml(float x) {
    const float h = f16r(x);
    const float u = 0.79788456080286535587989211986876f * h * (1.0f + 0.044715f * h * h);
    const float y = f16r(0.5f * h * (1.0f + tanhf(u)));
    return x <= -10.0f ? 0.0f : (x >= 10.0f ? x : y);
}

constexpr int C0_FRAMES = 128;   // conv0 output frames per block

// ---- WavLM conv0 (1 -> C channels, kernel K <= 16, stride S, no bias; ggml_conv_1d rounds both operands to fp16,
//      src/wavlm-extractor.cpp:664-667) evaluated on the fly from the waveform: the per-channel statistics of the
//      one-group-per-channel GroupNorm over time (:669-679) come from wav_autocorr_kernel below, this kernel computes the
//      convolution, normalises, applies affine + GELU and writes fp16 rows (generic K <= 16 / stride; the WavLM shape
//      K = 10, S = 5 runs wav_conv0_apply_fast_kernel).  Nothing fp32 of size [frames, C] ever touches HBM. ----
__global__ void __launch_bounds__(256) wav_conv0_kernel(const float * __restrict__ wav, const long long * __restrict__ wav_off,
                                                        const int * __restrict__ n_frames, const int * __restrict__ row0,
                                                        const float * __restrict__ w, int C, int K, int S,
                                                        const float2 * __restrict__ mean_rstd, const float * __restrict__ gw,
                                                        const float * __restrict__ gb, __half * __restrict__ out) {
    MIO_PDL_PROLOGUE();
    __shared__ float xs[C0_FRAMES * 8 + 32];
    const int clip = blockIdx.y, chunk = blockIdx.x;
    const int L = n_frames[clip];
    const int t0 = chunk * C0_FRAMES;
    if (t0 >= (L + C0_FRAMES - 1) / C0_FRAMES * C0_FRAMES) return;   // past this clip's (chunk-padded) extent
    const int nt = min(C0_FRAMES, L - t0);           // valid frames of this chunk
    const float * x = wav + wav_off[clip];
    if (nt > 0) {
        const int span = (nt - 1) * S + K;
        for (int i = threadIdx.x; i < span; i += blockDim.x) xs[i] = f16r(x[(long long) t0 * S + i]);
    }
    __syncthreads();
    for (int c = threadIdx.x; c < C; c += blockDim.x) {
        float wk[16];
#pragma unroll
        for (int j = 0; j < 16; ++j) wk[j] = j < K ? f16r(w[c * K + j]) : 0.0f;
        const float2 mr = mean_rstd[(size_t) clip * C + c];
        const float a = gw[c] * mr.y, b = gb[c] - mr.x * mr.y * gw[c];
        __half * o = out + ((size_t) row0[clip] + t0) * C + c;
        for (int t = 0; t < C0_FRAMES; ++t) {
            float v = 0.0f;
            if (t < nt) {
                float y = 0.0f;
#pragma unroll
                for (int j = 0; j < 16; ++j) if (j < K) y = fmaf(wk[j], xs[t * S + j], y);
                v = gelu_ggml(fmaf(y, a, b));
            }
            o[(size_t) t * C] = __float2half_rn(v);   // padding rows of the clip are written as zeros
        }
    }
}

// ---- conv0 statistics without evaluating the convolution: y_c[t] = sum_k w_c[k] x[S t + k], so over the frames of a clip
//      sum_t y_c = w_c . s            with s[k]    = sum_t x[S t + k]
//      sum_t y_c^2 = w_c^T R w_c      with R[k,k'] = sum_t x[S t + k] x[S t + k']
//      -- K + K*K numbers per clip (fp16-rounded x like the convolution's operand, accumulated in double) instead of a
//      pass of 2*K*C flops per frame.  partial[clip][chunk][a], a < K*K + K: a < K -> s[a], else R[(a-K)/K][(a-K)%K]. ----
__global__ void __launch_bounds__(128) wav_autocorr_kernel(const float * __restrict__ wav, const long long * __restrict__ wav_off,
                                                           const int * __restrict__ n_frames, int K, int S,
                                                           double * __restrict__ partial, int n_chunks) {
    MIO_PDL_PROLOGUE();
    __shared__ float xs[C0_FRAMES * 8 + 32];
    const int clip = blockIdx.y, chunk = blockIdx.x;
    const int L = n_frames[clip];
    const int t0 = chunk * C0_FRAMES;
    if (t0 >= L) return;
    const int nt = min(C0_FRAMES, L - t0);
    const float * x = wav + wav_off[clip];
    const int span = (nt - 1) * S + K;
    for (int i = threadIdx.x; i < span; i += blockDim.x) xs[i] = f16r(x[(long long) t0 * S + i]);
    __syncthreads();
    const int NA = K * K + K;
    double * p = partial + ((size_t) clip * n_chunks + chunk) * NA;
    for (int a = threadIdx.x; a < NA; a += blockDim.x) {
        double acc = 0.0;
        if (a < K) {
            for (int t = 0; t < nt; ++t) acc += (double) xs[t * S + a];
        } else {
            const int k = (a - K) / K, k2 = (a - K) - k * K;
            for (int t = 0; t < nt; ++t) acc += (double) xs[t * S + k] * (double) xs[t * S + k2];
        }
        p[a] = acc;
    }
}

// (s, R) partials of a clip, summed in chunk order (deterministic) -> (mean, rstd) per channel;
// ggml_group_norm: biased variance, eps inside the sqrt (ggml-cpu/ops.cpp:3955-4028)
__global__ void __launch_bounds__(256) wav_conv0_moments_kernel(const double * __restrict__ partial, const int * __restrict__ n_frames,
                                                                const float * __restrict__ w, int C, int K, int n_chunks, float eps,
                                                                float2 * __restrict__ mean_rstd) {
    MIO_PDL_PROLOGUE();
    __shared__ double acc[16 * 16 + 16];
    const int clip = blockIdx.x;
    const int NA = K * K + K;
    const int used = (n_frames[clip] + C0_FRAMES - 1) / C0_FRAMES;
    for (int a = threadIdx.x; a < NA; a += blockDim.x) {
        double v = 0.0;
        for (int k = 0; k < used; ++k) v += partial[((size_t) clip * n_chunks + k) * NA + a];
        acc[a] = v;
    }
    __syncthreads();
    const double n = (double) max(1, n_frames[clip]);
    for (int c = threadIdx.x; c < C; c += blockDim.x) {
        double wk[16];
        for (int j = 0; j < K; ++j) wk[j] = (double) f16r(w[c * K + j]);
        double s = 0.0, q = 0.0;
        for (int j = 0; j < K; ++j) {
            s += wk[j] * acc[j];
            double row = 0.0;
            for (int j2 = 0; j2 < K; ++j2) row += wk[j2] * acc[K + j * K + j2];
            q += wk[j] * row;
        }
        const double mean = s / n;
        const double var = fmax(q / n - mean * mean, 0.0);
        mean_rstd[(size_t) clip * C + c] = make_float2((float) mean, (float) (1.0 / sqrt(var + (double) eps)));
    }
}

// ---- conv0 + GroupNorm affine + GELU, register-tiled: a thread owns two adjacent channels (packed fp32x2 FMAs, one
//      half2 store per frame) and walks the chunk four frames at a time (the 25 samples those frames touch come from
//      shared memory as seven 16-byte loads).  K = 10, S = 5 (WavLM's first layer); other shapes use wav_conv0_kernel. ----
typedef unsigned long long u64c;
__device__ __forceinline__ float2 ffma2(float2 a, float2 b, float2 c) {
    u64c d, ra, rb, rc;
    asm("mov.b64 %0, {%1, %2};" : "=l"(ra) : "f"(a.x), "f"(a.y));
    asm("mov.b64 %0, {%1, %2};" : "=l"(rb) : "f"(b.x), "f"(b.y));
    asm("mov.b64 %0, {%1, %2};" : "=l"(rc) : "f"(c.x), "f"(c.y));
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(ra), "l"(rb), "l"(rc));
    float2 r;
    asm("mov.b64 {%0, %1}, %2;" : "=f"(r.x), "=f"(r.y) : "l"(d));
    return r;
}
// the GEMM epilogue's GELU (act_apply, gemm_tcgen05.cu): ggml's fp16 table reproduced by two roundings, SFU tanh
__device__ __forceinline__ float gelu_fast(float x) {
    const float h = f16r(x);
    const float u = 0.79788456080286535587989211986876f * h * (1.0f + 0.044715f * h * h);
    const float t = 1.0f - __fdividef(2.0f, 1.0f + __expf(2.0f * u));
    const float y = f16r(0.5f * h * (1.0f + t));
    return x <= -10.0f ? 0.0f : (x >= 10.0f ? x : y);
}

__global__ void __launch_bounds__(256) wav_conv0_apply_fast_kernel(const float * __restrict__ wav, const long long * __restrict__ wav_off,
                                                                   const int * __restrict__ n_frames, const int * __restrict__ row0,
                                                                   const float * __restrict__ w, int C,
                                                                   const float2 * __restrict__ mean_rstd, const float * __restrict__ gw,
                                                                   const float * __restrict__ gb, __half * __restrict__ out) {
    MIO_PDL_PROLOGUE();
    constexpr int K = 10, S = 5, TF = 4;
    __shared__ __align__(16) float xs[C0_FRAMES * S + 32];
    const int clip = blockIdx.y, chunk = blockIdx.x;
    const int L = n_frames[clip];
    const int t0 = chunk * C0_FRAMES;
    if (t0 >= (L + C0_FRAMES - 1) / C0_FRAMES * C0_FRAMES) return;   // past this clip's (chunk-padded) extent
    const int nt = max(0, min(C0_FRAMES, L - t0));                    // valid frames of this chunk
    const float * x = wav + wav_off[clip];
    const int span = nt > 0 ? (nt - 1) * S + K : 0;
    for (int i = threadIdx.x; i < C0_FRAMES * S + 32; i += blockDim.x) xs[i] = i < span ? f16r(x[(long long) t0 * S + i]) : 0.0f;
    __syncthreads();
    for (int cp = threadIdx.x; cp < (C >> 1); cp += blockDim.x) {
        const int c = cp * 2;
        float2 wk[K];
#pragma unroll
        for (int j = 0; j < K; ++j) wk[j] = make_float2(f16r(w[c * K + j]), f16r(w[(c + 1) * K + j]));
        const float2 m0 = mean_rstd[(size_t) clip * C + c], m1 = mean_rstd[(size_t) clip * C + c + 1];
        const float2 sc = make_float2(gw[c] * m0.y, gw[c + 1] * m1.y);
        const float2 sh = make_float2(gb[c] - m0.x * m0.y * gw[c], gb[c + 1] - m1.x * m1.y * gw[c + 1]);
        __half2 * o = reinterpret_cast<__half2 *>(out + ((size_t) row0[clip] + t0) * C + c);
        for (int t = 0; t < C0_FRAMES; t += TF) {
            float xv[28];
#pragma unroll
            for (int i = 0; i < 7; ++i) {
                const float4 v = *reinterpret_cast<const float4 *>(xs + t * S + 4 * i);
                xv[4 * i] = v.x; xv[4 * i + 1] = v.y; xv[4 * i + 2] = v.z; xv[4 * i + 3] = v.w;
            }
#pragma unroll
            for (int f = 0; f < TF; ++f) {
                float2 y = make_float2(0.0f, 0.0f);
#pragma unroll
                for (int j = 0; j < K; ++j) y = ffma2(wk[j], make_float2(xv[f * S + j], xv[f * S + j]), y);
                y = ffma2(y, sc, sh);
                const bool live = t + f < nt;                  // padding rows of the clip are written as zeros
                o[(size_t) (t + f) * (C >> 1)] = __floats2half2_rn(live ? gelu_fast(y.x) : 0.0f, live ? gelu_fast(y.y) : 0.0f);
            }
        }
    }
}

// ---- row relayout + cast: dst row r takes src row map[r] (or zeros when map[r] < 0), scaled, as fp32 and/or 16-bit ----
__global__ void relayout_kernel(const float * __restrict__ src, const int * __restrict__ map, int rows, int dim, float scale,
                                float * __restrict__ dst32, void * __restrict__ dst16, int dtype) {
    MIO_PDL_PROLOGUE();
    const size_t i = (size_t) blockIdx.x * blockDim.x + threadIdx.x;    // float4 index
    const int d4 = dim >> 2;
    if (i >= (size_t) rows * d4) return;
    const int r = (int) (i / d4), c4 = (int) (i - (size_t) r * d4);
    const int s = map[r];
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (s >= 0) {
        v = reinterpret_cast<const float4 *>(src + (size_t) s * dim)[c4];
        v.x *= scale; v.y *= scale; v.z *= scale; v.w *= scale;
    }
    if (dst32) reinterpret_cast<float4 *>(dst32 + (size_t) r * dim)[c4] = v;
    if (dst16) {
        uint2 o;
        o.x = (uint32_t) to16(v.x, dtype) | ((uint32_t) to16(v.y, dtype) << 16);
        o.y = (uint32_t) to16(v.z, dtype) | ((uint32_t) to16(v.w, dtype) << 16);
        reinterpret_cast<uint2 *>((uint16_t *) dst16 + (size_t) r * dim)[c4] = o;
    }
}

// ---- acc = first ? x : acc + x  (mean over the layer outputs, src/wavlm-extractor.cpp:853-865) ----
__global__ void accumulate_kernel(const float * __restrict__ x, float * __restrict__ acc, size_t n4, int first) {
    MIO_PDL_PROLOGUE();
    const size_t i = (size_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n4) return;
    float4 v = reinterpret_cast<const float4 *>(x)[i];
    if (!first) {
        const float4 a = reinterpret_cast<const float4 *>(acc)[i];
        v.x += a.x; v.y += a.y; v.z += a.z; v.w += a.w;
    }
    reinterpret_cast<float4 *>(acc)[i] = v;
}

// ---- WavLM gated relative-position gate (src/wavlm-extractor.cpp:762-790).  For (time t, head h) of a clip:
//      g = Wg x[t, h*hd:(h+1)*hd] + bg (8 values), a = sigmoid(g0+..+g3), b = sigmoid(g4+..+g7).  The reference lays
//      these out head-fastest (j = h + H*t) and then REINTERPRETS the vector as [head', time'] with time fastest, so
//      the value lands at (q = j % seq, h' = j / seq) and is combined with gru_const[h']:
//      gate[h'][q] = a * (b * const[h'] - 1) + 2.  One warp per (t, h). ----
__global__ void wavlm_gate_kernel(const float * __restrict__ x, const int * __restrict__ row0, const int * __restrict__ seq_len,
                                  const int * __restrict__ gate_off, int n_clips, int max_seq, int H, int hd,
                                  const float * __restrict__ gw, const float * __restrict__ gb,
                                  const float * __restrict__ gconst, float * __restrict__ gate) {
    MIO_PDL_PROLOGUE();
    const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    const int per_clip = max_seq * H;
    if (warp >= n_clips * per_clip) return;
    const int clip = warp / per_clip, j = warp - clip * per_clip;
    const int seq = seq_len[clip];
    if (j >= seq * H) return;
    const int t = j / H, h = j - t * H;
    const float * xr = x + ((size_t) row0[clip] + t) * (size_t) (H * hd) + (size_t) h * hd;
    float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    for (int d = lane; d < hd; d += 32) {
        const float xv = xr[d];
#pragma unroll
        for (int o = 0; o < 8; ++o) acc[o] = fmaf(gw[o * hd + d], xv, acc[o]);
    }
#pragma unroll
    for (int o = 0; o < 8; ++o) acc[o] = warp_sum(acc[o]) + gb[o];
    if (lane == 0) {
        const float a = 1.0f / (1.0f + expf(-((acc[0] + acc[1]) + (acc[2] + acc[3]))));
        const float b = 1.0f / (1.0f + expf(-((acc[4] + acc[5]) + (acc[6] + acc[7]))));
        const int q = j % seq, hh = j / seq;
        gate[(size_t) gate_off[clip] + (size_t) hh * seq + q] = a * (b * gconst[hh] - 1.0f) + 2.0f;
    }
}

// ---- ConvNeXt block front half (src/miocodec-decoder.cpp:886-895): depthwise conv over time (kernel KW, "same"
//      padding inside the clip, fp16-rounded operands like ggml_conv_1d_dw_ph) + bias, then LayerNorm(eps) with
//      affine over channels -> 16-bit GEMM operand.  One warp per row; gap rows are written as zeros. ----
__global__ void dwconv_ln_kernel(const float * __restrict__ x, const int * __restrict__ row_clip, const int * __restrict__ row_pos,
                                 const int * __restrict__ seq_len, int rows, int dim, int KW, const float * __restrict__ w,
                                 const float * __restrict__ b, const float * __restrict__ nw, const float * __restrict__ nb,
                                 float eps, void * __restrict__ out16, int dtype) {
    MIO_PDL_PROLOGUE();
    extern __shared__ float dsm[];                       // [warps][dim]
    const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (row >= rows) return;
    const int lane = threadIdx.x & 31;
    float * y = dsm + (size_t) (threadIdx.x >> 5) * dim;
    uint16_t * o = (uint16_t *) out16 + (size_t) row * dim;
    const int clip = row_clip[row];
    if (clip < 0) { for (int c = lane; c < dim; c += 32) o[c] = 0; return; }
    const int pos = row_pos[row], T = seq_len[clip], half = KW / 2;
    float s = 0.0f;
    for (int c = lane; c < dim; c += 32) {
        float acc = 0.0f;
        for (int j = 0; j < KW; ++j) {
            const int p = pos + j - half;
            if (p >= 0 && p < T) acc = fmaf(f16r(w[c * KW + j]), f16r(x[(size_t) (row + j - half) * dim + c]), acc);
        }
        acc += b[c];
        y[c] = acc;
        s += acc;
    }
    const float mean = warp_sum(s) / (float) dim;
    float q = 0.0f;
    for (int c = lane; c < dim; c += 32) { const float d = y[c] - mean; q += d * d; }
    const float rstd = 1.0f / sqrtf(warp_sum(q) / (float) dim + eps);
    for (int c = lane; c < dim; c += 32) o[c] = to16((y[c] - mean) * rstd * nw[c] + nb[c], dtype);
}

// ---- attentive statistics pooling (src/miocodec-decoder.cpp:913-929): per (clip, channel) softmax over TIME of the
//      attention logits, then mean = sum alpha x, std = sqrt(clamp(sum alpha x^2 - mean^2, 1e-4, 1e4));
//      stat[clip] = [mean(dim) | std(dim)].  One warp per (clip, channel). ----
__global__ void attn_stats_pool_kernel(const float * __restrict__ logits, const float * __restrict__ x,
                                       const int * __restrict__ row0, const int * __restrict__ seq_len, int n_clips, int dim,
                                       float * __restrict__ stat) {
    MIO_PDL_PROLOGUE();
    const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (warp >= n_clips * dim) return;
    const int clip = warp / dim, c = warp - clip * dim;
    const int T = seq_len[clip];
    const float * a = logits + (size_t) row0[clip] * dim + c;
    const float * xv = x + (size_t) row0[clip] * dim + c;
    float mx = -INFINITY;
    for (int t = lane; t < T; t += 32) mx = fmaxf(mx, a[(size_t) t * dim]);
    mx = warp_max(mx);
    float se = 0.0f, s1 = 0.0f, s2 = 0.0f;
    for (int t = lane; t < T; t += 32) {
        const float e = expf(a[(size_t) t * dim] - mx);
        const float v = xv[(size_t) t * dim];
        se += e; s1 += e * v; s2 += e * v * v;
    }
    se = warp_sum(se); s1 = warp_sum(s1); s2 = warp_sum(s2);
    if (lane == 0) {
        const float mean = s1 / se, m2 = s2 / se;
        const float var = fminf(fmaxf(m2 - mean * mean, 1e-4f), 1e4f);
        stat[(size_t) clip * 2 * dim + c] = mean;
        stat[(size_t) clip * 2 * dim + dim + c] = sqrtf(var);
    }
}

// ---- embedding = LayerNorm(W stat + b) (:935-937), zeros if anything is non-finite (:1048-1061).  One block per clip. ----
__global__ void pool_proj_ln_kernel(const float * __restrict__ stat, int in_dim, const float * __restrict__ w,
                                    const float * __restrict__ b, const float * __restrict__ nw, const float * __restrict__ nb,
                                    float eps, int out_dim, float * __restrict__ emb) {
    MIO_PDL_PROLOGUE();
    extern __shared__ float psm[];                       // [out_dim] + [2]
    const int clip = blockIdx.x;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, n_warps = blockDim.x >> 5;
    const float * s = stat + (size_t) clip * in_dim;
    for (int o = warp; o < out_dim; o += n_warps) {
        float acc = 0.0f;
        for (int k = lane; k < in_dim; k += 32) acc = fmaf(w[(size_t) o * in_dim + k], s[k], acc);
        acc = warp_sum(acc);
        if (lane == 0) psm[o] = acc + b[o];
    }
    __syncthreads();
    if (warp == 0) {
        float t = 0.0f;
        for (int o = lane; o < out_dim; o += 32) t += psm[o];
        const float mean = warp_sum(t) / (float) out_dim;
        float q = 0.0f;
        for (int o = lane; o < out_dim; o += 32) { const float d = psm[o] - mean; q += d * d; }
        const float rstd = 1.0f / sqrtf(warp_sum(q) / (float) out_dim + eps);
        int bad = 0;
        for (int o = lane; o < out_dim; o += 32) {
            const float v = (psm[o] - mean) * rstd * nw[o] + nb[o];
            psm[o] = v;
            bad |= !isfinite(v);
        }
        bad = __any_sync(0xffffffffu, bad);
        for (int o = lane; o < out_dim; o += 32) emb[(size_t) clip * out_dim + o] = bad ? 0.0f : psm[o];
    }
}

// ---- front end (src/wavlm-extractor.cpp:205-240): peak = max|x| per clip, then x / (peak + 1e-8) resampled linearly
//      to the WavLM rate with double-precision positions, exactly like normalize_audio + resample_linear ----
__global__ void wav_peak_kernel(const float * __restrict__ src, const long long * __restrict__ src_off,
                                const int * __restrict__ n_src, unsigned int * __restrict__ peak_bits) {
    MIO_PDL_PROLOGUE();
    const int clip = blockIdx.y;
    const float * x = src + src_off[clip];
    const int n = n_src[clip];
    float m = 0.0f;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) m = fmaxf(m, fabsf(x[i]));
    m = warp_max(m);
    if ((threadIdx.x & 31) == 0 && m > 0.0f) atomicMax(peak_bits + clip, __float_as_uint(m));   // non-negative floats order as uints
}

__global__ void wav_resample_kernel(const float * __restrict__ src, const long long * __restrict__ src_off,
                                    const int * __restrict__ n_src, float * __restrict__ wav,
                                    const long long * __restrict__ wav_off, const int * __restrict__ n_wav,
                                    const unsigned int * __restrict__ peak_bits, int rate_in, int rate_out) {
    MIO_PDL_PROLOGUE();
    const int clip = blockIdx.y;
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_wav[clip]) return;
    const float * x = src + src_off[clip];
    const float m = __uint_as_float(peak_bits[clip]) + 1e-8f;
    float v;
    if (rate_in == rate_out) {
        v = x[i] / m;
    } else {
        const double ratio = (double) rate_out / (double) rate_in;
        const double pos = (double) i / ratio;
        const long long i0 = (long long) floor(pos);
        const long long i1 = min((long long) n_src[clip] - 1, i0 + 1);
        const double t = pos - (double) i0;
        v = (float) ((1.0 - t) * (double) (x[max(0ll, i0)] / m) + t * (double) (x[i1] / m));
    }
    wav[wav_off[clip] + i] = v;
}

inline unsigned blocks_for(size_t n, int threads) { return (unsigned) ((n + threads - 1) / threads); }

} // namespace

void launch_wav_prepare(const float * src, const long long * src_off, const int * n_src, float * wav, const long long * wav_off,
                        const int * n_wav, unsigned int * peak_bits, int rate_in, int rate_out, int n_clips, int max_n_src,
                        int max_n_wav, cudaStream_t st) {
    MIO_CUDA(cudaMemsetAsync(peak_bits, 0, (size_t) n_clips * sizeof(unsigned int), st));
    const unsigned gx = std::min<unsigned>(64, blocks_for((size_t) std::max(1, max_n_src), 256));
    launch_pdl(wav_peak_kernel, dim3(gx, n_clips), dim3(256), 0, st, src, src_off, n_src, peak_bits);
    launch_pdl(wav_resample_kernel, dim3(blocks_for((size_t) std::max(1, max_n_wav), 256), n_clips), dim3(256), 0, st, src,
               src_off, n_src, wav, wav_off, n_wav, (const unsigned int *) peak_bits, rate_in, rate_out);
    count_launch(2);
}

int wav_conv0_chunk_frames() { return C0_FRAMES; }

void launch_wav_conv0_moments(const float * wav, const long long * wav_off, const int * n_frames, const float * w, int C, int K,
                              int S, double * partial, int n_chunks, float eps, float2 * mean_rstd, int n_clips, cudaStream_t st) {
    if (K > 16 || (C0_FRAMES - 1) * S + K > C0_FRAMES * 8 + 32) throw cuda_error("wavlm conv0: unsupported kernel/stride");
    launch_pdl(wav_autocorr_kernel, dim3(n_chunks, n_clips), dim3(128), 0, st, wav, wav_off, n_frames, K, S, partial, n_chunks);
    launch_pdl(wav_conv0_moments_kernel, dim3(n_clips), dim3(256), 0, st, (const double *) partial, n_frames, w, C, K, n_chunks, eps,
               mean_rstd);
    count_launch(2);
}

void launch_wav_conv0_apply(const float * wav, const long long * wav_off, const int * n_frames, const int * row0,
                            const float * w, int C, int K, int S, int n_chunks, const float2 * mean_rstd, const float * gw,
                            const float * gb, void * out16, int n_clips, cudaStream_t st) {
    if (K == 10 && S == 5 && (C & 1) == 0)
        launch_pdl(wav_conv0_apply_fast_kernel, dim3(n_chunks, n_clips), dim3(256), 0, st, wav, wav_off, n_frames, row0, w, C,
                   mean_rstd, gw, gb, (__half *) out16);
    else
        launch_pdl(wav_conv0_kernel, dim3(n_chunks, n_clips), dim3(256), 0, st, wav, wav_off, n_frames, row0, w, C, K, S,
                   mean_rstd, gw, gb, (__half *) out16);
    count_launch();
}

void launch_relayout(const float * src, const int * map, int rows, int dim, float scale, float * dst32, void * dst16,
                     int dtype, cudaStream_t st) {
    if (rows <= 0) return;
    if (dim & 3) throw cuda_error("relayout: dim must be a multiple of 4");
    launch_pdl(relayout_kernel, dim3(blocks_for((size_t) rows * (dim >> 2), 256)), dim3(256), 0, st, src, map, rows, dim, scale,
               dst32, dst16, dtype);
    count_launch();
}

void launch_accumulate(const float * x, float * acc, size_t n, bool first, cudaStream_t st) {
    if (n == 0) return;
    if (n & 3) throw cuda_error("accumulate: element count must be a multiple of 4");
    launch_pdl(accumulate_kernel, dim3(blocks_for(n >> 2, 256)), dim3(256), 0, st, x, acc, n >> 2, first ? 1 : 0);
    count_launch();
}

void launch_wavlm_gate(const float * x, const int * row0, const int * seq_len, const int * gate_off, int n_clips, int max_seq,
                       int H, int hd, const float * gw, const float * gb, const float * gconst, float * gate, cudaStream_t st) {
    const size_t warps = (size_t) n_clips * max_seq * H;
    if (warps == 0) return;
    launch_pdl(wavlm_gate_kernel, dim3(blocks_for(warps * 32, 256)), dim3(256), 0, st, x, row0, seq_len, gate_off, n_clips,
               max_seq, H, hd, gw, gb, gconst, gate);
    count_launch();
}

void launch_dwconv_ln(const float * x, const int * row_clip, const int * row_pos, const int * seq_len, int rows, int dim, int KW,
                      const float * w, const float * b, const float * nw, const float * nb, float eps, void * out16, int dtype,
                      cudaStream_t st) {
    if (rows <= 0) return;
    const size_t smem = (size_t) 8 * dim * sizeof(float);
    if (smem > 48 * 1024) throw cuda_error("dwconv_ln: channel count too large");
    launch_pdl(dwconv_ln_kernel, dim3(blocks_for((size_t) rows, 8)), dim3(256), smem, st, x, row_clip, row_pos, seq_len, rows,
               dim, KW, w, b, nw, nb, eps, out16, dtype);
    count_launch();
}

void launch_attn_stats_pool(const float * logits, const float * x, const int * row0, const int * seq_len, int n_clips, int dim,
                            float * stat, cudaStream_t st) {
    launch_pdl(attn_stats_pool_kernel, dim3(blocks_for((size_t) n_clips * dim * 32, 256)), dim3(256), 0, st, logits, x, row0,
               seq_len, n_clips, dim, stat);
    count_launch();
}

void launch_pool_proj_ln(const float * stat, int in_dim, const float * w, const float * b, const float * nw, const float * nb,
                         float eps, int out_dim, float * emb, int n_clips, cudaStream_t st) {
    const size_t smem = (size_t) (out_dim + 2) * sizeof(float);
    launch_pdl(pool_proj_ln_kernel, dim3(n_clips), dim3(256), smem, st, stat, in_dim, w, b, nw, nb, eps, out_dim, emb);
    count_launch();
}

} // namespace mio
