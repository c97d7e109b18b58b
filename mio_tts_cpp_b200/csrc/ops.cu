// Coalesced, vectorised warp-shuffle kernels for the HBM-bound pieces of the MioCodec decode path:
// codebook gather, AdaLN conditioning vectors, LayerNorm(+affine / AdaLN), GroupNorm (two-phase) + SiLU,
// bilinear frame interpolation, RoPE table, transposed-conv fold + SnakeBeta (wave-upsampler branch).
// Reference semantics are cited per kernel (paths relative to /root/reference).
#include "common.h"

#include <cuda_bf16.h>
#include <cuda_fp16.h>

namespace mio {
namespace {

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

__device__ __forceinline__ uint32_t pack16(float a, float b, int dtype) {
    if (dtype == OP_BF16) {
        __nv_bfloat162 v = __floats2bfloat162_rn(a, b);
        return *reinterpret_cast<uint32_t *>(&v);
    }
    return pack_f16x2_sat(a, b);
}

__device__ __forceinline__ float silu_exact(float x) { return x / (1.0f + expf(-x)); }

// ---- A1: ggml_get_rows(token_embd, codes), src/miocodec-decoder.cpp:1360.  Pure copy: bit-exact. ----------
__global__ void gather_kernel(const float * __restrict__ table, const int * __restrict__ codes,
                              float * __restrict__ out, int n_rows, int dim) {
    MIO_PDL_PROLOGUE();
    const int warps_per_block = blockDim.x >> 5;
    const int row = blockIdx.x * warps_per_block + (threadIdx.x >> 5);
    if (row >= n_rows) return;
    const int lane = threadIdx.x & 31;
    const float * src = table + (size_t) codes[row] * dim;
    float * dst = out + (size_t) row * dim;
    if ((dim & 3) == 0) {
        const float4 * s4 = reinterpret_cast<const float4 *>(src);
        float4 * d4 = reinterpret_cast<float4 *>(dst);
        for (int i = lane; i < (dim >> 2); i += 32) d4[i] = __ldg(s4 + i);
    } else {
        for (int i = lane; i < dim; i += 32) dst[i] = __ldg(src + i);
    }
}

// ---- AdaLN conditioning: p = W * silu(emb) + b for every conditioned norm at once
//      (src/miocodec-decoder.cpp:1164-1181 cond = silu(embedding); :1204-1210, :1271-1277, :1390-1400) ----
// Block = 64 utterances x 64 output rows staged in shared memory (silu(emb) once per block); a thread owns a 4 x 4
// (utterance, row) tile and walks k four at a time with 16-byte loads: 8 loads per 64 FMAs instead of 5 per 4.  Every
// accumulator still sums k = 0 .. adim-1 in order, like the reference's dot product.
constexpr int COND_UTT = 64, COND_ROWS = 64;
__global__ void __launch_bounds__(256) cond_kernel(const float * __restrict__ w, const float * __restrict__ b,
                                                   const float * __restrict__ emb, float * __restrict__ out, int n_out,
                                                   int adim, int n_utt) {
    MIO_PDL_PROLOGUE();
    extern __shared__ __align__(16) float cond_smem[];
    const int ld = adim + 4;                                 // +4 floats: 16-byte rows that do not all start in bank 0
    float * es = cond_smem;                                  // [COND_UTT][ld]
    float * ws = cond_smem + COND_UTT * ld;                  // [COND_ROWS][ld]
    const int utt0 = blockIdx.y * COND_UTT, j0 = blockIdx.x * COND_ROWS;
    // staging: 16-byte loads, several in flight per thread (one scalar load per iteration left the block latency-bound)
    const int a4 = adim >> 2;
    const float4 zero4 = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
#pragma unroll 4
    for (int i = threadIdx.x; i < COND_UTT * a4; i += blockDim.x) {
        const int u = i / a4, k4 = i - u * a4;
        float4 v = utt0 + u < n_utt ? __ldg(reinterpret_cast<const float4 *>(emb + (size_t) (utt0 + u) * adim) + k4) : zero4;
        if (utt0 + u < n_utt) { v.x = silu_exact(v.x); v.y = silu_exact(v.y); v.z = silu_exact(v.z); v.w = silu_exact(v.w); }
        *reinterpret_cast<float4 *>(es + u * ld + 4 * k4) = v;
    }
#pragma unroll 4
    for (int i = threadIdx.x; i < COND_ROWS * a4; i += blockDim.x) {
        const int r = i / a4, k4 = i - r * a4;
        *reinterpret_cast<float4 *>(ws + r * ld + 4 * k4) =
            j0 + r < n_out ? __ldg(reinterpret_cast<const float4 *>(w + (size_t) (j0 + r) * adim) + k4) : zero4;
    }
    __syncthreads();
    const int ug = threadIdx.x & 15, rg = threadIdx.x >> 4;  // utterances ug + 16*u (conflict-free 16-byte reads: ld/4 is odd), rows 4*rg .. +3
    float acc[4][4];
#pragma unroll
    for (int u = 0; u < 4; ++u)
#pragma unroll
        for (int r = 0; r < 4; ++r) acc[u][r] = 0.0f;
    for (int k = 0; k < adim; k += 4) {
        float4 e4[4], w4[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) e4[u] = *reinterpret_cast<const float4 *>(es + (ug + 16 * u) * ld + k);
#pragma unroll
        for (int r = 0; r < 4; ++r) w4[r] = *reinterpret_cast<const float4 *>(ws + (4 * rg + r) * ld + k);
#pragma unroll
        for (int u = 0; u < 4; ++u)
#pragma unroll
            for (int r = 0; r < 4; ++r) {
                acc[u][r] = fmaf(w4[r].x, e4[u].x, acc[u][r]);
                acc[u][r] = fmaf(w4[r].y, e4[u].y, acc[u][r]);
                acc[u][r] = fmaf(w4[r].z, e4[u].z, acc[u][r]);
                acc[u][r] = fmaf(w4[r].w, e4[u].w, acc[u][r]);
            }
    }
    // stage the 64 x 64 result tile (over the embeddings, which are dead now) and write 256-byte row segments: the
    // per-thread float4s would land 100 KB apart
    __syncthreads();
    float * os = cond_smem;                                  // [COND_UTT][COND_ROWS + 4]
    constexpr int OLD = COND_ROWS + 4;
#pragma unroll
    for (int u = 0; u < 4; ++u)
        *reinterpret_cast<float4 *>(os + (ug + 16 * u) * OLD + 4 * rg) = make_float4(acc[u][0], acc[u][1], acc[u][2], acc[u][3]);
    __syncthreads();
    for (int i = threadIdx.x; i < COND_UTT * (COND_ROWS / 4); i += blockDim.x) {
        const int u = i / (COND_ROWS / 4), c4 = i - u * (COND_ROWS / 4);
        const int utt = utt0 + u, j = j0 + 4 * c4;
        if (utt >= n_utt || j >= n_out) continue;
        const float4 v = *reinterpret_cast<const float4 *>(os + u * OLD + 4 * c4);
        float * o = out + (size_t) utt * n_out + j;
        if (j + 4 <= n_out && (n_out & 3) == 0) {
            *reinterpret_cast<float4 *>(o) = make_float4(v.x + b[j], v.y + b[j + 1], v.z + b[j + 2], v.w + b[j + 3]);
        } else {
            const float vv[4] = {v.x, v.y, v.z, v.w};
            for (int r = 0; r < 4; ++r) if (j + r < n_out) o[r] = vv[r] + b[j + r];
        }
    }
}

// ---- LayerNorm (ggml_norm, biased variance, ggml-cpu/ops.cpp:3642-3687) + affine (layer_norm(),
//      src/miocodec-decoder.cpp:136-150) or AdaLN xn*(1+scale)+shift (adaln_apply :1183-1188).
//      One warp per row, row kept in registers (dim = 128*NV), fp32 statistics, 16-bit or fp32 output. ----
template <int NV, int RW>
__global__ void __launch_bounds__(256) layernorm_kernel(const float * __restrict__ x, int rows, float eps, const float * __restrict__ w,
                                                        const float * __restrict__ b, const float * __restrict__ cond, int cond_ld,
                                                        int shift_off, int scale_off, const int * __restrict__ row_utt, void * out16,
                                                        int out_dtype, float * out32, int rev) {
    MIO_PDL_PROLOGUE();
    constexpr int DIM = NV * 128;
    // RW rows per warp, all their loads in flight together; the affine / AdaLN parameters of the first row's utterance
    // are fetched before the reductions (the other rows almost always share them) so nothing waits on them afterwards
    const int row0 = (MIO_BLOCK_X(rev) * (blockDim.x >> 5) + (threadIdx.x >> 5)) * RW;
    if (row0 >= rows) return;
    const int lane = threadIdx.x & 31;
    float4 v[RW][NV];
    int utt[RW];
#pragma unroll
    for (int r = 0; r < RW; ++r) {
        const int row = min(row0 + r, rows - 1);
        utt[r] = row_utt ? row_utt[row] : 0;
        const float4 * xr = reinterpret_cast<const float4 *>(x + (size_t) row * DIM);
#pragma unroll
        for (int i = 0; i < NV; ++i) v[r][i] = xr[lane + 32 * i];
    }
    // hoisting costs 8*NV registers: wider rows fetch each parameter where it is used instead (occupancy matters more there)
    constexpr bool HOIST = NV <= 4;
    float4 pa[HOIST ? NV : 1], pb[HOIST ? NV : 1];      // (scale, shift) or (weight, bias)
    auto param_ptrs = [&](int u, const float4 *& a4, const float4 *& b4) {
        a4 = cond ? reinterpret_cast<const float4 *>(cond + (size_t) max(u, 0) * cond_ld + scale_off) : reinterpret_cast<const float4 *>(w);
        b4 = cond ? reinterpret_cast<const float4 *>(cond + (size_t) max(u, 0) * cond_ld + shift_off) : reinterpret_cast<const float4 *>(b);
    };
    auto load_params = [&](int u) {
        if constexpr (HOIST) {
            const float4 * a4, * b4;
            param_ptrs(u, a4, b4);
#pragma unroll
            for (int i = 0; i < NV; ++i) { pa[i] = __ldg(a4 + lane + 32 * i); pb[i] = __ldg(b4 + lane + 32 * i); }
        }
    };
    const bool affine = cond != nullptr || w != nullptr;
    if (HOIST && affine) load_params(utt[0]);
    float rstd[RW];
#pragma unroll
    for (int r = 0; r < RW; ++r) {
        float s = 0.0f;
#pragma unroll
        for (int i = 0; i < NV; ++i) s += (v[r][i].x + v[r][i].y) + (v[r][i].z + v[r][i].w);
        const float mean = warp_sum(s) * (1.0f / DIM);
        float q = 0.0f;
#pragma unroll
        for (int i = 0; i < NV; ++i) {
            v[r][i].x -= mean; v[r][i].y -= mean; v[r][i].z -= mean; v[r][i].w -= mean;
            q += (v[r][i].x * v[r][i].x + v[r][i].y * v[r][i].y) + (v[r][i].z * v[r][i].z + v[r][i].w * v[r][i].w);
        }
        const float var = warp_sum(q) * (1.0f / DIM);
        rstd[r] = 1.0f / sqrtf(var + eps);
    }
#pragma unroll
    for (int r = 0; r < RW; ++r) {
        const int row = row0 + r;
        if (row >= rows) break;
        if (HOIST && r > 0 && cond && utt[r] != utt[r - 1]) load_params(utt[r]);
        const float4 * a4 = nullptr, * b4 = nullptr;
        if (!HOIST && affine) param_ptrs(utt[r], a4, b4);
#pragma unroll
        for (int i = 0; i < NV; ++i) {
            const int c4 = lane + 32 * i;
            float4 y = make_float4(v[r][i].x * rstd[r], v[r][i].y * rstd[r], v[r][i].z * rstd[r], v[r][i].w * rstd[r]);
            float4 A = make_float4(0.f, 0.f, 0.f, 0.f), B = A;
            if (affine) {
                if constexpr (HOIST) { A = pa[i]; B = pb[i]; }
                else { A = __ldg(a4 + c4); B = __ldg(b4 + c4); }
            }
            if (cond) {
                y.x = y.x * (A.x + 1.0f) + B.x; y.y = y.y * (A.y + 1.0f) + B.y;
                y.z = y.z * (A.z + 1.0f) + B.z; y.w = y.w * (A.w + 1.0f) + B.w;
            } else if (w) {
                y.x = y.x * A.x + B.x; y.y = y.y * A.y + B.y;
                y.z = y.z * A.z + B.z; y.w = y.w * A.w + B.w;
            }
            if (utt[r] < 0) y = make_float4(0.f, 0.f, 0.f, 0.f);   // gap row of the layout
            if (out32) reinterpret_cast<float4 *>(out32 + (size_t) row * DIM)[c4] = y;
            if (out16) {
                uint2 o;
                o.x = pack16(y.x, y.y, out_dtype);
                o.y = pack16(y.z, y.w, out_dtype);
                reinterpret_cast<uint2 *>((uint16_t *) out16 + (size_t) row * DIM)[c4] = o;
            }
        }
    }
}

// generic-width fallback (dim % 128 != 0 or dim > 1024): one warp per row, three passes over the row
__global__ void layernorm_generic_kernel(const float * __restrict__ x, int rows, int dim, float eps,
                                         const float * __restrict__ w, const float * __restrict__ b,
                                         const float * __restrict__ cond, int cond_ld, int shift_off, int scale_off,
                                         const int * __restrict__ row_utt, void * out16, int out_dtype,
                                         float * out32, int rev) {
    MIO_PDL_PROLOGUE();
    const int row = MIO_BLOCK_X(rev) * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (row >= rows) return;
    const int lane = threadIdx.x & 31;
    const int utt = row_utt ? row_utt[row] : 0;
    const float * xr = x + (size_t) row * dim;
    float s = 0.0f;
    for (int i = lane; i < dim; i += 32) s += xr[i];
    const float mean = warp_sum(s) / (float) dim;
    float q = 0.0f;
    for (int i = lane; i < dim; i += 32) { const float d = xr[i] - mean; q += d * d; }
    const float rstd = 1.0f / sqrtf(warp_sum(q) / (float) dim + eps);
    const float * shift = cond ? cond + (size_t) max(utt, 0) * cond_ld + shift_off : nullptr;
    const float * scale = cond ? cond + (size_t) max(utt, 0) * cond_ld + scale_off : nullptr;
    for (int i = lane; i < dim; i += 32) {
        float y = (xr[i] - mean) * rstd;
        if (cond) y = y * (scale[i] + 1.0f) + shift[i];
        else if (w) y = y * w[i] + b[i];
        if (utt < 0) y = 0.0f;
        if (out32) out32[(size_t) row * dim + i] = y;
        if (out16) ((uint16_t *) out16)[(size_t) row * dim + i] = (uint16_t) (pack16(y, 0.f, out_dtype) & 0xFFFF);
    }
}

// Both casts are persistent (one resident wave) with four 16-byte loads in flight per thread: with one float4 per thread and
// a block per 4 KB the 460 800-row stages of the wave upsampler ran at 1.6-2.4 TB/s.
constexpr int CAST_UNROLL = 4;
__global__ void __launch_bounds__(256) cast16_kernel(const float * __restrict__ x, void * out16, int out_dtype, size_t n4, int rev) {
    MIO_PDL_PROLOGUE();
    const size_t per_block = (size_t) blockDim.x * CAST_UNROLL;
    const size_t n_chunks = (n4 + per_block - 1) / per_block;
    for (size_t c = blockIdx.x; c < n_chunks; c += gridDim.x) {
        const size_t base = (rev ? n_chunks - 1 - c : c) * per_block + threadIdx.x;
        float4 v[CAST_UNROLL];
#pragma unroll
        for (int u = 0; u < CAST_UNROLL; ++u) {
            const size_t i = base + (size_t) u * blockDim.x;
            if (i < n4) v[u] = reinterpret_cast<const float4 *>(x)[i];
        }
#pragma unroll
        for (int u = 0; u < CAST_UNROLL; ++u) {
            const size_t i = base + (size_t) u * blockDim.x;
            if (i < n4) reinterpret_cast<uint2 *>(out16)[i] = make_uint2(pack16(v[u].x, v[u].y, out_dtype), pack16(v[u].z, v[u].w, out_dtype));
        }
    }
}

// The same cast for an operand whose NEIGHBOUR rows a multi-tap GEMM reads (polyphase transposed convolution): rows
// that carry no data (row_utt < 0: the gaps between utterances, which in-place residual updates leave dirty) become 0.
// A block has (rows per pass) x c4 threads (c4 = float4 slots per row), so a thread keeps its column and only its row advances.
__global__ void cast16_rows_kernel(const float * __restrict__ x, void * out16, int out_dtype, int rows, int c4,
                                                          const int * __restrict__ row_utt, int rev) {
    MIO_PDL_PROLOGUE();
    const int rpp = blockDim.x / c4;                       // rows per pass
    const int col = threadIdx.x % c4, r_in = threadIdx.x / c4;
    const int rows_per_chunk = rpp * CAST_UNROLL;
    const int n_chunks = (rows + rows_per_chunk - 1) / rows_per_chunk;
    for (int c = blockIdx.x; c < n_chunks; c += gridDim.x) {
        const int row0 = (rev ? n_chunks - 1 - c : c) * rows_per_chunk + r_in;
        float4 v[CAST_UNROLL];
        bool live[CAST_UNROLL];
#pragma unroll
        for (int u = 0; u < CAST_UNROLL; ++u) {
            const int row = row0 + u * rpp;
            live[u] = row < rows && __ldg(row_utt + row) >= 0;
            if (live[u]) v[u] = reinterpret_cast<const float4 *>(x)[(size_t) row * c4 + col];
        }
#pragma unroll
        for (int u = 0; u < CAST_UNROLL; ++u) {
            const int row = row0 + u * rpp;
            if (row >= rows) continue;
            uint2 o = make_uint2(0u, 0u);
            if (live[u]) { o.x = pack16(v[u].x, v[u].y, out_dtype); o.y = pack16(v[u].z, v[u].w, out_dtype); }
            reinterpret_cast<uint2 *>(out16)[(size_t) row * c4 + col] = o;
        }
    }
}

// ---- GroupNorm phase 1: per (utterance, group) sum and sum of squares over (channels-in-group x ALL frames)
//      (ggml_group_norm, ggml-cpu/ops.cpp:3955-4028: double accumulators).  One block per <=64-row tile. ----
template <int VEC>
__global__ void gn_stats_kernel(const float * __restrict__ x, int dim, int groups, const row_tile * __restrict__ tiles,
                                double * __restrict__ stats, int rev) {
    MIO_PDL_PROLOGUE();
    extern __shared__ double s_acc[];   // [groups][2]
    const row_tile t = tiles[MIO_BLOCK_X(rev)];
    for (int i = threadIdx.x; i < groups * 2; i += blockDim.x) s_acc[i] = 0.0;
    __syncthreads();
    const int cpg = dim / groups;
    const int cols = dim / VEC;                       // column slots per row
    const int rows_per_pass = blockDim.x / cols > 0 ? blockDim.x / cols : 1;
    const int col = threadIdx.x % cols;
    const int ry = threadIdx.x / cols;
    if (ry < rows_per_pass) {
        float s = 0.0f, q = 0.0f;
        for (int r = ry; r < t.n; r += rows_per_pass) {
            const float * p = x + (size_t) (t.row0 + r) * dim + col * VEC;
            if constexpr (VEC == 4) {
                const float4 v = *reinterpret_cast<const float4 *>(p);
                s += (v.x + v.y) + (v.z + v.w);
                q += (v.x * v.x + v.y * v.y) + (v.z * v.z + v.w * v.w);
            } else {
                const float v = *p;
                s += v; q += v * v;
            }
        }
        const int g = (col * VEC) / cpg;
        atomicAdd(&s_acc[2 * g], (double) s);
        atomicAdd(&s_acc[2 * g + 1], (double) q);
    }
    __syncthreads();
    for (int i = threadIdx.x; i < groups * 2; i += blockDim.x)
        atomicAdd(&stats[(size_t) t.utt * groups * 2 + i], s_acc[i]);
}

// (sum, sumsq) of one (utterance, group) -> (mean, rstd); the arithmetic of ggml_group_norm's tail in double
__device__ __forceinline__ float2 gn_mean_rstd(const double * __restrict__ stats, const int * __restrict__ utt_len, int utt,
                                               int groups, int g, int cpg, float eps) {
    const double n = (double) __ldg(utt_len + utt) * (double) cpg;
    const double * sp = stats + ((size_t) utt * groups + g) * 2;
    const double mean = sp[0] / n;
    double var = sp[1] / n - mean * mean;
    if (var < 0.0) var = 0.0;
    return make_float2((float) mean, 1.0f / sqrtf((float) var + eps));
}

// ---- GroupNorm phase 2: normalise, affine, SiLU, round to the conv operand type (fp16, as the reference's
//      im2col does: llama.cpp/ggml/src/ggml.c:4409-4426), zero the layout's gap rows. ----
__global__ void gn_apply_kernel(const float * __restrict__ x, int rows, int dim, int groups,
                                const double * __restrict__ stats, const int * __restrict__ utt_len, float eps,
                                const float * __restrict__ w,
                                const float * __restrict__ b, const int * __restrict__ row_utt, void * out16,
                                int out_dtype, int rev) {
    MIO_PDL_PROLOGUE();
    const int c4n = dim >> 2;
    const size_t i = (size_t) MIO_BLOCK_X(rev) * blockDim.x + threadIdx.x;
    if (i >= (size_t) rows * c4n) return;
    const int row = (int) (i / c4n);
    const int c4 = (int) (i % c4n);
    const int utt = row_utt[row];
    uint2 o = make_uint2(0u, 0u);
    if (utt >= 0) {
        const int cpg = dim / groups;
        const float4 v = reinterpret_cast<const float4 *>(x)[i];
        const float4 ww = __ldg(reinterpret_cast<const float4 *>(w) + c4);
        const float4 bb = __ldg(reinterpret_cast<const float4 *>(b) + c4);
        float y[4] = {v.x, v.y, v.z, v.w};
        const float wv[4] = {ww.x, ww.y, ww.z, ww.w};
        const float bv[4] = {bb.x, bb.y, bb.z, bb.w};
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const float2 m = gn_mean_rstd(stats, utt_len, utt, groups, (c4 * 4 + k) / cpg, cpg, eps);
            const float t = ((y[k] - m.x) * m.y) * wv[k] + bv[k];
            y[k] = silu_exact(t);
        }
        o.x = pack16(y[0], y[1], out_dtype);
        o.y = pack16(y[2], y[3], out_dtype);
    }
    reinterpret_cast<uint2 *>(out16)[i] = o;
}

// fast path: channels-per-group % 4 == 0 (one statistics lookup per float4) and 256 % (dim/4) == 0, so a thread keeps
// its channel quad (affine weights in registers) and walks rows; SiLU on the SFU (output is rounded to fp16 anyway)
// Persistent form: a block owns a CONTIGUOUS run of 8-row chunks (so a thread's rows almost always share one utterance
// and the double-precision (mean, rstd) tail is derived once per thread, not once per chunk) and the next chunk's loads
// are in flight while the current one is normalised.
template <int ROWS_PER_THREAD>
__global__ void __launch_bounds__(256, 4) gn_apply_fast_kernel(const float * __restrict__ x, int rows, int dim, int groups,
                                     const double * __restrict__ stats, const int * __restrict__ utt_len, float eps,
                                     const float * __restrict__ w,
                                     const float * __restrict__ b, const int * __restrict__ row_utt, void * out16,
                                     int out_dtype, int rev) {
    MIO_PDL_PROLOGUE();
    const int c4n = dim >> 2;
    const int rpb = blockDim.x / c4n;                 // rows per block pass
    const int c4 = threadIdx.x % c4n;
    const int g = (c4 * 4) / (dim / groups);
    const float4 ww = __ldg(reinterpret_cast<const float4 *>(w) + c4);
    const float4 bb = __ldg(reinterpret_cast<const float4 *>(b) + c4);
    const int rows_per_chunk = rpb * ROWS_PER_THREAD;
    const int n_chunks = (rows + rows_per_chunk - 1) / rows_per_chunk;
    const int blk = MIO_BLOCK_X(rev);
    const int c_lo = (int) ((long long) blk * n_chunks / gridDim.x), c_hi = (int) ((long long) (blk + 1) * n_chunks / gridDim.x);
    float4 v[ROWS_PER_THREAD], vn[ROWS_PER_THREAD];
    int utt[ROWS_PER_THREAD], uttn[ROWS_PER_THREAD];
    auto load = [&](int chunk, float4 (&vv)[ROWS_PER_THREAD], int (&uu)[ROWS_PER_THREAD]) {
        const int row_base = chunk * rows_per_chunk + threadIdx.x / c4n;
#pragma unroll
        for (int it = 0; it < ROWS_PER_THREAD; ++it) {
            const int row = row_base + it * rpb;
            uu[it] = row < rows ? __ldg(row_utt + row) : -1;
            if (uu[it] >= 0) vv[it] = reinterpret_cast<const float4 *>(x)[(size_t) row * c4n + c4];
        }
    };
    int m_utt = -1;
    float2 m = make_float2(0.f, 1.f);
    if (c_lo < c_hi) load(rev ? c_hi - 1 : c_lo, v, utt);
    for (int i = 0; i < c_hi - c_lo; ++i) {
        const int chunk = rev ? c_hi - 1 - i : c_lo + i;
        if (i + 1 < c_hi - c_lo) load(rev ? chunk - 1 : chunk + 1, vn, uttn);
        const int row_base = chunk * rows_per_chunk + threadIdx.x / c4n;
#pragma unroll
        for (int it = 0; it < ROWS_PER_THREAD; ++it) {
            const int row = row_base + it * rpb;
            if (row >= rows) continue;
            uint2 o = make_uint2(0u, 0u);
            if (utt[it] >= 0) {
                if (utt[it] != m_utt) { m_utt = utt[it]; m = gn_mean_rstd(stats, utt_len, m_utt, groups, g, dim / groups, eps); }
                float y[4] = {v[it].x, v[it].y, v[it].z, v[it].w};
                const float wv[4] = {ww.x, ww.y, ww.z, ww.w};
                const float bv[4] = {bb.x, bb.y, bb.z, bb.w};
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    const float t = ((y[k] - m.x) * m.y) * wv[k] + bv[k];
                    y[k] = __fdividef(t, 1.0f + __expf(-t));
                }
                o.x = pack16(y[0], y[1], out_dtype);
                o.y = pack16(y[2], y[3], out_dtype);
            }
            reinterpret_cast<uint2 *>(out16)[(size_t) row * c4n + c4] = o;
        }
#pragma unroll
        for (int it = 0; it < ROWS_PER_THREAD; ++it) { v[it] = vn[it]; utt[it] = uttn[it]; }
    }
}

// ---- ggml_interpolate(BILINEAR) along frames only: half-pixel centres, clamped taps
//      (ggml-cpu/ops.cpp:7577-7617, call src/miocodec-decoder.cpp:1373); identity when lengths match. ----
__global__ void interp_kernel(const float * __restrict__ src, const int * __restrict__ src_row0,
                              const int * __restrict__ src_len, float * __restrict__ dst, int rows, int dim,
                              const int * __restrict__ row_utt, const int * __restrict__ row_pos,
                              const int * __restrict__ utt_len) {
    MIO_PDL_PROLOGUE();
    const int c4n = dim >> 2;
    const size_t i = (size_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (size_t) rows * c4n) return;
    const int row = (int) (i / c4n);
    const int c4 = (int) (i % c4n);
    const int utt = row_utt[row];
    float4 o = make_float4(0.f, 0.f, 0.f, 0.f);
    if (utt >= 0) {
        const int L = src_len[utt], n_out = utt_len[utt], p = row_pos[row];
        const float4 * s4 = reinterpret_cast<const float4 *>(src);
        if (L == n_out) {
            o = s4[(size_t) (src_row0[utt] + p) * c4n + c4];
        } else {
            const float sf = (float) n_out / (float) L;
            const float xs = ((float) p + 0.5f) / sf - 0.5f;
            int x0 = (int) floorf(xs);
            int x1 = x0 + 1;
            x0 = min(max(x0, 0), L - 1);
            x1 = min(max(x1, 0), L - 1);
            float dx = xs - (float) x0;
            dx = fminf(fmaxf(dx, 0.0f), 1.0f);
            const float4 a = s4[(size_t) (src_row0[utt] + x0) * c4n + c4];
            const float4 c = s4[(size_t) (src_row0[utt] + x1) * c4n + c4];
            o.x = a.x * (1.0f - dx) + c.x * dx; o.y = a.y * (1.0f - dx) + c.y * dx;
            o.z = a.z * (1.0f - dx) + c.z * dx; o.w = a.w * (1.0f - dx) + c.w * dx;
        }
    }
    reinterpret_cast<float4 *>(dst)[i] = o;
}

// ---- RoPE cache: theta_i = pos * s^i by repeated fp32 multiplication (ggml-cpu/ops.cpp:5616-5625) ----
__global__ void rope_table_kernel(float2 * __restrict__ table, int n_pos, int half_dim, float theta_scale) {
    MIO_PDL_PROLOGUE();
    const int pos = blockIdx.x * blockDim.x + threadIdx.x;
    if (pos >= n_pos) return;
    float theta = (float) pos;
    for (int i = 0; i < half_dim; ++i) {
        table[(size_t) pos * half_dim + i] = make_float2(cosf(theta), sinf(theta));
        theta *= theta_scale;
    }
}

// ---- wave-upsampler stage: fold (col2im) of ggml_conv_transpose_1d computed as a [rows, k*Cout] GEMM,
//      + bias, crop pad=(k-s)/2 per side, SnakeBeta (src/miocodec-decoder.cpp:1410-1442, :1332-1343). ----
__global__ void convt_fold_snake_kernel(const float * __restrict__ y, int ld_y, const int * __restrict__ src_row0,
                                        const int * __restrict__ src_len, float * __restrict__ dst, int rows, int cout,
                                        int ksize, int stride, int pad, const float * __restrict__ bias,
                                        const float * __restrict__ snake_a, const float * __restrict__ snake_inv_b,
                                        const int * __restrict__ row_utt, const int * __restrict__ row_pos) {
    MIO_PDL_PROLOGUE();
    const size_t i = (size_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (size_t) rows * cout) return;
    const int row = (int) (i / cout);
    const int co = (int) (i % cout);
    const int utt = row_utt[row];
    float o = 0.0f;
    if (utt >= 0) {
        const int t = row_pos[row] + pad;
        const int L = src_len[utt];
        float acc = 0.0f;
        // out[t] = sum over (l, kk) with l*stride + kk == t; ascending l like the reference's accumulation
        const int l_hi = min(L - 1, t / stride);
        int l_lo = (t - ksize + stride) / stride;   // ceil((t - ksize + 1) / stride) for t-ksize+1 >= 0
        if (t - ksize + 1 <= 0) l_lo = 0;
        for (int l = l_lo; l <= l_hi; ++l) {
            const int kk = t - l * stride;
            if (kk >= 0 && kk < ksize) acc += y[(size_t) (src_row0[utt] + l) * ld_y + (size_t) kk * cout + co];
        }
        acc += bias[co];
        const float s = sinf(acc * snake_a[co]);
        o = acc + (s * s) * snake_inv_b[co];
    }
    dst[i] = o;
}

__global__ void snake_kernel(float * __restrict__ x, int rows, int dim, const float * __restrict__ snake_a,
                             const float * __restrict__ snake_inv_b) {
    MIO_PDL_PROLOGUE();
    const size_t i = (size_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (size_t) rows * dim) return;
    const int c = (int) (i % dim);
    const float v = x[i];
    const float s = sinf(v * snake_a[c]);
    x[i] = v + (s * s) * snake_inv_b[c];
}

inline unsigned blocks_for(size_t n, int threads) { return (unsigned) ((n + threads - 1) / threads); }

} // namespace

void launch_gather(const float * table, const int * codes, float * out, int n_rows, int dim, cudaStream_t st) {
    if (n_rows <= 0) return;
    launch_pdl(gather_kernel, dim3(blocks_for((size_t) n_rows, 8)), dim3(256), 0, st, table, codes, out, n_rows, dim);
    MIO_CUDA(cudaGetLastError());
    count_launch();
}

void launch_cond(const float * w, const float * b, const float * emb, float * out, int n_out, int adim, int n_utt,
                 cudaStream_t st) {
    if (n_out <= 0 || n_utt <= 0) return;
    dim3 grid(blocks_for((size_t) n_out, COND_ROWS), (unsigned) ((n_utt + COND_UTT - 1) / COND_UTT));
    if (adim & 3) throw cuda_error("cond: adanorm dimension must be a multiple of 4");
    const size_t smem = (size_t) (COND_UTT + COND_ROWS) * (adim + 4) * sizeof(float);
    if (smem > 200 * 1024) throw cuda_error("cond: adanorm dimension too large");
    static bool configured[64] = {false};
    int dev = 0;
    MIO_CUDA(cudaGetDevice(&dev));
    if (smem > 48 * 1024 && !configured[dev & 63]) {
        MIO_CUDA(cudaFuncSetAttribute(cond_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
        configured[dev & 63] = true;
    }
    launch_pdl(cond_kernel, dim3(grid), dim3(256), smem, st, w, b, emb, out, n_out, adim, n_utt);
    MIO_CUDA(cudaGetLastError());
    count_launch();
}

void launch_layernorm(const float * x, int rows, int dim, float eps, const float * w, const float * b,
                      const float * cond, int cond_ld, int shift_off, int scale_off, const int * row_utt,
                      void * out16, int out_dtype, float * out32, cudaStream_t st) {
    if (rows <= 0) return;
    const unsigned grid = blocks_for((size_t) rows, 8);
    const unsigned grid2 = blocks_for((size_t) rows, 16);      // two rows per warp
    // two rows per warp up to 512 columns; wider rows already keep enough bytes in flight and two of them cost occupancy
    // (measured: 768 columns 21.4 us with one row per warp, 24.7 us with two)
#define LN_CASE(NV)                                                                                              \
    case NV:                                                                                                     \
        if (NV <= 4)                                                                                             \
            launch_pdl(layernorm_kernel<NV, 2>, dim3(grid2), dim3(256), 0, st, x, rows, eps, w, b, cond, cond_ld, shift_off,         \
                       scale_off, row_utt, out16, out_dtype, out32, traversal_reverse());                                            \
        else                                                                                                     \
            launch_pdl(layernorm_kernel<NV, 1>, dim3(grid), dim3(256), 0, st, x, rows, eps, w, b, cond, cond_ld, shift_off,          \
                       scale_off, row_utt, out16, out_dtype, out32, traversal_reverse());                                            \
        break;
    if ((dim % 128) == 0 && dim <= 1024 && (!cond || ((cond_ld | shift_off | scale_off) & 3) == 0)) {
        switch (dim / 128) {
            LN_CASE(1) LN_CASE(2) LN_CASE(3) LN_CASE(4) LN_CASE(5) LN_CASE(6) LN_CASE(7) LN_CASE(8)
        }
    } else {
        launch_pdl(layernorm_generic_kernel, dim3(grid), dim3(256), 0, st, x, rows, dim, eps, w, b, cond, cond_ld, shift_off, scale_off,
                                                       row_utt, out16, out_dtype, out32, traversal_reverse());
    }
#undef LN_CASE
    MIO_CUDA(cudaGetLastError());
    count_launch();
}

static unsigned one_wave_blocks(int blocks_per_sm) {
    static int n_sm[64] = {0};
    int dev = 0;
    MIO_CUDA(cudaGetDevice(&dev));
    if (n_sm[dev & 63] == 0) MIO_CUDA(cudaDeviceGetAttribute(&n_sm[dev & 63], cudaDevAttrMultiProcessorCount, dev));
    return (unsigned) (n_sm[dev & 63] * blocks_per_sm);
}

void launch_cast16(const float * x, void * out16, int out_dtype, size_t n, cudaStream_t st) {
    if (n == 0) return;
    if (n & 3) throw cuda_error("cast16: element count must be a multiple of 4");
    const unsigned chunks = blocks_for(n >> 2, 256 * CAST_UNROLL);
    launch_pdl(cast16_kernel, dim3(std::min(chunks, one_wave_blocks(8))), dim3(256), 0, st, x, out16, out_dtype, n >> 2, traversal_reverse());
    MIO_CUDA(cudaGetLastError());
    count_launch();
}

void launch_cast16_rows(const float * x, void * out16, int out_dtype, int rows, int cols, const int * row_utt, cudaStream_t st) {
    if (rows <= 0) return;
    const int c4 = cols >> 2;
    if ((cols & 3) || c4 > 1024) throw cuda_error("cast16: the column count must be a multiple of 4, at most 4096");
    const int rpp = c4 >= 256 ? 1 : 256 / c4;                  // whole rows per block pass: the block has rpp * c4 <= 1024 threads
    const unsigned chunks = blocks_for((size_t) rows, rpp * CAST_UNROLL);
    launch_pdl(cast16_rows_kernel, dim3(std::min(chunks, one_wave_blocks(c4 > 256 ? 2 : 8))), dim3(rpp * c4), 0, st, x, out16, out_dtype, rows,
               c4, row_utt, traversal_reverse());
    MIO_CUDA(cudaGetLastError());
    count_launch();
}

void launch_gn_stats(const float * x, int dim, int groups, const row_tile * tiles, int n_tiles, double * stats,
                     cudaStream_t st) {
    if (n_tiles <= 0) return;
    const int cpg = dim / groups;
    const size_t smem = (size_t) groups * 2 * sizeof(double);
    if ((cpg & 3) == 0) {
        const int cols = dim / 4;
        const int threads = cols >= 256 ? ((cols + 31) / 32) * 32 : 256;
        launch_pdl(gn_stats_kernel<4>, dim3(n_tiles), dim3(threads > 1024 ? 1024 : threads), smem, st, x, dim, groups, tiles, stats, traversal_reverse());
    } else {
        const int threads = dim >= 256 ? ((dim + 31) / 32) * 32 : 256;
        launch_pdl(gn_stats_kernel<1>, dim3(n_tiles), dim3(threads > 1024 ? 1024 : threads), smem, st, x, dim, groups, tiles, stats, traversal_reverse());
    }
    MIO_CUDA(cudaGetLastError());
    count_launch();
}

void launch_gn_apply(const float * x, int rows, int dim, int groups, const double * stats, const int * utt_len, float eps,
                     const float * w, const float * b, const int * row_utt, void * out16, int out_dtype, cudaStream_t st) {
    if (rows <= 0) return;
    if (dim & 3) throw cuda_error("group norm: channel count must be a multiple of 4");
    const int c4n = dim >> 2, cpg = dim / groups;
    if ((cpg & 3) == 0 && c4n <= 256 && (256 % c4n) == 0) {
        constexpr int RPT = 4;
        const int rows_per_block = (256 / c4n) * RPT;
        static int n_sm[64] = {0};
        int dev = 0;
        MIO_CUDA(cudaGetDevice(&dev));
        if (n_sm[dev & 63] == 0) MIO_CUDA(cudaDeviceGetAttribute(&n_sm[dev & 63], cudaDevAttrMultiProcessorCount, dev));
        const unsigned chunks = blocks_for((size_t) rows, rows_per_block);
        const unsigned grid = std::min<unsigned>(chunks, (unsigned) n_sm[dev & 63] * 4u);     // one resident wave (4 blocks per SM)
        launch_pdl(gn_apply_fast_kernel<RPT>, dim3(grid), dim3(256), 0, st, 
            x, rows, dim, groups, stats, utt_len, eps, w, b, row_utt, out16, out_dtype, traversal_reverse());
    } else {
        launch_pdl(gn_apply_kernel, dim3(blocks_for((size_t) rows * (dim >> 2), 256)), dim3(256), 0, st, 
            x, rows, dim, groups, stats, utt_len, eps, w, b, row_utt, out16, out_dtype, traversal_reverse());
    }
    MIO_CUDA(cudaGetLastError());
    count_launch();
}

void launch_interp(const float * src, const int * src_row0, const int * src_len, float * dst, int rows, int dim,
                   const int * row_utt, const int * row_pos, const int * utt_len, cudaStream_t st) {
    if (rows <= 0) return;
    if (dim & 3) throw cuda_error("interp: channel count must be a multiple of 4");
    launch_pdl(interp_kernel, dim3(blocks_for((size_t) rows * (dim >> 2), 256)), dim3(256), 0, st, src, src_row0, src_len, dst, rows, dim,
                                                                               row_utt, row_pos, utt_len);
    MIO_CUDA(cudaGetLastError());
    count_launch();
}

void launch_rope_table(float2 * table, int n_pos, int half_dim, float theta_scale, cudaStream_t st) {
    if (n_pos <= 0) return;
    launch_pdl(rope_table_kernel, dim3(blocks_for((size_t) n_pos, 128)), dim3(128), 0, st, table, n_pos, half_dim, theta_scale);
    MIO_CUDA(cudaGetLastError());
    count_launch();
}

void launch_convt_fold_snake(const float * y, int ld_y, const int * src_row0, const int * src_len, float * dst,
                             int rows, int cout, int ksize, int stride, int pad, const float * bias,
                             const float * snake_a, const float * snake_inv_b, const int * row_utt,
                             const int * row_pos, cudaStream_t st) {
    if (rows <= 0) return;
    launch_pdl(convt_fold_snake_kernel, dim3(blocks_for((size_t) rows * cout, 256)), dim3(256), 0, st, 
        y, ld_y, src_row0, src_len, dst, rows, cout, ksize, stride, pad, bias, snake_a, snake_inv_b, row_utt, row_pos);
    MIO_CUDA(cudaGetLastError());
    count_launch();
}

void launch_snake(float * x, int rows, int dim, const float * snake_a, const float * snake_inv_b, cudaStream_t st) {
    if (rows <= 0) return;
    launch_pdl(snake_kernel, dim3(blocks_for((size_t) rows * dim, 256)), dim3(256), 0, st, x, rows, dim, snake_a, snake_inv_b);
    MIO_CUDA(cudaGetLastError());
    count_launch();
}

} // namespace mio
