// Full (unwindowed) multi-head attention with WavLM's gated relative-position bias.
//
// Reference: softmax(Q K^T / sqrt(hd) + gate[q,h] * rel[bucket(k - q), h]) V with the bias materialised as a dense
// [seq, seq, heads] mask (/root/reference/src/wavlm-extractor.cpp:744-760 raw bias, :791 mask, :817-835 attention).
// Here the bias is never materialised: `rel_d[h][d + max_seq - 1]` holds the bucketed embedding per relative distance
// d = k - q (built on the host with the reference's float formulas, wavlm-extractor.cpp:257-279) and
// `gate[clip][h][q]` comes from wavlm_gate_kernel.  A CTA owns one 64-query tile of one head of one clip and streams
// 64-key tiles through shared memory with an online softmax (flash-attention recurrence); Q K^T and P V run on
// tensor cores (mma.sync m16n8k16, fp32 accumulate).  The 1/sqrt(hd) scale is folded into Wq/bq at load.
#include "refenc.h"

#include <cuda_bf16.h>
#include <cuda_fp16.h>

namespace mio {
namespace {

constexpr int FQ = 64, FK = 64, FHD = 64, FLDS = 72;

template <bool BF16>
__device__ __forceinline__ void mma16816(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
    if constexpr (BF16) {
        asm volatile(
            "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
            : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
            : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
    } else {
        asm volatile(
            "mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
            : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
            : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
    }
}
template <bool BF16> __device__ __forceinline__ uint32_t pack2(float a, float b) {
    if constexpr (BF16) {
        __nv_bfloat162 v = __floats2bfloat162_rn(a, b);
        return *reinterpret_cast<uint32_t *>(&v);
    } else {
        return pack_f16x2_sat(a, b);
    }
}

template <bool BF16>
__global__ void __launch_bounds__(128) attention_relbias_kernel(const uint16_t * __restrict__ qkv, uint16_t * __restrict__ out,
                                                                int dim, const int * __restrict__ row0,
                                                                const int * __restrict__ seq_len,
                                                                const int * __restrict__ gate_off,
                                                                const float * __restrict__ gate,
                                                                const float * __restrict__ rel_d, int max_seq,
                                                                const int2 * __restrict__ work) {
    MIO_PDL_PROLOGUE();
    __shared__ __align__(16) uint16_t Qs[FQ * FLDS];
    __shared__ __align__(16) uint16_t Ks[FK * FLDS];
    __shared__ __align__(16) uint16_t Vs[FK * FLDS];
    __shared__ float rels[FQ + FK];

    const int2 wk = work[blockIdx.x];            // (clip, first query of the tile)
    const int clip = wk.x, q0 = wk.y, head = blockIdx.y;
    const int S = seq_len[clip], r0 = row0[clip];
    const int ld = 3 * dim;
    const float * rel_h = rel_d + (size_t) head * (2 * max_seq - 1) + (max_seq - 1);   // index by d = k - q
    const float * gate_h = gate + (size_t) gate_off[clip] + (size_t) head * S;

    auto cp16 = [](uint16_t * dst, const uint16_t * src, bool valid) {
        const uint32_t d = (uint32_t) __cvta_generic_to_shared(dst);
        const int n = valid ? 16 : 0;
        asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(d), "l"(src), "r"(n) : "memory");
    };
    for (int c = threadIdx.x; c < FQ * 8; c += blockDim.x) {
        const int i = c >> 3, ch = c & 7;
        const bool ok = q0 + i < S;
        cp16(Qs + i * FLDS + ch * 8, qkv + (size_t) (r0 + (ok ? q0 + i : 0)) * ld + head * FHD + ch * 8, ok);
    }
    asm volatile("cp.async.commit_group;" ::: "memory");

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int g = lane >> 2, tq = lane & 3;
    const int qw = warp * 16;                    // first query (in tile) of this warp
    const int qi0 = qw + g, qi1 = qw + g + 8;
    const bool qv0 = q0 + qi0 < S, qv1 = q0 + qi1 < S;
    const float gt0 = qv0 ? gate_h[q0 + qi0] : 0.0f, gt1 = qv1 ? gate_h[q0 + qi1] : 0.0f;

    float o[8][4];
#pragma unroll
    for (int nd = 0; nd < 8; ++nd) o[nd][0] = o[nd][1] = o[nd][2] = o[nd][3] = 0.0f;
    float m0 = -INFINITY, m1 = -INFINITY, l0 = 0.0f, l1 = 0.0f;
    uint32_t qa[4][4];

    for (int k0 = 0; k0 < S; k0 += FK) {
        __syncthreads();                         // previous tile's K/V/rels fully consumed
        for (int c = threadIdx.x; c < FK * 8; c += blockDim.x) {
            const int j = c >> 3, ch = c & 7;
            const bool ok = k0 + j < S;
            const uint16_t * p = qkv + (size_t) (r0 + (ok ? k0 + j : 0)) * ld + head * FHD + ch * 8;
            cp16(Ks + j * FLDS + ch * 8, p + dim, ok);
            cp16(Vs + j * FLDS + ch * 8, p + 2 * dim, ok);
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
        // rel embedding for d = (k0 + kj) - (q0 + qi), d - dmin in [0, FQ + FK - 2]
        const int dmin = k0 - q0 - (FQ - 1);
        for (int i = threadIdx.x; i < FQ + FK - 1; i += blockDim.x) {
            const int d = dmin + i;
            rels[i] = (d > -max_seq && d < max_seq) ? rel_h[d] : 0.0f;
        }
        asm volatile("cp.async.wait_group 0;" ::: "memory");
        __syncthreads();
        if (k0 == 0) {
#pragma unroll
            for (int kk = 0; kk < 4; ++kk) {
                const uint16_t * base = Qs + (qw + g) * FLDS + kk * 16 + 2 * tq;
                qa[kk][0] = *reinterpret_cast<const uint32_t *>(base);
                qa[kk][1] = *reinterpret_cast<const uint32_t *>(base + 8 * FLDS);
                qa[kk][2] = *reinterpret_cast<const uint32_t *>(base + 8);
                qa[kk][3] = *reinterpret_cast<const uint32_t *>(base + 8 * FLDS + 8);
            }
        }
        // ---- S = Q K^T (16 x 64 per warp) + bias ----
        float s[8][4];
        float mx0 = m0, mx1 = m1;
#pragma unroll
        for (int nt = 0; nt < 8; ++nt) {
            s[nt][0] = s[nt][1] = s[nt][2] = s[nt][3] = 0.0f;
            const uint16_t * kb = Ks + (nt * 8 + g) * FLDS + 2 * tq;
#pragma unroll
            for (int kk = 0; kk < 4; ++kk) {
                const uint32_t b0 = *reinterpret_cast<const uint32_t *>(kb + kk * 16);
                const uint32_t b1 = *reinterpret_cast<const uint32_t *>(kb + kk * 16 + 8);
                mma16816<BF16>(s[nt], qa[kk], b0, b1);
            }
#pragma unroll
            for (int e = 0; e < 2; ++e) {
                const int kj = nt * 8 + 2 * tq + e;
                const bool kv = k0 + kj < S;
                // d - dmin = kj - qi + FQ - 1
                s[nt][e] = (qv0 && kv) ? s[nt][e] + gt0 * rels[kj - qi0 + FQ - 1] : -INFINITY;
                s[nt][2 + e] = (qv1 && kv) ? s[nt][2 + e] + gt1 * rels[kj - qi1 + FQ - 1] : -INFINITY;
                mx0 = fmaxf(mx0, s[nt][e]);
                mx1 = fmaxf(mx1, s[nt][2 + e]);
            }
        }
        mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 1));
        mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 2));
        mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 1));
        mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 2));
        const float base0 = mx0 == -INFINITY ? 0.0f : mx0, base1 = mx1 == -INFINITY ? 0.0f : mx1;
        const float c0 = __expf(m0 - base0), c1 = __expf(m1 - base1);   // m = -inf -> 0
        m0 = mx0; m1 = mx1;
        float ps0 = 0.0f, ps1 = 0.0f;
#pragma unroll
        for (int nt = 0; nt < 8; ++nt) {
#pragma unroll
            for (int e = 0; e < 2; ++e) {
                s[nt][e] = __expf(s[nt][e] - base0);
                s[nt][2 + e] = __expf(s[nt][2 + e] - base1);
                ps0 += s[nt][e];
                ps1 += s[nt][2 + e];
            }
        }
        ps0 += __shfl_xor_sync(0xffffffffu, ps0, 1);
        ps0 += __shfl_xor_sync(0xffffffffu, ps0, 2);
        ps1 += __shfl_xor_sync(0xffffffffu, ps1, 1);
        ps1 += __shfl_xor_sync(0xffffffffu, ps1, 2);
        l0 = l0 * c0 + ps0;
        l1 = l1 * c1 + ps1;
#pragma unroll
        for (int nd = 0; nd < 8; ++nd) { o[nd][0] *= c0; o[nd][1] *= c0; o[nd][2] *= c1; o[nd][3] *= c1; }
        // ---- O += P V ----
#pragma unroll
        for (int kk = 0; kk < 4; ++kk) {
            uint32_t pa[4];
            pa[0] = pack2<BF16>(s[2 * kk][0], s[2 * kk][1]);
            pa[1] = pack2<BF16>(s[2 * kk][2], s[2 * kk][3]);
            pa[2] = pack2<BF16>(s[2 * kk + 1][0], s[2 * kk + 1][1]);
            pa[3] = pack2<BF16>(s[2 * kk + 1][2], s[2 * kk + 1][3]);
            const int mi = lane >> 3, mr = lane & 7;
            const int key = kk * 16 + (mi & 1) * 8 + mr;
#pragma unroll
            for (int nd = 0; nd < 8; nd += 2) {
                const uint32_t addr = (uint32_t) __cvta_generic_to_shared(Vs + key * FLDS + (nd + (mi >> 1)) * 8);
                uint32_t b0, b1, b2, b3;
                asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
                             : "=r"(b0), "=r"(b1), "=r"(b2), "=r"(b3)
                             : "r"(addr));
                mma16816<BF16>(o[nd], pa, b0, b1);
                mma16816<BF16>(o[nd + 1], pa, b2, b3);
            }
        }
    }

    const float inv0 = l0 > 0.0f ? 1.0f / l0 : 0.0f;
    const float inv1 = l1 > 0.0f ? 1.0f / l1 : 0.0f;
#pragma unroll
    for (int nd = 0; nd < 8; ++nd) {
        const int col = head * FHD + nd * 8 + 2 * tq;
        if (qv0)
            *reinterpret_cast<uint32_t *>(out + (size_t) (r0 + q0 + qi0) * dim + col) = pack2<BF16>(o[nd][0] * inv0, o[nd][1] * inv0);
        if (qv1)
            *reinterpret_cast<uint32_t *>(out + (size_t) (r0 + q0 + qi1) * dim + col) = pack2<BF16>(o[nd][2] * inv1, o[nd][3] * inv1);
    }
}

} // namespace

void launch_attention_relbias(const void * qkv16, void * out16, int dtype, int dim, int heads, const int * row0,
                              const int * seq_len, const int * gate_off, const float * gate, const float * rel_d, int max_seq,
                              const int2 * work, int n_work, cudaStream_t st) {
    if (n_work <= 0) return;
    if (dim / heads != FHD || (dim % 8) != 0) throw cuda_error("wavlm attention: head_dim must be 64");
    dim3 grid((unsigned) n_work, (unsigned) heads);
    if (dtype == OP_BF16)
        launch_pdl(attention_relbias_kernel<true>, grid, dim3(128), 0, st, (const uint16_t *) qkv16, (uint16_t *) out16, dim, row0,
                   seq_len, gate_off, gate, rel_d, max_seq, work);
    else
        launch_pdl(attention_relbias_kernel<false>, grid, dim3(128), 0, st, (const uint16_t *) qkv16, (uint16_t *) out16, dim, row0,
                   seq_len, gate_off, gate, rel_d, max_seq, work);
    count_launch();
}

} // namespace mio
