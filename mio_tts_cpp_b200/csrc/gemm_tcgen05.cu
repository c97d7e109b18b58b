// tcgen05 / TMEM / TMA GEMM for the MioCodec decoder's dense contractions (sm_100a only).
//
//   D[M, N] = sum_{tap < taps} A[row + tap - taps/2, 0:K] * W[N, tap*K : (tap+1)*K]^T      (fp32 accumulate)
//
// taps = 1: the linears of transformer_block_eval / linear()  (/root/reference/src/miocodec-decoder.cpp:128-134,
//           1190-1291), the x2 transposed conv as one [.,2*C] GEMM (:1368-1372) and the iSTFT head (:1460).
// taps = 3: ggml_conv_1d_ph k=3 of resnet_block_eval (:1293-1330) as an implicit GEMM: three row-shifted TMA
//           loads of the same [rows, C] activation; zero gap rows of the row layout give the "same" padding.
//
// Two SMs of a TPC work as a CTA pair (cluster 2x1x1, tcgen05 cta_group::2) on one 256 x BLOCK_N tile: each CTA
// stages its own 128 rows of A and HALF of the W tile, the leader's single MMA thread drives both tensor cores
// (UMMA M=256), and each CTA keeps the accumulator of its 128 rows in its own TMEM.  Per-SM L2->smem traffic per
// FLOP drops by a third against the 128 x 256 single-CTA tile and one more pipeline stage fits.
// (NCTA=1 keeps the single-CTA form for BLOCK_N=128 and for A/B testing: MIO_TTS_B200_GEMM_CTA=1.)
//
// Structure (one persistent CTA per SM):
//   warp 0      TMA producer: cp.async.bulk.tensor.2d -> 128B-swizzled smem stages, mbarrier complete_tx
//   warp 1      allocates TMEM; one lane issues tcgen05.mma.cta_group::1.kind::f16 (M=128, N=BLOCK_N, K=16),
//               tcgen05.commit releases smem stages / publishes the accumulator
//   warps 2..   epilogue (4*EW warps; a warp may only touch TMEM lanes 32*(warp%4)..+31): per 32-row x 32/64-column
//               chunk tcgen05.ld 32x32b.x32 hands each thread one ROW; the fused bias / AdaLN gate / residual /
//               SwiGLU / RoPE math runs in that layout, the result goes to a swizzled smem box and ONE lane issues a
//               TMA store (cp.async.bulk.tensor .global.shared::cta), so output rows are written as full 64/128-byte
//               lines without per-thread address arithmetic.  The fp32 residual (always the in-place output) is
//               TMA-LOADED into the same box two chunks ahead.  Two TMEM accumulators overlap epilogue(i) with MMA(i+1).
#include "common.h"

#include <cuda_bf16.h>
#include <cuda_fp16.h>

#include <atomic>
#include <cstdlib>
#include <cstring>
#include <mutex>

namespace mio {

std::atomic<unsigned long long> g_launch_count{0};

namespace {

constexpr int BLOCK_M = 128;
constexpr int BLOCK_K = 64;      // 16-bit elements -> 128-byte rows = one SWIZZLE_128B atom
constexpr int UMMA_K = 16;
constexpr int B_BOX_ROWS = 128;      // rows of W per TMA box of a whole tile: ONE request per stage and CTA (with four 32-row boxes the
                                     // producer's request rate, not the bytes, bounded the short-K shapes: QKV 82 -> 76 us, step -2 %)
constexpr int B_TAIL_BOX_ROWS = 32;  // ... of a split-tail sub-tile (map_b_tail): the narrowest one (BLOCK_N / 4) of a CTA pair takes one
constexpr uint32_t PEER_BIT_MASK = 0xFEFFFFFFu;   // clears the CTA-pair rank bit of a shared::cluster address
constexpr int EPI_NBUF = 3;      // output boxes per epilogue warp (residual prefetch distance = 2 chunks)

// per-epilogue chunk geometry: CW accumulator columns -> one 32-row box of ROW_BYTES per row
template <int EPI> struct epi_cfg {
    static constexpr int CW = EPI == EPI_SWIGLU16 ? 64 : 32;
    static constexpr int ROW_BYTES = EPI == EPI_STORE32 ? 128 : 64;
    static constexpr int BUF_BYTES = 32 * ROW_BYTES;
    static constexpr int ROPE_BYTES = 0;      // (the (cos,sin) values of a row live in its thread's registers)
    static constexpr int WARP_BYTES = EPI_NBUF * BUF_BYTES + ROPE_BYTES;
};

template <int BN, int EPI, int EW, int NCTA> struct tile_cfg {   // EW = epilogue warps per TMEM lane quadrant
    static constexpr int THREADS = 64 + 128 * EW;
    static constexpr int A_BYTES = BLOCK_M * BLOCK_K * 2;
    static constexpr int B_ROWS = BN / NCTA;               // rows of the W tile staged by ONE CTA
    static constexpr int B_BYTES = B_ROWS * BLOCK_K * 2;
    static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
    static constexpr int EPI_BYTES = 4 * EW * epi_cfg<EPI>::WARP_BYTES;
    static constexpr int BAR_BYTES = 768;                  // pipeline barriers [0,256) + per-warp residual barriers
    static constexpr int STAGES_FIT = (227 * 1024 - 1024 - BAR_BYTES - EPI_BYTES) / STAGE_BYTES;
    static constexpr int STAGES = STAGES_FIT > 8 ? 8 : STAGES_FIT;
    static constexpr int TMEM_COLS = 2 * BN;
    static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + 1024 /*alignment slack*/ + EPI_BYTES + BAR_BYTES;
    static_assert(B_ROWS % B_BOX_ROWS == 0, "W stage must be a whole number of TMA boxes");
    static_assert(2 * STAGES + 5 <= 32, "pipeline barrier block is 256 bytes");
    static_assert(256 + 4 * EW * EPI_NBUF * 8 <= BAR_BYTES, "residual barrier block");
    static_assert(STAGES >= 2, "not enough shared memory for the operand pipeline");
};

__device__ __forceinline__ uint32_t smem_u32(const void * p) { return (uint32_t) __cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    uint32_t done = 0;
    long long t0 = 0;
    while (true) {
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(done)
            : "r"(bar), "r"(parity)
            : "memory");
        if (done) break;
        // a protocol bug must not hang the GPU: trap after ~2 s of waiting
        if (t0 == 0) t0 = clock64();
        else if (clock64() - t0 > 4000000000ll) __trap();
    }
}
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap * map, uint32_t bar, int c0, int c1) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
        ::"r"(dst), "l"((uint64_t) map), "r"(bar), "r"(c0), "r"(c1)
        : "memory");
}
// CTA-pair forms: the load lands in the issuing CTA's smem, its bytes are counted on the LEADER's mbarrier
__device__ __forceinline__ void tma_load_2d_pair(uint32_t dst, const CUtensorMap * map, uint32_t bar, int c0, int c1) {
    asm volatile(
        "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
        ::"r"(dst), "l"((uint64_t) map), "r"(bar & PEER_BIT_MASK), "r"(c0), "r"(c1)
        : "memory");
}
__device__ __forceinline__ void mbar_arrive_leader(uint32_t bar) {
    asm volatile("mbarrier.arrive.shared::cluster.b64 _, [%0];" ::"r"(bar & PEER_BIT_MASK) : "memory");
}
__device__ __forceinline__ uint32_t cluster_ctarank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ void tma_store_2d(const CUtensorMap * map, uint32_t src, int c0, int c1) {
    asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];"
                 ::"l"((uint64_t) map), "r"(src), "r"(c0), "r"(c1)
                 : "memory");
}
// out[box] += smem[box], the addition performed by the L2 (one rounding per element, like a read-add-write in the SM)
__device__ __forceinline__ void tma_reduce_add_2d(const CUtensorMap * map, uint32_t src, int c0, int c1) {
    asm volatile("cp.reduce.async.bulk.tensor.2d.global.shared::cta.add.tile.bulk_group [%0, {%2, %3}], [%1];"
                 ::"l"((uint64_t) map), "r"(src), "r"(c0), "r"(c1)
                 : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int N> __device__ __forceinline__ void bulk_wait_read() {   // smem of all but the N newest stores is reusable
    asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory");
}
__device__ __forceinline__ void bulk_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ float4 lds128(uint32_t addr) {
    float4 v;
    asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(addr));
    return v;
}
__device__ __forceinline__ void sts128(uint32_t addr, float4 v) {
    asm volatile("st.shared.v4.f32 [%0], {%1,%2,%3,%4};" ::"r"(addr), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}
__device__ __forceinline__ void sts128u(uint32_t addr, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
    asm volatile("st.shared.v4.b32 [%0], {%1,%2,%3,%4};" ::"r"(addr), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}
__device__ __forceinline__ void tcgen05_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tcgen05_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
template <int NCTA> __device__ __forceinline__ void tcgen05_commit(uint32_t bar) {
    if constexpr (NCTA == 1) {
        asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
    } else {   // arrives on the barrier at this offset in BOTH CTAs of the pair
        asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
                     ::"r"(bar), "h"((uint16_t) 3)
                     : "memory");
    }
}
template <int NCTA>
__device__ __forceinline__ void tcgen05_mma_f16(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc,
                                                uint32_t accumulate) {
    if constexpr (NCTA == 1) {
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "setp.ne.b32 p, %4, 0;\n\t"
            "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
            ::"r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
            : "memory");
    } else {
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "setp.ne.b32 p, %4, 0;\n\t"
            "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}"
            ::"r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
            : "memory");
    }
}
__device__ __forceinline__ void tmem_ld_32x32b_x32(uint32_t taddr, uint32_t (&r)[32]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
          "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
          "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
          "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(taddr)
        : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// K-major, SWIZZLE_128B shared-memory matrix descriptor (cute::UMMA::SmemDescriptor bit layout):
// start>>4 [0,14) | LBO>>4 [16,30) | SBO>>4 [32,46) | version=1 [46,48) | layout SWIZZLE_128B=2 [61,64)
__device__ __forceinline__ uint64_t umma_desc_sw128(uint32_t smem_addr) {
    uint64_t d = 0;
    d |= (uint64_t) ((smem_addr >> 4) & 0x3FFF);
    d |= (uint64_t) 1 << 16;                 // leading byte offset (unused for swizzled K-major), 16 B
    d |= (uint64_t) (1024 >> 4) << 32;       // stride byte offset: 8 rows x 128 B
    d |= (uint64_t) 1 << 46;                 // descriptor version (Blackwell)
    d |= (uint64_t) 2 << 61;                 // SWIZZLE_128B
    return d;
}

__device__ __forceinline__ uint32_t pack16(float a, float b, int dtype) {
    if (dtype == OP_BF16) {
        __nv_bfloat162 v = __floats2bfloat162_rn(a, b);
        return *reinterpret_cast<uint32_t *>(&v);
    }
    return pack_f16x2_sat(a, b);
}

// x * sigmoid(x) on the SFU (ex2.approx + rcp.approx, ~2 ulp; the result is rounded to 16 bits anyway)
__device__ __forceinline__ float silu_f(float x) { return __fdividef(x, 1.0f + __expf(-x)); }

// ---- epilogue math of one chunk, row-per-thread (tcgen05.ld 32x32b: lane = row, 32 consecutive columns) ----------
// Output boxes in shared memory use the TMA swizzle of their tensor map so that a quarter-warp's 16-byte accesses
// hit 32 distinct banks: SWIZZLE_128B for 128-byte rows (fp32), SWIZZLE_64B for 64-byte rows (16-bit).

// ggml_gelu on the reference's CPU backend is a table over fp16 inputs: gelu(x) = fp16(gelu_tanh(fp16(x)))
// (ggml-cpu/vec.h, GGML_GELU_FP16); the same two roundings are applied here so both sides agree to the fp16 ulp.
__device__ __forceinline__ float act_apply(float x, int act) {
    if (act == 1) {
        const float h = __half2float(__float2half_rn(x));
        const float u = 0.79788456080286535587989211986876f * h * (1.0f + 0.044715f * h * h);
        const float t = 1.0f - __fdividef(2.0f, 1.0f + __expf(2.0f * u));      // tanh(u)
        const float y = __half2float(__float2half_rn(0.5f * h * (1.0f + t)));
        return x <= -10.0f ? 0.0f : (x >= 10.0f ? x : y);
    }
    if (act == 2) return tanhf(x);
    return x;
}

// SnakeBeta (/root/reference/src/miocodec-decoder.cpp:1332-1343): x + sin^2(alpha x) / beta, the sine on the SFU after a
// two-constant reduction of the argument to [-pi, pi] (absolute error ~1e-6)
__device__ __forceinline__ float snake_f(float x, float al, float ib) {
    const float t = x * al;
    const float kf = rintf(t * 0.15915494309189535f);
    float r = fmaf(-kf, 6.2831854820251465f, t);
    r = fmaf(-kf, -1.7484555e-7f, r);
    const float sn = __sinf(r);
    return fmaf(sn * sn, ib, x);
}

// fp32 box: out = act(acc + bias) * gate + residual, residual read from / result written to the same box slot.
// BIAS/GATE/RES are compile-time so an absent term costs no issue slots.  The bias / gate values of the chunk (bv, gv)
// were fetched by the caller one chunk ahead (zeros past N: those columns are clipped by the TMA store).
// STATS: sq = {sum, sum of squares} of this row's columns 0-15 and 16-31 of the chunk (final values, residual included).
// SNAKE: act 3; gv holds the chunk's alpha values (fetched like a gate), the 1/beta values are read here.
template <bool BIAS, bool GATE, bool RES, bool STATS = false, bool SNAKE = false>
__device__ __forceinline__ void chunk_store32(const gemm_args & a, const uint32_t (&raw)[32], uint32_t buf, int lane,
                                              const float4 (&bv)[8], const float4 (&gv)[8], float (&sq)[4], int n = 0) {
    const uint32_t row = buf + lane * 128;
    const uint32_t sw = (uint32_t) (lane & 7) << 4;
    if (STATS) { sq[0] = sq[1] = sq[2] = sq[3] = 0.0f; }
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        const uint32_t addr = row + (((uint32_t) j << 4) ^ sw);
        float4 o = make_float4(__uint_as_float(raw[4 * j]), __uint_as_float(raw[4 * j + 1]), __uint_as_float(raw[4 * j + 2]),
                               __uint_as_float(raw[4 * j + 3]));
        if (BIAS) { o.x += bv[j].x; o.y += bv[j].y; o.z += bv[j].z; o.w += bv[j].w; }
        if constexpr (SNAKE) {
            const float4 ib = __ldg(reinterpret_cast<const float4 *>(a.snake_ib + n) + j);
            o.x = snake_f(o.x, gv[j].x, ib.x); o.y = snake_f(o.y, gv[j].y, ib.y);
            o.z = snake_f(o.z, gv[j].z, ib.z); o.w = snake_f(o.w, gv[j].w, ib.w);
        } else if (a.act != 0) {   // warp-uniform
            o.x = act_apply(o.x, a.act); o.y = act_apply(o.y, a.act); o.z = act_apply(o.z, a.act); o.w = act_apply(o.w, a.act);
        }
        if (GATE) { o.x *= gv[j].x; o.y *= gv[j].y; o.z *= gv[j].z; o.w *= gv[j].w; }
        if (RES) {
            const float4 q = lds128(addr);
            o.x += q.x; o.y += q.y; o.z += q.z; o.w += q.w;
        }
        if (STATS) {
            sq[2 * (j >> 2)] += (o.x + o.y) + (o.z + o.w);
            sq[2 * (j >> 2) + 1] += (o.x * o.x + o.y * o.y) + (o.z * o.z + o.w * o.w);
        }
        sts128(addr, o);
    }
}

// warp-uniform dispatch to the term combinations in use; has_res: the residual sits in the box and is added here
// (false in reduce-add mode, where the box carries the new term alone)
__device__ __forceinline__ void chunk_store32_any(const gemm_args & a, const uint32_t (&raw)[32], uint32_t buf, int lane,
                                                  const float4 (&bv)[8], const float4 (&gv)[8], bool has_res, float (&sq)[4], int n) {
    if (a.act == 3) {   // polyphase transposed convolution of the wave upsampler: bias, SnakeBeta, plain store
        chunk_store32<true, false, false, false, true>(a, raw, buf, lane, bv, gv, sq, n);
        return;
    }
    if (a.gn_stats) {   // the k=3 convolutions of the resnet blocks: bias, no gate; conv2 with the residual in the box
        if (has_res) chunk_store32<true, false, true, true>(a, raw, buf, lane, bv, gv, sq);
        else chunk_store32<true, false, false, true>(a, raw, buf, lane, bv, gv, sq);
        return;
    }
    if (a.gate) {
        if (has_res) { if (a.bias) chunk_store32<true, true, true>(a, raw, buf, lane, bv, gv, sq);
                       else chunk_store32<false, true, true>(a, raw, buf, lane, bv, gv, sq); }
        else { if (a.bias) chunk_store32<true, true, false>(a, raw, buf, lane, bv, gv, sq);
               else chunk_store32<false, true, false>(a, raw, buf, lane, bv, gv, sq); }
    } else if (has_res) {
        if (a.bias) chunk_store32<true, false, true>(a, raw, buf, lane, bv, gv, sq);
        else chunk_store32<false, false, true>(a, raw, buf, lane, bv, gv, sq);
    } else {
        if (a.bias) chunk_store32<true, false, false>(a, raw, buf, lane, bv, gv, sq);
        else chunk_store32<false, false, false>(a, raw, buf, lane, bv, gv, sq);
    }
}

template <bool BF16> __device__ __forceinline__ uint32_t pack16t(float a, float b) {
    if constexpr (BF16) {
        __nv_bfloat162 v = __floats2bfloat162_rn(a, b);
        return *reinterpret_cast<uint32_t *>(&v);
    } else {
        return pack_f16x2_sat(a, b);
    }
}

// 16-bit box of 32 columns (+ bias); ROPE rotates (even, odd) column pairs by this row's (cos, sin) from the staging tile
template <bool ROPE, bool BIAS, bool BF16>
__device__ __forceinline__ void chunk_store16(const gemm_args & a, const uint32_t (&raw)[32], uint32_t buf, const float4 (&rcs)[8],
                                              int lane, int n) {
    const uint32_t row = buf + lane * 64;
    const uint32_t sw = (uint32_t) ((lane >> 1) & 3) << 4;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        float x[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) x[e] = __uint_as_float(raw[8 * j + e]);
        if (BIAS && n + 8 * j < a.N) {
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                const float4 b = __ldg(reinterpret_cast<const float4 *>(a.bias + n) + 2 * j + h);
                x[4 * h] += b.x; x[4 * h + 1] += b.y; x[4 * h + 2] += b.z; x[4 * h + 3] += b.w;
            }
        }
        if (!ROPE && a.act == 3) {   // warp-uniform
            if (n + 8 * j < a.N) {
#pragma unroll
                for (int h = 0; h < 2; ++h) {
                    const float4 al = __ldg(reinterpret_cast<const float4 *>(a.snake_a + n) + 2 * j + h);
                    const float4 ib = __ldg(reinterpret_cast<const float4 *>(a.snake_ib + n) + 2 * j + h);
                    x[4 * h] = snake_f(x[4 * h], al.x, ib.x); x[4 * h + 1] = snake_f(x[4 * h + 1], al.y, ib.y);
                    x[4 * h + 2] = snake_f(x[4 * h + 2], al.z, ib.z); x[4 * h + 3] = snake_f(x[4 * h + 3], al.w, ib.w);
                }
            }
        } else if (!ROPE && a.act != 0) {   // warp-uniform
#pragma unroll
            for (int e = 0; e < 8; ++e) x[e] = act_apply(x[e], a.act);
        }
        if constexpr (ROPE) {
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                const float4 cs = rcs[2 * j + h];                                        // (cos,sin) of 2 pairs
                const float x0 = x[4 * h], x1 = x[4 * h + 1], y0 = x[4 * h + 2], y1 = x[4 * h + 3];
                x[4 * h] = x0 * cs.x - x1 * cs.y;
                x[4 * h + 1] = x0 * cs.y + x1 * cs.x;
                x[4 * h + 2] = y0 * cs.z - y1 * cs.w;
                x[4 * h + 3] = y0 * cs.w + y1 * cs.z;
            }
        }
        sts128u(row + (((uint32_t) j << 4) ^ sw), pack16t<BF16>(x[0], x[1]), pack16t<BF16>(x[2], x[3]),
                pack16t<BF16>(x[4], x[5]), pack16t<BF16>(x[6], x[7]));
    }
}

template <bool ROPE>
__device__ __forceinline__ void chunk_store16_any(const gemm_args & a, const uint32_t (&raw)[32], uint32_t buf, const float4 (&rcs)[8],
                                                  int lane, int n) {
    const bool bf16 = a.out_dtype == OP_BF16;   // warp-uniform
    if (a.bias) { if (bf16) chunk_store16<ROPE, true, true>(a, raw, buf, rcs, lane, n);
                  else chunk_store16<ROPE, true, false>(a, raw, buf, rcs, lane, n); }
    else { if (bf16) chunk_store16<ROPE, false, true>(a, raw, buf, rcs, lane, n);
           else chunk_store16<ROPE, false, false>(a, raw, buf, rcs, lane, n); }
}

// SwiGLU: 32 accumulator columns (gate_j, up_j interleaved) -> 16 outputs = 16-byte slots 2*half, 2*half+1 of the row
template <bool BF16>
__device__ __forceinline__ void chunk_swiglu16(const uint32_t (&raw)[32], uint32_t buf, int lane, int half) {
    const uint32_t row = buf + lane * 64;
    const uint32_t sw = (uint32_t) ((lane >> 1) & 3) << 4;
#pragma unroll
    for (int j = 0; j < 2; ++j) {
        uint32_t o[4];
#pragma unroll
        for (int e = 0; e < 4; ++e) {
            const float h0 = silu_f(__uint_as_float(raw[16 * j + 4 * e])) * __uint_as_float(raw[16 * j + 4 * e + 1]);
            const float h1 = silu_f(__uint_as_float(raw[16 * j + 4 * e + 2])) * __uint_as_float(raw[16 * j + 4 * e + 3]);
            o[e] = pack16t<BF16>(h0, h1);
        }
        sts128u(row + (((uint32_t) (2 * half + j) << 4) ^ sw), o[0], o[1], o[2], o[3]);
    }
}

template <int BN, int EPI, int EW, int NCTA>
__global__ void __launch_bounds__(64 + 128 * EW, 1) gemm_tcgen05_kernel(const __grid_constant__ gemm_args a) {
    using C = tile_cfg<BN, EPI, EW, NCTA>;
    MIO_PDL_TRIGGER();   // the next kernel's CTAs may be scheduled as ours retire; it waits for our completion itself
    extern __shared__ uint8_t smem_raw[];
    const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    const uint32_t bar_base = smem_base + C::STAGES * C::STAGE_BYTES + C::EPI_BYTES;   // [stages][epilogue boxes][barriers]
    // barriers: full[STAGES], empty[STAGES], tmem_full[2], tmem_empty[2]; then the TMEM base address.
    // In a CTA pair both CTAs hold the same block; full[] and tmem_empty[] are only used in the leader (rank 0).
    auto full_bar = [&](int s) { return bar_base + 8u * s; };
    auto empty_bar = [&](int s) { return bar_base + 8u * (C::STAGES + s); };
    auto tfull_bar = [&](int s) { return bar_base + 8u * (2 * C::STAGES + s); };
    auto tempty_bar = [&](int s) { return bar_base + 8u * (2 * C::STAGES + 2 + s); };
    const uint32_t tmem_slot = bar_base + 8u * (2 * C::STAGES + 4);
    volatile uint32_t * tmem_slot_ptr =
        reinterpret_cast<volatile uint32_t *>(smem_raw + (tmem_slot - smem_u32(smem_raw)));

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const int cta_rank = NCTA == 2 ? (int) cluster_ctarank() : 0;
    const int unit = blockIdx.x / NCTA;            // CTA (pair) index = persistent tile-loop lane
    const int n_units = gridDim.x / NCTA;

    constexpr int TILE_M = BLOCK_M * NCTA;
    const int num_m_tiles = (a.M + TILE_M - 1) / TILE_M;
    const int num_n_tiles = (a.N + BN - 1) / BN;
    const int num_tiles = num_m_tiles * num_n_tiles;
    const int k_blocks = (a.K + BLOCK_K - 1) / BLOCK_K;
    const int k_iters = k_blocks * a.taps;
    // Work items: tiles [0, split_from) whole; the tiles of the last, partial wave are cut along N into
    // (1 << split_shift) sub-tiles of BN >> split_shift columns so that the tail occupies more CTA pairs.  The K order of
    // every output element is unchanged, so results do not depend on the split.
    const int split_from = a.split_from, split_shift = a.split_shift;
    const int num_items = split_from + ((num_tiles - split_from) << split_shift);
    auto item_tile = [&](int w) { return w < split_from ? w : split_from + ((w - split_from) >> split_shift); };
    auto item_width = [&](int w) { return w < split_from ? BN : (BN >> split_shift); };
    auto item_n0 = [&](int w) {   // first column of the item
        const int t = item_tile(w);
        return (t % num_n_tiles) * BN + (w < split_from ? 0 : ((w - split_from) & ((1 << split_shift) - 1)) * (BN >> split_shift));
    };
    auto item_m0 = [&](int w) {     // reverse_m: row tiles from the last to the first (ping-pong traversal, common.h)
        const int mt = item_tile(w) / num_n_tiles;
        return (a.reverse_m ? num_m_tiles - 1 - mt : mt) * TILE_M;
    };

    if (threadIdx.x == 0) {
        for (int s = 0; s < C::STAGES; ++s) { mbar_init(full_bar(s), 1); mbar_init(empty_bar(s), 1); }
        for (int s = 0; s < 2; ++s) { mbar_init(tfull_bar(s), 1); mbar_init(tempty_bar(s), 4 * EW * NCTA); }
        for (int s = 0; s < 4 * EW * EPI_NBUF; ++s) mbar_init(bar_base + 256u + 8u * s, 1);   // residual boxes
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) {
        if constexpr (NCTA == 1) {
            asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot),
                         "r"((uint32_t) C::TMEM_COLS)
                         : "memory");
            asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
        } else {
            asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot),
                         "r"((uint32_t) C::TMEM_COLS)
                         : "memory");
            asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
        }
    }
    tcgen05_fence_before();
    if constexpr (NCTA == 2) cluster_sync_all(); else __syncthreads();
    tcgen05_fence_after();
    const uint32_t tmem_base = *tmem_slot_ptr;
    MIO_PDL_WAIT();      // barriers and TMEM are set up; from here on we touch memory the previous kernel wrote

    if (warp == 0) {
        // ===== TMA producer (every CTA stages its own A rows and its share of the W tile) =====
        if (lane == 0) {
            uint32_t kit = 0;
            for (int t = unit; t < num_items; t += n_units) {
                const int b_rows = item_width(t) / NCTA;               // W rows this CTA stages for the item
                const bool tail_item = t >= split_from && split_shift > 0;
                const int box_rows = tail_item ? B_TAIL_BOX_ROWS : B_BOX_ROWS;
                const CUtensorMap * wmap = tail_item ? &a.map_b_tail : &a.map_b;
                const int n_boxes = b_rows / box_rows;
                const uint32_t stage_tx = C::A_BYTES + b_rows * BLOCK_K * 2;
                const int m0 = item_m0(t) + cta_rank * BLOCK_M;
                const int n0 = item_n0(t) + cta_rank * b_rows;
                for (int tap = 0; tap < a.taps; ++tap) {
                    const int arow = m0 + a.tap_row0 + (tap >> a.tap_shift);
                    const int acol = a.a_col0 + (tap & ((1 << a.tap_shift) - 1)) * a.K;
                    for (int kb = 0; kb < k_blocks; ++kb, ++kit) {
                        const int s = kit % C::STAGES;
                        mbar_wait(empty_bar(s), ((kit / C::STAGES) & 1) ^ 1);
                        const uint32_t sa = smem_base + s * C::STAGE_BYTES;
                        if constexpr (NCTA == 1) {
                            mbar_expect_tx(full_bar(s), stage_tx);
                            tma_load_2d(sa, &a.map_a, full_bar(s), acol + kb * BLOCK_K, arow);
                            for (int bx = 0; bx < n_boxes; ++bx)
                                tma_load_2d(sa + C::A_BYTES + bx * (box_rows * BLOCK_K * 2), wmap, full_bar(s),
                                            tap * a.K + kb * BLOCK_K, n0 + bx * box_rows);
                        } else {
                            if (cta_rank == 0) mbar_expect_tx(full_bar(s), 2 * stage_tx);   // both CTAs' bytes
                            tma_load_2d_pair(sa, &a.map_a, full_bar(s), acol + kb * BLOCK_K, arow);
                            for (int bx = 0; bx < n_boxes; ++bx)
                                tma_load_2d_pair(sa + C::A_BYTES + bx * (box_rows * BLOCK_K * 2), wmap, full_bar(s),
                                                 tap * a.K + kb * BLOCK_K, n0 + bx * box_rows);
                        }
                    }
                }
            }
        }
    } else if (warp == 1) {
        // ===== MMA issuer (leader CTA only) =====
        if (lane == 0 && cta_rank == 0) {
            // instruction descriptor (cute::UMMA::InstrDescriptor): c=f32 [4,6), a/b format [7,10)/[10,13),
            // K-major a/b (bits 15,16 = 0), N>>3 [17,23), M>>4 [24,29); M counts the rows of the whole pair
            const uint32_t idesc0 = (1u << 4) | ((uint32_t) a.dtype << 7) | ((uint32_t) a.dtype << 10) |
                                    ((uint32_t) (TILE_M >> 4) << 24);
            uint32_t kit = 0, it = 0;
            for (int t = unit; t < num_items; t += n_units, ++it) {
                const uint32_t idesc = idesc0 | ((uint32_t) (item_width(t) >> 3) << 17);
                const int acc = it & 1;
                mbar_wait(tempty_bar(acc), ((it >> 1) & 1) ^ 1);
                tcgen05_fence_after();
                const uint32_t tmem_d = tmem_base + acc * BN;
                for (int ki = 0; ki < k_iters; ++ki, ++kit) {
                    const int s = kit % C::STAGES;
                    mbar_wait(full_bar(s), (kit / C::STAGES) & 1);
                    tcgen05_fence_after();
                    const uint32_t sa = smem_base + s * C::STAGE_BYTES;
                    const uint64_t da = umma_desc_sw128(sa);
                    const uint64_t db = umma_desc_sw128(sa + C::A_BYTES);
#pragma unroll
                    for (int k = 0; k < BLOCK_K / UMMA_K; ++k) {
                        // advance 32 bytes (16 elements) inside the 128B swizzle atom: +2 in the >>4 address field
                        tcgen05_mma_f16<NCTA>(tmem_d, da + 2 * k, db + 2 * k, idesc, (ki | k) != 0);
                    }
                    tcgen05_commit<NCTA>(empty_bar(s));   // frees the smem stage (in both CTAs) once these MMAs retire
                }
                tcgen05_commit<NCTA>(tfull_bar(acc));     // accumulator complete (signalled in both CTAs)
            }
        }
    } else {
        // ===== epilogue warps (each CTA drains the accumulator of its own 128 rows) =====
        using E = epi_cfg<EPI>;
        const int quad = warp & 3;                   // TMEM lane quadrant this warp may access
        const int ew = warp - 2;                     // epilogue warp index; column group = ew / 4
        const int cgrp = ew >> 2;
        const uint32_t ebuf = smem_base + C::STAGES * C::STAGE_BYTES + ew * E::WARP_BYTES;
        auto rbar = [&](int b) { return bar_base + 256u + 8u * (ew * EPI_NBUF + b); };
        // in-place fp32 residual: either read back through TMA, added in registers and stored (has_res), or -- red_add --
        // left where it is and updated by a TMA reduce-add of the new term alone (no residual bytes cross L2 -> SM)
        const bool red_add = EPI == EPI_STORE32 && a.res != nullptr && a.red_add != 0;
        const bool has_res = EPI == EPI_STORE32 && a.res != nullptr && !red_add;

        auto tile_row0 = [&](int t) { return item_m0(t) + cta_rank * BLOCK_M + quad * 32; };
        auto chunk_ok = [&](int t, int c) {   // chunk c (of E::CW accumulator columns) of item t holds output
            return c * E::CW < item_width(t) && item_n0(t) + c * E::CW < a.N && tile_row0(t) < a.M;
        };
        // residual prefetch cursor: walks this warp's chunks in processing order, across tiles
        int pf_t = unit, pf_c = cgrp;
        uint32_t n_ld = 0;
        auto pf_seek = [&]() { while (pf_t < num_items && !chunk_ok(pf_t, pf_c)) { pf_t += n_units; pf_c = cgrp; } };
        auto pf_issue = [&]() {   // lane 0: TMA-load the residual box of the cursor's chunk, advance the cursor
            if (pf_t >= num_items) return;
            const int b = n_ld % EPI_NBUF;
            mbar_expect_tx(rbar(b), E::BUF_BYTES);
            tma_load_2d(ebuf + b * E::BUF_BYTES, &a.map_out, rbar(b), item_n0(pf_t) + pf_c * E::CW, tile_row0(pf_t));
            ++n_ld;
            pf_c += EW;
            pf_seek();
        };
        if (has_res && lane == 0) {
            pf_seek();
            pf_issue();
            pf_issue();
        }
        __syncwarp();

        uint32_t it = 0, nch = 0;                    // tiles / chunks processed by this warp
        // ROPE16: this lane's row position, loaded one item ahead so that its latency never sits in front of a tile
        int pos_next = 0;
        if constexpr (EPI == EPI_ROPE16) {
            if (unit < num_items) { const int r = tile_row0(unit) + lane; pos_next = r < a.M ? __ldg(a.row_pos + r) : 0; }
        }
        // STORE32 with a per-utterance gate: likewise this lane's row -> utterance index
        auto utt_of = [&](int w) {
            const int r = tile_row0(w) + lane;
            if (a.gn_stats != nullptr) return (w < num_items && r < a.M) ? __ldg(a.row_utt + r) : -1;   // -1: not a data row
            return (a.gate != nullptr && a.row_utt != nullptr && w < num_items && r < a.M) ? __ldg(a.row_utt + r) : 0;
        };
        int utt_next = EPI == EPI_STORE32 ? utt_of(unit) : 0;
        for (int t = unit; t < num_items; t += n_units, ++it) {
            const int n0 = item_n0(t);
            const int acc = it & 1;
            const int row0 = tile_row0(t);
            // per-row operands of this tile: the AdaLN gate vector of the row's utterance (STORE32), or the rope-table
            // rows this lane fetches for the transposed (cos,sin) staging (ROPE16: row (lane>>3)+4u, 16-byte slot lane&7)
            const float * gate_row = nullptr;
            const float4 * rope_row = nullptr;           // ROPE16: this thread's row of the (cos, sin) table
            int utt_row = 0;
            if constexpr (EPI == EPI_STORE32) {
                if (a.gate) gate_row = a.gate + (size_t) max(utt_next, 0) * a.gate_ld;
                if (a.act == 3) gate_row = a.snake_a;          // SnakeBeta alpha per column, fetched like a gate
                utt_row = utt_next;
                utt_next = utt_of(t + n_units);
            } else if constexpr (EPI == EPI_ROPE16) {
                const int pos = max(pos_next, 0);
                if (t + n_units < num_items) { const int r = tile_row0(t + n_units) + lane; pos_next = r < a.M ? __ldg(a.row_pos + r) : 0; }
                rope_row = reinterpret_cast<const float4 *>(a.rope + (size_t) pos * (a.head_dim >> 1));
            }
            const uint32_t taddr = tmem_base + ((uint32_t) (quad * 32) << 16) + acc * BN;
            // RoPE: a thread owns a row, and the 16 (cos, sin) pairs of a 32-column chunk depend only on the row's position and
            // on where the chunk sits inside its head -- with head_dim 64 and two column groups a warp's chunks of a tile all
            // sit in the same half of a head, so the 128 bytes are read ONCE per tile (straight from the table: 4 full
            // sectors per thread, L2-resident) and stay in registers; no staging tile, no shared-memory transpose.
            float4 cs[EPI == EPI_ROPE16 ? 8 : 1];
            int cs_p0 = -1;
            auto rope_fetch = [&](int c) {
                if constexpr (EPI == EPI_ROPE16) {
                    const int n = n0 + c * E::CW;
                    if (chunk_ok(t, c) && n < a.rope_cols) {
                        const int hd = a.head_dim;
                        const int p0 = (((hd & (hd - 1)) == 0 ? (n & (hd - 1)) : (n % hd))) >> 2;   // first pair of the chunk, in float4 (2-pair) units
                        if (p0 != cs_p0) {               // warp-uniform
                            cs_p0 = p0;
#pragma unroll
                            for (int u = 0; u < 8; ++u) cs[u] = __ldg(rope_row + p0 + u);
                        }
                    }
                }
            };
            rope_fetch(cgrp);
            // STORE32: bias and gate values of a chunk, fetched the same way (they were 4-8 serialised L2 round trips
            // in front of the math of every chunk: 40 % of the epilogue samples of the O-projections)
            float4 bv[EPI == EPI_STORE32 ? 8 : 1], gv[EPI == EPI_STORE32 ? 8 : 1];
            auto bg_fetch = [&](int c) {
                if constexpr (EPI == EPI_STORE32) {
                    if (!chunk_ok(t, c)) return;
                    const int n = n0 + c * E::CW;
                    const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
                    if (a.bias) {
#pragma unroll
                        for (int j = 0; j < 8; ++j) bv[j] = n + 4 * j < a.N ? __ldg(reinterpret_cast<const float4 *>(a.bias + n) + j) : z;
                    }
                    if (gate_row) {
#pragma unroll
                        for (int j = 0; j < 8; ++j) gv[j] = n + 4 * j < a.N ? __ldg(reinterpret_cast<const float4 *>(gate_row + n) + j) : z;
                    }
                }
            };
            bg_fetch(cgrp);
            mbar_wait(tfull_bar(acc), (it >> 1) & 1);
            tcgen05_fence_after();
#pragma unroll 1
            for (int c = cgrp; chunk_ok(t, c) && !(a.dbg & 1); c += EW, ++nch) {
                const int n = n0 + c * E::CW;
                const int b = nch % EPI_NBUF;
                const uint32_t buf = ebuf + b * E::BUF_BYTES;
                uint32_t raw[32];
                if constexpr (EPI == EPI_ROPE16) {
                    const bool rotate = n < a.rope_cols;
                    tmem_ld_32x32b_x32(taddr + c * 32, raw);
                    rope_fetch(c);                       // a no-op unless the chunk sits elsewhere in its head than the last one
                    if (lane == 0) bulk_wait_read<EPI_NBUF - 1>();
                    __syncwarp();
                    tmem_ld_wait();
                    if (rotate) chunk_store16_any<true>(a, raw, buf, cs, lane, n);
                    else chunk_store16_any<false>(a, raw, buf, cs, lane, n);
                } else if constexpr (EPI == EPI_SWIGLU16) {
                    tmem_ld_32x32b_x32(taddr + c * 64, raw);
                    if (lane == 0) bulk_wait_read<EPI_NBUF - 1>();
                    __syncwarp();
                    tmem_ld_wait();
                    if (a.out_dtype == OP_BF16) {
                        chunk_swiglu16<true>(raw, buf, lane, 0);
                        tmem_ld_32x32b_x32(taddr + c * 64 + 32, raw);
                        tmem_ld_wait();
                        chunk_swiglu16<true>(raw, buf, lane, 1);
                    } else {
                        chunk_swiglu16<false>(raw, buf, lane, 0);
                        tmem_ld_32x32b_x32(taddr + c * 64 + 32, raw);
                        tmem_ld_wait();
                        chunk_swiglu16<false>(raw, buf, lane, 1);
                    }
                } else if constexpr (EPI == EPI_STORE16) {
                    tmem_ld_32x32b_x32(taddr + c * 32, raw);
                    if (lane == 0) bulk_wait_read<EPI_NBUF - 1>();
                    __syncwarp();
                    tmem_ld_wait();
                    { const float4 none[8] = {}; chunk_store16_any<false>(a, raw, buf, none, lane, n); }
                } else {
                    tmem_ld_32x32b_x32(taddr + c * 32, raw);
                    if (has_res) mbar_wait(rbar(b), (nch / EPI_NBUF) & 1);       // residual box landed
                    else { if (lane == 0) bulk_wait_read<EPI_NBUF - 1>(); __syncwarp(); }
                    tmem_ld_wait();
                    float sq[4];
                    chunk_store32_any(a, raw, buf, lane, bv, gv, has_res, sq, n);
                    bg_fetch(c + EW);
                    if (a.gn_stats != nullptr) {
                        // GroupNorm statistics of the consumer: this row's two 16-column halves of the chunk
                        const int cpg = a.N / a.gn_groups;                       // 16, or a multiple of 32
                        const int g0 = n / cpg, g1 = (n + 16) / cpg;
                        if (g0 == g1) { sq[0] += sq[2]; sq[1] += sq[3]; }
                        const int u0 = __shfl_sync(0xffffffffu, utt_row, 0);
                        if (__all_sync(0xffffffffu, utt_row == u0 && u0 >= 0)) {  // one utterance: reduce over the 32 rows first
                            // in DOUBLE: which rows share a warp depends on where the utterance sits in the batch layout, and
                            // an utterance must decode to the same bits alone and inside any batch -- fp32 partial sums over
                            // rows would make the statistics depend on that grouping
                            double d[4] = {(double) sq[0], (double) sq[1], (double) sq[2], (double) sq[3]};
#pragma unroll
                            for (int o = 16; o > 0; o >>= 1) {
#pragma unroll
                                for (int e = 0; e < 4; ++e) d[e] += __shfl_xor_sync(0xffffffffu, d[e], o);
                            }
                            const double dv = lane == 0 ? d[0] : (lane == 1 ? d[1] : (lane == 2 ? d[2] : d[3]));
                            if (lane < 4 && (lane < 2 || g0 != g1))
                                atomicAdd(a.gn_stats + ((size_t) u0 * a.gn_groups + (lane < 2 ? g0 : g1)) * 2 + (lane & 1), dv);
                        } else if (utt_row >= 0) {                                // utterance boundary or gap rows in the warp
                            double * sp = a.gn_stats + (size_t) utt_row * a.gn_groups * 2;
                            atomicAdd(sp + 2 * g0, (double) sq[0]);
                            atomicAdd(sp + 2 * g0 + 1, (double) sq[1]);
                            if (g0 != g1) { atomicAdd(sp + 2 * g1, (double) sq[2]); atomicAdd(sp + 2 * g1 + 1, (double) sq[3]); }
                        }
                    }
                }
                fence_proxy_async();
                __syncwarp();
                if (lane == 0) {
                    if (red_add) tma_reduce_add_2d(&a.map_out, buf, n, row0);
                    else tma_store_2d(&a.map_out, buf, EPI == EPI_SWIGLU16 ? (n >> 1) : n, row0);
                    bulk_commit();
                    if (has_res) {
                        // the box of chunk nch-1 is the next prefetch target: its store must have read smem
                        bulk_wait_read<1>();
                        pf_issue();
                    }
                }
            }
            tcgen05_fence_before();
            __syncwarp();
            if (lane == 0) {
                if constexpr (NCTA == 1) mbar_arrive(tempty_bar(acc));
                else mbar_arrive_leader(tempty_bar(acc));
            }
        }
        if (lane == 0) bulk_wait_all();              // smem must outlive the last stores' reads
        __syncwarp();
    }

    tcgen05_fence_before();
    if constexpr (NCTA == 2) cluster_sync_all(); else __syncthreads();
    if (warp == 1) {
        tcgen05_fence_after();
        if constexpr (NCTA == 1) {
            asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base),
                         "r"((uint32_t) C::TMEM_COLS)
                         : "memory");
        } else {
            asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base),
                         "r"((uint32_t) C::TMEM_COLS)
                         : "memory");
        }
    }
}

#include "gemm_ln.inc"
#include "gemm_astat.inc"

typedef CUresult (*encode_tiled_fn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                    const cuuint64_t *, const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave,
                                    CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

encode_tiled_fn get_encode_fn() {
    static encode_tiled_fn fn = nullptr;
    static std::once_flag once;
    std::call_once(once, [] {
        void * p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess) {
            fn = (encode_tiled_fn) p;
        }
    });
    if (!fn) throw cuda_error("cuTensorMapEncodeTiled is not available from the CUDA driver");
    return fn;
}

// Split of the partial last wave: R = tiles % units tiles would keep R of the `units` CTA (pairs) busy for a whole tile
// time.  Cutting each of them into s column sub-tiles costs about (0.25 + 0.75 / s) of a tile time per pass (the A rows
// are staged again by every sub-tile; only W and the MMAs shrink), ceil(R * s / units) passes.
int tail_split_shift(int tiles, int units, int bn, int ncta) {
    static const int mode = [] { const char * e = std::getenv("MIO_TTS_B200_GEMM_SPLIT"); return e ? std::atoi(e) : -1; }();
    const int r = tiles % units;
    if (r == 0 || mode == 0) return 0;
    int best = 0;
    float best_cost = 1.0f;
    for (int sh = 1; sh <= 2; ++sh) {
        const int s = 1 << sh;
        if ((bn >> sh) / ncta < B_TAIL_BOX_ROWS || (bn >> sh) < 64) break;
        const float cost = (float) ((r * s + units - 1) / units) * (0.25f + 0.75f / (float) s);
        if (cost < best_cost - 0.05f) { best_cost = cost; best = sh; }
    }
    if (mode > 0 && mode <= 2) best = ((bn >> mode) / ncta >= B_TAIL_BOX_ROWS && (bn >> mode) >= 64) ? mode : 0;   // forced (tests)
    return best;
}

template <int BN, int EPI, int EW, int NCTA> void launch_t(const gemm_args & a0, int num_sms, cudaStream_t st) {
    using C = tile_cfg<BN, EPI, EW, NCTA>;
    static bool configured[64] = {false};   // per-device function attribute
    int dev = 0;
    MIO_CUDA(cudaGetDevice(&dev));
    if (!configured[dev & 63]) {
        MIO_CUDA(cudaFuncSetAttribute(gemm_tcgen05_kernel<BN, EPI, EW, NCTA>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                      C::SMEM_BYTES));
        configured[dev & 63] = true;
    }
    const int tiles = ((a0.M + BLOCK_M * NCTA - 1) / (BLOCK_M * NCTA)) * ((a0.N + BN - 1) / BN);
    const int units = num_sms / NCTA;
    gemm_args a = a0;
    a.split_shift = tail_split_shift(tiles, units, BN, NCTA);
    a.split_from = a.split_shift ? tiles - tiles % units : tiles;
    const int items = a.split_from + ((tiles - a.split_from) << a.split_shift);
    cudaLaunchConfig_t cfg;
    std::memset(&cfg, 0, sizeof(cfg));
    cfg.gridDim = dim3((unsigned) (NCTA * (items < units ? items : units)));
    cfg.blockDim = dim3(C::THREADS);
    cfg.dynamicSmemBytes = C::SMEM_BYTES;
    cfg.stream = st;
    cudaLaunchAttribute attr[2];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = NCTA; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
    attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[1].val.programmaticStreamSerializationAllowed = pdl_enabled() ? 1 : 0;
    cfg.attrs = attr;
    cfg.numAttrs = 2;
    MIO_CUDA(cudaLaunchKernelEx(&cfg, gemm_tcgen05_kernel<BN, EPI, EW, NCTA>, a));
    count_launch();
}

template <int BN, int EW, int NCTA> void launch_bn(const gemm_args & a, int num_sms, cudaStream_t st) {
    switch (a.epilogue) {
        case EPI_STORE32: launch_t<BN, EPI_STORE32, EW, NCTA>(a, num_sms, st); break;
        case EPI_STORE16: launch_t<BN, EPI_STORE16, EW, NCTA>(a, num_sms, st); break;
        case EPI_SWIGLU16: launch_t<BN, EPI_SWIGLU16, EW, NCTA>(a, num_sms, st); break;
        case EPI_ROPE16: launch_t<BN, EPI_ROPE16, EW, NCTA>(a, num_sms, st); break;
        default: throw cuda_error("gemm: bad epilogue");
    }
}

} // namespace

namespace { thread_local int t_traversal_reverse = 0; }
int traversal_reverse() { return t_traversal_reverse; }
void set_traversal_reverse(int rev) {
    static const bool on = [] { const char * e = std::getenv("MIO_TTS_B200_PINGPONG"); return !(e && e[0] == '0'); }();
    t_traversal_reverse = on ? (rev ? 1 : 0) : 0;
}

bool pdl_enabled() {
    static int v = -1;
    if (v < 0) { const char * e = std::getenv("MIO_TTS_B200_PDL"); v = (e && e[0] == '0') ? 0 : 1; }
    return v == 1;
}

int gemm_block_n(int N) { return N >= 256 ? 256 : 128; }
int gemm_w_box_rows() { return B_BOX_ROWS; }
int gemm_w_tail_box_rows() { return B_TAIL_BOX_ROWS; }

CUtensorMap make_tmap_16(const void * ptr, uint64_t rows, uint64_t cols, uint32_t box_rows) {
    if ((cols % 8) != 0) throw cuda_error("tensor map: row pitch must be a multiple of 16 bytes");
    CUtensorMap m;
    const cuuint64_t dims[2] = {cols, rows};
    const cuuint64_t strides[1] = {cols * 2};
    const cuuint32_t box[2] = {(cuuint32_t) BLOCK_K, box_rows};   // W maps: box_rows = gemm_w_box_rows()
    const cuuint32_t estr[2] = {1, 1};
    const CUresult r = get_encode_fn()(&m, CU_TENSOR_MAP_DATA_TYPE_UINT16, 2, const_cast<void *>(ptr), dims, strides,
                                       box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                                       CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) throw cuda_error("cuTensorMapEncodeTiled failed: " + std::to_string((int) r));
    return m;
}

CUtensorMap make_tmap_out(const void * ptr, uint64_t rows, uint64_t cols, uint64_t ld_elems, int elem_bytes) {
    if (elem_bytes != 2 && elem_bytes != 4) throw cuda_error("output tensor map: element size must be 2 or 4");
    if (((uintptr_t) ptr & 15) != 0 || (ld_elems * elem_bytes) % 16 != 0)
        throw cuda_error("output tensor map: base and row pitch must be multiples of 16 bytes");
    CUtensorMap m;
    const cuuint64_t dims[2] = {cols, rows};
    const cuuint64_t strides[1] = {ld_elems * (uint64_t) elem_bytes};
    const cuuint32_t box[2] = {32, 32};    // one epilogue chunk: 32 rows x 32 elements (128 B fp32 / 64 B 16-bit rows)
    const cuuint32_t estr[2] = {1, 1};
    const CUresult r = get_encode_fn()(&m, elem_bytes == 4 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_UINT16, 2,
                                       const_cast<void *>(ptr), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                                       elem_bytes == 4 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_64B,
                                       CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) throw cuda_error("cuTensorMapEncodeTiled (output) failed: " + std::to_string((int) r));
    return m;
}

void gemm_launch(const gemm_args & a, int num_sms, cudaStream_t st) {
    if (a.M <= 0 || a.N <= 0 || a.K <= 0) return;
    if (a.epilogue == EPI_ROPE16 && (a.head_dim % 32) != 0) throw cuda_error("gemm: rope epilogue needs head_dim % 32 == 0");
    if (a.res != nullptr && a.res != a.out) throw cuda_error("gemm: the residual must be the (in-place) output");
    if (a.act == 3 && (a.snake_a == nullptr || a.snake_ib == nullptr || a.bias == nullptr || a.gate != nullptr || a.res != nullptr ||
                       a.gn_stats != nullptr || (a.epilogue != EPI_STORE32 && a.epilogue != EPI_STORE16)))
        throw cuda_error("gemm: the SnakeBeta epilogue needs bias and the per-column alpha / 1/beta vectors, and no gate, residual or statistics");
    static const bool use_red = [] { const char * e = std::getenv("MIO_TTS_B200_GEMM_RED"); return !(e && e[0] == '0'); }();
    static int ew = -1, ncta = -1;
    if (ew < 0) {   // epilogue warps per TMEM lane quadrant
        const char * e = std::getenv("MIO_TTS_B200_GEMM_EW");
        ew = e ? std::atoi(e) : 2;
        if (ew != 1 && ew != 2) ew = 2;
        const char * c = std::getenv("MIO_TTS_B200_GEMM_CTA");   // 1 = single-CTA tiles everywhere (A/B testing)
        ncta = (c && std::atoi(c) == 1) ? 1 : 2;
    }
    gemm_args g = a;
    static const int dbg = [] { const char * e = std::getenv("MIO_TTS_B200_GEMM_DBG"); return e ? std::atoi(e) : 0; }();
    g.dbg = dbg;       // 1: timing experiment only -- epilogues release the accumulators without reading them (wrong results)
    g.reverse_m = traversal_reverse();
    if (a.epilogue == EPI_STORE32_LN) {
        if (a.N != 512 || a.taps != 1 || a.res == nullptr || a.ln_a == nullptr || a.ln_b == nullptr || (a.K % BLOCK_K) != 0 ||
            a.gn_stats != nullptr || (a.gate != nullptr && a.row_utt == nullptr))
            throw cuda_error("gemm: the LayerNorm-fused epilogue needs N = 512, one tap, the in-place residual and LN parameters");
        static const bool ln_pf = [] { const char * e = std::getenv("MIO_TTS_B200_LN_PREFETCH"); return !(e && e[0] == '0'); }();
        g.ln_prefetch = ln_pf ? 1 : 0;
        launch_ln(g, num_sms, st);
        return;
    }
    if (g.gn_stats != nullptr && (a.epilogue != EPI_STORE32 || a.row_utt == nullptr || a.bias == nullptr || a.gate != nullptr ||
                                  !gemm_gn_stats_ok(a.N, a.gn_groups)))
        throw cuda_error("gemm: GroupNorm statistics need the fp32-store epilogue with bias, row_utt, no gate, and a supported group width");
    // the statistics are taken over the FINAL values, so the residual has to pass through the SM (no L2 reduce-add)
    g.red_add = (use_red && a.res != nullptr && a.epilogue == EPI_STORE32 && g.gn_stats == nullptr) ? 1 : 0;
    if (ncta == 2 && astat_applies(g)) { launch_astat(g, num_sms, st); return; }
    if (gemm_block_n(a.N) == 256) {
        if (ncta == 2) { if (ew == 1) launch_bn<256, 1, 2>(g, num_sms, st); else launch_bn<256, 2, 2>(g, num_sms, st); }
        else { if (ew == 1) launch_bn<256, 1, 1>(g, num_sms, st); else launch_bn<256, 2, 1>(g, num_sms, st); }
    } else {
        if (ew == 1) launch_bn<128, 1, 1>(g, num_sms, st); else launch_bn<128, 2, 1>(g, num_sms, st);
    }
}

} // namespace mio
