// tcgen05 / TMEM / TMA GEMM for the MioCodec decoder's dense contractions (sm_100a only).
//
//   D[M, N] = sum_{tap < taps} A[row + tap - taps/2, 0:K] * W[N, tap*K : (tap+1)*K]^T      (fp32 accumulate)
//
// taps = 1: the linears of transformer_block_eval / linear()  (/root/reference/src/miocodec-decoder.cpp:128-134,
//           1190-1291), the x2 transposed conv as one [.,2*C] GEMM (:1368-1372) and the iSTFT head (:1460).
// taps = 3: ggml_conv_1d_ph k=3 of resnet_block_eval (:1293-1330) as an implicit GEMM: three row-shifted TMA
//           loads of the same [rows, C] activation; zero gap rows of the row layout give the "same" padding.
//
// Structure (one persistent CTA per SM, 192 threads):
//   warp 0      TMA producer: cp.async.bulk.tensor.2d -> 128B-swizzled smem stages, mbarrier complete_tx
//   warp 1      allocates TMEM; one lane issues tcgen05.mma.cta_group::1.kind::f16 (M=128, N=BLOCK_N, K=16),
//               tcgen05.commit releases smem stages / publishes the accumulator
//   warps 2..5  epilogue: tcgen05.ld 32x32b.x32 -> registers -> fused bias / gate / residual / SwiGLU / RoPE
//               -> vectorised global stores.  Two TMEM accumulator buffers overlap epilogue(i) with MMA(i+1).
#include "common.h"

#include <cuda_bf16.h>
#include <cuda_fp16.h>

#include <atomic>
#include <cstdlib>
#include <mutex>

namespace mio {

std::atomic<unsigned long long> g_launch_count{0};

namespace {

constexpr int BLOCK_M = 128;
constexpr int BLOCK_K = 64;      // 16-bit elements -> 128-byte rows = one SWIZZLE_128B atom
constexpr int UMMA_K = 16;
constexpr int EPI_LD = 33;      // padded row of the epilogue transpose tile (conflict-free scalar STS/LDS)

template <int BN, int EW> struct tile_cfg {   // EW = epilogue warps per TMEM lane quadrant
    static constexpr int THREADS = 64 + 128 * EW;
    static constexpr int A_BYTES = BLOCK_M * BLOCK_K * 2;
    static constexpr int B_BYTES = BN * BLOCK_K * 2;
    static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
    static constexpr int STAGES = BN == 256 ? (EW >= 4 ? 3 : 4) : (EW >= 4 ? 4 : 6);
    static constexpr int TMEM_COLS = 2 * BN;
    static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + 1024 /*alignment slack*/ + 256 /*barriers*/ +
                                      4 * EW * 32 * EPI_LD * 4 /*epilogue transpose tiles*/;
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
__device__ __forceinline__ void tcgen05_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tcgen05_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tcgen05_commit(uint32_t bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void tcgen05_mma_f16(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc,
                                                uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
        : "memory");
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
    __half2 v = __floats2half2_rn(a, b);
    return *reinterpret_cast<uint32_t *>(&v);
}

__device__ __forceinline__ float silu_f(float x) { return x / (1.0f + __expf(-x)); }

// ---- epilogue of one 32x32 accumulator chunk (one warp: TMEM lane quadrant x 32 columns) ----------------
// tcgen05.ld hands each thread one ROW (32 consecutive columns).  Storing that directly makes every warp-wide
// access touch 32 different 128-byte lines, so the chunk is first transposed through a padded smem tile
// (stride 36 floats: conflict-free for both the row-wise STS.128 and the LDS.128 below) and then processed in a
// coalesced layout: lane -> (row = lane/8 + 4u, 4 consecutive columns), i.e. 4 full 128-byte row segments per
// warp instruction for the residual / gate / bias loads and the stores.

// Per-chunk operands in the coalesced layout, issued BEFORE the TMEM load so their latency overlaps it:
// the residual rows (EPI_STORE32) or the RoPE (cos,sin) pairs of the rows' positions (EPI_ROPE16).
// rmeta[u] = utterance (gate) or position (rope) of unit row u, loaded once per tile.
template <int EPI>
__device__ __forceinline__ void epilogue_prefetch(const gemm_args & a, int row0, int lane, int n, const int (&rmeta)[8],
                                                  float4 (&pre)[8]) {
    const int col = n + (lane & 7) * 4;
    if constexpr (EPI == EPI_STORE32) {
        const bool ok = a.res != nullptr && (a.ldr & 3) == 0 && col + 4 <= a.N;
#pragma unroll
        for (int u = 0; u < 8; ++u) {
            const int r = row0 + (lane >> 3) + 4 * u;
            pre[u] = (ok && r < a.M) ? *reinterpret_cast<const float4 *>(a.res + (size_t) r * a.ldr + col)
                                     : make_float4(0.f, 0.f, 0.f, 0.f);
        }
    } else if constexpr (EPI == EPI_ROPE16) {
        const int half = a.head_dim >> 1;
        const int pair = (col % a.head_dim) >> 1;
#pragma unroll
        for (int u = 0; u < 8; ++u)
            pre[u] = col < a.rope_cols ? __ldg(reinterpret_cast<const float4 *>(a.rope + (size_t) rmeta[u] * half + pair))
                                       : make_float4(1.f, 0.f, 1.f, 0.f);
    }
}

template <int EPI>
__device__ __forceinline__ void epilogue_chunk(const gemm_args & a, int row0, int lane, int n, uint32_t (&raw)[32],
                                               float * stage, const int (&rmeta)[8], const float4 (&pre)[8]) {
#pragma unroll
    for (int j = 0; j < 32; ++j) stage[lane * EPI_LD + j] = __uint_as_float(raw[j]);
    __syncwarp();
    const int c = (lane & 7) * 4;
    const int col = n + c;
#pragma unroll
    for (int u = 0; u < 8; ++u) {
        const int rr = (lane >> 3) + 4 * u;
        const int r = row0 + rr;
        if (r >= a.M || col >= a.N) continue;
        const float * sp = stage + rr * EPI_LD + c;
        float x[4] = {sp[0], sp[1], sp[2], sp[3]};
        const float4 v = make_float4(x[0], x[1], x[2], x[3]);
        const bool full = col + 4 <= a.N;

        if constexpr (EPI == EPI_STORE32) {
            float * orow = (float *) a.out + (size_t) r * a.ldo + col;
            const float * gate = a.gate ? a.gate + (size_t) max(rmeta[u], 0) * a.gate_ld + col : nullptr;
            const float * res = a.res ? a.res + (size_t) r * a.ldr + col : nullptr;
            if (full && (a.ldo & 3) == 0 && (!res || (a.ldr & 3) == 0)) {
                float4 o = v;
                if (a.bias) {
                    const float4 b = __ldg(reinterpret_cast<const float4 *>(a.bias + col));
                    o.x += b.x; o.y += b.y; o.z += b.z; o.w += b.w;
                }
                if (gate) {
                    const float4 g = __ldg(reinterpret_cast<const float4 *>(gate));
                    o.x *= g.x; o.y *= g.y; o.z *= g.z; o.w *= g.w;
                }
                if (res) {
                    const float4 q = pre[u];
                    o.x += q.x; o.y += q.y; o.z += q.z; o.w += q.w;
                }
                *reinterpret_cast<float4 *>(orow) = o;
            } else {
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                    if (col + e < a.N) {
                        float o = x[e];
                        if (a.bias) o += a.bias[col + e];
                        if (gate) o *= gate[e];
                        if (res) o += res[e];
                        orow[e] = o;
                    }
                }
            }
        } else if constexpr (EPI == EPI_STORE16 || EPI == EPI_ROPE16) {
            if (a.bias) {
#pragma unroll
                for (int e = 0; e < 4; ++e) if (col + e < a.N) x[e] += a.bias[col + e];
            }
            if constexpr (EPI == EPI_ROPE16) {
                if (col < a.rope_cols) {   // q and k columns; (col, col+1), (col+2, col+3) are rotation pairs
                    const float4 cs = pre[u];   // (cos,sin) x 2
                    const float s = col < a.q_cols ? a.q_scale : 1.0f;
                    const float x0 = x[0], x1 = x[1], y0 = x[2], y1 = x[3];
                    x[0] = (x0 * cs.x - x1 * cs.y) * s;
                    x[1] = (x0 * cs.y + x1 * cs.x) * s;
                    x[2] = (y0 * cs.z - y1 * cs.w) * s;
                    x[3] = (y0 * cs.w + y1 * cs.z) * s;
                }
            }
            uint16_t * orow = (uint16_t *) a.out + (size_t) r * a.ldo + col;
            if (full && (a.ldo & 3) == 0) {
                *reinterpret_cast<uint2 *>(orow) = make_uint2(pack16(x[0], x[1], a.out_dtype), pack16(x[2], x[3], a.out_dtype));
            } else {
#pragma unroll
                for (int e = 0; e < 4; ++e)
                    if (col + e < a.N) orow[e] = (uint16_t) (pack16(x[e], 0.0f, a.out_dtype) & 0xFFFF);
            }
        } else {   // EPI_SWIGLU16: columns (2j, 2j+1) = (gate_j, up_j); N is even
            uint16_t * orow = (uint16_t *) a.out + (size_t) r * a.ldo + (col >> 1);
            const float h0 = silu_f(x[0]) * x[1];
            const float h1 = silu_f(x[2]) * x[3];
            if (full && (a.ldo & 1) == 0) {
                *reinterpret_cast<uint32_t *>(orow) = pack16(h0, h1, a.out_dtype);
            } else {
                orow[0] = (uint16_t) (pack16(h0, 0.0f, a.out_dtype) & 0xFFFF);
                if (col + 3 < a.N) orow[1] = (uint16_t) (pack16(h1, 0.0f, a.out_dtype) & 0xFFFF);
            }
        }
    }
    __syncwarp();
}

template <int BN, int EPI, int EW>
__global__ void __launch_bounds__(64 + 128 * EW, 1) gemm_tcgen05_kernel(const __grid_constant__ gemm_args a) {
    using C = tile_cfg<BN, EW>;
    extern __shared__ uint8_t smem_raw[];
    const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    const uint32_t bar_base = smem_base + C::STAGES * C::STAGE_BYTES;
    // barriers: full[STAGES], empty[STAGES], tmem_full[2], tmem_empty[2]; then the TMEM base address
    auto full_bar = [&](int s) { return bar_base + 8u * s; };
    auto empty_bar = [&](int s) { return bar_base + 8u * (C::STAGES + s); };
    auto tfull_bar = [&](int s) { return bar_base + 8u * (2 * C::STAGES + s); };
    auto tempty_bar = [&](int s) { return bar_base + 8u * (2 * C::STAGES + 2 + s); };
    const uint32_t tmem_slot = bar_base + 8u * (2 * C::STAGES + 4);
    volatile uint32_t * tmem_slot_ptr =
        reinterpret_cast<volatile uint32_t *>(smem_raw + (tmem_slot - smem_u32(smem_raw)));

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;

    const int num_m_tiles = (a.M + BLOCK_M - 1) / BLOCK_M;
    const int num_n_tiles = (a.N + BN - 1) / BN;
    const int num_tiles = num_m_tiles * num_n_tiles;
    const int k_blocks = (a.K + BLOCK_K - 1) / BLOCK_K;
    const int k_iters = k_blocks * a.taps;

    if (threadIdx.x == 0) {
        for (int s = 0; s < C::STAGES; ++s) { mbar_init(full_bar(s), 1); mbar_init(empty_bar(s), 1); }
        for (int s = 0; s < 2; ++s) { mbar_init(tfull_bar(s), 1); mbar_init(tempty_bar(s), 4 * EW); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot),
                     "r"((uint32_t) C::TMEM_COLS)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tcgen05_fence_before();
    __syncthreads();
    tcgen05_fence_after();
    const uint32_t tmem_base = *tmem_slot_ptr;

    if (warp == 0) {
        // ===== TMA producer =====
        if (lane == 0) {
            uint32_t kit = 0;
            for (int t = blockIdx.x; t < num_tiles; t += gridDim.x) {
                const int m0 = (t / num_n_tiles) * BLOCK_M;
                const int n0 = (t % num_n_tiles) * BN;
                for (int tap = 0; tap < a.taps; ++tap) {
                    const int arow = m0 + tap - (a.taps >> 1);
                    for (int kb = 0; kb < k_blocks; ++kb, ++kit) {
                        const int s = kit % C::STAGES;
                        mbar_wait(empty_bar(s), ((kit / C::STAGES) & 1) ^ 1);
                        mbar_expect_tx(full_bar(s), C::STAGE_BYTES);
                        const uint32_t sa = smem_base + s * C::STAGE_BYTES;
                        tma_load_2d(sa, &a.map_a, full_bar(s), kb * BLOCK_K, arow);
                        tma_load_2d(sa + C::A_BYTES, &a.map_b, full_bar(s), tap * a.K + kb * BLOCK_K, n0);
                    }
                }
            }
        }
    } else if (warp == 1) {
        // ===== MMA issuer =====
        if (lane == 0) {
            // instruction descriptor (cute::UMMA::InstrDescriptor): c=f32 [4,6), a/b format [7,10)/[10,13),
            // K-major a/b (bits 15,16 = 0), N>>3 [17,23), M>>4 [24,29)
            const uint32_t idesc = (1u << 4) | ((uint32_t) a.dtype << 7) | ((uint32_t) a.dtype << 10) |
                                   ((uint32_t) (BN >> 3) << 17) | ((uint32_t) (BLOCK_M >> 4) << 24);
            uint32_t kit = 0, it = 0;
            for (int t = blockIdx.x; t < num_tiles; t += gridDim.x, ++it) {
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
                        tcgen05_mma_f16(tmem_d, da + 2 * k, db + 2 * k, idesc, (ki | k) != 0);
                    }
                    tcgen05_commit(empty_bar(s));   // frees the smem stage once these MMAs retire
                }
                tcgen05_commit(tfull_bar(acc));     // accumulator complete
            }
        }
    } else {
        // ===== epilogue warps =====
        const int quad = warp & 3;                   // TMEM lane quadrant this warp may access
        const int ew = warp - 2;                     // epilogue warp index; column group = ew / 4
        const int cgrp = ew >> 2;
        float * stage = reinterpret_cast<float *>(smem_raw + (bar_base + 256 - smem_u32(smem_raw))) + ew * (32 * EPI_LD);
        uint32_t it = 0;
        for (int t = blockIdx.x; t < num_tiles; t += gridDim.x, ++it) {
            const int m0 = (t / num_n_tiles) * BLOCK_M;
            const int n0 = (t % num_n_tiles) * BN;
            const int acc = it & 1;
            const int row0 = m0 + quad * 32;
            int rmeta[8];
            {
                const int * mp = EPI == EPI_ROPE16 ? a.row_pos : (EPI == EPI_STORE32 && a.gate ? a.row_utt : nullptr);
#pragma unroll
                for (int u = 0; u < 8; ++u) {
                    const int r = row0 + (lane >> 3) + 4 * u;
                    rmeta[u] = (mp != nullptr && r < a.M) ? __ldg(mp + r) : 0;
                }
            }
            const uint32_t taddr = tmem_base + ((uint32_t) (quad * 32) << 16) + acc * BN;
            mbar_wait(tfull_bar(acc), (it >> 1) & 1);
            tcgen05_fence_after();
#pragma unroll 1
            for (int c = cgrp; c < BN / 32; c += EW) {
                const int n = n0 + c * 32;
                if (n >= a.N) break;                 // warp-uniform
                float4 pre[8];
                epilogue_prefetch<EPI>(a, row0, lane, n, rmeta, pre);
                uint32_t raw[32];
                tmem_ld_32x32b_x32(taddr + c * 32, raw);
                tmem_ld_wait();
                if (row0 < a.M) epilogue_chunk<EPI>(a, row0, lane, n, raw, stage, rmeta, pre);   // warp-uniform
            }
            tcgen05_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(tempty_bar(acc));
        }
    }

    tcgen05_fence_before();
    __syncthreads();
    if (warp == 1) {
        tcgen05_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base),
                     "r"((uint32_t) C::TMEM_COLS)
                     : "memory");
    }
}

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

template <int BN, int EPI, int EW> void launch_t(const gemm_args & a, int num_sms, cudaStream_t st) {
    using C = tile_cfg<BN, EW>;
    static bool configured[64] = {false};   // per-device function attribute
    int dev = 0;
    MIO_CUDA(cudaGetDevice(&dev));
    if (!configured[dev & 63]) {
        MIO_CUDA(cudaFuncSetAttribute(gemm_tcgen05_kernel<BN, EPI, EW>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                      C::SMEM_BYTES));
        configured[dev & 63] = true;
    }
    const int tiles = ((a.M + BLOCK_M - 1) / BLOCK_M) * ((a.N + BN - 1) / BN);
    const int grid = tiles < num_sms ? tiles : num_sms;
    gemm_tcgen05_kernel<BN, EPI, EW><<<grid, C::THREADS, C::SMEM_BYTES, st>>>(a);
    MIO_CUDA(cudaGetLastError());
    count_launch();
}

template <int BN, int EW> void launch_bn(const gemm_args & a, int num_sms, cudaStream_t st) {
    switch (a.epilogue) {
        case EPI_STORE32: launch_t<BN, EPI_STORE32, EW>(a, num_sms, st); break;
        case EPI_STORE16: launch_t<BN, EPI_STORE16, EW>(a, num_sms, st); break;
        case EPI_SWIGLU16: launch_t<BN, EPI_SWIGLU16, EW>(a, num_sms, st); break;
        case EPI_ROPE16: launch_t<BN, EPI_ROPE16, EW>(a, num_sms, st); break;
        default: throw cuda_error("gemm: bad epilogue");
    }
}

} // namespace

int gemm_block_n(int N) { return N >= 256 ? 256 : 128; }

CUtensorMap make_tmap_16(const void * ptr, uint64_t rows, uint64_t cols, uint32_t box_rows) {
    if ((cols % 8) != 0) throw cuda_error("tensor map: row pitch must be a multiple of 16 bytes");
    CUtensorMap m;
    const cuuint64_t dims[2] = {cols, rows};
    const cuuint64_t strides[1] = {cols * 2};
    const cuuint32_t box[2] = {(cuuint32_t) BLOCK_K, box_rows};
    const cuuint32_t estr[2] = {1, 1};
    const CUresult r = get_encode_fn()(&m, CU_TENSOR_MAP_DATA_TYPE_UINT16, 2, const_cast<void *>(ptr), dims, strides,
                                       box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                                       CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) throw cuda_error("cuTensorMapEncodeTiled failed: " + std::to_string((int) r));
    return m;
}

void gemm_launch(const gemm_args & a, int num_sms, cudaStream_t st) {
    if (a.M <= 0 || a.N <= 0 || a.K <= 0) return;
    if (a.epilogue == EPI_ROPE16 && (a.head_dim % 32) != 0) throw cuda_error("gemm: rope epilogue needs head_dim % 32 == 0");
    static int ew = -1;
    if (ew < 0) {   // epilogue warps per TMEM quadrant: memory-level parallelism of the residual read-modify-write
        const char * e = std::getenv("MIO_TTS_B200_GEMM_EW");
        ew = e ? std::atoi(e) : 4;
        if (ew != 1 && ew != 2 && ew != 4) ew = 4;
    }
    if (gemm_block_n(a.N) == 256) {
        if (ew == 1) launch_bn<256, 1>(a, num_sms, st);
        else if (ew == 2) launch_bn<256, 2>(a, num_sms, st);
        else launch_bn<256, 4>(a, num_sms, st);
    } else {
        if (ew == 1) launch_bn<128, 1>(a, num_sms, st);
        else if (ew == 2) launch_bn<128, 2>(a, num_sms, st);
        else launch_bn<128, 4>(a, num_sms, st);
    }
}

} // namespace mio
