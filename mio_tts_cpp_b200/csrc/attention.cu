// Banded (local-window) multi-head attention for the MioCodec transformer blocks.
//
// Reference: softmax(Q K^T / sqrt(hd) + mask) V with mask 0 iff |q - k| <= window/2 else -inf, materialised as a
// dense N x N matrix (/root/reference/src/miocodec-decoder.cpp:1251-1257, mask fill :1501-1525).  Here only the
// band is computed: a CTA owns one 64-query tile of one head of one utterance and the <=128 keys it can see;
// QK^T and PV run on tensor cores (mma.sync m16n8k16, fp32 accumulate) with the softmax in registers
// (S accumulator fragments are re-used as the A operand of PV).  RoPE and the 1/sqrt(hd) scale were already
// applied to q/k by the QKV GEMM epilogue.  A generic SIMT kernel covers other head sizes / wider windows.
#include "common.h"

#include <cuda_bf16.h>
#include <cuda_fp16.h>

#include <algorithm>
#include <cstdlib>
#include <cstring>

namespace mio {
namespace {

constexpr int TQ = 64;        // queries per tile
constexpr int TK = 128;       // keys visible to a tile (TQ + 2*32)
constexpr int HD = 64;        // head dim of the fast path
constexpr int LDS = 72;       // padded smem row (elements): 144 B keeps ldmatrix / fragment loads conflict-free

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
__global__ void __launch_bounds__(128) attention_band_kernel(const uint16_t * __restrict__ qkv,
                                                             uint16_t * __restrict__ out, int dim, int half_window,
                                                             const row_tile * __restrict__ tiles) {
    MIO_PDL_PROLOGUE();
    __shared__ __align__(16) uint16_t Qs[TQ * LDS];
    __shared__ __align__(16) uint16_t Ks[TK * LDS];
    __shared__ __align__(16) uint16_t Vs[TK * LDS];

    const row_tile t = tiles[blockIdx.x];
    const int head = blockIdx.y;
    const int ld = 3 * dim;
    const int kpos0 = t.pos0 - 32;     // position of key index 0

    // ---- stage Q, K, V tiles with cp.async (16-byte chunks in flight together; rows outside the utterance are
    //      zero-filled through src-size 0) ----
    auto cp16 = [](uint16_t * dst, const uint16_t * src, bool valid) {
        const uint32_t d = (uint32_t) __cvta_generic_to_shared(dst);
        const int n = valid ? 16 : 0;
        asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(d), "l"(src), "r"(n) : "memory");
    };
    for (int c = threadIdx.x; c < TQ * 8; c += blockDim.x) {
        const int i = c >> 3, ch = c & 7;
        const bool ok = i < t.n;
        cp16(Qs + i * LDS + ch * 8, qkv + (size_t) (t.row0 + (ok ? i : 0)) * ld + head * HD + ch * 8, ok);
    }
    for (int c = threadIdx.x; c < TK * 8; c += blockDim.x) {
        const int j = c >> 3, ch = c & 7;
        const int pos = kpos0 + j;
        const bool ok = pos >= 0 && pos < t.utt_len;
        const uint16_t * p = qkv + (size_t) (t.utt_row0 + (ok ? pos : 0)) * ld + head * HD + ch * 8;
        cp16(Ks + j * LDS + ch * 8, p + dim, ok);
        cp16(Vs + j * LDS + ch * 8, p + 2 * dim, ok);
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    __syncthreads();

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int g = lane >> 2, tq = lane & 3;
    const int q0 = warp * 16;          // first query of this warp; its keys are indices [q0, q0 + 80)

    // ---- S = Q K^T (16 x 80); A and B fragments through ldmatrix (4 + 20 instructions instead of 96 scalar loads) ----
    const int mi = lane >> 3, mr = lane & 7;
    uint32_t qa[4][4];
#pragma unroll
    for (int kk = 0; kk < 4; ++kk) {
        const uint32_t addr = (uint32_t) __cvta_generic_to_shared(Qs + (q0 + (mi & 1) * 8 + mr) * LDS + kk * 16 + (mi >> 1) * 8);
        asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
                     : "=r"(qa[kk][0]), "=r"(qa[kk][1]), "=r"(qa[kk][2]), "=r"(qa[kk][3]) : "r"(addr));
    }
    float s[10][4];
#pragma unroll
    for (int nt = 0; nt < 10; ++nt) {
        s[nt][0] = s[nt][1] = s[nt][2] = s[nt][3] = 0.0f;
#pragma unroll
        for (int kp = 0; kp < 2; ++kp) {      // head-dim columns [32 kp, 32 kp + 32): fragments (b0, b1) of kk = 2 kp and 2 kp + 1
            const uint32_t addr = (uint32_t) __cvta_generic_to_shared(Ks + (q0 + nt * 8 + mr) * LDS + kp * 32 + mi * 8);
            uint32_t b0, b1, b2, b3;
            asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
                         : "=r"(b0), "=r"(b1), "=r"(b2), "=r"(b3) : "r"(addr));
            mma16816<BF16>(s[nt], qa[2 * kp], b0, b1);
            mma16816<BF16>(s[nt], qa[2 * kp + 1], b2, b3);
        }
    }

    // ---- band mask + softmax (rows g and g+8 of this warp's 16 queries) ----
    // In warp-local key index jj = nt*8 + 2*tq + e (key position kpos0 + q0 + jj), row r of the warp sees
    // jj in [r + 32 - hw, r + 32 + hw], clipped to the utterance: one unsigned compare per element.
    const int qi0 = q0 + g, qi1 = q0 + g + 8;             // query index in tile
    const bool qv0 = qi0 < t.n, qv1 = qi1 < t.n;
    const int j_min = -(kpos0 + q0), j_max = t.utt_len - 1 - kpos0 - q0;
    int lo0 = max(g + 32 - half_window, j_min), hi0 = min(g + 32 + half_window, j_max);
    int lo1 = max(g + 40 - half_window, j_min), hi1 = min(g + 40 + half_window, j_max);
    if (!qv0 || hi0 < lo0) { lo0 = 1 << 20; hi0 = lo0; }
    if (!qv1 || hi1 < lo1) { lo1 = 1 << 20; hi1 = lo1; }
    const unsigned span0 = (unsigned) (hi0 - lo0), span1 = (unsigned) (hi1 - lo1);
    const int d0 = 2 * tq - lo0, d1 = 2 * tq - lo1;
    // full 65-wide window, all 16 queries present, no utterance edge inside the 80 keys: tiles 2..7 are unmasked
    const bool interior = half_window == 32 && j_min <= 0 && j_max >= 79 && q0 + 15 < t.n;
    float m0 = -INFINITY, m1 = -INFINITY;
#pragma unroll
    for (int nt = 0; nt < 10; ++nt) {
        if (nt < 2 || nt > 7 || !interior) {
#pragma unroll
            for (int e = 0; e < 2; ++e) {
                const int c = nt * 8 + e;
                s[nt][e] = (unsigned) (c + d0) <= span0 ? s[nt][e] : -INFINITY;
                s[nt][2 + e] = (unsigned) (c + d1) <= span1 ? s[nt][2 + e] : -INFINITY;
            }
        }
        m0 = fmaxf(m0, fmaxf(s[nt][0], s[nt][1]));
        m1 = fmaxf(m1, fmaxf(s[nt][2], s[nt][3]));
    }
    m0 = fmaxf(m0, __shfl_xor_sync(0xffffffffu, m0, 1));
    m0 = fmaxf(m0, __shfl_xor_sync(0xffffffffu, m0, 2));
    m1 = fmaxf(m1, __shfl_xor_sync(0xffffffffu, m1, 1));
    m1 = fmaxf(m1, __shfl_xor_sync(0xffffffffu, m1, 2));
    if (m0 == -INFINITY) m0 = 0.0f;
    if (m1 == -INFINITY) m1 = 0.0f;
    float l0 = 0.0f, l1 = 0.0f;
#pragma unroll
    for (int nt = 0; nt < 10; ++nt) {
#pragma unroll
        for (int e = 0; e < 2; ++e) {
            s[nt][e] = __expf(s[nt][e] - m0);
            s[nt][2 + e] = __expf(s[nt][2 + e] - m1);
            l0 += s[nt][e];
            l1 += s[nt][2 + e];
        }
    }
    l0 += __shfl_xor_sync(0xffffffffu, l0, 1);
    l0 += __shfl_xor_sync(0xffffffffu, l0, 2);
    l1 += __shfl_xor_sync(0xffffffffu, l1, 1);
    l1 += __shfl_xor_sync(0xffffffffu, l1, 2);

    // ---- O = P V (16 x 64), P fragments straight from the S accumulators ----
    float o[8][4];
#pragma unroll
    for (int nd = 0; nd < 8; ++nd) o[nd][0] = o[nd][1] = o[nd][2] = o[nd][3] = 0.0f;
#pragma unroll
    for (int kk = 0; kk < 5; ++kk) {
        uint32_t pa[4];
        pa[0] = pack2<BF16>(s[2 * kk][0], s[2 * kk][1]);
        pa[1] = pack2<BF16>(s[2 * kk][2], s[2 * kk][3]);
        pa[2] = pack2<BF16>(s[2 * kk + 1][0], s[2 * kk + 1][1]);
        pa[3] = pack2<BF16>(s[2 * kk + 1][2], s[2 * kk + 1][3]);
        const int key = q0 + kk * 16 + (mi & 1) * 8 + mr;
#pragma unroll
        for (int nd = 0; nd < 8; nd += 2) {
            const uint32_t addr = (uint32_t) __cvta_generic_to_shared(Vs + key * LDS + (nd + (mi >> 1)) * 8);
            uint32_t b0, b1, b2, b3;
            asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
                         : "=r"(b0), "=r"(b1), "=r"(b2), "=r"(b3)
                         : "r"(addr));
            mma16816<BF16>(o[nd], pa, b0, b1);
            mma16816<BF16>(o[nd + 1], pa, b2, b3);
        }
    }

    // ---- normalise, stage the warp's 16 x 64 tile in its own (now dead) Q rows, store 128-byte rows coalesced ----
    const float inv0 = l0 > 0.0f ? __fdividef(1.0f, l0) : 0.0f;
    const float inv1 = l1 > 0.0f ? __fdividef(1.0f, l1) : 0.0f;
    __syncwarp();
#pragma unroll
    for (int nd = 0; nd < 8; ++nd) {
        *reinterpret_cast<uint32_t *>(Qs + qi0 * LDS + nd * 8 + 2 * tq) = pack2<BF16>(o[nd][0] * inv0, o[nd][1] * inv0);
        *reinterpret_cast<uint32_t *>(Qs + qi1 * LDS + nd * 8 + 2 * tq) = pack2<BF16>(o[nd][2] * inv1, o[nd][3] * inv1);
    }
    __syncwarp();
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int c = lane + 32 * i, row = q0 + (c >> 3), ch = c & 7;
        if (row < t.n) {
            const uint4 v = *reinterpret_cast<const uint4 *>(Qs + row * LDS + ch * 8);
            *reinterpret_cast<uint4 *>(out + (size_t) (t.row0 + row) * dim + head * HD + ch * 8) = v;
        }
    }
}

// =====================================================================================================================
// Band attention, persistent + TMA-fed (the default fast path).
//
// One CTA per SM walks (tile, head) items; tiles are the layout's runs of <= 64 rows inside ONE utterance, counted from
// the utterance's first row, so a row's place in its tile -- and with it the order in which the MMAs sum its keys --
// does not depend on where the utterance sits in the batch (bit-identical alone and in any batch, like every other
// kernel of the path).  Warp 0 is the producer: per item it issues
// five cp.async.bulk.tensor loads -- Q rows [r0, r0+64), K and V rows [r0-32, r0+96) of the head's 64 columns, 128-byte
// rows in SWIZZLE_128B boxes, out-of-range rows zero-filled by the TMA unit -- into a ring of ATT_STAGES x 40 KB stages
// guarded by full/empty mbarriers.  Three consumer groups of four warps take the items round-robin: ldmatrix from the
// swizzled stage (conflict-free without padding), QK^T and PV on mma.sync m16n8k16 with the softmax in registers, the
// stage released as soon as the last V fragment is read, and each warp's 16 x 64 output tile goes back through its
// own swizzled box and ONE lane's TMA store.  Loads of up to ATT_STAGES - 3 items are in flight while three are being
// computed: the kernel streams q/k/v once at HBM speed instead of the load -> compute -> store round per CTA of the
// first version (kept below as attention_band_kernel: MIO_TTS_B200_ATTN=v1).
// =====================================================================================================================
constexpr int ATT_STAGES = 5;
constexpr int ATT_GROUPS = 3;                       // consumer groups of 4 warps
constexpr int ATT_THREADS = 32 + ATT_GROUPS * 128;
constexpr int ATT_Q_BYTES = TQ * 128, ATT_KV_BYTES = TK * 128;
constexpr int ATT_STAGE_BYTES = ATT_Q_BYTES + 2 * ATT_KV_BYTES;          // 40 KB
constexpr int ATT_OUT_BYTES = 16 * 128;                                   // one warp's output box
constexpr int ATT_SMEM_BYTES = ATT_STAGES * ATT_STAGE_BYTES + ATT_GROUPS * 4 * ATT_OUT_BYTES + 128 + 1024;

struct att_args {
    CUtensorMap map_qkv;     // [rows, 3*dim] 16-bit, box 64 cols x 64 rows, SWIZZLE_128B
    CUtensorMap map_out;     // [rows, dim] 16-bit, box 64 cols x 16 rows, SWIZZLE_128B
    int rows, dim, heads, half_window, n_items;
    const row_tile * tiles;
    uint16_t * out;          // for the partial last tile of an utterance (rows past its end belong to someone else)
};

__device__ __forceinline__ uint32_t att_smem_u32(const void * p) { return (uint32_t) __cvta_generic_to_shared(p); }
__device__ __forceinline__ void att_mbar_wait(uint32_t bar, uint32_t parity) {
    uint32_t done = 0;
    long long t0 = 0;
    while (true) {
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(done) : "r"(bar), "r"(parity) : "memory");
        if (done) break;
        if (t0 == 0) t0 = clock64();
        else if (clock64() - t0 > 4000000000ll) __trap();     // a protocol bug must not hang the GPU
    }
}
__device__ __forceinline__ void att_tma_load(uint32_t dst, const CUtensorMap * map, uint32_t bar, int c0, int c1) {
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
                 ::"r"(dst), "l"((uint64_t) map), "r"(bar), "r"(c0), "r"(c1) : "memory");
}
// swizzled address of 16-byte chunk c of 128-byte row r in a 1024-aligned SWIZZLE_128B tile
__device__ __forceinline__ uint32_t att_sw(uint32_t base, int r, int c) { return base + r * 128 + (((c ^ r) & 7) << 4); }

template <bool BF16>
__global__ void __launch_bounds__(ATT_THREADS, 1) attention_band_tma_kernel(const __grid_constant__ att_args a) {
    MIO_PDL_TRIGGER();
    extern __shared__ uint8_t att_smem_raw[];
    const uint32_t base = (att_smem_u32(att_smem_raw) + 1023u) & ~1023u;
    const uint32_t out_base = base + ATT_STAGES * ATT_STAGE_BYTES;
    const uint32_t bar_base = out_base + ATT_GROUPS * 4 * ATT_OUT_BYTES;
    auto full_bar = [&](int s) { return bar_base + 8u * s; };
    auto empty_bar = [&](int s) { return bar_base + 8u * (ATT_STAGES + s); };
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

    if (threadIdx.x == 0) {
        for (int s = 0; s < ATT_STAGES; ++s) {
            asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(full_bar(s)), "r"(1) : "memory");
            asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(empty_bar(s)), "r"(4) : "memory");
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    MIO_PDL_WAIT();

    const int n_local = a.n_items > (int) blockIdx.x ? (a.n_items - (int) blockIdx.x + (int) gridDim.x - 1) / (int) gridDim.x : 0;

    if (warp == 0) {
        // ===== producer =====
        if (lane == 0) {
            for (int i = 0; i < n_local; ++i) {
                const int item = blockIdx.x + i * gridDim.x;
                const int tile = item / a.heads, head = item - tile * a.heads;
                const int r0 = __ldg(&a.tiles[tile].row0), c0 = head * HD;
                const int s = i % ATT_STAGES;
                att_mbar_wait(empty_bar(s), ((i / ATT_STAGES) & 1) ^ 1);
                const uint32_t st = base + s * ATT_STAGE_BYTES;
                asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(full_bar(s)), "r"(ATT_STAGE_BYTES) : "memory");
                att_tma_load(st, &a.map_qkv, full_bar(s), c0, r0);
                att_tma_load(st + ATT_Q_BYTES, &a.map_qkv, full_bar(s), a.dim + c0, r0 - 32);
                att_tma_load(st + ATT_Q_BYTES + 8192, &a.map_qkv, full_bar(s), a.dim + c0, r0 + 32);
                att_tma_load(st + ATT_Q_BYTES + ATT_KV_BYTES, &a.map_qkv, full_bar(s), 2 * a.dim + c0, r0 - 32);
                att_tma_load(st + ATT_Q_BYTES + ATT_KV_BYTES + 8192, &a.map_qkv, full_bar(s), 2 * a.dim + c0, r0 + 32);
            }
        }
        return;
    }

    // ===== consumers =====
    const int cw = warp - 1;                    // 0 .. 4*ATT_GROUPS-1
    const int grp = cw >> 2, w = cw & 3;
    const int q0 = w * 16;                      // this warp's queries in the tile; its keys are tile-key indices [q0, q0 + 80)
    const int g = lane >> 2, tq = lane & 3;
    const int mi = lane >> 3, mr = lane & 7;
    const int hw = a.half_window;
    const uint32_t obuf = out_base + cw * ATT_OUT_BYTES;
    bool store_pending = false;

    for (int i = grp; i < n_local; i += ATT_GROUPS) {
        const int item = blockIdx.x + i * gridDim.x;
        const int tile = item / a.heads, head = item - tile * a.heads;
        const int2 * tp = reinterpret_cast<const int2 *>(a.tiles + tile);              // 24-byte records, 8-byte aligned
        const int2 t01 = __ldg(tp), t23 = __ldg(tp + 1), t45 = __ldg(tp + 2);          // (row0, n) (utt, pos0) (utt_row0, utt_len)
        const int utt_len = t45.y;
        const int r0 = t01.x, tn = t01.y, kpos0 = t23.y - 32;                          // position of tile-key index 0
        // band limits of rows g and g+8 in warp-local key index jj (key position kpos0 + q0 + jj): the row sees
        // jj in [r + 32 - hw, r + 32 + hw], clipped to the utterance
        const int j_min = -(kpos0 + q0), j_max = utt_len - 1 - kpos0 - q0;
        int lo[2], hi[2];
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            lo[h] = max(g + 8 * h + 32 - hw, j_min);
            hi[h] = min(g + 8 * h + 32 + hw, j_max);
            if (q0 + g + 8 * h >= tn || hi[h] < lo[h]) { lo[h] = 1 << 20; hi[h] = lo[h]; }
        }
        // full 65-wide window, all 16 queries present, no utterance edge inside the 80 keys: key tiles 2..7 are unmasked
        const bool interior = hw == 32 && j_min <= 0 && j_max >= 79 && q0 + 15 < tn;

        const int s = i % ATT_STAGES;
        const uint32_t Qs = base + s * ATT_STAGE_BYTES, Ks = Qs + ATT_Q_BYTES, Vs = Ks + ATT_KV_BYTES;
        att_mbar_wait(full_bar(s), (i / ATT_STAGES) & 1);

        // ---- S = Q K^T (16 x 80) ----
        uint32_t qa[4][4];
#pragma unroll
        for (int kk = 0; kk < 4; ++kk) {
            const uint32_t addr = att_sw(Qs, q0 + (mi & 1) * 8 + mr, kk * 2 + (mi >> 1));
            asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
                         : "=r"(qa[kk][0]), "=r"(qa[kk][1]), "=r"(qa[kk][2]), "=r"(qa[kk][3]) : "r"(addr));
        }
        float sc[10][4];
#pragma unroll
        for (int nt = 0; nt < 10; ++nt) {
            sc[nt][0] = sc[nt][1] = sc[nt][2] = sc[nt][3] = 0.0f;
#pragma unroll
            for (int kp = 0; kp < 2; ++kp) {
                const uint32_t addr = att_sw(Ks, q0 + nt * 8 + mr, kp * 4 + mi);
                uint32_t b0, b1, b2, b3;
                asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
                             : "=r"(b0), "=r"(b1), "=r"(b2), "=r"(b3) : "r"(addr));
                mma16816<BF16>(sc[nt], qa[2 * kp], b0, b1);
                mma16816<BF16>(sc[nt], qa[2 * kp + 1], b2, b3);
            }
        }

        // ---- band mask + softmax (rows g and g+8): one unsigned compare per element; tiles 2..7 unmasked when interior ----
        const unsigned span0 = (unsigned) (hi[0] - lo[0]), span1 = (unsigned) (hi[1] - lo[1]);
        const int d0 = 2 * tq - lo[0], d1 = 2 * tq - lo[1];
        float m0 = -INFINITY, m1 = -INFINITY;
#pragma unroll
        for (int nt = 0; nt < 10; ++nt) {
            if (nt < 2 || nt > 7 || !interior) {
#pragma unroll
                for (int e = 0; e < 2; ++e) {
                    const int c = nt * 8 + e;
                    sc[nt][e] = (unsigned) (c + d0) <= span0 ? sc[nt][e] : -INFINITY;
                    sc[nt][2 + e] = (unsigned) (c + d1) <= span1 ? sc[nt][2 + e] : -INFINITY;
                }
            }
            m0 = fmaxf(m0, fmaxf(sc[nt][0], sc[nt][1]));
            m1 = fmaxf(m1, fmaxf(sc[nt][2], sc[nt][3]));
        }
        m0 = fmaxf(m0, __shfl_xor_sync(0xffffffffu, m0, 1));
        m0 = fmaxf(m0, __shfl_xor_sync(0xffffffffu, m0, 2));
        m1 = fmaxf(m1, __shfl_xor_sync(0xffffffffu, m1, 1));
        m1 = fmaxf(m1, __shfl_xor_sync(0xffffffffu, m1, 2));
        if (m0 == -INFINITY) m0 = 0.0f;
        if (m1 == -INFINITY) m1 = 0.0f;
        float l0 = 0.0f, l1 = 0.0f;
#pragma unroll
        for (int nt = 0; nt < 10; ++nt) {
#pragma unroll
            for (int e = 0; e < 2; ++e) {
                sc[nt][e] = __expf(sc[nt][e] - m0);
                sc[nt][2 + e] = __expf(sc[nt][2 + e] - m1);
                l0 += sc[nt][e];
                l1 += sc[nt][2 + e];
            }
        }
        l0 += __shfl_xor_sync(0xffffffffu, l0, 1);
        l0 += __shfl_xor_sync(0xffffffffu, l0, 2);
        l1 += __shfl_xor_sync(0xffffffffu, l1, 1);
        l1 += __shfl_xor_sync(0xffffffffu, l1, 2);

        // ---- O = P V (16 x 64) ----
        float o[8][4];
#pragma unroll
        for (int nd = 0; nd < 8; ++nd) o[nd][0] = o[nd][1] = o[nd][2] = o[nd][3] = 0.0f;
#pragma unroll
        for (int kk = 0; kk < 5; ++kk) {
            uint32_t pa[4];
            pa[0] = pack2<BF16>(sc[2 * kk][0], sc[2 * kk][1]);
            pa[1] = pack2<BF16>(sc[2 * kk][2], sc[2 * kk][3]);
            pa[2] = pack2<BF16>(sc[2 * kk + 1][0], sc[2 * kk + 1][1]);
            pa[3] = pack2<BF16>(sc[2 * kk + 1][2], sc[2 * kk + 1][3]);
            const int key = q0 + kk * 16 + (mi & 1) * 8 + mr;
#pragma unroll
            for (int nd = 0; nd < 8; nd += 2) {
                const uint32_t addr = att_sw(Vs, key, nd + (mi >> 1));
                uint32_t b0, b1, b2, b3;
                asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
                             : "=r"(b0), "=r"(b1), "=r"(b2), "=r"(b3) : "r"(addr));
                mma16816<BF16>(o[nd], pa, b0, b1);
                mma16816<BF16>(o[nd + 1], pa, b2, b3);
            }
        }
        // the stage is free once every lane's fragment loads have been issued and consumed by the MMAs above
        __syncwarp();
        if (lane == 0) asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(empty_bar(s)) : "memory");

        // ---- normalise -> this warp's swizzled output box -> TMA store (rows past the layout are clipped) ----
        const float inv0 = l0 > 0.0f ? __fdividef(1.0f, l0) : 0.0f;
        const float inv1 = l1 > 0.0f ? __fdividef(1.0f, l1) : 0.0f;
        if (store_pending) {
            if (lane == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
            __syncwarp();
        }
#pragma unroll
        for (int nd = 0; nd < 8; ++nd) {
            const uint32_t p0 = pack2<BF16>(o[nd][0] * inv0, o[nd][1] * inv0);
            const uint32_t p1 = pack2<BF16>(o[nd][2] * inv1, o[nd][3] * inv1);
            asm volatile("st.shared.b32 [%0], %1;" ::"r"(att_sw(obuf, g, nd) + 4 * tq), "r"(p0) : "memory");
            asm volatile("st.shared.b32 [%0], %1;" ::"r"(att_sw(obuf, g + 8, nd) + 4 * tq), "r"(p1) : "memory");
        }
        if (q0 + 16 <= tn) {
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            __syncwarp();
            if (lane == 0) {
                asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];"
                             ::"l"((uint64_t) &a.map_out), "r"(obuf), "r"(head * HD), "r"(r0 + q0) : "memory");
                asm volatile("cp.async.bulk.commit_group;" ::: "memory");
            }
        } else if (q0 < tn) {
            // the utterance ends inside this warp's 16 rows: plain 16-byte stores of the rows that exist
            __syncwarp();
#pragma unroll
            for (int it = 0; it < 4; ++it) {
                const int c = lane + 32 * it, row = c >> 3, ch = c & 7;
                if (q0 + row < tn) {
                    uint4 v;
                    asm volatile("ld.shared.v4.b32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(att_sw(obuf, row, ch)));
                    *reinterpret_cast<uint4 *>(a.out + (size_t) (r0 + q0 + row) * a.dim + head * HD + ch * 8) = v;
                }
            }
            __syncwarp();
        }
        store_pending = true;
    }
    if (lane == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
    __syncwarp();
}

// ---- generic fallback: one warp per (query, head); any head_dim, any window ----
template <bool BF16> __device__ __forceinline__ float ld16(const uint16_t * p) {
    if constexpr (BF16) return __bfloat162float(*reinterpret_cast<const __nv_bfloat16 *>(p));
    else return __half2float(*reinterpret_cast<const __half *>(p));
}
template <bool BF16> __device__ __forceinline__ uint16_t st16(float v) {
    if constexpr (BF16) { __nv_bfloat16 h = __float2bfloat16_rn(v); return *reinterpret_cast<uint16_t *>(&h); }
    else { __half h = __float2half_rn(v); return *reinterpret_cast<uint16_t *>(&h); }
}

template <bool BF16>
__global__ void attention_generic_kernel(const uint16_t * __restrict__ qkv, uint16_t * __restrict__ out, int dim,
                                         int head_dim, int half_window, const row_tile * __restrict__ tiles) {
    MIO_PDL_PROLOGUE();
    extern __shared__ float sc[];   // [warps][2*half_window+1]
    const row_tile t = tiles[blockIdx.x];
    const int head = blockIdx.y;
    const int warps = blockDim.x >> 5, warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int ld = 3 * dim;
    float * my = sc + (size_t) warp * (2 * half_window + 1);
    for (int i = warp; i < t.n; i += warps) {
        const int qp = t.pos0 + i;
        const int k_lo = max(0, qp - half_window), k_hi = min(t.utt_len - 1, qp + half_window);
        const int nk = k_hi - k_lo + 1;
        const uint16_t * q = qkv + (size_t) (t.row0 + i) * ld + head * head_dim;
        float mx = -INFINITY;
        for (int j = lane; j < nk; j += 32) {
            const uint16_t * k = qkv + (size_t) (t.utt_row0 + k_lo + j) * ld + dim + head * head_dim;
            float acc = 0.0f;
            for (int d = 0; d < head_dim; ++d) acc += ld16<BF16>(q + d) * ld16<BF16>(k + d);
            my[j] = acc;
            mx = fmaxf(mx, acc);
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
        float sum = 0.0f;
        for (int j = lane; j < nk; j += 32) { const float p = __expf(my[j] - mx); my[j] = p; sum += p; }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
        __syncwarp();
        const float inv = 1.0f / sum;
        for (int d = lane; d < head_dim; d += 32) {
            float acc = 0.0f;
            for (int j = 0; j < nk; ++j)
                acc += my[j] * ld16<BF16>(qkv + (size_t) (t.utt_row0 + k_lo + j) * ld + 2 * dim + head * head_dim + d);
            out[(size_t) (t.row0 + i) * dim + head * head_dim + d] = st16<BF16>(acc * inv);
        }
        __syncwarp();
    }
}

} // namespace

bool attention_tma_ok(int dim, int heads, int window) {
    return attention_mode() != 0 && heads > 0 && dim / heads == HD && window / 2 <= 32 && (dim % 8) == 0;
}

void launch_attention_tma(const CUtensorMap & map_qkv, const CUtensorMap & map_out, void * out16, int dtype, int dim, int heads,
                          int window, int rows, const row_tile * tiles, int n_tiles, int num_sms, cudaStream_t st) {
    if (rows <= 0 || n_tiles <= 0) return;
    att_args a;
    a.map_qkv = map_qkv; a.map_out = map_out;
    a.rows = rows; a.dim = dim; a.heads = heads; a.half_window = window / 2;
    a.n_items = n_tiles * heads;
    a.tiles = tiles; a.out = (uint16_t *) out16;
    static bool configured[64] = {false};
    int dev = 0;
    MIO_CUDA(cudaGetDevice(&dev));
    if (!configured[dev & 63]) {
        MIO_CUDA(cudaFuncSetAttribute(attention_band_tma_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, ATT_SMEM_BYTES));
        MIO_CUDA(cudaFuncSetAttribute(attention_band_tma_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, ATT_SMEM_BYTES));
        configured[dev & 63] = true;
    }
    const unsigned grid = (unsigned) std::min(a.n_items, num_sms);
    if (dtype == OP_BF16) launch_pdl(attention_band_tma_kernel<true>, dim3(grid), dim3(ATT_THREADS), (size_t) ATT_SMEM_BYTES, st, a);
    else launch_pdl(attention_band_tma_kernel<false>, dim3(grid), dim3(ATT_THREADS), (size_t) ATT_SMEM_BYTES, st, a);
    MIO_CUDA(cudaGetLastError());
    count_launch();
}

void launch_attention(const void * qkv16, void * out16, int dtype, int dim, int heads, int window,
                      const row_tile * tiles, int n_tiles, bool force_generic, cudaStream_t st) {
    if (n_tiles <= 0) return;
    const int head_dim = dim / heads;
    const int half = window / 2;
    dim3 grid((unsigned) n_tiles, (unsigned) heads);
    const uint16_t * q = (const uint16_t *) qkv16;
    uint16_t * o = (uint16_t *) out16;
    if (!force_generic && head_dim == HD && half <= 32 && (dim % 8) == 0) {
        if (dtype == OP_BF16) launch_pdl(attention_band_kernel<true>, dim3(grid), dim3(128), 0, st, q, o, dim, half, tiles);
        else launch_pdl(attention_band_kernel<false>, dim3(grid), dim3(128), 0, st, q, o, dim, half, tiles);
    } else {
        const size_t smem = 4 * (size_t) (2 * half + 1) * sizeof(float);
        if (smem > 48 * 1024) throw cuda_error("attention: window too large for the generic kernel");
        if (dtype == OP_BF16) launch_pdl(attention_generic_kernel<true>, dim3(grid), dim3(128), smem, st, q, o, dim, head_dim, half, tiles);
        else launch_pdl(attention_generic_kernel<false>, dim3(grid), dim3(128), smem, st, q, o, dim, head_dim, half, tiles);
    }
    MIO_CUDA(cudaGetLastError());
    count_launch();
}

} // namespace mio
