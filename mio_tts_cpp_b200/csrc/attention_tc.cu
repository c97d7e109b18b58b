// Band attention on the 5th-generation tensor cores (tcgen05 / TMEM / TMA), sm_100a only.
//
// Reference: softmax(Q K^T / sqrt(hd) + band mask) V over a dense N x N matrix
// (/root/reference/src/miocodec-decoder.cpp:1219-1257, mask :1501-1525).  Here, per (128-query tile, head):
//
//   producer warp   three cp.async.bulk.tensor loads: Q rows [r0, r0+128), K and V rows [r0-32, r0+160) of the head's 64
//                   columns (128-byte rows, SWIZZLE_128B, rows outside the tensor zero-filled) into a 3-stage ring
//   MMA thread      S[128 x 192] = Q K^T: four tcgen05.mma (M 128, N 192, K 16), both operands K-major from smem,
//                   accumulator in TMEM;  O[128 x 64] = P V: twelve tcgen05.mma with the A operand (P, 16-bit) read FROM
//                   TMEM and V as an MN-major smem operand (the [key, dim] rows are used as they are: no transpose)
//   softmax warps   two sets of four warps (one TMEM lane quadrant each, sets alternate items): tcgen05.ld hands every
//                   thread ONE query row -- the 96 score columns its 32-row quadrant can see -- band mask, max, exp,
//                   sum in registers; P goes back into TMEM over the S columns (tcgen05.st, two 16-bit values per
//                   column, zeros outside the quadrant's window); after the PV MMAs the same thread scales its output
//                   row by 1 / sum, and each warp stores its 32 x 64 box with one TMA store
//
// QK^T of item i+1 is issued before PV of item i, so the tensor pipe works under the softmax of the other set.  Tiles
// are runs of <= 128 rows of ONE utterance counted from its first row: a row's place in its tile does not depend on the
// batch, so results are bit-identical alone and inside any batch.  The kernel streams q/k/v once (K/V halo 1.5x
// through L2) and is bound by HBM; the mma.sync kernels in attention.cu (MIO_TTS_B200_ATTN=v2 / v1) were bound by the
// legacy HMMA issue rate.
#include "common.h"

#include <cuda_bf16.h>
#include <cuda_fp16.h>

#include <algorithm>
#include <cstring>

namespace mio {
namespace {

constexpr int AT_Q = 128;                      // queries per tile
constexpr int AT_K = AT_Q + 64;                // keys visible to a tile
constexpr int AT_HD = 64;
constexpr int AT_STAGES = 3;
constexpr int AT_Q_BYTES = AT_Q * 128, AT_KV_BYTES = AT_K * 128;
constexpr int AT_STAGE_BYTES = AT_Q_BYTES + 2 * AT_KV_BYTES;        // 64 KB
constexpr int AT_OUT_WARP_BYTES = 32 * 128;                         // one warp's 32 x 64 output box
constexpr int AT_THREADS = 640;                                     // warp 0 TMA, warp 1 MMA, warps 4-11 / 12-19 softmax sets
constexpr int AT_SMEM_BYTES = AT_STAGES * AT_STAGE_BYTES + 8 * AT_OUT_WARP_BYTES + 256 + 1024;
constexpr int AT_TMEM_COLS = 512;                                   // two sets x (S 192 + O 64)

struct attc_args {
    CUtensorMap map_q;       // [rows, 3*dim] 16-bit, box 64 cols x 128 rows, SWIZZLE_128B
    CUtensorMap map_kv;      // same tensor, box 64 cols x 192 rows
    CUtensorMap map_out;     // [rows, dim] 16-bit, box 64 cols x 32 rows, SWIZZLE_128B
    int rows, dim, heads, half_window, n_items, dtype, n_tiles, reverse;
    const row_tile * tiles;  // <= 128-row runs inside one utterance
    uint16_t * out;
};

__device__ __forceinline__ uint32_t s_u32(const void * p) { return (uint32_t) __cvta_generic_to_shared(p); }
__device__ __forceinline__ void bar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void bar_arrive(uint32_t bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void bar_wait(uint32_t bar, uint32_t parity) {
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
__device__ __forceinline__ bool bar_test(uint32_t bar, uint32_t parity) {     // non-blocking
    uint32_t done = 0;
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                 : "=r"(done) : "r"(bar), "r"(parity) : "memory");
    return done != 0;
}
__device__ __forceinline__ void tma_load(uint32_t dst, const CUtensorMap * map, uint32_t bar, int c0, int c1) {
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
                 ::"r"(dst), "l"((uint64_t) map), "r"(bar), "r"(c0), "r"(c1) : "memory");
}
__device__ __forceinline__ void fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void mma_commit(uint32_t bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
// D[tmem] (+)= A[smem desc] * B[smem desc]
__device__ __forceinline__ void mma_ss(uint32_t d, uint64_t da, uint64_t db, uint32_t idesc, uint32_t acc) {
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                 "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
                 ::"r"(d), "l"(da), "l"(db), "r"(idesc), "r"(acc) : "memory");
}
// D[tmem] (+)= A[tmem] * B[smem desc]
__device__ __forceinline__ void mma_ts(uint32_t d, uint32_t a_tmem, uint64_t db, uint32_t idesc, uint32_t acc) {
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                 "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}"
                 ::"r"(d), "r"(a_tmem), "l"(db), "r"(idesc), "r"(acc) : "memory");
}
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&r)[32]) {
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
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t (&r)[16]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
          "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
        : "r"(taddr)
        : "memory");
}
__device__ __forceinline__ void tmem_st8(uint32_t taddr, const uint32_t (&r)[8]) {
    asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};"
                 ::"r"(taddr), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7])
                 : "memory");
}
__device__ __forceinline__ void tmem_st1(uint32_t taddr, uint32_t v) {
    asm volatile("tcgen05.st.sync.aligned.32x32b.x1.b32 [%0], {%1};" ::"r"(taddr), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t tmem_ld1(uint32_t taddr) {     // completes before returning
    uint32_t v;
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x1.b32 {%0}, [%1];" : "=r"(v) : "r"(taddr) : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
    return v;
}
__device__ __forceinline__ void tmem_st16(uint32_t taddr, const uint32_t (&r)[16]) {
    asm volatile(
        "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], "
        "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};"
        ::"r"(taddr), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]),
          "r"(r[8]), "r"(r[9]), "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15])
        : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

// shared-memory matrix descriptors (cute::UMMA::SmemDescriptor): start>>4 [0,14) | LBO>>4 [16,30) | SBO>>4 [32,46) |
// version 1 [46,48) | SWIZZLE_128B = 2 [61,64).  Both operand forms used here are rows of 128 bytes in 1024-byte atoms of
// 8 rows: K-major (row = M/N index, 64 K elements per row) and MN-major (row = K index, 64 N elements per row).
__device__ __forceinline__ uint64_t desc_sw128(uint32_t addr) {
    return (uint64_t) ((addr >> 4) & 0x3FFF) | ((uint64_t) 1 << 16) | ((uint64_t) (1024 >> 4) << 32) | ((uint64_t) 1 << 46) |
           ((uint64_t) 2 << 61);
}

__device__ __forceinline__ float ex2_fast(float x) {      // MUFU.EX2, flush-to-zero: one instruction (2^-inf = 0)
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}

template <bool BF16> __device__ __forceinline__ uint32_t pk(float a, float b) {
    if constexpr (BF16) {
        __nv_bfloat162 v = __floats2bfloat162_rn(a, b);
        return *reinterpret_cast<uint32_t *>(&v);
    } else {
        return pack_f16x2_sat(a, b);
    }
}

template <bool BF16>
__global__ void __launch_bounds__(AT_THREADS, 1) attention_band_tc_kernel(const __grid_constant__ attc_args a) {
    MIO_PDL_TRIGGER();
    extern __shared__ uint8_t at_smem_raw[];
    const uint32_t base = (s_u32(at_smem_raw) + 1023u) & ~1023u;
    const uint32_t out_base = base + AT_STAGES * AT_STAGE_BYTES;
    const uint32_t bar_base = out_base + 8 * AT_OUT_WARP_BYTES;
    auto load_full = [&](int s) { return bar_base + 8u * s; };
    auto load_empty = [&](int s) { return bar_base + 8u * (AT_STAGES + s); };
    auto s_full = [&](int b) { return bar_base + 8u * (2 * AT_STAGES + b); };
    auto p_full = [&](int b) { return bar_base + 8u * (2 * AT_STAGES + 2 + b); };
    auto o_full = [&](int b) { return bar_base + 8u * (2 * AT_STAGES + 4 + b); };
    const uint32_t tmem_slot = bar_base + 8u * (2 * AT_STAGES + 6);
    volatile uint32_t * tmem_slot_ptr = reinterpret_cast<volatile uint32_t *>(at_smem_raw + (tmem_slot - s_u32(at_smem_raw)));
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

    if (threadIdx.x == 0) {
        for (int s = 0; s < AT_STAGES; ++s) { bar_init(load_full(s), 1); bar_init(load_empty(s), 1); }
        for (int b = 0; b < 2; ++b) { bar_init(s_full(b), 1); bar_init(p_full(b), 8); bar_init(o_full(b), 1); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "r"((uint32_t) AT_TMEM_COLS) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    fence_before();
    __syncthreads();
    fence_after();
    const uint32_t tmem = *tmem_slot_ptr;
    MIO_PDL_WAIT();

    const int n_local = a.n_items > (int) blockIdx.x ? (a.n_items - (int) blockIdx.x + (int) gridDim.x - 1) / (int) gridDim.x : 0;

    if (warp == 0) {
        // ===== producer: the warp fetches the first rows of its next 32 items together, lane 0 issues the loads =====
        const int step_t = (int) gridDim.x / a.heads, step_h = (int) gridDim.x % a.heads;
        int tile = (int) blockIdx.x / a.heads, head = (int) blockIdx.x % a.heads;
        for (int i0 = 0; i0 < n_local; i0 += 32) {
            int my_tile = tile + lane * step_t, my_head = head + lane * step_h;
            my_tile += my_head / a.heads; my_head %= a.heads;
            const int my_r0 = i0 + lane < n_local ? __ldg(&a.tiles[a.reverse ? a.n_tiles - 1 - my_tile : my_tile].row0) : 0;
            const int n_here = min(32, n_local - i0);
            for (int l = 0; l < n_here; ++l) {
                const int r0 = __shfl_sync(0xffffffffu, my_r0, l);
                const int c0 = __shfl_sync(0xffffffffu, my_head, l) * AT_HD;
                if (lane == 0) {
                    const int i = i0 + l, s = i % AT_STAGES;
                    bar_wait(load_empty(s), ((i / AT_STAGES) & 1) ^ 1);
                    const uint32_t st = base + s * AT_STAGE_BYTES;
                    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(load_full(s)), "r"(AT_STAGE_BYTES) : "memory");
                    tma_load(st, &a.map_q, load_full(s), c0, r0);
                    tma_load(st + AT_Q_BYTES, &a.map_kv, load_full(s), a.dim + c0, r0 - 32);
                    tma_load(st + AT_Q_BYTES + AT_KV_BYTES, &a.map_kv, load_full(s), 2 * a.dim + c0, r0 - 32);
                }
                __syncwarp();
            }
            tile += 32 * step_t; head += 32 * step_h;
            tile += head / a.heads; head %= a.heads;
        }
    } else if (warp == 1) {
        // ===== MMA issuer =====
        if (lane == 0) {
            // instruction descriptors (cute::UMMA::InstrDescriptor): c = f32 [4,6), a/b format [7,10)/[10,13), a/b major
            // bits 15/16 (1 = MN-major), N>>3 [17,23), M>>4 [24,29)
            const uint32_t fmt = (1u << 4) | ((uint32_t) a.dtype << 7) | ((uint32_t) a.dtype << 10) | ((uint32_t) (AT_Q >> 4) << 24);
            const uint32_t idesc_s = fmt | ((uint32_t) (AT_K >> 3) << 17);                 // S = Q K^T: N = 192, both K-major
            const uint32_t idesc_o = fmt | ((uint32_t) (AT_HD >> 3) << 17) | (1u << 16);   // O = P V: N = 64, V MN-major
            // Two streams of work for one thread: QK^T of item q as soon as its stage has landed (and PV of item q-2, which
            // reads the P that aliases this S buffer, has been issued: the tensor pipe runs MMAs in issue order), PV of item
            // p as soon as its softmax has written P.  Neither may wait for the other, so both barriers are polled.
            int q = 0, pv = 0;
            long long t0 = 0;
            while (pv < n_local) {
                bool progressed = false;
                if (q < n_local && q - pv < 2 && bar_test(load_full(q % AT_STAGES), (q / AT_STAGES) & 1)) {
                    fence_after();
                    const uint32_t st = base + (q % AT_STAGES) * AT_STAGE_BYTES;
                    const uint64_t dq = desc_sw128(st), dk = desc_sw128(st + AT_Q_BYTES);
#pragma unroll
                    for (int k = 0; k < AT_HD / 16; ++k)                 // +32 bytes inside the 128-byte swizzle atom per step
                        mma_ss(tmem + (q & 1) * 256, dq + 2 * k, dk + 2 * k, idesc_s, k != 0);
                    mma_commit(s_full(q & 1));
                    ++q;
                    progressed = true;
                }
                if (pv < q && bar_test(p_full(pv & 1), (pv >> 1) & 1)) {
                    fence_after();
                    const int b = pv & 1, s = pv % AT_STAGES;
                    const uint32_t p_tmem = tmem + b * 256;              // P over the S columns: 2 keys per column
                    const uint32_t o_tmem = tmem + b * 256 + AT_K;
                    const uint64_t dv = desc_sw128(base + s * AT_STAGE_BYTES + AT_Q_BYTES + AT_KV_BYTES);
#pragma unroll
                    for (int k = 0; k < AT_K / 16; ++k)                  // 16 keys per step: 8 P columns, 16 V rows (2048 B)
                        mma_ts(o_tmem, p_tmem + 8 * k, dv + (uint64_t) (128 * k), idesc_o, k != 0);
                    mma_commit(o_full(b));
                    mma_commit(load_empty(s));                           // Q, K, V of the item are no longer needed
                    ++pv;
                    progressed = true;
                }
                if (progressed) t0 = 0;
                else if (t0 == 0) t0 = clock64();
                else if (clock64() - t0 > 4000000000ll) __trap();
            }
        }
    } else if (warp >= 4) {
        // ===== softmax / output warps: two sets (alternate items) x two halves x four TMEM lane quadrants =====
        // A query row is shared by TWO threads (same lane of warps `quad` of the two halves): half h owns score columns
        // idx [48h, 48h+48) of the 96 its quadrant can see, P words [24h, 24h+24) of the quadrant's 48, and output columns
        // [32h, 32h+32).  Row max and row sum cross the pair through four spare TMEM columns of the row's own lane (the
        // set's O columns, idle until the PV MMAs that this pair's p_full arrivals release) and a 64-thread named barrier.
        // Sixteen softmax warps (four per scheduler) hide the TMEM / MUFU / barrier latencies that eight could not.
        const int cw = warp - 4;
        const int set = cw >> 3, half = (cw >> 2) & 1, quad = warp & 3;
        const int pair_bar = 1 + set * 4 + quad;                       // named barrier of the two warps sharing these rows
        const int hw = a.half_window;
        const uint32_t obuf = out_base + (set * 4 + quad) * AT_OUT_WARP_BYTES;
        const uint32_t lane_addr = (uint32_t) (quad * 32) << 16;
        const int gstep_t = (2 * (int) gridDim.x) / a.heads, gstep_h = (2 * (int) gridDim.x) % a.heads;
        int tile, head;
        {
            const int item = (int) blockIdx.x + set * (int) gridDim.x;
            tile = item / a.heads; head = item % a.heads;
        }
        // (tcgen05 accesses of the two warps are ordered across the barrier by the before / after thread-sync fences)
        auto pair_sync = [&]() {
            fence_before();
            asm volatile("bar.sync %0, 64;" ::"r"(pair_bar) : "memory");
            fence_after();
        };
        bool store_pending = false;
        for (int i = set; i < n_local; i += 2) {
            const int2 * tp = reinterpret_cast<const int2 *>(a.tiles + (a.reverse ? a.n_tiles - 1 - tile : tile));   // 24-byte records
            const int2 t01 = __ldg(tp), t23 = __ldg(tp + 1), t45 = __ldg(tp + 2);  // (row0, n) (utt, pos0) (utt_row0, utt_len)
            const int cur_head = head;
            tile += gstep_t; head += gstep_h;
            if (head >= a.heads) { head -= a.heads; ++tile; }
            const int r0 = t01.x, tn = t01.y, pos0 = t23.y, utt_len = t45.y;
            const int k = i >> 1;
            // this thread's row: query r = 32 quad + lane of the tile, position pos0 + r; the quadrant sees score columns
            // [32 quad, 32 quad + 96) as idx 0..95; the row's own key is idx lane + 32; its band is idx in [lo, hi]
            const int r = quad * 32 + lane;
            const int pos = pos0 + r;
            int lo = lane + 32 - min(hw, pos), hi = lane + 32 + min(hw, utt_len - 1 - pos);
            if (r >= tn || hi < lo) { lo = 1 << 20; hi = lo; }
            const unsigned span = (unsigned) (hi - lo);
            const int d0 = 48 * half - lo;                          // idx of this half's element e is 48 half + e

            const uint32_t s_tmem = tmem + set * 256 + lane_addr;
            bar_wait(s_full(set), k & 1);
            fence_after();
            float m = -INFINITY, l = 0.0f;
            uint32_t pw[24];
            {
                uint32_t raw[3][16];
#pragma unroll
                for (int c = 0; c < 3; ++c) tmem_ld16(s_tmem + quad * 32 + half * 48 + c * 16, raw[c]);
                tmem_ld_wait();
#pragma unroll
                for (int c = 0; c < 3; ++c) {
#pragma unroll
                    for (int e = 0; e < 16; ++e) {
                        const float v = (unsigned) (c * 16 + e + d0) <= span ? __uint_as_float(raw[c][e]) : -INFINITY;
                        raw[c][e] = __float_as_uint(v);
                        m = fmaxf(m, v);
                    }
                }
                tmem_st1(s_tmem + AT_K + half, __float_as_uint(m));
                tmem_st_wait();
                pair_sync();
                m = fmaxf(m, __uint_as_float(tmem_ld1(s_tmem + AT_K + (half ^ 1))));
                if (m == -INFINITY) m = 0.0f;
                const float ml2 = -m * 1.4426950408889634f;       // exp(s - m) = 2^(s log2e - m log2e): one FFMA + MUFU.EX2
#pragma unroll
                for (int c = 0; c < 3; ++c) {
#pragma unroll
                    for (int e = 0; e < 16; e += 2) {
                        const float p0 = ex2_fast(fmaf(__uint_as_float(raw[c][e]), 1.4426950408889634f, ml2));
                        const float p1 = ex2_fast(fmaf(__uint_as_float(raw[c][e + 1]), 1.4426950408889634f, ml2));
                        l += p0 + p1;
                        pw[(c * 16 + e) >> 1] = pk<BF16>(p0, p1);
                    }
                }
                tmem_st1(s_tmem + AT_K + 2 + half, __float_as_uint(l));
            }
            // P over the S columns of this row: word w holds keys 2w, 2w+1; the quadrant's 48 words sit at column 16 quad
            // (this half's 24 at + 24 half); the other 48 of the row's 96 words are zero, 24 written by each half
            {
                uint32_t zero8[8];
#pragma unroll
                for (int e = 0; e < 8; ++e) zero8[e] = 0u;
                const uint32_t (*pw8)[8] = reinterpret_cast<const uint32_t (*)[8]>(pw);
#pragma unroll
                for (int j = 0; j < 3; ++j) tmem_st8(s_tmem + quad * 16 + half * 24 + j * 8, pw8[j]);
#pragma unroll
                for (int j = 0; j < 3; ++j) {
                    int w0 = quad * 16 + 48 + half * 24 + j * 8;   // words [16 quad + 48, 16 quad + 96) mod 96
                    if (w0 >= 96) w0 -= 96;
                    tmem_st8(s_tmem + w0, zero8);
                }
            }
            tmem_st_wait();
            if (store_pending) {      // half 0's lane 0 issued the previous store of this box; the pair barrier below orders it
                if (lane == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
                store_pending = false;
            }
            pair_sync();              // partner's row sum is in TMEM; the previous store has read the box
            l += __uint_as_float(tmem_ld1(s_tmem + AT_K + 2 + (half ^ 1)));   // read BEFORE p_full: PV overwrites these columns
            const float inv = l > 0.0f ? __fdividef(1.0f, l) : 0.0f;
            fence_before();
            __syncwarp();
            if (lane == 0) bar_arrive(p_full(set));

            // ---- O row = (P V) / l -> the pair's swizzled output box -> one TMA store ----
            bar_wait(o_full(set), k & 1);
            fence_after();
            {
                uint32_t o[32];
                tmem_ld32(s_tmem + AT_K + half * 32, o);
                tmem_ld_wait();
                const uint32_t row = obuf + lane * 128;
#pragma unroll
                for (int c = 0; c < 4; ++c) {                   // 16-byte chunk 4 half + c = head-dim columns [32 half + 8c, +8)
                    const uint32_t * s8 = &o[c * 8];
                    const uint32_t w0 = pk<BF16>(__uint_as_float(s8[0]) * inv, __uint_as_float(s8[1]) * inv);
                    const uint32_t w1 = pk<BF16>(__uint_as_float(s8[2]) * inv, __uint_as_float(s8[3]) * inv);
                    const uint32_t w2 = pk<BF16>(__uint_as_float(s8[4]) * inv, __uint_as_float(s8[5]) * inv);
                    const uint32_t w3 = pk<BF16>(__uint_as_float(s8[6]) * inv, __uint_as_float(s8[7]) * inv);
                    asm volatile("st.shared.v4.b32 [%0], {%1,%2,%3,%4};" ::"r"(row + (((uint32_t) ((4 * half + c) ^ lane) & 7u) << 4)),
                                 "r"(w0), "r"(w1), "r"(w2), "r"(w3) : "memory");
                }
            }
            fence_before();      // the TMEM reads above are ordered before the next MMA that overwrites this set's columns
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            pair_sync();         // both halves of every row are in the box
            if (quad * 32 + 32 <= tn) {
                if (half == 0) {
                    if (lane == 0) {
                        asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];"
                                     ::"l"((uint64_t) &a.map_out), "r"(obuf), "r"(cur_head * AT_HD), "r"(r0 + quad * 32) : "memory");
                        asm volatile("cp.async.bulk.commit_group;" ::: "memory");
                    }
                    store_pending = true;
                }
            } else if (quad * 32 < tn) {
                // the utterance ends inside these 32 rows: plain 16-byte stores of the rows that exist (4 of 8 passes per half)
#pragma unroll
                for (int it = 0; it < 4; ++it) {
                    const int c = lane + 32 * (it + 4 * half), rr = c >> 3, ch = c & 7;
                    if (quad * 32 + rr < tn) {
                        uint4 v;
                        asm volatile("ld.shared.v4.b32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
                                     : "r"(obuf + rr * 128 + (((uint32_t) (ch ^ rr) & 7u) << 4)));
                        *reinterpret_cast<uint4 *>(a.out + (size_t) (r0 + quad * 32 + rr) * a.dim + cur_head * AT_HD + ch * 8) = v;
                    }
                }
                pair_sync();     // the box may be rewritten only after both halves have read it
            }
        }
        if (lane == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
        __syncwarp();
    }

    fence_before();
    __syncthreads();
    if (warp == 1) {
        fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"((uint32_t) AT_TMEM_COLS) : "memory");
    }
}

} // namespace

int attention_mode() {   // 0 = v1 (cp.async + mma.sync), 1 = v2 (TMA ring + mma.sync), 2 = tcgen05 (default)
    static const int mode = [] {
        const char * e = std::getenv("MIO_TTS_B200_ATTN");
        if (e && std::strcmp(e, "v1") == 0) return 0;
        if (e && std::strcmp(e, "v2") == 0) return 1;
        return 2;
    }();
    return mode;
}

bool attention_tc_ok(int dim, int heads, int window) {
    return attention_mode() == 2 && heads > 0 && dim / heads == AT_HD && window / 2 <= 32 && (dim % 8) == 0;
}

void launch_attention_tc(const CUtensorMap & map_q, const CUtensorMap & map_kv, const CUtensorMap & map_out, void * out16,
                         int dtype, int dim, int heads, int window, int rows, const row_tile * tiles128, int n_tiles,
                         int num_sms, cudaStream_t st) {
    if (rows <= 0 || n_tiles <= 0) return;
    attc_args a;
    a.map_q = map_q; a.map_kv = map_kv; a.map_out = map_out;
    a.rows = rows; a.dim = dim; a.heads = heads; a.half_window = window / 2; a.dtype = dtype;
    a.n_items = n_tiles * heads; a.n_tiles = n_tiles; a.reverse = traversal_reverse();
    a.tiles = tiles128; a.out = (uint16_t *) out16;
    static bool configured[64] = {false};
    int dev = 0;
    MIO_CUDA(cudaGetDevice(&dev));
    if (!configured[dev & 63]) {
        MIO_CUDA(cudaFuncSetAttribute(attention_band_tc_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, AT_SMEM_BYTES));
        MIO_CUDA(cudaFuncSetAttribute(attention_band_tc_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, AT_SMEM_BYTES));
        configured[dev & 63] = true;
    }
    const unsigned grid = (unsigned) std::min(a.n_items, num_sms);
    if (dtype == OP_BF16) launch_pdl(attention_band_tc_kernel<true>, dim3(grid), dim3(AT_THREADS), (size_t) AT_SMEM_BYTES, st, a);
    else launch_pdl(attention_band_tc_kernel<false>, dim3(grid), dim3(AT_THREADS), (size_t) AT_SMEM_BYTES, st, a);
    MIO_CUDA(cudaGetLastError());
    count_launch();
}

} // namespace mio
