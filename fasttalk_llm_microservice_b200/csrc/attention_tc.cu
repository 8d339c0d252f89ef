// K6 on tcgen05 / TMEM — causal chunked-prefill attention over the paged KV cache (head_dim 64, GQA group 4).
//
// One CTA = one (request, 64-query tile, KV head): the tile's 64 tokens x the 4 q heads of the group are 256
// query rows, handled as two 128-row halves (32 tokens x 4 heads each, row = token * 4 + head) so that every
// K/V tile that reaches shared memory feeds 256 rows of tensor-core work (the mma.sync kernel re-fetched each
// tile per q head).  Per KV tile of 64 keys and per half:
//
//   S  = Q K^T   tcgen05.mma  M128 N64 K64   A = Q half (TMA 3-D box -> 128-byte-swizzled rows in smem),
//                                           B = the K slab exactly as it lies in the cache: [64 keys][64 d] fp16
//                                           with chunk ^ (row & 7) IS the K-major SWIZZLE_128B operand layout
//   softmax      4 warps, one thread per row: tcgen05.ld of the row's 64 scores, running max / sum in fp32,
//                exp2, P -> fp16 -> TENSOR MEMORY (tcgen05.st, row = lane: the A operand of the next product)
//   O += P V     tcgen05.mma  M128 N64 K64   B = the V slab as it lies in the cache: [64 keys][64 d] is the
//                                           MN-major SWIZZLE_128B layout (d contiguous, 8-key groups 1 KB apart)
//
// S (two buffers per half, so the scores of tile j + 1 are ready before the softmax of tile j ends) and O live in
// TMEM together with two P buffers per half (all 512 columns: S 4 x 64, O 2 x 64, P 4 x 32).  O is rescaled lazily: the row maximum used for scaling only moves when
// the running maximum exceeds it by more than 2^8 (P stays far inside fp16 range, the accumulation is fp32, the
// final normalisation uses the same reference — exact arithmetic, not an approximation), so the TMEM round trip
// of a rescale happens on the first tiles only.  The two halves ping-pong: while the softmax warps of one half
// work on their scores, the tensor core runs the other half's products.
//
// Warp roles (320 threads): warp 0 = TMA producer (Q once, then a ring of 16 KB K|V tiles, one bulk copy each),
// warp 1 = TMEM owner + MMA issuer (one elected thread), warps 2..5 = softmax / epilogue of half 0, warps 6..9 of
// half 1 (warp w may touch TMEM lanes 32 * (w % 4) .. +31), one thread per query row.
//
// What bounds it (in-kernel clock64 timeline + ncu, profiles/r02_attn_prefill_tc.md): per 64-key tile the softmax
// warps spend ~1200 clk in the exponentials — 64 x 256 of them at the SFU's 16 / clk / SM — and ~1300 clk in the
// latency-bound rest (TMEM load ~450, row maximum, hand-over); both halves go through these phases in lock step, so
// the SFU is busy half of the time and the tensor pipe 14 %.  Tried on top and not kept: two threads per row (16
// softmax warps), prefetching the next tile's scores into a second register set, starting the halves half a tile
// apart — each within 7 % either way.  The SFU bound itself is ~1024 clk per tile, i.e. about twice this speed.
#include "common.cuh"
#include "ft_internal.h"

#include <mutex>

namespace ft {

constexpr int TC_THREADS = 320;
constexpr int TC_ROWS = 128;                 // rows per half (UMMA M); 32 tokens x G heads of them are real
constexpr int TC_TOK = 64;                   // query tokens per CTA
constexpr int TC_SUB_BYTES = kKvTile * 128;  // one [64 tokens][64 dims] fp16 sub-slab = 8 KB; also 64 rows of a Q block
constexpr int TC_QBLK_BYTES = TC_ROWS * 128; // one 128-row x 64-dim block of Q = 16 KB
constexpr float TC_RESCALE_THRESHOLD = 8.f;  // log2 units

template <int D>
struct TcCfg {
    static constexpr int KH = D / 64;                                   // 64-dim blocks of the head
    static constexpr int TILE_BYTES = 2 * KH * TC_SUB_BYTES;             // K slab + V slab: 16 KB (d 64) / 32 KB (d 128)
    static constexpr int STAGES = D == 64 ? 6 : 4;
    static constexpr int Q_BYTES = 2 * KH * TC_QBLK_BYTES;               // both halves
    static constexpr size_t SMEM = 1024 + Q_BYTES + (size_t)STAGES * TILE_BYTES + 512;
    // TMEM columns: S, two buffers per half, 64 fp32 columns each; P (fp16, 32 columns) is written over the first half
    // of the S buffer it was computed from; O of the two halves behind them
    static constexpr int COL_S = 0, COL_O = 256, COLS = 512;
    static_assert(256 + 2 * D <= 512, "tensor memory");
};

__device__ __forceinline__ void tmem_ld16_nowait(uint32_t taddr, uint32_t* r) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
          "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
        : "r"(taddr)
        : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void tmem_st16(uint32_t taddr, const uint32_t* r) {
    asm volatile(
        "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], "
        "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};" ::"r"(taddr),
        "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]), "r"(r[9]),
        "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15])
        : "memory");
}
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

// 3-D tiled tensor-map copy (d, head, token) -> shared memory
__device__ __forceinline__ void tma_tile_3d(void* dst_smem, const CUtensorMap* map, int c0, int c1, int c2, uint64_t* bar) {
    asm volatile(
        "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];" ::"r"(
            smem_u32(dst_smem)),
        "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2)
        : "memory");
}
// MN-major SWIZZLE_128B operand (V slab): 8-row groups (8 keys x 128 B) 1024 B apart along K; one 64-element MN block
__device__ __forceinline__ uint64_t umma_desc_mn128(uint32_t smem_addr) {
    uint64_t d = 0;
    d |= (uint64_t)((smem_addr & 0x3ffff) >> 4);
    d |= (uint64_t)(TC_SUB_BYTES >> 4) << 16;   // LBO: stride between 64-element MN blocks = one sub-slab (d = 128: two blocks)
    d |= (uint64_t)(1024 >> 4) << 32;       // SBO: stride between 8-row K groups
    d |= (uint64_t)1 << 46;
    d |= (uint64_t)2 << 61;                 // SWIZZLE_128B
    return d;
}
// D[tmem] (+)= A[tmem] * B[smem desc]: the A operand (P, fp16 pairs packed into 32-bit columns, row = lane) is read
// from tensor memory, so only V crosses the shared-memory port — with both operands in shared memory an
// M128 N64 product asks for 192 B/clk, more than the port's 128 B/clk
__device__ __forceinline__ void umma_f16_ts(uint32_t d_tmem, uint32_t a_tmem, uint64_t b_desc, uint32_t idesc,
                                            uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}\n" ::"r"(d_tmem),
        "r"(a_tmem), "l"(b_desc), "r"(idesc), "r"(accumulate)
        : "memory");
}
// 2^x on the SFU, one instruction: the arguments here are <= 2^8 by construction (lazy rescale) and anything below
// -126 may flush to zero, so exp2f()'s range fix-up (a compare and two multiplies per element) buys nothing
__device__ __forceinline__ float ex2_fast(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
__device__ __forceinline__ void mbar_wait_sleep(uint64_t* bar, uint32_t parity) {
    while (!mbar_try_wait(bar, parity)) __nanosleep(20);
}

template <int D, int G>
__global__ void __launch_bounds__(TC_THREADS, 1)
attn_prefill_tc_kernel(const __grid_constant__ CUtensorMap map_q, const uint8_t* __restrict__ kv_layer,
                       const int* __restrict__ page_table, int max_pages, const int* __restrict__ req_q_start,
                       const int* __restrict__ req_slot, const int* __restrict__ req_pos0,
                       const int* __restrict__ tile_req, const int* __restrict__ tile_q0,
                       __nv_bfloat16* __restrict__ out, __nv_bfloat16* __restrict__ out_lo, int nq, int nkv,
                       float scale_log2, int stagger_on) {
    using C = TcCfg<D>;
    constexpr int KH = C::KH;
    constexpr int R = 32 * G;                                 // real rows of a half (row = token * G + head)
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* q_s = smem;                                     // [2 halves][KH dim blocks][128 rows][128 B]
    uint8_t* kv_s = smem + C::Q_BYTES;                       // [STAGES][K: KH sub-slabs | V: KH sub-slabs]
    uint64_t* bars = reinterpret_cast<uint64_t*>(kv_s + (size_t)C::STAGES * C::TILE_BYTES);
    uint64_t* kv_full = bars;                                // [STAGES]
    uint64_t* kv_empty = kv_full + C::STAGES;                // [STAGES]
    uint64_t* q_full = kv_empty + C::STAGES;                 // [1]
    uint64_t* s_full = q_full + 1;                           // [2][2] scores of half h, buffer j & 1, are in TMEM
    uint64_t* p_full = s_full + 4;                           // [2][2] probabilities of half h, tile parity j & 1, are in TMEM
    uint64_t* pv_done = p_full + 4;                          // [2][2] O += P V of half h, tile parity j & 1, retired
    uint64_t* stagger = pv_done + 4;                         // [1] half 0 knows its first row maxima: half 1 may start
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(stagger + 1);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;

    if (threadIdx.x == 0) {
        tma_prefetch_desc(&map_q);
        for (int s = 0; s < C::STAGES; ++s) {
            mbar_init(&kv_full[s], 1);
            mbar_init(&kv_empty[s], 1);
        }
        mbar_init(q_full, 1);
        mbar_init(stagger, TC_ROWS / 32);
        for (int h = 0; h < 2; ++h) {
            mbar_init(&s_full[2 * h], 1);
            mbar_init(&s_full[2 * h + 1], 1);
            mbar_init(&p_full[2 * h], TC_ROWS / 32);         // one arrival per softmax warp
            mbar_init(&p_full[2 * h + 1], TC_ROWS / 32);
            mbar_init(&pv_done[2 * h], 1);
            mbar_init(&pv_done[2 * h + 1], 1);
        }
        fence_mbar_init();
    }
    if (warp == 1) {
        tmem_alloc(tmem_slot, C::COLS);
        tmem_relinquish();
    }
    if (R < TC_ROWS && warp >= 2) {
        // a GQA group of 3 fills 96 of a half's 128 rows: the other 32 rows of every Q block are zeroed so that their
        // (never stored) scores are finite.  Generic-proxy stores, made visible to the tensor core before the barrier.
        for (int i = threadIdx.x - 64; i < 2 * KH * (TC_ROWS - R) * 8; i += TC_THREADS - 64) {
            const int blk = i / ((TC_ROWS - R) * 8), rem = i % ((TC_ROWS - R) * 8);
            *reinterpret_cast<uint4*>(q_s + (size_t)blk * TC_QBLK_BYTES + (R + rem / 8) * 128 + (rem % 8) * 16) = make_uint4(0u, 0u, 0u, 0u);
        }
        fence_proxy_async();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    pdl_wait();
    pdl_launch_dependents();

    // the engine lists a request's q tiles in order, so the LAST tiles see the most keys: take the list back to
    // front, otherwise the longest CTAs start last and the grid ends in a long tail
    const int tile_idx = (int)gridDim.x - 1 - (int)blockIdx.x;
    const int req = tile_req[tile_idx];
    const int q0 = tile_q0[tile_idx];
    const int q_row0 = req_q_start[req];
    const int q_len = req_q_start[req + 1] - q_row0;
    const int pos0 = req_pos0[req];
    const int kvh = blockIdx.y;
    const int rows_here = q_len - q0 < TC_TOK ? q_len - q0 : TC_TOK;
    const int last_pos = pos0 + q0 + rows_here - 1;
    const int n_tiles = last_pos / kKvTile + 1;
    const int* pt = page_table + (size_t)req_slot[req] * max_pages;

    if (warp == 0) {
        // ---------------------------------------------------------------- TMA producer
        if (lane == 0) {
            mbar_expect_tx(q_full, 2 * KH * R * 128);
            for (int h = 0; h < 2; ++h)
                for (int kh = 0; kh < KH; ++kh)
                    tma_tile_3d(q_s + (size_t)(h * KH + kh) * TC_QBLK_BYTES, &map_q, kh * 64, kvh * G, q_row0 + q0 + 32 * h, q_full);
        }
        const uint64_t pol = policy_evict_last();            // re-read by the other KV heads' CTAs / later q tiles
        int stage = 0;
        uint32_t phase = 0;
        for (int t0 = 0; t0 < n_tiles; t0 += 32) {
            const int mine = t0 + lane;
            const int my_page = mine < n_tiles ? pt[mine] : 0;
            const int n = n_tiles - t0 < 32 ? n_tiles - t0 : 32;
            for (int i = 0; i < n; ++i) {
                const int page = __shfl_sync(0xffffffffu, my_page, i);
                if (lane == 0) {
                    mbar_wait_sleep(&kv_empty[stage], phase ^ 1);
                    mbar_expect_tx(&kv_full[stage], C::TILE_BYTES);
                    tma_bulk_g2s(kv_s + (size_t)stage * C::TILE_BYTES,
                                 kv_layer + ((size_t)page * nkv + kvh) * C::TILE_BYTES, C::TILE_BYTES, &kv_full[stage], pol);
                }
                if (++stage == C::STAGES) { stage = 0; phase ^= 1; }
            }
        }
    } else if (warp == 1) {
        // ---------------------------------------------------------------- MMA issuer
        if (elect_one()) {
            constexpr uint32_t idesc_qk = umma_idesc_f16(TC_ROWS, kKvTile, false);
            constexpr uint32_t idesc_pv = umma_idesc_f16(TC_ROWS, D, true);
            const uint32_t q_addr = smem_u32(q_s);
            auto issue_qk = [&](int h, int buf, uint32_t k_addr) {
#pragma unroll
                for (int kh = 0; kh < KH; ++kh) {
                    const uint64_t a = umma_desc_k128(q_addr + (h * KH + kh) * TC_QBLK_BYTES);
                    const uint64_t b = umma_desc_k128(k_addr + kh * TC_SUB_BYTES);
#pragma unroll
                    for (int k = 0; k < 4; ++k)
                        umma_bf16(tmem_base + C::COL_S + (2 * h + buf) * 64, a + 2 * k, b + 2 * k, idesc_qk, (kh | k) != 0);
                }
                umma_commit(&s_full[2 * h + buf]);
            };
            auto stage_of = [&](int j) { return j % C::STAGES; };
            auto k_addr_of = [&](int j) { return smem_u32(kv_s + (size_t)stage_of(j) * C::TILE_BYTES); };
            mbar_wait(q_full, 0);
            // scores run two tiles ahead of the softmax (two S buffers per half): tile j + 2 is issued as soon as the
            // product of tile j — which reads its P out of the same buffer — has been issued in front of it
            for (int j = 0; j < 2 && j < n_tiles; ++j) {
                mbar_wait(&kv_full[stage_of(j)], (j / C::STAGES) & 1);
                tc_fence_after();
                issue_qk(0, j & 1, k_addr_of(j));
                if (!stagger_on) issue_qk(1, j & 1, k_addr_of(j));
            }
            if (stagger_on) {        // half 1 starts when half 0 is about to enter its SFU-bound phase, so that the two halves'
                                     // exponentials do not fall on the SFU at the same time (profiles/r02_attn_prefill_tc.md)
                mbar_wait(stagger, 0);
                for (int j = 0; j < 2 && j < n_tiles; ++j) issue_qk(1, j & 1, k_addr_of(j));
            }
            for (int j = 0; j < n_tiles; ++j) {
                const uint32_t v_addr = k_addr_of(j) + KH * TC_SUB_BYTES;
                const bool more = j + 2 < n_tiles;
                for (int h = 0; h < 2; ++h) {
                    // Every hand-over has one barrier PER BUFFER (tile parity): the softmax of one half may run up to two
                    // tiles ahead of what this thread has observed (it is busy with the other half), and a single barrier
                    // completing two phases unobserved would look "not yet complete" again — a deadlock.
                    mbar_wait(&p_full[2 * h + (j & 1)], (j >> 1) & 1);   // P_h(j) written, S_h(j) consumed, O_h rescaled if needed
                    tc_fence_after();
                    const uint32_t p_tmem = tmem_base + C::COL_S + (2 * h + (j & 1)) * 64;    // 64 fp16 = 32 columns, over S
#pragma unroll
                    for (int k = 0; k < kKvTile / 16; ++k)
                        umma_f16_ts(tmem_base + C::COL_O + h * D, p_tmem + 8 * k, umma_desc_mn128(v_addr + k * 2048), idesc_pv,
                                    (j | k) != 0);
                    umma_commit(&pv_done[2 * h + (j & 1)]);
                    if (h == 1) umma_commit(&kv_empty[stage_of(j)]);   // every product that reads this stage has been issued
                    if (more) {
                        if (h == 0) {
                            mbar_wait(&kv_full[stage_of(j + 2)], ((j + 2) / C::STAGES) & 1);
                            tc_fence_after();
                        }
                        issue_qk(h, j & 1, k_addr_of(j + 2));    // in order behind the product that read P out of this buffer
                    }
                }
            }
        }
    } else {
        // ---------------------------------------------------------------- softmax + epilogue (one thread per row)
        const int h = (warp - 2) >> 2;
        const int quarter = warp & 3;
        const int row = quarter * 32 + lane;                     // TMEM lane == row of the half
        const int tok = h * 32 + row / G;                        // query token within the CTA's 64
        const int head = kvh * G + row % G;
        const int pos = pos0 + q0 + tok;                         // absolute position of this row's query
        const bool valid = row < R && q0 + tok < q_len;
        const uint32_t lane_base = tmem_base + ((uint32_t)(quarter * 32) << 16);
        const uint32_t s_taddr0 = lane_base + C::COL_S + 2 * h * 64, o_taddr = lane_base + C::COL_O + h * D;
        float m_ref = -INFINITY, l_run = 0.f;                    // m_ref in scaled (log2) units
        for (int j = 0; j < n_tiles; ++j) {
            mbar_wait(&s_full[2 * h + (j & 1)], (j >> 1) & 1);
            tc_fence_after();
            const uint32_t s_taddr = s_taddr0 + (j & 1) * 64;
            uint32_t sr[64];
#pragma unroll
            for (int c = 0; c < 4; ++c) tmem_ld16_nowait(s_taddr + 16 * c, sr + 16 * c);
            tmem_ld_wait();
            float t4[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};      // four chains instead of one of 64
            if (j * kKvTile + kKvTile - 1 > pos) {               // diagonal tile: causal mask
#pragma unroll
                for (int c = 0; c < 64; ++c) {
                    float v = __uint_as_float(sr[c]);
                    v = j * kKvTile + c <= pos ? v : -INFINITY;
                    sr[c] = __float_as_uint(v);
                    t4[c & 3] = fmaxf(t4[c & 3], v);
                }
            } else {
#pragma unroll
                for (int c = 0; c < 64; ++c) t4[c & 3] = fmaxf(t4[c & 3], __uint_as_float(sr[c]));
            }
            const float tmax = fmaxf(fmaxf(t4[0], t4[1]), fmaxf(t4[2], t4[3]));
            if (stagger_on && j == 0 && h == 0) {
                __syncwarp();
                if (lane == 0) mbar_arrive(stagger);
            }
            const float m_new = fmaxf(m_ref, tmax * scale_log2);
            const bool grow = m_new > m_ref + TC_RESCALE_THRESHOLD;
            if (j == 0) {
                m_ref = m_new;
            } else if (__any_sync(0xffffffffu, grow)) {          // tcgen05.ld / st are warp-wide: rescale all 32 rows
                mbar_wait(&pv_done[2 * h + ((j - 1) & 1)], ((j - 1) >> 1) & 1);     // O_h complete up to tile j - 1
                tc_fence_after();
                const float alpha = exp2f(m_ref - m_new);        // rows that did not grow: m_new >= m_ref, alpha <= 1
                l_run *= alpha;
                m_ref = m_new;
#pragma unroll
                for (int c = 0; c < D / 16; ++c) {
                    uint32_t orr[16];
                    tmem_ld16_nowait(o_taddr + 16 * c, orr);
                    tmem_ld_wait();
#pragma unroll
                    for (int k = 0; k < 16; ++k) orr[k] = __float_as_uint(__uint_as_float(orr[k]) * alpha);
                    tmem_st16(o_taddr + 16 * c, orr);
                }
                tmem_st_wait();
            }
            const float mu = m_ref == -INFINITY ? 0.f : m_ref;
            float ps[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
            for (int c16 = 0; c16 < 2; ++c16) {                  // 32 keys -> 16 packed columns per store
                uint32_t pk[16];
#pragma unroll
                for (int e = 0; e < 16; ++e) {
                    const float p0 = ex2_fast(fmaf(__uint_as_float(sr[c16 * 32 + 2 * e]), scale_log2, -mu));
                    const float p1 = ex2_fast(fmaf(__uint_as_float(sr[c16 * 32 + 2 * e + 1]), scale_log2, -mu));
                    ps[(2 * e) & 3] += p0;
                    ps[(2 * e + 1) & 3] += p1;
                    __half2 t = __floats2half2_rn(p0, p1);
                    pk[e] = *reinterpret_cast<uint32_t*>(&t);
                }
                tmem_st16(s_taddr + 16 * c16, pk);               // P over the scores it came from (this row has read them)
            }
            tmem_st_wait();
            l_run += (ps[0] + ps[1]) + (ps[2] + ps[3]);
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&p_full[2 * h + (j & 1)]);
        }
        // ---- epilogue: O / l -> bf16 hi + lo planes
        mbar_wait(&pv_done[2 * h + ((n_tiles - 1) & 1)], ((n_tiles - 1) >> 1) & 1);
        tc_fence_after();
        const float inv = 1.f / l_run;
        const size_t off = ((size_t)(q_row0 + q0 + tok) * nq + head) * D;
#pragma unroll
        for (int c = 0; c < D / 16; ++c) {
            uint32_t orr[16];
            tmem_ld16_nowait(o_taddr + 16 * c, orr);
            tmem_ld_wait();
            if (valid) {
                uint32_t hi[8], lo[8];
#pragma unroll
                for (int k = 0; k < 8; ++k) {
                    const float x = __uint_as_float(orr[2 * k]) * inv, y = __uint_as_float(orr[2 * k + 1]) * inv;
                    hi[k] = pack_bf16(x, y);
                    lo[k] = pack_bf16(x - bf16_lo(hi[k]), y - bf16_hi(hi[k]));
                }
                uint4* dst = reinterpret_cast<uint4*>(out + off + 16 * c);
                dst[0] = make_uint4(hi[0], hi[1], hi[2], hi[3]);
                dst[1] = make_uint4(hi[4], hi[5], hi[6], hi[7]);
                if (out_lo) {
                    uint4* dl = reinterpret_cast<uint4*>(out_lo + off + 16 * c);
                    dl[0] = make_uint4(lo[0], lo[1], lo[2], lo[3]);
                    dl[1] = make_uint4(lo[4], lo[5], lo[6], lo[7]);
                }
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, C::COLS);
    }
}

// ------------------------------------------------------------------ host side
bool attn_prefill_tc_supported(int nq, int nkv, int d) {
    static const bool off = getenv("FT_ATTN_PREFILL_MMA_SYNC") != nullptr;        // A/B switch (scripts/, tests)
    return !off && (d == 64 || d == 128) && nkv >= 1 && (nq == 4 * nkv || nq == 3 * nkv);
}

// q tensor maps are cached per (device, pointer, rows, heads, head_dim, group): the engine's q buffer never moves
struct QMapEntry { int dev; const void* q; int rows, nq, d, g; CUtensorMap map; };
static std::mutex g_qmap_mu;
static QMapEntry g_qmaps[16];
static int g_qmap_n = 0, g_qmap_next = 0;

template <int D, int G>
static int launch_tc(const CUtensorMap& map, const void* kv_layer, const int* page_table, int max_pages,
                     const int* req_q_start, const int* req_slot, const int* req_pos0, const int* tile_req,
                     const int* tile_q0, int n_tiles, void* out, void* out_lo, int nq, int nkv, bool pdl, cudaStream_t stream) {
    using C = TcCfg<D>;
    static PerDeviceFlag attr_set;
    if (!attr_set.done()) {
        FT_CHECK_CUDA(cudaFuncSetAttribute(attn_prefill_tc_kernel<D, G>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)C::SMEM));
        attr_set.mark();
    }
    const float scale_log2 = 1.4426950408889634f / sqrtf((float)D);
    static const int stagger_on = getenv("FT_TC_STAGGER") ? atoi(getenv("FT_TC_STAGGER")) : 1;       // A/B switch; measured: 4–6 % at d = 64
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(n_tiles, nkv);
    cfg.blockDim = dim3(TC_THREADS);
    cfg.dynamicSmemBytes = C::SMEM;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl ? 1 : 0;
    FT_CHECK_CUDA(cudaLaunchKernelEx(&cfg, attn_prefill_tc_kernel<D, G>, map, (const uint8_t*)kv_layer, page_table, max_pages,
                                     req_q_start, req_slot, req_pos0, tile_req, tile_q0, (__nv_bfloat16*)out,
                                     (__nv_bfloat16*)out_lo, nq, nkv, scale_log2, stagger_on));
    return 0;
}

int launch_attn_prefill_tc(const void* q, int q_rows, const void* kv_layer, const int* page_table, int max_pages,
                           const int* req_q_start, const int* req_slot, const int* req_pos0, const int* tile_req,
                           const int* tile_q0, int n_tiles, void* out, void* out_lo, int nq, int nkv, int d, bool pdl,
                           cudaStream_t stream) {
    FT_REQUIRE(q_rows >= 1, "attn_prefill (tcgen05): the q buffer's row count is needed for its tensor map");
    const int g = nq / nkv;
    CUtensorMap map;
    {
        std::lock_guard<std::mutex> lk(g_qmap_mu);
        const int dev = current_device_slot();
        const QMapEntry* hit = nullptr;
        for (int i = 0; i < g_qmap_n; ++i)
            if (g_qmaps[i].dev == dev && g_qmaps[i].q == q && g_qmaps[i].rows == q_rows && g_qmaps[i].nq == nq &&
                g_qmaps[i].d == d && g_qmaps[i].g == g)
                hit = &g_qmaps[i];
        if (!hit) {
            QMapEntry e{dev, q, q_rows, nq, d, g, {}};
            if (int r = make_tmap_f16_3d(&e.map, q, d, nq, q_rows, 64, g, 32)) return r;
            g_qmaps[g_qmap_next] = e;
            hit = &g_qmaps[g_qmap_next];
            g_qmap_next = (g_qmap_next + 1) % 16;
            if (g_qmap_n < 16) ++g_qmap_n;
        }
        map = hit->map;
    }
#define FT_TC_CASE(D_, G_)                                                                                        \
    if (d == D_ && g == G_)                                                                                       \
        return launch_tc<D_, G_>(map, kv_layer, page_table, max_pages, req_q_start, req_slot, req_pos0, tile_req,  \
                                 tile_q0, n_tiles, out, out_lo, nq, nkv, pdl, stream);
    FT_TC_CASE(64, 4)
    FT_TC_CASE(64, 3)
    FT_TC_CASE(128, 4)
    FT_TC_CASE(128, 3)
#undef FT_TC_CASE
    FT_REQUIRE(false, "attn_prefill (tcgen05): head_dim %d with a GQA group of %d is not built", d, g);
}

}  // namespace ft
