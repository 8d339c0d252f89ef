// K5 / K6 — attention over the paged KV cache.
//
// KV cache layer layout: [page][kv_head][K|V][64 tokens][D] fp16, one page = 64 tokens (q, K, V
// and the softmax probabilities are fp16 with fp32 accumulation: same bytes as bf16, three more
// mantissa bits, and the dtype the reference runs its engine with, docker-compose.vllm.yml:45).  The
// K and V slabs of a (page, head) are adjacent, so one attention tile is ONE contiguous
// 2*64*D*2-byte block that a single bulk-TMA copy (cp.async.bulk, SASS UBLKCP) drops into a
// shared-memory stage.  The 16-byte chunks of every token row are XOR-swizzled already in HBM
// (rope_kv_append writes them that way), so ldmatrix reads the landed tile conflict-free
// without a tensor map.
//
// K5 decode: persistent CTAs, stream-K over the flattened (sequence, kv head, tile) list so
// every CTA streams the same number of 64-token tiles no matter how ragged the batch is.
// A (sequence, head) whose tiles span several CTAs is finished inside the same launch by the
// last of those CTAs to arrive (per-segment arrival counters in the scratch buffer).
// The math is mma.sync m16n8k16 with the GQA group (<= 8 q heads) in the M rows: decode
// attention is HBM-bound (SURVEY.md §8a' a'5), tensor throughput is irrelevant here and
// tcgen05's minimum M of 64 would be 94 % padding.
#include "common.cuh"
#include "ft_internal.h"

#include <mutex>

namespace ft {

constexpr int ATTN_COMPUTE_WARPS = 4;
constexpr int ATTN_THREADS = (ATTN_COMPUTE_WARPS + 1) * 32;
constexpr int ATTN_GMAX = 8;                       // q heads per kv head supported (rows 0..7)
constexpr int ATTN_MAX_B_SMEM = 1024;              // batches up to this size keep the tile plan in smem
constexpr float LOG2E = 1.4426950408889634f;

template <int D>
struct AttnCfg {
    static constexpr int SLAB_BYTES = kKvTile * D * 2;       // K (or V) of one page+head
    static constexpr int TILE_BYTES = 2 * SLAB_BYTES;
    static constexpr int STAGES = D == 64 ? 6 : 3;           // 96 KB in flight per CTA
    static constexpr int REC = D + 2;                        // O[D], m, l  (floats per head row)
    static constexpr size_t SMEM = (size_t)STAGES * TILE_BYTES + 2 * STAGES * 8 +
                                   (size_t)ATTN_COMPUTE_WARPS * ATTN_GMAX * REC * 4 + 128;
};

// Prefill attention is bound by its per-warp ldmatrix -> mma -> exp2 chains, not by the KV stream: a short ring
// (64 KB at d = 64) and <= 136 registers let THREE CTAs share an SM (12 compute warps instead of 8).
template <int D>
struct PrefillCfg {
    static constexpr int TILE_BYTES = AttnCfg<D>::TILE_BYTES;
    static constexpr int SLAB_BYTES = AttnCfg<D>::SLAB_BYTES;
    static constexpr int STAGES = D == 64 ? 4 : 3;     // d = 128 keeps two CTAs per SM (its rows need the registers)
    static constexpr size_t SMEM = (size_t)STAGES * TILE_BYTES + 2 * STAGES * 8 + 128;
};

__device__ __forceinline__ void compute_bar_sync() {
    asm volatile("bar.sync 1, %0;" ::"n"(ATTN_COMPUTE_WARPS * 32) : "memory");
}

struct DecodeSeg {
    int b, h, tile_begin, tile_end, tiles_b;
};

// Next (sequence, head) segment of the flat tile range [f, f_end).  tile_prefix[b] = number of
// tiles of sequences < b; flat index = tile_prefix[b]*nkv + h*tiles_b + tile.
__device__ __forceinline__ bool next_decode_seg(long long& f, long long f_end,
                                                const int* __restrict__ tile_prefix, int B, int nkv,
                                                int& b, DecodeSeg& s) {
    if (f >= f_end) return false;
    while (b < B && (long long)tile_prefix[b + 1] * nkv <= f) ++b;
    const int base = tile_prefix[b];
    const int tiles_b = tile_prefix[b + 1] - base;
    const long long rem = f - (long long)base * nkv;
    s.b = b;
    s.h = (int)(rem / tiles_b);
    s.tiles_b = tiles_b;
    s.tile_begin = (int)(rem % tiles_b);
    const long long seg_left = tiles_b - s.tile_begin;
    const long long take = (f_end - f) < seg_left ? (f_end - f) : seg_left;
    s.tile_end = s.tile_begin + (int)take;
    f += take;
    return true;
}

__device__ __forceinline__ int find_seq(const int* __restrict__ tile_prefix, int B, int nkv,
                                        long long f) {
    int lo = 0, hi = B - 1;      // largest b with tile_prefix[b]*nkv <= f
    while (lo < hi) {
        const int mid = (lo + hi + 1) >> 1;
        if ((long long)tile_prefix[mid] * nkv <= f) lo = mid; else hi = mid - 1;
    }
    return lo;
}

__global__ void attn_decode_plan_kernel(const int* __restrict__ ctx_lens, int* __restrict__ tile_prefix,
                                        int B, int* __restrict__ seg_counters, int n_counters) {
    pdl_wait();
    pdl_launch_dependents();
    for (int i = threadIdx.x; i < n_counters; i += blockDim.x) seg_counters[i] = 0;
    if (threadIdx.x == 0) {
        int acc = 0;
        tile_prefix[0] = 0;
        for (int b = 0; b < B; ++b) {
            acc += (ctx_lens[b] + kKvTile - 1) / kKvTile;
            tile_prefix[b + 1] = acc;
        }
    }
}

// Always-on per-CTA timeline of the most recent decode-attention launch (globaltimer ns):
// [cta][0] entry, [1] first tile landed, [2] last tile consumed, [3] exit.  Read back with
// ft_debug_attn_trace (tuning aid, scripts/trace_attn.py).
constexpr int ATTN_TRACE_CTAS = 1024;
__device__ unsigned long long g_attn_trace[ATTN_TRACE_CTAS * 4];
__device__ __forceinline__ void attn_trace(int slot) {
    if (blockIdx.x < ATTN_TRACE_CTAS) {
        unsigned long long t;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
        g_attn_trace[blockIdx.x * 4 + slot] = t;
    }
}

template <int D>
__global__ void __launch_bounds__(ATTN_THREADS)
attn_decode_kernel(const __half* __restrict__ q, const uint8_t* __restrict__ kv_layer,
                   const int* __restrict__ page_table, int max_pages, const int* __restrict__ ctx_lens,
                   const int* __restrict__ tile_prefix, __nv_bfloat16* __restrict__ out,
                   __nv_bfloat16* __restrict__ out_lo, int* __restrict__ seg_counters,
                   float* __restrict__ scratch, int B, int nq, int nkv, float scale_log2, int early) {
    using C = AttnCfg<D>;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 127) & ~uintptr_t(127));
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + (size_t)C::STAGES * C::TILE_BYTES);
    uint64_t* empty_bar = full_bar + C::STAGES;
    float* merge_buf = reinterpret_cast<float*>(empty_bar + C::STAGES);
    __shared__ int s_last;
    __shared__ int s_prefix[ATTN_MAX_B_SMEM + 1];     // tile plan, searched many times per CTA
    // `early` (engine, layers >= 1): the producer may start streaming KV tiles BEFORE the
    // programmatic-dependency wait, because everything it reads then — the tile plan, the page
    // table and every tile except the one that receives this step's token — was written by
    // kernels that completed before the previous layer's attention released its dependents
    // (this kernel triggers its dependents only AFTER its own wait, which is what makes that
    // chain of completions hold).  The QKV projection's latency then hides behind the first
    // 96 KB of KV per CTA.

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const int G = nq / nkv;
    if (threadIdx.x == 0) { attn_trace(0); g_attn_trace[(blockIdx.x < ATTN_TRACE_CTAS ? blockIdx.x : 0) * 4 + 1] = 0; }

    if (threadIdx.x == 0) {
        for (int s = 0; s < C::STAGES; ++s) {
            mbar_init(&full_bar[s], 1);
            mbar_init(&empty_bar[s], ATTN_COMPUTE_WARPS);
        }
        fence_mbar_init();
    }
    const bool is_producer = warp == ATTN_COMPUTE_WARPS;
    bool waited = false;
    if (!early) {                       // the tile plan itself may come from the previous kernel
        pdl_wait();
        pdl_launch_dependents();
        waited = true;
    }
    // one coalesced read of the tile plan instead of a dependent global load per search step
    // (the start-up chain was 3.3 us of idle HBM per launch, scripts/trace_attn.py)
    if (B <= ATTN_MAX_B_SMEM) {
        for (int i = threadIdx.x; i <= B; i += ATTN_THREADS) s_prefix[i] = tile_prefix[i];
        tile_prefix = s_prefix;
    }
    __syncthreads();
    if (!waited && !is_producer) {
        pdl_wait();
        pdl_launch_dependents();
        waited = true;
    }

    // a sequence with no context has no tiles and so no CTA of its own: its output rows are defined as zero
    if (blockIdx.x == 0 && !is_producer) {
        for (int b = 0; b < B; ++b) {
            if (tile_prefix[b + 1] != tile_prefix[b]) continue;
            for (int e = threadIdx.x; e < nq * D; e += ATTN_COMPUTE_WARPS * 32) {
                out[(size_t)b * nq * D + e] = __float2bfloat16(0.f);
                if (out_lo) out_lo[(size_t)b * nq * D + e] = __float2bfloat16(0.f);
            }
        }
    }
    const long long total = (long long)tile_prefix[B] * nkv;
    const long long per = (total + gridDim.x - 1) / gridDim.x;
    long long f = per * blockIdx.x;
    long long f_end = f + per < total ? f + per : total;
    if (f >= total) {
        if (!waited) pdl_wait();
        return;
    }
    int b_cur = find_seq(tile_prefix, B, nkv, f);
    DecodeSeg seg;
    int stage = 0;
    uint32_t phase = 0;

    if (is_producer) {
        // ---------------- producer warp: one bulk-TMA copy per tile
        const uint64_t pol = policy_evict_first();       // every KV byte is read once per step
        while (next_decode_seg(f, f_end, tile_prefix, B, nkv, b_cur, seg)) {
            const int* pt = page_table + (size_t)seg.b * max_pages;
            for (int t0 = seg.tile_begin; t0 < seg.tile_end; t0 += 32) {
                const int mine = t0 + lane;
                const int my_page = mine < seg.tile_end ? pt[mine] : 0;
                const int n = seg.tile_end - t0 < 32 ? seg.tile_end - t0 : 32;
                for (int i = 0; i < n; ++i) {
                    const int page = __shfl_sync(0xffffffffu, my_page, i);
                    if (!waited && t0 + i == seg.tiles_b - 1) {
                        pdl_wait();          // the sequence's last tile holds the token appended this step
                        waited = true;
                    }
                    if (lane == 0) {
                        mbar_wait(&empty_bar[stage], phase ^ 1);
                        mbar_expect_tx(&full_bar[stage], C::TILE_BYTES);
                        tma_bulk_g2s(smem + (size_t)stage * C::TILE_BYTES,
                                     kv_layer + ((size_t)page * nkv + seg.h) * C::TILE_BYTES,
                                     C::TILE_BYTES, &full_bar[stage], pol);
                    }
                    if (++stage == C::STAGES) { stage = 0; phase ^= 1; }
                }
            }
        }
        if (!waited) pdl_wait();
        return;
    }

    // ---------------- compute warps: warp w owns token rows 16w..16w+15 of every tile
    const int g = lane >> 2;        // mma row  = q head within the GQA group
    const int tq = lane & 3;        // mma column pair
    const int r0 = warp * 16;
    const int cta_slot = blockIdx.x;

    while (next_decode_seg(f, f_end, tile_prefix, B, nkv, b_cur, seg)) {
        const int ctx = ctx_lens[seg.b];
        // Q fragments (A operand, rows >= G are zero padding)
        uint32_t qa[D / 16][2];
        {
            const __half* qrow = q + ((size_t)seg.b * nq + (size_t)seg.h * G + g) * D;
#pragma unroll
            for (int ks = 0; ks < D / 16; ++ks) {
                qa[ks][0] = g < G ? *reinterpret_cast<const uint32_t*>(qrow + ks * 16 + 2 * tq) : 0u;
                qa[ks][1] = g < G ? *reinterpret_cast<const uint32_t*>(qrow + ks * 16 + 8 + 2 * tq) : 0u;
            }
        }
        float o[D / 8][4];
#pragma unroll
        for (int nt = 0; nt < D / 8; ++nt) o[nt][0] = o[nt][1] = o[nt][2] = o[nt][3] = 0.f;
        float m_run = -INFINITY, l_run = 0.f;

        for (int tile = seg.tile_begin; tile < seg.tile_end; ++tile) {
            mbar_wait(&full_bar[stage], phase);
            if (threadIdx.x == 0 && blockIdx.x < ATTN_TRACE_CTAS && g_attn_trace[blockIdx.x * 4 + 1] == 0) attn_trace(1);
            const uint32_t kbase = smem_u32(smem + (size_t)stage * C::TILE_BYTES);
            const uint32_t vbase = kbase + C::SLAB_BYTES;

            // S[16 x 16 tokens] = Q K^T
            float s[2][4];
#pragma unroll
            for (int j = 0; j < 2; ++j) {
                s[j][0] = s[j][1] = s[j][2] = s[j][3] = 0.f;
                const int row = r0 + 8 * j + (lane & 7);
#pragma unroll
                for (int kp = 0; kp < D / 32; ++kp) {
                    uint32_t b0, b1, b2, b3;
                    ldmatrix_x4(kbase + kv_swizzled_chunk<D>(row, kp * 4 + (lane >> 3)), b0, b1, b2, b3);
                    mma_f16_16816(s[j], qa[2 * kp][0], 0u, qa[2 * kp][1], 0u, b0, b1);
                    mma_f16_16816(s[j], qa[2 * kp + 1][0], 0u, qa[2 * kp + 1][1], 0u, b2, b3);
                }
            }
            // mask + online softmax for row g (this lane: tokens 2tq, 2tq+1 of each 8-token group)
            const int tok0 = tile * kKvTile + r0 + 2 * tq;
            float p[2][2];
            float tmax = -INFINITY;
#pragma unroll
            for (int j = 0; j < 2; ++j)
#pragma unroll
                for (int e = 0; e < 2; ++e) {
                    const float v = (tok0 + 8 * j + e < ctx) ? s[j][e] * scale_log2 : -INFINITY;
                    p[j][e] = v;
                    tmax = fmaxf(tmax, v);
                }
            tmax = fmaxf(tmax, __shfl_xor_sync(0xffffffffu, tmax, 1));
            tmax = fmaxf(tmax, __shfl_xor_sync(0xffffffffu, tmax, 2));
            const float m_new = fmaxf(m_run, tmax);
            const float m_use = m_new == -INFINITY ? 0.f : m_new;
            const float alpha = exp2f(m_run - m_use);
            float psum = 0.f;
#pragma unroll
            for (int j = 0; j < 2; ++j)
#pragma unroll
                for (int e = 0; e < 2; ++e) {
                    p[j][e] = exp2f(p[j][e] - m_use);
                    psum += p[j][e];
                }
            l_run = l_run * alpha + psum;
            m_run = m_new;
            const uint32_t pa0 = pack_f16_sat(p[0][0], p[0][1]);
            const uint32_t pa2 = pack_f16_sat(p[1][0], p[1][1]);

            // O[16 x D] = alpha * O + P V
#pragma unroll
            for (int np = 0; np < D / 16; ++np) {
                uint32_t b0, b1, b2, b3;
                const int row = r0 + ((lane >> 3) & 1) * 8 + (lane & 7);
                ldmatrix_x4_trans(vbase + kv_swizzled_chunk<D>(row, 2 * np + (lane >> 4)), b0, b1, b2, b3);
                o[2 * np][0] *= alpha; o[2 * np][1] *= alpha;
                o[2 * np + 1][0] *= alpha; o[2 * np + 1][1] *= alpha;
                mma_f16_16816(o[2 * np], pa0, 0u, pa2, 0u, b0, b1);
                mma_f16_16816(o[2 * np + 1], pa0, 0u, pa2, 0u, b2, b3);
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(&empty_bar[stage]);
            if (++stage == C::STAGES) { stage = 0; phase ^= 1; }
        }

        if (threadIdx.x == 0) attn_trace(2);
        // ---- combine the four warps' partial states for this (sequence, head) segment
        l_run += __shfl_xor_sync(0xffffffffu, l_run, 1);
        l_run += __shfl_xor_sync(0xffffffffu, l_run, 2);
        if (g < G) {
            float* rec = merge_buf + ((size_t)warp * ATTN_GMAX + g) * C::REC;
#pragma unroll
            for (int nt = 0; nt < D / 8; ++nt) {
                rec[nt * 8 + 2 * tq] = o[nt][0];
                rec[nt * 8 + 2 * tq + 1] = o[nt][1];
            }
            if (tq == 0) { rec[D] = m_run; rec[D + 1] = l_run; }
        }
        compute_bar_sync();
        const bool whole = seg.tile_begin == 0 && seg.tile_end == seg.tiles_b;
        const int seg_idx = seg.b * nkv + seg.h;
        float* part = scratch + (size_t)(cta_slot + seg_idx) * ATTN_GMAX * C::REC;
        const size_t out_row = ((size_t)seg.b * nq + (size_t)seg.h * G) * D;
        for (int e = threadIdx.x; e < G * D; e += ATTN_COMPUTE_WARPS * 32) {
            const int hg = e / D, dim = e % D;
            float mw[ATTN_COMPUTE_WARPS], M = -INFINITY;
#pragma unroll
            for (int w = 0; w < ATTN_COMPUTE_WARPS; ++w) {
                mw[w] = merge_buf[((size_t)w * ATTN_GMAX + hg) * C::REC + D];
                M = fmaxf(M, mw[w]);
            }
            const float Mu = M == -INFINITY ? 0.f : M;
            float L = 0.f, O = 0.f;
#pragma unroll
            for (int w = 0; w < ATTN_COMPUTE_WARPS; ++w) {
                const float sc = exp2f(mw[w] - Mu);
                const float* rec = merge_buf + ((size_t)w * ATTN_GMAX + hg) * C::REC;
                L += rec[D + 1] * sc;
                O += rec[dim] * sc;
            }
            if (whole) {
                const float r = O / L;
                const __nv_bfloat16 hb = __float2bfloat16(r);
                out[out_row + e] = hb;
                if (out_lo) out_lo[out_row + e] = __float2bfloat16(r - __bfloat162float(hb));
            } else {
                part[(size_t)hg * C::REC + dim] = O;
                if (dim == 0) { part[(size_t)hg * C::REC + D] = M; part[(size_t)hg * C::REC + D + 1] = L; }
            }
        }
        if (!whole) {
            // The last CTA to deliver a part of this (sequence, head) merges all of them.
            const long long F0 = (long long)tile_prefix[seg.b] * nkv + (long long)seg.h * seg.tiles_b;
            const int c_first = (int)(F0 / per), c_last = (int)((F0 + seg.tiles_b - 1) / per);
            compute_bar_sync();
            if (threadIdx.x == 0) {
                __threadfence();            // cumulative: publishes the whole CTA's partial record
                const int old = atomicAdd(&seg_counters[seg_idx], 1);
                const int last = old == c_last - c_first;
                if (last) seg_counters[seg_idx] = 0;      // re-arm for the next layer's launch
                s_last = last;
            }
            compute_bar_sync();
            if (s_last) {
                __threadfence();
                for (int e = threadIdx.x; e < G * D; e += ATTN_COMPUTE_WARPS * 32) {
                    const int hg = e / D, dim = e % D;
                    // all parts' (m, l, o) requested before the first is used: chained, a two-part merge
                    // was four L2 round trips at the very end of the kernel
                    float M = -INFINITY, L = 0.f, O = 0.f;
                    for (int cb = c_first; cb <= c_last; cb += 4) {
                        float pm[4], pl[4], po[4];
#pragma unroll
                        for (int k = 0; k < 4; ++k) {
                            const float* rec = scratch + ((size_t)(cb + k + seg_idx) * ATTN_GMAX + hg) * C::REC;
                            const bool on = cb + k <= c_last;
                            pm[k] = on ? __ldcg(rec + D) : -INFINITY;
                            pl[k] = on ? __ldcg(rec + D + 1) : 0.f;
                            po[k] = on ? __ldcg(rec + dim) : 0.f;
                        }
                        float Mn = M;
#pragma unroll
                        for (int k = 0; k < 4; ++k) Mn = fmaxf(Mn, pm[k]);
                        const float Mu = Mn == -INFINITY ? 0.f : Mn;
                        const float keep = exp2f(M - Mu);              // rescale what earlier groups of parts gave
                        L *= keep;
                        O *= keep;
#pragma unroll
                        for (int k = 0; k < 4; ++k) {
                            const float sc = exp2f(pm[k] - Mu);
                            L += pl[k] * sc;
                            O += po[k] * sc;
                        }
                        M = Mn;
                    }
                    const float r = O / L;
                    const __nv_bfloat16 hb = __float2bfloat16(r);
                    out[out_row + e] = hb;
                    if (out_lo) out_lo[out_row + e] = __float2bfloat16(r - __bfloat162float(hb));
                }
            }
        }
        compute_bar_sync();
    }
    if (threadIdx.x == 0) attn_trace(3);
}

// ------------------------------------------------------------------ K6 chunked prefill
// grid = (q tiles, q heads).  A q tile = 64 consecutive query tokens of one request; warp w
// owns query rows 16w..16w+15 and walks every KV tile up to the causal diagonal.  Keys/values
// come from the paged cache (the chunk's own K/V were appended by rope_kv_append before), so a
// cached prefix and the new tokens take the same path.
template <int D>
__global__ void __launch_bounds__(ATTN_THREADS, D == 64 ? 3 : 2)
attn_prefill_kernel(const __half* __restrict__ q, const uint8_t* __restrict__ kv_layer,
                    const int* __restrict__ page_table, int max_pages,
                    const int* __restrict__ req_q_start, const int* __restrict__ req_slot,
                    const int* __restrict__ req_pos0, const int* __restrict__ tile_req,
                    const int* __restrict__ tile_q0, __nv_bfloat16* __restrict__ out,
                    __nv_bfloat16* __restrict__ out_lo, int nq, int nkv, float scale_log2) {
    using C = PrefillCfg<D>;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 127) & ~uintptr_t(127));
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + (size_t)C::STAGES * C::TILE_BYTES);
    uint64_t* empty_bar = full_bar + C::STAGES;

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    if (threadIdx.x == 0) {
        for (int s = 0; s < C::STAGES; ++s) {
            mbar_init(&full_bar[s], 1);
            mbar_init(&empty_bar[s], ATTN_COMPUTE_WARPS);
        }
        fence_mbar_init();
    }
    __syncthreads();
    pdl_wait();
    pdl_launch_dependents();

    const int req = tile_req[blockIdx.x];
    const int q0 = tile_q0[blockIdx.x];                       // first query of the tile, chunk-relative
    const int q_len = req_q_start[req + 1] - req_q_start[req];
    const int pos0 = req_pos0[req];                           // absolute position of chunk query 0
    const int head = blockIdx.y;
    const int kvh = head / (nq / nkv);
    const int rows_here = q_len - q0 < 64 ? q_len - q0 : 64;
    const int last_pos = pos0 + q0 + rows_here - 1;           // newest key any row of the tile sees
    const int n_tiles = last_pos / kKvTile + 1;
    const int* pt = page_table + (size_t)req_slot[req] * max_pages;

    if (warp == ATTN_COMPUTE_WARPS) {
        const uint64_t pol = policy_evict_last();             // re-read by the other heads / q tiles
        int stage = 0;
        uint32_t phase = 0;
        for (int t0 = 0; t0 < n_tiles; t0 += 32) {
            const int mine = t0 + lane;
            const int my_page = mine < n_tiles ? pt[mine] : 0;
            const int n = n_tiles - t0 < 32 ? n_tiles - t0 : 32;
            for (int i = 0; i < n; ++i) {
                const int page = __shfl_sync(0xffffffffu, my_page, i);
                if (lane == 0) {
                    mbar_wait(&empty_bar[stage], phase ^ 1);
                    mbar_expect_tx(&full_bar[stage], C::TILE_BYTES);
                    tma_bulk_g2s(smem + (size_t)stage * C::TILE_BYTES,
                                 kv_layer + ((size_t)page * nkv + kvh) * C::TILE_BYTES, C::TILE_BYTES,
                                 &full_bar[stage], pol);
                }
                if (++stage == C::STAGES) { stage = 0; phase ^= 1; }
            }
        }
        return;
    }

    const int g = lane >> 2, tq = lane & 3;
    const int row_a = q0 + warp * 16 + g, row_b = row_a + 8;  // chunk-relative query rows
    const int pos_a = pos0 + row_a, pos_b = pos0 + row_b;
    const bool va = row_a < q_len, vb = row_b < q_len;
    uint32_t qa[D / 16][4];
    {
        const __half* qa_ptr = q + ((size_t)(req_q_start[req] + row_a) * nq + head) * D;
        const __half* qb_ptr = q + ((size_t)(req_q_start[req] + row_b) * nq + head) * D;
#pragma unroll
        for (int ks = 0; ks < D / 16; ++ks) {
            qa[ks][0] = va ? *reinterpret_cast<const uint32_t*>(qa_ptr + ks * 16 + 2 * tq) : 0u;
            qa[ks][1] = vb ? *reinterpret_cast<const uint32_t*>(qb_ptr + ks * 16 + 2 * tq) : 0u;
            qa[ks][2] = va ? *reinterpret_cast<const uint32_t*>(qa_ptr + ks * 16 + 8 + 2 * tq) : 0u;
            qa[ks][3] = vb ? *reinterpret_cast<const uint32_t*>(qb_ptr + ks * 16 + 8 + 2 * tq) : 0u;
        }
    }
    float o[D / 8][4];
#pragma unroll
    for (int nt = 0; nt < D / 8; ++nt) o[nt][0] = o[nt][1] = o[nt][2] = o[nt][3] = 0.f;
    float m_a = -INFINITY, m_b = -INFINITY, l_a = 0.f, l_b = 0.f;

    int stage = 0;
    uint32_t phase = 0;
    for (int tile = 0; tile < n_tiles; ++tile) {
        mbar_wait(&full_bar[stage], phase);
        const uint32_t kbase = smem_u32(smem + (size_t)stage * C::TILE_BYTES);
        const uint32_t vbase = kbase + C::SLAB_BYTES;
        float s[8][4];
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            s[j][0] = s[j][1] = s[j][2] = s[j][3] = 0.f;
            const int row = 8 * j + (lane & 7);
#pragma unroll
            for (int kp = 0; kp < D / 32; ++kp) {
                uint32_t b0, b1, b2, b3;
                ldmatrix_x4(kbase + kv_swizzled_chunk<D>(row, kp * 4 + (lane >> 3)), b0, b1, b2, b3);
                mma_f16_16816(s[j], qa[2 * kp][0], qa[2 * kp][1], qa[2 * kp][2], qa[2 * kp][3], b0, b1);
                mma_f16_16816(s[j], qa[2 * kp + 1][0], qa[2 * kp + 1][1], qa[2 * kp + 1][2],
                               qa[2 * kp + 1][3], b2, b3);
            }
        }
        const int key0 = tile * kKvTile + 2 * tq;
        float ta = -INFINITY, tb = -INFINITY;
#pragma unroll
        for (int j = 0; j < 8; ++j)
#pragma unroll
            for (int e = 0; e < 2; ++e) {
                const int key = key0 + 8 * j + e;
                s[j][e] = key <= pos_a ? s[j][e] * scale_log2 : -INFINITY;
                s[j][2 + e] = key <= pos_b ? s[j][2 + e] * scale_log2 : -INFINITY;
                ta = fmaxf(ta, s[j][e]);
                tb = fmaxf(tb, s[j][2 + e]);
            }
        ta = fmaxf(ta, __shfl_xor_sync(0xffffffffu, ta, 1));
        ta = fmaxf(ta, __shfl_xor_sync(0xffffffffu, ta, 2));
        tb = fmaxf(tb, __shfl_xor_sync(0xffffffffu, tb, 1));
        tb = fmaxf(tb, __shfl_xor_sync(0xffffffffu, tb, 2));
        const float na = fmaxf(m_a, ta), nb = fmaxf(m_b, tb);
        const float ua = na == -INFINITY ? 0.f : na, ub = nb == -INFINITY ? 0.f : nb;
        const float alpha_a = exp2f(m_a - ua), alpha_b = exp2f(m_b - ub);
        float sa = 0.f, sb = 0.f;
#pragma unroll
        for (int j = 0; j < 8; ++j)
#pragma unroll
            for (int e = 0; e < 2; ++e) {
                s[j][e] = exp2f(s[j][e] - ua);
                s[j][2 + e] = exp2f(s[j][2 + e] - ub);
                sa += s[j][e];
                sb += s[j][2 + e];
            }
        l_a = l_a * alpha_a + sa;
        l_b = l_b * alpha_b + sb;
        m_a = na;
        m_b = nb;
#pragma unroll
        for (int nt = 0; nt < D / 8; ++nt) {
            o[nt][0] *= alpha_a; o[nt][1] *= alpha_a;
            o[nt][2] *= alpha_b; o[nt][3] *= alpha_b;
        }
#pragma unroll
        for (int kk = 0; kk < 4; ++kk) {               // 16 keys per k-step
            const uint32_t pa0 = pack_f16_sat(s[2 * kk][0], s[2 * kk][1]);
            const uint32_t pa1 = pack_f16_sat(s[2 * kk][2], s[2 * kk][3]);
            const uint32_t pa2 = pack_f16_sat(s[2 * kk + 1][0], s[2 * kk + 1][1]);
            const uint32_t pa3 = pack_f16_sat(s[2 * kk + 1][2], s[2 * kk + 1][3]);
            const int row = kk * 16 + ((lane >> 3) & 1) * 8 + (lane & 7);
#pragma unroll
            for (int np = 0; np < D / 16; ++np) {
                uint32_t b0, b1, b2, b3;
                ldmatrix_x4_trans(vbase + kv_swizzled_chunk<D>(row, 2 * np + (lane >> 4)), b0, b1, b2, b3);
                mma_f16_16816(o[2 * np], pa0, pa1, pa2, pa3, b0, b1);
                mma_f16_16816(o[2 * np + 1], pa0, pa1, pa2, pa3, b2, b3);
            }
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(&empty_bar[stage]);
        if (++stage == C::STAGES) { stage = 0; phase ^= 1; }
    }
    l_a += __shfl_xor_sync(0xffffffffu, l_a, 1);
    l_a += __shfl_xor_sync(0xffffffffu, l_a, 2);
    l_b += __shfl_xor_sync(0xffffffffu, l_b, 1);
    l_b += __shfl_xor_sync(0xffffffffu, l_b, 2);
    const float ia = 1.f / l_a, ib = 1.f / l_b;
    const size_t off_a = ((size_t)(req_q_start[req] + row_a) * nq + head) * D;
    const size_t off_b = ((size_t)(req_q_start[req] + row_b) * nq + head) * D;
    auto store = [&](size_t off, float x, float y) {
        const uint32_t h = pack_bf16(x, y);
        *reinterpret_cast<uint32_t*>(out + off) = h;
        if (out_lo) *reinterpret_cast<uint32_t*>(out_lo + off) = pack_bf16(x - bf16_lo(h), y - bf16_hi(h));
    };
#pragma unroll
    for (int nt = 0; nt < D / 8; ++nt) {
        if (va) store(off_a + nt * 8 + 2 * tq, o[nt][0] * ia, o[nt][1] * ia);
        if (vb) store(off_b + nt * 8 + 2 * tq, o[nt][2] * ib, o[nt][3] * ib);
    }
}

// ------------------------------------------------------------------ launchers
// scratch = [B*nkv arrival counters (padded to 64 ints)] [partial records]
static size_t counter_floats(int B, int nkv) { return (size_t)((B * nkv + 63) / 64) * 64; }
size_t attn_decode_scratch_floats(int B, int nkv, int group, int d, int n_ctas) {
    (void)group;
    return counter_floats(B, nkv) + (size_t)(n_ctas + B * nkv) * ATTN_GMAX * (d + 2);
}

template <int D>
static int decode_ctas_per_sm_d() {
    using C = AttnCfg<D>;
    FT_CHECK_CUDA(cudaFuncSetAttribute(attn_decode_kernel<D>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       (int)C::SMEM));
    int n = 0;
    FT_CHECK_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, attn_decode_kernel<D>, ATTN_THREADS,
                                                                C::SMEM));
    return n;
}
int attn_decode_ctas_per_sm(int d) {
    if (d == 64) return decode_ctas_per_sm_d<64>();
    if (d == 128) return decode_ctas_per_sm_d<128>();
    set_last_error("attn_decode: head_dim %d not supported", d);
    return -2;
}

static void fill_cfg(cudaLaunchConfig_t& cfg, cudaLaunchAttribute* attr, dim3 grid, dim3 block,
                     size_t smem, bool pdl, cudaStream_t stream) {
    cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = smem;
    cfg.stream = stream;
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl ? 1 : 0;
}

int launch_attn_decode_plan(const int* ctx_lens, int* tile_prefix, int B, int nkv, float* scratch, bool pdl,
                            cudaStream_t stream) {
    cudaLaunchConfig_t cfg;
    cudaLaunchAttribute attr[1];
    fill_cfg(cfg, attr, dim3(1), dim3(256), 0, pdl, stream);
    FT_CHECK_CUDA(cudaLaunchKernelEx(&cfg, attn_decode_plan_kernel, ctx_lens, tile_prefix, B,
                                     reinterpret_cast<int*>(scratch), (int)counter_floats(B, nkv)));
    return 0;
}

template <int D>
static int launch_attn_decode_d(const void* q, const void* kv_layer, const int* page_table, int max_pages,
                                const int* ctx_lens, const int* tile_prefix, void* out, void* out_lo,
                                float* scratch, int B, int nq, int nkv, int n_ctas, bool pdl,
                                cudaStream_t stream, cudaEvent_t ev0, cudaEvent_t ev1, bool early) {
    using C = AttnCfg<D>;
    static PerDeviceFlag attr_set;
    if (!attr_set.done()) {
        FT_CHECK_CUDA(cudaFuncSetAttribute(attn_decode_kernel<D>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                           (int)C::SMEM));
        attr_set.mark();
    }
    const float scale_log2 = LOG2E / sqrtf((float)D);
    cudaLaunchConfig_t cfg;
    cudaLaunchAttribute attr[1];
    fill_cfg(cfg, attr, dim3(n_ctas), dim3(ATTN_THREADS), C::SMEM, pdl && !ev0, stream);
    if (ev0) FT_CHECK_CUDA(cudaEventRecord(ev0, stream));
    FT_CHECK_CUDA(cudaLaunchKernelEx(&cfg, attn_decode_kernel<D>, (const __half*)q,
                                     (const uint8_t*)kv_layer, page_table, max_pages, ctx_lens, tile_prefix,
                                     (__nv_bfloat16*)out, (__nv_bfloat16*)out_lo, reinterpret_cast<int*>(scratch),
                                     scratch + counter_floats(B, nkv), B, nq, nkv, scale_log2,
                                     (int)(early && pdl && !ev0)));
    if (ev1) FT_CHECK_CUDA(cudaEventRecord(ev1, stream));
    return 0;
}

int launch_attn_decode(const void* q, const void* kv_layer, const int* page_table, int max_pages,
                       const int* ctx_lens, const int* tile_prefix, void* out, void* out_lo, float* scratch,
                       int B, int nq, int nkv, int d, int n_ctas, bool pdl, cudaStream_t stream,
                       cudaEvent_t ev0, cudaEvent_t ev1, bool early) {
    FT_REQUIRE(d == 64 || d == 128, "attn_decode: head_dim %d not supported (64 or 128)", d);
    FT_REQUIRE(nq % nkv == 0 && nq / nkv <= ATTN_GMAX, "attn_decode: GQA group %d/%d not supported", nq, nkv);
    FT_REQUIRE(B >= 1 && n_ctas >= 1, "attn_decode: B=%d n_ctas=%d", B, n_ctas);
    if (d == 64)
        return launch_attn_decode_d<64>(q, kv_layer, page_table, max_pages, ctx_lens, tile_prefix, out, out_lo,
                                        scratch, B, nq, nkv, n_ctas, pdl, stream, ev0, ev1, early);
    return launch_attn_decode_d<128>(q, kv_layer, page_table, max_pages, ctx_lens, tile_prefix, out, out_lo,
                                     scratch, B, nq, nkv, n_ctas, pdl, stream, ev0, ev1, early);
}

template <int D>
static int launch_attn_prefill_d(const void* q, const void* kv_layer, const int* page_table, int max_pages,
                                 const int* req_q_start, const int* req_slot, const int* req_pos0,
                                 const int* tile_req, const int* tile_q0, int n_tiles, void* out, void* out_lo,
                                 int nq, int nkv, bool pdl, cudaStream_t stream) {
    using C = PrefillCfg<D>;
    static PerDeviceFlag attr_set;
    if (!attr_set.done()) {
        FT_CHECK_CUDA(cudaFuncSetAttribute(attn_prefill_kernel<D>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                           (int)C::SMEM));
        FT_CHECK_CUDA(cudaFuncSetAttribute(attn_prefill_kernel<D>, cudaFuncAttributePreferredSharedMemoryCarveout,
                                           cudaSharedmemCarveoutMaxShared));
        attr_set.mark();
    }
    const float scale_log2 = LOG2E / sqrtf((float)D);
    cudaLaunchConfig_t cfg;
    cudaLaunchAttribute attr[1];
    fill_cfg(cfg, attr, dim3(n_tiles, nq), dim3(ATTN_THREADS), C::SMEM, pdl, stream);
    FT_CHECK_CUDA(cudaLaunchKernelEx(&cfg, attn_prefill_kernel<D>, (const __half*)q,
                                     (const uint8_t*)kv_layer, page_table, max_pages, req_q_start, req_slot,
                                     req_pos0, tile_req, tile_q0, (__nv_bfloat16*)out, (__nv_bfloat16*)out_lo, nq,
                                     nkv, scale_log2));
    return 0;
}

int launch_attn_prefill(const void* q, const void* kv_layer, const int* page_table, int max_pages,
                        const int* req_q_start, const int* req_slot, const int* req_pos0,
                        const int* tile_req, const int* tile_q0, int n_tiles, void* out, void* out_lo, int nq,
                        int nkv, int d, bool pdl, cudaStream_t stream, int q_rows) {
    FT_REQUIRE(d == 64 || d == 128, "attn_prefill: head_dim %d not supported (64 or 128)", d);
    FT_REQUIRE(nq % nkv == 0, "attn_prefill: nq %% nkv != 0");
    FT_REQUIRE(n_tiles >= 1, "attn_prefill: no tiles");
    if (q_rows > 0 && attn_prefill_tc_supported(nq, nkv, d))        // tcgen05 / TMEM kernel (attention_tc.cu)
        return launch_attn_prefill_tc(q, q_rows, kv_layer, page_table, max_pages, req_q_start, req_slot, req_pos0,
                                      tile_req, tile_q0, n_tiles, out, out_lo, nq, nkv, d, pdl, stream);
    if (d == 64)
        return launch_attn_prefill_d<64>(q, kv_layer, page_table, max_pages, req_q_start, req_slot, req_pos0,
                                         tile_req, tile_q0, n_tiles, out, out_lo, nq, nkv, pdl, stream);
    return launch_attn_prefill_d<128>(q, kv_layer, page_table, max_pages, req_q_start, req_slot, req_pos0,
                                      tile_req, tile_q0, n_tiles, out, out_lo, nq, nkv, pdl, stream);
}

}  // namespace ft

// ------------------------------------------------------------------ C ABI (kernel level)
extern "C" int ft_debug_attn_trace(unsigned long long* host_out, int n_ctas) {
    if (n_ctas > ft::ATTN_TRACE_CTAS) n_ctas = ft::ATTN_TRACE_CTAS;
    FT_CHECK_CUDA(cudaDeviceSynchronize());
    FT_CHECK_CUDA(cudaMemcpyFromSymbol(host_out, ft::g_attn_trace, sizeof(unsigned long long) * 4 * n_ctas));
    return 0;
}

extern "C" int ft_attn_decode_default_ctas(int d) {
    int dev = 0, sms = 0;
    if (cudaGetDevice(&dev) != cudaSuccess ||
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) {
        ft::set_last_error("ft_attn_decode_default_ctas: no CUDA device");
        return -1;
    }
    const int per = ft::attn_decode_ctas_per_sm(d);
    return per <= 0 ? per : per * sms;
}

extern "C" long long ft_attn_decode_scratch_bytes(int B, int nkv, int group, int d, int n_ctas) {
    return (long long)ft::attn_decode_scratch_floats(B, nkv, group, d, n_ctas) * 4;
}

extern "C" int ft_attn_decode(const void* q, const void* kv_layer, const int* page_table, int max_pages,
                              const int* ctx_lens, int* tile_prefix, void* out, void* out_lo, float* scratch,
                              int B, int nq, int nkv, int d, int n_ctas, void* stream) {
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (int r = ft::launch_attn_decode_plan(ctx_lens, tile_prefix, B, nkv, scratch, false, st)) return r;
    return ft::launch_attn_decode(q, kv_layer, page_table, max_pages, ctx_lens, tile_prefix, out, out_lo,
                                  scratch, B, nq, nkv, d, n_ctas, false, st, nullptr, nullptr, false);
}

extern "C" int ft_attn_prefill(const void* q, const void* kv_layer, const int* page_table, int max_pages,
                               const int* req_q_start, const int* req_slot, const int* req_pos0,
                               const int* tile_req, const int* tile_q0, int n_tiles, void* out, void* out_lo,
                               int nq, int nkv, int d, int q_rows, void* stream) {
    return ft::launch_attn_prefill(q, kv_layer, page_table, max_pages, req_q_start, req_slot, req_pos0,
                                   tile_req, tile_q0, n_tiles, out, out_lo, nq, nkv, d, false,
                                   static_cast<cudaStream_t>(stream), q_rows);
}

// ------------------------------------------------------------------ decode attention without allocations
// Some hosts hand this library device buffers but must not let it (or themselves) allocate while
// they are serving: a voice pipeline whose GPU memory is carved up once at start-up (the
// reference's deployment gives the engine a fixed share of the card, docker-compose.vllm.yml:44
// `--gpu-memory-utilization`), or a latency-critical audio thread for which cudaMalloc's implicit
// device-wide synchronisation is a dropout.  ft_attn_decode makes such callers bring their own
// work buffers; ft_attn_decode_noalloc needs none: its tile-prefix table, per-segment arrival
// counters and partial-result records live in the fixed array below, which every GPU that loads
// this library carries from module load until the process exits (about 6.1 MiB).
//
// The array is one per device and is shared by every call: it still holds whatever the previous
// call left in it.  Nothing is assumed about those contents — the planning kernel rewrites the
// prefix table and zeroes the arrival counters at the start of every call, and partial records
// are always written before they are read — but two calls that overlap on the same GPU
// (different streams, or different host threads) would trample each other's counters and
// records.  Issue them so that one finishes before the next starts.
namespace ft {

constexpr int NOALLOC_MAX_SEQS = 1024;        // B
constexpr int NOALLOC_MAX_SEGMENTS = 1024;    // B * nkv
constexpr int NOALLOC_MAX_CTAS = 512;
constexpr size_t NOALLOC_SCRATCH_FLOATS =
    NOALLOC_MAX_SEGMENTS + (size_t)(NOALLOC_MAX_CTAS + NOALLOC_MAX_SEGMENTS) * ATTN_GMAX * (128 + 2);
constexpr size_t NOALLOC_PREFIX_INTS = NOALLOC_MAX_SEQS + 64;

static __device__ __align__(256) float g_noalloc_scratch[NOALLOC_SCRATCH_FLOATS + NOALLOC_PREFIX_INTS];

// In host code `g_noalloc_scratch` is only a handle for the device symbol, not an address a
// kernel can use; the device address is looked up once per GPU, with that GPU current.
constexpr int NOALLOC_MAX_DEVICES = 64;
static float* g_noalloc_dev_ptr[NOALLOC_MAX_DEVICES] = {nullptr};
static std::mutex g_noalloc_mu;

static int noalloc_scratch_for_current_device(float** out) {
    int dev = -1;
    FT_CHECK_CUDA(cudaGetDevice(&dev));
    FT_REQUIRE(dev >= 0 && dev < NOALLOC_MAX_DEVICES, "ft_attn_decode_noalloc: device ordinal %d not supported", dev);
    std::lock_guard<std::mutex> lk(g_noalloc_mu);
    if (!g_noalloc_dev_ptr[dev]) {
        float* p = nullptr;
        FT_CHECK_CUDA(cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_noalloc_scratch));
        FT_REQUIRE(p != nullptr, "ft_attn_decode_noalloc: cudaGetSymbolAddress returned a null address");
        g_noalloc_dev_ptr[dev] = p;
    }
    *out = g_noalloc_dev_ptr[dev];
    return 0;
}

}  // namespace ft

extern "C" int ft_attn_decode_noalloc(const void* q, const void* kv_layer, const int* page_table, int max_pages,
                                      const int* ctx_lens, void* out, void* out_lo, int B, int nq, int nkv, int d,
                                      int n_ctas, void* stream) {
    using namespace ft;
    FT_REQUIRE(q && kv_layer && page_table && ctx_lens && out, "ft_attn_decode_noalloc: null argument");
    FT_REQUIRE(d == 64 || d == 128, "ft_attn_decode_noalloc: head_dim %d not supported (64 or 128)", d);
    FT_REQUIRE(B >= 1 && nkv >= 1 && nq >= nkv, "ft_attn_decode_noalloc: bad shape B=%d nq=%d nkv=%d", B, nq, nkv);
    if (n_ctas <= 0) {
        n_ctas = ft_attn_decode_default_ctas(d);
        if (n_ctas <= 0) return n_ctas;
    }
    FT_REQUIRE(B <= NOALLOC_MAX_SEQS && (long long)B * nkv <= NOALLOC_MAX_SEGMENTS && n_ctas <= NOALLOC_MAX_CTAS,
               "ft_attn_decode_noalloc: B=%d, B*nkv=%lld, n_ctas=%d exceed the built-in work area "
               "(at most %d sequences, %d (sequence, kv-head) pairs, %d CTAs); use ft_attn_decode with "
               "caller-provided scratch for larger batches",
               B, (long long)B * nkv, n_ctas, NOALLOC_MAX_SEQS, NOALLOC_MAX_SEGMENTS, NOALLOC_MAX_CTAS);
    const size_t need = attn_decode_scratch_floats(B, nkv, nq / nkv, d, n_ctas);
    FT_REQUIRE(need <= NOALLOC_SCRATCH_FLOATS,
               "ft_attn_decode_noalloc: needs %zu floats of work area, the built-in one holds %zu", need,
               (size_t)NOALLOC_SCRATCH_FLOATS);
    float* scratch = nullptr;
    if (int r = noalloc_scratch_for_current_device(&scratch)) return r;
    int* tile_prefix = reinterpret_cast<int*>(scratch + NOALLOC_SCRATCH_FLOATS);
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    // the planning kernel rewrites the prefix table and zeroes the arrival counters: nothing a
    // previous call left behind is relied upon
    if (int r = launch_attn_decode_plan(ctx_lens, tile_prefix, B, nkv, scratch, false, st)) return r;
    return launch_attn_decode(q, kv_layer, page_table, max_pages, ctx_lens, tile_prefix, out, out_lo, scratch, B,
                              nq, nkv, d, n_ctas, false, st, nullptr, nullptr, false);
}
