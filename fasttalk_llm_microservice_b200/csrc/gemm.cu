// K3 — bf16 projection GEMM on tcgen05 / TMEM, TMA-fed.   Y[t, n] = sum_k X[t, k] * W[n, k]
//
// The weight matrix is the UMMA "A" operand (128 output features fill the 128 TMEM lanes),
// the activations are the "B" operand (tokens = UMMA N, 16..256), so the decode case
// (T <= 64 tokens) is pure weight streaming: every CTA pulls its 128 x K_range slab of W
// through a ring of TMA stages exactly once (measured ~6 TB/s in steady state).  Split-K over
// gridDim.z gives one full wave of CTAs when n_out/128 is small.
//
// Activations come as TWO bf16 planes, hi = bf16(x) and lo = bf16(x - hi): both are multiplied
// by the same weight tile into the same TMEM accumulator (D += W*hi; D += W*lo).  That keeps
// ~16 mantissa bits of every activation through a pure bf16 x bf16 -> fp32 tensor-core product;
// the weight tile is already in shared memory, so in the HBM-bound decode regime the second
// product is free (DESIGN.md "precision").  lo may be absent (HILO = false).
//
// Epilogue (8 warps): phase 1 moves the fp32 accumulator tile TMEM -> [token][row] staging
// (shared memory, or — when K is split — this CTA's slab of an L2-resident scratch buffer);
// phase 2, for split K, is a rendezvous of the row tile's CTAs on a global arrival counter (all
// CTAs of the one-wave grid are co-resident); phase 3 walks tokens one per warp with float4
// rows, sums the split-K slabs for this CTA's share of the tokens and finishes the layer step
// right there (ft_gemm_epilogue.mode): residual add + next-norm statistics, 1/rms + RoPE + paged
// KV append, 1/rms + SwiGLU, or logits.  RMSNorm's 1/rms is applied AFTER the projection, so no
// stand-alone reduce / norm / RoPE kernel sits between the GEMMs of a layer.
//
// Warp roles (320 threads): warp 0 = TMA producer, warp 1 = TMEM owner + MMA issuer,
// warps 2..9 = epilogue.
#include "common.cuh"
#include "ft_internal.h"

#include <stdlib.h>
#include <string.h>

namespace ft {

constexpr int GEMM_BM = 128;      // W rows per CTA  (UMMA M)
constexpr int GEMM_BK = 64;       // bf16 per 128-byte swizzle row
constexpr int GEMM_A_BYTES = GEMM_BM * GEMM_BK * 2;
constexpr int GEMM_EPI_WARPS = 8;
constexpr int GEMM_EPI_THREADS = 32 * GEMM_EPI_WARPS;
constexpr int GEMM_THREADS = 64 + GEMM_EPI_THREADS;

__device__ __forceinline__ void epi_bar_sync() {
    asm volatile("bar.sync 2, %0;" ::"n"(GEMM_EPI_THREADS) : "memory");
}
__device__ __forceinline__ unsigned ld_acquire_sys_u32(const unsigned* p) {
    unsigned v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release_sys_u32(unsigned* p, unsigned v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned ld_acquire_u32(const unsigned* p) {
    unsigned v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

// Always-on, nearly free timeline of CTA (0,0,0) of the most recent GEMM launch (globaltimer ns):
// read back with ft_debug_gemm_trace (tuning aid, scripts/trace_gemm.py).
__device__ unsigned long long g_gemm_trace[16];
// ... and of the last 256 launches (ft_debug_gemm_trace_hist): row = launch number mod 256.  Launches
// overlap under PDL, so a launch's row index is fixed at entry and carried in shared memory.
constexpr int GEMM_TRACE_HIST = 256;
__device__ unsigned long long g_gemm_trace_hist[GEMM_TRACE_HIST * 16];
__device__ unsigned int g_gemm_trace_n;
__device__ __forceinline__ int trace_begin() {       // thread 0 of every CTA; returns -1 for untraced CTAs
    if (blockIdx.x != 0 || blockIdx.y != 0 || blockIdx.z != 0) return -1;
    return (int)(atomicAdd(&g_gemm_trace_n, 1u) % GEMM_TRACE_HIST);
}
__device__ __forceinline__ void trace(int slot, int base) {
    if (base >= 0) {
        unsigned long long t;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
        g_gemm_trace[slot] = t;
        g_gemm_trace_hist[base * 16 + slot] = t;
    }
}

__device__ __forceinline__ float4 f4_mul(float4 a, float s) { return make_float4(a.x * s, a.y * s, a.z * s, a.w * s); }
__device__ __forceinline__ float4 f4_add(float4 a, float4 b) {
    return make_float4(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w);
}
__device__ __forceinline__ float4 f4_shfl_xor(float4 a, int m) {
    return make_float4(__shfl_xor_sync(0xffffffffu, a.x, m), __shfl_xor_sync(0xffffffffu, a.y, m),
                       __shfl_xor_sync(0xffffffffu, a.z, m), __shfl_xor_sync(0xffffffffu, a.w, m));
}
// four fp32 -> bf16 hi plane (8 bytes) and bf16 lo plane = rounding residual (8 bytes)
__device__ __forceinline__ void store_hi_lo4(__nv_bfloat16* hi, __nv_bfloat16* lo, size_t idx, float4 y) {
    uint2 h;
    h.x = pack_bf16(y.x, y.y);
    h.y = pack_bf16(y.z, y.w);
    *reinterpret_cast<uint2*>(hi + idx) = h;
    if (lo) {
        uint2 l;
        l.x = pack_bf16(y.x - bf16_lo(h.x), y.y - bf16_hi(h.x));
        l.y = pack_bf16(y.z - bf16_lo(h.y), y.w - bf16_hi(h.y));
        *reinterpret_cast<uint2*>(lo + idx) = l;
    }
}
// four fp32 -> one saturating fp16 plane (8 bytes): the single-plane input of the large prefill projections
__device__ __forceinline__ void store_f16x4(__half* p, size_t idx, float4 y) {
    uint2 h;
    h.x = pack_f16_sat(y.x, y.y);
    h.y = pack_f16_sat(y.z, y.w);
    *reinterpret_cast<uint2*>(p + idx) = h;
}
// explicit shared-memory loads: the staging / rstd pointers reach the epilogue as generic pointers, and
// generic loads interleaved with global stores cost ~0.5 us per token row here (scripts/trace_decode_gemms.py)
__device__ __forceinline__ float4 lds_f4(uint32_t a) {
    float4 v;
    asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(a));
    return v;
}
__device__ __forceinline__ float2 lds_f2(uint32_t a) {
    float2 v;
    asm volatile("ld.shared.v2.f32 {%0,%1}, [%2];" : "=f"(v.x), "=f"(v.y) : "r"(a));
    return v;
}
__device__ __forceinline__ float lds_f(uint32_t a) {
    float v;
    asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(a));
    return v;
}
__device__ __forceinline__ float silu(float g) { return __fdividef(g, 1.f + __expf(-g)); }

// Sum of one token row's four floats over the n_split partial slabs.  All loads of a group of eight are
// issued before the first add: with a run-time trip count the compiler chained load -> add -> load, one L2
// round trip per slab (a 6-way QKV split spent ~3 us here).
template <int BN>
__device__ __forceinline__ float4 sum_slabs(const float* first, int n_split) {
    const float4* src = reinterpret_cast<const float4*>(first);
    constexpr size_t STRIDE = (size_t)BN * GEMM_BM / 4;
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int p0 = 0; p0 < n_split; p0 += 8) {
        float4 v[8];
#pragma unroll
        for (int k = 0; k < 8; ++k)
            v[k] = p0 + k < n_split ? __ldcg(src + (size_t)(p0 + k) * STRIDE) : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
        for (int k = 0; k < 8; ++k) acc = f4_add(acc, v[k]);     // fixed order: slab 0, 1, 2, ...
    }
    return acc;
}

// Where a CTA's output tile sits and how its K range was split (the stand-alone kernel takes these
// from the block index, the persistent kernel from its tile loop).
struct EpiCtx {
    int row_tile, tok_tile, split, n_split, tile_id;
    unsigned cta_linear;     // arrival word of this CTA in the tensor-parallel exchange
    unsigned n_ctas;         // CTAs of the launch (tiles x splits)
    int n_out, T;
    bool has_acc;            // false: K range empty, the tile is all zeros
    int trace_base;          // row of the launch in the trace history, -1 = this CTA is not traced
};

// per-token 1/rms of the activation's source (RMSNorm deferred past the projection), all BN tokens of the tile
template <int BN>
__device__ __forceinline__ void epilogue_rstd(const ft_gemm_epilogue& ep, const EpiCtx& c, float* rstd_s) {
    const int et = threadIdx.x - 64;
    const int t0 = c.tok_tile * BN, T = c.T;
    // per-token 1/rms of the activation's source (RMSNorm deferred past the projection)
    for (int j = et; j < BN; j += GEMM_EPI_THREADS) {
        float r = 1.f;
        const int t = t0 + j;
        if (ep.ssq_in && t < T) {
            float s = 0.f;
            const float* sp = ep.ssq_in + t;
            int p = 0;
            for (; p + 8 <= ep.ssq_parts; p += 8) {       // eight loads in flight, summed in the same order as one by one
                float v[8];
#pragma unroll
                for (int k = 0; k < 8; ++k) v[k] = __ldcg(sp + (size_t)(p + k) * ep.ssq_stride);
#pragma unroll
                for (int k = 0; k < 8; ++k) s += v[k];
            }
            for (; p < ep.ssq_parts; ++p) s += __ldcg(sp + (size_t)p * ep.ssq_stride);
            r = rsqrtf(s * ep.inv_h + ep.eps);
        }
        rstd_s[j] = r;
    }
}

// The epilogue (warps 2..9) for token columns [col0, col0 + CB) of the tile: accumulator out of TMEM,
// split-K rendezvous (CB == BN only), fused layer step.  stage_f: [CB][128] fp32 of shared memory
// nobody else touches while this runs; rstd_s from epilogue_rstd; the caller has waited for the
// accumulator (tmem_acc = its first column).
template <int BN, int CB, int MODE>
__device__ __forceinline__ void gemm_epilogue_m(const ft_gemm_epilogue& ep, const EpiCtx& c, uint32_t tmem_acc,
                                                float* stage_f, const float* rstd_s, int col0) {
    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    constexpr int mode = MODE;      // compile-time: the token loop carries no dispatch (see gemm_epilogue below)
    const int row0 = c.row_tile * GEMM_BM, t0 = c.tok_tile * BN;
    const int split = c.split, n_split = c.n_split, n_out = c.n_out, T = c.T;
    const int ew = warp - 2;                // 0..7
    const int q = warp & 3;                 // TMEM lane quarter this warp may read (warp_id % 4)
    const int et = threadIdx.x - 64;        // 0..255
    const bool reduce = (mode == FT_EPI_RESID || mode == FT_EPI_QKV_ROPE) && n_split > 1;

    // ---- this CTA's share of the tile's tokens, and the global operands of the first token batch: residual
    // rows, or position -> RoPE angles / KV page.  They depend on nothing this launch computes, so they are
    // requested BEFORE the accumulator is touched and their L2 latency hides behind phases 1 and 2 (1.5 us of
    // the 3.4 us a QKV tile's token loop took, scripts/trace_decode_gemms.py).  Tokens are taken U at a time
    // per warp; later batches are requested while the current one is being finished.
    const int jbeg = reduce ? BN * split / n_split : col0;
    const int jend = reduce ? BN * (split + 1) / n_split : col0 + CB;
    const int r4 = 4 * lane;
    const int row = row0 + r4;
    constexpr int U = BN >= 128 ? 4 : 1;
    constexpr bool is_resid = mode == FT_EPI_RESID, is_rope = mode == FT_EPI_QKV_ROPE;
    // (head_dim is 64 or 128, checked at launch: masks and shifts instead of divisions in the token loop)
    const int d_ = ep.head_dim, half_ = d_ >> 1, d_shift_ = d_ == 128 ? 7 : 6;
    const int q_dim_ = ep.nq * d_, kv_dim_ = ep.nkv * d_;
    const int region = row0 < q_dim_ ? 0 : row0 < q_dim_ + kv_dim_ ? 1 : 2;   // tile-uniform
    const int i_ = is_rope ? row & (d_ - 1) : 0;
    const int ih_ = i_ & (half_ - 1);                                         // rotate_half: cos / sin column
    // K / V rows: slab of (page, kv head, K|V) = kv_layer + page * page_bytes_ + head_off_; inside it the row's 16-byte
    // chunk is swizzled with the token's row (kv_swizzled_chunk in common.cuh)
    const size_t slab_bytes_ = (size_t)kKvTile * d_ * 2;
    const size_t page_bytes_ = slab_bytes_ * 2 * ep.nkv;
    const size_t head_off_ = region == 0 ? 0 : ((size_t)((row - q_dim_ - (region == 2 ? kv_dim_ : 0)) >> d_shift_) * 2 + (region == 2 ? 1 : 0)) * slab_bytes_;
    const int kv_off_ = (i_ >> 6) * (kKvTile * 128) + (i_ & 7) * 2, kv_chunk_ = (i_ >> 3) & 7;
    float4 pre_r[U], pre_c[U], pre_s[U];
    int pre_pos[U], pre_page[U];
    auto prefetch = [&](int j0) {
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const int j = j0 + u * GEMM_EPI_WARPS, t = t0 + j;
            pre_pos[u] = 0;
            if (j < jend && t < T) {
                if (is_resid) pre_r[u] = __ldcg(reinterpret_cast<const float4*>(ep.resid + (size_t)t * n_out + row));
                if (is_rope) pre_pos[u] = ep.positions[t];
            }
        }
        if (is_rope) {
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const int j = j0 + u * GEMM_EPI_WARPS, t = t0 + j;
                if (j < jend && t < T) {
                    if (region < 2) {
                        pre_c[u] = *reinterpret_cast<const float4*>(ep.rope_cos + (size_t)pre_pos[u] * half_ + ih_);
                        pre_s[u] = *reinterpret_cast<const float4*>(ep.rope_sin + (size_t)pre_pos[u] * half_ + ih_);
                    }
                    if (region > 0)
                        pre_page[u] = ep.page_table[(size_t)ep.seq_rows[t] * ep.max_pages + pre_pos[u] / kKvTile];
                }
            }
        }
    };
    if (mode != FT_EPI_PARTIAL) prefetch(jbeg + ew);

    // ---- phase 1: accumulator tile TMEM -> [token][row] staging
    const int tile_id = c.tile_id;
    float* slabs = reduce ? ep.scratch + (size_t)tile_id * n_split * (BN * GEMM_BM) : nullptr;
    {
        constexpr int CPW = CB >= 32 ? CB / 2 : 16;     // columns per warp: two warps share a lane quarter
        const int cbeg = (ew >> 2) * CPW;
        const int rl = q * 32 + lane;                   // tile row this thread reads out of TMEM
#pragma unroll 1
        for (int c0 = cbeg; c0 < cbeg + CPW && c0 < CB; c0 += 16) {
            if (t0 + col0 + c0 >= T) break;
            float v[16];
            if (c.has_acc) {
                tmem_ld16(tmem_acc + ((uint32_t)(q * 32) << 16) + col0 + c0, v);
            } else {
#pragma unroll
                for (int j = 0; j < 16; ++j) v[j] = 0.f;
            }
            // three store flavours, each with compile-time strides so that no address
            // arithmetic is left in the unrolled body
            if (mode == FT_EPI_PARTIAL) {
                float* p = reinterpret_cast<float*>(ep.out) + ((size_t)split * T + t0 + col0 + c0) * n_out + row0 + rl;
                const int valid = T - (t0 + col0 + c0);
#pragma unroll
                for (int j = 0; j < 16; ++j) {
                    if (j < valid) *p = v[j];
                    p += n_out;
                }
            } else if (reduce) {
                float* p = slabs + ((size_t)split * BN + c0) * GEMM_BM + rl;
#pragma unroll
                for (int j = 0; j < 16; ++j) __stcg(p + j * GEMM_BM, v[j]);
            } else {
                const uint32_t a = smem_u32(stage_f) + (uint32_t)(c0 * GEMM_BM + rl) * 4;
#pragma unroll
                for (int j = 0; j < 16; ++j)
                    asm volatile("st.shared.f32 [%0], %1;" ::"r"(a + j * GEMM_BM * 4), "f"(v[j]) : "memory");
            }
        }
    }
    if (et == 0) trace(7, c.trace_base);

    if (mode != FT_EPI_PARTIAL) {
        // ---- phase 2: make the tile visible (CTA barrier, or split-K rendezvous in global memory)
        // Arrival word of the row tile.  sync_mode 0: word 0 counts arrivals and re-arms itself (word 1 counts
        // CTAs that are done; the last one zeroes both) — any launch order works.  sync_mode 1 / 2: launches that
        // share the buffer alternate between word 0 and word 1 and every CTA zeroes the OTHER word on its way
        // out of the rendezvous, which takes the returning atomic off the kernel's tail (engine).
        unsigned* cnt = reinterpret_cast<unsigned*>(ep.sync) + 2 * tile_id;
        const int par = ep.sync_mode;
        unsigned* arrive = par ? cnt + (par - 1) : cnt;
        if (reduce) {
            epi_bar_sync();                  // every warp's slab stores precede thread 0's release below
            if (et == 0) {
                __threadfence();             // cumulative over the CTA (bar.sync above), as in attention.cu
                atomicAdd(arrive, 1u);
                unsigned spins = 0;
                while (ld_acquire_u32(arrive) < (unsigned)n_split) {
                    __nanosleep(32);
                    if (++spins > (1u << 25)) __trap();      // a peer never arrived: fail loudly, do not hang
                }
                if (par)                     // nobody looks at the other words during this launch: zero ALL of them,
                    for (unsigned k = c.cta_linear; k < (unsigned)ep.sync_slots; k += c.n_ctas)   // not just this tile's
                        reinterpret_cast<unsigned*>(ep.sync)[2 * k + (2 - par)] = 0;
            }
            epi_bar_sync();
        } else {
            epi_bar_sync();
        }

        // ---- phase 3: one token per warp, four consecutive rows per lane
        if (et == 0) trace(11, c.trace_base);      // rendezvous passed / tile staged: token loop starts
        const uint32_t stage_a = smem_u32(stage_f), rstd_a = smem_u32(rstd_s);
        float4 gam = make_float4(1.f, 1.f, 1.f, 1.f);
        if (mode == FT_EPI_RESID) {
            const uint2 gv = *reinterpret_cast<const uint2*>(reinterpret_cast<const __nv_bfloat16*>(ep.gamma) + row);
            gam = make_float4(bf16_lo(gv.x), bf16_hi(gv.x), bf16_lo(gv.y), bf16_hi(gv.y));
        }
        // ---- tensor-parallel exchange (RESID, 2 GPUs): this CTA's partial rows go to the peer over
        // NVLink, the peer's twin CTA (same block index, same tile and token share) sends its own
        const bool tp = mode == FT_EPI_RESID && ep.tp_peer_xchg != nullptr;
        if (tp) {
            for (int j = jbeg + ew; j < jend; j += GEMM_EPI_WARPS) {
                const int t = t0 + j;
                if (t >= T) break;
                float4 y;
                if (reduce) {
                    y = sum_slabs<BN>(slabs + (size_t)j * GEMM_BM + r4, n_split);
                } else {
                    y = lds_f4(stage_a + (uint32_t)((j - col0) * GEMM_BM + r4) * 4);
                }
                *reinterpret_cast<float4*>(ep.tp_peer_xchg + (size_t)t * n_out + row) = y;
            }
            epi_bar_sync();                  // every warp's peer stores precede thread 0's system-scope release
            if (et == 0) {
                const unsigned cta = c.cta_linear;
                __threadfence_system();      // cumulative over the CTA (bar.sync above)
                st_release_sys_u32(ep.tp_peer_flags + cta, ep.tp_epoch);
                unsigned spins = 0;
                while ((int)(ld_acquire_sys_u32(ep.tp_flags + cta) - ep.tp_epoch) < 0) {
                    __nanosleep(32);
                    if (++spins > (1u << 25)) __trap();      // the peer GPU never ran this launch: fail loudly
                }
            }
            epi_bar_sync();
        }
        for (int j0 = jbeg + ew; j0 < jend; j0 += GEMM_EPI_WARPS * U) {
#pragma unroll
            for (int u = 0; u < U; ++u) {
            const int j = j0 + u * GEMM_EPI_WARPS;
            const int t = t0 + j;
            if (j >= jend || t >= T) break;
            float4 y;
            if (reduce) {
                y = sum_slabs<BN>(slabs + (size_t)j * GEMM_BM + r4, n_split);
            } else {
                y = lds_f4(stage_a + (uint32_t)((j - col0) * GEMM_BM + r4) * 4);
            }
            const float rs = lds_f(rstd_a + (uint32_t)j * 4);
            if (mode == FT_EPI_RESID) {
                const size_t idx = (size_t)t * n_out + row;
                if (tp) y = f4_add(y, __ldcg(reinterpret_cast<const float4*>(ep.tp_xchg + idx)));   // a + b == b + a
                float4 r = f4_add(pre_r[u], y);
                *reinterpret_cast<float4*>(ep.resid + idx) = r;
                const float4 rg = make_float4(r.x * gam.x, r.y * gam.y, r.z * gam.z, r.w * gam.w);
                if (ep.out) store_hi_lo4(reinterpret_cast<__nv_bfloat16*>(ep.out), reinterpret_cast<__nv_bfloat16*>(ep.out_lo), idx, rg);
                if (ep.out_f16) store_f16x4(reinterpret_cast<__half*>(ep.out_f16), idx, rg);
                const float ss = warp_sum(r.x * r.x + r.y * r.y + r.z * r.z + r.w * r.w);
                if (lane == 0) ep.ssq_out[(size_t)c.row_tile * ep.ssq_stride + t] = ss;
            } else if (mode == FT_EPI_QKV_ROPE) {
                const int half = half_, q_dim = q_dim_, i = i_;
                const int pos = pre_pos[u];
                float4 val = f4_mul(y, rs);
                if (region < 2) {
                    // rotate_half partner (i +- d/2) sits half/4 lanes away in the same warp
                    const float4 b = f4_shfl_xor(val, half >> 2);
                    const float4 c = pre_c[u], s = pre_s[u];
                    const float sg = i < half ? -1.f : 1.f;
                    val = make_float4(val.x * c.x + sg * b.x * s.x, val.y * c.y + sg * b.y * s.y,
                                      val.z * c.z + sg * b.z * s.z, val.w * c.w + sg * b.w * s.w);
                }
                uint2 h4;                       // q, K and V leave as saturating fp16 (attention dtype)
                h4.x = pack_f16_sat(val.x, val.y);
                h4.y = pack_f16_sat(val.z, val.w);
                if (region == 0) {
                    *reinterpret_cast<uint2*>(reinterpret_cast<__half*>(ep.q_out) + (size_t)t * q_dim + row) = h4;
                } else {
                    const int r = pos % kKvTile;
                    uint8_t* slab = reinterpret_cast<uint8_t*>(ep.kv_layer) + (size_t)pre_page[u] * page_bytes_ + head_off_;
                    // sub-slab of 64 dims, 128-byte rows, chunks swizzled with the row
                    const int off = kv_off_ + r * 128 + ((kv_chunk_ ^ (r & 7)) << 4);
                    *reinterpret_cast<uint2*>(slab + off) = h4;
                }
            } else if (mode == FT_EPI_SWIGLU) {
                // rows 0..63 of the tile are gate features, rows 64..127 the matching up features:
                // every lane finishes two features (all 32 lanes busy, no shuffles)
                const float2 g2 = lds_f2(stage_a + (uint32_t)((j - col0) * GEMM_BM + 2 * lane) * 4);
                const float2 u2 = lds_f2(stage_a + (uint32_t)((j - col0) * GEMM_BM + 64 + 2 * lane) * 4);
                const float a0 = silu(g2.x * rs) * (u2.x * rs), a1 = silu(g2.y * rs) * (u2.y * rs);
                const uint32_t h = pack_bf16(a0, a1);
                const size_t idx = (size_t)t * (n_out >> 1) + c.row_tile * 64 + 2 * lane;
                if (ep.out) *reinterpret_cast<uint32_t*>(reinterpret_cast<__nv_bfloat16*>(ep.out) + idx) = h;
                if (ep.out_lo)
                    *reinterpret_cast<uint32_t*>(reinterpret_cast<__nv_bfloat16*>(ep.out_lo) + idx) =
                        pack_bf16(a0 - bf16_lo(h), a1 - bf16_hi(h));
                if (ep.out_f16) *reinterpret_cast<uint32_t*>(reinterpret_cast<__half*>(ep.out_f16) + idx) = pack_f16_sat(a0, a1);
            } else if (mode == FT_EPI_LOGITS) {
                const size_t idx = (size_t)t * (ep.out_stride ? ep.out_stride : n_out) + ep.out_col0 + row;
                const float4 lg = f4_mul(y, rs);
                *reinterpret_cast<float4*>(reinterpret_cast<float*>(ep.out) + idx) = lg;
                if (ep.tp_peer_out) *reinterpret_cast<float4*>(reinterpret_cast<float*>(ep.tp_peer_out) + idx) = lg;
            } else {   // FT_EPI_BF16
                store_hi_lo4(reinterpret_cast<__nv_bfloat16*>(ep.out), reinterpret_cast<__nv_bfloat16*>(ep.out_lo),
                             (size_t)t * n_out + row, f4_mul(y, rs));
            }
        }
        if (j0 + GEMM_EPI_WARPS * U < jend) prefetch(j0 + GEMM_EPI_WARPS * U);
        }
        if (reduce && !par) {
            // the last CTA of the row tile to finish re-arms the counters for the next launch
            epi_bar_sync();
            if (et == 0 && atomicAdd(cnt + 1, 1u) == (unsigned)n_split - 1) {
                cnt[0] = 0;
                cnt[1] = 0;
            }
        }
    }
}

// One specialised copy of the epilogue per mode, chosen once per tile: with the mode a run-time value
// the token loop went through a jump table in constant memory on every row, and with ~0.7 KB of kernel
// parameters re-read by uniform loads around it the loop ran at ~0.4 us per token row
// (scripts/trace_decode_gemms.py: 3.4 us for 8 SwiGLU rows with no global access in them).
template <int BN, int CB>
__device__ __forceinline__ void gemm_epilogue(const ft_gemm_epilogue& ep, const EpiCtx& c, uint32_t tmem_acc,
                                              float* stage_f, const float* rstd_s, int col0) {
    switch (ep.mode) {
        case FT_EPI_PARTIAL: gemm_epilogue_m<BN, CB, FT_EPI_PARTIAL>(ep, c, tmem_acc, stage_f, rstd_s, col0); break;
        case FT_EPI_BF16: gemm_epilogue_m<BN, CB, FT_EPI_BF16>(ep, c, tmem_acc, stage_f, rstd_s, col0); break;
        case FT_EPI_SWIGLU: gemm_epilogue_m<BN, CB, FT_EPI_SWIGLU>(ep, c, tmem_acc, stage_f, rstd_s, col0); break;
        case FT_EPI_RESID: gemm_epilogue_m<BN, CB, FT_EPI_RESID>(ep, c, tmem_acc, stage_f, rstd_s, col0); break;
        case FT_EPI_QKV_ROPE: gemm_epilogue_m<BN, CB, FT_EPI_QKV_ROPE>(ep, c, tmem_acc, stage_f, rstd_s, col0); break;
        default: gemm_epilogue_m<BN, CB, FT_EPI_LOGITS>(ep, c, tmem_acc, stage_f, rstd_s, col0); break;
    }
}

template <int BN, bool HILO, bool F16 = false>
__global__ void __launch_bounds__(GEMM_THREADS, 1)
gemm_bf16_tcgen05_kernel(const __grid_constant__ CUtensorMap map_w,
                         const __grid_constant__ CUtensorMap map_x,
                         const __grid_constant__ CUtensorMap map_x_lo,
                         const __grid_constant__ ft_gemm_epilogue ep, int n_out, int T, int k_chunks,
                         int stages) {
    constexpr int B_PLANE = BN * GEMM_BK * 2;
    constexpr int B_BYTES = (HILO ? 2 : 1) * B_PLANE;
    constexpr int STAGE_BYTES = GEMM_A_BYTES + B_BYTES;
    constexpr uint32_t TMEM_COLS = BN < 32 ? 32 : BN;

    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                               ~uintptr_t(1023));
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + (size_t)stages * STAGE_BYTES);
    uint64_t* empty_bar = full_bar + stages;
    uint64_t* tmem_full_bar = empty_bar + stages;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_full_bar + 1);
    float* rstd_s = reinterpret_cast<float*>(tmem_slot + 2);     // [BN] per-token 1/rms

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const int row0 = blockIdx.x * GEMM_BM;
    const int t0 = blockIdx.y * BN;
    const int split = blockIdx.z;
    const int n_split = gridDim.z;
    const int kc_begin = (int)(((long long)k_chunks * split) / n_split);
    const int kc_end = (int)(((long long)k_chunks * (split + 1)) / n_split);
    const int n_iter = kc_end - kc_begin;
    int& s_trace_base = *reinterpret_cast<int*>(tmem_slot + 1);     // spare word of the dynamic shared memory
    if (threadIdx.x == 0) {
        s_trace_base = trace_begin();
        trace(0, s_trace_base);
        if (s_trace_base >= 0) g_gemm_trace_hist[s_trace_base * 16 + 10] = ((unsigned long long)ep.mode << 32) | (unsigned)n_out;
    }

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&map_w);
        tma_prefetch_desc(&map_x);
        if (HILO) tma_prefetch_desc(&map_x_lo);
        for (int s = 0; s < stages; ++s) {
            mbar_init(&full_bar[s], 1);
            mbar_init(&empty_bar[s], 1);
        }
        mbar_init(tmem_full_bar, 1);
        fence_mbar_init();
    }
    if (warp == 1) {
        tmem_alloc(tmem_slot, TMEM_COLS);
        tmem_relinquish();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    if (threadIdx.x == 0) trace(1, s_trace_base);

    pdl_launch_dependents();

    if (warp == 0) {
        if (elect_one()) {
            // One token tile (decode): weights are streamed once per step, the activations are re-read
            // by every row tile.  Several token tiles (prefill): row tiles vary fastest, so a wave of
            // CTAs shares one or two activation tiles and sweeps ALL of W, which the next token tile
            // sweeps again — W is the L2-resident set (67 MB at most for the 1B gate/up projection).
            const bool multi = gridDim.y > 1;
            const uint64_t pol_w = multi ? policy_evict_last() : policy_evict_first();
            const uint64_t pol_x = multi ? policy_evict_first() : policy_evict_last();
            // Weights do not depend on the previous kernel: start streaming them before the
            // programmatic-dependency wait, then add the activation tiles.
            const int pre = n_iter < stages ? n_iter : stages;
            for (int i = 0; i < pre; ++i) {
                mbar_expect_tx(&full_bar[i], STAGE_BYTES);
                tma_tile_2d(smem + (size_t)i * STAGE_BYTES, &map_w, (kc_begin + i) * GEMM_BK, row0,
                            &full_bar[i], pol_w);
            }
            // Keep HBM busy across the kernel boundary: pull this CTA's share of the NEXT
            // projection's weights into L2 (126 MB) while this kernel waits on its producer.
            if (ep.prefetch_bytes > 0) {
                const unsigned long long n_cta = (unsigned long long)gridDim.x * gridDim.y * gridDim.z;
                const unsigned long long cta = ((unsigned long long)blockIdx.z * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x;
                const unsigned long long total = (unsigned long long)ep.prefetch_bytes;
                const unsigned long long per = ((total / n_cta) + 16383ull) & ~16383ull;
                const unsigned long long beg = per * cta;
                const unsigned long long end = beg + per < total ? beg + per : total;
                const uint8_t* base = reinterpret_cast<const uint8_t*>(ep.prefetch);
                for (unsigned long long off = beg; off < end; off += 16384ull) {
                    const uint32_t sz = (uint32_t)(end - off < 16384ull ? end - off : 16384ull) & ~15u;
                    if (sz) asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(base + off), "r"(sz) : "memory");
                }
            }
            trace(2, s_trace_base);
            pdl_wait();
            trace(3, s_trace_base);
            for (int i = 0; i < pre; ++i) {
                uint8_t* bs = smem + (size_t)i * STAGE_BYTES + GEMM_A_BYTES;
                tma_tile_2d(bs, &map_x, (kc_begin + i) * GEMM_BK, t0, &full_bar[i], pol_x);
                if (HILO) tma_tile_2d(bs + B_PLANE, &map_x_lo, (kc_begin + i) * GEMM_BK, t0, &full_bar[i], pol_x);
            }
            int s = 0;
            uint32_t ph = 0;      // parity of the completed "empty" phase we wait for
            for (int i = pre; i < n_iter; ++i) {
                mbar_wait(&empty_bar[s], ph);
                mbar_expect_tx(&full_bar[s], STAGE_BYTES);
                uint8_t* st = smem + (size_t)s * STAGE_BYTES;
                tma_tile_2d(st, &map_w, (kc_begin + i) * GEMM_BK, row0, &full_bar[s], pol_w);
                tma_tile_2d(st + GEMM_A_BYTES, &map_x, (kc_begin + i) * GEMM_BK, t0, &full_bar[s], pol_x);
                if (HILO)
                    tma_tile_2d(st + GEMM_A_BYTES + B_PLANE, &map_x_lo, (kc_begin + i) * GEMM_BK, t0,
                                &full_bar[s], pol_x);
                if (++s == stages) { s = 0; ph ^= 1; }
            }
        }
    } else if (warp == 1) {
        if (elect_one()) {
            constexpr uint32_t idesc = F16 ? umma_idesc_f16(GEMM_BM, BN) : umma_idesc_bf16(GEMM_BM, BN);
            int s = 0;
            uint32_t ph = 0;
            for (int i = 0; i < n_iter; ++i) {
                mbar_wait(&full_bar[s], ph);
                if (i == 0) trace(4, s_trace_base);
                tc_fence_after();
                const uint32_t a_addr = smem_u32(smem + (size_t)s * STAGE_BYTES);
                const uint64_t a_desc = umma_desc_k128(a_addr);
                const uint64_t b_desc = umma_desc_k128(a_addr + GEMM_A_BYTES);
#pragma unroll
                for (int k = 0; k < GEMM_BK / 16; ++k)   // UMMA K = 16 bf16 = 32 B -> +2 in addr>>4
                    umma_bf16(tmem_base, a_desc + 2 * k, b_desc + 2 * k, idesc, (i | k) != 0);
                if (HILO) {
                    const uint64_t l_desc = umma_desc_k128(a_addr + GEMM_A_BYTES + B_PLANE);
#pragma unroll
                    for (int k = 0; k < GEMM_BK / 16; ++k)
                        umma_bf16(tmem_base, a_desc + 2 * k, l_desc + 2 * k, idesc, 1u);
                }
                umma_commit(&empty_bar[s]);               // frees the smem stage when MMAs retire
                if (++s == stages) { s = 0; ph ^= 1; }
            }
            trace(5, s_trace_base);
            umma_commit(tmem_full_bar);
        }
    } else {
        // ================================================================ epilogue warps (2..9)
        pdl_wait();                             // ssq_in / resid / positions come from earlier kernels
        EpiCtx c;
        c.row_tile = blockIdx.x; c.tok_tile = blockIdx.y; c.split = split; c.n_split = n_split;
        c.tile_id = blockIdx.y * gridDim.x + blockIdx.x;
        c.cta_linear = (blockIdx.z * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x;
        c.n_ctas = gridDim.x * gridDim.y * gridDim.z;
        c.n_out = n_out; c.T = T; c.has_acc = n_iter > 0; c.trace_base = s_trace_base;
        epilogue_rstd<BN>(ep, c, rstd_s);
        if (n_iter > 0) {
            mbar_wait(tmem_full_bar, 0);
            tc_fence_after();
        }
        if (threadIdx.x == 64) trace(6, s_trace_base);
        // the pipeline stages are idle once the MMAs retired: the staging tile aliases them
        gemm_epilogue<BN, BN>(ep, c, tmem_base, reinterpret_cast<float*>(smem), rstd_s, 0);
        if (threadIdx.x == 64) trace(8, s_trace_base);
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, TMEM_COLS);
    }
    if (threadIdx.x == 0) trace(9, s_trace_base);
}

// ------------------------------------------------------------------ persistent variant (multi-wave grids)
// Grids with more tiles than SMs — the prefill projections (up to 8192 tiles) and the LM head (1002
// row tiles) — run as one CTA per SM that walks tiles (row tiles fastest).  What the walk buys over
// one CTA per tile: the TMA ring never drains between tiles, TMEM holds TWO accumulators so the MMAs
// of tile i+1 run while the epilogue warps are still finishing tile i, and barriers / TMEM are set up
// once.  The epilogue works on 64 token columns at a time through its own 32 KB staging buffer (the
// ring is busy with the next tile), which leaves four 48 KB stages for 128-token tiles.  K is never
// split here (every tile's epilogue is self-contained).
template <int BN>
__global__ void __launch_bounds__(GEMM_THREADS, 1)
gemm_persistent_kernel(const __grid_constant__ CUtensorMap map_w, const __grid_constant__ CUtensorMap map_x,
                       const __grid_constant__ CUtensorMap map_x_lo, const __grid_constant__ ft_gemm_epilogue ep,
                       int n_out, int T, int k_chunks, int stages) {
    constexpr int CB = BN < 64 ? BN : 64;
    constexpr int B_PLANE = BN * GEMM_BK * 2;
    constexpr int STAGE_BYTES = GEMM_A_BYTES + 2 * B_PLANE;
    constexpr uint32_t TMEM_COLS = 2 * BN < 32 ? 32 : 2 * BN;

    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    float* stage_f = reinterpret_cast<float*>(smem + (size_t)stages * STAGE_BYTES);
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(reinterpret_cast<uint8_t*>(stage_f) + CB * GEMM_BM * 4);
    uint64_t* empty_bar = full_bar + stages;
    uint64_t* tmem_full_bar = empty_bar + stages;        // [2]
    uint64_t* tmem_empty_bar = tmem_full_bar + 2;        // [2]
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_empty_bar + 2);
    float* rstd_s = reinterpret_cast<float*>(tmem_slot + 2);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const int n_row_tiles = n_out / GEMM_BM;
    const int n_tok_tiles = (T + BN - 1) / BN;
    const int n_tiles = n_row_tiles * n_tok_tiles;

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&map_w);
        tma_prefetch_desc(&map_x);
        tma_prefetch_desc(&map_x_lo);
        for (int s = 0; s < stages; ++s) {
            mbar_init(&full_bar[s], 1);
            mbar_init(&empty_bar[s], 1);
        }
        for (int a = 0; a < 2; ++a) {
            mbar_init(&tmem_full_bar[a], 1);
            mbar_init(&tmem_empty_bar[a], 1);
        }
        fence_mbar_init();
    }
    if (warp == 1) {
        tmem_alloc(tmem_slot, TMEM_COLS);
        tmem_relinquish();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    pdl_launch_dependents();

    if (warp == 0) {
        if (elect_one()) {
            // several token tiles: W is the L2-resident set; one token tile (LM head): W streams once
            const bool multi = n_tok_tiles > 1;
            const uint64_t pol_w = multi ? policy_evict_last() : policy_evict_first();
            const uint64_t pol_x = multi ? policy_evict_first() : policy_evict_last();
            int it = 0;
            bool waited = false;
            for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
                const int row0 = (tile % n_row_tiles) * GEMM_BM, t0 = (tile / n_row_tiles) * BN;
                int i = 0;
                if (!waited) {
                    // weights do not depend on the previous kernel: request the first ring-full before the wait
                    const int pre = k_chunks < stages ? k_chunks : stages;
                    for (int j = 0; j < pre; ++j) {
                        mbar_expect_tx(&full_bar[j], STAGE_BYTES);
                        tma_tile_2d(smem + (size_t)j * STAGE_BYTES, &map_w, j * GEMM_BK, row0, &full_bar[j], pol_w);
                    }
                    pdl_wait();
                    waited = true;
                    for (int j = 0; j < pre; ++j) {
                        uint8_t* bs = smem + (size_t)j * STAGE_BYTES + GEMM_A_BYTES;
                        tma_tile_2d(bs, &map_x, j * GEMM_BK, t0, &full_bar[j], pol_x);
                        tma_tile_2d(bs + B_PLANE, &map_x_lo, j * GEMM_BK, t0, &full_bar[j], pol_x);
                    }
                    i = pre;
                    it = pre;
                }
                for (; i < k_chunks; ++i, ++it) {
                    const int slot = it % stages;
                    if (it >= stages) mbar_wait(&empty_bar[slot], ((it / stages) - 1) & 1);
                    mbar_expect_tx(&full_bar[slot], STAGE_BYTES);
                    uint8_t* st = smem + (size_t)slot * STAGE_BYTES;
                    tma_tile_2d(st, &map_w, i * GEMM_BK, row0, &full_bar[slot], pol_w);
                    tma_tile_2d(st + GEMM_A_BYTES, &map_x, i * GEMM_BK, t0, &full_bar[slot], pol_x);
                    tma_tile_2d(st + GEMM_A_BYTES + B_PLANE, &map_x_lo, i * GEMM_BK, t0, &full_bar[slot], pol_x);
                }
            }
            if (!waited) pdl_wait();
        }
    } else if (warp == 1) {
        if (elect_one()) {
            constexpr uint32_t idesc = umma_idesc_bf16(GEMM_BM, BN);
            int ct = 0;
            uint32_t use = 0;
            for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, ++use) {
                const uint32_t acc = use & 1;
                if (use >= 2) {                       // the epilogue must have emptied this accumulator
                    mbar_wait(&tmem_empty_bar[acc], ((use >> 1) - 1) & 1);
                    tc_fence_after();
                }
                const uint32_t d_tmem = tmem_base + acc * BN;
                for (int i = 0; i < k_chunks; ++i, ++ct) {
                    const int slot = ct % stages;
                    mbar_wait(&full_bar[slot], (ct / stages) & 1);
                    tc_fence_after();
                    const uint32_t a_addr = smem_u32(smem + (size_t)slot * STAGE_BYTES);
                    const uint64_t a_desc = umma_desc_k128(a_addr);
                    const uint64_t b_desc = umma_desc_k128(a_addr + GEMM_A_BYTES);
                    const uint64_t l_desc = umma_desc_k128(a_addr + GEMM_A_BYTES + B_PLANE);
#pragma unroll
                    for (int k = 0; k < GEMM_BK / 16; ++k)
                        umma_bf16(d_tmem, a_desc + 2 * k, b_desc + 2 * k, idesc, (i | k) != 0);
#pragma unroll
                    for (int k = 0; k < GEMM_BK / 16; ++k)
                        umma_bf16(d_tmem, a_desc + 2 * k, l_desc + 2 * k, idesc, 1u);
                    umma_commit(&empty_bar[slot]);
                }
                umma_commit(&tmem_full_bar[acc]);
            }
        }
    } else {
        const int et = threadIdx.x - 64;
        pdl_wait();
        uint32_t use = 0;
        for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, ++use) {
            const uint32_t acc = use & 1;
            EpiCtx c;
            c.row_tile = tile % n_row_tiles; c.tok_tile = tile / n_row_tiles; c.split = 0; c.n_split = 1;
            c.tile_id = tile; c.cta_linear = (unsigned)tile; c.n_ctas = (unsigned)n_tiles;
            c.n_out = n_out; c.T = T; c.has_acc = true; c.trace_base = -1;
            epilogue_rstd<BN>(ep, c, rstd_s);
            mbar_wait(&tmem_full_bar[acc], (use >> 1) & 1);
            tc_fence_after();
            for (int col0 = 0; col0 < BN; col0 += CB) {
                if (c.tok_tile * BN + col0 >= T) break;
                gemm_epilogue<BN, CB>(ep, c, tmem_base + acc * BN, stage_f, rstd_s, col0);
                tc_fence_before();
                epi_bar_sync();                       // staging (and, after the last half, the accumulator) is free
            }
            if (et == 0) mbar_arrive(&tmem_empty_bar[acc]);
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, TMEM_COLS);
    }
}

// ------------------------------------------------------------------ CTA-pair variant (cta_group::2) for prefill
// Two SMs of a TPC form a cluster and issue ONE 256 x 128 MMA per step: each CTA holds its own 128 weight
// rows (A) and only HALF of the 128-token activation tile (B, 64 tokens x hi/lo); the tensor core reads the
// other half from the peer's shared memory.  Per flop that is half the activation traffic into each SM's
// shared memory, 32 KB instead of 48 KB per stage (six stages instead of four), and one MMA issue stream for
// two SMs.  Everything else is the persistent kernel: tile walk (here over row-tile PAIRS), two accumulators
// per CTA, per-mode epilogue on the CTA's own 128 rows.  Barriers: TMA of both CTAs counts on the LEADER's
// full barrier; tcgen05.commit multicasts "stage free" / "accumulator ready" to both; both epilogues report
// "accumulator drained" to the leader.
template <int BN, bool HILO, bool F16 = false>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(GEMM_THREADS, 1)
gemm_pair_kernel(const __grid_constant__ CUtensorMap map_w, const __grid_constant__ CUtensorMap map_x_half,
                 const __grid_constant__ CUtensorMap map_x_lo_half, const __grid_constant__ ft_gemm_epilogue ep,
                 int n_out, int T, int k_chunks, int stages) {
    constexpr int HALF = BN / 2, CB = 64;      // BN = 128 or 256 tokens per tile, half of them in each CTA
    constexpr int B_PLANE = HALF * GEMM_BK * 2;
    constexpr int STAGE_BYTES = GEMM_A_BYTES + (HILO ? 2 : 1) * B_PLANE;          // per CTA
    constexpr uint32_t TMEM_COLS = 2 * BN;

    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    float* stage_f = reinterpret_cast<float*>(smem + (size_t)stages * STAGE_BYTES);
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(reinterpret_cast<uint8_t*>(stage_f) + CB * GEMM_BM * 4);
    uint64_t* empty_bar = full_bar + stages;
    uint64_t* tmem_full_bar = empty_bar + stages;        // [2]
    uint64_t* tmem_empty_bar = tmem_full_bar + 2;        // [2], the leader's copy counts both CTAs
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_empty_bar + 2);
    float* rstd_s = reinterpret_cast<float*>(tmem_slot + 2);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const int rank = (int)cluster_ctarank();
    const int n_row_pairs = n_out / (2 * GEMM_BM);
    const int n_tok_tiles = (T + BN - 1) / BN;
    const int cl = (int)cluster_id_x(), n_cl = (int)gridDim.x / 2;
    // Work list.  MIXED (256-token tiles of ONE fp16 plane): whole rounds of the clusters take 256-token units; what is
    // left over — less than one round — is cut into 128-token halves when those fit one round of their own, so the
    // last round costs half a tile instead of a whole one (1B down projection at 8192 tokens: 256 units on 74
    // clusters = 3 rounds + 34 units -> 68 halves).  map_x_lo_half is then the 64-token box of the same plane.
    constexpr bool MIXED = BN == 256 && F16 && !HILO;
    const int n_units = n_row_pairs * n_tok_tiles;
    int n_big = n_units, n_small = 0;
    if (MIXED) {
        const int rem = n_units % n_cl;
        if (rem > 0 && 2 * rem <= n_cl) {
            n_big = n_units - rem;
            n_small = 2 * rem;
        }
    }
    const int n_work = n_big + n_small;
    // unit w -> row pair, first token, 128-token half?
    auto unit = [&](int w, int& row_pair, int& tok0) -> bool {
        if (!MIXED || w < n_big) {
            row_pair = w % n_row_pairs;
            tok0 = (w / n_row_pairs) * BN;
            return false;
        }
        const int s = w - n_big, b = n_big + (s >> 1);
        row_pair = b % n_row_pairs;
        tok0 = (b / n_row_pairs) * BN + (s & 1) * HALF;
        return true;
    };

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&map_w);
        tma_prefetch_desc(&map_x_half);
        if (HILO || MIXED) tma_prefetch_desc(&map_x_lo_half);
        for (int s = 0; s < stages; ++s) {
            mbar_init(&full_bar[s], 1);
            mbar_init(&empty_bar[s], 1);
        }
        for (int a = 0; a < 2; ++a) {
            mbar_init(&tmem_full_bar[a], 1);
            mbar_init(&tmem_empty_bar[a], 2);
        }
        fence_mbar_init();
    }
    if (warp == 1) {
        tmem_alloc_pair(tmem_slot, TMEM_COLS);
        tmem_relinquish_pair();
    }
    tc_fence_before();
    cluster_sync_all();                     // both CTAs' barriers and TMEM exist before anything crosses over
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    pdl_launch_dependents();

    if (warp == 0) {
        if (elect_one()) {
            const bool multi = n_tok_tiles > 1;
            const uint64_t pol_w = multi ? policy_evict_last() : policy_evict_first();
            const uint64_t pol_x = multi ? policy_evict_first() : policy_evict_last();
            pdl_wait();
            int it = 0;
            for (int w = cl; w < n_work; w += n_cl) {
                int row_pair, tok_first;
                const bool small = unit(w, row_pair, tok_first);
                const int row0 = (2 * row_pair + rank) * GEMM_BM;
                const int tok0 = tok_first + rank * (small ? HALF / 2 : HALF);   // this CTA's half of the token tile
                const CUtensorMap* mx = small ? &map_x_lo_half : &map_x_half;
                const uint32_t bytes = small ? GEMM_A_BYTES + B_PLANE / 2 : STAGE_BYTES;
                for (int i = 0; i < k_chunks; ++i, ++it) {
                    const int slot = it % stages;
                    if (it >= stages) mbar_wait(&empty_bar[slot], ((it / stages) - 1) & 1);
                    if (rank == 0) mbar_expect_tx(&full_bar[slot], 2 * bytes);            // both CTAs' bytes
                    uint8_t* st = smem + (size_t)slot * STAGE_BYTES;
                    tma_tile_2d_pair(st, &map_w, i * GEMM_BK, row0, &full_bar[slot], pol_w);
                    tma_tile_2d_pair(st + GEMM_A_BYTES, mx, i * GEMM_BK, tok0, &full_bar[slot], pol_x);
                    if (HILO)
                        tma_tile_2d_pair(st + GEMM_A_BYTES + B_PLANE, &map_x_lo_half, i * GEMM_BK, tok0,
                                         &full_bar[slot], pol_x);
                }
            }
        }
    } else if (warp == 1 && rank == 0) {
        if (elect_one()) {
            constexpr uint32_t idesc_full = F16 ? umma_idesc_f16(2 * GEMM_BM, BN) : umma_idesc_bf16(2 * GEMM_BM, BN);
            constexpr uint32_t idesc_half = umma_idesc_f16(2 * GEMM_BM, HALF);
            int ct = 0;
            uint32_t use = 0;
            for (int w = cl; w < n_work; w += n_cl, ++use) {
                const uint32_t acc = use & 1;
                const uint32_t idesc = MIXED && w >= n_big ? idesc_half : idesc_full;
                if (use >= 2) {
                    mbar_wait(&tmem_empty_bar[acc], ((use >> 1) - 1) & 1);
                    tc_fence_after();
                }
                const uint32_t d_tmem = tmem_base + acc * BN;
                for (int i = 0; i < k_chunks; ++i, ++ct) {
                    const int slot = ct % stages;
                    mbar_wait(&full_bar[slot], (ct / stages) & 1);
                    tc_fence_after();
                    const uint32_t a_addr = smem_u32(smem + (size_t)slot * STAGE_BYTES);
                    const uint64_t a_desc = umma_desc_k128(a_addr);
                    const uint64_t b_desc = umma_desc_k128(a_addr + GEMM_A_BYTES);
#pragma unroll
                    for (int k = 0; k < GEMM_BK / 16; ++k)
                        umma_bf16_pair(d_tmem, a_desc + 2 * k, b_desc + 2 * k, idesc, (i | k) != 0);
                    if (HILO) {
                        const uint64_t l_desc = umma_desc_k128(a_addr + GEMM_A_BYTES + B_PLANE);
#pragma unroll
                        for (int k = 0; k < GEMM_BK / 16; ++k)
                            umma_bf16_pair(d_tmem, a_desc + 2 * k, l_desc + 2 * k, idesc, 1u);
                    }
                    umma_commit_pair(&empty_bar[slot]);
                }
                umma_commit_pair(&tmem_full_bar[acc]);
            }
        }
    } else if (warp >= 2) {
        const int et = threadIdx.x - 64;
        pdl_wait();
        uint32_t use = 0;
        for (int w = cl; w < n_work; w += n_cl, ++use) {
            const uint32_t acc = use & 1;
            EpiCtx c;
            int row_pair, tok_first;
            const bool small = unit(w, row_pair, tok_first);
            c.row_tile = 2 * row_pair + rank; c.tok_tile = tok_first / (small ? HALF : BN); c.split = 0; c.n_split = 1;
            c.tile_id = (tok_first / BN) * (n_out / GEMM_BM) + c.row_tile; c.cta_linear = (unsigned)c.tile_id;
            c.n_ctas = (unsigned)(n_work * 2);
            c.n_out = n_out; c.T = T; c.has_acc = true; c.trace_base = -1;
            if (MIXED && small) {
                epilogue_rstd<HALF>(ep, c, rstd_s);
                mbar_wait(&tmem_full_bar[acc], (use >> 1) & 1);
                tc_fence_after();
                for (int col0 = 0; col0 < HALF; col0 += CB) {
                    if (tok_first + col0 >= T) break;
                    gemm_epilogue<HALF, CB>(ep, c, tmem_base + acc * BN, stage_f, rstd_s, col0);
                    tc_fence_before();
                    epi_bar_sync();
                }
            } else {
                epilogue_rstd<BN>(ep, c, rstd_s);
                mbar_wait(&tmem_full_bar[acc], (use >> 1) & 1);
                tc_fence_after();
                for (int col0 = 0; col0 < BN; col0 += CB) {
                    if (tok_first + col0 >= T) break;
                    gemm_epilogue<BN, CB>(ep, c, tmem_base + acc * BN, stage_f, rstd_s, col0);
                    tc_fence_before();
                    epi_bar_sync();
                }
            }
            if (et == 0) mbar_arrive_leader(&tmem_empty_bar[acc]);
        }
    }

    tc_fence_before();
    cluster_sync_all();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc_pair(tmem_base, TMEM_COLS);
    }
}

// ------------------------------------------------------------------ host side
typedef CUresult (*PFN_encodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                    const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                    const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                    CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static PFN_encodeTiled get_encode_fn() {
    static PFN_encodeTiled fn = nullptr;
    if (!fn) {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult qres;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) ==
                cudaSuccess &&
            qres == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<PFN_encodeTiled>(p);
    }
    return fn;
}

// rows x K bf16 row-major matrix; box = box_rows x 64 columns, 128-byte swizzle.
int make_tmap_bf16_k64(CUtensorMap* map, const void* base, int rows, int K, int box_rows) {
    PFN_encodeTiled enc = get_encode_fn();
    FT_REQUIRE(enc != nullptr, "cuTensorMapEncodeTiled not available from the driver");
    FT_REQUIRE((reinterpret_cast<uintptr_t>(base) & 15) == 0, "tensor map base not 16-byte aligned");
    cuuint64_t dims[2] = {(cuuint64_t)K, (cuuint64_t)rows};
    cuuint64_t strides[1] = {(cuuint64_t)K * 2};
    cuuint32_t box[2] = {(cuuint32_t)GEMM_BK, (cuuint32_t)box_rows};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = enc(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims,
                     strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                     CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    FT_REQUIRE(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled failed (%d) rows=%d K=%d box_rows=%d",
               (int)r, rows, K, box_rows);
    return 0;
}

// fp16 tensor [d2][d1][d0] (d0 contiguous), box b0 x b1 x b2, 128-byte swizzle (b0 * 2 bytes must be 128).
int make_tmap_f16_3d(CUtensorMap* map, const void* base, int d0, int d1, int d2, int b0, int b1, int b2) {
    PFN_encodeTiled enc = get_encode_fn();
    FT_REQUIRE(enc != nullptr, "cuTensorMapEncodeTiled not available from the driver");
    FT_REQUIRE((reinterpret_cast<uintptr_t>(base) & 15) == 0 && b0 * 2 == 128, "3-D tensor map: bad base / box");
    cuuint64_t dims[3] = {(cuuint64_t)d0, (cuuint64_t)d1, (cuuint64_t)d2};
    cuuint64_t strides[2] = {(cuuint64_t)d0 * 2, (cuuint64_t)d0 * d1 * 2};
    cuuint32_t box[3] = {(cuuint32_t)b0, (cuuint32_t)b1, (cuuint32_t)b2};
    cuuint32_t estr[3] = {1, 1, 1};
    CUresult r = enc(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 3, const_cast<void*>(base), dims, strides, box, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    FT_REQUIRE(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled (3-D) failed (%d) dims=%d,%d,%d", (int)r, d0, d1, d2);
    return 0;
}

static int device_sms() {          // of the CURRENT device
    static int sms[kMaxDevices] = {};
    const int slot = current_device_slot();
    if (!sms[slot]) {
        int dev = 0, n = 0;
        if (cudaGetDevice(&dev) != cudaSuccess ||
            cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0)
            n = 148;
        sms[slot] = n;
    }
    return sms[slot];
}

// Fewest 128 x 128 tiles a launch needs before the persistent / CTA-pair kernels take it.  One token tile: more tiles
// than SMs.  Two or more token tiles (mixed decode + prefill steps of a few hundred rows): from 32 tiles on — a pair
// loads 24 KB per k-step and CTA instead of 32 KB, which is what bounds these single-wave grids (1B, 56 decode rows +
// 968 prompt tokens: O and down projection 32 -> 27 us each, the step 4.42 -> 4.13 ms; profiles/r02_mixed_step.md).
static int pair_min_tiles(int T) {
    if (T <= 128) return device_sms() + 1;
    if (const char* ov = getenv("FT_GEMM_PAIR_MIN_TILES")) return atoi(ov);       // A/B switch (scripts/)
    return 32;
}

bool gemm_pair_eligible(int n_out, int T) {
    static const bool off = getenv("FT_GEMM_NO_PAIR") != nullptr || getenv("FT_GEMM_NO_PERSISTENT") != nullptr;
    const int bn = gemm_pick_bn(T, true);
    return !off && bn == 128 && n_out % (2 * GEMM_BM) == 0 && (n_out / GEMM_BM) * ((T + bn - 1) / bn) >= pair_min_tiles(T);
}

// Does launch_gemm have a one-fp16-plane form (x_planes = 3) of this launch?  128-token tiles, K unsplit, and either the
// CTA-pair kernel or the tile-per-CTA kernel would run it (the persistent tile walk is hi + lo only).
bool gemm_f16_path_ok(int n_out, int T, int K, int splits) {
    static const bool no_persistent = getenv("FT_GEMM_NO_PERSISTENT") != nullptr;
    const int bn = gemm_pick_bn(T, true);
    if (bn != 128 || splits != 1) return false;
    if (gemm_pair_eligible(n_out, T)) return true;
    const bool multiwave = !no_persistent && (n_out / GEMM_BM) * ((T + bn - 1) / bn) > device_sms();
    return !(multiwave && K <= 4096);
}

// Token tile: with both activation planes a 256-token tile would leave only two pipeline
// stages in 227 KB of shared memory, so the hi/lo path tops out at 128.
int gemm_pick_bn(int T, bool hilo) {
    const int cap = hilo ? 128 : 256;
    const int bn = T <= 16 ? 16 : T <= 32 ? 32 : T <= 64 ? 64 : T <= 128 ? 128 : 256;
    return bn > cap ? cap : bn;
}

// Split K so that (row tiles x token tiles x splits) fills ONE wave of CTAs (one CTA per SM):
// a second, partial wave costs a whole extra prologue + epilogue on these microsecond kernels
// (profiles/r01_launches_decode_v1.md: 160- and 256-CTA grids ran in two waves).
int gemm_pick_splits(int n_out, int T, int K, int num_sms, bool hilo) {
    const int bn = gemm_pick_bn(T, hilo);
    const int tiles = (n_out / GEMM_BM) * ((T + bn - 1) / bn);
    const int kc = K / GEMM_BK;
    if (tiles >= num_sms) return 1;
    int s = num_sms / tiles;
    const int max_s = kc / 2 > 0 ? kc / 2 : 1;      // at least two K-chunks per CTA
    if (s > max_s) s = max_s;
    if (s > GEMM_MAX_SPLITS) s = GEMM_MAX_SPLITS;
    return s < 1 ? 1 : s;
}

template <int BN, bool HILO>
static int stage_plan(int n_out, int T, int K, int splits, int* stages_out, size_t* smem_out) {
    constexpr int STAGE_BYTES = GEMM_A_BYTES + (HILO ? 2 : 1) * BN * GEMM_BK * 2;
    const int kc = K / GEMM_BK;
    const int per_cta = (kc + splits - 1) / splits;
    const int tiles = (n_out / GEMM_BM) * ((T + BN - 1) / BN) * splits;
    // One CTA per SM with the deepest ring that fits: measured on the 8192-token prefill step, two
    // co-resident CTAs with two-stage rings ran the projections 1.45x slower than one CTA with four
    // stages (the ring, not epilogue overlap, is what keeps the tensor pipe fed).
    int budget = 208 * 1024;
    if (const char* ov = getenv("FT_GEMM_MW_KB"))         // tuning experiments: smem budget of multi-wave grids
        if (tiles > device_sms() && atoi(ov) >= 64 && atoi(ov) <= 208) budget = atoi(ov) * 1024;
    int stages = budget / STAGE_BYTES;
    if (stages > 12) stages = 12;
    if (const char* ov = getenv("FT_GEMM_STAGES")) {      // tuning experiments (scripts/bench_gemm.py)
        const int v = atoi(ov);
        if (v >= 2 && v <= stages) stages = v;
    }
    if (stages > per_cta) stages = per_cta;
    // A CTA that streams at most six chunks gains nothing from a deep ring, and a shallow one (66 KB) lets a
    // decode-attention CTA (96 KB ring) become resident on the same SM while this launch is still running:
    // the attention after the QKV projection starts streaming KV early, the O projection after it loads its
    // weights before the attention's last CTAs leave (1.724 -> 1.702 ms per 64-session step).
    // (only when there is an attention worth overlapping: with a handful of rows the deep ring wins, it has all
    //  of the CTA's chunks in flight at once)
    if (per_cta <= 6 && T > 16 && stages > 2) stages = 2;
    if (stages < 2) stages = 2;
    // the epilogue parks a [BN][128] fp32 tile in the (idle) pipeline memory
    while ((size_t)stages * STAGE_BYTES < (size_t)BN * GEMM_BM * 4) ++stages;
    *stages_out = stages;
    *smem_out = (size_t)stages * STAGE_BYTES + 1024 + (2 * stages + 1) * 8 + 16 + BN * 4 + 16;
    FT_REQUIRE(*smem_out <= 227 * 1024, "gemm: %zu bytes of shared memory needed", *smem_out);
    return 0;
}

template <int BN, bool HILO, bool F16 = false>
static int launch_gemm_bn(const CUtensorMap& mw, const CUtensorMap& mx, const CUtensorMap& mxl, int n_out,
                          int T, int K, int splits, const ft_gemm_epilogue& epi, bool pdl, cudaStream_t stream) {
    int stages = 0;
    size_t smem = 0;
    if (int r = stage_plan<BN, HILO>(n_out, T, K, splits, &stages, &smem)) return r;
    static PerDeviceFlag attr_set;
    if (!attr_set.done()) {
        FT_CHECK_CUDA(cudaFuncSetAttribute(gemm_bf16_tcgen05_kernel<BN, HILO, F16>,
                                           cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
        // without this the driver sizes the shared-memory carve-out for ONE CTA of a ~100 KB launch and
        // the second CTA the multi-wave plan counts on never becomes resident (ncu: occupancy limit 1)
        FT_CHECK_CUDA(cudaFuncSetAttribute(gemm_bf16_tcgen05_kernel<BN, HILO, F16>,
                                           cudaFuncAttributePreferredSharedMemoryCarveout,
                                           cudaSharedmemCarveoutMaxShared));
        attr_set.mark();
    }
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(n_out / GEMM_BM, (T + BN - 1) / BN, splits);
    cfg.blockDim = dim3(GEMM_THREADS);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = stream;
    cudaLaunchAttribute attr[2];
    int na = 0;
    if (pdl) {
        attr[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        attr[na].val.programmaticStreamSerializationAllowed = 1;
        ++na;
    }
    cfg.attrs = attr;
    cfg.numAttrs = na;
    FT_CHECK_CUDA(cudaLaunchKernelEx(&cfg, gemm_bf16_tcgen05_kernel<BN, HILO, F16>, mw, mx, mxl, epi, n_out, T,
                                     K / GEMM_BK, stages));
    return 0;
}

template <int BN>
static int launch_persistent_bn(const CUtensorMap& mw, const CUtensorMap& mx, const CUtensorMap& mxl, int n_out,
                                int T, int K, const ft_gemm_epilogue& epi, bool pdl, cudaStream_t stream) {
    constexpr int CB = BN < 64 ? BN : 64;
    constexpr int STAGE_BYTES = GEMM_A_BYTES + 2 * BN * GEMM_BK * 2;
    constexpr int STAGING = CB * GEMM_BM * 4;
    const size_t fixed = 1024 + 256 + BN * 4 + STAGING;
    int stages = (int)((227 * 1024 - fixed) / STAGE_BYTES);
    if (stages > 12) stages = 12;
    if (stages > K / GEMM_BK && K / GEMM_BK >= 2) stages = K / GEMM_BK;
    FT_REQUIRE(stages >= 2, "gemm (persistent): no room for a two-stage ring");
    const size_t smem = (size_t)stages * STAGE_BYTES + STAGING + 1024 + (2 * stages + 4) * 8 + 16 + BN * 4 + 16;
    FT_REQUIRE(smem <= 227 * 1024, "gemm (persistent): %zu bytes of shared memory needed", smem);
    static PerDeviceFlag attr_set;
    if (!attr_set.done()) {
        FT_CHECK_CUDA(cudaFuncSetAttribute(gemm_persistent_kernel<BN>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                           227 * 1024));
        attr_set.mark();
    }
    const int tiles = (n_out / GEMM_BM) * ((T + BN - 1) / BN);
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(tiles < device_sms() ? tiles : device_sms());
    cfg.blockDim = dim3(GEMM_THREADS);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    int na = 0;
    if (pdl) {
        attr[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        attr[na].val.programmaticStreamSerializationAllowed = 1;
        ++na;
    }
    cfg.attrs = attr;
    cfg.numAttrs = na;
    FT_CHECK_CUDA(cudaLaunchKernelEx(&cfg, gemm_persistent_kernel<BN>, mw, mx, mxl, epi, n_out, T, K / GEMM_BK, stages));
    return 0;
}

template <int BN, bool HILO, bool F16 = false>
static int launch_pair(const CUtensorMap& mw, const CUtensorMap& mx_half, const CUtensorMap& mxl_half, int n_out, int T,
                       int K, const ft_gemm_epilogue& epi, bool pdl, cudaStream_t stream) {
    constexpr int STAGE_BYTES = GEMM_A_BYTES + (HILO ? 2 : 1) * (BN / 2) * GEMM_BK * 2;
    constexpr int STAGING = 64 * GEMM_BM * 4;
    const size_t fixed = 1024 + 256 + BN * 4 + STAGING;
    int stages = (int)((227 * 1024 - fixed) / STAGE_BYTES);
    if (stages > 8) stages = 8;
    if (stages > K / GEMM_BK && K / GEMM_BK >= 2) stages = K / GEMM_BK;
    const size_t smem = (size_t)stages * STAGE_BYTES + STAGING + 1024 + (2 * stages + 4) * 8 + 16 + BN * 4 + 16;
    FT_REQUIRE(smem <= 227 * 1024, "gemm (pair): %zu bytes of shared memory needed", smem);
    static int max_clusters_dev[kMaxDevices] = {};
    int& max_clusters = max_clusters_dev[current_device_slot()];
    if (!max_clusters) {
        FT_CHECK_CUDA(cudaFuncSetAttribute(gemm_pair_kernel<BN, HILO, F16>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
        cudaLaunchConfig_t q = {};
        q.gridDim = dim3(device_sms() & ~1);
        q.blockDim = dim3(GEMM_THREADS);
        q.dynamicSmemBytes = smem;
        cudaLaunchAttribute qa[1];
        qa[0].id = cudaLaunchAttributeClusterDimension;
        qa[0].val.clusterDim.x = 2; qa[0].val.clusterDim.y = 1; qa[0].val.clusterDim.z = 1;
        q.attrs = qa;
        q.numAttrs = 1;
        int n = 0;
        if (cudaOccupancyMaxActiveClusters(&n, gemm_pair_kernel<BN, HILO, F16>, &q) != cudaSuccess || n <= 0) n = device_sms() / 2;
        max_clusters = n;
    }
    const int n_work = (n_out / (2 * GEMM_BM)) * ((T + BN - 1) / BN);
    const int clusters = n_work < max_clusters ? n_work : max_clusters;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(2 * clusters);     // (the kernel derives its work list — and the MIXED split — from the grid)
    cfg.blockDim = dim3(GEMM_THREADS);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    int na = 0;
    if (pdl) {
        attr[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        attr[na].val.programmaticStreamSerializationAllowed = 1;
        ++na;
    }
    cfg.attrs = attr;
    cfg.numAttrs = na;
    FT_CHECK_CUDA(cudaLaunchKernelEx(&cfg, gemm_pair_kernel<BN, HILO, F16>, mw, mx_half, mxl_half, epi, n_out, T, K / GEMM_BK, stages));
    return 0;
}

size_t gemm_scratch_floats(int n_out, int T, int splits, bool hilo) {
    const int bn = gemm_pick_bn(T, hilo);
    return (size_t)(n_out / GEMM_BM) * ((T + bn - 1) / bn) * splits * bn * GEMM_BM;
}

int launch_gemm(const CUtensorMap& mw, const CUtensorMap& mx, const CUtensorMap* mx_lo, int bn, int n_out,
                int T, int K, int splits, const ft_gemm_epilogue& epi, bool pdl, cudaStream_t stream,
                const CUtensorMap* mx_half, const CUtensorMap* mx_lo_half) {
    FT_REQUIRE(n_out % GEMM_BM == 0, "gemm: n_out=%d must be a multiple of %d", n_out, GEMM_BM);
    FT_REQUIRE(K % GEMM_BK == 0, "gemm: K=%d must be a multiple of %d", K, GEMM_BK);
    FT_REQUIRE(T >= 1, "gemm: T=%d", T);
    FT_REQUIRE(splits >= 1 && splits <= GEMM_MAX_SPLITS && splits <= K / GEMM_BK,
               "gemm: bad split count %d", splits);
    const int mode = epi.mode;
    FT_REQUIRE(mode >= FT_EPI_PARTIAL && mode <= FT_EPI_LOGITS, "gemm: bad epilogue mode %d", mode);
    const bool staged = mode == FT_EPI_RESID || mode == FT_EPI_QKV_ROPE;
    FT_REQUIRE(mode == FT_EPI_PARTIAL || staged || splits == 1, "gemm: this epilogue needs splits == 1");
    const bool hilo = mx_lo != nullptr;
    FT_REQUIRE(mode <= FT_EPI_BF16 || hilo, "gemm: the fused epilogues need both activation planes");
    if (staged && splits > 1) {
        // the split-K rendezvous spins on peers of the same row tile: every CTA must be resident
        const int ctas = (n_out / GEMM_BM) * ((T + bn - 1) / bn) * splits;
        FT_REQUIRE(ctas <= device_sms(), "gemm: %d CTAs do not fit one wave of %d SMs (split-K rendezvous)", ctas,
                   device_sms());
        FT_REQUIRE(epi.scratch && epi.sync, "gemm: split-K fused epilogues need scratch and sync buffers");
        FT_REQUIRE(epi.sync_mode >= 0 && epi.sync_mode <= 2 &&
                       (epi.sync_mode == 0 || epi.sync_slots >= (n_out / GEMM_BM) * ((T + bn - 1) / bn)),
                   "gemm: sync_mode %d needs sync_slots covering the launch's tiles", epi.sync_mode);
    }
    FT_REQUIRE(!(hilo && bn > 128), "gemm: hi/lo activations support token tiles up to 128");
    if (epi.tp_peer_xchg)
        FT_REQUIRE(mode == FT_EPI_RESID && epi.tp_xchg && epi.tp_peer_flags && epi.tp_flags,
                   "gemm: the tensor-parallel exchange belongs to the RESID epilogue and needs all four buffers");
    if (mode == FT_EPI_LOGITS && epi.out_stride)
        FT_REQUIRE(epi.out_stride % 4 == 0 && epi.out_col0 % 4 == 0 && epi.out_col0 + n_out <= epi.out_stride,
                   "gemm: logits slice [%d, %d) does not fit rows of %d", epi.out_col0, epi.out_col0 + n_out,
                   epi.out_stride);
    if (mode == FT_EPI_QKV_ROPE) {
        FT_REQUIRE(epi.head_dim == 64 || epi.head_dim == 128, "gemm: head_dim %d not supported", epi.head_dim);
        FT_REQUIRE((epi.nq * epi.head_dim) % GEMM_BM == 0 && (epi.nkv * epi.head_dim) % GEMM_BM == 0 &&
                       n_out == (epi.nq + 2 * epi.nkv) * epi.head_dim,
                   "gemm: q / kv widths must be multiples of 128 and add up to n_out");
    }
    // more tiles than SMs, K unsplit: one persistent CTA per SM walks the tiles
    static const bool no_persistent = getenv("FT_GEMM_NO_PERSISTENT") != nullptr;     // A/B switch (scripts/)
    // (not for K > 4096: a 128-chunk mainloop hides nothing behind its epilogue, and the ring competing
    //  with the overlapped epilogue's traffic ran the 1B down projection 1.2x slower than tile-per-CTA)
    const bool multiwave = hilo && splits == 1 && mode != FT_EPI_PARTIAL && !no_persistent &&
                           (n_out / GEMM_BM) * ((T + bn - 1) / bn) >= pair_min_tiles(T);
    // 128-token tiles with the 64-token maps at hand: CTA pairs (cta_group::2), two row tiles per MMA
    static const bool no_pair = getenv("FT_GEMM_NO_PAIR") != nullptr;                 // A/B switch (scripts/)
    if (multiwave && !no_pair && bn == 128 && mx_half && mx_lo_half && n_out % (2 * GEMM_BM) == 0) {
        // 256-token tiles (each CTA holds 128 of them: the bn == 128 maps ARE the half maps) when there are enough tokens
        // — the 256 x 256 MMA reads each weight tile once per 256 tokens (gate/up: 753 -> 706 us), but halves the
        // number of work units: only when at least eight rounds of them remain per cluster (measured: with 3–5
        // rounds the down and QKV projections lose more to the last, partly empty round than the shape gains)
        const long long work256 = (long long)(n_out / (2 * GEMM_BM)) * ((T + 255) / 256);
        // x_planes == 1: the multi-wave (prefill) tiles multiply the hi plane only — half the tensor work; the
        // latency-bound decode tiles, where the second product is free, always use both
        const bool one = epi.x_planes == 1;
        if (epi.x_planes == 3) {      // ONE fp16 activation plane against fp16 weights (the caller passed those maps)
            // 256-token tiles also from two rounds on when K is long: the tile's epilogue then hides behind the next
            // mainloop, and the kernel cuts a last partial round into 128-token halves (MIXED in gemm_pair_kernel).
            // Short K with a heavy epilogue (QKV, O) stays on 128-token tiles: there the epilogue sets the pace.
            static const bool no_mixed = getenv("FT_GEMM_NO_MIXED_TILES") != nullptr;          // A/B switch (scripts/)
            const long long rounds256 = work256 / (device_sms() / 2);
            if (rounds256 >= 8 || (!no_mixed && K >= 4096 && rounds256 >= 2))
                return launch_pair<256, false, true>(mw, mx, *mx_half, n_out, T, K, epi, pdl, stream);
            return launch_pair<128, false, true>(mw, *mx_half, *mx_lo_half, n_out, T, K, epi, pdl, stream);
        }
        if (work256 >= 8LL * (device_sms() / 2))
            return one ? launch_pair<256, false>(mw, mx, *mx_lo, n_out, T, K, epi, pdl, stream)
                       : launch_pair<256, true>(mw, mx, *mx_lo, n_out, T, K, epi, pdl, stream);
        return one ? launch_pair<128, false>(mw, *mx_half, *mx_lo_half, n_out, T, K, epi, pdl, stream)
                   : launch_pair<128, true>(mw, *mx_half, *mx_lo_half, n_out, T, K, epi, pdl, stream);
    }
    if (epi.x_planes == 3) {          // one fp16 plane x fp16 weights on the tile-per-CTA kernel (grids of at most one wave)
        FT_REQUIRE(bn == 128 && splits == 1 && !(multiwave && K <= 4096),
                   "gemm: no fp16 single-plane form of this launch (check gemm_f16_path_ok)");
        return launch_gemm_bn<128, false, true>(mw, mx, mx, n_out, T, K, 1, epi, pdl, stream);
    }
    if (multiwave && K <= 4096) {
        switch (bn) {
            case 16: return launch_persistent_bn<16>(mw, mx, *mx_lo, n_out, T, K, epi, pdl, stream);
            case 32: return launch_persistent_bn<32>(mw, mx, *mx_lo, n_out, T, K, epi, pdl, stream);
            case 64: return launch_persistent_bn<64>(mw, mx, *mx_lo, n_out, T, K, epi, pdl, stream);
            case 128: return launch_persistent_bn<128>(mw, mx, *mx_lo, n_out, T, K, epi, pdl, stream);
        }
    }
#define FT_GEMM_CASE(BN_)                                                                              \
    case BN_:                                                                                          \
        return hilo ? launch_gemm_bn<BN_, true>(mw, mx, *mx_lo, n_out, T, K, splits, epi, pdl, stream)  \
                    : launch_gemm_bn<BN_, false>(mw, mx, mx, n_out, T, K, splits, epi, pdl, stream);
    switch (bn) {
        FT_GEMM_CASE(16)
        FT_GEMM_CASE(32)
        FT_GEMM_CASE(64)
        FT_GEMM_CASE(128)
        case 256:
            return launch_gemm_bn<256, false>(mw, mx, mx, n_out, T, K, splits, epi, pdl, stream);
    }
#undef FT_GEMM_CASE
    FT_REQUIRE(false, "gemm: unsupported BN %d", bn);
}

}  // namespace ft

// ------------------------------------------------------------------ C ABI (kernel level)
extern "C" int ft_gemm_fused(const void* w, const void* x, const void* x_lo, int n_out, int T, int K,
                             int n_splits, const ft_gemm_epilogue* epi, void* stream) {
    using namespace ft;
    FT_REQUIRE(w && x && epi, "ft_gemm_fused: null argument");
    FT_REQUIRE(n_out % GEMM_BM == 0 && K % GEMM_BK == 0 && T >= 1,
               "ft_gemm_fused: need n_out %% 128 == 0, K %% 64 == 0, T >= 1 (got %d, %d, %d)", n_out, K, T);
    const int sms = device_sms();
    if (epi->x_planes == 3) {
        // w and x are FP16, one activation plane (what the engine's large prefill steps run): 128-token tiles, K unsplit
        FT_REQUIRE(!x_lo && gemm_f16_path_ok(n_out, T, K, 1) && epi->mode != FT_EPI_PARTIAL,
                   "ft_gemm_fused: x_planes = 3 needs x_lo = NULL, T > 64 and a fused epilogue (n_out %d, T %d, K %d)", n_out, T, K);
        CUtensorMap mw16, mx16, mx16h;
        if (int r = make_tmap_bf16_k64(&mw16, w, n_out, K, GEMM_BM)) return r;      // 2-byte elements: the same boxes
        if (int r = make_tmap_bf16_k64(&mx16, x, T, K, 128)) return r;
        if (int r = make_tmap_bf16_k64(&mx16h, x, T, K, 64)) return r;
        return launch_gemm(mw16, mx16, &mx16, 128, n_out, T, K, 1, *epi, false, static_cast<cudaStream_t>(stream), &mx16h,
                           &mx16h);
    }
    const bool hilo = x_lo != nullptr;
    const bool staged = epi->mode == FT_EPI_RESID || epi->mode == FT_EPI_QKV_ROPE;
    const int bn = gemm_pick_bn(T, hilo);
    if (n_splits <= 0)
        n_splits = epi->mode == FT_EPI_PARTIAL || staged ? gemm_pick_splits(n_out, T, K, sms, hilo) : 1;
    CUtensorMap mw, mx, mxl;
    if (int r = make_tmap_bf16_k64(&mw, w, n_out, K, GEMM_BM)) return r;
    if (int r = make_tmap_bf16_k64(&mx, x, T, K, bn)) return r;
    if (hilo)
        if (int r = make_tmap_bf16_k64(&mxl, x_lo, T, K, bn)) return r;
    CUtensorMap mxh, mxlh;
    const bool halves = hilo && bn == 128;
    if (halves) {
        if (int r = make_tmap_bf16_k64(&mxh, x, T, K, 64)) return r;
        if (int r = make_tmap_bf16_k64(&mxlh, x_lo, T, K, 64)) return r;
    }
    return launch_gemm(mw, mx, hilo ? &mxl : nullptr, bn, n_out, T, K, n_splits, *epi, false,
                       static_cast<cudaStream_t>(stream), halves ? &mxh : nullptr, halves ? &mxlh : nullptr);
}

extern "C" int ft_gemm_bf16(const void* w, const void* x, const void* x_lo, void* out, void* out_lo,
                            int n_out, int T, int K, int n_splits, int out_mode, void* stream) {
    FT_REQUIRE(out_mode >= 0 && out_mode <= 2, "ft_gemm_bf16: out_mode %d (use ft_gemm_fused for the rest)",
               out_mode);
    ft_gemm_epilogue epi = {};
    epi.mode = out_mode;
    epi.out = out;
    epi.out_lo = out_lo;
    return ft_gemm_fused(w, x, x_lo, n_out, T, K, n_splits, &epi, stream);
}

extern "C" int ft_debug_gemm_trace(unsigned long long* host_out16) {
    FT_CHECK_CUDA(cudaDeviceSynchronize());
    FT_CHECK_CUDA(cudaMemcpyFromSymbol(host_out16, ft::g_gemm_trace, sizeof(unsigned long long) * 16));
    return 0;
}

extern "C" int ft_debug_gemm_trace_hist(unsigned long long* host_out, int n_rows) {
    FT_REQUIRE(host_out && n_rows >= 1 && n_rows <= ft::GEMM_TRACE_HIST, "ft_debug_gemm_trace_hist: bad arguments");
    FT_CHECK_CUDA(cudaDeviceSynchronize());
    unsigned int n = 0;
    FT_CHECK_CUDA(cudaMemcpyFromSymbol(&n, ft::g_gemm_trace_n, sizeof(n)));
    static unsigned long long all[ft::GEMM_TRACE_HIST * 16];
    FT_CHECK_CUDA(cudaMemcpyFromSymbol(all, ft::g_gemm_trace_hist, sizeof(all)));
    for (int i = 0; i < n_rows; ++i) {          // oldest first: the last n_rows launches
        const unsigned row = (n - (unsigned)n_rows + (unsigned)i) % ft::GEMM_TRACE_HIST;
        memcpy(host_out + (size_t)i * 16, all + (size_t)row * 16, 16 * sizeof(unsigned long long));
    }
    return (int)(n < (unsigned)n_rows ? n : (unsigned)n_rows);
}

extern "C" long long ft_gemm_scratch_bytes(int n_out, int T, int K, int n_splits) {
    if (n_splits <= 0) n_splits = ft::gemm_pick_splits(n_out, T, K, ft::device_sms(), true);
    return (long long)ft::gemm_scratch_floats(n_out, T, n_splits, true) * 4;
}

extern "C" int ft_gemm_pick_splits(int n_out, int T, int K, int hilo) {
    return ft::gemm_pick_splits(n_out, T, K, ft::device_sms(), hilo != 0);
}
