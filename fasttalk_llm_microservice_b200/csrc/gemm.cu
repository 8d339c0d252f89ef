// K3 — bf16 projection GEMM on tcgen05 / TMEM, TMA-fed.   Y[t, n] = sum_k X[t, k] * W[n, k]
//
// The weight matrix is the UMMA "A" operand (128 output features fill the 128 TMEM lanes),
// the activations are the "B" operand (tokens = UMMA N, 16..256), so the decode case
// (T <= 64 tokens) is pure weight streaming: every CTA pulls its 128 x K_range slab of W
// through a ring of TMA stages exactly once.  Split-K over gridDim.z gives enough CTAs to
// keep HBM busy when n_out/128 is small.
//
// Activations come as TWO bf16 planes, hi = bf16(x) and lo = bf16(x - hi): both are multiplied
// by the same weight tile into the same TMEM accumulator (D += W*hi; D += W*lo).  That keeps
// ~16 mantissa bits of every activation through a pure bf16 x bf16 -> fp32 tensor-core product;
// the weight tile is already in shared memory, so in the HBM-bound decode regime the second
// product is free (DESIGN.md "precision").  lo may be absent (HILO = false).
//
// Epilogues (ft_gemm_epilogue.mode, include/fasttalk_b200.h):
//   PARTIAL / BF16 / SWIGLU / LOGITS  straight from TMEM to global memory;
//   RESID / QKV_ROPE                  the split-K CTAs of one row tile form a thread-block
//     CLUSTER: each parks its fp32 partial tile in its own shared memory, and after a cluster
//     barrier every CTA reduces a slice of the tokens across all peers through distributed
//     shared memory (ld.shared::cluster) and finishes the layer step there — residual add +
//     next-norm statistics, or RoPE + paged KV append — so no stand-alone reduce / norm / RoPE
//     kernel sits between the projections (RMSNorm's 1/rms is applied after the next GEMM).
//
// Warp roles (192 threads): warp 0 = TMA producer, warp 1 = TMEM owner + MMA issuer,
// warps 2..5 = epilogue (tcgen05.ld -> global).
#include "common.cuh"
#include "ft_internal.h"

namespace ft {

constexpr int GEMM_BM = 128;      // W rows per CTA  (UMMA M)
constexpr int GEMM_BK = 64;       // bf16 per 128-byte swizzle row
constexpr int GEMM_A_BYTES = GEMM_BM * GEMM_BK * 2;
constexpr int GEMM_THREADS = 192;
constexpr int GEMM_MAX_CLUSTER = 8;

__device__ __forceinline__ void epi_bar_sync() { asm volatile("bar.sync 2, 128;" ::: "memory"); }

__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire;" ::: "memory");
}
__device__ __forceinline__ float ld_dsmem_f32(uint32_t local_addr, uint32_t rank) {
    uint32_t remote;
    float v;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(remote) : "r"(local_addr), "r"(rank));
    asm volatile("ld.shared::cluster.f32 %0, [%1];" : "=f"(v) : "r"(remote) : "memory");
    return v;
}

struct HiLo {
    __nv_bfloat16 hi, lo;
};
__device__ __forceinline__ HiLo split_hi_lo(float x) {
    HiLo r;
    r.hi = __float2bfloat16(x);
    r.lo = __float2bfloat16(x - __bfloat162float(r.hi));
    return r;
}

template <int BN, bool HILO>
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
    float* red_s = rstd_s + BN;                                  // [4][BN] cross-warp sums

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const int row0 = blockIdx.x * GEMM_BM;
    const int t0 = blockIdx.y * BN;
    const int split = blockIdx.z;
    const int n_split = gridDim.z;
    const int kc_begin = (int)(((long long)k_chunks * split) / n_split);
    const int kc_end = (int)(((long long)k_chunks * (split + 1)) / n_split);
    const int n_iter = kc_end - kc_begin;
    const int mode = ep.mode;
    const bool staged = mode == FT_EPI_RESID || mode == FT_EPI_QKV_ROPE;

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

    pdl_launch_dependents();

    if (warp == 0) {
        if (elect_one()) {
            const uint64_t pol_w = policy_evict_first();   // weights: streamed once per step
            const uint64_t pol_x = policy_evict_last();    // activations: re-read by every row tile
            // Weights do not depend on the previous kernel: start streaming them before the
            // programmatic-dependency wait, then add the activation tiles.
            const int pre = n_iter < stages ? n_iter : stages;
            for (int i = 0; i < pre; ++i) {
                mbar_expect_tx(&full_bar[i], STAGE_BYTES);
                tma_tile_2d(smem + (size_t)i * STAGE_BYTES, &map_w, (kc_begin + i) * GEMM_BK, row0,
                            &full_bar[i], pol_w);
            }
            pdl_wait();
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
            constexpr uint32_t idesc = umma_idesc_bf16(GEMM_BM, BN);
            int s = 0;
            uint32_t ph = 0;
            for (int i = 0; i < n_iter; ++i) {
                mbar_wait(&full_bar[s], ph);
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
            umma_commit(tmem_full_bar);
        }
    } else {
        // ---------------- epilogue warps.  A warp may only touch the TMEM lane quarter (warp_id % 4).
        const int q = warp & 3;
        const int tid = q * 32 + lane;          // tile row owned by this thread
        const int row = row0 + tid;
        pdl_wait();                             // ssq_in / resid / positions come from earlier kernels
        // per-token 1/rms of the activation's source (RMSNorm deferred past the projection)
        for (int j = tid; j < BN; j += 128) {
            float r = 1.f;
            const int t = t0 + j;
            if (ep.ssq_in && t < T) {
                float s = 0.f;
                for (int p = 0; p < ep.ssq_parts; ++p) s += __ldcg(ep.ssq_in + (size_t)p * ep.ssq_stride + t);
                r = rsqrtf(s * ep.inv_h + ep.eps);
            }
            rstd_s[j] = r;
        }
        epi_bar_sync();
        if (n_iter > 0) {
            mbar_wait(tmem_full_bar, 0);
            tc_fence_after();
        }
        const size_t split_off = (size_t)split * T;
        float* stage_f = reinterpret_cast<float*>(smem);   // pipeline stages are idle once the MMAs retired
#pragma unroll 1
        for (int c0 = 0; c0 < BN; c0 += 16) {
            if (t0 + c0 >= T) break;
            float v[16];
            if (n_iter > 0) {
                tmem_ld16(tmem_base + ((uint32_t)(q * 32) << 16) + c0, v);
            } else {
#pragma unroll
                for (int j = 0; j < 16; ++j) v[j] = 0.f;
            }
            if (staged) {
                // park the fp32 partial tile as [token][row] for the cluster-wide reduction
#pragma unroll
                for (int j = 0; j < 16; ++j) stage_f[(c0 + j) * GEMM_BM + tid] = v[j];
            } else if (mode == FT_EPI_PARTIAL || mode == FT_EPI_LOGITS) {
                float* o = reinterpret_cast<float*>(ep.out);
#pragma unroll
                for (int j = 0; j < 16; ++j) {
                    const int t = t0 + c0 + j;
                    if (t < T) o[(split_off + t) * n_out + row] = v[j] * rstd_s[c0 + j];
                }
            } else if (mode == FT_EPI_BF16) {
                __nv_bfloat16* o = reinterpret_cast<__nv_bfloat16*>(ep.out);
                __nv_bfloat16* ol = reinterpret_cast<__nv_bfloat16*>(ep.out_lo);
#pragma unroll
                for (int j = 0; j < 16; ++j) {
                    const int t = t0 + c0 + j;
                    if (t < T) {
                        const HiLo hl = split_hi_lo(v[j] * rstd_s[c0 + j]);
                        o[(size_t)t * n_out + row] = hl.hi;
                        if (ol) ol[(size_t)t * n_out + row] = hl.lo;
                    }
                }
            } else {
                // fused SwiGLU: lanes 0..63 hold gate features, lanes 64..127 the matching up
                // features; up warps hand their values to the gate warps through smem
                const int f_half = n_out >> 1;
                if (q >= 2) {
#pragma unroll
                    for (int j = 0; j < 16; ++j) stage_f[((q - 2) * 32 + lane) * 17 + j] = v[j];
                }
                epi_bar_sync();
                if (q < 2) {
                    __nv_bfloat16* o = reinterpret_cast<__nv_bfloat16*>(ep.out);
                    __nv_bfloat16* ol = reinterpret_cast<__nv_bfloat16*>(ep.out_lo);
                    const int feat = blockIdx.x * 64 + q * 32 + lane;
#pragma unroll
                    for (int j = 0; j < 16; ++j) {
                        const int t = t0 + c0 + j;
                        if (t < T) {
                            const float rs = rstd_s[c0 + j];
                            const float g = v[j] * rs, u = stage_f[(q * 32 + lane) * 17 + j] * rs;
                            const HiLo hl = split_hi_lo(g / (1.f + __expf(-g)) * u);
                            o[(size_t)t * f_half + feat] = hl.hi;
                            if (ol) ol[(size_t)t * f_half + feat] = hl.lo;
                        }
                    }
                }
                epi_bar_sync();
            }
        }
    }

    if (staged) {
        // ---------------- cluster-wide split-K reduction through distributed shared memory
        __syncwarp();
        if (n_split > 1) cluster_sync_all(); else __syncthreads();
        if (warp >= 2) {
            const int q = warp & 3;
            const int tid = q * 32 + lane;
            const int row = row0 + tid;
            float* stage_f = reinterpret_cast<float*>(smem);
            const uint32_t stage_addr = smem_u32(stage_f);
            // this CTA finishes tokens [j0, j1) of the tile
            const int j0 = BN * split / n_split, j1 = BN * (split + 1) / n_split;
            if (mode == FT_EPI_RESID) {
                float* resid = ep.resid;
                const float gam = __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(ep.gamma)[row]);
                __nv_bfloat16* o = reinterpret_cast<__nv_bfloat16*>(ep.out);
                __nv_bfloat16* ol = reinterpret_cast<__nv_bfloat16*>(ep.out_lo);
                for (int j = j0; j < j1; ++j) {
                    const int t = t0 + j;
                    float ss = 0.f;
                    if (t < T) {
                        float y = 0.f;
                        if (n_split == 1) {
                            y = stage_f[j * GEMM_BM + tid];
                        } else {
                            for (int p = 0; p < n_split; ++p)
                                y += ld_dsmem_f32(stage_addr + (uint32_t)(j * GEMM_BM + tid) * 4, (uint32_t)p);
                        }
                        const size_t idx = (size_t)t * n_out + row;
                        const float r = resid[idx] + y;
                        resid[idx] = r;
                        const HiLo hl = split_hi_lo(r * gam);
                        o[idx] = hl.hi;
                        ol[idx] = hl.lo;
                        ss = r * r;
                    }
                    ss = warp_sum(ss);
                    if (lane == 0) red_s[q * BN + j] = ss;
                }
                epi_bar_sync();
                for (int j = j0 + tid; j < j1; j += 128) {
                    const int t = t0 + j;
                    if (t < T)
                        ep.ssq_out[(size_t)blockIdx.x * ep.ssq_stride + t] =
                            (red_s[j] + red_s[BN + j]) + (red_s[2 * BN + j] + red_s[3 * BN + j]);
                }
            } else {
                // QKV_ROPE.  pass A: reduce, apply 1/rms, leave the value in my own staging slot
                // (slots of my token slice are read by nobody else)
                for (int j = j0; j < j1; ++j) {
                    if (t0 + j >= T) break;
                    float y = 0.f;
                    if (n_split == 1) {
                        y = stage_f[j * GEMM_BM + tid];
                    } else {
                        for (int p = 0; p < n_split; ++p)
                            y += ld_dsmem_f32(stage_addr + (uint32_t)(j * GEMM_BM + tid) * 4, (uint32_t)p);
                    }
                    stage_f[j * GEMM_BM + tid] = y * rstd_s[j];
                }
                epi_bar_sync();
                // pass B: rotate with the partner row (i +- d/2 lives in the same 128-row tile)
                const int d = ep.head_dim, half = d >> 1;
                const int q_dim = ep.nq * d, kv_dim = ep.nkv * d;
                const int i = row % d, fi = i % half;
                const int region = row < q_dim ? 0 : row < q_dim + kv_dim ? 1 : 2;
                for (int j = j0; j < j1; ++j) {
                    const int t = t0 + j;
                    if (t >= T) break;
                    const float a = stage_f[j * GEMM_BM + tid];
                    float val = a;
                    const int pos = ep.positions[t];
                    if (region < 2) {
                        const float b = stage_f[j * GEMM_BM + (tid ^ half)];
                        const float c = ep.rope_cos[(size_t)pos * half + fi], s = ep.rope_sin[(size_t)pos * half + fi];
                        val = i < half ? a * c - b * s : a * c + b * s;
                    }
                    if (region == 0) {
                        reinterpret_cast<__nv_bfloat16*>(ep.q_out)[(size_t)t * q_dim + row] = __float2bfloat16(val);
                    } else {
                        const int kvh = (row - q_dim - (region == 2 ? kv_dim : 0)) / d;
                        const int page = ep.page_table[(size_t)ep.seq_rows[t] * ep.max_pages + pos / kKvTile];
                        const int r = pos % kKvTile;
                        uint8_t* slab = reinterpret_cast<uint8_t*>(ep.kv_layer) +
                                        (((size_t)page * ep.nkv + kvh) * 2 + (region == 2 ? 1 : 0)) *
                                            (size_t)(kKvTile * d * 2);
                        const int off = r * (d * 2) + ((((i >> 3) ^ (r & 7))) << 4) + (i & 7) * 2;
                        *reinterpret_cast<__nv_bfloat16*>(slab + off) = __float2bfloat16(val);
                    }
                }
            }
        }
        __syncwarp();
        if (n_split > 1) cluster_sync_all();     // peers may still be reading my staging tile
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, TMEM_COLS);
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

static int device_sms() {
    static int sms = 0;
    if (!sms) {
        int dev = 0;
        if (cudaGetDevice(&dev) != cudaSuccess ||
            cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || sms <= 0)
            sms = 148;
    }
    return sms;
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
    // multi-wave grids: two CTAs per SM so one CTA's epilogue overlaps the other's streaming
    const int budget = tiles > device_sms() ? 108 * 1024 : 208 * 1024;
    int stages = budget / STAGE_BYTES;
    if (stages > 12) stages = 12;
    if (stages > per_cta) stages = per_cta;
    if (stages < 2) stages = 2;
    // the staged epilogues park a [BN][128] fp32 tile in the (idle) pipeline memory
    while ((size_t)stages * STAGE_BYTES < (size_t)BN * GEMM_BM * 4) ++stages;
    *stages_out = stages;
    *smem_out = (size_t)stages * STAGE_BYTES + 1024 + (2 * stages + 1) * 8 + 16 + 5 * BN * 4 + 16;
    FT_REQUIRE(*smem_out <= 227 * 1024, "gemm: %zu bytes of shared memory needed", *smem_out);
    return 0;
}

template <int BN, bool HILO>
static int launch_gemm_bn(const CUtensorMap& mw, const CUtensorMap& mx, const CUtensorMap& mxl, int n_out,
                          int T, int K, int splits, const ft_gemm_epilogue& epi, bool pdl, cudaStream_t stream) {
    int stages = 0;
    size_t smem = 0;
    if (int r = stage_plan<BN, HILO>(n_out, T, K, splits, &stages, &smem)) return r;
    static bool attr_set = false;
    if (!attr_set) {
        FT_CHECK_CUDA(cudaFuncSetAttribute(gemm_bf16_tcgen05_kernel<BN, HILO>,
                                           cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
        attr_set = true;
    }
    const bool staged = epi.mode == FT_EPI_RESID || epi.mode == FT_EPI_QKV_ROPE;
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
    if (staged && splits > 1) {
        attr[na].id = cudaLaunchAttributeClusterDimension;
        attr[na].val.clusterDim.x = 1;
        attr[na].val.clusterDim.y = 1;
        attr[na].val.clusterDim.z = splits;
        ++na;
    }
    cfg.attrs = attr;
    cfg.numAttrs = na;
    FT_CHECK_CUDA(cudaLaunchKernelEx(&cfg, gemm_bf16_tcgen05_kernel<BN, HILO>, mw, mx, mxl, epi, n_out, T,
                                     K / GEMM_BK, stages));
    return 0;
}

// How many clusters of `splits` CTAs of the hi/lo kernel can be resident at once.
template <int BN>
static int max_active_clusters(int splits, size_t smem) {
    cudaFuncSetAttribute(gemm_bf16_tcgen05_kernel<BN, true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                         227 * 1024);
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(1, 1, splits);
    cfg.blockDim = dim3(GEMM_THREADS);
    cfg.dynamicSmemBytes = smem;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 1;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = splits;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    int n = 0;
    if (cudaOccupancyMaxActiveClusters(&n, gemm_bf16_tcgen05_kernel<BN, true>, &cfg) != cudaSuccess) {
        cudaGetLastError();
        return 0;
    }
    return n;
}

int gemm_pick_cluster_splits(int n_out, int T, int K, int num_sms) {
    const int bn = gemm_pick_bn(T, true);
    const int tiles = (n_out / GEMM_BM) * ((T + bn - 1) / bn);
    const int kc = K / GEMM_BK;
    if (tiles >= num_sms) return 1;
    int s = num_sms / tiles;
    if (s > GEMM_MAX_CLUSTER) s = GEMM_MAX_CLUSTER;
    if (s > kc / 2) s = kc / 2 > 0 ? kc / 2 : 1;
    static int cache[4][GEMM_MAX_CLUSTER + 1];      // [bn index][splits] -> resident clusters (0 = unknown)
    const int bi = bn == 16 ? 0 : bn == 32 ? 1 : bn == 64 ? 2 : 3;
    for (; s > 1; --s) {
        if (!cache[bi][s]) {
            int stages = 0;
            size_t smem = 0;
            int n = 0;
            switch (bn) {
                case 16: if (!stage_plan<16, true>(n_out, T, K, s, &stages, &smem)) n = max_active_clusters<16>(s, smem); break;
                case 32: if (!stage_plan<32, true>(n_out, T, K, s, &stages, &smem)) n = max_active_clusters<32>(s, smem); break;
                case 64: if (!stage_plan<64, true>(n_out, T, K, s, &stages, &smem)) n = max_active_clusters<64>(s, smem); break;
                default: if (!stage_plan<128, true>(n_out, T, K, s, &stages, &smem)) n = max_active_clusters<128>(s, smem); break;
            }
            cache[bi][s] = n > 0 ? n : -1;
        }
        if (cache[bi][s] >= tiles) return s;       // every cluster of the grid is resident in one wave
    }
    return 1;
}

int launch_gemm(const CUtensorMap& mw, const CUtensorMap& mx, const CUtensorMap* mx_lo, int bn, int n_out,
                int T, int K, int splits, const ft_gemm_epilogue& epi, bool pdl, cudaStream_t stream) {
    FT_REQUIRE(n_out % GEMM_BM == 0, "gemm: n_out=%d must be a multiple of %d", n_out, GEMM_BM);
    FT_REQUIRE(K % GEMM_BK == 0, "gemm: K=%d must be a multiple of %d", K, GEMM_BK);
    FT_REQUIRE(T >= 1, "gemm: T=%d", T);
    FT_REQUIRE(splits >= 1 && splits <= GEMM_MAX_SPLITS && splits <= K / GEMM_BK,
               "gemm: bad split count %d", splits);
    const int mode = epi.mode;
    FT_REQUIRE(mode >= FT_EPI_PARTIAL && mode <= FT_EPI_LOGITS, "gemm: bad epilogue mode %d", mode);
    const bool staged = mode == FT_EPI_RESID || mode == FT_EPI_QKV_ROPE;
    FT_REQUIRE(mode == FT_EPI_PARTIAL || staged || splits == 1, "gemm: this epilogue needs splits == 1");
    FT_REQUIRE(!staged || splits <= GEMM_MAX_CLUSTER, "gemm: cluster-reduced epilogues split at most %d ways",
               GEMM_MAX_CLUSTER);
    const bool hilo = mx_lo != nullptr;
    FT_REQUIRE(!staged || hilo, "gemm: the fused RESID / QKV_ROPE epilogues need both activation planes");
    FT_REQUIRE(!(hilo && bn > 128), "gemm: hi/lo activations support token tiles up to 128");
    if (mode == FT_EPI_QKV_ROPE) {
        FT_REQUIRE(epi.head_dim == 64 || epi.head_dim == 128, "gemm: head_dim %d not supported", epi.head_dim);
        FT_REQUIRE((epi.nq * epi.head_dim) % GEMM_BM == 0 && (epi.nkv * epi.head_dim) % GEMM_BM == 0 &&
                       n_out == (epi.nq + 2 * epi.nkv) * epi.head_dim,
                   "gemm: q / kv widths must be multiples of 128 and add up to n_out");
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
    const bool hilo = x_lo != nullptr;
    const bool staged = epi->mode == FT_EPI_RESID || epi->mode == FT_EPI_QKV_ROPE;
    const int bn = gemm_pick_bn(T, hilo);
    if (n_splits <= 0)
        n_splits = epi->mode == FT_EPI_PARTIAL ? gemm_pick_splits(n_out, T, K, sms, hilo)
                   : staged                    ? gemm_pick_cluster_splits(n_out, T, K, sms)
                                               : 1;
    CUtensorMap mw, mx, mxl;
    if (int r = make_tmap_bf16_k64(&mw, w, n_out, K, GEMM_BM)) return r;
    if (int r = make_tmap_bf16_k64(&mx, x, T, K, bn)) return r;
    if (hilo)
        if (int r = make_tmap_bf16_k64(&mxl, x_lo, T, K, bn)) return r;
    return launch_gemm(mw, mx, hilo ? &mxl : nullptr, bn, n_out, T, K, n_splits, *epi, false,
                       static_cast<cudaStream_t>(stream));
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

extern "C" int ft_gemm_pick_splits(int n_out, int T, int K, int hilo) {
    return ft::gemm_pick_splits(n_out, T, K, ft::device_sms(), hilo != 0);
}
