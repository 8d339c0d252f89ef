// K3 — bf16 projection GEMM on tcgen05 / TMEM, TMA-fed.   Y[t, n] = sum_k X[t, k] * W[n, k]
//
// The weight matrix is the UMMA "A" operand (128 output features fill the 128 TMEM lanes),
// the activations are the "B" operand (tokens = UMMA N, 16..256), so the decode case
// (T <= 64 tokens) is pure weight streaming: every CTA pulls its 128 x K_range slab of W
// through a ring of TMA stages exactly once.  Split-K over gridDim.z gives enough CTAs to
// keep HBM busy when n_out/128 is small; the fp32 partial sums are reduced by whichever
// fused kernel consumes them (add_rmsnorm / rope_kv_append).
//
// Activations come as TWO bf16 planes, hi = bf16(x) and lo = bf16(x - hi): both are multiplied
// by the same weight tile into the same TMEM accumulator (D += W*hi; D += W*lo).  That keeps
// ~16 mantissa bits of every activation through a pure bf16 x bf16 -> fp32 tensor-core product;
// the weight tile is already in shared memory, so in the HBM-bound decode regime the second
// product is free (DESIGN.md "precision").  lo may be absent (HILO = false).
//
// Epilogues: fp32 partials | bf16 (hi[/lo]) | fused SwiGLU for the gate/up projection
// (weight rows interleaved in blocks of 64: gate rows land in TMEM lanes 0..63, the matching
// up rows in lanes 64..127, so silu(g)*u is formed in the epilogue and the 2f-wide
// intermediate never touches memory).
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

__device__ __forceinline__ void epi_bar_sync() {
    asm volatile("bar.sync 2, 128;" ::: "memory");
}

template <int BN, bool HILO>
__global__ void __launch_bounds__(GEMM_THREADS, 1)
gemm_bf16_tcgen05_kernel(const __grid_constant__ CUtensorMap map_w,
                         const __grid_constant__ CUtensorMap map_x,
                         const __grid_constant__ CUtensorMap map_x_lo, void* __restrict__ out,
                         void* __restrict__ out_lo, int n_out, int T, int k_chunks, int stages,
                         int out_mode) {
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

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const int row0 = blockIdx.x * GEMM_BM;
    const int t0 = blockIdx.y * BN;
    const int split = blockIdx.z;
    const int kc_begin = (int)(((long long)k_chunks * split) / gridDim.z);
    const int kc_end = (int)(((long long)k_chunks * (split + 1)) / gridDim.z);
    const int n_iter = kc_end - kc_begin;

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
        // Epilogue: a warp may only touch the TMEM lane quarter (warp_id % 4).
        const int q = warp & 3;
        const int row = row0 + q * 32 + lane;
        if (n_iter > 0) {
            mbar_wait(tmem_full_bar, 0);
            tc_fence_after();
        }
        const size_t split_off = (size_t)split * T;
        float* exch = reinterpret_cast<float*>(smem);      // stage 0 is idle once the MMAs retired
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
            if (out_mode == 0) {
                float* o = reinterpret_cast<float*>(out);
#pragma unroll
                for (int j = 0; j < 16; ++j) {
                    const int t = t0 + c0 + j;
                    if (t < T) o[(split_off + t) * n_out + row] = v[j];
                }
            } else if (out_mode == 1) {
                __nv_bfloat16* o = reinterpret_cast<__nv_bfloat16*>(out);
                __nv_bfloat16* ol = reinterpret_cast<__nv_bfloat16*>(out_lo);
#pragma unroll
                for (int j = 0; j < 16; ++j) {
                    const int t = t0 + c0 + j;
                    if (t < T) {
                        const __nv_bfloat16 h = __float2bfloat16(v[j]);
                        o[(size_t)t * n_out + row] = h;
                        if (ol) ol[(size_t)t * n_out + row] = __float2bfloat16(v[j] - __bfloat162float(h));
                    }
                }
            } else {
                // fused SwiGLU: lanes 0..63 hold gate features, lanes 64..127 the matching up
                // features; up warps hand their values to the gate warps through smem
                const int f_half = n_out >> 1;
                if (q >= 2) {
#pragma unroll
                    for (int j = 0; j < 16; ++j) exch[((q - 2) * 32 + lane) * 17 + j] = v[j];
                }
                epi_bar_sync();
                if (q < 2) {
                    __nv_bfloat16* o = reinterpret_cast<__nv_bfloat16*>(out);
                    __nv_bfloat16* ol = reinterpret_cast<__nv_bfloat16*>(out_lo);
                    const int feat = blockIdx.x * 64 + q * 32 + lane;
#pragma unroll
                    for (int j = 0; j < 16; ++j) {
                        const int t = t0 + c0 + j;
                        if (t < T) {
                            const float g = v[j], u = exch[(q * 32 + lane) * 17 + j];
                            const float a = g / (1.f + __expf(-g)) * u;
                            const __nv_bfloat16 h = __float2bfloat16(a);
                            o[(size_t)t * f_half + feat] = h;
                            if (ol) ol[(size_t)t * f_half + feat] = __float2bfloat16(a - __bfloat162float(h));
                        }
                    }
                }
                epi_bar_sync();
            }
        }
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
static int launch_gemm_bn(const CUtensorMap& mw, const CUtensorMap& mx, const CUtensorMap& mxl, void* out,
                          void* out_lo, int n_out, int T, int K, int splits, int out_mode, bool pdl,
                          cudaStream_t stream) {
    constexpr int STAGE_BYTES = GEMM_A_BYTES + (HILO ? 2 : 1) * BN * GEMM_BK * 2;
    const int kc = K / GEMM_BK;
    const int per_cta = (kc + splits - 1) / splits;
    const int tiles = (n_out / GEMM_BM) * ((T + BN - 1) / BN) * splits;
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    // multi-wave grids: two CTAs per SM so one CTA's epilogue overlaps the other's streaming
    const int budget = tiles > sms ? 108 * 1024 : 208 * 1024;
    int stages = budget / STAGE_BYTES;
    if (stages > 12) stages = 12;
    if (stages > per_cta) stages = per_cta;
    if (stages < 2) stages = 2;
    const size_t smem = (size_t)stages * STAGE_BYTES + 1024 + (2 * stages + 1) * 8 + 16;
    FT_REQUIRE(smem <= 227 * 1024, "gemm: %zu bytes of shared memory needed", smem);
    static bool attr_set = false;
    if (!attr_set) {
        FT_CHECK_CUDA(cudaFuncSetAttribute(gemm_bf16_tcgen05_kernel<BN, HILO>,
                                           cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
        attr_set = true;
    }
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(n_out / GEMM_BM, (T + BN - 1) / BN, splits);
    cfg.blockDim = dim3(GEMM_THREADS);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl ? 1 : 0;
    FT_CHECK_CUDA(cudaLaunchKernelEx(&cfg, gemm_bf16_tcgen05_kernel<BN, HILO>, mw, mx, mxl, out, out_lo, n_out,
                                     T, kc, stages, out_mode));
    return 0;
}

int launch_gemm(const CUtensorMap& mw, const CUtensorMap& mx, const CUtensorMap* mx_lo, int bn, void* out,
                void* out_lo, int n_out, int T, int K, int splits, int out_mode, bool pdl,
                cudaStream_t stream) {
    FT_REQUIRE(n_out % GEMM_BM == 0, "gemm: n_out=%d must be a multiple of %d", n_out, GEMM_BM);
    FT_REQUIRE(K % GEMM_BK == 0, "gemm: K=%d must be a multiple of %d", K, GEMM_BK);
    FT_REQUIRE(T >= 1, "gemm: T=%d", T);
    FT_REQUIRE(splits >= 1 && splits <= GEMM_MAX_SPLITS && splits <= K / GEMM_BK,
               "gemm: bad split count %d", splits);
    FT_REQUIRE(out_mode == 0 || splits == 1, "gemm: bf16 / SwiGLU output needs splits == 1");
    FT_REQUIRE(out_mode >= 0 && out_mode <= 2, "gemm: bad out_mode %d", out_mode);
    const bool hilo = mx_lo != nullptr;
    FT_REQUIRE(!(hilo && bn > 128), "gemm: hi/lo activations support token tiles up to 128");
#define FT_GEMM_CASE(BN_)                                                                                  \
    case BN_:                                                                                              \
        return hilo ? launch_gemm_bn<BN_, true>(mw, mx, *mx_lo, out, out_lo, n_out, T, K, splits, out_mode, \
                                                pdl, stream)                                               \
                    : launch_gemm_bn<BN_, false>(mw, mx, mx, out, out_lo, n_out, T, K, splits, out_mode,    \
                                                 pdl, stream);
    switch (bn) {
        FT_GEMM_CASE(16)
        FT_GEMM_CASE(32)
        FT_GEMM_CASE(64)
        FT_GEMM_CASE(128)
        case 256:
            return launch_gemm_bn<256, false>(mw, mx, mx, out, out_lo, n_out, T, K, splits, out_mode, pdl, stream);
    }
#undef FT_GEMM_CASE
    FT_REQUIRE(false, "gemm: unsupported BN %d", bn);
}

}  // namespace ft

// ------------------------------------------------------------------ C ABI (kernel level)
extern "C" int ft_gemm_bf16(const void* w, const void* x, const void* x_lo, void* out, void* out_lo,
                            int n_out, int T, int K, int n_splits, int out_mode, void* stream) {
    using namespace ft;
    int dev = 0, sms = 148;
    FT_CHECK_CUDA(cudaGetDevice(&dev));
    FT_CHECK_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    FT_REQUIRE(n_out % GEMM_BM == 0 && K % GEMM_BK == 0 && T >= 1,
               "ft_gemm_bf16: need n_out %% 128 == 0, K %% 64 == 0, T >= 1 (got %d, %d, %d)", n_out,
               K, T);
    const bool hilo = x_lo != nullptr;
    const int bn = gemm_pick_bn(T, hilo);
    if (n_splits <= 0) n_splits = out_mode == 0 ? gemm_pick_splits(n_out, T, K, sms, hilo) : 1;
    CUtensorMap mw, mx, mxl;
    if (int r = make_tmap_bf16_k64(&mw, w, n_out, K, GEMM_BM)) return r;
    if (int r = make_tmap_bf16_k64(&mx, x, T, K, bn)) return r;
    if (hilo)
        if (int r = make_tmap_bf16_k64(&mxl, x_lo, T, K, bn)) return r;
    return launch_gemm(mw, mx, hilo ? &mxl : nullptr, bn, out, out_lo, n_out, T, K, n_splits, out_mode, false,
                       static_cast<cudaStream_t>(stream));
}

extern "C" int ft_gemm_pick_splits(int n_out, int T, int K, int hilo) {
    int dev = 0, sms = 148;
    if (cudaGetDevice(&dev) == cudaSuccess)
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    return ft::gemm_pick_splits(n_out, T, K, sms, hilo != 0);
}
