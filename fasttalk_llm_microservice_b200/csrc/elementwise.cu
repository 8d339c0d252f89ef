// K1/K2/K4 + SwiGLU: the small HBM-light fused kernels between the projection GEMMs.  Each one
// also folds in the split-K reduction of the fp32 partial sums its producer GEMM wrote, so the
// partials never make a separate pass.
#include "common.cuh"
#include "ft_internal.h"

#include <stdarg.h>
#include <stdio.h>

namespace ft {

// ------------------------------------------------------------------ error string (host)
static thread_local char g_err[512] = "";
void set_last_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}
const char* last_error() { return g_err; }

// ------------------------------------------------------------------ K1 + K2
// row i of the output <- row r = row_idx ? row_idx[i] : i of the residual stream.
//   embed_tokens != null : resid[r] = embed[token[r]]                       (layer-0 entry)
//   else                 : resid[r] += sum_s partials[s][r]                 (n_splits may be 0)
//   y = resid[r] * rsqrt(mean(resid[r]^2) + eps) * w;  xn[i] = bf16(y);  xn_lo[i] = bf16(y - xn[i])
constexpr int NORM_THREADS = 256;
constexpr int NORM_MAXV = 8;      // float4 per thread -> h <= 8192

__global__ void __launch_bounds__(NORM_THREADS)
add_rmsnorm_kernel(float* __restrict__ resid, const float* __restrict__ partials, int n_splits,
                   size_t split_stride, const int* __restrict__ embed_tokens,
                   const __nv_bfloat16* __restrict__ embed, const __nv_bfloat16* __restrict__ w,
                   __nv_bfloat16* __restrict__ xn, __nv_bfloat16* __restrict__ xn_lo,
                   const int* __restrict__ row_idx, int h, float eps) {
    pdl_wait();
    pdl_launch_dependents();
    const int i = blockIdx.x;
    const int r = row_idx ? row_idx[i] : i;
    const int nv = h >> 2;
    float4 a[NORM_MAXV];
    float ss = 0.f;
    float* rrow = resid + (size_t)r * h;
#pragma unroll
    for (int v = 0; v < NORM_MAXV; ++v) {
        const int c = threadIdx.x + v * NORM_THREADS;
        if (c < nv) {
            float4 x;
            if (embed_tokens) {
                const uint2 e = *reinterpret_cast<const uint2*>(embed + (size_t)embed_tokens[r] * h + c * 4);
                x = make_float4(bf16_lo(e.x), bf16_hi(e.x), bf16_lo(e.y), bf16_hi(e.y));
            } else {
                x = reinterpret_cast<const float4*>(rrow)[c];
                const float4* pp = reinterpret_cast<const float4*>(partials + (size_t)r * h) + c;
                const size_t ps = split_stride >> 2;
                int s = 0;
                for (; s + 4 <= n_splits; s += 4) {        // four independent loads in flight
                    const float4 p0 = __ldcg(pp + (s + 0) * ps), p1 = __ldcg(pp + (s + 1) * ps);
                    const float4 p2 = __ldcg(pp + (s + 2) * ps), p3 = __ldcg(pp + (s + 3) * ps);
                    x.x += (p0.x + p1.x) + (p2.x + p3.x); x.y += (p0.y + p1.y) + (p2.y + p3.y);
                    x.z += (p0.z + p1.z) + (p2.z + p3.z); x.w += (p0.w + p1.w) + (p2.w + p3.w);
                }
                for (; s < n_splits; ++s) {
                    const float4 p = __ldcg(pp + s * ps);
                    x.x += p.x; x.y += p.y; x.z += p.z; x.w += p.w;
                }
            }
            if (embed_tokens || n_splits > 0) reinterpret_cast<float4*>(rrow)[c] = x;
            a[v] = x;
            ss += x.x * x.x + x.y * x.y + x.z * x.z + x.w * x.w;
        }
    }
    __shared__ float red[NORM_THREADS / 32];
    ss = warp_sum(ss);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = ss;
    __syncthreads();
    float tot = 0.f;
#pragma unroll
    for (int k = 0; k < NORM_THREADS / 32; ++k) tot += red[k];
    const float rstd = rsqrtf(tot / (float)h + eps);
#pragma unroll
    for (int v = 0; v < NORM_MAXV; ++v) {
        const int c = threadIdx.x + v * NORM_THREADS;
        if (c < nv) {
            const uint2 wv = *reinterpret_cast<const uint2*>(w + c * 4);
            const float y0 = a[v].x * rstd * bf16_lo(wv.x), y1 = a[v].y * rstd * bf16_hi(wv.x);
            const float y2 = a[v].z * rstd * bf16_lo(wv.y), y3 = a[v].w * rstd * bf16_hi(wv.y);
            uint2 o;
            o.x = pack_bf16(y0, y1);
            o.y = pack_bf16(y2, y3);
            *reinterpret_cast<uint2*>(xn + (size_t)i * h + c * 4) = o;
            if (xn_lo) {
                uint2 l;
                l.x = pack_bf16(y0 - bf16_lo(o.x), y1 - bf16_hi(o.x));
                l.y = pack_bf16(y2 - bf16_lo(o.y), y3 - bf16_hi(o.y));
                *reinterpret_cast<uint2*>(xn_lo + (size_t)i * h + c * 4) = l;
            }
        }
    }
}

// ------------------------------------------------------------------ norm-deferred pipeline entry / row gather
// resid[t] = embed[token[t]];  planes = hi/lo(resid * gamma);  ssq[0][t] = sum resid^2, ssq[p>0][t] = 0
__global__ void __launch_bounds__(NORM_THREADS)
embed_planes_kernel(const int* __restrict__ tokens, const __nv_bfloat16* __restrict__ embed,
                    const __nv_bfloat16* __restrict__ gamma, float* __restrict__ resid,
                    __nv_bfloat16* __restrict__ out, __nv_bfloat16* __restrict__ out_lo,
                    float* __restrict__ ssq, int ssq_parts, int ssq_stride, int h) {
    pdl_wait();
    pdl_launch_dependents();
    const int t = blockIdx.x;
    const int nv = h >> 2;
    const __nv_bfloat16* erow = embed + (size_t)tokens[t] * h;
    float ss = 0.f;
    for (int c = threadIdx.x; c < nv; c += NORM_THREADS) {
        const uint2 e = *reinterpret_cast<const uint2*>(erow + c * 4);
        const uint2 g = *reinterpret_cast<const uint2*>(gamma + c * 4);
        const float4 x = make_float4(bf16_lo(e.x), bf16_hi(e.x), bf16_lo(e.y), bf16_hi(e.y));
        reinterpret_cast<float4*>(resid + (size_t)t * h)[c] = x;
        ss += x.x * x.x + x.y * x.y + x.z * x.z + x.w * x.w;
        const float y0 = x.x * bf16_lo(g.x), y1 = x.y * bf16_hi(g.x), y2 = x.z * bf16_lo(g.y), y3 = x.w * bf16_hi(g.y);
        uint2 o, l;
        o.x = pack_bf16(y0, y1);
        o.y = pack_bf16(y2, y3);
        l.x = pack_bf16(y0 - bf16_lo(o.x), y1 - bf16_hi(o.x));
        l.y = pack_bf16(y2 - bf16_lo(o.y), y3 - bf16_hi(o.y));
        *reinterpret_cast<uint2*>(out + (size_t)t * h + c * 4) = o;
        *reinterpret_cast<uint2*>(out_lo + (size_t)t * h + c * 4) = l;
    }
    __shared__ float red[NORM_THREADS / 32];
    ss = warp_sum(ss);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = ss;
    __syncthreads();
    if (threadIdx.x < ssq_parts) {
        float tot = 0.f;
        if (threadIdx.x == 0)
            for (int k = 0; k < NORM_THREADS / 32; ++k) tot += red[k];
        ssq[(size_t)threadIdx.x * ssq_stride + t] = tot;
    }
}

// out[i] = x[row_idx[i]] for both planes, and the rows' sums of squares.
__global__ void __launch_bounds__(NORM_THREADS)
gather_planes_kernel(const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ x_lo,
                     const float* __restrict__ ssq, int ssq_parts, int ssq_stride,
                     const int* __restrict__ row_idx, int h, __nv_bfloat16* __restrict__ out,
                     __nv_bfloat16* __restrict__ out_lo, float* __restrict__ ssq_out, int ssq_out_stride) {
    pdl_wait();
    pdl_launch_dependents();
    const int i = blockIdx.x;
    const int r = row_idx[i];
    const int nv = h >> 3;              // 16-byte chunks
    for (int c = threadIdx.x; c < nv; c += NORM_THREADS) {
        reinterpret_cast<uint4*>(out + (size_t)i * h)[c] = reinterpret_cast<const uint4*>(x + (size_t)r * h)[c];
        reinterpret_cast<uint4*>(out_lo + (size_t)i * h)[c] = reinterpret_cast<const uint4*>(x_lo + (size_t)r * h)[c];
    }
    if (threadIdx.x < ssq_parts)
        ssq_out[(size_t)threadIdx.x * ssq_out_stride + i] = __ldcg(ssq + (size_t)threadIdx.x * ssq_stride + r);
}

// ------------------------------------------------------------------ K4
// One warp per (token, head) of the fused QKV projection output:
//   q heads : reduce partials, RoPE (rotate_half pairs i, i+d/2), write fp16 q[t][head][d]
//   k heads : reduce, RoPE, write into the paged cache (swizzled 16-byte chunks)
//   v heads : reduce, write into the paged cache
// KV cache layer layout: [page][kv_head][K|V][64 tokens][d] fp16.
template <int D>
__global__ void __launch_bounds__(128)
rope_kv_append_kernel(const float* __restrict__ qkv, int n_splits, size_t split_stride,
                      const float* __restrict__ cos_t, const float* __restrict__ sin_t,
                      const int* __restrict__ positions, const int* __restrict__ seq_slots,
                      const int* __restrict__ page_table, int max_pages,
                      __half* __restrict__ q_out, __half* __restrict__ kv_layer, int T,
                      int nq, int nkv) {
    pdl_wait();
    pdl_launch_dependents();
    constexpr int HALF = D / 2;
    constexpr int PAIRS = HALF / 32;          // rotation pairs per lane
    const int heads = nq + 2 * nkv;
    const int item = blockIdx.x * 4 + (threadIdx.x >> 5);
    if (item >= T * heads) return;
    const int t = item / heads;
    const int hd = item % heads;
    const int lane = threadIdx.x & 31;
    const int pos = positions[t];
    const size_t row_off = (size_t)t * heads * D + (size_t)hd * D;

    float lo[PAIRS], hi[PAIRS];
#pragma unroll
    for (int p = 0; p < PAIRS; ++p) {
        const int i = lane + p * 32;
        float a = 0.f, b = 0.f;
#pragma unroll 4
        for (int s = 0; s < n_splits; ++s) {
            a += __ldcg(qkv + s * split_stride + row_off + i);
            b += __ldcg(qkv + s * split_stride + row_off + i + HALF);
        }
        lo[p] = a;
        hi[p] = b;
    }
    const bool is_v = hd >= nq + nkv;
    if (!is_v) {
#pragma unroll
        for (int p = 0; p < PAIRS; ++p) {
            const int i = lane + p * 32;
            const float c = cos_t[(size_t)pos * HALF + i];
            const float s = sin_t[(size_t)pos * HALF + i];
            const float a = lo[p], b = hi[p];
            lo[p] = a * c - b * s;
            hi[p] = b * c + a * s;
        }
    }
    if (hd < nq) {
        __half* o = q_out + (size_t)t * nq * D + (size_t)hd * D;
        auto sat = [](float x) { return __float2half_rn(fminf(fmaxf(x, -65504.f), 65504.f)); };
#pragma unroll
        for (int p = 0; p < PAIRS; ++p) {
            o[lane + p * 32] = sat(lo[p]);
            o[lane + p * 32 + HALF] = sat(hi[p]);
        }
    } else {
        const int kvh = is_v ? hd - nq - nkv : hd - nq;
        const int page = page_table[(size_t)seq_slots[t] * max_pages + pos / kKvTile];
        const int r = pos % kKvTile;
        uint8_t* slab = reinterpret_cast<uint8_t*>(kv_layer) +
                        (((size_t)page * nkv + kvh) * 2 + (is_v ? 1 : 0)) * (size_t)(kKvTile * D * 2);
#pragma unroll
        for (int p = 0; p < PAIRS; ++p) {
            const int e0 = lane + p * 32, e1 = e0 + HALF;
            auto sat = [](float x) { return __float2half_rn(fminf(fmaxf(x, -65504.f), 65504.f)); };
            *reinterpret_cast<__half*>(slab + kv_swizzled_chunk<D>(r, e0 >> 3) + (e0 & 7) * 2) = sat(lo[p]);
            *reinterpret_cast<__half*>(slab + kv_swizzled_chunk<D>(r, e1 >> 3) + (e1 & 7) * 2) = sat(hi[p]);
        }
    }
}

// ------------------------------------------------------------------ SwiGLU
// act[t][j] = bf16( silu(sum_s gu[s][t][j]) * sum_s gu[s][t][f + j] )
__global__ void __launch_bounds__(256)
silu_mul_kernel(const float* __restrict__ gu, int n_splits, size_t split_stride,
                __nv_bfloat16* __restrict__ act, __nv_bfloat16* __restrict__ act_lo, int T, int f) {
    pdl_wait();
    pdl_launch_dependents();
    const int fv = f >> 2;
    const size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= (size_t)T * fv) return;
    const int t = (int)(idx / fv);
    const int c = (int)(idx % fv);
    float4 g = make_float4(0, 0, 0, 0), u = make_float4(0, 0, 0, 0);
    for (int s = 0; s < n_splits; ++s) {
        const float4* row = reinterpret_cast<const float4*>(gu + s * split_stride + (size_t)t * 2 * f);
        const float4 a = row[c], b = row[fv + c];
        g.x += a.x; g.y += a.y; g.z += a.z; g.w += a.w;
        u.x += b.x; u.y += b.y; u.z += b.z; u.w += b.w;
    }
    auto silu = [](float x) { return x / (1.f + __expf(-x)); };
    const float y0 = silu(g.x) * u.x, y1 = silu(g.y) * u.y, y2 = silu(g.z) * u.z, y3 = silu(g.w) * u.w;
    uint2 o;
    o.x = pack_bf16(y0, y1);
    o.y = pack_bf16(y2, y3);
    *reinterpret_cast<uint2*>(act + (size_t)t * f + c * 4) = o;
    if (act_lo) {
        uint2 l;
        l.x = pack_bf16(y0 - bf16_lo(o.x), y1 - bf16_hi(o.x));
        l.y = pack_bf16(y2 - bf16_lo(o.y), y3 - bf16_hi(o.y));
        *reinterpret_cast<uint2*>(act_lo + (size_t)t * f + c * 4) = l;
    }
}

// ------------------------------------------------------------------ launchers
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

int launch_add_rmsnorm(float* resid, const float* partials, int n_splits, size_t split_stride,
                       const int* embed_tokens, const void* embed_table, const void* norm_w,
                       void* xn_out, void* xn_lo, const int* row_idx, int T, int h, float eps, bool pdl,
                       cudaStream_t stream) {
    FT_REQUIRE(h % 4 == 0 && h <= 4 * NORM_THREADS * NORM_MAXV, "add_rmsnorm: unsupported h=%d", h);
    FT_REQUIRE(T >= 1, "add_rmsnorm: T=%d", T);
    cudaLaunchConfig_t cfg;
    cudaLaunchAttribute attr[1];
    fill_cfg(cfg, attr, dim3(T), dim3(NORM_THREADS), 0, pdl, stream);
    FT_CHECK_CUDA(cudaLaunchKernelEx(&cfg, add_rmsnorm_kernel, resid, partials, n_splits, split_stride,
                                     embed_tokens, (const __nv_bfloat16*)embed_table,
                                     (const __nv_bfloat16*)norm_w, (__nv_bfloat16*)xn_out,
                                     (__nv_bfloat16*)xn_lo, row_idx, h, eps));
    return 0;
}

int launch_rope_kv_append(const float* qkv, int n_splits, size_t split_stride, const float* cos_t,
                          const float* sin_t, const int* positions, const int* seq_slots,
                          const int* page_table, int max_pages, void* q_out, void* kv_layer, int T,
                          int nq, int nkv, int d, bool pdl, cudaStream_t stream) {
    FT_REQUIRE(d == 64 || d == 128, "rope_kv_append: head_dim %d not supported (64 or 128)", d);
    const int items = T * (nq + 2 * nkv);
    cudaLaunchConfig_t cfg;
    cudaLaunchAttribute attr[1];
    fill_cfg(cfg, attr, dim3((items + 3) / 4), dim3(128), 0, pdl, stream);
    if (d == 64) {
        FT_CHECK_CUDA(cudaLaunchKernelEx(&cfg, rope_kv_append_kernel<64>, qkv, n_splits, split_stride,
                                         cos_t, sin_t, positions, seq_slots, page_table, max_pages,
                                         (__half*)q_out, (__half*)kv_layer, T, nq, nkv));
    } else {
        FT_CHECK_CUDA(cudaLaunchKernelEx(&cfg, rope_kv_append_kernel<128>, qkv, n_splits, split_stride,
                                         cos_t, sin_t, positions, seq_slots, page_table, max_pages,
                                         (__half*)q_out, (__half*)kv_layer, T, nq, nkv));
    }
    return 0;
}

int launch_embed_planes(const int* tokens, const void* embed, const void* gamma, float* resid, void* out,
                        void* out_lo, float* ssq, int ssq_parts, int ssq_stride, int T, int h, bool pdl,
                        cudaStream_t stream) {
    FT_REQUIRE(h % 4 == 0 && ssq_parts >= 1 && ssq_parts <= NORM_THREADS && T >= 1,
               "embed_planes: bad shape h=%d parts=%d T=%d", h, ssq_parts, T);
    cudaLaunchConfig_t cfg;
    cudaLaunchAttribute attr[1];
    fill_cfg(cfg, attr, dim3(T), dim3(NORM_THREADS), 0, pdl, stream);
    FT_CHECK_CUDA(cudaLaunchKernelEx(&cfg, embed_planes_kernel, tokens, (const __nv_bfloat16*)embed,
                                     (const __nv_bfloat16*)gamma, resid, (__nv_bfloat16*)out,
                                     (__nv_bfloat16*)out_lo, ssq, ssq_parts, ssq_stride, h));
    return 0;
}

int launch_gather_planes(const void* x, const void* x_lo, const float* ssq, int ssq_parts, int ssq_stride,
                         const int* row_idx, int n_rows, int h, void* out, void* out_lo, float* ssq_out,
                         int ssq_out_stride, bool pdl, cudaStream_t stream) {
    FT_REQUIRE(h % 8 == 0 && ssq_parts >= 1 && ssq_parts <= NORM_THREADS && n_rows >= 1,
               "gather_planes: bad shape h=%d parts=%d rows=%d", h, ssq_parts, n_rows);
    cudaLaunchConfig_t cfg;
    cudaLaunchAttribute attr[1];
    fill_cfg(cfg, attr, dim3(n_rows), dim3(NORM_THREADS), 0, pdl, stream);
    FT_CHECK_CUDA(cudaLaunchKernelEx(&cfg, gather_planes_kernel, (const __nv_bfloat16*)x,
                                     (const __nv_bfloat16*)x_lo, ssq, ssq_parts, ssq_stride, row_idx, h,
                                     (__nv_bfloat16*)out, (__nv_bfloat16*)out_lo, ssq_out, ssq_out_stride));
    return 0;
}

int launch_silu_mul(const float* gu, int n_splits, size_t split_stride, void* act, void* act_lo, int T,
                    int f, bool pdl, cudaStream_t stream) {
    FT_REQUIRE(f % 4 == 0, "silu_mul: f=%d must be a multiple of 4", f);
    const size_t n = (size_t)T * (f / 4);
    cudaLaunchConfig_t cfg;
    cudaLaunchAttribute attr[1];
    fill_cfg(cfg, attr, dim3((unsigned)((n + 255) / 256)), dim3(256), 0, pdl, stream);
    FT_CHECK_CUDA(cudaLaunchKernelEx(&cfg, silu_mul_kernel, gu, n_splits, split_stride,
                                     (__nv_bfloat16*)act, (__nv_bfloat16*)act_lo, T, f));
    return 0;
}

// fp16(hi + lo) of two bf16 planes (lo may be null: plain bf16 -> fp16, used once for the weights).  Large prefill
// projections multiply ONE fp16 activation plane against fp16 copies of the weights (engine, prefill_planes = 3):
// 11 mantissa bits in one tensor-core pass instead of 16 in two.  Saturating, eight elements per thread.
__global__ void __launch_bounds__(256)
planes_to_f16_kernel(const uint4* __restrict__ hi, const uint4* __restrict__ lo, uint4* __restrict__ out, size_t n8) {
    pdl_wait();
    pdl_launch_dependents();
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n8; i += (size_t)gridDim.x * blockDim.x) {
        const uint4 a = hi[i];
        uint4 b = make_uint4(0u, 0u, 0u, 0u);
        if (lo) b = lo[i];
        uint4 o;
        o.x = pack_f16_sat(bf16_lo(a.x) + bf16_lo(b.x), bf16_hi(a.x) + bf16_hi(b.x));
        o.y = pack_f16_sat(bf16_lo(a.y) + bf16_lo(b.y), bf16_hi(a.y) + bf16_hi(b.y));
        o.z = pack_f16_sat(bf16_lo(a.z) + bf16_lo(b.z), bf16_hi(a.z) + bf16_hi(b.z));
        o.w = pack_f16_sat(bf16_lo(a.w) + bf16_lo(b.w), bf16_hi(a.w) + bf16_hi(b.w));
        out[i] = o;
    }
}

int launch_planes_to_f16(const void* hi, const void* lo, void* out, size_t n_elems, bool pdl, cudaStream_t stream) {
    FT_REQUIRE(n_elems % 8 == 0, "planes_to_f16: element count must be a multiple of 8");
    const size_t n8 = n_elems / 8;
    size_t blocks = (n8 + 255) / 256;
    if (blocks > 148 * 16) blocks = 148 * 16;
    if (blocks == 0) return 0;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)blocks);
    cfg.blockDim = dim3(256);
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl ? 1 : 0;
    FT_CHECK_CUDA(cudaLaunchKernelEx(&cfg, planes_to_f16_kernel, (const uint4*)hi, (const uint4*)lo, (uint4*)out, n8));
    return 0;
}

}  // namespace ft

// ------------------------------------------------------------------ C ABI (kernel level)
extern "C" const char* ft_last_error(void) { return ft::last_error(); }

extern "C" int ft_add_rmsnorm(float* resid, const float* partials, int n_splits, long long split_stride,
                              const int* embed_tokens, const void* embed_table, const void* norm_w,
                              void* xn_out, void* xn_lo, const int* row_idx, int T, int h, float eps,
                              void* stream) {
    return ft::launch_add_rmsnorm(resid, partials, n_splits, (size_t)split_stride, embed_tokens,
                                  embed_table, norm_w, xn_out, xn_lo, row_idx, T, h, eps, false,
                                  static_cast<cudaStream_t>(stream));
}

extern "C" int ft_rope_kv_append(const float* qkv, int n_splits, long long split_stride,
                                 const float* cos_t, const float* sin_t, const int* positions,
                                 const int* seq_slots, const int* page_table, int max_pages, void* q_out,
                                 void* kv_layer, int T, int nq, int nkv, int d, void* stream) {
    return ft::launch_rope_kv_append(qkv, n_splits, (size_t)split_stride, cos_t, sin_t, positions,
                                     seq_slots, page_table, max_pages, q_out, kv_layer, T, nq, nkv, d,
                                     false, static_cast<cudaStream_t>(stream));
}

extern "C" int ft_embed_planes(const int* tokens, const void* embed_table, const void* gamma, float* resid,
                               void* out, void* out_lo, float* ssq, int ssq_parts, int ssq_stride, int T, int h,
                               void* stream) {
    return ft::launch_embed_planes(tokens, embed_table, gamma, resid, out, out_lo, ssq, ssq_parts, ssq_stride, T,
                                   h, false, static_cast<cudaStream_t>(stream));
}

extern "C" int ft_gather_planes(const void* x, const void* x_lo, const float* ssq, int ssq_parts, int ssq_stride,
                                const int* row_idx, int n_rows, int h, void* out, void* out_lo, float* ssq_out,
                                int ssq_out_stride, void* stream) {
    return ft::launch_gather_planes(x, x_lo, ssq, ssq_parts, ssq_stride, row_idx, n_rows, h, out, out_lo, ssq_out,
                                    ssq_out_stride, false, static_cast<cudaStream_t>(stream));
}

extern "C" int ft_silu_mul(const float* gu, int n_splits, long long split_stride, void* act, void* act_lo,
                           int T, int f, void* stream) {
    return ft::launch_silu_mul(gu, n_splits, (size_t)split_stride, act, act_lo, T, f, false,
                               static_cast<cudaStream_t>(stream));
}
