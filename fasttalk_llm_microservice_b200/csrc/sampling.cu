// K7 — fused temperature / top-k / top-p / argmax sampling over fp32 logits [B, V].
//
// One 1024-thread CTA per sequence.  Greedy (temperature <= 0) is an argmax with
// lowest-index tie-break (torch.argmax semantics).  Otherwise e_i = exp((x_i - max)/T) and the
// kept set is found WITHOUT sorting: a 3-level radix select over the (monotone) float bit
// pattern of e_i finds first the k-th largest value (counts), then the nucleus threshold
// (masses) in 3 passes each; the draw walks the kept set in index order.  The row (513 KB for
// V = 128256) stays L2-resident between passes.
#include "common.cuh"
#include "ft_internal.h"

namespace ft {

constexpr int SAMPLE_THREADS = 1024;
constexpr int SAMPLE_WARPS = SAMPLE_THREADS / 32;

struct SampleSmem {
    float hist[2048];
    float warp_val[SAMPLE_WARPS];
    int warp_idx[SAMPLE_WARPS];
    float bc_f[4];
    int bc_i[4];
};

__device__ __forceinline__ unsigned long long splitmix64(unsigned long long x) {
    x += 0x9E3779B97F4A7C15ull;
    x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
    x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
    return x ^ (x >> 31);
}

__device__ __forceinline__ float block_max(float v, SampleSmem& sm) {
    v = warp_max(v);
    if ((threadIdx.x & 31) == 0) sm.warp_val[threadIdx.x >> 5] = v;
    __syncthreads();
    float r = sm.warp_val[0];
#pragma unroll
    for (int w = 1; w < SAMPLE_WARPS; ++w) r = fmaxf(r, sm.warp_val[w]);
    __syncthreads();
    return r;
}

__device__ __forceinline__ float weight_of(float x, float maxv, float inv_t) {
    return __expf((x - maxv) * inv_t);
}

// Largest key K such that weight{key >= K, key >= min_key} >= target (weight = count or mass).
// Returns K; *w_ge = weight{key >= K}.  All threads must call; result is uniform.
__device__ uint32_t radix_threshold(const float* __restrict__ x, int V, float maxv, float inv_t,
                                    uint32_t min_key, float target, bool by_mass, SampleSmem& sm,
                                    float* w_ge) {
    uint32_t prefix = 0, mask = 0;
    float above = 0.f, at = 0.f;
    const int shifts[3] = {20, 9, 0};
    const int bits[3] = {11, 11, 9};
    for (int lvl = 0; lvl < 3; ++lvl) {
        const int shift = shifts[lvl], nb = 1 << bits[lvl];
        for (int i = threadIdx.x; i < 2048; i += SAMPLE_THREADS) sm.hist[i] = 0.f;
        __syncthreads();
        for (int i = threadIdx.x; i < V; i += SAMPLE_THREADS) {
            const float e = weight_of(x[i], maxv, inv_t);
            const uint32_t key = __float_as_uint(e);
            if (key >= min_key && (key & mask) == prefix)
                atomicAdd(&sm.hist[(key >> shift) & (nb - 1)], by_mass ? e : 1.f);
        }
        __syncthreads();
        // warp w sums bins [64w, 64w+64)
        {
            const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
            float s = sm.hist[w * 64 + l] + sm.hist[w * 64 + 32 + l];
            s = warp_sum(s);
            if (l == 0) sm.warp_val[w] = s;
        }
        __syncthreads();
        if (threadIdx.x == 0) {
            float acc = above;
            int bin = 0;
            float acc_above = above, at_bin = 0.f;
            bool found = false;
            for (int w = SAMPLE_WARPS - 1; w >= 0 && !found; --w) {
                if (acc + sm.warp_val[w] >= target) {
                    for (int b = w * 64 + 63; b >= w * 64; --b) {
                        if (acc + sm.hist[b] >= target) {
                            bin = b; acc_above = acc; at_bin = sm.hist[b]; found = true;
                            break;
                        }
                        acc += sm.hist[b];
                    }
                    if (!found) {      // rounding: the warp total said yes, the walk said no
                        bin = w * 64; acc_above = acc - sm.hist[bin]; at_bin = sm.hist[bin]; found = true;
                    }
                } else {
                    acc += sm.warp_val[w];
                }
            }
            if (!found) {              // target above the total weight: keep everything
                bin = 0; acc_above = acc - sm.hist[0]; at_bin = sm.hist[0];
            }
            sm.bc_i[0] = bin;
            sm.bc_f[0] = acc_above;
            sm.bc_f[1] = at_bin;
        }
        __syncthreads();
        prefix |= (uint32_t)sm.bc_i[0] << shift;
        mask |= (uint32_t)(nb - 1) << shift;
        above = sm.bc_f[0];
        at = sm.bc_f[1];
        __syncthreads();
    }
    *w_ge = above + at;
    return prefix < min_key ? min_key : prefix;
}

__global__ void __launch_bounds__(SAMPLE_THREADS)
sample_kernel(const float* __restrict__ logits, int V, const float* __restrict__ temperature,
              const int* __restrict__ top_k, const float* __restrict__ top_p,
              const unsigned long long* __restrict__ seeds, int* __restrict__ out_tokens) {
    __shared__ SampleSmem sm;
    pdl_wait();
    pdl_launch_dependents();
    const int b = blockIdx.x;
    const float* x = logits + (size_t)b * V;
    const float temp = temperature ? temperature[b] : 0.f;

    // ---- max (+ argmax, lowest index on ties)
    float best = -INFINITY;
    int best_i = 0x7fffffff;
    if ((V & 3) == 0 && (reinterpret_cast<uintptr_t>(x) & 15) == 0) {
        // 16-byte loads, four in flight per thread: the row is 513 KB behind ONE CTA, so it is the
        // number of outstanding loads, not arithmetic, that sets this pass's time
        const float4* x4 = reinterpret_cast<const float4*>(x);
        const int n4 = V >> 2;
        auto take = [&](const float4& v, int i4) {      // ascending index order inside the thread
            if (v.x > best) { best = v.x; best_i = 4 * i4; }
            if (v.y > best) { best = v.y; best_i = 4 * i4 + 1; }
            if (v.z > best) { best = v.z; best_i = 4 * i4 + 2; }
            if (v.w > best) { best = v.w; best_i = 4 * i4 + 3; }
        };
        int i = threadIdx.x;
        for (; i + 3 * SAMPLE_THREADS < n4; i += 4 * SAMPLE_THREADS) {
            const float4 a = __ldcg(x4 + i), b4 = __ldcg(x4 + i + SAMPLE_THREADS),
                         c4 = __ldcg(x4 + i + 2 * SAMPLE_THREADS), d4 = __ldcg(x4 + i + 3 * SAMPLE_THREADS);
            take(a, i); take(b4, i + SAMPLE_THREADS); take(c4, i + 2 * SAMPLE_THREADS); take(d4, i + 3 * SAMPLE_THREADS);
        }
        for (; i < n4; i += SAMPLE_THREADS) take(__ldcg(x4 + i), i);
    } else {
        for (int i = threadIdx.x; i < V; i += SAMPLE_THREADS) {
            const float v = x[i];
            if (v > best) { best = v; best_i = i; }      // ascending i per thread: first max wins
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const float ov = __shfl_xor_sync(0xffffffffu, best, o);
        const int oi = __shfl_xor_sync(0xffffffffu, best_i, o);
        if (ov > best || (ov == best && oi < best_i)) { best = ov; best_i = oi; }
    }
    if ((threadIdx.x & 31) == 0) { sm.warp_val[threadIdx.x >> 5] = best; sm.warp_idx[threadIdx.x >> 5] = best_i; }
    __syncthreads();
    if (threadIdx.x == 0) {
        float bv = sm.warp_val[0];
        int bi = sm.warp_idx[0];
        for (int w = 1; w < SAMPLE_WARPS; ++w)
            if (sm.warp_val[w] > bv || (sm.warp_val[w] == bv && sm.warp_idx[w] < bi)) {
                bv = sm.warp_val[w]; bi = sm.warp_idx[w];
            }
        sm.bc_f[2] = bv;
        sm.bc_i[2] = bi;
    }
    __syncthreads();
    const float maxv = sm.bc_f[2];
    const int argmax = sm.bc_i[2];
    __syncthreads();
    if (temp <= 1e-6f) {
        if (threadIdx.x == 0) out_tokens[b] = argmax;
        return;
    }
    const float inv_t = 1.f / temp;
    const int k = top_k ? top_k[b] : 0;
    const float p = top_p ? top_p[b] : 1.f;

    uint32_t thr = 0;           // keep keys >= thr
    float kept = 0.f;           // mass of the kept set
    bool have_mass = false;
    if (k > 0 && k < V) {
        float cnt;
        thr = radix_threshold(x, V, maxv, inv_t, 0u, (float)k, false, sm, &cnt);
    }
    if (p < 1.f) {
        // mass of the top-k survivors (or of everything)
        float z = 0.f;
        for (int i = threadIdx.x; i < V; i += SAMPLE_THREADS) {
            const float e = weight_of(x[i], maxv, inv_t);
            if (__float_as_uint(e) >= thr) z += e;
        }
        z = warp_sum(z);
        if ((threadIdx.x & 31) == 0) sm.warp_val[threadIdx.x >> 5] = z;
        __syncthreads();
        float Z = 0.f;
        for (int w = 0; w < SAMPLE_WARPS; ++w) Z += sm.warp_val[w];
        __syncthreads();
        thr = radix_threshold(x, V, maxv, inv_t, thr, p * Z, true, sm, &kept);
        have_mass = false;      // recompute exactly below (histogram sums are order-dependent)
    }
    (void)have_mass;

    // ---- draw: warp w owns the contiguous index range [w*span, (w+1)*span)
    const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
    const int span = (V + SAMPLE_WARPS - 1) / SAMPLE_WARPS;
    const int lo = w * span, hi = lo + span < V ? lo + span : V;
    float mine = 0.f;
    for (int i = lo + l; i < hi; i += 32) {
        const float e = weight_of(x[i], maxv, inv_t);
        if (__float_as_uint(e) >= thr) mine += e;
    }
    mine = warp_sum(mine);
    if (l == 0) sm.warp_val[w] = mine;
    __syncthreads();
    if (threadIdx.x == 0) {
        float tot = 0.f;
        for (int i = 0; i < SAMPLE_WARPS; ++i) tot += sm.warp_val[i];
        const unsigned long long r = splitmix64(seeds ? seeds[b] : 0x1234ull + b);
        const float u = (float)(r >> 40) * (1.0f / 16777216.0f) * tot;      // [0, tot)
        float acc = 0.f;
        int ww = SAMPLE_WARPS - 1;
        for (int i = 0; i < SAMPLE_WARPS; ++i) {
            if (acc + sm.warp_val[i] > u) { ww = i; break; }
            acc += sm.warp_val[i];
        }
        // if rounding pushed u past the total, fall back to the last warp that holds mass
        while (ww > 0 && sm.warp_val[ww] == 0.f) --ww;
        sm.bc_i[0] = ww;
        sm.bc_f[0] = u - acc;
        sm.bc_i[1] = argmax;       // final fallback
    }
    __syncthreads();
    if (w == sm.bc_i[0]) {
        float need = sm.bc_f[0];
        int chosen = -1, last_kept = -1;
        float run = 0.f;
        for (int i0 = lo; i0 < hi && chosen < 0; i0 += 32) {
            const int i = i0 + l;
            float e = 0.f;
            if (i < hi) {
                e = weight_of(x[i], maxv, inv_t);
                if (__float_as_uint(e) < thr) e = 0.f;
            }
            float inc = e;          // inclusive warp scan
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const float n = __shfl_up_sync(0xffffffffu, inc, o);
                if (l >= o) inc += n;
            }
            const bool hit = e > 0.f && run + inc > need;
            const unsigned ballot = __ballot_sync(0xffffffffu, hit);
            const unsigned kept_mask = __ballot_sync(0xffffffffu, e > 0.f);
            if (kept_mask) last_kept = i0 + 31 - __clz(kept_mask);
            if (ballot) chosen = i0 + __ffs(ballot) - 1;
            run += __shfl_sync(0xffffffffu, inc, 31);
        }
        if (l == 0) out_tokens[b] = chosen >= 0 ? chosen : (last_kept >= 0 ? last_kept : sm.bc_i[1]);
    }
}

int launch_sample(const float* logits, int B, int V, const float* temperature, const int* top_k,
                  const float* top_p, const unsigned long long* seeds, int* out_tokens, bool pdl,
                  cudaStream_t stream) {
    FT_REQUIRE(B >= 1 && V >= 1, "sample: B=%d V=%d", B, V);
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(B);
    cfg.blockDim = dim3(SAMPLE_THREADS);
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl ? 1 : 0;
    FT_CHECK_CUDA(cudaLaunchKernelEx(&cfg, sample_kernel, logits, V, temperature, top_k, top_p, seeds,
                                     out_tokens));
    return 0;
}

}  // namespace ft

extern "C" int ft_sample(const float* logits, int B, int V, const float* temperature, const int* top_k,
                         const float* top_p, const unsigned long long* seeds, int* out_tokens,
                         void* stream) {
    return ft::launch_sample(logits, B, V, temperature, top_k, top_p, seeds, out_tokens, false,
                             static_cast<cudaStream_t>(stream));
}
