// Internal (C++) declarations shared by the .cu translation units.  The public C ABI is
// include/fasttalk_b200.h.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/fasttalk_b200.h"

namespace ft {

constexpr int GEMM_MAX_SPLITS = 16;

// Per-device bookkeeping.  cudaFuncSetAttribute and the SM count belong to a DEVICE; a host may drive
// several GPUs from one process (the C ABI allows it), so "done once" flags are kept per device ordinal.
constexpr int kMaxDevices = 64;
inline int current_device_slot() {
    int d = 0;
    if (cudaGetDevice(&d) != cudaSuccess || d < 0) d = 0;
    return d % kMaxDevices;
}
// Usage: `if (!flag.done()) { ...set the attributes...; flag.mark(); }` — the flag is raised only AFTER the
// set-up succeeded, so a second thread either sees it done or repeats the (idempotent) set-up itself.
struct PerDeviceFlag {
    volatile unsigned char flags[kMaxDevices] = {};
    bool done() const { return flags[current_device_slot()] != 0; }
    void mark() {
        __sync_synchronize();
        flags[current_device_slot()] = 1;
    }
};

// ---- gemm.cu
int make_tmap_bf16_k64(CUtensorMap* map, const void* base, int rows, int K, int box_rows);
int make_tmap_f16_3d(CUtensorMap* map, const void* base, int d0, int d1, int d2, int b0, int b1, int b2);
int gemm_pick_bn(int T, bool hilo);
bool gemm_pair_eligible(int n_out, int T);
bool gemm_f16_path_ok(int n_out, int T, int K, int splits);      // is there a one-fp16-plane form (x_planes = 3) of this launch?      // would launch_gemm run this (hi/lo, unsplit) projection on the CTA-pair kernel?
int gemm_pick_splits(int n_out, int T, int K, int num_sms, bool hilo);
size_t gemm_scratch_floats(int n_out, int T, int splits, bool hilo);   // split-K slabs of one launch
// Modes and fields: ft_gemm_epilogue in include/fasttalk_b200.h.
// mx_half / mx_lo_half: the same activations with 64-row boxes (only used with bn == 128, for the CTA-pair kernel)
int launch_gemm(const CUtensorMap& mw, const CUtensorMap& mx, const CUtensorMap* mx_lo, int bn, int n_out,
                int T, int K, int splits, const ft_gemm_epilogue& epi, bool pdl, cudaStream_t stream,
                const CUtensorMap* mx_half = nullptr, const CUtensorMap* mx_lo_half = nullptr);

// ---- elementwise.cu
int launch_add_rmsnorm(float* resid, const float* partials, int n_splits, size_t split_stride,
                       const int* embed_tokens, const void* embed_table, const void* norm_w,
                       void* xn_out, void* xn_lo, const int* row_idx, int T, int h, float eps, bool pdl,
                       cudaStream_t stream);
int launch_rope_kv_append(const float* qkv, int n_splits, size_t split_stride, const float* cos_t,
                          const float* sin_t, const int* positions, const int* seq_slots,
                          const int* page_table, int max_pages, void* q_out, void* kv_layer, int T,
                          int nq, int nkv, int d, bool pdl, cudaStream_t stream);
int launch_embed_planes(const int* tokens, const void* embed, const void* gamma, float* resid, void* out,
                        void* out_lo, float* ssq, int ssq_parts, int ssq_stride, int T, int h, bool pdl,
                        cudaStream_t stream);
int launch_gather_planes(const void* x, const void* x_lo, const float* ssq, int ssq_parts, int ssq_stride,
                         const int* row_idx, int n_rows, int h, void* out, void* out_lo, float* ssq_out,
                         int ssq_out_stride, bool pdl, cudaStream_t stream);
int launch_planes_to_f16(const void* hi, const void* lo, void* out, size_t n_elems, bool pdl, cudaStream_t stream);
int launch_silu_mul(const float* gu, int n_splits, size_t split_stride, void* act, void* act_lo, int T,
                    int f, bool pdl, cudaStream_t stream);

// ---- attention.cu
size_t attn_decode_scratch_floats(int B, int nkv, int group, int d, int n_ctas);
int attn_decode_ctas_per_sm(int d);      // resident CTAs of the decode kernel per SM (<= 0: error)
const char* last_error();
int launch_attn_decode_plan(const int* ctx_lens, int* tile_prefix, int B, int nkv, float* scratch, bool pdl,
                            cudaStream_t stream);
int launch_attn_decode(const void* q, const void* kv_layer, const int* page_table, int max_pages,
                       const int* ctx_lens, const int* tile_prefix, void* out, void* out_lo, float* scratch,
                       int B, int nq, int nkv, int d, int n_ctas, bool pdl, cudaStream_t stream,
                       cudaEvent_t ev0 = nullptr, cudaEvent_t ev1 = nullptr, bool early = false);
// q_rows: rows (tokens) the q / out buffers hold — the tcgen05 kernel builds a tensor map over q; <= 0 = unknown,
// which selects the mma.sync kernel
int launch_attn_prefill(const void* q, const void* kv_layer, const int* page_table, int max_pages,
                        const int* req_q_start, const int* req_slot, const int* req_pos0,
                        const int* tile_req, const int* tile_q0, int n_tiles, void* out, void* out_lo, int nq,
                        int nkv, int d, bool pdl, cudaStream_t stream, int q_rows = 0);

// ---- attention_tc.cu (tcgen05 / TMEM prefill attention: head_dim 64 / 128, GQA group 3 / 4)
bool attn_prefill_tc_supported(int nq, int nkv, int d);
int launch_attn_prefill_tc(const void* q, int q_rows, const void* kv_layer, const int* page_table, int max_pages,
                           const int* req_q_start, const int* req_slot, const int* req_pos0, const int* tile_req,
                           const int* tile_q0, int n_tiles, void* out, void* out_lo, int nq, int nkv, int d, bool pdl,
                           cudaStream_t stream);

// ---- sampling.cu
int launch_sample(const float* logits, int B, int V, const float* temperature, const int* top_k,
                  const float* top_p, const unsigned long long* seeds, int* out_tokens, bool pdl,
                  cudaStream_t stream);

}  // namespace ft
