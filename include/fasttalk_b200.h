/* fasttalk_b200 — C ABI of the B200-native generation backend.
 *
 * The reference (Berkay2002/fasttalk-llm-microservice) has no FFI of its own: its hot path is
 * an HTTP hop to Ollama (`app/core/ollama_handler.py:142-177`, POST /api/chat) or vLLM
 * (`app/core/vllm_handler.py:177,271`, POST /v1/chat/completions).  This library replaces what
 * sits behind that hop.  The Python class that binds it
 * (`fasttalk_llm_microservice_b200/handler.py: NativeHandler`) duck-types
 * `VLLMHandler` (`app/core/vllm_handler.py:24-338`) and `OllamaHandler`
 * (`app/core/ollama_handler.py:24-458`); INTEGRATION.md shows the binding.
 *
 * Conventions: every function returns 0 (or a non-negative value where documented) on
 * success and a negative value on error, with a message in ft_last_error() (thread local).
 * No exceptions cross the ABI.  All `const void*` tensor arguments are DEVICE pointers to
 * bf16 unless stated otherwise; `stream` is a cudaStream_t (0 = default stream).
 * The library never falls back to the CPU.
 */
#ifndef FASTTALK_B200_H
#define FASTTALK_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

const char* ft_last_error(void);

/* ------------------------------------------------------------------------------------------
 * Engine level — what replaces the engine behind `VLLMHandler.generate_stream_async`
 * (vllm_handler.py:216-308) and `OllamaHandler.generate_stream` (ollama_handler.py:110-231):
 * tokens in, sampled tokens out, continuous batching over concurrent sessions.
 * ---------------------------------------------------------------------------------------- */
typedef struct ft_engine ft_engine;

typedef struct {
    int hidden, ffn, layers, n_q_heads, n_kv_heads, head_dim, vocab;
    float rms_eps;
    int max_batch;           /* sequences per step; vLLM --max-num-seqs (docker-compose.vllm.yml:47) */
    int max_batched_tokens;  /* tokens per step; vLLM --max-num-batched-tokens (docker-compose.vllm.yml:48) */
    int num_pages;           /* KV pages of 64 tokens */
    int max_context;         /* context window in tokens; vLLM --max-model-len (docker-compose.vllm.yml:43) */
    int max_sessions;        /* sessions whose KV may stay cached between turns */
    int n_eos;
    int eos_ids[4];          /* stop tokens (<|eot_id|> 128009, <|end_of_text|> 128001) */
    int use_pdl;             /* programmatic dependent launch between the step's kernels */
    int debug_logit_rows;    /* rows of the logits buffer (>= max_batch); more only for ft_debug_* */
    /* Tensor parallelism (Llama-3.1-8B, TP=2 over NVLink).  tp_size 0/1 = off, 2 = this
     * engine is one of two ranks, one process per GPU.  n_q_heads, n_kv_heads and ffn are then this
     * rank's SHARE (half), vocab stays the full vocabulary, and the weights are the rank's slices:
     * wqkv rows (q | k | v of the local heads), wo columns of the local heads, wgu rows of the local
     * half of ffn (then interleaved), wd columns of that half, lm_head rows [rank * vocab/2, ...).
     * Both ranks must be fed the same requests in the same order and stepped together; call
     * ft_engine_tp_export / ft_engine_tp_attach after creation, before the first step. */
    int tp_size, tp_rank;
    /* Prompt tokens a step may admit WHILE other sequences are decoding (they all wait for the step): the
     * latency side of vLLM's --max-num-batched-tokens trade-off (docker-compose.vllm.yml:48).  0 = default
     * (896), < 0 = no cap (every step may fill max_batched_tokens).  The admitted count is rounded up so that decode
     * rows + prompt tokens end on a 128-token tile of the projections (those rows cost nothing more).  The allowance
     * grows with the number of sequences waiting for their first token per decoding sequence (cap x waiting /
     * (2 x decoding) once more than two wait per decoding one): one early bird ahead of a burst of 63 arrivals does not
     * hold the burst back, 35 arrivals beside 29 decoding sequences get the cap.  With nothing decoding a step uses the
     * full max_batched_tokens. */
    int decode_prefill_cap;
    /* How the large prefill projections (more tiles than SMs, 128-token tiles) multiply their activations:
     * 0 = default, 1 = bf16 hi plane only, 2 = bf16 hi + lo planes (two MMAs per tile), 3 = ONE fp16 plane
     * (fp16(hi + lo), staged by a small kernel) against fp16 copies of the weights the engine makes at creation
     * (+ 2 bytes per projection weight).  Decode-sized steps always use hi + lo. */
    int prefill_planes;
} ft_engine_config;

typedef struct {
    /* bf16, row-major [out, in]; per-layer tensors stacked on a leading L dimension */
    const void* embed;       /* [vocab, hidden] */
    const void* lm_head;     /* [vocab, hidden] (== embed when tied) */
    const void* final_norm;  /* [hidden] */
    const void* attn_norm;   /* [L, hidden] */
    const void* wqkv;        /* [L, (nq + 2 nkv) * d, hidden]   rows: q | k | v */
    const void* wo;          /* [L, hidden, nq * d] */
    const void* mlp_norm;    /* [L, hidden] */
    const void* wgu;         /* [L, 2 * ffn, hidden]  rows interleaved in blocks of 64: gate rows
                                64i..64i+63, then up rows 64i..64i+63 (for the fused SwiGLU epilogue) */
    const void* wd;          /* [L, hidden, ffn] */
    const float* rope_cos;   /* fp32 [max_context, d/2] */
    const float* rope_sin;   /* fp32 [max_context, d/2] */
} ft_weights;

typedef struct {
    long long steps, decode_steps, prefill_tokens, decode_tokens, kernel_launches;
    long long prefix_hit_tokens;      /* prompt tokens served from a session's cached KV */
    long long h2d_bytes, d2h_bytes;   /* per-step metadata uploaded / sampled tokens read back */
    long long decode_ctx_tokens;      /* sum over decode steps of the rows' context lengths */
    long long attn_decode_launches;   /* decode-attention launches timed while profiling is on */
    long long attn_decode_kv_tokens;  /* context tokens those launches streamed (per launch: one layer) */
    double attn_decode_ms;            /* their summed CUDA-event time */
    int active_requests, waiting_requests, live_sessions, free_pages, total_pages;
    int last_batch, last_prefill_tokens;
    double last_step_ms;              /* host wall time of the last ft_engine_step */
} ft_engine_stats;

/* Tensor-parallel wiring.  ft_engine_tp_export returns this engine's three peer-visible device
 * buffers (exchange area, arrival words, logits) so the host can hand them to the other rank —
 * ft_ipc_export / ft_ipc_open turn a device pointer into a 64-byte handle another process can map
 * (cudaIpc*; the peer GPU must be reachable over NVLink / P2P).  ft_engine_tp_attach takes the
 * peer's three buffers as mapped in THIS process. */
int ft_engine_tp_export(ft_engine* e, void** xchg, void** flags, void** logits);
int ft_engine_tp_attach(ft_engine* e, void* peer_xchg, void* peer_flags, void* peer_logits);
int ft_ipc_export(void* device_ptr, unsigned char handle_out[64]);
int ft_ipc_open(const unsigned char handle[64], void** device_ptr_out);
int ft_ipc_close(void* device_ptr);

/* Weights stay owned by the caller and must outlive the engine. */
ft_engine* ft_engine_create(const ft_engine_config* cfg, const ft_weights* w);
void ft_engine_destroy(ft_engine* e);

/* Queue one generation.  `tokens` is a HOST array: the whole templated conversation so far
 * (the reference re-sends the full history every turn, conversation_manager.py:54-56).
 * `session_key` identifies the conversation (the reference's request_id == session id,
 * websocket_server_vllm.py:528-534); the longest common prefix with the KV cached for that key
 * is reused and only the remainder is prefilled.  temperature <= 0 means greedy.
 * top_k <= 0 and top_p >= 1 disable those filters.  Returns a request id >= 0. */
long long ft_engine_submit(ft_engine* e, unsigned long long session_key, const int* tokens, int n_tokens,
                           float temperature, int top_k, float top_p, int max_new_tokens,
                           unsigned long long seed, int ignore_eos);

/* One continuous-batching iteration (admit -> chunked prefill + decode -> sample).  Blocks until
 * the step's tokens are on the host.  Returns the number of tokens produced (0 = idle). */
int ft_engine_step(ft_engine* e);

/* The engine's own stepping thread, instead of calling ft_engine_step from the host language: ft_engine_start spawns a
 * native thread that runs one step after the other while anything is runnable (and sleeps on a condition variable
 * otherwise), so that nothing — no interpreter, no lock of the host language — sits between two steps; the owner submits,
 * cancels and closes as before (all thread-safe), blocks in ft_engine_wait until events are queued and drains them.
 * ft_engine_step is refused while the thread runs.  ft_engine_stop joins it (ft_engine_destroy does so too); start / stop
 * may alternate.  Not for tensor-parallel engines (their hosts step both ranks in lock step).
 * ft_engine_wait: blocks up to timeout_ms; returns the number of queued events (0 = timeout), or < 0 when the stepping
 * thread died on an error (ft_last_error has it) and nothing is left to drain, or when the engine is being destroyed
 * (ft_engine_destroy sends a thread that is still blocked here away before it frees anything).
 * No counterpart in the reference: its engines (Ollama, vLLM) are separate server processes. */
int ft_engine_start(ft_engine* e);
int ft_engine_stop(ft_engine* e);
int ft_engine_wait(ft_engine* e, int timeout_ms);

/* Drain the tokens produced since the last drain, over all requests: parallel HOST arrays of
 * (request id, token, finish flag: 0 running, 1 stop token, 2 length, 3 cancelled, 4 error).
 * Returns the number of entries written (<= cap). */
int ft_engine_drain(ft_engine* e, long long* req_ids, int* tokens, int* finish, int cap);

/* The reference's cancel_generation (vllm_handler.py:310-326).  Frees the batch slot; the
 * session's KV stays cached. */
int ft_engine_cancel(ft_engine* e, long long req);
/* Forget a session and free its KV pages (session end / disconnect). */
int ft_engine_session_close(ft_engine* e, unsigned long long session_key);
int ft_engine_stats_get(ft_engine* e, ft_engine_stats* out);
/* Change ft_engine_config.decode_prefill_cap of a live engine (same encoding: 0 default, < 0 uncapped);
 * takes effect at the next step. */
int ft_engine_set_decode_prefill_cap(ft_engine* e, int cap);
/* Measurement hooks (bench.py): CUDA events on the engine's own stream.  ft_engine_timer(e, 0, 0)
 * records a start event; ft_engine_timer(e, 1, &ms) records a stop event, waits for it and returns
 * the device time between the two.  With profiling on, every decode-attention launch is bracketed
 * by events and accumulated into ft_engine_stats (this breaks PDL overlap around that kernel). */
int ft_engine_timer(ft_engine* e, int stop, float* ms_out);
/* One row of four floats per step for the most recent steps, oldest first, at most `cap_rows` (the engine
 * keeps 4096) -> HOST float [rows][4]: device ms (CUDA events on the engine's stream: metadata upload ..
 * sampled-token copy), decode rows, prompt tokens, host wall ms of the step.  Returns the rows written and
 * forgets them.  ft_engine_timer's total minus the sum of the device times is what the GPU waited for the
 * host between steps. */
int ft_engine_step_log(ft_engine* e, float* out, int cap_rows);
int ft_engine_set_profile(ft_engine* e, int on);
/* fp32 logits [vocab] of `req` from the most recent step it took part in -> HOST buffer. */
int ft_engine_get_logits(ft_engine* e, long long req, float* host_out);
/* Test hook: teacher-forced logits of every position of `tokens` (chunked prefill with chunk
 * `chunk` tokens through the same kernels) -> HOST fp32 [n_tokens, vocab].  Engine must be idle. */
int ft_debug_prefill_logits(ft_engine* e, const int* tokens, int n_tokens, int chunk, float* host_out);

/* ------------------------------------------------------------------------------------------
 * Kernel level — one entry point per kernel so each can be checked against the oracle.
 * ---------------------------------------------------------------------------------------- */

/* K3: Y[t, n] = sum_k (X[t, k] + Xlo[t, k]) W[n, k]  (tcgen05/TMEM, TMA-fed).  x_lo (may be NULL) is the
 * bf16 residual plane of the activations: x = bf16(a), x_lo = bf16(a - x).
 * out_mode 0: fp32 partial sums out[split][T][n_out] (n_splits <= 0 picks a split count;
 *             ft_gemm_pick_splits tells which);
 * out_mode 1: bf16 out[T][n_out] (+ residual plane out_lo if not NULL);
 * out_mode 2: fused SwiGLU, w rows interleaved gate|up in blocks of 64 (see ft_weights.wgu):
 *             out[T][n_out/2] = bf16(silu(g) * u) (+ out_lo).
 * n_out % 128 == 0, K % 64 == 0. */
int ft_gemm_bf16(const void* w, const void* x, const void* x_lo, void* out, void* out_lo, int n_out, int T,
                 int K, int n_splits, int out_mode, void* stream);
int ft_gemm_pick_splits(int n_out, int T, int K, int hilo);
/* Tuning aid: globaltimer (ns) timeline of CTA (0,0,0) of the most recent GEMM launch -> HOST
 * uint64[16]: 0 entry, 1 set-up done, 2 weights requested, 3 dependency wait passed, 4 first stage
 * landed, 5 last MMA issued, 6 accumulator ready, 7 epilogue stores issued, 8 reduce done, 9 exit. */
int ft_debug_gemm_trace(unsigned long long* host_out16);
/* The same timeline for the last n_rows (<= 256) tile-per-CTA GEMM launches, oldest first -> HOST
 * uint64 [n_rows][16]; word 10 of a row = (epilogue mode << 32) | n_out.  Returns rows filled. */
int ft_debug_gemm_trace_hist(unsigned long long* host_out, int n_rows);

/* K3 with the fused epilogues the engine uses.  RMSNorm is applied AFTER the projection: the
 * activation planes hold resid * gamma (un-normalised), the per-token 1/rms is a scalar that
 * commutes with the GEMM, so it is applied in the consumer's epilogue from per-row-tile partial
 * sums of squares.  This removes every stand-alone norm / RoPE kernel from the decode chain.
 *   FT_EPI_SWIGLU  (2): as out_mode 2, gate and up scaled by rstd[t] first (ssq_in may be NULL)
 *   FT_EPI_RESID   (3): the split-K CTAs of a row tile park their fp32 partial tiles in an
 *                       L2-resident scratch buffer, rendezvous on a global arrival counter (the
 *                       grid is one wave, so all of them are resident) and each finishes a share
 *                       of the tokens:
 *                       resid[t][n] += y;  out/out_lo[t][n] = hi/lo(resid * gamma[n]);
 *                       ssq_out[row_tile][t] = sum over the tile's 128 rows of resid^2
 *   FT_EPI_QKV_ROPE(4): same reduction, y *= rstd[t], RoPE on q and k rows, q -> q_out
 *                       [T][nq*d] (fp16), k/v -> paged fp16 KV cache layer (layout of ft_rope_kv_append)
 *   FT_EPI_LOGITS  (5): out fp32 [T][n_out] = y * rstd[t]
 * rstd[t] = rsqrt(sum_p ssq_in[p * ssq_stride + t] * inv_h + eps).  x_lo is required. */
enum { FT_EPI_PARTIAL = 0, FT_EPI_BF16 = 1, FT_EPI_SWIGLU = 2, FT_EPI_RESID = 3, FT_EPI_QKV_ROPE = 4,
       FT_EPI_LOGITS = 5 };
typedef struct {
    int mode;
    void* out;                 /* see the mode */
    void* out_lo;
    float* resid;              /* RESID: fp32 [T][n_out], updated in place */
    const void* gamma;         /* RESID: bf16 [n_out], weight of the NEXT RMSNorm */
    float* ssq_out;            /* RESID: fp32 [n_out/128][ssq_stride] */
    const float* ssq_in;       /* SWIGLU / QKV_ROPE / LOGITS: fp32 [ssq_parts][ssq_stride] or NULL */
    int ssq_parts, ssq_stride;
    float inv_h, eps;
    const float* rope_cos;     /* QKV_ROPE: fp32 [pos][d/2] */
    const float* rope_sin;
    const int* positions;      /* QKV_ROPE: int32 [T] */
    const int* seq_rows;       /* QKV_ROPE: int32 [T], page-table row of each token */
    const int* page_table;     /* QKV_ROPE: int32 [rows][max_pages] */
    int max_pages;
    void* q_out;               /* QKV_ROPE: fp16 [T][nq*d] */
    void* kv_layer;            /* QKV_ROPE: paged KV cache layer */
    int nq, nkv, head_dim;
    float* scratch;            /* RESID / QKV_ROPE with n_splits > 1: ft_gemm_scratch_bytes() of device memory */
    int* sync;                 /* 〃: int32 [2 * row tiles * token tiles], zero before first use (self re-arming) */
    const void* prefetch;      /* optional: a contiguous region (the NEXT projection's weights) the CTAs pull */
    long long prefetch_bytes;  /*   into L2 while this kernel sits in its dependency / epilogue latency */
    /* Tensor parallelism over NVLink peer memory, fused into the epilogue (2 GPUs; NULL / 0 = off).
     * RESID: K is split across the two GPUs (row-parallel O / down projection).  Each CTA stores its
     * fp32 partial rows straight into the PEER's exchange buffer, raises the twin CTA's flag there
     * (release at system scope), waits for its own flag and adds what the peer stored here: the
     * all-reduce is part of this kernel, tile by tile, no collective call.  a + b == b + a, so both
     * GPUs end with bit-identical residuals.
     * LOGITS: the vocabulary is split; rows go to out[t * out_stride + out_col0 + n] locally and to the
     * same place in the peer's logits buffer (tp_peer_out). */
    float* tp_peer_xchg;       /* RESID: peer's exchange buffer [T][n_out] fp32 (this launch's parity half) */
    const float* tp_xchg;      /* RESID: this GPU's exchange buffer, written by the peer */
    unsigned* tp_peer_flags;   /* RESID: peer's per-CTA arrival words (>= grid size) */
    const unsigned* tp_flags;  /* RESID: this GPU's arrival words */
    unsigned tp_epoch;         /* RESID: value of this launch's arrivals; must grow launch over launch */
    void* tp_peer_out;         /* LOGITS: peer's logits buffer (same layout as out) or NULL */
    int out_stride, out_col0;  /* LOGITS: row stride / first column of out (0, 0 = n_out, 0) */
    int sync_mode;             /* split-K arrival words: 0 = self re-arming (any launch order); 1 / 2 = this launch
                                  counts on word 0 / 1 of each pair and zeroes word 1 / 0 of ALL sync_slots pairs —
                                  launches sharing `sync` must alternate 1, 2, 1, ... and each must complete before
                                  the next one's epilogue starts (stream order + the epilogue's dependency wait) */
    int sync_slots;            /* sync_mode 1 / 2: counter pairs in `sync` */
    void* out_f16;             /* RESID / SWIGLU, optional: the same values as out (+ out_lo) once more as ONE saturating fp16
                                  plane — the input of a following projection that runs in its single-plane form.  With
                                  out_f16 set, out (and out_lo) may be NULL: the bf16 planes are then not written */
    int x_planes;              /* 0 / 2: multiply both activation planes everywhere; 1: grids of more tiles than SMs with
                                  128-token tiles (prefill) multiply the hi plane only — half the tensor work there;
                                  3: x and w are FP16 tensors, ONE activation plane (x_lo = NULL), T > 64, K unsplit, any
                                  fused epilogue — the form the engine's large prefill steps run (CTA-pair kernels; long-K
                                  launches of two or more rounds use 256-token tiles and cut the last partial round into
                                  128-token halves) */
} ft_gemm_epilogue;
long long ft_gemm_scratch_bytes(int n_out, int T, int K, int n_splits);
int ft_gemm_fused(const void* w, const void* x, const void* x_lo, int n_out, int T, int K, int n_splits,
                  const ft_gemm_epilogue* epi, void* stream);

/* Layer-0 entry of the norm-deferred pipeline: resid[t] = embed[token[t]] (fp32);
 * out/out_lo[t] = hi/lo(resid * gamma); ssq[0][t] = sum resid^2, ssq[p>0][t] = 0. */
int ft_embed_planes(const int* tokens, const void* embed_table, const void* gamma, float* resid, void* out,
                    void* out_lo, float* ssq, int ssq_parts, int ssq_stride, int T, int h, void* stream);
/* Row gather of activation planes + their sums of squares (last-token rows for the LM head). */
int ft_gather_planes(const void* x, const void* x_lo, const float* ssq, int ssq_parts, int ssq_stride,
                     const int* row_idx, int n_rows, int h, void* out, void* out_lo, float* ssq_out,
                     int ssq_out_stride, void* stream);

/* K1+K2: resid[r] (fp32) = embed[token[r]]  or  resid[r] += sum_s partials[s][r];
 * y = rmsnorm(resid[r]) * w, r = row_idx ? row_idx[i] : i;  xn[i] = bf16(y), xn_lo[i] = bf16(y - xn[i])
 * (xn_lo may be NULL). */
int ft_add_rmsnorm(float* resid, const float* partials, int n_splits, long long split_stride,
                   const int* embed_tokens, const void* embed_table, const void* norm_w, void* xn_out,
                   void* xn_lo, const int* row_idx, int T, int h, float eps, void* stream);

/* K4: split-K reduce of the fused QKV projection, RoPE on q and k (cos/sin tables fp32
 * [pos][d/2]), q -> FP16 [T][nq][d], k/v -> paged cache layer
 * [page][kv_head][K|V][d/64][64 tokens][64 dims] FP16 (saturating): every K (V) slab is d/64 sub-slabs with
 * 128-byte token rows whose eight 16-byte chunks are XOR-swizzled with (row & 7) — conflict-free for ldmatrix and
 * exactly the 128-byte-swizzled UMMA operand layout (tests: kernels.kv_slab_logical / kv_slab_store).  Attention
 * runs in fp16 with fp32 accumulation: same bytes as bf16, three more mantissa bits, and the
 * dtype the reference deploys its engine with (docker-compose.vllm.yml:45 --dtype float16). */
int ft_rope_kv_append(const float* qkv, int n_splits, long long split_stride, const float* cos_t,
                      const float* sin_t, const int* positions, const int* seq_rows,
                      const int* page_table, int max_pages, void* q_out, void* kv_layer, int T, int nq,
                      int nkv, int d, void* stream);

/* SwiGLU on un-interleaved partials: a = silu(sum_s gu[s][t][j]) * sum_s gu[s][t][f+j];
 * act = bf16(a), act_lo = bf16(a - act) (may be NULL).  The engine uses the GEMM's fused epilogue. */
int ft_silu_mul(const float* gu, int n_splits, long long split_stride, void* act, void* act_lo, int T, int f,
                void* stream);

/* K5: paged GQA decode attention, one query token per sequence.  q FP16 [B][nq][d], KV cache FP16,
 * out bf16 [B][nq*d] (+ residual plane out_lo, may be NULL); a sequence with ctx_lens[b] == 0 gets zero rows.  `tile_prefix` (int32 [B+1]) and
 * `scratch` (ft_attn_decode_scratch_bytes) are device work buffers the call fills. */
long long ft_attn_decode_scratch_bytes(int B, int nkv, int group, int d, int n_ctas);
/* Tuning aid: per-CTA globaltimer (ns) timeline of the most recent ft_attn_decode launch ->
 * HOST uint64 [n_ctas][4]: entry, first tile landed, last tile consumed, exit. */
int ft_debug_attn_trace(unsigned long long* host_out, int n_ctas);
/* persistent grid size the engine uses: SM count x resident CTAs of the decode kernel */
int ft_attn_decode_default_ctas(int d);
int ft_attn_decode(const void* q, const void* kv_layer, const int* page_table, int max_pages,
                   const int* ctx_lens, int* tile_prefix, void* out, void* out_lo, float* scratch, int B, int nq,
                   int nkv, int d, int n_ctas, void* stream);

/* K5 without allocations — no counterpart in the reference (its engines sit behind HTTP and own
 * their memory).  Same computation as ft_attn_decode for hosts that must not allocate device
 * memory while serving (GPU memory carved up once at start-up; real-time audio threads where
 * cudaMalloc's device-wide synchronisation is a dropout): the tile-prefix table, arrival counters
 * and partial-result records live in a fixed ~6.1 MiB array inside the library, present on every
 * GPU that loads it until the process exits.  The caller passes only inputs and outputs.
 * Limits: B <= 1024 sequences, B * nkv <= 1024, n_ctas <= 512 (n_ctas <= 0 picks the default);
 * anything larger is refused with an error — use ft_attn_decode with your own scratch instead.
 * The first call on each GPU looks up the array's device address (cudaGetSymbolAddress, with
 * that GPU current); make that first call before starting any stream capture.  No cudaMalloc /
 * cudaFree / allocating library routine is called.  All calls on a GPU share the one array and
 * it keeps the previous call's contents: every call re-initialises what it needs (prefix table,
 * zeroed counters), but calls that overlap in time on the same GPU race on it — serialise them. */
int ft_attn_decode_noalloc(const void* q, const void* kv_layer, const int* page_table, int max_pages,
                           const int* ctx_lens, void* out, void* out_lo, int B, int nq, int nkv, int d,
                           int n_ctas, void* stream);

/* K6: causal chunked-prefill attention over the paged cache.  Request r owns query rows
 * [req_q_start[r], req_q_start[r+1]) of q/out, page-table row req_slot[r], and its first query
 * sits at absolute position req_pos0[r].  Tiles (tile_req, tile_q0) enumerate 64-query blocks.
 * q_rows = number of rows (tokens) the q and out buffers hold.  With a GQA group of 3 or 4 (Llama-3.2-1B / 3B,
 * Llama-3.1-8B), head_dim 64 or 128 and q_rows > 0 this runs on tcgen05 / TMEM (one CTA per 64 queries x KV head, S, P
 * and O in tensor memory, K/V slabs consumed as UMMA operands exactly as they lie in the cache); other shapes, or
 * q_rows <= 0, run the mma.sync kernel. */
int ft_attn_prefill(const void* q, const void* kv_layer, const int* page_table, int max_pages,
                    const int* req_q_start, const int* req_slot, const int* req_pos0, const int* tile_req,
                    const int* tile_q0, int n_tiles, void* out, void* out_lo, int nq, int nkv, int d,
                    int q_rows, void* stream);

/* K7: per-row temperature / top-k / top-p sampling (argmax when temperature <= 0). */
int ft_sample(const float* logits, int B, int V, const float* temperature, const int* top_k,
              const float* top_p, const unsigned long long* seeds, int* out_tokens, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* FASTTALK_B200_H */
