// The engine: KV page allocator, session table with prefix reuse, continuous-batching
// scheduler (decode rows + chunked-prefill rows in the same step) and the model forward that
// strings the kernels together.  Host logic only — every FLOP runs in the kernels of
// gemm.cu / elementwise.cu / attention.cu / sampling.cu.  There is no CPU fallback.
#include "common.cuh"
#include "ft_internal.h"
#include "../../include/fasttalk_b200.h"

#include <algorithm>
#include <chrono>
#include <condition_variable>
#include <deque>
#include <mutex>
#include <new>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <string>
#include <thread>
#include <unordered_map>
#include <vector>

namespace ft {

static inline int ceil_div(int a, int b) { return (a + b - 1) / b; }

struct Session {
    unsigned long long key = 0;
    std::vector<int> pages;      // KV pages in position order
    std::vector<int> tokens;     // tokens whose K/V sit in the cache (positions 0..size-1)
    bool busy = false;
    bool close_pending = false;  // session_close arrived while its request was inside a running step
    long long last_used = 0;
};

enum ReqState { REQ_WAITING = 0, REQ_PREFILL = 1, REQ_DECODE = 2, REQ_DONE = 3 };

struct Request {
    long long id = 0;
    Session* sess = nullptr;
    std::vector<int> prompt;     // full token list of this turn
    int prefill_pos = 0;         // next prompt index to run through the model
    float temperature = 0.f, top_p = 1.f;
    int top_k = 0, max_new = 0, n_generated = 0, ignore_eos = 0;
    unsigned long long seed = 0;
    int state = REQ_WAITING;
    int last_token = -1;         // sampled but not yet fed back
    int logits_row = -1;         // row of the logits buffer in the latest step
    bool in_flight = false;      // part of the step that is on the GPU right now (the engine lock is released
                                 // while the step runs, so submit / cancel / close do not wait for the GPU)
    bool cancel_pending = false; // cancelled while in flight: finished as "cancelled" when the step returns
};

struct OutTok { long long req; int token; int finish; };

// host-pinned + device mirror of everything a step's kernels index
struct StepMeta {
    int* h = nullptr;
    int* d = nullptr;
    size_t ints = 0;
    // offsets (in ints).  Fixed-size arrays come first; the page table and the per-token arrays
    // are packed behind them at per-step offsets so one H2D copy moves only what the step uses.
    size_t o_ctx, o_qstart, o_rslot, o_rpos0, o_treq, o_tq0, o_lrow, o_temp, o_topk, o_topp, o_seed,
        o_pt;                                     // fixed
    size_t o_tokens, o_pos, o_rows, upload_ints;  // set by layout_step()
    size_t o_tprefix, o_sampled;                  // device-side work areas (never uploaded)
};

static inline size_t align4(size_t n) { return (n + 3) & ~size_t(3); }
static void layout_step(StepMeta& m, int n_req, int max_pages, int T) {
    m.o_tokens = m.o_pt + align4((size_t)n_req * max_pages);
    m.o_pos = m.o_tokens + align4(T);
    m.o_rows = m.o_pos + align4(T);
    m.upload_ints = m.o_rows + align4(T);
}

struct Engine {
    ft_engine_config cfg;
    ft_weights w;
    int dev = 0, sms = 148;
    cudaStream_t stream = nullptr;
    std::mutex mu;               // scheduler state; NOT held while a step's kernels run
    std::mutex step_mu;          // one step (or debug / timer call that uses the stream's buffers) at a time

    int q_dim, kv_dim, qkv_dim, group, max_pages, t_max, r_max, logit_rows;
    int attn_ctas = 0;
    int decode_prefill_cap = 0;            // prompt tokens admitted per step while sequences are decoding (0 = no cap)
    int prefill_planes = 2;                // activation planes of the multi-wave prefill GEMMs (1 = hi only, 3 = one fp16 plane)
    // prefill_planes == 3: fp16 copies of the four projections' weights and fp16 staging planes of their inputs
    void *w16_qkv = nullptr, *w16_o = nullptr, *w16_gu = nullptr, *w16_d = nullptr;
    void *x16_xn = nullptr, *x16_attn = nullptr, *x16_act = nullptr;
    std::vector<CUtensorMap> m16_wqkv, m16_wo, m16_wgu, m16_wd;
    CUtensorMap m16_xn[2], m16_attn[2], m16_act[2];      // token boxes of 128 and 64 rows
    size_t kv_layer_bytes = 0;

    // device buffers
    uint8_t* kv = nullptr;
    float* resid = nullptr;
    float* ssq = nullptr;          // [ssq_parts][t_max] partial sums of squares of the residual rows
    float* ssq_last = nullptr;     // [ssq_parts][rows_last] gathered for the LM-head rows
    int ssq_parts = 0, rows_last = 0;
    float* gemm_scratch = nullptr; // split-K slabs of the fused-epilogue GEMMs (one wave of CTAs)
    int* gemm_sync = nullptr;      // their arrival counters
    int sync_parity = 0;           // which word of each counter pair the next split-K launch uses
    // bf16 activations feeding the GEMMs are kept as hi + lo planes (see gemm.cu)
    void *xn = nullptr, *q = nullptr, *attn_out = nullptr, *act = nullptr, *xn_last = nullptr;
    void *xn_lo = nullptr, *attn_out_lo = nullptr, *act_lo = nullptr, *xn_last_lo = nullptr;
    float* logits = nullptr;
    float* attn_scratch = nullptr;
    StepMeta meta;

    // tensor parallelism (2 ranks): buffers the peer writes into, and the peer's as mapped here
    int tp = 1, vocab_local = 0;
    float* tp_xchg = nullptr;          // [2 parities][t_max][hidden] fp32 partial rows from the peer
    unsigned* tp_flags = nullptr;      // [kTpFlags] arrival words written by the peer
    float* peer_xchg = nullptr;
    unsigned* peer_flags = nullptr;
    float* peer_logits = nullptr;
    unsigned tp_epoch = 0;             // advances once per exchanging launch, identically on both ranks

    // tensor maps: weights per layer, activations per BN
    std::vector<CUtensorMap> m_wqkv, m_wo, m_wgu, m_wd;
    CUtensorMap m_lm;
    CUtensorMap m_xn[4], m_attn[4], m_act[4], m_last[4];              // token tiles 16/32/64/128
    CUtensorMap m_xn_lo[4], m_attn_lo[4], m_act_lo[4], m_last_lo[4];

    // scheduler state
    std::vector<int> free_pages;
    std::unordered_map<unsigned long long, Session*> sessions;
    std::unordered_map<long long, Request*> requests;
    std::deque<Request*> waiting;
    std::vector<Request*> active;
    std::deque<OutTok> out;
    long long next_req = 0, clock = 0;
    ft_engine_stats st = {};
    bool pdl = false;
    // optional CUDA-event timing of the decode-attention kernel (bench roofline)
    bool profile = false;
    std::vector<cudaEvent_t> prof_ev;      // 2 per layer
    int prof_used = 0;
    long long prof_ctx_sum = 0;            // sum of context lengths of the step's decode rows
    cudaEvent_t t0 = nullptr, t1 = nullptr;
    // device time of every step (events on the engine stream around upload .. sampled-token copy)
    cudaEvent_t s0 = nullptr, s1 = nullptr;
    struct StepRec { float gpu_ms, n_decode, n_prefill, host_ms; };
    std::deque<StepRec> step_log;          // most recent kStepTimes steps, oldest first

    // the engine's own stepping thread (ft_engine_start): steps whenever something is runnable, so that no host
    // language sits between two steps; the owner only submits, waits for events and drains them
    std::thread loop_thread;
    bool loop_on = false, loop_stop = false;
    int loop_error = 0;
    std::string loop_error_msg;
    std::condition_variable cv_work;       // submit / cancel / stop -> the loop
    std::condition_variable cv_out;        // events queued, loop failed or stopped -> ft_engine_wait
    int waiters = 0;                       // threads inside ft_engine_wait
    bool closing = false;                  // ft_engine_destroy has begun: waiters leave, nobody new comes in
};
constexpr size_t kStepTimes = 4096;
constexpr int kDefaultPrefillPlanes = 3;      // one fp16 plane x fp16 weight copies in the large prefill GEMMs (measured: DESIGN.md §2)
constexpr int kGemmTokenTile = 128;             // token tile of the multi-tile projections (gemm_pick_bn)
constexpr int kDefaultDecodePrefillCap = 896;    // <= 128 decode rows + 896 -> 1024 rows: a 1B step of ~4.3 ms (1280 rows: 6.6 ms)

// Every entry point that touches CUDA makes the engine's device current for its duration: the caller may be a
// thread whose current device is another GPU (a fresh thread starts on device 0).
// Does `dev` have a live primary context in this process?  (driver API through the runtime's entry-point lookup; on
// any doubt: yes)
static bool primary_ctx_active(int dev) {
    typedef CUresult (*Fn)(CUdevice, unsigned int*, int*);
    static Fn fn = [] {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuDevicePrimaryCtxGetState", &p, cudaEnableDefault, &q) != cudaSuccess ||
            q != cudaDriverEntryPointSuccess)
            p = nullptr;
        return reinterpret_cast<Fn>(p);
    }();
    unsigned int flags = 0;
    int active = 1;
    if (!fn || fn((CUdevice)dev, &flags, &active) != CUDA_SUCCESS) return true;
    return active != 0;
}

struct DeviceGuard {
    int prev = -1, dev;
    explicit DeviceGuard(int d) : dev(d) {
        if (cudaGetDevice(&prev) != cudaSuccess) prev = -1;
        if (prev != dev) cudaSetDevice(dev);
    }
    ~DeviceGuard() {
        // Back to the caller's device — unless that is only the default of a thread that never touched a GPU: since
        // CUDA 12 cudaSetDevice creates the device's context, and "restoring" device 0 from a process that lives on
        // another GPU would build one there (0.4 s and ~0.5 GB on a GPU that belongs to another rank).
        if (prev >= 0 && prev != dev && primary_ctx_active(prev)) cudaSetDevice(prev);
    }
};
// The split-K rendezvous and the tensor-parallel flag wait spin inside kernels and need every CTA of a launch
// resident at once.  One engine alone on a GPU guarantees that (grids are sized to one wave); two engines in one
// process on the SAME GPU (two models behind one server) could each hold part of the SMs and wait for CTAs that
// can never be scheduled.  Steps are synchronous (they end with a stream synchronise), so serialising the steps
// of engines that share a device removes the hazard without costing a lone engine anything.
static std::mutex g_device_step_mu[kMaxDevices];

constexpr int kTpFlags = 16384;        // >= CTAs of any exchanging GEMM launch; the last word is the step barrier

// Step barrier between the two ranks before sampling: both halves of the logits are in place.
__global__ void tp_barrier_kernel(unsigned* peer_flag, const unsigned* flag, unsigned epoch) {
    if (threadIdx.x == 0) {
        __threadfence_system();
        asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(peer_flag), "r"(epoch) : "memory");
        unsigned v, spins = 0;
        do {
            asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(flag) : "memory");
            if ((int)(v - epoch) >= 0) break;
            __nanosleep(64);
            if (++spins > (1u << 25)) __trap();
        } while (true);
    }
}

static int bn_index(int bn) { return bn == 16 ? 0 : bn == 32 ? 1 : bn == 64 ? 2 : 3; }
static const int kBns[4] = {16, 32, 64, 128};

// ------------------------------------------------------------------ pages / sessions
static void release_pages_beyond(Engine* e, Session* s, int keep_tokens) {
    const int keep = ceil_div(keep_tokens, kKvTile);
    while ((int)s->pages.size() > keep) {
        e->free_pages.push_back(s->pages.back());
        s->pages.pop_back();
    }
    if ((int)s->tokens.size() > keep_tokens) s->tokens.resize(keep_tokens);
}

static void drop_session(Engine* e, Session* s) {
    release_pages_beyond(e, s, 0);
    e->sessions.erase(s->key);
    delete s;
}

// Free pages by evicting idle sessions, least recently used first.
static bool reclaim_pages(Engine* e, int need) {
    while ((int)e->free_pages.size() < need) {
        Session* victim = nullptr;
        for (auto& kv : e->sessions)
            if (!kv.second->busy && !kv.second->pages.empty() &&
                (!victim || kv.second->last_used < victim->last_used))
                victim = kv.second;
        if (!victim) return false;
        release_pages_beyond(e, victim, 0);
    }
    return true;
}

static bool ensure_pages(Engine* e, Session* s, int n_tokens) {
    const int need = ceil_div(n_tokens, kKvTile) - (int)s->pages.size();
    if (need <= 0) return true;
    if ((int)e->free_pages.size() < need && !reclaim_pages(e, need)) return false;
    for (int i = 0; i < need; ++i) {
        s->pages.push_back(e->free_pages.back());
        e->free_pages.pop_back();
    }
    return true;
}

static void finish_request(Engine* e, Request* r, int reason, int token) {
    e->out.push_back({r->id, token, reason});
    r->state = REQ_DONE;
    if (r->sess) {
        r->sess->busy = false;
        r->sess->last_used = ++e->clock;
    }
}

// ------------------------------------------------------------------ forward
struct Batch {
    int n_decode = 0;                 // rows [0, n_decode): one token per sequence
    int T = 0;                        // total token rows
    int n_req = 0;                    // page-table rows (decode first, then prefill requests)
    int n_prefill_req = 0;
    int n_tiles = 0;                  // prefill attention q tiles
    int n_logit_rows = 0;
    bool logit_rows_identity = false; // logits wanted for rows 0..n_logit_rows-1 exactly (no gather)
};

#define FT_TRY(x)                                   \
    do {                                            \
        int _r = (x);                               \
        if (_r) return _r;                          \
    } while (0)

// One projection with a fused epilogue.  `staged` epilogues (RESID / QKV_ROPE) split K across
// one wave of CTAs that rendezvous in the epilogue; the others run unsplit.
struct F16Path {                 // the fp16 single-plane form of one projection (prefill_planes == 3); null maps = not available
    const CUtensorMap* mw = nullptr;
    const CUtensorMap* mx = nullptr;         // [2]: 128- and 64-row boxes
    const void *hi = nullptr, *lo = nullptr;
    void* x16 = nullptr;
    bool ready = false;                      // the producer's epilogue already wrote x16: no staging kernel
};

static int run_gemm(Engine* e, const CUtensorMap& mw, const CUtensorMap* xmaps, const CUtensorMap* xmaps_lo,
                    int T, int n_out, int K, ft_gemm_epilogue epi, const F16Path* f16 = nullptr) {
    const int bn = gemm_pick_bn(T, true);
    const bool staged = epi.mode == FT_EPI_RESID || epi.mode == FT_EPI_QKV_ROPE;
    const int splits = staged ? gemm_pick_splits(n_out, T, K, e->sms, true) : 1;
    epi.scratch = e->gemm_scratch;
    epi.sync = e->gemm_sync;
    epi.x_planes = e->prefill_planes == 3 ? 2 : e->prefill_planes;
    if (staged && splits > 1) {              // launches are stream-ordered and each epilogue waits for its predecessor
        epi.sync_mode = 1 + e->sync_parity;
        epi.sync_slots = e->sms;
        e->sync_parity ^= 1;
    }
    if (e->tp > 1 && epi.mode == FT_EPI_RESID) {
        // row-parallel projection: the all-reduce over the two ranks happens inside the epilogue
        const unsigned epoch = ++e->tp_epoch;
        const size_t half = (size_t)e->t_max * e->cfg.hidden;
        epi.tp_peer_xchg = e->peer_xchg + (epoch & 1) * half;
        epi.tp_xchg = e->tp_xchg + (epoch & 1) * half;
        epi.tp_peer_flags = e->peer_flags;
        epi.tp_flags = e->tp_flags;
        epi.tp_epoch = epoch;
        const int ctas = (n_out / 128) * ((T + bn - 1) / bn) * splits;
        FT_REQUIRE(ctas < kTpFlags, "engine: %d CTAs exceed the tensor-parallel arrival words", ctas);
    }
    e->st.kernel_launches++;
    if (f16 && f16->mw && e->prefill_planes == 3 && gemm_f16_path_ok(n_out, T, K, splits)) {
        // large prefill projection: ONE fp16 plane (= fp16(hi + lo), staged by a small kernel) against fp16 weights
        if (!f16->ready) {
            FT_TRY(launch_planes_to_f16(f16->hi, f16->lo, f16->x16, (size_t)T * K, e->pdl, e->stream));
            e->st.kernel_launches++;
        }
        epi.x_planes = 3;
        return launch_gemm(*f16->mw, f16->mx[0], &f16->mx[0], bn, n_out, T, K, 1, epi, e->pdl, e->stream, &f16->mx[1],
                           &f16->mx[1]);
    }
    return launch_gemm(mw, xmaps[bn_index(bn)], &xmaps_lo[bn_index(bn)], bn, n_out, T, K, splits, epi, e->pdl,
                       e->stream, &xmaps[bn_index(64)], &xmaps_lo[bn_index(64)]);
}

// The norm-deferred layer chain (5 kernels per layer):
//   QKV GEMM [x 1/rms, RoPE, KV append] -> attention -> O GEMM [resid +=, planes = resid*gamma_mlp, ssq]
//   -> gate/up GEMM [x 1/rms, SwiGLU] -> down GEMM [resid +=, planes = resid*gamma_next, ssq]
// `xn` planes always hold resid * gamma of the NEXT norm; `ssq` its per-row-tile sums of squares.
static int forward(Engine* e, const Batch& b) {
    const ft_engine_config& c = e->cfg;
    StepMeta& m = e->meta;
    int* d = m.d;
    const int T = b.T;
    const size_t H = c.hidden;
    auto bf16_off = [](const void* p, size_t elems) {
        return (const void*)((const uint8_t*)p + elems * 2);
    };
    FT_TRY(launch_embed_planes(d + m.o_tokens, e->w.embed, e->w.attn_norm, e->resid, e->xn, e->xn_lo, e->ssq,
                               e->ssq_parts, e->t_max, T, c.hidden, e->pdl, e->stream));
    e->st.kernel_launches++;
    if (b.n_decode > 0) {
        FT_TRY(launch_attn_decode_plan(d + m.o_ctx, d + m.o_tprefix, b.n_decode, c.n_kv_heads, e->attn_scratch,
                                       e->pdl, e->stream));
        e->st.kernel_launches++;
    }
    // L2 prefetch of the next projection's weights only pays when the GEMMs are latency-bound
    const bool decode_only = b.n_tiles == 0 && T <= 128;
    const bool use16 = e->prefill_planes == 3 && e->w16_qkv != nullptr && T > 128;
    // which projections take the one-fp16-plane form this step: their producers write that plane in their epilogues
    // (O -> gate/up input, down -> next QKV input, gate/up -> down input); layer 0's QKV input (embedding kernel) and the
    // attention output are staged by planes_to_f16 instead
    static const bool direct = getenv("FT_F16_STAGE_ALL") == nullptr;          // A/B switch: stage every plane with planes_to_f16
    const bool f16_qkv = direct && use16 && gemm_f16_path_ok(e->qkv_dim, T, c.hidden, gemm_pick_splits(e->qkv_dim, T, c.hidden, e->sms, true));
    const bool f16_gu = direct && use16 && gemm_f16_path_ok(2 * c.ffn, T, c.hidden, 1);
    const bool f16_d = direct && use16 && gemm_f16_path_ok(c.hidden, T, c.ffn, gemm_pick_splits(c.hidden, T, c.ffn, e->sms, true));
    ft_gemm_epilogue rstd = {};            // fields shared by every epilogue that applies 1/rms
    rstd.ssq_in = e->ssq;
    rstd.ssq_parts = e->ssq_parts;
    rstd.ssq_stride = e->t_max;
    rstd.inv_h = 1.f / (float)c.hidden;
    rstd.eps = c.rms_eps;
    for (int l = 0; l < c.layers; ++l) {
        uint8_t* kv_l = e->kv + (size_t)l * e->kv_layer_bytes;
        {
            ft_gemm_epilogue ep = rstd;
            ep.mode = FT_EPI_QKV_ROPE;
            ep.rope_cos = e->w.rope_cos;
            ep.rope_sin = e->w.rope_sin;
            ep.positions = d + m.o_pos;
            ep.seq_rows = d + m.o_rows;
            ep.page_table = d + m.o_pt;
            ep.max_pages = e->max_pages;
            ep.q_out = e->q;
            ep.kv_layer = kv_l;
            ep.nq = c.n_q_heads;
            ep.nkv = c.n_kv_heads;
            ep.head_dim = c.head_dim;
            const F16Path f = {use16 ? &e->m16_wqkv[l] : nullptr, e->m16_xn, e->xn, e->xn_lo, e->x16_xn, l > 0 && f16_qkv};
            FT_TRY(run_gemm(e, e->m_wqkv[l], e->m_xn, e->m_xn_lo, T, e->qkv_dim, c.hidden, ep, &f));
        }
        if (b.n_decode > 0) {
            cudaEvent_t ev0 = nullptr, ev1 = nullptr;
            if (e->profile) {
                ev0 = e->prof_ev[2 * l];
                ev1 = e->prof_ev[2 * l + 1];
                e->prof_used = l + 1;
            }
            FT_TRY(launch_attn_decode(e->q, kv_l, d + m.o_pt, e->max_pages, d + m.o_ctx, d + m.o_tprefix,
                                      e->attn_out, e->attn_out_lo, e->attn_scratch, b.n_decode, c.n_q_heads,
                                      c.n_kv_heads, c.head_dim, e->attn_ctas, e->pdl, e->stream, ev0, ev1,
                                      /*early=*/l > 0 && b.n_tiles == 0));
            e->st.kernel_launches++;
        }
        if (b.n_tiles > 0) {
            FT_TRY(launch_attn_prefill(e->q, kv_l, d + m.o_pt, e->max_pages, d + m.o_qstart, d + m.o_rslot,
                                       d + m.o_rpos0, d + m.o_treq, d + m.o_tq0, b.n_tiles, e->attn_out,
                                       e->attn_out_lo, c.n_q_heads, c.n_kv_heads, c.head_dim, e->pdl,
                                       e->stream, e->t_max));
            e->st.kernel_launches++;
        }
        {
            ft_gemm_epilogue ep = {};
            ep.mode = FT_EPI_RESID;
            ep.resid = e->resid;
            ep.gamma = bf16_off(e->w.mlp_norm, (size_t)l * H);
            ep.out = e->xn;
            ep.out_lo = e->xn_lo;
            ep.ssq_out = e->ssq;
            ep.ssq_stride = e->t_max;
            if (f16_gu) {           // the gate/up projection reads the fp16 plane alone: no hi / lo planes to write
                ep.out_f16 = e->x16_xn;
                ep.out = ep.out_lo = nullptr;
            }
            if (decode_only) {      // stream the gate/up weights into L2 behind this latency-bound kernel
                ep.prefetch = bf16_off(e->w.wgu, (size_t)l * 2 * c.ffn * H);
                ep.prefetch_bytes = (long long)2 * c.ffn * H * 2;
            }
            const F16Path f = {use16 ? &e->m16_wo[l] : nullptr, e->m16_attn, e->attn_out, e->attn_out_lo, e->x16_attn};
            FT_TRY(run_gemm(e, e->m_wo[l], e->m_attn, e->m_attn_lo, T, c.hidden, e->q_dim, ep, &f));
        }
        {   // gate/up projection: 1/rms and SwiGLU in the epilogue, act = silu(g) * u
            ft_gemm_epilogue ep = rstd;
            ep.mode = FT_EPI_SWIGLU;
            ep.out = e->act;
            ep.out_lo = e->act_lo;
            if (f16_d) {
                ep.out_f16 = e->x16_act;
                ep.out = ep.out_lo = nullptr;
            }
            if (decode_only) {
                ep.prefetch = bf16_off(e->w.wd, (size_t)l * c.ffn * H);
                ep.prefetch_bytes = (long long)c.ffn * H * 2;
            }
            const F16Path f = {use16 ? &e->m16_wgu[l] : nullptr, e->m16_xn, e->xn, e->xn_lo, e->x16_xn, f16_gu};
            FT_TRY(run_gemm(e, e->m_wgu[l], e->m_xn, e->m_xn_lo, T, 2 * c.ffn, c.hidden, ep, &f));
        }
        {
            ft_gemm_epilogue ep = {};
            ep.mode = FT_EPI_RESID;
            ep.resid = e->resid;
            ep.gamma = l + 1 < c.layers ? bf16_off(e->w.attn_norm, (size_t)(l + 1) * H) : e->w.final_norm;
            ep.out = e->xn;
            ep.out_lo = e->xn_lo;
            ep.ssq_out = e->ssq;
            ep.ssq_stride = e->t_max;
            if (f16_qkv && l + 1 < c.layers) {     // (the last layer's planes feed the LM head's row gather)
                ep.out_f16 = e->x16_xn;
                ep.out = ep.out_lo = nullptr;
            }
            if (decode_only && l + 1 < c.layers) {
                ep.prefetch = bf16_off(e->w.wqkv, (size_t)(l + 1) * e->qkv_dim * H);
                ep.prefetch_bytes = (long long)e->qkv_dim * H * 2;
            }
            const F16Path f = {use16 ? &e->m16_wd[l] : nullptr, e->m16_act, e->act, e->act_lo, e->x16_act, f16_d};
            FT_TRY(run_gemm(e, e->m_wd[l], e->m_act, e->m_act_lo, T, c.hidden, c.ffn, ep, &f));
        }
    }
    if (b.n_logit_rows > 0) {
        // LM head on the rows that need logits (all of them in a pure decode step)
        const CUtensorMap* xm = e->m_xn;
        const CUtensorMap* xl = e->m_xn_lo;
        ft_gemm_epilogue ep = rstd;
        ep.mode = FT_EPI_LOGITS;
        ep.out = e->logits;
        if (!b.logit_rows_identity) {
            FT_TRY(launch_gather_planes(e->xn, e->xn_lo, e->ssq, e->ssq_parts, e->t_max, d + m.o_lrow,
                                        b.n_logit_rows, c.hidden, e->xn_last, e->xn_last_lo, e->ssq_last,
                                        e->rows_last, e->pdl, e->stream));
            e->st.kernel_launches++;
            xm = e->m_last;
            xl = e->m_last_lo;
            ep.ssq_in = e->ssq_last;
            ep.ssq_stride = e->rows_last;
        }
        if (e->tp > 1) {                 // vocabulary-parallel: each rank fills its columns on both GPUs
            ep.out_stride = c.vocab;
            ep.out_col0 = c.tp_rank * e->vocab_local;
            ep.tp_peer_out = e->peer_logits;
        }
        FT_TRY(run_gemm(e, e->m_lm, xm, xl, b.n_logit_rows, e->vocab_local, c.hidden, ep));
        if (e->tp > 1) {
            const unsigned epoch = ++e->tp_epoch;
            tp_barrier_kernel<<<1, 32, 0, e->stream>>>(e->peer_flags + kTpFlags - 1, e->tp_flags + kTpFlags - 1, epoch);
            FT_CHECK_CUDA(cudaGetLastError());
            e->st.kernel_launches++;
        }
    }
    return 0;
}

static int upload_meta(Engine* e) {
    FT_CHECK_CUDA(cudaMemcpyAsync(e->meta.d, e->meta.h, e->meta.upload_ints * sizeof(int),
                                  cudaMemcpyHostToDevice, e->stream));
    e->st.h2d_bytes += (long long)(e->meta.upload_ints * sizeof(int));
    return 0;
}

static unsigned long long mix64(unsigned long long x) {
    x += 0x9E3779B97F4A7C15ull;
    x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
    x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
    return x ^ (x >> 31);
}

// ------------------------------------------------------------------ one scheduler step
static int engine_step(Engine* e, std::unique_lock<std::mutex>& lk) {
    const ft_engine_config& c = e->cfg;
    StepMeta& m = e->meta;
    const auto t_begin = std::chrono::steady_clock::now();

    // retire finished requests from the active list
    e->active.erase(std::remove_if(e->active.begin(), e->active.end(),
                                   [&](Request* r) { return r->state == REQ_DONE; }),
                    e->active.end());
    // admit
    while (!e->waiting.empty() && (int)e->active.size() < c.max_batch) {
        Request* r = e->waiting.front();
        if (r->state == REQ_DONE) { e->waiting.pop_front(); continue; }
        if (!ensure_pages(e, r->sess, (int)r->prompt.size() + 1)) {
            if (e->active.empty()) {            // nothing will ever free pages for it
                e->waiting.pop_front();
                finish_request(e, r, 4, -1);
                continue;
            }
            break;
        }
        e->waiting.pop_front();
        r->state = REQ_PREFILL;
        e->active.push_back(r);
    }
    if (e->active.empty()) return 0;

    // ---- build the batch: decode rows first, then prefill chunks up to the token budget
    Batch b;
    int* h = m.h;
    std::vector<Request*> dec, pre;
    std::vector<int> pre_len;
    for (Request* r : e->active)
        if (r->state == REQ_DECODE) dec.push_back(r);
    int budget = c.max_batched_tokens - (int)dec.size();
    // Sessions that are streaming wait for the whole step: beside decode rows a step admits about
    // `decode_prefill_cap` prompt tokens (a ~1k-token chunk adds a few ms to a 1B step instead of the ~20 ms a full
    // 8192-token budget costs).  Once more than two sessions wait for their first token PER streaming session the
    // allowance grows in that proportion (cap x waiting / (2 x streaming)): a burst of 63 arrivals behind one early
    // bird gets the whole budget (one session waits 20 ms, 63 get their first token soonest), 56 arrivals beside 8
    // streaming sessions 3.5 x the cap, 35 beside 29 the cap — no cliff at "more waiting than streaming" that would
    // stall half of the sessions for a full-budget step.
    int n_pending = 0;
    for (Request* r : e->active) n_pending += r->state == REQ_PREFILL;
    if (e->decode_prefill_cap > 0 && !dec.empty()) {
        const long long n_dec = (long long)dec.size();
        const long long allow = (long long)e->decode_prefill_cap * std::max<long long>(n_pending, 2 * n_dec) / (2 * n_dec);
        // the projections work on 128-token tiles, so the rows up to the next tile boundary ride for free
        const long long rows = n_dec + allow;
        const long long tiled = (rows + kGemmTokenTile - 1) / kGemmTokenTile * kGemmTokenTile - n_dec;
        budget = (int)std::min<long long>(budget, tiled);
    }
    for (Request* r : e->active) {
        if (r->state != REQ_PREFILL || budget <= 0) continue;
        const int remaining = (int)r->prompt.size() - r->prefill_pos;
        const int take = std::min(remaining, budget);
        pre.push_back(r);
        pre_len.push_back(take);
        budget -= take;
    }
    // page growth for decode rows (evicting idle sessions if needed)
    for (size_t i = 0; i < dec.size();) {
        Request* r = dec[i];
        if (!ensure_pages(e, r->sess, (int)r->sess->tokens.size() + 1)) {
            finish_request(e, r, 2, -1);
            dec.erase(dec.begin() + i);
        } else {
            ++i;
        }
    }
    b.n_decode = (int)dec.size();
    b.n_prefill_req = (int)pre.size();
    b.n_req = b.n_decode + b.n_prefill_req;
    if (b.n_req == 0) return 0;
    {
        int total = b.n_decode;
        for (int len : pre_len) total += len;
        layout_step(m, b.n_req, e->max_pages, total);
    }

    int row = 0, lrow = 0;
    long long ctx_sum = 0;
    for (int i = 0; i < b.n_decode; ++i, ++row) {
        Request* r = dec[i];
        Session* s = r->sess;
        const int pos = (int)s->tokens.size();
        h[m.o_tokens + row] = r->last_token;
        h[m.o_pos + row] = pos;
        h[m.o_rows + row] = i;
        h[m.o_ctx + i] = pos + 1;
        ctx_sum += pos + 1;
        for (size_t p = 0; p < s->pages.size(); ++p) h[m.o_pt + (size_t)i * e->max_pages + p] = s->pages[p];
        h[m.o_lrow + lrow] = row;
        r->logits_row = lrow++;
    }
    e->prof_ctx_sum = ctx_sum;
    e->st.decode_ctx_tokens += ctx_sum;
    int tiles = 0;
    h[m.o_qstart] = row;
    for (int j = 0; j < b.n_prefill_req; ++j) {
        Request* r = pre[j];
        Session* s = r->sess;
        const int pt_row = b.n_decode + j;
        const int pos0 = r->prefill_pos;
        for (int k = 0; k < pre_len[j]; ++k, ++row) {
            h[m.o_tokens + row] = r->prompt[pos0 + k];
            h[m.o_pos + row] = pos0 + k;
            h[m.o_rows + row] = pt_row;
        }
        h[m.o_qstart + j + 1] = row;
        h[m.o_rslot + j] = pt_row;
        h[m.o_rpos0 + j] = pos0;
        for (size_t p = 0; p < s->pages.size(); ++p)
            h[m.o_pt + (size_t)pt_row * e->max_pages + p] = s->pages[p];
        for (int q0 = 0; q0 < pre_len[j]; q0 += 64, ++tiles) {
            h[m.o_treq + tiles] = j;
            h[m.o_tq0 + tiles] = q0;
        }
        r->logits_row = -1;
        if (pos0 + pre_len[j] == (int)r->prompt.size()) {      // prompt complete: sample from last row
            h[m.o_lrow + lrow] = row - 1;
            r->logits_row = lrow++;
        }
    }
    b.T = row;
    b.n_tiles = tiles;
    b.n_logit_rows = lrow;
    b.logit_rows_identity = lrow == b.T && b.n_prefill_req == 0;      // pure decode: row i wants logits row i
    // sampling parameters per logit row
    float* hf = reinterpret_cast<float*>(h);
    unsigned long long* hseed = reinterpret_cast<unsigned long long*>(h + m.o_seed);
    auto put_sampling = [&](Request* r) {
        const int lr = r->logits_row;
        hf[m.o_temp + lr] = r->temperature;
        h[m.o_topk + lr] = r->top_k;
        hf[m.o_topp + lr] = r->top_p;
        hseed[lr] = mix64(r->seed * 0x100000001B3ull + (unsigned long long)r->n_generated);
    };
    for (Request* r : dec) put_sampling(r);
    for (Request* r : pre) if (r->logits_row >= 0) put_sampling(r);

    e->prof_used = 0;
    FT_CHECK_CUDA(cudaEventRecord(e->s0, e->stream));
    FT_TRY(upload_meta(e));
    FT_TRY(forward(e, b));
    if (b.n_logit_rows > 0) {
        const float* dtemp = reinterpret_cast<const float*>(m.d + m.o_temp);
        const float* dtopp = reinterpret_cast<const float*>(m.d + m.o_topp);
        const unsigned long long* dseed = reinterpret_cast<const unsigned long long*>(m.d + m.o_seed);
        FT_TRY(launch_sample(e->logits, b.n_logit_rows, c.vocab, dtemp, m.d + m.o_topk, dtopp, dseed,
                             m.d + m.o_sampled, e->pdl, e->stream));
        e->st.kernel_launches++;
        FT_CHECK_CUDA(cudaMemcpyAsync(h + m.o_sampled, m.d + m.o_sampled, b.n_logit_rows * sizeof(int),
                                      cudaMemcpyDeviceToHost, e->stream));
    }
    FT_CHECK_CUDA(cudaEventRecord(e->s1, e->stream));
    // The GPU works for milliseconds from here: give the scheduler lock back so that submit / cancel / close /
    // drain / stats from other threads return at once (a submit used to wait for the whole step, which
    // serialised 64 sessions arriving together into 64 steps).  What they may do meanwhile is limited to state
    // this step does not use: requests in flight are only MARKED (cancel_pending / close_pending).
    for (Request* r : dec) r->in_flight = true;
    for (Request* r : pre) r->in_flight = true;
    lk.unlock();
    const cudaError_t sync_rc = cudaStreamSynchronize(e->stream);
    lk.lock();
    for (Request* r : dec) r->in_flight = false;
    for (Request* r : pre) r->in_flight = false;
    FT_CHECK_CUDA(sync_rc);
    e->st.d2h_bytes += (long long)b.n_logit_rows * sizeof(int);
    {
        float ms = 0.f;
        if (cudaEventElapsedTime(&ms, e->s0, e->s1) == cudaSuccess) {
            if (e->step_log.size() >= kStepTimes) e->step_log.pop_front();
            const float host_ms =
                std::chrono::duration<float, std::milli>(std::chrono::steady_clock::now() - t_begin).count();
            e->step_log.push_back({ms, (float)b.n_decode, (float)(b.T - b.n_decode), host_ms});
        }
    }
    for (int l = 0; l < e->prof_used; ++l) {
        float ms = 0.f;
        if (cudaEventElapsedTime(&ms, e->prof_ev[2 * l], e->prof_ev[2 * l + 1]) == cudaSuccess) {
            e->st.attn_decode_ms += ms;
            e->st.attn_decode_launches++;
            e->st.attn_decode_kv_tokens += b.n_decode > 0 ? e->prof_ctx_sum : 0;
        }
    }

    // ---- bookkeeping: the K/V of every token we just ran is now cached
    int produced = 0;
    auto emit = [&](Request* r) {
        const int tok = h[m.o_sampled + r->logits_row];
        r->last_token = tok;
        r->n_generated++;
        produced++;
        bool is_eos = false;
        if (!r->ignore_eos)
            for (int k = 0; k < c.n_eos; ++k) is_eos |= tok == c.eos_ids[k];
        const int ctx_after = (int)r->sess->tokens.size() + 1;      // if this token were fed back
        if (is_eos) finish_request(e, r, 1, tok);
        else if (r->n_generated >= r->max_new || ctx_after >= c.max_context) finish_request(e, r, 2, tok);
        else e->out.push_back({r->id, tok, 0});
    };
    std::vector<Session*> to_close;
    auto settle = [&](Request* r) -> bool {      // true: the request was cancelled / its session closed mid-step
        if (r->sess && r->sess->close_pending && std::find(to_close.begin(), to_close.end(), r->sess) == to_close.end())
            to_close.push_back(r->sess);
        if (!r->cancel_pending) return false;
        r->cancel_pending = false;
        finish_request(e, r, 3, -1);
        return true;
    };
    for (Request* r : dec) {
        r->sess->tokens.push_back(r->last_token);
        if (!settle(r)) emit(r);
    }
    for (int j = 0; j < b.n_prefill_req; ++j) {
        Request* r = pre[j];
        Session* s = r->sess;
        s->tokens.insert(s->tokens.end(), r->prompt.begin() + r->prefill_pos,
                         r->prompt.begin() + r->prefill_pos + pre_len[j]);
        r->prefill_pos += pre_len[j];
        e->st.prefill_tokens += pre_len[j];
        if (settle(r)) continue;
        if (r->logits_row >= 0) {
            r->state = REQ_DECODE;
            emit(r);
        }
    }
    for (Session* sx : to_close) {
        for (auto& kv : e->requests)
            if (kv.second->sess == sx) kv.second->sess = nullptr;
        drop_session(e, sx);
    }
    e->st.steps++;
    if (b.n_decode > 0) e->st.decode_steps++;
    e->st.decode_tokens += b.n_decode;
    e->st.last_batch = b.n_req;
    e->st.last_prefill_tokens = b.T - b.n_decode;
    e->st.last_step_ms =
        std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t_begin).count();
    return produced;
}

// ------------------------------------------------------------------ construction
static int engine_init(Engine* e) {
    const ft_engine_config& c = e->cfg;
    FT_REQUIRE(c.head_dim == 64 || c.head_dim == 128, "engine: head_dim %d not supported", c.head_dim);
    FT_REQUIRE(c.n_q_heads % c.n_kv_heads == 0 && c.n_q_heads / c.n_kv_heads <= 8,
               "engine: GQA group must divide and be <= 8");
    FT_REQUIRE(c.hidden % 128 == 0 && c.ffn % 64 == 0 && c.vocab % 128 == 0,
               "engine: hidden/vocab must be multiples of 128 and ffn of 64");
    FT_REQUIRE(c.max_batch >= 1 && c.max_batched_tokens >= c.max_batch, "engine: bad batch limits");
    FT_REQUIRE(c.max_context >= 64 && c.num_pages >= 1, "engine: bad context/page limits");
    FT_CHECK_CUDA(cudaGetDevice(&e->dev));
    FT_CHECK_CUDA(cudaDeviceGetAttribute(&e->sms, cudaDevAttrMultiProcessorCount, e->dev));
    int cc_major = 0;
    FT_CHECK_CUDA(cudaDeviceGetAttribute(&cc_major, cudaDevAttrComputeCapabilityMajor, e->dev));
    FT_REQUIRE(cc_major == 10, "engine: needs an sm_100 GPU (found compute capability %d.x)", cc_major);
    FT_CHECK_CUDA(cudaStreamCreateWithFlags(&e->stream, cudaStreamNonBlocking));

    e->q_dim = c.n_q_heads * c.head_dim;
    e->kv_dim = c.n_kv_heads * c.head_dim;
    e->qkv_dim = e->q_dim + 2 * e->kv_dim;
    FT_REQUIRE(e->qkv_dim % 128 == 0 && e->q_dim % 64 == 0, "engine: projection dims must tile");
    e->group = c.n_q_heads / c.n_kv_heads;
    e->max_pages = ceil_div(c.max_context, kKvTile);
    e->t_max = c.max_batched_tokens;
    e->r_max = c.max_batch;
    e->logit_rows = std::max(c.max_batch, c.debug_logit_rows);
    e->pdl = c.use_pdl != 0;
    e->prefill_planes = c.prefill_planes >= 1 && c.prefill_planes <= 3 ? c.prefill_planes : kDefaultPrefillPlanes;
    if (const char* ov = getenv("FT_PREFILL_PLANES"))       // A/B switch for experiments (scripts/)
        if (atoi(ov) >= 1 && atoi(ov) <= 3) e->prefill_planes = atoi(ov);
    e->decode_prefill_cap = c.decode_prefill_cap == 0 ? kDefaultDecodePrefillCap : c.decode_prefill_cap < 0 ? 0 : c.decode_prefill_cap;
    e->tp = c.tp_size > 1 ? c.tp_size : 1;
    FT_REQUIRE(e->tp <= 2 && c.tp_rank >= 0 && c.tp_rank < e->tp, "engine: tensor parallelism is built for 2 ranks");
    FT_REQUIRE(c.vocab % (128 * e->tp) == 0, "engine: vocab must split into 128-row tiles per rank");
    e->vocab_local = c.vocab / e->tp;
    e->kv_layer_bytes = (size_t)c.num_pages * c.n_kv_heads * 2 * kKvTile * c.head_dim * 2;

    const size_t T = e->t_max;
    FT_CHECK_CUDA(cudaMalloc(&e->kv, e->kv_layer_bytes * c.layers));
    FT_CHECK_CUDA(cudaMemsetAsync(e->kv, 0, e->kv_layer_bytes * c.layers, e->stream));
    FT_CHECK_CUDA(cudaMalloc(&e->resid, T * c.hidden * 4));
    const size_t rows_last = std::max<size_t>(e->logit_rows, 256);
    FT_CHECK_CUDA(cudaMalloc(&e->q, T * e->q_dim * 2));
    FT_CHECK_CUDA(cudaMalloc(&e->logits, (size_t)e->logit_rows * c.vocab * 4));
    if (e->tp > 1) {
        // plain cudaMalloc: these three are exported to the peer process through cudaIpc handles
        FT_CHECK_CUDA(cudaMalloc(&e->tp_xchg, 2 * T * c.hidden * 4));
        FT_CHECK_CUDA(cudaMalloc(&e->tp_flags, kTpFlags * sizeof(unsigned)));
        FT_CHECK_CUDA(cudaMemsetAsync(e->tp_flags, 0, kTpFlags * sizeof(unsigned), e->stream));
    }
    auto alloc_zero = [&](void** p, size_t bytes) -> int {
        FT_CHECK_CUDA(cudaMalloc(p, bytes));
        FT_CHECK_CUDA(cudaMemsetAsync(*p, 0, bytes, e->stream));
        return 0;
    };
    FT_TRY(alloc_zero(&e->xn, T * c.hidden * 2));
    FT_TRY(alloc_zero(&e->xn_lo, T * c.hidden * 2));
    FT_TRY(alloc_zero(&e->attn_out, T * e->q_dim * 2));
    FT_TRY(alloc_zero(&e->attn_out_lo, T * e->q_dim * 2));
    FT_TRY(alloc_zero(&e->act, T * c.ffn * 2));
    FT_TRY(alloc_zero(&e->act_lo, T * c.ffn * 2));
    FT_TRY(alloc_zero(&e->xn_last, rows_last * c.hidden * 2));
    FT_TRY(alloc_zero(&e->xn_last_lo, rows_last * c.hidden * 2));
    // per-row-tile sums of squares of the residual rows (RMSNorm statistics, see forward())
    e->ssq_parts = c.hidden / 128;
    e->rows_last = (int)rows_last;
    {
        void* p = nullptr;
        FT_TRY(alloc_zero(&p, (size_t)e->ssq_parts * T * 4));
        e->ssq = static_cast<float*>(p);
        FT_TRY(alloc_zero(&p, (size_t)e->ssq_parts * rows_last * 4));
        e->ssq_last = static_cast<float*>(p);
        // a split-K grid never exceeds one CTA per SM, each parking at most a 128 x 128 fp32 tile
        FT_TRY(alloc_zero(&p, (size_t)e->sms * 128 * 128 * 4));
        e->gemm_scratch = static_cast<float*>(p);
        FT_TRY(alloc_zero(&p, (size_t)2 * e->sms * 4));
        e->gemm_sync = static_cast<int*>(p);
    }

    // decode attention: persistent grid = SMs x resident CTAs
    {
        int per_sm = attn_decode_ctas_per_sm(c.head_dim);
        FT_REQUIRE(per_sm >= 1, "engine: decode attention kernel does not fit on an SM (%s)", last_error());
        e->attn_ctas = e->sms * per_sm;
    }
    FT_CHECK_CUDA(cudaMalloc(&e->attn_scratch,
                             attn_decode_scratch_floats(e->r_max, c.n_kv_heads, e->group, c.head_dim,
                                                        e->attn_ctas) * 4));

    // step metadata block
    StepMeta& m = e->meta;
    size_t o = 0;
    auto take = [&](size_t n) { size_t r = o; o += align4(n); return r; };
    const size_t R = e->r_max, LR = rows_last;
    const size_t max_tiles = T / 64 + R + 1;
    m.o_ctx = take(R); m.o_qstart = take(R + 1); m.o_rslot = take(R); m.o_rpos0 = take(R);
    m.o_treq = take(max_tiles); m.o_tq0 = take(max_tiles);
    m.o_lrow = take(LR); m.o_temp = take(LR); m.o_topk = take(LR); m.o_topp = take(LR);
    m.o_seed = take(2 * LR);
    m.o_pt = take(0);
    o += align4(R * e->max_pages) + 3 * align4(T);      // page table + tokens | positions | rows
    m.o_tprefix = take(R + 1);
    m.o_sampled = take(LR);
    m.ints = o;
    layout_step(m, (int)R, e->max_pages, (int)T);
    FT_CHECK_CUDA(cudaMallocHost(&m.h, o * sizeof(int)));
    FT_CHECK_CUDA(cudaMalloc(&m.d, o * sizeof(int)));
    memset(m.h, 0, o * sizeof(int));
    FT_CHECK_CUDA(cudaMemsetAsync(m.d, 0, o * sizeof(int), e->stream));

    // tensor maps
    const size_t L = c.layers;
    e->m_wqkv.resize(L); e->m_wo.resize(L); e->m_wgu.resize(L); e->m_wd.resize(L);
    auto at = [](const void* p, size_t elems) { return (const void*)((const uint8_t*)p + elems * 2); };
    for (size_t l = 0; l < L; ++l) {
        FT_TRY(make_tmap_bf16_k64(&e->m_wqkv[l], at(e->w.wqkv, l * e->qkv_dim * c.hidden), e->qkv_dim, c.hidden, 128));
        FT_TRY(make_tmap_bf16_k64(&e->m_wo[l], at(e->w.wo, l * (size_t)c.hidden * e->q_dim), c.hidden, e->q_dim, 128));
        FT_TRY(make_tmap_bf16_k64(&e->m_wgu[l], at(e->w.wgu, l * 2 * (size_t)c.ffn * c.hidden), 2 * c.ffn, c.hidden, 128));
        FT_TRY(make_tmap_bf16_k64(&e->m_wd[l], at(e->w.wd, l * (size_t)c.hidden * c.ffn), c.hidden, c.ffn, 128));
    }
    FT_TRY(make_tmap_bf16_k64(&e->m_lm, e->w.lm_head, e->vocab_local, c.hidden, 128));
    for (int i = 0; i < 4; ++i) {
        // activation maps span the whole buffer; rows past the step's T hold stale (finite) data
        // whose output columns the GEMM epilogue masks
        FT_TRY(make_tmap_bf16_k64(&e->m_xn[i], e->xn, (int)T, c.hidden, kBns[i]));
        FT_TRY(make_tmap_bf16_k64(&e->m_xn_lo[i], e->xn_lo, (int)T, c.hidden, kBns[i]));
        FT_TRY(make_tmap_bf16_k64(&e->m_attn[i], e->attn_out, (int)T, e->q_dim, kBns[i]));
        FT_TRY(make_tmap_bf16_k64(&e->m_attn_lo[i], e->attn_out_lo, (int)T, e->q_dim, kBns[i]));
        FT_TRY(make_tmap_bf16_k64(&e->m_act[i], e->act, (int)T, c.ffn, kBns[i]));
        FT_TRY(make_tmap_bf16_k64(&e->m_act_lo[i], e->act_lo, (int)T, c.ffn, kBns[i]));
        FT_TRY(make_tmap_bf16_k64(&e->m_last[i], e->xn_last, (int)rows_last, c.hidden, kBns[i]));
        FT_TRY(make_tmap_bf16_k64(&e->m_last_lo[i], e->xn_last_lo, (int)rows_last, c.hidden, kBns[i]));
    }
    if (e->prefill_planes == 3) {
        // fp16 copies of the projection weights (bf16 -> fp16 is exact for every weight above fp16's subnormal range) and
        // fp16 staging planes of their inputs
        auto to16 = [&](void** dst, const void* src, size_t elems) -> int {
            FT_CHECK_CUDA(cudaMalloc(dst, elems * 2));
            return launch_planes_to_f16(src, nullptr, *dst, elems, false, e->stream);
        };
        FT_TRY(to16(&e->w16_qkv, e->w.wqkv, L * e->qkv_dim * c.hidden));
        FT_TRY(to16(&e->w16_o, e->w.wo, L * (size_t)c.hidden * e->q_dim));
        FT_TRY(to16(&e->w16_gu, e->w.wgu, L * 2 * (size_t)c.ffn * c.hidden));
        FT_TRY(to16(&e->w16_d, e->w.wd, L * (size_t)c.hidden * c.ffn));
        FT_TRY(alloc_zero(&e->x16_xn, T * c.hidden * 2));
        FT_TRY(alloc_zero(&e->x16_attn, T * e->q_dim * 2));
        FT_TRY(alloc_zero(&e->x16_act, T * c.ffn * 2));
        e->m16_wqkv.resize(L); e->m16_wo.resize(L); e->m16_wgu.resize(L); e->m16_wd.resize(L);
        for (size_t l = 0; l < L; ++l) {
            FT_TRY(make_tmap_bf16_k64(&e->m16_wqkv[l], at(e->w16_qkv, l * e->qkv_dim * c.hidden), e->qkv_dim, c.hidden, 128));
            FT_TRY(make_tmap_bf16_k64(&e->m16_wo[l], at(e->w16_o, l * (size_t)c.hidden * e->q_dim), c.hidden, e->q_dim, 128));
            FT_TRY(make_tmap_bf16_k64(&e->m16_wgu[l], at(e->w16_gu, l * 2 * (size_t)c.ffn * c.hidden), 2 * c.ffn, c.hidden, 128));
            FT_TRY(make_tmap_bf16_k64(&e->m16_wd[l], at(e->w16_d, l * (size_t)c.hidden * c.ffn), c.hidden, c.ffn, 128));
        }
        const int boxes[2] = {128, 64};
        for (int i = 0; i < 2; ++i) {         // 16-bit tiles: the tensor map only moves bits, the MMA descriptor names the type
            FT_TRY(make_tmap_bf16_k64(&e->m16_xn[i], e->x16_xn, (int)T, c.hidden, boxes[i]));
            FT_TRY(make_tmap_bf16_k64(&e->m16_attn[i], e->x16_attn, (int)T, e->q_dim, boxes[i]));
            FT_TRY(make_tmap_bf16_k64(&e->m16_act[i], e->x16_act, (int)T, c.ffn, boxes[i]));
        }
    }
    e->prof_ev.resize(2 * L);
    for (auto& ev : e->prof_ev) FT_CHECK_CUDA(cudaEventCreate(&ev));
    FT_CHECK_CUDA(cudaEventCreate(&e->t0));
    FT_CHECK_CUDA(cudaEventCreate(&e->t1));
    FT_CHECK_CUDA(cudaEventCreate(&e->s0));
    FT_CHECK_CUDA(cudaEventCreate(&e->s1));
    e->free_pages.reserve(c.num_pages);
    for (int p = c.num_pages - 1; p >= 0; --p) e->free_pages.push_back(p);
    e->st.total_pages = c.num_pages;
    FT_CHECK_CUDA(cudaStreamSynchronize(e->stream));
    return 0;
}

static void engine_free(Engine* e) {
    if (e->stream) cudaStreamSynchronize(e->stream);
    for (auto& kv : e->sessions) delete kv.second;
    for (auto& kv : e->requests) delete kv.second;
    cudaFree(e->kv); cudaFree(e->resid); cudaFree(e->ssq); cudaFree(e->ssq_last); cudaFree(e->gemm_scratch); cudaFree(e->gemm_sync); cudaFree(e->xn); cudaFree(e->q);
    cudaFree(e->attn_out); cudaFree(e->act); cudaFree(e->xn_last); cudaFree(e->logits);
    cudaFree(e->tp_xchg); cudaFree(e->tp_flags);
    cudaFree(e->xn_lo); cudaFree(e->attn_out_lo); cudaFree(e->act_lo); cudaFree(e->xn_last_lo);
    cudaFree(e->attn_scratch); cudaFree(e->meta.d);
    cudaFree(e->w16_qkv); cudaFree(e->w16_o); cudaFree(e->w16_gu); cudaFree(e->w16_d);
    cudaFree(e->x16_xn); cudaFree(e->x16_attn); cudaFree(e->x16_act);
    if (e->meta.h) cudaFreeHost(e->meta.h);
    for (auto ev : e->prof_ev) if (ev) cudaEventDestroy(ev);
    if (e->t0) cudaEventDestroy(e->t0);
    if (e->t1) cudaEventDestroy(e->t1);
    if (e->s0) cudaEventDestroy(e->s0);
    if (e->s1) cudaEventDestroy(e->s1);
    if (e->stream) cudaStreamDestroy(e->stream);
    delete e;
}

}  // namespace ft

// ------------------------------------------------------------------ C ABI (engine level)
using namespace ft;

extern "C" ft_engine* ft_engine_create(const ft_engine_config* cfg, const ft_weights* w) {
    if (!cfg || !w) { set_last_error("ft_engine_create: null argument"); return nullptr; }
    Engine* e = new (std::nothrow) Engine();
    if (!e) { set_last_error("ft_engine_create: out of host memory"); return nullptr; }
    e->cfg = *cfg;
    e->w = *w;
    if (e->cfg.max_sessions < e->cfg.max_batch) e->cfg.max_sessions = e->cfg.max_batch * 4;
    if (engine_init(e) != 0) { engine_free(e); return nullptr; }
    return reinterpret_cast<ft_engine*>(e);
}

extern "C" void ft_engine_destroy(ft_engine* h) {
    if (!h) return;
    Engine* e = reinterpret_cast<Engine*>(h);
    ft_engine_stop(h);
    {   // a thread still blocked in ft_engine_wait (a daemon thread of the host at interpreter exit) leaves first
        std::unique_lock<std::mutex> lk(e->mu);
        e->closing = true;
        e->cv_out.notify_all();
        e->cv_out.wait_for(lk, std::chrono::seconds(2), [&] { return e->waiters == 0; });
    }
    DeviceGuard dg(e->dev);
    engine_free(e);
}

extern "C" long long ft_engine_submit(ft_engine* h, unsigned long long key, const int* tokens, int n,
                                      float temperature, int top_k, float top_p, int max_new,
                                      unsigned long long seed, int ignore_eos) {
    Engine* e = reinterpret_cast<Engine*>(h);
    FT_REQUIRE(e && tokens && n >= 1, "ft_engine_submit: bad arguments");
    FT_REQUIRE(max_new >= 1, "ft_engine_submit: max_new_tokens must be >= 1");
    FT_REQUIRE(n + 1 <= e->cfg.max_context, "ft_engine_submit: prompt of %d tokens does not fit the %d-token context window", n, e->cfg.max_context);
    for (int i = 0; i < n; ++i)
        FT_REQUIRE(tokens[i] >= 0 && tokens[i] < e->cfg.vocab, "ft_engine_submit: token %d out of range at %d", tokens[i], i);
    std::lock_guard<std::mutex> lk(e->mu);
    Session* s = nullptr;
    auto it = e->sessions.find(key);
    if (it != e->sessions.end()) {
        s = it->second;
        FT_REQUIRE(!s->busy, "ft_engine_submit: session %llu already has a generation in flight", key);
    } else {
        if ((int)e->sessions.size() >= e->cfg.max_sessions) {      // drop the stalest idle session
            Session* victim = nullptr;
            for (auto& kv : e->sessions)
                if (!kv.second->busy && (!victim || kv.second->last_used < victim->last_used)) victim = kv.second;
            FT_REQUIRE(victim, "ft_engine_submit: all %d sessions are busy", e->cfg.max_sessions);
            drop_session(e, victim);
        }
        s = new Session();
        s->key = key;
        e->sessions[key] = s;
    }
    // longest common prefix with the cached conversation; at least one token must run
    int lcp = 0;
    const int lim = std::min((int)s->tokens.size(), n - 1);
    while (lcp < lim && s->tokens[lcp] == tokens[lcp]) ++lcp;
    release_pages_beyond(e, s, lcp);
    e->st.prefix_hit_tokens += lcp;
    s->busy = true;
    s->last_used = ++e->clock;
    Request* r = new Request();
    r->id = e->next_req++;
    r->sess = s;
    r->prompt.assign(tokens, tokens + n);
    r->prefill_pos = lcp;
    r->temperature = temperature;
    r->top_k = top_k;
    r->top_p = top_p;
    r->max_new = max_new;
    r->seed = seed;
    r->ignore_eos = ignore_eos;
    e->requests[r->id] = r;
    e->waiting.push_back(r);
    e->cv_work.notify_one();
    return r->id;
}

extern "C" int ft_engine_tp_export(ft_engine* h, void** xchg, void** flags, void** logits) {
    Engine* e = reinterpret_cast<Engine*>(h);
    FT_REQUIRE(e && xchg && flags && logits, "ft_engine_tp_export: bad arguments");
    FT_REQUIRE(e->tp > 1, "ft_engine_tp_export: the engine was created without tensor parallelism");
    DeviceGuard dg(e->dev);
    FT_CHECK_CUDA(cudaStreamSynchronize(e->stream));       // the arrival words are zeroed
    *xchg = e->tp_xchg;
    *flags = e->tp_flags;
    *logits = e->logits;
    return 0;
}

extern "C" int ft_engine_tp_attach(ft_engine* h, void* peer_xchg, void* peer_flags, void* peer_logits) {
    Engine* e = reinterpret_cast<Engine*>(h);
    FT_REQUIRE(e && peer_xchg && peer_flags && peer_logits, "ft_engine_tp_attach: bad arguments");
    FT_REQUIRE(e->tp > 1, "ft_engine_tp_attach: the engine was created without tensor parallelism");
    std::lock_guard<std::mutex> lk(e->mu);
    e->peer_xchg = static_cast<float*>(peer_xchg);
    e->peer_flags = static_cast<unsigned*>(peer_flags);
    e->peer_logits = static_cast<float*>(peer_logits);
    return 0;
}

extern "C" int ft_ipc_export(void* device_ptr, unsigned char handle_out[64]) {
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "cudaIpcMemHandle_t is 64 bytes");
    FT_REQUIRE(device_ptr && handle_out, "ft_ipc_export: bad arguments");
    cudaIpcMemHandle_t hd;
    FT_CHECK_CUDA(cudaIpcGetMemHandle(&hd, device_ptr));
    memcpy(handle_out, &hd, 64);
    return 0;
}

extern "C" int ft_ipc_open(const unsigned char handle[64], void** device_ptr_out) {
    FT_REQUIRE(handle && device_ptr_out, "ft_ipc_open: bad arguments");
    cudaIpcMemHandle_t hd;
    memcpy(&hd, handle, 64);
    FT_CHECK_CUDA(cudaIpcOpenMemHandle(device_ptr_out, hd, cudaIpcMemLazyEnablePeerAccess));
    return 0;
}

extern "C" int ft_ipc_close(void* device_ptr) {
    FT_REQUIRE(device_ptr, "ft_ipc_close: null pointer");
    FT_CHECK_CUDA(cudaIpcCloseMemHandle(device_ptr));
    return 0;
}

extern "C" int ft_engine_step(ft_engine* h) {
    Engine* e = reinterpret_cast<Engine*>(h);
    FT_REQUIRE(e, "ft_engine_step: null engine");
    FT_REQUIRE(e->tp == 1 || e->peer_xchg, "ft_engine_step: tensor-parallel engine has no peer attached");
    std::lock_guard<std::mutex> sk(e->step_mu);
    std::lock_guard<std::mutex> dk(g_device_step_mu[e->dev % kMaxDevices]);
    DeviceGuard dg(e->dev);
    std::unique_lock<std::mutex> lk(e->mu);
    FT_REQUIRE(!e->loop_on, "ft_engine_step: the engine runs its own stepping thread (ft_engine_start); stop it first");
    return engine_step(e, lk);
}

// ---- the engine's own stepping thread
static bool runnable(const Engine* e) {           // under e->mu
    for (const Request* r : e->waiting)
        if (r->state != REQ_DONE) return true;
    for (const Request* r : e->active)
        if (r->state != REQ_DONE) return true;
    return false;
}

static void loop_main(Engine* e) {
    // a fresh thread's current device is 0: make the engine's GPU current for the thread's whole life, so that the
    // per-step DeviceGuard has nothing to restore (restoring "device 0" from a rank that lives on another GPU creates a
    // context there — 0.4 s the first time)
    cudaSetDevice(e->dev);
    for (;;) {
        {
            std::unique_lock<std::mutex> lk(e->mu);
            e->cv_work.wait_for(lk, std::chrono::milliseconds(20), [&] { return e->loop_stop || runnable(e); });
            if (e->loop_stop) break;
            if (!runnable(e)) continue;
        }
        int rc;
        bool stepped;
        {
            std::lock_guard<std::mutex> sk(e->step_mu);
            std::lock_guard<std::mutex> dk(g_device_step_mu[e->dev % kMaxDevices]);
            DeviceGuard dg(e->dev);
            std::unique_lock<std::mutex> lk(e->mu);
            const long long before = e->st.steps;
            rc = engine_step(e, lk);
            stepped = e->st.steps != before;
            if (rc < 0) {
                e->loop_error = rc;
                e->loop_error_msg = ft_last_error();
            }
        }
        e->cv_out.notify_all();
        if (rc < 0) break;
        if (!stepped) std::this_thread::sleep_for(std::chrono::microseconds(200));    // waiting on pages / admission
    }
    {
        std::lock_guard<std::mutex> lk(e->mu);
        e->loop_on = false;
    }
    e->cv_out.notify_all();
}

extern "C" int ft_engine_start(ft_engine* h) {
    Engine* e = reinterpret_cast<Engine*>(h);
    FT_REQUIRE(e, "ft_engine_start: null engine");
    FT_REQUIRE(e->tp == 1, "ft_engine_start: tensor-parallel ranks are stepped in lock step by their host (tp.py)");
    {
        std::lock_guard<std::mutex> lk(e->mu);
        if (e->loop_on) return 0;
        FT_REQUIRE(e->loop_error == 0, "ft_engine_start: the stepping thread failed earlier: %s", e->loop_error_msg.c_str());
    }
    if (e->loop_thread.joinable()) e->loop_thread.join();       // a loop that ended by itself
    {
        std::lock_guard<std::mutex> lk(e->mu);
        e->loop_stop = false;
        e->loop_on = true;
    }
    e->loop_thread = std::thread(loop_main, e);
    return 0;
}

extern "C" int ft_engine_stop(ft_engine* h) {
    Engine* e = reinterpret_cast<Engine*>(h);
    FT_REQUIRE(e, "ft_engine_stop: null engine");
    {
        std::lock_guard<std::mutex> lk(e->mu);
        e->loop_stop = true;
    }
    e->cv_work.notify_all();
    if (e->loop_thread.joinable()) e->loop_thread.join();
    return 0;
}

extern "C" int ft_engine_wait(ft_engine* h, int timeout_ms) {
    Engine* e = reinterpret_cast<Engine*>(h);
    FT_REQUIRE(e, "ft_engine_wait: null engine");
    std::unique_lock<std::mutex> lk(e->mu);
    FT_REQUIRE(!e->closing, "ft_engine_wait: the engine is being destroyed");
    ++e->waiters;
    e->cv_out.wait_for(lk, std::chrono::milliseconds(timeout_ms < 0 ? 0 : timeout_ms),
                       [&] { return !e->out.empty() || e->loop_error != 0 || !e->loop_on || e->closing; });
    --e->waiters;
    if (e->closing) {
        e->cv_out.notify_all();              // ft_engine_destroy waits for the last waiter to leave
        set_last_error("ft_engine_wait: the engine is being destroyed");
        return -1;
    }
    if (e->loop_error != 0 && e->out.empty()) {
        set_last_error("ft_engine_wait: the stepping thread stopped: %s", e->loop_error_msg.c_str());
        return -1;
    }
    return (int)e->out.size();
}

extern "C" int ft_engine_drain(ft_engine* h, long long* req_ids, int* tokens, int* finish, int cap) {
    Engine* e = reinterpret_cast<Engine*>(h);
    FT_REQUIRE(e && req_ids && tokens && finish, "ft_engine_drain: bad arguments");
    std::lock_guard<std::mutex> lk(e->mu);
    int n = 0;
    while (n < cap && !e->out.empty()) {
        const OutTok& o = e->out.front();
        req_ids[n] = o.req; tokens[n] = o.token; finish[n] = o.finish;
        if (o.finish) {                 // the request is over: forget it
            auto it = e->requests.find(o.req);
            if (it != e->requests.end() && it->second->state == REQ_DONE) {
                Request* r = it->second;
                e->active.erase(std::remove(e->active.begin(), e->active.end(), r), e->active.end());
                e->waiting.erase(std::remove(e->waiting.begin(), e->waiting.end(), r), e->waiting.end());
                delete r;
                e->requests.erase(it);
            }
        }
        e->out.pop_front();
        ++n;
    }
    return n;
}

extern "C" int ft_engine_cancel(ft_engine* h, long long req) {
    Engine* e = reinterpret_cast<Engine*>(h);
    FT_REQUIRE(e, "ft_engine_cancel: null engine");
    std::lock_guard<std::mutex> lk(e->mu);
    auto it = e->requests.find(req);
    if (it == e->requests.end() || it->second->state == REQ_DONE) return 1;    // nothing to cancel
    if (it->second->in_flight) {         // inside the running step: finished as "cancelled" when it returns
        it->second->cancel_pending = true;
        return 0;
    }
    finish_request(e, it->second, 3, -1);
    e->cv_out.notify_all();
    return 0;
}

extern "C" int ft_engine_session_close(ft_engine* h, unsigned long long key) {
    Engine* e = reinterpret_cast<Engine*>(h);
    FT_REQUIRE(e, "ft_engine_session_close: null engine");
    std::lock_guard<std::mutex> lk(e->mu);
    auto it = e->sessions.find(key);
    if (it == e->sessions.end()) return 1;
    Session* s = it->second;
    if (s->busy)
        for (auto& kv : e->requests)
            if (kv.second->sess == s && kv.second->state != REQ_DONE) {
                if (kv.second->in_flight) {      // its pages are being read and written right now: close after the step
                    kv.second->cancel_pending = true;
                    s->close_pending = true;
                    return 0;
                }
                finish_request(e, kv.second, 3, -1);
            }
    for (auto& kv : e->requests)
        if (kv.second->sess == s) kv.second->sess = nullptr;
    drop_session(e, s);
    return 0;
}

extern "C" int ft_engine_stats_get(ft_engine* h, ft_engine_stats* out) {
    Engine* e = reinterpret_cast<Engine*>(h);
    FT_REQUIRE(e && out, "ft_engine_stats_get: bad arguments");
    std::lock_guard<std::mutex> lk(e->mu);
    e->st.active_requests = 0;
    for (Request* r : e->active) e->st.active_requests += r->state != REQ_DONE;
    e->st.waiting_requests = (int)e->waiting.size();
    e->st.live_sessions = (int)e->sessions.size();
    e->st.free_pages = (int)e->free_pages.size();
    *out = e->st;
    return 0;
}

extern "C" int ft_engine_set_decode_prefill_cap(ft_engine* h, int cap) {
    Engine* e = reinterpret_cast<Engine*>(h);
    FT_REQUIRE(e, "ft_engine_set_decode_prefill_cap: null engine");
    std::lock_guard<std::mutex> lk(e->mu);
    e->decode_prefill_cap = cap == 0 ? kDefaultDecodePrefillCap : cap < 0 ? 0 : cap;
    return 0;
}

extern "C" int ft_engine_set_profile(ft_engine* h, int on) {
    Engine* e = reinterpret_cast<Engine*>(h);
    FT_REQUIRE(e, "ft_engine_set_profile: null engine");
    std::lock_guard<std::mutex> lk(e->mu);
    e->profile = on != 0;
    return 0;
}

extern "C" int ft_engine_timer(ft_engine* h, int stop, float* ms_out) {
    Engine* e = reinterpret_cast<Engine*>(h);
    FT_REQUIRE(e, "ft_engine_timer: null engine");
    std::lock_guard<std::mutex> sk(e->step_mu);
    std::lock_guard<std::mutex> lk(e->mu);
    DeviceGuard dg(e->dev);
    if (!stop) {
        FT_CHECK_CUDA(cudaEventRecord(e->t0, e->stream));
        return 0;
    }
    FT_REQUIRE(ms_out, "ft_engine_timer: null output");
    FT_CHECK_CUDA(cudaEventRecord(e->t1, e->stream));
    FT_CHECK_CUDA(cudaEventSynchronize(e->t1));
    FT_CHECK_CUDA(cudaEventElapsedTime(ms_out, e->t0, e->t1));
    return 0;
}

extern "C" int ft_engine_step_log(ft_engine* h, float* out, int cap_rows) {
    Engine* e = reinterpret_cast<Engine*>(h);
    FT_REQUIRE(e && out && cap_rows >= 0, "ft_engine_step_log: bad arguments");
    std::lock_guard<std::mutex> lk(e->mu);
    while ((int)e->step_log.size() > cap_rows) e->step_log.pop_front();
    int n = 0;
    for (const auto& r : e->step_log) {
        out[4 * n + 0] = r.gpu_ms; out[4 * n + 1] = r.n_decode; out[4 * n + 2] = r.n_prefill; out[4 * n + 3] = r.host_ms;
        ++n;
    }
    e->step_log.clear();
    return n;
}

extern "C" int ft_engine_get_logits(ft_engine* h, long long req, float* host_out) {
    Engine* e = reinterpret_cast<Engine*>(h);
    FT_REQUIRE(e && host_out, "ft_engine_get_logits: bad arguments");
    std::lock_guard<std::mutex> sk(e->step_mu);       // the logits buffer belongs to the step while one runs
    std::lock_guard<std::mutex> lk(e->mu);
    DeviceGuard dg(e->dev);
    auto it = e->requests.find(req);
    FT_REQUIRE(it != e->requests.end() && it->second->logits_row >= 0,
               "ft_engine_get_logits: request %lld has no logits from the last step", req);
    FT_CHECK_CUDA(cudaMemcpy(host_out, e->logits + (size_t)it->second->logits_row * e->cfg.vocab,
                             (size_t)e->cfg.vocab * 4, cudaMemcpyDeviceToHost));
    return 0;
}

extern "C" int ft_debug_prefill_logits(ft_engine* h, const int* tokens, int n, int chunk, float* host_out) {
    Engine* e = reinterpret_cast<Engine*>(h);
    FT_REQUIRE(e && tokens && host_out && n >= 1, "ft_debug_prefill_logits: bad arguments");
    std::lock_guard<std::mutex> sk(e->step_mu);
    std::lock_guard<std::mutex> dk(g_device_step_mu[e->dev % kMaxDevices]);
    std::lock_guard<std::mutex> lk(e->mu);
    DeviceGuard dg(e->dev);
    FT_REQUIRE(e->active.empty() && e->waiting.empty(), "ft_debug_prefill_logits: engine must be idle");
    FT_REQUIRE(chunk >= 1 && chunk <= e->logit_rows && chunk <= e->t_max,
               "ft_debug_prefill_logits: chunk %d exceeds debug_logit_rows %d / max_batched_tokens %d", chunk,
               e->logit_rows, e->t_max);
    FT_REQUIRE(n + 1 <= e->cfg.max_context, "ft_debug_prefill_logits: too long");
    Session tmp;
    if (!ensure_pages(e, &tmp, n)) { set_last_error("ft_debug_prefill_logits: out of KV pages"); return -3; }
    StepMeta& m = e->meta;
    int* hh = m.h;
    int rc = 0;
    for (int p0 = 0; p0 < n && rc == 0; p0 += chunk) {
        const int len = std::min(chunk, n - p0);
        Batch b;
        b.n_decode = 0; b.T = len; b.n_req = 1; b.n_prefill_req = 1; b.n_logit_rows = len;
        b.logit_rows_identity = true;
        layout_step(m, 1, e->max_pages, len);
        for (int k = 0; k < len; ++k) {
            hh[m.o_tokens + k] = tokens[p0 + k];
            hh[m.o_pos + k] = p0 + k;
            hh[m.o_rows + k] = 0;
            hh[m.o_lrow + k] = k;
        }
        hh[m.o_qstart] = 0; hh[m.o_qstart + 1] = len;
        hh[m.o_rslot] = 0; hh[m.o_rpos0] = p0;
        for (size_t p = 0; p < tmp.pages.size(); ++p) hh[m.o_pt + p] = tmp.pages[p];
        int tiles = 0;
        for (int q0 = 0; q0 < len; q0 += 64, ++tiles) { hh[m.o_treq + tiles] = 0; hh[m.o_tq0 + tiles] = q0; }
        b.n_tiles = tiles;
        rc = upload_meta(e);
        if (!rc) rc = forward(e, b);
        if (!rc && cudaMemcpyAsync(host_out + (size_t)p0 * e->cfg.vocab, e->logits, (size_t)len * e->cfg.vocab * 4,
                                   cudaMemcpyDeviceToHost, e->stream) != cudaSuccess) {
            set_last_error("ft_debug_prefill_logits: logits copy failed");
            rc = -1;
        }
        if (!rc && cudaStreamSynchronize(e->stream) != cudaSuccess) {
            set_last_error("ft_debug_prefill_logits: %s", cudaGetErrorString(cudaGetLastError()));
            rc = -1;
        }
    }
    release_pages_beyond(e, &tmp, 0);
    return rc;
}
