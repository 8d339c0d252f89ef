#!/usr/bin/env python
"""bench.py — streamed tok/s of the B200-native generation backend on BASELINE.json's metric.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]

Workload (BASELINE.json configs[1]): Llama-3.2-1B bf16, 64 concurrent sessions per GPU,
2k-token history, continuous batching.  A "step" is one decode step of the engine over the
64-session batch at context ~2048 (one pass of the hot path over one batch).

  value     K timed decode steps, CUDA events on the engine's stream, max over ranks;
            KV cache and weights resident in HBM (the engine still uploads its ~2 KB of
            per-step metadata and reads the 64 sampled token ids back every step).
  e2e       the same metric through the reference-facing plugin: 64 concurrent
            NativeHandler.generate_stream_async calls (the handler interface the reference's WebSocket
            server calls, websocket_server_vllm.py:528-539) with HOST buffers — `messages` in, one `str`
            chunk per token out.  Each session re-sends a ~2k-token history of which ~224 tokens are
            new (the rest is served from the session's KV), prefill rides along with decode, 32 tokens
            are generated; chunks / wall time, with TTFT and inter-token percentiles.
  e2e_engine  the same turn one level down, at the C ABI (ft_engine_submit, the engine's own stepping thread, ft_engine_wait /
            drain; host token lists).
  ws        the same turn one level up: the reference's UNCHANGED WebSocketLLMServer (baseline/_ref) in
            front of NativeHandler, 64 `websockets` clients in their own process, timed at the client.
  roofline  the dominant kernel (paged decode attention): algorithmic KV bytes per launch /
            its CUDA-event time, measured in a second pass over K more steps with events
            around every launch (events break PDL overlap, so they stay out of `value`).
  cpu_baseline / --impl reference   the oracle port of the same decode step on the host cores.

N > 1: independent replicas (one engine + 64 sessions per GPU, no collective on the data
path); torch.distributed only provides the barrier and the max-over-ranks reduction.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

MODEL = "llama-3.2-1b"
SESSIONS = 64
CTX = 2048
TURN_NEW = 224          # new history tokens per turn (SURVEY.md §8d cfg 2)
TURN_GEN = 32           # generated tokens per turn
METRIC = "Llama-3.2-1B streamed tok/s at 64 sessions/GPU; p50 TTFT & inter-token latency"
FALLBACK_HBM_GBS = 6650.0


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            with open(p) as fh:
                return float(json.load(fh)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return FALLBACK_HBM_GBS, "fallback (B200_PROFILING.md 6.65 TB/s)"


class ClockSampler:
    """SM clock and clock-event reasons sampled DURING the timed region.

    In-process NVML (nvidia_ml_py) on a thread, one sample every 5 ms; the NVML calls release the GIL
    and never touch the CUDA stream.  Round 1 spawned `nvidia-smi -lms 100` right before the timer:
    the tool's start-up (driver attach, ~0.3 s of host and PCIe activity) overlapped the whole 40 ms
    window of a 20-step run and only two samples ever landed.  start() is now called well before the
    timed region; mark()/unmark() bracket the region so only samples taken inside it are reported.
    Falls back to the nvidia-smi subprocess (also started early) when NVML cannot be loaded."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
    NAMES = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]

    def __init__(self, index: int, period_s: float = 0.005):
        self.index, self.period = index, period_s
        self.samples = []               # (t, sm_mhz, max_mhz, [reasons], power_w)
        self.proc, self.thread, self.run, self.how = None, None, False, None
        self.t_in, self.t_out = None, None

    def _physical_index(self):
        vis = os.environ.get("CUDA_VISIBLE_DEVICES")
        if vis:
            ids = [v.strip() for v in vis.split(",") if v.strip()]
            if self.index < len(ids) and ids[self.index].isdigit():
                return int(ids[self.index])
        return self.index

    def start(self):
        try:
            import pynvml as nv
            nv.nvmlInit()
            h = nv.nvmlDeviceGetHandleByIndex(self._physical_index())
            mx = float(nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM))
            bits = [(nv.nvmlClocksEventReasonHwSlowdown, "hw_slowdown"),
                    (nv.nvmlClocksEventReasonHwThermalSlowdown, "hw_thermal_slowdown"),
                    (nv.nvmlClocksEventReasonSwThermalSlowdown, "sw_thermal_slowdown"),
                    (nv.nvmlClocksEventReasonSwPowerCap, "sw_power_cap")]

            def loop():
                while self.run:
                    try:
                        sm = float(nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM))
                        r = int(nv.nvmlDeviceGetCurrentClocksEventReasons(h))
                        pw = nv.nvmlDeviceGetPowerUsage(h) / 1000.0
                        self.samples.append((time.perf_counter(), sm, mx, [n for b, n in bits if r & b], pw))
                    except Exception:
                        pass
                    time.sleep(self.period)
            self.run, self.how = True, "nvml in-process, 5 ms period"
            self.thread = threading.Thread(target=loop, daemon=True)
            self.thread.start()
            return
        except Exception:
            self.run = False
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-i",
                 str(self._physical_index()), "-lms", "20"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.how = "nvidia-smi -lms 20 subprocess"
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            f = [x.strip() for x in line.split(",")]
            if len(f) < 6:
                continue
            try:
                self.samples.append((time.perf_counter(), float(f[0]), float(f[1]),
                                     [n for n, v in zip(self.NAMES, f[2:6]) if v.lower().startswith("active")], None))
            except ValueError:
                continue

    def mark(self):
        self.t_in = time.perf_counter()

    def unmark(self):
        self.t_out = time.perf_counter()

    def stop(self):
        self.run = False
        if self.thread:
            self.thread.join(timeout=1.0)
        if self.proc:
            self.proc.terminate()

    def report(self, t_in=None, t_out=None):
        """Samples inside [t_in, t_out] (default: the mark()/unmark() window); the sampler may still be running."""
        lo = t_in if t_in is not None else self.t_in if self.t_in is not None else 0.0
        hi = t_out if t_out is not None else self.t_out if self.t_out is not None else float("inf")
        inside = [x for x in list(self.samples) if lo <= x[0] <= hi]
        if not inside and self.samples:                                 # window shorter than one period: nearest samples
            mid = 0.5 * (lo + (hi if hi != float("inf") else lo))
            inside = sorted(self.samples, key=lambda x: abs(x[0] - mid))[:2]
        sm = [x[1] for x in inside]
        reasons = sorted({r for x in inside for r in x[3]})
        pw = [x[4] for x in inside if x[4] is not None]
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_min_mhz": min(sm) if sm else None,
                "sm_max_mhz": max(x[2] for x in inside) if inside else None, "reasons": reasons,
                "samples": len(sm), "samples_total": len(self.samples), "power_w_max": max(pw) if pw else None,
                "how": self.how}


def pct(xs, q):
    xs = sorted(xs)
    return xs[min(len(xs) - 1, int(q * len(xs)))] if xs else None


# ----------------------------------------------------------------------------- reference arm
def run_reference(args, rank, world):
    """The reference's CPU path (Ollama on host cores) cannot be installed here; this arm times
    the oracle port of the same decode step on the host cores (oracle/cpu_baseline.py)."""
    if rank != 0:
        return
    from fasttalk_llm_microservice_b200.model_spec import get_spec
    from oracle import cpu_baseline as cb
    spec = get_spec(MODEL)
    budget_s = 150.0
    batch = SESSIONS
    wts = cb.make_weights(spec)
    # bounded sample: the largest session count whose projected run fits the budget
    for cand in (64, 16, 4, 1):
        _, ms, _, _ = cb.run(spec, cand, CTX, 1, 1, weights=wts)
        batch = cand
        if ms * 1e-3 * (args.steps + args.warmup) <= budget_s:
            break
    tps, ms, cores, desc = cb.run(spec, batch, CTX, args.steps, args.warmup, weights=wts)
    line = {
        "impl": "reference", "metric": METRIC, "value": tps, "unit": "tok/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"{MODEL} decode step, {batch} of {SESSIONS} sessions at ctx {CTX} on host cores",
                   "model": MODEL, "sessions": SESSIONS, "ctx": CTX},
        "cpu_baseline": {"value": tps, "unit": "tok/s", "cores": cores, "kind": "port", "sample": desc},
        "e2e": {"value": tps, "unit": "tok/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def claim_stdout():
    """stdout carries exactly ONE line, the JSON record.  Everything else this process prints goes to stderr — Python's
    prints (the reference server's loggers write to sys.stdout at every level) and what native libraries write to file
    descriptor 1 (NCCL prints its version line there).  Returns the file object the record is written to."""
    sys.stdout.flush()
    real = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    sys.stdout = sys.stderr
    return real


# ----------------------------------------------------------------------------- our arm
def run_native(args, rank, world, local_rank):
    import torch
    real_stdout = claim_stdout()
    from fasttalk_llm_microservice_b200 import build
    from fasttalk_llm_microservice_b200.engine import Engine, EngineLimits
    from fasttalk_llm_microservice_b200.model_spec import get_spec
    from fasttalk_llm_microservice_b200.weights import init_weights

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — the native path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    if rank == 0:
        build.build()
    if dist:
        dist.barrier()
    spec = get_spec(MODEL)
    K, W = args.steps, args.warmup
    w = init_weights(spec, seed=rank, device="cuda")
    max_new_long = 2 * K + W + 96       # prefill steps + warm-up + timed pass + profiled pass
    max_ctx = CTX + max_new_long + 128
    eng = Engine(spec, w, EngineLimits(max_batch=SESSIONS, max_batched_tokens=8192, max_context=max_ctx,
                                       decode_prefill_cap=-1),      # set-up prefill unthrottled; default cap before e2e
                 eos_ids=())
    gens = [torch.Generator().manual_seed(1234 + rank * 1000 + i) for i in range(SESSIONS)]
    rand_tokens = lambda i, n: torch.randint(0, 128000, (n,), generator=gens[i]).tolist()

    def barrier():
        torch.cuda.synchronize()
        if dist:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x: float) -> float:
        if not dist:
            return x
        t = torch.tensor([x], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    # ---- setup: 64 sessions prefilled to CTX tokens, all decoding (untimed)
    hist = [rand_tokens(i, CTX) for i in range(SESSIONS)]
    rids = [eng.submit(f"s{i}", hist[i], max_new_tokens=max_new_long, ignore_eos=True)
            for i in range(SESSIONS)]
    while eng.stats()["last_prefill_tokens"] > 0 or eng.stats()["steps"] == 0:
        eng.step()
    eng.drain()

    # ---- settle: the set-up above prefilled 131k tokens on the tensor cores at the board's power limit; the
    # power controller keeps SM clocks capped for a while afterwards.  Start the clock sampler here and give the
    # board SETTLE_S of idle before the warm-up, so that the timed steps below run at the clocks a server that has
    # been decoding for a while sees.  Nothing is timed here; warm-up and the timed region are as passed.
    clocks = ClockSampler(local_rank)
    clocks.start()
    time.sleep(args.settle)

    # ---- value: K decode steps, CUDA events on the engine stream
    for _ in range(W):
        eng.step()
    eng.drain()
    eng.step_times_ms()
    st0 = eng.stats()
    barrier()
    clocks.mark()
    torch.cuda.cudart().cudaProfilerStart()      # `ncu --profile-from-start off` sees exactly the timed steps
    eng.timer_start()
    for _ in range(K):
        eng.step()
    ms = eng.timer_stop_ms()
    torch.cuda.cudart().cudaProfilerStop()
    clocks.unmark()
    barrier()
    clk = clocks.report()               # (the sampler keeps running: the e2e turn below reports its own window)
    per_step = eng.step_times_ms()[-K:]
    st1 = eng.stats()
    ms_rank = ms
    ms = max_over_ranks(ms)
    rank_ms = None
    if dist:
        t = torch.zeros(world, device="cuda", dtype=torch.float64)
        t[rank] = ms_rank / K
        dist.all_reduce(t)
        rank_ms = [round(float(x), 4) for x in t.tolist()]
    tokens = st1["decode_tokens"] - st0["decode_tokens"]
    assert tokens == SESSIONS * K, (tokens, SESSIONS, K)
    value = world * tokens / (ms * 1e-3)
    launches = st1["kernel_launches"] - st0["kernel_launches"]
    ctx_tokens = st1["decode_ctx_tokens"] - st0["decode_ctx_tokens"]
    step_bytes = spec.weight_bytes() + ctx_tokens / K * spec.kv_bytes_per_token()
    peak, peak_src = measured_peaks()
    step_achieved = step_bytes * K / (ms * 1e-3) / 1e9
    eng.drain()
    step_stats = None
    if per_step:
        sp = sorted(per_step)
        step_stats = {"min": sp[0], "median": sp[len(sp) // 2], "max": sp[-1], "sum": sum(per_step),
                      "host_gap_ms_per_step": (ms_rank - sum(per_step)) / K,
                      "first5": [round(x, 4) for x in per_step[:5]],
                      "what": "device ms of each timed step (events around upload..sampled-token copy), rank 0; "
                              "host_gap = (timed total - their sum) / K = GPU idle between steps"}

    if args.tune:
        if rank == 0:
            print(json.dumps({"ms_per_step": ms / K, "step_frac": step_achieved / peak, "step_ms": step_stats,
                              "clocks": clk}), file=real_stdout, flush=True)
        eng.close()
        return

    # ---- roofline of the dominant kernel: second pass, events around every launch
    eng.set_profile(True)
    p0 = eng.stats()
    eng.timer_start()
    for _ in range(K):
        eng.step()
    ms_prof = eng.timer_stop_ms()
    p1 = eng.stats()
    eng.set_profile(False)
    eng.drain()
    n_l = p1["attn_decode_launches"] - p0["attn_decode_launches"]
    attn_ms = p1["attn_decode_ms"] - p0["attn_decode_ms"]
    kv_tok = p1["attn_decode_kv_tokens"] - p0["attn_decode_kv_tokens"]
    layer_kv_bytes = spec.kv_bytes_per_token() // spec.layers
    attn_bytes_per_launch = kv_tok / max(n_l, 1) * layer_kv_bytes
    attn_achieved = attn_bytes_per_launch / (attn_ms / max(n_l, 1) * 1e-3) / 1e9 if attn_ms > 0 else 0.0
    # DRAM traffic of that kernel from the newest committed `ncu --set full` capture of the same workload
    # (profiles/rNN_attn_traffic.json, written from profiles/rNN_ncu_decode.md); null when there is none for this
    # shape.  It is a per-round capture, not a live counter: the kernel's source hash is recorded beside it.
    traffic, traffic_src = None, None
    import glob
    cands = sorted(glob.glob(os.path.join(os.path.dirname(os.path.abspath(__file__)), "profiles", "r*_attn_traffic.json")))
    tpath = cands[-1] if cands else ""
    if MODEL == "llama-3.2-1b" and SESSIONS == 64 and CTX == 2048 and os.path.exists(tpath):
        with open(tpath) as f:
            tj = json.load(f)
        traffic, traffic_src = tj["dram_bytes_per_launch"], tj["source"]

    # ---- finish the long requests, then e2e: one whole turn per session through the public API
    for r in rids:
        eng.cancel(r)
    eng.step(); eng.drain()
    eng.set_decode_prefill_cap(args.prefill_cap)      # serving configuration from here on (0 = the engine default, 896)
    # warm the sessions with a known history (prefix reuse needs the cached tokens == re-sent prefix)
    for i in range(SESSIONS):
        eng.close_session(f"s{i}")
    base = [rand_tokens(i, CTX - TURN_NEW) for i in range(SESSIONS)]
    rr = [eng.submit(f"t{i}", base[i], max_new_tokens=1, ignore_eos=True) for i in range(SESSIONS)]
    live = set(rr)
    while live:
        eng.step()
        for r, t, fin in eng.drain():
            if fin:
                live.discard(r)
    e0 = eng.stats()
    eng.step_log()
    native_loop = not args.no_native_loop
    if not native_loop:
        os.environ["FASTTALK_NATIVE_LOOP"] = "0"       # the handler and WebSocket levels step from Python as well
    if native_loop:
        eng.start_loop()               # the engine steps on its own thread; this one submits, waits and drains
    barrier()
    t_start = time.perf_counter()
    import numpy as np
    rids_t = []
    for i in range(SESSIONS):          # host token buffers in; the engine uploads what it needs
        full = base[i] + rand_tokens(i, TURN_NEW)
        rids_t.append(eng.submit(f"t{i}", full, max_new_tokens=TURN_GEN, ignore_eos=True))
    rid0 = min(rids_t)
    assert rids_t == list(range(rid0, rid0 + SESSIONS))
    # per-request arrival stamps kept in arrays: the loop below is what a streaming host does per step
    t_first = np.full(SESSIONS, -1.0)
    t_last = np.zeros(SESSIONS)
    gaps, n_out, n_live = [], 0, SESSIONS
    n_timeouts = 0
    while n_live:
        if native_loop:
            if eng.wait_events(100) <= 0:
                n_timeouts += 1
                continue
        else:
            eng.step()
        now = time.perf_counter()
        n, rq, tk, fn = eng.drain_arrays()
        idx = (rq[:n] - rid0).astype(np.int64)
        got = idx[tk[:n] >= 0]
        n_out += int(got.size)
        seen = t_first[got] >= 0
        gaps.extend((now - t_last[got[seen]]).tolist())
        t_first[got[~seen]] = now - t_start
        t_last[got] = now
        n_live -= int(np.count_nonzero(fn[:n]))
    first = {i: float(t_first[i]) for i in range(SESSIONS)}
    torch.cuda.synchronize()
    e2e_s = time.perf_counter() - t_start
    print(f"[rank {rank}] e2e_engine turn: {e2e_s * 1e3:.1f} ms, {n_timeouts} wait time-outs, largest gap "
          f"{1e3 * max(gaps):.1f} ms, last first token at {1e3 * float(t_first.max()):.1f} ms", file=sys.stderr, flush=True)
    if native_loop:
        eng.stop_loop()
    e_clk = clocks.report(t_start, time.perf_counter())
    clocks.stop()
    e2e_s = max_over_ranks(e2e_s)
    e1 = eng.stats()
    e_log = eng.step_log()
    # the steps streaming sessions waited longest for: [device ms, decode rows, prompt tokens, host ms]
    e_mixed = sorted((r for r in e_log.tolist() if r[1] > 0), key=lambda r: -r[0])[:4]
    # the turn's pure decode steps: device ms and host ms (launch .. bookkeeping) per step
    e_dec = [r for r in e_log.tolist() if r[1] >= SESSIONS and r[2] == 0]
    e_dec_ms = [statistics.median(r[0] for r in e_dec), statistics.median(r[3] for r in e_dec)] if e_dec else None
    e_steps = e1["steps"] - e0["steps"]
    e2e_value = world * n_out / e2e_s

    # ---- cold TTFT: one 2048-token conversation nobody has cached, alone on the GPU
    def ttft_once(key, toks):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        rid = eng.submit(key, toks, max_new_tokens=1, ignore_eos=True)
        while True:
            eng.step()
            if any(r == rid and t >= 0 for r, t, fin in eng.drain()):
                return 1e3 * (time.perf_counter() - t0)
    gcold = torch.Generator().manual_seed(99 + rank)
    cold = sorted(ttft_once(f"cold{j}", torch.randint(0, 128000, (CTX,), generator=gcold).tolist()) for j in range(5))
    ttft_cold_ms = cold[len(cold) // 2]

    for i in range(SESSIONS):
        eng.close_session(f"t{i}")
    for j in range(5):
        eng.close_session(f"cold{j}")

    # ---- e2e through the reference-facing plugin: 64 concurrent NativeHandler.generate_stream_async calls,
    # `messages` (host strings) in, one `str` chunk per token out — the call websocket_server_vllm.py:528-539 makes.
    # Inside the timed region: chat template + tokenisation, submit, the engine thread's steps (metadata H2D, sampled
    # ids D2H), detokenisation, one call_soon_threadsafe per step, the asyncio queues.
    import asyncio
    import gc
    import random
    # This process holds ~10^7 Python objects by now (token lists of the set-up): a full collection takes tens of
    # milliseconds and, in the middle of a turn, shows up as a 60 ms inter-token gap at the WebSocket level.  Collect once
    # and take what exists out of the collector's sight, as a long-running server does after start-up.
    gc.collect()
    gc.freeze()
    from fasttalk_llm_microservice_b200.handler import EngineRunner, NativeHandler
    from fasttalk_llm_microservice_b200.tokenizer import ByteTokenizer
    tok = ByteTokenizer(spec.vocab)
    runner = EngineRunner(eng, tok)
    handler = NativeHandler(model=MODEL, runner=runner)
    rnd = random.Random(4321 + rank)
    words = ["alpha", "bravo", "charlie", "delta", "echo", "foxtrot", "golf", "hotel", "india", "juliet", "kilo", "lima"]

    def text_of(n_chars):
        out = []
        while sum(len(x) + 1 for x in out) < n_chars:
            out.append(rnd.choice(words))
        return " ".join(out)[:n_chars]

    def conv_sized(target_tokens, tail):
        """messages whose templated length is exactly target_tokens: system + user, then `tail` messages."""
        sys_len = max(16, target_tokens - 64)
        for _ in range(4):
            m = [{"role": "system", "content": text_of(sys_len // 2)}, {"role": "user", "content": text_of(sys_len - sys_len // 2)}] + tail
            sys_len += target_tokens - len(tok.apply_chat_template(m))
        m[1]["content"] = m[1]["content"] + "x" * max(0, target_tokens - len(tok.apply_chat_template(m)))
        return m

    async def handler_turn(convs, max_tokens, tag):
        t0 = time.perf_counter()
        async def one(i, msgs):
            stamps = []
            async for chunk in handler.generate_stream_async(msgs, temperature=0.0, max_tokens=max_tokens, top_p=1.0,
                                                             request_id=f"{tag}{i}"):
                stamps.append(time.perf_counter() - t0)
            return stamps
        res = await asyncio.gather(*[one(i, m) for i, m in enumerate(convs)])
        return res, time.perf_counter() - t0

    conv1 = [conv_sized(CTX - TURN_NEW, []) for _ in range(SESSIONS)]
    asyncio.run(handler_turn(conv1, 1, "h"))                                     # turn 1 (untimed): histories cached
    conv2 = [m + [{"role": "assistant", "content": "ok"}, {"role": "user", "content": text_of(TURN_NEW - 16)}] for m in conv1]
    h0 = eng.stats()
    barrier()
    h_res, h_wall = asyncio.run(handler_turn(conv2, TURN_GEN, "h"))
    torch.cuda.synchronize()
    h_wall = max_over_ranks(h_wall)
    h1 = eng.stats()
    h_steps = max(1, h1["steps"] - h0["steps"])
    h_tokens = (h1["decode_tokens"] - h0["decode_tokens"]) + SESSIONS          # engine-side count: decode rows + first tokens
    h_chunks = sum(len(r) for r in h_res)
    h_first = [r[0] for r in h_res if r]
    h_gaps = [b - a for r in h_res for a, b in zip(r, r[1:])]
    e2e_handler = {"value": world * h_chunks / h_wall, "unit": "tok/s",
                   "h2d_bytes_per_step": (h1["h2d_bytes"] - h0["h2d_bytes"]) / h_steps,
                   "d2h_bytes_per_step": (h1["d2h_bytes"] - h0["d2h_bytes"]) / h_steps,
                   "what": f"{SESSIONS} concurrent NativeHandler.generate_stream_async calls (the reference's handler "
                           f"interface): messages in, one str chunk per token out; ~{CTX}-token templated history "
                           f"(~{TURN_NEW} new, rest served from the session's KV), {TURN_GEN} generated; byte tokenizer",
                   "chunks": h_chunks, "engine_tokens": h_tokens, "steps": h_steps, "wall_s": h_wall,
                   "prompt_tokens_mean": sum(len(tok.apply_chat_template(m)) for m in conv2) / SESSIONS,
                   "prefix_hit_tokens": h1["prefix_hit_tokens"] - h0["prefix_hit_tokens"],
                   "ttft_ms_p50": 1e3 * pct(h_first, 0.5), "ttft_ms_p99": 1e3 * pct(h_first, 0.99),
                   "itl_ms_p50": 1e3 * pct(h_gaps, 0.5), "itl_ms_p99": 1e3 * pct(h_gaps, 0.99),
                   "decode_prefill_cap": args.prefill_cap or 896}
    for i in range(SESSIONS):
        eng.close_session(f"h{i}")

    # ---- WebSocket level (SURVEY.md §8d "two levels"): the reference's UNCHANGED WebSocketLLMServer from
    # baseline/_ref, uvicorn on 127.0.0.1, 64 `websockets` clients in their own process, two turns; turn 2 timed
    # client-side.  Rank 0 only, N = 1 only (one server per GPU in a deployment; the clients need host cores).
    ws_level = None
    if rank == 0 and world == 1 and not args.no_ws:
        sys.path.insert(0, os.path.join(ROOT, "tests"))
        try:
            import ws_harness as WS
            if WS.find_reference() is None:
                ws_level = {"unavailable": "no reference checkout (baseline/_ref): run scripts/install_reference.sh"}
            else:
                srv = WS.ReferenceServer(runner, env={"DEFAULT_MAX_TOKENS": str(TURN_GEN), "VLLM_MODEL": MODEL})
                try:
                    n1 = CTX - TURN_NEW - TURN_GEN - 40
                    specs = [{"system_prompt": text_of(n1 // 2), "turns": [text_of(n1 - n1 // 2), text_of(TURN_NEW - 16)]}
                             for _ in range(SESSIONS)]
                    w0 = eng.stats()
                    res = WS.run_clients_subprocess(srv.port, specs)
                    w1 = eng.stats()
                    ws_level = WS.turn_metrics(res, 1)
                    ws_level.update({"unit": "tok/s", "value": ws_level.pop("tok_s"),
                                     "what": "unchanged reference WebSocketLLMServer (websocket_server_vllm.py) + NativeHandler, "
                                             f"{SESSIONS} websockets clients in a separate process, turn 2 of 2 "
                                             f"(~{CTX}-token history re-sent, {TURN_GEN} tokens generated), timed at the client",
                                     "turn1": {k: v for k, v in WS.turn_metrics(res, 0).items() if k in ("tok_s", "ttft_ms_p50", "ttft_ms_p99")},
                                     "prefix_hit_tokens": w1["prefix_hit_tokens"] - w0["prefix_hit_tokens"],
                                     "reference_from": os.path.relpath(srv.ref, ROOT) if srv.ref.startswith(ROOT) else srv.ref})
                finally:
                    srv.stop()
        except Exception as e:  # noqa: BLE001 - the WS level is reported, never allowed to lose the bench line
            ws_level = {"unavailable": f"{type(e).__name__}: {e}"[:300]}
    runner.shutdown()

    cpu_base = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        from oracle import cpu_baseline as cb
        wts = cb.make_weights(spec)
        _, ms1, _, _ = cb.run(spec, SESSIONS, CTX, 1, 1, weights=wts)
        n_cpu = max(2, min(20, int(15000 / max(ms1, 1))))
        tps, cms, cores, desc = cb.run(spec, SESSIONS, CTX, n_cpu, 1, weights=wts)
        cpu_base = {"value": tps, "unit": "tok/s", "cores": cores, "kind": "port", "sample": desc,
                    "ms_per_step": cms}

    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": "tok/s", "n_gpus": world, "steps": K, "warmup": W,
            "ms_per_step": ms / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "bf16", "data": "synthetic",
            "config": {"workload": f"{MODEL} bf16, {SESSIONS} sessions/GPU decoding at ctx {CTX}..{CTX + K + W}, "
                                   "continuous batching (BASELINE configs[1])",
                       "model": MODEL, "sessions_per_gpu": SESSIONS, "ctx": CTX, "parallelism": f"replicas x{world}",
                       "l2": "no flush needed: each step streams 6.8 GB >> 126 MB L2",
                       "weights": "random-init (seed per rank, std 0.02)", "sampling": "greedy, EOS ignored"},
            "roofline": {"bound": "hbm", "kernel": "attn_decode_kernel<64>", "achieved": attn_achieved, "peak": peak,
                         "unit": "GB/s", "frac": attn_achieved / peak, "traffic": traffic, "traffic_source": traffic_src,
                         "peak_source": peak_src,
                         "launches_timed": n_l, "bytes_per_launch": attn_bytes_per_launch,
                         "us_per_launch": 1e3 * attn_ms / max(n_l, 1),
                         "pass": "second pass of K steps with CUDA events around each launch",
                         "ms_per_step_profiled": ms_prof / K},
            "step_roofline": {"bound": "hbm", "bytes_per_step": step_bytes, "achieved": step_achieved, "peak": peak,
                              "unit": "GB/s", "frac": step_achieved / peak,
                              "note": "whole decode step: weights + KV bytes (SURVEY.md §8d) / device time"},
            "cpu_baseline": cpu_base,
            "e2e": e2e_handler,
            "e2e_engine": {"value": e2e_value, "unit": "tok/s",
                    "h2d_bytes_per_step": (e1["h2d_bytes"] - e0["h2d_bytes"]) / max(e_steps, 1),
                    "d2h_bytes_per_step": (e1["d2h_bytes"] - e0["d2h_bytes"]) / max(e_steps, 1),
                    "what": f"one turn/session via the C ABI (ft_engine_submit / {'ft_engine_start + wait' if not args.no_native_loop else 'step'} / drain) with host token lists: "
                            f"{CTX}-token history ({TURN_NEW} new, rest prefix-cached), {TURN_GEN} generated",
                    "steps": e_steps, "wall_s": e2e_s, "tokens": n_out,
                    "ttft_ms_p50": 1e3 * pct(list(first.values()), 0.5), "ttft_ms_p99": 1e3 * pct(list(first.values()), 0.99),
                    "itl_ms_p50": 1e3 * pct(gaps, 0.5), "itl_ms_p99": 1e3 * pct(gaps, 0.99),
                    "prefix_hit_tokens": e1["prefix_hit_tokens"] - e0["prefix_hit_tokens"],
                    "slowest_steps_with_decode_rows": [[round(x, 3) for x in r] for r in e_mixed],
                    "decode_steps_in_turn_ms_p50": {"device": e_dec_ms[0], "host": e_dec_ms[1], "n": len(e_dec)} if e_dec_ms else None,
                    "clocks_in_turn": {k: e_clk[k] for k in ("sm_mhz", "sm_min_mhz", "reasons", "samples", "power_w_max")},
                    "first_step": [round(x, 3) for x in e_log[0].tolist()] if len(e_log) else None},
            "ws": ws_level,
            "latency": {"itl_ms_p50_decode": ms / K, "ttft_ms_cold_2048_single": ttft_cold_ms,
                        "ttft_ms_warm_p50_64_sessions": 1e3 * pct(list(first.values()), 0.5)},
            "gpu_launches": launches,
            "clocks": clk,
            "step_ms": step_stats,
            "rank_ms_per_step": rank_ms,
            "settle_s": args.settle,
        }
        print(json.dumps(line), file=real_stdout, flush=True)
    eng.close()
    if dist:
        dist.barrier()
        dist.destroy_process_group()


def run_tp(args, rank, world, local_rank):
    """`--tp 2` (BASELINE configs[3]): Llama-3.1-8B at a 7936-token history, one engine over two GPUs with the all-reduce
    fused into the row-parallel GEMMs (csrc/gemm.cu), against the same engine on one GPU.  Launch under torchrun with
    2 ranks.  One JSON line from rank 0; `value` = streamed tok/s at 8 sessions with TP=2."""
    import torch
    import torch.distributed as dist
    assert world == 2, "--tp 2 needs exactly two ranks (torchrun --nproc-per-node 2)"
    real_stdout = claim_stdout()
    torch.cuda.set_device(local_rank)
    dist.init_process_group("gloo")
    sys.path.insert(0, os.path.join(ROOT, "scripts"))
    import bench_tp
    from fasttalk_llm_microservice_b200 import build, tp
    from fasttalk_llm_microservice_b200.engine import Engine, EngineLimits
    from fasttalk_llm_microservice_b200.model_spec import get_spec
    from fasttalk_llm_microservice_b200.weights import init_weights
    if rank == 0:
        build.build()
    dist.barrier()
    spec = get_spec("llama-3.1-8b")
    w = init_weights(spec, seed=0, device="cuda")
    lim = EngineLimits(max_batch=8, max_batched_tokens=8192, max_context=8192, decode_prefill_cap=-1)
    ctx, steps = 7936, max(args.steps, 8)
    eng = Engine(spec, w, lim, eos_ids=(), tp_rank=rank, tp_size=2)
    tp.connect(eng)
    tp2 = []
    for sessions in (1, 8):
        dist.barrier()
        tp2.append(bench_tp.run(eng, sessions, ctx, steps, spec, f"a{sessions}_", 0.5))
    dist.barrier()
    eng.close()
    if rank == 0:
        one = Engine(spec, w, lim, eos_ids=())
        tp1 = [bench_tp.run(one, sessions, ctx, steps, spec, f"b{sessions}_", 1.0) for sessions in (1, 8)]
        one.close()
        line = {"metric": "Llama-3.1-8B streamed tok/s, 8192-token window, TP=2 over NVLink (BASELINE configs[3])",
                "value": tp2[1]["tok_s"], "unit": "tok/s", "n_gpus": 2, "steps": steps, "warmup": 8,
                "ms_per_step": tp2[1]["ms_per_step"], "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
                "dtype": "bf16", "data": "synthetic",
                "config": {"workload": "llama-3.1-8b bf16, 8 sessions at a 7936-token history, one engine over 2 GPUs",
                           "parallelism": "tp2, all-reduce fused into the row-parallel GEMM epilogues (no collective call)"},
                "tp2": tp2, "tp1_same_box": tp1,
                "tp2_over_tp1": [a["tok_s"] / b["tok_s"] for a, b in zip(tp2, tp1)],
                "exchange_ab": "profiles/r02_tp2.md (fused GEMM + exchange vs GEMM + ncclAllReduce)"}
        print(json.dumps(line), file=real_stdout, flush=True)
    dist.barrier()
    dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="native", choices=["native", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-ws", action="store_true", help="skip the WebSocket-level run")
    ap.add_argument("--tp", type=int, default=1, help="2 = the Llama-3.1-8B tensor-parallel configuration (torchrun, 2 ranks)")
    ap.add_argument("--tune", action="store_true", help="kernel tuning: print ms/step of the timed decode steps and exit")
    ap.add_argument("--prefill-cap", type=int, default=0,
                    help="engine decode_prefill_cap for the e2e runs (0 = engine default 896, < 0 = uncapped)")
    ap.add_argument("--no-native-loop", action="store_true",
                    help="step the engine from Python in the e2e levels instead of ft_engine_start's native thread (A/B)")
    ap.add_argument("--settle", type=float, default=2.0,
                    help="idle seconds between the set-up prefill and the warm-up (power-cap recovery; untimed)")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "native" else args.warmup
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank, world)
        return
    if args.tp == 2:
        run_tp(args, rank, world, local_rank)
        return
    if world == 1 and args.gpus > 1:
        # launched without torchrun: re-exec under it so `python bench.py --gpus N` works too
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={args.gpus}",
               "--master-addr", "127.0.0.1", "--master-port", "29531", os.path.abspath(__file__),
               "--gpus", str(args.gpus), "--steps", str(args.steps), "--warmup", str(args.warmup)]
        raise SystemExit(subprocess.call(cmd))
    run_native(args, rank, world, local_rank)


if __name__ == "__main__":
    main()
