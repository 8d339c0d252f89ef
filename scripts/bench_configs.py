"""The other BASELINE.json configurations, measured like bench.py measures configs[1] (they are not
bench lines; their numbers go into DESIGN.md §6 and profiles/):

  cfg1  Llama-3.2-1B, ONE session, 128-token prompt, 256 greedy tokens (the reference's CPU-runnable case)
  cfg3  Llama-3.2-3B, 256 sessions at a 2k history on one GPU, and the 32-sessions-per-GPU share of an 8-GPU node
  cfg5  Llama-3.1-8B agent turns through the OpenAI wire (boundary B): 16 concurrent agents, two completions per
        turn — 960-token prompt (system + tool schemas + history) -> 48 tokens, + 256-token tool result -> 128 tokens

    python scripts/bench_configs.py [cfg1] [cfg3] [cfg5] [models64]
"""
import asyncio, json, os, socket, sys, threading, time
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from fasttalk_llm_microservice_b200 import build
from fasttalk_llm_microservice_b200.engine import Engine, EngineLimits
from fasttalk_llm_microservice_b200.model_spec import get_spec
from fasttalk_llm_microservice_b200.weights import init_weights

PEAK = 6545e9


def rand(n, seed):
    return torch.randint(0, 128000, (n,), generator=torch.Generator().manual_seed(seed)).tolist()


def decode_run(eng, spec, sessions, ctx, steps, tag):
    rids = [eng.submit(f"{tag}{i}", rand(ctx, 100 + i), max_new_tokens=steps + 40, ignore_eos=True) for i in range(sessions)]
    t0 = time.perf_counter()
    while True:
        eng.step(); eng.drain()
        st = eng.stats()
        if st["last_prefill_tokens"] == 0 and st["last_batch"] == sessions:
            break
    torch.cuda.synchronize()
    prefill_s = time.perf_counter() - t0
    for _ in range(8):
        eng.step(); eng.drain()
    eng.timer_start()
    for _ in range(steps):
        eng.step()
    ms = eng.timer_stop_ms() / steps
    eng.drain()
    for r in rids:
        eng.cancel(r)
    eng.step(); eng.drain()
    for i in range(sessions):
        eng.close_session(f"{tag}{i}")
    b = spec.weight_bytes() + sessions * (ctx + 8 + steps // 2) * spec.kv_bytes_per_token()
    return {"sessions": sessions, "ctx": ctx, "ms_per_step": ms, "tok_s": sessions / ms * 1e3,
            "cold_prefill_s": prefill_s, "prefill_tok_s": sessions * ctx / prefill_s,
            "bytes_per_step": b, "hbm_frac": b / (ms * 1e-3) / PEAK}


def cfg1():
    spec = get_spec("llama-3.2-1b")
    eng = Engine(spec, init_weights(spec, seed=0, device="cuda"), EngineLimits(max_batch=4, max_batched_tokens=512, max_context=1024), eos_ids=())
    eng.generate([rand(128, 1)], 16)                      # warm-up
    eng.close_session(0)
    prompt = rand(128, 2)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    rid = eng.submit("one", prompt, max_new_tokens=256, ignore_eos=True)
    stamps = []
    while len(stamps) < 256:
        eng.step()
        for r, t, fin in eng.drain():
            if t >= 0:
                stamps.append(time.perf_counter())
    total = stamps[-1] - t0
    gaps = sorted(b - a for a, b in zip(stamps, stamps[1:]))
    b = spec.weight_bytes() + 256 * spec.kv_bytes_per_token()
    res = {"config": "cfg1: llama-3.2-1b, 1 session, 128-token prompt, 256 greedy tokens", "tok_s": 256 / total,
           "ttft_ms": 1e3 * (stamps[0] - t0), "itl_ms_p50": 1e3 * gaps[len(gaps) // 2], "itl_ms_p99": 1e3 * gaps[int(len(gaps) * 0.99)],
           "hbm_frac_decode": b / gaps[len(gaps) // 2] / PEAK}
    eng.close()
    return res


def cfg3():
    spec = get_spec("llama-3.2-3b")
    w = init_weights(spec, seed=0, device="cuda")
    out = {"config": "cfg3: llama-3.2-3b, 2k histories; 256 sessions on one GPU and the 32/GPU share of 8 GPUs"}
    eng = Engine(spec, w, EngineLimits(max_batch=256, max_batched_tokens=8192, max_context=2048 + 256, decode_prefill_cap=-1), eos_ids=())
    out["n1_256_sessions"] = decode_run(eng, spec, 256, 2048, 32, "a")
    out["n8_share_32_sessions"] = decode_run(eng, spec, 32, 2048, 64, "b")
    eng.close()
    return out


def models64():
    """Not a BASELINE configuration: the bench's own shape (64 sessions at a 2k history) on the two larger models, to
    show where the step sits against the roofline when the projections are bigger."""
    out = {"config": "64 sessions x 2k history on the larger models (same shape as bench.py's cfg 2)"}
    for name in ("llama-3.2-3b", "llama-3.1-8b"):
        spec = get_spec(name)
        w = init_weights(spec, seed=0, device="cuda")
        eng = Engine(spec, w, EngineLimits(max_batch=64, max_batched_tokens=8192, max_context=2048 + 256, decode_prefill_cap=-1), eos_ids=())
        out[name] = decode_run(eng, spec, 64, 2048, 48, "m")
        eng.close()
        del w
        torch.cuda.empty_cache()
    return out


def cfg5():
    import httpx, openai, uvicorn
    from fasttalk_llm_microservice_b200.handler import EngineRunner, NativeHandler
    from fasttalk_llm_microservice_b200.openai_server import create_app
    from fasttalk_llm_microservice_b200.tokenizer import ByteTokenizer
    spec = get_spec("llama-3.1-8b")
    eng = Engine(spec, init_weights(spec, seed=0, device="cuda"), EngineLimits(max_batch=16, max_batched_tokens=8192, max_context=4096, decode_prefill_cap=-1), eos_ids=())
    tok = ByteTokenizer(spec.vocab)
    runner = EngineRunner(eng, tok)
    h = NativeHandler(model="llama3.1:8b", runner=runner)
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    server = uvicorn.Server(uvicorn.Config(create_app(h), host="127.0.0.1", port=port, log_level="error", lifespan="off"))
    server.install_signal_handlers = lambda: None
    th = threading.Thread(target=server.run, daemon=True); th.start()
    while not server.started:
        time.sleep(0.02)
    # byte tokenizer: one ASCII character = one token, so text length = token count
    tools = [{"type": "function", "function": {"name": f"tool_{i}", "description": "d" * 40, "parameters": {
        "type": "object", "properties": {"q": {"type": "string"}}, "required": ["q"]}}} for i in range(2)]
    n_agents = 16

    async def agent(i, ac, stats):
        sysm = {"role": "system", "content": f"agent {i:02d} " + "s" * 50}
        user = {"role": "user", "content": "h" * 500}
        t0 = time.perf_counter()
        first = None
        n1 = 0
        stream = await ac.chat.completions.create(model="m", messages=[sysm, user], tools=tools, stream=True,
                                                  temperature=0.0, max_tokens=48)
        async for ch in stream:
            if ch.choices and ch.choices[0].delta.content:
                n1 += 1
                first = first or time.perf_counter()
        t1 = time.perf_counter()
        msgs = [sysm, user, {"role": "assistant", "tool_calls": [{"id": "c1", "type": "function", "function": {
            "name": "tool_0", "arguments": json.dumps({"q": "x" * 20})}}]}, {"role": "tool", "tool_call_id": "c1", "content": "r" * 256}]
        n2 = 0
        first2 = None
        stream = await ac.chat.completions.create(model="m", messages=msgs, tools=tools, stream=True,
                                                  temperature=0.0, max_tokens=128)
        async for ch in stream:
            if ch.choices and ch.choices[0].delta.content:
                n2 += 1
                first2 = first2 or time.perf_counter()
        t2 = time.perf_counter()
        stats.append({"ttft1": first - t0, "round1": t1 - t0, "ttft2": first2 - t1, "round2": t2 - t1, "chunks": n1 + n2})

    async def main():
        ac = openai.AsyncOpenAI(base_url=f"http://127.0.0.1:{port}/v1", api_key="x", timeout=600, max_retries=0,
                                http_client=httpx.AsyncClient(limits=httpx.Limits(max_connections=64), timeout=600))
        stats = []
        await agent(99, ac, [])                         # warm-up turn
        t0 = time.perf_counter()
        await asyncio.gather(*[agent(i, ac, stats) for i in range(n_agents)])
        wall = time.perf_counter() - t0
        await ac.close()
        return stats, wall
    p0 = eng.stats()
    stats, wall = asyncio.run(main())
    p1 = eng.stats()
    med = lambda k: sorted(s[k] for s in stats)[len(stats) // 2]
    res = {"config": "cfg5: llama-3.1-8b, 16 agents, OpenAI wire with tools, 2 completions per turn (48 + 128 tokens)",
           "turn_wall_s": wall, "generated_tokens": p1["decode_tokens"] - p0["decode_tokens"] + 2 * (n_agents + 1),
           "chunks_streamed": sum(s["chunks"] for s in stats), "tok_s": sum(s["chunks"] for s in stats) / wall,
           "ttft_round1_ms_p50": 1e3 * med("ttft1"), "ttft_round2_ms_p50": 1e3 * med("ttft2"),
           "round1_s_p50": med("round1"), "round2_s_p50": med("round2"),
           "prefix_hit_tokens": p1["prefix_hit_tokens"] - p0["prefix_hit_tokens"],
           "prefill_tokens": p1["prefill_tokens"] - p0["prefill_tokens"]}
    server.should_exit = True
    th.join(timeout=5)
    runner.shutdown(); eng.close()
    return res


def _http_client_proc(port, n_streams, hist_chars, gen, q, tag):
    """One client process: n_streams concurrent SSE streams, counts chunks without parsing JSON."""
    import asyncio, httpx, time

    async def one(i, hc, out):
        body = {"model": "m", "stream": True, "temperature": 0.0, "max_tokens": gen,
                "messages": [{"role": "system", "content": f"{tag}-{i:03d}"}, {"role": "user", "content": "h" * hist_chars}]}
        t0 = time.perf_counter(); first = None; n = 0
        async with hc.stream("POST", f"http://127.0.0.1:{port}/v1/chat/completions", json=body) as r:
            async for line in r.aiter_lines():
                if line.startswith("data: {") and '"content":"' in line and '"content":""' not in line:
                    n += 1
                    first = first or time.perf_counter() - t0
        out.append((n, first, time.perf_counter() - t0))

    async def main():
        out = []
        async with httpx.AsyncClient(timeout=600, limits=httpx.Limits(max_connections=n_streams + 4)) as hc:
            await asyncio.gather(*[one(i, hc, out) for i in range(n_streams)])
        return out
    q.put(asyncio.run(main()))


def cfg2_http():
    """cfg 2 through boundary B: 64 concurrent SSE streams (4 client processes x 16) against ONE server process."""
    import multiprocessing as mp, uvicorn
    from fasttalk_llm_microservice_b200.handler import EngineRunner, NativeHandler
    from fasttalk_llm_microservice_b200.openai_server import create_app
    from fasttalk_llm_microservice_b200.tokenizer import ByteTokenizer
    spec = get_spec("llama-3.2-1b")
    eng = Engine(spec, init_weights(spec, seed=0, device="cuda"), EngineLimits(max_batch=64, max_batched_tokens=8192, max_context=4096, decode_prefill_cap=-1), eos_ids=())
    runner = EngineRunner(eng, ByteTokenizer(spec.vocab))
    h = NativeHandler(model="llama3.2:1b", runner=runner)
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    server = uvicorn.Server(uvicorn.Config(create_app(h), host="127.0.0.1", port=port, log_level="error", lifespan="off"))
    server.install_signal_handlers = lambda: None
    th = threading.Thread(target=server.run, daemon=True); th.start()
    while not server.started:
        time.sleep(0.02)
    ctx = mp.get_context("spawn")
    res = {}
    for tag, gen in (("warm", 8), ("run", 256)):          # first pass warms the server; histories differ per pass
        q = ctx.Queue()
        ps = [ctx.Process(target=_http_client_proc, args=(port, 16, 2000, gen, q, f"{tag}{k}")) for k in range(4)]
        t0 = time.perf_counter()
        for p_ in ps: p_.start()
        rows = [r for _ in ps for r in q.get(timeout=600)]
        for p_ in ps: p_.join()
        wall = time.perf_counter() - t0
        res[tag] = (rows, wall)
    rows, wall = res["run"]
    chunks = sum(r[0] for r in rows)
    stream_s = max(r[2] for r in rows)
    ttft = sorted(r[1] for r in rows)
    out = {"config": "cfg2 over the OpenAI wire: llama-3.2-1b, 64 SSE streams (4 client processes), ~2k-token cold histories, 256 tokens each",
           "chunks": chunks, "slowest_stream_s": stream_s, "tok_s_streamed": chunks / stream_s,
           "ttft_ms_p50": 1e3 * ttft[len(ttft) // 2], "wall_incl_process_spawn_s": wall}
    server.should_exit = True
    th.join(timeout=5)
    runner.shutdown(); eng.close()
    return out

if __name__ == "__main__":
    build.build()
    which = sys.argv[1:] or ["cfg1", "cfg3", "cfg5"]
    for name in which:
        r = {"cfg1": cfg1, "cfg3": cfg3, "cfg5": cfg5, "cfg2_http": cfg2_http, "models64": models64}[name]()
        print("CFG " + json.dumps(r), flush=True)
        torch.cuda.empty_cache()
