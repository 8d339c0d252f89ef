"""Boundary A on the GPU (round-1 verdict, missing #1): the reference's UNCHANGED WebSocketLLMServer
(`app/core/websocket_server_vllm.py:111-123, 528-539`) with NativeHandler swapped in, over the REAL engine,
64 concurrent `websockets` clients, byte tokenizer, DEFAULT_TEMPERATURE=0.0.  Every session's token frames
must equal what the handler streams for the very conversation the server built.

The reference comes from `baseline/_ref` (scripts/install_reference.sh; git-ignored, travels to the GPU box)."""
import asyncio
import os

import numpy as np
import pytest

import ws_harness as W

pytestmark = [pytest.mark.gpu,
              pytest.mark.skipif(W.find_reference() is None,
                                 reason="baseline/_ref missing: run scripts/install_reference.sh before gpurun")]
GOLD = os.path.join(os.path.dirname(__file__), "golden")


def test_64_websocket_clients_through_the_unchanged_reference_server():
    from fasttalk_llm_microservice_b200 import build
    from fasttalk_llm_microservice_b200.engine import Engine, EngineLimits
    from fasttalk_llm_microservice_b200.handler import EngineRunner, NativeHandler
    from fasttalk_llm_microservice_b200.model_spec import SPECS
    from fasttalk_llm_microservice_b200.tokenizer import ByteTokenizer
    from fasttalk_llm_microservice_b200.weights import init_weights
    build.build()
    g = np.load(os.path.join(GOLD, "hf_tiny-d64.npz"))
    spec = SPECS["tiny-d64"]
    w = init_weights(spec, seed=int(g["seed"]), std=float(g["std"]), norm_jitter=float(g["norm_jitter"]))
    tok = ByteTokenizer(spec.vocab)
    eng = Engine(spec, w, EngineLimits(max_batch=64, max_batched_tokens=2048, max_context=2048), eos_ids=())
    runner = EngineRunner(eng, tok)
    srv = W.ReferenceServer(runner, env={"DEFAULT_MAX_TOKENS": "24"})
    try:
        specs = [{"system_prompt": f"you are assistant number {i}" + " be brief" * (i % 7),
                  "turns": [f"question {i} " + "x" * (5 * i), "and a follow-up " + "y" * (3 * i)]} for i in range(64)]
        st0 = eng.stats()
        res = W.run_clients(srv.port, specs)
        st1 = eng.stats()
        assert len(res) == 64 and len(srv.calls) == 128
        # the second turn re-sent the history and hit the session's KV (request_id == session id)
        assert st1["prefix_hit_tokens"] - st0["prefix_hit_tokens"] > 0
        # replay, at handler level, the exact calls the server made (same messages, same sampling keywords) under
        # fresh session ids, all sessions of a turn concurrently as over the wire, and compare the streams
        h = NativeHandler(model="tiny-d64", runner=runner)
        by_sid = {}
        for c in srv.calls:
            by_sid.setdefault(c["kw"]["request_id"], []).append(c)

        async def replay(turn):
            async def one(r):
                c = by_sid[r["session_id"]][turn]
                kw = dict(c["kw"], request_id="replay-" + r["session_id"])
                return [p async for p in h.generate_stream_async(c["messages"], **kw)]
            return await asyncio.gather(*[one(r) for r in res])
        diverged = 0
        for turn in (0, 1):
            want = asyncio.run(replay(turn))
            for r, exp in zip(res, want):
                t = r["turns"][turn]
                assert t["end"]["type"] == "response_complete", t["end"]
                assert t["end"]["stats"]["tokens_generated"] == len(t["chunks"])
                assert 1 <= len(t["chunks"]) <= 24 and all(isinstance(c, str) and c for c in t["chunks"])
                if t["chunks"] != exp:
                    # The decode-attention work split depends on which sessions share a step, so two runs of one
                    # conversation can differ in the last float bits; a token flips only at an exact near-tie.
                    # Such a session must still agree on a prefix, and there may be at most two of them.
                    diverged += 1
                    assert t["chunks"][:1] == exp[:1], (turn, r["session_id"])
        assert diverged <= 2, diverged
        m = W.turn_metrics(res, 1)
        print(f"\n[ws] 64 sessions, tiny-d64: turn 2 {m['tok_s']:.0f} tok/s, TTFT p50 {m['ttft_ms_p50']:.1f} ms, "
              f"ITL p50 {m['itl_ms_p50']:.2f} ms; {diverged} session-turns diverged at a near-tie")
    finally:
        srv.stop()
        runner.shutdown()
        eng.close()
