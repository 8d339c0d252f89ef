"""Boundary B over the REAL engine on the GPU: the `openai` client streaming from
openai_server.py gets, chunk for chunk, what NativeHandler yields for the same conversation,
and a second turn is served from the session's KV cache."""
import os
import socket
import threading
import time

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(__file__), "golden")


@pytest.fixture(scope="module")
def served():
    import uvicorn
    from fasttalk_llm_microservice_b200 import build
    from fasttalk_llm_microservice_b200.engine import Engine, EngineLimits
    from fasttalk_llm_microservice_b200.handler import EngineRunner, NativeHandler
    from fasttalk_llm_microservice_b200.model_spec import SPECS
    from fasttalk_llm_microservice_b200.openai_server import create_app
    from fasttalk_llm_microservice_b200.tokenizer import ByteTokenizer
    from fasttalk_llm_microservice_b200.weights import init_weights
    build.build()
    g = np.load(os.path.join(GOLD, "hf_tiny-d64.npz"))
    spec = SPECS["tiny-d64"]
    w = init_weights(spec, seed=int(g["seed"]), std=float(g["std"]), norm_jitter=float(g["norm_jitter"]))
    tok = ByteTokenizer(spec.vocab)
    eng = Engine(spec, w, EngineLimits(max_batch=16, max_batched_tokens=512, max_context=2048), eos_ids=())
    runner = EngineRunner(eng, tok)
    h = NativeHandler(model="tiny-d64", runner=runner)
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    server = uvicorn.Server(uvicorn.Config(create_app(h), host="127.0.0.1", port=port, log_level="warning", lifespan="off"))
    server.install_signal_handlers = lambda: None
    th = threading.Thread(target=server.run, daemon=True)
    th.start()
    for _ in range(500):
        if server.started:
            break
        time.sleep(0.02)
    assert server.started
    yield h, eng, f"http://127.0.0.1:{port}/v1"
    server.should_exit = True
    th.join(timeout=5)
    runner.shutdown()
    eng.close()


def test_openai_stream_equals_handler_stream_and_reuses_kv(served):
    import openai
    h, eng, base = served
    msgs = [{"role": "system", "content": "be brief"}, {"role": "user", "content": "tell me about B200 " * 3}]
    direct = list(h.generate_stream(msgs, temperature=0.0, max_tokens=24, request_id="direct"))
    c = openai.OpenAI(base_url=base, api_key="x", timeout=60.0, max_retries=0)
    try:
        assert [m.id for m in c.models.list().data] == ["tiny-d64"]
        chunks = list(c.chat.completions.create(model="tiny-d64", messages=msgs, stream=True, temperature=0.0,
                                                max_tokens=24, stream_options={"include_usage": True}))
        wire = [x.choices[0].delta.content for x in chunks if x.choices and x.choices[0].delta.content]
        assert wire == direct
        fin = [x.choices[0].finish_reason for x in chunks if x.choices and x.choices[0].finish_reason]
        assert fin == ["length"] and chunks[-1].usage.completion_tokens == 24
        hits0 = eng.stats()["prefix_hit_tokens"]
        msgs2 = msgs + [{"role": "assistant", "content": "".join(wire)}, {"role": "user", "content": "go on"}]
        r = c.chat.completions.create(model="tiny-d64", messages=msgs2, temperature=0.0, max_tokens=8)
        assert r.usage.completion_tokens == 8 and r.choices[0].finish_reason == "length"
        assert eng.stats()["prefix_hit_tokens"] - hits0 >= chunks[-1].usage.prompt_tokens - 1
    finally:
        c.close()
    assert h.get_active_requests() == {} and eng.stats()["active_requests"] == 0


def test_ollama_ndjson_wire_over_the_real_engine(served):
    """§8f-3 on the GPU: the Ollama wire (`POST /api/chat`, NDJSON; ollama_handler.py:142-177, 268-296) answered by
    the same process over the real engine — raw NDJSON first, then, when the reference is on this box
    (baseline/_ref), the reference's UNCHANGED OllamaHandler as the client.  Both must give what NativeHandler
    yields for the conversation."""
    import json
    import sys
    import httpx
    import ws_harness as W
    h, eng, base = served
    root = base[:-3]
    msgs = [{"role": "system", "content": "be brief"}, {"role": "user", "content": "say something about NVLink"}]
    direct = list(h.generate_stream(msgs, temperature=0.0, max_tokens=20, top_p=0.9, top_k=40, request_id="ollama-direct"))
    assert httpx.get(root + "/").status_code == 200
    with httpx.stream("POST", root + "/api/chat", timeout=60.0, json={
            "model": "tiny-d64", "messages": msgs, "stream": True,
            "options": {"temperature": 0.0, "num_predict": 20, "top_p": 0.9, "top_k": 40}}) as r:
        assert r.status_code == 200
        lines = [json.loads(ln) for ln in r.iter_lines() if ln.strip()]
    assert [ln["done"] for ln in lines[:-1]] == [False] * (len(lines) - 1) and lines[-1]["done"] is True
    assert [ln["message"]["content"] for ln in lines[:-1]] == direct
    assert lines[-1]["done_reason"] == "length" and lines[-1]["eval_count"] == 20 and lines[-1]["prompt_eval_count"] > 0
    ref = W.find_reference()
    if ref is None:
        pytest.skip("raw NDJSON checked; no reference checkout for the unchanged OllamaHandler part")
    for m in [k for k in sys.modules if k == "app" or k.startswith("app.")]:
        del sys.modules[m]
    sys.path.insert(0, ref)
    try:
        from app.core.ollama_handler import OllamaHandler
        oh = OllamaHandler(base_url=root, model="tiny-d64")
        assert oh.check_connection() is True
        out = list(oh.generate_stream(msgs, temperature=0.0, max_tokens=20, top_p=0.9, top_k=40, request_id="oh-1"))
        assert out == direct
        oh.close()
    finally:
        sys.path.remove(ref)
    assert h.get_active_requests() == {} and eng.stats()["active_requests"] == 0
