"""Boundary A over a real socket on CPU: the reference's unchanged WebSocketLLMServer, served by uvicorn,
NativeHandler swapped in, 16 concurrent `websockets` clients; engine = tests/fake_engine.py.  The GPU twin
(real engine, 64 clients) is tests/test_ws_gpu.py."""
import pytest

from fasttalk_llm_microservice_b200.handler import EngineRunner
from fasttalk_llm_microservice_b200.tokenizer import ByteTokenizer
from fake_engine import FakeEngine
import ws_harness as W

pytestmark = pytest.mark.skipif(W.find_reference() is None, reason="no reference checkout / baseline/_ref on this box")


def test_sixteen_websocket_sessions_two_turns_each():
    tok = ByteTokenizer()
    eng = FakeEngine(tok, reply=lambda toks: f"echo {len(toks)} tokens")
    runner = EngineRunner(eng, tok)
    srv = W.ReferenceServer(runner, env={"DEFAULT_MAX_TOKENS": "64"})
    try:
        specs = [{"system_prompt": f"system {i}", "turns": [f"hello {i}", "and again"]} for i in range(16)]
        res = W.run_clients(srv.port, specs)
    finally:
        srv.stop()
        runner.shutdown()
    by_sid = {}
    for s, toks, kw in eng.submitted:
        by_sid.setdefault(s, []).append((toks, kw))
    assert len(res) == 16 and len(by_sid) == 16
    for r in res:
        (t1, kw1), (t2, _) = by_sid[r["session_id"]]                 # request_id == session id (websocket_server_vllm.py:528-534)
        for turn, toks in zip(r["turns"], (t1, t2)):
            assert turn["end"]["type"] == "response_complete"
            assert turn["end"]["stats"]["tokens_generated"] == len(turn["chunks"])
            assert "".join(turn["chunks"]) == f"echo {len(toks)} tokens"
        assert t2[:len(t1)] == t1                                    # full history re-sent: turn 1's prompt is a prefix
        assert kw1["temperature"] == 0.0 and kw1["max_new_tokens"] == 64
    # the server called the handler with the reference's keyword set
    assert all(set(c["kw"]) >= {"temperature", "max_tokens", "request_id"} for c in srv.calls)
    m = W.turn_metrics(res, 1)
    assert m["sessions"] == 16 and m["tokens"] == sum(len(r["turns"][1]["chunks"]) for r in res) and m["tok_s"] > 0


def test_legacy_ollama_provider_path_of_the_unchanged_server():
    """LLM_PROVIDER=ollama: the same unchanged server takes its legacy branch (`_init_ollama_handler`,
    `_generate_with_ollama`, websocket_server_vllm.py:124-137, 543-600) and drives the SYNCHRONOUS `generate_stream` with
    Ollama's keyword set (top_k included); NativeHandler stands in for OllamaHandler."""
    tok = ByteTokenizer()
    eng = FakeEngine(tok, reply=lambda toks: f"legacy {len(toks)}")
    runner = EngineRunner(eng, tok)
    srv = W.ReferenceServer(runner, env={"DEFAULT_MAX_TOKENS": "32", "DEFAULT_TOP_K": "40"}, provider="ollama")
    try:
        res = W.run_clients(srv.port, [{"system_prompt": f"s{i}", "turns": [f"hi {i}", "again"]} for i in range(4)])
    finally:
        srv.stop()
        runner.shutdown()
    assert len(res) == 4 and all(c.get("sync") for c in srv.calls) and len(srv.calls) == 8
    assert all({"temperature", "max_tokens", "top_p", "top_k", "request_id"} <= set(c["kw"]) for c in srv.calls)
    by_sid = {}
    for s_, toks, kw in eng.submitted:
        by_sid.setdefault(s_, []).append((toks, kw))
    for r in res:
        (t1, kw1), (t2, _) = by_sid[r["session_id"]]
        assert ["".join(t["chunks"]) for t in r["turns"]] == [f"legacy {len(t1)}", f"legacy {len(t2)}"]
        assert all(t["end"]["type"] == "response_complete" and t["end"]["stats"]["provider"] == "ollama" for t in r["turns"])
        assert kw1["top_k"] == 40 and t2[:len(t1)] == t1
