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
