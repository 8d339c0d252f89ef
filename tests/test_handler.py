"""Host logic of boundary A on CPU: tokenizer, chat template, NativeHandler streaming /
cancel / stop / errors (engine replaced by tests/fake_engine.py), signature parity with the
reference handlers, and the UNCHANGED reference WebSocket server driving NativeHandler."""
import asyncio
import inspect
import os
import sys

import pytest

from fasttalk_llm_microservice_b200.errors import LLMServiceError
from fasttalk_llm_microservice_b200.handler import EngineRunner, NativeHandler
from fasttalk_llm_microservice_b200.tokenizer import ByteTokenizer
from fake_engine import FakeEngine, FakeLoopEngine

REF = os.environ.get("FASTTALK_REFERENCE_PATH", "/root/reference")
HAVE_REF = os.path.exists(os.path.join(REF, "app", "core", "vllm_handler.py"))


def make_handler(reply="Hello from the fake engine.", **kw):
    tok = ByteTokenizer()
    eng = FakeEngine(tok, reply, **kw)
    h = NativeHandler(base_url="http://ignored/v1", model="llama3.2:1b", runner=EngineRunner(eng, tok))
    return h, eng, tok


# ------------------------------------------------------------------ tokenizer
def test_byte_tokenizer_roundtrip_and_template():
    tok = ByteTokenizer()
    text = "héllo wörld ✓ 你好"
    ids = tok.encode(text)
    assert all(0 <= i < 256 for i in ids) and tok.decode(ids) == text
    # ids >= 256 are a bijection with private-use code points: generated text re-encodes exactly
    gen = [300, 65, 127999, 70000, 66, 128100]
    s = tok.decode(gen)
    assert tok.encode(s) == gen and len(s) == len(gen)
    msgs = [{"role": "system", "content": "be brief"}, {"role": "user", "content": "hi"}]
    t = tok.apply_chat_template(msgs)
    assert t[0] == 128000 and t.count(128009) == 2 and t.count(128006) == 3
    assert bytes(t[-2:]) == b"\n\n" and t[-3] == 128007
    # a longer conversation keeps the earlier one as an exact token prefix (KV reuse relies on it)
    t2 = tok.apply_chat_template(msgs + [{"role": "assistant", "content": s}, {"role": "user", "content": "more"}])
    assert t2[:len(t)] == t
    assert t2[len(t):len(t) + len(gen)] == gen


def test_stream_decoder_multibyte():
    tok = ByteTokenizer()
    dec = tok.stream_decoder()
    pieces = [dec.push(b) for b in "a✓b".encode()]
    assert pieces == ["a", "", "", "✓", "b"]
    assert dec.push(tok.eot) == ""
    small = ByteTokenizer(1024)
    assert small.eot == 1023 and small.decode(small.encode("ok")) == "ok"


# ------------------------------------------------------------------ handler
def test_sync_stream_one_chunk_per_token():
    h, eng, tok = make_handler()
    msgs = [{"role": "user", "content": "hi"}]
    out = list(h.generate_stream(msgs, temperature=0.0, max_tokens=64, top_p=0.9, top_k=40, request_id="sess-1"))
    assert "".join(out) == "Hello from the fake engine."
    assert len(out) == len("Hello from the fake engine.")            # websocket_server_vllm.py:538 counts chunks
    sess, tokens, kw = eng.submitted[0]
    assert sess == "sess-1" and tokens == tok.apply_chat_template(msgs)
    assert kw == dict(temperature=0.0, top_k=40, top_p=0.9, max_new_tokens=64)
    assert h.get_active_requests() == {}


def test_async_stream_concurrent_sessions():
    h, eng, tok = make_handler(reply=lambda toks: f"reply-{len(toks)}")

    async def one(i):
        msgs = [{"role": "user", "content": "x" * i}]
        return "".join([c async for c in h.generate_stream_async(msgs, max_tokens=50, request_id=f"s{i}")])

    async def main():
        return await asyncio.gather(*[one(i) for i in range(1, 9)])

    outs = asyncio.run(main())
    for i, o in enumerate(outs, start=1):
        assert o == f"reply-{len(tok.apply_chat_template([{'role': 'user', 'content': 'x' * i}]))}"


def test_max_tokens_and_stop_sequences():
    h, eng, tok = make_handler(reply="one two three four")
    assert "".join(h.generate_stream([{"role": "user", "content": "q"}], max_tokens=5)) == "one t"
    assert "".join(h.generate_stream([{"role": "user", "content": "q"}], stop=["three"])) == "one two "
    assert h.get_active_requests() == {}


def test_cancel_and_generator_close_release_the_request():
    h, eng, tok = make_handler(reply="x" * 1000)
    gen = h.generate_stream([{"role": "user", "content": "q"}], request_id="c1")
    assert next(gen) == "x"
    assert list(h.get_active_requests()) == ["c1"]
    assert h.cancel_generation("c1") is True
    rest = list(gen)
    assert len(rest) < 999 and h.get_active_requests() == {}
    assert h.cancel_generation("c1") is False

    async def disconnect():
        agen = h.generate_stream_async([{"role": "user", "content": "q"}], request_id="c2")
        assert await agen.__anext__() == "x"
        await agen.aclose()                     # what a WebSocket disconnect does to the generator
    asyncio.run(disconnect())
    assert h.get_active_requests() == {} and not eng.reqs


def test_errors_follow_the_reference_convention():
    h, eng, tok = make_handler(max_context=32)
    with pytest.raises(LLMServiceError) as ei:
        list(h.generate_stream([{"role": "user", "content": "y" * 100}]))
    d = ei.value.to_dict()
    assert set(d) == {"message", "category", "severity", "recoverable", "retry_after"}

    def boom(*a, **k):
        raise RuntimeError("engine exploded")
    eng.submit = boom
    with pytest.raises(LLMServiceError) as ei:
        list(h.generate_stream([{"role": "user", "content": "hi"}]))
    assert ei.value.category.value == "processing" and ei.value.severity.value == "high"
    assert h.check_connection() is True
    info = h.get_model_info()
    assert info["current_model"] == "llama3.2:1b" and info["models"] == ["llama3.2:1b"]


# ------------------------------------------------------------------ against the reference itself
@pytest.mark.skipif(not HAVE_REF, reason="reference checkout not on this box")
def test_signature_parity_with_reference_handlers():
    sys.path.insert(0, REF)
    try:
        from app.core.ollama_handler import OllamaHandler
        from app.core.vllm_handler import VLLMHandler
    finally:
        sys.path.remove(REF)

    def params(f):
        return [p for p in inspect.signature(f).parameters if p != "self"]

    for name in ("__init__", "generate_stream", "generate_stream_async", "check_connection", "get_model_info",
                 "cancel_generation", "get_active_requests"):
        theirs, ours = params(getattr(VLLMHandler, name)), params(getattr(NativeHandler, name))
        assert set(theirs) <= set(ours), (name, theirs, ours)
    for name in ("__init__", "generate_stream", "check_connection", "get_model_info", "cancel_generation",
                 "get_active_requests", "cleanup_stale_requests", "close"):
        theirs, ours = params(getattr(OllamaHandler, name)), params(getattr(NativeHandler, name))
        assert set(theirs) <= set(ours), (name, theirs, ours)
    assert inspect.isasyncgenfunction(NativeHandler.generate_stream_async)
    assert inspect.isgeneratorfunction(NativeHandler.generate_stream)


@pytest.mark.skipif(not HAVE_REF, reason="reference checkout not on this box")
def test_unchanged_reference_server_drives_native_handler(monkeypatch):
    """WebSocketLLMServer (websocket_server_vllm.py) is imported unmodified from the reference
    checkout; only `app.core.vllm_handler.VLLMHandler` is swapped before construction."""
    monkeypatch.syspath_prepend(REF)
    monkeypatch.setenv("LLM_PROVIDER", "vllm")
    monkeypatch.setenv("ENABLE_PYDANTIC_AI", "false")
    monkeypatch.setenv("DEFAULT_TEMPERATURE", "0.0")
    monkeypatch.setenv("DEFAULT_MAX_TOKENS", "64")
    monkeypatch.setenv("LLM_MAX_CONNECTIONS", "64")
    for m in [k for k in sys.modules if k == "app" or k.startswith("app.")]:
        monkeypatch.delitem(sys.modules, m)
    from starlette.testclient import TestClient
    import app.core.vllm_handler as vh
    from app.utils.config import Config

    tok = ByteTokenizer()
    eng = FakeEngine(tok, reply=lambda toks: f"echo {len(toks)} tokens")
    runner = EngineRunner(eng, tok)
    made = []

    class Patched(NativeHandler):
        def __init__(self, **kw):
            super().__init__(runner=runner, **kw)
            made.append(kw)

    monkeypatch.setattr(vh, "VLLMHandler", Patched)
    from app.core.websocket_server_vllm import WebSocketLLMServer
    server = WebSocketLLMServer(Config())
    assert set(made[0]) == {"base_url", "model", "api_key", "timeout"}
    with TestClient(server.app) as client:
        assert client.get("/health").status_code == 200
        with client.websocket_connect("/ws/llm") as ws:
            started = ws.receive_json()
            assert started["type"] == "session_started"
            sid = started["session_id"]
            ws.send_json({"type": "start_session", "config": {"system_prompt": "be brief"}})
            assert ws.receive_json()["type"] == "session_configured"
            replies = []
            for turn in ("hello", "and again"):
                ws.send_json({"type": "user_message", "text": turn})
                chunks = []
                while True:
                    msg = ws.receive_json()
                    if msg["type"] == "token":
                        chunks.append(msg["data"])
                    else:
                        break
                assert msg["type"] == "response_complete", msg
                assert msg["stats"]["tokens_generated"] == len(chunks)
                replies.append("".join(chunks))
            ws.send_json({"type": "end_session"})
            assert ws.receive_json()["type"] == "session_ended"
    # both turns reached the engine under the session id, with the default sampling knobs
    assert [s for s, _, _ in eng.submitted] == [sid, sid]
    t1, t2 = eng.submitted[0][1], eng.submitted[1][1]
    assert replies[0] == f"echo {len(t1)} tokens"
    # turn 2 re-sends the history: turn 1's prompt (minus the generation header) is its prefix
    assert t2[:len(t1)] == t1
    assert eng.submitted[0][2]["temperature"] == 0.0 and eng.submitted[0][2]["max_new_tokens"] == 64


def test_async_generation_times_out_when_the_engine_goes_quiet():
    """One idle watchdog per generation (not a wait_for per token): a stream that gets no token for `timeout`
    seconds fails with LLMServiceError like the reference's HTTP timeout does, and its batch slot is freed;
    a stream that keeps receiving tokens is never cut, however long it runs."""
    import time as _time
    tok = ByteTokenizer()

    class Stalling(FakeEngine):
        def step(self):
            _time.sleep(0.01)
            return 0

    eng = Stalling(tok)
    runner = EngineRunner(eng, tok)
    h = NativeHandler(model="m", runner=runner, timeout=0.3)

    async def go():
        t0 = _time.perf_counter()
        with pytest.raises(LLMServiceError):
            async for _ in h.generate_stream_async([{"role": "user", "content": "hi"}], request_id="quiet"):
                pass
        return _time.perf_counter() - t0
    took = asyncio.run(go())
    assert 0.25 <= took < 2.0, took
    assert h.get_active_requests() == {} and eng.stats()["active_requests"] == 0
    runner.shutdown()

    class Slow(FakeEngine):                       # a token every 50 ms for ~1 s: longer than the timeout in total
        def step(self):
            _time.sleep(0.05)
            return super().step()

    eng2 = Slow(tok, reply="abcdefghijklmnopqrst")
    runner2 = EngineRunner(eng2, tok)
    h2 = NativeHandler(model="m", runner=runner2, timeout=0.3)

    async def go2():
        return [c async for c in h2.generate_stream_async([{"role": "user", "content": "hi"}], request_id="slow")]
    assert "".join(asyncio.run(go2())) == "abcdefghijklmnopqrst"
    runner2.shutdown()


def test_runner_leaves_the_stepping_to_an_engine_that_steps_itself():
    """Engines with a stepping thread of their own (ft_engine_start) are never stepped from Python: the runner waits for
    events, drains and dispatches; paused() lends the engine out (its thread stopped) and starts it again; shutdown
    stops it."""
    tok = ByteTokenizer()
    eng = FakeLoopEngine(tok, "Stepped by the engine itself.")
    runner = EngineRunner(eng, tok)
    h = NativeHandler(base_url="http://ignored/v1", model="llama3.2:1b", runner=runner)
    msgs = [{"role": "user", "content": "hi"}]
    assert "".join(h.generate_stream(msgs, request_id="a")) == "Stepped by the engine itself."

    async def many():
        async def one(i):
            return "".join([c async for c in h.generate_stream_async(msgs, request_id=f"s{i}")])
        return await asyncio.gather(*[one(i) for i in range(8)])
    assert asyncio.run(many()) == ["Stepped by the engine itself."] * 8
    assert eng.loop_on and eng.stepped_by_owner == 0 and eng.starts == 1
    with runner.paused() as lent:
        assert lent is eng and not eng.loop_on
        rid = eng.submit("direct", tok.encode("x"), max_new_tokens=3)
        while not any(r == rid and fin for r, t, fin in eng.drain()):
            eng.step()                                   # the borrower steps it
    assert eng.loop_on and eng.starts == 2
    assert "".join(h.generate_stream(msgs, request_id="b")) == "Stepped by the engine itself."
    runner.shutdown()
    assert not eng.loop_on and eng.stepped_by_owner == 0


def test_native_loop_can_be_switched_off(monkeypatch):
    monkeypatch.setenv("FASTTALK_NATIVE_LOOP", "0")
    tok = ByteTokenizer()
    eng = FakeLoopEngine(tok, "Stepped from Python.")
    runner = EngineRunner(eng, tok)
    h = NativeHandler(base_url="http://ignored/v1", model="llama3.2:1b", runner=runner)
    assert "".join(h.generate_stream([{"role": "user", "content": "hi"}], request_id="a")) == "Stepped from Python."
    assert eng.starts == 0 and eng.steps > 0
    runner.shutdown()


class _SlowTokenizer(ByteTokenizer):
    """Stands for a tokenizer.json: its encoder runs without the GIL (here: sleeps) for a while on long histories."""
    releases_gil = True

    def __init__(self, delay):
        super().__init__()
        self.delay, self.threads = delay, set()

    def apply_chat_template(self, messages, add_generation_prompt=True):
        import threading
        import time
        self.threads.add(threading.current_thread().name)
        time.sleep(self.delay)
        return super().apply_chat_template(messages, add_generation_prompt)


def test_long_histories_are_tokenised_off_the_event_loop():
    """With a tokenizer that releases the GIL a burst of long conversations is tokenised on worker threads, in parallel,
    and the loop stays free; short ones stay on the loop; a caller that goes away mid-tokenisation leaves nothing behind."""
    import time
    tok = _SlowTokenizer(0.05)
    eng = FakeEngine(tok, "ok")
    h = NativeHandler(base_url="http://ignored/v1", model="llama3.2:1b", runner=EngineRunner(eng, tok))
    long_msgs = [{"role": "user", "content": "x" * 3000}]

    async def burst():
        ticks = []

        async def heartbeat():
            while True:
                ticks.append(time.monotonic())
                await asyncio.sleep(0.005)
        hb = asyncio.ensure_future(heartbeat())

        async def one(i):
            return "".join([c async for c in h.generate_stream_async(long_msgs, request_id=f"b{i}")])
        t0 = time.monotonic()
        outs = await asyncio.gather(*[one(i) for i in range(8)])
        wall = time.monotonic() - t0
        hb.cancel()
        return outs, wall, max(b - a for a, b in zip(ticks, ticks[1:]))
    outs, wall, worst_gap = asyncio.run(burst())
    assert outs == ["ok"] * 8
    assert wall < 8 * 0.05 * 0.75, wall                     # 8 x 50 ms in parallel, not one after the other
    assert worst_gap < 0.04, worst_gap                      # the loop kept running while they were tokenised
    assert all(n.startswith("fasttalk-b200-tokenise") for n in tok.threads), tok.threads
    tok.threads.clear()
    assert "".join(h.generate_stream([{"role": "user", "content": "hi"}], request_id="s")) == "ok"

    async def short_one():
        return "".join([c async for c in h.generate_stream_async([{"role": "user", "content": "hi"}], request_id="t")])
    assert asyncio.run(short_one()) == "ok"
    assert not any(n.startswith("fasttalk-b200-tokenise") for n in tok.threads), tok.threads

    async def goes_away():
        agen = h.generate_stream_async(long_msgs, request_id="gone")
        task = asyncio.ensure_future(agen.__anext__())
        await asyncio.sleep(0.01)                           # tokenisation is under way on a worker thread
        task.cancel()
        with pytest.raises(asyncio.CancelledError):
            await task
        await asyncio.sleep(0.15)                           # _start returns, the request it submitted is released
    asyncio.run(goes_away())
    assert h.get_active_requests() == {}
    h.runner.shutdown()
