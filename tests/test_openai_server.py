"""Boundary B on CPU: the OpenAI-compatible front (openai_server.py) served by uvicorn on
127.0.0.1 and driven by the real `openai` client — the same client the reference's
`VLLMHandler` wraps (`app/core/vllm_handler.py:53-62`) — and, when the reference checkout is
here, by the UNCHANGED `VLLMHandler` / `VLLMWithToolsHandler` themselves.  The engine is
tests/fake_engine.py (scripted replies); the GPU twin is tests/test_openai_server_gpu.py."""
import asyncio
import json
import os
import socket
import sys
import threading
import time

import httpx
import pytest

from fasttalk_llm_microservice_b200.chat_format import HermesToolParser, normalise_messages, ChatFormatError
from fasttalk_llm_microservice_b200.handler import EngineRunner, NativeHandler
from fasttalk_llm_microservice_b200.openai_server import create_app
from fasttalk_llm_microservice_b200.tokenizer import ByteTokenizer
from fake_engine import FakeEngine

REF = os.environ.get("FASTTALK_REFERENCE_PATH", "/root/reference")
HAVE_REF = os.path.exists(os.path.join(REF, "app", "core", "vllm_handler.py"))

WEATHER_TOOL = {"type": "function", "function": {
    "name": "get_weather", "description": "Weather for a city",
    "parameters": {"type": "object", "properties": {"city": {"type": "string"}}, "required": ["city"]}}}
CALL_TEXT = 'Let me check.\n<tool_call>\n{"name": "get_weather", "arguments": {"city": "Paris"}}\n</tool_call>'


def scripted(tok):
    """Reply depends on the prompt so one server serves every test."""
    def reply(tokens):
        text = tok.decode(tokens)
        if "<tool_response>" in text:
            return "It is sunny in Paris."
        if "<tools>" in text:
            return CALL_TEXT
        if "count" in text:
            return "one two three four five six seven eight nine ten"
        return "Hello from the fake engine."
    return reply


class Served:
    def __init__(self):
        import uvicorn
        self.tok = ByteTokenizer()
        self.eng = FakeEngine(self.tok, scripted(self.tok), max_context=4096)
        self.runner = EngineRunner(self.eng, self.tok)
        self.handler = NativeHandler(model="llama3.2:1b", runner=self.runner)
        self.app = create_app(self.handler, "llama3.2:1b")
        s = socket.socket()
        s.bind(("127.0.0.1", 0))
        self.port = s.getsockname()[1]
        s.close()
        cfg = uvicorn.Config(self.app, host="127.0.0.1", port=self.port, log_level="warning", lifespan="off")
        self.server = uvicorn.Server(cfg)
        self.server.install_signal_handlers = lambda: None
        self.thread = threading.Thread(target=self.server.run, daemon=True)
        self.thread.start()
        for _ in range(200):
            if self.server.started:
                break
            time.sleep(0.02)
        assert self.server.started
        self.base = f"http://127.0.0.1:{self.port}/v1"

    def stop(self):
        self.server.should_exit = True
        self.thread.join(timeout=5)
        self.runner.shutdown()


@pytest.fixture(scope="module")
def srv():
    s = Served()
    yield s
    s.stop()


@pytest.fixture()
def client(srv):
    import openai
    c = openai.OpenAI(base_url=srv.base, api_key="not-needed", timeout=20.0, max_retries=0)
    yield c
    c.close()


# ------------------------------------------------------------------ pure host logic
def test_hermes_parser_streams_content_and_calls_across_any_split():
    text = 'Sure.<tool_call>{"name": "a", "arguments": {"x": 1}}</tool_call>\n<tool_call>\n{"name": "b", "arguments": {}}\n</tool_call>'
    for step in (1, 2, 3, 7, 1000):
        p = HermesToolParser()
        ev = []
        for i in range(0, len(text), step):
            ev += p.feed(text[i:i + step])
        ev += p.flush()
        content = "".join(e[1] for e in ev if e[0] == "content")
        calls = [(e[1], e[3], json.loads(e[4])) for e in ev if e[0] == "tool_call"]
        assert content == "Sure." and calls == [(0, "a", {"x": 1}), (1, "b", {})], (step, ev)
    # text that only looks like a tag, a body that is not a call, and a call cut off by max_tokens
    p = HermesToolParser()
    assert "".join(e[1] for e in p.feed("a < b <tool and <tool_call>oops</tool_call> c") + p.flush()) == \
        "a < b <tool and <tool_call>oops</tool_call> c"
    p = HermesToolParser()
    ev = p.feed('<tool_call>{"name": "f", "arguments": {"k": "v"}}') + p.flush()
    assert [e[0] for e in ev] == ["tool_call"] and ev[0][3] == "f" and json.loads(ev[0][4]) == {"k": "v"}
    p = HermesToolParser(enabled=False)
    assert p.feed("<tool_call>x") == [("content", "<tool_call>x")] and p.flush() == []


def test_normalise_messages_tools_and_tool_turns():
    msgs = [{"role": "system", "content": "be brief"},
            {"role": "user", "content": [{"type": "text", "text": "weather "}, {"type": "text", "text": "in Paris?"}]},
            {"role": "assistant", "content": None, "tool_calls": [
                {"id": "c1", "type": "function", "function": {"name": "get_weather", "arguments": '{"city": "Paris"}'}}]},
            {"role": "tool", "tool_call_id": "c1", "content": "sunny"}]
    out = normalise_messages(msgs, [WEATHER_TOOL])
    assert [m["role"] for m in out] == ["system", "user", "assistant", "tool"]
    assert out[0]["content"].startswith("be brief\n\n") and "<tools>" in out[0]["content"] and "get_weather" in out[0]["content"]
    assert out[1]["content"] == "weather in Paris?"
    assert out[2]["content"] == '<tool_call>\n{"name": "get_weather", "arguments": {"city": "Paris"}}\n</tool_call>'
    assert out[3]["content"] == "<tool_response>\nsunny\n</tool_response>"
    # what the parser accepts is what history rendering reproduces (KV prefix reuse across the tool round)
    ev = HermesToolParser().feed(out[2]["content"])
    assert ev[0][3] == "get_weather" and json.loads(ev[0][4]) == {"city": "Paris"}
    assert normalise_messages([{"role": "user", "content": "x"}], [WEATHER_TOOL])[0]["role"] == "system"
    assert "<tools>" not in normalise_messages(msgs[:2], [WEATHER_TOOL], "none")[0]["content"]
    for bad in ([], None, [{"content": "no role"}], [{"role": "alien", "content": "x"}],
                [{"role": "user", "content": [{"type": "image_url", "image_url": {"url": "x"}}]}]):
        with pytest.raises(ChatFormatError):
            normalise_messages(bad)


# ------------------------------------------------------------------ the wire, through the openai client
def test_health_and_models(srv, client):
    assert httpx.get(srv.base.replace("/v1", "/health")).status_code == 200    # vllm_handler.py:84
    ids = [m.id for m in client.models.list().data]                            # vllm_handler.py:105-108
    assert ids == ["llama3.2:1b"]


def test_streaming_chunks_follow_the_openai_shape(srv, client):
    stream = client.chat.completions.create(model="llama3.2:1b", stream=True, temperature=0.0, max_tokens=64,
                                            top_p=0.9, messages=[{"role": "user", "content": "hi"}],
                                            stream_options={"include_usage": True})
    chunks = list(stream)
    assert chunks[0].choices[0].delta.role == "assistant" and chunks[0].choices[0].delta.content == ""
    body = [c for c in chunks[1:] if c.choices and c.choices[0].delta.content]
    assert "".join(c.choices[0].delta.content for c in body) == "Hello from the fake engine."
    assert len(body) == len("Hello from the fake engine.")        # one chunk per token (websocket_server_vllm.py:538)
    assert all(c.choices[0].finish_reason is None for c in body)
    fin = [c for c in chunks if c.choices and c.choices[0].finish_reason]
    assert len(fin) == 1 and fin[0].choices[0].finish_reason == "stop"
    usage = chunks[-1].usage
    assert chunks[-1].choices == [] and usage.completion_tokens == len(body)
    n_prompt = len(srv.tok.apply_chat_template([{"role": "user", "content": "hi"}]))
    assert usage.prompt_tokens == n_prompt and usage.total_tokens == n_prompt + len(body)
    assert len({c.id for c in chunks}) == 1 and all(c.object == "chat.completion.chunk" for c in chunks)
    sent = srv.eng.submitted[-1][2]
    assert sent["temperature"] == 0.0 and sent["top_p"] == pytest.approx(0.9) and sent["max_new_tokens"] == 64


def test_non_streaming_length_and_stop(srv, client):
    r = client.chat.completions.create(model="llama3.2:1b", messages=[{"role": "user", "content": "count"}],
                                       max_tokens=7)
    assert r.choices[0].message.content == "one two" and r.choices[0].finish_reason == "length"
    assert r.usage.completion_tokens == 7 and r.object == "chat.completion"
    r = client.chat.completions.create(model="llama3.2:1b", messages=[{"role": "user", "content": "count"}],
                                       stop=["four", "zzz"])
    assert r.choices[0].message.content == "one two three " and r.choices[0].finish_reason == "stop"
    r = client.chat.completions.create(model="llama3.2:1b", messages=[{"role": "user", "content": "count"}], stop="six")
    assert r.choices[0].message.content == "one two three four five "


def test_tool_call_deltas_accumulate_like_the_reference_does(srv, client):
    msgs = [{"role": "user", "content": "weather in Paris?"}]
    stream = client.chat.completions.create(model="llama3.2:1b", stream=True, messages=msgs, tools=[WEATHER_TOOL])
    # the accumulation loop of vllm_handler.py:394-408, restated
    calls, current, text, finish = {}, None, [], None
    for ch in stream:
        d = ch.choices[0].delta
        if d.tool_calls:
            for tc in d.tool_calls:
                if tc.id:
                    current = tc.id
                    calls.setdefault(current, {"name": None, "arguments": ""})
                if tc.function and current:
                    if tc.function.name:
                        calls[current]["name"] = tc.function.name
                    if tc.function.arguments:
                        calls[current]["arguments"] += tc.function.arguments
        elif d.content:
            text.append(d.content)
        finish = ch.choices[0].finish_reason or finish
    assert "".join(text) == "Let me check.\n" and finish == "tool_calls"
    (cid, call), = calls.items()
    assert cid.startswith("chatcmpl-tool-") and call["name"] == "get_weather"
    assert json.loads(call["arguments"]) == {"city": "Paris"}
    prompt = srv.tok.decode(srv.eng.submitted[-1][1])
    assert "<tools>" in prompt and '"get_weather"' in prompt
    # second round: the tool result goes back, the model answers in text
    msgs += [{"role": "assistant", "tool_calls": [{"id": cid, "type": "function", "function": {
                 "name": "get_weather", "arguments": call["arguments"]}}]},
             {"role": "tool", "tool_call_id": cid, "content": "sunny, 21C"}]
    r = client.chat.completions.create(model="llama3.2:1b", messages=msgs, tools=[WEATHER_TOOL])
    assert r.choices[0].message.content == "It is sunny in Paris." and r.choices[0].message.tool_calls is None
    prompt2 = srv.tok.decode(srv.eng.submitted[-1][1])
    assert "<tool_response>\nsunny, 21C\n</tool_response>" in prompt2 and "<tool_call>" in prompt2
    # non-streaming form of a tool call
    r = client.chat.completions.create(model="llama3.2:1b", messages=msgs[:1], tools=[WEATHER_TOOL])
    m = r.choices[0].message
    assert r.choices[0].finish_reason == "tool_calls" and m.tool_calls[0].function.name == "get_weather"
    assert json.loads(m.tool_calls[0].function.arguments) == {"city": "Paris"} and m.content == "Let me check.\n"
    # tool_choice="none": the tools stay out of the prompt and the text is not parsed
    r = client.chat.completions.create(model="llama3.2:1b", messages=msgs[:1], tools=[WEATHER_TOOL], tool_choice="none")
    assert r.choices[0].message.tool_calls is None


def test_request_errors_are_http_400(srv, client):
    import openai
    with pytest.raises(openai.BadRequestError):
        client.chat.completions.create(model="llama3.2:1b", messages=[])
    with pytest.raises(openai.BadRequestError):
        client.chat.completions.create(model="llama3.2:1b", messages=[{"role": "user", "content": "x"}], n=2)
    with pytest.raises(openai.BadRequestError):
        client.chat.completions.create(model="llama3.2:1b", messages=[{"role": "user", "content": "x"}], temperature=9)
    with pytest.raises(openai.BadRequestError) as ei:       # longer than the context window
        client.chat.completions.create(model="llama3.2:1b", messages=[{"role": "user", "content": "x" * 5000}])
    assert "context window" in str(ei.value)
    r = httpx.post(srv.base + "/chat/completions", content=b"{not json")
    assert r.status_code == 400 and "error" in r.json()
    assert srv.handler.get_active_requests() == {}


def test_turns_of_one_conversation_share_a_session_and_concurrent_ones_do_not(srv, client):
    import openai
    n0 = len(srv.eng.submitted)
    m1 = [{"role": "system", "content": "sys"}, {"role": "user", "content": "first"}]
    a1 = client.chat.completions.create(model="llama3.2:1b", messages=m1).choices[0].message.content
    m2 = m1 + [{"role": "assistant", "content": a1}, {"role": "user", "content": "second"}]
    client.chat.completions.create(model="llama3.2:1b", messages=m2)
    s1, s2 = srv.eng.submitted[n0][0], srv.eng.submitted[n0 + 1][0]
    assert s1 == s2                                   # same engine session → KV prefix reuse
    t1, t2 = srv.eng.submitted[n0][1], srv.eng.submitted[n0 + 1][1]
    assert t2[:len(t1)] == t1

    async def many():
        ac = openai.AsyncOpenAI(base_url=srv.base, api_key="x", timeout=20.0, max_retries=0)
        try:
            rs = await asyncio.gather(*[ac.chat.completions.create(model="llama3.2:1b", messages=m1) for _ in range(8)])
        finally:
            await ac.close()
        return rs
    n1 = len(srv.eng.submitted)
    rs = asyncio.run(many())
    assert all(r.choices[0].message.content == a1 for r in rs)
    assert len({s for s, _, _ in srv.eng.submitted[n1:]}) > 1       # overlapping requests got separate slots
    # an explicit session id wins over the content hash
    httpx.post(srv.base + "/chat/completions", headers={"X-Session-Id": "abc"},
               json={"model": "m", "messages": [{"role": "user", "content": "x"}]}).raise_for_status()
    assert srv.eng.submitted[-1][0] == "abc#0"


def test_client_disconnect_frees_the_request(srv):
    async def go():
        async with httpx.AsyncClient(timeout=10.0) as hc:
            async with hc.stream("POST", srv.base + "/chat/completions", json={
                    "model": "m", "stream": True, "messages": [{"role": "user", "content": "count"}]}) as r:
                assert r.status_code == 200 and r.headers["content-type"].startswith("text/event-stream")
                got = 0
                async for line in r.aiter_lines():
                    if line.startswith("data: "):
                        got += 1
                        if got == 3:
                            break           # leave mid-stream
    asyncio.run(go())
    for _ in range(300):
        if not srv.handler.get_active_requests() and srv.eng.stats()["active_requests"] == 0:
            break
        time.sleep(0.01)
    assert srv.handler.get_active_requests() == {} and srv.eng.stats()["active_requests"] == 0
    # ... and the conversation key it held: the next turn of the same conversation gets slot #0 again
    httpx.post(srv.base + "/chat/completions",
               json={"model": "m", "messages": [{"role": "user", "content": "count"}]}).raise_for_status()
    assert srv.eng.submitted[-1][0].endswith("#0")


def test_session_key_is_released_when_the_first_token_never_comes():
    """anything but LLMServiceError before the first token (a client that hangs up early raises
    CancelledError there) used to leave the key busy for good; later turns then got base#1, base#2, ...
    and lost their KV prefix.  Both wires."""
    seen = []

    class Broken:
        model = "m"

        def check_connection(self):
            return True

        async def generate_stream_async(self, messages, request_id=None, **kw):
            seen.append(request_id)
            raise RuntimeError("boom")
            yield ""                                  # pragma: no cover

    app = create_app(Broken(), "m")

    async def go():
        tr = httpx.ASGITransport(app=app, raise_app_exceptions=False)
        async with httpx.AsyncClient(transport=tr, base_url="http://t") as c:
            body = {"model": "m", "messages": [{"role": "user", "content": "hi"}]}
            for path in ("/v1/chat/completions", "/v1/chat/completions", "/api/chat", "/api/chat"):
                r = await c.post(path, json=body)
                assert r.status_code == 500
    asyncio.run(go())
    assert len(seen) == 4 and all(k.endswith("#0") for k in seen), seen


# ------------------------------------------------------------------ the unchanged reference client side
@pytest.mark.skipif(not HAVE_REF, reason="reference checkout not on this box")
def test_unchanged_reference_vllm_handler_talks_to_the_server(srv, monkeypatch):
    monkeypatch.syspath_prepend(REF)
    for m in [k for k in sys.modules if k == "app" or k.startswith("app.")]:
        monkeypatch.delitem(sys.modules, m)
    from app.core.vllm_handler import VLLMHandler, VLLMWithToolsHandler
    h = VLLMHandler(base_url=srv.base, model="llama3.2:1b")
    assert h.check_connection() is True
    assert h.get_model_info() == {"models": ["llama3.2:1b"], "current_model": "llama3.2:1b"}
    msgs = [{"role": "user", "content": "hi"}]
    out = list(h.generate_stream(msgs, temperature=0.0, max_tokens=32, top_p=0.9, request_id="r1"))
    assert "".join(out) == "Hello from the fake engine." and len(out) == len("Hello from the fake engine.")

    async def run_async():
        return [c async for c in h.generate_stream_async(messages=msgs, temperature=0.0, max_tokens=5, request_id="r2")]
    assert "".join(asyncio.run(run_async())) == "Hello"
    assert h.get_active_requests() == {}

    seen = []

    def get_weather(city):
        seen.append(city)
        return "sunny, 21C in " + city
    th = VLLMWithToolsHandler(base_url=srv.base, model="llama3.2:1b")
    text = "".join(th.generate_stream_with_tools([{"role": "user", "content": "weather in Paris?"}], [WEATHER_TOOL],
                                                 {"get_weather": get_weather}, temperature=0.0, max_tokens=200))
    assert seen == ["Paris"] and text == "Let me check.\nIt is sunny in Paris."


# ------------------------------------------------------------------ the Ollama wire on the same server
def test_ollama_ndjson_wire(srv):
    root = srv.base[:-3]
    assert httpx.get(root + "/").status_code == 200                          # ollama_handler.py:76-78
    assert "details" in httpx.post(root + "/api/show", json={"name": "llama3.2:1b"}).json()
    assert httpx.get(root + "/api/tags").json()["models"][0]["name"] == "llama3.2:1b"
    with httpx.stream("POST", root + "/api/chat", json={
            "model": "llama3.2:1b", "messages": [{"role": "user", "content": "count"}], "stream": True,
            "options": {"temperature": 0.0, "num_predict": 9, "top_p": 0.9, "top_k": 40}, "keep_alive": "5m"}) as r:
        assert r.status_code == 200
        lines = [json.loads(ln) for ln in r.iter_lines() if ln.strip()]
    assert [ln["done"] for ln in lines] == [False] * 9 + [True]
    assert "".join(ln["message"]["content"] for ln in lines) == "one two t"
    assert lines[-1]["done_reason"] == "length" and lines[-1]["eval_count"] == 9 and lines[-1]["prompt_eval_count"] > 0
    sent = srv.eng.submitted[-1][2]
    assert sent == {"temperature": 0.0, "top_k": 40, "top_p": pytest.approx(0.9), "max_new_tokens": 9}
    # Ollama's own sampling defaults when `options` is silent; non-streaming form; stop strings
    r = httpx.post(root + "/api/chat", json={"model": "m", "stream": False, "options": {"stop": ["three"]},
                                             "messages": [{"role": "user", "content": "count"}]}).json()
    assert r["message"]["content"] == "one two " and r["done"] is True and r["done_reason"] == "stop"
    sent = srv.eng.submitted[-1][2]
    assert (sent["temperature"], sent["top_k"], sent["top_p"]) == (pytest.approx(0.8), 40, pytest.approx(0.9))
    assert httpx.post(root + "/api/chat", json={"model": "m", "messages": []}).status_code == 400


@pytest.mark.skipif(not HAVE_REF, reason="reference checkout not on this box")
def test_unchanged_reference_ollama_handler_talks_to_the_server(srv, monkeypatch):
    monkeypatch.syspath_prepend(REF)
    for m in [k for k in sys.modules if k == "app" or k.startswith("app.")]:
        monkeypatch.delitem(sys.modules, m)
    from app.core.ollama_handler import OllamaHandler
    h = OllamaHandler(base_url=srv.base[:-3], model="llama3.2:1b")
    try:
        assert h.check_connection() is True
        assert h.get_model_info()["details"]["family"] == "llama"
        out = list(h.generate_stream([{"role": "user", "content": "hi"}], temperature=0.0, max_tokens=256,
                                     top_p=0.9, top_k=40, request_id="o1"))
        assert "".join(out) == "Hello from the fake engine." and len(out) == len("Hello from the fake engine.")
        assert list(h.get_active_requests()) == []
    finally:
        h.close()
