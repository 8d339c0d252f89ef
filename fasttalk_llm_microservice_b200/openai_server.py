"""OpenAI-compatible HTTP front over the in-process engine (boundary B).

With this app listening where the reference expects vLLM, nothing in the reference is patched:
`VLLM_BASE_URL=http://host:port/v1` and the unchanged `VLLMHandler`
(`app/core/vllm_handler.py:53-62` — `openai.OpenAI/AsyncOpenAI`), `VLLMWithToolsHandler`
(`:341-457`) and PydanticAI `VoiceAgent` (`app/agents/voice_agent.py:127-139`) talk to the B200
engine.  Endpoints they use:

  GET  /health                      `vllm_handler.py:84`, `voice_agent.py:325` (200 = ready)
  GET  /v1/models                   `vllm_handler.py:105-108` (`models.list()` → `.data[].id`)
  POST /v1/chat/completions         `vllm_handler.py:245-271` (`stream=True`, temperature,
                                    max_tokens, top_p, stop, tools) → SSE of
                                    `chat.completion.chunk`; `delta.content` per token
                                    (`:285-288`), `finish_reason` ends it (`:291-295`),
                                    `delta.tool_calls[*]` carry id / function.name /
                                    function.arguments (`:394-408`)

Run:  python -m fasttalk_llm_microservice_b200.openai_server --model llama3.2:1b --port 8000
"""
from __future__ import annotations

import argparse
import hashlib
import json
import os
import threading
import time
import uuid
from typing import Any, AsyncGenerator, Dict, List, Optional

from starlette.applications import Starlette
from starlette.requests import Request
from starlette.background import BackgroundTask
from starlette.responses import JSONResponse, StreamingResponse
from starlette.routing import Route

from .chat_format import ChatFormatError, HermesToolParser, normalise_messages
from .errors import ErrorCategory, LLMServiceError
from .handler import NativeHandler


def _error(status: int, message: str, etype: str = "invalid_request_error", param: Optional[str] = None,
           code: Optional[str] = None) -> JSONResponse:
    return JSONResponse({"error": {"message": message, "type": etype, "param": param, "code": code}},
                        status_code=status)


async def _abandon(agen, keys: "_SessionKeys", key: str) -> None:
    """Close a generation and give its conversation key back (idempotent on both counts)."""
    try:
        await agen.aclose()
    except BaseException:  # noqa: BLE001 - closing must not mask the error being propagated
        pass
    keys.release(key)


class _SessionKeys:
    """The wire is stateless but the engine's KV reuse is per session.  A conversation is
    recognised by its opening (system prompt + first user turn, or an explicit `user` /
    `X-Session-Id`); a second request for a key that is still generating gets its own slot."""

    def __init__(self):
        self.lock = threading.Lock()
        self.busy: Dict[str, int] = {}

    def acquire(self, body: Dict[str, Any], messages: List[Dict[str, str]], header: Optional[str]) -> str:
        base = header or body.get("user")
        if not base:
            h = hashlib.sha1()
            for m in messages[:2]:
                h.update(m["role"].encode() + b"\0" + m["content"].encode() + b"\0")
            base = "conv-" + h.hexdigest()
        with self.lock:
            n = 0
            while f"{base}#{n}" in self.busy:
                n += 1
            key = f"{base}#{n}"
            self.busy[key] = 1
            return key

    def release(self, key: str):
        with self.lock:
            self.busy.pop(key, None)


def _sampling(body: Dict[str, Any]) -> Dict[str, Any]:
    def num(name, lo, hi, cast=float):
        v = body.get(name)
        if v is None:
            return None
        try:
            v = cast(v)
        except (TypeError, ValueError):
            raise ChatFormatError(f"{name} must be a number", name)
        if not (lo <= v <= hi):
            raise ChatFormatError(f"{name} must be in [{lo}, {hi}]", name)
        return v
    if body.get("n") not in (None, 1):
        raise ChatFormatError("only n=1 is supported", "n")
    stop = body.get("stop")
    if isinstance(stop, str):
        stop = [stop]
    if stop is not None and (not isinstance(stop, list) or not all(isinstance(s, str) for s in stop)):
        raise ChatFormatError("stop must be a string or a list of strings", "stop")
    max_tokens = num("max_completion_tokens", 1, 1 << 30, int) or num("max_tokens", 1, 1 << 30, int)
    top_k = num("top_k", -1, 1 << 30, int)              # vLLM extension; -1 / 0 = off
    return {"temperature": num("temperature", 0.0, 2.0), "top_p": num("top_p", 0.0, 1.0),
            "top_k": top_k if top_k and top_k > 0 else None, "max_tokens": max_tokens, "stop": stop or None,
            "seed": num("seed", -(1 << 63), 1 << 63, int)}


def create_app(handler: NativeHandler, served_model: Optional[str] = None) -> Starlette:
    served = served_model or handler.model
    keys = _SessionKeys()
    created = int(time.time())

    async def health(request: Request):
        ok = handler.check_connection()
        return JSONResponse({"status": "ok" if ok else "unavailable"}, status_code=200 if ok else 503)

    async def models(request: Request):
        return JSONResponse({"object": "list", "data": [
            {"id": served, "object": "model", "created": created, "owned_by": "fasttalk_b200", "root": served,
             "max_model_len": getattr(getattr(handler.runner.engine, "limits", None), "max_context", None)}]})

    async def chat(request: Request):
        try:
            body = await request.json()
            if not isinstance(body, dict):
                raise ValueError("body must be a JSON object")
        except Exception as e:  # noqa: BLE001
            return _error(400, f"invalid JSON body: {e}")
        try:
            tools, tool_choice = body.get("tools") or None, body.get("tool_choice")
            messages = normalise_messages(body.get("messages"), tools, tool_choice)
            samp = _sampling(body)
        except ChatFormatError as e:
            return _error(400, str(e), param=e.param)
        use_tools = bool(tools) and tool_choice != "none"
        model = body.get("model") or served
        stream = bool(body.get("stream"))
        include_usage = bool((body.get("stream_options") or {}).get("include_usage"))
        cid = "chatcmpl-" + uuid.uuid4().hex
        key = keys.acquire(body, messages, request.headers.get("x-session-id"))
        meta: Dict[str, Any] = {}

        def chunk(delta: Dict[str, Any], finish: Optional[str] = None, usage=None, no_choice=False) -> bytes:
            obj = {"id": cid, "object": "chat.completion.chunk", "created": created, "model": model,
                   "choices": [] if no_choice else [{"index": 0, "delta": delta, "logprobs": None,
                                                     "finish_reason": finish}]}
            if usage is not None:
                obj["usage"] = usage
            return b"data: " + json.dumps(obj, ensure_ascii=False, separators=(",", ":")).encode() + b"\n\n"

        def usage_of() -> Dict[str, int]:
            p, c = int(meta.get("prompt_tokens", 0)), int(meta.get("completion_tokens", 0))
            return {"prompt_tokens": p, "completion_tokens": c, "total_tokens": p + c}

        # start the generation before answering so that request errors become HTTP errors
        agen = handler.generate_stream_async(messages, temperature=samp["temperature"], max_tokens=samp["max_tokens"],
                                             top_p=samp["top_p"], top_k=samp["top_k"], stop=samp["stop"],
                                             request_id=key, seed=samp["seed"], _meta=meta)
        first: Optional[str] = None
        try:
            first = await agen.__anext__()
        except StopAsyncIteration:
            first = None
        except LLMServiceError as e:
            keys.release(key)
            bad = e.category == ErrorCategory.VALIDATION
            return _error(400 if bad else 500, e.message if hasattr(e, "message") else str(e),
                          "invalid_request_error" if bad else "internal_error")
        except BaseException:
            # anything else — CancelledError when the client hangs up before the first token included —
            # must not leave the conversation key marked busy (later turns would get base#1, base#2, ...
            # and lose their KV prefix) nor the generation running
            await _abandon(agen, keys, key)
            raise

        async def events() -> AsyncGenerator[tuple, None]:
            """content / tool_call events of this completion, then ('end', finish_reason)."""
            parser = HermesToolParser(enabled=use_tools)
            try:
                if first is not None:
                    for ev in parser.feed(first):
                        yield ev
                    async for piece in agen:
                        for ev in parser.feed(piece):
                            yield ev
                for ev in parser.flush():
                    yield ev
                await agen.aclose()             # fills meta
                yield ("end", "tool_calls" if parser.n_calls else meta.get("finish_reason", "stop"))
            finally:
                await agen.aclose()
                keys.release(key)

        if not stream:
            text, calls, finish = [], [], "stop"
            try:
                async for ev in events():
                    if ev[0] == "content":
                        text.append(ev[1])
                    elif ev[0] == "tool_call":
                        calls.append({"id": ev[2], "type": "function", "function": {"name": ev[3], "arguments": ev[4]}})
                    else:
                        finish = ev[1]
            except LLMServiceError as e:
                return _error(500, str(e), "internal_error")
            msg: Dict[str, Any] = {"role": "assistant", "content": "".join(text) if (text or not calls) else None}
            if calls:
                msg["tool_calls"] = calls
            return JSONResponse({"id": cid, "object": "chat.completion", "created": created, "model": model,
                                 "choices": [{"index": 0, "message": msg, "logprobs": None, "finish_reason": finish}],
                                 "usage": usage_of()})

        async def sse() -> AsyncGenerator[bytes, None]:
            yield chunk({"role": "assistant", "content": ""})
            try:
                async for ev in events():
                    if ev[0] == "content":
                        yield chunk({"content": ev[1]})
                    elif ev[0] == "tool_call":
                        _, idx, call_id, name, args = ev
                        yield chunk({"tool_calls": [{"index": idx, "id": call_id, "type": "function",
                                                     "function": {"name": name, "arguments": ""}}]})
                        yield chunk({"tool_calls": [{"index": idx, "function": {"arguments": args}}]})
                    else:
                        yield chunk({}, finish=ev[1])
                        if include_usage:
                            yield chunk({}, usage=usage_of(), no_choice=True)
            except LLMServiceError as e:        # mid-stream failure: the wire has no status left to change
                yield b"data: " + json.dumps({"error": {"message": str(e), "type": "internal_error"}}).encode() + b"\n\n"
            yield b"data: [DONE]\n\n"

        # the background task runs after the response is over, also when the body was never iterated
        # (client gone before the first byte): the generation is closed and the key released exactly once
        return StreamingResponse(sse(), media_type="text/event-stream",
                                 headers={"Cache-Control": "no-cache", "X-Accel-Buffering": "no"},
                                 background=BackgroundTask(_abandon, agen, keys, key))

    from .ollama_front import ollama_routes
    return Starlette(routes=[
        Route("/health", health, methods=["GET"]),
        Route("/v1/models", models, methods=["GET"]),
        Route("/v1/chat/completions", chat, methods=["POST"]),
    ] + ollama_routes(handler, served, keys))     # the same process answers Ollama's wire (§8f-3)


def main(argv=None):
    ap = argparse.ArgumentParser(description="OpenAI-compatible server over the fasttalk B200 engine")
    ap.add_argument("--model", default=os.environ.get("VLLM_MODEL", os.environ.get("FASTTALK_MODEL", "llama3.2:1b")))
    ap.add_argument("--served-model-name", default=None)
    ap.add_argument("--host", default="0.0.0.0")
    ap.add_argument("--port", type=int, default=8000)
    ap.add_argument("--gpu", type=int, default=None, help="CUDA device of this replica (one process per GPU)")
    a = ap.parse_args(argv)
    if a.gpu is not None:
        os.environ["FASTTALK_GPU"] = str(a.gpu)
    import uvicorn
    handler = NativeHandler(model=a.model)
    if not handler.check_connection():
        raise SystemExit("fasttalk_b200: engine did not start (no CUDA device or extension not built)")
    uvicorn.run(create_app(handler, a.served_model_name or a.model), host=a.host, port=a.port, log_level="info")


if __name__ == "__main__":
    main()
