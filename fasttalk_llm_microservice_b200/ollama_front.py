"""Ollama-compatible NDJSON routes over the in-process engine.

With these routes listening where the reference expects Ollama (`OLLAMA_BASE_URL`), the unchanged
`OllamaHandler` (`app/core/ollama_handler.py`) drives the B200 engine:

  GET  /            `ollama_handler.py:76-78` (`check_connection`: any 2xx)
  POST /api/show    `ollama_handler.py:98-102` (`get_model_info`: the JSON is returned as is)
  GET  /api/tags    model list (not called by the reference; Ollama clients probe it)
  POST /api/chat    `ollama_handler.py:142-177`: `{model, messages, stream, options{temperature,
                    num_predict, top_p, top_k, stop}, keep_alive}` → one JSON object per line,
                    `message.content` per token (`:287-290`), `"done": true` ends it (`:293-296`),
                    `{"error": …}` reports a failure (`:277-285`)
"""
from __future__ import annotations

import datetime
import json
import time
from typing import Any, AsyncGenerator, Dict, List

from starlette.background import BackgroundTask
from starlette.requests import Request
from starlette.responses import JSONResponse, PlainTextResponse, StreamingResponse
from starlette.routing import Route

from .chat_format import ChatFormatError, normalise_messages
from .errors import LLMServiceError

# Ollama's documented sampling defaults, used when `options` leaves a knob out
OLLAMA_DEFAULTS = {"temperature": 0.8, "top_k": 40, "top_p": 0.9}


def _now() -> str:
    return datetime.datetime.now(datetime.timezone.utc).isoformat().replace("+00:00", "Z")


def ollama_routes(handler, served: str, keys) -> List[Route]:
    from .openai_server import _abandon      # close a generation + release its conversation key (idempotent)

    async def root(request: Request):
        ok = handler.check_connection()
        return PlainTextResponse("Ollama is running" if ok else "engine unavailable", status_code=200 if ok else 503)

    def model_card() -> Dict[str, Any]:
        eng = handler.runner.engine
        spec = getattr(eng, "spec", None)
        info = {"general.architecture": "llama"}
        if spec is not None:
            info.update({"llama.block_count": spec.layers, "llama.embedding_length": spec.hidden,
                         "llama.attention.head_count": spec.n_q_heads, "llama.attention.head_count_kv": spec.n_kv_heads,
                         "llama.context_length": getattr(getattr(eng, "limits", None), "max_context", None),
                         "general.parameter_count": spec.param_count()})
        return {"modelfile": f"FROM {served}", "parameters": "", "template": "llama3",
                "details": {"family": "llama", "format": "fasttalk_b200", "quantization_level": "BF16"},
                "model_info": info}

    async def show(request: Request):
        return JSONResponse(model_card())

    async def tags(request: Request):
        return JSONResponse({"models": [{"name": served, "model": served, "modified_at": _now(),
                                         "details": model_card()["details"]}]})

    async def chat(request: Request):
        try:
            body = await request.json()
            messages = normalise_messages(body.get("messages"), body.get("tools") or None)
            opt = body.get("options") or {}
            if not isinstance(opt, dict):
                raise ChatFormatError("options must be an object")
            stop = opt.get("stop")
            stop = [stop] if isinstance(stop, str) else stop
            num_predict = opt.get("num_predict")
            kw = dict(temperature=float(opt.get("temperature", OLLAMA_DEFAULTS["temperature"])),
                      top_p=float(opt.get("top_p", OLLAMA_DEFAULTS["top_p"])),
                      top_k=int(opt.get("top_k", OLLAMA_DEFAULTS["top_k"])),
                      max_tokens=int(num_predict) if num_predict is not None and int(num_predict) > 0 else None,
                      stop=stop or None, seed=opt.get("seed"))
        except ChatFormatError as e:
            return JSONResponse({"error": str(e)}, status_code=400)
        except Exception as e:  # noqa: BLE001
            return JSONResponse({"error": f"invalid request: {e}"}, status_code=400)
        model = body.get("model") or served
        stream = body.get("stream", True) is not False
        key = keys.acquire(body, messages, request.headers.get("x-session-id"))
        meta: Dict[str, Any] = {}
        t0 = time.perf_counter_ns()
        agen = handler.generate_stream_async(messages, request_id=key, _meta=meta, **kw)
        try:
            first = await agen.__anext__()
        except StopAsyncIteration:
            first = None
        except LLMServiceError as e:
            keys.release(key)
            return JSONResponse({"error": str(e)}, status_code=400)
        except BaseException:               # client gone before the first token, or any other failure
            await _abandon(agen, keys, key)
            raise
        t_first = time.perf_counter_ns()

        def line(content: str, done: bool) -> bytes:
            obj = {"model": model, "created_at": _now(), "message": {"role": "assistant", "content": content}, "done": done}
            if done:
                t1 = time.perf_counter_ns()
                obj.update({"done_reason": meta.get("finish_reason", "stop"), "total_duration": t1 - t0,
                            "load_duration": 0, "prompt_eval_count": int(meta.get("prompt_tokens", 0)),
                            "prompt_eval_duration": t_first - t0, "eval_count": int(meta.get("completion_tokens", 0)),
                            "eval_duration": t1 - t_first})
            return json.dumps(obj, ensure_ascii=False).encode() + b"\n"

        async def pieces() -> AsyncGenerator[str, None]:
            try:
                if first is not None:
                    yield first
                    async for p in agen:
                        yield p
            finally:
                await agen.aclose()
                keys.release(key)

        if not stream:
            try:
                text = "".join([p async for p in pieces()])
            except LLMServiceError as e:
                return JSONResponse({"error": str(e)}, status_code=500)
            return JSONResponse(json.loads(line(text, True)))

        async def ndjson() -> AsyncGenerator[bytes, None]:
            try:
                async for p in pieces():
                    yield line(p, False)
                yield line("", True)
            except LLMServiceError as e:
                yield json.dumps({"error": str(e)}).encode() + b"\n"

        return StreamingResponse(ndjson(), media_type="application/x-ndjson",
                                 background=BackgroundTask(_abandon, agen, keys, key))

    return [Route("/", root, methods=["GET", "HEAD"]), Route("/api/show", show, methods=["POST"]),
            Route("/api/tags", tags, methods=["GET"]), Route("/api/chat", chat, methods=["POST"])]
