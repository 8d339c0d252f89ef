"""NativeHandler — the drop-in for the reference's backend handlers (boundary A).

Duck-types both `VLLMHandler` (`app/core/vllm_handler.py:24-338`) and `OllamaHandler`
(`app/core/ollama_handler.py:24-458`): same constructor keywords, same
`generate_stream` / `generate_stream_async` signatures and yield type (one `str` chunk per
generated token), same `check_connection`, `get_model_info`, `cancel_generation`,
`get_active_requests`, `cleanup_stale_requests`, `close`, and the same error convention
(`LLMServiceError` raised from inside the generator, `check_connection` returns False instead
of raising).  Instead of an HTTP hop to an engine container it tokenises the conversation,
submits it to the in-process B200 engine and streams sampled tokens back.

The continuous-batching loop runs on the engine's own native thread (`ft_engine_start`); one Python
thread waits for its events (GIL released) and a second one detokenises and hands each event loop ONE
wake-up carrying all of a step's tokens, so 64 sessions cost one
`call_soon_threadsafe` per decode step rather than 64.  (Tensor-parallel leaders and
`FASTTALK_NATIVE_LOOP=0` step from the Python thread instead.)
"""
from __future__ import annotations

import asyncio
import concurrent.futures
import contextlib
import functools
import logging
import os
import queue
import random
import threading
import time
import uuid
from typing import Any, AsyncGenerator, Dict, Generator, List, Optional

from .errors import ErrorCategory, ErrorSeverity, LLMServiceError
from .tokenizer import load_tokenizer

try:  # the reference's structured logger, when we run inside the reference service
    from app.utils.logger import get_logger  # type: ignore
    logger = get_logger(__name__)
except Exception:  # noqa: BLE001
    logger = logging.getLogger(__name__)

_DONE = object()
_TIMEOUT = object()


def _arm_idle_watchdog(sink: "_Sink", loop, timeout: float) -> None:
    """One timer per generation instead of one `asyncio.wait_for` per token (64 sessions x ~600 tokens/s each
    made the per-chunk timeout handles a measurable share of the event loop): when it fires it either
    finds that nothing arrived for `timeout` seconds and wakes the consumer with _TIMEOUT, or re-arms
    itself for what is left of the current idle period."""
    def check():
        idle = time.monotonic() - sink.last
        if idle >= timeout:
            sink.aq.put_nowait(_TIMEOUT)
        else:
            sink.timer = loop.call_later(timeout - idle, check)
    sink.timer = loop.call_later(timeout, check)


class _Sink:
    """Where one request's text chunks go: an asyncio queue on its loop, or a thread queue."""
    __slots__ = ("loop", "aq", "tq", "decoder", "text", "stop", "finished", "error", "rid", "start_time",
                 "cancelled", "n_tokens", "finish", "n_prompt", "last", "timer")

    def __init__(self, loop, decoder, stop):
        self.loop = loop
        self.aq = asyncio.Queue() if loop is not None else None
        self.tq = queue.Queue() if loop is None else None
        self.decoder, self.stop = decoder, stop or []
        self.text, self.finished, self.error, self.rid = "", False, None, -1
        self.start_time, self.cancelled, self.n_tokens = time.time(), False, 0
        self.finish, self.n_prompt = None, 0      # engine finish reason ("stop" | "length" | ...), prompt tokens
        self.last, self.timer = time.monotonic(), None     # async consumers: last item seen / idle watchdog handle


class EngineRunner:
    """Owns the threads around the engine.  `engine` needs submit/step/drain/cancel/close_session/stats
    (fasttalk_llm_microservice_b200.engine.Engine, or a test double)."""

    def __init__(self, engine, tokenizer):
        self.engine, self.tokenizer = engine, tokenizer
        self._lock = threading.Lock()
        self._sinks: Dict[int, _Sink] = {}
        self._wake = threading.Event()
        self._stop = False
        self.last_step_time = time.time()
        self.error: Optional[BaseException] = None
        # Detokenising and posting a step's tokens is Python work (a few hundred microseconds at 64 sessions);
        # it runs on its own thread so that the engine thread is back inside ft_engine_step — which releases
        # the GIL — right away, and the GPU does not idle while tokens are being delivered.
        # Engines with a stepping thread of their own (ft_engine_start; not the tensor-parallel leader, not test doubles)
        # are not stepped from Python at all; FASTTALK_NATIVE_LOOP=0 keeps the Python loop (A/B, debugging).
        self._native_loop = (hasattr(engine, "start_loop") and getattr(engine, "tp_size", 1) == 1
                             and os.environ.get("FASTTALK_NATIVE_LOOP", "1") != "0")
        self._paused = False
        self._dq: "queue.SimpleQueue" = queue.SimpleQueue()
        self._dthread = threading.Thread(target=self._dispatch_loop, name="fasttalk-b200-dispatch", daemon=True)
        self._dthread.start()
        self._thread = threading.Thread(target=self._run, name="fasttalk-b200-engine", daemon=True)
        self._thread.start()

    # ---- called from request threads / event loops
    def submit(self, sink: _Sink, session, tokens, **kw) -> int:
        with self._lock:
            rid = self.engine.submit(session, tokens, **kw)
            sink.rid = rid
            self._sinks[rid] = sink
        self._wake.set()
        return rid

    def cancel(self, sink: _Sink):
        sink.cancelled = True
        with self._lock:
            if sink.rid in self._sinks and not sink.finished:
                try:
                    self.engine.cancel(sink.rid)
                except Exception:  # noqa: BLE001
                    pass
        self._wake.set()

    @contextlib.contextmanager
    def paused(self):
        """Hand the engine to the caller for the duration (parity checks step it themselves and read the logits of every
        step): the engine's own stepping thread is stopped and started again afterwards.  Only while no request of this
        runner is in flight."""
        with self._lock:
            assert not self._sinks, "EngineRunner.paused(): requests are in flight"
        if not self._native_loop:
            yield self.engine                   # the Python loop only steps while it has requests of its own
            return
        self._paused = True
        self.engine.stop_loop()
        try:
            yield self.engine
        finally:
            self.engine.start_loop()
            self._paused = False

    def alive(self) -> bool:
        return self._thread.is_alive() and self.error is None

    def shutdown(self):
        self._stop = True
        self._wake.set()
        self._thread.join(timeout=5)
        self._dq.put(None)
        self._dthread.join(timeout=5)

    def _dispatch_loop(self):
        while True:
            events = self._dq.get()
            if events is None:
                return
            try:
                self._dispatch(events)
            except Exception as e:  # noqa: BLE001 - a broken consumer must not stop delivery to the others
                logger.error(f"token dispatch failed: {e}")

    # ---- engine thread
    def _run(self):
        try:
            dev = getattr(self.engine, "device", None)
            if dev is not None:                 # a fresh thread starts on device 0: make the engine's GPU current
                import torch
                torch.cuda.set_device(dev)
            if self._native_loop:
                # the engine steps on its own native thread; this one only waits for events (GIL released) and hands
                # them to the dispatch thread — no Python between two steps
                self.engine.start_loop()
                try:
                    while not self._stop:
                        if self._paused:                        # the engine is on loan (paused()): nothing to wait for
                            time.sleep(0.005)
                            continue
                        if self.engine.wait_events(50) > 0:
                            self.last_step_time = time.time()
                            events = self.engine.drain()
                            if events:
                                self._dq.put(events)
                        else:
                            with self._lock:
                                if not self._sinks:
                                    self.last_step_time = time.time()      # idle, not stuck
                finally:
                    self.engine.stop_loop()
                return
            while not self._stop:
                with self._lock:
                    busy = bool(self._sinks)
                if not busy:
                    self._wake.wait(0.05)
                    self._wake.clear()
                    continue
                produced = self.engine.step()
                self.last_step_time = time.time()
                events = self.engine.drain()
                if events:
                    self._dq.put(events)
                if not produced and not events:
                    time.sleep(0.0005)          # waiting on admission / nothing runnable yet
        except BaseException as e:  # noqa: BLE001 - surface engine failures to every waiter
            self.error = e
            logger.error(f"engine thread died: {e}")
            with self._lock:
                sinks = list(self._sinks.values())
                self._sinks.clear()
            for s in sinks:
                s.error = e
                self._post({id(s.loop): (s.loop, [(s, _DONE)])})

    def _dispatch(self, events):
        per_loop: Dict[int, Any] = {}
        for rid, tok, fin in events:
            with self._lock:
                sink = self._sinks.get(rid)
            if sink is None:
                continue
            items = per_loop.setdefault(id(sink.loop), (sink.loop, []))[1]
            if tok >= 0 and not (fin == "stop"):
                piece = sink.decoder.push(tok)
                sink.n_tokens += 1
                if piece:
                    items.append((sink, piece))
            if fin is not None:
                tail = sink.decoder.flush()
                if tail:
                    items.append((sink, tail))
                sink.finish = fin
                if fin == "error":
                    sink.error = RuntimeError("engine could not serve the request (out of KV pages)")
                sink.finished = True
                items.append((sink, _DONE))
                with self._lock:
                    self._sinks.pop(rid, None)
        self._post(per_loop)

    @staticmethod
    def _deliver(items):
        for sink, item in items:
            sink.aq.put_nowait(item)

    def _post(self, per_loop):
        for loop, items in per_loop.values():
            if not items:
                continue
            if loop is None:
                for sink, item in items:
                    sink.tq.put(item)
            else:
                try:
                    loop.call_soon_threadsafe(self._deliver, items)     # one wake-up per step per loop
                except RuntimeError:
                    pass                                                # loop closed: consumer is gone


_shared_lock = threading.Lock()
_shared: Dict[str, EngineRunner] = {}

# conversations shorter than this are tokenised on the caller's event loop (a thread hop costs more than it saves)
_OFF_LOOP_MIN_CHARS = 2048
_tok_pool: Optional[concurrent.futures.ThreadPoolExecutor] = None


def _tokenise_pool() -> concurrent.futures.ThreadPoolExecutor:
    global _tok_pool
    with _shared_lock:
        if _tok_pool is None:
            _tok_pool = concurrent.futures.ThreadPoolExecutor(max_workers=max(4, min(8, os.cpu_count() or 4)),
                                                              thread_name_prefix="fasttalk-b200-tokenise")
        return _tok_pool


def _env_int(name, default):
    try:
        return int(os.environ.get(name, default))
    except ValueError:
        return default


def build_runner(model: str) -> EngineRunner:
    """Create the engine for `model` on GPU $FASTTALK_GPU: weights from $FASTTALK_CHECKPOINT (a
    Hugging Face Llama directory) or $FASTTALK_WEIGHTS (stacked layout), else random-init."""
    import torch
    from .engine import Engine, EngineLimits
    from .model_spec import get_spec
    from .weights import init_weights
    dev = _env_int("FASTTALK_GPU", 0)
    if not torch.cuda.is_available():
        raise LLMServiceError("fasttalk_b200: no CUDA device; the native backend has no CPU fallback",
                              category=ErrorCategory.GPU, severity=ErrorSeverity.CRITICAL, recoverable=False)
    torch.cuda.set_device(dev)
    ckpt = os.environ.get("FASTTALK_CHECKPOINT")
    tok_path = None
    if ckpt:                            # a Hugging Face Llama directory: config.json + *.safetensors (+ tokenizer.json)
        from .checkpoint import load_hf_checkpoint
        spec, w = load_hf_checkpoint(ckpt, device=f"cuda:{dev}", name=model)
        if os.path.exists(os.path.join(ckpt, "tokenizer.json")):
            tok_path = os.path.join(ckpt, "tokenizer.json")
    else:
        spec = get_spec(os.environ.get("FASTTALK_MODEL", model))
        wpath = os.environ.get("FASTTALK_WEIGHTS")
        if wpath:                       # one safetensors file already in weights.py's stacked layout
            from safetensors.torch import load_file
            w = load_file(wpath, device=f"cuda:{dev}")
            if "lm_head" not in w:
                w["lm_head"] = w["embed"]
        else:
            w = init_weights(spec, seed=_env_int("FASTTALK_SEED", 0), device=f"cuda:{dev}")
    tok = load_tokenizer(spec.vocab, tok_path, require=bool(ckpt))
    lim = EngineLimits(max_batch=_env_int("FASTTALK_MAX_SESSIONS", 64),
                       max_batched_tokens=_env_int("FASTTALK_MAX_BATCHED_TOKENS", 8192),
                       max_context=_env_int("FASTTALK_MAX_CONTEXT", _env_int("DEFAULT_CONTEXT_WINDOW", 8192)),
                       decode_prefill_cap=_env_int("FASTTALK_DECODE_PREFILL_CAP", 0))
    eng = Engine(spec, w, lim, eos_ids=tok.eos_ids)
    # Loading a checkpoint and a tokenizer leaves millions of long-lived Python objects behind; a full collection over
    # them later takes tens of milliseconds and lands in the middle of somebody's stream.  Collect once and move what
    # exists out of the collector's generations (FASTTALK_GC_FREEZE=0 leaves the collector alone).
    if os.environ.get("FASTTALK_GC_FREEZE", "1") != "0":
        import gc
        gc.collect()
        gc.freeze()
    return EngineRunner(eng, tok)


class NativeHandler:
    """Drop-in for VLLMHandler / OllamaHandler backed by the in-process B200 engine."""

    def __init__(self, base_url: str = "native://b200", model: str = "llama3.2:1b", api_key: str = "not-needed",
                 timeout: float = 600.0, keep_alive: str = "5m", runner: Optional[EngineRunner] = None):
        # base_url / api_key / keep_alive are accepted for signature parity and ignored
        self.base_url = (base_url or "").rstrip("/")
        self.model = model
        self.api_key = api_key
        self.keep_alive = keep_alive
        self.timeout = timeout
        self._runner = runner
        self._active_requests: Dict[str, Dict[str, Any]] = {}
        self._requests_lock = threading.Lock()
        self._connection_ok = False
        logger.info(f"NativeHandler initialized: model={model} (in-process B200 engine)")

    # ------------------------------------------------------------------ engine access
    @property
    def runner(self) -> EngineRunner:
        if self._runner is None:
            with _shared_lock:
                if self.model not in _shared or not _shared[self.model].alive():
                    _shared[self.model] = build_runner(self.model)
                self._runner = _shared[self.model]
        return self._runner

    # ------------------------------------------------------------------ lifecycle (a4 / a5)
    def check_connection(self) -> bool:
        """vllm_handler.py:75-95 / ollama_handler.py:68-85: True when the backend answers;
        never raises."""
        try:
            ok = self.runner.alive()
            if ok:
                self.runner.engine.stats()
            self._connection_ok = ok
            return ok
        except Exception as e:  # noqa: BLE001
            logger.error(f"native engine unavailable: {e}")
            self._connection_ok = False
            return False

    def get_model_info(self) -> Dict[str, Any]:
        """vllm_handler.py:97-115 shape: {"models": [...], "current_model": ...}."""
        try:
            st = self.runner.engine.stats()
            return {"models": [self.model], "current_model": self.model, "backend": "fasttalk_b200", "engine": st}
        except Exception as e:  # noqa: BLE001
            raise LLMServiceError(f"Failed to get model info: {e}", category=ErrorCategory.CONNECTION,
                                  severity=ErrorSeverity.MEDIUM)

    def cancel_generation(self, request_id: Optional[str] = None) -> bool:
        """vllm_handler.py:310-326 (one id) and ollama_handler.py:353-384 (None = all)."""
        with self._requests_lock:
            ids = list(self._active_requests) if request_id is None else [request_id]
            hit = False
            for rid in ids:
                info = self._active_requests.get(rid)
                if info is None:
                    continue
                info["cancelled"] = True
                self.runner.cancel(info["sink"])
                hit = True
        if not hit and request_id is not None:
            logger.warning(f"Request {request_id} not found for cancellation")
        return hit

    def get_active_requests(self) -> Dict[str, Dict[str, Any]]:
        """vllm_handler.py:328-338 (dict view; iterating it gives the ids like
        ollama_handler.py:412-415's list)."""
        with self._requests_lock:
            now = time.time()
            return {rid: {"start_time": i["start_time"], "duration_s": now - i["start_time"],
                          "cancelled": i.get("cancelled", False)} for rid, i in self._active_requests.items()}

    def cleanup_stale_requests(self, timeout_seconds: int = 300) -> int:
        """ollama_handler.py:417-445."""
        now = time.time()
        with self._requests_lock:
            stale = [r for r, i in self._active_requests.items() if now - i["start_time"] > timeout_seconds]
        return sum(1 for r in stale if self.cancel_generation(r))

    def close(self):
        """ollama_handler.py:447-458."""
        self.cancel_generation(None)

    def end_session(self, request_id: str):
        """Free the KV a conversation holds (not part of the reference interface; call it on
        session end to return pages early — idle sessions are evicted LRU anyway)."""
        try:
            self.runner.engine.close_session(request_id)
        except Exception:  # noqa: BLE001
            pass

    # ------------------------------------------------------------------ generation (a1 / a2 / a3)
    def _start(self, loop, messages, temperature, max_tokens, top_p, top_k, stop, request_id, prefix,
               tools=None, seed=None):
        req_id = request_id or f"{prefix}-{uuid.uuid4()}"
        runner = self.runner
        tok = runner.tokenizer
        if tools or any(not isinstance(m.get("content"), str) or m.get("role") == "tool" for m in messages):
            from .chat_format import ChatFormatError, normalise_messages
            try:                # tool schemas / tool turns in the history (vllm_handler.py:245-260, :425-440)
                messages = normalise_messages(messages, tools)
            except ChatFormatError as e:
                raise LLMServiceError(str(e), category=ErrorCategory.VALIDATION, severity=ErrorSeverity.MEDIUM)
        tokens = tok.apply_chat_template(messages)
        eng = runner.engine
        max_ctx = getattr(getattr(eng, "limits", None), "max_context", 1 << 30)
        room = max_ctx - len(tokens) - 1
        if room < 1:
            raise LLMServiceError(
                f"conversation of {len(tokens)} tokens does not fit the {max_ctx}-token context window",
                category=ErrorCategory.VALIDATION, severity=ErrorSeverity.MEDIUM)
        max_new = min(int(max_tokens), room) if max_tokens else room
        sink = _Sink(loop, tok.stream_decoder(), stop)
        sink.n_prompt = len(tokens)
        with self._requests_lock:
            self._active_requests[req_id] = {"start_time": time.time(), "cancelled": False, "sink": sink}
        try:
            runner.submit(sink, req_id, tokens,
                          temperature=1.0 if temperature is None else float(temperature),
                          top_k=int(top_k or 0), top_p=1.0 if top_p is None else float(top_p),
                          max_new_tokens=max_new,
                          seed=(random.getrandbits(32) if seed is None else int(seed)) & 0xFFFFFFFF)
        except Exception:
            with self._requests_lock:
                self._active_requests.pop(req_id, None)
            raise
        return req_id, sink

    def _tokenise_off_loop(self, messages) -> bool:
        tok = self.runner.tokenizer
        if not getattr(tok, "releases_gil", False) or os.environ.get("FASTTALK_TOKENISE_OFF_LOOP", "1") == "0":
            return False
        n = 0
        for m in messages:
            c = m.get("content")
            n += len(c) if isinstance(c, str) else 0
        return n >= _OFF_LOOP_MIN_CHARS

    def _abandon_started(self, fut):
        if fut.cancelled() or fut.exception() is not None:
            return
        req_id, sink = fut.result()
        self._finish(req_id, sink)

    def _finish(self, req_id, sink, meta=None, hit_stop=False):
        if meta is not None:                    # what the OpenAI front reports (openai_server.py)
            meta["finish_reason"] = "stop" if hit_stop or sink.finish != "length" else "length"
            meta["prompt_tokens"], meta["completion_tokens"] = sink.n_prompt, sink.n_tokens
        if not sink.finished:
            self.runner.cancel(sink)
        with self._requests_lock:
            self._active_requests.pop(req_id, None)

    @staticmethod
    def _apply_stop(sink: _Sink, piece: str):
        """Returns (text to yield now, hit_stop).  Text that could still turn out to be the start
        of a stop string is held back in sink.text until it is disambiguated."""
        if not sink.stop:
            return piece, False
        buf = sink.text + piece
        cut = min((i for i in (buf.find(s) for s in sink.stop if s) if i >= 0), default=-1)
        if cut >= 0:
            sink.text = ""
            return buf[:cut], True
        hold = 0
        for s_ in sink.stop:
            for k in range(min(len(s_) - 1, len(buf)), 0, -1):
                if buf.endswith(s_[:k]):
                    hold = max(hold, k)
                    break
        sink.text = buf[len(buf) - hold:] if hold else ""
        return buf[:len(buf) - hold], False

    def generate_stream(self, messages: List[Dict[str, str]], temperature: Optional[float] = None,
                        max_tokens: Optional[int] = None, top_p: Optional[float] = None,
                        top_k: Optional[int] = None, stop: Optional[List[str]] = None,
                        tools: Optional[List[Dict]] = None, request_id: Optional[str] = None,
                        seed: Optional[int] = None, _meta: Optional[Dict[str, Any]] = None,
                        ) -> Generator[str, None, None]:
        """Synchronous streaming generation: vllm_handler.py:117-214 (has `tools`) and
        ollama_handler.py:110-231 (has `top_k`) in one signature."""
        req_id = sink = None
        hit = False
        try:
            req_id, sink = self._start(None, messages, temperature, max_tokens, top_p, top_k, stop, request_id,
                                       "native", tools, seed)
            deadline = time.time() + self.timeout
            while True:
                try:
                    item = sink.tq.get(timeout=max(0.0, min(1.0, deadline - time.time())))
                except queue.Empty:
                    if time.time() >= deadline:
                        raise TimeoutError(f"no token within {self.timeout}s")
                    if not self.runner.alive():
                        raise RuntimeError(f"engine stopped: {self.runner.error}")
                    continue
                if item is _DONE:
                    if sink.error:
                        raise sink.error
                    if sink.text:               # held-back text that never became a stop string
                        yield sink.text
                    break
                if sink.cancelled:              # cancel_generation: stop at once, like the
                    break                       # reference's per-chunk flag check (vllm_handler.py:181-187)
                text, hit = self._apply_stop(sink, item)
                if text:
                    yield text
                if hit:
                    break
        except LLMServiceError:
            raise
        except Exception as e:  # noqa: BLE001
            logger.error(f"[{req_id}] Generation error: {e}")
            raise LLMServiceError(f"native generation error: {e}", category=ErrorCategory.PROCESSING,
                                  severity=ErrorSeverity.HIGH)
        finally:
            if sink is not None:
                self._finish(req_id, sink, _meta, hit)

    async def generate_stream_async(self, messages: List[Dict[str, str]], temperature: Optional[float] = None,
                                    max_tokens: Optional[int] = None, top_p: Optional[float] = None,
                                    stop: Optional[List[str]] = None, tools: Optional[List[Dict]] = None,
                                    request_id: Optional[str] = None, top_k: Optional[int] = None,
                                    seed: Optional[int] = None, _meta: Optional[Dict[str, Any]] = None,
                                    ) -> AsyncGenerator[str, None]:
        """Async streaming generation: vllm_handler.py:216-308."""
        req_id = sink = None
        hit = False
        try:
            loop = asyncio.get_running_loop()
            start = functools.partial(self._start, loop, messages, temperature, max_tokens, top_p,
                                      top_k, stop, request_id, "native-async", tools, seed)
            if self._tokenise_off_loop(messages):
                # A real tokenizer.json spends ~1 ms on a 2k-token history with the GIL released: on a worker thread the
                # loop keeps delivering other sessions' tokens meanwhile, and a burst of arrivals tokenises in parallel.
                fut = loop.run_in_executor(_tokenise_pool(), start)
                try:
                    req_id, sink = await asyncio.shield(fut)
                except asyncio.CancelledError:
                    # the caller went away mid-tokenisation: whatever _start submits is released when it returns
                    fut.add_done_callback(self._abandon_started)
                    raise
            else:
                req_id, sink = start()
            _arm_idle_watchdog(sink, loop, self.timeout)
            while True:
                item = await sink.aq.get()
                sink.last = time.monotonic()
                if item is _TIMEOUT:
                    raise TimeoutError(f"no token within {self.timeout}s")
                if item is _DONE:
                    if sink.error:
                        raise sink.error
                    if sink.text:
                        yield sink.text
                    break
                if sink.cancelled:
                    break
                text, hit = self._apply_stop(sink, item)
                if text:
                    yield text
                if hit:
                    break
        except LLMServiceError:
            raise
        except (asyncio.CancelledError, GeneratorExit):
            raise
        except Exception as e:  # noqa: BLE001
            logger.error(f"[{req_id}] Async generation error: {e}")
            raise LLMServiceError(f"native async generation error: {e}", category=ErrorCategory.PROCESSING,
                                  severity=ErrorSeverity.HIGH)
        finally:
            # disconnects finalise the generator here: the batch slot is freed
            if sink is not None:
                if sink.timer is not None:
                    sink.timer.cancel()
                self._finish(req_id, sink, _meta, hit)


def install(into_vllm: bool = True, into_ollama: bool = True) -> None:
    """Swap NativeHandler in for the reference's handler classes without editing the reference
    (SURVEY.md §8b): `WebSocketLLMServer` imports `VLLMHandler` lazily
    (websocket_server_vllm.py:113), the legacy server imports `OllamaHandler` at module top
    (websocket_server.py:22), so call this BEFORE importing `app.core.websocket_server*`."""
    if into_vllm:
        import app.core.vllm_handler as vh  # type: ignore
        vh.VLLMHandler = NativeHandler
    if into_ollama:
        import app.core.ollama_handler as oh  # type: ignore
        oh.OllamaHandler = NativeHandler
