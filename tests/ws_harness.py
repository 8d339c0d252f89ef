"""Boundary A for real: the reference's UNCHANGED WebSocket server in front of NativeHandler.

The reference's hot loop is `app/core/websocket_server_vllm.py:528-539`
(`async for token in vllm_handler.generate_stream_async(...)` -> `send_json` per token).  This module
starts that server class — imported unmodified from the reference checkout (`/root/reference`, or the
git-ignored copy scripts/install_reference.sh puts in `baseline/_ref`, which is what the GPU box has) —
with only `app.core.vllm_handler.VLLMHandler` swapped for a NativeHandler subclass bound to a given
EngineRunner (handler.install() does the same swap for a deployment), serves it with uvicorn on
127.0.0.1, and drives it with N concurrent `websockets` clients speaking the reference's protocol
(`session_started` / `start_session` / `user_message` / `token` ... `response_complete`).

Test and bench infrastructure only.  As a script it is the client side:

    python tests/ws_harness.py --client --port P --spec spec.json      -> one JSON line with per-session results
"""
from __future__ import annotations

import asyncio
import json
import os
import socket
import sys
import threading
import time
from typing import Dict, List, Optional

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def find_reference() -> Optional[str]:
    """Directory that holds the reference's `app` package, or None."""
    for p in (os.environ.get("FASTTALK_REFERENCE_PATH"), "/root/reference", os.path.join(ROOT, "baseline", "_ref")):
        if p and os.path.exists(os.path.join(p, "app", "core", "websocket_server_vllm.py")):
            return p
    return None


class ReferenceServer:
    """The unchanged WebSocketLLMServer over `runner` (an EngineRunner), on 127.0.0.1:<free port>."""

    def __init__(self, runner, env: Optional[Dict[str, str]] = None, max_connections: int = 128, provider: str = "vllm"):
        """provider "vllm": the server's `_generate_with_vllm` loop over `generate_stream_async`; "ollama": its legacy
        `_generate_with_ollama` loop over the synchronous `generate_stream` (OllamaHandler swapped instead)."""
        import uvicorn
        ref = find_reference()
        if ref is None:
            raise RuntimeError("no reference checkout (run scripts/install_reference.sh where /root/reference exists)")
        self.ref = ref
        self._env_old = {}
        settings = {"LLM_PROVIDER": provider, "ENABLE_PYDANTIC_AI": "false", "DEFAULT_TEMPERATURE": "0.0",
                    "DEFAULT_MAX_TOKENS": "32", "LLM_MAX_CONNECTIONS": str(max_connections), "LOG_LEVEL": "ERROR",
                    "VLLM_MODEL": "llama3.2:1b", **(env or {})}
        for k, v in settings.items():
            self._env_old[k] = os.environ.get(k)
            os.environ[k] = v
        for m in [k for k in sys.modules if k == "app" or k.startswith("app.")]:
            del sys.modules[m]
        sys.path.insert(0, ref)
        try:
            import app.core.vllm_handler as vh                          # the reference's modules, unmodified
            import app.core.ollama_handler as oh
            from app.utils.config import Config
            from fasttalk_llm_microservice_b200.handler import NativeHandler
            self.calls: List[Dict] = []                                 # what the server asked the handler for
            calls = self.calls

            class Bound(NativeHandler):
                def __init__(self, **kw):
                    super().__init__(runner=runner, **kw)

                def generate_stream_async(self, messages, **kw):
                    calls.append({"messages": [dict(m) for m in messages], "kw": dict(kw)})
                    return super().generate_stream_async(messages, **kw)

                def generate_stream(self, messages, **kw):
                    calls.append({"messages": [dict(m) for m in messages], "kw": dict(kw), "sync": True})
                    return super().generate_stream(messages, **kw)

            vh.VLLMHandler = Bound                                      # == handler.install(), bound to this runner
            oh.OllamaHandler = Bound
            from app.core.websocket_server_vllm import WebSocketLLMServer
            self.server_obj = WebSocketLLMServer(Config())
        finally:
            sys.path.remove(ref)
        s = socket.socket()
        s.bind(("127.0.0.1", 0))
        self.port = s.getsockname()[1]
        s.close()
        cfg = uvicorn.Config(self.server_obj.app, host="127.0.0.1", port=self.port, log_level="error", lifespan="off",
                             ws="websockets", ws_ping_interval=None)
        self.uv = uvicorn.Server(cfg)
        self.uv.install_signal_handlers = lambda: None
        self.thread = threading.Thread(target=self.uv.run, daemon=True, name="ref-ws-server")
        self.thread.start()
        for _ in range(500):
            if self.uv.started:
                break
            time.sleep(0.02)
        assert self.uv.started, "the reference server did not start"

    def stop(self):
        self.uv.should_exit = True
        self.thread.join(timeout=10)
        for k, v in self._env_old.items():
            if v is None:
                os.environ.pop(k, None)
            else:
                os.environ[k] = v


# ----------------------------------------------------------------------------- client side
async def _one_session(port: int, spec: Dict, barrier: "_TurnBarrier", idx: int) -> Dict:
    import websockets
    out = {"turns": []}
    async with websockets.connect(f"ws://127.0.0.1:{port}/ws/llm", max_size=None, ping_interval=None,
                                  open_timeout=60) as ws:
        started = json.loads(await ws.recv())
        assert started["type"] == "session_started", started
        out["session_id"] = started["session_id"]
        await ws.send(json.dumps({"type": "start_session", "config": {"system_prompt": spec["system_prompt"]}}))
        assert json.loads(await ws.recv())["type"] == "session_configured"
        for t, text in enumerate(spec["turns"]):
            await barrier.wait(t)                      # all sessions send turn t together
            t0 = time.perf_counter()
            await ws.send(json.dumps({"type": "user_message", "text": text}))
            chunks, stamps = [], []
            while True:
                msg = json.loads(await ws.recv())
                if msg["type"] == "token":
                    chunks.append(msg["data"])
                    stamps.append(time.perf_counter() - t0)
                else:
                    break
            out["turns"].append({"chunks": chunks, "stamps": stamps, "end": msg, "t_send": t0,
                                 "t_done": time.perf_counter()})
        await ws.send(json.dumps({"type": "end_session"}))
        try:
            await asyncio.wait_for(ws.recv(), timeout=5)
        except Exception:  # noqa: BLE001
            pass
    return out


class _TurnBarrier:
    def __init__(self, n: int, n_turns: int):
        self.n, self.count = n, [0] * n_turns
        self.events = [asyncio.Event() for _ in range(n_turns)]

    async def wait(self, t: int):
        self.count[t] += 1
        if self.count[t] == self.n:
            self.events[t].set()
        await self.events[t].wait()


async def run_clients_async(port: int, specs: List[Dict]) -> List[Dict]:
    n_turns = max(len(s["turns"]) for s in specs)
    assert all(len(s["turns"]) == n_turns for s in specs)
    barrier = _TurnBarrier(len(specs), n_turns)
    return await asyncio.gather(*[_one_session(port, s, barrier, i) for i, s in enumerate(specs)])


def run_clients(port: int, specs: List[Dict]) -> List[Dict]:
    return asyncio.run(run_clients_async(port, specs))


def run_clients_subprocess(port: int, specs: List[Dict], timeout: float = 600.0) -> List[Dict]:
    """The clients in their own process (their JSON parsing must not share the server's GIL)."""
    import subprocess
    import tempfile
    with tempfile.NamedTemporaryFile("w", suffix=".json", delete=False) as f:
        json.dump(specs, f)
        path = f.name
    try:
        r = subprocess.run([sys.executable, os.path.abspath(__file__), "--client", "--port", str(port), "--spec", path],
                           capture_output=True, text=True, timeout=timeout)
        if r.returncode != 0:
            raise RuntimeError("ws clients failed: " + r.stderr[-2000:])
        return json.loads(r.stdout.strip().splitlines()[-1])
    finally:
        os.unlink(path)


def turn_metrics(results: List[Dict], turn: int) -> Dict:
    """tok/s over the whole turn (first send .. last completion), TTFT and inter-token percentiles."""
    def pct(xs, q):
        xs = sorted(xs)
        return xs[min(len(xs) - 1, int(q * len(xs)))] if xs else None
    ts = [r["turns"][turn] for r in results]
    n_tok = sum(len(t["chunks"]) for t in ts)
    wall = max(t["t_done"] for t in ts) - min(t["t_send"] for t in ts)
    ttft = [t["stamps"][0] for t in ts if t["stamps"]]
    gaps = [b - a for t in ts for a, b in zip(t["stamps"], t["stamps"][1:])]
    return {"tokens": n_tok, "wall_s": wall, "tok_s": n_tok / wall if wall > 0 else 0.0,
            "ttft_ms_p50": 1e3 * pct(ttft, 0.5), "ttft_ms_p99": 1e3 * pct(ttft, 0.99),
            "itl_ms_p50": 1e3 * pct(gaps, 0.5) if gaps else None, "itl_ms_p99": 1e3 * pct(gaps, 0.99) if gaps else None,
            "sessions": len(ts)}


if __name__ == "__main__":
    import argparse
    ap = argparse.ArgumentParser()
    ap.add_argument("--client", action="store_true")
    ap.add_argument("--port", type=int, required=True)
    ap.add_argument("--spec", required=True)
    a = ap.parse_args()
    with open(a.spec) as fh:
        specs_ = json.load(fh)
    print(json.dumps(run_clients(a.port, specs_)), flush=True)
