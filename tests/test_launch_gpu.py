"""§8f-4 on the GPU: launch.py really starts one engine replica per GPU (separate processes, each seeing one device),
every replica answers the OpenAI wire, a conversation keeps landing on the replica that holds its KV, and the group
goes down together on SIGTERM.  Uses every GPU the box has (one on the round-end box, 2/4/8 under `gpurun --gpus N`)."""
import os
import signal
import socket
import subprocess
import sys
import time

import httpx
import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_one_replica_per_gpu_serves_and_stops():
    from fasttalk_llm_microservice_b200.launch import replica_for
    n = min(torch.cuda.device_count(), 8)
    s = socket.socket(); s.bind(("127.0.0.1", 0)); base = s.getsockname()[1]; s.close()
    env = dict(os.environ, FASTTALK_MODEL="tiny-d64", FASTTALK_MAX_SESSIONS="8", FASTTALK_MAX_CONTEXT="1024",
               FASTTALK_MAX_BATCHED_TOKENS="256", PYTHONPATH=ROOT)
    p = subprocess.Popen([sys.executable, "-m", "fasttalk_llm_microservice_b200.launch", "--gpus", str(n), "--model", "tiny-d64",
                          "--host", "127.0.0.1", "--port-base", str(base)], cwd=ROOT, env=env,
                         stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    try:
        urls = [f"http://127.0.0.1:{base + i}" for i in range(n)]
        deadline = time.time() + 180
        up = set()
        while time.time() < deadline and len(up) < n:
            assert p.poll() is None, p.stdout.read()[-3000:]
            for u in urls:
                if u not in up:
                    try:
                        if httpx.get(u + "/health", timeout=2.0).status_code == 200:
                            up.add(u)
                    except Exception:  # noqa: BLE001
                        pass
            time.sleep(0.5)
        assert len(up) == n, f"only {len(up)} of {n} replicas came up"
        msgs = [{"role": "user", "content": "hello there"}]
        texts = []
        for sid in ("session-a", "session-b", "session-c"):
            u = urls[replica_for(sid, n)]                      # sticky: both turns of a session hit the same replica
            r1 = httpx.post(u + "/v1/chat/completions", timeout=60.0, headers={"X-Session-Id": sid},
                            json={"model": "tiny-d64", "messages": msgs, "temperature": 0.0, "max_tokens": 8}).json()
            a1 = r1["choices"][0]["message"]["content"]
            r2 = httpx.post(u + "/v1/chat/completions", timeout=60.0, headers={"X-Session-Id": sid},
                            json={"model": "tiny-d64", "temperature": 0.0, "max_tokens": 8,
                                  "messages": msgs + [{"role": "assistant", "content": a1}, {"role": "user", "content": "more"}]}).json()
            assert r1["usage"]["completion_tokens"] == 8 and r2["usage"]["completion_tokens"] == 8
            texts.append(a1)
        assert len(set(texts)) == 1                            # same weights (seed) on every replica: same greedy answer
    finally:
        p.send_signal(signal.SIGTERM)
        try:
            p.wait(timeout=30)
        except subprocess.TimeoutExpired:
            p.kill()
    assert p.returncode is not None
