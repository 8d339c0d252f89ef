"""One engine process per GPU (sessions are independent, replicas only, no
collective).  Each replica is the OpenAI/Ollama-compatible server of openai_server.py bound to
its own GPU and port; the reference's service instances (one per replica, unchanged) point
`VLLM_BASE_URL` / `OLLAMA_BASE_URL` at them.  A conversation must keep going to the replica that
holds its KV cache, so assignment is sticky per session: `replica_for(session_id, n)`.

    python -m fasttalk_llm_microservice_b200.launch --gpus 8 --model llama3.2:1b --port-base 8000
"""
from __future__ import annotations

import argparse
import hashlib
import os
import signal
import subprocess
import sys
import time
from typing import Dict, List, Tuple


def replica_for(session_id: str, n_replicas: int) -> int:
    """Sticky session → replica map (stable across processes, unlike hash())."""
    return int.from_bytes(hashlib.sha1(str(session_id).encode()).digest()[:8], "little") % max(1, n_replicas)


def replica_commands(n_gpus: int, model: str, host: str, port_base: int, extra_env: Dict[str, str] = None
                     ) -> List[Tuple[List[str], Dict[str, str]]]:
    out = []
    for i in range(n_gpus):
        env = dict(os.environ)
        env.update(extra_env or {})
        # the replica sees exactly one device: no accidental contexts on its neighbours' GPUs
        env["CUDA_VISIBLE_DEVICES"] = str(i)
        env["FASTTALK_GPU"] = "0"
        cmd = [sys.executable, "-m", "fasttalk_llm_microservice_b200.openai_server", "--model", model,
               "--host", host, "--port", str(port_base + i)]
        out.append((cmd, env))
    return out


def main(argv=None) -> int:
    ap = argparse.ArgumentParser(description="one fasttalk B200 engine replica per GPU")
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--model", default=os.environ.get("VLLM_MODEL", "llama3.2:1b"))
    ap.add_argument("--host", default="0.0.0.0")
    ap.add_argument("--port-base", type=int, default=8000)
    a = ap.parse_args(argv)
    procs = [subprocess.Popen(cmd, env=env) for cmd, env in replica_commands(a.gpus, a.model, a.host, a.port_base)]
    for i in range(a.gpus):
        print(f"replica {i}: GPU {i}  http://{a.host}:{a.port_base + i}/v1", flush=True)

    def stop(*_):
        for p in procs:
            if p.poll() is None:
                p.terminate()
    signal.signal(signal.SIGINT, stop)
    signal.signal(signal.SIGTERM, stop)
    rc = 0
    try:
        while procs:
            for p in list(procs):
                r = p.poll()
                if r is not None:
                    procs.remove(p)
                    if r != 0:              # one replica died: take the group down, report it
                        rc = r
                        stop()
            time.sleep(0.2)
    finally:
        stop()
    return rc


if __name__ == "__main__":
    sys.exit(main())
