"""Decode-step workload for ncu: 64 sessions at ctx ~2048 (BASELINE configs[1]), a few decode
steps inside a cudaProfilerStart/Stop window so `ncu --profile-from-start off` sees only them.

    python scripts/profile_decode.py [--steps 2] [--sessions 64] [--ctx 2048]
"""
import argparse
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from fasttalk_llm_microservice_b200 import build  # noqa: E402
from fasttalk_llm_microservice_b200.engine import Engine, EngineLimits  # noqa: E402
from fasttalk_llm_microservice_b200.model_spec import get_spec  # noqa: E402
from fasttalk_llm_microservice_b200.weights import init_weights  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--steps", type=int, default=2)
ap.add_argument("--sessions", type=int, default=64)
ap.add_argument("--ctx", type=int, default=2048)
ap.add_argument("--model", default="llama-3.2-1b")
ap.add_argument("--no-pdl", action="store_true")
a = ap.parse_args()

build.build()
spec = get_spec(a.model)
w = init_weights(spec, seed=0, device="cuda")
eng = Engine(spec, w, EngineLimits(max_batch=a.sessions, max_batched_tokens=8192, max_context=a.ctx + 256, decode_prefill_cap=-1,
                                   use_pdl=not a.no_pdl), eos_ids=())
g = torch.Generator().manual_seed(1)
for i in range(a.sessions):
    eng.submit(f"s{i}", torch.randint(0, 128000, (a.ctx,), generator=g).tolist(), max_new_tokens=a.steps + 64, ignore_eos=True)
while eng.stats()["last_prefill_tokens"] > 0 or eng.stats()["steps"] == 0:
    eng.step()
for _ in range(3):
    eng.step()
torch.cuda.synchronize()
torch.cuda.cudart().cudaProfilerStart()
eng.timer_start()
for _ in range(a.steps):
    eng.step()
ms = eng.timer_stop_ms()
torch.cuda.cudart().cudaProfilerStop()
st = eng.stats()
print(f"decode: {ms / a.steps:.3f} ms/step, batch {st['last_batch']}, launches total {st['kernel_launches']}")
eng.close()
