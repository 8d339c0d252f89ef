"""Per-CTA timeline (ft_debug_attn_trace) of the LAST decode-attention launch of real engine steps:
entry, first tile landed, last tile consumed, exit — relative to the earliest entry."""
import os, sys
import numpy as np
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from fasttalk_llm_microservice_b200 import build, native as N
from fasttalk_llm_microservice_b200.engine import Engine, EngineLimits
from fasttalk_llm_microservice_b200.model_spec import get_spec
from fasttalk_llm_microservice_b200.weights import init_weights
build.build()
lib = N.load()
spec = get_spec("llama-3.2-1b")
sessions, ctx = int(os.environ.get("SESSIONS", 64)), 2048
eng = Engine(spec, init_weights(spec, seed=0, device="cuda"), EngineLimits(max_batch=sessions, max_batched_tokens=8192, max_context=ctx + 256), eos_ids=())
g = torch.Generator().manual_seed(1)
for i in range(sessions):
    eng.submit(f"s{i}", torch.randint(0, 128000, (ctx,), generator=g).tolist(), max_new_tokens=200, ignore_eos=True)
while eng.stats()["last_prefill_tokens"] > 0 or eng.stats()["steps"] == 0:
    eng.step()
nc = lib.ft_attn_decode_default_ctas(spec.head_dim)
p = lambda a: f"min {a.min() / 1e3:6.2f} p50 {np.median(a) / 1e3:6.2f} p90 {np.percentile(a, 90) / 1e3:6.2f} max {a.max() / 1e3:6.2f}"
for rep in range(3):
    for _ in range(5):
        eng.step()
    tr = np.zeros((nc, 4), dtype=np.uint64)
    lib.ft_debug_attn_trace(tr.ctypes.data, nc)
    t = tr.astype(np.int64)
    t0 = t[:, 0].min()
    print(f"CTAs {nc}: entry {p(t[:, 0] - t0)} | first tile after entry {p(t[:, 1] - t[:, 0])} | stream done {p(t[:, 2] - t0)} | "
          f"exit {p(t[:, 3] - t0)} | exit - stream done {p(t[:, 3] - t[:, 2])} us", flush=True)
eng.close()
