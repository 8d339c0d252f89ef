"""In-kernel timelines (ft_debug_gemm_trace_hist) of the projections of real decode steps: for each GEMM of the
last layers, microseconds from kernel entry of CTA (0,0,0) to each phase, and the gaps between launches."""
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
for _ in range(10):
    eng.step()
n = 64                                        # one step = 1 + 16*4 traced launches (the LM head is the persistent kernel)
tr = np.zeros((n, 16), dtype=np.uint64)
got = lib.ft_debug_gemm_trace_hist(tr.ctypes.data, n)
t = tr.astype(np.int64)
names = ["setup", "Wreq", "dep", "stage0", "mma_iss", "acc", "epi_st", "reduced", "exit", "loop0"]
modes = {3: "RESID", 4: "QKV", 2: "SWIGLU", 5: "LOGITS"}
print("mode n_out | us from entry: " + " ".join(names) + " | entry after prev entry / prev exit")
agg = {}
for i in range(8, n):
    mode, n_out = int(t[i, 10] >> 32), int(t[i, 10] & 0xffffffff)
    rel = np.concatenate([(t[i, 1:10] - t[i, 0]) / 1e3, [(t[i, 11] - t[i, 0]) / 1e3]])
    if i == n - 3 or i == n - 2 or i == n - 1: print('iter stamps', modes.get(mode, mode), [(int(t[i, k]) - int(t[i, 11])) for k in (12, 13, 14, 15)], 'reduced', int(t[i, 8]) - int(t[i, 11]))
    gap_entry = (t[i, 0] - t[i - 1, 0]) / 1e3
    gap_exit = (t[i, 0] - t[i - 1, 9]) / 1e3
    key = (modes.get(mode, mode), n_out)
    agg.setdefault(key, []).append(np.concatenate([rel, [gap_entry, gap_exit]]))
for key, rows in agg.items():
    m = np.median(np.array(rows), axis=0)
    print(f"{key[0]:6s} {key[1]:6d} | " + " ".join(f"{v:7.2f}" for v in m[:10]) + f" | {m[10]:7.2f} {m[11]:7.2f}   (n={len(rows)})")
eng.close()
