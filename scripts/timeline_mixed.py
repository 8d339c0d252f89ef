"""Kernel timeline (torch.profiler / CUPTI, no serialisation) of one MIXED step: 56 sessions decoding at ctx ~2k while
eight re-send a 2048-token history of which 224 tokens are new, `CAP` prompt tokens admitted per step.  Writes
gpurun_out/mixed_timeline.txt (start, duration, gap to the previous kernel's end) and prints, per kernel and grid, the
time it adds to the step (end - max(previous end, start)) next to its own duration.

    CAP=896 python scripts/timeline_mixed.py
"""
import os, sys, json
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from fasttalk_llm_microservice_b200 import build
from fasttalk_llm_microservice_b200.engine import Engine, EngineLimits
from fasttalk_llm_microservice_b200.model_spec import get_spec
from fasttalk_llm_microservice_b200.weights import init_weights
from torch.profiler import profile, ProfilerActivity
build.build()
spec = get_spec("llama-3.2-1b")
w = init_weights(spec, seed=0, device="cuda")
eng = Engine(spec, w, EngineLimits(max_batch=64, max_batched_tokens=8192, max_context=2048 + 256, decode_prefill_cap=-1), eos_ids=())
g = torch.Generator().manual_seed(1)
hist = [torch.randint(0, 128000, (2048,), generator=g).tolist() for _ in range(64)]
for i in range(56):
    eng.submit(f"s{i}", hist[i], max_new_tokens=200, ignore_eos=True)
warm = [eng.submit(f"s{i}", hist[i][:1824], max_new_tokens=1, ignore_eos=True) for i in range(56, 64)]
for _ in range(40):
    eng.step(); eng.drain()
CAP = int(os.environ.get("CAP", "768"))
eng.set_decode_prefill_cap(CAP)
for i in range(56, 64):
    eng.submit(f"s{i}", hist[i], max_new_tokens=8, ignore_eos=True)
torch.cuda.synchronize()
eng.step_log()
with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
    eng.step(); eng.step()
    torch.cuda.synchronize()
print("steps:", [r.tolist() for r in eng.step_log()[-2:]])
prof.export_chrome_trace("gpurun_out/mixed_trace.json")
ev = json.load(open("gpurun_out/mixed_trace.json"))["traceEvents"]
ks = sorted([e for e in ev if e.get("cat") == "kernel"], key=lambda e: e["ts"])
t0 = ks[0]["ts"]
out = []
for e in ks:
    out.append((e["ts"] - t0, e["dur"], e["name"][:50]))
with open("gpurun_out/mixed_timeline.txt", "w") as fh:
    prev_end = 0
    for ts, dur, name in out:
        fh.write(f"{ts:9.1f} {dur:7.1f} gap={ts-prev_end:6.1f} {name}\n")
        prev_end = max(prev_end, ts + dur)
import collections
agg = collections.OrderedDict(); prev_end = 0; first = None
for e in ks:
    ts, dur = e["ts"] - t0, e["dur"]
    if ts > 2000 and first is None: first = ts
    end = ts + dur
    if 1000 < ts < 3000:
        key = e["name"].split("(")[0][-40:] + str(e["args"].get("grid"))
        a = agg.setdefault(key, [0, 0.0, 0.0]); a[0] += 1; a[1] += end - max(prev_end, ts); a[2] += dur
    prev_end = max(prev_end, end)
print("CAP", CAP, "PAIR_MIN", os.environ.get("FT_GEMM_PAIR_MIN_TILES"))
for k, (n, eff, dur) in agg.items():
    print(f"  n={n:3d} eff/launch={eff/n:7.1f} dur={dur/n:7.1f}  {k}")
eng.close()
