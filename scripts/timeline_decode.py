"""Kernel timeline of real decode steps (PDL on, no serialisation) through CUPTI / torch.profiler:
per-kernel start, duration and the gap to the previous kernel's end."""
import argparse, os, sys, collections
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from fasttalk_llm_microservice_b200 import build
from fasttalk_llm_microservice_b200.engine import Engine, EngineLimits
from fasttalk_llm_microservice_b200.model_spec import get_spec
from fasttalk_llm_microservice_b200.weights import init_weights

ap = argparse.ArgumentParser()
ap.add_argument("--steps", type=int, default=3)
ap.add_argument("--sessions", type=int, default=64)
ap.add_argument("--ctx", type=int, default=2048)
ap.add_argument("--no-pdl", action="store_true")
a = ap.parse_args()
build.build()
spec = get_spec("llama-3.2-1b")
w = init_weights(spec, seed=0, device="cuda")
eng = Engine(spec, w, EngineLimits(max_batch=a.sessions, max_batched_tokens=8192, max_context=a.ctx + 256, decode_prefill_cap=-1,
                                   use_pdl=not a.no_pdl), eos_ids=())
g = torch.Generator().manual_seed(1)
for i in range(a.sessions):
    eng.submit(f"s{i}", torch.randint(0, 128000, (a.ctx,), generator=g).tolist(), max_new_tokens=a.steps + 64, ignore_eos=True)
while eng.stats()["last_prefill_tokens"] > 0 or eng.stats()["steps"] == 0:
    eng.step()
for _ in range(5):
    eng.step()
torch.cuda.synchronize()
from torch.profiler import profile, ProfilerActivity
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    for _ in range(a.steps):
        eng.step()
    torch.cuda.synchronize()
evs = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA and "emcpy" not in e.name]
evs.sort(key=lambda e: e.time_range.start)
print(f"{len(evs)} kernel events")
# the middle step: from the 2nd embed kernel to the 3rd
starts = [i for i, e in enumerate(evs) if "embed_planes" in e.name]
lo, hi = (starts[1], starts[2]) if len(starts) >= 3 else (0, len(evs))
step = evs[lo:hi]
t0 = step[0].time_range.start
agg = collections.OrderedDict()
prev_end = t0
print("  start_us   dur_us   gap_us  kernel")
for i, e in enumerate(step):
    s, d = e.time_range.start - t0, e.time_range.end - e.time_range.start
    gap = e.time_range.start - prev_end
    name = e.name.split("(")[0].replace("void ft::", "").replace("ft::", "")[:48]
    if i < 14 or i > len(step) - 6:
        print(f"{s:10.1f} {d:8.1f} {gap:8.1f}  {name}")
    k = agg.setdefault(name + (f" grid~{i % 5}" if "gemm" in name and 2 <= i < len(step) - 3 else ""), [0, 0.0, 0.0])
    k[0] += 1; k[1] += d; k[2] += gap
    prev_end = max(prev_end, e.time_range.end)
print(f"step span {step[-1].time_range.end - t0:.1f} us")
if len(starts) >= 3:
    gaps = [evs[starts[i + 1]].time_range.start - max(e.time_range.end for e in evs[starts[i]:starts[i + 1]]) for i in range(len(starts) - 1)]
    print("GPU idle between steps (last kernel end -> next embed start), us:", [round(g, 1) for g in gaps])
for n, (c, d, gp) in agg.items():
    print(f"{n:60s} n={c:3d} dur avg {d / c:7.2f}  gap-before avg {gp / c:6.2f}")
eng.close()
