"""Kernel timeline of prefill steps through CUPTI / torch.profiler.
  cold: `--sessions 4 --ctx 2048` → one 8192-token step of cold prompts
  warm: `--sessions 64 --ctx 2048 --new 224` → the bench's e2e turn (prefix-cached history + 224 new tokens each)"""
import argparse, os, sys, collections, time
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from fasttalk_llm_microservice_b200 import build
from fasttalk_llm_microservice_b200.engine import Engine, EngineLimits
from fasttalk_llm_microservice_b200.model_spec import get_spec
from fasttalk_llm_microservice_b200.weights import init_weights

ap = argparse.ArgumentParser()
ap.add_argument("--sessions", type=int, default=4)
ap.add_argument("--ctx", type=int, default=2048)
ap.add_argument("--new", type=int, default=0, help="warm mode: tokens appended to a cached history of ctx-new")
ap.add_argument("--model", default="llama-3.2-1b")
ap.add_argument("--budget", type=int, default=8192)
ap.add_argument("--no-trace", action="store_true", help="skip the CUPTI pass (when running under ncu)")
a = ap.parse_args()
build.build()
spec = get_spec(a.model)
w = init_weights(spec, seed=0, device="cuda")
eng = Engine(spec, w, EngineLimits(max_batch=max(a.sessions, 4), max_batched_tokens=a.budget, max_context=a.ctx + 256), eos_ids=())
g = torch.Generator().manual_seed(1)
toks = [torch.randint(0, 128000, (a.ctx,), generator=g).tolist() for _ in range(a.sessions)]


def run_all(prompts, tag):
    live = set(eng.submit(f"s{i}", p, max_new_tokens=1, ignore_eos=True) for i, p in enumerate(prompts))
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    n = 0
    while live:
        eng.step(); n += 1
        for r, t, fin in eng.drain():
            if fin:
                live.discard(r)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    tok = sum(len(p) for p in prompts)
    print(f"{tag}: {n} steps, {dt * 1e3:.2f} ms wall, {tok} prompt tokens submitted")
    return dt


# warm-up pass (also fills the cache for warm mode)
hist = [t[:a.ctx - a.new] for t in toks] if a.new else toks
run_all(hist, "warmup")
if not a.new:
    for i in range(a.sessions):
        eng.close_session(f"s{i}")
    run_all(hist, "cold (untraced)")
    for i in range(a.sessions):
        eng.close_session(f"s{i}")
if a.no_trace:
    run_all(toks, "untraced")
    eng.close()
    sys.exit(0)
from torch.profiler import profile, ProfilerActivity
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    run_all(toks, "traced")
evs = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA and "emcpy" not in e.name]
evs.sort(key=lambda e: e.time_range.start)
agg = collections.OrderedDict()
for e in evs:
    name = e.name.split("(")[0].replace("void ft::", "").replace("ft::", "")[:60]
    k = agg.setdefault(name, [0, 0.0])
    k[0] += 1; k[1] += e.time_range.end - e.time_range.start
span = evs[-1].time_range.end - evs[0].time_range.start
tot = sum(v[1] for v in agg.values())
print(f"{len(evs)} kernels, span {span:.0f} us, sum of durations {tot:.0f} us")
for n, (c, d) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"{n:62s} n={c:4d} total {d:9.1f} us  avg {d / c:8.2f}  {100 * d / tot:5.1f}%")
# first layer of the first traced step in order
print("first 12 kernels:")
t0 = evs[0].time_range.start
for e in evs[:12]:
    print(f"{e.time_range.start - t0:10.1f} {e.time_range.end - e.time_range.start:9.1f}  {e.name[:90]}")
eng.close()
