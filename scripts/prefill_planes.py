"""Round-1 verdict, next #8: does prefill still need both activation planes now that attention is fp16?

For each model and for prefill_planes = 2 (hi + lo, round 1) and 1 (hi only in the multi-wave prefill GEMMs):
  * worst-position logit error of a 2048-token prompt prefilled in ONE 2048-token step (CTA-pair kernel) against
    the fp32 oracle (max|delta| / max|logit| per position; north_star tolerance 2e-2),
  * teacher-forced decode error and greedy parity of 16 tokens generated AFTER that prefill (the KV cache the
    single-plane projections wrote is what decode then reads),
  * device time of a cold 8192-token prefill step (4 x 2048) and of the e2e-shaped mixed step.

    python scripts/prefill_planes.py [--models llama-3.2-1b,llama-3.2-3b,llama-3.1-8b]
"""
import argparse
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from fasttalk_llm_microservice_b200 import build  # noqa: E402
from fasttalk_llm_microservice_b200.engine import Engine, EngineLimits  # noqa: E402
from fasttalk_llm_microservice_b200.model_spec import get_spec  # noqa: E402
from fasttalk_llm_microservice_b200.weights import init_weights  # noqa: E402
from oracle import llama_oracle as O  # noqa: E402
from oracle import parity as P  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--models", default="llama-3.2-1b,llama-3.2-3b,llama-3.1-8b")
ap.add_argument("--n", type=int, default=2048)
a = ap.parse_args()
build.build()
rows = []
for name in a.models.split(","):
    spec = get_spec(name)
    w = init_weights(spec, seed=0, device="cuda")
    w32 = {k: v.float() for k, v in w.items() if k != "lm_head"}
    w32["lm_head"] = w32["embed"] if spec.tied_embeddings else w["lm_head"].float()
    prompt = torch.randint(0, 128000, (a.n,), generator=torch.Generator().manual_seed(11))
    with torch.no_grad():
        ref_last = O.forward(w32, spec, prompt)[-256:].cpu()          # last 256 positions: deepest context
        ref_tokens, ref_logits = O.greedy_generate(w32, spec, prompt, 16, return_logits=True)
    del w32
    torch.cuda.empty_cache()
    for planes in (2, 3, 1):
        eng = Engine(spec, w, EngineLimits(max_batch=4, max_batched_tokens=8192, max_context=a.n + 256,
                                           debug_logit_rows=a.n, prefill_planes=planes, decode_prefill_cap=-1), eos_ids=())
        got = eng.debug_prefill_logits(prompt.tolist(), chunk=a.n)[-256:]
        err = ((got - ref_last).abs().amax(-1) / ref_last.abs().amax(-1))
        job = P.Job(f"{name} planes={planes} decode after {a.n}-token prefill", "pp", prompt.tolist(), ref_tokens, ref_logits)
        P.run(eng, [job])
        eng.close_session("pp")
        # timing: cold 8192-token step
        times = []
        for rep in range(3):
            ps = [torch.randint(0, 128000, (2048,), generator=torch.Generator().manual_seed(100 + rep * 4 + i)).tolist() for i in range(4)]
            for i, p in enumerate(ps):
                eng.submit(f"t{rep}-{i}", p, max_new_tokens=1, ignore_eos=True)
            eng.step_log()
            while eng.stats()["active_requests"] or eng.stats()["waiting_requests"]:
                eng.step()
                eng.drain()
            log = eng.step_log()
            times.append(float(log[0][0]))
            for i in range(4):
                eng.close_session(f"t{rep}-{i}")
        row = {"model": name, "planes": planes, "prefill_logit_err_max": float(err.max()), "prefill_logit_err_mean": float(err.mean()),
               "decode_err_teacher_forced": job.worst_err, "greedy_exact": job.exact, "waived": len(job.waived),
               "waived_margins": [round(m, 6) for _, m, _, _ in job.waived], "cold_8192_step_ms": min(times)}
        print(json.dumps(row), flush=True)
        rows.append(row)
        eng.close()
    del w
    torch.cuda.empty_cache()
print("| model | planes | prefill logit err max (mean) | decode err after prefill | greedy 16 | cold 8192-token step |")
print("|---|---|---|---|---|---|")
for r in rows:
    print(f"| {r['model']} | {r['planes']} | {r['prefill_logit_err_max']:.2e} ({r['prefill_logit_err_mean']:.2e}) | "
          f"{r['decode_err_teacher_forced']:.2e} | {r['greedy_exact']}/16 (+{r['waived']} waived) | {r['cold_8192_step_ms']:.1f} ms |")
