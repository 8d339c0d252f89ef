"""Device time of MIXED steps: `ND` sessions decoding at ctx ~2k while the others re-send histories whose last ~1040 tokens
are new, for each prompt-token cap in `CAPS` (min over three rounds of three equal steps, from ft_engine_step_log).

    CAPS=200,456,768,840,968 [FT_GEMM_PAIR_MIN_TILES=149] [FT_PREFILL_PLANES=2] python scripts/mixed_step.py
"""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from fasttalk_llm_microservice_b200 import build
from fasttalk_llm_microservice_b200.engine import Engine, EngineLimits
from fasttalk_llm_microservice_b200.model_spec import get_spec
from fasttalk_llm_microservice_b200.weights import init_weights
build.build()
spec = get_spec(os.environ.get("MODEL", "llama-3.2-1b"))
w = init_weights(spec, seed=0, device="cuda")
eng = Engine(spec, w, EngineLimits(max_batch=64, max_batched_tokens=8192, max_context=2048 + 256, decode_prefill_cap=-1), eos_ids=())
g = torch.Generator().manual_seed(1)
hist = [torch.randint(0, 128000, (2048,), generator=g).tolist() for _ in range(64)]
ND = int(os.environ.get("ND", "56"))
for i in range(ND):
    eng.submit(f"s{i}", hist[i], max_new_tokens=400, ignore_eos=True)
for i in range(ND, 64):
    eng.submit(f"s{i}", hist[i][:1000], max_new_tokens=1, ignore_eos=True)
for _ in range(40):
    eng.step(); eng.drain()
res = {}
for cap in [int(c) for c in os.environ.get("CAPS", "768").split(",")]:
    eng.set_decode_prefill_cap(cap)
    best = None
    for rep in range(3):
        for i in range(ND, 64):
            eng.submit(f"s{i}", hist[i][:1000] + hist[i][1000 + rep:2040], max_new_tokens=1, ignore_eos=True)
        torch.cuda.synchronize()
        eng.step_log()
        for _ in range(3):
            eng.step()
        eng.drain()
        rows = [r.tolist() for r in eng.step_log()]
        full = [r[0] for r in rows if int(r[2]) == rows[0][2]]
        best = min(full) if best is None else min(best, min(full))
        eng.set_decode_prefill_cap(-1)
        for _ in range(4):
            eng.step(); eng.drain()
        eng.set_decode_prefill_cap(cap)
    res[cap] = (rows[0][1], rows[0][2], round(best, 3))
print("MIXED", os.environ.get("FT_GEMM_PAIR_MIN_TILES"), os.environ.get("FT_PREFILL_PLANES"), res)
eng.close()
