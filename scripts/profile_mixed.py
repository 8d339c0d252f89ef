"""One MIXED step for ncu: ~56 sessions decoding at ctx ~2k while four sessions re-send a 2048-token history of which 224
tokens are new (prefill cap 768 -> decode rows + prompt tokens fill seven 128-token tiles), inside a cudaProfilerStart/Stop window.

    ncu --profile-from-start off --metrics gpu__time_duration.sum --csv python scripts/profile_mixed.py
"""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from fasttalk_llm_microservice_b200 import build
from fasttalk_llm_microservice_b200.engine import Engine, EngineLimits
from fasttalk_llm_microservice_b200.model_spec import get_spec
from fasttalk_llm_microservice_b200.weights import init_weights
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
assert eng.stats()["last_prefill_tokens"] == 0 and eng.stats()["last_batch"] == 56, eng.stats()
eng.set_decode_prefill_cap(768)
for i in range(56, 64):
    eng.submit(f"s{i}", hist[i], max_new_tokens=8, ignore_eos=True)
torch.cuda.synchronize()
eng.step_log()
torch.cuda.cudart().cudaProfilerStart()
eng.step()
torch.cuda.cudart().cudaProfilerStop()
print("mixed step [device ms, decode rows, prompt tokens, host ms]:", eng.step_log()[-1].tolist())
eng.close()
