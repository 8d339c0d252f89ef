"""Quick numerics + timing check of the tcgen05 prefill attention against the fp32 oracle (development aid).
    python scripts/attn_tc_check.py"""
import os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from fasttalk_llm_microservice_b200 import build, kernels as K
from oracle import llama_oracle as O
import test_kernels_gpu as TK
build.build()
DEV = "cuda"
for nq, nkv, d, reqs in [(8, 2, 64, [(0, 5)]), (8, 2, 64, [(0, 64), (0, 65), (10, 1)]), (32, 8, 64, [(100, 130), (0, 200), (64, 64)]),
                         (32, 8, 64, [(0, 2048)]), (32, 8, 64, [(1824, 224)] * 3),
                         (3, 1, 128, [(250, 20)]), (24, 8, 128, [(33, 70), (0, 129)]), (32, 8, 128, [(0, 1024)]), (24, 8, 128, [(900, 224)] * 2),
                         (6, 2, 64, [(100, 130), (0, 65)])]:
    g = torch.Generator(device=DEV).manual_seed(len(reqs) * 31 + nq + d)
    totals = [p + n for p, n in reqs]
    kv, pt, ks, vs = TK.build_paged_kv(K, totals, nkv, d, g)
    qs = [0]
    for _, n in reqs:
        qs.append(qs[-1] + n)
    q = torch.randn(qs[-1], nq, d, device=DEV, generator=g).half()
    for tc in (True, False):
        hi, lo = K.attn_prefill(q, kv, pt, qs, list(range(len(reqs))), [p for p, _ in reqs], tensor_cores=tc)
        torch.cuda.synchronize()
        out = hi.float() + lo.float()
        worst = 0.0
        for r, (p0, n) in enumerate(reqs):
            ref = O.attention(q[qs[r]:qs[r + 1]].float(), ks[r].float(), vs[r].float(), p0)
            got = out[qs[r]:qs[r + 1]]
            worst = max(worst, ((got - ref).abs().max() / ref.abs().max()).item())
        print(f"nq={nq} nkv={nkv} reqs={reqs[:3]}{'...' if len(reqs) > 3 else ''} tensor_cores={tc}: rel err {worst:.2e} finite={bool(torch.isfinite(out).all())}", flush=True)
