"""Microbenchmark of the tcgen05 GEMM on decode shapes: achieved weight-streaming bandwidth per
launch, rotating over enough weight copies that nothing is served from L2.

    python scripts/bench_gemm.py
"""
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from fasttalk_llm_microservice_b200 import build, kernels as K, native as N  # noqa: E402

build.build()
lib = N.load()
dev = "cuda"
torch.manual_seed(0)


def bench(name, n_out, Kd, T, hilo, splits, mode="partials", copies=None, iters=60):
    copies = copies or max(2, int(300e6 // (n_out * Kd * 2)) + 1)
    ws = [(torch.randn(n_out, Kd, device=dev) * 0.02).bfloat16() for _ in range(copies)]
    x, x_lo = K.split_hi_lo(torch.randn(T, Kd, device=dev))
    xl = x_lo if hilo else None
    s = splits or lib.ft_gemm_pick_splits(n_out, T, Kd, int(hilo))
    if mode == "partials":
        out = torch.empty(s, T, n_out, device=dev)
        run = lambda w: N.check(lib.ft_gemm_bf16(N.ptr(w), N.ptr(x), N.ptr(xl), N.ptr(out), 0, n_out, T, Kd, s, 0,
                                                 N.cur_stream()), "gemm")
    else:
        out = torch.empty(T, n_out // 2, device=dev, dtype=torch.bfloat16)
        out2 = torch.empty_like(out)
        run = lambda w: N.check(lib.ft_gemm_bf16(N.ptr(w), N.ptr(x), N.ptr(xl), N.ptr(out), N.ptr(out2), n_out, T, Kd,
                                                 1, 2, N.cur_stream()), "gemm")
    for i in range(5):
        run(ws[i % copies])
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(iters):
        run(ws[i % copies])
    e1.record()
    torch.cuda.synchronize()
    us = e0.elapsed_time(e1) * 1e3 / iters
    gbs = n_out * Kd * 2 / us / 1e3
    print(f"{name:10s} n_out={n_out:6d} K={Kd:5d} T={T:3d} hilo={int(hilo)} splits={s:2d} mode={mode:8s} "
          f"stages={os.environ.get('FT_GEMM_STAGES', 'auto'):>4s}: {us:7.2f} us/launch  {gbs:7.0f} GB/s of W", flush=True)
    return us


if __name__ == "__main__":
    shapes = [("qkv", 3072, 2048), ("o", 2048, 2048), ("gate_up", 16384, 2048), ("down", 2048, 8192),
              ("lm_head", 128256, 2048)]
    for hilo in (True, False):
        for name, n, k in shapes:
            bench(name, n, k, 64, hilo, 0)
    bench("gate_up", 16384, 2048, 64, True, 1, mode="swiglu")
    for st in ("2", "3", "4", "6"):
        os.environ["FT_GEMM_STAGES"] = st
        bench("gate_up", 16384, 2048, 64, True, 1)
        bench("down", 2048, 8192, 64, True, 9)
    os.environ.pop("FT_GEMM_STAGES")
    for s in (1, 2):
        bench("gate_up", 16384, 2048, 64, True, s)
    for s in (4, 6, 9, 12, 16):
        bench("down", 2048, 8192, 64, True, s)
    for T in (1, 16, 32, 128):
        bench("gate_up", 16384, 2048, T, True, 1)
