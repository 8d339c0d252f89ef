"""Decode attention alone at the BASELINE configs[1] shape: launch time and per-CTA timeline."""
import os, sys
import numpy as np
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from fasttalk_llm_microservice_b200 import build, kernels as K, native as N
build.build()
lib = N.load()
dev = "cuda"
nq, nkv, d, B, ctxl = 32, 8, 64, 64, 2048
for n_ctas in (0, 148, 444, 592):
    g = torch.Generator(device=dev).manual_seed(1)
    n_layers = 6                                    # rotate over > L2 worth of KV
    pages = B * 33
    kvs = [torch.randn(pages, nkv, 2, 64, d, device=dev, generator=g).bfloat16() for _ in range(n_layers)]
    pt = torch.randperm(pages, device=dev, generator=g).int().reshape(B, 33).contiguous()
    q = torch.randn(B, nq, d, device=dev, generator=g).bfloat16()
    ctx = torch.full((B,), ctxl + 10, device=dev, dtype=torch.int32)
    nc = n_ctas or lib.ft_attn_decode_default_ctas(d)
    scratch = torch.empty(lib.ft_attn_decode_scratch_bytes(B, nkv, nq // nkv, d, nc) // 4, device=dev)
    tp = torch.empty(B + 1, device=dev, dtype=torch.int32)
    out = torch.empty(B, nq * d, device=dev, dtype=torch.bfloat16); out_lo = torch.empty_like(out)
    run = lambda kv: N.check(lib.ft_attn_decode(N.ptr(q), N.ptr(kv), N.ptr(pt), 33, N.ptr(ctx), N.ptr(tp), N.ptr(out),
                                                N.ptr(out_lo), N.ptr(scratch), B, nq, nkv, d, nc, N.cur_stream()), "attn")
    for i in range(4):
        run(kvs[i % n_layers])
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    iters = 30
    e0.record()
    for i in range(iters):
        run(kvs[i % n_layers])
    e1.record(); torch.cuda.synchronize()
    us = e0.elapsed_time(e1) * 1e3 / iters
    bytes_ = B * (ctxl + 10) * nkv * d * 2 * 2
    tr = np.zeros((nc, 4), dtype=np.uint64)
    lib.ft_debug_attn_trace(tr.ctypes.data, nc)
    t = tr.astype(np.int64); t0 = t[:, 0].min()
    rel = t - t0
    p = lambda a: f"min {a.min():6d} p50 {int(np.median(a)):6d} max {a.max():6d}"
    print(f"n_ctas={nc}: {us:6.1f} us/launch (plan+attn) {bytes_ / us / 1e3:6.0f} GB/s | entry {p(rel[:,0])} | first tile {p(rel[:,1]-rel[:,0])} after entry | "
          f"stream done {p(rel[:,2])} | exit {p(rel[:,3])}", flush=True)
