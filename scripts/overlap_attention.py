"""Experiment: in a mixed step the decode-attention kernel (HBM-bound) and the tcgen05 prefill-attention kernel (latency /
SFU-bound, one CTA per SM) run one after the other.  Would two streams overlap them?  Shapes of the 1B mixed step:
58 decode rows at ctx 2048 + five prefill chunks (4 x 224 + 70 rows) over 1824 cached tokens.

    python scripts/overlap_attention.py        -> us per pair of launches: back to back on one stream / on two streams
"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from fasttalk_llm_microservice_b200 import build, native as N  # noqa: E402

build.build()
lib = N.load()
DEV = "cuda"
nq, nkv, d = 32, 8, 64
g = torch.Generator(device=DEV).manual_seed(1)
i32 = lambda v: torch.tensor(v, device=DEV, dtype=torch.int32)

# decode side
B, ctx = 58, 2048
ppseq = ctx // 64
pages_d = B * ppseq
qd = torch.randn(B, nq, d, device=DEV, generator=g).half()
ctxs = torch.full((B,), ctx, device=DEV, dtype=torch.int32)
pt_d = torch.randperm(pages_d, device=DEV, generator=g).int().reshape(B, ppseq).contiguous()
kvs_d = [torch.randn(pages_d, nkv, 2, 64, d, device=DEV, generator=g).half() for _ in range(3)]
n_ctas = lib.ft_attn_decode_default_ctas(d)
scratch = torch.empty(lib.ft_attn_decode_scratch_bytes(B, nkv, nq // nkv, d, n_ctas) // 4, device=DEV)
tpfx = torch.empty(B + 1, device=DEV, dtype=torch.int32)
out_d = torch.empty(B, nq * d, device=DEV, dtype=torch.bfloat16)
out_d_lo = torch.empty_like(out_d)

# prefill side
reqs = [(1824, 224)] * 4 + [(1824, 70)]
totals = [p + n for p, n in reqs]
n_pages = [(t + 63) // 64 for t in totals]
maxp = max(n_pages)
kv_p = torch.randn(sum(n_pages), nkv, 2, 64, d, device=DEV, generator=g).half()
pt_p = torch.zeros(len(reqs), maxp, dtype=torch.int32, device=DEV)
off = 0
for r, n in enumerate(n_pages):
    pt_p[r, :n] = torch.arange(off, off + n, device=DEV, dtype=torch.int32)
    off += n
qs = [0]
for _, n in reqs:
    qs.append(qs[-1] + n)
T = qs[-1]
qp = torch.randn(T, nq, d, device=DEV, generator=g).half()
tile_req, tile_q0 = [], []
for r, (_, n) in enumerate(reqs):
    for q0 in range(0, n, 64):
        tile_req.append(r); tile_q0.append(q0)
t_qs, t_slot, t_pos0, t_treq, t_tq0 = i32(qs), i32(list(range(len(reqs)))), i32([p for p, _ in reqs]), i32(tile_req), i32(tile_q0)
out_p = torch.zeros(T, nq * d, device=DEV, dtype=torch.bfloat16)
out_p_lo = torch.zeros_like(out_p)


CTAS = [n_ctas]


def decode(kv, stream):
    N.check(lib.ft_attn_decode(N.ptr(qd), N.ptr(kv), N.ptr(pt_d), ppseq, N.ptr(ctxs), N.ptr(tpfx), N.ptr(out_d), N.ptr(out_d_lo),
                               N.ptr(scratch), B, nq, nkv, d, CTAS[0], stream), "ft_attn_decode")


def prefill(stream):
    N.check(lib.ft_attn_prefill(N.ptr(qp), N.ptr(kv_p), N.ptr(pt_p), maxp, N.ptr(t_qs), N.ptr(t_slot), N.ptr(t_pos0), N.ptr(t_treq),
                                N.ptr(t_tq0), len(tile_req), N.ptr(out_p), N.ptr(out_p_lo), nq, nkv, d, T, stream), "ft_attn_prefill")


s0 = torch.cuda.current_stream()
s1 = torch.cuda.Stream()
h0, h1 = s0.cuda_stream, s1.cuda_stream


def timed(fn, iters=40):
    for i in range(5):
        fn(i)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(s0)
    for i in range(iters):
        fn(i)
    e1.record(s0)
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e3 / iters


def serial(i):
    decode(kvs_d[i % 3], h0)
    prefill(h0)


def forked(prefill_first):
    def fn(i):
        fork, join = torch.cuda.Event(), torch.cuda.Event()
        fork.record(s0)
        s1.wait_event(fork)
        if prefill_first:
            prefill(h1)
            decode(kvs_d[i % 3], h0)
        else:
            decode(kvs_d[i % 3], h0)
            prefill(h1)
        join.record(s1)
        s0.wait_event(join)
    return fn


print(f"decode alone            {timed(lambda i: decode(kvs_d[i % 3], h0)):7.1f} us")
print(f"prefill alone           {timed(lambda i: prefill(h0)):7.1f} us  ({T} rows, {len(tile_req)} x {nkv} CTAs)")
print(f"one stream, back to back {timed(serial):7.1f} us")
print(f"two streams, prefill launched first {timed(forked(True)):7.1f} us")
print(f"two streams, decode launched first  {timed(forked(False)):7.1f} us")
for c in (148, 222):
    CTAS[0] = c
    print(f"decode with {c} CTAs: alone {timed(lambda i: decode(kvs_d[i % 3], h0)):7.1f} us, two streams (prefill first) "
          f"{timed(forked(True)):7.1f} us, (decode first) {timed(forked(False)):7.1f} us")
