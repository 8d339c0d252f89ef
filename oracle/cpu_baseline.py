"""ORACLE-side CPU baseline — test/bench infrastructure only (never on the product path).

Batched decode step of the oracle arithmetic (oracle/llama_oracle.py) on the host cores:
the CPU stand-in for the reference's CPU path, which is Ollama/llama.cpp on host cores
(`app/core/ollama_handler.py:142-177` with COMPUTE_DEVICE=cpu, docker-compose.cpu.yml:11) and
cannot be installed here.  fp32 PyTorch, all host threads, weights read once
per step for the whole batch (what any batching CPU engine does), per-sequence attention over
a contiguous fp32 KV cache.  It is NOT llama.cpp: a quantised GGUF would be faster.
"""
from __future__ import annotations

import math
import os
import time
from typing import Dict, List

import torch

from . import llama_oracle as O


class CpuDecodeBench:
    """B sequences, each with `ctx` cached tokens (synthetic KV of the right shape), stepping
    one token at a time.  KV is preallocated so a step costs what a real decode step costs."""

    def __init__(self, spec, weights_f32: Dict[str, torch.Tensor], batch: int, ctx: int, max_new: int, seed: int = 0):
        self.spec, self.w, self.B = spec, weights_f32, batch
        g = torch.Generator().manual_seed(seed)
        cap = ctx + max_new
        # synthetic KV: values do not change the cost of a step; one random sequence tiled over
        # the batch keeps the set-up short
        one = (spec.layers, 1, cap, spec.n_kv_heads, spec.head_dim)
        self.k = (torch.randn(one, generator=g) * 0.5).repeat(1, batch, 1, 1, 1)
        self.v = (torch.randn(one, generator=g) * 0.5).repeat(1, batch, 1, 1, 1)
        self.len = ctx
        self.tok = torch.randint(0, min(spec.vocab, 128000), (batch,), generator=g)

    @torch.no_grad()
    def step(self) -> torch.Tensor:
        s, w, B = self.spec, self.w, self.B
        pos = self.len
        cos, sin = O.rope_cos_sin(s, torch.full((1,), pos))
        x = w["embed"][self.tok]
        scale = 1.0 / math.sqrt(s.head_dim)
        for l in range(s.layers):
            xn = O.rmsnorm(x, w["attn_norm"][l], s.rms_eps)
            qkv = xn @ w["wqkv"][l].T
            q = qkv[:, : s.q_dim].reshape(B, s.n_q_heads, s.head_dim)
            k = qkv[:, s.q_dim: s.q_dim + s.kv_dim].reshape(B, s.n_kv_heads, s.head_dim)
            v = qkv[:, s.q_dim + s.kv_dim:].reshape(B, s.n_kv_heads, s.head_dim)
            q, k = O.apply_rope(q, cos, sin), O.apply_rope(k, cos, sin)
            self.k[l, :, pos] = k
            self.v[l, :, pos] = v
            kk = self.k[l, :, : pos + 1]                      # [B, ctx, nkv, d]
            vv = self.v[l, :, : pos + 1]
            qg = q.reshape(B, s.n_kv_heads, s.group, s.head_dim)
            sc = torch.einsum("bhgd,bthd->bhgt", qg, kk) * scale
            p = torch.softmax(sc, -1)
            a = torch.einsum("bhgt,bthd->bhgd", p, vv).reshape(B, s.q_dim)
            x = x + a @ w["wo"][l].T
            xn = O.rmsnorm(x, w["mlp_norm"][l], s.rms_eps)
            x = x + O.swiglu(xn @ w["wgu"][l].T) @ w["wd"][l].T
        logits = O.rmsnorm(x, w["final_norm"], s.rms_eps) @ w["lm_head"].T
        self.tok = torch.argmax(logits, -1)
        self.len += 1
        return self.tok


def make_weights(spec, seed: int = 0) -> Dict[str, torch.Tensor]:
    """fp32 copies of bf16-rounded normal(0, 0.02) weights (same recipe as the product's)."""
    g = torch.Generator().manual_seed(seed)
    L, h = spec.layers, spec.hidden

    def rnd(*shape):
        # one random row block tiled over the leading dims: cost-equivalent, quick to build
        base = (torch.randn(*shape[-2:], generator=g) * 0.02).to(torch.bfloat16).float()
        return base.repeat(*shape[:-2], 1, 1) if len(shape) > 2 else base

    w = {"embed": rnd(spec.vocab, h), "final_norm": torch.ones(h), "attn_norm": torch.ones(L, h),
         "mlp_norm": torch.ones(L, h), "wqkv": rnd(L, spec.qkv_dim, h), "wo": rnd(L, h, spec.q_dim),
         "wgu": rnd(L, 2 * spec.ffn, h), "wd": rnd(L, h, spec.ffn)}
    w["lm_head"] = w["embed"] if spec.tied_embeddings else rnd(spec.vocab, h)
    return w


def run(spec, batch: int, ctx: int, steps: int, warmup: int, threads: int = 0, seed: int = 0, weights=None):
    """Returns (tokens_per_s, ms_per_step, cores, description)."""
    threads = threads or (os.cpu_count() or 1)
    torch.set_num_threads(threads)
    w = weights if weights is not None else make_weights(spec, seed)
    b = CpuDecodeBench(spec, w, batch, ctx, steps + warmup + 1, seed)
    for _ in range(warmup):
        b.step()
    t0 = time.perf_counter()
    for _ in range(steps):
        b.step()
    dt = time.perf_counter() - t0
    desc = (f"{steps} decode steps of {batch} sessions at ctx {ctx} (synthetic fp32 KV), "
            f"fp32 PyTorch oracle port, {threads} threads")
    return batch * steps / dt, 1e3 * dt / steps, threads, desc
