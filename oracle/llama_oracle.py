"""ORACLE — test infrastructure only.  Never imported by the product path.

fp32 PyTorch restatement of the generation arithmetic the reference delegates to
Ollama / vLLM (`app/core/ollama_handler.py:142-177`, `app/core/vllm_handler.py:177,271`).
The reference holds no model code and no golden vectors for this path
(SURVEY.md §4, §8c), and neither engine is installable here (Ollama: unpinned
`ollama/ollama:latest`, docker-compose.cpu.yml:11; vLLM: unpinned
`vllm/vllm-openai:latest`, docker-compose.vllm.yml:12), so against the reference
itself **parity is unpinned**.  What pins this file instead: tests/test_oracle.py
checks it against transformers 5.5.0 `LlamaForCausalLM` (the public
implementation of the named architecture) on the same weights, and against the
committed vectors in tests/golden/ that tests/golden/make_golden.py generated
from that transformers model.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl
reference legs may import this module.

Architecture followed (public Llama-3): pre-norm decoder; RMSNorm with fp32
accumulation, eps 1e-5; GQA; RoPE in the "rotate_half" layout with llama3
frequency scaling; SwiGLU down(silu(gate(x)) * up(x)); no biases; causal mask;
fp32 logits.  Weights are the product's bf16 tensors upcast to fp32, so weight
rounding is not charged to the kernels under test.
"""
from __future__ import annotations

import math
from typing import Dict, List, Optional, Tuple

import torch


# --------------------------------------------------------------------------- rope
def rope_inv_freq(head_dim: int, theta: float, factor: float, low: float, high: float,
                  orig_max: int) -> torch.Tensor:
    """llama3 rope frequency scaling (public Llama-3.1 recipe)."""
    inv = 1.0 / (theta ** (torch.arange(0, head_dim, 2, dtype=torch.float32) / head_dim))
    out = inv.clone()
    for i in range(inv.numel()):
        wl = 2 * math.pi / inv[i].item()
        if wl > orig_max / low:            # low-frequency band: fully scaled
            out[i] = inv[i] / factor
        elif wl >= orig_max / high:        # medium band: smooth interpolation
            s = (orig_max / wl - low) / (high - low)
            out[i] = (1 - s) * inv[i] / factor + s * inv[i]
    return out


def rope_cos_sin(spec, positions: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
    inv = rope_inv_freq(spec.head_dim, spec.rope_theta, spec.rope_factor,
                        spec.rope_low_freq_factor, spec.rope_high_freq_factor,
                        spec.rope_original_max_pos)
    ang = torch.outer(positions.to(torch.float32).cpu(), inv)      # [T, d/2], always on the host
    return ang.cos().to(positions.device), ang.sin().to(positions.device)


def apply_rope(x: torch.Tensor, cos: torch.Tensor, sin: torch.Tensor) -> torch.Tensor:
    """x [T, heads, d]; rotate_half layout: pairs are (i, i + d/2)."""
    d2 = x.shape[-1] // 2
    x1, x2 = x[..., :d2], x[..., d2:]
    c, s = cos[:, None, :], sin[:, None, :]
    return torch.cat([x1 * c - x2 * s, x2 * c + x1 * s], dim=-1)


# --------------------------------------------------------------------------- ops
def rmsnorm(x: torch.Tensor, w: torch.Tensor, eps: float) -> torch.Tensor:
    x = x.to(torch.float32)
    return x * torch.rsqrt(x.pow(2).mean(-1, keepdim=True) + eps) * w.to(torch.float32)


def attention(q: torch.Tensor, k: torch.Tensor, v: torch.Tensor, q_pos0: int) -> torch.Tensor:
    """q [Tq, nq, d], k/v [Tk, nkv, d]; query i sits at absolute position q_pos0+i and
    sees keys 0..q_pos0+i.  Returns [Tq, nq*d] fp32."""
    Tq, nq, d = q.shape
    Tk, nkv, _ = k.shape
    g = nq // nkv
    k = k.repeat_interleave(g, dim=1)
    v = v.repeat_interleave(g, dim=1)
    s = torch.einsum("qhd,khd->hqk", q, k) / math.sqrt(d)
    qi = torch.arange(Tq, device=q.device)[:, None] + q_pos0
    ki = torch.arange(Tk, device=q.device)[None, :]
    s = s.masked_fill((ki > qi)[None], float("-inf"))
    p = torch.softmax(s, dim=-1)
    return torch.einsum("hqk,khd->qhd", p, v).reshape(Tq, nq * d)


def swiglu(gu: torch.Tensor) -> torch.Tensor:
    f = gu.shape[-1] // 2
    return torch.nn.functional.silu(gu[..., :f]) * gu[..., f:]


# --------------------------------------------------------------------------- model
class OracleState:
    """Contiguous fp32 KV cache of one sequence."""

    def __init__(self, spec):
        self.k: List[Optional[torch.Tensor]] = [None] * spec.layers
        self.v: List[Optional[torch.Tensor]] = [None] * spec.layers
        self.len = 0


def forward(weights: Dict[str, torch.Tensor], spec, tokens: torch.Tensor,
            state: Optional[OracleState] = None, last_only: bool = False) -> torch.Tensor:
    """Run `tokens` (1-D int64) through the model after whatever `state` holds.
    Returns fp32 logits [T, V] (or [1, V] with last_only)."""
    if state is None:
        state = OracleState(spec)
    f32 = lambda t: t.to(torch.float32)
    T = tokens.numel()
    pos0 = state.len
    dev = weights["embed"].device
    tokens = tokens.to(dev)
    cos, sin = rope_cos_sin(spec, torch.arange(pos0, pos0 + T, device=dev))
    x = f32(weights["embed"][tokens])
    for l in range(spec.layers):
        xn = rmsnorm(x, weights["attn_norm"][l], spec.rms_eps)
        qkv = xn @ f32(weights["wqkv"][l]).T
        q = qkv[:, : spec.q_dim].reshape(T, spec.n_q_heads, spec.head_dim)
        k = qkv[:, spec.q_dim: spec.q_dim + spec.kv_dim].reshape(T, spec.n_kv_heads, spec.head_dim)
        v = qkv[:, spec.q_dim + spec.kv_dim:].reshape(T, spec.n_kv_heads, spec.head_dim)
        q, k = apply_rope(q, cos, sin), apply_rope(k, cos, sin)
        state.k[l] = k if state.k[l] is None else torch.cat([state.k[l], k])
        state.v[l] = v if state.v[l] is None else torch.cat([state.v[l], v])
        a = attention(q, state.k[l], state.v[l], pos0)
        x = x + a @ f32(weights["wo"][l]).T
        xn = rmsnorm(x, weights["mlp_norm"][l], spec.rms_eps)
        x = x + swiglu(xn @ f32(weights["wgu"][l]).T) @ f32(weights["wd"][l]).T
    state.len = pos0 + T
    if last_only:
        x = x[-1:]
    xn = rmsnorm(x, weights["final_norm"], spec.rms_eps)
    return xn @ f32(weights["lm_head"]).T


@torch.no_grad()
def greedy_generate(weights, spec, prompt: torch.Tensor, n_new: int,
                    return_logits: bool = False):
    """Greedy decode (argmax, lowest index on ties like torch.argmax) ignoring EOS."""
    state = OracleState(spec)
    logits = forward(weights, spec, prompt, state, last_only=True)
    out, all_logits = [], []
    for _ in range(n_new):
        all_logits.append(logits[0])
        tok = int(torch.argmax(logits[0]))
        out.append(tok)
        logits = forward(weights, spec, torch.tensor([tok]), state, last_only=True)
    all_logits = [l.cpu() for l in all_logits]
    if return_logits:
        return out, torch.stack(all_logits)
    return out


# --------------------------------------------------------------------------- sampling
def sampling_support(logits: torch.Tensor, temperature: float, top_k: int, top_p: float):
    """The probability vector a temperature / top-k / top-p sampler draws from
    (vLLM / Ollama semantics: scale by 1/T, keep the k largest, then the smallest
    prefix of the sorted distribution whose mass reaches top_p, renormalise)."""
    x = logits.to(torch.float64) / max(temperature, 1e-6)
    p = torch.softmax(x, -1)
    order = torch.argsort(p, descending=True, stable=True)
    keep = torch.ones_like(p, dtype=torch.bool)
    if top_k and top_k > 0:
        keep[order[top_k:]] = False
    ps = torch.where(keep, p, torch.zeros_like(p))
    ps = ps / ps.sum()
    if top_p < 1.0:
        sp = ps[order]
        csum = torch.cumsum(sp, 0)
        cut = int(torch.searchsorted(csum, torch.tensor(top_p, dtype=csum.dtype)).item())
        drop = order[cut + 1:]
        ps[drop] = 0
        ps = ps / ps.sum()
    return ps
