"""Deterministic random-init weights of the named architecture (no checkpoints on
the box: BASELINE.json north_star, SURVEY.md §8d "Synthetic inputs").

Layout (all bf16, row-major, `[out_features, in_features]` like nn.Linear):
  embed      [V, h]
  lm_head    [V, h]            (same storage as embed when tied)
  final_norm [h]
  attn_norm  [L, h]     wqkv [L, q_dim + 2*kv_dim, h]   (rows: q | k | v)
  wo         [L, h, q_dim]
  mlp_norm   [L, h]     wgu  [L, 2*ffn, h]              (rows: gate | up)
  wd         [L, h, ffn]
"""
from __future__ import annotations

from typing import Dict

import torch

from .model_spec import ModelSpec


def init_weights(spec: ModelSpec, seed: int = 0, device: str = "cpu",
                 std: float = 0.02, norm_jitter: float = 0.0) -> Dict[str, torch.Tensor]:
    """normal(0, std) projections (HF `initializer_range`), norm weights 1 (+ optional
    jitter so tests can tell a missing norm-weight multiply), rounded to bf16.

    The generator runs on `device`; CPU and CUDA streams differ, so parity tests
    create weights once and share the same tensors between oracle and engine.
    """
    g = torch.Generator(device=device)
    g.manual_seed(seed)

    def rnd(*shape):
        return (torch.randn(*shape, generator=g, device=device, dtype=torch.float32) * std).to(torch.bfloat16)

    def norm(*shape):
        w = torch.ones(*shape, device=device, dtype=torch.float32)
        if norm_jitter:
            w = w + norm_jitter * torch.randn(*shape, generator=g, device=device, dtype=torch.float32)
        return w.to(torch.bfloat16)

    L, h = spec.layers, spec.hidden
    w = {
        "embed": rnd(spec.vocab, h),
        "final_norm": norm(h),
        "attn_norm": norm(L, h),
        "mlp_norm": norm(L, h),
    }
    # per-layer generation keeps the peak fp32 temporary small for the 8B config
    w["wqkv"] = torch.stack([rnd(spec.qkv_dim, h) for _ in range(L)])
    w["wo"] = torch.stack([rnd(h, spec.q_dim) for _ in range(L)])
    w["wgu"] = torch.stack([rnd(2 * spec.ffn, h) for _ in range(L)])
    w["wd"] = torch.stack([rnd(h, spec.ffn) for _ in range(L)])
    w["lm_head"] = w["embed"] if spec.tied_embeddings else rnd(spec.vocab, h)
    return w


def weights_to(w: Dict[str, torch.Tensor], device) -> Dict[str, torch.Tensor]:
    out = {k: v.to(device) for k, v in w.items() if k != "lm_head"}
    out["lm_head"] = out["embed"] if w["lm_head"] is w["embed"] else w["lm_head"].to(device)
    return out


def rope_tables(spec: ModelSpec, max_pos: int, device="cpu"):
    """cos/sin tables [max_pos, head_dim/2] fp32 with llama3 frequency scaling.

    Same arithmetic as the oracle (oracle/llama_oracle.py: rope_inv_freq), done
    once on the host side so the CUDA kernel reads exactly the oracle's angles.
    """
    inv_freq = rope_inv_freq(spec)
    pos = torch.arange(max_pos, dtype=torch.float32)
    ang = torch.outer(pos, inv_freq)            # fp32, like HF's rotary embedding
    return ang.cos().to(device).contiguous(), ang.sin().to(device).contiguous()


def rope_inv_freq(spec: ModelSpec) -> torch.Tensor:
    """llama3 rope scaling of the base inverse frequencies (public Llama-3.1 recipe)."""
    import math
    d = spec.head_dim
    inv = 1.0 / (spec.rope_theta ** (torch.arange(0, d, 2, dtype=torch.float32) / d))
    low_wl = spec.rope_original_max_pos / spec.rope_low_freq_factor
    high_wl = spec.rope_original_max_pos / spec.rope_high_freq_factor
    wl = 2 * math.pi / inv
    scaled = torch.where(wl > low_wl, inv / spec.rope_factor, inv)
    smooth = (spec.rope_original_max_pos / wl - spec.rope_low_freq_factor) / (
        spec.rope_high_freq_factor - spec.rope_low_freq_factor)
    mid = (1 - smooth) * scaled / spec.rope_factor + smooth * scaled
    is_mid = (wl >= high_wl) & (wl <= low_wl)
    return torch.where(is_mid, mid, scaled)


def interleave_gate_up(wgu: torch.Tensor) -> torch.Tensor:
    """[..., 2f, h] rows (gate | up) -> rows interleaved in blocks of 64 (gate 64i.., up 64i..), the
    layout the GEMM's fused SwiGLU epilogue reads (include/fasttalk_b200.h: ft_weights.wgu)."""
    f2, h = wgu.shape[-2], wgu.shape[-1]
    f = f2 // 2
    assert f % 64 == 0, "ffn must be a multiple of 64"
    lead = wgu.shape[:-2]
    g = wgu[..., :f, :].reshape(*lead, f // 64, 1, 64, h)
    u = wgu[..., f:, :].reshape(*lead, f // 64, 1, 64, h)
    return torch.cat([g, u], dim=-3).reshape(*lead, f2, h).contiguous()
