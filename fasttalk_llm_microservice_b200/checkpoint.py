"""Load a Hugging Face Llama checkpoint directory (config.json + *.safetensors) into the engine's
stacked weight layout (weights.py), and write one.

The reference names models only by tag and lets the engine container fetch them
(`app/utils/config.py:86`, `docker-compose.vllm.yml:39`); an in-process engine has to read the
files itself.  Tensors are copied one at a time straight into pre-allocated device storage, so
an 8B checkpoint never exists twice in host memory.

    FASTTALK_CHECKPOINT=/models/Llama-3.2-1B-Instruct python -m fasttalk_llm_microservice_b200.openai_server
"""
from __future__ import annotations

import json
import os
from typing import Dict, Tuple

import torch

from .model_spec import ModelSpec


def spec_from_hf_config(cfg: Dict, name: str = "checkpoint") -> ModelSpec:
    arch = (cfg.get("architectures") or ["LlamaForCausalLM"])[0]
    if cfg.get("model_type", "llama") != "llama" or "Llama" not in arch:
        raise ValueError(f"only Llama checkpoints are supported (got model_type={cfg.get('model_type')!r}, {arch})")
    for flag in ("attention_bias", "mlp_bias"):
        if cfg.get(flag):
            raise ValueError(f"{flag}=true is not supported")
    if cfg.get("quantization_config"):
        raise ValueError("quantised checkpoints are not supported: the engine computes on bf16 weights")
    heads = int(cfg["num_attention_heads"])
    head_dim = int(cfg.get("head_dim") or cfg["hidden_size"] // heads)
    rope = cfg.get("rope_parameters") or cfg.get("rope_scaling") or {}
    rtype = rope.get("rope_type", rope.get("type", "default"))
    if rtype not in ("llama3", "default"):
        raise ValueError(f"rope type {rtype!r} is not supported (llama3 / default only)")
    theta = float(rope.get("rope_theta", cfg.get("rope_theta", 500000.0)))
    max_pos = int(cfg.get("max_position_embeddings", 131072))
    llama3 = rtype == "llama3"
    return ModelSpec(
        name=name, hidden=int(cfg["hidden_size"]), ffn=int(cfg["intermediate_size"]),
        layers=int(cfg["num_hidden_layers"]), n_q_heads=heads,
        n_kv_heads=int(cfg.get("num_key_value_heads", heads)), head_dim=head_dim, vocab=int(cfg["vocab_size"]),
        tied_embeddings=bool(cfg.get("tie_word_embeddings", False)), rope_theta=theta,
        # factor 1 with equal low/high factors leaves every frequency untouched (weights.rope_inv_freq)
        rope_factor=float(rope.get("factor", 1.0)) if llama3 else 1.0,
        rope_low_freq_factor=float(rope.get("low_freq_factor", 1.0)) if llama3 else 1.0,
        rope_high_freq_factor=float(rope.get("high_freq_factor", 4.0)) if llama3 else 4.0,
        rope_original_max_pos=int(rope.get("original_max_position_embeddings", 8192)) if llama3 else max_pos,
        rms_eps=float(cfg.get("rms_norm_eps", 1e-5)), max_position=max_pos)


def _tensor_index(path: str) -> Dict[str, str]:
    """tensor name → file, for a single-file or sharded safetensors checkpoint."""
    idx = os.path.join(path, "model.safetensors.index.json")
    if os.path.exists(idx):
        with open(idx) as f:
            return {k: os.path.join(path, v) for k, v in json.load(f)["weight_map"].items()}
    from safetensors import safe_open
    files = sorted(f for f in os.listdir(path) if f.endswith(".safetensors"))
    if not files:
        raise FileNotFoundError(f"no .safetensors file in {path}")
    out = {}
    for fn in files:
        with safe_open(os.path.join(path, fn), framework="pt") as f:
            for k in f.keys():
                out[k] = os.path.join(path, fn)
    return out


def load_hf_checkpoint(path: str, device: str = "cpu", name: str = None) -> Tuple[ModelSpec, Dict[str, torch.Tensor]]:
    from safetensors import safe_open
    with open(os.path.join(path, "config.json")) as f:
        spec = spec_from_hf_config(json.load(f), name or os.path.basename(os.path.normpath(path)))
    for dim, what in ((spec.hidden, "hidden_size"), (spec.ffn, "intermediate_size"), (spec.q_dim, "heads*head_dim"),
                      (spec.kv_dim, "kv_heads*head_dim")):
        if dim % 64:
            raise ValueError(f"{what}={dim} must be a multiple of 64 for the sm_100a tile shapes")
    if spec.head_dim not in (64, 128):
        raise ValueError(f"head_dim={spec.head_dim}: the attention kernels are built for 64 and 128")
    index = _tensor_index(path)
    L, h, q, kv, f_ = spec.layers, spec.hidden, spec.q_dim, spec.kv_dim, spec.ffn
    bf = dict(dtype=torch.bfloat16, device=device)
    w = {"embed": torch.empty(spec.vocab, h, **bf), "final_norm": torch.empty(h, **bf),
         "attn_norm": torch.empty(L, h, **bf), "mlp_norm": torch.empty(L, h, **bf),
         "wqkv": torch.empty(L, q + 2 * kv, h, **bf), "wo": torch.empty(L, h, q, **bf),
         "wgu": torch.empty(L, 2 * f_, h, **bf), "wd": torch.empty(L, h, f_, **bf)}
    handles: Dict[str, object] = {}

    def put(dst: torch.Tensor, key: str):
        if key not in index:
            raise KeyError(f"checkpoint has no tensor {key!r}")
        fn = index[key]
        if fn not in handles:
            handles[fn] = safe_open(fn, framework="pt")
        t = handles[fn].get_tensor(key)
        if tuple(t.shape) != tuple(dst.shape):
            raise ValueError(f"{key}: shape {tuple(t.shape)} does not match config ({tuple(dst.shape)})")
        dst.copy_(t)                                    # casts fp16/fp32 checkpoints to bf16

    put(w["embed"], "model.embed_tokens.weight")
    put(w["final_norm"], "model.norm.weight")
    for l in range(L):
        p = f"model.layers.{l}."
        put(w["attn_norm"][l], p + "input_layernorm.weight")
        put(w["mlp_norm"][l], p + "post_attention_layernorm.weight")
        put(w["wqkv"][l, :q], p + "self_attn.q_proj.weight")
        put(w["wqkv"][l, q:q + kv], p + "self_attn.k_proj.weight")
        put(w["wqkv"][l, q + kv:], p + "self_attn.v_proj.weight")
        put(w["wo"][l], p + "self_attn.o_proj.weight")
        put(w["wgu"][l, :f_], p + "mlp.gate_proj.weight")
        put(w["wgu"][l, f_:], p + "mlp.up_proj.weight")
        put(w["wd"][l], p + "mlp.down_proj.weight")
    if spec.tied_embeddings or "lm_head.weight" not in index:
        w["lm_head"] = w["embed"]
    else:
        w["lm_head"] = torch.empty(spec.vocab, h, **bf)
        put(w["lm_head"], "lm_head.weight")
    handles.clear()
    return spec, w


def hf_config_of(spec: ModelSpec) -> Dict:
    return {"architectures": ["LlamaForCausalLM"], "model_type": "llama", "vocab_size": spec.vocab,
            "hidden_size": spec.hidden, "intermediate_size": spec.ffn, "num_hidden_layers": spec.layers,
            "num_attention_heads": spec.n_q_heads, "num_key_value_heads": spec.n_kv_heads, "head_dim": spec.head_dim,
            "rms_norm_eps": spec.rms_eps, "max_position_embeddings": spec.max_position,
            "tie_word_embeddings": spec.tied_embeddings, "attention_bias": False, "mlp_bias": False,
            "hidden_act": "silu", "torch_dtype": "bfloat16", "rope_theta": spec.rope_theta,
            "rope_scaling": {"rope_type": "llama3", "factor": spec.rope_factor,
                             "low_freq_factor": spec.rope_low_freq_factor, "high_freq_factor": spec.rope_high_freq_factor,
                             "original_max_position_embeddings": spec.rope_original_max_pos}}


def save_hf_checkpoint(spec: ModelSpec, w: Dict[str, torch.Tensor], path: str, shards: int = 1) -> None:
    """The inverse of load_hf_checkpoint (tests; exporting random-init weights for other engines)."""
    from safetensors.torch import save_file
    os.makedirs(path, exist_ok=True)
    with open(os.path.join(path, "config.json"), "w") as f:
        json.dump(hf_config_of(spec), f, indent=1)
    q, kv, f_ = spec.q_dim, spec.kv_dim, spec.ffn
    sd = {"model.embed_tokens.weight": w["embed"], "model.norm.weight": w["final_norm"]}
    for l in range(spec.layers):
        p = f"model.layers.{l}."
        sd[p + "input_layernorm.weight"] = w["attn_norm"][l]
        sd[p + "post_attention_layernorm.weight"] = w["mlp_norm"][l]
        sd[p + "self_attn.q_proj.weight"] = w["wqkv"][l][:q]
        sd[p + "self_attn.k_proj.weight"] = w["wqkv"][l][q:q + kv]
        sd[p + "self_attn.v_proj.weight"] = w["wqkv"][l][q + kv:]
        sd[p + "self_attn.o_proj.weight"] = w["wo"][l]
        sd[p + "mlp.gate_proj.weight"] = w["wgu"][l][:f_]
        sd[p + "mlp.up_proj.weight"] = w["wgu"][l][f_:]
        sd[p + "mlp.down_proj.weight"] = w["wd"][l]
    if not spec.tied_embeddings:
        sd["lm_head.weight"] = w["lm_head"]
    sd = {k: v.detach().cpu().contiguous() for k, v in sd.items()}
    if shards <= 1:
        save_file(sd, os.path.join(path, "model.safetensors"), metadata={"format": "pt"})
        return
    names = sorted(sd)
    per = -(-len(names) // shards)
    weight_map = {}
    for i in range(shards):
        part = names[i * per:(i + 1) * per]
        if not part:
            continue
        fn = f"model-{i + 1:05d}-of-{shards:05d}.safetensors"
        save_file({k: sd[k] for k in part}, os.path.join(path, fn), metadata={"format": "pt"})
        weight_map.update({k: fn for k in part})
    with open(os.path.join(path, "model.safetensors.index.json"), "w") as f:
        json.dump({"metadata": {}, "weight_map": weight_map}, f)
