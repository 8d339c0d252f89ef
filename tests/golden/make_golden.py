"""Generate tests/golden/*.npz from transformers' LlamaForCausalLM (run in the
authoring container; the GPU box only reads the committed vectors).

    python tests/golden/make_golden.py

The reference repo has no model code or golden vectors for generation
(SURVEY.md §4), so the public transformers implementation of the named
architecture is the pin for oracle/llama_oracle.py.
"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

from fasttalk_llm_microservice_b200.model_spec import SPECS, TINY_D64, TINY_D128  # noqa: E402
from fasttalk_llm_microservice_b200.weights import init_weights  # noqa: E402


def hf_model(spec, w):
    from transformers import LlamaConfig, LlamaForCausalLM
    cfg = LlamaConfig(
        vocab_size=spec.vocab, hidden_size=spec.hidden, intermediate_size=spec.ffn,
        num_hidden_layers=spec.layers, num_attention_heads=spec.n_q_heads,
        num_key_value_heads=spec.n_kv_heads, head_dim=spec.head_dim, rms_norm_eps=spec.rms_eps,
        max_position_embeddings=spec.max_position, tie_word_embeddings=spec.tied_embeddings,
        attention_bias=False, mlp_bias=False,
        rope_parameters={"rope_type": "llama3", "rope_theta": spec.rope_theta,
                         "factor": spec.rope_factor, "low_freq_factor": spec.rope_low_freq_factor,
                         "high_freq_factor": spec.rope_high_freq_factor,
                         "original_max_position_embeddings": spec.rope_original_max_pos},
    )
    cfg._attn_implementation = "eager"
    m = LlamaForCausalLM(cfg).to(torch.float32).eval()
    sd = {"model.embed_tokens.weight": w["embed"].float(), "model.norm.weight": w["final_norm"].float()}
    q, kv, f = spec.q_dim, spec.kv_dim, spec.ffn
    for l in range(spec.layers):
        p = f"model.layers.{l}."
        sd[p + "input_layernorm.weight"] = w["attn_norm"][l].float()
        sd[p + "post_attention_layernorm.weight"] = w["mlp_norm"][l].float()
        sd[p + "self_attn.q_proj.weight"] = w["wqkv"][l][:q].float()
        sd[p + "self_attn.k_proj.weight"] = w["wqkv"][l][q:q + kv].float()
        sd[p + "self_attn.v_proj.weight"] = w["wqkv"][l][q + kv:].float()
        sd[p + "self_attn.o_proj.weight"] = w["wo"][l].float()
        sd[p + "mlp.gate_proj.weight"] = w["wgu"][l][:f].float()
        sd[p + "mlp.up_proj.weight"] = w["wgu"][l][f:].float()
        sd[p + "mlp.down_proj.weight"] = w["wd"][l].float()
    sd["lm_head.weight"] = w["lm_head"].float()
    missing, unexpected = m.load_state_dict(sd, strict=False)
    assert not unexpected, unexpected
    assert all("rotary" in k or "inv_freq" in k for k in missing), missing
    return m


@torch.no_grad()
def main():
    out_dir = os.path.dirname(os.path.abspath(__file__))
    for spec, seed, T, n_new in ((TINY_D64, 11, 40, 24), (TINY_D128, 12, 33, 24)):
        # std 0.08 keeps the tiny model's logits well separated; jittered norm weights
        w = init_weights(spec, seed=seed, std=0.08, norm_jitter=0.1)
        g = torch.Generator().manual_seed(seed + 100)
        prompt = torch.randint(0, spec.vocab, (T,), generator=g)
        m = hf_model(spec, w)
        logits = m(prompt[None]).logits[0].float()
        gen = m.generate(prompt[None], max_new_tokens=n_new, do_sample=False,
                         eos_token_id=None, pad_token_id=0)[0, T:]
        np.savez_compressed(
            os.path.join(out_dir, f"hf_{spec.name}.npz"),
            seed=seed, std=0.08, norm_jitter=0.1, prompt=prompt.numpy(),
            logits=logits.numpy(), greedy=gen.numpy())
        print(spec.name, "logits", tuple(logits.shape), "greedy", gen.tolist()[:8], "...")

    # llama3-scaled inverse frequencies of the three named configs, from transformers
    from transformers.modeling_rope_utils import ROPE_INIT_FUNCTIONS
    from transformers import LlamaConfig
    inv = {}
    for name in ("llama-3.2-1b", "llama-3.2-3b", "llama-3.1-8b"):
        s = SPECS[name]
        cfg = LlamaConfig(hidden_size=s.hidden, num_attention_heads=s.n_q_heads, head_dim=s.head_dim,
                          max_position_embeddings=s.max_position,
                          rope_parameters={"rope_type": "llama3", "rope_theta": s.rope_theta,
                                           "factor": s.rope_factor, "low_freq_factor": 1.0,
                                           "high_freq_factor": 4.0,
                                           "original_max_position_embeddings": 8192})
        f, _ = ROPE_INIT_FUNCTIONS["llama3"](cfg, "cpu")
        inv[name] = f.float().numpy()
    np.savez_compressed(os.path.join(out_dir, "hf_rope_inv_freq.npz"), **inv)
    print("rope inv_freq saved")


if __name__ == "__main__":
    main()
