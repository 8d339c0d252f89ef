"""Real-model plumbing on CPU: a checkpoint written by transformers'
`save_pretrained` loads into the stacked layout and the oracle reproduces transformers' logits on
it; a real `tokenizer.json` (byte-level BPE built here) streams UTF-8 safely."""
import json
import os
import sys

import pytest
import torch

from fasttalk_llm_microservice_b200.checkpoint import (load_hf_checkpoint, save_hf_checkpoint, spec_from_hf_config,
                                                        hf_config_of)
from fasttalk_llm_microservice_b200.model_spec import SPECS, TINY_D64, TINY_D128
from fasttalk_llm_microservice_b200.tokenizer import HFTokenizer, load_tokenizer
from fasttalk_llm_microservice_b200.weights import init_weights

sys.path.insert(0, os.path.join(os.path.dirname(__file__), "golden"))


@pytest.mark.parametrize("spec,shard", [(TINY_D64, None), (TINY_D128, "300KB")])
def test_transformers_checkpoint_roundtrip(tmp_path, spec, shard):
    from make_golden import hf_model
    from oracle import llama_oracle as O
    w = init_weights(spec, seed=5, std=0.08, norm_jitter=0.1)
    m = hf_model(spec, w).to(torch.bfloat16)
    m.save_pretrained(tmp_path, safe_serialization=True, **({"max_shard_size": shard} if shard else {}))
    if shard:
        assert os.path.exists(tmp_path / "model.safetensors.index.json")
    spec2, w2 = load_hf_checkpoint(str(tmp_path), name=spec.name)
    assert spec2 == spec
    for k in ("embed", "final_norm", "attn_norm", "mlp_norm", "wqkv", "wo", "wgu", "wd", "lm_head"):
        assert torch.equal(w2[k], w[k]), k
    assert (w2["lm_head"] is w2["embed"]) == spec.tied_embeddings
    prompt = torch.randint(0, spec.vocab, (19,), generator=torch.Generator().manual_seed(1))
    with torch.no_grad():
        want = hf_model(spec, w)(prompt[None]).logits[0]      # fresh fp32 model (the bf16 cast above rounded inv_freq)
        got = O.forward(w2, spec2, prompt)
    assert (got - want).abs().max().item() < 1e-4 * want.abs().max().item()


def test_save_load_inverse_and_config_mapping(tmp_path):
    spec = TINY_D128
    w = init_weights(spec, seed=6)
    save_hf_checkpoint(spec, w, str(tmp_path), shards=3)
    spec2, w2 = load_hf_checkpoint(str(tmp_path), name=spec.name)
    assert spec2 == spec and all(torch.equal(w2[k], w[k]) for k in w)
    for name in ("llama-3.2-1b", "llama-3.2-3b", "llama-3.1-8b"):
        assert spec_from_hf_config(hf_config_of(SPECS[name]), name) == SPECS[name]
    # the public Llama-3.2-1B config.json, field for field
    cfg = {"architectures": ["LlamaForCausalLM"], "attention_bias": False, "head_dim": 64, "hidden_size": 2048,
           "intermediate_size": 8192, "max_position_embeddings": 131072, "mlp_bias": False, "model_type": "llama",
           "num_attention_heads": 32, "num_hidden_layers": 16, "num_key_value_heads": 8, "rms_norm_eps": 1e-05,
           "rope_scaling": {"factor": 32.0, "high_freq_factor": 4.0, "low_freq_factor": 1.0,
                            "original_max_position_embeddings": 8192, "rope_type": "llama3"},
           "rope_theta": 500000.0, "tie_word_embeddings": True, "vocab_size": 128256}
    assert spec_from_hf_config(cfg, "llama-3.2-1b") == SPECS["llama-3.2-1b"]
    for bad in ({"model_type": "mistral", "architectures": ["MistralForCausalLM"]}, {"attention_bias": True},
                {"quantization_config": {"quant_method": "awq"}}, {"rope_scaling": {"rope_type": "yarn", "factor": 2}}):
        with pytest.raises(ValueError):
            spec_from_hf_config({**cfg, **bad})
    os.remove(tmp_path / "model-00001-of-00003.safetensors")
    with pytest.raises(Exception):
        load_hf_checkpoint(str(tmp_path))


def test_real_tokenizer_json_streams_utf8(tmp_path):
    from tokenizers import Tokenizer, decoders, models, pre_tokenizers, trainers
    tk = Tokenizer(models.BPE())
    tk.pre_tokenizer = pre_tokenizers.ByteLevel(add_prefix_space=False)
    tk.decoder = decoders.ByteLevel()
    specials = ["<|begin_of_text|>", "<|end_of_text|>", "<|start_header_id|>", "<|end_header_id|>", "<|eot_id|>"]
    tk.train_from_iterator(["hello world, héllo wörld ✓ 你好 " * 4, "system user assistant be brief"],
                           trainers.BpeTrainer(vocab_size=400, special_tokens=specials,
                                               initial_alphabet=pre_tokenizers.ByteLevel.alphabet()))
    path = str(tmp_path / "tokenizer.json")
    tk.save(path)
    tok = load_tokenizer(0, path)
    assert isinstance(tok, HFTokenizer) and tok.eos_ids == (tok.eot, tok.end_of_text)
    text = "héllo wörld ✓ 你好 🙂 done"
    ids = tok.encode(text)
    assert tok.decode(ids) == text
    dec = tok.stream_decoder()
    pieces = [dec.push(t) for t in ids] + [dec.flush()]
    assert "".join(pieces) == text and all("�" not in p for p in pieces)
    t = tok.apply_chat_template([{"role": "system", "content": "be brief"}, {"role": "user", "content": "hello"}])
    assert t[0] == tok.bos and t.count(tok.eot) == 2 and t.count(tok.start_header) == 3
    assert tok.decode(t) == "system\n\nbe briefuser\n\nhelloassistant\n\n"
    assert dec.push(tok.eot) == ""
    # incremental: a long reply streams to exactly its full decoding, the decoder keeps only a short window of ids
    # (re-decoding the whole reply per token was O(n^2) on the dispatch thread)
    long_text = ("héllo wörld ✓ 你好 🙂 hello world, " * 60).strip()
    ids = tok.encode(long_text)
    dec = tok.stream_decoder()
    out = "".join(dec.push(t) for t in ids) + dec.flush()
    assert out == long_text and len(dec.ids) <= 80 and len(ids) > 600
    # a checkpoint without a tokenizer.json is refused instead of silently served with the byte tokenizer
    with pytest.raises(FileNotFoundError):
        load_tokenizer(128256, None, require=True)


def test_per_gpu_replica_commands_and_sticky_assignment():
    from fasttalk_llm_microservice_b200.launch import replica_commands, replica_for
    cmds = replica_commands(4, "llama3.2:1b", "127.0.0.1", 9000)
    assert [e["CUDA_VISIBLE_DEVICES"] for _, e in cmds] == ["0", "1", "2", "3"]
    assert [c[c.index("--port") + 1] for c, _ in cmds] == ["9000", "9001", "9002", "9003"]
    assert all(e["FASTTALK_GPU"] == "0" and "fasttalk_llm_microservice_b200.openai_server" in c for c, e in cmds)
    ids = [f"session-{i}" for i in range(4000)]
    assign = [replica_for(s, 8) for s in ids]
    assert assign == [replica_for(s, 8) for s in ids]                       # sticky
    counts = [assign.count(r) for r in range(8)]
    assert min(counts) > 400 and max(counts) < 600                          # and balanced
