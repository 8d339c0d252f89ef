"""Tensor parallelism, host side (CPU): the weight slices recombine to the full model's logits, and
the leader/follower protocol keeps two ranks' engines in lock step (gloo, world_size 2).  The
two-GPU run of the fused GEMM + NVLink exchange is tests/test_tp_gpu.py."""
import os
import socket
import sys

import pytest
import torch
import torch.multiprocessing as mp

from fasttalk_llm_microservice_b200.model_spec import SPECS, LLAMA_3_1_8B
from fasttalk_llm_microservice_b200.tp import shard_spec, shard_weights
from fasttalk_llm_microservice_b200.weights import init_weights


def tp_forward(w, spec, tokens, tp=2):
    """The math the two ranks perform, in fp32 torch: column-parallel q/k/v and gate/up, attention
    on the local heads, row-parallel o / down whose partial outputs are SUMMED (the all-reduce),
    vocabulary-parallel logits that are CONCATENATED."""
    from oracle import llama_oracle as O
    loc = shard_spec(spec, tp)
    ws = [shard_weights(spec, w, r, tp) for r in range(tp)]
    f32 = lambda t: t.float()
    T = tokens.numel()
    cos, sin = O.rope_cos_sin(spec, torch.arange(T))
    x = f32(w["embed"][tokens])
    for l in range(spec.layers):
        xn = O.rmsnorm(x, w["attn_norm"][l], spec.rms_eps)
        part = []
        for r in range(tp):
            qkv = xn @ f32(ws[r]["wqkv"][l]).T
            q = qkv[:, :loc.q_dim].reshape(T, loc.n_q_heads, loc.head_dim)
            k = qkv[:, loc.q_dim:loc.q_dim + loc.kv_dim].reshape(T, loc.n_kv_heads, loc.head_dim)
            v = qkv[:, loc.q_dim + loc.kv_dim:].reshape(T, loc.n_kv_heads, loc.head_dim)
            a = O.attention(O.apply_rope(q, cos, sin), O.apply_rope(k, cos, sin), v, 0)
            part.append(a @ f32(ws[r]["wo"][l]).T)
        x = x + sum(part)
        xn = O.rmsnorm(x, w["mlp_norm"][l], spec.rms_eps)
        x = x + sum(O.swiglu(xn @ f32(ws[r]["wgu"][l]).T) @ f32(ws[r]["wd"][l]).T for r in range(tp))
    xn = O.rmsnorm(x, w["final_norm"], spec.rms_eps)
    return torch.cat([xn @ f32(ws[r]["lm_head"]).T for r in range(tp)], dim=-1)


def test_weight_slices_recombine_to_the_full_model():
    from oracle import llama_oracle as O
    spec = SPECS["tiny-tp"]
    w = init_weights(spec, seed=3, std=0.08, norm_jitter=0.1)
    toks = torch.randint(0, spec.vocab, (21,), generator=torch.Generator().manual_seed(2))
    with torch.no_grad():
        want = O.forward(w, spec, toks)
        got = tp_forward(w, spec, toks)
    assert (got - want).abs().max().item() < 2e-5 * want.abs().max().item()
    loc = shard_spec(spec, 2)
    assert (loc.n_q_heads, loc.n_kv_heads, loc.ffn, loc.vocab, loc.hidden) == (4, 2, 256, 1024, 256)
    s0 = shard_weights(spec, w, 0, 2)
    assert s0["wqkv"].shape == (2, loc.qkv_dim, 256) and s0["wo"].shape == (2, 256, loc.q_dim)
    assert s0["wgu"].shape == (2, 2 * loc.ffn, 256) and s0["wd"].shape == (2, 256, loc.ffn)
    assert s0["lm_head"].shape == (512, 256) and s0["embed"] is w["embed"]
    assert shard_weights(spec, w, 0, 1) is w and shard_spec(spec, 1) is spec
    # the 8B configuration of BASELINE.json splits; shapes that cannot are refused
    l8 = shard_spec(LLAMA_3_1_8B, 2)
    assert (l8.n_q_heads, l8.n_kv_heads, l8.ffn) == (16, 4, 7168)
    for bad in ("tiny-d64", "tiny-d128"):           # one kv head per rank is not a whole 128-row tile / odd heads
        with pytest.raises(ValueError):
            shard_spec(SPECS[bad], 2)
    with pytest.raises(ValueError):
        shard_spec(spec, 4)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _lockstep_worker(rank, port, q):
    import torch.distributed as dist
    sys.path.insert(0, os.path.dirname(__file__))
    from fake_engine import FakeEngine
    from fasttalk_llm_microservice_b200.tokenizer import ByteTokenizer
    from fasttalk_llm_microservice_b200.tp import TPLeader, follower_loop
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=2)
    tok = ByteTokenizer()
    eng = FakeEngine(tok, "abcdefgh")
    if rank == 1:
        steps = follower_loop(eng)
        q.put(("follower", steps, eng.submitted, eng.closed, eng.stats()["active_requests"]))
    else:
        lead = TPLeader(eng)
        r0 = lead.submit("s0", [1, 2, 3], temperature=0.0, max_new_tokens=5)
        out = []
        for i in range(3):
            lead.step()
            out += lead.drain()
        r1 = lead.submit("s1", [4, 5], temperature=0.7, top_p=0.9, max_new_tokens=50)
        lead.step()
        lead.cancel(r1)
        lead.close_session("s1")
        while eng.stats()["active_requests"]:
            lead.step()
            out += lead.drain()
        lead.shutdown()
        q.put(("leader", eng.steps, eng.submitted, eng.closed, [t for r, t, f in out if r == r0]))
    dist.destroy_process_group()


def test_leader_follower_stay_in_lock_step():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    ps = [ctx.Process(target=_lockstep_worker, args=(r, port, q)) for r in range(2)]
    for p in ps:
        p.start()
    got = dict((x[0], x[1:]) for x in (q.get(timeout=120), q.get(timeout=120)))
    for p in ps:
        p.join(timeout=60)
        assert p.exitcode == 0
    l_steps, l_sub, l_closed, l_tokens = got["leader"]
    f_steps, f_sub, f_closed, f_active = got["follower"]
    assert f_steps == l_steps and f_sub == l_sub and f_closed == l_closed == ["s1"] and f_active == 0
    assert l_tokens == list(b"abcde")
