"""One rank of the two-GPU tensor-parallel check; launched by tests/test_tp_gpu.py (or by hand):

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 \
        --master-port 29551 tests/tp_worker.py [--model tiny-tp]

Both ranks build the same full weights, keep their slices, swap buffer handles, then run the same
prompts.  Checked on every rank: prefill logits vs the fp32 oracle (2e-2 relative, north_star),
greedy tokens vs the oracle's, bit-identical logits on the two ranks, and — rank 0 — that a
TP=1 engine on the same GPU gives the same tokens."""
import argparse
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="tiny-tp")
    ap.add_argument("--new", type=int, default=24)
    a = ap.parse_args()
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    assert world == 2
    torch.cuda.set_device(local)
    dist.init_process_group("gloo")                       # control plane only; the data path is the fused kernel
    from fasttalk_llm_microservice_b200 import build, tp
    from fasttalk_llm_microservice_b200.engine import Engine, EngineLimits
    from fasttalk_llm_microservice_b200.model_spec import get_spec
    from fasttalk_llm_microservice_b200.weights import init_weights
    from oracle import llama_oracle as O
    build.build()
    spec = get_spec(a.model)
    tiny = spec.hidden <= 512
    w = init_weights(spec, seed=21, std=0.08 if tiny else 0.02, norm_jitter=0.1 if tiny else 0.0)   # CPU generator: same on both ranks
    # the named models also get one long prompt in a single step: > 148 tiles per projection, i.e. the CTA-pair
    # kernel with the exchange in its epilogue
    lim = EngineLimits(max_batch=8, max_batched_tokens=256 if tiny else 2048, max_context=1024 if tiny else 4096,
                       debug_logit_rows=64)
    eng = Engine(spec, w, lim, eos_ids=(), tp_rank=rank, tp_size=2)
    tp.connect(eng)
    g = torch.Generator().manual_seed(5)
    prompts = [torch.randint(0, spec.vocab, (n,), generator=g).tolist() for n in ((40, 7, 130) if tiny else (40, 7, 130, 1500))]

    # ---- prefill logits vs the oracle, chunked (several exchanging launches per layer) and in one go
    wd = {k: v.cuda() for k, v in w.items()}
    with torch.no_grad():
        ref = O.forward(wd, spec, torch.tensor(prompts[0])).cpu()
    worst = 0.0
    for chunk in (64, 16):
        got = eng.debug_prefill_logits(prompts[0], chunk=chunk)
        err = ((got - ref).abs().amax(-1) / ref.abs().amax(-1)).max().item()
        worst = max(worst, err)
        assert err < 2e-2, f"rank {rank}: logits rel err {err:.3e} (chunk {chunk})"
    both = [None, None]
    dist.all_gather_object(both, got.numpy().tobytes())
    assert both[0] == both[1], "the two ranks' logits differ bitwise"

    # ---- batched greedy generation: continuous batching, decode through the split-K rendezvous + exchange
    # every position of every sequence against the oracle, teacher-forced, near-ties re-aligned (oracle/parity.py);
    # both ranks take the same decisions because their logits are bit-identical
    from oracle import parity as P
    jobs = []
    for i, p in enumerate(prompts):
        with torch.no_grad():
            ref_t, ref_l = O.greedy_generate(wd, spec, torch.tensor(p), a.new, return_logits=True)
        jobs.append(P.Job(f"TP=2 {spec.name} rank {rank} prompt {i} ({len(p)} tokens)", f"par{i}", p, ref_t, ref_l))
    P.run(eng, jobs)
    lines = [P.finish(j) for j in jobs]
    for i in range(len(prompts)):
        eng.close_session(f"par{i}")
    out = eng.generate(prompts, a.new)
    toks = [None, None]
    dist.all_gather_object(toks, out)
    assert toks[0] == toks[1], "the two ranks sampled different tokens"
    st = eng.stats()
    dist.barrier()
    # ---- the serving arrangement: rank 0 drives through TPLeader, rank 1 only replays (tp.follower_loop)
    if rank == 0:
        lead = tp.TPLeader(eng)
        rids = [lead.submit(f"lead{i}", p, max_new_tokens=a.new, ignore_eos=True) for i, p in enumerate(prompts)]
        got = {r: [] for r in rids}
        live = set(rids)
        lead.cancel(rids[1])                       # a cancel travels with the next step like everything else
        while live:
            lead.step()
            for r, t, fin in lead.drain():
                if t >= 0:
                    got[r].append(t)
                if fin:
                    live.discard(r)
        lead.shutdown()
        assert got[rids[0]] == out[0] and got[rids[2]] == out[2], "leader/follower run differs from the lock-step run"
        assert got[rids[1]] == []
    else:
        n_steps = tp.follower_loop(eng)
        assert n_steps > 0
    dist.barrier()
    if rank == 0:
        one = Engine(spec, w, lim, eos_ids=())
        same = one.generate(prompts, a.new)
        agree = sum(int(x == y) for o1, o2 in zip(out, same) for x, y in zip(o1, o2))
        total = sum(len(o) for o in out)
        for ln in lines:
            print("TP_PARITY " + ln, flush=True)
        print(f"TP_OK model={spec.name} logits_rel_err={worst:.3e} tokens={total} tp1_agreement={agree}/{total} "
              f"launches={st['kernel_launches']}", flush=True)
        one.close()
    dist.barrier()
    eng.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
