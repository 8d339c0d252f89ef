"""BASELINE configs[3]: Llama-3.1-8B, 8192-token window, TP=2 over NVLink vs TP=1, decode latency.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 \
        --master-port 29552 scripts/bench_tp.py [--model llama-3.1-8b] [--ctx 7936] [--steps 64]

Rank 0 prints one JSON line: ms/step and tok/s for 1 and 8 sessions at ctx `--ctx`, TP=2 and TP=1
side by side, with the HBM roofline of each (weights + KV bytes per GPU per step)."""
import argparse, json, os, sys, time
import torch
import torch.distributed as dist
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def run(eng, sessions, ctx, steps, spec, tag, share):
    g = torch.Generator().manual_seed(7)
    rids = [eng.submit(f"{tag}{i}", torch.randint(0, 128000, (ctx,), generator=g).tolist(),
                       max_new_tokens=steps + 40, ignore_eos=True) for i in range(sessions)]
    t0 = time.perf_counter()
    while eng.stats()["last_prefill_tokens"] > 0 or eng.stats()["steps"] == 0 or eng.stats()["last_batch"] < sessions:
        eng.step(); eng.drain()
    torch.cuda.synchronize()
    prefill_s = time.perf_counter() - t0
    for _ in range(8):
        eng.step(); eng.drain()
    eng.timer_start()
    for _ in range(steps):
        eng.step()
    ms = eng.timer_stop_ms() / steps
    eng.drain()
    for r in rids:
        eng.cancel(r)
    eng.step(); eng.drain()
    for i in range(sessions):
        eng.close_session(f"{tag}{i}")
    ctx_now = ctx + 8 + steps // 2
    bytes_gpu = (spec.weight_bytes() - 2 * spec.vocab * spec.hidden * (0 if spec.tied_embeddings else 1)) * share \
        + sessions * ctx_now * spec.kv_bytes_per_token() * share
    return {"sessions": sessions, "ms_per_step": ms, "tok_s": sessions / ms * 1e3, "prefill_s": prefill_s,
            "bytes_per_gpu_per_step": bytes_gpu, "hbm_frac": bytes_gpu / (ms * 1e-3) / 6545e9}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="llama-3.1-8b")
    ap.add_argument("--ctx", type=int, default=7936)
    ap.add_argument("--steps", type=int, default=64)
    a = ap.parse_args()
    rank, local = int(os.environ["RANK"]), int(os.environ.get("LOCAL_RANK", os.environ["RANK"]))
    torch.cuda.set_device(local)
    dist.init_process_group("gloo")
    from fasttalk_llm_microservice_b200 import build, tp
    from fasttalk_llm_microservice_b200.engine import Engine, EngineLimits
    from fasttalk_llm_microservice_b200.model_spec import get_spec
    from fasttalk_llm_microservice_b200.weights import init_weights
    build.build()
    spec = get_spec(a.model)
    w = init_weights(spec, seed=0, device="cuda")           # Philox on identical GPUs: the same weights on both ranks
    chk = [None, None]
    dist.all_gather_object(chk, float(w["wd"][-1].float().sum()))
    assert chk[0] == chk[1], "ranks drew different weights"
    lim = EngineLimits(max_batch=8, max_batched_tokens=8192, max_context=8192, decode_prefill_cap=-1)
    res = {"model": spec.name, "ctx": a.ctx, "steps": a.steps, "tp2": [], "tp1": []}
    eng = Engine(spec, w, lim, eos_ids=(), tp_rank=rank, tp_size=2)
    tp.connect(eng)
    for s in (1, 8):
        dist.barrier()
        res["tp2"].append(run(eng, s, a.ctx, a.steps, spec, f"a{s}_", 0.5))
    dist.barrier()
    eng.close()
    if rank == 0:
        one = Engine(spec, w, lim, eos_ids=())
        for s in (1, 8):
            res["tp1"].append(run(one, s, a.ctx, a.steps, spec, f"b{s}_", 1.0))
        one.close()
        print("TP_BENCH " + json.dumps(res), flush=True)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
