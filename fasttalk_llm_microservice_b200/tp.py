"""Tensor parallelism over two GPUs for the 8B-class configuration (SURVEY.md §8e, §8a′12).

The reference gets this from vLLM's `--tensor-parallel-size` when one GPU is not enough
(`docker-compose.vllm.yml:49`); here it is a latency optimisation: each GPU streams half of the
weights and half of the KV cache per step.  One process per GPU:

* `shard_spec` / `shard_weights` — Megatron-style slices: q/k/v and gate/up by output rows
  (column-parallel), o and down by input columns (row-parallel), lm_head by vocabulary rows.
* The all-reduce after the row-parallel projections is NOT a collective call: it is fused into the
  GEMM epilogue (csrc/gemm.cu, `ft_gemm_epilogue.tp_*`) — every CTA stores its partial tile into
  the peer GPU's exchange buffer over NVLink and adds what the peer's twin CTA stored here.
  `connect()` only swaps the cudaIpc handles of those buffers once, over `torch.distributed`.
* Both ranks must see the same requests in the same order.  `TPLeader` (rank 0, what the handler
  talks to) forwards every call to `follower_loop` (rank 1) bundled with the next step.
"""
from __future__ import annotations

import ctypes as C
import threading
from dataclasses import replace
from typing import Dict, List, Optional

import torch

from .model_spec import ModelSpec


def shard_spec(spec: ModelSpec, tp: int) -> ModelSpec:
    """The per-rank model: half the heads and half the FFN; hidden and vocab stay whole."""
    if tp == 1:
        return spec
    if tp != 2:
        raise ValueError("tensor parallelism is built for 2 ranks")
    if spec.n_q_heads % tp or spec.n_kv_heads % tp or (spec.ffn // tp) % 64 or spec.vocab % (128 * tp) \
            or (spec.n_kv_heads // tp * spec.head_dim) % 128:
        raise ValueError(f"{spec.name}: heads / ffn / vocab do not split over {tp} ranks in whole tiles")
    return replace(spec, name=f"{spec.name}/tp{tp}", n_q_heads=spec.n_q_heads // tp,
                   n_kv_heads=spec.n_kv_heads // tp, ffn=spec.ffn // tp)


def shard_weights(spec: ModelSpec, w: Dict[str, torch.Tensor], rank: int, tp: int) -> Dict[str, torch.Tensor]:
    """Slices of the full weights (weights.py layout, gate|up NOT yet interleaved) for `rank`.
    Kv heads go with the q heads of their GQA groups, so attention stays local to a rank."""
    if tp == 1:
        return w
    shard_spec(spec, tp)                      # validates divisibility
    q, kv, f, V = spec.q_dim, spec.kv_dim, spec.ffn, spec.vocab
    ql, kvl, fl, vl = q // tp, kv // tp, f // tp, V // tp
    r = rank
    wqkv = torch.cat([w["wqkv"][:, r * ql:(r + 1) * ql], w["wqkv"][:, q + r * kvl:q + (r + 1) * kvl],
                      w["wqkv"][:, q + kv + r * kvl:q + kv + (r + 1) * kvl]], dim=1)
    wgu = torch.cat([w["wgu"][:, r * fl:(r + 1) * fl], w["wgu"][:, f + r * fl:f + (r + 1) * fl]], dim=1)
    out = {"embed": w["embed"], "final_norm": w["final_norm"], "attn_norm": w["attn_norm"], "mlp_norm": w["mlp_norm"],
           "wqkv": wqkv.contiguous(), "wo": w["wo"][:, :, r * ql:(r + 1) * ql].contiguous(),
           "wgu": wgu.contiguous(), "wd": w["wd"][:, :, r * fl:(r + 1) * fl].contiguous(),
           "lm_head": w["lm_head"][r * vl:(r + 1) * vl].contiguous()}
    return out


def connect(engine, group=None) -> None:
    """Swap the peer-visible buffers of the two ranks' engines (call on both ranks, once)."""
    import torch.distributed as dist
    from . import native as N
    lib = N.load()
    rank, world = dist.get_rank(group), dist.get_world_size(group)
    assert world == 2, "tensor parallelism is built for 2 ranks"
    ptrs = [C.c_void_p() for _ in range(3)]
    N.check(lib.ft_engine_tp_export(engine._h, *[C.byref(p) for p in ptrs]), "ft_engine_tp_export")
    handles = []
    for p in ptrs:
        buf = (C.c_ubyte * 64)()
        N.check(lib.ft_ipc_export(p, buf), "ft_ipc_export")
        handles.append(bytes(buf))
    gathered: List[Optional[List[bytes]]] = [None] * world
    dist.all_gather_object(gathered, handles, group=group)
    peer = []
    for h in gathered[1 - rank]:
        out = C.c_void_p()
        N.check(lib.ft_ipc_open((C.c_ubyte * 64).from_buffer_copy(h), C.byref(out)), "ft_ipc_open")
        peer.append(out)
    N.check(lib.ft_engine_tp_attach(engine._h, *peer), "ft_engine_tp_attach")
    engine._tp_peer = peer                     # mapped for the engine's lifetime
    dist.barrier(group=group)


# ---------------------------------------------------------------------------------------------
# Keeping the two ranks' schedulers in lock step
# ---------------------------------------------------------------------------------------------
class TPLeader:
    """Rank 0's engine front: same submit/step/drain/cancel/close_session/stats surface as Engine
    (what handler.EngineRunner needs).  Calls that change scheduler state are applied locally at
    once and queued; the queue travels to the follower with the next `step`, which the follower
    replays in the same order before stepping, so both schedulers build identical batches."""

    def __init__(self, engine, group=None):
        self.engine, self.group = engine, group
        self.limits = getattr(engine, "limits", None)
        self.spec = getattr(engine, "spec", None)
        self._lock = threading.Lock()
        self._pending: list = []

    def _send(self, cmds):
        import torch.distributed as dist
        dist.broadcast_object_list([cmds], src=0, group=self.group)

    def submit(self, session, tokens, **kw) -> int:
        with self._lock:
            rid = self.engine.submit(session, tokens, **kw)
            self._pending.append(("submit", session, list(tokens), kw))
            return rid

    def cancel(self, rid: int) -> bool:
        with self._lock:
            self._pending.append(("cancel", rid))
            return self.engine.cancel(rid)

    def close_session(self, session) -> bool:
        with self._lock:
            self._pending.append(("close_session", session))
            return self.engine.close_session(session)

    def step(self) -> int:
        with self._lock:
            cmds, self._pending = self._pending + [("step",)], []
            self._send(cmds)
            return self.engine.step()

    def drain(self):
        return self.engine.drain()

    def stats(self):
        return self.engine.stats()

    def shutdown(self):
        with self._lock:
            self._send(self._pending + [("stop",)])
            self._pending = []

    def close(self):
        self.shutdown()
        if hasattr(self.engine, "close"):
            self.engine.close()


def follower_loop(engine, group=None) -> int:
    """Rank 1: replay the leader's calls until it says stop.  Returns the number of steps run."""
    import torch.distributed as dist
    steps = 0
    while True:
        box = [None]
        dist.broadcast_object_list(box, src=0, group=group)
        for cmd in box[0]:
            op = cmd[0]
            if op == "submit":
                engine.submit(cmd[1], cmd[2], **cmd[3])
            elif op == "cancel":
                engine.cancel(cmd[1])
            elif op == "close_session":
                engine.close_session(cmd[1])
            elif op == "step":
                engine.step()
                engine.drain()              # the leader streams the tokens; identical here by construction
                steps += 1
            elif op == "stop":
                return steps
            else:
                raise ValueError(f"unknown tensor-parallel command {op!r}")
