"""Python face of the native engine (ft_engine_* in include/fasttalk_b200.h).

The engine replaces what the reference reaches over HTTP: the Ollama / vLLM process
behind `app/core/ollama_handler.py:142-177` and `app/core/vllm_handler.py:177,271`.
PyTorch allocates the weights; everything else (KV pages, scheduler, kernels) is native.
"""
from __future__ import annotations

import ctypes as C
import hashlib
from dataclasses import dataclass
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np
import torch

from . import native as N
from .model_spec import ModelSpec
from .weights import interleave_gate_up, rope_tables

FINISH_REASONS = {0: None, 1: "stop", 2: "length", 3: "cancelled", 4: "error"}
LLAMA3_EOS = (128009, 128001)


def session_key(request_id) -> int:
    """Stable 64-bit key of a conversation id (the reference's request_id == session id,
    websocket_server_vllm.py:528-534)."""
    if isinstance(request_id, int):
        return request_id & 0xFFFFFFFFFFFFFFFF
    return int.from_bytes(hashlib.blake2b(str(request_id).encode(), digest_size=8).digest(), "little")


@dataclass
class EngineLimits:
    max_batch: int = 64                 # vLLM --max-num-seqs (docker-compose.vllm.yml:47; 64 for cfg 2)
    max_batched_tokens: int = 8192      # vLLM --max-num-batched-tokens (docker-compose.vllm.yml:48)
    max_context: int = 8192             # vLLM --max-model-len (docker-compose.vllm.yml:43)
    num_pages: int = 0                  # 0 = enough for max_batch full contexts (+25 %)
    max_sessions: int = 0
    use_pdl: bool = True
    decode_prefill_cap: int = 0         # prompt tokens per step while sequences decode: 0 = 896, < 0 = uncapped
    prefill_planes: int = 0             # large prefill GEMMs: 0 = default, 1 = bf16 hi only, 2 = bf16 hi + lo, 3 = one fp16 plane x fp16 weight copies
    debug_logit_rows: int = 0


class Engine:
    def __init__(self, spec: ModelSpec, weights: Dict[str, torch.Tensor], limits: Optional[EngineLimits] = None,
                 eos_ids: Sequence[int] = LLAMA3_EOS, device: Optional[torch.device] = None,
                 tp_rank: int = 0, tp_size: int = 1):
        """`weights` is always the FULL model (weights.py layout); with tp_size == 2 this rank keeps
        only its slices (tp.shard_weights) and `tp.connect(engine)` must follow on both ranks."""
        if not torch.cuda.is_available():
            raise N.NativeError("fasttalk_b200 Engine needs a CUDA device (sm_100a); there is no CPU fallback")
        self.lib = N.load()
        self.full_spec, self.tp_rank, self.tp_size = spec, tp_rank, tp_size
        self.limits = lim = limits or EngineLimits()
        self.device = device or torch.device("cuda", torch.cuda.current_device())
        torch.cuda.set_device(self.device)
        if tp_size > 1:
            from .tp import shard_spec, shard_weights
            weights = shard_weights(spec, weights, tp_rank, tp_size)
            spec = shard_spec(spec, tp_size)
        self.spec = spec
        w = {k: v.to(self.device).contiguous() for k, v in weights.items() if k != "lm_head"}
        w["lm_head"] = (w["embed"] if weights["lm_head"] is weights["embed"]
                        else weights["lm_head"].to(self.device).contiguous())
        for k, v in w.items():
            assert v.dtype == torch.bfloat16, f"weight {k} must be bf16"
        # engine-side layout of the gate/up projection (fused SwiGLU epilogue); one layer at a time
        # so the temporary stays small
        wgu = torch.empty_like(w["wgu"])
        for l in range(spec.layers):
            wgu[l] = interleave_gate_up(w["wgu"][l])
        w["wgu"] = wgu
        cos, sin = rope_tables(spec, lim.max_context, device=self.device)
        self._keep = (w, cos, sin)      # the engine only borrows these pointers
        pages_per_ctx = -(-lim.max_context // 64)
        num_pages = lim.num_pages or int(lim.max_batch * pages_per_ctx * 1.25) + 8
        cfg = N.EngineConfig()
        cfg.hidden, cfg.ffn, cfg.layers = spec.hidden, spec.ffn, spec.layers
        cfg.n_q_heads, cfg.n_kv_heads, cfg.head_dim, cfg.vocab = spec.n_q_heads, spec.n_kv_heads, spec.head_dim, spec.vocab
        cfg.rms_eps = spec.rms_eps
        cfg.max_batch, cfg.max_batched_tokens = lim.max_batch, max(lim.max_batched_tokens, lim.max_batch)
        cfg.num_pages, cfg.max_context = num_pages, lim.max_context
        cfg.max_sessions = lim.max_sessions or 4 * lim.max_batch
        eos = [e for e in eos_ids if 0 <= e < spec.vocab][:4]
        cfg.n_eos = len(eos)
        for i, e in enumerate(eos):
            cfg.eos_ids[i] = e
        cfg.use_pdl = int(lim.use_pdl)
        cfg.debug_logit_rows = lim.debug_logit_rows
        cfg.tp_size, cfg.tp_rank = tp_size, tp_rank
        cfg.decode_prefill_cap = lim.decode_prefill_cap
        cfg.prefill_planes = lim.prefill_planes
        wt = N.Weights()
        for name in ("embed", "lm_head", "final_norm", "attn_norm", "wqkv", "wo", "mlp_norm", "wgu", "wd"):
            setattr(wt, name, w[name].data_ptr())
        wt.rope_cos, wt.rope_sin = cos.data_ptr(), sin.data_ptr()
        self.num_pages = num_pages
        self._h = self.lib.ft_engine_create(C.byref(cfg), C.byref(wt))
        if not self._h:
            raise N.NativeError(f"ft_engine_create failed: {N.last_error()}")
        cap = 4 * lim.max_batch + 64
        self._cap = cap
        self._d_req = np.empty(cap, dtype=np.int64)
        self._d_tok = np.empty(cap, dtype=np.int32)
        self._d_fin = np.empty(cap, dtype=np.int32)

    # ------------------------------------------------------------------ lifecycle
    def close(self):
        if getattr(self, "_h", None):
            self.lib.ft_engine_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ------------------------------------------------------------------ requests
    def submit(self, session, tokens: Sequence[int], temperature: float = 0.0, top_k: int = 0, top_p: float = 1.0,
               max_new_tokens: int = 256, seed: int = 0, ignore_eos: bool = False) -> int:
        arr = np.ascontiguousarray(tokens, dtype=np.int32)
        rid = self.lib.ft_engine_submit(self._h, session_key(session), arr.ctypes.data, int(arr.size),
                                        float(temperature), int(top_k or 0), float(1.0 if top_p is None else top_p),
                                        int(max_new_tokens), int(seed) & 0xFFFFFFFFFFFFFFFF, int(ignore_eos))
        if rid < 0:
            raise N.NativeError(f"ft_engine_submit failed ({rid}): {N.last_error()}")
        return rid

    def step(self) -> int:
        """One continuous-batching iteration; returns tokens produced (0 when idle)."""
        return N.check(self.lib.ft_engine_step(self._h), "ft_engine_step")

    # the engine's own stepping thread (ft_engine_start): the owner submits, waits and drains; nothing in Python sits
    # between two steps
    def start_loop(self):
        N.check(self.lib.ft_engine_start(self._h), "ft_engine_start")

    def stop_loop(self):
        N.check(self.lib.ft_engine_stop(self._h), "ft_engine_stop")

    def wait_events(self, timeout_ms: int = 50) -> int:
        """Block (GIL released) until the stepping thread has queued events; returns how many (0 = timeout)."""
        return N.check(self.lib.ft_engine_wait(self._h, int(timeout_ms)), "ft_engine_wait")

    def drain(self) -> List[Tuple[int, int, Optional[str]]]:
        """(request id, token, finish_reason|None) for everything produced since the last drain."""
        out = []
        while True:
            n = N.check(self.lib.ft_engine_drain(self._h, self._d_req.ctypes.data, self._d_tok.ctypes.data,
                                                 self._d_fin.ctypes.data, self._cap), "ft_engine_drain")
            for i in range(n):
                out.append((int(self._d_req[i]), int(self._d_tok[i]), FINISH_REASONS[int(self._d_fin[i])]))
            if n < self._cap:
                return out

    def drain_arrays(self):
        """Same as drain() without building Python tuples (hot loops)."""
        n = N.check(self.lib.ft_engine_drain(self._h, self._d_req.ctypes.data, self._d_tok.ctypes.data,
                                             self._d_fin.ctypes.data, self._cap), "ft_engine_drain")
        return n, self._d_req, self._d_tok, self._d_fin

    def cancel(self, rid: int) -> bool:
        return N.check(self.lib.ft_engine_cancel(self._h, rid), "ft_engine_cancel") == 0

    def close_session(self, session) -> bool:
        return N.check(self.lib.ft_engine_session_close(self._h, session_key(session)), "ft_engine_session_close") == 0

    def stats(self) -> Dict:
        st = N.EngineStats()
        N.check(self.lib.ft_engine_stats_get(self._h, C.byref(st)), "ft_engine_stats_get")
        return {n: getattr(st, n) for n, _ in st._fields_}

    def timer_start(self):
        N.check(self.lib.ft_engine_timer(self._h, 0, None), "ft_engine_timer")

    def timer_stop_ms(self) -> float:
        ms = C.c_float(0)
        N.check(self.lib.ft_engine_timer(self._h, 1, C.byref(ms)), "ft_engine_timer")
        return float(ms.value)

    def step_log(self, cap: int = 4096) -> np.ndarray:
        """float32 [steps, 4] since the last call, oldest first: device ms, decode rows, prompt tokens, host ms
        (the engine keeps 4096)."""
        buf = np.empty((cap, 4), dtype=np.float32)
        n = N.check(self.lib.ft_engine_step_log(self._h, buf.ctypes.data, cap), "ft_engine_step_log")
        return buf[:n].copy()

    def step_times_ms(self, cap: int = 4096) -> List[float]:
        """Device time of each step since the last call."""
        return self.step_log(cap)[:, 0].tolist()

    def set_decode_prefill_cap(self, cap: int):
        """Prompt tokens admitted per step while sequences decode (0 = default 896, < 0 = uncapped; rounded up so that decode rows + prompt tokens fill 128-token tiles)."""
        N.check(self.lib.ft_engine_set_decode_prefill_cap(self._h, int(cap)), "ft_engine_set_decode_prefill_cap")

    def set_profile(self, on: bool):
        N.check(self.lib.ft_engine_set_profile(self._h, int(on)), "ft_engine_set_profile")

    def logits(self, rid: int) -> torch.Tensor:
        out = np.empty(self.spec.vocab, dtype=np.float32)
        N.check(self.lib.ft_engine_get_logits(self._h, rid, out.ctypes.data), "ft_engine_get_logits")
        return torch.from_numpy(out)

    def debug_prefill_logits(self, tokens: Sequence[int], chunk: int = 64) -> torch.Tensor:
        arr = np.ascontiguousarray(tokens, dtype=np.int32)
        out = np.empty((arr.size, self.spec.vocab), dtype=np.float32)
        N.check(self.lib.ft_debug_prefill_logits(self._h, arr.ctypes.data, int(arr.size), int(chunk),
                                                 out.ctypes.data), "ft_debug_prefill_logits")
        return torch.from_numpy(out)

    # ------------------------------------------------------------------ convenience
    def generate(self, prompts: Sequence[Sequence[int]], max_new_tokens: int, sessions: Optional[Sequence] = None,
                 own_thread: bool = False, **sampling) -> List[List[int]]:
        """Run a set of prompts to completion (tests / smoke / offline use): returns generated token lists.
        own_thread: let the engine step on its native thread (ft_engine_start) as the serving path does, instead of
        calling ft_engine_step from this loop — no Python between two steps."""
        sessions = sessions or [f"gen-{i}" for i in range(len(prompts))]
        rids = [self.submit(s, p, max_new_tokens=max_new_tokens, **sampling) for s, p in zip(sessions, prompts)]
        out = {r: [] for r in rids}
        live = set(rids)
        if own_thread:
            self.start_loop()
            try:
                quiet = 0
                while live:
                    quiet = 0 if self.wait_events(100) > 0 else quiet + 1
                    if quiet > 600:
                        raise N.NativeError("engine produced nothing for 60 s")
                    for rid, tok, fin in self.drain():
                        if tok >= 0 and rid in out:
                            out[rid].append(tok)
                        if fin is not None:
                            live.discard(rid)
            finally:
                self.stop_loop()
            return [out[r] for r in rids]
        idle = 0
        while live:
            idle = 0 if self.step() > 0 else idle + 1
            if idle > 1000:
                raise N.NativeError("engine made no progress for 1000 steps")
            for rid, tok, fin in self.drain():
                if tok >= 0:
                    out[rid].append(tok)
                if fin is not None:
                    live.discard(rid)
        return [out[r] for r in rids]
