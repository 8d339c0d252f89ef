"""torch-tensor front ends of the kernel-level C ABI (include/fasttalk_b200.h).  PyTorch only
supplies device memory and the stream; all arithmetic happens in the CUDA library."""
from __future__ import annotations

from typing import Optional

import torch

from . import native as N

KV_TILE = 64


def _chk_bf16(t, name):
    assert t.is_cuda and t.dtype == torch.bfloat16 and t.is_contiguous(), f"{name}: need contiguous CUDA bf16"


def split_hi_lo(x: torch.Tensor):
    """fp32 -> (bf16 hi, bf16 lo) planes with hi + lo ~= x to ~16 mantissa bits."""
    hi = x.to(torch.bfloat16)
    lo = (x.float() - hi.float()).to(torch.bfloat16)
    return hi.contiguous(), lo.contiguous()


def gemm_bf16(w: torch.Tensor, x: torch.Tensor, x_lo: Optional[torch.Tensor] = None, n_splits: int = 0,
              mode: str = "partials"):
    """Y[t, n] = sum_k (X + Xlo)[t, k] W[n, k].
    mode "partials": fp32 [S, T, n_out]; "bf16": (hi, lo) bf16 [T, n_out];
    "swiglu": w rows interleaved gate|up per 64 -> (hi, lo) bf16 [T, n_out/2]."""
    _chk_bf16(w, "w"); _chk_bf16(x, "x")
    if x_lo is not None:
        _chk_bf16(x_lo, "x_lo")
    n_out, K = w.shape
    T = x.shape[0]
    assert x.shape[1] == K
    lib = N.load()
    st = N.cur_stream()
    if mode == "partials":
        if n_splits <= 0:
            n_splits = lib.ft_gemm_pick_splits(n_out, T, K, int(x_lo is not None))
        out = torch.empty(n_splits, T, n_out, device=x.device, dtype=torch.float32)
        N.check(lib.ft_gemm_bf16(N.ptr(w), N.ptr(x), N.ptr(x_lo), N.ptr(out), 0, n_out, T, K, n_splits, 0, st),
                "ft_gemm_bf16")
        return out
    width = n_out if mode == "bf16" else n_out // 2
    out = torch.empty(T, width, device=x.device, dtype=torch.bfloat16)
    out_lo = torch.empty_like(out)
    N.check(lib.ft_gemm_bf16(N.ptr(w), N.ptr(x), N.ptr(x_lo), N.ptr(out), N.ptr(out_lo), n_out, T, K, 1,
                             1 if mode == "bf16" else 2, st), "ft_gemm_bf16")
    return out, out_lo


def gemm_fused(w: torch.Tensor, x: torch.Tensor, x_lo: torch.Tensor, mode: int, n_splits: int = 0, **f) -> None:
    """ft_gemm_fused with an ft_gemm_epilogue built from keyword tensors / scalars (see
    include/fasttalk_b200.h).  Outputs are whatever tensors the caller put in the epilogue."""
    import ctypes as C
    if f.get("x_planes") == 3:       # the fp16 single-plane form of the large prefill projections: fp16 weights, x_lo = None
        assert x_lo is None and all(t.is_cuda and t.dtype == torch.float16 and t.is_contiguous() for t in (w, x)), \
            "x_planes = 3: need contiguous CUDA fp16 w and x, x_lo = None"
    else:
        _chk_bf16(w, "w"); _chk_bf16(x, "x"); _chk_bf16(x_lo, "x_lo")
    n_out, K = w.shape
    T = x.shape[0]
    ep = N.GemmEpilogue()
    ep.mode = mode
    lib = N.load()
    if mode in (N.EPI_RESID, N.EPI_QKV_ROPE) and "scratch" not in f:
        # split-K work buffers: slabs for the partial tiles + self re-arming arrival counters
        f = dict(f)
        f["scratch"] = torch.empty(max(1, lib.ft_gemm_scratch_bytes(n_out, T, K, n_splits) // 4), device=x.device,
                                   dtype=torch.float32)
        f["sync"] = torch.zeros(2 * (n_out // 128) * ((T + 15) // 16), device=x.device, dtype=torch.int32)
    for name, _ in N.GemmEpilogue._fields_:
        if name == "mode" or name not in f:
            continue
        v = f[name]
        setattr(ep, name, v.data_ptr() if isinstance(v, torch.Tensor) else v)
    N.check(N.load().ft_gemm_fused(N.ptr(w), N.ptr(x), N.ptr(x_lo), n_out, T, K, n_splits, C.byref(ep),
                                   N.cur_stream()), "ft_gemm_fused")


def embed_planes(tokens: torch.Tensor, table: torch.Tensor, gamma: torch.Tensor, ssq_parts: int):
    """resid = table[tokens] (fp32); planes = hi/lo(resid * gamma); ssq [parts, T]."""
    T, h = tokens.numel(), table.shape[1]
    resid = torch.empty(T, h, device=table.device, dtype=torch.float32)
    hi = torch.empty(T, h, device=table.device, dtype=torch.bfloat16)
    lo = torch.empty_like(hi)
    ssq = torch.empty(ssq_parts, T, device=table.device, dtype=torch.float32)
    N.check(N.load().ft_embed_planes(N.ptr(tokens), N.ptr(table), N.ptr(gamma), N.ptr(resid), N.ptr(hi), N.ptr(lo),
                                     N.ptr(ssq), ssq_parts, T, T, h, N.cur_stream()), "ft_embed_planes")
    return resid, hi, lo, ssq


def gather_planes(x: torch.Tensor, x_lo: torch.Tensor, ssq: torch.Tensor, row_idx: torch.Tensor):
    n, h = row_idx.numel(), x.shape[1]
    hi = torch.empty(n, h, device=x.device, dtype=torch.bfloat16)
    lo = torch.empty_like(hi)
    so = torch.empty(ssq.shape[0], n, device=x.device, dtype=torch.float32)
    N.check(N.load().ft_gather_planes(N.ptr(x), N.ptr(x_lo), N.ptr(ssq), ssq.shape[0], ssq.stride(0), N.ptr(row_idx),
                                      n, h, N.ptr(hi), N.ptr(lo), N.ptr(so), n, N.cur_stream()), "ft_gather_planes")
    return hi, lo, so


def add_rmsnorm(resid: torch.Tensor, norm_w: torch.Tensor, eps: float, partials: Optional[torch.Tensor] = None,
                embed_tokens: Optional[torch.Tensor] = None, embed_table: Optional[torch.Tensor] = None,
                row_idx: Optional[torch.Tensor] = None):
    """In-place residual update + RMSNorm; returns the bf16 (hi, lo) planes [n_rows, h]."""
    assert resid.dtype == torch.float32 and resid.is_contiguous()
    h = resid.shape[1]
    T = row_idx.numel() if row_idx is not None else resid.shape[0]
    xn = torch.empty(T, h, device=resid.device, dtype=torch.bfloat16)
    xn_lo = torch.empty_like(xn)
    ns, stride = 0, 0
    if partials is not None:
        assert partials.dtype == torch.float32 and partials.is_contiguous()
        ns, stride = partials.shape[0], partials.stride(0)
    N.check(N.load().ft_add_rmsnorm(N.ptr(resid), N.ptr(partials), ns, stride, N.ptr(embed_tokens),
                                    N.ptr(embed_table), N.ptr(norm_w), N.ptr(xn), N.ptr(xn_lo), N.ptr(row_idx),
                                    T, h, eps, N.cur_stream()), "ft_add_rmsnorm")
    return xn, xn_lo


def kv_cache_alloc(num_pages: int, nkv: int, d: int, device) -> torch.Tensor:
    """One layer of paged KV cache: [page][kv_head][K|V][64][d] fp16 (swizzled rows)."""
    return torch.zeros(num_pages, nkv, 2, KV_TILE, d, device=device, dtype=torch.float16)


def rope_kv_append(qkv_partials: torch.Tensor, cos: torch.Tensor, sin: torch.Tensor, positions: torch.Tensor,
                   seq_rows: torch.Tensor, page_table: torch.Tensor, kv_layer: torch.Tensor,
                   nq: int, nkv: int, d: int) -> torch.Tensor:
    S, T, _ = qkv_partials.shape
    q = torch.empty(T, nq, d, device=qkv_partials.device, dtype=torch.float16)
    N.check(N.load().ft_rope_kv_append(N.ptr(qkv_partials), S, qkv_partials.stride(0), N.ptr(cos), N.ptr(sin),
                                       N.ptr(positions), N.ptr(seq_rows), N.ptr(page_table),
                                       page_table.shape[1], N.ptr(q), N.ptr(kv_layer), T, nq, nkv, d,
                                       N.cur_stream()), "ft_rope_kv_append")
    return q


def silu_mul(gu_partials: torch.Tensor):
    S, T, f2 = gu_partials.shape
    act = torch.empty(T, f2 // 2, device=gu_partials.device, dtype=torch.bfloat16)
    act_lo = torch.empty_like(act)
    N.check(N.load().ft_silu_mul(N.ptr(gu_partials), S, gu_partials.stride(0), N.ptr(act), N.ptr(act_lo), T,
                                 f2 // 2, N.cur_stream()), "ft_silu_mul")
    return act, act_lo


def attn_decode(q: torch.Tensor, kv_layer: torch.Tensor, page_table: torch.Tensor, ctx_lens: torch.Tensor,
                n_ctas: int = 0):
    """q fp16 [B, nq, d], fp16 KV cache -> bf16 (hi, lo) planes [B, nq*d]."""
    assert q.dtype == torch.float16 and kv_layer.dtype == torch.float16
    B, nq, d = q.shape
    nkv = kv_layer.shape[1]
    lib = N.load()
    if n_ctas <= 0:
        n_ctas = N.check(lib.ft_attn_decode_default_ctas(d), "ft_attn_decode_default_ctas")
    scratch = torch.empty(lib.ft_attn_decode_scratch_bytes(B, nkv, nq // nkv, d, n_ctas) // 4,
                          device=q.device, dtype=torch.float32)
    tile_prefix = torch.empty(B + 1, device=q.device, dtype=torch.int32)
    out = torch.empty(B, nq * d, device=q.device, dtype=torch.bfloat16)
    out_lo = torch.empty_like(out)
    N.check(lib.ft_attn_decode(N.ptr(q), N.ptr(kv_layer), N.ptr(page_table), page_table.shape[1],
                               N.ptr(ctx_lens), N.ptr(tile_prefix), N.ptr(out), N.ptr(out_lo), N.ptr(scratch), B,
                               nq, nkv, d, n_ctas, N.cur_stream()), "ft_attn_decode")
    return out, out_lo


def attn_decode_noalloc(q: torch.Tensor, kv_layer: torch.Tensor, page_table: torch.Tensor, ctx_lens: torch.Tensor,
                        out: torch.Tensor, out_lo: torch.Tensor, n_ctas: int = 0) -> None:
    """Decode attention for callers that may not allocate device memory while serving (fixed
    start-up memory budget, real-time audio thread): the library's built-in work area replaces
    the scratch buffers attn_decode() allocates, and the caller supplies `out` / `out_lo`
    (bf16 [B, nq*d]) itself, so this call allocates nothing anywhere.  Calls on one GPU share
    that work area: do not overlap them.  Raises NativeError beyond B*nkv = 1024 / n_ctas = 512."""
    B, nq, d = q.shape
    nkv = kv_layer.shape[1]
    N.check(N.load().ft_attn_decode_noalloc(N.ptr(q), N.ptr(kv_layer), N.ptr(page_table), page_table.shape[1],
                                            N.ptr(ctx_lens), N.ptr(out), N.ptr(out_lo), B, nq, nkv, d, n_ctas,
                                            N.cur_stream()), "ft_attn_decode_noalloc")


def attn_prefill(q: torch.Tensor, kv_layer: torch.Tensor, page_table: torch.Tensor, req_q_start, req_slot,
                 req_pos0, tensor_cores: bool = True):
    """q fp16 [T, nq, d]; request r owns rows req_q_start[r]:req_q_start[r+1].  Returns (hi, lo)."""
    assert q.dtype == torch.float16 and kv_layer.dtype == torch.float16
    T, nq, d = q.shape
    nkv = kv_layer.shape[1]
    dev = q.device
    tile_req, tile_q0 = [], []
    for r in range(len(req_slot)):
        for q0 in range(0, req_q_start[r + 1] - req_q_start[r], 64):
            tile_req.append(r); tile_q0.append(q0)
    i32 = lambda v: torch.tensor(v, device=dev, dtype=torch.int32)
    t_qs, t_slot, t_pos0, t_treq, t_tq0 = i32(req_q_start), i32(req_slot), i32(req_pos0), i32(tile_req), i32(tile_q0)
    out = torch.zeros(T, nq * d, device=dev, dtype=torch.bfloat16)
    out_lo = torch.zeros_like(out)
    N.check(N.load().ft_attn_prefill(N.ptr(q), N.ptr(kv_layer), N.ptr(page_table), page_table.shape[1],
                                     N.ptr(t_qs), N.ptr(t_slot), N.ptr(t_pos0), N.ptr(t_treq), N.ptr(t_tq0),
                                     len(tile_req), N.ptr(out), N.ptr(out_lo), nq, nkv, d, T if tensor_cores else 0,
                                     N.cur_stream()),
            "ft_attn_prefill")
    return out, out_lo


def sample(logits: torch.Tensor, temperature=None, top_k=None, top_p=None, seeds=None) -> torch.Tensor:
    B, V = logits.shape
    assert logits.dtype == torch.float32 and logits.is_contiguous()
    out = torch.empty(B, device=logits.device, dtype=torch.int32)
    N.check(N.load().ft_sample(N.ptr(logits), B, V, N.ptr(temperature), N.ptr(top_k), N.ptr(top_p),
                               N.ptr(seeds), N.ptr(out), N.cur_stream()), "ft_sample")
    return out


# ---- KV cache (un)swizzle helpers for tests: logical [tokens, d] <-> stored slab
# A stored slab is d / 64 sub-slabs of [64 tokens][64 dims]; inside a sub-slab the eight 16-byte chunks of a row are
# XOR-swizzled with the row index (csrc/common.cuh: kv_swizzled_chunk).  The XOR is an involution.
def _kv_perm(x: torch.Tensor) -> torch.Tensor:
    h, rows = x.shape[0], x.shape[1]
    r = torch.arange(rows, device=x.device)[None, :, None]
    c = torch.arange(8, device=x.device)[None, None, :]
    src = (c ^ (r & 7)).expand(h, rows, 8)
    return torch.gather(x, 2, src[..., None].expand(h, rows, 8, 8))


def kv_slab_logical(slab: torch.Tensor) -> torch.Tensor:
    """slab [64, d] as stored (memory order [d/64][64][64]) -> logical [64, d]."""
    rows, d = slab.shape
    x = slab.reshape(d // 64, rows, 8, 8)
    return _kv_perm(x).permute(1, 0, 2, 3).reshape(rows, d)


def kv_slab_store(logical: torch.Tensor) -> torch.Tensor:
    """inverse of kv_slab_logical: logical [64, d] -> the [64, d]-shaped view of the stored bytes."""
    rows, d = logical.shape
    x = logical.reshape(rows, d // 64, 8, 8).permute(1, 0, 2, 3).contiguous()
    return _kv_perm(x).reshape(rows, d)
