"""Per-kernel parity: every CUDA kernel, called through the C ABI, against the fp32 oracle
(oracle/llama_oracle.py) or plain fp32 torch on the same seeded inputs.

Tolerances: the kernels take bf16 inputs and accumulate in fp32, so where the reference is
fed the same bf16-rounded inputs the only differences are summation order (GEMM, norms:
1e-3 relative to the tensor's max) and bf16 rounding of outputs / of the softmax
probabilities fed to the PV product (attention runs in fp16: 2e-3 relative to max |out|).  Activations
leave every kernel as bf16 hi + lo planes; hi + lo must match the fp32 reference to ~5e-5.
"""
import math

import pytest
import torch

from oracle import llama_oracle as O

pytestmark = pytest.mark.gpu

DEV = "cuda"


@pytest.fixture(scope="module")
def K():
    from fasttalk_llm_microservice_b200 import build, kernels
    build.build()
    torch.backends.cuda.matmul.allow_tf32 = False
    return kernels


def rel_err(a, b):
    return (a.float() - b.float()).abs().max().item() / max(b.float().abs().max().item(), 1e-12)


# ------------------------------------------------------------------ K3 GEMM
def hilo(a, b):
    return a.float() + b.float()


@pytest.mark.parametrize("n_out,T,Kd,splits", [
    (128, 1, 64, 1), (128, 5, 64, 1), (256, 16, 128, 1), (256, 17, 128, 2),
    (3072, 64, 2048, 0), (2048, 64, 2048, 0), (16384, 64, 2048, 0), (2048, 64, 8192, 0),
    (2048, 1, 8192, 0), (1024, 100, 512, 3), (1024, 300, 512, 1), (512, 700, 256, 2),
    (5120, 33, 3072, 0), (128256, 8, 256, 1),
])
@pytest.mark.parametrize("with_lo", [False, True])
def test_gemm_tcgen05(K, n_out, T, Kd, splits, with_lo):
    g = torch.Generator(device=DEV).manual_seed(n_out * 7 + T * 3 + Kd)
    w = (torch.randn(n_out, Kd, device=DEV, generator=g) * 0.05).bfloat16()
    xf = torch.randn(T, Kd, device=DEV, generator=g)
    x, x_lo = K.split_hi_lo(xf)
    part = K.gemm_bf16(w, x, x_lo if with_lo else None, n_splits=splits)
    torch.cuda.synchronize()
    ref = (hilo(x, x_lo) if with_lo else x.float()) @ w.float().T
    got = part.sum(0)
    assert got.shape == ref.shape
    assert rel_err(got, ref) < 1e-3, (part.shape, rel_err(got, ref))
    if with_lo:      # the two-plane product tracks the fp32 activations ~8 bits closer than bf16 alone
        exact = xf @ w.float().T
        assert rel_err(got, exact) < 5e-5 + rel_err(got, ref)


def test_gemm_bf16_out(K):
    g = torch.Generator(device=DEV).manual_seed(5)
    w = (torch.randn(512, 256, device=DEV, generator=g) * 0.05).bfloat16()
    x, x_lo = K.split_hi_lo(torch.randn(70, 256, device=DEV, generator=g))
    hi, lo = K.gemm_bf16(w, x, x_lo, mode="bf16")
    ref = hilo(x, x_lo) @ w.float().T
    assert rel_err(hi, ref) < 8e-3           # one bf16 rounding of the output
    assert rel_err(hilo(hi, lo), ref) < 1e-4


@pytest.mark.parametrize("f,T,h", [(8192, 64, 2048), (512, 5, 256), (1024, 200, 384)])
def test_gemm_fused_swiglu(K, f, T, h):
    from fasttalk_llm_microservice_b200.weights import interleave_gate_up
    g = torch.Generator(device=DEV).manual_seed(f + T)
    wgu = (torch.randn(2 * f, h, device=DEV, generator=g) * 0.05).bfloat16()
    x, x_lo = K.split_hi_lo(torch.randn(T, h, device=DEV, generator=g))
    hi, lo = K.gemm_bf16(interleave_gate_up(wgu), x, x_lo, mode="swiglu")
    ref = O.swiglu(hilo(x, x_lo) @ wgu.float().T)
    assert hi.shape == (T, f)
    assert rel_err(hilo(hi, lo), ref) < 2e-4, rel_err(hilo(hi, lo), ref)


@pytest.mark.parametrize("n_out,T,Kd,mode,what", [
    (16384, 2600, 2048, "swiglu", "gemm_pair_kernel<256>: 64 row pairs x 11 token tiles of 256 >= 8 rounds per cluster"),
    (2048, 2600, 2048, "bf16", "gemm_pair_kernel<128>: too few 256-token work units, 128-token tiles"),
    (16384, 1300, 2048, "swiglu", "gemm_pair_kernel<128> with a ragged last token tile"),
])
def test_gemm_cta_pair_kernels(K, n_out, T, Kd, mode, what):
    """The cta_group::2 prefill kernels at kernel level (round 1 only reached the 256-token instantiation through
    the 8300-token engine test): integer-valued bf16 operands make the fp32 result exact, so every tile, the
    peer-CTA half of every activation tile and the ragged tail are checked bit for bit; then random data
    against fp32 torch."""
    from fasttalk_llm_microservice_b200.weights import interleave_gate_up
    g = torch.Generator(device=DEV).manual_seed(n_out + T)
    w = torch.randint(-3, 4, (n_out, Kd), device=DEV, generator=g).bfloat16()
    x = torch.randint(-2, 3, (T, Kd), device=DEV, generator=g).bfloat16()
    x_lo = torch.randint(-1, 2, (T, Kd), device=DEV, generator=g).bfloat16()
    if mode == "bf16":
        wi = torch.randint(-1, 2, (n_out, Kd), device=DEV, generator=g).bfloat16()      # |y| small enough for exact bf16 hi+lo
        hi, lo = K.gemm_bf16(wi, x, x_lo, mode="bf16")
        ref = hilo(x, x_lo) @ wi.float().T
        assert torch.equal(hilo(hi, lo), ref)
    else:
        hi, lo = K.gemm_bf16(interleave_gate_up(w), x, x_lo, mode="swiglu")
        ref = O.swiglu(hilo(x, x_lo) @ w.float().T)
        assert hi.shape == (T, n_out // 2)
        assert rel_err(hilo(hi, lo), ref) < 1e-4, rel_err(hilo(hi, lo), ref)
    wr = (torch.randn(n_out, Kd, device=DEV, generator=g) * 0.05).bfloat16()
    xr, xr_lo = K.split_hi_lo(torch.randn(T, Kd, device=DEV, generator=g))
    if mode == "bf16":
        hi, lo = K.gemm_bf16(wr, xr, xr_lo, mode="bf16")
        assert rel_err(hilo(hi, lo), hilo(xr, xr_lo) @ wr.float().T) < 1e-4
    else:
        hi, lo = K.gemm_bf16(interleave_gate_up(wr), xr, xr_lo, mode="swiglu")
        assert rel_err(hilo(hi, lo), O.swiglu(hilo(xr, xr_lo) @ wr.float().T)) < 2e-4


@pytest.mark.parametrize("n_out,T,Kd,what", [
    (2048, 8192, 4096, "256-token tiles, 256 units on 74 pairs: 3 rounds + 34 units cut into 68 halves"),
    (2048, 5000, 4096, "ragged tail: 160 units -> 2 rounds + 12 units cut into 24 halves, the last half beyond T"),
    (2048, 1000, 4096, "less than two rounds of 256-token units: 128-token tiles"),
    (1024, 700, 2048, "single wave of pairs (32 tiles or more, two or more token tiles)"),
])
def test_gemm_fp16_single_plane_form(K, n_out, T, Kd, what):
    """The form the engine's large prefill steps run (x_planes = 3): ONE fp16 activation plane against fp16 weights on the
    CTA-pair kernels, residual epilogue with all three outputs.  Integer-valued operands make the fp32 accumulation exact:
    every tile, both work-list kinds and the ragged tail are compared bit for bit with torch."""
    from fasttalk_llm_microservice_b200 import native as N
    g = torch.Generator(device=DEV).manual_seed(n_out + T)
    w = torch.randint(-2, 3, (n_out, Kd), device=DEV, generator=g).half()
    x = torch.randint(-2, 3, (T, Kd), device=DEV, generator=g).half()
    resid0 = torch.randint(-8, 9, (T, n_out), device=DEV, generator=g).float()
    resid = resid0.clone()
    gamma = torch.ones(n_out, device=DEV, dtype=torch.bfloat16)
    hi = torch.full((T, n_out), 7.0, device=DEV, dtype=torch.bfloat16)
    lo = torch.full_like(hi, 7.0)
    f16 = torch.full((T, n_out), 7.0, device=DEV, dtype=torch.float16)
    ssq = torch.zeros(n_out // 128, T, device=DEV)
    K.gemm_fused(w, x, None, N.EPI_RESID, n_splits=1, x_planes=3, resid=resid, gamma=gamma, out=hi, out_lo=lo, out_f16=f16,
                 ssq_out=ssq, ssq_stride=T)
    torch.cuda.synchronize()
    ref = resid0 + x.float() @ w.float().T
    assert torch.equal(resid, ref)
    assert torch.equal(hilo(hi, lo), ref)                       # |values| < 2^16: hi + lo is exact
    assert torch.equal(f16.float(), ref.clamp(-65504, 65504).half().float())
    assert torch.allclose(ssq.sum(0), (ref * ref).sum(1), rtol=1e-5)
    # the bf16 planes are optional next to the fp16 plane
    resid2, f16b = resid0.clone(), torch.zeros_like(f16)
    K.gemm_fused(w, x, None, N.EPI_RESID, n_splits=1, x_planes=3, resid=resid2, gamma=gamma, out_f16=f16b, ssq_out=ssq, ssq_stride=T)
    torch.cuda.synchronize()
    assert torch.equal(resid2, ref) and torch.equal(f16b, f16)


def test_attn_decode_empty_context_rows_are_zero(K):
    """A sequence with ctx_lens == 0 has no KV tile and so no CTA works on it: its output rows are defined
    (zero) instead of left as whatever the buffer held (round-1 verdict, weak #11)."""
    g = torch.Generator(device=DEV).manual_seed(4)
    nq, nkv, d = 8, 2, 64
    kv = torch.randn(8, nkv, 2, 64, d, device=DEV, generator=g).half()
    pt = torch.arange(8, device=DEV, dtype=torch.int32).reshape(4, 2).contiguous()
    ctx = torch.tensor([70, 0, 5, 0], device=DEV, dtype=torch.int32)
    q = torch.randn(4, nq, d, device=DEV, generator=g).half()
    for _ in range(2):                      # poison the blocks the caching allocator will hand to the wrapper's outputs
        junk = [torch.full((4, nq * d), 7.0, device=DEV, dtype=torch.bfloat16) for _ in range(2)]
        del junk
    hi, lo = K.attn_decode(q, kv, pt, ctx)
    out = hilo(hi, lo)
    assert torch.isfinite(out).all()
    assert out[1].abs().max().item() == 0 and out[3].abs().max().item() == 0
    assert out[0].abs().max().item() > 0 and out[2].abs().max().item() > 0


def test_gemm_linearity(K):
    """Size-independent property at a full config shape: W(x1 + x2) == W x1 + W x2 on
    bf16-exact inputs (small integers), bit-exact in fp32."""
    g = torch.Generator(device=DEV).manual_seed(9)
    w = torch.randint(-3, 4, (8192, 2048), device=DEV, generator=g).bfloat16()
    x1 = torch.randint(-2, 3, (64, 2048), device=DEV, generator=g).bfloat16()
    x2 = torch.randint(-2, 3, (64, 2048), device=DEV, generator=g).bfloat16()
    y = lambda x: K.gemm_bf16(w, x).sum(0)
    assert torch.equal(y(x1 + x2), y(x1) + y(x2))
    assert torch.equal(y(x1), x1.float() @ w.float().T)
    assert torch.equal(K.gemm_bf16(w, x1, x2).sum(0), y(x1) + y(x2))      # second plane == second product


# ------------------------------------------------------------------ K3 fused epilogues (norm-deferred chain)
@pytest.mark.parametrize("h,T,Kd,splits", [(2048, 64, 2048, 0), (2048, 64, 8192, 0), (256, 5, 512, 1),
                                           (384, 37, 256, 2), (2048, 200, 2048, 0), (2048, 16, 1024, 8)])
def test_gemm_resid_epilogue(K, h, T, Kd, splits):
    """O / down projection: split-K rendezvous, resid += y, planes of resid * gamma_next,
    per-row-tile sums of squares."""
    from fasttalk_llm_microservice_b200 import native as N
    g = torch.Generator(device=DEV).manual_seed(h + T + Kd)
    w = (torch.randn(h, Kd, device=DEV, generator=g) * 0.05).bfloat16()
    x, x_lo = K.split_hi_lo(torch.randn(T, Kd, device=DEV, generator=g))
    resid = torch.randn(T, h, device=DEV, generator=g)
    gamma = (1 + 0.1 * torch.randn(h, device=DEV, generator=g)).bfloat16()
    ref_resid = resid + hilo(x, x_lo) @ w.float().T
    out, out_lo = torch.zeros(T, h, device=DEV, dtype=torch.bfloat16), torch.zeros(T, h, device=DEV, dtype=torch.bfloat16)
    parts = h // 128
    ssq = torch.full((parts, T + 3), -1.0, device=DEV)
    r2 = resid.clone()
    K.gemm_fused(w, x, x_lo, N.EPI_RESID, n_splits=splits, resid=r2, gamma=gamma, out=out, out_lo=out_lo,
                 ssq_out=ssq, ssq_stride=T + 3)
    torch.cuda.synchronize()
    assert rel_err(r2, ref_resid) < 1e-3
    assert rel_err(hilo(out, out_lo), ref_resid * gamma.float()) < 1e-3
    ref_ssq = (ref_resid ** 2).reshape(T, parts, 128).sum(-1).T
    assert rel_err(ssq[:, :T], ref_ssq) < 2e-3
    assert (ssq[:, T:] == -1).all()


@pytest.mark.parametrize("nq,nkv,d,T,Kd,splits", [(32, 8, 64, 64, 2048, 0), (24, 8, 128, 9, 3072, 0),
                                                  (8, 2, 64, 5, 256, 1), (3, 1, 128, 70, 384, 2),
                                                  (32, 8, 128, 130, 512, 0)])
def test_gemm_qkv_rope_epilogue(K, nq, nkv, d, T, Kd, splits):
    """QKV projection: split-K rendezvous, 1/rms, RoPE, q out and swizzled paged K/V append."""
    from fasttalk_llm_microservice_b200 import native as N
    from fasttalk_llm_microservice_b200.model_spec import ModelSpec
    from fasttalk_llm_microservice_b200.weights import rope_tables
    spec = ModelSpec("t", hidden=Kd, ffn=128, layers=1, n_q_heads=nq, n_kv_heads=nkv, head_dim=d)
    g = torch.Generator(device=DEV).manual_seed(nq + d + T)
    n_out = (nq + 2 * nkv) * d
    w = (torch.randn(n_out, Kd, device=DEV, generator=g) * 0.05).bfloat16()
    x, x_lo = K.split_hi_lo(torch.randn(T, Kd, device=DEV, generator=g))
    parts = 3
    ssq = torch.rand(parts, T, device=DEV, generator=g) * Kd
    rstd = torch.rsqrt(ssq.sum(0) / Kd + 1e-5)
    n_seq, max_pages, num_pages = 3, 8, 32
    rows = torch.randint(0, n_seq, (T,), device=DEV, generator=g, dtype=torch.int32)
    # distinct (sequence, position) per token
    pos = torch.zeros(T, dtype=torch.int32, device=DEV)
    seen = {}
    for t in range(T):
        r = int(rows[t])
        seen[r] = seen.get(r, int(torch.randint(0, 60, (1,), generator=torch.Generator().manual_seed(r)))) + 1
        pos[t] = seen[r]
    pt = torch.randperm(num_pages, device=DEV, generator=g)[:n_seq * max_pages].reshape(n_seq, max_pages).int().contiguous()
    cos, sin = rope_tables(spec, 512, device=DEV)
    kv = K.kv_cache_alloc(num_pages, nkv, d, DEV)
    q = torch.zeros(T, nq, d, device=DEV, dtype=torch.float16)
    K.gemm_fused(w, x, x_lo, N.EPI_QKV_ROPE, n_splits=splits, ssq_in=ssq, ssq_parts=parts, ssq_stride=T,
                 inv_h=1.0 / Kd, eps=1e-5, rope_cos=cos, rope_sin=sin, positions=pos, seq_rows=rows, page_table=pt,
                 max_pages=max_pages, q_out=q, kv_layer=kv, nq=nq, nkv=nkv, head_dim=d)
    torch.cuda.synchronize()
    full = (hilo(x, x_lo) @ w.float().T) * rstd[:, None]
    qr = full[:, :nq * d].reshape(T, nq, d)
    kr = full[:, nq * d:(nq + nkv) * d].reshape(T, nkv, d)
    vr = full[:, (nq + nkv) * d:].reshape(T, nkv, d)
    c, s = O.rope_cos_sin(spec, pos.cpu().long())
    c, s = c.to(DEV), s.to(DEV)
    assert rel_err(q, O.apply_rope(qr, c, s)) < 1e-3
    kref = O.apply_rope(kr, c, s)
    for t in range(T):
        page = pt[rows[t], pos[t] // 64].item()
        r = pos[t].item() % 64
        for hk in range(nkv):
            assert rel_err(K.kv_slab_logical(kv[page, hk, 0])[r], kref[t, hk]) < 1e-3
            assert rel_err(K.kv_slab_logical(kv[page, hk, 1])[r], vr[t, hk]) < 1e-3


def test_gemm_logits_and_swiglu_with_rstd(K):
    from fasttalk_llm_microservice_b200 import native as N
    from fasttalk_llm_microservice_b200.weights import interleave_gate_up
    g = torch.Generator(device=DEV).manual_seed(3)
    T, h, V, f = 33, 256, 1024, 512
    x, x_lo = K.split_hi_lo(torch.randn(T, h, device=DEV, generator=g))
    ssq = torch.rand(2, T, device=DEV, generator=g) * h
    rstd = torch.rsqrt(ssq.sum(0) / h + 1e-5)
    w = (torch.randn(V, h, device=DEV, generator=g) * 0.05).bfloat16()
    logits = torch.zeros(T, V, device=DEV)
    K.gemm_fused(w, x, x_lo, N.EPI_LOGITS, out=logits, ssq_in=ssq, ssq_parts=2, ssq_stride=T, inv_h=1.0 / h, eps=1e-5)
    assert rel_err(logits, (hilo(x, x_lo) @ w.float().T) * rstd[:, None]) < 1e-3
    wgu = (torch.randn(2 * f, h, device=DEV, generator=g) * 0.05).bfloat16()
    act, act_lo = torch.zeros(T, f, device=DEV, dtype=torch.bfloat16), torch.zeros(T, f, device=DEV, dtype=torch.bfloat16)
    K.gemm_fused(interleave_gate_up(wgu), x, x_lo, N.EPI_SWIGLU, out=act, out_lo=act_lo, ssq_in=ssq, ssq_parts=2,
                 ssq_stride=T, inv_h=1.0 / h, eps=1e-5)
    ref = O.swiglu((hilo(x, x_lo) @ wgu.float().T) * rstd[:, None])
    assert rel_err(hilo(act, act_lo), ref) < 2e-4


def test_embed_and_gather_planes(K):
    g = torch.Generator(device=DEV).manual_seed(6)
    V, h, T, parts = 1000, 2048, 19, 16
    table = torch.randn(V, h, device=DEV, generator=g).bfloat16()
    gamma = (1 + 0.1 * torch.randn(h, device=DEV, generator=g)).bfloat16()
    toks = torch.randint(0, V, (T,), device=DEV, generator=g, dtype=torch.int32)
    resid, hi, lo, ssq = K.embed_planes(toks, table, gamma, parts)
    ref = table[toks.long()].float()
    assert torch.equal(resid, ref)
    assert rel_err(hilo(hi, lo), ref * gamma.float()) < 5e-5
    assert rel_err(ssq.sum(0), (ref ** 2).sum(-1)) < 1e-5 and (ssq[1:] == 0).all()
    rows = torch.tensor([18, 0, 7], device=DEV, dtype=torch.int32)
    ghi, glo, gss = K.gather_planes(hi, lo, ssq, rows)
    assert torch.equal(ghi, hi[rows.long()]) and torch.equal(glo, lo[rows.long()])
    assert torch.equal(gss, ssq[:, rows.long()])


# ------------------------------------------------------------------ K1 / K2
def test_add_rmsnorm_partials(K):
    g = torch.Generator(device=DEV).manual_seed(1)
    T, h, S = 37, 2048, 3
    resid = torch.randn(T, h, device=DEV, generator=g)
    part = torch.randn(S, T, h, device=DEV, generator=g)
    w = (1 + 0.1 * torch.randn(h, device=DEV, generator=g)).bfloat16()
    ref_res = resid + part.sum(0)
    ref = O.rmsnorm(ref_res, w, 1e-5)
    r2 = resid.clone()
    xn, xn_lo = K.add_rmsnorm(r2, w, 1e-5, partials=part)
    assert rel_err(r2, ref_res) < 1e-6
    assert rel_err(xn, ref) < 6e-3
    assert rel_err(hilo(xn, xn_lo), ref) < 5e-5
    part9 = torch.randn(9, T, h, device=DEV, generator=g)          # unrolled-by-4 reduce + tail
    r3 = resid.clone()
    K.add_rmsnorm(r3, w, 1e-5, partials=part9)
    assert rel_err(r3, resid + part9.sum(0)) < 1e-6


def test_add_rmsnorm_embed_and_rows(K):
    g = torch.Generator(device=DEV).manual_seed(2)
    V, h, T = 1000, 3072, 19
    table = torch.randn(V, h, device=DEV, generator=g).bfloat16()
    toks = torch.randint(0, V, (T,), device=DEV, generator=g, dtype=torch.int32)
    w = (1 + 0.1 * torch.randn(h, device=DEV, generator=g)).bfloat16()
    resid = torch.zeros(T, h, device=DEV)
    xn, xn_lo = K.add_rmsnorm(resid, w, 1e-5, embed_tokens=toks, embed_table=table)
    assert torch.equal(resid, table[toks.long()].float())
    assert rel_err(xn, O.rmsnorm(resid, w, 1e-5)) < 6e-3
    assert rel_err(hilo(xn, xn_lo), O.rmsnorm(resid, w, 1e-5)) < 5e-5
    rows = torch.tensor([18, 0, 7], device=DEV, dtype=torch.int32)
    before = resid.clone()
    xr, _ = K.add_rmsnorm(resid, w, 1e-5, row_idx=rows)
    assert torch.equal(resid, before)
    assert rel_err(xr, O.rmsnorm(before[rows.long()], w, 1e-5)) < 6e-3


# ------------------------------------------------------------------ K4
@pytest.mark.parametrize("nq,nkv,d", [(32, 8, 64), (24, 8, 128), (3, 1, 128)])
def test_rope_kv_append(K, nq, nkv, d):
    from fasttalk_llm_microservice_b200.model_spec import ModelSpec
    from fasttalk_llm_microservice_b200.weights import rope_tables
    spec = ModelSpec("t", hidden=nq * d, ffn=128, layers=1, n_q_heads=nq, n_kv_heads=nkv, head_dim=d)
    g = torch.Generator(device=DEV).manual_seed(3)
    S, max_pages, num_pages = 2, 4, 16
    # three sequences with tokens at assorted positions (incl. page boundaries)
    pos = torch.tensor([0, 1, 63, 64, 65, 130, 5, 200], device=DEV, dtype=torch.int32)
    rows = torch.tensor([0, 0, 0, 0, 0, 1, 2, 2], device=DEV, dtype=torch.int32)
    T = pos.numel()
    pt = torch.randperm(num_pages, device=DEV, generator=g)[:12].reshape(3, 4).int().contiguous()
    qkv = torch.randn(S, T, (nq + 2 * nkv) * d, device=DEV, generator=g)
    cos, sin = rope_tables(spec, 256, device=DEV)
    kv = K.kv_cache_alloc(num_pages, nkv, d, DEV)
    q = K.rope_kv_append(qkv, cos, sin, pos, rows, pt, kv, nq, nkv, d)
    torch.cuda.synchronize()
    full = qkv.sum(0)
    qr = full[:, :nq * d].reshape(T, nq, d)
    kr = full[:, nq * d:(nq + nkv) * d].reshape(T, nkv, d)
    vr = full[:, (nq + nkv) * d:].reshape(T, nkv, d)
    c, s = O.rope_cos_sin(spec, pos.cpu().long())
    c, s = c.to(DEV), s.to(DEV)
    assert rel_err(q, O.apply_rope(qr, c, s)) < 6e-3
    kref = O.apply_rope(kr, c, s)
    for t in range(T):
        page = pt[rows[t], pos[t] // 64].item()
        r = pos[t].item() % 64
        for hk in range(nkv):
            kl = K.kv_slab_logical(kv[page, hk, 0])
            vl = K.kv_slab_logical(kv[page, hk, 1])
            assert rel_err(kl[r], kref[t, hk]) < 6e-3
            assert rel_err(vl[r], vr[t, hk]) < 6e-3


def test_silu_mul(K):
    g = torch.Generator(device=DEV).manual_seed(4)
    gu = torch.randn(2, 9, 2 * 8192, device=DEV, generator=g)
    act, act_lo = K.silu_mul(gu)
    assert rel_err(act, O.swiglu(gu.sum(0))) < 6e-3
    assert rel_err(hilo(act, act_lo), O.swiglu(gu.sum(0))) < 5e-5


# ------------------------------------------------------------------ K5 decode attention
def build_paged_kv(K, ctx_lens, nkv, d, g, extra_pages=5):
    """Random bf16 K/V per sequence stored through a shuffled page table."""
    B = len(ctx_lens)
    n_pages = [max(1, -(-c // 64)) for c in ctx_lens]
    total = sum(n_pages) + extra_pages
    perm = torch.randperm(total, device=DEV, generator=g).int()
    max_pages = max(n_pages) + 1
    pt = torch.zeros(B, max_pages, dtype=torch.int32, device=DEV)
    kv = K.kv_cache_alloc(total, nkv, d, DEV)
    # poison: garbage beyond ctx must be ignored (finite, as in a recycled page)
    kv.copy_(torch.randn(kv.shape, device=DEV, generator=g).half() * 50)
    ks, vs, off = [], [], 0
    for b, c in enumerate(ctx_lens):
        k = torch.randn(c, nkv, d, device=DEV, generator=g).half()
        v = torch.randn(c, nkv, d, device=DEV, generator=g).half()
        ks.append(k); vs.append(v)
        for p in range(n_pages[b]):
            page = perm[off + p].item()
            pt[b, p] = page
            n = min(64, c - p * 64)
            for hk in range(nkv):
                kslab = K.kv_slab_logical(kv[page, hk, 0]).clone()
                vslab = K.kv_slab_logical(kv[page, hk, 1]).clone()
                kslab[:n] = k[p * 64:p * 64 + n, hk]
                vslab[:n] = v[p * 64:p * 64 + n, hk]
                kv[page, hk, 0] = K.kv_slab_store(kslab)
                kv[page, hk, 1] = K.kv_slab_store(vslab)
        off += n_pages[b]
    return kv, pt, ks, vs


@pytest.mark.parametrize("nq,nkv,d,ctx_lens,n_ctas", [
    (32, 8, 64, [1], 0),
    (32, 8, 64, [63, 64, 65, 1, 200], 0),
    (32, 8, 64, [63, 64, 65, 1, 200], 3),          # CTAs span several (seq, head) segments
    (32, 8, 64, [700], 0),                          # one sequence split over many CTAs
    (8, 2, 64, [129, 640, 77, 300], 7),
    (24, 8, 128, [130, 5, 257], 0),                 # 3B shape: group of 3, d=128
    (32, 8, 128, [513], 5),                         # 8B shape
    (8, 1, 64, [100, 31], 2),                       # group of 8
    (4, 4, 64, [90, 10], 0),                        # MHA (group of 1)
])
def test_attn_decode(K, nq, nkv, d, ctx_lens, n_ctas):
    g = torch.Generator(device=DEV).manual_seed(sum(ctx_lens) + nq + d)
    kv, pt, ks, vs = build_paged_kv(K, ctx_lens, nkv, d, g)
    B = len(ctx_lens)
    q = torch.randn(B, nq, d, device=DEV, generator=g).half()
    ctx = torch.tensor(ctx_lens, device=DEV, dtype=torch.int32)
    out, out_lo = K.attn_decode(q, kv, pt, ctx, n_ctas=n_ctas)
    torch.cuda.synchronize()
    assert rel_err(out, hilo(out, out_lo)) < 5e-3         # lo is the rounding residual of hi
    out = hilo(out, out_lo)
    for b in range(B):
        ref = O.attention(q[b:b + 1].float(), ks[b].float(), vs[b].float(), ctx_lens[b] - 1)
        assert rel_err(out[b:b + 1], ref) < 2e-3, (b, rel_err(out[b:b + 1], ref))
    # repeat on the same scratch: the per-segment arrival counters must have re-armed themselves
    again = hilo(*K.attn_decode(q, kv, pt, ctx, n_ctas=n_ctas))
    assert torch.equal(again, out)


def test_attn_decode_cfg2_shape(K):
    """BASELINE config 2 shape: 64 sequences x 2048 context, Llama-3.2-1B heads.  Checked by a
    size-independent property as well: softmax weights sum to 1, so with V == const the
    output is that constant, whatever the split / merge schedule."""
    g = torch.Generator(device=DEV).manual_seed(77)
    nq, nkv, d, B, ctxl = 32, 8, 64, 64, 2048
    pages = B * 32
    kv = (torch.randn(pages, nkv, 2, 64, d, device=DEV, generator=g)).half()
    kv[:, :, 1] = 0.5
    pt = torch.randperm(pages, device=DEV, generator=g).int().reshape(B, 32).contiguous()
    q = torch.randn(B, nq, d, device=DEV, generator=g).half()
    ctx = torch.full((B,), ctxl, device=DEV, dtype=torch.int32)
    ctx[3] = 1999
    out = hilo(*K.attn_decode(q, kv, pt, ctx))
    assert (out - 0.5).abs().max().item() < 4e-3
    # and against the oracle on two of the sequences
    kv2 = (torch.randn(pages, nkv, 2, 64, d, device=DEV, generator=g)).half()
    out2 = hilo(*K.attn_decode(q, kv2, pt, ctx))
    for b in (3, 40):
        n = int(ctx[b])
        pg = pt[b].long()
        kl = torch.stack([torch.stack([K.kv_slab_logical(kv2[p, h, 0]) for h in range(nkv)], 1) for p in pg]).reshape(-1, nkv, d)[:n]
        vl = torch.stack([torch.stack([K.kv_slab_logical(kv2[p, h, 1]) for h in range(nkv)], 1) for p in pg]).reshape(-1, nkv, d)[:n]
        ref = O.attention(q[b:b + 1].float(), kl.float(), vl.float(), n - 1)
        assert rel_err(out2[b:b + 1], ref) < 2e-3


# ------------------------------------------------------------------ K6 prefill attention
@pytest.mark.parametrize("nq,nkv,d,reqs", [
    (8, 2, 64, [(0, 5)]),                       # (cached prefix length, new tokens)
    (8, 2, 64, [(0, 64), (0, 65), (10, 1)]),
    (32, 8, 64, [(100, 130), (0, 200), (64, 64)]),
    (24, 8, 128, [(33, 70), (0, 129)]),
    (3, 1, 128, [(250, 20)]),
])
def test_attn_prefill(K, nq, nkv, d, reqs):
    g = torch.Generator(device=DEV).manual_seed(len(reqs) * 31 + nq + d)
    totals = [p + n for p, n in reqs]
    kv, pt, ks, vs = build_paged_kv(K, totals, nkv, d, g)
    q_start = [0]
    for _, n in reqs:
        q_start.append(q_start[-1] + n)
    T = q_start[-1]
    q = torch.randn(T, nq, d, device=DEV, generator=g).half()
    out = hilo(*K.attn_prefill(q, kv, pt, q_start, list(range(len(reqs))), [p for p, _ in reqs]))
    torch.cuda.synchronize()
    for r, (p0, n) in enumerate(reqs):
        ref = O.attention(q[q_start[r]:q_start[r + 1]].float(), ks[r].float(), vs[r].float(), p0)
        got = out[q_start[r]:q_start[r + 1]]
        assert rel_err(got, ref) < 2e-3, (r, rel_err(got, ref))


@pytest.mark.parametrize("nq,nkv,d", [(8, 2, 64), (6, 2, 64), (32, 8, 64), (32, 8, 128), (24, 8, 128), (3, 1, 128)])
def test_attn_prefill_tcgen05_against_mma_sync_and_oracle(K, nq, nkv, d):
    """The tcgen05 / TMEM prefill kernel (GQA groups of 3 and 4, head_dim 64 and 128 — the 1B, 3B and 8B shapes) on the
    same inputs as the round-1 mma.sync kernel (`tensor_cores=False` selects it): cold and prefix-cached requests, ragged
    query tiles, a single query, a context long enough for the lazy rescale and the ring to wrap several times.  Both
    against the fp32 oracle, and against each other."""
    reqs = [(0, 130), (700, 224), (64, 1), (5, 64), (1500, 33)]
    g = torch.Generator(device=DEV).manual_seed(nq * 13 + d)
    totals = [p + n for p, n in reqs]
    kv, pt, ks, vs = build_paged_kv(K, totals, nkv, d, g)
    q_start = [0]
    for _, n in reqs:
        q_start.append(q_start[-1] + n)
    q = (torch.randn(q_start[-1], nq, d, device=DEV, generator=g) * 1.5).half()      # some spread in the scores
    args = (q, kv, pt, q_start, list(range(len(reqs))), [p for p, _ in reqs])
    tc = hilo(*K.attn_prefill(*args, tensor_cores=True))
    ms = hilo(*K.attn_prefill(*args, tensor_cores=False))
    torch.cuda.synchronize()
    assert torch.isfinite(tc).all()
    assert rel_err(tc, ms) < 1.5e-3, rel_err(tc, ms)
    for r, (p0, n) in enumerate(reqs):
        ref = O.attention(q[q_start[r]:q_start[r + 1]].float(), ks[r].float(), vs[r].float(), p0)
        assert rel_err(tc[q_start[r]:q_start[r + 1]], ref) < 2e-3, (r, rel_err(tc[q_start[r]:q_start[r + 1]], ref))
        assert rel_err(ms[q_start[r]:q_start[r + 1]], ref) < 2e-3


# ------------------------------------------------------------------ K7 sampling
def test_sample_argmax_matches_torch(K):
    g = torch.Generator(device=DEV).manual_seed(8)
    logits = torch.randn(64, 128256, device=DEV, generator=g)
    logits[5, 1000] = logits[5, 77] = 50.0          # tie: lowest index wins
    logits[6, 128255] = 60.0
    tok = K.sample(logits)
    assert torch.equal(tok.long(), torch.argmax(logits, -1))
    assert tok[5].item() == 77 and tok[6].item() == 128255
    t0 = torch.zeros(64, device=DEV)
    assert torch.equal(K.sample(logits, temperature=t0).long(), torch.argmax(logits, -1))


@pytest.mark.parametrize("temp,top_k,top_p", [(1.0, 0, 1.0), (0.7, 0, 0.9), (0.7, 40, 0.9), (1.3, 5, 1.0), (0.5, 0, 0.3)])
def test_sample_distribution(K, temp, top_k, top_p):
    """Draw 8192 samples of one logits row with different seeds; the empirical distribution
    must stay inside the oracle's support and within 4 % total variation of it."""
    g = torch.Generator(device=DEV).manual_seed(10)
    V, n = 4096, 8192
    row = torch.randn(V, device=DEV, generator=g) * 2.0
    logits = row[None].repeat(n, 1).contiguous()
    seeds = torch.arange(n, device=DEV, dtype=torch.int64) * 7919 + 13
    tok = K.sample(logits, temperature=torch.full((n,), temp, device=DEV),
                   top_k=torch.full((n,), top_k, device=DEV, dtype=torch.int32),
                   top_p=torch.full((n,), top_p, device=DEV), seeds=seeds)
    ref = O.sampling_support(row.cpu(), temp, top_k, top_p)
    emp = torch.bincount(tok.cpu().long(), minlength=V).double() / n
    assert emp[ref == 0].sum().item() < 1e-9, "sampled a token outside the kept set"
    tv = 0.5 * (emp - ref).abs().sum().item()
    # expected TV of an n-sample multinomial with k-token support is ~ sqrt(k / (2 pi n))
    k = int((ref > 0).sum())
    assert tv < 0.04 + math.sqrt(k / (2 * math.pi * n)), (tv, k)


def test_sample_full_vocab_row(K):
    g = torch.Generator(device=DEV).manual_seed(12)
    logits = torch.randn(4, 128256, device=DEV, generator=g) * 3
    seeds = torch.tensor([1, 2, 3, 4], device=DEV, dtype=torch.int64)
    tok = K.sample(logits, temperature=torch.full((4,), 0.7, device=DEV),
                   top_k=torch.zeros(4, device=DEV, dtype=torch.int32),
                   top_p=torch.full((4,), 0.9, device=DEV), seeds=seeds)
    for b in range(4):
        ref = O.sampling_support(logits[b].cpu(), 0.7, 0, 0.9)
        assert ref[tok[b].item()] > 0


def test_gemm_split_k_alternating_arrival_words(K):
    """ft_gemm_epilogue.sync_mode 1 / 2 (what the engine uses): launches that share one `sync` buffer alternate
    between the two words of each counter pair and clear the other for ALL slots, whatever their tile counts —
    a 24-tile launch, two 16-tile launches, 24 tiles again, ... must all match the self re-arming mode."""
    from fasttalk_llm_microservice_b200 import native as N
    g = torch.Generator(device=DEV).manual_seed(77)
    T = 64
    shapes = [(3072, 2048), (2048, 2048), (2048, 8192), (3072, 2048), (2048, 8192), (2048, 2048), (3072, 2048)]
    sync = torch.zeros(2 * 148, device=DEV, dtype=torch.int32)
    scratch = torch.empty(148 * 128 * 128, device=DEV, dtype=torch.float32)
    par = 0
    for rep, (n_out, Kd) in enumerate(shapes * 2):
        w = (torch.randn(n_out, Kd, device=DEV, generator=g) * 0.05).bfloat16()
        x, x_lo = K.split_hi_lo(torch.randn(T, Kd, device=DEV, generator=g))
        gamma = (1 + 0.1 * torch.randn(n_out, device=DEV, generator=g)).bfloat16()
        resid0 = torch.randn(T, n_out, device=DEV, generator=g)
        outs = []
        for mode in ("alt", "self"):
            z = lambda: torch.zeros(T, n_out, device=DEV, dtype=torch.bfloat16)
            o = dict(resid=resid0.clone(), out=z(), out_lo=z(), ssq_out=torch.zeros(n_out // 128, T, device=DEV))
            extra = dict(sync=sync, scratch=scratch, sync_mode=1 + par, sync_slots=148) if mode == "alt" else {}
            K.gemm_fused(w, x, x_lo, N.EPI_RESID, gamma=gamma, ssq_stride=T, **o, **extra)
            torch.cuda.synchronize()
            outs.append(o)
        par ^= 1
        for k in ("resid", "out", "out_lo", "ssq_out"):
            assert torch.equal(outs[0][k], outs[1][k]), (rep, k)


def test_gemm_multiwave_prefill_shapes(K):
    """More tiles than SMs (the prefill regime): the persistent / CTA-pair kernels against torch, 128-token
    tiles with a ragged last one, RESID + SWIGLU + LOGITS epilogues."""
    from fasttalk_llm_microservice_b200 import native as N
    from fasttalk_llm_microservice_b200.weights import interleave_gate_up
    g = torch.Generator(device=DEV).manual_seed(91)
    T, h, f = 1100, 2048, 2048
    x, x_lo = K.split_hi_lo(torch.randn(T, h, device=DEV, generator=g))
    xf = hilo(x, x_lo)
    ssq = torch.rand(h // 128, T, device=DEV, generator=g) * 128
    rstd = torch.rsqrt(ssq.sum(0) / h + 1e-5)
    rs = dict(ssq_in=ssq, ssq_parts=h // 128, ssq_stride=T, inv_h=1.0 / h, eps=1e-5)
    # RESID: 16 row tiles x 9 token tiles
    w = (torch.randn(h, h, device=DEV, generator=g) * 0.03).bfloat16()
    gamma = (1 + 0.1 * torch.randn(h, device=DEV, generator=g)).bfloat16()
    resid = torch.randn(T, h, device=DEV, generator=g)
    ref = resid + xf @ w.float().T
    out, out_lo = torch.zeros(T, h, device=DEV, dtype=torch.bfloat16), torch.zeros(T, h, device=DEV, dtype=torch.bfloat16)
    so = torch.zeros(h // 128, T, device=DEV)
    K.gemm_fused(w, x, x_lo, N.EPI_RESID, n_splits=1, resid=resid, gamma=gamma, out=out, out_lo=out_lo, ssq_out=so,
                 ssq_stride=T)
    torch.cuda.synchronize()
    assert rel_err(resid, ref) < 1e-3 and rel_err(hilo(out, out_lo), ref * gamma.float()) < 1e-3
    assert rel_err(so, (ref ** 2).reshape(T, h // 128, 128).sum(-1).T) < 2e-3
    # SWIGLU: 32 row tiles x 9 token tiles
    wgu = (torch.randn(2 * f, h, device=DEV, generator=g) * 0.03).bfloat16()
    act, act_lo = torch.zeros(T, f, device=DEV, dtype=torch.bfloat16), torch.zeros(T, f, device=DEV, dtype=torch.bfloat16)
    K.gemm_fused(interleave_gate_up(wgu), x, x_lo, N.EPI_SWIGLU, out=act, out_lo=act_lo, **rs)
    assert rel_err(hilo(act, act_lo), O.swiglu((xf @ wgu.float().T) * rstd[:, None])) < 5e-4
    # LOGITS: 40 row tiles x 9 token tiles
    wl = (torch.randn(5120, h, device=DEV, generator=g) * 0.03).bfloat16()
    logits = torch.zeros(T, 5120, device=DEV)
    K.gemm_fused(wl, x, x_lo, N.EPI_LOGITS, out=logits, **rs)
    assert rel_err(logits, (xf @ wl.float().T) * rstd[:, None]) < 1e-3


# ------------------------------------------------------------------ out-of-bounds writes (own guard bands)
class Guarded:
    """A tensor inside a larger allocation whose margins hold a canary pattern.  compute-sanitizer is not available
    on this GPU pool, so writes outside a kernel's buffers are caught this way: every output / work buffer of the
    kernels with hand-rolled cross-CTA protocols gets guard bands that must come back untouched."""
    PAD = 4096

    def __init__(self, numel, dtype, fill=None):
        self.esz = torch.empty((), dtype=dtype).element_size()
        n = numel * self.esz
        self.raw = torch.full((n + 2 * self.PAD,), 0xA5, device=DEV, dtype=torch.uint8)
        self.t = self.raw[self.PAD:self.PAD + n].view(dtype)
        if fill is not None:
            self.t.fill_(fill)
        self.n = n

    def intact(self):
        return bool((self.raw[:self.PAD] == 0xA5).all() and (self.raw[self.PAD + self.n:] == 0xA5).all())


@pytest.mark.parametrize("h,T,Kd", [(2048, 64, 2048), (2048, 64, 8192), (2048, 37, 2048)])
def test_split_k_resid_epilogue_stays_inside_its_buffers(K, h, T, Kd):
    from fasttalk_llm_microservice_b200 import native as N
    g = torch.Generator(device=DEV).manual_seed(h + T + Kd)
    lib = N.load()
    w = (torch.randn(h, Kd, device=DEV, generator=g) * 0.05).bfloat16()
    x, x_lo = K.split_hi_lo(torch.randn(T, Kd, device=DEV, generator=g))
    splits = lib.ft_gemm_pick_splits(h, T, Kd, 1)
    assert splits > 1
    resid0 = torch.randn(T, h, device=DEV, generator=g)
    resid = Guarded(T * h, torch.float32); resid.t.copy_(resid0.reshape(-1))
    out, out_lo = Guarded(T * h, torch.bfloat16, 0), Guarded(T * h, torch.bfloat16, 0)
    ssq = Guarded((h // 128) * T, torch.float32, 0)
    scratch = Guarded(lib.ft_gemm_scratch_bytes(h, T, Kd, splits) // 4, torch.float32, 0)
    sync = Guarded(2 * (h // 128) * ((T + 15) // 16), torch.int32, 0)
    gamma = torch.ones(h, device=DEV, dtype=torch.bfloat16)
    for _ in range(3):                      # re-armed arrival counters: launch after launch on the same buffers
        K.gemm_fused(w, x, x_lo, N.EPI_RESID, n_splits=splits, resid=resid.t, gamma=gamma, out=out.t, out_lo=out_lo.t,
                     ssq_out=ssq.t, ssq_stride=T, scratch=scratch.t, sync=sync.t)
    torch.cuda.synchronize()
    for name, buf in (("resid", resid), ("out", out), ("out_lo", out_lo), ("ssq", ssq), ("scratch", scratch), ("sync", sync)):
        assert buf.intact(), f"{name}: a write landed outside the buffer"
    ref = resid0 + 3 * (hilo(x, x_lo) @ w.float().T)
    assert rel_err(resid.t.view(T, h), ref) < 1e-3
    assert int(sync.t.abs().sum()) == 0      # the counters re-armed themselves


def test_decode_attention_stays_inside_its_buffers(K):
    from fasttalk_llm_microservice_b200 import native as N
    lib = N.load()
    g = torch.Generator(device=DEV).manual_seed(12)
    nq, nkv, d, ctx_lens = 32, 8, 64, [700, 63, 64, 65, 1, 2048]
    kv, pt, ks, vs = build_paged_kv(K, ctx_lens, nkv, d, g)
    B = len(ctx_lens)
    q = torch.randn(B, nq, d, device=DEV, generator=g).half()
    ctx = torch.tensor(ctx_lens, device=DEV, dtype=torch.int32)
    for n_ctas in (lib.ft_attn_decode_default_ctas(d), 3, 1):
        scratch = Guarded(lib.ft_attn_decode_scratch_bytes(B, nkv, nq // nkv, d, n_ctas) // 4, torch.float32, 0)
        tp = Guarded(B + 1, torch.int32, 0)
        out, out_lo = Guarded(B * nq * d, torch.bfloat16, 0), Guarded(B * nq * d, torch.bfloat16, 0)
        for _ in range(2):
            N.check(lib.ft_attn_decode(N.ptr(q), N.ptr(kv), N.ptr(pt), pt.shape[1], N.ptr(ctx), N.ptr(tp.t), N.ptr(out.t),
                                       N.ptr(out_lo.t), N.ptr(scratch.t), B, nq, nkv, d, n_ctas, N.cur_stream()), "ft_attn_decode")
        torch.cuda.synchronize()
        for name, buf in (("scratch", scratch), ("tile_prefix", tp), ("out", out), ("out_lo", out_lo)):
            assert buf.intact(), f"{name} (n_ctas={n_ctas}): a write landed outside the buffer"
        got = hilo(out.t, out_lo.t).view(B, nq * d)
        for b in range(B):
            ref = O.attention(q[b:b + 1].float(), ks[b].float(), vs[b].float(), ctx_lens[b] - 1)
            assert rel_err(got[b:b + 1], ref) < 2e-3
