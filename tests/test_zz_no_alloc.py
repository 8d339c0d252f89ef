"""ft_attn_decode_noalloc — decode attention for hosts that cannot allocate device memory while
they serve (memory carved up once at start-up, or a real-time audio thread that must not hit
cudaMalloc's device-wide synchronisation).  It must give what ft_attn_decode gives with
caller-provided scratch, call after call, even though all calls share one built-in work area
that still holds the previous call's leftovers.

This file is named to be collected last, so that a run that stops at the first failure has
already run everything else; it passes on its own in a fresh process as well.
"""
import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = "cuda"
TOL = 2e-3          # the decode-attention tolerance of tests/test_kernels_gpu.py


@pytest.fixture(scope="module")
def K():
    from fasttalk_llm_microservice_b200 import build, kernels
    build.build()
    return kernels


def rel_err(a, b):
    return (a.float() - b.float()).abs().max().item() / max(b.float().abs().max().item(), 1e-12)


def make_case(K, B, nq, nkv, d, ctx_lens, seed):
    g = torch.Generator(device=DEV).manual_seed(seed)
    n_pages = [-(-c // 64) for c in ctx_lens]
    total = sum(n_pages) + 3
    kv = (torch.randn(total, nkv, 2, 64, d, device=DEV, generator=g)).half()
    perm = torch.randperm(total, device=DEV, generator=g).int()
    pt = torch.zeros(B, max(n_pages) + 1, dtype=torch.int32, device=DEV)
    off = 0
    for b, n in enumerate(n_pages):
        pt[b, :n] = perm[off:off + n]
        off += n
    q = torch.randn(B, nq, d, device=DEV, generator=g).half()
    ctx = torch.tensor(ctx_lens, device=DEV, dtype=torch.int32)
    return q, kv, pt, ctx


def test_noalloc_matches_scratch_variant_twice_in_a_row(K):
    # two different batches back to back: the second call finds the first one's prefix table,
    # counters and partial records in the shared work area and must not be affected by them
    cases = [
        make_case(K, 64, 32, 8, 64, [2048 - 7 * i for i in range(64)], seed=1),
        make_case(K, 5, 32, 8, 64, [63, 64, 65, 1, 700], seed=2),
    ]
    for n_ctas in (0, 3):                        # default persistent grid, and a tiny one (many merges)
        for q, kv, pt, ctx in cases:
            want_hi, want_lo = K.attn_decode(q, kv, pt, ctx, n_ctas=n_ctas)
            got_hi = torch.empty_like(want_hi)   # outputs are the caller's: allocated before the call
            got_lo = torch.empty_like(want_lo)
            torch.cuda.synchronize()
            K.attn_decode_noalloc(q, kv, pt, ctx, got_hi, got_lo, n_ctas=n_ctas)
            torch.cuda.synchronize()
            want = want_hi.float() + want_lo.float()
            got = got_hi.float() + got_lo.float()
            assert rel_err(got, want) < TOL, rel_err(got, want)
            assert torch.isfinite(got).all()


def test_noalloc_d128_and_repeat_same_inputs(K):
    q, kv, pt, ctx = make_case(K, 3, 24, 8, 128, [130, 5, 257], seed=3)
    want_hi, want_lo = K.attn_decode(q, kv, pt, ctx)
    want = want_hi.float() + want_lo.float()
    for _ in range(2):
        hi, lo = torch.empty_like(want_hi), torch.empty_like(want_lo)
        K.attn_decode_noalloc(q, kv, pt, ctx, hi, lo)
        torch.cuda.synchronize()
        assert rel_err(hi.float() + lo.float(), want) < TOL


def test_noalloc_refuses_oversized_requests(K):
    from fasttalk_llm_microservice_b200.native import NativeError
    q, kv, pt, ctx = make_case(K, 2, 8, 2, 64, [10, 20], seed=4)
    hi = torch.empty(2, 8 * 64, device=DEV, dtype=torch.bfloat16)
    lo = torch.empty_like(hi)
    with pytest.raises(NativeError, match="exceed the built-in work area"):
        K.attn_decode_noalloc(q, kv, pt, ctx, hi, lo, n_ctas=513)
    # B * nkv = 2048 > 1024 (shape check happens before anything is launched)
    big_q = torch.empty(256, 32, 64, device=DEV, dtype=torch.float16)
    big_ctx = torch.ones(256, device=DEV, dtype=torch.int32)
    big_pt = torch.zeros(256, 1, device=DEV, dtype=torch.int32)
    kv8 = torch.zeros(1, 8, 2, 64, 64, device=DEV, dtype=torch.float16)
    out = torch.empty(256, 32 * 64, device=DEV, dtype=torch.bfloat16)
    with pytest.raises(NativeError, match="exceed the built-in work area"):
        K.attn_decode_noalloc(big_q, kv8, big_pt, big_ctx, out, torch.empty_like(out))
