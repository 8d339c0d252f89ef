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


def test_noalloc_against_the_oracle_and_allocates_nothing(K):
    """Checked directly against the fp32 oracle (not only against the scratch-taking entry point), and the claim
    in its name is measured: the device's free memory (cudaMemGetInfo) is the same after the calls as before,
    and so is torch's own allocator state — for the callers this entry point exists for, a hidden cudaMalloc
    would be a device-wide stall in the middle of a conversation."""
    from oracle import llama_oracle as O
    nq, nkv, d, ctx_lens = 32, 8, 64, [200, 64, 1, 333]
    g = torch.Generator(device=DEV).manual_seed(9)
    n_pages = [-(-c // 64) for c in ctx_lens]
    kv = torch.zeros(sum(n_pages), nkv, 2, 64, d, device=DEV, dtype=torch.float16)
    pt = torch.zeros(len(ctx_lens), max(n_pages), dtype=torch.int32, device=DEV)
    ks, vs, off = [], [], 0
    for b, c in enumerate(ctx_lens):
        k = torch.randn(c, nkv, d, device=DEV, generator=g).half()
        v = torch.randn(c, nkv, d, device=DEV, generator=g).half()
        ks.append(k); vs.append(v)
        for p in range(n_pages[b]):
            pt[b, p] = off + p
            n = min(64, c - p * 64)
            for h in range(nkv):
                slab_k = torch.zeros(64, d, device=DEV, dtype=torch.float16); slab_k[:n] = k[p * 64:p * 64 + n, h]
                slab_v = torch.zeros(64, d, device=DEV, dtype=torch.float16); slab_v[:n] = v[p * 64:p * 64 + n, h]
                kv[off + p, h, 0] = K.kv_slab_store(slab_k)
                kv[off + p, h, 1] = K.kv_slab_store(slab_v)
        off += n_pages[b]
    q = torch.randn(len(ctx_lens), nq, d, device=DEV, generator=g).half()
    ctx = torch.tensor(ctx_lens, device=DEV, dtype=torch.int32)
    hi = torch.empty(len(ctx_lens), nq * d, device=DEV, dtype=torch.bfloat16)
    lo = torch.empty_like(hi)
    K.attn_decode_noalloc(q, kv, pt, ctx, hi, lo)            # first call on this device: symbol-address lookup
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info()
    reserved0, allocated0 = torch.cuda.memory_reserved(), torch.cuda.memory_allocated()
    for _ in range(3):
        K.attn_decode_noalloc(q, kv, pt, ctx, hi, lo)
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info()
    assert free1 == free0, (free0, free1)
    assert torch.cuda.memory_reserved() == reserved0 and torch.cuda.memory_allocated() == allocated0
    got = hi.float() + lo.float()
    for b, c in enumerate(ctx_lens):
        ref = O.attention(q[b:b + 1].float(), ks[b].float(), vs[b].float(), c - 1)
        assert rel_err(got[b:b + 1], ref) < TOL, (b, rel_err(got[b:b + 1], ref))


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
