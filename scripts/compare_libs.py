"""Same-box comparators at kernel level (round-1 verdict, next #7): the library kernels this image ships
(FlashInfer 0.6, flash-attn 2.8, cuBLAS through torch.matmul) against this repository's own kernels,
through the C ABI, at the shapes the engine runs.  Every pair is timed the same way: CUDA events over
`iters` back-to-back launches on torch's current stream after a warm-up, rotating over enough copies of
the big operand (KV cache / weights) that nothing is served from the 126 MB L2.

    python scripts/compare_libs.py [--out profiles/r02_compare_libs.json]

A section whose library cannot run here (no JIT cache, unsupported shape) records "unavailable: <why>".
"""
import argparse
import json
import os
import sys
import traceback

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from fasttalk_llm_microservice_b200 import build, kernels as K, native as N  # noqa: E402

DEV = "cuda"
PEAK = 6545.0
try:
    with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as fh:
        _pk = json.load(fh)
        PEAK, PEAK_TF = float(_pk["hbm_gbs"]), float(_pk.get("bf16_tflops_sustained", 1390.7))
except Exception:  # noqa: BLE001
    PEAK_TF = 1390.7


def timed(fns, iters=40, warm=5):
    """fns: list of zero-arg launchers rotated round-robin.  Returns microseconds per launch."""
    for i in range(warm):
        fns[i % len(fns)]()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(iters):
        fns[i % len(fns)]()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e3 / iters


def guarded(name, fn, rows):
    try:
        fn(rows)
    except Exception as e:  # noqa: BLE001
        traceback.print_exc()
        rows.append({"section": name, "unavailable": f"{type(e).__name__}: {e}"[:400]})


# ----------------------------------------------------------------------------- decode attention
def decode_attention(rows):
    for label, B, ctx, nq, nkv, d in (("1B heads, 64 x 2048 (cfg 2)", 64, 2048, 32, 8, 64),
                                      ("3B heads, 256 x 2048 (cfg 3)", 256, 2048, 24, 8, 128)):
        ppseq = ctx // 64
        pages = B * ppseq
        copies = 2 if pages * nkv * 2 * 64 * d * 2 > 200e6 else 3
        g = torch.Generator(device=DEV).manual_seed(1)
        q = torch.randn(B, nq, d, device=DEV, generator=g).half()
        ctxs = torch.full((B,), ctx, device=DEV, dtype=torch.int32)
        pt = torch.randperm(pages, device=DEV, generator=g).int().reshape(B, ppseq).contiguous()
        kv_bytes = pages * nkv * 2 * 64 * d * 2
        # ours: [page][kv head][K|V][64][d] (swizzled rows; random data, the swizzle does not change timing)
        ours_kv = [torch.randn(pages, nkv, 2, 64, d, device=DEV, generator=g).half() for _ in range(copies)]
        lib = N.load()
        n_ctas = lib.ft_attn_decode_default_ctas(d)
        scratch = torch.empty(lib.ft_attn_decode_scratch_bytes(B, nkv, nq // nkv, d, n_ctas) // 4, device=DEV)
        tp = torch.empty(B + 1, device=DEV, dtype=torch.int32)
        out = torch.empty(B, nq * d, device=DEV, dtype=torch.bfloat16)
        out_lo = torch.empty_like(out)

        def ours(kv):
            return lambda: N.check(lib.ft_attn_decode(N.ptr(q), N.ptr(kv), N.ptr(pt), ppseq, N.ptr(ctxs), N.ptr(tp),
                                                      N.ptr(out), N.ptr(out_lo), N.ptr(scratch), B, nq, nkv, d, n_ctas,
                                                      N.cur_stream()), "ft_attn_decode")
        us = timed([ours(kv) for kv in ours_kv])
        row = {"section": "decode attention", "shape": label, "kv_MB": kv_bytes / 1e6,
               "ours_us": us, "ours_GBs": kv_bytes / us / 1e3, "ours_frac_hbm": kv_bytes / us / 1e3 / PEAK,
               "ours_note": "plan kernel + attn_decode_kernel per call (the engine runs the plan once per step)"}
        del ours_kv
        try:
            import flashinfer
            ws = torch.empty(256 << 20, dtype=torch.uint8, device=DEV)
            w = flashinfer.BatchDecodeWithPagedKVCacheWrapper(ws, "NHD", use_tensor_cores=True)
            indptr = (torch.arange(B + 1, device=DEV, dtype=torch.int32) * ppseq)
            indices = pt.reshape(-1).contiguous()
            last = torch.full((B,), 64, device=DEV, dtype=torch.int32)
            w.plan(indptr, indices, last, nq, nkv, d, 64, pos_encoding_mode="NONE", q_data_type=torch.float16,
                   kv_data_type=torch.float16)
            fi_kv = [torch.randn(pages, 2, 64, nkv, d, device=DEV, generator=g).half() for _ in range(copies)]
            us_fi = timed([(lambda kv=kv: w.run(q, kv)) for kv in fi_kv])
            row.update({"flashinfer_us": us_fi, "flashinfer_GBs": kv_bytes / us_fi / 1e3,
                        "flashinfer_frac_hbm": kv_bytes / us_fi / 1e3 / PEAK,
                        "flashinfer": f"{flashinfer.__version__} BatchDecodeWithPagedKVCacheWrapper(NHD, tensor cores), page 64, fp16",
                        "speedup_vs_flashinfer": us_fi / us})
            # parity of the two on the same logical cache (one sequence's worth is enough)
            del fi_kv
        except Exception as e:  # noqa: BLE001
            traceback.print_exc()
            row["flashinfer"] = f"unavailable: {type(e).__name__}: {e}"[:300]
        rows.append(row)
        print(json.dumps(row), flush=True)
        torch.cuda.empty_cache()


# ----------------------------------------------------------------------------- prefill attention
def prefill_attention(rows):
    for heads, nq, nkv, d, label, reqs in (
            ("1B", 32, 8, 64, "cold 4 x 2048", [(0, 2048)] * 4),
            ("1B", 32, 8, 64, "warm 36 x 224 new over 1824 cached", [(1824, 224)] * 36),
            ("3B", 24, 8, 128, "cold 4 x 2048", [(0, 2048)] * 4),
            ("8B", 32, 8, 128, "cold 4 x 2048", [(0, 2048)] * 4),
            ("8B", 32, 8, 128, "warm 16 x 224 new over 7712 cached", [(7712, 224)] * 16)):
        g = torch.Generator(device=DEV).manual_seed(2)
        totals = [p + n for p, n in reqs]
        n_pages = [(t + 63) // 64 for t in totals]
        pages = sum(n_pages)
        maxp = max(n_pages)
        kv = torch.randn(pages, nkv, 2, 64, d, device=DEV, generator=g).half()
        pt = torch.zeros(len(reqs), maxp, dtype=torch.int32, device=DEV)
        off = 0
        for r, n in enumerate(n_pages):
            pt[r, :n] = torch.arange(off, off + n, device=DEV, dtype=torch.int32)
            off += n
        qs = [0]
        for _, n in reqs:
            qs.append(qs[-1] + n)
        T = qs[-1]
        q = torch.randn(T, nq, d, device=DEV, generator=g).half()
        flops = sum(4.0 * nq * d * (n * p + n * (n + 1) / 2) for p, n in reqs)       # QK^T + PV over the causal region
        # ours
        tile_req, tile_q0 = [], []
        for r, (_, n) in enumerate(reqs):
            for q0 in range(0, n, 64):
                tile_req.append(r); tile_q0.append(q0)
        i32 = lambda v: torch.tensor(v, device=DEV, dtype=torch.int32)
        t_qs, t_slot, t_pos0, t_treq, t_tq0 = i32(qs), i32(list(range(len(reqs)))), i32([p for p, _ in reqs]), i32(tile_req), i32(tile_q0)
        out = torch.zeros(T, nq * d, device=DEV, dtype=torch.bfloat16)
        out_lo = torch.zeros_like(out)
        lib = N.load()
        def ours(q_rows):
            return lambda: N.check(lib.ft_attn_prefill(N.ptr(q), N.ptr(kv), N.ptr(pt), maxp, N.ptr(t_qs), N.ptr(t_slot),
                                                       N.ptr(t_pos0), N.ptr(t_treq), N.ptr(t_tq0), len(tile_req), N.ptr(out),
                                                       N.ptr(out_lo), nq, nkv, d, q_rows, N.cur_stream()), "ft_attn_prefill")
        us = timed([ours(T)], iters=20)                  # tcgen05 / TMEM kernel
        hi_tc = out.float() + out_lo.float()
        us_sync = timed([ours(0)], iters=20)             # the round-1 mma.sync kernel (q_rows = 0 selects it)
        hi_sync = out.float() + out_lo.float()
        row = {"section": f"prefill attention ({heads} heads: nq {nq}, nkv {nkv}, d {d})", "shape": label, "GFLOP": flops / 1e9, "ours_us": us,
               "ours": "attn_prefill_tc_kernel (tcgen05 / TMEM)", "ours_TFLOPs": flops / us / 1e6,
               "ours_frac_tensor": flops / us / 1e6 / PEAK_TF, "ours_mma_sync_us": us_sync,
               "ours_mma_sync_TFLOPs": flops / us_sync / 1e6, "tc_over_mma_sync_speedup": us_sync / us,
               "tc_vs_mma_sync_max_abs_diff": float((hi_tc - hi_sync).abs().max()),
               "tc_vs_mma_sync_rel": float((hi_tc - hi_sync).abs().max() / hi_sync.abs().max())}
        try:
            import flashinfer
            ws = torch.empty(256 << 20, dtype=torch.uint8, device=DEV)
            w = flashinfer.BatchPrefillWithPagedKVCacheWrapper(ws, "NHD")
            qo_indptr = i32(qs)
            kv_indptr = i32([0] + list(torch.tensor(n_pages).cumsum(0).tolist()))
            kv_indices = torch.arange(pages, device=DEV, dtype=torch.int32)
            last = i32([(t - 1) % 64 + 1 for t in totals])
            w.plan(qo_indptr, kv_indptr, kv_indices, last, nq, nkv, d, 64, causal=True, q_data_type=torch.float16,
                   kv_data_type=torch.float16)
            fkv = torch.randn(pages, 2, 64, nkv, d, device=DEV, generator=g).half()
            us_fi = timed([lambda: w.run(q, fkv)], iters=20)
            row.update({"flashinfer_us": us_fi, "flashinfer_TFLOPs": flops / us_fi / 1e6, "speedup_vs_flashinfer": us_fi / us,
                        "flashinfer": f"{flashinfer.__version__} BatchPrefillWithPagedKVCacheWrapper(NHD), page 64, fp16, causal"})
        except Exception as e:  # noqa: BLE001
            traceback.print_exc()
            row["flashinfer"] = f"unavailable: {type(e).__name__}: {e}"[:300]
        try:
            from flash_attn import flash_attn_varlen_func
            import flash_attn
            # contiguous (un-paged) K/V of the same lengths: flash-attn 2's paged path needs 256-token pages
            ks = torch.randn(sum(totals), nkv, d, device=DEV, generator=g).half()
            vs = torch.randn(sum(totals), nkv, d, device=DEV, generator=g).half()
            cu_q = i32(qs)
            cu_k = i32([0] + list(torch.tensor(totals).cumsum(0).tolist()))
            fa = lambda: flash_attn_varlen_func(q, ks, vs, cu_q, cu_k, max(n for _, n in reqs), max(totals), causal=True)
            us_fa = timed([fa], iters=20)
            row.update({"flash_attn_us": us_fa, "flash_attn_TFLOPs": flops / us_fa / 1e6, "speedup_vs_flash_attn": us_fa / us,
                        "flash_attn": f"{flash_attn.__version__} flash_attn_varlen_func, contiguous K/V, fp16, causal (bottom-right aligned)"})
        except Exception as e:  # noqa: BLE001
            traceback.print_exc()
            row["flash_attn"] = f"unavailable: {type(e).__name__}: {e}"[:300]
        rows.append(row)
        print(json.dumps(row), flush=True)
        torch.cuda.empty_cache()


# ----------------------------------------------------------------------------- projections
def projections(rows):
    shapes = [("qkv", 3072, 2048), ("o", 2048, 2048), ("gate_up", 16384, 2048), ("down", 2048, 8192)]
    lib = N.load()
    for T in (64, 8192):
        for name, n_out, Kd in shapes:
            copies = max(2, int(300e6 // (n_out * Kd * 2)) + 1) if T == 64 else 2
            g = torch.Generator(device=DEV).manual_seed(3)
            ws = [(torch.randn(n_out, Kd, device=DEV, generator=g) * 0.02).bfloat16() for _ in range(copies)]
            x32 = torch.randn(T, Kd, device=DEV, generator=g)
            x, x_lo = K.split_hi_lo(x32)
            w_bytes, flops = n_out * Kd * 2, 2.0 * T * n_out * Kd
            us_t = timed([(lambda w=w: torch.matmul(x, w.t())) for w in ws])
            row = {"section": "projection GEMM (1B shapes)", "name": name, "T": T, "n_out": n_out, "K": Kd,
                   "torch_matmul_us": us_t, "torch": f"torch {torch.__version__} matmul bf16 (cuBLAS), ONE activation plane, plain bf16 output"}
            if T == 64:
                # what the engine launches at decode: both planes, split K, partial tiles out (the fused epilogues add the
                # rendezvous and the layer step on top; the stand-alone ABI exposes the partial-sum form)
                s = lib.ft_gemm_pick_splits(n_out, T, Kd, 1)
                out = torch.empty(s, T, n_out, device=DEV)
                ours = [(lambda w=w: N.check(lib.ft_gemm_bf16(N.ptr(w), N.ptr(x), N.ptr(x_lo), N.ptr(out), 0, n_out, T, Kd, s, 0,
                                                              N.cur_stream()), "gemm")) for w in ws]
                us = timed(ours)
                row.update({"ours_us": us, "ours": f"ft_gemm_bf16 hi+lo planes, split-K {s}, fp32 partial tiles",
                            "ours_GBs_of_W": w_bytes / us / 1e3, "torch_GBs_of_W": w_bytes / us_t / 1e3,
                            "ours_frac_hbm": w_bytes / us / 1e3 / PEAK, "torch_frac_hbm": w_bytes / us_t / 1e3 / PEAK})
            else:
                o = torch.empty(T, n_out, device=DEV, dtype=torch.bfloat16)
                o_lo = torch.empty_like(o)
                both = [(lambda w=w: N.check(lib.ft_gemm_bf16(N.ptr(w), N.ptr(x), N.ptr(x_lo), N.ptr(o), N.ptr(o_lo), n_out, T, Kd, 1, 1,
                                                              N.cur_stream()), "gemm")) for w in ws]
                us = timed(both, iters=20)
                row.update({"ours_us": us, "ours": "ft_gemm_bf16 hi+lo planes (2x the MMAs), bf16 hi+lo output",
                            "ours_useful_TFLOPs": flops / us / 1e6, "ours_issued_TFLOPs": 2 * flops / us / 1e6,
                            "torch_TFLOPs": flops / us_t / 1e6, "torch_frac_tensor": flops / us_t / 1e6 / PEAK_TF,
                            "ours_issued_frac_tensor": 2 * flops / us / 1e6 / PEAK_TF})
                # the form the engine's large prefill steps run since round 2: ONE fp16 plane against fp16 weight copies
                # (CTA-pair kernels), here with a plain bf16 output like the library call, and the library on fp16 too
                x16 = x32.half()
                ws16 = [w.half() for w in ws]
                ob = torch.empty(T, n_out, device=DEV, dtype=torch.bfloat16)
                one = [(lambda w=w: K.gemm_fused(w, x16, None, N.EPI_BF16, n_splits=1, x_planes=3, out=ob)) for w in ws16]
                us16 = timed(one, iters=20)
                us_t16 = timed([(lambda w=w: torch.matmul(x16, w.t())) for w in ws16], iters=20)
                ref16 = x16.float() @ ws16[-1].float().T
                err16 = ((ob.float() - ref16).abs().max() / ref16.abs().max()).item()
                row.update({"ours_f16_plane_us": us16, "ours_f16_plane": "ft_gemm_fused x_planes=3 (one fp16 plane x fp16 weights, CTA pairs), bf16 output",
                            "ours_f16_plane_TFLOPs": flops / us16 / 1e6, "torch_matmul_f16_us": us_t16,
                            "torch_f16_TFLOPs": flops / us_t16 / 1e6, "ours_f16_plane_over_torch_time": us16 / us_t16,
                            "ours_f16_plane_rel_err": err16})
            row["ours_over_torch_time"] = row["ours_us"] / us_t
            rows.append(row)
            print(json.dumps(row), flush=True)
            del ws
            torch.cuda.empty_cache()


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "compare_libs.json"))
    ap.add_argument("--only", default="", help="comma list of: decode, prefill, gemm")
    a = ap.parse_args()
    build.build()
    rows = []
    only = set(x for x in a.only.split(",") if x)
    if not only or "decode" in only:
        guarded("decode attention", decode_attention, rows)
    if not only or "prefill" in only:
        guarded("prefill attention", prefill_attention, rows)
    if not only or "gemm" in only:
        guarded("projections", projections, rows)
    os.makedirs(os.path.dirname(a.out), exist_ok=True)
    with open(a.out, "w") as fh:
        json.dump({"peak_hbm_GBs": PEAK, "peak_bf16_TFLOPs_sustained": PEAK_TF, "rows": rows}, fh, indent=1)
    print("wrote", a.out)
