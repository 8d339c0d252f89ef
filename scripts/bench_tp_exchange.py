"""A/B of the tensor-parallel all-reduce (round-1 verdict, next #9): the row-parallel projection with the exchange
fused into its epilogue (P2P stores over NVLink + flags, csrc/gemm.cu) against the same projection followed by
`ncclAllReduce` (torch.distributed, NCCL) on its output — Llama-3.1-8B's two row-parallel shapes at TP=2.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29553 \
        scripts/bench_tp_exchange.py

Per shape and token count, microseconds per call (CUDA events, 200 calls after warm-up, both ranks in lock step):
  fused     ft_gemm_fused(RESID, tp_*): GEMM + split-K rendezvous + exchange + residual/norm epilogue, ONE kernel
  gemm      the same launch without the exchange (what a collective-based design runs first)
  nccl      ncclAllReduce of the [T, 4096] fp32 partial (what it runs next; a finishing kernel would follow)
Also checks that the fused result equals local + peer partial sums (bit-identical on both ranks).
"""
import ctypes as C
import json
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    rank, local = int(os.environ["RANK"]), int(os.environ.get("LOCAL_RANK", os.environ["RANK"]))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    from fasttalk_llm_microservice_b200 import build, kernels as K, native as N
    if rank == 0:
        build.build()
    dist.barrier()
    lib = N.load()
    rt = C.CDLL("libcudart.so.12")
    rt.cudaMalloc.argtypes = [C.POINTER(C.c_void_p), C.c_size_t]
    rt.cudaMemset.argtypes = [C.c_void_p, C.c_int, C.c_size_t]

    def raw_alloc(nbytes):
        p = C.c_void_p()
        assert rt.cudaMalloc(C.byref(p), nbytes) == 0
        assert rt.cudaMemset(p, 0, nbytes) == 0
        return p

    def swap(ptr):
        """hand `ptr` to the peer, get the peer's as mapped here (cudaIpc over ft_ipc_*)"""
        buf = (C.c_ubyte * 64)()
        N.check(lib.ft_ipc_export(ptr, buf), "ft_ipc_export")
        got = [None, None]
        dist.all_gather_object(got, bytes(buf))
        out = C.c_void_p()
        N.check(lib.ft_ipc_open((C.c_ubyte * 64).from_buffer_copy(got[1 - rank]), C.byref(out)), "ft_ipc_open")
        return out

    H, TMAX = 4096, 64
    xchg = raw_alloc(2 * TMAX * H * 4)
    flags = raw_alloc(16384 * 4)
    peer_xchg, peer_flags = swap(xchg), swap(flags)
    dist.barrier()
    rows = []
    epoch = 0
    for name, Kd in (("o_proj (K = 4096 / 2)", 2048), ("down_proj (K = 14336 / 2)", 7168)):
        g = torch.Generator(device="cuda").manual_seed(5 + rank)
        w = (torch.randn(H, Kd, device="cuda", generator=g) * 0.02).bfloat16()
        for T in (1, 8, 64):
            x, x_lo = K.split_hi_lo(torch.randn(T, Kd, device="cuda", generator=g))
            gamma = torch.ones(H, device="cuda", dtype=torch.bfloat16)
            resid = torch.zeros(T, H, device="cuda")
            out, out_lo = torch.empty(T, H, device="cuda", dtype=torch.bfloat16), torch.empty(T, H, device="cuda", dtype=torch.bfloat16)
            ssq = torch.zeros(H // 128, T, device="cuda")
            splits = lib.ft_gemm_pick_splits(H, T, Kd, 1)
            scratch = torch.empty(max(1, lib.ft_gemm_scratch_bytes(H, T, Kd, splits) // 4), device="cuda")
            sync = torch.zeros(2 * (H // 128) * 4, device="cuda", dtype=torch.int32)
            half = TMAX * H * 4

            def fused():
                nonlocal epoch
                epoch += 1
                off = (epoch & 1) * half
                K.gemm_fused(w, x, x_lo, N.EPI_RESID, n_splits=splits, resid=resid, gamma=gamma, out=out, out_lo=out_lo,
                             ssq_out=ssq, ssq_stride=T, scratch=scratch, sync=sync,
                             tp_peer_xchg=peer_xchg.value + off, tp_xchg=xchg.value + off, tp_peer_flags=peer_flags.value,
                             tp_flags=flags.value, tp_epoch=epoch)

            def plain():
                K.gemm_fused(w, x, x_lo, N.EPI_RESID, n_splits=splits, resid=resid, gamma=gamma, out=out, out_lo=out_lo,
                             ssq_out=ssq, ssq_stride=T, scratch=scratch, sync=sync)

            part = torch.zeros(T, H, device="cuda")

            def nccl():
                dist.all_reduce(part)

            # correctness of the fused exchange: resid == local partial + peer partial, identical on both ranks
            resid.zero_()
            dist.barrier()
            fused()
            torch.cuda.synchronize()
            mine = (x.float() + x_lo.float()) @ w.float().T
            both = mine.clone()
            dist.all_reduce(both)
            err = ((resid - both).abs().max() / both.abs().max()).item()
            chk = [None, None]
            dist.all_gather_object(chk, float(resid.double().sum()))
            assert err < 1e-3 and chk[0] == chk[1], (err, chk)

            def timed(fn, iters=200):
                for _ in range(10):
                    fn()
                torch.cuda.synchronize()
                dist.barrier()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                for _ in range(iters):
                    fn()
                e1.record()
                torch.cuda.synchronize()
                t = torch.tensor([e0.elapsed_time(e1) * 1e3 / iters], device="cuda")
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
                return float(t)
            t_fused, t_plain, t_nccl = timed(fused), timed(plain), timed(nccl)
            rows.append({"shape": name, "T": T, "splits": splits, "fused_us": t_fused, "gemm_only_us": t_plain,
                         "nccl_allreduce_us": t_nccl, "gemm_plus_nccl_us": t_plain + t_nccl,
                         "exchange_bytes_each_way": T * H * 4, "exchange_cost_us": t_fused - t_plain,
                         "fused_speedup_over_gemm_plus_nccl": (t_plain + t_nccl) / t_fused, "rel_err": err})
            if rank == 0:
                print(json.dumps(rows[-1]), flush=True)
    if rank == 0:
        print("TP_EXCHANGE " + json.dumps({"nccl": torch.cuda.nccl.version(), "rows": rows}), flush=True)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
