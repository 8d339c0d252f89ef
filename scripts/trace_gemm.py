"""Print the in-kernel timeline (ft_debug_gemm_trace) of the decode-shape GEMMs."""
import os, sys
import numpy as np
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from fasttalk_llm_microservice_b200 import build, kernels as K, native as N
build.build()
lib = N.load()
dev = "cuda"
names = ["entry", "setup", "W req", "dep ok", "stage0", "mma iss", "acc rdy", "epi st", "reduced", "exit"]
def run(name, n_out, Kd, T, hilo, splits, mode=0):
    ws = [(torch.randn(n_out, Kd, device=dev) * 0.02).bfloat16() for _ in range(max(2, int(300e6 // (n_out * Kd * 2)) + 1))]
    x, x_lo = K.split_hi_lo(torch.randn(T, Kd, device=dev))
    s = splits or lib.ft_gemm_pick_splits(n_out, T, Kd, int(hilo))
    out = torch.empty(s * T * n_out, device=dev)
    out2 = torch.empty(T * n_out, device=dev)
    tr = np.zeros(16, dtype=np.uint64)
    rows = []
    for i in range(6):
        N.check(lib.ft_gemm_bf16(N.ptr(ws[i % len(ws)]), N.ptr(x), N.ptr(x_lo if hilo else None), N.ptr(out), N.ptr(out2),
                                 n_out, T, Kd, s, mode, N.cur_stream()), "gemm")
        lib.ft_debug_gemm_trace(tr.ctypes.data)
        rows.append((tr[:10].astype(np.int64) - int(tr[0])).tolist())
    med = np.median(np.array(rows[2:]), axis=0)
    print(f"{name:8s} hilo={int(hilo)} S={s:2d} mode={mode}: " + "  ".join(f"{n}={int(v)}" for n, v in zip(names, med)), flush=True)
for hilo in (True, False):
    run("o", 2048, 2048, 64, hilo, 0)
    run("qkv", 3072, 2048, 64, hilo, 0)
    run("down", 2048, 8192, 64, hilo, 0)
    run("gate_up", 16384, 2048, 64, hilo, 1)
run("gate_up", 16384, 2048, 64, True, 1, mode=2)
run("o_T1", 2048, 2048, 1, True, 0)
