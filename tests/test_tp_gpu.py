"""Two-GPU tensor parallelism: the row-parallel projections' all-reduce runs INSIDE the GEMM
epilogue over NVLink peer memory (csrc/gemm.cu, ft_gemm_epilogue.tp_*).  Needs two GPUs in the
box (`gpurun --gpus 2`); skipped on a single-GPU box."""
import os
import socket
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs")
def test_tp2_matches_the_oracle_and_tp1():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
           "--master-addr", "127.0.0.1", "--master-port", str(port), os.path.join(ROOT, "tests", "tp_worker.py")]
    r = subprocess.run(cmd, cwd=ROOT, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert "TP_OK" in r.stdout, r.stdout[-2000:]
    print(r.stdout[r.stdout.index("TP_OK"):].splitlines()[0])
    from oracle.parity import REPORT
    REPORT.extend(ln[len("TP_PARITY "):] for ln in r.stdout.splitlines() if ln.startswith("TP_PARITY "))
