"""The C-ABI library loads and exports every symbol include/fasttalk_b200.h declares
(no compute calls: this runs without a GPU)."""
import os
import re

import pytest

from fasttalk_llm_microservice_b200 import build, native

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib():
    build.build()
    return native.load()


def declared_symbols():
    src = open(os.path.join(ROOT, "include", "fasttalk_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(ft_[a-z0-9_]+)\s*\(", src)))


def test_header_symbols_exported(lib):
    syms = declared_symbols()
    assert len(syms) >= 20
    for s in syms:
        assert hasattr(lib, s), f"{s} declared in include/fasttalk_b200.h but not exported"
        assert s in native.SIGNATURES, f"{s} has no ctypes signature in native.py"


def test_no_extra_bindings():
    assert set(native.SIGNATURES) - set(declared_symbols()) == set()


def test_last_error_is_a_string(lib):
    assert isinstance(native.last_error(), str)


def test_pick_splits_is_host_only(lib):
    # pure host arithmetic; falls back to 148 SMs when no device is present
    s = lib.ft_gemm_pick_splits(3072, 64, 2048, 1)
    assert 1 <= s <= 16 and 24 * s <= 148          # one wave of CTAs (gemm.cu: gemm_pick_splits)
    assert lib.ft_gemm_pick_splits(128256, 64, 2048, 1) == 1


def test_engine_refuses_without_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from fasttalk_llm_microservice_b200.engine import Engine
    from fasttalk_llm_microservice_b200.model_spec import TINY_D64
    with pytest.raises(native.NativeError):
        Engine(TINY_D64, {})


def test_header_is_plain_c():
    """include/fasttalk_b200.h is the boundary a C / cgo / JNI host binds: it must compile as C99 on its own."""
    import shutil
    import subprocess
    if shutil.which("gcc") is None:
        pytest.skip("no gcc on this box")
    hdr = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "include", "fasttalk_b200.h")
    r = subprocess.run(["gcc", "-fsyntax-only", "-x", "c", "-std=c99", "-Wall", "-Werror", hdr],
                       capture_output=True, text=True)
    assert r.returncode == 0, r.stderr


def test_a_c_host_links_and_runs(tmp_path):
    """gcc + the header + the .so is all a compiled host needs; the ctypes struct mirrors have the C sizes."""
    import ctypes as C
    import shutil
    import subprocess
    if shutil.which("gcc") is None:
        pytest.skip("no gcc on this box")
    build.build()
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    libdir = os.path.dirname(native.LIB_PATH)
    exe = str(tmp_path / "host")
    r = subprocess.run(["gcc", "-std=c99", "-Wall", "-Werror", os.path.join(root, "tests", "c_host", "host.c"),
                        "-I", os.path.join(root, "include"), "-L", libdir, "-lfasttalk_b200",
                        "-Wl,-rpath," + libdir, "-o", exe], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    r = subprocess.run([exe], capture_output=True, text=True, timeout=120)
    assert r.returncode == 0 and "C_HOST_OK" in r.stdout, r.stdout + r.stderr
    sizes = dict(kv.split("=") for kv in r.stdout.split() if kv.startswith(("epilogue=", "config=")))
    assert int(sizes["epilogue"]) == C.sizeof(native.GemmEpilogue)
    assert int(sizes["config"]) == C.sizeof(native.EngineConfig)
