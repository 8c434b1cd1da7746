"""CPU-side checks: the C-ABI library builds/loads, exports every symbol the header declares,
the ops refuse CPU tensors (no fallback), and the chain arithmetic is exact on the host."""
import os
import re
import subprocess
import sys

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def built():
    from mercury_b200 import _build

    return _build.build_library()


def test_library_exports_every_declared_symbol(built):
    from mercury_b200 import _lib

    hdr = open(os.path.join(ROOT, "include", "mercury_b200.h")).read()
    declared = set(re.findall(r"\b(mercury_[a-z0-9_]+)\s*\(", hdr))
    declared -= {"mercury_update_args"}
    assert declared, "no declarations found"
    lib = _lib.load()
    for name in sorted(declared):
        assert hasattr(lib, name), f"{name} is declared in include/mercury_b200.h but not exported"
        assert name in _lib.SIGNATURES, f"{name} has no ctypes signature"
    assert set(_lib.SIGNATURES) == declared
    assert lib.mercury_abi_version() == _lib.ABI_VERSION
    # size queries are pure host functions and may be called without a GPU
    assert lib.mercury_cdf_workspace_bytes(1 << 20) > 0
    assert lib.mercury_update_samples_workspace_bytes(1 << 20, 128) % 256 == 0
    assert lib.mercury_sample_topk_workspace_bytes(1 << 20, 256) >= (1 << 22)


def test_sass_contains_bulk_copies(built):
    """The staged K1 kernels really use the TMA bulk-copy path (SASS: UBLKCP)."""
    out = subprocess.run(["cuobjdump", "-sass", built], capture_output=True, text=True)
    if out.returncode != 0:
        pytest.skip("cuobjdump unavailable")
    assert "UBLKCP" in out.stdout
    assert "sm_100a" in out.stdout or "SM100a" in out.stdout.upper() or "sm_100" in out.stdout


def test_ops_refuse_cpu_tensors(built):
    from mercury_b200 import ops

    with pytest.raises(RuntimeError, match="CUDA tensor"):
        ops.score(torch.randn(4, 10), torch.zeros(4, dtype=torch.int64))
    with pytest.raises(RuntimeError, match="CUDA tensor"):
        ops.multinomial(torch.ones(4), 2, True)
    with pytest.raises(RuntimeError, match="CUDA tensor"):
        ops.cdf(torch.ones(4))


def test_missing_library_fails_loudly(tmp_path, monkeypatch):
    from mercury_b200 import _lib

    monkeypatch.setattr(_lib, "_lib", None)
    monkeypatch.setattr(_lib, "SO_PATH", str(tmp_path / "nope.so"))
    with pytest.raises(ImportError, match="no CPU or PyTorch fallback"):
        _lib.load()


def test_argument_errors_are_reported(built):
    from mercury_b200 import _lib

    lib = _lib.load()
    rc = lib.mercury_cdf(None, (1 << 24) + 1, 0, None, None, None, None, 0, None)
    assert rc == _lib.ERR_ARG and "2^24" in _lib.last_error()
    rc = lib.mercury_score(None, 7, None, 4, 10, 10, -100, None, None, None)
    assert rc == _lib.ERR_ARG and "dtype" in _lib.last_error()
    rc = lib.mercury_sample_topk(None, 10, None, 11, None, None, None, 0, None)
    assert rc == _lib.ERR_ARG and "without replacement" in _lib.last_error()


def test_chain_arithmetic_on_host():
    """chain.cuh (shared with the kernels) reproduces the serial fp32 sum bit for bit."""
    exe = os.path.join(ROOT, "build", "chain_host_test")
    os.makedirs(os.path.dirname(exe), exist_ok=True)
    subprocess.check_call(["g++", "-O2", "-ffp-contract=off", "-std=c++17", "-I", os.path.join(ROOT, "mercury_b200", "csrc"),
                           os.path.join(ROOT, "tests", "host", "chain_host_test.cpp"), "-o", exe])
    out = subprocess.run([exe, "300000"], capture_output=True, text=True)
    assert out.returncode == 0, out.stdout[-2000:]
    assert "ALL OK" in out.stdout


def test_header_is_plain_c(built):
    """include/mercury_b200.h is the C-ABI contract: it must compile as C99 (what a cgo / JNI / FFI binding would include)
    and link against the built library from a C program that makes no compute call."""
    from mercury_b200 import _lib

    src = os.path.join(ROOT, "build", "abi_probe.c")
    os.makedirs(os.path.dirname(src), exist_ok=True)
    with open(src, "w") as f:
        f.write('#include <stdio.h>\n#include "mercury_b200.h"\n'
                'int main(void) {\n'
                '    mercury_update_args a = {0};\n'
                '    (void)a;\n'
                '    printf("%d %zu\\n", mercury_abi_version(), mercury_update_samples_workspace_bytes(1000000, 128));\n'
                '    return mercury_cdf(NULL, -1, 0, NULL, NULL, NULL, NULL, 0, NULL) == MERCURY_ERR_ARG ? 0 : 1;\n'
                '}\n')
    subprocess.check_call(["gcc", "-std=c99", "-Wall", "-Wextra", "-pedantic", "-Werror", "-fsyntax-only", "-I", os.path.join(ROOT, "include"), src])
    exe = os.path.join(ROOT, "build", "abi_probe")
    libdir = os.path.join(ROOT, "mercury_b200")
    subprocess.check_call(["gcc", "-std=c99", "-I", os.path.join(ROOT, "include"), src, "-o", exe, "-L", libdir, "-lmercury_b200",
                           "-Wl,-rpath," + libdir])
    out = subprocess.run([exe], capture_output=True, text=True)
    assert out.returncode == 0, out.stderr
    ver, ws = out.stdout.split()
    assert int(ver) == _lib.ABI_VERSION and int(ws) > 0
