"""mercury_cdf_static: the CDF entry point for hosts that cannot allocate device memory.

Some users of this library have no allocator to call: their process runs inside a runtime that
took the whole GPU for its own pool, or they bind the C ABI from a small host program and do
not want to own a workspace.  For them the library carries its scratch as a built-in device
array.  This test drives that entry point straight through the C ABI, twice in a row with
different pools (the second call finds the first call's leftovers in the shared array), and
requires the same bits as `mercury_cdf` with a caller-provided workspace -- the tolerance the
other exact-CDF tests use is bit equality.  The file sorts last on purpose: if it fails under
`-x`, everything else has already run.  It must pass alone in a fresh process as well.
"""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

DEV = "cuda:0"


def _weights(n, seed):
    g = torch.Generator().manual_seed(seed)
    w = -torch.log(torch.rand(n, generator=g) + 1e-12) + 1.15
    return (w / w.sum()).to(DEV)


def _with_workspace(lib, w, mode, stream):
    n = w.numel()
    ws = torch.empty(lib.mercury_cdf_workspace_bytes(n), dtype=torch.uint8, device=DEV)
    cdf = torch.empty_like(w)
    total = torch.empty(1, device=DEV)
    status = torch.zeros(1, dtype=torch.int32, device=DEV)
    rc = lib.mercury_cdf(w.data_ptr(), n, mode, cdf.data_ptr(), total.data_ptr(), status.data_ptr(), ws.data_ptr(),
                         ws.numel(), stream)
    assert rc == 0, lib.mercury_last_error()
    return cdf, total, status


def _static(lib, w, mode, stream):
    n = w.numel()
    cdf = torch.empty_like(w)  # the caller still owns inputs and outputs
    total = torch.empty(1, device=DEV)
    status = torch.zeros(1, dtype=torch.int32, device=DEV)
    rc = lib.mercury_cdf_static(w.data_ptr(), n, mode, cdf.data_ptr(), total.data_ptr(), status.data_ptr(), stream)
    assert rc == 0, lib.mercury_last_error()
    return cdf, total, status


@pytest.mark.parametrize("mode", [0, 1])  # exact serial-fp32, fp64
def test_static_scratch_matches_workspace_version_twice_in_a_row(mode):
    assert torch.cuda.is_available()
    from mercury_b200 import _lib

    lib = _lib.load()
    stream = torch.cuda.current_stream().cuda_stream
    # large pool first, then a small one: the second call sees stale boundaries of the first
    for n, seed in [(1_000_003, 1), (5_000, 2), (_lib.STATIC_MAX_N, 3)]:
        w = _weights(n, seed)
        want, want_total, _ = _with_workspace(lib, w, mode, stream)
        got, got_total, status = _static(lib, w, mode, stream)
        torch.cuda.synchronize()
        assert int(status.item()) == 0
        assert np.array_equal(got.cpu().numpy().view(np.uint32), want.cpu().numpy().view(np.uint32))
        assert np.array_equal(got_total.cpu().numpy().view(np.uint32), want_total.cpu().numpy().view(np.uint32))
        if mode == 0:  # and both are the serial fp32 chain itself
            chain = np.cumsum(w.cpu().numpy(), dtype=np.float32)
            assert np.array_equal(got.cpu().numpy().view(np.uint32), (chain / chain[-1]).astype(np.float32).view(np.uint32))


def test_static_scratch_rejects_pools_beyond_its_limit():
    from mercury_b200 import _lib

    lib = _lib.load()
    n = _lib.STATIC_MAX_N + 1
    w = torch.ones(n, device=DEV)
    out = torch.empty_like(w)
    total = torch.empty(1, device=DEV)
    status = torch.zeros(1, dtype=torch.int32, device=DEV)
    rc = lib.mercury_cdf_static(w.data_ptr(), n, 0, out.data_ptr(), total.data_ptr(), status.data_ptr(),
                                torch.cuda.current_stream().cuda_stream)
    assert rc == _lib.ERR_ARG
    msg = _lib.last_error()
    assert "exceeds the built-in scratch" in msg and str(_lib.STATIC_MAX_N) in msg
