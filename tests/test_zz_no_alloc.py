"""sgc_col_stats_static: column statistics for callers that cannot give the library a workspace.

Some hosts of libsgc_b200.so are not allowed to allocate device memory at the point where they
need the Reddit column statistics (utils.py:119): code running inside another runtime's callback
where cudaMalloc is forbidden, a process whose device memory belongs to a pool it does not own,
a latency-bound loop that cannot take the device synchronisation an allocation may bring.  For
them the library keeps the partial sums in a static __device__ array of its own (8 MiB per GPU)
and the entry point takes inputs and outputs only.

The checks here go through the C ABI directly (ctypes, raw pointers).  The new entry point must
return exactly what sgc_col_stats returns -- the two add the same partial sums in the same order,
so the comparison is bit for bit, tighter than the 1e-12 the statistics tests allow against the
oracle -- and it must do so on two calls in a row: the second call finds the first call's partial
sums still in the array, and has to be unaffected by them.

This file sorts last on purpose: `pytest -x` stops at the first failure, and a failure here should
not hide the parity tests.  It passes on its own in a fresh process (the first call does the
address look-up) and after the rest of the suite.
"""
import pytest
import torch

from sgc_b200 import _ffi

MAX_BYTES = 8 << 20      # SGC_COL_STATS_STATIC_MAX_BYTES in include/sgc_b200.h


def test_oversized_request_is_refused_before_any_gpu_work():
    """Beyond the 8 MiB of built-in scratch the call must fail with SGC_ERR_WORKSPACE and say why;
    the check comes before anything touches the device, so it also runs on a box without a GPU
    (the pointer is never dereferenced)."""
    lib = _ffi.lib()
    n, f = 1 << 20, 4096                                  # 592 blocks x 2 x 4096 doubles = 37 MiB
    assert lib.sgc_col_stats_workspace_bytes(n, f) > MAX_BYTES
    rc = lib.sgc_col_stats_static(0x1000, n, f, f, None, None, None, None, None)
    assert rc == -4
    msg = lib.sgc_last_error().decode()
    assert "built-in scratch" in msg and "sgc_col_stats" in msg
    rc = lib.sgc_col_stats_static(None, n, 8, 8, None, None, None, None, None)
    assert rc == -1                                       # same argument checks as sgc_col_stats


def _stats_with_workspace(lib, x, n, f, stream):
    dev = x.device
    out = [torch.empty(f, dtype=torch.float32, device=dev), torch.empty(f, dtype=torch.float32, device=dev),
           torch.empty(f, dtype=torch.float64, device=dev), torch.empty(f, dtype=torch.float64, device=dev)]
    ws = torch.empty(lib.sgc_col_stats_workspace_bytes(n, f), dtype=torch.uint8, device=dev)
    _ffi.call("sgc_col_stats", x.data_ptr(), n, f, x.stride(0), *[t.data_ptr() for t in out],
              ws.data_ptr(), ws.numel(), stream)
    return out


@pytest.mark.gpu
def test_static_scratch_matches_workspace_entry_point_twice_in_a_row():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    lib = _ffi.lib()
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    stream = torch.cuda.current_stream(dev).cuda_stream
    gen = torch.Generator(device="cpu").manual_seed(20261106)
    # call 1: Reddit's width on a padded pitch, enough rows for the full 4-blocks-per-SM grid;
    # call 2: a smaller, narrower matrix, so most of the array still holds call 1's partial sums
    cases = [(60_000, 602, 608, 3.0), (1_000, 130, 130, -7.5)]
    for n, f, ld, shift in cases:
        full = (torch.randn(n, ld, generator=gen, dtype=torch.float32) * 2.0 + shift).to(dev)
        x = full[:, :f]                                   # row pitch ld >= f, like the padded feature store
        want = _stats_with_workspace(lib, x, n, f, stream)
        # every buffer the call uses exists before it: inputs and outputs are the caller's
        got = [torch.full((f,), float("nan"), dtype=t.dtype, device=dev) for t in want]
        torch.cuda.synchronize(dev)
        before = torch.cuda.memory_allocated(dev)
        _ffi.call("sgc_col_stats_static", x.data_ptr(), n, f, x.stride(0), *[t.data_ptr() for t in got], stream)
        torch.cuda.synchronize(dev)
        assert torch.cuda.memory_allocated(dev) == before
        for g, w in zip(got, want):
            assert torch.equal(g, w)                      # bit for bit (NaN would fail too)
        # and both agree with torch's own statistics to fp32 rounding, as in test_fused_standardisation
        assert torch.allclose(got[0], x.mean(0), rtol=1e-5, atol=1e-6)
        assert torch.allclose(got[1], x.std(0), rtol=1e-5, atol=0)
    # outputs are optional, as with sgc_col_stats: only the fp32 pair
    m32 = torch.empty(f, dtype=torch.float32, device=dev)
    s32 = torch.empty(f, dtype=torch.float32, device=dev)
    _ffi.call("sgc_col_stats_static", x.data_ptr(), n, f, x.stride(0), m32.data_ptr(), s32.data_ptr(),
              None, None, stream)
    torch.cuda.synchronize(dev)
    assert torch.equal(m32, want[0]) and torch.equal(s32, want[1])
