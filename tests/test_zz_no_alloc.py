"""gnnb_colsum_static: the column-sum entry point for callers that are not allowed to allocate.

Some users of the library compute the GCN bias gradient from inside a step that runs under a fixed memory plan, or from
a process that must never call into an allocator while its streams are busy; they can hand over input and output
buffers but no workspace.  For them the library carries its own 2 MiB scratch array.  This file checks, through the C ABI
and on two calls in a row, that the allocation-free entry point returns what the workspace-taking one returns, that the
array's left-over contents from the first call do not leak into the second, and that oversized requests are refused with
a clear message.  It is collected last (file name), so a run that stops at the first failure has already run the rest.
"""
import ctypes

import pytest
import torch

from gnn_b200 import _lib

pytestmark = pytest.mark.gpu


def _p(t):
    return ctypes.c_void_p(t.data_ptr())


def _stream():
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def colsum_with_workspace(A):
    n, F = A.shape
    out = torch.empty(F, dtype=torch.float32, device=A.device)
    ws = torch.empty(_lib.workspace_bytes("gnnb_colsum_workspace_bytes", F), dtype=torch.uint8, device=A.device)
    _lib.call("gnnb_colsum", _p(A), A.stride(0), n, F, _p(out), _p(ws), ws.numel(), _stream())
    return out


def colsum_static(A, out):
    n, F = A.shape
    _lib.call("gnnb_colsum_static", _p(A), A.stride(0), n, F, _p(out), _stream())
    return out


def test_static_scratch_matches_workspace_version_twice_in_a_row():
    torch.manual_seed(0)
    first = torch.randn(70_000, 100, device="cuda") * 3.0            # fills all 1024 partial rows of the scratch
    second = torch.randn(777, 37, device="cuda") - 5.0                # smaller: reads far fewer partials than the first wrote
    out1 = torch.empty(100, dtype=torch.float32, device="cuda")
    out2 = torch.empty(37, dtype=torch.float32, device="cuda")
    before = torch.cuda.memory_allocated()
    colsum_static(first, out1)                                         # first call on this device also resolves the symbol
    colsum_static(second, out2)
    torch.cuda.synchronize()
    assert torch.cuda.memory_allocated() == before                    # nothing was taken from torch's allocator either
    want1, want2 = colsum_with_workspace(first), colsum_with_workspace(second)
    torch.cuda.synchronize()
    # same kernels, same fixed summation order: the tolerance of the other fp32 tests (1e-4 relative) is met exactly
    torch.testing.assert_close(out1, want1, rtol=1e-4, atol=1e-4)
    torch.testing.assert_close(out2, want2, rtol=1e-4, atol=1e-4)
    assert torch.equal(out1, want1) and torch.equal(out2, want2)
    torch.testing.assert_close(out2.double(), second.double().sum(0), rtol=1e-4, atol=1e-3)
    # and once more with the widest shape the scratch admits, a strided input and an empty one
    wide = torch.randn(300, 600, device="cuda")[:, :512]
    out3 = torch.empty(512, dtype=torch.float32, device="cuda")
    colsum_static(wide, out3)
    assert torch.equal(out3, colsum_with_workspace(wide))
    # an empty input right after a call that filled every partial row of the same width: the sums must be zero, i.e. no
    # partial the first call left in the shared array is read back
    colsum_static(first, out1)
    empty = torch.empty((0, 100), device="cuda")
    out4 = torch.full((100,), 7.0, device="cuda")
    colsum_static(empty, out4)
    torch.cuda.synchronize()
    assert torch.count_nonzero(out4) == 0


def test_static_scratch_refuses_oversized_requests():
    A = torch.zeros((4, 513), device="cuda")
    out = torch.empty(513, device="cuda")
    with pytest.raises(_lib.GnnbError, match="exceeds the built-in scratch"):
        colsum_static(A, out)
    with pytest.raises(_lib.GnnbError, match="bad argument"):
        _lib.call("gnnb_colsum_static", None, 8, 4, 8, _p(out), _stream())          # rows but no input pointer
