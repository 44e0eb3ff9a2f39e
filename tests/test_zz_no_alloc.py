"""dgs_l2norm_noalloc: the L2 norm for callers that cannot allocate device memory.

Some users of libdgs.so have nothing but the buffers they were handed: a plugin host binding the library
over cgo / JNI / ctypes owns no device allocator, and a trainer whose caching allocator already holds all
of HBM cannot count on a raw cudaMalloc succeeding.  For them the scratch of the norm's first pass is a
fixed __device__ array inside the library (shared by every call, never cleared, 2^26 elements at most).
This test calls the entry point through the C ABI, twice in a row, and checks both results against
dgs_l2norm() -- the existing entry point with caller-provided scratch -- with the tolerance the other norm
tests use.  The second call is smaller than the first, so the array still holds the first call's partials
beyond the slots the second call writes: they must not leak into the result.

The file name sorts after every other test module so that pytest collects it last; it does not depend on
any other test having run (it builds/loads the library itself through the `lib` fixture).
"""
import ctypes

import numpy as np
import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

from determined_clone_b200 import _lib as L  # noqa: E402

CODE = {torch.float32: L.F32, torch.bfloat16: L.BF16}


def _stream():
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def _ptr(t):
    return ctypes.c_void_p(t.data_ptr())


def _reference_norm(x):
    """The existing entry point: caller allocates the partials."""
    n = x.numel()
    partials = torch.empty(L.calls.dgs_l2norm_grid(n), dtype=torch.float32, device="cuda")
    out = torch.zeros(1, device="cuda")
    L.calls.dgs_l2norm(_ptr(x), CODE[x.dtype], n, _ptr(partials), _ptr(out), _stream())
    return out


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_noalloc_norm_matches_existing_entry_point_twice_in_a_row(lib, dtype):
    gen = torch.Generator(device="cuda").manual_seed(11)
    first = (torch.randn(3_000_003, device="cuda", generator=gen) * 0.3).to(dtype)   # 92 partials
    second = (torch.randn(70_001, device="cuda", generator=gen) * 5.0).to(dtype)     # 3 partials; 89 stale ones stay behind
    for x in (first, second):
        out = torch.full((1,), -1.0, device="cuda")
        before = torch.cuda.memory_allocated()
        L.calls.dgs_l2norm_noalloc(_ptr(x), CODE[dtype], x.numel(), _ptr(out), _stream())
        assert torch.cuda.memory_allocated() == before  # nothing was taken from torch's allocator either
        torch.cuda.synchronize()
        want = _reference_norm(x)
        torch.cuda.synchronize()
        np.testing.assert_allclose(float(out), float(want), rtol=2e-6)
        exact = np.sqrt(np.sum(np.square(x.float().cpu().numpy().astype(np.float64))))
        np.testing.assert_allclose(float(out), exact, rtol=2e-6)


def test_noalloc_norm_edge_sizes_and_limit(lib):
    out = torch.full((1,), -1.0, device="cuda")
    x = torch.ones(8, device="cuda")
    L.calls.dgs_l2norm_noalloc(_ptr(x), L.F32, 0, _ptr(out), _stream())  # empty input: norm 0
    torch.cuda.synchronize()
    assert float(out) == 0.0
    L.calls.dgs_l2norm_noalloc(_ptr(x), L.F32, 8, _ptr(out), _stream())
    torch.cuda.synchronize()
    np.testing.assert_allclose(float(out), np.sqrt(8.0), rtol=2e-6)
    # one element past the limit is refused with a message that names the limit; nothing is launched
    launches = L.calls.dgs_launch_count()
    with pytest.raises(L.DgsError, match="exceeds DGS_NOALLOC_MAX_ELEMS"):
        L.calls.dgs_l2norm_noalloc(_ptr(x), L.F32, L.NOALLOC_MAX_ELEMS + 1, _ptr(out), _stream())
    assert L.calls.dgs_launch_count() == launches
    with pytest.raises(L.DgsError, match="out_norm is NULL"):
        L.calls.dgs_l2norm_noalloc(_ptr(x), L.F32, 8, None, _stream())
