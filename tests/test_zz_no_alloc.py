"""GPU: bfCudaGmresProjectStatic — the allocation-free variant of bfCudaGmresProject (static device scratch,
no cudaMalloc on the call path; no counterpart in the reference).

Called through the C ABI and compared with the existing entry point (caller-provided scratch) on the same
inputs, on two calls in a row (the shared array keeps the first call's contents: the second call — fewer basis
vectors, other data — must not see them).  The two entry points launch the same kernels with the same fixed-order
sums, so the results are compared to the tolerance the Gram-Schmidt tests use AND found bit-identical.

Collected last on purpose (file name).  Passes alone in a fresh process and after
the other tests (the scratch address is looked up lazily, per device, on first use)."""
import ctypes as C

import numpy as np
import pytest

from butterfly_b200 import capi

pytestmark = pytest.mark.gpu
MAX_M = 62  # BF_CUDA_GMRES_PROJECT_STATIC_MAX_M (include/bf_cuda.h)


def test_static_scratch_project_matches_the_scratch_buffer_entry_point():
    import torch

    lib = capi.load_library()
    rng = np.random.default_rng(31)
    n = 20011
    Vh = rng.standard_normal((MAX_M, n)) + 1j * rng.standard_normal((MAX_M, n))
    V = torch.from_numpy(Vh).cuda()
    scratch = torch.empty(int(lib.bfCudaGmresScratchBytes(MAX_M)), dtype=torch.uint8, device="cuda")
    p = lambda t: C.c_void_p(t.data_ptr())
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream or 1)
    # everything the calls touch exists before the first call: the new entry point must not need an allocation
    ws = [torch.from_numpy(rng.standard_normal(n) + 1j * rng.standard_normal(n)).cuda() for _ in range(3)]
    want = torch.zeros(MAX_M + 1, dtype=torch.complex128, device="cuda")
    got = torch.full((MAX_M + 1,), complex(np.nan, np.nan), dtype=torch.complex128, device="cuda")
    torch.cuda.synchronize()
    before = torch.cuda.memory_allocated()
    # two calls in a row (then a third): most vectors first, so that stale partials of a larger call are there to be misread
    for call, (m, w) in enumerate(zip((MAX_M, 9, 0), ws)):
        want.zero_()
        got.fill_(complex(np.nan, np.nan))
        capi.check(lib.bfCudaGmresProject(p(V), n, m, p(w), n, p(want), p(scratch), st), "bfCudaGmresProject")
        capi.check(lib.bfCudaGmresProjectStatic(p(V), n, m, p(w), n, p(got), st), "bfCudaGmresProjectStatic")
        torch.cuda.synchronize()
        a, b = got.cpu().numpy(), want.cpu().numpy()
        assert np.allclose(a[: m + 1], b[: m + 1], rtol=1e-13, atol=1e-11), f"call {call}: m = {m}"
        assert np.array_equal(a[: m + 1], b[: m + 1]), f"call {call}: same kernels, same order of sums"
        assert np.isnan(a[m + 1:]).all(), "wrote past the m + 1 results"
        wh = w.cpu().numpy()
        assert np.allclose(a[:m], np.conj(Vh[:m]) @ wh, rtol=1e-13, atol=1e-11)
        assert abs(a[m].real - np.vdot(wh, wh).real) <= 1e-12 * n and a[m].imag == 0
    assert torch.cuda.memory_allocated() == before


def test_static_scratch_project_rejects_sizes_beyond_its_maximum():
    import torch

    lib = capi.load_library()
    n = 64
    V = torch.zeros((MAX_M + 1, n), dtype=torch.complex128, device="cuda")
    w = torch.ones(n, dtype=torch.complex128, device="cuda")
    out = torch.zeros(MAX_M + 2, dtype=torch.complex128, device="cuda")
    p = lambda t: C.c_void_p(t.data_ptr())
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream or 1)
    with pytest.raises(capi.BfCudaError) as e:
        capi.check(lib.bfCudaGmresProjectStatic(p(V), n, MAX_M + 1, p(w), n, p(out), st), "bfCudaGmresProjectStatic")
    assert e.value.status == 1 and "maximum 62" in str(e.value)  # BF_CUDA_ERROR_INVALID_ARGUMENT, clear message
    torch.cuda.synchronize()
    assert not out.abs().any(), "a refused call must not launch anything"
    # and the array is still usable afterwards
    capi.check(lib.bfCudaGmresProjectStatic(p(V), n, 1, p(w), n, p(out), st), "bfCudaGmresProjectStatic")
    torch.cuda.synchronize()
    assert out[1].real.item() == n
