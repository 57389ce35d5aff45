"""GPU: device-resident GMRES (bfCudaSolveGMRES) against the reference's own bfSolveGMRES
(src/linalg.c:47) on the same second-kind system K diag(w) + I/2 built by the reference
(examples/multiple_scattering/multiple_scattering_context.c:456-487) — same iteration count, same
iterate; and bfSolveGMRES itself running on the GPU through the swapped vtable."""
import numpy as np
import pytest

from butterfly_b200 import capi
from butterfly_b200.plan import Plan
from oracle import portlib, reflib

pytestmark = pytest.mark.gpu
needs_ref = pytest.mark.skipif(not reflib.attach_available(), reason="oracle/_ref + attach layer not built")
# the iterate is a linear combination of <= 50 basis vectors with O(1) coefficients and the system
# is well conditioned (second kind): agreement is at the level of the apply tolerance times ~10
X_TOL = 1e-11


@needs_ref
@pytest.mark.parametrize("n,k,tol,max_iter", [(2048, 50.0, 1e-12, 50), (4096, 100.0, 1e-8, 50), (2048, 50.0, 1e-12, 4)])
def test_device_gmres_matches_bfSolveGMRES(n, k, tol, max_iter):
    h = reflib.Helm2(reflib.circle_points(n), k)
    try:
        h.make_second_kind()
        b = reflib.randn_complex(n, 0)
        x_ref, it_ref, _ = h.solve_gmres(b, tol, max_iter)
        with Plan(h.flatten()) as plan:
            x, it, res, applies = plan.solve_gmres(b, None, tol, max_iter)
        assert it == it_ref, (it, it_ref)
        assert applies == min(it + 1, max_iter)
        assert np.linalg.norm(x - x_ref) / np.linalg.norm(x_ref) <= X_TOL
        # and the reference's own GMRES driving the GPU apply through the swapped vtable
        h.attach()
        x_att, it_att, _ = h.solve_gmres(b, tol, max_iter)
        # ... and the attach layer's bfSolveGMRES-shaped entry point (Arnoldi on the device)
        x_dev, it_dev, _ = h.solve_gmres_cuda(b, tol, max_iter)
        h.detach()
        assert it_att == it_ref and it_dev == it_ref
        assert np.linalg.norm(x_att - x_ref) / np.linalg.norm(x_ref) <= X_TOL
        assert np.linalg.norm(x_dev - x_ref) / np.linalg.norm(x_ref) <= X_TOL
    finally:
        h.close()


@needs_ref
@pytest.mark.parametrize("tol,max_iter", [(1e-12, 50), (1e-6, 50), (1e-12, 3)])
def test_left_preconditioned_gmres_matches_bfSolveGMRES(tol, max_iter):
    """bfSolveGMRES(A, B, NULL, tol, maxNumIter, &numIter, M) with the block Jacobi LU preconditioner of
    multiple_scattering_context.c:544-600,803-810, against bfCudaSolveGMRESMat with the same arguments:
    the operator applies run on the GPU, bfMatSolve(M, .) on the host as in the reference."""
    n = 2048
    h = reflib.Helm2(reflib.circle_points(n), 50.0)
    try:
        h.make_second_kind()
        M = h.make_block_jacobi(8)
        b = reflib.randn_complex(n, 0)
        x_ref, it_ref, _ = h.solve_gmres_precond(b, M, tol, max_iter)
        h.attach()
        x_att, it_att, _ = h.solve_gmres_precond(b, M, tol, max_iter)             # reference GMRES, GPU applies
        x_dev, it_dev, _ = h.solve_gmres_precond(b, M, tol, max_iter, cuda=True)  # Arnoldi on the device
        h.detach()
        assert it_att == it_ref and it_dev == it_ref, (it_ref, it_att, it_dev)
        assert np.linalg.norm(x_att - x_ref) / np.linalg.norm(x_ref) <= X_TOL
        assert np.linalg.norm(x_dev - x_ref) / np.linalg.norm(x_ref) <= X_TOL
        x_plain, _, _ = h.solve_gmres(b, tol, max_iter)
        assert not np.array_equal(x_plain, x_ref)   # the preconditioner changed the iteration
    finally:
        h.close()


def test_device_gmres_on_synthetic_operator_with_initial_iterate():
    """No reference needed: A = 0.5 I + (small HODBF), x0 != 0; the result must solve the system as
    checked with the oracle apply, and a zero right-hand side returns zero at iteration 0."""
    raw = portlib_safe_operator()
    rng = np.random.default_rng(2)
    n = raw.num_rows
    b = rng.standard_normal(n) + 1j * rng.standard_normal(n)
    x0 = 0.1 * (rng.standard_normal(n) + 1j * rng.standard_normal(n))
    with Plan(raw) as plan:
        x, it, res, applies = plan.solve_gmres(b, x0, 1e-12, 60)
        assert 0 < it < 60 and res < 1e-12
        r = portlib.apply(raw, x[:, None], capi.OP_N)[:, 0] - b
        assert np.linalg.norm(r) / np.linalg.norm(b) < 1e-9  # one uncounted column behind (reference semantics)
        xz, itz, _, _ = plan.solve_gmres(np.zeros(n, np.complex128), None, 1e-12, 10)
        assert itz == 0 and not xz.any()


def test_gram_schmidt_building_blocks_against_numpy():
    """bfCudaGmresProject / Subtract / Scale on random data: zdotc projections, the update, its norm."""
    import ctypes as C

    import torch

    lib = capi.load_library()
    rng = np.random.default_rng(2)
    n, mmax = 10007, 19
    Vh = rng.standard_normal((mmax, n)) + 1j * rng.standard_normal((mmax, n))
    wh = rng.standard_normal(n) + 1j * rng.standard_normal(n)
    V, w = torch.from_numpy(Vh).cuda(), torch.from_numpy(wh).cuda()
    scratch = torch.empty(int(lib.bfCudaGmresScratchBytes(mmax)), dtype=torch.uint8, device="cuda")
    out = torch.zeros(mmax + 1, dtype=torch.complex128, device="cuda")
    nrm2 = torch.zeros(1, dtype=torch.complex128, device="cuda")
    p = lambda t: C.c_void_p(t.data_ptr())
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream or 1)
    for m in (0, 1, 8, 9, 19):
        capi.check(lib.bfCudaGmresProject(p(V), n, m, p(w), n, p(out), p(scratch), st), "project")
        got = out.cpu().numpy()
        want = np.conj(Vh[:m]) @ wh
        assert np.allclose(got[:m], want, rtol=1e-13, atol=1e-11)
        assert abs(got[m].real - np.vdot(wh, wh).real) <= 1e-12 * n and got[m].imag == 0
        w2 = w.clone()
        capi.check(lib.bfCudaGmresSubtract(p(V), n, m, p(out), p(w2), n, p(nrm2), p(scratch), st), "subtract")
        wn = wh - Vh[:m].T @ got[:m]
        assert np.allclose(w2.cpu().numpy(), wn, rtol=1e-13, atol=1e-10)
        assert abs(nrm2.cpu().numpy()[0].real - np.vdot(wn, wn).real) <= 1e-12 * np.vdot(wn, wn).real
        d1, d2 = torch.zeros_like(w), torch.zeros_like(w)
        capi.check(lib.bfCudaGmresScale(p(w2), n, p(nrm2), p(d1), p(d2), st), "scale")
        assert np.allclose(d1.cpu().numpy(), wn / np.linalg.norm(wn), rtol=1e-13, atol=1e-15)
        assert torch.equal(d1, d2)
    # run-to-run determinism of the two-stage sums
    a = torch.zeros_like(out)
    capi.check(lib.bfCudaGmresProject(p(V), n, mmax, p(w), n, p(a), p(scratch), st), "project")
    capi.check(lib.bfCudaGmresProject(p(V), n, mmax, p(w), n, p(out), p(scratch), st), "project")
    assert torch.equal(a, out)


def portlib_safe_operator():
    from butterfly_b200 import synth

    return synth.shifted_identity_hodbf(1024, seed=5)


@pytest.mark.parametrize("orth", ["mgs", "cgs", "cgs2"])
def test_sharded_gmres_matches_device_gmres_world1(orth):
    """butterfly_b200/gmres.py (row-distributed GMRES; here one rank on one GPU, stage kernels through
    bfCudaApplyStages, Gram-Schmidt through bfCudaGmresProject / Subtract / Scale) against
    bfCudaSolveGMRES on the same operator: same iteration count and iterate."""
    import os

    import torch
    import torch.distributed as dist

    from butterfly_b200 import gmres, rowshard

    raw = portlib_safe_operator()
    rng = np.random.default_rng(8)
    n = raw.num_rows
    b = rng.standard_normal(n) + 1j * rng.standard_normal(n)
    with Plan(raw) as plan:
        x_ref, it_ref, _, _ = plan.solve_gmres(b, None, 1e-12, 60)
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", "29541")
    created = not dist.is_initialized()
    if created:
        dist.init_process_group("gloo", rank=0, world_size=1)
    try:
        op = rowshard.RowShardedOperator(raw, 1, 0, rowshard.GpuBackend)
        x, it, res, applies = gmres.solve_gmres_sharded(op, torch.from_numpy(b).cuda(), None, 1e-12, 60, orth)
        assert abs(it - it_ref) <= (0 if orth == "mgs" else 1)
        assert np.linalg.norm(x.cpu().numpy() - x_ref) / np.linalg.norm(x_ref) <= (1e-11 if orth != "cgs" else 1e-9)
        op.backend.close()
    finally:
        if created:
            dist.destroy_process_group()
