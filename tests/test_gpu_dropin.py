"""GPU: the drop-in itself.  The reference's own bfMatMul(A, X) is called twice on the same
fac_helm2 object — once as built (CPU, OpenBLAS) and once after bfCudaMatAttach swapped the
object's vtable — and the two results must agree to rel. l2 <= 1e-12 per column (north star).
Also the golden butterfly block (values + the reference's output) through the C ABI."""
from pathlib import Path

import numpy as np
import pytest

from butterfly_b200 import capi
from butterfly_b200.plan import Plan
from butterfly_b200.rawops import RawOps
from oracle import portlib, reflib

pytestmark = pytest.mark.gpu
G = Path(__file__).resolve().parent / "golden"
TOL = 1e-12


def test_golden_block_on_gpu():
    raw = RawOps.load(G / "helm2_block.npz")
    io = np.load(G / "helm2_block_io.npz")
    with Plan(raw, build_transpose=3) as plan:
        y = plan.apply_host(io["x"])
        assert np.linalg.norm(y - io["y"]) / np.linalg.norm(io["y"]) <= TOL
        X = np.stack([io["x"], 1j * io["x"]], axis=1)
        Y = plan.apply_host(X)
        assert np.linalg.norm(Y[:, 1] - 1j * io["y"]) / np.linalg.norm(io["y"]) <= TOL
        z = np.conj(io["y"])
        for op in (capi.OP_T, capi.OP_H):
            assert np.linalg.norm(plan.apply_host(z, op) - portlib.apply(raw, z, op)) <= TOL * np.linalg.norm(z) * 10


needs_ref = pytest.mark.skipif(not reflib.attach_available(), reason="oracle/_ref + attach layer not built")


@needs_ref
@pytest.mark.parametrize("n,k", [(2048, 50.0), (4096, 100.0)])
def test_bfmatmul_through_swapped_vtable(n, k):
    h = reflib.Helm2(reflib.circle_points(n), k)
    try:
        x = reflib.randn_complex(n, 0)
        X = np.stack([x, np.conj(x), 1j * np.roll(x, 3), x * x, np.abs(x) + 0j], axis=1)
        y_cpu, _ = h.matmul(x)
        Y_cpu, _ = h.matmul(X)
        h.attach()
        y_gpu, _ = h.matmul(x)      # same reference call, now on the B200
        Y_gpu, _ = h.matmul(X)
        h.detach()
        y_back, _ = h.matmul(x)
        assert np.linalg.norm(y_gpu - y_cpu) / np.linalg.norm(y_cpu) <= TOL
        for c in range(X.shape[1]):
            assert np.linalg.norm(Y_gpu[:, c] - Y_cpu[:, c]) / np.linalg.norm(Y_cpu[:, c]) <= TOL
        assert np.array_equal(y_back, y_cpu), "detach must restore the CPU path"
        if n == 2048:
            m = np.load(G / "helm2_n2048_meta.npz")
            assert np.linalg.norm(y_gpu[:, 0] - m["y"]) / np.linalg.norm(m["y"]) <= TOL
    finally:
        h.close()


@needs_ref
def test_configs0_same_factors_cpu_vs_gpu():
    """BASELINE.json configs[0] at its own size: fac_helm2 on 16 384 circle points, k = 400
    (examples/simple/bf_all_blocks.c:96-147).  The reference builds the factors ONCE; its bfMatMul runs on
    the CPU and, after bfCudaMatAttach, on the B200 — same factors, same operands — at 1, 64 and 256 RHS,
    rel. l2 <= 1e-12 per column (north star)."""
    n = 16384
    h = reflib.Helm2(reflib.circle_points(n), 400.0)
    try:
        x = reflib.randn_complex(n, 0)
        rng = np.random.default_rng(7)
        X = rng.standard_normal((n, 256)) + 1j * rng.standard_normal((n, 256))
        X[:, 0] = x
        y_cpu, _ = h.matmul(x)
        Y64_cpu, _ = h.matmul(X[:, :64].copy())
        # 256 columns on one core take the reference ~10 s: the last 192 are checked on a spread of 24 of them
        probe = np.arange(64, 256, 8)
        Yp_cpu, _ = h.matmul(X[:, probe].copy())
        h.attach()
        y_gpu, _ = h.matmul(x)
        Y64_gpu, _ = h.matmul(X[:, :64].copy())
        Y256_gpu, _ = h.matmul(X)
        h.detach()
        assert np.linalg.norm(y_gpu - y_cpu) / np.linalg.norm(y_cpu) <= TOL
        for c in range(64):
            assert np.linalg.norm(Y64_gpu[:, c] - Y64_cpu[:, c]) / np.linalg.norm(Y64_cpu[:, c]) <= TOL
            assert np.linalg.norm(Y256_gpu[:, c] - Y64_cpu[:, c]) / np.linalg.norm(Y64_cpu[:, c]) <= TOL
        for i, c in enumerate(probe):
            assert np.linalg.norm(Y256_gpu[:, c] - Yp_cpu[:, i]) / np.linalg.norm(Yp_cpu[:, i]) <= TOL
    finally:
        h.close()


@needs_ref
def test_corrected_system_matrix_through_swapped_vtable():
    """The multiple-scattering system matrix (multiple_scattering_context.c:456-487): Kapur-Rokhlin
    correction, column weights, + I/2, and a sparse MatCooComplex term in a MatSum next to the
    butterflies (src/mat_block_dense.c:493-499) — the reference's bfMatMul before and after attaching."""
    n = 2048
    h = reflib.Helm2(reflib.circle_points(n), 50.0)
    try:
        h.apply_kr_correction(6)
        h.make_second_kind()
        rng = np.random.default_rng(5)
        rows, cols = rng.integers(0, n, 1500), rng.integers(0, n, 1500)
        rows = np.concatenate([rows, np.repeat(np.arange(100, 140), 3)])
        cols = np.concatenate([cols, 1500 + rng.integers(0, 40, 120)])
        vals = rng.standard_normal(len(rows)) + 1j * rng.standard_normal(len(rows))
        h.sum_with_coo(rows, cols, vals)
        x = reflib.randn_complex(n, 0)
        X = np.stack([x, np.conj(x), 1j * np.roll(x, 3)], axis=1)
        # one column per reference call: bfMatCooComplexMul corrupts the heap for wider operands
        Y_cpu = np.concatenate([h.matmul(X[:, c].copy())[0] for c in range(X.shape[1])], axis=1)
        h.attach()
        y_gpu, _ = h.matmul(x)
        Y_gpu, _ = h.matmul(X)
        h.detach()
        assert np.linalg.norm(y_gpu[:, 0] - Y_cpu[:, 0]) / np.linalg.norm(Y_cpu[:, 0]) <= TOL
        for c in range(X.shape[1]):
            assert np.linalg.norm(Y_gpu[:, c] - Y_cpu[:, c]) / np.linalg.norm(Y_cpu[:, c]) <= TOL
        # the sparse term applied as the matrix it stores (entries of a row summed)
        D = np.zeros((n, n), np.complex128)
        np.add.at(D, (rows, cols), vals)
        raw_acc = h.flatten(flags=1)
        with Plan(raw_acc) as plan:
            y_acc = plan.apply_host(x)
        want = portlib.apply(raw_acc, x[:, None])[:, 0]
        assert np.linalg.norm(y_acc - want) / np.linalg.norm(want) <= TOL
    finally:
        h.close()


@needs_ref
def test_fused_tree_permutation_matches_bfMatPermuteRows():
    """Callers permute around the apply (examples/simple/helm2_bie.c:70-81,139); fused into the
    plan the result must be bit-identical to permuting with the reference's own routine."""
    n = 2048
    h = reflib.Helm2(reflib.circle_points(n), 50.0)
    try:
        raw = h.flatten()
        perm, rev = h.perm(), reflib.reverse_perm(h.perm())
        x = reflib.randn_complex(n, 0)
        with Plan(raw) as plan:
            y_plain = plan.apply_host(reflib.permute_rows_complex(x, rev))
            want = reflib.permute_rows_complex(y_plain, perm)
            plan.set_perms(rev, perm)
            assert np.array_equal(plan.apply_host(x), want)
    finally:
        h.close()


# ---- real path (configs[3]: LBO-style streamed butterfly, forward and transpose) ------------------
@pytest.mark.parametrize("name", ["lbo_stream_n600", "lbo_sphere_fixture"])
def test_golden_stream_on_gpu(name):
    """The streamed real operators with the reference's factor values (spherical harmonics on 600 points;
    the reference's own tests/sphere_{Lam,Phi}.txt fixture), against the reference's own bfMatMulVec /
    bfMatRmulVec outputs (tests/golden, no reference needed at run time)."""
    raw = RawOps.load(G / f"{name}.npz")
    io = np.load(G / f"{name}_io.npz")
    with Plan(raw, build_transpose=1) as plan:
        y = plan.apply_host(io["x"], capi.OP_N)
        z = plan.apply_host(io["v"], capi.OP_T)
        assert np.linalg.norm(y - io["y"]) / np.linalg.norm(io["y"]) <= TOL
        assert np.linalg.norm(z - io["z"]) / np.linalg.norm(io["z"]) <= TOL
        # a block of RHS (no reference equivalent: bfMatDenseRealMul is BF_DIE): columns independent
        X = np.stack([io["x"], -2.0 * io["x"], np.roll(io["x"], 5)], axis=1)
        Y = plan.apply_host(X, capi.OP_N)
        assert np.linalg.norm(Y[:, 0] - io["y"]) / np.linalg.norm(io["y"]) <= TOL
        assert np.linalg.norm(Y[:, 1] + 2.0 * io["y"]) / np.linalg.norm(io["y"]) <= 2 * TOL


@needs_ref
def test_mulvec_rmulvec_through_swapped_vtable():
    """bfMatMulVec(Phi, x) / bfMatRmulVec(Phi, v) (examples/covariance/lbo_cov.c:41,51,56) on a live
    bfFacStreamer factorization: CPU reference vs the same calls after bfCudaMatAttach."""
    s = reflib.sphere_stream(1500, 15)
    rng = np.random.default_rng(3)
    x, v = rng.standard_normal(s.num_cols), rng.standard_normal(s.num_rows)
    y_cpu, _ = s.mulvec(x)
    z_cpu, _ = s.mulvec(v, transposed=True)
    s.attach(build_transpose=1)
    y_gpu, _ = s.mulvec(x)
    z_gpu, _ = s.mulvec(v, transposed=True)
    s.detach()
    y_back, _ = s.mulvec(x)
    assert np.linalg.norm(y_gpu - y_cpu) / np.linalg.norm(y_cpu) <= TOL
    assert np.linalg.norm(z_gpu - z_cpu) / np.linalg.norm(z_cpu) <= TOL
    assert np.array_equal(y_back, y_cpu)


def test_reference_streamed_fixture_parity():
    """configs[3] above toy size: the operator tools/make_lbo_fixture.py had the unmodified reference stream
    (bfFacStreamer on sphere points x real spherical harmonics), applied on the GPU, against the reference's own
    bfMatMulVec / bfMatRmulVec results stored with it.  The fixture is built offline (minutes) and is not in git."""
    from pathlib import Path

    from butterfly_b200.rawops import RawOps

    fx = Path(__file__).resolve().parent.parent / "fixtures" / "lbo_sphere"
    if not (Path(str(fx) + ".raw.npz").exists() and Path(str(fx) + ".ref.npz").exists()):
        pytest.skip("fixtures/lbo_sphere.* not built (python tools/make_lbo_fixture.py)")
    raw = RawOps.load(str(fx) + ".raw.npz")
    ref = np.load(str(fx) + ".ref.npz")
    with Plan(raw, build_transpose=1) as plan:
        for op, xk, yk in ((capi.OP_N, "X", "Y"), (capi.OP_T, "XT", "YT")):
            X, Yref = ref[xk], ref[yk]
            for j in range(X.shape[1]):
                y = plan.apply_host(np.ascontiguousarray(X[:, j:j + 1]), op)[:, 0]
                assert np.linalg.norm(y - Yref[:, j]) / np.linalg.norm(Yref[:, j]) <= TOL
            Y = plan.apply_host(np.ascontiguousarray(X), op)  # the same columns as one block
            assert np.linalg.norm(Y - Yref) / np.linalg.norm(Yref) <= TOL
