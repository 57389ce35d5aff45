"""GPU: every vtable slot of an ATTACHED reference object that touches the apply path, against the same
call on the unattached object (the reference's own CPU code, oracle/_ref):

  * bfMatRmul (X A, src/mat_product.c:282-310) through the swapped vtable
  * bfMatMulVec / bfMatRmulVec with a strided BfVecReal view
  * bfMatMulVec with a BfVecComplex (an extension: the reference's block types refuse it)
  * the aliasing slots (GetView / Copy / Delete of a view) and the mutating slots (Transpose,
    ScaleCols + AddInplace) — VERDICT r1 #6, ADVICE r1: a view must never own the plan, and an attached
    object must never apply a stale snapshot.
"""
import numpy as np
import pytest

from oracle import reflib

pytestmark = [pytest.mark.gpu, pytest.mark.skipif(not reflib.attach_available(), reason="oracle/_ref + attach layer not built")]
TOL = 1e-12


def rel(a, b):
    return np.linalg.norm(a - b) / np.linalg.norm(b)


@pytest.fixture()
def helm():
    h = reflib.Helm2(reflib.circle_points(2048), 50.0)
    yield h
    h.close()


@pytest.fixture()
def helm_dense():
    """N = 512: every leaf of the hierarchy is dense (no MatBlockCoo / MatBlockDiag factors), the one
    family of Helmholtz operators on which the reference's own bfMatRmul and bfMatTranspose run at all:
    MatBlockDiag / MatBlockCoo have no Rmul slot and MatBlockCoo no Transpose slot (their vtables,
    src/mat_block_diag.c:31-51, src/mat_block_coo.c:28-47), so on a butterflied operator the reference
    jumps through NULL."""
    h = reflib.Helm2(reflib.circle_points(512), 12.5)
    yield h
    h.close()


def test_complex_rmul_vs_the_reference(helm_dense):
    """One operand row: the reference's bfMatBlockDenseRmul accumulates into column-range VIEWS with a
    stride-blind loop (src/mat_block_dense.c:676, src/mat_dense_complex.c:1571-1588), which is only right
    for a single row."""
    rng = np.random.default_rng(11)
    X = rng.standard_normal((1, 512)) + 1j * rng.standard_normal((1, 512))
    want, _ = helm_dense.rmul(X)                 # reference CPU: X A
    K = helm_dense.dense_kernel()
    assert rel(want, X @ K) <= 1e-13
    helm_dense.attach(build_transpose=1)
    got, _ = helm_dense.rmul(X)
    helm_dense.detach()
    assert got.shape == want.shape == (1, 512)
    assert rel(got[0], want[0]) <= TOL


def test_complex_rmul_on_a_butterflied_operator(helm):
    """Where the reference cannot go (see helm_dense): X A = (A^T X^T)^T against the oracle port of A^T."""
    from butterfly_b200 import capi
    from oracle import portlib

    rng = np.random.default_rng(12)
    X = rng.standard_normal((3, 2048)) + 1j * rng.standard_normal((3, 2048))
    raw = helm.flatten()
    want = portlib.apply(raw, np.ascontiguousarray(X.T), capi.OP_T).T
    helm.attach(build_transpose=1)
    got, _ = helm.rmul(X)
    helm.detach()
    assert got.shape == (3, 2048)
    for r in range(3):
        assert rel(got[r], want[r]) <= TOL


def test_mulvec_complex_is_an_extension_that_matches_bfmatmul(helm):
    x = reflib.randn_complex(2048, 0)
    want, _ = helm.matmul(x)                     # reference CPU, one-column BfMatDenseComplex
    helm.attach(build_transpose=1)
    got = helm.mulvec_complex(x)                 # BfVecComplex, stride 1
    got3 = helm.mulvec_complex(x, stride=3)      # strided view
    gt = helm.mulvec_complex(x, transposed=True)  # RmulVec: A^T x
    want_t, _ = helm.rmul(x[None, :])            # x^T A = (A^T x)^T through the GPU Rmul (pinned above)
    helm.detach()
    assert rel(got, want[:, 0]) <= TOL
    assert np.array_equal(got3, got)
    assert rel(gt, want_t[0]) <= TOL


def test_strided_bfvecreal_through_swapped_vtable():
    s = reflib.sphere_stream(1500, 15)
    rng = np.random.default_rng(4)
    x, v = rng.standard_normal(s.num_cols), rng.standard_normal(s.num_rows)
    y_cpu, _ = s.mulvec_strided(x, 3)
    z_cpu, _ = s.mulvec_strided(v, 2, transposed=True)
    y_dense, _ = s.mulvec(x)
    assert np.array_equal(y_cpu, y_dense)        # the reference itself: a strided view changes nothing
    s.attach(build_transpose=1)
    y_gpu, _ = s.mulvec_strided(x, 3)
    z_gpu, _ = s.mulvec_strided(v, 2, transposed=True)
    y_gpu1, _ = s.mulvec(x)
    s.detach()
    assert rel(y_gpu, y_cpu) <= TOL and rel(z_gpu, z_cpu) <= TOL
    assert np.array_equal(y_gpu, y_gpu1)


@pytest.mark.parametrize("direct", [False, True])
def test_deleting_a_view_leaves_the_attached_owner_usable(helm, direct):
    """multiple_scattering_context.c:616 puts bfMatGet(M, BF_POLICY_VIEW) into a product and deletes it later."""
    x = reflib.randn_complex(2048, 0)
    want, _ = helm.matmul(x)
    helm.attach()
    view = helm.get_view(direct=direct)
    y_view = helm.matmul_on(view, x)             # vtable GetView: a plain reference view (CPU); direct: carries the GPU slots
    helm.delete_mat(view)
    y_after, _ = helm.matmul(x)                  # the owner still has its plan
    y_again, _ = helm.matmul(x)
    helm.detach()
    y_back, _ = helm.matmul(x)
    assert rel(y_view, want) <= TOL and rel(y_after, want) <= TOL
    assert np.array_equal(y_after, y_again)
    assert np.array_equal(y_back, want)


def test_copy_of_an_attached_object_is_a_plain_reference_object(helm):
    x = reflib.randn_complex(2048, 0)
    want, _ = helm.matmul(x)
    helm.attach()
    c = helm.copy()
    helm.detach()
    y = helm.matmul_on(c, x)                     # CPU path on an independent copy
    helm.delete_mat(c)
    assert np.array_equal(y, want)


@pytest.mark.parametrize("build_transpose", [0, 3])
def test_transpose_after_attach_matches_the_reference(helm_dense, build_transpose):
    """bfMatTranspose on an attached object (src/mat_block_dense.c:950-986; complex leaves: the CONJUGATE
    transpose, src/mat_dense_complex.c:1475-1478): with the reverse direction compiled only a flag flips,
    without it the snapshot is retaken from the transposed host object."""
    helm = helm_dense
    x = reflib.randn_complex(512, 0)
    X = np.stack([x, 1j * np.roll(x, 7)], axis=1)
    helm.transpose()
    want, _ = helm.matmul(X)                     # reference CPU: A^H X
    want_r, _ = helm.rmul(x[None, :].copy())     # x^T A^H (one row, see test_complex_rmul_vs_the_reference)
    helm.transpose()
    back, _ = helm.matmul(X)
    helm.attach(build_transpose=build_transpose)
    helm.transpose()
    got, _ = helm.matmul(X)
    got_v = helm.mulvec_complex(x)
    got_r, _ = helm.rmul(x[None, :].copy()) if build_transpose else (None, None)
    helm.transpose()
    got_back, _ = helm.matmul(X)
    helm.detach()
    for c in range(2):
        assert rel(got[:, c], want[:, c]) <= TOL
        assert rel(got_back[:, c], back[:, c]) <= TOL
    assert rel(got_v, want[:, 0]) <= TOL
    if got_r is not None:
        assert rel(got_r[0], want_r[0]) <= TOL


def test_inplace_edits_after_attach_refresh_the_snapshot(helm):
    """bfMatScaleCols + bfMatAddInplace (K diag(w) + I/2, multiple_scattering_context.c:456-487) AFTER the
    attach: the next bfMatMul must apply the edited operator, not the snapshot."""
    x = reflib.randn_complex(2048, 0)
    y0, _ = helm.matmul(x)
    helm.attach()
    y_gpu0, _ = helm.matmul(x)
    helm.make_second_kind()                      # edits through the attached vtable
    y_gpu1, _ = helm.matmul(x)
    helm.detach()
    y_cpu1, _ = helm.matmul(x)                   # the same (edited) object on the CPU
    assert rel(y_gpu0, y0) <= TOL
    assert rel(y_gpu1, y_cpu1) <= TOL
    assert rel(y_cpu1, y0) > 1e-3                # the edit is visible at all
