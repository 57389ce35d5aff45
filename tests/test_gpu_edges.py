"""GPU parity on the edges: empty operators, zero right-hand sides, single leaves of awkward shapes
(1 x 1, one past every tiling boundary, much taller / wider than a tile), stored-transposed leaves
with padded strides, leaves overlapping on the same output rows, and a dense matrix against numpy."""
import numpy as np
import pytest

from butterfly_b200 import capi
from butterfly_b200.plan import Plan
from butterfly_b200.synth import _Builder
from oracle import portlib

pytestmark = pytest.mark.gpu
TOL = 1e-12


def _x(rng, n, nrhs, cplx):
    x = rng.standard_normal((n, nrhs))
    return x + 1j * rng.standard_normal((n, nrhs)) if cplx else x


def _rel(y, y_ref):
    d = np.linalg.norm(y_ref)
    return np.linalg.norm(y - y_ref) / d if d > 0 else np.linalg.norm(y)


@pytest.mark.parametrize("dtype", [capi.REAL_F64, capi.COMPLEX_F64])
def test_empty_operator_is_zero(dtype):
    b = _Builder(dtype, True, np.random.default_rng(0))
    raw = b.finish(37, 21)  # no leaves at all: Y = 0 (bfMatZerosLike, src/mat_block_coo.c:401)
    cplx = dtype == capi.COMPLEX_F64
    with Plan(raw, build_transpose=1) as plan:
        for nrhs in (1, 5):
            y = plan.apply_host(_x(np.random.default_rng(1), 21, nrhs, cplx))
            assert y.shape == (37, nrhs) and not y.any()
        assert not plan.apply_host(_x(np.random.default_rng(1), 37, 1, cplx), capi.OP_T).any()
        assert plan.info["numEntries"] == 0


def test_zero_columns_is_a_no_op(lib):
    import ctypes as C

    b = _Builder(capi.REAL_F64, True, np.random.default_rng(0))
    b.dense(0, 0, 1, 0, 9, 7)
    raw = b.finish(9, 7)
    with Plan(raw) as plan:
        x = np.zeros((7, 1))
        y = np.full((9, 1), 3.0)
        assert lib.bfCudaApplyHost(plan.handle, capi.OP_N, x.ctypes.data, 1, 0, y.ctypes.data, 1) == 0
        assert (y == 3.0).all()


SHAPES = [(1, 1), (1, 200), (200, 1), (7, 9), (8, 4), (9, 5), (31, 33), (32, 32), (33, 31), (65, 129), (17, 300), (300, 17)]


@pytest.mark.parametrize("dtype", [capi.REAL_F64, capi.COMPLEX_F64])
@pytest.mark.parametrize("trans", [False, True])
def test_single_leaf_shapes(dtype, trans):
    cplx = dtype == capi.COMPLEX_F64
    rng = np.random.default_rng(5)
    for m, k in SHAPES:
        b = _Builder(dtype, True, rng)
        b.dense(0, 0, 1, 0, m, k, trans=trans, pad=3 if (m + k) % 2 else 0)
        raw = b.finish(m, k)
        with Plan(raw, build_transpose=3) as plan:
            for nrhs in (1, 2, 9, 40):
                x = _x(rng, k, nrhs, cplx)
                assert _rel(plan.apply_host(x), portlib.apply(raw, x)) <= TOL, (m, k, nrhs)
            xt = _x(rng, m, 3, cplx)
            for op in (capi.OP_T,) + ((capi.OP_H,) if cplx else ()):
                assert _rel(plan.apply_host(xt, op), portlib.apply(raw, xt, op)) <= TOL, (m, k, op)


@pytest.mark.parametrize("dtype", [capi.REAL_F64, capi.COMPLEX_F64])
def test_large_dense_leaf_against_numpy(dtype):
    """One 1500 x 2300 leaf: ~50 items x ~100 tiles each; checked against numpy's matmul directly."""
    cplx = dtype == capi.COMPLEX_F64
    rng = np.random.default_rng(9)
    m, k = 1500, 2300
    b = _Builder(dtype, True, rng)
    b.dense(0, 0, 1, 0, m, k)
    raw = b.finish(m, k)
    A = raw.values.view(np.complex128).reshape(m, k) if cplx else raw.values.reshape(m, k)
    with Plan(raw, build_transpose=1) as plan:
        for nrhs in (1, 64):
            x = _x(rng, k, nrhs, cplx)
            assert _rel(plan.apply_host(x), A @ x) <= TOL
        xt = _x(rng, m, 1, cplx)
        assert _rel(plan.apply_host(xt, capi.OP_T), A.T @ xt) <= TOL


def test_overlapping_leaves_accumulate_in_order():
    """Several leaves on the same output rows, partially overlapping, through a product: the sums must
    follow the op order (src/mat_block_coo.c:404-418) — checked against the oracle, and bit-for-bit
    between two runs."""
    rng = np.random.default_rng(3)
    b = _Builder(capi.COMPLEX_F64, True, rng)
    t = b.new_buf(50)
    b.dense(0, 0, t, 0, 50, 40)
    for (r0, m, c0, k) in [(0, 30, 0, 50), (10, 40, 5, 45), (25, 20, 0, 20), (0, 60, 10, 40)]:
        b.dense(t, c0, 1, r0, m, k)
    b.identity(0, 0, 1, 5, 40, 40)
    b.diag(0, 3, 1, 20, 30, 30)
    raw = b.finish(60, 40)
    with Plan(raw) as plan:
        for nrhs in (1, 6):
            x = _x(rng, 40, nrhs, True)
            y = plan.apply_host(x)
            assert _rel(y, portlib.apply(raw, x)) <= TOL
            assert np.array_equal(y, plan.apply_host(x))
