"""GPU parity: the CUDA path, called through the C ABI, against the CPU oracle on the same seeded
inputs.  Tolerance: relative l2 <= 1e-12 per the north star (fp64; summation order inside a leaf
differs from the oracle's, so bitwise equality is not expected); index handling (permutations,
gaps, strides) is checked exactly where it is exact."""
import numpy as np
import pytest

from butterfly_b200 import capi, synth
from butterfly_b200.plan import Plan
from oracle import portlib

pytestmark = pytest.mark.gpu
TOL = 1e-12


def _x(rng, n, nrhs, cplx):
    x = rng.standard_normal((n, nrhs))
    return x + 1j * rng.standard_normal((n, nrhs)) if cplx else x


def _relerr(y, y_ref):
    return np.linalg.norm(y - y_ref) / max(np.linalg.norm(y_ref), 1e-300)


CASES = {
    "hodbf_c": lambda: synth.random_hodbf(n=3000, dtype=capi.COMPLEX_F64, seed=3),
    "hodbf_r": lambda: synth.random_hodbf(n=2100, dtype=capi.REAL_F64, seed=4),
    "rect_r": lambda: synth.random_rect(700, 300, capi.REAL_F64, seed=5),
    "rect_c": lambda: synth.random_rect(333, 517, capi.COMPLEX_F64, seed=6),
    "plain_c": lambda: synth.random_hodbf(n=4096, dtype=capi.COMPLEX_F64, seed=7, exotic=False, leaf=64, rank_lo=16, rank_hi=40),
}


@pytest.fixture(scope="module", params=list(CASES))
def case(request):
    raw = CASES[request.param]()
    plan = Plan(raw, build_transpose=3)
    yield raw, plan
    plan.close()


@pytest.mark.parametrize("nrhs", [1, 2, 7, 33, 256])
def test_forward(case, nrhs):
    raw, plan = case
    cplx = raw.dtype == capi.COMPLEX_F64
    x = _x(np.random.default_rng(nrhs), raw.num_cols, nrhs, cplx)
    y = plan.apply_host(x, capi.OP_N)
    y_ref = portlib.apply(raw, x, capi.OP_N)
    assert _relerr(y, y_ref) <= TOL
    for c in range(0, nrhs, max(1, nrhs // 5)):
        assert _relerr(y[:, c], y_ref[:, c]) <= TOL


@pytest.mark.parametrize("op", [capi.OP_T, capi.OP_H])
@pytest.mark.parametrize("nrhs", [1, 5])
def test_transpose(case, op, nrhs):
    raw, plan = case
    cplx = raw.dtype == capi.COMPLEX_F64
    if op == capi.OP_H and not cplx:
        pytest.skip("OP_H is complex only")
    x = _x(np.random.default_rng(7), raw.num_rows, nrhs, cplx)
    y = plan.apply_host(x, op)
    assert _relerr(y, portlib.apply(raw, x, op)) <= TOL


def test_strided_host_buffers(case):
    raw, plan = case
    cplx = raw.dtype == capi.COMPLEX_F64
    rng = np.random.default_rng(1)
    for nrhs, ld in ((1, 3), (4, 9)):
        xb = _x(rng, raw.num_cols, ld, cplx)
        yb = np.full((raw.num_rows, ld + 2), 7.0, xb.dtype)
        plan.apply_host(xb[:, :nrhs], capi.OP_N, out=yb[:, :nrhs])
        y_ref = portlib.apply(raw, np.ascontiguousarray(xb[:, :nrhs]), capi.OP_N)
        assert _relerr(yb[:, :nrhs], y_ref) <= TOL
        assert np.all(yb[:, nrhs:] == 7.0), "wrote outside the result columns"


def test_fused_permutations_bit_exact(case):
    """out[perm[i]] = in[i] semantics (src/mat_dense_complex.c:741-774) fused into the first
    gather / last scatter: identical bits to permuting on the host around an unpermuted apply."""
    raw, plan = case
    cplx = raw.dtype == capi.COMPLEX_F64
    rng = np.random.default_rng(3)
    pin, pout = rng.permutation(raw.num_cols), rng.permutation(raw.num_rows)
    for nrhs in (1, 6):
        x = _x(rng, raw.num_cols, nrhs, cplx)
        plan.set_perms(None, None)
        y_plain = plan.apply_host(portlib.permute_rows(pin, x), capi.OP_N)
        plan.set_perms(pin, pout)
        y_fused = plan.apply_host(x, capi.OP_N)
        assert np.array_equal(y_fused, portlib.permute_rows(pout, y_plain))
        # transposed direction: (P_out A P_in)^T
        xt = _x(rng, raw.num_rows, nrhs, cplx)
        yt_fused = plan.apply_host(xt, capi.OP_T)
        plan.set_perms(None, None)
        inv_out, inv_in = np.argsort(pout), np.argsort(pin)
        yt_plain = plan.apply_host(portlib.permute_rows(inv_out, xt), capi.OP_T)
        assert np.array_equal(yt_fused, portlib.permute_rows(inv_in, yt_plain))
    plan.set_perms(None, None)


def test_linearity_and_determinism(case):
    raw, plan = case
    cplx = raw.dtype == capi.COMPLEX_F64
    rng = np.random.default_rng(9)
    x1, x2 = _x(rng, raw.num_cols, 1, cplx), _x(rng, raw.num_cols, 1, cplx)
    y1, y2 = plan.apply_host(x1), plan.apply_host(x2)
    y12 = plan.apply_host(2.0 * x1 - 3.0 * x2)
    assert _relerr(y12, 2.0 * y1 - 3.0 * y2) <= 1e-13
    assert np.array_equal(plan.apply_host(x1), y1), "no atomics: repeated applies are bit-identical"
    # a column of a wide apply equals... the same operator: 1-RHS and wide kernels agree
    X = np.concatenate([x1, x2, x1 + x2], axis=1)
    Y = plan.apply_host(X)
    assert _relerr(Y[:, :1], y1) <= 1e-13 and _relerr(Y[:, 1:2], y2) <= 1e-13


def test_synthetic_fill_matches_host_restatement():
    raw = synth.random_hodbf(n=1500, dtype=capi.COMPLEX_F64, seed=12, with_values=False)
    with Plan(raw, build_transpose=1) as plan:
        plan.fill_synthetic(42)
        host = portlib.materialize(raw, 42)
        rng = np.random.default_rng(0)
        x = _x(rng, raw.num_cols, 3, True)
        assert _relerr(plan.apply_host(x), portlib.apply(host, x)) <= TOL
        xt = _x(rng, raw.num_rows, 1, True)
        assert _relerr(plan.apply_host(xt, capi.OP_T), portlib.apply(host, xt, capi.OP_T)) <= TOL


def test_device_api_and_launch_count():
    import torch

    raw = synth.random_hodbf(n=2048, dtype=capi.COMPLEX_F64, seed=5)
    with Plan(raw) as plan:
        x = _x(np.random.default_rng(2), raw.num_cols, 1, True)
        xd = torch.from_numpy(x).cuda()
        yd = torch.empty((raw.num_rows, 1), dtype=torch.complex128, device="cuda")
        torch.cuda.synchronize()
        plan.apply_device(xd, yd, stream=torch.cuda.current_stream().cuda_stream)
        torch.cuda.synchronize()
        assert _relerr(yd.cpu().numpy(), portlib.apply(raw, x)) <= TOL
        # one cooperative launch for the whole apply, or one launch per stage
        assert plan.last_launch_count == (1 if plan.is_fused() else plan.info["numStagesN"])
        plan.set_fused(not plan.is_fused())
        y2 = torch.empty_like(yd)
        plan.apply_device(xd, y2, stream=torch.cuda.current_stream().cuda_stream)
        torch.cuda.synchronize()
        assert torch.equal(y2, yd)


def test_errors_do_not_abort(lib):
    raw = synth.random_rect(50, 40, capi.REAL_F64, seed=1)
    with Plan(raw) as plan:
        with pytest.raises(capi.BfCudaError):
            plan.apply_host(np.zeros((50, 1)), capi.OP_T)  # direction not built
        with pytest.raises(ValueError):
            plan.apply_host(np.zeros((41, 1)))


@pytest.mark.parametrize("op", [capi.OP_N, capi.OP_T, capi.OP_H])
def test_single_launch_apply_is_bit_identical_to_per_stage_launches(case, op):
    """The 1-RHS apply as ONE cooperative launch (fused tables: flat workspace, readiness counters instead of
    the stage barrier) forms every sum from the same operands in the same order as the per-stage launches.
    Repeated: the two counter sets alternate between applies."""
    raw, plan = case
    cplx = raw.dtype == capi.COMPLEX_F64
    if op == capi.OP_H and not cplx:
        pytest.skip("OP_H is complex only")
    n_in = raw.num_cols if op == capi.OP_N else raw.num_rows
    x = _x(np.random.default_rng(21), n_in, 1, cplx)
    plan.set_fused(False)
    y_staged = plan.apply_host(x, op).copy()
    stages = plan.last_launch_count
    plan.set_fused(True)
    assert plan.is_fused(op), "fused tables missing for a single-GPU plan"
    for rep in range(5):
        y = plan.apply_host(x * (rep + 1), op)
        assert plan.last_launch_count <= 2  # the apply itself (+ the operand staging copy where one is needed)
        assert np.array_equal(y, y_staged * (rep + 1)) or _relerr(y, y_staged * (rep + 1)) <= 1e-15
        if rep == 0:
            assert np.array_equal(y, y_staged), "fused and per-stage results differ"
    assert stages >= plan.last_launch_count
    assert _relerr(y_staged, portlib.apply(raw, x, op)) <= TOL


def test_synthetic_fill_right_after_create_with_several_live_plans():
    """Regression: the plan's stream is non-blocking, so nothing of plan creation or of the synthetic fill may be
    left on the legacy stream (a cudaMemset of the pool, pageable cudaMemcpy of the tables) when the fill kernel
    or the first apply starts.  Six plans created and filled back to back once left one of them with a partly
    zeroed pool.  Every plan must agree with the first one, and a re-fill must change nothing."""
    import torch

    raw = synth.random_hodbf(n=24000, dtype=capi.COMPLEX_F64, seed=19, leaf=64, rank_lo=16, rank_hi=40, with_values=False, exotic=False)
    x = torch.from_numpy(_x(np.random.default_rng(4), raw.num_cols, 1, True)).cuda()
    stream = torch.cuda.current_stream().cuda_stream
    plans = []
    try:
        for _ in range(5):
            p = Plan(raw)
            p.fill_synthetic(11)
            plans.append(p)
        ys = []
        for p in plans:
            y = torch.empty_like(x)
            for _ in range(3):
                p.apply_device(x, y, stream=stream)
            torch.cuda.synchronize()
            ys.append(y)
        for y in ys[1:]:
            assert torch.equal(y, ys[0])
        plans[0].fill_synthetic(11)
        y = torch.empty_like(x)
        plans[0].apply_device(x, y, stream=stream)
        torch.cuda.synchronize()
        assert torch.equal(y, ys[0])
    finally:
        for p in plans:
            p.close()
