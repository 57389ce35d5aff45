"""CPU: the plan compiler's tables, interpreted with numpy, reproduce the oracle.  Pins stage
assignment, item / tile row ranges, workspace aliasing, queue partition, transposition and
conjugation and pool packing without a GPU."""
import numpy as np
import pytest

from butterfly_b200 import capi, synth
from oracle import portlib
from table_interp import Tables


def _x(rng, n, nrhs, cplx):
    x = rng.standard_normal((n, nrhs))
    return x + 1j * rng.standard_normal((n, nrhs)) if cplx else x


CASES = [
    ("hodbf_c", lambda: synth.random_hodbf(n=1500, dtype=capi.COMPLEX_F64, seed=3)),
    ("hodbf_r", lambda: synth.random_hodbf(n=1100, dtype=capi.REAL_F64, seed=4)),
    ("rect_r", lambda: synth.random_rect(700, 300, capi.REAL_F64, seed=5)),
    ("rect_c", lambda: synth.random_rect(333, 517, capi.COMPLEX_F64, seed=6)),
    ("plain_c", lambda: synth.random_hodbf(n=900, dtype=capi.COMPLEX_F64, seed=7, exotic=False)),
]


@pytest.mark.parametrize("name,make", CASES)
@pytest.mark.parametrize("op", [capi.OP_N, capi.OP_T, capi.OP_H])
def test_tables_match_oracle(name, make, op):
    raw = make()
    cplx = raw.dtype == capi.COMPLEX_F64
    if op == capi.OP_H and not cplx:
        pytest.skip("OP_H is complex only")
    rng = np.random.default_rng(11)
    n_in = raw.num_cols if op == capi.OP_N else raw.num_rows
    x = _x(rng, n_in, 3, cplx)
    y_ref = portlib.apply(raw, x, op)
    t = Tables(raw, op, num_queues=37)
    try:
        t.check_queues()
        y = t.apply(x)
    finally:
        t.close()
    assert not np.isnan(y).any(), "a result row is never written"
    err = np.linalg.norm(y - y_ref) / np.linalg.norm(y_ref)
    assert err < 1e-13, err


@pytest.mark.parametrize("name,make", CASES)
@pytest.mark.parametrize("op", [capi.OP_N, capi.OP_T])
def test_fused_tables_under_adversarial_schedules(name, make, op):
    """The single-launch tables (readiness counters instead of stage barriers): bit-identical to the stage-major
    interpretation whatever order the queues run in, no deadlock, no read of an unwritten workspace row."""
    raw = make()
    cplx = raw.dtype == capi.COMPLEX_F64
    rng = np.random.default_rng(12)
    x = _x(rng, raw.num_cols if op == capi.OP_N else raw.num_rows, 2, cplx)
    t = Tables(raw, op, num_queues=23)
    try:
        assert t.fused
        y = t.apply(x)
        for order, seed in (("forward", 0), ("reverse", 0), ("random", 1), ("random", 2)):
            yf, waits = t.apply_fused(x, order, seed)
            assert not np.isnan(yf).any()
            assert np.array_equal(yf, y), order
            if order == "reverse":
                assert waits > 0  # the schedule really ran consumers ahead of their producers
    finally:
        t.close()


def test_queue_count_does_not_change_the_result():
    raw = synth.random_hodbf(n=1024, dtype=capi.COMPLEX_F64, seed=9)
    x = _x(np.random.default_rng(1), raw.num_cols, 1, True)
    ys = []
    for q in (1, 2, 8, 1184):
        t = Tables(raw, capi.OP_N, num_queues=q)
        t.check_queues()
        ys.append(t.apply(x))
        t.close()
    # dealing the same items to more queues changes nothing, bit for bit ...
    assert np.array_equal(ys[0], ys[1])
    # ... and when a queue's share gets small enough that long items are cut into fewer rows (other
    # tile shapes, another summation order) the result moves by rounding only
    for y in ys[2:]:
        assert np.linalg.norm(y - ys[0]) <= 1e-14 * np.linalg.norm(ys[0])


def test_oracle_transpose_is_adjoint():
    """<A x, y> == <x, A^H y> for the oracle itself (independent sanity of the checker)."""
    raw = synth.random_hodbf(n=700, dtype=capi.COMPLEX_F64, seed=2)
    rng = np.random.default_rng(5)
    x, y = _x(rng, 700, 1, True), _x(rng, 700, 1, True)
    lhs = np.vdot(y, portlib.apply(raw, x, capi.OP_N))
    rhs = np.vdot(portlib.apply(raw, y, capi.OP_H), x)
    assert abs(lhs - rhs) / abs(lhs) < 1e-12


def _heavy_row_operator(seed=31):
    """A HODLR-of-butterflies operator plus one block row whose near-field leaves (X -> Y) are far longer than a
    queue's share of stage 0: the shape that left one warp working alone at N = 16 384 (bf_plan.cpp, step 4b)."""
    from butterfly_b200.synth import _Builder, _hodlr

    n = 1536
    rng = np.random.default_rng(seed)
    b = _Builder(capi.COMPLEX_F64, True, rng)
    _hodlr(b, 0, n, 48, 9, 23, False)
    for j in range(12):  # rows 200..239 get 12 x 120 more columns
        b.dense(0, 40 + 120 * j, 1, 200, 40, 120, trans=(j % 3 == 1))
    return b.finish(n, n)


def _stage_entries(t):
    """matrix entries streamed per stage (dense tiles only)"""
    out = []
    for s in range(t.num_stages):
        tl = t.tiles[t.stage_tile[s]:t.stage_tile[s + 1]]
        kc, rows = tl["kc"].astype(np.int64), tl["rows"].astype(np.int64)
        out.append(int((rows * kc)[kc < 0xFFFD].sum()))
    return out


@pytest.mark.parametrize("op", [capi.OP_N, capi.OP_T])
def test_long_result_rows_are_spread_over_later_stages(op, monkeypatch):
    raw = _heavy_row_operator()
    x = _x(np.random.default_rng(5), raw.num_cols if op == capi.OP_N else raw.num_rows, 2, True)
    y_ref = portlib.apply(raw, x, op)
    monkeypatch.setenv("BF_PLAN_DEFER", "0")
    t0 = Tables(raw, op, num_queues=96)
    monkeypatch.setenv("BF_PLAN_DEFER", "1")
    t1 = Tables(raw, op, num_queues=96)
    try:
        for t in (t0, t1):
            t.check_queues()
            y = t.apply(x)
            assert not np.isnan(y).any()
            assert np.linalg.norm(y - y_ref) <= 1e-13 * np.linalg.norm(y_ref)
        c0, c1 = _stage_entries(t0), _stage_entries(t1)
        assert sum(c0) == sum(c1), "deferral moves leaves between stages, it does not add or drop entries"
        if op == capi.OP_N:  # (transposed, the long rows become long columns: many short items, nothing to move)
            assert c1[0] < c0[0], "nothing was moved out of stage 0"
        assert t1.fused
        y1 = t1.apply(x)
        for order, seed in (("reverse", 0), ("random", 3)):
            yf, _ = t1.apply_fused(x, order, seed)
            assert np.array_equal(yf, y1), order
    finally:
        t0.close()
        t1.close()


def _stacked_leaves_operator(dtype, seed=41):
    """The end of a streamed factorization (bfFacStreamer's Psi): a block column of many leaves of a few rows each
    over shared column ranges, followed by a run of identity leaves; plus leaves side by side in the operand rows,
    one stored transposed and one with a padded row stride (neither of which may be merged blindly)."""
    from butterfly_b200.synth import _Builder

    rng = np.random.default_rng(seed)
    b = _Builder(dtype, True, rng)
    n_in, mid, n_out = 300, 260, 411
    t = b.new_buf(mid)      # X -> t: leaves side by side (k = 5..20 each) over three row blocks
    u = b.new_buf(n_out)    # t -> u: vertical stacks over two column ranges
    r0 = 0
    for rows in (90, 100, 70):
        c0 = 0
        while c0 < n_in:
            k = int(min(n_in - c0, rng.integers(5, 21)))
            b.dense(0, c0, t, r0, rows, k, trans=(c0 == 0 and rows == 100), pad=(3 if c0 > 200 else 0))
            c0 += k
        r0 += rows
    r = 0
    while r < n_out:
        m = int(min(n_out - r, rng.integers(3, 13)))
        b.dense(t, 0, u, r, m, 200)
        b.dense(t, 200, u, r, m, 60, pad=(2 if r % 2 else 0))
        r += m
    r = 0
    while r < n_out:  # u -> Y: a run of identity leaves with the same cuts
        m = int(min(n_out - r, rng.integers(3, 13)))
        b.identity(u, r, 1, r, m, m)
        r += m
    return b.finish(n_out, n_in)


@pytest.mark.parametrize("dtype", [capi.COMPLEX_F64, capi.REAL_F64])
@pytest.mark.parametrize("op", [capi.OP_N, capi.OP_T])
def test_stacks_of_small_leaves_are_merged(dtype, op, monkeypatch):
    raw = _stacked_leaves_operator(dtype)
    cplx = dtype == capi.COMPLEX_F64
    x = _x(np.random.default_rng(6), raw.num_cols if op == capi.OP_N else raw.num_rows, 3, cplx)
    y_ref = portlib.apply(raw, x, op)
    counts = {}
    for merge in ("0", "1"):
        monkeypatch.setenv("BF_PLAN_MERGE", merge)
        t = Tables(raw, op, num_queues=29)
        try:
            t.check_queues()
            y = t.apply(x)
            assert not np.isnan(y).any()
            assert np.linalg.norm(y - y_ref) <= 1e-13 * np.linalg.norm(y_ref), merge
            kc = t.tiles["kc"].astype(np.int64)
            counts[merge] = (int((kc < 0xFFFD).sum()), len(t.items), sum(_stage_entries(t)))
            if merge == "1" and t.fused:
                yf, _ = t.apply_fused(x, "reverse", 0)
                assert np.array_equal(yf, y)
        finally:
            t.close()
    # the run of identity leaves u -> Y is ONE copy of a whole temporary once merged: elided, its writers store into Y
    monkeypatch.setenv("BF_PLAN_MERGE", "1")
    stages = {}
    for elide in ("0", "1"):
        monkeypatch.setenv("BF_PLAN_ELIDE", elide)
        t = Tables(raw, op, num_queues=29)
        try:
            stages[elide] = t.num_stages
            y = t.apply(x)
            assert np.linalg.norm(y - y_ref) <= 1e-13 * np.linalg.norm(y_ref), elide
        finally:
            t.close()
    assert stages["1"] == stages["0"] - 1, stages
    assert counts["1"][2] == counts["0"][2], "merging moves no entry in or out of the operator"
    assert counts["1"][0] < 0.6 * counts["0"][0], counts  # far fewer (and wider / taller) tiles
    assert counts["1"][1] < counts["0"][1], counts        # and fewer items
