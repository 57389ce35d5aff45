"""CPU, only where oracle/_ref exists: the flattener + port oracle + plan tables against the
reference's own bfMatMul on a live fac_helm2 factorization (the pin SURVEY.md §8c names:
examples/simple/bf_all_blocks.c:136-153)."""
from pathlib import Path

import numpy as np
import pytest

from butterfly_b200 import capi
from oracle import portlib, reflib
from table_interp import Tables

pytestmark = [
    pytest.mark.reference,
    pytest.mark.skipif(not (reflib.available() and capi.ATTACH_LIB_PATH.exists()), reason="oracle/_ref or the attach layer is not built"),
]
G = Path(__file__).resolve().parent / "golden"


@pytest.fixture(scope="module")
def helm():
    h = reflib.Helm2(reflib.circle_points(2048), 50.0)
    yield h
    h.close()


def test_flatten_matches_golden_structure(helm):
    m = np.load(G / "helm2_n2048_meta.npz")
    raw = helm.flatten()
    assert np.array_equal(raw.fingerprint(), m["fingerprint"]), "flattened index tables changed"
    assert np.array_equal(raw.buf_rows, m["buf_rows"])
    assert raw.num_entries == int(m["num_entries"])
    assert np.array_equal(helm.perm(), m["perm"])


def test_reference_reproduces_its_golden_output(helm):
    m = np.load(G / "helm2_n2048_meta.npz")
    assert np.array_equal(reflib.randn_complex(2048, 0), m["x"])  # bfSeed(0) + bfComplexRandn
    y, _ = helm.matmul(m["x"])
    assert np.linalg.norm(y[:, 0] - m["y"]) / np.linalg.norm(m["y"]) < 1e-14


@pytest.mark.parametrize("nrhs", [1, 3])
def test_port_oracle_and_tables_match_reference(helm, nrhs):
    m = np.load(G / "helm2_n2048_meta.npz")
    x = m["x"][:, None] if nrhs == 1 else m["x3"]
    y_ref, _ = helm.matmul(x)
    raw = helm.flatten()
    y_port = portlib.apply(raw, x)
    assert np.linalg.norm(y_port - y_ref) / np.linalg.norm(y_ref) < 1e-13
    t = Tables(raw, capi.OP_N, 148 * 8)
    try:
        y_tab = t.apply(x)
    finally:
        t.close()
    assert np.linalg.norm(y_tab - y_ref) / np.linalg.norm(y_ref) < 1e-13


def test_butterfly_vs_dense_kernel_pin(helm):
    """The reference's own accuracy check (bf_all_blocks.c:149-153): butterfly vs dense kernel."""
    x = reflib.randn_complex(2048, 0)
    y, _ = helm.matmul(x)
    yd = helm.dense_kernel() @ x
    assert np.linalg.norm(y[:, 0] - yd) / np.linalg.norm(yd) < 1e-11


def _correction_entries(n, rng):
    """Scattered entries (most land in butterflied blocks) plus a cluster with several entries per row."""
    r = rng.integers(0, n, 1500)
    c = rng.integers(0, n, 1500)
    r2 = np.repeat(np.arange(100, 140), 3)
    c2 = 1500 + rng.integers(0, 40, len(r2))
    rows, cols = np.concatenate([r, r2]), np.concatenate([c, c2])
    return rows, cols, rng.standard_normal(len(rows)) + 1j * rng.standard_normal(len(rows))


def test_kapur_rokhlin_corrected_operator_and_sparse_sum_term():
    """The corrected system matrix of examples/multiple_scattering (multiple_scattering_context.c:456-487):
    bfHelm2ApplyKrCorrectionTree edits the near-field leaves in place; a MatSum [operator, MatCooComplex]
    carries sparse entries next to butterflied blocks (src/mat_block_dense.c:493-499).  Reference
    bfMatMul vs flattener + port oracle + compiled tables; then the same sparse term with its entries
    summed (BF_CUDA_FLATTEN_COO_ACCUMULATE) against dense arithmetic."""
    n = 2048
    h = reflib.Helm2(reflib.circle_points(n), 50.0)
    try:
        x = reflib.randn_complex(n, 0)
        y_plain, _ = h.matmul(x)
        h.apply_kr_correction(6)
        y_kr, _ = h.matmul(x)
        assert np.linalg.norm(y_kr - y_plain) > 1e-3 * np.linalg.norm(y_plain)   # the correction did something
        raw_kr = h.flatten()
        y0 = portlib.apply(raw_kr, x[:, None])
        assert np.linalg.norm(y0 - y_kr) / np.linalg.norm(y_kr) < 1e-13

        rows, cols, vals = _correction_entries(n, np.random.default_rng(5))
        h.sum_with_coo(rows, cols, vals)
        y_ref, _ = h.matmul(x)                       # bfMatSumMul -> bfMatCooComplexMul (last entry of a row wins)
        raw = h.flatten()
        assert raw.num_ops > raw_kr.num_ops
        assert np.linalg.norm(portlib.apply(raw, x[:, None]) - y_ref) / np.linalg.norm(y_ref) < 1e-13
        t = Tables(raw, capi.OP_N, 148 * 12)
        try:
            t.check_queues()
            assert np.linalg.norm(t.apply(x) - y_ref) / np.linalg.norm(y_ref) < 1e-13
        finally:
            t.close()

        raw_acc = h.flatten(flags=1)                 # BF_CUDA_FLATTEN_COO_ACCUMULATE
        D = np.zeros((n, n), np.complex128)
        np.add.at(D, (rows, cols), vals)
        want = y0 + (D @ x)[:, None]
        assert np.linalg.norm(portlib.apply(raw_acc, x[:, None]) - want) / np.linalg.norm(want) < 1e-13
    finally:
        h.close()


# ---- real (streamed) path: live bfFacStreamer factorization --------------------------------------
@pytest.fixture(scope="module")
def stream():
    return reflib.sphere_stream(600, 11)


def test_stream_flatten_matches_golden_structure(stream):
    io = np.load(G / "lbo_stream_n600_io.npz")
    raw = stream.flatten()
    assert np.array_equal(raw.fingerprint(), io["fingerprint"]), "flattened index tables changed"
    assert np.array_equal(stream.row_perm(), io["row_perm"])
    y, _ = stream.mulvec(io["x"])
    z, _ = stream.mulvec(io["v"], transposed=True)
    assert np.linalg.norm(y - io["y"]) <= 1e-14 * np.linalg.norm(io["y"])
    assert np.linalg.norm(z - io["z"]) <= 1e-14 * np.linalg.norm(io["z"])


def test_stream_port_and_tables_match_reference_mulvec_rmulvec(stream):
    rng = np.random.default_rng(5)
    raw = stream.flatten()
    for op, n_in, transposed in ((capi.OP_N, stream.num_cols, False), (capi.OP_T, stream.num_rows, True)):
        x = rng.standard_normal(n_in)
        y_ref, _ = stream.mulvec(x, transposed=transposed)
        y_port = portlib.apply(raw, x[:, None], op)[:, 0]
        assert np.linalg.norm(y_port - y_ref) / np.linalg.norm(y_ref) < 1e-13
        t = Tables(raw, op, 148 * 12)
        try:
            y_tab = t.apply(x)[:, 0]
        finally:
            t.close()
        assert np.linalg.norm(y_tab - y_ref) / np.linalg.norm(y_ref) < 1e-13


def test_stream_vs_dense_band_pin(stream):
    """The streamer's own self-check (src/fac_streamer.c:286-301): butterfly vs the dense band it was
    fed, to about the SVD truncation tolerance (1e-3)."""
    B, _ = reflib.real_sph_harm_band(stream.points, range(12))
    rev = reflib.reverse_perm(stream.row_perm())
    Bp = np.zeros_like(B)
    Bp[rev] = B  # rows into row-tree order: out[rev[i]] = in[i]
    x = np.random.default_rng(1).standard_normal(stream.num_cols)
    y, _ = stream.mulvec(x)
    assert np.abs(y - Bp @ x).max() / np.abs(Bp @ x).max() < 5e-3


def test_sphere_fixture_golden_is_reproducible():
    """The committed lbo_sphere_fixture golden is what the reference produces today from its own test data."""
    s = reflib.sphere_fixture_stream()
    raw = s.flatten()
    io = np.load(G / "lbo_sphere_fixture_io.npz")
    assert np.array_equal(raw.fingerprint(), io["fingerprint"])
    y, _ = s.mulvec(io["x"])
    assert np.linalg.norm(y - io["y"]) / np.linalg.norm(io["y"]) < 1e-14
    assert np.array_equal(s.row_perm(), io["row_perm"])


def test_lbo_fixture_generator_round_trip(tmp_path):
    """tools/make_lbo_fixture.py (the offline producer of the bench's reference-structured real operator): the raw
    ops it writes reproduce, through the port oracle, the reference results it stores beside them."""
    import subprocess
    import sys

    from butterfly_b200.rawops import RawOps

    root = Path(__file__).resolve().parent.parent
    out = tmp_path / "lbo_tiny"
    r = subprocess.run([sys.executable, str(root / "tools" / "make_lbo_fixture.py"), "--n", "600", "--lmax", "11", "--vectors", "2", "--out", str(out)],
                       capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    raw = RawOps.load(str(out) + ".raw.npz")
    ref = np.load(str(out) + ".ref.npz")
    assert raw.dtype == capi.REAL_F64 and raw.num_rows == 600 and raw.num_cols == 144
    assert float(ref["butterfly_vs_dense_rel_l2"]) < 5e-3  # the streamer's own tolerance (1e-3) on the whole matrix
    y = portlib.apply(raw, ref["X"], capi.OP_N)
    yt = portlib.apply(raw, ref["XT"], capi.OP_T)
    assert np.linalg.norm(y - ref["Y"]) <= 1e-13 * np.linalg.norm(ref["Y"])
    assert np.linalg.norm(yt - ref["YT"]) <= 1e-13 * np.linalg.norm(ref["YT"])
