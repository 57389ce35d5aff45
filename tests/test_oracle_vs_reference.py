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
