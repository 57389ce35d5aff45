"""CPU: pin the oracle (oracle/plan_oracle.c) and the plan compiler against golden vectors that
the UNMODIFIED reference produced (oracle/make_golden.py; bfMatMul, bfMatPermuteRows,
bfPermGetReversePerm of the reference build).  Needs no reference at run time."""
from pathlib import Path

import numpy as np
import pytest

from butterfly_b200 import capi
from butterfly_b200.rawops import RawOps
from oracle import portlib
from table_interp import Tables

G = Path(__file__).resolve().parent / "golden"


@pytest.fixture(scope="module")
def block():
    raw = RawOps.load(G / "helm2_block.npz")
    io = np.load(G / "helm2_block_io.npz")
    return raw, io["x"], io["y"]


def test_port_oracle_reproduces_reference_bfmatmul(block):
    raw, x, y = block
    assert raw.dtype == capi.COMPLEX_F64 and raw.num_ops == 11 and raw.num_entries == 11184
    yo = portlib.apply(raw, x)
    assert np.linalg.norm(yo - y) / np.linalg.norm(y) < 1e-13


def test_compiled_tables_reproduce_reference_bfmatmul(block):
    raw, x, y = block
    t = Tables(raw, capi.OP_N, 148 * 8)
    try:
        t.check_queues()
        yt = t.apply(x)[:, 0]
    finally:
        t.close()
    assert np.linalg.norm(yt - y) / np.linalg.norm(y) < 1e-13
    assert t.num_stages == 3  # BlockDiag, BlockCoo, BlockDiag (src/fac_helm2.c:692-695)


def test_permutations_bit_exact_with_reference():
    m = np.load(G / "helm2_n2048_meta.npz")
    perm, rev = m["perm"], m["rev_perm"]
    # bfPermGetReversePerm (src/perm.c:195-220) is the inverse permutation
    assert np.array_equal(np.argsort(perm), rev)
    # out[perm[i]] = in[i] (src/mat_dense_complex.c:741-774)
    assert np.array_equal(portlib.permute_rows(rev, m["x"]), m["x_permuted"])
    assert np.array_equal(portlib.permute_rows(perm, m["y"]), m["y_permuted"])


def test_raw_ops_file_roundtrip(block, lib, tmp_path):
    import ctypes as C

    raw, x, y = block
    d = raw.desc()
    path = str(tmp_path / "block.bfraw").encode()
    assert lib.bfCudaRawOpsSave(C.byref(d), path) == 0
    out = C.POINTER(capi.PlanDesc)()
    assert lib.bfCudaRawOpsLoad(path, C.byref(out)) == 0
    try:
        back = RawOps.from_desc(out.contents)
    finally:
        lib.bfCudaRawOpsFree(out)
    assert np.array_equal(back.fingerprint(), raw.fingerprint())
    assert np.array_equal(back.buf_rows, raw.buf_rows)
    assert np.array_equal(portlib.apply(back, x), portlib.apply(raw, x))
    assert lib.bfCudaRawOpsLoad(str(tmp_path / "missing").encode(), C.byref(out)) == 6  # BF_CUDA_ERROR_IO
    # truncated and corrupted files are refused, not read out of bounds
    blob = (tmp_path / "block.bfraw").read_bytes()
    for bad in (blob[: len(blob) - 64], blob[: 56 + 8], blob + b"\0" * 8):
        (tmp_path / "bad.bfraw").write_bytes(bad)
        assert lib.bfCudaRawOpsLoad(str(tmp_path / "bad.bfraw").encode(), C.byref(out)) == 6
    ops_at = 56 + 8 * len(raw.buf_rows)            # first RawOp; its `val` offset is the last 8 bytes of 72
    forged = bytearray(blob)
    forged[ops_at + 64 : ops_at + 72] = (2 ** 40).to_bytes(8, "little")
    (tmp_path / "bad.bfraw").write_bytes(bytes(forged))
    assert lib.bfCudaRawOpsLoad(str(tmp_path / "bad.bfraw").encode(), C.byref(out)) == 6
    # a value-less dense leaf (GPU-built operator) would reload as zeros: Save refuses it
    d2 = raw.desc()
    ops = C.cast(d2.ops, C.POINTER(capi.RawOp))
    ops[0].val = None
    assert lib.bfCudaRawOpsSave(C.byref(d2), str(tmp_path / "noval.bfraw").encode()) == 1  # INVALID_ARGUMENT


# ---- real (streamed, LBO-style) path: bfMatMulVec / bfMatRmulVec of the reference ----------------
@pytest.fixture(scope="module")
def stream():
    raw = RawOps.load(G / "lbo_stream_n600.npz")
    io = np.load(G / "lbo_stream_n600_io.npz")
    return raw, io


def test_port_oracle_reproduces_reference_mulvec_and_rmulvec(stream):
    raw, io = stream
    assert raw.dtype == capi.REAL_F64 and (raw.num_rows, raw.num_cols) == (600, 144)
    assert np.array_equal(raw.fingerprint(), io["fingerprint"])
    y = portlib.apply(raw, io["x"][:, None], capi.OP_N)[:, 0]
    z = portlib.apply(raw, io["v"][:, None], capi.OP_T)[:, 0]
    assert np.linalg.norm(y - io["y"]) / np.linalg.norm(io["y"]) < 1e-13   # bfMatMulVec, src/mat_product.c:247
    assert np.linalg.norm(z - io["z"]) / np.linalg.norm(io["z"]) < 1e-13   # bfMatRmulVec, src/mat_product.c:312


@pytest.mark.parametrize("op,xin,yout", [(capi.OP_N, "x", "y"), (capi.OP_T, "v", "z")])
def test_compiled_tables_reproduce_reference_real_path(stream, op, xin, yout):
    raw, io = stream
    t = Tables(raw, op, 148 * 12)
    try:
        t.check_queues()
        y = t.apply(io[xin])[:, 0]
    finally:
        t.close()
    assert np.linalg.norm(y - io[yout]) / np.linalg.norm(io[yout]) < 1e-13


@pytest.fixture(scope="module")
def fixture_stream():
    return RawOps.load(G / "lbo_sphere_fixture.npz"), np.load(G / "lbo_sphere_fixture_io.npz")


@pytest.mark.parametrize("op,xin,yout", [(capi.OP_N, "x", "y"), (capi.OP_T, "v", "z")])
def test_reference_sphere_fixture_factorization(fixture_stream, op, xin, yout):
    """bfFacStreamer run on the reference's own LBO test data (tests/sphere.obj, sphere_{Lam,Phi}.txt,
    the eigenpairs of tests/test_linalg.c:21-22): port oracle and compiled tables against the reference's
    bfMatMulVec / bfMatRmulVec, and both within the streamer's tolerance of the dense eigenvector matrix."""
    raw, io = fixture_stream
    assert raw.dtype == capi.REAL_F64 and (raw.num_rows, raw.num_cols) == (500, 32)
    assert np.array_equal(raw.fingerprint(), io["fingerprint"])
    y = portlib.apply(raw, io[xin][:, None], op)[:, 0]
    assert np.linalg.norm(y - io[yout]) / np.linalg.norm(io[yout]) < 1e-13
    t = Tables(raw, op, 148 * 12)
    try:
        t.check_queues()
        yt = t.apply(io[xin])[:, 0]
    finally:
        t.close()
    assert np.linalg.norm(yt - io[yout]) / np.linalg.norm(io[yout]) < 1e-13
    dense = io[yout + "_dense"]
    assert np.linalg.norm(yt - dense) / np.linalg.norm(dense) < 1e-3   # tol handed to bfFacStreamerInit


def test_streamed_operator_contains_identity_leaves(stream):
    raw, _ = stream
    kinds = set(raw.ops["kind"].tolist())
    assert capi.LEAF_IDENTITY in kinds and capi.LEAF_DENSE in kinds  # pass-through Psi / W blocks, src/fac.c:649-676
