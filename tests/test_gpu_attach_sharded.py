"""GPU: bfCudaMatAttachSharded — the reference's own bfMatMul / bfMatMulVec / bfSolveGMRES on an object whose
factors live on SEVERAL GPUs (block rows: peer stores between the stage kernels; RHS columns: replicas), against
the same calls on the unattached object (reference CPU code, oracle/_ref).  One host thread drives all GPUs.  On a
one-GPU box the ranks share device 0 (same kernels, same protocol)."""
import numpy as np
import pytest

from butterfly_b200 import capi
from butterfly_b200.sharded import Comm
from oracle import reflib

pytestmark = [pytest.mark.gpu, pytest.mark.skipif(not reflib.attach_available(), reason="oracle/_ref + attach layer not built")]
TOL = 1e-12


def _comm(world):
    import torch

    n = torch.cuda.device_count()
    return Comm(list(range(world)) if n >= world else [0] * world)


def rel(a, b):
    return np.linalg.norm(a - b) / np.linalg.norm(b)


@pytest.mark.parametrize("mode", [capi.SHARD_BLOCK_ROWS, capi.SHARD_RHS_COLS], ids=["block_rows", "rhs_cols"])
@pytest.mark.parametrize("world", [2, 3])
def test_bfmatmul_on_a_sharded_attachment(world, mode):
    n = 4096
    h = reflib.Helm2(reflib.circle_points(n), 100.0)
    comm = _comm(world)
    try:
        x = reflib.randn_complex(n, 0)
        X = np.stack([x, np.conj(x), 1j * np.roll(x, 3), x * x, np.abs(x) + 0j], axis=1)
        y_cpu, _ = h.matmul(x)
        Y_cpu, _ = h.matmul(X)
        h.attach_sharded(comm, mode)
        y_gpu, _ = h.matmul(x)
        Y_gpu, _ = h.matmul(X)
        y_vec = h.mulvec_complex(x)
        h.detach()
        y_back, _ = h.matmul(x)
        assert rel(y_gpu, y_cpu) <= TOL and rel(y_vec, y_cpu[:, 0]) <= TOL
        for c in range(X.shape[1]):
            assert rel(Y_gpu[:, c], Y_cpu[:, c]) <= TOL
        assert np.array_equal(y_back, y_cpu)
    finally:
        h.close()
        comm.close()


def test_reference_gmres_drives_the_sharded_operator():
    """bfSolveGMRES (src/linalg.c:47) itself, unchanged, on K diag(w) + I / 2 — built by the reference AFTER the
    attach (ScaleCols + AddInplace mark the snapshot stale; the next product re-shards the edited operator)."""
    n = 2048
    h = reflib.Helm2(reflib.circle_points(n), 50.0)
    comm = _comm(2)
    try:
        h.attach_sharded(comm, capi.SHARD_BLOCK_ROWS)
        h.make_second_kind()
        b = reflib.randn_complex(n, 1)
        x_gpu, it_gpu, _ = h.solve_gmres(b, 1e-12, 50)
        h.detach()
        x_cpu, it_cpu, _ = h.solve_gmres(b, 1e-12, 50)
        assert it_gpu == it_cpu
        assert rel(x_gpu, x_cpu) <= 1e-10
    finally:
        h.close()
        comm.close()


def test_real_mulvec_on_a_sharded_attachment():
    s = reflib.sphere_stream(1500, 15)
    comm = _comm(2)
    rng = np.random.default_rng(3)
    x, v = rng.standard_normal(s.num_cols), rng.standard_normal(s.num_rows)
    y_cpu, _ = s.mulvec(x)
    z_cpu, _ = s.mulvec(v, transposed=True)
    s.attach_sharded(comm, capi.SHARD_BLOCK_ROWS)
    y_rows, _ = s.mulvec(x)
    s.detach()
    s.attach_sharded(comm, capi.SHARD_RHS_COLS, build_transpose=1)
    y_cols, _ = s.mulvec(x)
    z_cols, _ = s.mulvec(v, transposed=True)
    s.detach()
    comm.close()
    assert rel(y_rows, y_cpu) <= TOL and rel(y_cols, y_cpu) <= TOL and rel(z_cols, z_cpu) <= TOL
