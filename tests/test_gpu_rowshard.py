"""GPU (one device): block-row sharding through the real kernels.  All ranks' shares are compiled in
this process (leaves of the other ranks flagged BF_CUDA_LEAF_REMOTE), run stage by stage with
bfCudaApplyStages on NaN-filled workspaces, and the exchanged rows are moved with tensor copies in
place of all_to_all_single (the collective itself is covered by the gloo tests and bench.py --shard
rows on >= 2 GPUs).  The assembled Y must match the oracle."""
import numpy as np
import pytest
import torch

from butterfly_b200 import capi, rowshard, synth
from butterfly_b200.rawops import RawOps
from oracle import portlib

pytestmark = pytest.mark.gpu


class NanWsBackend(rowshard.GpuBackend):
    def begin(self, x, y):
        ws = super().begin(x, y)
        ws.fill_(float("nan"))
        return ws


@pytest.mark.parametrize("world,nrhs", [(2, 1), (3, 1), (3, 4), (5, 33)])
def test_emulated_ranks_match_oracle(world, nrhs):
    raw = synth.random_hodbf(n=3072, dtype=capi.COMPLEX_F64, seed=9, exotic=True)
    rng = np.random.default_rng(4)
    x = rng.standard_normal((raw.num_cols, nrhs)) + 1j * rng.standard_normal((raw.num_cols, nrhs))
    y_ref = portlib.apply(raw, x, capi.OP_N)
    shares = rowshard.RowShardedOperator.build_all_local(raw, world, NanWsBackend)
    xt = torch.from_numpy(x).cuda()
    ys = [torch.full((raw.num_rows, nrhs), float("nan"), dtype=torch.complex128, device="cuda") for _ in range(world)]
    rowshard.emulate_apply(shares, xt, ys)
    torch.cuda.synchronize()
    y = np.full_like(y_ref, np.nan)
    for sh, yr in zip(shares, ys):
        r0, r1 = sh.y_range
        y[r0:r1] = yr[r0:r1].cpu().numpy()
    assert np.linalg.norm(y - y_ref) / np.linalg.norm(y_ref) <= 1e-12
    assert sum(sh.local_entries for sh in shares) == raw.num_entries
    if world > 2:
        assert sum(sh.exchanged_rows for sh in shares) > 0
    for sh in shares:
        sh.backend.close()


def test_real_streamed_operator_sharded():
    from pathlib import Path

    G = Path(__file__).resolve().parent / "golden"
    raw = RawOps.load(G / "lbo_stream_n600.npz")
    io = np.load(G / "lbo_stream_n600_io.npz")
    shares = rowshard.RowShardedOperator.build_all_local(raw, 3, NanWsBackend)
    xt = torch.from_numpy(io["x"][:, None].copy()).cuda()
    ys = [torch.full((raw.num_rows, 1), float("nan"), dtype=torch.float64, device="cuda") for _ in range(3)]
    rowshard.emulate_apply(shares, xt, ys)
    y = np.concatenate([yr[sh.y_range[0]:sh.y_range[1], 0].cpu().numpy() for sh, yr in zip(shares, ys)])
    assert np.linalg.norm(y - io["y"]) / np.linalg.norm(io["y"]) <= 1e-12
    for sh in shares:
        sh.backend.close()
