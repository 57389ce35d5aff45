"""GPU: one operator over several GPUs through the C ABI (bfCudaCommInit / bfCudaPlanShard /
bfCudaShardedApplyHost; csrc/bf_comm.cu, bf_shard.cpp, the MODE-2 stage kernel).

On a one-GPU box the ranks SHARE device 0 (bfCudaCommInit accepts a repeated ordinal): the real stage kernels, the
real peer stores (into another rank's workspace on the same device) and the real flag hand-over run, the kernels of
all ranks one after the other on one stream.  Where two or more GPUs are visible the same cases run with one rank
per GPU over NVLink peer memory.  Against the oracle port, rel. l2 <= 1e-12 (north star)."""
import numpy as np
import pytest

from butterfly_b200 import capi, synth
from butterfly_b200.plan import Plan
from butterfly_b200.sharded import Comm, ShardedPlan
from oracle import portlib

pytestmark = pytest.mark.gpu
TOL = 1e-12


def _gpus():
    import torch

    return torch.cuda.device_count()


def _devices(world, spread):
    if not spread:
        return [0] * world
    if _gpus() < world:
        pytest.skip(f"needs {world} GPUs")
    return list(range(world))


def _x(rng, n, nrhs, cplx):
    x = rng.standard_normal((n, nrhs))
    return x + 1j * rng.standard_normal((n, nrhs)) if cplx else x


CASES = {
    "hodbf_c": lambda: synth.random_hodbf(n=3000, dtype=capi.COMPLEX_F64, seed=3),
    "hodbf_r": lambda: synth.random_hodbf(n=2100, dtype=capi.REAL_F64, seed=4),
    "plain_c": lambda: synth.random_hodbf(n=4096, dtype=capi.COMPLEX_F64, seed=7, exotic=False, leaf=64, rank_lo=16, rank_hi=40),
}


@pytest.mark.parametrize("spread", [False, True], ids=["one_gpu", "multi_gpu"])
@pytest.mark.parametrize("world,fractions", [(2, None), (3, [0.2, 0.5, 0.3]), (2, [0.37, 0.63]), (4, None)])
@pytest.mark.parametrize("name", list(CASES))
def test_block_rows_match_the_oracle(name, world, fractions, spread):
    raw = CASES[name]()
    cplx = raw.dtype == capi.COMPLEX_F64
    comm = Comm(_devices(world, spread))
    rng = np.random.default_rng(3)
    with ShardedPlan(raw, comm, capi.SHARD_BLOCK_ROWS, fractions) as sp:
        info = sp.info
        assert info["world"] == world and info["numEntries"] == raw.num_entries
        assert info["maxRankEntries"] < raw.num_entries  # every rank stores only a share of the factors
        if fractions is not None:
            assert info["sentRowsPerApply"] > 0  # unequal shares cut through butterflies: rows cross ranks
        for rep in range(4):  # consecutive applies: the flags carry the apply count
            x = _x(rng, raw.num_cols, 1, cplx)
            y = sp.apply_host(x)
            y_ref = portlib.apply(raw, x, capi.OP_N)
            assert np.linalg.norm(y - y_ref) <= TOL * np.linalg.norm(y_ref), rep
        X = _x(rng, raw.num_cols, 3, cplx)  # a block goes column by column
        Y = sp.apply_host(X)
        Y_ref = portlib.apply(raw, X, capi.OP_N)
        for c in range(3):
            assert np.linalg.norm(Y[:, c] - Y_ref[:, c]) <= TOL * np.linalg.norm(Y_ref[:, c])
        with pytest.raises(capi.BfCudaError):
            sp.apply_host(np.zeros(raw.num_rows, sp.np_dtype), capi.OP_T)  # forward only
    comm.close()


@pytest.mark.parametrize("spread", [False, True], ids=["one_gpu", "multi_gpu"])
def test_block_rows_device_resident_and_deterministic(spread):
    import torch

    raw = CASES["hodbf_c"]()
    world = 2
    devs = _devices(world, spread)
    comm = Comm(devs)
    with ShardedPlan(raw, comm, capi.SHARD_BLOCK_ROWS, [0.45, 0.55]) as sp, Plan(raw) as full:
        x = _x(np.random.default_rng(9), raw.num_cols, 1, True)
        xs = [torch.from_numpy(x).to(f"cuda:{d}") for d in devs]
        ys = [torch.full((raw.num_rows, 1), float("nan"), dtype=torch.complex128, device=f"cuda:{d}") for d in devs]
        torch.cuda.synchronize()
        outs = []
        for _ in range(3):
            sp.apply_device(xs, ys)
            sp.synchronize()
            cuts = sp.info["yCuts"]
            outs.append(np.concatenate([ys[r][cuts[r]:cuts[r + 1]].cpu().numpy() for r in range(world)]))
        assert np.array_equal(outs[0], outs[1]) and np.array_equal(outs[1], outs[2])  # no atomics: run-to-run identical
        y_full = full.apply_host(x)
        assert np.linalg.norm(outs[0] - y_full) <= 1e-14 * np.linalg.norm(y_full)  # same factors as the unsharded plan
        for r in range(world):  # a rank writes its own rows of the result only
            other = np.ones(raw.num_rows, bool)
            other[cuts[r]:cuts[r + 1]] = False
            assert torch.isnan(ys[r][torch.from_numpy(other).to(ys[r].device)].real).all()
    comm.close()


@pytest.mark.parametrize("spread", [False, True], ids=["one_gpu", "multi_gpu"])
@pytest.mark.parametrize("name", ["hodbf_c", "hodbf_r"])
def test_rhs_columns_replicated_plans(name, spread):
    raw = CASES[name]()
    cplx = raw.dtype == capi.COMPLEX_F64
    comm = Comm(_devices(2, spread))
    with ShardedPlan(raw, comm, capi.SHARD_RHS_COLS, build_transpose=1) as sp:
        for nrhs in (1, 5, 64):
            X = _x(np.random.default_rng(nrhs), raw.num_cols, nrhs, cplx)
            Y = sp.apply_host(X)
            Y_ref = portlib.apply(raw, X, capi.OP_N)
            for c in range(0, nrhs, max(1, nrhs // 4)):
                assert np.linalg.norm(Y[:, c] - Y_ref[:, c]) <= TOL * np.linalg.norm(Y_ref[:, c])
        Z = _x(np.random.default_rng(2), raw.num_rows, 4, cplx)
        assert np.linalg.norm(sp.apply_host(Z, capi.OP_T) - portlib.apply(raw, Z, capi.OP_T)) <= TOL * np.linalg.norm(Z) * 30
    comm.close()


def test_single_rank_shard_is_an_ordinary_plan():
    raw = CASES["hodbf_r"]()
    comm = Comm([0])
    with ShardedPlan(raw, comm, capi.SHARD_BLOCK_ROWS) as sp:
        x = _x(np.random.default_rng(1), raw.num_cols, 2, False)
        y_ref = portlib.apply(raw, x, capi.OP_N)
        assert np.linalg.norm(sp.apply_host(x) - y_ref) <= TOL * np.linalg.norm(y_ref)
    comm.close()
