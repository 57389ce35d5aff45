"""GPU: persistence of the COMPILED plan (bfCudaPlanSave / bfCudaPlanLoad / bfCudaShardRankLoad, csrc/bf_planio.cu;
SURVEY.md §8f-1 — the reference has Dump writers without a reader, src/mat.c:67-73).  save -> destroy -> load ->
apply must be BIT-identical: the file holds the tables and the pool bytes the kernels consumed."""
import ctypes as C

import numpy as np
import pytest

from butterfly_b200 import capi, synth
from butterfly_b200.plan import Plan
from oracle import portlib

pytestmark = pytest.mark.gpu


def _x(rng, n, nrhs, cplx):
    x = rng.standard_normal((n, nrhs))
    return x + 1j * rng.standard_normal((n, nrhs)) if cplx else x


@pytest.mark.parametrize("dtype", [capi.COMPLEX_F64, capi.REAL_F64], ids=["c128", "f64"])
def test_save_load_is_bit_identical_all_directions(dtype, tmp_path):
    raw = synth.random_hodbf(n=3000, dtype=dtype, seed=3)
    cplx = dtype == capi.COMPLEX_F64
    rng = np.random.default_rng(1)
    perm_in, perm_out = rng.permutation(raw.num_cols), rng.permutation(raw.num_rows)
    ops = [capi.OP_N, capi.OP_T] + ([capi.OP_H] if cplx else [])
    xs = {(op, k): _x(rng, raw.num_cols if op == capi.OP_N else raw.num_rows, k, cplx) for op in ops for k in (1, 5)}
    path = tmp_path / "plan.bfplan"
    with Plan(raw, build_transpose=3 if cplx else 1) as plan:
        plan.set_perms(perm_in, perm_out)
        want = {key: plan.apply_host(x, key[0]).copy() for key, x in xs.items()}
        info = plan.info
        plan.save(path)
    assert path.stat().st_size > info["poolBytes"]
    with Plan.load(path) as back:
        for k2 in ("numRows", "numCols", "numOps", "numEntries", "numStagesN", "numStagesT", "numItemsN", "numTilesN", "poolBytes", "workspaceRowsN"):
            assert back.info[k2] == info[k2], k2
        for key, x in xs.items():
            assert np.array_equal(back.apply_host(x, key[0]), want[key]), key
        # the single-launch tables travelled too
        back.set_fused(True)
        assert back.is_fused(capi.OP_N)
        assert np.array_equal(back.apply_host(xs[(capi.OP_N, 1)]), want[(capi.OP_N, 1)])
        with pytest.raises(capi.BfCudaError) as e:
            back.fill_synthetic(1)
        assert e.value.status == 2  # NOT_IMPLEMENTED: no per-leaf data behind a loaded plan
    # against the oracle as well (perms: Xp[inPerm[i]] = X[i]; Yp[outPerm[i]] = Y[i])
    x = xs[(capi.OP_N, 1)]
    xp = np.empty_like(x); xp[perm_in] = x
    yp = np.empty_like(want[(capi.OP_N, 1)]); yp[perm_out] = portlib.apply(raw, xp, capi.OP_N)
    assert np.linalg.norm(want[(capi.OP_N, 1)] - yp) <= 1e-12 * np.linalg.norm(yp)


def test_values_that_only_exist_on_the_device(tmp_path):
    """A value-less structure filled ON the GPU (the bench operators, bfCudaPlanBuildHelm2 / FillSynthetic): nothing
    on the host could rebuild these bytes, the file must carry them."""
    raw = synth.random_hodbf(n=4096, dtype=capi.COMPLEX_F64, seed=7, exotic=False, leaf=64, rank_lo=16, rank_hi=40, with_values=False)
    x = _x(np.random.default_rng(2), raw.num_cols, 3, True)
    with Plan(raw) as plan:
        plan.fill_synthetic(99)
        want = plan.apply_host(x).copy()
        plan.save(tmp_path / "p.bfplan")
    assert np.abs(want).max() > 0
    with Plan.load(tmp_path / "p.bfplan") as back:
        assert np.array_equal(back.apply_host(x), want)


def test_damaged_files_are_refused(tmp_path, lib):
    raw = synth.random_rect(333, 517, capi.COMPLEX_F64, seed=6)
    with Plan(raw) as plan:
        plan.save(tmp_path / "ok.bfplan")
    blob = (tmp_path / "ok.bfplan").read_bytes()
    h = C.c_void_p()
    cases = {"truncated_pool": blob[:-4096], "truncated_tables": blob[:200], "trailing": blob + b"\0" * 16,
             "magic": b"BFPLAN01" + blob[8:], "empty": b""}
    forged = bytearray(blob)  # a tile's operand address pushed far outside the operand: must be caught by the bounds check
    # (find the first stage's tile array by scanning for a plausible spot is brittle; corrupt a wide band of the table area instead)
    for off in range(400, min(len(forged), 4000), 8):
        forged[off + 3] = 0x3F
    cases["forged_tables"] = bytes(forged)
    for name, data in cases.items():
        (tmp_path / "bad.bfplan").write_bytes(data)
        st = lib.bfCudaPlanLoad(str(tmp_path / "bad.bfplan").encode(), -1, C.byref(h))
        assert st == 6, (name, st)  # BF_CUDA_ERROR_IO
    assert lib.bfCudaPlanLoad(str(tmp_path / "missing").encode(), -1, C.byref(h)) == 6


def test_sharded_rank_plans_round_trip(tmp_path, lib):
    """Two ranks of a block-row sharded operator (both on GPU 0, one stream): build, apply, save each rank's plan,
    destroy everything, load the ranks back, connect, apply: identical rows."""
    import torch

    raw = synth.random_hodbf(n=3000, dtype=capi.COMPLEX_F64, seed=3)
    world = 2
    d = raw.desc()
    sd = C.c_void_p()
    fr = (C.c_double * world)(0.4, 0.6)
    capi.check(lib.bfCudaShardDescsCreate(C.byref(d), world, fr, C.byref(sd)), "bfCudaShardDescsCreate")
    cuts = (C.c_int64 * (world + 1))()
    capi.check(lib.bfCudaShardDescsGetLayout(sd, None, cuts, None, None, None), "layout")
    x = torch.from_numpy(_x(np.random.default_rng(4), raw.num_cols, 1, True)).cuda()
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream or 1)

    def connect_and_apply(ranks):
        hs = (capi.ShardHandle * world)()
        for r in range(world):
            capi.check(lib.bfCudaShardRankGetHandle(ranks[r], C.byref(hs[r])), "handle")
        for r in range(world):
            capi.check(lib.bfCudaShardRankConnect(ranks[r], hs), "connect")
        out = []
        for rep in range(2):
            ys = [torch.zeros(raw.num_rows, 1, dtype=torch.complex128, device="cuda") for _ in range(world)]
            nS = lib.bfCudaPlanGetFused  # noqa: F841 (keep lib alive in closures)
            info = capi.PlanInfo()
            capi.check(lib.bfCudaPlanGetInfo(lib.bfCudaShardRankGetPlan(ranks[0]), C.byref(info)), "info")
            for s in range(info.numStagesN):  # stage-major: ranks share the GPU and the stream
                for r in range(world):
                    capi.check(lib.bfCudaShardRankApplyStages(ranks[r], C.c_void_p(x.data_ptr()), C.c_void_p(ys[r].data_ptr()), s, s + 1, stream), "apply")
            torch.cuda.synchronize()
            out.append(np.concatenate([ys[r][cuts[r]:cuts[r + 1]].cpu().numpy() for r in range(world)]))
        assert np.array_equal(out[0], out[1])
        return out[0]

    ranks = []
    for r in range(world):
        h = C.c_void_p()
        capi.check(lib.bfCudaShardRankCreate(lib.bfCudaShardDescsGet(sd, r), C.byref(h)), "bfCudaShardRankCreate")
        ranks.append(h)
    lib.bfCudaShardDescsFree(sd)
    y0 = connect_and_apply(ranks)
    y_ref = portlib.apply(raw, x.cpu().numpy(), capi.OP_N)
    assert np.linalg.norm(y0 - y_ref) <= 1e-12 * np.linalg.norm(y_ref)
    for r in range(world):
        capi.check(lib.bfCudaPlanSave(lib.bfCudaShardRankGetPlan(ranks[r]), str(tmp_path / f"rank{r}.bfplan").encode()), "bfCudaPlanSave")
    for h in ranks:
        lib.bfCudaShardRankDestroy(h)
    loaded = []
    for r in range(world):
        h = C.c_void_p()
        capi.check(lib.bfCudaShardRankLoad(str(tmp_path / f"rank{r}.bfplan").encode(), -1, C.byref(h)), "bfCudaShardRankLoad")
        loaded.append(h)
    y1 = connect_and_apply(loaded)
    for h in loaded:
        lib.bfCudaShardRankDestroy(h)
    assert np.array_equal(y1, y0)
