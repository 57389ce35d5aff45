"""CPU: block-row sharding behind the C ABI (csrc/bf_shard.cpp, the shard half of csrc/bf_plan.cpp).

1. bfCudaShardDescsCreate decides who applies which leaf; the Python restatement butterfly_b200/rowshard.shard
   (round 1, exercised by the gloo tests) must agree leaf for leaf: same cuts, same owners, same Y ranges.
2. Every rank's descriptor is compiled (bfCudaDebugCompile, no GPU) and ALL ranks are run together by the numpy
   table interpreter the way the stage kernels run: stage by stage, each rank on its OWN NaN-initialised flat
   workspace, an item's rows copied into the workspaces of the ranks in its itemPeers mask (the kernel's peer
   stores).  A row some rank reads but nobody pushed stays NaN and poisons the result; the assembled Y (every
   rank its own rows) must match the oracle.  The per-stage receive masks — what a stage kernel waits for — must
   cover every push."""
import ctypes as C

import numpy as np
import pytest

from butterfly_b200 import capi, rowshard, synth
from butterfly_b200.rawops import RawOps
from oracle import portlib
from table_interp import Tables


def shard_descs(raw, world, fractions=None):
    lib = capi.load_library()
    d = raw.desc()
    h = C.c_void_p()
    fr = None if fractions is None else (C.c_double * world)(*fractions)
    capi.check(lib.bfCudaShardDescsCreate(C.byref(d), world, fr, C.byref(h)), "bfCudaShardDescsCreate")
    n = C.c_int64()
    capi.check(lib.bfCudaShardDescsGetLayout(h, C.byref(n), None, None, None, None), "layout")
    cuts = (C.c_int64 * (world + 1))()
    owner = np.zeros(n.value, np.int32)
    oo, orow = np.zeros(n.value, np.int64), np.zeros(n.value, np.int64)
    capi.check(lib.bfCudaShardDescsGetLayout(h, None, cuts, owner.ctypes.data_as(C.POINTER(C.c_int32)),
                                             oo.ctypes.data_as(C.POINTER(C.c_int64)), orow.ctypes.data_as(C.POINTER(C.c_int64))), "layout")
    ranks = [RawOps.from_desc(lib.bfCudaShardDescsGet(h, r).contents) for r in range(world)]
    lib.bfCudaShardDescsFree(h)
    return ranks, list(cuts), owner, oo, orow


CASES = {
    "hodbf_c": lambda: synth.random_hodbf(n=1500, dtype=capi.COMPLEX_F64, seed=3),
    "hodbf_r": lambda: synth.random_hodbf(n=1100, dtype=capi.REAL_F64, seed=4),
    "rect_r": lambda: synth.random_rect(700, 300, capi.REAL_F64, seed=5),
    "shifted": lambda: synth.shifted_identity_hodbf(1024, seed=5),
}


@pytest.mark.parametrize("name", list(CASES))
@pytest.mark.parametrize("world,fractions", [(2, None), (3, None), (2, [0.37, 0.63]), (4, [1, 2, 1, 3])])
def test_native_ownership_matches_the_python_restatement(name, world, fractions):
    raw = CASES[name]()
    ranks, cuts, owner, oo, orow = shard_descs(raw, world, fractions)
    raw_py, owner_py, cuts_py = rowshard.shard(raw, world, fractions)
    assert cuts == [int(c) for c in cuts_py]
    assert np.array_equal(owner, owner_py)
    fp = raw_py.fingerprint()
    oi, orr = rowshard.origin_of(raw_py)
    assert np.array_equal(oo, oi) and np.array_equal(orow, orr)
    for r in range(world):
        got = ranks[r].fingerprint()
        want = fp.copy()
        want[:, 1] |= (owner_py.astype(np.int64) << 8) | np.where(owner_py != r, 8, 0)  # owner bits + REMOTE
        assert np.array_equal(got, want)


@pytest.mark.parametrize("name", list(CASES))
@pytest.mark.parametrize("world,fractions", [(2, None), (3, None), (2, [0.37, 0.63])])
def test_all_ranks_emulated_with_peer_pushes(name, world, fractions):
    raw = CASES[name]()
    cplx = raw.dtype == capi.COMPLEX_F64
    ranks, cuts, owner, _, _ = shard_descs(raw, world, fractions)
    rng = np.random.default_rng(5)
    x = rng.standard_normal((raw.num_cols, 2)) + (1j * rng.standard_normal((raw.num_cols, 2)) if cplx else 0)
    y_ref = portlib.apply(raw, x, capi.OP_N)
    tabs = [Tables(ranks[r], capi.OP_N, num_queues=19, shard=(world, r)) for r in range(world)]
    try:
        nS = tabs[0].num_stages
        assert all(t.num_stages == nS and t.ws_rows == tabs[0].ws_rows for t in tabs), "ranks disagree on the structure"
        assert all(not t.fused for t in tabs), "a shard has no single-launch tables"
        spaces = [t.new_space(x) for t in tabs]
        pushed_rows = 0
        for s in range(nS):
            for r, t in enumerate(tabs):
                t.run_stage(spaces[r], s)
            for r, t in enumerate(tabs):  # the peer stores of stage s
                i0, i1 = t.stage_item[s], t.stage_item[s + 1]
                seen = 0
                for it, peers in zip(t.items[i0:i1], t.item_peers[i0:i1]):
                    if not peers:
                        continue
                    assert int(it["outAddr"]) >> 30 == capi.SPACE_W and not (int(peers) >> r) & 1
                    row, rows = int(it["outAddr"]) & capi.ROW_MASK, int(it["rows"])
                    for q in range(world):
                        if (int(peers) >> q) & 1:
                            spaces[q][2][row:row + rows] = spaces[r][2][row:row + rows]
                            assert (int(tabs[q].recv_mask[s]) >> r) & 1, "a push the reader does not wait for"
                            pushed_rows += rows
                    seen |= int(peers)
                assert seen == int(t.send_mask[s])
        # the two sides of the hand-over are derived independently (sender: items -> readers; reader: remote leaves):
        # rank q waits for rank r's flag of stage s exactly when r posts one to q
        for s in range(nS):
            for r in range(world):
                for q in range(world):
                    assert ((int(tabs[r].send_mask[s]) >> q) & 1) == ((int(tabs[q].recv_mask[s]) >> r) & 1), (s, r, q)
        y = np.full_like(y_ref, np.nan)
        for r in range(world):
            y[cuts[r]:cuts[r + 1]] = spaces[r][1][cuts[r]:cuts[r + 1]]
        assert not np.isnan(y).any(), "a rank read a row nobody pushed (or a result row has no owner)"
        assert np.linalg.norm(y - y_ref) <= 1e-13 * np.linalg.norm(y_ref)
        if fractions is not None and name.startswith("hodbf"):
            assert pushed_rows > 0  # unequal shares cut through the butterflies: the exchange is exercised
    finally:
        for t in tabs:
            t.close()
