"""CPU, gloo, world_size 2: the host logic of the RHS-column sharding.  Each rank
applies the operator to its slab of columns (the table interpreter stands in for the GPU plan: same
tables, same order), slabs are gathered, and the result must equal the unsharded apply (columns are
independent; the interpreter's BLAS may pick a different summation order per block width, hence
1e-14 rather than bitwise) and the oracle (1e-12)."""
import os
import socket
import sys
from pathlib import Path

import numpy as np
import pytest

ROOT = Path(__file__).resolve().parent.parent


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, nrhs, q):
    try:
        sys.path.insert(0, str(ROOT))
        sys.path.insert(0, str(ROOT / "tests"))
        os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
        import torch
        import torch.distributed as dist

        from butterfly_b200 import capi, shard, synth
        from oracle import portlib
        from table_interp import Tables

        dist.init_process_group("gloo", rank=rank, world_size=world)
        raw = synth.random_hodbf(n=512, dtype=capi.COMPLEX_F64, seed=3)
        rng = np.random.default_rng(11)  # same block on every rank
        x = rng.standard_normal((raw.num_cols, nrhs)) + 1j * rng.standard_normal((raw.num_cols, nrhs))
        T = Tables(raw)
        c0, c1, y_loc = shard.apply_column_shard(lambda xs: T.apply(np.ascontiguousarray(xs)), x, world, rank)
        if y_loc is None:
            y_loc = np.zeros((raw.num_rows, 0), np.complex128)
        y = shard.gather_columns(torch.from_numpy(np.ascontiguousarray(y_loc)), raw.num_rows, nrhs).numpy()
        y_full = T.apply(x)
        y_ref = portlib.apply(raw, x, capi.OP_N)
        ok_bits = bool(np.linalg.norm(y - y_full) <= 1e-14 * np.linalg.norm(y_full))
        err = float(np.linalg.norm(y - y_ref) / np.linalg.norm(y_ref))
        dist.barrier()
        dist.destroy_process_group()
        q.put((rank, c0, c1, ok_bits, err))
    except Exception as e:  # pragma: no cover
        q.put((rank, -1, -1, False, repr(e)))


@pytest.mark.parametrize("nrhs", [1, 5, 8])
def test_rhs_column_sharding_two_ranks(nrhs):
    import torch.multiprocessing as mp

    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, nrhs, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=240) for _ in procs)
    for p in procs:
        p.join(timeout=60)
    assert [r[0] for r in res] == [0, 1]
    for rank, c0, c1, ok_bits, err in res:
        assert ok_bits, (rank, err)
        assert err <= 1e-12
    # the slabs tile [0, nrhs) exactly
    assert res[0][1] == 0 and res[0][2] == res[1][1] and res[1][2] == nrhs


def test_column_range_partitions():
    from butterfly_b200.shard import column_range

    for nrhs in (0, 1, 7, 256, 257):
        for world in (1, 2, 3, 8):
            spans = [column_range(nrhs, world, r) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == nrhs
            assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
            widths = [b - a for a, b in spans]
            assert max(widths) - min(widths) <= 1
