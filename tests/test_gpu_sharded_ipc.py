"""GPU (>= 2 GPUs): the one-process-per-GPU form of block-row sharding — bfCudaShardRank* with the peers'
workspaces and flags mapped through CUDA IPC (what bench.py does under torchrun).  Two ranks are launched with
torch.distributed.run; each writes the rows of the result it owns; together they must match the oracle."""
import subprocess
import sys
from pathlib import Path

import numpy as np
import pytest

from butterfly_b200 import capi, synth
from oracle import portlib

pytestmark = pytest.mark.gpu
ROOT = Path(__file__).resolve().parent.parent


def test_two_processes_two_gpus_ipc_peers(tmp_path):
    import torch

    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs (two processes on ONE GPU are time-sliced: their spinning stage kernels would wait out whole time slices)")
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
                        "--master-port", "29577", str(ROOT / "tests" / "_sharded_ipc_worker.py"), str(tmp_path)],
                       capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, (r.stdout + r.stderr)[-3000:]
    raw = synth.random_hodbf(n=3000, dtype=capi.COMPLEX_F64, seed=3)
    rng = np.random.default_rng(17)
    parts = [np.load(tmp_path / f"rank{k}.npz") for k in range(2)]
    assert int(parts[0]["exchanged"]) + int(parts[1]["exchanged"]) > 0  # unequal shares: rows really cross the GPUs
    assert parts[0]["rows"][1] == parts[1]["rows"][0] and parts[0]["rows"][0] == 0 and parts[1]["rows"][1] == raw.num_rows
    for rep in range(3):
        xh = rng.standard_normal((raw.num_cols, 1)) + 1j * rng.standard_normal((raw.num_cols, 1))
        y = np.concatenate([p[f"y{rep}"] for p in parts])
        y_ref = portlib.apply(raw, xh, capi.OP_N)
        assert not np.isnan(y).any()
        assert np.linalg.norm(y - y_ref) <= 1e-12 * np.linalg.norm(y_ref), rep
