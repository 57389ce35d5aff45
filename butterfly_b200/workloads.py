"""Benchmark workloads: flattened operators of the named sizes (structure only; values are filled
on the device by bfCudaPlanFillSynthetic).  Input generation, not part of the apply path."""
from __future__ import annotations

import os
import subprocess
import sys
import tempfile
from pathlib import Path

import numpy as np

from . import capi, synth
from .rawops import RawOps

ROOT = Path(__file__).resolve().parent.parent
STUB = ROOT / "oracle" / "_ref" / "libbf_shape_stub.so"


def helm2_circle(n: int, k: float | None = None, cache_dir: str | None = None):
    """Structure of bfFacHelm2MakeMultilevel on n unit-circle points at wavenumber k (default: the
    example's ~41 points per wavelength, k = n / 40.96, run_bf_all_blocks_test.sh:8-9), laid out
    by the reference's own code with value computation skipped (tools/helm2_structure.py).
    Returns (RawOps, meta).  Falls back to synth.random_hodbf — and says so in meta — only if the
    prebuilt reference libraries are absent."""
    k = float(k) if k else n / 40.96
    meta = {"workload": f"helm2_circle N={n} k={k:g}", "structure": "reference fac_helm2 (values synthetic)"}
    if not STUB.exists():
        raw = synth.random_hodbf(n=n, dtype=capi.COMPLEX_F64, seed=1, leaf=64, rank_lo=16, rank_hi=40,
                                 with_values=False, exotic=False)
        meta["structure"] = "FALLBACK synth.random_hodbf (oracle/_ref absent)"
        return raw, meta
    cache_dir = cache_dir or tempfile.gettempdir()
    path = Path(cache_dir) / f"bf_helm2_circle_{n}_{k:g}.npz"
    if not path.exists():
        tmp = str(path) + f".{os.getpid()}.tmp.npz"
        subprocess.run([sys.executable, str(ROOT / "tools" / "helm2_structure.py"), str(n), repr(k), tmp],
                       check=True, stdout=subprocess.DEVNULL)
        os.replace(tmp, path)
    z = np.load(path, allow_pickle=False)
    raw = RawOps(int(z["dtype"]), int(z["num_rows"]), int(z["num_cols"]), z["buf_rows"],
                 z["ops"].astype(capi.RAWOP_DTYPE), None)
    meta["perm"] = z["perm"]
    meta["structure_seconds"] = float(z["seconds"])
    return raw, meta


def lbo_like(num_rows: int = 500_000, num_eigs: int = 4096):
    """configs[3] stand-in: real fp64 eigenbasis butterfly of the streamed shape (synth.lbo_like)."""
    raw = synth.lbo_like(num_rows, num_eigs)
    return raw, {"workload": f"lbo_like {num_rows} vertices x {num_eigs} eigenvectors, 16 bands, depth 5, ranks 20-44",
                 "structure": "synthetic streamed-butterfly shape (values synthetic)"}
