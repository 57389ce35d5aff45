"""Benchmark workloads: flattened operators of the named sizes.  Input generation for bench.py and the
tools — NOT part of the product package: the generators run the reference's own layout code through the
oracle build (oracle/_ref), which only the bench and the tests may touch."""
from __future__ import annotations

import os
import subprocess
import sys
import tempfile
from pathlib import Path

import numpy as np

from butterfly_b200 import capi, synth
from butterfly_b200.rawops import RawOps

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


def helm2_circle_recorded(n: int, k: float | None = None, cache_dir: str | None = None):
    """The same operator laid out INSIDE a recording (tools/helm2_record.py, csrc/bf_helm2_record.c): the
    structure plus the recipes bfCudaPlanBuildHelm2 needs to compute the REAL Helmholtz values on the GPU.
    Returns (RawOps, meta, helm2.Recording); raises if the recorder or the reference build is absent."""
    from butterfly_b200 import helm2

    k = float(k) if k else n / 40.96
    if not (STUB.parent / "libbutterfly_ref.so").exists() or not helm2.HELM2_LIB_PATH.exists():
        raise FileNotFoundError("oracle/_ref or libbfcuda_helm2.so not built")
    cache_dir = cache_dir or tempfile.gettempdir()
    path = Path(cache_dir) / f"bf_helm2_circle_rec_{n}_{k:g}.npz"
    if not path.exists():
        tmp = str(path) + f".{os.getpid()}.tmp.npz"
        subprocess.run([sys.executable, str(ROOT / "tools" / "helm2_record.py"), str(n), repr(k), tmp],
                       check=True, stdout=subprocess.DEVNULL)
        os.replace(tmp, path)
    z = np.load(path, allow_pickle=False)
    raw = RawOps(int(z["dtype"]), int(z["num_rows"]), int(z["num_cols"]), z["buf_rows"],
                 z["ops"].astype(capi.RAWOP_DTYPE), None)
    meta = {"workload": f"helm2_circle N={n} k={k:g}",
            "structure": "reference fac_helm2 laid out inside a recording; values = the real Helmholtz factors, computed on the GPU (bfCudaPlanBuildHelm2)",
            "perm": z["perm"], "points": z["points"], "k_wave": k, "structure_seconds": float(z["seconds"])}
    return raw, meta, helm2.Recording.load_from(z)


def dense_rows_check(meta, y_tree_order, x, num_rows: int = 64, seed: int = 1) -> float:
    """rel. l2 error of an apply against `num_rows` sampled rows of the dense kernel matrix i/4 H0(k|x_i - x_j|)
    in tree order (scipy's Hankel function): the reference's own accuracy check (bf_all_blocks.c:104-153) at a
    size where the matrix itself (4 TB at N = 524288) cannot be formed."""
    from scipy.special import hankel1

    tp = meta["points"][meta["perm"]]
    n = len(tp)
    rows = np.random.default_rng(seed).choice(n, min(num_rows, n), replace=False)
    d = np.hypot(tp[rows, None, 0] - tp[None, :, 0], tp[rows, None, 1] - tp[None, :, 1])
    with np.errstate(invalid="ignore", divide="ignore"):
        K = 0.25j * hankel1(0, meta["k_wave"] * d)
    K[d == 0] = 0
    want = K @ np.asarray(x).reshape(-1)
    return float(np.linalg.norm(np.asarray(y_tree_order).reshape(-1)[rows] - want) / np.linalg.norm(want))


def lbo_like(num_rows: int = 500_000, num_eigs: int = 4096):
    """configs[3] stand-in: real fp64 eigenbasis butterfly of the streamed shape (synth.lbo_like)."""
    raw = synth.lbo_like(num_rows, num_eigs)
    return raw, {"workload": f"lbo_like {num_rows} vertices x {num_eigs} eigenvectors, 16 bands, depth 5, ranks 20-44",
                 "structure": "synthetic streamed-butterfly shape (values synthetic)"}
