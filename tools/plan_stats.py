"""Tile statistics of a compiled plan (development aid): row / kc histograms and the fraction of the
DMMA fragment work (rows rounded up to 8, kc to 4) that is useful."""
import sys, time
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
sys.path.insert(0, str(Path(__file__).resolve().parent.parent / "tests"))
import numpy as np
from butterfly_b200 import capi, synth, workloads
from table_interp import Tables

n = int(sys.argv[1]) if len(sys.argv) > 1 else 65536
kind = sys.argv[2] if len(sys.argv) > 2 else "helm2"
t0 = time.time()
if kind == "helm2":
    raw, meta = workloads.helm2_circle(n)
else:
    raw = synth.random_hodbf(n=n, dtype=capi.COMPLEX_F64, seed=1, leaf=64, rank_lo=16, rank_hi=40, with_values=False, exotic=False)
print(f"{kind} n={n}: ops {raw.num_ops} E {raw.num_entries} ({time.time()-t0:.1f}s)")
T = Tables(raw)
tiles = T.tiles
kc = tiles["kc"].astype(np.int64); rows = tiles["rows"].astype(np.int64)
dense = kc < 0xFFFE
kc, rows = kc[dense], rows[dense]
useful = (rows * kc).sum()
padded = (((rows + 7) // 8 * 8) * ((kc + 3) // 4 * 4)).sum()
print(f"tiles {len(kc)}  useful {useful}  dmma-padded {padded}  efficiency {useful/padded:.3f}")
print("rows hist:", np.bincount(rows, minlength=33)[1:])
print("kc quantiles:", np.percentile(kc, [0, 10, 25, 50, 75, 90, 100]))
print("mean tile bytes:", (rows * kc * 16).mean())
sib = T.stage_item_begin if hasattr(T, "stage_item_begin") else None
