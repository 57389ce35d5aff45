"""Tile statistics of a compiled plan (development aid): row / kc histograms and the fraction of the
DMMA fragment work (rows rounded up to 8, kc to 4) that is useful."""
import sys, time
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
sys.path.insert(0, str(Path(__file__).resolve().parent.parent / "tests"))
import numpy as np
from butterfly_b200 import capi, synth
from tools import workloads
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
# per-stage DMMA fragment efficiency
for s in range(T.num_stages):
    tl = T.tiles[T.stage_tile[s]:T.stage_tile[s + 1]]
    k = tl["kc"].astype(np.int64); r = tl["rows"].astype(np.int64)
    d = k < 0xFFFE
    k, r = k[d], r[d]
    u = (r * k).sum(); p = (((r + 7) // 8 * 8) * ((k + 3) // 4 * 4)).sum()
    it = T.items[T.stage_item[s]:T.stage_item[s + 1]]
    print(f"stage {s}: tiles {len(k)} items {len(it)} useful/padded {u / max(p, 1):.3f} mean rows {r.mean():.1f} mean kc {k.mean():.1f} tiles/item {len(k) / max(len(it), 1):.1f} rows%8==0 {np.mean(r % 8 == 0):.2f}")
