"""Per-kernel totals of an `ncu --metrics gpu__time_duration.sum --csv` launch list, written under profiles/
(development aid).  usage: python tools/launch_list_summary.py LAUNCHES.csv OUT.csv "header note" """
import collections
import csv
import sys

src, out, note = sys.argv[1], sys.argv[2], sys.argv[3] if len(sys.argv) > 3 else ""
rows = [r for r in csv.reader(l for l in open(src) if not l.startswith("==")) if len(r) > 10]
hdr, rows = rows[0], rows[1:]
ik, iv, iu = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
scale = {"ns": 1e-3, "us": 1.0, "usecond": 1.0, "nsecond": 1e-3, "ms": 1e3, "msecond": 1e3, "s": 1e6, "second": 1e6}
tot = collections.OrderedDict()
for r in rows:
    k = r[ik].split("(")[0]
    n, t = tot.get(k, (0, 0.0))
    tot[k] = (n + 1, t + float(r[iv].replace(",", "")) * scale.get(r[iu], 1e-3))
allus = sum(t for _, t in tot.values())
with open(out, "w") as f:
    f.write(f"# {note}\n")
    f.write("# ncu --metrics gpu__time_duration.sum --clock-control none: cold-cache, serialised launches - compare SHARES, not absolute times\n")
    f.write("kernel,launches,total_us,share_of_captured\n")
    for k, (n, t) in sorted(tot.items(), key=lambda kv: -kv[1][1]):
        f.write(f"\"{k}\",{n},{t:.1f},{t / allus:.4f}\n")
    g = [float(r[iv].replace(",", "")) * scale.get(r[iu], 1e-3) for r in rows if "bfGemvStage" in r[ik]]
    if g:
        per = 8 if len(g) % 8 == 0 else len(g)
        f.write(f"# first 1-RHS apply = {per} bfGemvStage launches and nothing else: stage times us = {[round(v, 1) for v in g[:per]]}, sum {sum(g[:per]):.1f}\n")
print(open(out).read())
