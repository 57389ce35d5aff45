"""Per-kernel SASS mnemonic counts of libbfcuda.so (cuobjdump -sass), written under profiles/: the evidence that the
hot kernels are sm_100a code using the fp64 tensor pipe (DMMA), 1-D TMA bulk copies (UBLKCP) and mbarriers (SYNCS).
tcgen05 (UTC*MMA / LDTM) does not appear on purpose: it has no fp64.  usage: python tools/sass_summary.py OUT.txt"""
import collections
import re
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
lib = ROOT / "butterfly_b200" / "lib" / "libbfcuda.so"
out = Path(sys.argv[1]) if len(sys.argv) > 1 else ROOT / "profiles" / "sass_summary.txt"
sass = subprocess.run(["cuobjdump", "-sass", str(lib)], capture_output=True, text=True).stdout
WATCH = ["DMMA", "DFMA", "UBLKCP", "SYNCS", "UTMALDG", "UTCHMMA", "UTCQMMA", "LDTM", "ATOMG", "RED", "MEMBAR", "FENCE", "LDG", "STG", "LDS", "SHFL", "NANOSLEEP", "ACQBULK", "CCTL"]
arch, fn = None, None
counts = collections.OrderedDict()
for line in sass.splitlines():
    m = re.search(r"arch = (sm_\w+)", line)
    if m:
        arch = m.group(1)
    m = re.search(r"Function : (\S+)", line)
    if m:
        fn = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip().split("(")[0]
        counts[fn] = collections.Counter({"arch": arch})
        counts[fn]["arch"] = arch
        continue
    m = re.match(r"\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+)", line)
    if m and fn:
        op = m.group(1)
        counts[fn]["instructions"] += 1
        for w in WATCH:
            if op.startswith(w):
                counts[fn][w] += 1
with open(out, "w") as f:
    f.write("# cuobjdump -sass butterfly_b200/lib/libbfcuda.so : static instruction counts per kernel (tools/sass_summary.py)\n")
    f.write("# DMMA = fp64 tensor-core MMA (mma.sync m8n8k4.f64), UBLKCP = cp.async.bulk (1-D TMA), SYNCS = mbarrier ops; no UTC*MMA / LDTM: tcgen05 has no fp64\n")
    cols = ["instructions"] + WATCH
    f.write("kernel,arch," + ",".join(cols) + "\n")
    for fn, c in counts.items():
        f.write(f"\"{fn}\",{c['arch']}," + ",".join(str(c.get(k, 0)) for k in cols) + "\n")
print(open(out).read()[:3000])
