"""A/B of the 1-RHS apply: one cooperative launch over the fused tables vs one launch per stage, on the reference's
fac_helm2 structure (synthetic values).  usage: python tools/fused_ab.py N [N ...]   Prints one JSON line per N."""
import json
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import numpy as np
import torch

from butterfly_b200.plan import Plan
from tools import workloads

peak = json.loads((Path(__file__).resolve().parent.parent / "MEASURED_PEAKS.json").read_text())["hbm_gbs"] if (Path(__file__).resolve().parent.parent / "MEASURED_PEAKS.json").exists() else 6556.5
stream = torch.cuda.current_stream().cuda_stream
for n in [int(a) for a in sys.argv[1:]] or [16384, 65536]:
    raw, _ = workloads.helm2_circle(n)
    plan = Plan(raw)
    plan.fill_synthetic(7)
    E = plan.info["numEntries"]
    x = torch.randn(n, 1, dtype=torch.complex128, device="cuda")
    out = {"n": n, "E": int(E), "stages": int(plan.info["numStagesN"])}
    ys = {}
    for mode in ("staged", "fused"):
        plan.set_fused(mode == "fused")
        y = torch.empty(n, 1, dtype=torch.complex128, device="cuda")
        for _ in range(10):
            plan.apply_device(x, y, stream=stream)
        torch.cuda.synchronize()
        reps = 200 if n <= 131072 else 30
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            plan.apply_device(x, y, stream=stream)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / reps
        gbs = (16.0 * E + 32.0 * n) / ms / 1e6
        out[mode] = {"ms": round(ms, 5), "gbs": round(gbs, 1), "frac": round(gbs / peak, 4), "launches": plan.last_launch_count}
        ys[mode] = y.clone()
    out["bit_identical"] = bool(torch.equal(ys["staged"], ys["fused"]))
    print(json.dumps(out), flush=True)
    plan.close()
