"""A/B of plan-compiler knobs that are read per compile (BF_PLAN_DEFER, BF_PLAN_TILE_COST): one process, one plan per
variant and size, ms per 1-RHS apply with per-stage launches and with the single launch.
usage: python tools/plan_knob_ab.py VARIANT[,VARIANT...] [N ...]  with VARIANT = KNOB=v[:KNOB=v...]
e.g.  python tools/plan_knob_ab.py BF_PLAN_TILE_COST=384,BF_PLAN_TILE_COST=8192:BF_PLAN_DEFER=0 16384 65536"""
import json
import os
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch

from butterfly_b200.plan import Plan
from tools import workloads

values = sys.argv[1].split(",")
knob = "variant"
KNOBS = ("BF_PLAN_DEFER", "BF_PLAN_TILE_COST", "BF_PLAN_ITEM_COST", "BF_PLAN_SORT")
stream = torch.cuda.current_stream().cuda_stream
for n in [int(a) for a in sys.argv[2:]] or [16384, 65536]:
    raw, _ = workloads.helm2_circle(n)
    x = torch.randn(n, 1, dtype=torch.complex128, device="cuda")
    plans, ys = {}, {}
    for v in values:
        for k in KNOBS:
            os.environ.pop(k, None)
        for kv in v.split(":"):
            if kv:
                k, val = kv.split("=")
                os.environ[k] = val
        plans[v] = Plan(raw)
        plans[v].fill_synthetic(7)
    reps = 300 if n <= 131072 else 30
    for rnd in range(2):
        for v, plan in plans.items():
            out = {"n": n, knob: v, "round": rnd}
            for mode in ("staged", "fused"):
                plan.set_fused(mode == "fused")
                y = torch.empty_like(x)
                for _ in range(10):
                    plan.apply_device(x, y, stream=stream)
                torch.cuda.synchronize()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                for _ in range(reps):
                    plan.apply_device(x, y, stream=stream)
                e1.record()
                torch.cuda.synchronize()
                out[mode + "_ms"] = round(e0.elapsed_time(e1) / reps, 5)
                ys[v] = y.clone()
            print(json.dumps(out), flush=True)
    d = max(float((ys[v] - ys[values[0]]).norm() / ys[values[0]].norm()) for v in values)
    print(json.dumps({"n": n, "max_relative_difference_between_variants": d}), flush=True)
    for p in plans.values():
        p.close()
    for k in KNOBS:
        os.environ.pop(k, None)
