"""Development probe: block-row sharded apply (all ranks in this process) against the unsharded plan.
usage: python tools/shard_probe.py N WORLD [shared]   (shared: all ranks on GPU 0)"""
import json
import sys
import time
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import numpy as np
import torch

from butterfly_b200 import capi
from butterfly_b200.plan import Plan
from butterfly_b200.sharded import Comm, ShardedPlan
from tools import workloads

n, world = int(sys.argv[1]), int(sys.argv[2])
shared = len(sys.argv) > 3 and sys.argv[3] == "shared"
devs = [0] * world if shared else list(range(world))
raw, _ = workloads.helm2_circle(n)
out = {"n": n, "world": world, "devices": devs}
E = raw.num_entries


def timed(fn, sync, reps=20):
    for _ in range(3):
        fn()
    sync()
    t0 = time.perf_counter()
    for _ in range(reps):
        fn()
    sync()
    return (time.perf_counter() - t0) / reps * 1e3


if n <= 524288:
    with Plan(raw, device=0) as full:
        full.fill_synthetic(7)
        x0 = torch.randn(n, 1, dtype=torch.complex128, device="cuda:0")
        y0 = torch.empty_like(x0)
        for fused in (False, True):
            full.set_fused(fused)
            out["unsharded_ms_fused" if fused else "unsharded_ms_staged"] = timed(lambda: full.apply_device(x0, y0, stream=0), torch.cuda.synchronize)
        full.set_fused(False)
        full.set_profiling(True)
        full.apply_device(x0, y0, stream=0)
        torch.cuda.synchronize()
        out["unsharded_stage_ms"] = [round(float(v), 4) for v in full.stage_times_ms()]
        del x0, y0
comm = Comm(devs)
sp = ShardedPlan(raw, comm, capi.SHARD_BLOCK_ROWS)
for r in range(world):
    sp.rank_plan(r).fill_synthetic(7)
out["info"] = sp.info
xs = [torch.randn(n, 1, dtype=torch.complex128, device=f"cuda:{d}") for d in devs]
ys = [torch.empty(n, 1, dtype=torch.complex128, device=f"cuda:{d}") for d in devs]
out["sharded_ms_back_to_back"] = timed(lambda: sp.apply_device(xs, ys), sp.synchronize)
sp.synchronize()
t0 = time.perf_counter()
for _ in range(20):
    sp.apply_device(xs, ys)
out["host_enqueue_ms_per_apply"] = (time.perf_counter() - t0) / 20 * 1e3
sp.synchronize()
for r in range(world):
    sp.rank_plan(r).set_profiling(True)
for _ in range(6):
    sp.apply_device(xs, ys)
sp.synchronize()
out["sharded_stage_ms_back_to_back"] = [[round(float(v), 4) for v in sp.rank_plan(r).stage_times_ms()] for r in range(world)]
for r in range(world):
    sp.rank_plan(r).set_profiling(False)
out["sharded_ms_synced_each"] = timed(lambda: (sp.apply_device(xs, ys), sp.synchronize()), sp.synchronize)
# per-stage times of every rank (events on the ranks' own streams)
for r in range(world):
    sp.rank_plan(r).set_profiling(True)
sp.apply_device(xs, ys)
sp.synchronize()
out["sharded_stage_ms"] = [[round(float(v), 4) for v in sp.rank_plan(r).stage_times_ms()] for r in range(world)]
print(json.dumps(out))
