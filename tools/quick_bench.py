"""Ad-hoc timing of the stage kernels on a value-less synthetic operator (development aid)."""
import sys, time
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import numpy as np, torch
from butterfly_b200 import capi, synth
from butterfly_b200.plan import Plan

n = int(sys.argv[1]) if len(sys.argv) > 1 else 65536
nrhs_list = [int(a) for a in sys.argv[2].split(",")] if len(sys.argv) > 2 else [1, 256]
t0 = time.time()
if len(sys.argv) > 3 and sys.argv[3] == "helm2":
    from tools import workloads
    raw, _ = workloads.helm2_circle(n)
else:
    raw = synth.random_hodbf(n=n, dtype=capi.COMPLEX_F64, seed=1, leaf=64, rank_lo=16, rank_hi=40, with_values=False, exotic=False)
print(f"structure: ops {raw.num_ops} E {raw.num_entries} ({raw.num_entries*16/1e9:.2f} GB) built in {time.time()-t0:.1f}s", flush=True)
t0 = time.time()
plan = Plan(raw)
plan.fill_synthetic(7)
print("plan:", plan.info, f"{time.time()-t0:.1f}s", flush=True)
E = raw.num_entries
stream = torch.cuda.current_stream().cuda_stream
for nrhs in nrhs_list:
    x = torch.randn(n, nrhs, dtype=torch.complex128, device="cuda")
    y = torch.empty(n, nrhs, dtype=torch.complex128, device="cuda")
    for _ in range(3):
        plan.apply_device(x, y, stream=stream)
    torch.cuda.synchronize()
    K = 10
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(K + 1)]
    ev[0].record()
    for i in range(K):
        plan.apply_device(x, y, stream=stream)
        ev[i + 1].record()
    torch.cuda.synchronize()
    ts = np.array([ev[i].elapsed_time(ev[i + 1]) for i in range(K)])
    ms = float(np.median(ts))
    print(f"nrhs {nrhs}: {ms:.3f} ms/apply  {16*E/ms/1e6:.1f} GB/s pool  {8*E*nrhs/ms/1e9:.2f} TFLOP/s  y norm {float(y.abs().max()):.3g}", flush=True)
    plan.set_profiling(True)
    plan.apply_device(x, y, stream=stream)
    torch.cuda.synchronize()
    st = plan.stage_times_ms(); se = plan.stage_entries()
    plan.set_profiling(False)
    for s in range(len(st)):
        print(f"   stage {s}: {st[s]:.3f} ms  entries {se[s]}  {16*se[s]/max(st[s],1e-6)/1e6:.0f} GB/s  {8*se[s]*nrhs/max(st[s],1e-6)/1e9:.2f} TF/s")
