"""Time the real many-RHS kernel (bfWideStage<real>) on the synthetic 500k-vertex streamed shape: 64 and 256 columns,
against the DFMA peak measured on the device.  usage: python tools/real_wide_time.py"""
import ctypes as C
import json
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch

from butterfly_b200 import capi
from butterfly_b200.plan import Plan
from tools import workloads

raw, meta = workloads.lbo_like()
plan = Plan(raw)
plan.fill_synthetic(7)
E = plan.info["numEntries"]
dfma, dmma = C.c_double(), C.c_double()
capi.check(capi.load_library().bfCudaMeasureFp64Peak(0, C.byref(dfma), C.byref(dmma)), "bfCudaMeasureFp64Peak")
stream = torch.cuda.current_stream().cuda_stream
x1 = torch.randn(raw.num_cols, 1, dtype=torch.float64, device="cuda")
y1 = torch.empty(raw.num_rows, 1, dtype=torch.float64, device="cuda")
plan.apply_device(x1, y1, stream=stream)
torch.cuda.synchronize()
for nr in (64, 256):
    X = x1.repeat(1, nr).contiguous()
    Y = torch.empty(raw.num_rows, nr, dtype=torch.float64, device="cuda")
    for _ in range(2):
        plan.apply_device(X, Y, stream=stream)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5):
        plan.apply_device(X, Y, stream=stream)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 5
    tf = 2.0 * E * nr / (ms * 1e-3) / 1e12
    err = float((Y - y1).abs().max() / y1.abs().max())  # every column is the 1-RHS result
    print(json.dumps({"nrhs": nr, "ms": round(ms, 4), "tflops": round(tf, 2), "frac_of_dfma_peak": round(tf / dfma.value, 3),
                      "dfma_peak": round(dfma.value, 2), "max_abs_diff_vs_1rhs_over_max": err}), flush=True)
plan.close()
