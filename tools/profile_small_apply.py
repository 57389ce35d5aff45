"""Four per-stage 1-RHS applies of the reference-structured operator at N (default 16384), for an ncu capture of
the stage kernels:  ncu --set full --clock-control none --import-source on -k regex:bfGemvStage -s 12 -c 2 -o out python tools/profile_small_apply.py"""
import sys
sys.path.insert(0, "/root/repo")
import torch
from butterfly_b200.plan import Plan
from tools import workloads
n = int(sys.argv[1]) if len(sys.argv) > 1 else 16384
stream = torch.cuda.current_stream().cuda_stream
raw, _ = workloads.helm2_circle(n)
plan = Plan(raw); plan.fill_synthetic(7); plan.set_fused(False)
x = torch.randn(n, 1, dtype=torch.complex128, device="cuda"); y = torch.empty_like(x)
for _ in range(4): plan.apply_device(x, y, stream=stream)
torch.cuda.synchronize()
plan.close()
