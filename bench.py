#!/usr/bin/env python
"""bench.py — butterfly apply throughput on B200 (contract: see DESIGN.md §6).

A step = one apply Y = A X of the flattened Helmholtz HODBF (reference fac_helm2 structure on N
unit-circle points, synthetic values) to an n x nrhs complex-fp64 block.  `value` is whole-job
GFLOP/s (flops = 8 E nrhs) with X, Y resident in HBM; `e2e` is the same metric
through the host-buffer C-ABI call (bfCudaApplyHost: what the swapped bfMatMul vtable slot calls),
H2D and D2H inside the timed region.  N > 1 GPUs: one process per GPU, plan replicated, the RHS
columns sharded across ranks (no collective on the data path; weak scaling in RHS).
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

DEFAULT_N = 524288  # largest rung of the ladder (2^20, 2^19, 2^18) whose factors fit 180 GB of HBM
                    # at the example's 41 points per wavelength: 2^20 needs 186 GB
CPU_SAMPLE_N = 16384  # configs[0]: the reference's own CPU-runnable case (examples/simple, 16k circle points)


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--n", "--points", dest="n", type=int, default=int(os.environ.get("BF_BENCH_N", DEFAULT_N)))
    ap.add_argument("--k", type=float, default=0.0, help="wavenumber (default n / 40.96)")
    ap.add_argument("--nrhs", type=int, default=1, help="RHS columns PER GPU")
    ap.add_argument("--shard", default="cols", choices=["cols", "rows"],
                    help="cols: plan replicated, RHS columns sharded, no collective (default, weak scaling); "
                         "rows: ONE operator block-row sharded over the ranks, workspace rows pushed to their readers by the stage kernels (strong scaling)")
    ap.add_argument("--row-fractions", default="", help="--shard rows: comma-separated Y shares per rank (default equal); unequal shares cut "
                                                         "through the top-level butterflies and exercise the exchange")
    ap.add_argument("--solve", type=int, default=0, help="--shard rows: also time this many GMRES iterations over the sharded operator (configs[4])")
    ap.add_argument("--verify", action="store_true", help="--shard rows: also apply the unsharded plan (must fit one GPU) and compare own rows")
    ap.add_argument("--values", default="helm2", choices=["helm2", "synthetic"],
                    help="helm2: the real Helmholtz factor values, computed on the GPU (bfCudaPlanBuildHelm2); synthetic: hash fill")
    ap.add_argument("--truncation-scale", type=float, default=0.0,
                    help="--values helm2: threshold of the truncated-SVD least squares as a multiple of the reference's rule "
                         "(0 = the reference's rule; 4 gives a far more accurate operator, DESIGN.md 8c)")
    ap.add_argument("--no-extra", action="store_true", help="skip the 256-RHS side measurement")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-1m", action="store_true", help="multi-GPU runs (>= 4 GPUs): skip the 2^20-point block-row sharded side block")
    return ap.parse_args()


def peaks():
    p = ROOT / "MEASURED_PEAKS.json"
    if p.exists():
        d = json.loads(p.read_text())
        return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region."""

    Q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index):
        self.index, self.samples, self.stop, self.t = index, [], False, None

    def _run(self):
        while not self.stop:
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits"],
                                     capture_output=True, text=True, timeout=5).stdout.strip()
                if out:
                    self.samples.append([s.strip() for s in out.split(",")])
            except Exception:
                pass
            time.sleep(0.02)

    def __enter__(self):
        self.t = threading.Thread(target=self._run, daemon=True)
        self.t.start()
        return self

    def __exit__(self, *a):
        self.stop = True
        self.t.join(timeout=6)

    def summary(self):
        import statistics

        mhz = [int(s[0]) for s in self.samples if s and s[0].isdigit()]
        mx = [int(s[1]) for s in self.samples if len(s) > 1 and s[1].isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({names[i] for s in self.samples for i in range(4) if len(s) > 2 + i and s[2 + i].lower().startswith("active")})
        return {"sm_mhz": int(statistics.median(mhz)) if mhz else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(self.samples)}


def cpu_reference_leg(steps, warmup, sample_n=CPU_SAMPLE_N, keep=False):
    """The reference's own CPU bfMatMul (oracle/_ref = unmodified sources + OpenBLAS) on a bounded
    sample of the workload: the same generator at a smaller N, REAL factor values, 1 RHS.  keep: also
    return the live reference operator (for the same-factors GPU leg, configs0_block)."""
    import numpy as np

    from oracle import reflib

    if not reflib.available():
        if keep:
            return None, None
        # the oracle port on a synthetic operator of the same family
        from butterfly_b200 import capi, synth
        from oracle import portlib

        raw = synth.random_hodbf(n=sample_n, dtype=capi.COMPLEX_F64, seed=1, leaf=64, rank_lo=16, rank_hi=40, exotic=False)
        x = np.random.default_rng(0).standard_normal((sample_n, 1)) + 0j
        ts = []
        for i in range(warmup + steps):
            t0 = time.perf_counter(); portlib.apply(raw, x); ts.append(time.perf_counter() - t0)
        t = float(np.median(ts[warmup:]))
        return {"value": 8 * raw.num_entries / t / 1e9, "unit": "GFLOP/s", "cores": 1, "kind": "port",
                "sample": f"oracle port, synthetic HODBF n={sample_n}, 1 RHS, median of {steps}", "ms_per_step": t * 1e3, "entries": raw.num_entries}
    k = sample_n / 40.96
    t0 = time.perf_counter()
    h = reflib.Helm2(reflib.circle_points(sample_n), k)
    build_s = time.perf_counter() - t0
    E = h.flatten().num_entries if reflib.attach_available() else None
    x = reflib.randn_complex(sample_n, 0)
    best = None
    ncpu = os.cpu_count() or 1
    for threads in sorted({1, ncpu}):
        used = reflib.set_blas_threads(threads)
        ts = [h.matmul(x)[1] for _ in range(warmup + steps)]
        t = float(np.median(ts[warmup:]))
        if best is None or t < best[0]:
            best = (t, used)
    if not keep:
        h.close()
    t, used = best
    if E is None:
        E = int(2112 * sample_n)
    r = {"value": 8 * E / t / 1e9, "unit": "GFLOP/s", "cores": int(used), "kind": "reference",
         "sample": f"reference bfMatMul (unmodified sources, OpenBLAS) on fac_helm2 circle N={sample_n} k={k:g} (configs[0]: same generator at the reference's own CPU size; built in {build_s:.1f}s), 1 RHS, median of {steps}, best of 1 and {ncpu} BLAS threads",
         "ms_per_step": t * 1e3, "entries": int(E), "matvecs_per_s": 1.0 / t, "sample_n": sample_n}
    return (r, h) if keep else r


def configs0_block(local_rank):
    """BASELINE.json configs[0] measured, not quoted: the reference builds fac_helm2 on 16 384 circle points
    (k = 400) ONCE; the SAME object is then timed through the reference's own bfMatMul on the CPU and — after
    bfCudaMatAttach — on the GPU (malloc'ed result, pageable operand: what a reference caller gets), plus the
    device-resident apply of that plan against the HBM roofline, plus parity on the same factors."""
    import numpy as np
    import torch

    from butterfly_b200 import capi
    from butterfly_b200.plan import Plan
    from oracle import reflib

    cpu, h = cpu_reference_leg(steps=10, warmup=2, keep=True)
    if h is None:
        return None, None
    out = {"workload": f"helm2_circle N={CPU_SAMPLE_N} k={CPU_SAMPLE_N / 40.96:g} nrhs=1, factors built by the reference (same object on both sides)",
           "same_config": True, "cpu_ms_per_apply": cpu["ms_per_step"], "cpu_cores": cpu["cores"], "E": cpu["entries"]}
    try:
        n = CPU_SAMPLE_N
        x = reflib.randn_complex(n, 0)
        rng = np.random.default_rng(7)
        X = rng.standard_normal((n, 64)) + 1j * rng.standard_normal((n, 64))
        probe = np.arange(0, 64, 4)
        y_cpu, _ = h.matmul(x)
        Yp_cpu, _ = h.matmul(X[:, probe].copy())
        t0 = time.perf_counter()
        h.attach()
        out["attach_s"] = time.perf_counter() - t0  # flatten + compile + upload of 554 MB of factors
        y_gpu, _ = h.matmul(x)
        Y_gpu, _ = h.matmul(X)
        rel = lambda a, b: float(np.linalg.norm(a - b) / np.linalg.norm(b))
        out["parity"] = {"rel_l2_same_factors_1rhs": rel(y_gpu, y_cpu),
                         "rel_l2_same_factors_64rhs_max": max(rel(Y_gpu[:, c], Yp_cpu[:, i]) for i, c in enumerate(probe)),
                         "columns_checked_of_64": len(probe), "tolerance": 1e-12}
        if max(out["parity"]["rel_l2_same_factors_1rhs"], out["parity"]["rel_l2_same_factors_64rhs_max"]) > 1e-12:
            raise SystemExit(f"bench.py: configs[0] parity failed: {out['parity']}")
        ts = [h.matmul(x)[1] for _ in range(30)]
        out["gpu_bfMatMul_ms_per_apply"] = float(np.median(ts[5:])) * 1e3  # inside bfMatMul: H2D + apply + D2H + malloc
        out["speedup_bfMatMul_same_operator"] = out["cpu_ms_per_apply"] / out["gpu_bfMatMul_ms_per_apply"]
        plan = Plan.borrow(h.plan_handle(), n, n)
        stream = torch.cuda.current_stream().cuda_stream
        xd = torch.from_numpy(x).cuda().reshape(n, 1)
        yd = torch.empty_like(xd)
        for _ in range(10):
            plan.apply_device(xd, yd, stream=stream)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        reps = 200
        e0.record()
        for _ in range(reps):
            plan.apply_device(xd, yd, stream=stream)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / reps
        E = plan.info["numEntries"]
        byts = 16.0 * E + 32.0 * n
        out.update({"device_ms_per_apply": ms, "device_gflops": 8.0 * E / (ms * 1e-3) / 1e9, "device_gbs": byts / (ms * 1e-3) / 1e9,
                    "frac_of_hbm_peak": byts / (ms * 1e-3) / 1e9 / peaks()[0], "launches_per_apply": int(plan.last_launch_count),
                    "l2_note": "554 MB of factors per apply: larger than L2 (126 MB), no flush needed"})
        out["parity"]["device_vs_cpu_1rhs"] = rel(yd.cpu().numpy(), y_cpu)
        h.detach()
    except SystemExit:
        raise
    except Exception as e:
        out["error"] = str(e)[:300]
    finally:
        h.close()
    return cpu, out


def workload_label(n, k, nrhs):
    return f"helm2_circle N={n} k={k:g} nrhs={nrhs}/GPU"


def run_reference(args, rank, world):
    if rank != 0:
        return  # rank 0 alone runs the CPU arm
    r = cpu_reference_leg(args.steps, max(args.warmup, 1))
    k = args.k or args.n / 40.96
    # `config` is our arm's (the contract: the reference arm runs on OUR config, each step a bounded sample
    # of that workload); what was actually timed is spelled out in config.sample / cpu_baseline.sample.  The
    # same-problem figure (same N, same factors, CPU vs GPU through bfMatMul) is our arm's `configs0` block.
    line = {
        "impl": "reference", "metric": "butterfly apply GFLOP/s (complex fp64, 1 RHS)", "value": r["value"], "unit": "GFLOP/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": r["ms_per_step"],
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "c128 (complex f64)", "data": "synthetic",
        "config": {"workload": workload_label(args.n, k, 1),
                   "sample": f"timed on helm2_circle N={r.get('sample_n', CPU_SAMPLE_N)} (bounded sample: the reference needs ~11 min and 75 GB of mallocs to build N={args.n}); value is a GFLOP/s RATE"},
        "matvecs_per_s_on_sample": r.get("matvecs_per_s"),
        "cpu_baseline": {k2: r[k2] for k2 in ("value", "unit", "cores", "kind", "sample")},
        "e2e": {"value": r["value"], "unit": "GFLOP/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def run_rows(args, rank, local_rank, world, n=None, steps=None, warmup=None, verify=None, solve=None):
    """Block-row sharded apply (SURVEY.md §8e / configs[4]): every rank holds 1 / world of the factors of ONE
    operator, X is replicated, and the rows of an intermediate level that another rank reads are stored straight
    into that rank's workspace by the producing stage kernel (peer memory over NVLink, one flag per stage and peer:
    csrc/bf_comm.cu, the MODE-2 kernel of bf_kernels.cu) — no NCCL, no collective, no host round trip inside an
    apply.  Returns the result dict (rank 0) — strong scaling: the job is one apply of the operator."""
    import numpy as np
    import torch
    import torch.distributed as dist

    from butterfly_b200 import capi
    from butterfly_b200.sharded import ShardRankOperator
    from tools import workloads

    n = n or args.n
    steps, warmup = steps or args.steps, warmup or args.warmup
    verify = args.verify if verify is None else verify
    solve_iters = args.solve if solve is None else solve
    nrhs = 1
    if world > 1 and rank != 0:
        dist.barrier()
    rec = None
    if args.values == "helm2":
        try:
            raw, meta, rec = workloads.helm2_circle_recorded(n, None)
        except Exception:
            rec = None
    if rec is None:
        raw, meta = workloads.helm2_circle(n, None)
    if world > 1 and rank == 0:
        dist.barrier()
    t0 = time.time()
    fr = [float(v) for v in args.row_fractions.split(",")] if args.row_fractions else None

    def exchange(blob):
        if world == 1:
            return [blob]
        out = [None] * world
        dist.all_gather_object(out, blob)
        return out

    op = ShardRankOperator(raw, world, rank, local_rank, exchange, fractions=fr, synthetic_seed=None if rec is not None else 2024,
                           helm2_recording=rec)
    build_stats = op.build_stats
    build_s = time.time() - t0
    E = int(raw.num_entries)
    g = torch.Generator(device="cuda").manual_seed(1234)  # the operand is replicated: same seed everywhere
    x = torch.randn(n, nrhs, dtype=torch.complex128, device="cuda", generator=g)
    y = torch.zeros(n, nrhs, dtype=torch.complex128, device="cuda")

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(warmup):
        op.apply(x, y)
    barrier()
    verify_err = None
    if verify:
        # the same leaves (same values) as ONE unsharded plan on this GPU, where it fits: own rows must agree
        from butterfly_b200.plan import Plan

        try:
            full = Plan(op.raw, device=local_rank)
            if rec is not None:
                from butterfly_b200 import helm2

                helm2.build_plan(full, rec.remapped(op.origin_op, op.origin_row))
            else:
                full.fill_synthetic(2024)
            yf = torch.empty_like(y)
            full.apply_device(x, yf, stream=torch.cuda.current_stream().cuda_stream)
            torch.cuda.synchronize()
            r0v, r1v = op.y_range
            verify_err = float((y[r0v:r1v] - yf[r0v:r1v]).norm() / yf[r0v:r1v].norm()) if r1v > r0v else 0.0
            full.close()
            del yf
        except Exception as e:  # the unsharded operator does not fit one GPU
            verify_err = None
            if rank == 0:
                print(f"bench.py: row_shard verify skipped: {str(e)[:160]}", file=sys.stderr)
        torch.cuda.empty_cache()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    ev0.record()
    for _ in range(steps):
        op.apply(x, y)
    ev1.record()
    barrier()
    ms = ev0.elapsed_time(ev1) / steps
    solve = None
    if solve_iters > 0:
        # configs[4]: repeated applies inside GMRES over the sharded operator; the first-kind operator is not a
        # convergent system, so a fixed number of iterations is timed (tol = 0)
        from butterfly_b200 import gmres

        bvec = x[:, 0].clone()
        gmres.solve_gmres_sharded(op, bvec, None, 0.0, 2, "cgs")  # warm-up: communicators, allocator
        for orth in ("mgs", "cgs", "cgs2"):
            barrier()
            t0 = time.perf_counter()
            _, its, resid, napp = gmres.solve_gmres_sharded(op, bvec, None, 0.0, solve_iters, orth)
            barrier()
            dt = time.perf_counter() - t0
            solve = dict(solve or {}, **{orth: {"iterations": int(its), "applies": int(napp), "ms_per_iteration": dt / max(its, 1) * 1e3,
                                                "apply_share": ms * napp / (dt * 1e3)}})
    # end to end: pinned host operand in, own rows of the result out
    xh = x.cpu().pin_memory()
    r0, r1 = op.y_range
    yh = torch.empty(max(r1 - r0, 1), nrhs, dtype=torch.complex128).pin_memory()
    barrier()
    t0 = time.perf_counter()
    e2e_steps = max(3, min(steps, 10))
    for _ in range(e2e_steps):
        x.copy_(xh, non_blocking=True)
        op.apply(x, y)
        yh[: r1 - r0].copy_(y[r0:r1], non_blocking=True)
        torch.cuda.synchronize()
    barrier()
    e2e_s = (time.perf_counter() - t0) / e2e_steps
    # real values: the assembled result (rows gathered from their owners) against sampled dense kernel rows
    dense_check = None
    if build_stats is not None:
        mask = torch.zeros(n, 1, dtype=torch.float64, device="cuda")
        mask[r0:r1] = 1.0
        yfull = y * mask
        if world > 1:
            dist.all_reduce(yfull, op=dist.ReduceOp.SUM)
        if rank == 0:
            try:
                dense_check = {"rows": 64, "rel_l2_vs_dense_rows": workloads.dense_rows_check(meta, yfull[:, 0].cpu().numpy(), xh[:, 0].numpy(), 64)}
            except Exception as e:
                dense_check = {"error": str(e)[:200]}
        del yfull
    stats = torch.tensor([ms, e2e_s, float(op.local_entries), float(op.exchanged_rows), verify_err if verify_err is not None else -1.0], device="cuda", dtype=torch.float64)
    mx = stats.clone()
    if world > 1:
        dist.all_reduce(mx, op=dist.ReduceOp.MAX)
        dist.all_reduce(stats, op=dist.ReduceOp.SUM)
    ms, e2e_s = float(mx[0]), float(mx[1])
    res = {"workload": f"helm2_circle N={n} k={n / 40.96:g} nrhs={nrhs}, block-row sharded x{world}", "E": E,
           "pool_GB_per_gpu_max": float(mx[2]) * 16 / 1e9, "ms_per_apply": ms, "gflops": 8.0 * E * nrhs / (ms * 1e-3) / 1e9,
           "matvecs_per_s": nrhs / (ms * 1e-3), "gbs_aggregate": (16.0 * E + 32.0 * n * nrhs) / (ms * 1e-3) / 1e9,
           "frac_of_aggregate_hbm_peak": (16.0 * E + 32.0 * n * nrhs) / (ms * 1e-3) / 1e9 / (peaks()[0] * world),
           "load_imbalance_max_over_mean": float(mx[2]) / (float(stats[2]) / world), "exchanged_rows_per_apply": int(stats[3]),
           "stages": op.num_stages, "launches_per_apply_per_gpu": op.num_stages, "build_s": round(build_s, 1),
           "e2e_ms_per_apply": e2e_s * 1e3, "e2e_gflops": 8.0 * E * nrhs / e2e_s / 1e9,
           "values": "real Helmholtz factors computed on the GPU, each rank its own leaves" if build_stats is not None else "synthetic",
           "value_build": build_stats, "dense_check": dense_check,
           "gmres": solve, "y_cuts": [int(c) for c in op.y_cuts],
           "verify_rel_l2_vs_unsharded_max": (float(mx[4]) if float(mx[4]) >= 0 else None),
           "exchange": "peer stores from the stage kernels into the readers' workspaces (CUDA IPC-mapped, NVLink) + one flag per stage and peer; no collective inside the apply"}
    op.close()
    del op, x, y
    torch.cuda.empty_cache()
    return res


def run_ours(args, rank, local_rank, world):
    import numpy as np
    import torch
    import torch.distributed as dist

    from butterfly_b200 import capi
    from tools import workloads
    from butterfly_b200.plan import Plan

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product has no CPU fallback")
    torch.cuda.set_device(local_rank)
    if world > 1 or args.shard == "rows":
        os.environ.setdefault("MASTER_PORT", "29533")
        os.environ.setdefault("RANK", "0")
        os.environ.setdefault("WORLD_SIZE", "1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    n, nrhs = args.n, args.nrhs
    if args.shard == "rows":
        with ClockSampler(local_rank) as clk:
            r = run_rows(args, rank, local_rank, world)
        if rank == 0:
            line = {"metric": "butterfly apply GFLOP/s (complex fp64, 1 RHS)" if nrhs == 1 else f"butterfly apply GFLOP/s (complex fp64, {nrhs} RHS)",
                    "value": r["gflops"], "unit": "GFLOP/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
                    "ms_per_step": r["ms_per_apply"], "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
                    "dtype": "c128 (complex f64)", "data": "synthetic",
                    "config": {"workload": r["workload"], "structure": "reference fac_helm2", "values": r["values"], "E": r["E"],
                               "parallelism": f"block rows x{world} ({r['exchange']})", "l2": "inputs larger than L2"},
                    "matvecs_per_s": r["matvecs_per_s"], "clocks": clk.summary(),
                    "e2e": {"value": r["e2e_gflops"], "unit": "GFLOP/s", "h2d_bytes_per_step": int(16 * n * nrhs), "d2h_bytes_per_step": int(16 * n * nrhs / world),
                            "ms_per_step": r["e2e_ms_per_apply"], "api": "bfCudaShardRankApplyDevice per rank (peer stores, no collective), pinned host buffers"},
                    "gpu_launches": int(r["stages"] * args.steps),
                    "roofline": {"bound": "hbm", "kernel": "bfGemvStage<complex>", "achieved": r["gbs_aggregate"] / world, "peak": peaks()[0], "unit": "GB/s",
                                 "frac": r["frac_of_aggregate_hbm_peak"], "peak_source": peaks()[1], "traffic": None},
                    "cpu_baseline": None, "row_shard": r}
            print(json.dumps(line), flush=True)
        dist.destroy_process_group()
        return
    # structure: rank 0 of the node generates (cached on disk), the others wait and load it
    if world > 1 and rank != 0:
        dist.barrier()
    rec = None
    values_note = "synthetic (bfCudaPlanFillSynthetic)"
    if args.values == "helm2":
        try:
            raw, meta, rec = workloads.helm2_circle_recorded(n, args.k or None)
        except Exception as e:  # recorder / reference build absent: structure stub + synthetic values
            values_note = f"synthetic (helm2 recording unavailable: {str(e)[:80]})"
    if rec is None:
        raw, meta = workloads.helm2_circle(n, args.k or None)
    if world > 1 and rank == 0:
        dist.barrier()
    t0 = time.time()
    plan = Plan(raw, device=local_rank)
    build_stats = None
    if rec is not None:
        from butterfly_b200 import helm2

        tb = time.time()
        rec.truncation_scale = float(args.truncation_scale)
        try:
            build_stats = helm2.build_plan(plan, rec)
            build_stats["truncation_scale"] = rec.truncation_scale
            build_stats["seconds"] = time.time() - tb
            values_note = "real Helmholtz factors computed on the GPU (bfHelm2Pinv + bfHelm2Fill)"
        except Exception as e:  # e.g. out of device memory next to the pool: measure on synthetic values and say so
            build_stats = None
            values_note = f"synthetic (GPU value build failed: {str(e)[:120]})"
            plan.fill_synthetic(2024)
        del rec
    else:
        plan.fill_synthetic(2024)
    info = plan.info
    E = info["numEntries"]
    build_s = time.time() - t0
    stream = torch.cuda.current_stream().cuda_stream

    g = torch.Generator(device="cuda").manual_seed(rank)
    x = torch.randn(n, nrhs, dtype=torch.complex128, device="cuda", generator=g)
    y = torch.empty(n, nrhs, dtype=torch.complex128, device="cuda")

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(args.warmup):
        plan.apply_device(x, y, stream=stream)
    barrier()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with ClockSampler(local_rank) as clk:
        ev0.record()
        for _ in range(args.steps):
            plan.apply_device(x, y, stream=stream)
        ev1.record()
        barrier()
    ms_total = ev0.elapsed_time(ev1)
    launches = plan.last_launch_count * args.steps
    if world > 1:
        t = torch.tensor([ms_total], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms_total = float(t.item())
    ms_step = ms_total / args.steps
    total_rhs = nrhs * world
    gflops = 8.0 * E * total_rhs / (ms_step * 1e-3) / 1e9

    # per-stage times of one profiled apply (dominant kernel share, smallest/largest stage)
    plan.set_profiling(True)
    plan.apply_device(x, y, stream=stream)
    torch.cuda.synchronize()
    stage_ms = plan.stage_times_ms()
    plan.set_profiling(False)
    stage_entries = plan.stage_entries()

    # end to end through the host-buffer C-ABI call, pinned host memory
    xh = torch.randn(n, nrhs, dtype=torch.complex128).pin_memory()
    yh = torch.empty(n, nrhs, dtype=torch.complex128).pin_memory()
    xn, yn = xh.numpy(), yh.numpy()
    for _ in range(2):
        plan.apply_host(xn, capi.OP_N, out=yn)
    barrier()
    e2e_steps = max(3, min(args.steps, 10))
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        plan.apply_host(xn, capi.OP_N, out=yn)
    barrier()
    e2e_s = (time.perf_counter() - t0) / e2e_steps
    if world > 1:
        t = torch.tensor([e2e_s], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_s = float(t.item())
    e2e_gflops = 8.0 * E * total_rhs / e2e_s / 1e9
    checksum = float(np.abs(yn).sum())
    # full-size correctness: the e2e result against sampled rows of the dense kernel matrix (real values only)
    dense_check = None
    if build_stats is not None and rank == 0:
        try:
            dense_check = {"rows": 64, "rel_l2_vs_dense_rows": workloads.dense_rows_check(meta, yn[:, 0], xn[:, 0], 64),
                           "what": "bfCudaApplyHost result vs 64 rows of i/4 H0(k|x_i - x_j|) (scipy), tree order; the reference-built operator's own error at N = 131072 is 1.0e-10 and grows with N (DESIGN.md 8c)"}
        except Exception as e:
            dense_check = {"error": str(e)[:200]}

    peak, peak_src = peaks()
    alg_bytes = 16.0 * E + 16.0 * nrhs * (2 * n)  # per apply, per GPU
    # launches of the TIMED applies: 1 when the whole apply is one cooperative launch (bfGemvStage<., single launch>),
    # else one per stage; the per-stage times below come from one extra, profiled apply (always per stage)
    n_launch = max(int(launches // max(args.steps, 1)), 1)
    achieved = alg_bytes / (ms_step * 1e-3) / 1e9
    roof = {
        "bound": "hbm" if nrhs < 16 else "fp64",
        "kernel": ("bfGemvStage<complex, single launch>" if n_launch == 1 else "bfGemvStage<complex>") if nrhs == 1 else "bfWideMmaStage<2>",
        "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "peak_source": peak_src,
        "traffic": None,
        "peak_note": "peak = COPY bandwidth (read + write) measured by the driver; this kernel only reads the pool, and a read-only stream can run a little above a copy's rate, so frac may touch 1.0",
        "algorithmic_bytes_per_launch": alg_bytes / n_launch, "launches_per_step": n_launch,
        "avg_launch_ms": ms_step / n_launch,
        "stage_ms_of_a_per_stage_apply": [round(float(v), 4) for v in stage_ms],
        "stage_gbs": [round(16.0 * float(e) / max(float(m), 1e-9) / 1e6, 1) for e, m in zip(stage_entries, stage_ms)],
    }
    tr = ROOT / "profiles" / "traffic.json"
    if tr.exists():
        try:
            tj = json.loads(tr.read_text())
            # per launch, like `achieved`: the single-launch kernel moves the whole apply, a stage launch its stage
            roof["traffic"] = tj.get("gemv_fused_bytes_per_launch") if n_launch == 1 and n == DEFAULT_N else tj.get("gemv_bytes_per_launch")
        except Exception:
            pass

    extra = None
    if not args.no_extra and nrhs == 1:
        # side measurement: the 256-RHS block apply per GPU on the same operator (configs[2]), against
        # the fp64 ceilings measured in this run on this device (register-resident DFMA / DMMA loops)
        try:
            import ctypes as C

            dfma, dmma = C.c_double(), C.c_double()
            capi.check(capi.load_library().bfCudaMeasureFp64Peak(local_rank, C.byref(dfma), C.byref(dmma)), "bfCudaMeasureFp64Peak")
            dmma_sus = C.c_double()
            with ClockSampler(local_rank) as clk_peak:
                capi.check(capi.load_library().bfCudaMeasureFp64PeakSustained(local_rank, 2.0, C.byref(dmma_sus)), "bfCudaMeasureFp64PeakSustained")
            nr = 256
            X = torch.randn(n, nr, dtype=torch.complex128, device="cuda")
            Y = torch.empty(n, nr, dtype=torch.complex128, device="cuda")
            plan.apply_device(X, Y, stream=stream)
            barrier()
            a0, a1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            with ClockSampler(local_rank) as clk256:
                a0.record()
                reps = 3
                for _ in range(reps):
                    plan.apply_device(X, Y, stream=stream)
                a1.record()
                barrier()
            ms = a0.elapsed_time(a1) / reps
            if world > 1:
                t = torch.tensor([ms], device="cuda", dtype=torch.float64)
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
                ms = float(t.item())
            tf = 8.0 * E * nr / (ms * 1e-3) / 1e12  # per GPU
            # end to end for the block apply: bfCudaApplyHost on pinned host buffers (2 x 16 n nr bytes over PCIe)
            e2e256 = None
            try:
                Xh = torch.empty(n, nr, dtype=torch.complex128).pin_memory()
                Yh = torch.empty(n, nr, dtype=torch.complex128).pin_memory()
                Xh.copy_(X)
                plan.apply_host(Xh.numpy(), capi.OP_N, out=Yh.numpy())
                t0 = time.perf_counter()
                for _ in range(2):
                    plan.apply_host(Xh.numpy(), capi.OP_N, out=Yh.numpy())
                dt = (time.perf_counter() - t0) / 2
                e2e256 = {"ms_per_apply": dt * 1e3, "gflops": 8.0 * E * nr / dt / 1e9, "h2d_bytes_per_step": int(16 * n * nr),
                          "d2h_bytes_per_step": int(16 * n * nr), "api": "bfCudaApplyHost (pinned host buffers)",
                          "matches_device_result": bool(torch.equal(Yh.cuda(), Y))}
                del Xh, Yh
            except Exception as e:
                e2e256 = {"error": str(e)[:200]}
            extra = {"nrhs_per_gpu": nr, "ms_per_apply": ms, "gflops": tf * 1e3 * world, "matvecs_per_s": nr * world / (ms * 1e-3),
                     "launches_per_apply": int(plan.last_launch_count), "e2e": e2e256,
                     "clocks": clk256.summary(),
                     "roofline": {"bound": "fp64 tensor pipe (DMMA)", "kernel": "bfWideMmaStage<2>", "achieved": tf, "peak": dmma_sus.value,
                                  "unit": "TFLOP/s", "frac": tf / dmma_sus.value if dmma_sus.value else None,
                                  "peak_source": "measured in this run, SUSTAINED (bfCudaMeasureFp64PeakSustained: DMMA m8n8k4 loop back to back for 2 s, second half; the apply runs ~1.2 s)",
                                  "peak_burst": dmma.value, "frac_of_burst": tf / dmma.value if dmma.value else None,
                                  "peak_clocks": clk_peak.summary(), "dfma_peak_burst": dfma.value}}
            del X, Y
        except Exception as e:  # e.g. out of memory for the workspace
            extra = {"nrhs_per_gpu": 256, "error": str(e)[:200]}

    lbo = None
    if not args.no_extra and nrhs == 1 and world == 1:
        # side measurement, configs[3]: real fp64 eigenbasis butterfly, forward and transposed 1-RHS apply
        try:
            rawL, metaL = workloads.lbo_like()
            planL = Plan(rawL, device=local_rank, build_transpose=1)
            planL.fill_synthetic(7)
            EL = planL.info["numEntries"]
            lbo = {"workload": metaL["workload"], "structure": metaL["structure"], "E": int(EL), "dtype": "f64"}
            for name, op, n_in, n_out in (("forward", capi.OP_N, rawL.num_cols, rawL.num_rows), ("transpose", capi.OP_T, rawL.num_rows, rawL.num_cols)):
                xv = torch.randn(n_in, 1, dtype=torch.float64, device="cuda")
                yv = torch.empty(n_out, 1, dtype=torch.float64, device="cuda")
                for _ in range(3):
                    planL.apply_device(xv, yv, op=op, stream=stream)
                torch.cuda.synchronize()
                b0, b1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                reps = 20
                b0.record()
                for _ in range(reps):
                    planL.apply_device(xv, yv, op=op, stream=stream)
                b1.record()
                torch.cuda.synchronize()
                ms = b0.elapsed_time(b1) / reps
                byts = 8.0 * EL + 8.0 * (n_in + n_out)
                lbo[name] = {"ms_per_apply": ms, "gflops": 2.0 * EL / (ms * 1e-3) / 1e9, "matvecs_per_s": 1e3 / ms,
                             "gbs": byts / (ms * 1e-3) / 1e9, "frac_of_hbm_peak": byts / (ms * 1e-3) / 1e9 / peaks()[0],
                             "launches": int(planL.last_launch_count)}
            # the real many-RHS kernel (bfWideStage<real>: register-blocked DFMA; the reference has no real block apply at
            # all): 64 columns, against the DFMA ceiling measured on this device
            try:
                import ctypes as C

                dfma, dmma = C.c_double(), C.c_double()
                capi.check(capi.load_library().bfCudaMeasureFp64Peak(local_rank, C.byref(dfma), C.byref(dmma)), "bfCudaMeasureFp64Peak")
                nr = 64
                Xw = torch.randn(rawL.num_cols, nr, dtype=torch.float64, device="cuda")
                Yw = torch.empty(rawL.num_rows, nr, dtype=torch.float64, device="cuda")
                for _ in range(2):
                    planL.apply_device(Xw, Yw, stream=stream)
                torch.cuda.synchronize()
                b0, b1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                b0.record()
                for _ in range(5):
                    planL.apply_device(Xw, Yw, stream=stream)
                b1.record()
                torch.cuda.synchronize()
                ms = b0.elapsed_time(b1) / 5
                tf = 2.0 * EL * nr / (ms * 1e-3) / 1e12
                lbo["forward_64rhs"] = {"ms_per_apply": ms, "tflops": tf, "kernel": "bfWideStage<real> (DFMA)", "dfma_peak_tflops": dfma.value,
                                        "frac_of_dfma_peak": tf / dfma.value if dfma.value else None, "matvecs_per_s": nr / (ms * 1e-3)}
                del Xw, Yw
            except Exception as e:
                lbo["forward_64rhs"] = {"error": str(e)[:200]}
            planL.close()
        except Exception as e:
            lbo = {"error": str(e)[:200]}
        # the same path on a REFERENCE-structured real operator: bfFacStreamer (src/fac_streamer.c:386-518) run offline
        # by tools/make_lbo_fixture.py on sphere points x real spherical harmonics; the fixture carries the flattened
        # operator, seeded operands and the reference's own bfMatMulVec / bfMatRmulVec results and CPU times
        fx = ROOT / "fixtures" / "lbo_sphere"
        if lbo is not None and Path(str(fx) + ".raw.npz").exists() and Path(str(fx) + ".ref.npz").exists():
            try:
                from butterfly_b200.rawops import RawOps

                rawF = RawOps.load(str(fx) + ".raw.npz")
                ref = np.load(str(fx) + ".ref.npz")
                planF = Plan(rawF, device=local_rank, build_transpose=1)
                EF = planF.info["numEntries"]
                blk = {"workload": f"sphere n={int(ref['n'])} x real spherical harmonics l<={int(ref['lmax'])} ({rawF.num_rows} x {rawF.num_cols}), "
                                   f"streamed by the unmodified reference (tol {float(ref['tol']):g})",
                       "structure": "reference bfFacStreamer (fixture built offline by tools/make_lbo_fixture.py)", "E": int(EF), "dtype": "f64",
                       "entries_over_dense": EF / (rawF.num_rows * rawF.num_cols), "reference_build_s": float(ref["reference_build_s"]),
                       "butterfly_vs_dense_rel_l2": float(ref["butterfly_vs_dense_rel_l2"]), "tolerance": 1e-12}
                for name, op, xk, yk, ck in (("forward", capi.OP_N, "X", "Y", "cpu_forward_ms"), ("transpose", capi.OP_T, "XT", "YT", "cpu_transpose_ms")):
                    Xh, Yref = ref[xk], ref[yk]
                    worst = 0.0
                    for j in range(Xh.shape[1]):
                        yj = planF.apply_host(np.ascontiguousarray(Xh[:, j:j + 1]), op)
                        worst = max(worst, float(np.linalg.norm(yj[:, 0] - Yref[:, j]) / np.linalg.norm(Yref[:, j])))
                    Yb = planF.apply_host(np.ascontiguousarray(Xh), op)  # all stored columns as one block (the wide kernel)
                    blockerr = float(np.linalg.norm(Yb - Yref) / np.linalg.norm(Yref))
                    xv = torch.from_numpy(np.ascontiguousarray(Xh[:, :1])).cuda()
                    yv = torch.empty(Yref.shape[0], 1, dtype=torch.float64, device="cuda")
                    for _ in range(5):
                        planF.apply_device(xv, yv, op=op, stream=stream)
                    torch.cuda.synchronize()
                    b0, b1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    reps = 200
                    b0.record()
                    for _ in range(reps):
                        planF.apply_device(xv, yv, op=op, stream=stream)
                    b1.record()
                    torch.cuda.synchronize()
                    ms = b0.elapsed_time(b1) / reps
                    byts = 8.0 * EF + 8.0 * (Xh.shape[0] + Yref.shape[0])
                    t0 = time.perf_counter()
                    for _ in range(20):
                        planF.apply_host(np.ascontiguousarray(Xh[:, :1]), op)
                    host_ms = (time.perf_counter() - t0) / 20 * 1e3
                    blk[name] = {"rel_l2_vs_reference_1rhs_max": worst, "rel_l2_vs_reference_block": blockerr, "vectors": int(Xh.shape[1]),
                                 "device_ms_per_apply": ms, "gbs": byts / (ms * 1e-3) / 1e9, "frac_of_hbm_peak": byts / (ms * 1e-3) / 1e9 / peaks()[0],
                                 "l2_note": "operator smaller than L2: the pool stays L2-resident between applies, the HBM fraction is not a roofline here" if 8.0 * EF < 100e6 else "operator larger than L2",
                                 "host_buffers_ms_per_apply": host_ms, "reference_cpu_ms_per_apply": float(ref[ck]), "cpu_cores": 1,
                                 "speedup_host_buffers_vs_reference": float(ref[ck]) / host_ms, "launches": int(planF.last_launch_count)}
                planF.close()
                lbo["reference_structure"] = blk
            except Exception as e:
                lbo["reference_structure"] = {"error": str(e)[:300]}
        elif lbo is not None:
            lbo["reference_structure"] = {"unavailable": "fixtures/lbo_sphere.{raw,ref}.npz not present (python tools/make_lbo_fixture.py builds them: minutes of the reference's streamer)"}

    gpu_build = None
    if not args.no_extra and nrhs == 1 and world == 1:
        # side measurement, SURVEY §8f-4: the operator's VALUES computed on the GPU (bfCudaMatAttachHelm2) next to
        # the unmodified reference building the same operator on the host, both checked against sampled rows of
        # the dense kernel matrix (the reference's own yardstick, bf_all_blocks.c:149-153)
        try:
            nb = int(os.environ.get("BF_BENCH_BUILD_N", 65536))
            r = subprocess.run([sys.executable, str(ROOT / "tools" / "helm2_gpu_build_check.py"), str(nb), repr(nb / 40.96),
                                "--no-dense", "--sample-rows", "256", "--same-factors"], capture_output=True, text=True, timeout=900)
            gpu_build = json.loads(r.stdout.strip().splitlines()[-1]) if r.returncode == 0 else {"error": (r.stdout + r.stderr)[-300:]}
            par = gpu_build.get("parity")
            if par and max(par["rel_l2_same_factors_1rhs"], par["rel_l2_same_factors_256rhs_max"]) > 1e-12:
                raise SystemExit(f"bench.py: same-factors parity failed at N={nb}: {par}")
            if "reference_build_s" in gpu_build:
                gpu_build["speedup_vs_reference_build"] = gpu_build["reference_build_s"] / gpu_build["gpu_build_s"]
                gpu_build["workload"] = f"helm2_circle N={nb} k={nb / 40.96:g}: layout by the reference's fac_helm2 inside a recording, values by bfHelm2Pinv / bfHelm2Fill"
        except SystemExit:
            raise
        except Exception as e:
            gpu_build = {"error": str(e)[:200]}

    rows_extra = rows_1m = None
    if not args.no_extra and nrhs == 1 and world > 1:
        # side measurement, configs[4]: the SAME operator block-row sharded over the ranks (strong scaling)
        plan.close()
        del x, y
        torch.cuda.empty_cache()
        try:
            # real values; checked against the unsharded plan of the same leaves (fits one GPU at this N); a short
            # GMRES leg so that configs[4]'s caller is measured by the driver too
            rows_extra = run_rows(args, rank, local_rank, world, steps=min(args.steps, 20), warmup=3, verify=True,
                                  solve=args.solve or 10)
        except Exception as e:
            rows_extra = {"error": str(e)[:300]}
        # configs[1] / configs[4] at their own size: 2^20 points, 186 GB of factors — more than one B200 holds, so it
        # only exists block-row sharded (>= 4 GPUs: 46 GB per rank).  Real values, built per rank on the GPU; checked
        # against dense kernel rows (no unsharded plan to compare with: it does not fit).
        if world >= 4 and not args.no_1m:
            try:
                rows_1m = run_rows(args, rank, local_rank, world, n=1 << 20, steps=10, warmup=2, verify=False, solve=0)
            except Exception as e:
                rows_1m = {"error": str(e)[:300]}

    cpu = configs0 = None
    if rank == 0 and world == 1 and not args.no_cpu:
        plan.close()  # free 67 GB before the second operator is attached
        del x, y
        torch.cuda.empty_cache()
        cpu, configs0 = configs0_block(local_rank)
        if cpu is None:
            cpu = cpu_reference_leg(steps=10, warmup=2)
        cpu = {k2: cpu[k2] for k2 in ("value", "unit", "cores", "kind", "sample")}

    if rank == 0:
        k = args.k or n / 40.96
        line = {
            "metric": "butterfly apply GFLOP/s (complex fp64, 1 RHS)" if nrhs == 1 else f"butterfly apply GFLOP/s (complex fp64, {nrhs} RHS per GPU)",
            "value": gflops, "unit": "GFLOP/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "c128 (complex f64)", "data": "synthetic",
            "config": {"workload": workload_label(n, k, nrhs), "structure": meta["structure"],
                       "values": values_note, "value_build": build_stats, "dense_check": dense_check,
                       "E": int(E), "pool_GB": info["poolBytes"] / 1e9, "stages": int(info["numStagesN"]),
                       "parallelism": f"rhs-columns x{world} (plan replicated, no collective)",
                       "l2": "inputs larger than L2 (factor pool streamed once per apply)", "plan_build_s": round(build_s, 1)},
            "matvecs_per_s": total_rhs / (ms_step * 1e-3),
            "clocks": clk.summary(),
            "e2e": {"value": e2e_gflops, "unit": "GFLOP/s", "h2d_bytes_per_step": int(16 * n * nrhs), "d2h_bytes_per_step": int(16 * n * nrhs),
                    "ms_per_step": e2e_s * 1e3, "api": "bfCudaApplyHost (pinned host buffers)", "result_checksum": checksum},
            "gpu_launches": int(launches),
            "roofline": roof,
            "cpu_baseline": cpu,
            "configs0": configs0,
            "rhs256": extra,
            "lbo_real": lbo,
            "helm2_gpu_build": gpu_build,
            "row_shard": rows_extra,
            "row_shard_1M": rows_1m,
        }
        print(json.dumps(line), flush=True)
    plan.close()  # idempotent
    if world > 1:
        dist.destroy_process_group()


def main():
    args = parse()
    rank = int(os.environ.get("RANK", 0))
    local_rank = int(os.environ.get("LOCAL_RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    if args.impl == "reference":
        run_reference(args, rank, world)
    else:
        run_ours(args, rank, local_rank, world)


if __name__ == "__main__":
    main()
