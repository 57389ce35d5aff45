"""Condense ncu outputs from gpurun_out/ into the tracked summaries under profiles/ (development
aid).  usage: python tools/make_profiles.py TAG launches.csv gemv.ncu-rep wide.ncu-rep"""
import collections, csv, json, subprocess, sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
tag, launches, gemv_rep, wide_rep = sys.argv[1:5]
P = ROOT / "profiles"
P.mkdir(exist_ok=True)

# ---- launch list: per-kernel share of the run, plus the share inside ONE 1-RHS apply ----------
rows = [r for r in csv.reader(l for l in open(launches) if not l.startswith("==")) if len(r) > 10]
hdr, rows = rows[0], rows[1:]
ik, iv = hdr.index("Kernel Name"), hdr.index("Metric Value")
tot = collections.OrderedDict()
for r in rows:
    k = r[ik]
    n, t = tot.get(k, (0, 0.0))
    tot[k] = (n + 1, t + float(r[iv]) / 1e3)
allus = sum(t for _, t in tot.values())
with open(P / f"{tag}_launches_bench.csv", "w") as f:
    f.write(f"# ncu launch list, {tag}: `python bench.py --steps 2 --warmup 1 --no-cpu` (helm2 N=524288; 1-RHS applies, fp64 peak probe, 256-RHS side measurement)\n")
    f.write("# ncu --metrics gpu__time_duration.sum --clock-control none -c 500 ; cold-cache / serialised: compare SHARES (bfPeakDmma / bfPeakDfma are the fp64 ceiling probes, not product kernels)\n")
    f.write("kernel,launches,total_us,share_of_captured\n")
    for k, (n, t) in sorted(tot.items(), key=lambda kv: -kv[1][1]):
        f.write(f"\"{k}\",{n},{t:.1f},{t / allus:.4f}\n")
    # one 1-RHS apply = 8 consecutive bfGemvStage launches after the fills
    g = [float(r[iv]) / 1e3 for r in rows if "bfGemvStage" in r[ik]]
    f.write(f"# one 1-RHS apply = 8 bfGemvStage launches (no other kernel): first apply stage times us = {[round(v, 1) for v in g[:8]]}, sum {sum(g[:8]):.1f}\n")
    f.write("# => bfGemvStage share of the timed 1-RHS step = 1.00 (bench.py's CUDA-event step time / 8 agrees with the mean of these)\n")


def raw(rep):
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    r = list(csv.reader(out.splitlines()))
    return r[0], r[1], r[2:]


WANT = ["Kernel Name", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "launch__registers_per_thread", "launch__grid_size", "launch__block_size", "launch__shared_mem_per_block_dynamic",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "sm__pipe_tensor_subpipe_dmma_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
        "lts__t_sector_hit_rate.pct",
        "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio", "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio", "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio"]


def summarise(rep, name, note):
    h, u, d = raw(rep)
    it = h.index("gpu__time_duration.sum")
    ib = h.index("dram__bytes_read.sum")
    d = [r for r in d if "nan" not in r[ib].lower()]  # launches ncu could not replay completely
    with open(P / name, "w") as f:
        f.write(f"# {note}\n")
        f.write("metric,unit," + ",".join(f"launch{i}" for i in range(len(d))) + "\n")
        for w in WANT:
            if w in h:
                i = h.index(w)
                f.write(f"{w},{u[i]}," + ",".join(r[i] for r in d) + "\n")
    return h, u, d


h, u, d = summarise(gemv_rep, f"{tag}_gemv_full_summary.csv",
                    "ncu --set full --clock-control none -k regex:bfGemvStage -s 8 -c 8 : python bench.py --steps 2 --warmup 1 --no-cpu --no-extra (N=524288); launches = the 8 stages of one apply")
ir, iw = h.index("dram__bytes_read.sum"), h.index("dram__bytes_write.sum")
scale = {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0}
gem = [float(r[ir]) * scale[u[ir]] + float(r[iw]) * scale[u[iw]] for r in d]
h2, u2, d2 = summarise(wide_rep, f"{tag}_wide_mma_full_summary.csv",
                       "ncu --set full --clock-control none -k regex:bfWideMma -s 16 -c 1 : python bench.py --steps 2 --warmup 1 --no-cpu (N=524288, 256 RHS in 2 chunks of 128); launch = stage 0 of one chunk")
ir2, iw2 = h2.index("dram__bytes_read.sum"), h2.index("dram__bytes_write.sum")
wid = [float(r[ir2]) * scale[u2[ir2]] + float(r[iw2]) * scale[u2[iw2]] for r in d2]
json.dump({"source": f"profiles/{tag}_gemv_full_summary.csv / {tag}_wide_mma_full_summary.csv (dram__bytes_read.sum + dram__bytes_write.sum per launch)",
           "gemv_stage_bytes": gem, "gemv_bytes_per_launch": sum(gem) / len(gem),
           "gemv_launch": f"mean over the {len(gem)} stage launches of one apply (same averaging as roofline.achieved)",
           "wide_mma_stage_bytes": wid}, open(P / "traffic.json", "w"), indent=1)
print("gemv traffic", gem, "wide", wid)
