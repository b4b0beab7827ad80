"""Development aid: turns the output directory of tests/dev/ncu_r02.sh into the files committed under profiles/.

    python tests/dev/ncu_summarise.py gpurun_out/c27 r02e

-> profiles/<tag>_launches.csv (raw list), <tag>_launches_summary.csv (per-kernel averages and shares of the training step),
   <tag>_ncu_summary.csv (selected metrics of the --set full capture), dram_traffic.json (DRAM bytes per launch)."""
import collections
import csv
import io
import json
import os
import shutil
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
src, tag = sys.argv[1], sys.argv[2]
prof = os.path.join(ROOT, "profiles")
WHAT = "TCNN_B200_GRAPH=0 python bench.py --steps 20 --warmup 5 (cfg1, batch 2^18)"

shutil.copy(os.path.join(src, "launches.csv"), os.path.join(prof, f"{tag}_launches.csv"))
rows = list(csv.DictReader(io.StringIO("".join(l for l in open(os.path.join(src, "launches.csv")) if l.startswith('"')))))
per = collections.defaultdict(list)
for r in rows:
    if r["Metric Name"] == "gpu__time_duration.sum":
        per[r["Kernel Name"].split("(")[0].replace("b200::", "").replace("void ", "")].append(float(r["Metric Value"].replace(",", "")) / 1000)
out = sorted(((n, len(v), sum(v[len(v) // 3:]) / len(v[len(v) // 3:])) for n, v in per.items()), key=lambda t: -t[1] * t[2])
train = lambda n, c: c >= 20 and not n.startswith("mlp_kernel<0") and not n.startswith("at::")  # noqa: E731
total = sum(a for n, c, a in out if train(n, c))
with open(os.path.join(prof, f"{tag}_launches_summary.csv"), "w") as f:
    f.write(f"# ncu --metrics gpu__time_duration.sum --clock-control none -c 400, {WHAT}; per-launch times are cold-cache and serialised: shares, not absolutes.  Raw list: {tag}_launches.csv\n")
    f.write("kernel,launches,avg_us_of_the_last_two_thirds,share_of_training_step\n")
    for n, c, a in out:
        f.write(f'"{n}",{c},{a:.2f},{(a / total if train(n, c) else 0):.3f}\n')

raw = list(csv.reader(open(os.path.join(src, "full_raw.csv"))))
hdr, units, data = raw[0], raw[1], raw[2:]
want = ["Kernel Name", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "launch__block_size", "launch__grid_size", "launch__registers_per_thread",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "smsp__inst_executed.sum", "l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum", "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio", "smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio",
        "TPC.TriageCompute.sm__pipe_tensor_cycles_active_realtime.avg.pct_of_peak_sustained_elapsed"]
idx = [hdr.index(w) for w in want if w in hdr]
with open(os.path.join(prof, f"{tag}_ncu_summary.csv"), "w") as f:
    f.write(f'"# ncu --set full --clock-control none --import-source on, {WHAT}: one launch of each training kernel after 40 skipped launches; cold-cache, serialised (tests/dev/ncu_r02.sh)"\n')
    w = csv.writer(f)
    w.writerow([hdr[i] for i in idx]); w.writerow([units[i] for i in idx])
    for d in data:
        w.writerow([d[i][:70] for i in idx])
k, dr, dw = hdr.index("Kernel Name"), hdr.index("dram__bytes_read.sum"), hdr.index("dram__bytes_write.sum")
traffic = {}
for d in data:
    name = "grid_forward" if "grid_forward" in d[k] else "mlp_fused" if "mlp_kernel" in d[k] else "adam" if d[k].startswith("adam_kernel") else "mlp_finalize"
    traffic[name] = (float(d[dr]) + float(d[dw])) * 1e6
traffic["_source"] = f"ncu --set full, profiles/{tag}_ncu_summary.csv (bytes per launch, dram__bytes_read.sum + dram__bytes_write.sum)"
json.dump(traffic, open(os.path.join(prof, "dram_traffic.json"), "w"), indent=1)
print(open(os.path.join(prof, f"{tag}_launches_summary.csv")).read()[:900])
