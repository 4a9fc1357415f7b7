#!/usr/bin/env python
"""Condenses the ncu outputs of tools/profile_bench.sh into profiles/<tag>_summary.md:
the launch list per kernel (count, total / mean us, share of the timed launches) and the headline metrics of the full
capture.   python tools/summarize_profile.py <tag> [kernel substring] [raw csv of the full capture]"""
import collections
import csv
import json
import os
import sys

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
tag = sys.argv[1]
kern = sys.argv[2] if len(sys.argv) > 2 else "qot_kernel"
G = os.path.join(REPO, "gpurun_out")
rows = [r for r in csv.reader(open(os.path.join(G, f"{tag}_launches.csv"))) if len(r) > 14 and r[0].isdigit()]
agg = collections.OrderedDict()
for r in rows:
    name = r[4].split("(")[0]
    a = agg.setdefault(name, [0, 0.0])
    a[0] += 1
    a[1] += float(r[14]) / 1e3
tot = sum(a[1] for a in agg.values())
out = [f"# {tag}: ncu launch list and full capture of `{kern}` (bench.py default workload)", "",
       f"`ncu --metrics gpu__time_duration.sum --clock-control none -c 400` over `bench.py --steps 5 --warmup 3 "
       f"--no-cpu-baseline --no-workloads --e2e-steps 2` ({len(rows)} launches; times are cold-cache and serialised - "
       "shares, not absolutes).", "", "| kernel | launches | total us | mean us | share |", "|---|---:|---:|---:|---:|"]
for name, (n, us) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    out.append(f"| `{name}` | {n} | {us:.1f} | {us / n:.1f} | {100 * us / tot:.1f} % |")
rawp = sys.argv[3] if len(sys.argv) > 3 else os.path.join(G, f"{tag}_kernel_raw.csv")  # optional: another capture
raw = list(csv.reader(open(rawp)))
hdr = raw[0]
want = ["gpu__time_duration.sum", "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
        "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem", "launch__occupancy_limit_blocks",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
        "smsp__thread_inst_executed_per_inst_executed.ratio", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed.avg.per_cycle_elapsed", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "dram__throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_sector_hit_rate.pct", "l1tex__t_sector_hit_rate.pct",
        "lts__throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__throughput.avg.pct_of_peak_sustained_elapsed",
        "sass__inst_executed_local_loads", "sass__inst_executed_local_stores", "memory_l2_theoretical_sectors_local"]
for r in raw[2:]:
    d = dict(zip(hdr, r))
    if kern in d.get("Kernel Name", ""):
        out += ["", f"## `{d['Kernel Name']}` (`ncu --set full`, {os.path.basename(rawp)})", "", "| metric | value | unit |", "|---|---:|---|"]
        units = dict(zip(hdr, raw[1]))
        for k in want:
            if k in d:
                out.append(f"| {k} | {d[k]} | {units.get(k, '')} |")
        out += ["", "| stall reason (warps per issue) | value |", "|---|---:|"]
        st = [(k, float(v)) for k, v in d.items() if k.startswith("smsp__average_warps_issue_stalled_") and
              k.endswith("_per_issue_active.ratio") and "not_issued" not in k and v]
        for k, v in sorted(st, key=lambda kv: -kv[1])[:9]:
            out.append(f"| {k[len('smsp__average_warps_issue_stalled_'):-len('_per_issue_active.ratio')]} | {v:.2f} |")
        break
p = os.path.join(G, f"{tag}_prof_plain.json")
if os.path.exists(p):
    try:
        d = json.loads(open(p).read().strip().splitlines()[-1])
        out += ["", f"Plain run of the same command (no profiler): value {d['value']:.4g} {d['unit']}, "
                    f"{d['ms_per_step']:.4f} ms per step, dominant kernel {d['roofline']['kernel_ms']:.4f} ms "
                    f"({100 * d['roofline']['kernel_ms'] / d['ms_per_step']:.0f} % of the step; the launch list above must "
                    "agree on the share)."]
    except Exception as ex:
        out.append(f"(plain run unreadable: {ex})")
open(os.path.join(REPO, "profiles", f"{tag}_summary.md"), "w").write("\n".join(out) + "\n")
print("\n".join(out))
