#!/usr/bin/env python
"""Per-source-line attribution of an ncu capture.

ncu's source page is per SASS instruction; this joins it with `nvdisasm -g` line markers of the same
cubin so samples / executed instructions / stall reasons can be summed per file:line.

  python tools/ncu_lines.py REPORT.ncu-rep build/orlg_w4.o KERNEL_MANGLED_SUBSTRING [--by inst|samples|long_sb|idle_lanes] [--top N]

`lanes` = average active threads of the line's instructions, `idle%` = the line's share of all idle lane slots
(32 x instructions - thread instructions): where divergence costs.
"""
import argparse
import collections
import csv
import io
import os
import re
import subprocess
import tempfile


def disasm_lines(obj, kernel):
    tmp = tempfile.mkdtemp()
    subprocess.check_call(["cuobjdump", "-xelf", "all", os.path.abspath(obj)], cwd=tmp, stdout=subprocess.DEVNULL)
    cubin = [f for f in os.listdir(tmp) if f.endswith(".cubin")][0]
    txt = subprocess.run(["nvdisasm", "-g", "-c", os.path.join(tmp, cubin)], capture_output=True, text=True).stdout
    out, cur, inside = {}, ("?", 0), False
    for ln in txt.splitlines():
        if ln.startswith("//---") and ".text." in ln:
            inside = kernel in ln
            continue
        if not inside:
            continue
        m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
        if m:
            cur = (os.path.basename(m.group(1)), int(m.group(2)))
            continue
        m = re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+(.*?);", ln)
        if m:
            out[int(m.group(1), 16)] = (cur, m.group(2).strip())
    return out


def ncu_rows(report, kernel_idx):
    if report.endswith(".csv"):  # the source page exported on the GPU box (the .ncu-rep itself is too big to bring back)
        txt = open(report).read()
    else:
        txt = subprocess.run(["ncu", "-i", report, "--page", "source", "--csv", "--print-source", "sass"],
                             capture_output=True, text=True).stdout
    sections, cur = [], None
    for row in csv.reader(io.StringIO(txt)):
        if row and row[0] == "Kernel Name":
            cur = {"name": row[1], "hdr": None, "rows": []}
            sections.append(cur)
        elif cur is not None and cur["hdr"] is None:
            cur["hdr"] = row
        elif cur is not None and row:
            cur["rows"].append(row)
    return sections[kernel_idx]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("report")
    ap.add_argument("obj")
    ap.add_argument("kernel")
    ap.add_argument("--launch", type=int, default=0)
    ap.add_argument("--by", default="samples")
    ap.add_argument("--top", type=int, default=40)
    a = ap.parse_args()
    lines = disasm_lines(a.obj, a.kernel)
    sec = ncu_rows(a.report, a.launch)
    h = {n: i for i, n in enumerate(sec["hdr"])}
    base = None
    agg = collections.defaultdict(lambda: collections.Counter())
    tot = collections.Counter()
    for r in sec["rows"]:
        addr = int(r[h["Address"]], 16) if r[h["Address"]].startswith("0x") else int(r[h["Address"]])
        if base is None:
            base = addr
        key, _ = lines.get(addr - base, (("?", 0), ""))
        vals = {"samples": r[h["# Samples"]], "inst": r[h["Instructions Executed"]],
                "thread_inst": r[h["Thread Instructions Executed"]],
                "long_sb": r[h["stall_long_sb"]], "wait": r[h["stall_wait"]], "short_sb": r[h["stall_short_sb"]],
                "no_inst": r[h["stall_no_inst"]], "branch": r[h["stall_branch_resolving"]]}
        for k, v in vals.items():
            v = float(v.replace(",", "") or 0)
            agg[key][k] += v
            tot[k] += v
    for c in list(agg.values()) + [tot]:  # lane slots an instruction issued for but no thread used
        c["idle_lanes"] = 32 * c["inst"] - c["thread_inst"]
    print(sec["name"])
    print("totals:", {k: int(v) for k, v in tot.items()})
    print(f"average active threads per instruction: {tot['thread_inst'] / max(tot['inst'], 1):.2f}")
    print(f"{'file:line':34s} {'samples%':>8s} {'inst%':>7s} {'long_sb':>8s} {'wait':>6s} {'short':>6s} {'noinst':>6s} {'lanes':>6s} {'idle%':>6s}")
    for key, c in sorted(agg.items(), key=lambda kv: -kv[1][a.by])[: a.top]:
        print(f"{key[0] + ':' + str(key[1]):34s} {100 * c['samples'] / max(tot['samples'], 1):8.2f} "
              f"{100 * c['inst'] / max(tot['inst'], 1):7.2f} {int(c['long_sb']):8d} {int(c['wait']):6d} "
              f"{int(c['short_sb']):6d} {int(c['no_inst']):6d} {c['thread_inst'] / max(c['inst'], 1):6.1f} "
              f"{100 * c['idle_lanes'] / max(tot['idle_lanes'], 1):6.2f}")


if __name__ == "__main__":
    main()
