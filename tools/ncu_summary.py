#!/usr/bin/env python
"""Summarise an .ncu-rep (from `ncu --set full --import-source on`) into a small text file
for profiles/: headline metrics, stall reasons, opcode mix and the hottest SASS lines.

usage: tools/ncu_summary.py gpurun_out/prof.ncu-rep profiles/r01_eval_cfg2.txt "<command profiled>"
"""
import collections
import csv
import io
import subprocess
import sys

KEYS = ["gpu__time_duration.sum", "launch__grid_size", "launch__block_size",
        "launch__registers_per_thread", "launch__shared_mem_per_block_dynamic",
        "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem",
        "sm__warps_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
        "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "dram__throughput.avg.pct_of_peak_sustained_elapsed",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "sm__cycles_elapsed.max",
        "lts__t_bytes.sum", "sm__throughput.avg.pct_of_peak_sustained_elapsed"]


def page(rep, name):
    out = subprocess.run(["ncu", "-i", rep, "--page", name, "--csv"], capture_output=True,
                         text=True).stdout
    return list(csv.reader(io.StringIO(out)))


def main():
    rep, dst = sys.argv[1], sys.argv[2]
    cmd = sys.argv[3] if len(sys.argv) > 3 else ""
    lines = [f"ncu summary of {rep}", f"command: {cmd}", ""]
    raw = page(rep, "raw")
    hdr, units = raw[0], raw[1]
    for row in raw[2:]:
        name = row[hdr.index("Kernel Name")] if "Kernel Name" in hdr else "?"
        lines.append(f"== kernel: {name}")
        for h, u, v in zip(hdr, units, row):
            if h in KEYS:
                lines.append(f"  {h:72s} {v} {u}")
        lines.append("  stall reasons (warps per issue-active cycle):")
        for h, u, v in zip(hdr, units, row):
            if h.startswith("smsp__average_warps_issue_stalled_") and h.endswith("_per_issue_active.ratio"):
                try:
                    if float(v) >= 0.02:
                        lines.append(f"    {h[len('smsp__average_warps_issue_stalled_'):-len('_per_issue_active.ratio')]:24s} {float(v):.3f}")
                except ValueError:
                    pass
    src = page(rep, "source")
    # one section per profiled launch: ["Kernel Name", name], header row, data rows
    sections, cur = [], None
    for row in src:
        if row and row[0] == "Kernel Name":
            cur = {"name": row[1] if len(row) > 1 else "?", "rows": []}
            sections.append(cur)
        elif cur is not None:
            cur["rows"].append(row)
    for sec in sections:
        if len(sec["rows"]) < 2 or "# Samples" not in sec["rows"][0]:
            continue
        h = sec["rows"][0]
        si, ie = h.index("# Samples"), h.index("Instructions Executed")
        data = [r for r in sec["rows"][1:] if len(r) > max(si, ie)]
        tot = sum(int(r[si]) for r in data) or 1
        ti = sum(int(r[ie]) for r in data) or 1
        ops_s, ops_i = collections.Counter(), collections.Counter()
        for r in data:
            t = r[1].strip().split()
            if not t:
                continue
            op = (t[1] if t[0].startswith("@") and len(t) > 1 else t[0]).split(".")[0]
            ops_s[op] += int(r[si])
            ops_i[op] += int(r[ie])
        lines += ["", f"== source view: {sec['name']}",
                  f"opcode mix (warp instructions executed {ti}, stall samples {tot}):"]
        for op, n in ops_i.most_common(14):
            lines.append(f"  {op:8s} {100 * n / ti:5.1f}% of instructions   {100 * ops_s[op] / tot:5.1f}% of samples")
        lines += ["", "hottest SASS lines by stall samples:"]
        for i in sorted(sorted(range(len(data)), key=lambda i: -int(data[i][si]))[:25]):
            r = data[i]
            lines.append(f"  {r[1].strip()[:64]:64s} samples {r[si]:>7s}  executed {r[ie]}")
    open(dst, "w").write("\n".join(lines) + "\n")
    print("\n".join(lines[:40]))


if __name__ == "__main__":
    main()
