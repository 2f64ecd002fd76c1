#!/usr/bin/env python
"""Summarise ncu output brought back in gpurun_out/ into small tracked files under profiles/.

  python tools/ncu_summarise.py launches gpurun_out/launches_X.csv profiles/X_launches.md
  python tools/ncu_summarise.py full gpurun_out/X.ncu-rep profiles/X   (writes X_metrics.md, X_opcodes.md, X_summary.json)

Runs here (no GPU): `ncu -i ... --page raw/source --csv` only reads the report.
"""
import collections
import csv
import io
import json
import re
import subprocess
import sys

KEY = re.compile(
    r"^(gpu__time_duration\.sum|dram__bytes_(read|write)\.sum|lts__t_bytes\.sum|lts__t_sectors\.sum|lts__t_sector_hit_rate\.pct|"
    r"l1tex__t_sector_hit_rate\.pct|l1tex__t_(sectors|requests)_pipe_lsu_mem_global_op_ld\.sum|launch__registers_per_thread|"
    r"launch__grid_size|launch__block_size|launch__occupancy_limit_(registers|shared_mem|warps)|sm__warps_active\.avg\.pct_of_peak_sustained_active|"
    r"smsp__thread_inst_executed_per_inst_executed\.ratio|smsp__inst_executed\.sum|smsp__issue_active\.avg\.pct_of_peak_sustained_active|"
    r"sm__throughput\.avg\.pct_of_peak_sustained_elapsed|gpu__dram_throughput\.avg\.pct_of_peak_sustained_elapsed|"
    r"lts__throughput\.avg\.pct_of_peak_sustained_elapsed|l1tex__throughput\.avg\.pct_of_peak_sustained_elapsed|"
    r"smsp__warps_eligible\.avg\.per_cycle_active|smsp__average_warp_latency_per_inst_issued\.ratio|"
    r"smsp__inst_executed_pipe_(xu|fma|alu|lsu)\.sum|sm__cycles_elapsed\.max|launch__shared_mem_per_block_dynamic)$")


def to_ms(v, unit):
    v = float(v.replace(",", ""))
    return {"ns": v / 1e6, "nsecond": v / 1e6, "us": v / 1e3, "usecond": v / 1e3, "ms": v, "msecond": v, "s": v * 1e3, "second": v * 1e3}[unit]


def launches(src, dst):
    lines = [l for l in open(src) if not l.startswith("==")]
    agg = collections.OrderedDict()
    for row in csv.DictReader(lines):
        if row.get("Metric Name") != "gpu__time_duration.sum":
            continue
        k = re.sub(r"\(.*", "", row["Kernel Name"])
        a = agg.setdefault(k, [0, 0.0])
        a[0] += 1
        a[1] += to_ms(row["Metric Value"], row["Metric Unit"])
    tot = sum(v[1] for v in agg.values())
    with open(dst, "w") as f:
        f.write(f"ncu launch list `{src}` (gpu__time_duration.sum, --clock-control none; cold-cache, serialised: compare shares)\n\n")
        f.write("| kernel | launches | total ms | share |\n|---|---:|---:|---:|\n")
        for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1]):
            f.write(f"| `{k}` | {v[0]} | {v[1]:.3f} | {v[1] / tot:.3f} |\n")
    print(open(dst).read())


def ncu_csv(rep, page, extra=()):
    out = subprocess.run(["ncu", "-i", rep, "--page", page, "--csv", *extra], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True).stdout
    return list(csv.reader(io.StringIO(out)))


def full(rep, prefix):
    rows = ncu_csv(rep, "raw")
    hdr, units, data = rows[0], rows[1], rows[2:]
    name_i = hdr.index("Kernel Name")
    summary = {"report": rep, "launches": []}
    with open(prefix + "_metrics.md", "w") as f:
        f.write(f"`ncu --set full --clock-control none` capture `{rep}`: one column per captured launch\n\n")
        f.write("kernels: " + "; ".join(d[name_i][:60] for d in data) + "\n\n| metric | unit | " + " | ".join(f"#{i}" for i in range(len(data))) + " |\n")
        f.write("|---|---|" + "---:|" * len(data) + "\n")
        for i, h in enumerate(hdr):
            if KEY.match(h):
                f.write(f"| {h} | {units[i]} | " + " | ".join(d[i] for d in data) + " |\n")
    for d in data:
        g = lambda n: d[hdr.index(n)] if n in hdr else None
        u = lambda n: units[hdr.index(n)]
        rd, wr = float(g("dram__bytes_read.sum")), float(g("dram__bytes_write.sum"))
        scale = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
        summary["launches"].append({
            "kernel": d[name_i][:80], "ms": to_ms(g("gpu__time_duration.sum"), u("gpu__time_duration.sum")),
            "dram_bytes": rd * scale[u("dram__bytes_read.sum")] + wr * scale[u("dram__bytes_write.sum")],
            "threads_per_inst": float(g("smsp__thread_inst_executed_per_inst_executed.ratio")),
            "warp_inst": float(g("smsp__inst_executed.sum")), "regs": int(g("launch__registers_per_thread")),
            "l1_hit_pct": float(g("l1tex__t_sector_hit_rate.pct")), "l2_hit_pct": float(g("lts__t_sector_hit_rate.pct")),
            "issue_active_pct": float(g("smsp__issue_active.avg.pct_of_peak_sustained_active")),
            "warps_active_pct": float(g("sm__warps_active.avg.pct_of_peak_sustained_active")),
            "dram_throughput_pct": float(g("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed"))})
    summary["dram_bytes_per_launch"] = sum(l["dram_bytes"] for l in summary["launches"]) / max(1, len(summary["launches"]))
    json.dump(summary, open(prefix + "_summary.json", "w"), indent=1)
    # opcode mix + stall samples of the longest captured launch
    longest = max(range(len(data)), key=lambda i: summary["launches"][i]["ms"])
    rows = ncu_csv(rep, "source", ("--print-source", "sass", "--launch-skip", str(longest), "--launch-count", "1"))
    hdr = rows[1]
    ia, it, isrc, iss = (hdr.index(n) for n in ("Instructions Executed", "Thread Instructions Executed", "Source", "# Samples"))
    stall_cols = [i for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
    agg = collections.defaultdict(lambda: [0, 0, 0, 0])
    stalls = collections.Counter()
    tot = 0
    for d in rows[2:]:
        if d and d[0] == "Kernel Name":
            break   # the page lists every selected launch one after the other: keep the first
        if len(d) <= it or not d[ia].isdigit():
            continue
        toks = d[isrc].strip().split()
        op = (toks[1] if toks[0].startswith("@") else toks[0]).split(".")[0]
        a = agg[op]
        a[0] += int(d[ia]); a[1] += int(d[it]); a[2] += int(d[iss]); a[3] += 1
        tot += int(d[ia])
        for i in stall_cols:
            if d[i].isdigit():
                stalls[hdr[i]] += int(d[i])
    with open(prefix + "_opcodes.md", "w") as f:
        f.write(f"SASS opcode mix of captured launch #{longest} ({summary['launches'][longest]['kernel']}), from `ncu --page source`\n\n")
        f.write(f"total warp instructions {tot}\n\n| opcode | static | warp inst | share | avg active threads | stall samples |\n|---|---:|---:|---:|---:|---:|\n")
        for k, v in sorted(agg.items(), key=lambda kv: -kv[1][0])[:28]:
            f.write(f"| {k} | {v[3]} | {v[0]} | {v[0] / tot:.3f} | {v[1] / max(v[0], 1):.1f} | {v[2]} |\n")
        st = sum(stalls.values())
        f.write("\nwarp stall samples by reason: " + ", ".join(f"{k} {v / st:.3f}" for k, v in stalls.most_common(8)) + "\n")
    print(open(prefix + "_metrics.md").read())
    print(open(prefix + "_opcodes.md").read())


if __name__ == "__main__":
    {"launches": launches, "full": full}[sys.argv[1]](sys.argv[2], sys.argv[3])
