#!/usr/bin/env python
"""Per-source-line share of warp instructions / active threads / stall samples of one captured launch (reads a .ncu-rep here, no GPU)."""
import csv, io, subprocess, sys
rep, launch = sys.argv[1], sys.argv[2] if len(sys.argv) > 2 else "0"
top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass", "--launch-skip", launch, "--launch-count", "1"],
                     stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True).stdout
cur, rows, tot, smp = None, [], 0, 0
for d in csv.reader(io.StringIO(out)):
    if len(d) >= 2 and d[0] == "File Path":
        cur = d[1].split("/")[-1]
    elif len(d) > 20 and d[0].isdigit() and d[2] == "-":
        n, t, s = int(d[7]), int(d[8]), int(d[6])
        rows.append((n, t, s, cur, int(d[0]), d[1].strip()[:100])); tot += n; smp += s
print(f"total warp instructions {tot}, stall samples {smp}")
for n, t, s, f, l, src in sorted(rows, reverse=True)[:top]:
    print(f"{n / tot:6.3f} thr={t / max(n, 1):5.1f} smp={s / max(smp, 1):6.3f} {f}:{l}  {src}")
