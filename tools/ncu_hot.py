"""Hottest SASS instructions (warp-stall samples) of one kernel of an ncu report.
    python tools/ncu_hot.py report.ncu-rep kernel-regex [top]"""
import csv, io, subprocess, sys
rep, rx = sys.argv[1], sys.argv[2]
top_n = int(sys.argv[3]) if len(sys.argv) > 3 else 40
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", "regex:" + rx],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr = next(r for r in rows if "Source" in r and "# Samples" in r)
ci = {h: i for i, h in enumerate(hdr)}
data, seen = [], False
for r in rows:
    if r is hdr or (len(r) == len(hdr) and r[ci["# Samples"]] == "# Samples"):
        if seen and data:
            break          # first kernel instance only
        seen = True
        continue
    if seen and len(r) == len(hdr) and r[ci["# Samples"]].isdigit():
        data.append(r)
tot = sum(int(r[ci["# Samples"]]) for r in data) or 1
print("instructions", len(data), "samples", tot, "warp-instr executed", sum(int(r[ci["Instructions Executed"]]) for r in data))
reg = [(k, int(r[ci["# Samples"]]), r[ci["Source"]].strip(), int(r[ci["Instructions Executed"]])) for k, r in enumerate(data)]
for k, s, src, n in sorted(sorted(reg, key=lambda x: -x[1])[:top_n]):
    print(f"{k:5d} {s:6d} {100 * s / tot:5.1f}% exec {n:8d}  {src[:110]}")
