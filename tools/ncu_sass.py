"""Print the SASS of one kernel from an ncu report with executed-instruction counts and stall samples.
    python tools/ncu_sass.py rep.ncu-rep regex:blend_forward [min_share]
"""
import csv, io, subprocess, sys
rep, kern = sys.argv[1], sys.argv[2]
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", kern], capture_output=True, text=True).stdout
lines = raw.splitlines()
# first kernel only
start = [i for i, l in enumerate(lines) if l.startswith('"Kernel Name"')]
body = lines[start[0] + 1:(start[1] if len(start) > 1 else len(lines))]
rows = list(csv.reader(io.StringIO("\n".join(body))))
h = rows[0]
ci, cs, csmp = h.index("Instructions Executed"), h.index("Source"), h.index("# Samples")
tot = sum(int(r[ci]) for r in rows[1:])
tots = sum(int(r[csmp]) for r in rows[1:])
print(f"total warp instructions {tot}  samples {tots}")
for n, r in enumerate(rows[1:]):
    print(f"{n:4d} {int(r[ci]):10d} {100*int(r[ci])/tot:5.2f}% smp {100*int(r[csmp])/max(tots,1):5.2f}%  {r[cs].strip()}")
