"""Per-source-line instruction counts and stall samples of one kernel: joins `ncu --page source --csv` (SASS rows, in
address order) with `nvdisasm -g -c` line annotations of the same function.
usage: ncu_lines.py <report.ncu-rep> <kernel regex> <cubin> <mangled function> [min_pct]"""
import csv, re, subprocess, sys
from collections import defaultdict

rep, kre, cubin, func = sys.argv[1:5]
min_pct = float(sys.argv[5]) if len(sys.argv) > 5 else 0.5
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", f"regex:{kre}"], capture_output=True, text=True).stdout
rows = list(csv.reader(src.splitlines()))
h = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
hdr = rows[h]
iI, iS, iSrc = hdr.index("Instructions Executed"), hdr.index("# Samples"), hdr.index("Source")
body = []
for r in rows[h + 1:]:
    if r and r[0] == "Kernel Name":
        break  # first matching launch only
    if len(r) > iI:
        body.append(r)
dis = subprocess.run(["nvdisasm", "-g", "-c", cubin], capture_output=True, text=True).stdout.splitlines()
start = next(i for i, l in enumerate(dis) if l.startswith("//-") and f".text.{func} " in l)
lines, cur = [], ("?", 0)
for l in dis[start + 1:]:
    if l.startswith("//-") and ".text." in l:
        break
    m = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m:
        cur = (m.group(1).split("/")[-1], int(m.group(2)))
        continue
    if re.match(r"\s+/\*[0-9a-f]{4,}\*/", l):
        lines.append(cur)
n = min(len(lines), len(body))
if len(lines) != len(body):
    print(f"# warning: {len(lines)} disassembled vs {len(body)} profiled instructions", file=sys.stderr)
agg = defaultdict(lambda: [0.0, 0.0, 0])
for (f, ln), r in zip(lines[:n], body[:n]):
    a = agg[(f, ln)]
    a[0] += float(r[iI] or 0); a[1] += float(r[iS] or 0); a[2] += 1
tot_i = sum(a[0] for a in agg.values()) or 1
tot_s = sum(a[1] for a in agg.values()) or 1
print(f"# total warp instructions {tot_i:.0f}, samples {tot_s:.0f}, SASS instructions {n}")
for (f, ln), a in sorted(agg.items()):
    if a[0] / tot_i * 100 >= min_pct or a[1] / tot_s * 100 >= min_pct:
        print(f"{f}:{ln:<5d} inst {a[0] / tot_i * 100:5.1f}%  stall {a[1] / tot_s * 100:5.1f}%  sass {a[2]}")
