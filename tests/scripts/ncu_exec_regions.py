"""Where a kernel's dynamic instructions go: consecutive SASS instructions with the same executed count are grouped into regions
(ncu source page of one captured launch).  usage: python tests/scripts/ncu_exec_regions.py report.ncu-rep <kernel regex> <launch id> [min share %]"""
import csv
import subprocess
import sys

rep, rx, lid = sys.argv[1], sys.argv[2], sys.argv[3]
min_share = float(sys.argv[4]) if len(sys.argv) > 4 else 1.0
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-id", f"::regex:{rx}:{lid}"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hdr = rows[1]
ix = {h: i for i, h in enumerate(hdr)}
data = [r for r in rows[2:] if len(r) == len(hdr) and r[0] != "Address"]
ex = [int(r[ix["Instructions Executed"]]) for r in data]
smp = [int(r[ix["# Samples"]]) for r in data]
tot, tots = sum(ex), sum(smp)
print(f"# {rows[0][1][:110]}\n# {len(data)} SASS instructions, {tot} warp instructions executed, {tots} stall samples")
groups = []
for i, e in enumerate(ex):
    if groups and (groups[-1][2] == e or (e > 0 and abs(groups[-1][2] - e) / e < 0.02)):
        groups[-1][1] = i
    else:
        groups.append([i, i, e])
for a, b, e in groups:
    n = sum(ex[a:b + 1])
    s = sum(smp[a:b + 1])
    if 100.0 * n / tot >= min_share or 100.0 * s / tots >= min_share:
        print(f"sass {a:5d}-{b:5d} ({b - a + 1:4d} instr) x {e:8d} warps = {100.0 * n / tot:5.1f}% of instructions, {100.0 * s / tots:5.1f}% of samples   {data[a][ix['Source']].strip()[:60]}")
