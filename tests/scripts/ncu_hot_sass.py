"""Top SASS instructions by warp-stall samples from an ncu report's source page (one kernel).
usage: python tests/scripts/ncu_hot_sass.py report.ncu-rep <kernel regex> <launch id> [top N]"""
import csv
import subprocess
import sys

rep, rx, lid = sys.argv[1], sys.argv[2], sys.argv[3]
top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-id", f"::regex:{rx}:{lid}"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hdr = rows[1]
ix = {h: i for i, h in enumerate(hdr)}
data = [r for r in rows[2:] if len(r) == len(hdr) and r[0] != "Address"]
tot = sum(int(r[ix["# Samples"]]) for r in data)
print(f"# {rows[0][1][:120]}\n# {len(data)} SASS instructions, {tot} stall samples")
stall_cols = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
order = sorted(range(len(data)), key=lambda i: -int(data[i][ix["# Samples"]]))[:top]
for i in sorted(order):
    r = data[i]
    n = int(r[ix["# Samples"]])
    why = sorted(((int(r[ix[c]]), c) for c in stall_cols if r[ix[c]].isdigit() and int(r[ix[c]]) > 0), reverse=True)[:3]
    print(f"{i:5d} {100.0*n/tot:5.1f}%  exec={r[ix['Instructions Executed']]:>9s}  {r[ix['Source']].strip()[:70]:70s} {' '.join(f'{c[6:]}={v}' for v, c in why)}")
