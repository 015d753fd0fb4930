"""Top warp-stall SASS lines of an ncu report (source page).  usage: ncu_top_stalls.py report.ncu-rep [N]"""
import csv, subprocess, sys
out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
h = rows[1]
src, samp, ie = h.index("Source"), h.index("Warp Stall Sampling (All Samples)"), h.index("Instructions Executed")
L = []
for k, r in enumerate(rows[2:]):
    try: L.append((int(r[samp]), k, r[src][:90], r[ie]))
    except (ValueError, IndexError): pass
tot = sum(x[0] for x in L)
print("total samples", tot)
for v, k, s, n in sorted(L, reverse=True)[: int(sys.argv[2]) if len(sys.argv) > 2 else 30]:
    print(f"{v:7d} {100.0 * v / tot:5.1f}%  line {k:5d}  executed {n:>9s}  {s}")
