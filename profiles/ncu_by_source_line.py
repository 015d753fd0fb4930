"""Instructions executed / stall samples of one kernel of an ncu report, aggregated by CUDA source line (inlined
callees count at their own file:line).  The report's SASS page carries no line info in CSV form, so the mapping is
taken from `nvdisasm -g` of the cubin inside the shipped library.
usage: ncu_by_source_line.py report.ncu-rep kernel_name cubin [N]"""
import collections, csv, re, subprocess, sys
rep, kern, cubin = sys.argv[1:4]
N = int(sys.argv[4]) if len(sys.argv) > 4 else 40
dis = subprocess.run(["nvdisasm", "-g", "-c", cubin], capture_output=True, text=True).stdout
line_of, cur, infn, fn_re = {}, None, False, re.compile(r"\.text\.(\S+):")
for l in dis.splitlines():
    m = re.match(r"\s*//## File \"([^\"]+)\", line (\d+)", l)
    if m:
        cur = (m.group(1).split("/")[-1], int(m.group(2)))
        continue
    m = re.match(r"\s*\.section\s+\.text\.(\S+?),", l)
    if m:
        infn = kern in m.group(1)
        continue
    m = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(.*?);", l)
    if m and infn:
        line_of[int(m.group(1), 16)] = cur
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass", "--kernel-name",
                      "regex:" + kern, "--launch-count", "1"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
h = rows[1]
ia, ie, isamp = h.index("Address"), h.index("Instructions Executed"), h.index("Warp Stall Sampling (All Samples)")
base = int(rows[2][ia], 16)
agg = collections.defaultdict(lambda: [0, 0])
for r in rows[2:]:
    try:
        off = int(r[ia], 16) - base
        a = agg[line_of.get(off)]
        a[0] += int(r[ie]); a[1] += int(r[isamp])
    except (ValueError, IndexError):
        pass
tot = sum(a[0] for a in agg.values()); ts = sum(a[1] for a in agg.values())
print("total warp instructions %d, stall samples %d" % (tot, ts))
for k, a in sorted(agg.items(), key=lambda x: -x[1][0])[:N]:
    print("%6.2f %% instr %6.2f %% samples  %s" % (100.0 * a[0] / tot, 100.0 * a[1] / max(ts, 1), k))
