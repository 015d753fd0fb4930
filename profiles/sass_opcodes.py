"""Per-kernel counts of the Blackwell-only SASS opcodes in libgomoku_b200.so (cuobjdump -sass):
UTCHMMA = tcgen05.mma, LDTM / STTM = tcgen05.ld / tcgen05.st, UTCBAR = tcgen05.commit, UBLKCP = cp.async.bulk,
UTCATOMSWS = tensor-memory allocation, SYNCS = mbarrier ops.  usage: sass_opcodes.py [lib.so] > table.md"""
import collections
import re
import subprocess
import sys

so = sys.argv[1] if len(sys.argv) > 1 else "alphazero-gomoku_b200/libgomoku_b200.so"
out = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True).stdout
OPS = ["UTCHMMA", "LDTM", "STTM", "UTCBAR", "UBLKCP", "UTCATOMSWS", "SYNCS", "UTMALDG", "HMMA", "FFMA", "DFMA|DADD|DMUL"]
counts = collections.OrderedDict()
cur = None
for line in out.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        cur = m.group(1)
        counts[cur] = collections.Counter()
        continue
    if cur is None:
        continue
    m = re.search(r"^\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
    if m:
        op = m.group(1).split(".")[0]
        for pat in OPS:
            if re.fullmatch(pat, op):
                counts[cur][pat] += 1
        counts[cur]["_all"] += 1
demangle = subprocess.run(["c++filt"] + list(counts), capture_output=True, text=True).stdout.splitlines()
print("| kernel | instructions | " + " | ".join(o.replace("|", "/") for o in OPS) + " |")
print("|---|---|" + "---|" * len(OPS))
for (name, c), dm in sorted(zip(counts.items(), demangle), key=lambda x: -x[0][1]["UTCHMMA"]):
    short = re.sub(r"\(.*", "", dm).replace("void gz::", "").replace("gz::", "")
    print("| `%s` | %d | " % (short[:70], c["_all"]) + " | ".join(str(c[p]) if c[p] else "" for p in OPS) + " |")
