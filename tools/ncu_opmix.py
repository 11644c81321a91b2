"""Opcode mix and hottest SASS lines of one kernel from an ncu capture's source page:
    python tools/ncu_opmix.py gpurun_out/<capture>.ncu-rep [top_n]"""
import collections
import csv
import io
import subprocess
import sys

rep = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 25
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr = rows[1]
ci, ei, si = hdr.index("Source"), hdr.index("Instructions Executed"), hdr.index("Warp Stall Sampling (All Samples)")
ops, tot, lines = collections.Counter(), 0.0, []
for r in rows[2:]:
    try:
        n = float(r[ei])
    except (ValueError, IndexError):
        continue
    toks = r[ci].split()
    op = (toks[1] if toks[0].startswith("@") else toks[0]).split(".")[0]
    ops[op] += n
    tot += n
    lines.append((float(r[si] or 0), n, r[ci].strip()))
print("warp instructions executed: %d" % tot)
for k, v in ops.most_common(top):
    print("  %-10s %12d  %5.1f %%" % (k, v, 100 * v / tot))
print("hottest lines by stall samples:")
for s, n, src in sorted(lines, reverse=True)[:top]:
    print("  %7d samples  %10d exec  %s" % (s, n, src))
