"""Per-kernel register / stack / shared-memory usage of libOptB200.so (cuobjdump -res-usage), demangled, sorted by
stack bytes then registers.  Run here (no GPU needed): python tools/resource_usage.py > profiles/<round>_res_usage.txt
A non-zero STACK on a hot kernel means spills or a local array: look at it before spending GPU time."""
import os
import re
import subprocess
import sys

LIB = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "opt_b200", "libOptB200.so")


def main():
    out = subprocess.run(["cuobjdump", "-res-usage", LIB], capture_output=True, text=True, check=True).stdout
    names, rows = [], []
    for m in re.finditer(r"Function (\S+):\s*\n\s*REG:(\d+) STACK:(\d+) SHARED:(\d+) LOCAL:(\d+)", out):
        names.append(m.group(1))
        rows.append(tuple(int(m.group(i)) for i in range(2, 6)))
    dem = subprocess.run(["c++filt"], input="\n".join(names), capture_output=True, text=True, check=True).stdout.splitlines()
    pat = sys.argv[1] if len(sys.argv) > 1 else ""
    table = sorted(zip(dem, rows), key=lambda t: (-t[1][1], -t[1][0], t[0]))
    seen = set()
    print("%5s %6s %7s  kernel" % ("regs", "stack", "shared"))
    for d, (reg, stack, shared, _local) in table:
        if "cub::" in d or (pat and not re.search(pat, d)):
            continue
        d = re.sub(r"\)\s*$", "", re.sub(r"\((?!anonymous)[^()]*(\([^()]*\)[^()]*)*$", "", d))
        line = "%5d %6d %7d  %s" % (reg, stack, shared, d[:170])
        if line not in seen:
            seen.add(line)
            print(line)


if __name__ == "__main__":
    main()
