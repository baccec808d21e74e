"""Per-kernel SASS / resource summary of liboceanfft.so: registers, spills, shared memory (cuobjdump -res-usage) and
counts of the instructions that prove the Blackwell-native path (cuobjdump -sass): TMA (UTMALDG / UTMASTG / UBLKCP),
packed fp32 (FFMA2 / FADD2 / FMUL2), cp.async (LDGSTS), barriers, shuffles, shared-memory and global accesses.

    python variants/sass_summary.py [lib.so] > profiles/r2_sass_summary.txt      (runs here: no GPU needed)
"""
import os
import re
import subprocess
import sys
from collections import Counter, OrderedDict

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "ocean-simulation_b200", "liboceanfft.so")
OPS = ["UTMALDG", "UTMASTG", "UBLKCP", "UTMAPF", "SYNCS", "LDGSTS", "FFMA2", "FADD2", "FMUL2", "FFMA", "FADD", "FMUL", "MUFU",
       "BAR", "SHFL", "LDS", "STS", "LDG", "STG", "LDL", "STL", "ATOMG", "RED", "UTCHMMA", "HMMA"]


def demangle(names):
    out = subprocess.run(["c++filt"], input="\n".join(names), capture_output=True, text=True).stdout.split("\n")
    return dict(zip(names, out))


def main():
    res = subprocess.run(["cuobjdump", "-res-usage", LIB], capture_output=True, text=True).stdout
    usage = OrderedDict()
    cur = None
    for line in res.split("\n"):
        m = re.match(r"\s*Function (\S+):", line)
        if m:
            cur = m.group(1)
            continue
        if cur and "REG:" in line:
            usage[cur] = dict(re.findall(r"(\w+)(?:\[\d+\])?:(\d+)", line))
            cur = None
    sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True).stdout
    counts, total = {}, {}
    cur = None
    for line in sass.split("\n"):
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            cur = m.group(1)
            counts[cur], total[cur] = Counter(), 0
            continue
        m = re.match(r"\s*/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+)", line)
        if cur and m:
            op = m.group(1)
            total[cur] += 1
            for o in OPS:
                if op == o or op.startswith(o + ".") or (o in ("BAR", "SYNCS") and op.startswith(o)):
                    counts[cur][o] += 1
                    break
    names = demangle(list(usage))
    print("SASS / resource summary of %s (cuobjdump -res-usage, cuobjdump -sass; sm_100a)" % os.path.relpath(LIB, ROOT))
    print("STACK > 0 = local memory frame (spills or by-value arrays); LDL / STL = local loads / stores actually emitted\n")
    for fn, u in usage.items():
        c = counts.get(fn, Counter())
        short = re.sub(r"\(.*", "", names.get(fn, fn)).replace("void ", "").replace("ocean::", "")
        print("%s" % short)
        print("    regs %s  stack %s B  static smem %s B  const %s B  | %d SASS instructions" % (
            u.get("REG"), u.get("STACK"), u.get("SHARED"), u.get("CONSTANT", u.get("CONST", "?")), total.get(fn, 0)))
        print("    " + "  ".join("%s %d" % (o, c[o]) for o in OPS if c[o]))
    all_ = Counter()
    for c in counts.values():
        all_.update(c)
    print("\nwhole library: " + "  ".join("%s %d" % (o, all_[o]) for o in OPS if all_[o]))


if __name__ == "__main__":
    main()
