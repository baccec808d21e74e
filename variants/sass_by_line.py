"""Executed warp-instructions of one kernel by source line: merges `ncu --page source --csv` (per-SASS-instruction counts) with
`nvdisasm -gi` (inlined line info).  python variants/sass_by_line.py <rep.ncu-rep> <kernel regex> <mangled .text name> <cubin>"""
import collections, csv, os, re, subprocess, sys
rep, kre, text, cubin = sys.argv[1:5]
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "-k", "regex:" + kre], capture_output=True, text=True).stdout
rows = list(csv.reader(out.split("\n")))
hi = [i for i, r in enumerate(rows) if r and r[0] == "Address"][0]
h = rows[hi]; iE = h.index("Instructions Executed")
counts = []
for r in rows[hi + 1:]:
    if len(r) != len(h): break
    counts.append(int(r[iE]))
txt = subprocess.run(["nvdisasm", "-gi", "-c", cubin], capture_output=True, text=True).stdout.split("\n")
start = [i for i, l in enumerate(txt) if l.startswith(".text." + text + ":")][0]
seq, cur = [], []
for l in txt[start + 1:]:
    if l.startswith(".text.") or l.startswith("//-----"): break
    m = re.match(r'\s*//## File "([^"]+)", line (\d+)', l)
    if m: cur.append((os.path.basename(m.group(1)), int(m.group(2)))); continue
    if re.match(r"\s*/\*[0-9a-f]{4,}\*/", l): seq.append(list(cur)); cur = []
assert len(seq) == len(counts), (len(seq), len(counts), "report and library are different builds")
agg, last = collections.Counter(), []
for loc, c in zip(seq, counts):
    loc = loc or last; last = loc
    ours = [x for x in loc if os.path.exists(os.path.join(ROOT, "ocean-simulation_b200", "csrc", x[0])) or x[0] == "ocean_noise.h"]
    agg[ours[0] if ours else ("?", 0)] += c
tot = sum(agg.values())
print("executed warp-instructions: %d" % tot)
for (f, ln), c in agg.most_common(int(sys.argv[5]) if len(sys.argv) > 5 else 22):
    path = os.path.join(ROOT, "include" if f == "ocean_noise.h" else "ocean-simulation_b200/csrc", f)
    line = open(path).read().split("\n")[ln - 1].strip()[:100] if os.path.exists(path) else ""
    print("%5.1f%%  %s:%d  %s" % (100.0 * c / tot, f, ln, line))
