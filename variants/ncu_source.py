"""Per-instruction view of an `ncu --page source --csv` export: stall totals, hottest SASS lines, opcode mix."""
import collections, csv, sys
rows = list(csv.reader(open(sys.argv[1])))
h = rows[1]
data = []
for r in rows[2:]:
    if r and r[0] == 'Kernel Name':
        break  # first kernel section only
    if len(r) == len(h) and r[0] != 'Address':
        data.append(r)
iS, iSrc, iEx = h.index('# Samples'), h.index('Source'), h.index('Instructions Executed')
st = [i for i, x in enumerate(h) if x.startswith('stall_') and 'Not Issued' not in x]
tot = sum(int(r[iS]) for r in data)
print("samples", tot, "SASS lines", len(data), "warp-inst executed", sum(int(r[iEx]) for r in data))
agg = {h[i]: sum(int(r[i]) for r in data) for i in st}
print("stalls:", ", ".join("%s %.1f%%" % (k[6:], 100.0 * v / tot) for k, v in sorted(agg.items(), key=lambda x: -x[1])[:9]))
ntop = int(sys.argv[2]) if len(sys.argv) > 2 else 40
top = sorted(range(len(data)), key=lambda i: -int(data[i][iS]))[:ntop]
for i in sorted(top):
    r = data[i]
    s = sorted(((h[j][6:], int(r[j])) for j in st if int(r[j]) > 0), key=lambda x: -x[1])[:3]
    print("%5d %6s  %-72s %s" % (i, r[iS], r[iSrc].strip()[:72], s))
oc = collections.Counter()
for r in data:
    t = r[iSrc].split()
    op = t[1] if t[0].startswith('@') else t[0]
    oc[op.split('.')[0]] += int(r[iEx])
t = sum(oc.values())
print("opcode mix (%% of executed warp-inst):", [(k, round(100.0 * v / t, 1)) for k, v in oc.most_common(24)])
