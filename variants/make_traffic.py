"""profiles/traffic.json from ncu CSVs, stamped with the hash of the frame-kernel sources they were measured on (bench.py
refuses a stale file):
    python variants/make_traffic.py <warm.csv> <full_raw.csv> <tag>
 warm.csv     : ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum --cache-control none --clock-control none --csv over the
                bench loop (one pass, no replay: caches as the program leaves them) -> mean per launch of each frame kernel
 full_raw.csv : ncu --set full ... --page raw --csv (cold: L2 flushed before every launch, launches serialised)"""
import csv, json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import bench

def scale(v, unit):
    u = unit.lower()
    return float(v.replace(",", "")) * {"byte": 1, "kbyte": 1e3, "mbyte": 1e6, "gbyte": 1e9}.get(u, 1)

def warm(path):
    rows = [r for r in csv.reader(l for l in open(path) if l.startswith('"'))]
    hdr = rows[0]; k, m, u, v = (hdr.index(x) for x in ("Kernel Name", "Metric Name", "Metric Unit", "Metric Value"))
    acc = {}
    for r in rows[1:]:
        kind = "rows" if "ocean_rows" in r[k] else ("cols" if "ocean_cols" in r[k] else None)
        if kind and r[m] in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
            acc.setdefault((kind, r[m]), []).append(scale(r[v], r[u]))
    out = {}
    for kind in ("rows", "cols"):
        rd, wr = acc[(kind, "dram__bytes_read.sum")], acc[(kind, "dram__bytes_write.sum")]
        n = len(rd); rd, wr = rd[n // 4:], wr[n // 4:]   # drop the first quarter (cache warm-up of the profiled run)
        out[kind] = {"dram_read_bytes": sum(rd) / len(rd), "dram_write_bytes": sum(wr) / len(wr), "launches": len(rd)}
    return out

def cold(path):
    rows = list(csv.reader(open(path)))
    hdr, units = rows[0], rows[1]
    col = {n: hdr.index(n) for n in ("Kernel Name", "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__time_duration.sum")}
    acc = {}
    for r in rows[2:]:
        kind = "rows" if "ocean_rows" in r[col["Kernel Name"]] else ("cols" if "ocean_cols" in r[col["Kernel Name"]] else None)
        if kind:
            acc.setdefault(kind, []).append((scale(r[col["dram__bytes_read.sum"]], units[col["dram__bytes_read.sum"]]),
                                             scale(r[col["dram__bytes_write.sum"]], units[col["dram__bytes_write.sum"]]),
                                             float(r[col["gpu__time_duration.sum"]])))
    return {k: {"dram_read_bytes": sum(x[0] for x in v) / len(v), "dram_write_bytes": sum(x[1] for x in v) / len(v),
                "duration": sum(x[2] for x in v) / len(v), "duration_unit": units[col["gpu__time_duration.sum"]], "launches": len(v)} for k, v in acc.items()}

w, c, tag = warm(sys.argv[1]), cold(sys.argv[2]), sys.argv[3]
frame = sum(w[k][d] for k in ("rows", "cols") for d in ("dram_read_bytes", "dram_write_bytes"))
out = {"frame_dram_bytes": frame, "source_hash": bench.kernel_source_hash(), "tag": tag,
       "how": "one-pass ncu --cache-control none over the bench loop (profiles/%s_warm_traffic.csv), mean per launch, rows + cols" % tag,
       "warm": w,
       "cold": dict(c, note="ncu --set full: L2 flushed before every launch, launches serialised (profiles/%s_ncu_full_summary.txt): the column "
                            "pass re-reads the workspace from HBM and most of its output is still dirty in L2 when the launch ends" % tag)}
json.dump(out, open(os.path.join(ROOT, "profiles", "traffic.json"), "w"), indent=1)
print(json.dumps({k: out[k] for k in ("frame_dram_bytes", "source_hash")}))
