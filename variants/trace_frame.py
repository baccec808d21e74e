"""Per-CTA phase timeline of one frame (rows + cols) from a -DOCEAN_TRACE build of the library.
   python variants/trace_frame.py            # builds variants/lib_trace.so, runs, prints the timeline (GPU box)"""
import ctypes, os, subprocess, sys
import numpy as np
HERE = os.path.dirname(os.path.abspath(__file__)); ROOT = os.path.dirname(HERE); sys.path.insert(0, ROOT)
LIB = os.path.join(HERE, "lib_trace.so")
EXTRA = [a for a in sys.argv[1:] if a.startswith("-D")]

def build():
    import oceanpkg
    bm = oceanpkg.build_module(); bm.build()
    obj = os.path.join(HERE, "frame_trace.o")
    subprocess.check_call(["nvcc", "-ccbin", "g++"] + bm.ARCH + [f for f in bm.COMMON if f not in ("-Xptxas", "-v")] + ["-DOCEAN_TRACE"] + EXTRA + ["-c", os.path.join(bm.CSRC, "ocean_frame.cu"), "-o", obj])
    others = [os.path.join(bm.OBJ, o) for o in ("ocean_init.o", "ocean_host.o", "ocean_surface.o")]
    subprocess.check_call(["nvcc", "-ccbin", "g++"] + bm.ARCH + ["-shared", "-o", LIB, obj] + others)
    os.remove(obj)

def main():
    if "--build-only" in sys.argv:
        build(); return
    if not os.path.exists(LIB):
        build()
    os.environ["OCEANFFT_LIB"] = LIB
    import torch, oceanpkg
    pkg = oceanpkg.load(); lib = pkg._lib.load()
    cfg = pkg.scene.config(3); o = cfg["ocean"]; N = cfg["N"]
    st = torch.cuda.Stream(); torch.cuda.set_stream(st)
    bodies = []
    for i in range(4):
        b = pkg.WaterBody(windSpeed=o[0], windDirection=(o[1], o[2]), gravity=o[3], fetch=o[4], depth=o[5], texturesSize=N, cascades=cfg["cascades"], seed=3 + i).Awake()
        b.SetStream(st.cuda_stream); bodies.append(b)
    for f in range(24): bodies[f % 4].CalculateWavesTexturesAtTime(f / 60.0)
    torch.cuda.synchronize()
    if "--standalone" in sys.argv:  # one row pass and one column pass, each alone on the GPU
        bodies[0].StepPartial(0.5, 1); torch.cuda.synchronize()
        bodies[0].StepPartial(0.5, 2); torch.cuda.synchronize()
    CT, SL = 8192, 10
    times = np.zeros((2, CT, SL, 2), np.uint64); smid = np.zeros((2, CT), np.uint32)
    lib.ocean_debug_trace_copy.argtypes = [ctypes.c_void_p, ctypes.c_void_p]
    assert lib.ocean_debug_trace_copy(times.ctypes.data, smid.ctypes.data) == 0
    g = times[..., 0].astype(np.float64)
    names = [["start", "loaded", "evolved", "fftA", "pdl_wait", "storedA", "fftB", "end"],
             ["start", "pdl_wait", "A arrived", "fftA done", "disp stored", "B arrived", "fftB done", "tile end", "cta end"]]
    nr = int((g[0, :, 0] > 0).sum()); nc = int((g[1, :, 2] > 0).sum())
    t0 = g[0, :nr, 0].min()
    print("rows CTAs traced:", nr, " cols tiles traced:", nc, " (times in us from the first row CTA start; last frame)")
    for k, n in ((0, nr), (1, nc)):
        print("kernel", "rows" if k == 0 else "cols")
        for s, name in enumerate(names[k]):
            v = g[k, :n, s]; v = v[v > 0]
            if v.size == 0: continue
            v = (v - t0) / 1e3
            print("  %-12s n=%5d  min %7.2f  p10 %7.2f  med %7.2f  p90 %7.2f  max %7.2f" % (name, v.size, v.min(), np.percentile(v, 10), np.median(v), np.percentile(v, 90), v.max()))
    # per-CTA phase durations (clock64 on the same SM), rows
    c = times[..., 1].astype(np.float64)
    d = np.diff(c[0, :nr, :8], axis=1) / 1.965e3
    print("rows per-CTA phase durations (us, by clock64 @1.965 GHz): " + ", ".join("%s %.2f" % (names[0][i + 1], np.median(d[:, i])) for i in range(7)))
    first = g[1, :148, 2] > 0
    d = np.diff(c[1, :nc, 2:8], axis=1) / 1.965e3
    print("cols per-tile phase durations (us): " + ", ".join("%s %.2f" % (names[1][i + 3], np.median(d[:, i])) for i in range(5)))
    # per-layer hand-off: last row CTA end vs first / last "A arrived" of the layer's column tiles
    rows_per_layer = nr // 4
    tiles_per_layer = nc // 4
    for L in range(4):
        re = (g[0, L * rows_per_layer:(L + 1) * rows_per_layer, 7] - t0) / 1e3
        rs = (g[0, L * rows_per_layer:(L + 1) * rows_per_layer, 0] - t0) / 1e3
        ca = (g[1, L * tiles_per_layer:(L + 1) * tiles_per_layer, 2] - t0) / 1e3
        ce = (g[1, L * tiles_per_layer:(L + 1) * tiles_per_layer, 7] - t0) / 1e3
        print("  layer %d: rows start %.1f..%.1f end %.1f..%.1f | cols A arrived %.1f..%.1f  tile end %.1f..%.1f" % (L, rs.min(), rs.max(), re.min(), re.max(), ca.min(), ca.max(), ce.min(), ce.max()))
    # CTAs per SM
    sm = smid[0, :nr]; print("rows CTAs per SM: min %d max %d" % (np.bincount(sm).min(), np.bincount(sm).max()))
    # CTA dispatch: start times of the row CTAs of a few SMs, in order
    t0 = g[0, :nr, 0]; t0 = (t0 - t0.min()) / 1e3
    for s_ in (0, 1, 73, 147):
        v = np.sort(t0[sm == s_]); print("  SM %3d row CTA starts (us):" % s_, " ".join("%.2f" % x for x in v))
    te = g[0, :nr, 7]; te = (te - g[0, :nr, 0].min()) / 1e3
    print("  rows kernel span: first start 0.00, last start %.2f, first end %.2f, last end %.2f" % (t0.max(), te.min(), te.max()))

if __name__ == "__main__":
    main()
