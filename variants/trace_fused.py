"""Per-item timeline of the fused frame kernel (-DOCEAN_TRACE build): python variants/trace_fused.py [-D...]"""
import ctypes, os, subprocess, sys
import numpy as np
HERE = os.path.dirname(os.path.abspath(__file__)); ROOT = os.path.dirname(HERE); sys.path.insert(0, ROOT)
sys.argv = [sys.argv[0]] + [a for a in sys.argv[1:]]
import importlib.util
spec = importlib.util.spec_from_file_location("trace_frame", os.path.join(HERE, "trace_frame.py")); tf = importlib.util.module_from_spec(spec); spec.loader.exec_module(tf)

def main():
    if "--build-only" in sys.argv:
        tf.build(); return
    if not os.path.exists(tf.LIB): tf.build()
    os.environ["OCEANFFT_LIB"] = tf.LIB
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
    CT, SL = 8192, 10
    times = np.zeros((2, CT, SL, 2), np.uint64); smid = np.zeros((2, CT), np.uint32)
    lib.ocean_debug_trace_copy.argtypes = [ctypes.c_void_p, ctypes.c_void_p]
    assert lib.ocean_debug_trace_copy(times.ctypes.data, smid.ctypes.data) == 0
    g = times[0, :1024, :, 0].astype(np.float64); info = smid[1, :1024]
    t0 = g[:, 0][g[:, 0] > 0].min()
    g = (g - t0) / 1e3
    names = ["item start", "inputs in", "evolved", "A staged", "B in", "B staged"]
    for grp in range(8):
        sl = slice(grp * 128, (grp + 1) * 128)
        role, layer = info[sl][0] // 16, info[sl][0] % 16
        assert (info[sl] == info[sl][0]).all()
        line = "%s%d " % ("R" if role == 0 else "C", layer)
        for s in (0, 1, 2, 3, 4, 5):
            v = g[sl, s]; v = v[v > -1e6]
            if role == 0 and s == 4: continue
            if role == 1 and s == 2: continue
            line += "| %s %.1f..%.1f (med %.1f) " % (names[s], v.min(), v.max(), np.median(v))
        print(line)
    # per-item durations by role
    for role in (0, 1):
        m = (info // 16) == role
        d_wait = g[m, 1] - g[m, 0]; d_tot = g[m, 5] - g[m, 0]
        print("role %s: wait for inputs med %.2f p90 %.2f | item total med %.2f p90 %.2f | compute (inputs->B staged) med %.2f" % (
            "rows" if role == 0 else "cols", np.median(d_wait), np.percentile(d_wait, 90), np.median(d_tot), np.percentile(d_tot, 90), np.median(g[m, 5] - g[m, 1])))
    if True:
        m = (info // 16) == 1
        print("cols: fftA %.2f  wait B %.2f  fftB+epilogue %.2f" % (np.median(g[m, 3] - g[m, 1]), np.median(g[m, 4] - g[m, 3]), np.median(g[m, 5] - g[m, 4])))
        m = (info // 16) == 0
        print("rows: evolve %.2f  fftA+stage %.2f  fftB+stage %.2f" % (np.median(g[m, 2] - g[m, 1]), np.median(g[m, 3] - g[m, 2]), np.median(g[m, 5] - g[m, 3])))
    print("frame span (first item start -> last B staged): %.1f us" % g[:, 5].max())

if __name__ == "__main__":
    main()
