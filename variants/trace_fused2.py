"""Per-item timeline of the dynamically scheduled fused frame kernel (-DOCEAN_TRACE build, OCEAN_FUSED=2):
   python variants/trace_fused2.py [-D...]      (GPU box; prints a summary of the LAST frame of a pipelined run)"""
import ctypes, os, sys
import numpy as np
HERE = os.path.dirname(os.path.abspath(__file__)); ROOT = os.path.dirname(HERE); sys.path.insert(0, ROOT)
import importlib.util
spec = importlib.util.spec_from_file_location("trace_frame", os.path.join(HERE, "trace_frame.py")); tf = importlib.util.module_from_spec(spec); spec.loader.exec_module(tf)

def main():
    if "--build-only" in sys.argv:
        tf.build(); return
    if not os.path.exists(tf.LIB): tf.build()
    os.environ["OCEANFFT_LIB"] = tf.LIB
    os.environ["OCEAN_FUSED"] = "2"
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
    n = 1024
    g = times[0, :n, :, 0].astype(np.float64)
    sm, cta = smid[0, :n], smid[1, :n]
    t0 = g[:, 5][g[:, 5] > 0].min()
    g = (g - t0) / 1e3
    wait, start, mid, bin_, end = g[:, 5], g[:, 0], g[:, 2], g[:, 3], g[:, 4]
    role = np.arange(n) // 512
    layer = (np.arange(n) % 512) // 128
    print("items traced: %d; frame span (first wait start -> last item end): %.1f us" % ((start > -1e6).sum(), end.max()))
    for r, nm in ((0, "R"), (1, "C")):
        for l in range(4):
            m = (role == r) & (layer == l)
            print("%s%d | wait begins %.1f..%.1f (med %.1f) | input in %.1f..%.1f (med %.1f) | A staged med %.1f | end %.1f..%.1f (med %.1f)" % (
                nm, l, wait[m].min(), wait[m].max(), np.median(wait[m]), start[m].min(), start[m].max(), np.median(start[m]),
                np.median(mid[m]), end[m].min(), end[m].max(), np.median(end[m])))
    for r, nm in ((0, "rows"), (1, "cols")):
        m = role == r
        print("%s: wait for first input med %.2f p90 %.2f | compute (input in -> B staged) med %.2f p90 %.2f | first half med %.2f | second half med %.2f" % (
            nm, np.median((start - wait)[m]), np.percentile((start - wait)[m], 90), np.median((end - start)[m]), np.percentile((end - start)[m], 90),
            np.median((mid - start)[m]), np.median((end - mid)[m])))
    m = role == 1
    print("cols: A staged -> B tile in med %.2f (wait for the second tile)" % np.median((bin_ - mid)[m]))
    # per CTA: number of items, first wait start, last end, busy fraction
    ids = np.unique(cta)
    cnt = np.array([(cta == c).sum() for c in ids])
    first = np.array([wait[cta == c].min() for c in ids]); last = np.array([end[cta == c].max() for c in ids])
    busy = np.array([(end - start)[cta == c].sum() for c in ids])
    print("CTAs seen: %d; items per CTA min %d med %d max %d; CTA first wait begins %.1f..%.1f (med %.1f); CTA last end %.1f..%.1f (med %.1f)" % (
        len(ids), cnt.min(), np.median(cnt), cnt.max(), first.min(), first.max(), np.median(first), last.min(), last.max(), np.median(last)))
    print("per CTA: lifetime (first wait -> last end) med %.1f us, compute (sum of input-in -> B staged) med %.1f us, waiting med %.1f us" % (
        np.median(last - first), np.median(busy), np.median(last - first - busy)))
    print("SMs hosting items: %d" % len(np.unique(sm)))

if __name__ == "__main__":
    main()
