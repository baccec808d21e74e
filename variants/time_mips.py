"""Times the consumers at the headline size: mip chains (both arrays), surface sampling, displaced queries."""
import ctypes, sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import oceanpkg
pkg = oceanpkg.load(); L = pkg._lib
cfg = pkg.scene.config(3)
o = cfg["ocean"]
body = pkg.WaterBody(windSpeed=o[0], windDirection=(o[1], o[2]), gravity=o[3], fetch=o[4], depth=o[5], texturesSize=cfg["N"], cascades=cfg["cascades"], seed=3).Awake()
for i in range(3): body.CalculateWavesTexturesAtTime(i/60)
st = torch.cuda.ExternalStream(body.stream)
def timeit(fn, n=200):
    for _ in range(10): fn()
    body.Sync()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with torch.cuda.stream(st):
        e0.record()
        for _ in range(n): fn()
        e1.record()
    e1.synchronize()
    return e0.elapsed_time(e1) / n * 1e3
print("mip chains (2 arrays x 4 layers x 512^2): %.2f us" % timeit(lambda: body.GenerateMips()))
print("frame: %.2f us" % timeit(lambda: body.CalculateWavesTexturesAtTime(0.5)))
L.check(body._lib.ocean_set_auto_mips(body._h, 1))
print("frame + auto mips: %.2f us" % timeit(lambda: body.CalculateWavesTexturesAtTime(0.5)))
L.check(body._lib.ocean_set_auto_mips(body._h, 0))
n = 1 << 20
pts = (torch.rand(n, 2, device="cuda") - 0.5) * 4000
d = torch.empty(n, 3, device="cuda"); nr = torch.empty(n, 3, device="cuda"); f = torch.empty(n, device="cuda"); h = torch.empty(n, device="cuda")
P = lambda t: ctypes.c_void_p(t.data_ptr())
for lod in (0.0, 1.5):
    t = timeit(lambda: L.check(body._lib.ocean_sample_surface(body._h, P(pts), n, lod, P(d), P(nr), P(f), 1)), 50)
    print("sample_surface 1M points lod %.1f: %.1f us (%.2f Gpoints/s)" % (lod, t, n / t / 1e3))
for mode in (L.QUERY_BILINEAR_ALL_CASCADES, L.QUERY_DISPLACED):
    t = timeit(lambda: L.check(body._lib.ocean_query_heights(body._h, P(pts), n, mode, P(h), 1, 1)), 50)
    print("query mode %d 1M points: %.1f us (%.2f Gpoints/s)" % (mode, t, n / t / 1e3))
boxes = torch.empty(n, 6, device="cuda"); boxes[:, 0] = pts[:, 0]; boxes[:, 2] = pts[:, 1]; boxes[:, 1] = (torch.rand(n, device="cuda") - 0.5) * 4; boxes[:, 3:] = 1.0
out4 = torch.empty(n, 4, device="cuda")
for mode in (L.QUERY_NEAREST_REF, L.QUERY_BILINEAR_ALL_CASCADES, L.QUERY_DISPLACED):
    t = timeit(lambda: L.check(body._lib.ocean_buoyancy(body._h, P(boxes), n, mode, 1.0, -9.81, P(out4), 1, 1)), 50)
    print("buoyancy (BO:18-30) mode %d 1M boxes: %.1f us (%.2f Gboxes/s)" % (mode, t, n / t / 1e3))
