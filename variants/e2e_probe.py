import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch, oceanpkg
pkg = oceanpkg.load(); L = pkg._lib
cfg = pkg.scene.config(3); o = cfg["ocean"]
stream = torch.cuda.Stream()
bodies = []
for i in range(4):
    b = pkg.WaterBody(windSpeed=o[0], windDirection=(o[1], o[2]), gravity=o[3], fetch=o[4], depth=o[5], texturesSize=cfg["N"], cascades=cfg["cascades"], seed=3 + i).Awake()
    b.SetStream(stream.cuda_stream); bodies.append(b)
K = int(os.environ.get("K", 300))
def wall(fn):
    for f in range(50): fn(f)
    for b in bodies: b.Sync()
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for f in range(K): fn(f)
    t1 = time.perf_counter()
    for b in bodies: b.Sync()
    torch.cuda.synchronize(); t2 = time.perf_counter()
    return (t1 - t0) / K * 1e6, (t2 - t0) / K * 1e6
print("step only              : enqueue %.1f us/step, total %.1f us/step" % wall(lambda f: bodies[f % 4].CalculateWavesTexturesAtTime(f / 60)))
for mode, name in ((L.READBACK_HEIGHT, "update, height readback"), (L.READBACK_RGBA, "update, rgba readback  ")):
    for b in bodies: b.SetReadbackMode(mode)
    print("%s: enqueue %.1f us/step, total %.1f us/step" % ((name,) + wall(lambda f: bodies[f % 4].Update(f / 60))))
# raw D2H of 1 MiB and 4 MiB from a side stream
for mb in (1, 4):
    src = torch.empty(mb << 20, dtype=torch.uint8, device="cuda"); dst = torch.empty(mb << 20, dtype=torch.uint8).pin_memory()
    s2 = torch.cuda.Stream()
    with torch.cuda.stream(s2):
        for _ in range(20): dst.copy_(src, non_blocking=True)
        s2.synchronize(); t0 = time.perf_counter()
        for _ in range(500): dst.copy_(src, non_blocking=True)
        s2.synchronize(); dt = (time.perf_counter() - t0) / 500
    print("raw D2H %d MiB: %.1f us per copy (%.1f GB/s)" % (mb, dt * 1e6, (mb << 20) / dt / 1e9))
