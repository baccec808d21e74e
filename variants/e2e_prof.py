import os, sys, time, ctypes
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch, oceanpkg
pkg = oceanpkg.load(); L = pkg._lib
cfg = pkg.scene.config(3); o = cfg["ocean"]
stream = torch.cuda.Stream()
bodies = []
for i in range(4):
    b = pkg.WaterBody(windSpeed=o[0], windDirection=(o[1], o[2]), gravity=o[3], fetch=o[4], depth=o[5], texturesSize=cfg["N"], cascades=cfg["cascades"], seed=3 + i).Awake()
    b.SetStream(stream.cuda_stream); b.SetReadbackMode(L.READBACK_HEIGHT); bodies.append(b)
K = 1000
for f in range(100): bodies[f % 4].Update(f / 60)
for b in bodies: b.Sync()
torch.cuda.synchronize(); t0 = time.perf_counter()
for f in range(K): bodies[f % 4].Update(f / 60)
for b in bodies: b.Sync()
torch.cuda.synchronize()
print("OCEAN_DBG=%s: %.2f us/update" % (os.environ.get("OCEAN_DBG", "0"), (time.perf_counter() - t0) / K * 1e6))
