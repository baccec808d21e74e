import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch, oceanpkg
pkg = oceanpkg.load(); cfg = pkg.scene.config(3); o = cfg["ocean"]
b = pkg.WaterBody(windSpeed=o[0], windDirection=(o[1], o[2]), gravity=o[3], fetch=o[4], depth=o[5], texturesSize=cfg["N"], cascades=cfg["cascades"], seed=3).Awake()
ts = []
for i in range(20):
    torch.cuda.synchronize(); t0 = time.perf_counter(); b.CalculateInitialSpectrumTextures(); ts.append(time.perf_counter() - t0)
print("init spectrum 4 x 512^2 (host wall, incl. sync + H2D of consts): median %.1f us, min %.1f us" % (np.median(ts) * 1e6, min(ts) * 1e6))
print("h0 checksum", float(np.abs(b.initialSpectrumTextures.astype(np.float64)).sum()))
